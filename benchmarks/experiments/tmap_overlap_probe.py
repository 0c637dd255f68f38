"""Does cuTensorMapEncodeTiled accept a tensor map whose dimension-1 stride (32 B) is smaller than the extent of
dimension 0 (128 B)?  That is the 'overlapping windows' view of a space-to-depth image the stem convolution wants:
row (n, oh, ow) of the im2col matrix for filter row r' is the 128 contiguous bytes starting at pixel (oh + r', ow)."""
import torch
from cuda.bindings import driver as cu

torch.cuda.init()
x = torch.zeros(4, 115, 115, 16, dtype=torch.bfloat16, device="cuda")
U64, U32 = cu.cuuint64_t, cu.cuuint32_t
dims = [U64(64), U64(112), U64(115), U64(4)]
strides = [U64(32), U64(115 * 32), U64(115 * 115 * 32)]
box = [U32(64), U32(16), U32(8), U32(1)]
es = [U32(1)] * 4
res = cu.cuTensorMapEncodeTiled(cu.CUtensorMapDataType.CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, x.data_ptr(), dims, strides,
                                box, es, cu.CUtensorMapInterleave.CU_TENSOR_MAP_INTERLEAVE_NONE,
                                cu.CUtensorMapSwizzle.CU_TENSOR_MAP_SWIZZLE_128B,
                                cu.CUtensorMapL2promotion.CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                cu.CUtensorMapFloatOOBfill.CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)
print("overlapping-window map:", res[0])
