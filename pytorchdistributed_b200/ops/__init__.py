from . import functional, reference
from .functional import (affine_act, as_nhwc_bf16, batch_norm_act, conv2d, conv_bn_act, cross_entropy,
                         global_avg_pool, linear, max_pool_3x3_s2, small_linear, stem_input, to_nhwc4, to_s2d)

__all__ = ["functional", "reference", "affine_act", "as_nhwc_bf16", "batch_norm_act", "conv2d", "conv_bn_act",
           "cross_entropy", "global_avg_pool", "linear", "max_pool_3x3_s2", "small_linear", "stem_input", "to_nhwc4", "to_s2d"]
