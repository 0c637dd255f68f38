// Symmetric memory on the CUDA virtual-memory-management API, with an NVSwitch multicast (NVLS) mapping.
//
// A "block" is the same amount of physical device memory on every rank of a group (one process per GPU):
//   * cuMemCreate'd with a POSIX-file-descriptor export handle; every process imports every peer's handle and maps it,
//     so kernels address all N copies with plain ld/st (the "unicast" mappings) - what csrc/p2p.cu's flag barriers
//     and two-shot collectives use;
//   * optionally bound to ONE multicast object shared by the group (cuMulticastCreate / AddDevice / BindMem) and mapped
//     a second time: a `multimem.ld_reduce` on that mapping makes the switch fetch the element from all N copies and
//     return their sum, a `multimem.st` stores to all N copies - the in-switch reduction NCCL calls NVLS.
// File descriptors travel between the processes over a unix socket (SCM_RIGHTS, pytorchdistributed_b200/comm/p2p.py);
// this file only creates, exports, imports, binds and maps.  Every step can fail on a box whose driver, fabric
// manager or container does not expose the feature: the caller then falls back (multicast -> unicast VMM -> CUDA IPC)
// and reports which path it took.
//
// The driver entry points are fetched at run time (cudaGetDriverEntryPoint), so the library still links without
// libcuda and loads on a CPU-only box.
//
// Stands in for what the reference gets from NCCL's own NVLS / P2P buffer registration behind
// init_process_group("nccl") (R/02 DDP基本概念/ddp_gpus.py:22).
#include <string.h>
#include <unistd.h>

#include <deque>
#include <mutex>

#include "common.cuh"
#include "ptdb200.h"

namespace {

constexpr int MAX_WORLD = 8;

struct Driver {
    CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
    CUresult (*MemRelease)(CUmemGenericAllocationHandle);
    CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
    CUresult (*MemAddressFree)(CUdeviceptr, size_t);
    CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
    CUresult (*MemUnmap)(CUdeviceptr, size_t);
    CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
    CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
    CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType,
                                           unsigned long long);
    CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
    CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
    CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
    CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                                 unsigned long long);
    CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t);
    CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
    CUresult (*DeviceGet)(CUdevice*, int);
    CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
    CUresult (*GetErrorString)(CUresult, const char**);
    bool ok = false;
    bool mc_ok = false;
};

template <class F>
bool fetch(const char* name, F& fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess ||
        p == nullptr) {
        cudaGetLastError();
        fn = nullptr;
        return false;
    }
    fn = reinterpret_cast<F>(p);
    return true;
}

const Driver& drv() {
    static Driver d;
    static std::once_flag once;
    std::call_once(once, [] {
        bool ok = true;
        ok &= fetch("cuMemCreate", d.MemCreate);
        ok &= fetch("cuMemRelease", d.MemRelease);
        ok &= fetch("cuMemAddressReserve", d.MemAddressReserve);
        ok &= fetch("cuMemAddressFree", d.MemAddressFree);
        ok &= fetch("cuMemMap", d.MemMap);
        ok &= fetch("cuMemUnmap", d.MemUnmap);
        ok &= fetch("cuMemSetAccess", d.MemSetAccess);
        ok &= fetch("cuMemGetAllocationGranularity", d.MemGetAllocationGranularity);
        ok &= fetch("cuMemExportToShareableHandle", d.MemExportToShareableHandle);
        ok &= fetch("cuMemImportFromShareableHandle", d.MemImportFromShareableHandle);
        ok &= fetch("cuDeviceGet", d.DeviceGet);
        ok &= fetch("cuDeviceGetAttribute", d.DeviceGetAttribute);
        ok &= fetch("cuGetErrorString", d.GetErrorString);
        d.ok = ok;
        bool mc = ok;
        mc &= fetch("cuMulticastCreate", d.MulticastCreate);
        mc &= fetch("cuMulticastAddDevice", d.MulticastAddDevice);
        mc &= fetch("cuMulticastBindMem", d.MulticastBindMem);
        mc &= fetch("cuMulticastUnbind", d.MulticastUnbind);
        mc &= fetch("cuMulticastGetGranularity", d.MulticastGetGranularity);
        d.mc_ok = mc;
    });
    return d;
}

const char* cu_str(CUresult r) {
    const char* s = nullptr;
    if (drv().GetErrorString && drv().GetErrorString(r, &s) == CUDA_SUCCESS && s) return s;
    return "unknown CUDA driver error";
}

#define PTD_CU_CHECK(expr)                                                                          \
    do {                                                                                            \
        CUresult _r = (expr);                                                                       \
        if (_r != CUDA_SUCCESS) {                                                                   \
            ptd_set_error("%s:%d: %s -> %d (%s)", __FILE__, __LINE__, #expr, (int)_r, cu_str(_r)); \
            return PTD_ERR_CUDA;                                                                    \
        }                                                                                           \
    } while (0)

struct Block {
    bool live = false;
    int rank = 0, world = 1, dev = 0;
    CUdevice cudev = 0;
    size_t bytes = 0;                                       // padded size, identical on every rank
    CUmemGenericAllocationHandle local = 0;
    CUmemGenericAllocationHandle peer[MAX_WORLD] = {0};     // imported handles (peer[rank] unused)
    CUdeviceptr uc[MAX_WORLD] = {0};                        // this process's mapping of rank r's memory
    CUmemGenericAllocationHandle mc_handle = 0;
    CUdeviceptr mc = 0;                                     // multicast mapping (0 = none)
    bool mc_bound = false;
};

std::mutex g_mu;
std::deque<Block> g_blocks(1);      // id 0 is never handed out; a deque: references stay valid while blocks are added

Block* get(int id) {
    if (id <= 0 || id >= (int)g_blocks.size() || !g_blocks[id].live) {
        ptd_set_error("vmm: unknown block id %d", id);
        return nullptr;
    }
    return &g_blocks[id];
}

CUmemAllocationProp alloc_prop(int dev) {
    CUmemAllocationProp prop;
    memset(&prop, 0, sizeof(prop));
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = dev;
    prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return prop;
}

int map_rw(CUdeviceptr* va, size_t bytes, size_t align, CUmemGenericAllocationHandle h, int dev) {
    const Driver& d = drv();
    PTD_CU_CHECK(d.MemAddressReserve(va, bytes, align, 0, 0));
    CUresult r = d.MemMap(*va, bytes, 0, h, 0);
    if (r != CUDA_SUCCESS) {
        d.MemAddressFree(*va, bytes);
        *va = 0;
        ptd_set_error("vmm: cuMemMap failed: %d (%s)", (int)r, cu_str(r));
        return PTD_ERR_CUDA;
    }
    CUmemAccessDesc acc;
    memset(&acc, 0, sizeof(acc));
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = dev;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    r = d.MemSetAccess(*va, bytes, &acc, 1);
    if (r != CUDA_SUCCESS) {
        d.MemUnmap(*va, bytes);
        d.MemAddressFree(*va, bytes);
        *va = 0;
        ptd_set_error("vmm: cuMemSetAccess failed: %d (%s)", (int)r, cu_str(r));
        return PTD_ERR_CUDA;
    }
    return PTD_OK;
}

void unmap(CUdeviceptr& va, size_t bytes) {
    if (va == 0) return;
    drv().MemUnmap(va, bytes);
    drv().MemAddressFree(va, bytes);
    va = 0;
}

}  // namespace

// Capability bits of the current device: 1 = VMM allocations with POSIX-fd export, 2 = multicast objects.
extern "C" int ptd_vmm_caps(void) {
    const Driver& d = drv();
    if (!d.ok) return 0;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return 0; }
    if (cudaFree(nullptr) != cudaSuccess) { cudaGetLastError(); return 0; }      // primary context exists
    CUdevice cudev;
    if (d.DeviceGet(&cudev, dev) != CUDA_SUCCESS) return 0;
    int vmm = 0, fd = 0, mc = 0;
    d.DeviceGetAttribute(&vmm, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, cudev);
    d.DeviceGetAttribute(&fd, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, cudev);
    if (d.mc_ok) d.DeviceGetAttribute(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cudev);
    int caps = 0;
    if (vmm && fd) caps |= 1;
    if (vmm && fd && mc) caps |= 2;
    return caps;
}

// Allocate `bytes` (rounded up to the allocation and multicast granularity; the rounded size is the same on every
// rank because it only depends on `bytes`, `world` and the device model) of zeroed memory on the current device, map
// it, and export it as a file descriptor (*fd_out; the caller sends it to the peers and closes it).  Returns a block
// id > 0 or a negative error.
extern "C" int ptd_vmm_create(long long bytes, int rank, int world, long long* padded_out, int* fd_out) {
    const Driver& d = drv();
    PTD_REQUIRE(d.ok, "vmm: CUDA driver VMM entry points not available");
    PTD_REQUIRE(bytes > 0 && world >= 1 && world <= MAX_WORLD && rank >= 0 && rank < world,
                "vmm_create: bad arguments (bytes %lld rank %d world %d)", bytes, rank, world);
    int dev = 0;
    PTD_CUDA_CHECK(cudaGetDevice(&dev));
    PTD_CUDA_CHECK(cudaFree(nullptr));
    CUdevice cudev;
    PTD_CU_CHECK(d.DeviceGet(&cudev, dev));
    CUmemAllocationProp prop = alloc_prop(dev);
    size_t gran = 0;
    PTD_CU_CHECK(d.MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
    if (d.mc_ok && (ptd_vmm_caps() & 2)) {
        CUmulticastObjectProp mp;
        memset(&mp, 0, sizeof(mp));
        mp.numDevices = (unsigned)world;
        mp.size = (size_t)bytes;
        mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
        size_t mg = 0;
        if (d.MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > gran)
            gran = mg;
    }
    if (gran == 0) gran = 2u << 20;
    const size_t padded = ((size_t)bytes + gran - 1) / gran * gran;
    Block b;
    b.rank = rank; b.world = world; b.dev = dev; b.cudev = cudev; b.bytes = padded;
    PTD_CU_CHECK(d.MemCreate(&b.local, padded, &prop, 0));
    int rc = map_rw(&b.uc[rank], padded, gran, b.local, dev);
    if (rc) { d.MemRelease(b.local); return rc; }
    if (cudaMemset((void*)b.uc[rank], 0, padded) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) {
        ptd_set_error("vmm_create: clearing the new block failed: %s", cudaGetErrorString(cudaGetLastError()));
        unmap(b.uc[rank], padded);
        d.MemRelease(b.local);
        return PTD_ERR_CUDA;
    }
    int fd = -1;
    CUresult r = d.MemExportToShareableHandle(&fd, b.local, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) {
        ptd_set_error("vmm_create: cuMemExportToShareableHandle failed: %d (%s)", (int)r, cu_str(r));
        unmap(b.uc[rank], padded);
        d.MemRelease(b.local);
        return PTD_ERR_CUDA;
    }
    b.live = true;
    std::lock_guard<std::mutex> lk(g_mu);
    g_blocks.push_back(b);
    if (padded_out) *padded_out = (long long)padded;
    if (fd_out) *fd_out = fd;
    return (int)g_blocks.size() - 1;
}

// Import rank `peer`'s allocation from the file descriptor it exported and map it read-write for this device.
extern "C" int ptd_vmm_import_peer(int id, int peer, int fd) {
    const Driver& d = drv();
    std::lock_guard<std::mutex> lk(g_mu);
    Block* b = get(id);
    if (!b) return PTD_ERR_ARG;
    PTD_REQUIRE(peer >= 0 && peer < b->world && peer != b->rank && b->uc[peer] == 0, "vmm_import_peer: bad peer %d", peer);
    PTD_CU_CHECK(d.MemImportFromShareableHandle(&b->peer[peer], (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    int rc = map_rw(&b->uc[peer], b->bytes, 2u << 20, b->peer[peer], b->dev);
    if (rc) { d.MemRelease(b->peer[peer]); b->peer[peer] = 0; }
    return rc;
}

// Multicast set-up, in this order, with a group barrier between the steps (the driver blocks a bind until every
// device of the team has been added):
//   rank 0: ptd_vmm_mc_create -> fd to every peer;  peers: ptd_vmm_mc_import;  all: ptd_vmm_mc_add_device;
//   barrier;  all: ptd_vmm_mc_bind (binds the local memory at offset 0 and maps the multicast range).
extern "C" int ptd_vmm_mc_create(int id, int* fd_out) {
    const Driver& d = drv();
    PTD_REQUIRE(d.mc_ok, "vmm: multicast entry points not available");
    std::lock_guard<std::mutex> lk(g_mu);
    Block* b = get(id);
    if (!b) return PTD_ERR_ARG;
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = (unsigned)b->world;
    mp.size = b->bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    PTD_CU_CHECK(d.MulticastCreate(&b->mc_handle, &mp));
    int fd = -1;
    CUresult r = d.MemExportToShareableHandle(&fd, b->mc_handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) {
        d.MemRelease(b->mc_handle);
        b->mc_handle = 0;
        ptd_set_error("vmm_mc_create: export failed: %d (%s)", (int)r, cu_str(r));
        return PTD_ERR_CUDA;
    }
    *fd_out = fd;
    return PTD_OK;
}

extern "C" int ptd_vmm_mc_import(int id, int fd) {
    const Driver& d = drv();
    PTD_REQUIRE(d.mc_ok, "vmm: multicast entry points not available");
    std::lock_guard<std::mutex> lk(g_mu);
    Block* b = get(id);
    if (!b) return PTD_ERR_ARG;
    PTD_CU_CHECK(d.MemImportFromShareableHandle(&b->mc_handle, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
    return PTD_OK;
}

extern "C" int ptd_vmm_mc_add_device(int id) {
    const Driver& d = drv();
    std::lock_guard<std::mutex> lk(g_mu);
    Block* b = get(id);
    if (!b) return PTD_ERR_ARG;
    PTD_REQUIRE(b->mc_handle != 0, "vmm_mc_add_device: no multicast object on block %d", id);
    PTD_CU_CHECK(d.MulticastAddDevice(b->mc_handle, b->cudev));
    return PTD_OK;
}

extern "C" int ptd_vmm_mc_bind(int id) {
    const Driver& d = drv();
    Block* b;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        b = get(id);
    }
    if (!b) return PTD_ERR_ARG;
    PTD_REQUIRE(b->mc_handle != 0, "vmm_mc_bind: no multicast object on block %d", id);
    PTD_CU_CHECK(d.MulticastBindMem(b->mc_handle, 0, b->local, 0, b->bytes, 0));
    b->mc_bound = true;
    return map_rw(&b->mc, b->bytes, 2u << 20, b->mc_handle, b->dev);
}

// Drop a half-built multicast object (some rank failed a step): the block keeps working through its unicast mappings.
extern "C" int ptd_vmm_mc_drop(int id) {
    const Driver& d = drv();
    std::lock_guard<std::mutex> lk(g_mu);
    Block* b = get(id);
    if (!b) return PTD_ERR_ARG;
    unmap(b->mc, b->bytes);
    if (b->mc_handle != 0) {
        if (b->mc_bound) d.MulticastUnbind(b->mc_handle, b->cudev, 0, b->bytes);
        d.MemRelease(b->mc_handle);
    }
    b->mc_handle = 0;
    b->mc_bound = false;
    return PTD_OK;
}

// uc_out: `world` pointers (this process's mapping of every rank's memory; uc_out[rank] is the local one).
// mc_out: the multicast mapping or NULL.
extern "C" int ptd_vmm_ptrs(int id, void** uc_out, void** mc_out, long long* bytes_out) {
    std::lock_guard<std::mutex> lk(g_mu);
    Block* b = get(id);
    if (!b) return PTD_ERR_ARG;
    for (int r = 0; r < b->world; ++r) uc_out[r] = (void*)b->uc[r];
    if (mc_out) *mc_out = (void*)b->mc;
    if (bytes_out) *bytes_out = (long long)b->bytes;
    return PTD_OK;
}

// Unmap everything and release the handles.  The caller synchronises the device and the group first (no peer may
// still be reading); physical memory goes away when the last process holding a handle releases it.
extern "C" int ptd_vmm_destroy(int id) {
    const Driver& d = drv();
    std::lock_guard<std::mutex> lk(g_mu);
    Block* b = get(id);
    if (!b) return PTD_ERR_ARG;
    unmap(b->mc, b->bytes);
    if (b->mc_handle != 0) {
        if (b->mc_bound) d.MulticastUnbind(b->mc_handle, b->cudev, 0, b->bytes);
        d.MemRelease(b->mc_handle);
        b->mc_handle = 0;
    }
    for (int r = 0; r < b->world; ++r) {
        unmap(b->uc[r], b->bytes);
        if (r != b->rank && b->peer[r] != 0) { d.MemRelease(b->peer[r]); b->peer[r] = 0; }
    }
    if (b->local != 0) { d.MemRelease(b->local); b->local = 0; }
    b->live = false;
    return PTD_OK;
}
