// Symmetric heap on the CUDA virtual-memory-management API, with NVSwitch multicast (NVLS) objects on top.
//
// cudaMalloc + cudaIpc* (dp.cu) gives every rank a pointer to every peer's buffer, which is all that P2P loads and
// stores need.  It cannot give a MULTICAST address: one virtual address whose stores the NVSwitch replicates into
// every rank's copy (multimem.st) and whose loads return the sum of every rank's copy, reduced inside the switch
// (multimem.ld_reduce).  Those need physical allocations made with cuMemCreate, shared as POSIX file descriptors,
// bound into a cuMulticastCreate object and mapped with cuMemMap.  This file is that substrate:
//
//   b200_vmm_query        what the current device supports (VMM, fd handles, multicast) + the granularities
//   b200_vmm_create       cuMemCreate (pinned device memory, fd-exportable) -> generic handle + exported fd
//   b200_vmm_import       fd received from a peer (SCM_RIGHTS over a unix socket, parallel/symm.py) -> handle
//   b200_vmm_map          cuMemAddressReserve + cuMemMap + cuMemSetAccess(read/write for the current device);
//                         the same call maps local memory, a peer's memory (NVLink P2P) and a multicast object
//   b200_mc_create / _add_device / _bind   rank 0 creates the object for `world` devices; EVERY rank adds its
//                         device, then (after a barrier: all devices must be added first) binds its local memory
//   b200_vmm_unmap / _release
//
// Only generic loads/stores and multimem instructions address these mappings; TMA tensor maps in this library are
// always built over this GPU's own memory.
// Driver entry points are resolved through cudaGetDriverEntryPoint: the library keeps no link-time dependency on
// libcuda and still loads on a machine without a driver (the CPU build check).
// Replaces the communicator buffers NCCL registers for torch DDP (reference scripts/train_transformer.py:122-123;
// SURVEY T12).
#include <cuda.h>
#include <unistd.h>

#include "b200llm.h"
#include "common.cuh"

namespace {

template <typename F>
F drv(const char* name) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) return nullptr;
    return reinterpret_cast<F>(p);
}

#define DRV(var, name, ...)                                                            \
    typedef CUresult (*var##_t)(__VA_ARGS__);                                          \
    static var##_t var = drv<var##_t>(name);                                           \
    if (var == nullptr) {                                                              \
        b200_set_error("driver entry point %s is not available", name);                \
        return -2;                                                                     \
    }

#define CU_OK(expr, what)                                                              \
    do {                                                                               \
        CUresult r__ = (expr);                                                         \
        if (r__ != CUDA_SUCCESS) {                                                     \
            b200_set_error("%s failed: CUresult %d", what, (int)r__);                  \
            return (int)r__;                                                           \
        }                                                                              \
    } while (0)

int current_device(int* dev) {
    B200_CHECK_CUDA(cudaFree(0));          // make sure the primary context exists and is current
    B200_CHECK_CUDA(cudaGetDevice(dev));
    return 0;
}

CUmemAllocationProp device_prop(int dev) {
    CUmemAllocationProp prop;
    memset(&prop, 0, sizeof(prop));
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = dev;
    prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    return prop;
}

}  // namespace

// out[0] = VMM supported, out[1] = POSIX fd handles supported, out[2] = multicast supported,
// out[3] = allocation granularity (bytes), out[4] = multicast granularity for `world` devices (0 if unsupported)
extern "C" int b200_vmm_query(long long* out, int world) {
    int dev = 0;
    if (int rc = current_device(&dev)) return rc;
    DRV(cuDeviceGetAttribute_, "cuDeviceGetAttribute", int*, CUdevice_attribute, CUdevice);
    DRV(cuMemGetAllocationGranularity_, "cuMemGetAllocationGranularity", size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
    int v = 0;
    out[0] = out[1] = out[2] = out[3] = out[4] = 0;
    if (cuDeviceGetAttribute_(&v, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED, dev) == CUDA_SUCCESS) out[0] = v;
    if (cuDeviceGetAttribute_(&v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev) == CUDA_SUCCESS) out[1] = v;
    if (cuDeviceGetAttribute_(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) == CUDA_SUCCESS) out[2] = v;
    if (out[0]) {
        CUmemAllocationProp prop = device_prop(dev);
        size_t g = 0;
        CU_OK(cuMemGetAllocationGranularity_(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED), "cuMemGetAllocationGranularity");
        out[3] = (long long)g;
    }
    if (out[2] && world > 1) {
        typedef CUresult (*gran_t)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
        gran_t fn = drv<gran_t>("cuMulticastGetGranularity");
        if (fn != nullptr) {
            CUmulticastObjectProp mp;
            memset(&mp, 0, sizeof(mp));
            mp.numDevices = (unsigned)world;
            mp.size = (size_t)out[3];
            mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
            size_t g = 0;
            if (fn(&g, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS) out[4] = (long long)g;
            else out[2] = 0;
        } else {
            out[2] = 0;
        }
    }
    return 0;
}

// bytes must be a multiple of the granularity.  fd_out (nullable): a file descriptor of THIS process to be sent to the peers.
extern "C" int b200_vmm_create(long long bytes, unsigned long long* handle_out, int* fd_out) {
    int dev = 0;
    if (int rc = current_device(&dev)) return rc;
    DRV(cuMemCreate_, "cuMemCreate", CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
    DRV(cuMemExport_, "cuMemExportToShareableHandle", void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
    CUmemAllocationProp prop = device_prop(dev);
    CUmemGenericAllocationHandle h = 0;
    CU_OK(cuMemCreate_(&h, (size_t)bytes, &prop, 0), "cuMemCreate");
    *handle_out = (unsigned long long)h;
    if (fd_out != nullptr) {
        int fd = -1;
        CU_OK(cuMemExport_(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0), "cuMemExportToShareableHandle");
        *fd_out = fd;
    }
    return 0;
}

extern "C" int b200_vmm_import(int fd, unsigned long long* handle_out) {
    int dev = 0;
    if (int rc = current_device(&dev)) return rc;
    DRV(cuMemImport_, "cuMemImportFromShareableHandle", CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
    CUmemGenericAllocationHandle h = 0;
    CU_OK(cuMemImport_(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR), "cuMemImportFromShareableHandle");
    *handle_out = (unsigned long long)h;
    return 0;
}

// Map `bytes` of the allocation (or multicast object) `handle` into this process with read/write access for the
// current device.  Local memory, a peer's memory and a multicast object are all mapped by this one call.
extern "C" int b200_vmm_map(unsigned long long handle, long long bytes, void** ptr_out) {
    int dev = 0;
    if (int rc = current_device(&dev)) return rc;
    DRV(cuMemAddressReserve_, "cuMemAddressReserve", CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
    DRV(cuMemMap_, "cuMemMap", CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
    DRV(cuMemSetAccess_, "cuMemSetAccess", CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
    DRV(cuMemAddressFree_, "cuMemAddressFree", CUdeviceptr, size_t);
    CUdeviceptr va = 0;
    CU_OK(cuMemAddressReserve_(&va, (size_t)bytes, 0, 0, 0), "cuMemAddressReserve");
    CUresult r = cuMemMap_(va, (size_t)bytes, 0, (CUmemGenericAllocationHandle)handle, 0);
    if (r != CUDA_SUCCESS) {
        cuMemAddressFree_(va, (size_t)bytes);
        b200_set_error("cuMemMap failed: CUresult %d", (int)r);
        return (int)r;
    }
    CUmemAccessDesc acc;
    memset(&acc, 0, sizeof(acc));
    acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc.location.id = dev;
    acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    r = cuMemSetAccess_(va, (size_t)bytes, &acc, 1);
    if (r != CUDA_SUCCESS) {
        b200_set_error("cuMemSetAccess failed: CUresult %d", (int)r);
        return (int)r;
    }
    *ptr_out = (void*)va;
    return 0;
}

extern "C" int b200_vmm_unmap(void* ptr, long long bytes) {
    DRV(cuMemUnmap_, "cuMemUnmap", CUdeviceptr, size_t);
    DRV(cuMemAddressFree_, "cuMemAddressFree", CUdeviceptr, size_t);
    CU_OK(cuMemUnmap_((CUdeviceptr)ptr, (size_t)bytes), "cuMemUnmap");
    CU_OK(cuMemAddressFree_((CUdeviceptr)ptr, (size_t)bytes), "cuMemAddressFree");
    return 0;
}

extern "C" int b200_vmm_release(unsigned long long handle) {
    DRV(cuMemRelease_, "cuMemRelease", CUmemGenericAllocationHandle);
    CU_OK(cuMemRelease_((CUmemGenericAllocationHandle)handle), "cuMemRelease");
    return 0;
}

extern "C" int b200_vmm_close_fd(int fd) {
    if (fd >= 0) close(fd);
    return 0;
}

// ---- multicast objects --------------------------------------------------------------------------------------
extern "C" int b200_mc_create(long long bytes, int world, unsigned long long* handle_out, int* fd_out) {
    int dev = 0;
    if (int rc = current_device(&dev)) return rc;
    DRV(cuMulticastCreate_, "cuMulticastCreate", CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
    DRV(cuMemExport_, "cuMemExportToShareableHandle", void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = (unsigned)world;
    mp.size = (size_t)bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    CUmemGenericAllocationHandle h = 0;
    CU_OK(cuMulticastCreate_(&h, &mp), "cuMulticastCreate");
    *handle_out = (unsigned long long)h;
    int fd = -1;
    CU_OK(cuMemExport_(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0), "cuMemExportToShareableHandle(multicast)");
    *fd_out = fd;
    return 0;
}

extern "C" int b200_mc_add_device(unsigned long long mc_handle) {
    int dev = 0;
    if (int rc = current_device(&dev)) return rc;
    DRV(cuMulticastAddDevice_, "cuMulticastAddDevice", CUmemGenericAllocationHandle, CUdevice);
    CU_OK(cuMulticastAddDevice_((CUmemGenericAllocationHandle)mc_handle, (CUdevice)dev), "cuMulticastAddDevice");
    return 0;
}

extern "C" int b200_mc_bind(unsigned long long mc_handle, unsigned long long mem_handle, long long bytes) {
    DRV(cuMulticastBindMem_, "cuMulticastBindMem", CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
    CU_OK(cuMulticastBindMem_((CUmemGenericAllocationHandle)mc_handle, 0, (CUmemGenericAllocationHandle)mem_handle, 0, (size_t)bytes, 0),
          "cuMulticastBindMem");
    return 0;
}
