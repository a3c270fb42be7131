// libpcsb200: peer (NVLink / NVSwitch) memory for the slab-decomposed solvers.
//
// One process per GPU.  Each rank allocates its slab arrays (iterates with their halo rows, and a few flag words) with
// pcs_peer_alloc, publishes the 64-byte IPC handle through the torch.distributed store, and maps its two row
// neighbours' allocations with pcs_peer_open.  The fused iteration kernel (pcs_pds.cu) then stores the boundary rows
// of x', z' straight into the neighbours' halo rows and releases a flag word in their memory; the only thing the host
// ever does per iteration is launch the one kernel.
//
// These are the ONLY entry points of the library that own device memory (cudaMalloc'd blocks are the unit of CUDA IPC
// export; PyTorch's caching allocator sub-allocates and may use virtual-memory segments that cannot be exported).
#include "pcs_common.cuh"
#include <string.h>

using namespace pcs;

#define PCS_CUDA_TRY(expr, what)                                                 \
    do {                                                                         \
        cudaError_t e_ = (expr);                                                 \
        if (e_ != cudaSuccess) {                                                 \
            set_error("%s: %s", what, cudaGetErrorString(e_));                   \
            (void)cudaGetLastError();                                            \
            return PCS_ERR_CUDA;                                                 \
        }                                                                        \
    } while (0)

extern "C" int pcs_peer_handle_bytes(void) { return (int)sizeof(cudaIpcMemHandle_t); }

// Allocate `bytes` of zero-initialised device memory on the current device and export it.
extern "C" int pcs_peer_alloc(void** dptr_out, int64_t bytes, void* handle_out) {
    PCS_REQUIRE(dptr_out && handle_out && bytes > 0, "pcs_peer_alloc: bad arguments");
    void* p = nullptr;
    PCS_CUDA_TRY(cudaMalloc(&p, (size_t)bytes), "pcs_peer_alloc(cudaMalloc)");
    cudaError_t e = cudaMemset(p, 0, (size_t)bytes);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) {
        set_error("pcs_peer_alloc: %s", cudaGetErrorString(e));
        cudaFree(p);
        (void)cudaGetLastError();
        return PCS_ERR_CUDA;
    }
    memcpy(handle_out, &h, sizeof(h));
    *dptr_out = p;
    return PCS_OK;
}

// Map a neighbour's allocation into this process (peer access is enabled lazily by the driver).
extern "C" int pcs_peer_open(const void* handle, void** dptr_out) {
    PCS_REQUIRE(handle && dptr_out, "pcs_peer_open: bad arguments");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void* p = nullptr;
    PCS_CUDA_TRY(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess), "pcs_peer_open(cudaIpcOpenMemHandle)");
    *dptr_out = p;
    return PCS_OK;
}

extern "C" int pcs_peer_close(void* dptr) {
    if (!dptr) return PCS_OK;
    PCS_CUDA_TRY(cudaIpcCloseMemHandle(dptr), "pcs_peer_close");
    return PCS_OK;
}

extern "C" int pcs_peer_free(void* dptr) {
    if (!dptr) return PCS_OK;
    PCS_CUDA_TRY(cudaFree(dptr), "pcs_peer_free");
    return PCS_OK;
}

// 1 if `device` can load/store `peer`'s memory directly (NVLink / PCIe P2P), else 0; negative on error.
extern "C" int pcs_peer_can_access(int device, int peer) {
    int ok = 0;
    if (cudaDeviceCanAccessPeer(&ok, device, peer) != cudaSuccess) {
        (void)cudaGetLastError();
        return -1;
    }
    return ok;
}
