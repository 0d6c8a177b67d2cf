// Symmetric (peer-mapped) device memory for the fused compute+collective kernels.
//
// The reference gets its collectives from the Neuron runtime (libnccom over NeuronLink); its kernels never touch
// a peer's memory.  Here every rank of a group allocates the SAME layout with cudaMalloc, exports it with CUDA IPC
// and maps every peer's copy, so a kernel can load from / store to a peer's HBM over NVLink 5 / NVSwitch with
// ordinary global-memory instructions (or TMA) and synchronise through flags in a small peer-mapped signal pad.
// Handle exchange is done by the Python side over torch.distributed (ops/symm_mem.py).
#include "common.cuh"

NXD_API int nxd_symm_handle_bytes(void) { return (int)sizeof(cudaIpcMemHandle_t); }

// Allocate `bytes` (zero-filled) and export an IPC handle (64 bytes) for the peers.
NXD_API int nxd_symm_alloc(size_t bytes, void** ptr, unsigned char* handle_out) {
    NXD_REQUIRE(ptr != nullptr && handle_out != nullptr && bytes > 0, "symm_alloc: bad arguments");
    void* p = nullptr;
    NXD_CHECK_CUDA(cudaMalloc(&p, bytes));
    NXD_CHECK_CUDA(cudaMemset(p, 0, bytes));
    cudaIpcMemHandle_t h;
    NXD_CHECK_CUDA(cudaIpcGetMemHandle(&h, p));
    memcpy(handle_out, &h, sizeof(h));
    *ptr = p;
    return 0;
}

// Map a peer's allocation into this process (peer access is enabled lazily by the driver).
NXD_API int nxd_symm_open(const unsigned char* handle, void** ptr) {
    NXD_REQUIRE(ptr != nullptr && handle != nullptr, "symm_open: bad arguments");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void* p = nullptr;
    NXD_CHECK_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *ptr = p;
    return 0;
}

NXD_API int nxd_symm_close(void* ptr) {
    if (ptr) NXD_CHECK_CUDA(cudaIpcCloseMemHandle(ptr));
    return 0;
}

NXD_API int nxd_symm_free(void* ptr) {
    if (ptr) NXD_CHECK_CUDA(cudaFree(ptr));
    return 0;
}

// ---- device-side barrier over the signal pads (used by tests and by kernels that need a full rendezvous) -------
namespace {
__global__ void symm_barrier_kernel(uint32_t* const* sig_peers, uint32_t* sig_local, int world, int rank, int slot,
                                    uint32_t epoch) {
    // every rank bumps slot[rank] on every peer, then waits until all `world` entries of its own pad reached epoch
    const int t = threadIdx.x;
    if (t < world) {
        uint32_t* dst = sig_peers[t] + slot + rank;
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(dst), "r"(epoch) : "memory");
        const uint32_t* src = sig_local + slot + t;
        uint32_t v;
        do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(src) : "memory");
        } while ((int32_t)(v - epoch) < 0);
    }
}
}  // namespace

// sig_peers: device array of `world` pointers to the peers' signal pads (own pad included at index rank).
NXD_API int nxd_symm_barrier(void* const* sig_peers, void* sig_local, int world, int rank, int slot, unsigned epoch,
                             void* stream) {
    NXD_REQUIRE(world >= 1 && world <= 32, "symm_barrier: world %d out of range", world);
    symm_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>((uint32_t* const*)sig_peers, (uint32_t*)sig_local, world,
                                                            rank, slot, epoch);
    NXD_LAUNCH_CHECK();
    return 0;
}
