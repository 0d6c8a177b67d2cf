// Shared device/host helpers for the nxd_b200 sm_100a kernel library.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define NXD_API extern "C" __attribute__((visibility("default")))

// ---- error reporting (host) -------------------------------------------------
void nxd_set_error(const char* fmt, ...);

#define NXD_CHECK_CUDA(expr)                                                              \
    do {                                                                                  \
        cudaError_t _e = (expr);                                                          \
        if (_e != cudaSuccess) {                                                          \
            nxd_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return (int)_e;                                                               \
        }                                                                                 \
    } while (0)

#define NXD_REQUIRE(cond, ...)                   \
    do {                                         \
        if (!(cond)) {                           \
            nxd_set_error(__VA_ARGS__);          \
            return -1;                           \
        }                                        \
    } while (0)

// Every kernel launch of the library goes through NXD_LAUNCH_CHECK (or NXD_LAUNCH_CHECK_N when one check covers
// several launches), which also feeds the process-wide launch counter read by nxd_kernel_launches().
void nxd_count_launches(int n);

#define NXD_LAUNCH_CHECK() NXD_LAUNCH_CHECK_N(1)

#define NXD_LAUNCH_CHECK_N(n_launches)                                          \
    do {                                                                        \
        nxd_count_launches(n_launches);                                         \
        cudaError_t _e = cudaGetLastError();                                    \
        if (_e != cudaSuccess) {                                                \
            nxd_set_error("%s:%d: launch -> %s", __FILE__, __LINE__, cudaGetErrorString(_e)); \
            return (int)_e;                                                     \
        }                                                                       \
    } while (0)

// ---- programmatic dependent launch (PDL) ---------------------------------------
// The token-generation graph is a chain of short kernels; launch latency and prologues (barrier init, tensor-memory
// allocation, tensor-map fetch, the first weight tiles) are a large part of every one of them.  Kernels launched through
// nxd_launch_pdl may START while their predecessor in the stream is still running (as soon as all its CTAs have executed
// pdl_trigger() or exited); they must call pdl_wait() before they read anything the predecessor wrote and before ANY
// global write (the predecessor may still be reading a buffer the allocator has handed to this kernel).  Constant data
// (weights) may be fetched before pdl_wait().  Both instructions are no-ops in a launch without the attribute.
// NXD_PDL=0 turns the attribute off.
int nxd_pdl_enabled();

template <typename... KArgs, typename... Args>
inline cudaError_t nxd_launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                  Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = nxd_pdl_enabled();
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// ---- device helpers ---------------------------------------------------------
typedef __nv_bfloat16 bf16;
typedef __nv_bfloat162 bf162;

struct alignas(16) bf16x8 {
    bf162 v[4];
};

__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void prefetch_l2(const void* p) {
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}

__device__ __forceinline__ void unpack8(const bf16x8& p, float* f) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float2 t = __bfloat1622float2(p.v[i]);
        f[2 * i] = t.x;
        f[2 * i + 1] = t.y;
    }
}

__device__ __forceinline__ bf16x8 pack8(const float* f) {
    bf16x8 p;
#pragma unroll
    for (int i = 0; i < 4; ++i) p.v[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    return p;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Block-wide sum; `red` is >= 32 floats of shared memory. All threads get the result.
__device__ __forceinline__ float block_sum(float v, float* red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nwarps = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();  // protect `red` from a previous use
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float r = (lane < nwarps) ? red[lane] : 0.f;
    r = warp_sum(r);
    return r;
}

__device__ __forceinline__ float block_max(float v, float* red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nwarps = (blockDim.x + 31) >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float r = (lane < nwarps) ? red[lane] : -INFINITY;
    r = warp_max(r);
    return r;
}

static inline int nxd_num_sms() {
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (sms <= 0) sms = 148;
    }
    return sms;
}
