// RMSNorm forward / backward for sm_100a.
//
// Replaces the reference's fp64-upcast torch ops
// (examples/training/llama/modeling_llama_nxd.py:142-149) and the inference
// `RmsNorm` custom call (examples/inference/modules/custom_calls.py:5-16).
//
// Layout: x [rows, H] bf16 row-major, H % 8 == 0.  One CTA per row keeps the row
// in registers as fp32 between the sum-of-squares pass and the scale pass, so HBM
// traffic is exactly one read + one write of the activation (HBM-bound op).
// Optional fused residual add:  s = x + r;  y = rmsnorm(s);  s is also written.
//
// Backward: dx = rstd * (g - xhat * mean(g * xhat)),  g = dy * w,  xhat = x * rstd
//           dw = sum_rows(dy * xhat)   -> persistent CTAs accumulate per-thread
//           partials in registers, write [P, H] fp32 partials to scratch, and a
//           second kernel reduces the P partial rows.
#include "common.cuh"

namespace {

template <int VPT>  // 16-byte vectors per thread
__global__ void rmsnorm_fwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ res,
                                   const bf16* __restrict__ w, bf16* __restrict__ y,
                                   bf16* __restrict__ res_out, float* __restrict__ rstd_out, int H,
                                   float eps) {
    __shared__ float red[32];
    pdl_trigger();
    pdl_wait();
    const int row = blockIdx.x;
    const int nvec = H >> 3;
    const bf16x8* xr = reinterpret_cast<const bf16x8*>(x + (size_t)row * H);
    const bf16x8* rr = res ? reinterpret_cast<const bf16x8*>(res + (size_t)row * H) : nullptr;
    float v[VPT][8];
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
        const int c = threadIdx.x + i * blockDim.x;
        if (c < nvec) {
            bf16x8 p = xr[c];
            unpack8(p, v[i]);
            if (rr) {
                float r[8];
                bf16x8 q = rr[c];
                unpack8(q, r);
#pragma unroll
                for (int k = 0; k < 8; ++k) v[i][k] += r[k];
                if (res_out) {
                    // round the residual stream to bf16 once, and normalise the rounded value
                    bf16x8 s = pack8(v[i]);
                    reinterpret_cast<bf16x8*>(res_out + (size_t)row * H)[c] = s;
                    unpack8(s, v[i]);
                }
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) ss += v[i][k] * v[i][k];
        }
    }
    ss = block_sum(ss, red);
    const float rstd = rsqrtf(ss / (float)H + eps);
    if (threadIdx.x == 0 && rstd_out) rstd_out[row] = rstd;
    const bf16x8* wr = reinterpret_cast<const bf16x8*>(w);
    bf16x8* yr = reinterpret_cast<bf16x8*>(y + (size_t)row * H);
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
        const int c = threadIdx.x + i * blockDim.x;
        if (c < nvec) {
            float wv[8], o[8];
            bf16x8 p = wr[c];
            unpack8(p, wv);
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] = v[i][k] * rstd * wv[k];
            yr[c] = pack8(o);
        }
    }
}

template <int VPT>
__global__ void rmsnorm_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ x,
                                   const bf16* __restrict__ w, const float* __restrict__ rstd,
                                   const bf16* __restrict__ dres, bf16* __restrict__ dx,
                                   float* __restrict__ dw_partial, int rows, int H) {
    __shared__ float red[32];
    const int nvec = H >> 3;
    float wv[VPT][8], dwacc[VPT][8];
    const bf16x8* wr = reinterpret_cast<const bf16x8*>(w);
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
        const int c = threadIdx.x + i * blockDim.x;
#pragma unroll
        for (int k = 0; k < 8; ++k) dwacc[i][k] = 0.f;
        if (c < nvec) {
            bf16x8 p = wr[c];
            unpack8(p, wv[i]);
        }
    }
    // software pipeline over the rows of this CTA: the loads of row i + 1 are issued before the block-wide reduction
    // of row i, so the row loop is no longer one global-load latency plus two barriers per row (61 us -> see profiles)
    bf16x8 ndy[VPT], nx[VPT];
    auto fetch = [&](int row) {
        const bf16x8* dyr = reinterpret_cast<const bf16x8*>(dy + (size_t)row * H);
        const bf16x8* xr = reinterpret_cast<const bf16x8*>(x + (size_t)row * H);
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int c = threadIdx.x + i * blockDim.x;
            if (c < nvec) {
                ndy[i] = dyr[c];
                nx[i] = xr[c];
            }
        }
    };
    if ((int)blockIdx.x < rows) fetch(blockIdx.x);
    for (int row = blockIdx.x; row < rows; row += gridDim.x) {
        const float rs = rstd[row];
        float g[VPT][8], xh[VPT][8];
        float dot = 0.f;
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int c = threadIdx.x + i * blockDim.x;
            if (c < nvec) {
                float d[8], xv[8];
                unpack8(ndy[i], d);
                unpack8(nx[i], xv);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    xh[i][k] = xv[k] * rs;
                    g[i][k] = d[k] * wv[i][k];
                    dot += g[i][k] * xh[i][k];
                    dwacc[i][k] += d[k] * xh[i][k];
                }
            }
        }
        if (row + (int)gridDim.x < rows) fetch(row + gridDim.x);
        dot = block_sum(dot, red) / (float)H;
        bf16x8* dxr = reinterpret_cast<bf16x8*>(dx + (size_t)row * H);
        const bf16x8* drr = dres ? reinterpret_cast<const bf16x8*>(dres + (size_t)row * H) : nullptr;
#pragma unroll
        for (int i = 0; i < VPT; ++i) {
            const int c = threadIdx.x + i * blockDim.x;
            if (c < nvec) {
                float o[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) o[k] = rs * (g[i][k] - xh[i][k] * dot);
                if (drr) {  // fused add of the gradient flowing through the residual branch
                    float r[8];
                    bf16x8 q = drr[c];
                    unpack8(q, r);
#pragma unroll
                    for (int k = 0; k < 8; ++k) o[k] += r[k];
                }
                dxr[c] = pack8(o);
            }
        }
    }
    float* out = dw_partial + (size_t)blockIdx.x * H;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
        const int c = threadIdx.x + i * blockDim.x;
        if (c < nvec) {
            float4* o4 = reinterpret_cast<float4*>(out + (size_t)c * 8);
            o4[0] = make_float4(dwacc[i][0], dwacc[i][1], dwacc[i][2], dwacc[i][3]);
            o4[1] = make_float4(dwacc[i][4], dwacc[i][5], dwacc[i][6], dwacc[i][7]);
        }
    }
}

// dw[h] = sum_p partial[p, h]; optionally accumulates into an fp32 buffer.
__global__ void colsum_partials_kernel(const float* __restrict__ partial, int P, int H,
                                       bf16* __restrict__ dw_bf16, float* __restrict__ dw_f32,
                                       int accumulate) {
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= H) return;
    float s = 0.f;
    for (int p = 0; p < P; ++p) s += partial[(size_t)p * H + h];
    if (dw_f32) dw_f32[h] = accumulate ? dw_f32[h] + s : s;
    if (dw_bf16) dw_bf16[h] = __float2bfloat16(s);
}

struct NormCfg {
    int threads;
    int vpt;
};

NormCfg pick_cfg(int H) {
    const int nvec = H / 8;
    NormCfg c;
    if (nvec <= 128) {
        c.vpt = 1;
    } else if (nvec <= 512) {
        c.vpt = 2;
    } else if (nvec <= 2048) {
        c.vpt = 4;
    } else {
        c.vpt = 8;
    }
    int t = (nvec + c.vpt - 1) / c.vpt;
    t = ((t + 31) / 32) * 32;
    if (t > 1024) t = 1024;
    if (t < 32) t = 32;
    c.threads = t;
    return c;
}

}  // namespace

// Number of partial rows (persistent CTAs) the backward kernel uses; the scratch
// buffer passed to nxd_rmsnorm_bwd must hold nxd_rmsnorm_bwd_num_partials * H floats.
NXD_API int nxd_rmsnorm_bwd_num_partials(int rows) {
    int p = nxd_num_sms() * 4;
    if (p > rows) p = rows;
    if (p < 1) p = 1;
    return p;
}

NXD_API int nxd_rmsnorm_fwd(const void* x, const void* residual, const void* w, void* y, void* residual_out,
                            float* rstd, int rows, int H, float eps, void* stream) {
    NXD_REQUIRE(H % 8 == 0 && H <= 65536, "rmsnorm: H=%d must be a multiple of 8 and <= 65536", H);
    if (rows == 0) return 0;
    NormCfg c = pick_cfg(H);
    cudaStream_t s = (cudaStream_t)stream;
#define LAUNCH(V)                                                                                          \
    NXD_CHECK_CUDA(nxd_launch_pdl(rmsnorm_fwd_kernel<V>, dim3(rows), dim3(c.threads), 0, s, (const bf16*)x,   \
                                  (const bf16*)residual, (const bf16*)w, (bf16*)y, (bf16*)residual_out, rstd, H, eps))
    switch (c.vpt) {
        case 1: LAUNCH(1); break;
        case 2: LAUNCH(2); break;
        case 4: LAUNCH(4); break;
        default: LAUNCH(8); break;
    }
#undef LAUNCH
    NXD_LAUNCH_CHECK();
    return 0;
}

static int rmsnorm_bwd_impl(const void* dy, const void* x, const void* w, const float* rstd, const void* dresidual,
                            void* dx, void* dw_bf16, float* dw_f32, int accumulate, float* scratch, int rows,
                            int H, cudaStream_t s) {
    NormCfg c = pick_cfg(H);
    const int P = nxd_rmsnorm_bwd_num_partials(rows);
#define LAUNCH(V)                                                                                        \
    rmsnorm_bwd_kernel<V><<<P, c.threads, 0, s>>>((const bf16*)dy, (const bf16*)x, (const bf16*)w, rstd, \
                                                  (const bf16*)dresidual, (bf16*)dx, scratch, rows, H)
    switch (c.vpt) {
        case 1: LAUNCH(1); break;
        case 2: LAUNCH(2); break;
        case 4: LAUNCH(4); break;
        default: LAUNCH(8); break;
    }
#undef LAUNCH
    NXD_LAUNCH_CHECK();
    colsum_partials_kernel<<<(H + 255) / 256, 256, 0, s>>>(scratch, P, H, (bf16*)dw_bf16, dw_f32, accumulate);
    NXD_LAUNCH_CHECK();
    return 0;
}

// rows == 0: nothing to differentiate, but a caller that asked for a fresh dW (accumulate == 0) still gets zeros
// (the column-sum kernel with no partial rows reads nothing and stores 0).
static int rmsnorm_bwd_empty(void* dw_bf16, float* dw_f32, int accumulate, int H, cudaStream_t s) {
    if (accumulate || (dw_bf16 == nullptr && dw_f32 == nullptr)) return 0;
    colsum_partials_kernel<<<(H + 255) / 256, 256, 0, s>>>(nullptr, 0, H, (bf16*)dw_bf16, dw_f32, 0);
    NXD_LAUNCH_CHECK();
    return 0;
}

// scratch: caller-provided device buffer of nxd_rmsnorm_bwd_num_partials(rows) * H floats.
NXD_API int nxd_rmsnorm_bwd(const void* dy, const void* x, const void* w, const float* rstd,
                            const void* dresidual, void* dx, void* dw_bf16, float* dw_f32, int accumulate,
                            float* scratch, int rows, int H, void* stream) {
    NXD_REQUIRE(H % 8 == 0 && H <= 65536, "rmsnorm_bwd: H=%d must be a multiple of 8 and <= 65536", H);
    NXD_REQUIRE(scratch != nullptr, "rmsnorm_bwd: scratch buffer is required");
    if (rows == 0) return rmsnorm_bwd_empty(dw_bf16, dw_f32, accumulate, H, (cudaStream_t)stream);
    return rmsnorm_bwd_impl(dy, x, w, rstd, dresidual, dx, dw_bf16, dw_f32, accumulate, scratch, rows, H,
                            (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------------------------
// Allocation-free variant of the backward pass.
//
// nxd_rmsnorm_bwd() asks its caller for a scratch buffer (the per-CTA partial sums of dW).  Some callers have no
// way to provide one: a host program that handed the whole HBM pool to another allocator at start-up (so a later
// cudaMalloc fails or fragments the pool), a latency-critical loop that cannot tolerate the device-wide
// synchronisation cudaMalloc / cudaFree imply, or a plug-in running inside someone else's operator callback where
// device memory may only come from that host's allocator.  For them the partial sums live in the library image
// itself: a file-scope __device__ array of kStaticScratchFloats floats (4 MiB) that every GPU which loads the
// library keeps until the process exits.
//
//   * Limit: H <= kStaticMaxH (8192).  The number of partial rows is reduced to fit (at least 128 rows of partials
//     remain at the largest H), which changes the summation order of dW slightly but not the result beyond fp32
//     rounding.  Larger H is rejected with an error; use nxd_rmsnorm_bwd() with your own buffer instead.
//   * The first call on each device resolves the array's device address (cudaGetSymbolAddress with that device
//     current).  Make that first call before you start capturing a stream into a graph.
//   * All calls share the one array.  Nothing in it has to start at zero -- every partial row that is read is
//     written earlier in the same call -- and its contents after a call are meaningless.  Two calls whose kernels
//     can overlap on the GPU (different streams) race on it; serialise them.
//   * The call itself performs no cudaMalloc / cudaFree and calls no allocating library routine.
// ---------------------------------------------------------------------------------------------------------
namespace {

constexpr int kStaticScratchFloats = 1 << 20;
constexpr int kStaticMaxH = 8192;
constexpr int kStaticMaxDevices = 64;

static __device__ __align__(256) float g_rmsnorm_bwd_static_scratch[kStaticScratchFloats];

// In host code the name above is only a handle for the symbol, not an address the GPU can use: resolve it per device.
int static_scratch_for_current_device(float** out) {
    static float* ptrs[kStaticMaxDevices] = {nullptr};
    int dev = 0;
    NXD_CHECK_CUDA(cudaGetDevice(&dev));
    NXD_REQUIRE(dev >= 0 && dev < kStaticMaxDevices, "rmsnorm_bwd_static: device index %d out of range", dev);
    float* p = __atomic_load_n(&ptrs[dev], __ATOMIC_ACQUIRE);
    if (p == nullptr) {
        void* addr = nullptr;
        NXD_CHECK_CUDA(cudaGetSymbolAddress(&addr, g_rmsnorm_bwd_static_scratch));
        NXD_REQUIRE(addr != nullptr, "rmsnorm_bwd_static: could not resolve the scratch symbol on device %d", dev);
        p = static_cast<float*>(addr);
        __atomic_store_n(&ptrs[dev], p, __ATOMIC_RELEASE);
    }
    *out = p;
    return 0;
}

int rmsnorm_bwd_static_launch(const void* dy, const void* x, const void* w, const float* rstd, const void* dresidual,
                              void* dx, void* dw_bf16, float* dw_f32, int accumulate, float* scratch, int P, int rows,
                              int H, cudaStream_t s) {
    NormCfg c = pick_cfg(H);
#define LAUNCH(V)                                                                                        \
    rmsnorm_bwd_kernel<V><<<P, c.threads, 0, s>>>((const bf16*)dy, (const bf16*)x, (const bf16*)w, rstd, \
                                                  (const bf16*)dresidual, (bf16*)dx, scratch, rows, H)
    switch (c.vpt) {
        case 1: LAUNCH(1); break;
        case 2: LAUNCH(2); break;
        case 4: LAUNCH(4); break;
        default: LAUNCH(8); break;
    }
#undef LAUNCH
    NXD_LAUNCH_CHECK();
    colsum_partials_kernel<<<(H + 255) / 256, 256, 0, s>>>(scratch, P, H, (bf16*)dw_bf16, dw_f32, accumulate);
    NXD_LAUNCH_CHECK();
    return 0;
}

}  // namespace

NXD_API int nxd_rmsnorm_bwd_static_max_h(void) { return kStaticMaxH; }

NXD_API int nxd_rmsnorm_bwd_static(const void* dy, const void* x, const void* w, const float* rstd,
                                   const void* dresidual, void* dx, void* dw_bf16, float* dw_f32, int accumulate,
                                   int rows, int H, void* stream) {
    NXD_REQUIRE(H % 8 == 0 && H > 0, "rmsnorm_bwd_static: H=%d must be a positive multiple of 8", H);
    NXD_REQUIRE(H <= kStaticMaxH,
                "rmsnorm_bwd_static: H=%d exceeds the built-in scratch limit of %d columns; call nxd_rmsnorm_bwd with a "
                "caller-provided scratch buffer instead",
                H, kStaticMaxH);
    NXD_REQUIRE(rows >= 0, "rmsnorm_bwd_static: negative row count");
    if (rows == 0) return rmsnorm_bwd_empty(dw_bf16, dw_f32, accumulate, H, (cudaStream_t)stream);
    float* scratch = nullptr;
    if (int rc = static_scratch_for_current_device(&scratch)) return rc;
    int P = nxd_rmsnorm_bwd_num_partials(rows);
    const int fit = kStaticScratchFloats / H;
    if (P > fit) P = fit;
    return rmsnorm_bwd_static_launch(dy, x, w, rstd, dresidual, dx, dw_bf16, dw_f32, accumulate, scratch, P, rows, H,
                                     (cudaStream_t)stream);
}
