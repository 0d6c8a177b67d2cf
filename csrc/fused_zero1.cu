// ZeRO-1 optimizer step over NVLink peer memory: the gradient reduce-scatter and the parameter
// all-gather of the data-parallel group are folded into the kernels that consume / produce the data.
//
// Reference: the ZeRO-1 optimizer used by trainer/trainer.py:244-269 reduce-scatters every gradient bucket
// (xm.reduce_scatter), casts and scales the shard, steps the inner optimizer on the shard and all-gathers the updated
// parameters (xm.all_gather): four passes over the shard plus two collectives.  Here the flat gradient and parameter
// buffers of a bucket live in symmetric memory and the step is TWO kernels:
//
//   nxd_zero1_reduce_shard     my shard of the summed gradient = sum over ranks (in rank order, fp32) of the peers'
//                              slices, PULLED with 16-byte loads straight from their flat gradient buffers; written
//                              once as fp32 (the reduce-scatter, the bf16 -> fp32 cast and the "higher precision
//                              collective" option of the reference in one pass; no staging buffer, nothing rounded
//                              to bf16 between ranks);
//   nxd_zero1_adamw_allgather  AdamW on the shard (fp32 master / moments, clip coefficient read from device memory)
//                              whose bf16 parameter output is PUSHED into every rank's flat parameter buffer -- the
//                              all-gather is the store of the update itself.
//
// The global gradient norm sits between the two (it needs every rank's shard), so the step stays two-phase; the LOCAL
// part of the norm is not a pass of its own, though: the reduce kernel squares what it has just summed and leaves two
// partial sums per block (elements before / after `split_at`: tensor-parallel-sharded and replicated parameters are
// weighted differently in the norm), which nxd_zero1_sumsq_finish adds up in a fixed order (bitwise reproducible).
// Rendezvous: the caller brackets the pair with nxd_symm_barrier (all gradients complete before the first pull; all
// parameter slices landed before the next forward).
#include "common.cuh"

namespace {

constexpr int kZMaxDp = 8;

struct ZPeers {
    void* ptr[kZMaxDp];
    int dp, rank;
};

__device__ __forceinline__ uint4 ld_nc_16(const void* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

// grad dtype bf16: 8 elements per 16-byte load; two loads per peer in flight per thread
// block-level finish of the two running sums of squares: partials[2 * block] / [2 * block + 1]
__device__ __forceinline__ void store_block_sumsq(float ss0, float ss1, float* __restrict__ partials) {
    __shared__ float red[2][8];
    ss0 = warp_sum(ss0);
    ss1 = warp_sum(ss1);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
        red[0][warp] = ss0;
        red[1][warp] = ss1;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        float a = 0.f, b = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) {
            a += red[0][w];
            b += red[1][w];
        }
        partials[2 * blockIdx.x] = a;
        partials[2 * blockIdx.x + 1] = b;
    }
}

// acc[0 .. cnt) are local elements base .. base + cnt: squares go to ss0 before split_at, to ss1 from split_at on
template <int kCnt>
__device__ __forceinline__ void add_squares(const float* acc, size_t base, size_t split_at, float& ss0, float& ss1) {
    float t = 0.f;
#pragma unroll
    for (int e = 0; e < kCnt; ++e) t = fmaf(acc[e], acc[e], t);
    if (base + kCnt <= split_at) {
        ss0 += t;
    } else if (base >= split_at) {
        ss1 += t;
    } else {
#pragma unroll
        for (int e = 0; e < kCnt; ++e) {
            if (base + e < split_at) ss0 = fmaf(acc[e], acc[e], ss0);
            else ss1 = fmaf(acc[e], acc[e], ss1);
        }
    }
}

__global__ void __launch_bounds__(256) zero1_reduce_shard_bf16_kernel(const ZPeers g, size_t lo, size_t n, float* __restrict__ out,
                                                                      size_t split_at, float* __restrict__ partials) {
    const size_t nvec = n >> 3;
    float ss0 = 0.f, ss1 = 0.f;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (size_t)gridDim.x * blockDim.x) {
        uint4 v[kZMaxDp];
#pragma unroll
        for (int r = 0; r < kZMaxDp; ++r)
            if (r < g.dp) v[r] = ld_nc_16(reinterpret_cast<const bf16*>(g.ptr[r]) + lo + (i << 3));
        float acc[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = 0.f;
#pragma unroll
        for (int r = 0; r < kZMaxDp; ++r) {
            if (r < g.dp) {
                const __nv_bfloat162* p = reinterpret_cast<const __nv_bfloat162*>(&v[r]);
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float2 f = __bfloat1622float2(p[e]);
                    acc[2 * e] += f.x;
                    acc[2 * e + 1] += f.y;
                }
            }
        }
        float4* o = reinterpret_cast<float4*>(out + (i << 3));
        o[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
        o[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
        if (partials) add_squares<8>(acc, i << 3, split_at, ss0, ss1);
    }
    if (partials) store_block_sumsq(ss0, ss1, partials);
}

__global__ void __launch_bounds__(256) zero1_reduce_shard_f32_kernel(const ZPeers g, size_t lo, size_t n, float* __restrict__ out,
                                                                     size_t split_at, float* __restrict__ partials) {
    const size_t nvec = n >> 2;
    float ss0 = 0.f, ss1 = 0.f;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (size_t)gridDim.x * blockDim.x) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int r = 0; r < kZMaxDp; ++r) {
            if (r < g.dp) {
                const uint4 v = ld_nc_16(reinterpret_cast<const float*>(g.ptr[r]) + lo + (i << 2));
                acc.x += __uint_as_float(v.x); acc.y += __uint_as_float(v.y);
                acc.z += __uint_as_float(v.z); acc.w += __uint_as_float(v.w);
            }
        }
        reinterpret_cast<float4*>(out)[i] = acc;
        if (partials) {
            const float a4[4] = {acc.x, acc.y, acc.z, acc.w};
            add_squares<4>(a4, i << 2, split_at, ss0, ss1);
        }
    }
    if (partials) store_block_sumsq(ss0, ss1, partials);
}

// sums[0] += sum of partials[2 b], sums[1] += sum of partials[2 b + 1] over the blocks, in a fixed order
__global__ void __launch_bounds__(256) zero1_sumsq_finish_kernel(const float* __restrict__ partials, int blocks,
                                                                 float* __restrict__ sum_sharded,
                                                                 float* __restrict__ sum_replicated) {
    __shared__ float red[2][8];
    float a = 0.f, b = 0.f;
    for (int i = threadIdx.x; i < blocks; i += 256) {
        a += partials[2 * i];
        b += partials[2 * i + 1];
    }
    a = warp_sum(a);
    b = warp_sum(b);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
        red[0][warp] = a;
        red[1][warp] = b;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        float x = 0.f, y = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) {
            x += red[0][w];
            y += red[1][w];
        }
        sum_sharded[0] += x;
        sum_replicated[0] += y;
    }
}

// 8 elements per thread and step: fp32 state in 2 x float4, bf16 parameters out as one 16-byte store per rank
__global__ void __launch_bounds__(256) zero1_adamw_allgather_kernel(float* __restrict__ master, float* __restrict__ m,
                                                                    float* __restrict__ v, const float* __restrict__ grad,
                                                                    const ZPeers p, size_t lo, size_t n, float lr, float beta1,
                                                                    float beta2, float eps, float wd, float bc1, float bc2,
                                                                    float grad_scale, const float* __restrict__ clip_coef) {
    const float scale = grad_scale * (clip_coef ? clip_coef[0] : 1.f);
    const size_t nvec = n >> 3;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (size_t)gridDim.x * blockDim.x) {
        float gg[8], mm[8], vv[8], pp[8];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const float4 a = reinterpret_cast<const float4*>(grad)[2 * i + h];
            const float4 b = reinterpret_cast<const float4*>(m)[2 * i + h];
            const float4 c = reinterpret_cast<const float4*>(v)[2 * i + h];
            const float4 d = reinterpret_cast<const float4*>(master)[2 * i + h];
            gg[4 * h] = a.x; gg[4 * h + 1] = a.y; gg[4 * h + 2] = a.z; gg[4 * h + 3] = a.w;
            mm[4 * h] = b.x; mm[4 * h + 1] = b.y; mm[4 * h + 2] = b.z; mm[4 * h + 3] = b.w;
            vv[4 * h] = c.x; vv[4 * h + 1] = c.y; vv[4 * h + 2] = c.z; vv[4 * h + 3] = c.w;
            pp[4 * h] = d.x; pp[4 * h + 1] = d.y; pp[4 * h + 2] = d.z; pp[4 * h + 3] = d.w;
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const float gs = gg[e] * scale;
            mm[e] = beta1 * mm[e] + (1.f - beta1) * gs;
            vv[e] = beta2 * vv[e] + (1.f - beta2) * gs * gs;
            pp[e] = pp[e] * (1.f - lr * wd) - lr * (mm[e] / bc1) / (sqrtf(vv[e] / bc2) + eps);
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            reinterpret_cast<float4*>(m)[2 * i + h] = make_float4(mm[4 * h], mm[4 * h + 1], mm[4 * h + 2], mm[4 * h + 3]);
            reinterpret_cast<float4*>(v)[2 * i + h] = make_float4(vv[4 * h], vv[4 * h + 1], vv[4 * h + 2], vv[4 * h + 3]);
            reinterpret_cast<float4*>(master)[2 * i + h] = make_float4(pp[4 * h], pp[4 * h + 1], pp[4 * h + 2], pp[4 * h + 3]);
        }
        uint4 packed;
        __nv_bfloat162* q = reinterpret_cast<__nv_bfloat162*>(&packed);
#pragma unroll
        for (int e = 0; e < 4; ++e) q[e] = __floats2bfloat162_rn(pp[2 * e], pp[2 * e + 1]);
        // own copy first, then the peers in ring order (every rank starts with a different peer)
#pragma unroll
        for (int d = 0; d < kZMaxDp; ++d) {
            if (d < p.dp) {
                const int peer = (p.rank + d) % p.dp;
                reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(p.ptr[peer]) + lo)[i] = packed;
            }
        }
    }
}

int fill_peers(ZPeers& z, void* const* ptrs, int dp, int rank) {
    NXD_REQUIRE(dp >= 1 && dp <= kZMaxDp && rank >= 0 && rank < dp && ptrs, "zero1: dp %d / rank %d out of range", dp, rank);
    for (int i = 0; i < dp; ++i) z.ptr[i] = ptrs[i];
    z.dp = dp; z.rank = rank;
    return 0;
}

inline int zero1_blocks(size_t nvec) {
    size_t b = (nvec + 255) / 256;
    const size_t cap = (size_t)nxd_num_sms() * 8;
    if (b > cap) b = cap;
    return b < 1 ? 1 : (int)b;
}

}  // namespace

// out[0 .. n) (fp32, local) = sum over ranks r of grad_r[lo .. lo + n).  grad_ptrs: dp pointers to every rank's flat
// gradient buffer (symmetric memory; [rank] local), grad_dtype 0 = fp32, 1 = bf16; lo and n in ELEMENTS, multiples of 8.
static int zero1_reduce_impl(void* const* grad_ptrs, int grad_dtype, long long lo, long long n, float* out, int dp, int rank,
                             long long split_at, float* partials, cudaStream_t s) {
    NXD_REQUIRE(lo >= 0 && n >= 0 && lo % 8 == 0 && n % 8 == 0, "zero1_reduce_shard: lo / n must be multiples of 8");
    if (n == 0) return 0;
    ZPeers z{};
    if (int rc = fill_peers(z, grad_ptrs, dp, rank)) return rc;
    if (grad_dtype == 1)
        zero1_reduce_shard_bf16_kernel<<<zero1_blocks((size_t)n >> 3), 256, 0, s>>>(z, (size_t)lo, (size_t)n, out,
                                                                                    (size_t)split_at, partials);
    else
        zero1_reduce_shard_f32_kernel<<<zero1_blocks((size_t)n >> 2), 256, 0, s>>>(z, (size_t)lo, (size_t)n, out,
                                                                                   (size_t)split_at, partials);
    NXD_LAUNCH_CHECK();
    return 0;
}

NXD_API int nxd_zero1_reduce_shard(void* const* grad_ptrs, int grad_dtype, long long lo, long long n, float* out, int dp,
                                   int rank, void* stream) {
    return zero1_reduce_impl(grad_ptrs, grad_dtype, lo, n, out, dp, rank, 0, nullptr, (cudaStream_t)stream);
}

// Blocks the reduce kernel launches for a shard of n elements: the caller provides 2 * blocks floats of scratch.
NXD_API int nxd_zero1_reduce_blocks(long long n, int grad_dtype) {
    return zero1_blocks((size_t)n >> (grad_dtype == 1 ? 3 : 2));
}

// The same reduction with the local part of the gradient norm folded in: sum_sharded[0] += sum of out[i]^2 for
// i < split_at, sum_replicated[0] += the same for i >= split_at (both accumulate; out is the UNAVERAGED sum over the
// ranks).  partials: 2 * nxd_zero1_reduce_blocks(n, grad_dtype) floats of scratch.  0 <= split_at <= n.
NXD_API int nxd_zero1_reduce_shard_sumsq(void* const* grad_ptrs, int grad_dtype, long long lo, long long n, float* out,
                                         int dp, int rank, long long split_at, float* partials, float* sum_sharded,
                                         float* sum_replicated, void* stream) {
    NXD_REQUIRE(partials && sum_sharded && sum_replicated && split_at >= 0 && split_at <= n,
                "zero1_reduce_shard_sumsq: scratch / outputs missing or split_at %lld outside [0, %lld]", split_at, n);
    if (n == 0) return 0;
    if (int rc = zero1_reduce_impl(grad_ptrs, grad_dtype, lo, n, out, dp, rank, split_at, partials, (cudaStream_t)stream))
        return rc;
    zero1_sumsq_finish_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(partials, nxd_zero1_reduce_blocks(n, grad_dtype),
                                                                  sum_sharded, sum_replicated);
    NXD_LAUNCH_CHECK();
    return 0;
}

// AdamW (decoupled weight decay) on my shard; the bf16 result goes to param_r[lo .. lo + n) of EVERY rank r.
// step >= 1; grad is the fp32 shard from nxd_zero1_reduce_shard, multiplied by grad_scale * clip_coef[0].
NXD_API int nxd_zero1_adamw_allgather(float* master, float* exp_avg, float* exp_avg_sq, const float* grad,
                                      void* const* param_ptrs, long long lo, long long n, float lr, float beta1,
                                      float beta2, float eps, float weight_decay, int step, float grad_scale,
                                      const float* clip_coef, int dp, int rank, void* stream) {
    NXD_REQUIRE(lo >= 0 && n >= 0 && lo % 8 == 0 && n % 8 == 0 && step >= 1, "zero1_adamw_allgather: bad arguments");
    if (n == 0) return 0;
    ZPeers z{};
    if (int rc = fill_peers(z, param_ptrs, dp, rank)) return rc;
    const float bc1 = 1.f - powf(beta1, (float)step);
    const float bc2 = 1.f - powf(beta2, (float)step);
    zero1_adamw_allgather_kernel<<<zero1_blocks((size_t)n >> 3), 256, 0, (cudaStream_t)stream>>>(
        master, exp_avg, exp_avg_sq, grad, z, (size_t)lo, (size_t)n, lr, beta1, beta2, eps, weight_decay, bc1, bc2,
        grad_scale, clip_coef);
    NXD_LAUNCH_CHECK();
    return 0;
}
