// Token-generation projections with the neighbouring element-wise work folded in (north-star "decode
// latency"): one weight-streaming kernel per projection instead of RMSNorm + GEMV + RoPE + cache write (or + SwiGLU)
// as separate launches.
//
//   nxd_decode_qkv_rope_cache   q, k, v = W_qkv rmsnorm(x);  RoPE(q, k) at positions[b];  q -> q_out,  k / v -> KV cache
//   nxd_decode_gate_up_swiglu   out = silu(W_gate rmsnorm(x)) * (W_up rmsnorm(x))
//
// Reference: the token-generation graph of examples/inference/modules/model_base.py:459-518 (RmsNorm custom call,
// attention_base.py:172-200 q/k/v projections + RoPE + cache scatter, MLP) -- there every step is its own HLO op inside
// one compiled graph; here a decode layer is 6 launches (these two, attention split + merge, two row-parallel GEMVs
// with the residual in their epilogue).
//
// Structure (both entry points, B <= 8 rows): a CTA owns 8 column PAIRS, warp w owns pair w and streams the two weight
// rows with 16-byte loads (all k steps of a lane in flight), the activations sit normalised in shared memory.  A pair is
// exactly what the epilogue combines: (d, d + D/2) of one head for RoPE, (gate_j, up_j) for SwiGLU.  Lane 0 of the warp
// applies the epilogue to the reduced sums and stores.
#include "decode_common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kWarps = 8;

__device__ __forceinline__ float dot8f(const uint4& a, const uint4& b, float acc) {
    const bf162* pa = reinterpret_cast<const bf162*>(&a);
    const bf162* pb = reinterpret_cast<const bf162*>(&b);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float2 fa = __bfloat1622float2(pa[i]), fb = __bfloat1622float2(pb[i]);
        acc = fmaf(fa.x, fb.x, acc);
        acc = fmaf(fa.y, fb.y, acc);
    }
    return acc;
}

// Programmatic dependent launch (common.cuh): let the successor start, pull this warp's two weight rows towards L2 (they do
// not depend on the predecessor), then wait for the predecessor's activations.
__device__ __forceinline__ void prefetch_rows_then_wait(const bf16* w0, const bf16* w1, int K) {
    pdl_trigger();
    const int lane = threadIdx.x & 31;
    for (int k = lane * 64; k < K; k += 32 * 64) {      // one 128-byte line per lane and step
        prefetch_l2(w0 + k);
        prefetch_l2(w1 + k);
    }
    pdl_wait();
}

// xs[b][k] = bf16(x[b][k] * rsqrt(mean(x[b]^2) + eps) * norm_w[k])  (norm_w == nullptr: plain copy).  All threads.
template <int kB>
__device__ __forceinline__ void load_normalised(const bf16* __restrict__ x, long long ldx, const bf16* __restrict__ norm_w,
                                                float eps, int B, int K, bf16* xs, float* red) {
    const int k8 = K / 8;
    float ss[kB];
#pragma unroll
    for (int b = 0; b < kB; ++b) ss[b] = 0.f;
    for (int c = threadIdx.x; c < k8; c += kThreads) {
#pragma unroll
        for (int b = 0; b < kB; ++b) {
            uint4 v = make_uint4(0u, 0u, 0u, 0u);
            if (b < B) v = *reinterpret_cast<const uint4*>(x + (size_t)b * ldx + (size_t)c * 8);
            reinterpret_cast<uint4*>(xs)[b * k8 + c] = v;
            const bf162* p = reinterpret_cast<const bf162*>(&v);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float2 f = __bfloat1622float2(p[i]);
                ss[b] = fmaf(f.x, f.x, fmaf(f.y, f.y, ss[b]));
            }
        }
    }
    if (norm_w == nullptr) {
        __syncthreads();
        return;
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int b = 0; b < kB; ++b) {
        ss[b] = warp_sum(ss[b]);
        if (lane == 0) red[b * kWarps + warp] = ss[b];
    }
    __syncthreads();
    float rstd[kB];
#pragma unroll
    for (int b = 0; b < kB; ++b) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) t += red[b * kWarps + w];
        rstd[b] = rsqrtf(t / (float)K + eps);
    }
    for (int c = threadIdx.x; c < k8; c += kThreads) {      // every thread rescales the chunks it wrote itself
        float wv[8];
        unpack8(reinterpret_cast<const bf16x8*>(norm_w)[c], wv);
#pragma unroll
        for (int b = 0; b < kB; ++b) {
            float xv[8];
            unpack8(reinterpret_cast<const bf16x8*>(xs)[b * k8 + c], xv);
#pragma unroll
            for (int e = 0; e < 8; ++e) xv[e] = xv[e] * rstd[b] * wv[e];
            reinterpret_cast<bf16x8*>(xs)[b * k8 + c] = pack8(xv);
        }
    }
    __syncthreads();
}

// acc0[b] = <w0, xs[b]>, acc1[b] = <w1, xs[b]>, complete in every lane.
template <int kB>
__device__ __forceinline__ void pair_dot(const bf16* __restrict__ w0, const bf16* __restrict__ w1, const bf16* xs, int K,
                                         float (&acc0)[kB], float (&acc1)[kB]) {
    const int lane = threadIdx.x & 31;
    const int k8 = K / 8;
#pragma unroll
    for (int b = 0; b < kB; ++b) acc0[b] = acc1[b] = 0.f;
#pragma unroll (kB <= 2 ? 8 : kB <= 4 ? 4 : 2)
    for (int k = lane; k < k8; k += 32) {
        const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(w0) + k);
        const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(w1) + k);
#pragma unroll
        for (int b = 0; b < kB; ++b) {
            const uint4 xv = reinterpret_cast<const uint4*>(xs)[b * k8 + k];
            acc0[b] = dot8f(a0, xv, acc0[b]);
            acc1[b] = dot8f(a1, xv, acc1[b]);
        }
    }
#pragma unroll
    for (int b = 0; b < kB; ++b) {
        acc0[b] = warp_sum(acc0[b]);
        acc1[b] = warp_sum(acc1[b]);
    }
}

using nxd_decode::CacheDst;
using nxd_decode::cache_row;

// grid = (n_q + 2 n_kv) heads x (D / 16) pair groups.  w: [(n_q + 2 n_kv) * D, K], head-major rows (q | k | v).
template <int kB>
__global__ void __launch_bounds__(kThreads)
decode_qkv_kernel(const bf16* __restrict__ x, long long ldx, const bf16* __restrict__ norm_w, float eps,
                  const bf16* __restrict__ w, long long ldw, const float* __restrict__ cos_t,
                  const float* __restrict__ sin_t, const int* __restrict__ positions, bf16* __restrict__ q_out,
                  CacheDst cache, int B, int K, int n_q, int n_kv, int D) {
    extern __shared__ __align__(16) uint8_t dsm[];
    bf16* xs = reinterpret_cast<bf16*>(dsm);
    float* red = reinterpret_cast<float*>(dsm + (size_t)kB * K * sizeof(bf16));
    const int groups = D / 16;
    const int head = blockIdx.x / groups;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int d = (blockIdx.x % groups) * 8 + warp;          // 0 .. D/2 - 1: rotates with d + D/2
    const bf16* w0 = w + ((size_t)head * D + d) * ldw;
    const bf16* w1 = w0 + (size_t)(D / 2) * ldw;
    prefetch_rows_then_wait(w0, w1, K);
    load_normalised<kB>(x, ldx, norm_w, eps, B, K, xs, red);
    float a[kB], bq[kB];
    pair_dot<kB>(w0, w1, xs, K, a, bq);
    if (lane != 0) return;
#pragma unroll
    for (int b = 0; b < kB; ++b) {
        if (b >= B) break;
        const int pos = positions[b];
        float o0 = a[b], o1 = bq[b];
        if (head < n_q + n_kv) {                           // q and k heads rotate (values rounded to bf16 first, as
            o0 = __bfloat162float(__float2bfloat16(o0));   // the unfused projection output is)
            o1 = __bfloat162float(__float2bfloat16(o1));
            const float cs = cos_t[(size_t)pos * (D / 2) + d], sn = sin_t[(size_t)pos * (D / 2) + d];
            const float r0 = o0 * cs - o1 * sn, r1 = o1 * cs + o0 * sn;
            o0 = r0;
            o1 = r1;
        }
        bf16* dst;
        if (head < n_q) {
            dst = q_out + ((size_t)b * n_q + head) * D;
        } else if (head < n_q + n_kv) {
            dst = cache.k + cache_row(cache, b, head - n_q, pos, D);
        } else {
            dst = cache.v + cache_row(cache, b, head - n_q - n_kv, pos, D);
        }
        dst[d] = __float2bfloat16(o0);
        dst[d + D / 2] = __float2bfloat16(o1);
    }
}

// grid = I / 8.  w: [2 I, K] with the gate rows first, then the up rows.  out: [B, I].
template <int kB>
__global__ void __launch_bounds__(kThreads)
decode_gate_up_kernel(const bf16* __restrict__ x, long long ldx, const bf16* __restrict__ norm_w, float eps,
                      const bf16* __restrict__ w, long long ldw, bf16* __restrict__ out, long long ldo, int B, int K,
                      int I) {
    extern __shared__ __align__(16) uint8_t dsm[];
    bf16* xs = reinterpret_cast<bf16*>(dsm);
    float* red = reinterpret_cast<float*>(dsm + (size_t)kB * K * sizeof(bf16));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int j = blockIdx.x * 8 + warp;
    const bf16* w0 = w + (size_t)min(j, I - 1) * ldw;
    const bf16* w1 = w + ((size_t)I + min(j, I - 1)) * ldw;
    prefetch_rows_then_wait(w0, w1, K);
    load_normalised<kB>(x, ldx, norm_w, eps, B, K, xs, red);
    if (j >= I) return;
    float g[kB], u[kB];
    pair_dot<kB>(w0, w1, xs, K, g, u);
    if (lane != 0) return;
#pragma unroll
    for (int b = 0; b < kB; ++b) {
        if (b >= B) break;
        const float gv = __bfloat162float(__float2bfloat16(g[b])), uv = __bfloat162float(__float2bfloat16(u[b]));
        out[(size_t)b * ldo + j] = __float2bfloat16(gv / (1.f + __expf(-gv)) * uv);
    }
}

template <typename F>
int set_smem(F kernel, size_t smem) {
    NXD_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    return 0;
}

}  // namespace

// x: [B, K] rows of pitch ldx (the residual stream BEFORE the input norm); norm_w: [K] RMSNorm weight or NULL (x is used as
// is); w: fused projection [(n_q + 2 n_kv) * D, K]; cos / sin: [max_pos, D / 2] fp32 tables; positions: [B] int32;
// q_out: [B, n_q, D]; cache_k / cache_v: contiguous [lines, n_kv, S_max, D] (stride_b / stride_h in elements) or paged
// [pages, n_kv, page, D] with block_table [lines, max_pages]; seq_ids: [B] cache line of each row or NULL (= row index).
// B <= 8, K % 8 == 0, D in {64, 128}.
NXD_API int nxd_decode_qkv_rope_cache(const void* x, long long ldx, const void* norm_w, float eps, const void* w,
                                      long long ldw, const float* cos_t, const float* sin_t, const int* positions,
                                      void* q_out, void* cache_k, void* cache_v, const int* seq_ids,
                                      const int* block_table, long long stride_b, long long stride_h, int page_size,
                                      int max_pages, int B, int K, int n_q, int n_kv, int D, void* stream) {
    NXD_REQUIRE(B >= 1 && B <= 8, "decode_qkv: B %d not in [1, 8]", B);
    NXD_REQUIRE(K % 8 == 0 && ldx % 8 == 0 && ldw % 8 == 0, "decode_qkv: K / ldx / ldw must be multiples of 8");
    NXD_REQUIRE(D == 64 || D == 128, "decode_qkv: head dim %d not in {64, 128}", D);
    const int kb = B <= 1 ? 1 : B <= 2 ? 2 : B <= 4 ? 4 : 8;
    const size_t smem = (size_t)kb * K * 2 + (size_t)kb * kWarps * 4;
    NXD_REQUIRE(smem <= 200 * 1024, "decode_qkv: B %d x K %d does not fit shared memory", B, K);
    CacheDst c{(bf16*)cache_k, (bf16*)cache_v, block_table, seq_ids, stride_b, stride_h, page_size > 0 ? page_size : 1,
               max_pages};
    const int grid = (n_q + 2 * n_kv) * (D / 16);
    cudaStream_t st = (cudaStream_t)stream;
#define LAUNCH(KB)                                                                                                   \
    do {                                                                                                             \
        if (int rc = set_smem(decode_qkv_kernel<KB>, smem)) return rc;                                               \
        NXD_CHECK_CUDA(nxd_launch_pdl(decode_qkv_kernel<KB>, dim3(grid), dim3(kThreads), smem, st, (const bf16*)x, ldx, \
                                      (const bf16*)norm_w, eps, (const bf16*)w, ldw, cos_t, sin_t, positions,        \
                                      (bf16*)q_out, c, B, K, n_q, n_kv, D));                                         \
    } while (0)
    switch (kb) {
        case 1: LAUNCH(1); break;
        case 2: LAUNCH(2); break;
        case 4: LAUNCH(4); break;
        default: LAUNCH(8); break;
    }
#undef LAUNCH
    NXD_LAUNCH_CHECK();
    return 0;
}

// out[B, I] = silu(W_gate n(x)) * (W_up n(x)),  n = RMSNorm with norm_w (NULL: identity);  w: [2 I, K], gate rows first.
NXD_API int nxd_decode_gate_up_swiglu(const void* x, long long ldx, const void* norm_w, float eps, const void* w,
                                      long long ldw, void* out, long long ldo, int B, int K, int I, void* stream) {
    NXD_REQUIRE(B >= 1 && B <= 8, "decode_gate_up: B %d not in [1, 8]", B);
    NXD_REQUIRE(K % 8 == 0 && ldx % 8 == 0 && ldw % 8 == 0, "decode_gate_up: K / ldx / ldw must be multiples of 8");
    const int kb = B <= 1 ? 1 : B <= 2 ? 2 : B <= 4 ? 4 : 8;
    const size_t smem = (size_t)kb * K * 2 + (size_t)kb * kWarps * 4;
    NXD_REQUIRE(smem <= 200 * 1024, "decode_gate_up: B %d x K %d does not fit shared memory", B, K);
    const int grid = (I + 7) / 8;
    cudaStream_t st = (cudaStream_t)stream;
#define LAUNCH(KB)                                                                                                  \
    do {                                                                                                            \
        if (int rc = set_smem(decode_gate_up_kernel<KB>, smem)) return rc;                                          \
        NXD_CHECK_CUDA(nxd_launch_pdl(decode_gate_up_kernel<KB>, dim3(grid), dim3(kThreads), smem, st, (const bf16*)x, \
                                      ldx, (const bf16*)norm_w, eps, (const bf16*)w, ldw, (bf16*)out, ldo, B, K, I)); \
    } while (0)
    switch (kb) {
        case 1: LAUNCH(1); break;
        case 2: LAUNCH(2); break;
        case 4: LAUNCH(4); break;
        default: LAUNCH(8); break;
    }
#undef LAUNCH
    NXD_LAUNCH_CHECK();
    return 0;
}
