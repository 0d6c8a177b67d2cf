// Decode-path fused kernels over NVLink peer memory:
//
//   nxd_skinny_gemm_ar_bf16   out[B, N] = all_reduce_r( X_r[B, Kloc] * W_r[N, Kloc]^T )      B <= 16 (token generation)
//   nxd_allreduce_oneshot     out[n]    = all_reduce_r( in_r[n] )                            small messages
//
// Reference: RowParallelLinear forward (src/neuronx_distributed/parallel_layers/layers.py:452-468) is `matmul` then
// `reduce_from_tensor_model_parallel_region` (mappings.py:176-190, an XLA all-reduce): two dependent ops whose cost in
// token generation is pure latency (a [4, 3072] activation is 24 KB).  Here both halves are ONE launch:
//
//   * the weight matrix is streamed exactly once (the op is weight-bandwidth bound): every warp owns output columns,
//     the B activation rows sit in shared memory, fp32 accumulation;
//   * each CTA pushes its fp32 partial block [B, cols] into the SAME block of every peer's staging slot
//     (slot = [parity][source rank]) with 16-byte stores over NVLink, then stamps flag [cta][my rank] on every peer
//     with the call's epoch (st.release.sys) -- a "one-shot" all-reduce: one NVLink write latency + one flag
//     latency, no ring steps;
//   * the same CTA waits until its tp - 1 flags carry the epoch (stamps only grow, nothing is ever reset, calls with
//     different grids can share the flags), adds the blocks in RANK ORDER (so every rank produces bit-identical
//     sums, in fp32 -- NCCL's bf16 ring rounds at every hop) and writes bf16.
//
// No barrier is needed for buffer reuse: staging is double-buffered by call parity, and a rank cannot finish call
// n + 1 before every peer has ENTERED call n + 1 (it needs their partials), i.e. finished reading the buffer of call n,
// which is the one call n + 2 overwrites.  All CTAs of the grid are co-resident (the host sizes the grid from the
// occupancy query), so the spin-wait cannot starve the producers.
#include "common.cuh"

namespace {

constexpr int kMaxTp = 8;
constexpr int kArThreads = 256;
constexpr int kArWarps = kArThreads / 32;
constexpr int kMaxCtas = 512;            // flags reserved per staging region: kMaxCtas * kMaxTp

struct ArArgs {
    float* stage_peer[kMaxTp];     // every rank's staging base: [2][tp][slot_elems] floats (peer-mapped, [rank] local)
    uint32_t* sig_peer[kMaxTp];    // peers' signal pads
    uint32_t* sig_local;
    size_t slot_elems;             // floats per (parity, source) slot
    int slot;                      // first flag; flag [cta][source rank]
    int tp, rank;
    uint32_t* state;               // local device memory: [0] = calls completed on this staging region, [1] = CTAs done
};

__device__ __forceinline__ void ar_st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ar_ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 ar_ld_cg(const float4* p) {
    float4 v;
    asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

// Push `n4` float4 of `part` (shared memory) into block offset `off` of my slot on every peer, signal, wait for the
// peers' blocks, reduce in rank order.  Result (fp32) is left in `part`.
__device__ __forceinline__ void oneshot_exchange(const ArArgs& ar, uint32_t epoch, float* part, int n4, size_t off,
                                                 int cta) {
    const int parity = (int)(epoch & 1u);
    const size_t my_slot = ((size_t)parity * ar.tp + ar.rank) * ar.slot_elems + off;
    // (peer, vector) pairs spread over the CTA: all NVLink stores are in flight at once
    for (int i = threadIdx.x; i < n4 * (ar.tp - 1); i += blockDim.x) {
        const int d = i / n4 + 1, v = i % n4;
        const int peer = (ar.rank + d) % ar.tp;
        reinterpret_cast<float4*>(ar.stage_peer[peer] + my_slot)[v] = reinterpret_cast<const float4*>(part)[v];
    }
    __syncthreads();
    if (threadIdx.x < ar.tp - 1) {
        const int peer = (ar.rank + 1 + (int)threadIdx.x) % ar.tp;
        __threadfence_system();
        ar_st_release_sys(ar.sig_peer[peer] + ar.slot + cta * kMaxTp + ar.rank, epoch);
        const uint32_t* flag = ar.sig_local + ar.slot + cta * kMaxTp + peer;
        while ((int32_t)(ar_ld_acquire_sys(flag) - epoch) < 0) {}
    }
    __syncthreads();
    const float* mine = ar.stage_peer[ar.rank] + (size_t)parity * ar.tp * ar.slot_elems + off;
    for (int v = threadIdx.x; v < n4; v += blockDim.x) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int r = 0; r < ar.tp; ++r) {
            const float4 x = r == ar.rank ? reinterpret_cast<const float4*>(part)[v]
                                          : ar_ld_cg(reinterpret_cast<const float4*>(mine + (size_t)r * ar.slot_elems) + v);
            acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w;
        }
        reinterpret_cast<float4*>(part)[v] = acc;
    }
    __syncthreads();
}

// The call counter lives in device memory so that a CUDA graph that captured the launch replays correctly: every CTA
// reads it on entry, the last CTA to finish advances it (kernels of one stream do not overlap).
__device__ __forceinline__ uint32_t ar_begin(const ArArgs& ar) {
    __shared__ uint32_t s_epoch;
    if (threadIdx.x == 0) s_epoch = *reinterpret_cast<volatile uint32_t*>(ar.state) + 1u;
    __syncthreads();
    return s_epoch;
}
__device__ __forceinline__ void ar_end(const ArArgs& ar, uint32_t epoch) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(ar.state + 1, 1u) == gridDim.x - 1) {
            ar.state[1] = 0u;
            __threadfence();
            *reinterpret_cast<volatile uint32_t*>(ar.state) = epoch;
        }
    }
}

__device__ __forceinline__ float dot8(const uint4& a, const uint4& b, float acc) {
    const __nv_bfloat162* pa = reinterpret_cast<const __nv_bfloat162*>(&a);
    const __nv_bfloat162* pb = reinterpret_cast<const __nv_bfloat162*>(&b);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float2 fa = __bfloat1622float2(pa[i]), fb = __bfloat1622float2(pb[i]);
        acc = fmaf(fa.x, fb.x, acc);
        acc = fmaf(fa.y, fb.y, acc);
    }
    return acc;
}

// Shared memory: X tile [kB][K] bf16, then partial block [kB][cols] fp32.
template <int kB>
__global__ void __launch_bounds__(kArThreads) skinny_gemm_ar_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w,
                                                                    const bf16* __restrict__ residual, bf16* __restrict__ out,
                                                                    int B, int N, int K, long long ldx, long long ldw,
                                                                    long long ldo, int cols, const ArArgs ar) {
    extern __shared__ __align__(16) uint8_t ar_smem[];
    bf16* xs = reinterpret_cast<bf16*>(ar_smem);
    float* part = reinterpret_cast<float*>(ar_smem + (size_t)kB * K * sizeof(bf16));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int k8 = K / 8;
    const int c0 = (int)blockIdx.x * cols;
    const int ncols = min(cols, N - c0);
    // programmatic dependent launch: this kernel may start while its predecessor still runs.  The weights do not depend
    // on it: pull the rows of the first column pass towards L2, THEN wait for the activations.
    pdl_trigger();
    if (warp * 2 < ncols) {
        const bf16* p0 = w + (size_t)(c0 + warp * 2) * ldw;
        const bool two = warp * 2 + 1 < ncols;
        for (int ln = lane; ln * 64 < K; ln += 32) {
            prefetch_l2(p0 + ln * 64);
            if (two) prefetch_l2(p0 + ldw + ln * 64);
        }
    }
    pdl_wait();
    for (int i = threadIdx.x; i < kB * k8; i += kArThreads) {
        const int row = i / k8, c = i % k8;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (row < B) v = *reinterpret_cast<const uint4*>(x + (size_t)row * ldx + (size_t)c * 8);
        reinterpret_cast<uint4*>(xs)[i] = v;
    }
    __syncthreads();
    // two columns per warp pass: 2 x (K / 256) independent 16-byte weight loads in flight per lane
    for (int cc = warp * 2; cc < cols; cc += kArWarps * 2) {
        float acc0[kB], acc1[kB];
#pragma unroll
        for (int b = 0; b < kB; ++b) acc0[b] = acc1[b] = 0.f;
        const bool v0 = cc < ncols, v1 = cc + 1 < ncols;
        const bf16* w0 = w + (size_t)(c0 + cc) * ldw;
        const bf16* w1 = w0 + ldw;
        // issue the weight loads of several k steps back to back: the loop is bound by load latency, not by math
#pragma unroll (kB <= 2 ? 8 : kB <= 4 ? 4 : 2)
        for (int k = lane; k < k8; k += 32) {
            const uint4 a0 = v0 ? __ldg(reinterpret_cast<const uint4*>(w0) + k) : make_uint4(0u, 0u, 0u, 0u);
            const uint4 a1 = v1 ? __ldg(reinterpret_cast<const uint4*>(w1) + k) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
            for (int b = 0; b < kB; ++b) {
                const uint4 xv = reinterpret_cast<const uint4*>(xs)[b * k8 + k];
                acc0[b] = dot8(a0, xv, acc0[b]);
                acc1[b] = dot8(a1, xv, acc1[b]);
            }
        }
#pragma unroll
        for (int b = 0; b < kB; ++b) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                acc0[b] += __shfl_xor_sync(0xffffffffu, acc0[b], o);
                acc1[b] += __shfl_xor_sync(0xffffffffu, acc1[b], o);
            }
        }
        if (lane == 0) {
#pragma unroll
            for (int b = 0; b < kB; ++b) {
                part[b * cols + cc] = acc0[b];
                part[b * cols + cc + 1] = acc1[b];
            }
        }
    }
    __syncthreads();
    uint32_t epoch = 0;
    if (ar.tp > 1) {
        epoch = ar_begin(ar);
        oneshot_exchange(ar, epoch, part, kB * cols / 4, (size_t)blockIdx.x * kB * cols, (int)blockIdx.x);
    }
    for (int i = threadIdx.x; i < B * ncols; i += kArThreads) {
        const int b = i / ncols, c = i % ncols;
        float v = part[b * cols + c];
        if (residual) v += __bfloat162float(residual[(size_t)b * ldo + c0 + c]);
        out[(size_t)b * ldo + c0 + c] = __float2bfloat16(v);
    }
    if (ar.tp > 1) ar_end(ar, epoch);
}

// ---- single GPU, long K (down projection, K = intermediate size): with one column pair per warp a warp walks K in
// several dependent batches of loads and the kernel is latency bound (33 MB in ~22 us).  Here a CTA owns only two
// column pairs and its 8 warps split K four ways, so every warp has ONE batch of loads in flight and four times as
// many warps share the stream; the four partial sums per output meet in shared memory.
template <int kB>
__global__ void __launch_bounds__(kArThreads) skinny_splitk_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w,
                                                                   const bf16* __restrict__ residual, bf16* __restrict__ out,
                                                                   int B, int N, int K, long long ldx, long long ldw,
                                                                   long long ldo) {
    constexpr int kSplit = 4, kPairs = kArWarps / kSplit, kCols = 2 * kPairs;
    extern __shared__ __align__(16) uint8_t ar_smem[];
    bf16* xs = reinterpret_cast<bf16*>(ar_smem);
    float* part = reinterpret_cast<float*>(ar_smem + (size_t)kB * K * sizeof(bf16));   // [kSplit][kB][kCols]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int k8 = K / 8;
    const int pair = warp % kPairs, ks = warp / kPairs;
    const int c0 = (int)blockIdx.x * kCols + pair * 2;
    const bool v0 = c0 < N, v1 = c0 + 1 < N;
    const bf16* w0 = w + (size_t)(v0 ? c0 : 0) * ldw;
    const bf16* w1 = w + (size_t)(v1 ? c0 + 1 : 0) * ldw;
    const int per = (k8 + kSplit - 1) / kSplit;
    const int kbeg = ks * per, kend = min(k8, kbeg + per);
    // programmatic dependent launch: weights towards L2 first (they do not depend on the predecessor), then wait
    pdl_trigger();
    for (int k = kbeg + lane * 8; k < kend; k += 256) {      // one 128-byte line = 8 vectors of 8 bf16
        prefetch_l2(w0 + (size_t)k * 8);
        prefetch_l2(w1 + (size_t)k * 8);
    }
    pdl_wait();
    for (int i = threadIdx.x; i < kB * k8; i += kArThreads) {
        const int row = i / k8, c = i % k8;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (row < B) v = *reinterpret_cast<const uint4*>(x + (size_t)row * ldx + (size_t)c * 8);
        reinterpret_cast<uint4*>(xs)[i] = v;
    }
    __syncthreads();
    float acc0[kB], acc1[kB];
#pragma unroll
    for (int b = 0; b < kB; ++b) acc0[b] = acc1[b] = 0.f;
#pragma unroll (kB <= 2 ? 8 : 4)
    for (int k = kbeg + lane; k < kend; k += 32) {
        const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(w0) + k);
        const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(w1) + k);
#pragma unroll
        for (int b = 0; b < kB; ++b) {
            const uint4 xv = reinterpret_cast<const uint4*>(xs)[b * k8 + k];
            acc0[b] = dot8(a0, xv, acc0[b]);
            acc1[b] = dot8(a1, xv, acc1[b]);
        }
    }
#pragma unroll
    for (int b = 0; b < kB; ++b) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            acc0[b] += __shfl_xor_sync(0xffffffffu, acc0[b], o);
            acc1[b] += __shfl_xor_sync(0xffffffffu, acc1[b], o);
        }
    }
    if (lane == 0) {
#pragma unroll
        for (int b = 0; b < kB; ++b) {
            part[(ks * kB + b) * kCols + pair * 2] = acc0[b];
            part[(ks * kB + b) * kCols + pair * 2 + 1] = acc1[b];
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < B * kCols; i += kArThreads) {
        const int b = i / kCols, c = i % kCols;
        const int col = (int)blockIdx.x * kCols + c;
        if (col >= N) continue;
        float v = 0.f;
#pragma unroll
        for (int s2 = 0; s2 < kSplit; ++s2) v += part[(s2 * kB + b) * kCols + c];
        if (residual) v += __bfloat162float(residual[(size_t)b * ldo + col]);
        out[(size_t)b * ldo + col] = __float2bfloat16(v);
    }
}

// ---- 9 .. 16 rows: the FMA loop above re-reads every X row from shared memory per weight vector and becomes FMA /
// LDS bound; here the 16 x K activation tile feeds mma.sync m16n8k16 (M = 16 is exactly one fragment).  A CTA owns
// `cols` output columns and walks them 8 at a time; its 8 warps split K, every lane fetches 16 contiguous bytes of a
// weight row per step (8 k values that serve two MMAs -- the contraction order is free, so A is read from shared
// memory with the same permutation), the warps' partial 16 x 8 tiles are summed through shared memory.
__device__ __forceinline__ void mma_m16n8k16_bf16(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                                  uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// Shared memory: X tile [16][K + 8] bf16 (rows padded by 16 bytes: conflict-free 16-byte reads), warp partials
// [8][16][8] fp32, result block [16][cols] fp32.
__global__ void __launch_bounds__(kArThreads) skinny16_gemm_ar_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w,
                                                                      const bf16* __restrict__ residual, bf16* __restrict__ out,
                                                                      int B, int N, int K, long long ldx, long long ldw,
                                                                      long long ldo, int cols, const ArArgs ar) {
    extern __shared__ __align__(16) uint8_t ar_smem[];
    const int pitch = K + 8;                                           // elements
    bf16* xs = reinterpret_cast<bf16*>(ar_smem);
    float* wpart = reinterpret_cast<float*>(ar_smem + (size_t)16 * pitch * sizeof(bf16));     // [8 warps][16][8]
    float* part = wpart + kArWarps * 128;                                                      // [16][cols]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int k8 = K / 8;
    pdl_trigger();
    pdl_wait();
    for (int i = threadIdx.x; i < 16 * k8; i += kArThreads) {
        const int row = i / k8, c = i % k8;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (row < B) v = *reinterpret_cast<const uint4*>(x + (size_t)row * ldx + (size_t)c * 8);
        *reinterpret_cast<uint4*>(xs + (size_t)row * pitch + c * 8) = v;
    }
    __syncthreads();
    const int c0 = (int)blockIdx.x * cols;
    const int ncols = min(cols, N - c0);
    // K is cut into chunks of 32 (4 lanes x 8 values); warp w takes chunks w, w + 8, ...
    const int chunks = K / 32;
    const int r = lane >> 2, kq = (lane & 3) * 8;
    for (int g = 0; g < cols; g += 8) {
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        const int n = c0 + g + r;                                      // the weight row (output column) of this lane
        const bool nv = g + r < ncols;
        const bf16* wrow = w + (size_t)(nv ? n : c0) * ldw;
#pragma unroll 4
        for (int ch = warp; ch < chunks; ch += kArWarps) {
            const int k0 = ch * 32 + kq;
            uint4 bv = nv ? __ldg(reinterpret_cast<const uint4*>(wrow + k0)) : make_uint4(0u, 0u, 0u, 0u);
            const uint4 alo = *reinterpret_cast<const uint4*>(xs + (size_t)r * pitch + k0);          // row r
            const uint4 ahi = *reinterpret_cast<const uint4*>(xs + (size_t)(r + 8) * pitch + k0);    // row r + 8
            // 8 k values per lane = two MMAs: slots (k, k+1 | k+8, k+9) take values (0,1 | 2,3) then (4,5 | 6,7)
            mma_m16n8k16_bf16(acc, alo.x, ahi.x, alo.y, ahi.y, bv.x, bv.y);
            mma_m16n8k16_bf16(acc, alo.z, ahi.z, alo.w, ahi.w, bv.z, bv.w);
        }
        if ((K & 31) != 0 && warp == 0) {                              // K % 32 tail (K % 8 == 0): zero-padded chunk
            const int k0 = chunks * 32 + kq;
            const bool kv = k0 < K;
            uint4 z = make_uint4(0u, 0u, 0u, 0u);
            const uint4 bv = (nv && kv) ? __ldg(reinterpret_cast<const uint4*>(wrow + k0)) : z;
            const uint4 alo = kv ? *reinterpret_cast<const uint4*>(xs + (size_t)r * pitch + k0) : z;
            const uint4 ahi = kv ? *reinterpret_cast<const uint4*>(xs + (size_t)(r + 8) * pitch + k0) : z;
            mma_m16n8k16_bf16(acc, alo.x, ahi.x, alo.y, ahi.y, bv.x, bv.y);
            mma_m16n8k16_bf16(acc, alo.z, ahi.z, alo.w, ahi.w, bv.z, bv.w);
        }
        // C fragment: acc[0..1] = row r, columns 2 (lane % 4) + {0, 1}; acc[2..3] = row r + 8
        float* wp = wpart + warp * 128;
        const int cc = (lane & 3) * 2;
        wp[r * 8 + cc] = acc[0]; wp[r * 8 + cc + 1] = acc[1];
        wp[(r + 8) * 8 + cc] = acc[2]; wp[(r + 8) * 8 + cc + 1] = acc[3];
        __syncthreads();
        if (threadIdx.x < 128) {
            float sum = 0.f;
#pragma unroll
            for (int ww = 0; ww < kArWarps; ++ww) sum += wpart[ww * 128 + threadIdx.x];
            part[(threadIdx.x >> 3) * cols + g + (threadIdx.x & 7)] = sum;
        }
        __syncthreads();
    }
    uint32_t epoch = 0;
    if (ar.tp > 1) {
        epoch = ar_begin(ar);
        oneshot_exchange(ar, epoch, part, 16 * cols / 4, (size_t)blockIdx.x * 16 * cols, (int)blockIdx.x);
    }
    for (int i = threadIdx.x; i < B * ncols; i += kArThreads) {
        const int b = i / ncols, c = i % ncols;
        float v = part[b * cols + c];
        if (residual) v += __bfloat162float(residual[(size_t)b * ldo + c0 + c]);
        out[(size_t)b * ldo + c0 + c] = __float2bfloat16(v);
    }
    if (ar.tp > 1) ar_end(ar, epoch);
}

// Stand-alone one-shot all-reduce of n bf16 values (n % 8 == 0); every CTA owns `per_cta` elements (multiple of 8).
__global__ void __launch_bounds__(kArThreads) allreduce_oneshot_kernel(const bf16* __restrict__ in, bf16* __restrict__ out,
                                                                       long long n, int per_cta, const ArArgs ar) {
    extern __shared__ __align__(16) uint8_t ar_smem[];
    float* part = reinterpret_cast<float*>(ar_smem);
    const long long e0 = (long long)blockIdx.x * per_cta;
    const int cnt = (int)min((long long)per_cta, n - e0);
    for (int i = threadIdx.x; i < per_cta / 8; i += kArThreads) {
        float f[8];
        if (i * 8 < cnt) {
            const uint4 v = *reinterpret_cast<const uint4*>(in + e0 + (size_t)i * 8);
            const __nv_bfloat162* p = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 t = __bfloat1622float2(p[j]);
                f[2 * j] = t.x; f[2 * j + 1] = t.y;
            }
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] = 0.f;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) part[i * 8 + j] = f[j];
    }
    const uint32_t epoch = ar_begin(ar);
    oneshot_exchange(ar, epoch, part, per_cta / 4, (size_t)blockIdx.x * per_cta, (int)blockIdx.x);
    for (int i = threadIdx.x; i < cnt / 8; i += kArThreads) {
        uint4 v;
        __nv_bfloat162* p = reinterpret_cast<__nv_bfloat162*>(&v);
#pragma unroll
        for (int j = 0; j < 4; ++j) p[j] = __floats2bfloat162_rn(part[i * 8 + 2 * j], part[i * 8 + 2 * j + 1]);
        *reinterpret_cast<uint4*>(out + e0 + (size_t)i * 8) = v;
    }
    ar_end(ar, epoch);
}

int fill_args(ArArgs& ar, void* const* stage_ptrs, void* const* sig_ptrs, int tp, int rank, int slot, void* state,
              size_t slot_elems) {
    NXD_REQUIRE(tp >= 1 && tp <= kMaxTp && rank >= 0 && rank < tp, "oneshot all-reduce: tp %d / rank %d out of range", tp, rank);
    NXD_REQUIRE(tp == 1 || (stage_ptrs && sig_ptrs && state), "oneshot all-reduce: missing symmetric buffers / state");
    for (int i = 0; i < tp && tp > 1; ++i) {
        ar.stage_peer[i] = (float*)stage_ptrs[i];
        ar.sig_peer[i] = (uint32_t*)sig_ptrs[i];
    }
    ar.sig_local = tp > 1 ? (uint32_t*)sig_ptrs[rank] : nullptr;
    ar.slot_elems = slot_elems; ar.slot = slot; ar.tp = tp; ar.rank = rank; ar.state = (uint32_t*)state;
    return 0;
}

template <int kB>
int launch_skinny(const bf16* x, const bf16* w, const bf16* residual, bf16* out, int B, int N, int K, long long ldx,
                  long long ldw, long long ldo, const ArArgs& ar, cudaStream_t st) {
    if (ar.tp == 1 && K >= 4096 && kB <= 4) {     // long contraction on one GPU: split K inside the CTA
        const size_t smem_k = (size_t)kB * K * sizeof(bf16) + (size_t)4 * kB * 4 * sizeof(float);
        NXD_REQUIRE(smem_k <= 200 * 1024, "skinny_gemm_ar: B %d x K %d does not fit shared memory", B, K);
        NXD_CHECK_CUDA(cudaFuncSetAttribute(skinny_splitk_kernel<kB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_k));
        NXD_CHECK_CUDA(nxd_launch_pdl(skinny_splitk_kernel<kB>, dim3((N + 3) / 4), dim3(kArThreads), smem_k, st, x, w, residual,
                                      out, B, N, K, ldx, ldw, ldo));
        NXD_LAUNCH_CHECK();
        return 0;
    }
    // columns per CTA: multiple of 16 (8 warps x 2 columns), as few as co-residency allows
    int dev = 0, sms = 0;
    NXD_CHECK_CUDA(cudaGetDevice(&dev));
    NXD_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    int cols = 16, grid = 0;
    size_t smem = 0;
    for (;; cols += 16) {
        smem = (size_t)kB * K * sizeof(bf16) + (size_t)kB * cols * sizeof(float);
        NXD_REQUIRE(smem <= 200 * 1024, "skinny_gemm_ar: B %d x K %d does not fit shared memory", B, K);
        NXD_CHECK_CUDA(cudaFuncSetAttribute(skinny_gemm_ar_kernel<kB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int per_sm = 0;
        NXD_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, skinny_gemm_ar_kernel<kB>, kArThreads, smem));
        NXD_REQUIRE(per_sm >= 1, "skinny_gemm_ar: kernel does not fit an SM");
        grid = (N + cols - 1) / cols;
        if (ar.tp == 1) break;            // no spin-wait, no flags: any grid works, one column pair per warp
        if (grid <= min(min(per_sm, 2) * sms, kMaxCtas)) break;
        NXD_REQUIRE(cols < 8192, "skinny_gemm_ar: cannot make N %d co-resident", N);
    }
    NXD_REQUIRE(ar.tp == 1 || (size_t)grid * kB * cols <= ar.slot_elems,
                "skinny_gemm_ar: staging slot too small (%zu floats, need %zu)", ar.slot_elems, (size_t)grid * kB * cols);
    NXD_CHECK_CUDA(nxd_launch_pdl(skinny_gemm_ar_kernel<kB>, dim3(grid), dim3(kArThreads), smem, st, x, w, residual, out, B, N,
                                  K, ldx, ldw, ldo, cols, ar));
    NXD_LAUNCH_CHECK();
    return 0;
}

int launch_skinny16(const bf16* x, const bf16* w, const bf16* residual, bf16* out, int B, int N, int K, long long ldx,
                    long long ldw, long long ldo, const ArArgs& ar, cudaStream_t st) {
    int dev = 0, sms = 0;
    NXD_CHECK_CUDA(cudaGetDevice(&dev));
    NXD_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    int cols = 8, grid = 0;
    size_t smem = 0;
    for (;; cols += 8) {
        smem = (size_t)16 * (K + 8) * sizeof(bf16) + (size_t)kArWarps * 128 * sizeof(float) + (size_t)16 * cols * sizeof(float);
        NXD_REQUIRE(smem <= 200 * 1024, "skinny_gemm_ar: 16 x K %d does not fit shared memory", K);
        NXD_CHECK_CUDA(cudaFuncSetAttribute(skinny16_gemm_ar_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int per_sm = 0;
        NXD_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, skinny16_gemm_ar_kernel, kArThreads, smem));
        NXD_REQUIRE(per_sm >= 1, "skinny_gemm_ar: kernel does not fit an SM");
        grid = (N + cols - 1) / cols;
        if (ar.tp == 1 || grid <= min(min(per_sm, 2) * sms, kMaxCtas)) break;
        NXD_REQUIRE(cols < 8192, "skinny_gemm_ar: cannot make N %d co-resident", N);
    }
    NXD_REQUIRE(ar.tp == 1 || (size_t)grid * 16 * cols <= ar.slot_elems,
                "skinny_gemm_ar: staging slot too small (%zu floats, need %zu)", ar.slot_elems, (size_t)grid * 16 * cols);
    NXD_CHECK_CUDA(nxd_launch_pdl(skinny16_gemm_ar_kernel, dim3(grid), dim3(kArThreads), smem, st, x, w, residual, out, B, N, K,
                                  ldx, ldw, ldo, cols, ar));
    NXD_LAUNCH_CHECK();
    return 0;
}

}  // namespace

// uint32 flags a staging region needs (one per CTA and source rank).
NXD_API int nxd_oneshot_num_flags(void) { return kMaxCtas * kMaxTp; }

// out[B, N] (+ residual[B, N], same pitch as out) = sum over ranks of x[B, K] * w[N, K]^T.  B <= 16, K % 8 == 0,
// N % 4 == 0, rows of x / w 16-byte aligned.  tp == 1: plain skinny GEMM (stage_ptrs / sig_ptrs may be NULL).
// stage_ptrs: every rank's staging region of 2 * tp * slot_elems floats, slot_elems >= 16 * (N + 8192) is always
// enough; sig_ptrs: signal pads; slot: first of nxd_oneshot_num_flags() flags that belong to this staging region;
// state: 2 zero-initialised uint32 in LOCAL device memory owned by this staging region (call counter + CTA counter;
// kept on the device so that the launch can be captured in a CUDA graph and replayed).
NXD_API int nxd_skinny_gemm_ar_bf16(const void* x, const void* w, const void* residual, void* out, int B, int N, int K,
                                    long long ldx, long long ldw, long long ldo, void* const* stage_ptrs,
                                    void* const* sig_ptrs, size_t slot_elems, int tp, int rank, int slot, void* state,
                                    void* stream) {
    NXD_REQUIRE(B >= 1 && B <= 16, "skinny_gemm_ar: B %d not in [1, 16]", B);
    NXD_REQUIRE(K % 8 == 0 && N % 4 == 0 && ldx % 8 == 0 && ldw % 8 == 0,
                "skinny_gemm_ar: K / ldx / ldw must be multiples of 8 and N of 4");
    ArArgs ar{};
    if (int rc = fill_args(ar, stage_ptrs, sig_ptrs, tp, rank, slot, state, slot_elems)) return rc;
    const bf16 *xb = (const bf16*)x, *wb = (const bf16*)w, *rb = (const bf16*)residual;
    cudaStream_t st = (cudaStream_t)stream;
    if (B <= 1) return launch_skinny<1>(xb, wb, rb, (bf16*)out, B, N, K, ldx, ldw, ldo, ar, st);
    if (B <= 2) return launch_skinny<2>(xb, wb, rb, (bf16*)out, B, N, K, ldx, ldw, ldo, ar, st);
    if (B <= 4) return launch_skinny<4>(xb, wb, rb, (bf16*)out, B, N, K, ldx, ldw, ldo, ar, st);
    if (B <= 8) return launch_skinny<8>(xb, wb, rb, (bf16*)out, B, N, K, ldx, ldw, ldo, ar, st);
    return launch_skinny16(xb, wb, rb, (bf16*)out, B, N, K, ldx, ldw, ldo, ar, st);   // tensor cores (mma.sync)
}

// out[n] = sum over ranks of in[n] (bf16, fp32 accumulation in rank order).  n % 8 == 0, n <= slot_elems.
NXD_API int nxd_allreduce_oneshot(const void* in, void* out, long long n, void* const* stage_ptrs, void* const* sig_ptrs,
                                  size_t slot_elems, int tp, int rank, int slot, void* state, void* stream) {
    NXD_REQUIRE(n > 0 && n % 8 == 0, "allreduce_oneshot: n %lld must be a positive multiple of 8", n);
    ArArgs ar{};
    if (int rc = fill_args(ar, stage_ptrs, sig_ptrs, tp, rank, slot, state, slot_elems)) return rc;
    if (tp == 1) {
        if (in != out) NXD_CHECK_CUDA(cudaMemcpyAsync(out, in, (size_t)n * 2, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
        return 0;
    }
    int dev = 0, sms = 0;
    NXD_CHECK_CUDA(cudaGetDevice(&dev));
    NXD_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    // 2048 elements (8 KB of fp32) per CTA until the grid would exceed the SM count, then larger blocks
    int per_cta = 2048;
    while ((n + per_cta - 1) / per_cta > min(sms, kMaxCtas)) per_cta *= 2;
    NXD_REQUIRE((size_t)per_cta * 4 <= 160 * 1024, "allreduce_oneshot: n %lld too large for the one-shot path", n);
    const int grid = (int)((n + per_cta - 1) / per_cta);
    NXD_REQUIRE((size_t)grid * per_cta <= slot_elems, "allreduce_oneshot: staging slot too small");
    const size_t smem = (size_t)per_cta * 4;
    NXD_CHECK_CUDA(cudaFuncSetAttribute(allreduce_oneshot_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    allreduce_oneshot_kernel<<<grid, kArThreads, smem, (cudaStream_t)stream>>>((const bf16*)in, (bf16*)out, n, per_cta, ar);
    NXD_LAUNCH_CHECK();
    return 0;
}
