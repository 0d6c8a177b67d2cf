// On-device sampling for sm_100a.
//
// Reference: src/neuronx_distributed/utils/sampling.py:27-77 -- greedy when top_k == 1, else
// top-k -> softmax -> cumsum -> compare against a uniform draw (it avoids torch.multinomial so
// the whole thing stays inside the compiled graph).  Here it is ONE kernel per [B, V] logits
// matrix, one CTA per row:
//   1. exact k-th largest value by 3-round radix select on order-preserving float keys
//      (11 + 11 + 10 bits, shared-memory histograms);
//   2. candidates >= threshold are compacted into shared memory (ties broken by lower index),
//      sorted descending (k <= 1024), softmaxed with temperature (optional top-p cut) and the
//      token is drawn by inverse CDF against the caller's uniform number.
// Greedy (top_k == 1) is a plain block arg-max.
#include "common.cuh"

namespace {

constexpr int kMaxTopK = 1024;
constexpr int kThreads = 1024;

__device__ __forceinline__ uint32_t order_key(float f) {
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);  // larger float -> larger key
}

template <typename T>
__device__ __forceinline__ float load_logit(const T* p, size_t i);
template <>
__device__ __forceinline__ float load_logit<bf16>(const bf16* p, size_t i) { return __bfloat162float(p[i]); }
template <>
__device__ __forceinline__ float load_logit<float>(const float* p, size_t i) { return p[i]; }

template <typename T>
__global__ void argmax_kernel(const T* __restrict__ logits, int V, long long row_stride,
                              long long* __restrict__ out) {
    __shared__ float sv[32];
    __shared__ int si[32];
    const T* z = logits + (size_t)blockIdx.x * row_stride;
    float best = -INFINITY;
    int bi = 0x7fffffff;
    int done = 0;
    if (sizeof(T) == 2 && (V % 8) == 0 && (row_stride % 8) == 0 && (reinterpret_cast<uintptr_t>(logits) & 15) == 0) {
        // 16-byte loads, two in flight per thread: the scalar loop below is latency bound (58 us for one row of 128 k)
        const bf16x8* zv = reinterpret_cast<const bf16x8*>(z);
        const int nv = V / 8;
        for (int c = threadIdx.x; c < nv; c += 2 * blockDim.x) {
            const int c2 = c + blockDim.x;
            const bf16x8 p0 = zv[c];
            bf16x8 p1 = p0;
            if (c2 < nv) p1 = zv[c2];
            float f[8];
            unpack8(p0, f);
#pragma unroll
            for (int e = 0; e < 8; ++e)
                if (f[e] > best || (f[e] == best && c * 8 + e < bi)) { best = f[e]; bi = c * 8 + e; }
            if (c2 < nv) {
                unpack8(p1, f);
#pragma unroll
                for (int e = 0; e < 8; ++e)
                    if (f[e] > best || (f[e] == best && c2 * 8 + e < bi)) { best = f[e]; bi = c2 * 8 + e; }
            }
        }
        done = V;
    }
    for (int i = done + threadIdx.x; i < V; i += blockDim.x) {
        const float v = load_logit<T>(z, i);
        if (v > best || (v == best && i < bi)) { best = v; bi = i; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) { sv[warp] = best; si[warp] = bi; }
    __syncthreads();
    if (warp == 0) {
        const int nw = (blockDim.x + 31) >> 5;
        best = lane < nw ? sv[lane] : -INFINITY;
        bi = lane < nw ? si[lane] : 0x7fffffff;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
        }
        if (lane == 0) out[blockIdx.x] = bi;
    }
}

template <typename T>
__global__ void __launch_bounds__(kThreads) topk_sample_kernel(const T* __restrict__ logits, int V,
                                                               long long row_stride, int top_k, float top_p,
                                                               float temperature,
                                                               const float* __restrict__ uniform,
                                                               long long* __restrict__ out) {
    __shared__ uint32_t hist[2048];
    __shared__ uint32_t s_prefix, s_need;
    __shared__ int s_count;
    __shared__ float cand_v[kMaxTopK];
    __shared__ int cand_i[kMaxTopK];
    __shared__ float sort_v[kMaxTopK];
    __shared__ int sort_i[kMaxTopK];
    __shared__ float red[32];
    const T* z = logits + (size_t)blockIdx.x * row_stride;
    const int tid = threadIdx.x;

    // ---- radix select: find key K such that count(key > K) < k <= count(key >= K)
    uint32_t prefix = 0, mask = 0;
    uint32_t need = (uint32_t)top_k;
    const int shifts[3] = {21, 10, 0};
    const int bits[3] = {11, 11, 10};
    for (int round = 0; round < 3; ++round) {
        const int nb = 1 << bits[round];
        for (int i = tid; i < nb; i += blockDim.x) hist[i] = 0;
        __syncthreads();
        for (int i = tid; i < V; i += blockDim.x) {
            const uint32_t key = order_key(load_logit<T>(z, i));
            if ((key & mask) == prefix) atomicAdd(&hist[(key >> shifts[round]) & (nb - 1)], 1u);
        }
        __syncthreads();
        if (tid == 0) {
            uint32_t acc = 0;
            int b = nb - 1;
            for (; b > 0; --b) {
                if (acc + hist[b] >= need) break;
                acc += hist[b];
            }
            s_prefix = prefix | ((uint32_t)b << shifts[round]);
            s_need = need - acc;
        }
        __syncthreads();
        prefix = s_prefix;
        need = s_need;
        mask |= (uint32_t)(nb - 1) << shifts[round];
        __syncthreads();
    }
    const uint32_t kth = prefix;  // exact k-th largest key; `need` of the ties at kth are kept

    // ---- compact: all keys > kth, then the first `need` (by index) keys == kth
    if (tid == 0) s_count = 0;
    __syncthreads();
    for (int i = tid; i < V; i += blockDim.x) {
        const float v = load_logit<T>(z, i);
        if (order_key(v) > kth) {
            const int slot = atomicAdd(&s_count, 1);
            if (slot < kMaxTopK) { cand_v[slot] = v; cand_i[slot] = i; }
        }
    }
    __syncthreads();
    const int n_gt = s_count;
    __syncthreads();
    {   // ties at the threshold: keep the `need` lowest indices (deterministic).  Each thread owns
        // a contiguous index chunk, counts its ties, a block exclusive scan gives its offset.
        const int chunk = (V + blockDim.x - 1) / blockDim.x;
        const int lo = tid * chunk, hi = min(V, lo + chunk);
        int mine = 0;
        for (int i = lo; i < hi; ++i) mine += (order_key(load_logit<T>(z, i)) == kth);
        // exclusive scan of `mine` over the block
        const int lane = tid & 31, warp = tid >> 5;
        int incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        int* wsum = reinterpret_cast<int*>(hist);  // histogram is dead now
        if (lane == 31) wsum[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            int w = wsum[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += t;
            }
            wsum[32 + lane] = w;  // inclusive warp totals
        }
        __syncthreads();
        int offset = incl - mine + (warp > 0 ? wsum[32 + warp - 1] : 0);
        for (int i = lo; i < hi && offset < (int)need; ++i) {
            const float v = load_logit<T>(z, i);
            if (order_key(v) == kth) {
                const int slot = n_gt + offset;
                if (slot < kMaxTopK) { cand_v[slot] = v; cand_i[slot] = i; }
                ++offset;
            }
        }
        if (tid == 0) s_count = n_gt + (int)need;
    }
    __syncthreads();
    const int n = min(s_count, top_k);

    // ---- rank sort (descending value, ascending index on ties); n <= 1024 = blockDim
    if (tid < n) {
        const float v = cand_v[tid];
        const int idx = cand_i[tid];
        int rank = 0;
        for (int j = 0; j < n; ++j) {
            const float vj = cand_v[j];
            rank += (vj > v) || (vj == v && cand_i[j] < idx);
        }
        sort_v[rank] = v;
        sort_i[rank] = idx;
    }
    __syncthreads();

    // ---- softmax over the k candidates, optional top-p, inverse-CDF draw
    const float inv_t = 1.f / fmaxf(temperature, 1e-6f);
    const float mx = sort_v[0];
    float e = (tid < n) ? __expf((sort_v[tid] - mx) * inv_t) : 0.f;
    const float total = block_sum(e, red);
    __syncthreads();
    if (tid < n) cand_v[tid] = e / total;  // probabilities, descending
    __syncthreads();
    if (tid == 0) {
        int keep = n;
        if (top_p > 0.f && top_p < 1.f) {
            float c = 0.f;
            keep = 0;
            for (int j = 0; j < n; ++j) {
                c += cand_v[j];
                ++keep;
                if (c >= top_p) break;
            }
        }
        float mass = 0.f;
        for (int j = 0; j < keep; ++j) mass += cand_v[j];
        const float u = uniform[blockIdx.x] * mass;
        float c = 0.f;
        int pick = keep - 1;
        for (int j = 0; j < keep; ++j) {
            c += cand_v[j];
            if (u < c) { pick = j; break; }
        }
        out[blockIdx.x] = sort_i[pick];
    }
}

}  // namespace

// logits [B, V] (dtype 0 = fp32, 1 = bf16, row pitch row_stride elements) -> out [B] int64.
NXD_API int nxd_argmax(const void* logits, int dtype, int B, int V, long long row_stride, long long* out,
                       void* stream) {
    if (B == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == 0) argmax_kernel<float><<<B, 1024, 0, s>>>((const float*)logits, V, row_stride, out);
    else argmax_kernel<bf16><<<B, 1024, 0, s>>>((const bf16*)logits, V, row_stride, out);
    NXD_LAUNCH_CHECK();
    return 0;
}

// top_k in [1, 1024]; top_p <= 0 or >= 1 disables nucleus filtering; uniform [B] in [0, 1).
NXD_API int nxd_topk_sample(const void* logits, int dtype, int B, int V, long long row_stride, int top_k,
                            float top_p, float temperature, const float* uniform, long long* out,
                            void* stream) {
    NXD_REQUIRE(top_k >= 1 && top_k <= kMaxTopK, "topk_sample: top_k=%d must be in [1, %d]", top_k, kMaxTopK);
    if (B == 0) return 0;
    if (top_k > V) top_k = V;
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == 0)
        topk_sample_kernel<float><<<B, kThreads, 0, s>>>((const float*)logits, V, row_stride, top_k, top_p,
                                                         temperature, uniform, out);
    else
        topk_sample_kernel<bf16><<<B, kThreads, 0, s>>>((const bf16*)logits, V, row_stride, top_k, top_p,
                                                        temperature, uniform, out);
    NXD_LAUNCH_CHECK();
    return 0;
}
