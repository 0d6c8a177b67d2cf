// Vocab-parallel cross entropy for sm_100a.
//
// Reference: src/neuronx_distributed/parallel_layers/loss_functions.py:11-130 does three TP
// all-reduces on fp32/fp64 [T] vectors and keeps an fp32 softmax [T, V/tp] for backward.
// Here ONE pass over the bf16 logits produces, per row, the local
//   stats[t] = (max, sum exp(z - max), z[target] if target is local else 0, sum z)
// which the ranks exchange as a single packed message; backward recomputes
// softmax - onehot straight from the bf16 logits and the global log-sum-exp.
#include "common.cuh"

namespace {

__global__ void ce_stats_kernel(const bf16* __restrict__ logits, const long long* __restrict__ target,
                                float4* __restrict__ stats, int V, int vocab_start) {
    __shared__ float red[32];
    const size_t row = blockIdx.x;
    const bf16* z = logits + row * (size_t)V;
    const int nvec = V >> 3;
    float m = -INFINITY, s = 0.f, sum = 0.f;
    for (int c = threadIdx.x; c < nvec; c += blockDim.x) {
        float v[8];
        bf16x8 p = reinterpret_cast<const bf16x8*>(z)[c];
        unpack8(p, v);
        float lm = v[0];
#pragma unroll
        for (int k = 1; k < 8; ++k) lm = fmaxf(lm, v[k]);
        const float nm = fmaxf(m, lm);
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            acc += __expf(v[k] - nm);
            sum += v[k];
        }
        s = s * __expf(m - nm) + acc;
        m = nm;
    }
    for (int i = (nvec << 3) + threadIdx.x; i < V; i += blockDim.x) {  // tail (V % 8)
        const float v = __bfloat162float(z[i]);
        const float nm = fmaxf(m, v);
        s = s * __expf(m - nm) + __expf(v - nm);
        m = nm;
        sum += v;
    }
    const float gm = block_max(m, red);
    const float contrib = (m == -INFINITY) ? 0.f : s * __expf(m - gm);
    const float gs = block_sum(contrib, red);
    const float gsum = block_sum(sum, red);
    if (threadIdx.x == 0) {
        const long long t = target[row] - vocab_start;
        const float tl = (t >= 0 && t < V) ? __bfloat162float(z[t]) : 0.f;
        stats[row] = make_float4(gm, gs, tl, gsum);
    }
}

// grad[t, v] = g[t] * ( softmax - (1-a) * onehot - a / V_global ), zero row if target < 0 (ignored)
__global__ void ce_bwd_kernel(const bf16* __restrict__ logits, const long long* __restrict__ target,
                              const float* __restrict__ lse, const float* __restrict__ gout,
                              bf16* __restrict__ grad, int V, int vocab_start, float smoothing,
                              float inv_vocab_global) {
    const size_t row = blockIdx.x;
    const bf16* z = logits + row * (size_t)V;
    bf16* g = grad + row * (size_t)V;
    const long long tg = target[row];
    const float go = (tg < 0) ? 0.f : gout[row];
    const float l = lse[row];
    const long long t = tg - vocab_start;
    const int nvec = V >> 3;
    const float sm = smoothing * inv_vocab_global;
    for (int c = threadIdx.x; c < nvec; c += blockDim.x) {
        float v[8], o[8];
        bf16x8 p = reinterpret_cast<const bf16x8*>(z)[c];
        unpack8(p, v);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            float pr = __expf(v[k] - l) - sm;
            if ((long long)(c * 8 + k) == t) pr -= (1.f - smoothing);
            o[k] = go * pr;
        }
        reinterpret_cast<bf16x8*>(g)[c] = pack8(o);
    }
    for (int i = (nvec << 3) + threadIdx.x; i < V; i += blockDim.x) {
        float pr = __expf(__bfloat162float(z[i]) - l) - sm;
        if ((long long)i == t) pr -= (1.f - smoothing);
        g[i] = __float2bfloat16(go * pr);
    }
}

}  // namespace

// logits [T, V] bf16 (this rank's vocab slice, row pitch V, V % 8 == 0 for the vector path to
// be aligned); target [T] int64 (global vocab ids; <0 = ignore); stats [T, 4] fp32.
NXD_API int nxd_ce_stats(const void* logits, const long long* target, float* stats, long long T, int V,
                         int vocab_start, void* stream) {
    NXD_REQUIRE(V % 8 == 0, "ce_stats: V/tp=%d must be a multiple of 8 (pad the vocab)", V);
    if (T == 0) return 0;
    const int threads = V >= 8192 ? 1024 : 256;
    ce_stats_kernel<<<(unsigned)T, threads, 0, (cudaStream_t)stream>>>((const bf16*)logits, target,
                                                                       (float4*)stats, V, vocab_start);
    NXD_LAUNCH_CHECK();
    return 0;
}

NXD_API int nxd_ce_bwd(const void* logits, const long long* target, const float* lse, const float* grad_out,
                       void* grad_logits, long long T, int V, int vocab_start, float smoothing,
                       int vocab_global, void* stream) {
    NXD_REQUIRE(V % 8 == 0, "ce_bwd: V/tp=%d must be a multiple of 8 (pad the vocab)", V);
    if (T == 0) return 0;
    const int threads = V >= 8192 ? 1024 : 256;
    ce_bwd_kernel<<<(unsigned)T, threads, 0, (cudaStream_t)stream>>>(
        (const bf16*)logits, target, lse, grad_out, (bf16*)grad_logits, V, vocab_start, smoothing,
        1.f / (float)vocab_global);
    NXD_LAUNCH_CHECK();
    return 0;
}
