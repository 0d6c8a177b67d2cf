// Vocab-parallel embedding lookup for sm_100a.
//
// Reference: src/neuronx_distributed/parallel_layers/layers.py:215-238 masks ids outside this
// rank's vocab range, looks up, zeroes the masked rows (three passes) and all-reduces.  Here the
// masked gather is one kernel (rows outside the range are written as zeros); optionally the
// output is written already transposed to the sequence-first [S, B, H] layout so the following
// reduce-scatter produces the sequence-parallel shard directly.
// Backward scatters dY rows into an fp32 dW with red.global.add (vocab rows collide rarely).
#include "common.cuh"

namespace {

__global__ void embedding_fwd_kernel(const long long* __restrict__ ids, const bf16* __restrict__ w,
                                     bf16* __restrict__ out, size_t T, int H, int vocab_start, int vocab_local,
                                     int B, int S, int seq_first_out) {
    const int nvec = H >> 3;
    const size_t total = T * (size_t)nvec;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (size_t)gridDim.x * blockDim.x) {
        const size_t t = idx / nvec;
        const int c = (int)(idx % nvec);
        const long long id = ids[t] - vocab_start;
        bf16x8 val;
        if (id >= 0 && id < vocab_local) {
            val = reinterpret_cast<const bf16x8*>(w + (size_t)id * H)[c];
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) val.v[k] = __floats2bfloat162_rn(0.f, 0.f);
        }
        size_t orow = t;
        if (seq_first_out) {  // ids are [B, S] row-major; output row = s * B + b
            const size_t b = t / S, s = t % S;
            orow = s * B + b;
        }
        reinterpret_cast<bf16x8*>(out + orow * (size_t)H)[c] = val;
    }
}

__global__ void embedding_bwd_kernel(const long long* __restrict__ ids, const bf16* __restrict__ dy,
                                     float* __restrict__ dw, size_t T, int H, int vocab_start, int vocab_local,
                                     int B, int S, int seq_first_in) {
    const int nvec = H >> 3;
    const size_t total = T * (size_t)nvec;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (size_t)gridDim.x * blockDim.x) {
        const size_t t = idx / nvec;
        const int c = (int)(idx % nvec);
        const long long id = ids[t] - vocab_start;
        if (id < 0 || id >= vocab_local) continue;
        size_t irow = t;
        if (seq_first_in) {
            const size_t b = t / S, s = t % S;
            irow = s * B + b;
        }
        float v[8];
        bf16x8 p = reinterpret_cast<const bf16x8*>(dy + irow * (size_t)H)[c];
        unpack8(p, v);
        float* dst = dw + (size_t)id * H + (size_t)c * 8;
#pragma unroll
        for (int k = 0; k < 8; ++k) atomicAdd(dst + k, v[k]);
    }
}

}  // namespace

NXD_API int nxd_embedding_fwd(const long long* ids, const void* weight, void* out, long long T, int H,
                              int vocab_start, int vocab_local, int B, int S, int seq_first_out, void* stream) {
    NXD_REQUIRE(H % 8 == 0, "embedding: H %% 8 != 0");
    if (T == 0) return 0;
    size_t blocks = ((size_t)T * (H / 8) + 255) / 256;
    if (blocks > (size_t)nxd_num_sms() * 16) blocks = (size_t)nxd_num_sms() * 16;
    embedding_fwd_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        ids, (const bf16*)weight, (bf16*)out, (size_t)T, H, vocab_start, vocab_local, B, S, seq_first_out);
    NXD_LAUNCH_CHECK();
    return 0;
}

// dw: fp32 [vocab_local, H], accumulated into (caller zeroes it when needed).
NXD_API int nxd_embedding_bwd(const long long* ids, const void* dy, float* dw, long long T, int H,
                              int vocab_start, int vocab_local, int B, int S, int seq_first_in, void* stream) {
    NXD_REQUIRE(H % 8 == 0, "embedding_bwd: H %% 8 != 0");
    if (T == 0) return 0;
    size_t blocks = ((size_t)T * (H / 8) + 255) / 256;
    if (blocks > (size_t)nxd_num_sms() * 16) blocks = (size_t)nxd_num_sms() * 16;
    embedding_bwd_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        ids, (const bf16*)dy, dw, (size_t)T, H, vocab_start, vocab_local, B, S, seq_first_in);
    NXD_LAUNCH_CHECK();
    return 0;
}
