// Vocab-parallel embedding lookup for sm_100a.
//
// Reference: src/neuronx_distributed/parallel_layers/layers.py:215-238 masks ids outside this
// rank's vocab range, looks up, zeroes the masked rows (three passes) and all-reduces.  Here the
// masked gather is one kernel (rows outside the range are written as zeros); optionally the
// output is written already transposed to the sequence-first [S, B, H] layout so the following
// reduce-scatter produces the sequence-parallel shard directly.
// Backward scatters dY rows into an fp32 dW with red.global.add (vocab rows collide rarely).
//
// nxd_embedding_fwd_sp_push is the sequence-parallel forward as ONE kernel over NVLink peer memory (SURVEY F-list:
// "embedding -> reduce-scatter"): the reference (and the NCCL path) reduce-scatters an [S, B, H] tensor that is zero
// except for the rows whose id this rank owns (50 MB through the ring for 8k tokens of Llama-3.2-3B).  Exactly one
// rank owns each row, so here the OWNER writes the row straight into the output shard of the rank that holds that
// sequence position -- S*B*H*2/tp bytes leave each GPU instead of (tp-1)/tp of the dense tensor, and nothing is
// reduced.  The kernel ends with a flag exchange, so when it completes on a rank, that rank's shard is complete.
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

namespace {
constexpr int kEmbMaxTp = 8;
struct EmbPushArgs {
    bf16* out_peer[kEmbMaxTp];       // every rank's output shard [S/tp * B, H] (peer-mapped; [rank] = local)
    uint32_t* sig_peer[kEmbMaxTp];
    uint32_t* sig_local;
    uint32_t* done;                  // local device counter, zero between calls
    int slot;                        // flag [source rank]
    int tp, rank;
    uint32_t epoch;
};

__global__ void embedding_fwd_sp_push_kernel(const long long* __restrict__ ids, const bf16* __restrict__ w, size_t T,
                                             int H, int vocab_start, int vocab_local, int B, int S, const EmbPushArgs a) {
    const int nvec = H >> 3;
    const int rows_per_rank = S / a.tp;            // sequence positions per rank
    const int warps = (gridDim.x * blockDim.x) >> 5;
    const int lane = threadIdx.x & 31;
    for (size_t t = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; t < T; t += warps) {
        const long long id = ids[t] - vocab_start;
        if (id < 0 || id >= vocab_local) continue;       // another rank owns this row
        const size_t b = t / S, s = t % S;               // ids are [B, S] row-major
        const int dest = (int)(s / rows_per_rank);
        const size_t orow = (s % rows_per_rank) * B + b; // [S/tp, B, H]
        const bf16x8* src = reinterpret_cast<const bf16x8*>(w + (size_t)id * H);
        bf16x8* dst = reinterpret_cast<bf16x8*>(a.out_peer[dest] + orow * (size_t)H);
        for (int c = lane; c < nvec; c += 32) dst[c] = src[c];
    }
    // every CTA: make its rows visible system-wide, then count; the last CTA tells the peers and waits for them
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        if (atomicAdd(a.done, 1u) == gridDim.x - 1) {
            *a.done = 0u;
            __threadfence_system();
            for (int d = 1; d < a.tp; ++d) {
                const int peer = (a.rank + d) % a.tp;
                asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a.sig_peer[peer] + a.slot + a.rank), "r"(a.epoch)
                             : "memory");
            }
            for (int d = 1; d < a.tp; ++d) {
                const int peer = (a.rank + d) % a.tp;
                uint32_t v;
                do {
                    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(a.sig_local + a.slot + peer) : "memory");
                } while ((int32_t)(v - a.epoch) < 0);
            }
        }
    }
}
}  // namespace

// Sequence-parallel embedding forward without a reduce-scatter.  ids: ALL B*S token ids ([B, S] row-major, the same on
// every rank); weight: this rank's vocab rows [vocab_local, H]; out_ptrs: tp pointers to every rank's output shard
// [S/tp, B, H] in symmetric memory (the shard of THIS call: the caller double-buffers by epoch parity); sig_ptrs:
// signal pads; slot: first of tp flags; done: one zeroed uint32 of local device memory; epoch = 1, 2, 3, ...
// S % tp == 0, H % 8 == 0.  When the kernel completes, out_ptrs[rank] holds the complete shard.
NXD_API int nxd_embedding_fwd_sp_push(const long long* ids, const void* weight, void* const* out_ptrs,
                                      void* const* sig_ptrs, void* done, long long T, int H, int vocab_start,
                                      int vocab_local, int B, int S, int tp, int rank, int slot, unsigned epoch,
                                      void* stream) {
    NXD_REQUIRE(tp >= 2 && tp <= kEmbMaxTp && rank >= 0 && rank < tp, "embedding_sp_push: tp %d / rank %d out of range", tp, rank);
    NXD_REQUIRE(H % 8 == 0 && S % tp == 0 && T == (long long)B * S, "embedding_sp_push: H %% 8, S %% tp, T == B*S required");
    NXD_REQUIRE(out_ptrs && sig_ptrs && done && epoch > 0, "embedding_sp_push: missing buffers");
    EmbPushArgs a{};
    for (int i = 0; i < tp; ++i) {
        a.out_peer[i] = (bf16*)out_ptrs[i];
        a.sig_peer[i] = (uint32_t*)sig_ptrs[i];
    }
    a.sig_local = (uint32_t*)sig_ptrs[rank];
    a.done = (uint32_t*)done;
    a.slot = slot; a.tp = tp; a.rank = rank; a.epoch = epoch;
    // one warp per token; ~1/tp of the tokens are owned here.  All CTAs are co-resident (the last one spins).
    int dev = 0, sms = 0;
    NXD_CHECK_CUDA(cudaGetDevice(&dev));
    NXD_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    long long blocks = (T + 7) / 8;
    if (blocks > 2LL * sms) blocks = 2LL * sms;
    embedding_fwd_sp_push_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        ids, (const bf16*)weight, (size_t)T, H, vocab_start, vocab_local, B, S, a);
    NXD_LAUNCH_CHECK();
    return 0;
}

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
