// Token-generation (decode) attention over a KV cache, and the cache write, for sm_100a
// (SURVEY G5 + "KV-cache update").
//
// Reference: examples/inference/modules/attention/attention_base.py:141-170 computes scores
// against the cached K (masked) and the new K separately, a joint softmax, and two PV matmuls;
// examples/inference/modules/model_base.py:388-422 re-emits the whole cache through
// torch.scatter.  Here the new K/V row is first written in place (nxd_kv_cache_update) and one
// split-K kernel attends over positions [0, len) of the cache:
//   grid = (B * n_kv_heads, n_splits); a CTA serves the `group = n_q/n_kv` query heads that share
//   its KV head (native GQA, no repeat_kv), scoring keys thread-per-key and accumulating PV
//   thread-per-channel; partial (m, l, o) per split go to scratch and a second kernel merges.
// The cache is either contiguous [B, n_kv, S_max, D] or paged [n_pages, n_kv, page, D] with a
// block table [B, max_pages_per_seq].
#include "common.cuh"

namespace {

constexpr int kMaxGroup = 8;     // q heads per kv head handled by one CTA
constexpr int kChunk = 256;      // keys per inner chunk == threads per CTA

struct CacheView {
    const bf16* k;
    const bf16* v;
    const int* block_table;  // null -> contiguous
    long long stride_b;      // elements between batches (contiguous) / between pages (paged)
    long long stride_h;      // elements between kv heads
    int page_size;
    int max_pages;
};

__device__ __forceinline__ const bf16* row_ptr(const bf16* base, const CacheView& c, int b, int h, int pos,
                                               int D) {
    if (c.block_table) {
        const int page = c.block_table[(size_t)b * c.max_pages + pos / c.page_size];
        return base + (size_t)page * c.stride_b + (size_t)h * c.stride_h + (size_t)(pos % c.page_size) * D;
    }
    return base + (size_t)b * c.stride_b + (size_t)h * c.stride_h + (size_t)pos * D;
}

// Block-wide max / sum of kMaxGroup values at once: one shared-memory exchange (two barriers) for the whole group
// instead of one per query head.  red: kMaxGroup * 8 floats (8 warps).
template <bool kMax>
__device__ __forceinline__ void block_reduce_group(float (&v)[kMaxGroup], int group, float (*red)[kChunk / 32]) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int g = 0; g < kMaxGroup; ++g) {
        if (g < group) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float t = __shfl_xor_sync(0xffffffffu, v[g], o);
                v[g] = kMax ? fmaxf(v[g], t) : v[g] + t;
            }
        }
    }
    __syncthreads();   // protect `red` from its previous use
    if (lane == 0) {
#pragma unroll
        for (int g = 0; g < kMaxGroup; ++g)
            if (g < group) red[g][warp] = v[g];
    }
    __syncthreads();
#pragma unroll
    for (int g = 0; g < kMaxGroup; ++g) {
        if (g < group) {
            float r = red[g][0];
#pragma unroll
            for (int w = 1; w < kChunk / 32; ++w) r = kMax ? fmaxf(r, red[g][w]) : r + red[g][w];
            v[g] = r;
        }
    }
}

template <int D>
__global__ void __launch_bounds__(kChunk) decode_attn_split_kernel(
    const bf16* __restrict__ q,  // [B, n_q, D]
    CacheView cache, const int* __restrict__ lengths, float* __restrict__ part_o,  // [B, n_q, splits, D]
    float* __restrict__ part_ml,                                                   // [B, n_q, splits, 2]
    int n_q, int n_kv, int n_splits, float scale) {
    const int b = blockIdx.x / n_kv;
    const int hk = blockIdx.x % n_kv;
    const int split = blockIdx.y;
    const int group = n_q / n_kv;
    const int len = lengths[b];
    const int per = (len + n_splits - 1) / n_splits;
    const int start = split * per;
    const int end = min(len, start + per);
    const int tid = threadIdx.x;

    constexpr int kLanesPerRow = D / 2;                  // PV phase: a thread owns two channels (one bf162 load per key)
    constexpr int kParts = kChunk / kLanesPerRow;        // key partitions: 8 for D = 64, 4 for D = 128
    __shared__ float sq[kMaxGroup][D];
    __shared__ float sp[kMaxGroup][kChunk];
    __shared__ float red[kMaxGroup][kChunk / 32];
    __shared__ float2 so[kParts][kMaxGroup][kLanesPerRow];

    for (int i = tid; i < group * D; i += kChunk) {
        const int g = i / D, d = i % D;
        sq[g][d] = __bfloat162float(q[((size_t)b * n_q + hk * group + g) * D + d]) * scale;
    }
    __syncthreads();

    float m_run[kMaxGroup], l_run[kMaxGroup];
    float2 o_acc[kMaxGroup];
#pragma unroll
    for (int g = 0; g < kMaxGroup; ++g) { m_run[g] = -INFINITY; l_run[g] = 0.f; o_acc[g] = make_float2(0.f, 0.f); }
    const int d2 = tid % kLanesPerRow;
    const int part = tid / kLanesPerRow;

    for (int c0 = start; c0 < end; c0 += kChunk) {
        const int pos = c0 + tid;
        const bool valid = pos < end;
        float s[kMaxGroup];
#pragma unroll
        for (int g = 0; g < kMaxGroup; ++g) s[g] = 0.f;
        if (valid) {
            const bf16x8* kr = reinterpret_cast<const bf16x8*>(row_ptr(cache.k, cache, b, hk, pos, D));
#pragma unroll
            for (int c = 0; c < D / 8; ++c) {
                float kv[8];
                bf16x8 p = kr[c];
                unpack8(p, kv);
#pragma unroll
                for (int g = 0; g < kMaxGroup; ++g) {
                    if (g < group) {
#pragma unroll
                        for (int e = 0; e < 8; ++e) s[g] += kv[e] * sq[g][c * 8 + e];
                    }
                }
            }
        }
        // chunk softmax for all heads of the group (online across chunks)
        float cm[kMaxGroup], pv[kMaxGroup], alpha[kMaxGroup];
#pragma unroll
        for (int g = 0; g < kMaxGroup; ++g) cm[g] = valid ? s[g] : -INFINITY;
        block_reduce_group<true>(cm, group, red);
#pragma unroll
        for (int g = 0; g < kMaxGroup; ++g) {
            if (g < group) {
                const float nm = fmaxf(m_run[g], cm[g]);
                pv[g] = valid ? __expf(s[g] - nm) : 0.f;
                alpha[g] = __expf(m_run[g] - nm);
                m_run[g] = nm;
                sp[g][tid] = pv[g];
            }
        }
        block_reduce_group<false>(pv, group, red);        // its barriers also publish sp[][] to the block
#pragma unroll
        for (int g = 0; g < kMaxGroup; ++g) {
            if (g < group) {
                l_run[g] = l_run[g] * alpha[g] + pv[g];
                o_acc[g].x *= alpha[g];
                o_acc[g].y *= alpha[g];
            }
        }
        // PV: thread (part, d2) accumulates keys j = part, part + kParts, ... for channels 2 d2, 2 d2 + 1
        const int nk = min(kChunk, end - c0);
        // 8 V rows are requested before the first one is used: the loop is bound by load latency (a profile at
        // batch 16 showed ~0.6 us per serial iteration), not by bandwidth or math
        constexpr int kPf = 8;
        for (int j0 = part; j0 < nk; j0 += kParts * kPf) {
            bf162 raw[kPf];
#pragma unroll
            for (int u = 0; u < kPf; ++u) {
                const int j = j0 + u * kParts;
                raw[u] = j < nk ? reinterpret_cast<const bf162*>(row_ptr(cache.v, cache, b, hk, c0 + j, D))[d2]
                                : __floats2bfloat162_rn(0.f, 0.f);
            }
#pragma unroll
            for (int u = 0; u < kPf; ++u) {
                const int j = j0 + u * kParts;
                if (j < nk) {
                    const float2 vv = __bfloat1622float2(raw[u]);
#pragma unroll
                    for (int g = 0; g < kMaxGroup; ++g) {
                        if (g < group) {
                            const float w = sp[g][j];
                            o_acc[g].x += w * vv.x;
                            o_acc[g].y += w * vv.y;
                        }
                    }
                }
            }
        }
        __syncthreads();     // sp[][] is rewritten by the next chunk
    }
    // merge the kParts partial accumulators
    __shared__ float sml[kMaxGroup][2];
#pragma unroll
    for (int g = 0; g < kMaxGroup; ++g) {
        if (g < group) {
            so[part][g][d2] = o_acc[g];
            if (tid == 0) { sml[g][0] = m_run[g]; sml[g][1] = l_run[g]; }
        }
    }
    __syncthreads();
    for (int i = tid; i < group * kLanesPerRow; i += kChunk) {
        const int g = i / kLanesPerRow, dd = i % kLanesPerRow;
        float2 acc = make_float2(0.f, 0.f);
#pragma unroll
        for (int p = 0; p < kParts; ++p) { acc.x += so[p][g][dd].x; acc.y += so[p][g][dd].y; }
        const size_t base = (((size_t)b * n_q + hk * group + g) * n_splits + split);
        reinterpret_cast<float2*>(part_o + base * D)[dd] = acc;
        if (dd == 0) {
            part_ml[base * 2 + 0] = sml[g][0];
            part_ml[base * 2 + 1] = sml[g][1];
        }
    }
}

template <int D>
__global__ void decode_attn_merge_kernel(const float* __restrict__ part_o, const float* __restrict__ part_ml,
                                         bf16* __restrict__ out, int n_splits) {
    const size_t bh = blockIdx.x;  // b * n_q + h
    const int d = threadIdx.x;
    float m = -INFINITY;
    for (int s = 0; s < n_splits; ++s) m = fmaxf(m, part_ml[(bh * n_splits + s) * 2]);
    float l = 0.f, o = 0.f;
    for (int s = 0; s < n_splits; ++s) {
        const float ms = part_ml[(bh * n_splits + s) * 2];
        const float w = (ms == -INFINITY) ? 0.f : __expf(ms - m);
        l += part_ml[(bh * n_splits + s) * 2 + 1] * w;
        o += part_o[(bh * n_splits + s) * D + d] * w;
    }
    out[bh * D + d] = __float2bfloat16(l > 0.f ? o / l : 0.f);
}

// new_k/new_v: [B, n_kv, T_new, D]; positions: [B] start position (rows t -> positions[b] + t)
__global__ void kv_cache_update_kernel(const bf16* __restrict__ new_k, const bf16* __restrict__ new_v,
                                       bf16* __restrict__ cache_k, bf16* __restrict__ cache_v,
                                       const int* __restrict__ positions, const int* __restrict__ seq_ids,
                                       const int* __restrict__ block_table, long long stride_b,
                                       long long stride_h, int page_size, int max_pages, int B, int n_kv,
                                       int T_new, int D) {
    const int vec_per_row = D >> 3;
    const size_t total = (size_t)B * n_kv * T_new * vec_per_row;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (size_t)gridDim.x * blockDim.x) {
        const int c = (int)(idx % vec_per_row);
        size_t r = idx / vec_per_row;
        const int t = (int)(r % T_new); r /= T_new;
        const int h = (int)(r % n_kv);
        const int b = (int)(r / n_kv);
        const int slot = seq_ids ? seq_ids[b] : b;  // continuous batching: batch row -> cache line
        const int pos = positions[b] + t;
        size_t off;
        if (block_table) {
            const int page = block_table[(size_t)slot * max_pages + pos / page_size];
            off = (size_t)page * stride_b + (size_t)h * stride_h + (size_t)(pos % page_size) * D;
        } else {
            off = (size_t)slot * stride_b + (size_t)h * stride_h + (size_t)pos * D;
        }
        const size_t src = (((size_t)b * n_kv + h) * T_new + t) * D;
        reinterpret_cast<bf16x8*>(cache_k + off)[c] = reinterpret_cast<const bf16x8*>(new_k + src)[c];
        reinterpret_cast<bf16x8*>(cache_v + off)[c] = reinterpret_cast<const bf16x8*>(new_v + src)[c];
    }
}

}  // namespace

NXD_API int nxd_decode_attn_num_splits(int B, int n_kv, int max_len) {
    // enough CTAs to cover the machine, at least 256 keys per split
    int want = (2 * nxd_num_sms() + B * n_kv - 1) / (B * n_kv);
    int by_len = (max_len + 255) / 256;
    int s = want < by_len ? want : by_len;
    if (s < 1) s = 1;
    if (s > 64) s = 64;
    return s;
}

// scratch: B * n_q * n_splits * (D + 2) floats.
NXD_API int nxd_decode_attn(const void* q, const void* cache_k, const void* cache_v, const int* lengths,
                            const int* block_table, long long stride_b, long long stride_h, int page_size,
                            int max_pages, void* out, float* scratch, int B, int n_q, int n_kv, int D,
                            int n_splits, float scale, void* stream) {
    NXD_REQUIRE(D == 64 || D == 128, "decode_attn: head dim %d not in {64, 128}", D);
    NXD_REQUIRE(n_q % n_kv == 0 && n_q / n_kv <= kMaxGroup, "decode_attn: unsupported GQA group %d/%d", n_q, n_kv);
    NXD_REQUIRE(scratch != nullptr && n_splits >= 1, "decode_attn: scratch required");
    if (B == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    CacheView cv{(const bf16*)cache_k, (const bf16*)cache_v, block_table, stride_b, stride_h,
                 page_size > 0 ? page_size : 1, max_pages};
    float* part_o = scratch;
    float* part_ml = scratch + (size_t)B * n_q * n_splits * D;
    dim3 grid(B * n_kv, n_splits);
    if (D == 64) {
        decode_attn_split_kernel<64><<<grid, kChunk, 0, s>>>((const bf16*)q, cv, lengths, part_o, part_ml, n_q,
                                                             n_kv, n_splits, scale);
        NXD_LAUNCH_CHECK();
        decode_attn_merge_kernel<64><<<B * n_q, 64, 0, s>>>(part_o, part_ml, (bf16*)out, n_splits);
    } else {
        decode_attn_split_kernel<128><<<grid, kChunk, 0, s>>>((const bf16*)q, cv, lengths, part_o, part_ml, n_q,
                                                              n_kv, n_splits, scale);
        NXD_LAUNCH_CHECK();
        decode_attn_merge_kernel<128><<<B * n_q, 128, 0, s>>>(part_o, part_ml, (bf16*)out, n_splits);
    }
    NXD_LAUNCH_CHECK();
    return 0;
}

NXD_API int nxd_kv_cache_update(const void* new_k, const void* new_v, void* cache_k, void* cache_v,
                                const int* positions, const int* seq_ids, const int* block_table,
                                long long stride_b, long long stride_h, int page_size, int max_pages, int B,
                                int n_kv, int T_new, int D, void* stream) {
    NXD_REQUIRE(D % 8 == 0, "kv_cache_update: D %% 8 != 0");
    const size_t total = (size_t)B * n_kv * T_new * (D / 8);
    if (total == 0) return 0;
    size_t blocks = (total + 255) / 256;
    if (blocks > (size_t)nxd_num_sms() * 8) blocks = (size_t)nxd_num_sms() * 8;
    kv_cache_update_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        (const bf16*)new_k, (const bf16*)new_v, (bf16*)cache_k, (bf16*)cache_v, positions, seq_ids, block_table,
        stride_b, stride_h, page_size > 0 ? page_size : 1, max_pages, B, n_kv, T_new, D);
    NXD_LAUNCH_CHECK();
    return 0;
}
