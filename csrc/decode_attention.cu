// Token-generation (decode) attention over a KV cache, and the cache write, for sm_100a
// (SURVEY G5 + "KV-cache update").
//
// Reference: examples/inference/modules/attention/attention_base.py:141-170 computes scores
// against the cached K (masked) and the new K separately, a joint softmax, and two PV matmuls;
// examples/inference/modules/model_base.py:388-422 re-emits the whole cache through
// torch.scatter.  Here the new K/V row is first written in place (nxd_kv_cache_update) and one
// split-K kernel attends over positions [0, len) of the cache:
//   grid = (B * n_kv_heads * head chunks, n_splits); a CTA serves up to 4 of the `group = n_q/n_kv` query heads that
//   share its KV head (native GQA, no repeat_kv); every 8-lane group of a warp owns a strided subset of the keys with
//   its own online-softmax state (see the kernel); partial (m, l, o) per split go to scratch and a second kernel merges.
// The cache is either contiguous [B, n_kv, S_max, D] or paged [n_pages, n_kv, page, D] with a
// block table [B, max_pages_per_seq].
#include <stdlib.h>

#include "common.cuh"
#include "mma_sync.cuh"

namespace {

constexpr int kMaxGroup = 64;    // q heads per kv head (processed in chunks of up to 4 heads per CTA)

struct CacheView {
    const bf16* k;
    const bf16* v;
    const int* block_table;  // null -> contiguous
    const int* seq_ids;      // null -> cache line (contiguous) / block-table row (paged) = batch row
    long long stride_b;      // elements between batches (contiguous) / between pages (paged)
    long long stride_h;      // elements between kv heads
    int page_size;
    int max_pages;
};

__device__ __forceinline__ const bf16* row_ptr(const bf16* base, const CacheView& c, int b, int h, int pos,
                                               int D) {
    const int line = c.seq_ids ? c.seq_ids[b] : b;     // continuous batching: batch row -> cache line, no gathered copy
    if (c.block_table) {
        const int page = c.block_table[(size_t)line * c.max_pages + pos / c.page_size];
        return base + (size_t)page * c.stride_b + (size_t)h * c.stride_h + (size_t)(pos % c.page_size) * D;
    }
    return base + (size_t)line * c.stride_b + (size_t)h * c.stride_h + (size_t)pos * D;
}

// Split-K decode attention, one pass over the cache rows [start, end) of one (sequence, kv head, head chunk).
//
// Work decomposition (no shared memory and no block barrier inside the key loop):
//   * a row of K or V (D bf16) is read by kLanes = 8 adjacent lanes, 16 bytes each (D = 128: two 16-byte pieces per
//     lane), so a warp streams 4 consecutive rows = 512 (1024) contiguous bytes per load instruction;
//   * lane group `sub` (0..3) of warp `w` owns the keys  start + (w * 4 + sub) + 16 * i  and keeps its OWN online-softmax
//     state (m, l, o) for the G query heads of the chunk: the partial dot products are summed over the 8 lanes with
//     3 shuffles, after which all 8 lanes hold the same score and update the same state, each for its D / 8 channels;
//   * kUnroll keys per lane group are loaded before the first is used (memory-level parallelism: 8 KB per warp);
//   * at the end the 16 lane groups of the CTA merge through shared memory and write one (m, l, o) partial.
template <int D, int G>
__global__ void __launch_bounds__(128) decode_attn_split_kernel(
    const bf16* __restrict__ q,  // [B, n_q, D]
    CacheView cache, const int* __restrict__ lengths, float* __restrict__ part_o,  // [B, n_q, splits, D]
    float* __restrict__ part_ml,                                                   // [B, n_q, splits, 2]
    int n_q, int n_kv, int n_splits, float scale_log2) {
    constexpr int kLanes = 8;            // lanes per cache row
    constexpr int kE = D / kLanes;       // channels per lane: 8 (D = 64) or 16 (D = 128)
    constexpr int kV = kE / 8;           // 16-byte pieces per lane and row
    constexpr int kUnroll = 4;
    constexpr int kGroups = 16;          // lane groups per CTA (4 warps x 4)
    pdl_trigger();
    pdl_wait();     // q and the newest cache row come from the projection kernel right before this one
    const int group = n_q / n_kv;
    const int chunks = group / G;
    const int b = blockIdx.x / (n_kv * chunks);
    const int hk = (blockIdx.x / chunks) % n_kv;
    const int h0 = hk * group + (blockIdx.x % chunks) * G;     // first query head of this CTA
    const int split = blockIdx.y;
    const int len = lengths[b];
    const int per = (len + n_splits - 1) / n_splits;
    const int start = split * per;
    const int end = min(len, start + per);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane >> 3, l8 = lane & 7;
    const int gid = warp * 4 + sub;

    float qf[G][kE];
#pragma unroll
    for (int g = 0; g < G; ++g) {
        const bf16x8* qp = reinterpret_cast<const bf16x8*>(q + ((size_t)b * n_q + h0 + g) * D);
#pragma unroll
        for (int v = 0; v < kV; ++v) {
            float t[8];
            unpack8(qp[l8 + v * kLanes], t);
#pragma unroll
            for (int e = 0; e < 8; ++e) qf[g][v * 8 + e] = t[e] * scale_log2;
        }
    }
    float m_run[G], l_run[G], o[G][kE];
#pragma unroll
    for (int g = 0; g < G; ++g) {
        m_run[g] = -INFINITY;
        l_run[g] = 0.f;
#pragma unroll
        for (int e = 0; e < kE; ++e) o[g][e] = 0.f;
    }

    // the trip count is CTA-uniform (the shuffles below need every lane of the warp); a lane group whose keys fall
    // behind `end` carries -inf scores through the iteration
    for (int kb = start; kb < end; kb += kGroups * kUnroll) {
        const int k0 = kb + gid;
        bf16x8 kr[kUnroll][kV], vr[kUnroll][kV];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const int pos = k0 + u * kGroups;
            if (pos < end) {
                const bf16x8* kp = reinterpret_cast<const bf16x8*>(row_ptr(cache.k, cache, b, hk, pos, D));
                const bf16x8* vp = reinterpret_cast<const bf16x8*>(row_ptr(cache.v, cache, b, hk, pos, D));
#pragma unroll
                for (int v = 0; v < kV; ++v) {
                    kr[u][v] = kp[l8 + v * kLanes];
                    vr[u][v] = vp[l8 + v * kLanes];
                }
            }
        }
        float sc[kUnroll][G];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            float kf[kE];
#pragma unroll
            for (int v = 0; v < kV; ++v) unpack8(kr[u][v], kf + v * 8);
#pragma unroll
            for (int g = 0; g < G; ++g) {
                float acc = 0.f;
#pragma unroll
                for (int e = 0; e < kE; ++e) acc = fmaf(kf[e], qf[g][e], acc);
                sc[u][g] = acc;
            }
        }
        // sum the partial dot products over the 8 lanes of the group (all lanes of the warp take part)
#pragma unroll
        for (int u = 0; u < kUnroll; ++u)
#pragma unroll
            for (int g = 0; g < G; ++g) {
                float t = sc[u][g];
                t += __shfl_xor_sync(0xffffffffu, t, 1);
                t += __shfl_xor_sync(0xffffffffu, t, 2);
                t += __shfl_xor_sync(0xffffffffu, t, 4);
                sc[u][g] = (k0 + u * kGroups < end) ? t : -INFINITY;
            }
#pragma unroll
        for (int g = 0; g < G; ++g) {
            float mx = sc[0][g];
#pragma unroll
            for (int u = 1; u < kUnroll; ++u) mx = fmaxf(mx, sc[u][g]);
            const float m_new = fmaxf(m_run[g], mx);
            const float m_use = (m_new == -INFINITY) ? 0.f : m_new;     // no key of this lane group yet: all weights 0
            const float alpha = exp2f(m_run[g] - m_use);                // exp2(-inf) = 0 while the state is empty
            m_run[g] = m_new;
            float pw[kUnroll], ps = 0.f;
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) {
                pw[u] = exp2f(sc[u][g] - m_use);
                ps += pw[u];
            }
            l_run[g] = l_run[g] * alpha + ps;
#pragma unroll
            for (int e = 0; e < kE; ++e) o[g][e] *= alpha;
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) {
                if (k0 + u * kGroups < end) {
#pragma unroll
                    for (int v = 0; v < kV; ++v) {
                        float vf[8];
                        unpack8(vr[u][v], vf);
#pragma unroll
                        for (int e = 0; e < 8; ++e) o[g][v * 8 + e] = fmaf(pw[u], vf[e], o[g][v * 8 + e]);
                    }
                }
            }
        }
    }

    // merge the 16 lane groups: (m, l) per (group, head) and o per (group, head, channel) through shared memory
    __shared__ float s_m[kGroups][G], s_l[kGroups][G];
    __shared__ float s_o[kGroups][G][D + 4];
    if (l8 == 0) {
#pragma unroll
        for (int g = 0; g < G; ++g) { s_m[gid][g] = m_run[g]; s_l[gid][g] = l_run[g]; }
    }
#pragma unroll
    for (int g = 0; g < G; ++g)
#pragma unroll
        for (int v = 0; v < kV; ++v)
#pragma unroll
            for (int e = 0; e < 8; ++e) s_o[gid][g][(l8 + v * kLanes) * 8 + e] = o[g][v * 8 + e];
    __syncthreads();
    for (int i = threadIdx.x; i < G * D; i += blockDim.x) {
        const int g = i / D, d = i % D;
        float m = -INFINITY;
#pragma unroll
        for (int t = 0; t < kGroups; ++t) m = fmaxf(m, s_m[t][g]);
        float l = 0.f, acc = 0.f;
#pragma unroll
        for (int t = 0; t < kGroups; ++t) {
            const float w = (s_m[t][g] == -INFINITY) ? 0.f : exp2f(s_m[t][g] - m);
            l += s_l[t][g] * w;
            acc += s_o[t][g][d] * w;
        }
        const size_t base = (((size_t)b * n_q + h0 + g) * n_splits + split);
        part_o[base * D + d] = acc;
        if (d == 0) {
            part_ml[base * 2 + 0] = m;       // log2 domain, like the weights of the merge kernel
            part_ml[base * 2 + 1] = l;
        }
    }
}

// Tensor-core variant for GQA groups of 2 .. 16 query heads per KV head (every Llama / Mixtral configuration): the
// `group` query heads that share a KV head are the M = 16 rows of mma.sync m16n8k16 tiles (rows >= group are zero), so a
// key costs two tensor-core instructions per 8 keys instead of ~30 FMA-pipe instructions per key, and K / V stream
// through a per-warp cp.async ring, so the next two tiles are in flight while the current one is multiplied (the FMA
// kernel above issues its loads and then waits for them, every iteration).
//
//   grid = (B * n_kv, n_splits), 4 warps; warp w owns the 32-key tiles w, w + 4, ... of the split and keeps its own
//   online-softmax state (m, l, O[16 x D]) in registers; the four states merge through shared memory at the end.
//   Per tile: S[16 x 32] = Q K^T (D / 16 k-steps x 4 key blocks), P = exp2(scale * (S - m)), O += P V
//   (2 k-steps x D / 8 channel blocks, V fragments through ldmatrix.trans).
template <int D>
struct MmaDecodeCfg {
    static constexpr int kWarps = 4;
    static constexpr int kTile = 32;                          // keys per warp step
    static constexpr int kStages = 3;
    static constexpr int kTileBytes = kTile * D * 2;          // one of K or V
    static constexpr int kWarpBytes = kStages * 2 * kTileBytes;
    static constexpr int kSmemBytes = kWarps * kWarpBytes;    // 96 KB (D = 64) / 192 KB (D = 128)
    static_assert(kWarps * 16 * (D + 4) * 4 + kWarps * 32 * 4 <= kSmemBytes, "merge buffers alias the ring");
};

template <int D>
__global__ void __launch_bounds__(128) decode_attn_mma_kernel(
    const bf16* __restrict__ q,  // [B, n_q, D]
    CacheView cache, const int* __restrict__ lengths, float* __restrict__ part_o,  // [B, n_q, splits, D]
    float* __restrict__ part_ml,                                                   // [B, n_q, splits, 2]
    int n_q, int n_kv, int n_splits, float scale_log2) {
    using C = MmaDecodeCfg<D>;
    extern __shared__ __align__(128) uint8_t dsmem[];
    pdl_trigger();
    pdl_wait();     // q and the newest cache row come from the projection kernel right before this one
    const int group = n_q / n_kv;
    const int b = blockIdx.x / n_kv, hk = blockIdx.x % n_kv;
    const int h0 = hk * group;
    const int split = blockIdx.y;
    const int len = lengths[b];
    const int per = ((len + n_splits - 1) / n_splits + C::kTile - 1) / C::kTile * C::kTile;   // whole tiles per split
    const int start = split * per;
    const int end = min(len, start + per);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g4 = lane >> 2, t4 = lane & 3;
    const uint32_t ring = (uint32_t)__cvta_generic_to_shared(dsmem) + warp * C::kWarpBytes;

    // Q fragments: rows = query heads of this KV head (zero rows beyond the group)
    uint32_t qa[D / 16][4];
    {
        const uint32_t* q0 = reinterpret_cast<const uint32_t*>(q + ((size_t)b * n_q + h0 + g4) * D);
        const uint32_t* q1 = reinterpret_cast<const uint32_t*>(q + ((size_t)b * n_q + h0 + g4 + 8) * D);
        const bool r0 = g4 < group, r1 = g4 + 8 < group;
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
            qa[kk][0] = r0 ? q0[kk * 8 + t4] : 0u;
            qa[kk][1] = r1 ? q1[kk * 8 + t4] : 0u;
            qa[kk][2] = r0 ? q0[kk * 8 + 4 + t4] : 0u;
            qa[kk][3] = r1 ? q1[kk * 8 + 4 + t4] : 0u;
        }
    }
    float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
    float o[D / 8][4];
#pragma unroll
    for (int j = 0; j < D / 8; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;

    const int n_tiles = end > start ? (end - start + C::kTile - 1) / C::kTile : 0;
    const int my_tiles = n_tiles > warp ? (n_tiles - warp + C::kWarps - 1) / C::kWarps : 0;
    auto issue = [&](int i) {       // tile i of this warp -> ring stage i % kStages (an empty group beyond the last tile)
        if (i < my_tiles) {
            const int key0 = start + (warp + i * C::kWarps) * C::kTile;
            const uint32_t sk = ring + (i % C::kStages) * 2 * C::kTileBytes, sv = sk + C::kTileBytes;
            constexpr int CH = D / 8;
#pragma unroll
            for (int j = 0; j < C::kTile * CH / 32; ++j) {
                const int idx = lane + j * 32;
                const int r = idx / CH, c = idx % CH;
                const int pos = key0 + r;
                const bool ok = pos < end;
                const size_t off = row_ptr(cache.k, cache, b, hk, ok ? pos : start, D) - cache.k + c * 8;
                mma::cp_async16(mma::tile_addr<D>(sk, r, c), cache.k + off, ok);
                mma::cp_async16(mma::tile_addr<D>(sv, r, c), cache.v + off, ok);   // rows past the end: zeros (0 * NaN)
            }
        }
        mma::cp_async_commit();
    };
#pragma unroll
    for (int i = 0; i < C::kStages - 1; ++i) issue(i);

    for (int i = 0; i < my_tiles; ++i) {
        issue(i + C::kStages - 1);
        mma::cp_async_wait<C::kStages - 1>();
        __syncwarp();
        const int key0 = start + (warp + i * C::kWarps) * C::kTile;
        const uint32_t sk = ring + (i % C::kStages) * 2 * C::kTileBytes, sv = sk + C::kTileBytes;
        float s[4][4];
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
#pragma unroll
            for (int nb = 0; nb < 2; ++nb) {
                uint32_t kb[4];
                mma::load_b<D>(kb, sk, nb * 16, kk, lane);
                mma::mma16816(s[2 * nb], qa[kk], kb[0], kb[1]);
                mma::mma16816(s[2 * nb + 1], qa[kk], kb[2], kb[3]);
            }
        }
        if (key0 + C::kTile > end) {      // the last tile of the split: keys past the end do not take part
#pragma unroll
            for (int nt = 0; nt < 4; ++nt)
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    if (key0 + nt * 8 + 2 * t4 + (c & 1) >= end) s[nt][c] = -INFINITY;
        }
        float alpha[2], mneg[2];
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            float mx = fmaxf(s[0][2 * r], s[0][2 * r + 1]);
#pragma unroll
            for (int nt = 1; nt < 4; ++nt) mx = fmaxf(mx, fmaxf(s[nt][2 * r], s[nt][2 * r + 1]));
            mx = mma::quad_max(mx);
            const float m_new = fmaxf(m_run[r], mx);      // finite: the first key of a tile is always inside the split
            alpha[r] = exp2f((m_run[r] - m_new) * scale_log2);
            m_run[r] = m_new;
            mneg[r] = -m_new * scale_log2;
            l_run[r] *= alpha[r];
        }
        uint32_t pa[2][4];
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
            const float p0 = exp2f(fmaf(s[nt][0], scale_log2, mneg[0])), p1 = exp2f(fmaf(s[nt][1], scale_log2, mneg[0]));
            const float p2 = exp2f(fmaf(s[nt][2], scale_log2, mneg[1])), p3 = exp2f(fmaf(s[nt][3], scale_log2, mneg[1]));
            l_run[0] += p0 + p1;
            l_run[1] += p2 + p3;
            pa[nt >> 1][(nt & 1) * 2] = mma::pack2(p0, p1);
            pa[nt >> 1][(nt & 1) * 2 + 1] = mma::pack2(p2, p3);
        }
#pragma unroll
        for (int j = 0; j < D / 8; ++j) {
            o[j][0] *= alpha[0];
            o[j][1] *= alpha[0];
            o[j][2] *= alpha[1];
            o[j][3] *= alpha[1];
        }
#pragma unroll
        for (int ks = 0; ks < 2; ++ks) {
#pragma unroll
            for (int dj = 0; dj < D / 16; ++dj) {
                uint32_t vb[4];
                mma::load_bt<D>(vb, sv, ks * 16, dj, lane);
                mma::mma16816(o[2 * dj], pa[ks], vb[0], vb[1]);
                mma::mma16816(o[2 * dj + 1], pa[ks], vb[2], vb[3]);
            }
        }
        __syncwarp();       // every lane is done with this stage before the next issue() overwrites it
    }
    mma::cp_async_wait<0>();
    __syncthreads();        // the ring is dead: reuse it for the merge of the four warp states

    float* s_o = reinterpret_cast<float*>(dsmem);                          // [warps][16][D + 4]
    float* s_ml = s_o + C::kWarps * 16 * (D + 4);                           // [warps][16][2]
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const float l = mma::quad_sum(l_run[r]);
        const int row = g4 + 8 * r;
        if (t4 == 0) {
            s_ml[(warp * 16 + row) * 2] = m_run[r] * scale_log2;           // log2 domain, like the merge kernel expects
            s_ml[(warp * 16 + row) * 2 + 1] = l;
        }
#pragma unroll
        for (int j = 0; j < D / 8; ++j) {
            float* dst = s_o + (warp * 16 + row) * (D + 4) + j * 8 + 2 * t4;
            dst[0] = o[j][2 * r];
            dst[1] = o[j][2 * r + 1];
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < group * D; i += blockDim.x) {
        const int g = i / D, d = i % D;
        float m = -INFINITY;
#pragma unroll
        for (int w = 0; w < C::kWarps; ++w) m = fmaxf(m, s_ml[(w * 16 + g) * 2]);
        float l = 0.f, acc = 0.f;
#pragma unroll
        for (int w = 0; w < C::kWarps; ++w) {
            const float mw = s_ml[(w * 16 + g) * 2];
            const float wgt = (mw == -INFINITY) ? 0.f : exp2f(mw - m);
            l += s_ml[(w * 16 + g) * 2 + 1] * wgt;
            acc += s_o[(w * 16 + g) * (D + 4) + d] * wgt;
        }
        const size_t base = (((size_t)b * n_q + h0 + g) * n_splits + split);
        part_o[base * D + d] = acc;
        if (d == 0) {
            part_ml[base * 2 + 0] = m;
            part_ml[base * 2 + 1] = l;
        }
    }
}

template <int D>
__global__ void decode_attn_merge_kernel(const float* __restrict__ part_o, const float* __restrict__ part_ml,
                                         bf16* __restrict__ out, int n_splits) {
    pdl_trigger();
    pdl_wait();
    const size_t bh = blockIdx.x;  // b * n_q + h
    const int d = threadIdx.x;
    float m = -INFINITY;
    for (int s = 0; s < n_splits; ++s) m = fmaxf(m, part_ml[(bh * n_splits + s) * 2]);
    float l = 0.f, o = 0.f;
    for (int s = 0; s < n_splits; ++s) {
        const float ms = part_ml[(bh * n_splits + s) * 2];
        const float w = (ms == -INFINITY) ? 0.f : exp2f(ms - m);
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
    // enough CTAs to cover the machine, at least 128 keys per split (two iterations of the 64-key CTA step)
    int want = (2 * nxd_num_sms() + B * n_kv - 1) / (B * n_kv);
    int by_len = (max_len + 127) / 128;
    int s = want < by_len ? want : by_len;
    if (s < 1) s = 1;
    if (s > 64) s = 64;
    return s;
}

// scratch: B * n_q * n_splits * (D + 2) floats.  seq_ids: [B] int32 cache line (contiguous cache) or block-table row
// (paged cache) of every batch row, NULL = the row index (continuous batching reads the lines in place).
static int decode_attn_impl(const void* q, const void* cache_k, const void* cache_v, const int* lengths,
                            const int* block_table, const int* seq_ids, long long stride_b, long long stride_h,
                            int page_size, int max_pages, void* out, float* scratch, int B, int n_q, int n_kv, int D,
                            int n_splits, float scale, void* stream) {
    NXD_REQUIRE(D == 64 || D == 128, "decode_attn: head dim %d not in {64, 128}", D);
    NXD_REQUIRE(n_q % n_kv == 0 && n_q / n_kv <= kMaxGroup, "decode_attn: unsupported GQA group %d/%d", n_q, n_kv);
    NXD_REQUIRE(scratch != nullptr && n_splits >= 1, "decode_attn: scratch required");
    if (B == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    CacheView cv{(const bf16*)cache_k, (const bf16*)cache_v, block_table, seq_ids, stride_b, stride_h,
                 page_size > 0 ? page_size : 1, max_pages};
    float* part_o = scratch;
    float* part_ml = scratch + (size_t)B * n_q * n_splits * D;
    const int group = n_q / n_kv;
    const int G = group <= 4 ? group : (group % 4 == 0 ? 4 : (group % 3 == 0 ? 3 : (group % 2 == 0 ? 2 : 1)));
    dim3 grid(B * n_kv * (group / G), n_splits);
    const float scale_log2 = scale * 1.4426950408889634f;
#define NXD_DECODE_LAUNCH(DD, GG)                                                                                  \
    NXD_CHECK_CUDA(nxd_launch_pdl(decode_attn_split_kernel<DD, GG>, grid, dim3(128), 0, s, (const bf16*)q, cv, lengths, \
                                  part_o, part_ml, n_q, n_kv, n_splits, scale_log2))
#define NXD_DECODE_DISPATCH(DD)                 \
    switch (G) {                                \
        case 1: NXD_DECODE_LAUNCH(DD, 1); break; \
        case 2: NXD_DECODE_LAUNCH(DD, 2); break; \
        case 3: NXD_DECODE_LAUNCH(DD, 3); break; \
        default: NXD_DECODE_LAUNCH(DD, 4); break; \
    }
    // GQA groups of 2 .. 16 heads: tensor-core kernel (NXD_DECODE_ATTN=fma keeps the FMA kernel for comparison)
    static const bool force_fma = [] {
        const char* e = getenv("NXD_DECODE_ATTN");
        return e != nullptr && e[0] == 'f';
    }();
    const bool use_mma = group >= 2 && group <= 16 && !force_fma;
    if (use_mma) {
        dim3 mgrid(B * n_kv, n_splits);
        if (D == 64) {
            static bool configured = false;
            if (!configured) {
                NXD_CHECK_CUDA(cudaFuncSetAttribute(decode_attn_mma_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                    MmaDecodeCfg<64>::kSmemBytes));
                configured = true;
            }
            NXD_CHECK_CUDA(nxd_launch_pdl(decode_attn_mma_kernel<64>, mgrid, dim3(128), MmaDecodeCfg<64>::kSmemBytes, s,
                                          (const bf16*)q, cv, lengths, part_o, part_ml, n_q, n_kv, n_splits, scale_log2));
        } else {
            static bool configured = false;
            if (!configured) {
                NXD_CHECK_CUDA(cudaFuncSetAttribute(decode_attn_mma_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                    MmaDecodeCfg<128>::kSmemBytes));
                configured = true;
            }
            NXD_CHECK_CUDA(nxd_launch_pdl(decode_attn_mma_kernel<128>, mgrid, dim3(128), MmaDecodeCfg<128>::kSmemBytes, s,
                                          (const bf16*)q, cv, lengths, part_o, part_ml, n_q, n_kv, n_splits, scale_log2));
        }
        NXD_LAUNCH_CHECK();
        if (D == 64) {
            NXD_CHECK_CUDA(nxd_launch_pdl(decode_attn_merge_kernel<64>, dim3(B * n_q), dim3(64), 0, s, (const float*)part_o,
                                          (const float*)part_ml, (bf16*)out, n_splits));
        } else {
            NXD_CHECK_CUDA(nxd_launch_pdl(decode_attn_merge_kernel<128>, dim3(B * n_q), dim3(128), 0, s,
                                          (const float*)part_o, (const float*)part_ml, (bf16*)out, n_splits));
        }
        NXD_LAUNCH_CHECK();
        return 0;
    }
    if (D == 64) {
        NXD_DECODE_DISPATCH(64)
        NXD_LAUNCH_CHECK();
        NXD_CHECK_CUDA(nxd_launch_pdl(decode_attn_merge_kernel<64>, dim3(B * n_q), dim3(64), 0, s, (const float*)part_o,
                                      (const float*)part_ml, (bf16*)out, n_splits));
    } else {
        NXD_DECODE_DISPATCH(128)
        NXD_LAUNCH_CHECK();
        NXD_CHECK_CUDA(nxd_launch_pdl(decode_attn_merge_kernel<128>, dim3(B * n_q), dim3(128), 0, s, (const float*)part_o,
                                      (const float*)part_ml, (bf16*)out, n_splits));
    }
#undef NXD_DECODE_DISPATCH
#undef NXD_DECODE_LAUNCH
    NXD_LAUNCH_CHECK();
    return 0;
}

NXD_API int nxd_decode_attn(const void* q, const void* cache_k, const void* cache_v, const int* lengths,
                            const int* block_table, long long stride_b, long long stride_h, int page_size,
                            int max_pages, void* out, float* scratch, int B, int n_q, int n_kv, int D,
                            int n_splits, float scale, void* stream) {
    return decode_attn_impl(q, cache_k, cache_v, lengths, block_table, nullptr, stride_b, stride_h, page_size, max_pages, out,
                            scratch, B, n_q, n_kv, D, n_splits, scale, stream);
}

NXD_API int nxd_decode_attn_lines(const void* q, const void* cache_k, const void* cache_v, const int* lengths,
                                  const int* block_table, const int* seq_ids, long long stride_b, long long stride_h,
                                  int page_size, int max_pages, void* out, float* scratch, int B, int n_q, int n_kv,
                                  int D, int n_splits, float scale, void* stream) {
    return decode_attn_impl(q, cache_k, cache_v, lengths, block_table, seq_ids, stride_b, stride_h, page_size, max_pages, out,
                            scratch, B, n_q, n_kv, D, n_splits, scale, stream);
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
