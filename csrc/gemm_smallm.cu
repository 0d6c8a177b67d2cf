// Weight-streaming tcgen05 GEMM for 9..128 activation rows: batched token generation, the speculation verify step, the
// last pipeline micro-batches of serving (the reference runs its token-generation graph at batch sizes up
// to the bucket size, src/neuronx_distributed/trace/model_builder.py buckets; parallel linear layers of
// parallel_layers/layers.py at [batch, 1, hidden]).
//
//   out[B, N] = x[B, K] @ w[N, K]^T (+ residual[B, N])          bf16 in, fp32 accumulation, bf16 out
//
// and two variants with the neighbouring element-wise work of a decoder layer in the epilogue (the batched counterparts of
// decode_fused.cu, which covers <= 8 rows with FMA GEMVs):
//
//   QKV    w = fused [q | k | v] head-major rows: RoPE(q, k) at positions[b], q -> q_out, k / v -> KV cache (contiguous
//          or paged).  A 128-row weight slab is one head (D = 128) or two (D = 64), so the rotation partner (d, d + D/2)
//          is always inside the CTA's own tile.
//   SwiGLU w = [gate rows | up rows]: a CTA streams gate slab t AND up slab t (two accumulators), out = silu(g) * u.
//
// "Swap A/B": the WEIGHT rows are the UMMA M dimension (128 per CTA, the full tensor-core width), the padded token count
// is the UMMA N dimension (16 / 32 / 64 / 128), so a 16-row batch wastes nothing of M and the kernel is a pure stream of
// the weight matrix through TMA:
//
//   warp 0     TMA producer: per 64-wide k block one 128 x 64 weight box (16 KB) + the BT x 64 activation box
//   warp 1     tcgen05.mma issuer (cta_group::1, M = 128, N = BT), accumulator D^T[n, b] in tensor memory
//   warps 2-5  epilogue: tcgen05.ld -> fp32 staging in shared memory (aliases the drained pipeline stages)
//
// Split-K over a thread-block CLUSTER: narrow layers (N = 2048 -> 16 weight slabs) are cut into up to 8 K slices, one
// per CTA of a cluster, and the partial sums are added through distributed shared memory: token columns
// [r * tpr, (r + 1) * tpr) belong to CTA r of the cluster.  Every CTA lays its accumulator out per owner in its own shared
// memory (the drained pipeline stages) and sends each owner its block with ONE bulk asynchronous copy
// (cp.async.bulk.shared::cluster.shared::cta) that completes on the owner's mbarrier -- measured: scalar
// st.shared::cluster pushes of the same data took 4.8 us for 64 tokens, the copies take 0.5.  Each owner then adds its
// slots in rank order (deterministic), adds the residual and writes 16-byte bf16 vectors.  No workspace, no atomics, no
// cluster barrier after start-up, one launch.
#include <stdlib.h>

#include "decode_common.cuh"
#include "gemm_tcgen05.cuh"
#include "ptx.cuh"

namespace {

using nxd_decode::CacheDst;
using nxd_decode::cache_row;
using nxd_gemm::make_tmap_2d;

constexpr int kWRows = 128;     // weight rows per CTA and accumulator (UMMA M)
constexpr int kBK = 64;         // 64 bf16 = one SWIZZLE_128B row
constexpr int kThreads = 192;   // TMA warp, MMA warp, 4 epilogue warps
constexpr int kMaxSplits = 8;   // portable cluster size

enum { kEpiPlain = 0, kEpiQkvRope = 1, kEpiSwiglu = 2 };

struct EpiArgs {
    bf16* out;                 // plain: [B, N]; SwiGLU: [B, I]
    long long ldo;
    const bf16* residual;      // plain only, may be null
    long long ldr;
    // QKV + RoPE + cache write
    const float* cos_t;        // [max_pos, D / 2]
    const float* sin_t;
    const int* positions;      // [B]
    bf16* q_out;               // [B, n_q, D]
    CacheDst cache;
    int n_q, n_kv, D;
    int up_row0;               // SwiGLU: first "up" row of w (= I)
};

template <int BT, int EPI>
struct Cfg {
    static constexpr int kAcc = EPI == kEpiSwiglu ? 2 : 1;           // accumulators (weight slabs) per CTA
    static constexpr int kWBytes = kAcc * kWRows * kBK * 2;          // 16 / 32 KB
    static constexpr int kXBytes = BT * kBK * 2;                     // 2 .. 16 KB
    static constexpr int kStageBytes = kWBytes + kXBytes;
    // one accumulator: pipeline + staging <= ~110 KB -> two CTAs per SM (BT = 128: one CTA, deeper pipeline);
    // two accumulators: one CTA per SM with 128-160 KB of weights in flight
    static constexpr int kStages = kAcc == 2 ? (BT <= 32 ? 5 : 4) : (BT == 16 ? 5 : BT == 32 ? 4 : BT == 64 ? 3 : 4);
    static constexpr int kPipeBytes = kStages * kStageBytes;
    // receive buffer, fp32 [accumulator][slot = source rank * tokens_per_rank + local token][128]; NOT aliased with the
    // pipeline: peers copy into it while this CTA may still be streaming.  The send buffer (same size, [accumulator]
    // [owner][local token][128]) aliases the pipeline stages, which are dead once the last MMA has retired.
    static constexpr int kStagingBytes = kAcc * BT * kWRows * 4;
    static constexpr int kBarOffset = kPipeBytes + kStagingBytes;
    static constexpr int kSmemBytes = kBarOffset + (2 * kStages + 2) * 8 + 16 + 1024;
    static_assert(kStagingBytes <= kPipeBytes, "the send buffer aliases the pipeline stages");
    static constexpr int kTmemCols = kAcc * BT < 32 ? 32 : kAcc * BT;
    static_assert(kStageBytes % 1024 == 0, "stages must keep the 1024-byte swizzle alignment");
    static_assert(kSmemBytes <= 232448, "shared memory budget exceeded");
};

__device__ __forceinline__ uint32_t map_to_cta(uint32_t local_addr, uint32_t cta) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(cta));
    return remote;
}
// bytes of this CTA's shared memory -> shared memory of a CTA of the cluster, completing (complete_tx) on ITS mbarrier
__device__ __forceinline__ void bulk_copy_to_cta(uint32_t remote_dst, uint32_t local_src, uint32_t bytes,
                                                 uint32_t remote_bar) {
    asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     remote_dst),
                 "r"(local_src), "r"(bytes), "r"(remote_bar)
                 : "memory");
}
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
__device__ __forceinline__ void epilogue_bar() { asm volatile("bar.sync 1, 128;" ::: "memory"); }
__device__ __forceinline__ float round_bf16(float v) { return __bfloat162float(__float2bfloat16(v)); }

// Phase stamps of CTA 0 (SM clock) for kernel tuning: build with -DNXD_SMALLM_TRACE, read with nxd_smallm_trace().
#ifdef NXD_SMALLM_TRACE
__device__ long long g_smallm_trace[16];
#define TRACE_STAMP(i, cond)                                   \
    do {                                                       \
        if (blockIdx.x == 0 && (cond)) g_smallm_trace[i] = clock64(); \
    } while (0)
#else
#define TRACE_STAMP(i, cond) do { } while (0)
#endif

template <int BT, int EPI>
__global__ void __launch_bounds__(kThreads) gemm_smallm_kernel(const __grid_constant__ CUtensorMap tmap_w,
                                                              const __grid_constant__ CUtensorMap tmap_x,
                                                              const EpiArgs ep, int B, int N, int K, int splits) {
    using C = Cfg<BT, EPI>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = ptx::align_smem_1024(smem_raw);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + C::kBarOffset);
    uint64_t* empty_bar = full_bar + C::kStages;
    uint64_t* acc_bar = empty_bar + C::kStages;
    uint64_t* recv_bar = acc_bar + 1;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(recv_bar + 1);
    float* staging = reinterpret_cast<float*>(smem + C::kPipeBytes);      // receive buffer
    float* sendbuf = reinterpret_cast<float*>(smem);                      // aliases the pipeline

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int split = splits > 1 ? (int)ptx::cluster_ctarank() : 0;
    const int n0 = (blockIdx.x / splits) * kWRows;      // first weight row of the slab (SwiGLU: of the gate slab)
    const int nkb = (K + kBK - 1) / kBK;
    const int per = (nkb + splits - 1) / splits;
    const int kb_begin = split * per;
    const int kb_end = min(nkb, kb_begin + per);
    const bool has_k = kb_begin < kb_end;
    const int tpr = (B + splits - 1) / splits;          // token columns per owner CTA

    TRACE_STAMP(0, threadIdx.x == 0);
    pdl_trigger();
    // warp 0 lane 0 owns the barriers and the TMA queue: it starts the weight stream right after initialising them,
    // before the CTA (or, with split-K, the cluster) has met -- the weights do not depend on the previous kernel either
    auto load_w = [&](int kb, int stage) {
        uint8_t* sw = smem + stage * C::kStageBytes;
        ptx::mbar_arrive_expect_tx(&full_bar[stage], C::kStageBytes);
        ptx::tma_load_2d(&tmap_w, &full_bar[stage], sw, kb * kBK, n0);
        if constexpr (C::kAcc == 2)
            ptx::tma_load_2d(&tmap_w, &full_bar[stage], sw + kWRows * kBK * 2, kb * kBK, ep.up_row0 + n0);
    };
    auto load_x = [&](int kb, int stage) {
        ptx::tma_load_2d(&tmap_x, &full_bar[stage], smem + stage * C::kStageBytes + C::kWBytes, kb * kBK, 0);
    };
    const int pre = min(C::kStages, kb_end - kb_begin);
    if (warp == 0 && lane == 0) {
        ptx::prefetch_tensormap(&tmap_w);
        ptx::prefetch_tensormap(&tmap_x);
        for (int i = 0; i < C::kStages; ++i) {
            ptx::mbar_init(&full_bar[i], 1);
            ptx::mbar_init(&empty_bar[i], 1);
        }
        ptx::mbar_init(acc_bar, 1);
        ptx::mbar_init(recv_bar, 1);
        ptx::fence_barrier_init();
        // every CTA of the cluster sends this one a [tpr][128] fp32 block per accumulator
        if (splits > 1 && split * tpr < B) ptx::mbar_arrive_expect_tx(recv_bar, (uint32_t)(splits * C::kAcc * tpr * kWRows * 4));
        for (int i = 0; i < pre; ++i) load_w(kb_begin + i, i);
    }
    if (warp == 1) ptx::tmem_alloc<1>(tmem_ptr_smem, C::kTmemCols);
    ptx::tcgen05_fence_before();
    __syncthreads();
    // split-K: a peer may only send once this CTA's receive barrier exists.  Every thread arrives at the cluster barrier
    // now and waits for it after its main loop (measured: a full cluster barrier here costs ~1 us of start-up)
    if (splits > 1) asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    ptx::tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;
    TRACE_STAMP(1, threadIdx.x == 0);

    if (warp == 0) {
        if (lane == 0) {
            pdl_wait();         // the activations come from the previous kernel (programmatic dependent launch)
            TRACE_STAMP(2, true);
            for (int i = 0; i < pre; ++i) load_x(kb_begin + i, i);
            int stage = pre == C::kStages ? 0 : pre;
            uint32_t phase = pre == C::kStages ? 1 : 0;
            for (int kb = kb_begin + pre; kb < kb_end; ++kb) {
                ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                load_w(kb, stage);
                load_x(kb, stage);
                if (++stage == C::kStages) { stage = 0; phase ^= 1; }
            }
        }
        __syncwarp();
        if (splits > 1) cluster_wait();
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = ptx::make_idesc_bf16(kWRows, BT, 0, 0);
            int stage = 0;
            uint32_t phase = 0;
            for (int kb = kb_begin; kb < kb_end; ++kb) {
                ptx::mbar_wait(&full_bar[stage], phase);
                TRACE_STAMP(3, kb == kb_begin);
                ptx::tcgen05_fence_after();
                const uint32_t sw = ptx::smem_u32(smem + stage * C::kStageBytes);
                const uint32_t sx = sw + C::kWBytes;
#pragma unroll
                for (int k = 0; k < kBK / 16; ++k) {
                    const uint64_t db = ptx::make_smem_desc(sx + k * 32, 16, 1024);
                    const uint32_t accumulate = (kb > kb_begin || k > 0) ? 1u : 0u;
#pragma unroll
                    for (int a = 0; a < C::kAcc; ++a) {
                        const uint64_t da = ptx::make_smem_desc(sw + a * (kWRows * kBK * 2) + k * 32, 16, 1024);
                        ptx::umma_bf16<1>(tmem_base + a * BT, da, db, idesc, accumulate);
                    }
                }
                ptx::umma_commit(&empty_bar[stage]);
                if (++stage == C::kStages) { stage = 0; phase ^= 1; }
            }
            TRACE_STAMP(4, true);
            ptx::umma_commit(acc_bar);   // arrives when every MMA above has retired (immediately when there were none)
        }
        __syncwarp();
        if (splits > 1) cluster_wait();
    } else {
        // accumulator column b -> (owner b / tpr, local token b % tpr): straight into the receive buffer when there is
        // no split-K, else into the send buffer [owner][local token][128]; consecutive lanes = consecutive n:
        // conflict-free 128-byte rows per warp store
        const int q = warp & 3;
        pdl_wait();     // the epilogue reads the residual / positions and writes global memory
        ptx::mbar_wait(acc_bar, 0);
        TRACE_STAMP(5, threadIdx.x == 64);
        ptx::tcgen05_fence_after();
        const int nl = q * 32 + lane;
#pragma unroll
        for (int a = 0; a < C::kAcc; ++a) {
            float* region = (splits == 1 ? staging : sendbuf) + a * (BT * kWRows);
#pragma unroll
            for (int c = 0; c < (BT + 31) / 32; ++c) {
                if (c * 32 >= B) break;
                uint32_t r[32];
                if (has_k) {
                    ptx::tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(q * 32) << 16) + a * BT + c * 32, r);
                    ptx::tmem_ld_wait();
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) r[j] = 0u;
                }
                // token b sits at row b of the buffer in both cases: owner * tpr + local = b
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (c * 32 + j < B) region[(c * 32 + j) * kWRows + nl] = __uint_as_float(r[j]);
            }
        }
        TRACE_STAMP(6, threadIdx.x == 64);
        if (splits > 1) {
            ptx::fence_proxy_async();       // the stores above are read by the bulk-copy engine
            cluster_wait();                 // every peer has initialised its receive barrier
            epilogue_bar();
            if (threadIdx.x == 64) {
                const uint32_t bytes = (uint32_t)(tpr * kWRows * 4);
                for (int a = 0; a < C::kAcc; ++a) {
                    for (int o = 0; o < splits; ++o) {
                        if (o * tpr >= B) break;     // owners without tokens expect nothing ... see expect_tx below
                        const uint32_t src = ptx::smem_u32(sendbuf + a * (BT * kWRows) + o * tpr * kWRows);
                        const uint32_t dst = ptx::smem_u32(staging + a * (BT * kWRows) + (split * tpr) * kWRows);
                        bulk_copy_to_cta(map_to_cta(dst, (uint32_t)o), src, bytes, map_to_cta(ptx::smem_u32(recv_bar),
                                                                                                (uint32_t)o));
                    }
                }
                ptx::bulk_commit();
            }
            // the blocks of all peers have landed in this CTA's receive buffer
            if (split * tpr < B) ptx::mbar_wait(recv_bar, 0);
        }
    }

    TRACE_STAMP(7, threadIdx.x == 64);
    if (splits == 1) __syncthreads();

    if (warp >= 2) {
        // owner pass: 16 threads x 8 columns cover the 128 weight rows of one token, 8 tokens per sweep
        const int idx = threadIdx.x - 64;
        const int n8 = (idx & 15) * 8, jrow = idx >> 4;
        auto reduce8 = [&](int a, int j, float* acc) {
            const float* region = staging + a * (BT * kWRows);
            const float4* p = reinterpret_cast<const float4*>(region + j * kWRows + n8);
            const float4 lo = p[0], hi = p[1];
            acc[0] = lo.x; acc[1] = lo.y; acc[2] = lo.z; acc[3] = lo.w;
            acc[4] = hi.x; acc[5] = hi.y; acc[6] = hi.z; acc[7] = hi.w;
            for (int s = 1; s < splits; ++s) {
                const float4* ps = reinterpret_cast<const float4*>(region + (s * tpr + j) * kWRows + n8);
                const float4 l2 = ps[0], h2 = ps[1];
                acc[0] += l2.x; acc[1] += l2.y; acc[2] += l2.z; acc[3] += l2.w;
                acc[4] += h2.x; acc[5] += h2.y; acc[6] += h2.z; acc[7] += h2.w;
            }
        };
        const int n = n0 + n8;
        if constexpr (EPI == kEpiPlain) {
            const bf16* residual = ep.residual;
            bf16* out = ep.out;
            const long long ldo = ep.ldo, ldr = ep.ldr;
            const bool vec = ((ldo & 7) == 0) && ((reinterpret_cast<uintptr_t>(out) & 15) == 0) &&
                             (residual == nullptr ||
                              ((ldr & 7) == 0 && (reinterpret_cast<uintptr_t>(residual) & 15) == 0));
            for (int j = jrow; j < tpr; j += 8) {
                const int b = split * tpr + j;
                if (b >= B) break;
                float acc[8];
                reduce8(0, j, acc);
                if (vec && n + 8 <= N) {
                    if (residual != nullptr) {
                        float f[8];
                        unpack8(*reinterpret_cast<const bf16x8*>(residual + (long long)b * ldr + n), f);
#pragma unroll
                        for (int e = 0; e < 8; ++e) acc[e] += f[e];
                    }
                    *reinterpret_cast<bf16x8*>(out + (long long)b * ldo + n) = pack8(acc);
                } else {
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        if (n + e < N) {
                            float v = acc[e];
                            if (residual != nullptr) v += __bfloat162float(residual[(long long)b * ldr + n + e]);
                            out[(long long)b * ldo + n + e] = __float2bfloat16(v);
                        }
                    }
                }
            }
        } else if constexpr (EPI == kEpiSwiglu) {
            // N = I here (rows of ONE of the two stacked matrices); out [B, I], I % 8 == 0
            for (int j = jrow; j < tpr; j += 8) {
                const int b = split * tpr + j;
                if (b >= B) break;
                float g[8], u[8];
                reduce8(0, j, g);
                reduce8(1, j, u);
                if (n < N) {
#pragma unroll
                    for (int e = 0; e < 8; ++e) {      // the unfused path rounds the projection to bf16 first
                        const float gv = round_bf16(g[e]), uv = round_bf16(u[e]);
                        g[e] = gv / (1.f + __expf(-gv)) * uv;
                    }
                    *reinterpret_cast<bf16x8*>(ep.out + (long long)b * ep.ldo + n) = pack8(g);
                }
            }
        } else {
            // pass 1: reduced, bf16-rounded head values back into slot 0; pass 2: rotate with the partner half
            if (splits > 1) {
                for (int j = jrow; j < tpr; j += 8) {
                    if (split * tpr + j >= B) break;
                    float acc[8];
                    reduce8(0, j, acc);
                    float4* p = reinterpret_cast<float4*>(staging + j * kWRows + n8);
                    p[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
                    p[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
                }
                epilogue_bar();
            }
            const int D = ep.D, half = D >> 1;
            const int head = n / D, d = n - head * D;
            const bool second = d >= half;
            const int dd = second ? d - half : d;                       // rotation index 0 .. D/2 - 1
            const int partner = second ? n8 - half : n8 + half;         // tile-local column of the other half
            for (int j = jrow; j < tpr; j += 8) {
                const int b = split * tpr + j;
                if (b >= B) break;
                if (n >= N) continue;
                float v[8];
                {
                    const float4* p = reinterpret_cast<const float4*>(staging + j * kWRows + n8);
                    const float4 lo = p[0], hi = p[1];
                    v[0] = lo.x; v[1] = lo.y; v[2] = lo.z; v[3] = lo.w;
                    v[4] = hi.x; v[5] = hi.y; v[6] = hi.z; v[7] = hi.w;
                }
                const int pos = ep.positions[b];
                bf16* dst;
                if (head < ep.n_q + ep.n_kv) {
                    float o[8], cs[8], sn[8];
                    const float4* p = reinterpret_cast<const float4*>(staging + j * kWRows + partner);
                    const float4 lo = p[0], hi = p[1];
                    o[0] = lo.x; o[1] = lo.y; o[2] = lo.z; o[3] = lo.w;
                    o[4] = hi.x; o[5] = hi.y; o[6] = hi.z; o[7] = hi.w;
                    const float4* pc = reinterpret_cast<const float4*>(ep.cos_t + (size_t)pos * half + dd);
                    const float4* psn = reinterpret_cast<const float4*>(ep.sin_t + (size_t)pos * half + dd);
                    const float4 c0 = pc[0], c1 = pc[1], s0 = psn[0], s1 = psn[1];
                    cs[0] = c0.x; cs[1] = c0.y; cs[2] = c0.z; cs[3] = c0.w;
                    cs[4] = c1.x; cs[5] = c1.y; cs[6] = c1.z; cs[7] = c1.w;
                    sn[0] = s0.x; sn[1] = s0.y; sn[2] = s0.z; sn[3] = s0.w;
                    sn[4] = s1.x; sn[5] = s1.y; sn[6] = s1.z; sn[7] = s1.w;
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        const float x = round_bf16(v[e]), y = round_bf16(o[e]);
                        v[e] = second ? x * cs[e] + y * sn[e] : x * cs[e] - y * sn[e];
                    }
                    dst = head < ep.n_q ? ep.q_out + ((size_t)b * ep.n_q + head) * D
                                        : ep.cache.k + cache_row(ep.cache, b, head - ep.n_q, pos, D);
                } else {
                    dst = ep.cache.v + cache_row(ep.cache, b, head - ep.n_q - ep.n_kv, pos, D);
                }
                *reinterpret_cast<bf16x8*>(dst + d) = pack8(v);
            }
        }
    }

    // owners read only their own shared memory; a CTA leaves once its outgoing copies have read the send buffer and
    // (waited above) its incoming blocks have landed
    TRACE_STAMP(8, threadIdx.x == 64);
    if (splits > 1 && threadIdx.x == 64) ptx::bulk_wait_read<0>();   // outgoing copies have read the send buffer
    __syncthreads();
    TRACE_STAMP(9, threadIdx.x == 0);
    if (warp == 1) ptx::tmem_dealloc<1>(tmem_base, C::kTmemCols);
}

int pick_splits(int n_tiles, int nkb, int ctas_per_sm) {
    // the largest power of two <= 8 that keeps the grid within one resident wave and leaves two k blocks per slice
    // (measured at the Llama-3.2 decode shapes: 16-24 slabs -> 8, 64 -> 4, 128 -> 2, the 1002 slabs of the LM head -> 1)
    const int slots = nxd_num_sms() * ctas_per_sm;
    int s = 1;
    while (s < kMaxSplits && n_tiles * s * 2 <= slots && nkb / (s * 2) >= 2) s *= 2;
    return s;
}

template <int BT, int EPI>
int launch(const CUtensorMap& tw, const CUtensorMap& tx, const EpiArgs& ep, int B, int N, int K, int splits,
           cudaStream_t stream) {
    using C = Cfg<BT, EPI>;
    auto kern = gemm_smallm_kernel<BT, EPI>;
    static bool configured = false;
    if (!configured) {
        NXD_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
        configured = true;
    }
    const int n_tiles = (N + kWRows - 1) / kWRows;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(n_tiles * splits));
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = C::kSmemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)splits;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = nxd_pdl_enabled();
    cfg.attrs = attr;
    cfg.numAttrs = 2;
    NXD_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, tw, tx, ep, B, N, K, splits));
    NXD_LAUNCH_CHECK();
    return 0;
}

// n_rows: weight rows covered by the grid (N, or I for SwiGLU); w_rows: rows of the weight tensor map
template <int EPI>
int run(const void* x, const void* w, const EpiArgs& ep, int B, int n_rows, int w_rows, int K, long long ldx,
        long long ldw, int splits, cudaStream_t st) {
    NXD_REQUIRE(B >= 1 && B <= (EPI == kEpiSwiglu ? 64 : 128) && n_rows >= 1 && K >= 1,
                "gemm_smallm: B=%d out of range (1..%d), N=%d, K=%d", B, EPI == kEpiSwiglu ? 64 : 128, n_rows, K);
    NXD_REQUIRE(K % 8 == 0 && ldx % 8 == 0 && ldw % 8 == 0, "gemm_smallm: K=%d and the row pitches must be multiples of 8",
                K);
    NXD_REQUIRE(((uintptr_t)x & 15) == 0 && ((uintptr_t)w & 15) == 0, "gemm_smallm: operands must be 16-byte aligned");
    const int bt = B <= 16 ? 16 : B <= 32 ? 32 : B <= 64 ? 64 : 128;
    const int n_tiles = (n_rows + kWRows - 1) / kWRows, nkb = (K + kBK - 1) / kBK;
    if (splits <= 0) {
        static const int forced = [] {      // NXD_SMALLM_SPLITS=1|2|4|8: debugging / tuning override of the heuristic
            const char* e = getenv("NXD_SMALLM_SPLITS");
            return e != nullptr ? atoi(e) : 0;
        }();
        splits = forced > 0 ? forced : pick_splits(n_tiles, nkb, (EPI == kEpiSwiglu || bt == 128) ? 1 : 2);
        while (splits > 1 && nkb / splits < 1) splits >>= 1;
    }
    NXD_REQUIRE(splits == 1 || splits == 2 || splits == 4 || splits == 8, "gemm_smallm: splits=%d", splits);
    CUtensorMap tw, tx;
    int rc = make_tmap_2d(&tw, w, (uint64_t)K, (uint64_t)w_rows, (uint64_t)ldw, kBK, kWRows);
    if (rc) return rc;
    rc = make_tmap_2d(&tx, x, (uint64_t)K, (uint64_t)B, (uint64_t)ldx, kBK, (uint32_t)bt);
    if (rc) return rc;
    switch (bt) {
        case 16: return launch<16, EPI>(tw, tx, ep, B, n_rows, K, splits, st);
        case 32: return launch<32, EPI>(tw, tx, ep, B, n_rows, K, splits, st);
        case 64: return launch<64, EPI>(tw, tx, ep, B, n_rows, K, splits, st);
        default:
            if constexpr (EPI == kEpiSwiglu) return -1;
            else return launch<128, EPI>(tw, tx, ep, B, n_rows, K, splits, st);
    }
}

}  // namespace

// x [B, K] (pitch ldx), w [N, K] (pitch ldw), residual [B, N] (pitch ldr) or NULL, out [B, N] (pitch ldo); bf16.
// 1 <= B <= 128; K, ldx, ldw multiples of 8; x and w 16-byte aligned.  splits: 0 = choose, else 1 / 2 / 4 / 8.
NXD_API int nxd_gemm_smallm_bf16(const void* x, const void* w, const void* residual, void* out, int B, int N, int K,
                                 long long ldx, long long ldw, long long ldo, long long ldr, int splits, void* stream) {
    EpiArgs ep = {};
    ep.out = (bf16*)out;
    ep.ldo = ldo;
    ep.residual = (const bf16*)residual;
    ep.ldr = ldr;
    return run<kEpiPlain>(x, w, ep, B, N, N, K, ldx, ldw, splits, (cudaStream_t)stream);
}

// Batched form of nxd_decode_qkv_rope_cache WITHOUT the norm (x is the normalised activation): w fused projection
// [(n_q + 2 n_kv) * D, K] head-major; RoPE(q, k) at positions[b]; q -> q_out [B, n_q, D]; k / v -> cache (same cache
// arguments as nxd_decode_qkv_rope_cache).  B <= 128, D in {64, 128}.
NXD_API int nxd_smallm_qkv_rope_cache(const void* x, long long ldx, const void* w, long long ldw, const float* cos_t,
                                      const float* sin_t, const int* positions, void* q_out, void* cache_k,
                                      void* cache_v, const int* seq_ids, const int* block_table, long long stride_b,
                                      long long stride_h, int page_size, int max_pages, int B, int K, int n_q, int n_kv,
                                      int D, int splits, void* stream) {
    NXD_REQUIRE(D == 64 || D == 128, "smallm_qkv: head dim %d not in {64, 128}", D);
    NXD_REQUIRE(((uintptr_t)q_out & 15) == 0 && ((uintptr_t)cache_k & 15) == 0 && ((uintptr_t)cache_v & 15) == 0 &&
                    stride_b % 8 == 0 && stride_h % 8 == 0,
                "smallm_qkv: q_out / cache must be 16-byte aligned with strides that are multiples of 8");
    EpiArgs ep = {};
    ep.cos_t = cos_t;
    ep.sin_t = sin_t;
    ep.positions = positions;
    ep.q_out = (bf16*)q_out;
    ep.cache = CacheDst{(bf16*)cache_k, (bf16*)cache_v, block_table, seq_ids, stride_b, stride_h,
                        page_size > 0 ? page_size : 1, max_pages};
    ep.n_q = n_q;
    ep.n_kv = n_kv;
    ep.D = D;
    const int N = (n_q + 2 * n_kv) * D;
    return run<kEpiQkvRope>(x, w, ep, B, N, N, K, ldx, ldw, splits, (cudaStream_t)stream);
}

// out[B, I] = silu(x W_gate^T) * (x W_up^T); w: [2 I, K] with the gate rows first.  B <= 64, I % 8 == 0.
NXD_API int nxd_smallm_gate_up_swiglu(const void* x, long long ldx, const void* w, long long ldw, void* out,
                                      long long ldo, int B, int K, int I, int splits, void* stream) {
    NXD_REQUIRE(I % 8 == 0 && ldo % 8 == 0 && ((uintptr_t)out & 15) == 0,
                "smallm_gate_up: I=%d and ldo must be multiples of 8, out 16-byte aligned", I);
    EpiArgs ep = {};
    ep.out = (bf16*)out;
    ep.ldo = ldo;
    ep.up_row0 = I;
    return run<kEpiSwiglu>(x, w, ep, B, I, 2 * I, K, ldx, ldw, splits, (cudaStream_t)stream);
}

// Tuning aid: the phase stamps of CTA 0 of the last launch (zeros unless built with -DNXD_SMALLM_TRACE).
NXD_API int nxd_smallm_trace(long long* out16) {
#ifdef NXD_SMALLM_TRACE
    NXD_CHECK_CUDA(cudaMemcpyFromSymbol(out16, g_smallm_trace, sizeof(long long) * 16));
#else
    for (int i = 0; i < 16; ++i) out16[i] = 0;
#endif
    return 0;
}
