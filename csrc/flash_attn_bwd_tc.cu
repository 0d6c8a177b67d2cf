// Flash attention BACKWARD on 5th-gen tensor cores (tcgen05 + TMEM + TMA), sm_100a, head dim 128 or 64.
//
// Same contract as nxd_flash_attn_bwd in flash_attn.cu (reference: the external NKI `flash_attn_bwd` kernel behind
// src/neuronx_distributed/kernels/flash_attn.py:19-28,103-148).
//
// One CTA owns 128 K/V rows of one (batch, kv head) and walks over the q heads of its GQA group and their query
// tiles of 64 rows.  Everything that is accumulated over the walk lives in TMEM (512 columns, one CTA per SM):
//     dK [128 x 128]   dV [128 x 128]   S^T [128 x 64]   dP^T [128 x 64]   dQ^T [128(d) x 64(q)]
// Per query tile i, five MMAs (A/B are smem descriptors over tiles that are each loaded / written ONCE):
//     S^T  = K   Q_i^T          A = K    (K-major)     B = Q_i  (K-major)
//     dP^T = V   dO_i^T         A = V    (K-major)     B = dO_i (K-major)
//     dV  += P^T  dO_i          A = P^T  (K-major)     B = dO_i (MN-major: the same smem tile read "transposed")
//     dK  += dS^T Q_i           A = dS^T (K-major)     B = Q_i  (MN-major)
//     dQ^T = K^T dS^T           A = K    (MN-major)    B = dS^T (MN-major)
// S^T / dP^T are produced transposed so that one thread owns one K/V row (TMEM lane): it reads its 64 scores,
// forms P^T = exp2(S^T c - L_q) and dS^T = P^T (dP^T - Delta_q) and writes one 128-byte swizzled smem row of each --
// exactly the K-major A operand of dV / dK and, read MN-major, the B operand of dQ^T.  dQ^T tiles are drained
// TMEM -> smem (transposed back) and added into an fp32 dQ accumulator with TMA reduce-add; dK (x scale) and dV
// leave through smem + TMA store.  The GQA reduction of dK/dV over the q-head group happens in TMEM for free.
//
// Head dim 64 (Llama-3.2-1B) runs the same pipeline with one 64-column group per tile.  Its dQ^T MMA keeps M = 128:
// the MN-major A descriptor then walks 64 rows past the K tile into the V tile, which fills TMEM lanes 64..127 with
// finite garbage that nobody reads -- cheaper than a second code path for the M = 64 TMEM layout.
//
// warp 0: TMA producer (K, V once; Q_i, dO_i, L_i, Delta_i through a 3-stage ring)   warp 1: TMEM alloc + MMA issue
// warps 2-9: softmax-backward math, epilogue.  warps 10-13: dQ^T drain (TMEM -> smem -> TMA reduce-add).  Two math warps share every TMEM lane quadrant (warp % 4) and
// split the 64 query columns of a tile between them: half the serial work per warp and twice the warps to hide the
// TMEM-load / exp2 / shared-store latencies behind (the 4-warp version ran one warp per SM sub-partition and was
// latency bound); in the epilogue one half stores dV while the other stores dK.
#include "attn_tc_common.cuh"
#include "gemm_tcgen05.cuh"
#include "ptx.cuh"

#include <type_traits>

namespace {
using namespace nxd_attn;

constexpr int kBN = 128;  // K/V rows per CTA
constexpr int kBM = 64;   // query rows per tile
constexpr int kThreadsBwd = 448;      // TMA warp + MMA warp + 8 math warps + 4 dQ drain warps
constexpr int kMathThreads = 256;
constexpr int kDrainThreads = 128;
constexpr int kFirstDrainWarp = 10;
constexpr int kQStages = 3;          // Q_i / dO_i / L_i / Delta_i ring: a tile is requested one full iteration before its MMAs

struct BwdParams {
    const float* lse_pad;    // [B*Hq, Sq_pad] log2-domain LSE, +inf in the padding
    const float* delta_pad;  // [B*Hq, Sq_pad]
    int B, Hq, Hkv, Sq, Sk, Sq_pad;
    int nsplit;              // CTAs that share one (kv block, kv head): each walks 1/nsplit of the (q head, q tile) list
    int pos_q[3], pos_do[3], pos_k[3], pos_v[3], pos_dk[3], pos_dv[3], pos_dq[3], pos_acc[3];
    float scale, scale_log2;
    DropParams drop;
};

template <int kD>
struct BwdSmem {
    static constexpr int kKOff = 0;                       // [kD / 64 groups][128 rows][64]   32 KB at D = 128
    static constexpr int kVOff = kKOff + kBN * kD * 2;    // 32 KB
    static constexpr int kQOff = kVOff + kBN * kD * 2;    // kQStages x [2 groups][64 rows][64]   48 KB
    static constexpr int kDOOff = kQOff + kQStages * kBM * kD * 2;   // 48 KB
    static constexpr int kPtOff = kDOOff + kQStages * kBM * kD * 2;  // [128 rows][64] bf16   16 KB
    static constexpr int kDStOff = kPtOff + kBN * kBM * 2;    // 16 KB
    static constexpr int kDQOff = kDStOff + kBN * kBM * 2;    // [4 groups][64 q rows][32 fp32]   32 KB
    static constexpr int kLOff = kDQOff + kBM * kD * 4;       // kQStages x (64 L + 64 Delta) floats   1.5 KB
    static constexpr int kBarOff = kLOff + kQStages * 2 * kBM * 4;
    static constexpr int kNumBars = 1 + 2 * kQStages + 6;
    // at least 120 KB so that only ONE CTA fits an SM: each CTA allocates all 512 TMEM columns
    static constexpr int kUsed = kBarOff + kNumBars * 8 + 16 + 1024;
    static constexpr int kBytes = kUsed < 120 * 1024 ? 120 * 1024 : kUsed;
    static constexpr int kStageBytes = kBM * kD * 2;          // one Q (or dO) tile
    static constexpr int kTxPerStage = 2 * kStageBytes + 2 * kBM * 4;
};
static_assert(BwdSmem<128>::kBytes <= 232448, "shared memory budget exceeded");

template <int kD, bool CAUSAL, bool DROP>
__global__ void __launch_bounds__(kThreadsBwd, 1)
fa_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_do,
                 const __grid_constant__ CUtensorMap tmap_k, const __grid_constant__ CUtensorMap tmap_v,
                 const __grid_constant__ CUtensorMap tmap_dk, const __grid_constant__ CUtensorMap tmap_dv,
                 const __grid_constant__ CUtensorMap tmap_dq, const __grid_constant__ CUtensorMap tmap_dk_acc,
                 const __grid_constant__ CUtensorMap tmap_dv_acc, const BwdParams p) {
    using L = BwdSmem<kD>;
    constexpr uint32_t kColDK = 0, kColDV = kD, kColST = 2 * kD, kColDPT = 2 * kD + 64, kColDQT = 2 * kD + 128;
    // bf16 P^T and dS^T as TENSOR-MEMORY A operands of the dV / dK MMAs: 128 lanes x 64 q = 32 packed columns each
    constexpr uint32_t kColPT = 2 * kD + 192, kColDST = kColPT + 32;
    static_assert(kColDST + 32 <= 512, "tensor memory budget");
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = align_smem_1024(smem_raw);
    uint8_t* sK = smem + L::kKOff;
    uint8_t* sV = smem + L::kVOff;
    uint8_t* sQ = smem + L::kQOff;
    uint8_t* sDO = smem + L::kDOOff;
    uint8_t* sPt = smem + L::kPtOff;
    uint8_t* sDSt = smem + L::kDStOff;
    uint8_t* sDQ = smem + L::kDQOff;
    float* sLD = reinterpret_cast<float*>(smem + L::kLOff);  // [stage][0: L, 1: Delta][64]
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBarOff);
    uint64_t* kv_full = bars;
    uint64_t* qdo_full = bars + 1;                  // [kQStages]
    uint64_t* qdo_empty = qdo_full + kQStages;      // [kQStages]
    uint64_t* st_full = qdo_empty + kQStages;
    uint64_t* st_empty = st_full + 1;
    uint64_t* pds_full = st_full + 2;
    uint64_t* pds_empty = st_full + 3;
    uint64_t* dq_full = st_full + 4;
    uint64_t* dq_empty = st_full + 5;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + L::kNumBars);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // grid = (kv heads x splits, batch, K/V blocks): the block index is the SLOWEST launch dimension, so with a causal mask
    // the CTAs are handed to the SMs longest first (K/V block 0 meets every query tile, the last one only two)
    const int nbk = blockIdx.z, hkv = blockIdx.x / p.nsplit, split = blockIdx.x % p.nsplit, b = blockIdx.y;
    const int group = p.Hq / p.Hkv;
    const int n0 = nbk * kBN;
    const int offset = p.Sk - p.Sq;
    int mb_start = 0;
    if (CAUSAL) mb_start = max(0, (n0 - offset) / kBM);
    const int mb_end = (p.Sq + kBM - 1) / kBM;
    const int tiles_per_head = max(mb_end - mb_start, 0);
    // (q head of the group, q tile) pairs of this K/V block, flattened; this CTA owns the slice [it0, it0 + total)
    const int all_tiles = tiles_per_head * group;
    const int it0 = (int)((long long)all_tiles * split / p.nsplit);
    const int total = (int)((long long)all_tiles * (split + 1) / p.nsplit) - it0;
    if (p.nsplit > 1 && total == 0) return;   // nothing to add into the shared dK / dV accumulators

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tensormap(&tmap_q);
        ptx::prefetch_tensormap(&tmap_do);
        ptx::prefetch_tensormap(&tmap_k);
        ptx::prefetch_tensormap(&tmap_v);
        ptx::prefetch_tensormap(&tmap_dq);
        ptx::mbar_init(kv_full, 1);
        for (int i = 0; i < kQStages; ++i) {
            ptx::mbar_init(&qdo_full[i], 1);
            ptx::mbar_init(&qdo_empty[i], 1);
        }
        ptx::mbar_init(st_full, 1);
        ptx::mbar_init(st_empty, kMathThreads);
        ptx::mbar_init(pds_full, kMathThreads);
        ptx::mbar_init(pds_empty, 1);
        ptx::mbar_init(dq_full, 1);
        ptx::mbar_init(dq_empty, kDrainThreads);
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<1>(tmem_ptr_smem, 512);
    ptx::tcgen05_fence_before();
    __syncthreads();
    ptx::tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    if (warp == 0) {
        // ============================ TMA producer ============================
        if (lane == 0) {
            int c[4];
            ptx::mbar_arrive_expect_tx(kv_full, 2 * kBN * kD * 2);
#pragma unroll
            for (int g = 0; g < kD / 64; ++g) {
                c[0] = g * 64;
                coords(p.pos_k, n0, hkv, b, c);
                tma_load_4d(&tmap_k, kv_full, sK + g * (kBN * 128), c[0], c[1], c[2], c[3]);
                coords(p.pos_v, n0, hkv, b, c);
                tma_load_4d(&tmap_v, kv_full, sV + g * (kBN * 128), c[0], c[1], c[2], c[3]);
            }
            int th = it0 % tiles_per_head, hq = hkv * group + it0 / tiles_per_head;   // (q tile, q head) walk
            for (int it = 0; it < total; ++it) {
                const int st = it % kQStages;
                const uint32_t ph = (uint32_t)(it / kQStages) & 1u;
                const int m0 = (mb_start + th) * kBM;
                ptx::mbar_wait(&qdo_empty[st], ph ^ 1u);
                ptx::mbar_arrive_expect_tx(&qdo_full[st], L::kTxPerStage);
#pragma unroll
                for (int g = 0; g < kD / 64; ++g) {
                    c[0] = g * 64;
                    coords(p.pos_q, m0, hq, b, c);
                    tma_load_4d(&tmap_q, &qdo_full[st], sQ + st * L::kStageBytes + g * (kBM * 128), c[0], c[1], c[2], c[3]);
                    coords(p.pos_do, m0, hq, b, c);
                    tma_load_4d(&tmap_do, &qdo_full[st], sDO + st * L::kStageBytes + g * (kBM * 128), c[0], c[1], c[2],
                                c[3]);
                }
                const long long row = ((long long)b * p.Hq + hq) * p.Sq_pad + m0;
                bulk_load_1d(sLD + st * 2 * kBM, p.lse_pad + row, kBM * 4, &qdo_full[st]);
                bulk_load_1d(sLD + st * 2 * kBM + kBM, p.delta_pad + row, kBM * 4, &qdo_full[st]);
                if (++th == tiles_per_head) { th = 0; ++hq; }
            }
        }
    } else if (warp == 1) {
        // ============================ MMA issuer ============================
        if (lane == 0 && total > 0) {
            constexpr uint32_t idesc_st = ptx::make_idesc_bf16(kBN, kBM, 0, 0);   // S^T, dP^T: 128 x 64, K-major A/B
            constexpr uint32_t idesc_kv = ptx::make_idesc_bf16(kBN, kD, 0, 1);    // dV, dK: 128 x 128, B MN-major
            constexpr uint32_t idesc_dq = ptx::make_idesc_bf16(128, kBM, 1, 1);   // dQ^T: 128 x 64, A and B MN-major
            const uint32_t k_addr = ptx::smem_u32(sK), v_addr = ptx::smem_u32(sV), q_addr = ptx::smem_u32(sQ),
                           do_addr = ptx::smem_u32(sDO), pt_addr = ptx::smem_u32(sPt), dst_addr = ptx::smem_u32(sDSt);
            auto issue_scores = [&](int it) {
                const int st = it % kQStages;
                ptx::mbar_wait(&qdo_full[st], (uint32_t)(it / kQStages) & 1u);
                ptx::mbar_wait(st_empty, ((uint32_t)it & 1u) ^ 1u);  // S^T/dP^T of tile it-1 are in registers
                ptx::tcgen05_fence_after();
                const uint32_t qb = q_addr + st * L::kStageBytes, dob = do_addr + st * L::kStageBytes;
#pragma unroll
                for (int kk = 0; kk < kD / 16; ++kk) {
                    const uint32_t ko = (kk >> 2) * (kBN * 128) + (kk & 3) * 32;   // K-major A: [128][64] groups
                    const uint32_t qo = (kk >> 2) * (kBM * 128) + (kk & 3) * 32;   // K-major B: [64][64] groups
                    ptx::umma_bf16<1>(tmem_base + kColST, ptx::make_smem_desc(k_addr + ko, 16, 1024),
                                      ptx::make_smem_desc(qb + qo, 16, 1024), idesc_st, kk != 0 ? 1u : 0u);
                }
#pragma unroll
                for (int kk = 0; kk < kD / 16; ++kk) {
                    const uint32_t ko = (kk >> 2) * (kBN * 128) + (kk & 3) * 32;
                    const uint32_t qo = (kk >> 2) * (kBM * 128) + (kk & 3) * 32;
                    ptx::umma_bf16<1>(tmem_base + kColDPT, ptx::make_smem_desc(v_addr + ko, 16, 1024),
                                      ptx::make_smem_desc(dob + qo, 16, 1024), idesc_st, kk != 0 ? 1u : 0u);
                }
                ptx::umma_commit(st_full);
            };
            ptx::mbar_wait(kv_full, 0);
            issue_scores(0);
            for (int it = 0; it < total; ++it) {
                if (it + 1 < total) issue_scores(it + 1);
                const int st = it % kQStages;
                const uint32_t qb = q_addr + st * L::kStageBytes, dob = do_addr + st * L::kStageBytes;
                ptx::mbar_wait(pds_full, (uint32_t)it & 1u);
                ptx::tcgen05_fence_after();
                // dV += P^T dO ; dK += dS^T Q      (contraction over the 64 q rows: 4 MMAs of K = 16 each)
#pragma unroll
                for (int kk = 0; kk < kBM / 16; ++kk) {
                    ptx::umma_bf16_ts(tmem_base + kColDV, tmem_base + kColPT + kk * 8,
                                      ptx::make_smem_desc(dob + kk * 2048, kBM * 128, 1024), idesc_kv,
                                      (it | kk) != 0 ? 1u : 0u);
                }
#pragma unroll
                for (int kk = 0; kk < kBM / 16; ++kk) {
                    ptx::umma_bf16_ts(tmem_base + kColDK, tmem_base + kColDST + kk * 8,
                                      ptx::make_smem_desc(qb + kk * 2048, kBM * 128, 1024), idesc_kv,
                                      (it | kk) != 0 ? 1u : 0u);
                }
                ptx::umma_commit(&qdo_empty[st]);   // Q_i / dO_i no longer needed
                // dQ^T = K^T dS^T   (contraction over the 128 kv rows: 8 MMAs); wait until dQ^T(it-1) was drained
                ptx::mbar_wait(dq_empty, ((uint32_t)it & 1u) ^ 1u);
                ptx::tcgen05_fence_after();
#pragma unroll
                for (int kk = 0; kk < kBN / 16; ++kk) {
                    ptx::umma_bf16<1>(tmem_base + kColDQT, ptx::make_smem_desc(k_addr + kk * 2048, kBN * 128, 1024),
                                      ptx::make_smem_desc(dst_addr + kk * 2048, kBN * 128, 1024), idesc_dq,
                                      kk != 0 ? 1u : 0u);
                }
                ptx::umma_commit(pds_empty);
                ptx::umma_commit(dq_full);
            }
        }
    } else if (warp >= kFirstDrainWarp) {
        // ============================ dQ drain ============================
        // dQ^T(it) [d = TMEM lane][64 q] -> this warp's staging rows (no transpose: the fp32 dQ accumulator in global
        // memory is d-major) -> TMA reduce-add, 32 q columns per box.  Its own warps: the softmax warps never wait for
        // the staging fence, and the tensor core gets its dQ^T columns back as soon as they are in registers.  No
        // CTA-wide barrier: staging buffers, bulk groups and the waits on them are private to the warp.
        const int q4 = warp & 3;                 // TMEM lane quadrant this warp may access
        const uint32_t lane_off = (uint32_t)(q4 * 32) << 16;
        const bool has_rows = q4 * 32 < kD;      // head dim 64: TMEM lanes 64..127 of dQ^T are padding
        uint8_t* stage = sDQ + q4 * 8192;        // two 32 x 32 fp32 boxes (4 KB each) of this warp
        int th = it0 % tiles_per_head, hq = hkv * group + it0 / tiles_per_head;
        for (int it = 0; it < total; ++it) {
            const int m0 = (mb_start + th) * kBM;
            ptx::mbar_wait(dq_full, (uint32_t)it & 1u);
            ptx::tcgen05_fence_after();
            uint32_t v[64];
            if (has_rows) {
                ptx::tmem_ld_32x32b_x32(tmem_base + kColDQT + lane_off, v);
                ptx::tmem_ld_32x32b_x32(tmem_base + kColDQT + lane_off + 32, v + 32);
                ptx::tmem_ld_wait();
            }
            ptx::tcgen05_fence_before();
            ptx::mbar_arrive(dq_empty);
            if (has_rows) {
                if (lane == 0) ptx::bulk_wait_read<0>();     // the previous reduce-adds have read the staging rows
                __syncwarp();
#pragma unroll
                for (int h2 = 0; h2 < 2; ++h2) {
                    uint8_t* dst = stage + h2 * 4096 + lane * 128;
#pragma unroll
                    for (int c = 0; c < 8; ++c)
                        st_shared_v4(dst + ((c ^ (lane & 7)) << 4),
                                     __float_as_uint(__uint_as_float(v[h2 * 32 + 4 * c]) * p.scale),
                                     __float_as_uint(__uint_as_float(v[h2 * 32 + 4 * c + 1]) * p.scale),
                                     __float_as_uint(__uint_as_float(v[h2 * 32 + 4 * c + 2]) * p.scale),
                                     __float_as_uint(__uint_as_float(v[h2 * 32 + 4 * c + 3]) * p.scale));
                }
                ptx::fence_proxy_async();
                __syncwarp();
                if (lane == 0) {
                    int c[4];
                    coords(p.pos_dq, q4 * 32, hq, b, c);
                    tma_reduce_add_4d(&tmap_dq, stage, m0, c[1], c[2], c[3]);
                    tma_reduce_add_4d(&tmap_dq, stage + 4096, m0 + 32, c[1], c[2], c[3]);
                    ptx::bulk_commit();
                }
            }
            if (++th == tiles_per_head) { th = 0; ++hq; }
        }
        if (lane == 0) ptx::bulk_wait<0>();
    } else {
        // ============================ softmax backward / epilogue ============================
        const int q4 = warp & 3;
        const int row = q4 * 32 + lane;         // K/V row inside the tile == TMEM lane (and d index for dQ^T)
        const int krow = n0 + row;
        const uint32_t lane_off = (uint32_t)(q4 * 32) << 16;
        const int half = (warp - 2) >> 2;        // which 32 of the 64 query columns this warp handles
        const int c0h = half * 32;
        // P^T = exp2(S^T c - L_q), dS^T = P^T (dP^T - Delta_q) for this thread's K/V row and 32 query columns.
        // MASK is instantiated only for the tiles that touch the causal diagonal or the end of the K/V sequence.
        auto softmax_bwd = [&](auto mask_tag, const uint32_t* s, const uint32_t* dp, const float* Ls, const float* Ds,
                               int m0, int hq, uint32_t* pk, uint32_t* dk_) {
            constexpr bool MASK = decltype(mask_tag)::value;
            // masked <=> krow >= Sk, or (causal) krow > m0 + c0h + c + offset  <=>  c < first_ok
            int first_ok = 0;
            if (MASK) {
                first_ok = CAUSAL ? krow - offset - m0 - c0h : 0;
                if (krow >= p.Sk) first_ok = 64;
            }
#pragma unroll
            for (int c = 0; c < 32; c += 4) {
                const float4 l4 = ld_shared_f4(Ls + c0h + c);
                const float4 d4 = ld_shared_f4(Ds + c0h + c);
                const float lv[4] = {l4.x, l4.y, l4.z, l4.w}, dv[4] = {d4.x, d4.y, d4.z, d4.w};
                float pv[4], dsv[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    pv[e] = ex2_approx(fmaf(__uint_as_float(s[c + e]), p.scale_log2, -lv[e]));
                    if (MASK && c + e < first_ok) pv[e] = 0.f;
                    float dpe = __uint_as_float(dp[c + e]);
                    if (DROP) {
                        // forward multiplied P by keep / (1 - p) before P V: dV takes the dropped P, and dP (= dO V^T
                        // w.r.t. the dropped P) is masked the same way before it meets Delta
                        const uint32_t hs = drop_head_seed(p.drop.seed, (uint32_t)(b * p.Hq + hq));
                        const float km = drop_keep(hs, (uint32_t)(m0 + c0h + c + e), (uint32_t)krow, p.drop.threshold)
                                             ? p.drop.inv_keep : 0.f;
                        dpe *= km;
                        dsv[e] = pv[e] * (dpe - dv[e]);
                        pv[e] *= km;
                    } else {
                        dsv[e] = pv[e] * (dpe - dv[e]);
                    }
                }
                pk[c >> 1] = pack_bf16x2(pv[0], pv[1]);
                pk[(c >> 1) + 1] = pack_bf16x2(pv[2], pv[3]);
                dk_[c >> 1] = pack_bf16x2(dsv[0], dsv[1]);
                dk_[(c >> 1) + 1] = pack_bf16x2(dsv[2], dsv[3]);
            }
        };

        int th = it0 % tiles_per_head, hq = hkv * group + it0 / tiles_per_head;
        for (int it = 0; it < total; ++it) {
            const int st = it % kQStages;
            const int m0 = (mb_start + th) * kBM;
            ptx::mbar_wait(st_full, (uint32_t)it & 1u);
            ptx::tcgen05_fence_after();
            uint32_t s[32], dp[32];
            ptx::tmem_ld_32x32b_x32(tmem_base + kColST + lane_off + c0h, s);
            ptx::tmem_ld_32x32b_x32(tmem_base + kColDPT + lane_off + c0h, dp);
            ptx::tmem_ld_wait();
            ptx::tcgen05_fence_before();
            ptx::mbar_arrive(st_empty);
            // L / Delta of this tile were delivered with Q_i / dO_i (the MMA warp already saw qdo_full complete;
            // wait on it here too so that the bulk-copied bytes are visible to this thread)
            ptx::mbar_wait(&qdo_full[st], (uint32_t)(it / kQStages) & 1u);
            const float* Ls = sLD + st * 2 * kBM;
            const float* Ds = Ls + kBM;
            // warp-uniform: does any (K/V row of this warp, query column of this warp) need masking?
            const bool need_mask = (n0 + kBN > p.Sk) || (CAUSAL && (n0 + q4 * 32 + 31 > m0 + c0h + offset));
            uint32_t pk[16], dk_[16];
            if (need_mask) softmax_bwd(std::true_type{}, s, dp, Ls, Ds, m0, hq, pk, dk_);
            else softmax_bwd(std::false_type{}, s, dp, Ls, Ds, m0, hq, pk, dk_);
            if (it > 0) ptx::mbar_wait(pds_empty, ((uint32_t)(it - 1)) & 1u);  // MMAs of tile it-1 finished with sPt/sDSt
            // P^T and dS^T go to tensor memory (A operands of dV / dK: this thread's row, its 32 q columns = 16 packed
            // words each); only dS^T is also needed in shared memory, as the MN-major B operand of dQ^T
            tmem_st_32x32b_x16(tmem_base + kColPT + lane_off + half * 16, pk);
            tmem_st_32x32b_x16(tmem_base + kColDST + lane_off + half * 16, dk_);
            uint8_t* drow = sDSt + row * 128;
#pragma unroll
            for (int c = 0; c < 4; ++c) {                // this warp's 4 of the row's 8 sixteen-byte chunks
                const int sw = ((half * 4 + c) ^ (row & 7)) << 4;
                st_shared_v4(drow + sw, dk_[4 * c], dk_[4 * c + 1], dk_[4 * c + 2], dk_[4 * c + 3]);
            }
            tmem_st_wait();
            ptx::fence_proxy_async();
            ptx::tcgen05_fence_before();
            ptx::mbar_arrive(pds_full);
            if (++th == tiles_per_head) { th = 0; ++hq; }
        }
        if (total > 0) {
            // the commit behind pds_empty(total-1) follows the last dQ^T MMA and covers every earlier MMA: dK / dV are final
            ptx::mbar_wait(pds_empty, ((uint32_t)(total - 1)) & 1u);
            ptx::tcgen05_fence_after();
        }
        if (total == 0) ptx::mbar_wait(kv_full, 0);  // nothing consumed K/V: let their TMA land before the tiles are reused
        named_bar_sync(1, kMathThreads);              // every math warp is past its last use of the K / V / Q / dO tiles
        if (p.nsplit > 1) {
            // ---- split epilogue: this CTA holds PARTIAL dV / dK -> fp32 -> smem -> TMA reduce-add into the accumulators
            // (32 rows x 32 columns per warp and step, two 4 KB staging buffers per warp in the dead K / V / Q / dO
            // tiles: 64 KB contiguous from sK for either head dim); warps 2-5 add dV, warps 6-9 add dK
            uint8_t* stg = sK + (half * 4 + q4) * 8192;
            int buf = 0;
            {
                const int which = half;
                const uint32_t col = which == 0 ? kColDV : kColDK;
                const float mul = which == 0 ? 1.f : p.scale;
#pragma unroll 1
                for (int g = 0; g < kD / 32; ++g) {
                    uint32_t o[32];
                    ptx::tmem_ld_32x32b_x32(tmem_base + col + lane_off + g * 32, o);
                    ptx::tmem_ld_wait();
                    if (lane == 0) ptx::bulk_wait_read<1>();
                    __syncwarp();
                    uint8_t* dst = stg + buf * 4096 + lane * 128;
#pragma unroll
                    for (int c = 0; c < 8; ++c)
                        *reinterpret_cast<float4*>(dst + ((c ^ (lane & 7)) << 4)) =
                            make_float4(__uint_as_float(o[4 * c]) * mul, __uint_as_float(o[4 * c + 1]) * mul,
                                        __uint_as_float(o[4 * c + 2]) * mul, __uint_as_float(o[4 * c + 3]) * mul);
                    ptx::fence_proxy_async();
                    __syncwarp();
                    if (lane == 0) {
                        if (n0 + q4 * 32 < p.Sk) {
                            int c[4];
                            c[0] = g * 32;
                            coords(p.pos_acc, n0 + q4 * 32, hkv, b, c);
                            tma_reduce_add_4d(which == 0 ? &tmap_dv_acc : &tmap_dk_acc, stg + buf * 4096, c[0], c[1], c[2],
                                              c[3]);
                        }
                        ptx::bulk_commit();
                    }
                    buf ^= 1;
                }
            }
        } else {
            // ---- epilogue: dV (warps 2-5), dK x scale (warps 6-9) -> bf16 -> smem (the dead V / K tiles) -> TMA store ----
            {
                const int which = half;
                uint8_t* stage = which == 0 ? sV : sK;
                const uint32_t col = which == 0 ? kColDV : kColDK;
                const float mul = which == 0 ? 1.f : p.scale;
    #pragma unroll 1
                for (int g = 0; g < kD / 64; ++g) {
                    uint32_t o[64];
                    if (total > 0) {
                        ptx::tmem_ld_32x32b_x32(tmem_base + col + lane_off + g * 64, o);
                        ptx::tmem_ld_32x32b_x32(tmem_base + col + lane_off + g * 64 + 32, o + 32);
                        ptx::tmem_ld_wait();
                    } else {
    #pragma unroll
                        for (int e = 0; e < 64; ++e) o[e] = 0u;
                    }
                    uint8_t* dst = stage + g * (kBN * 128) + row * 128;
    #pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        uint32_t w[4];
    #pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            bf162 t = __floats2bfloat162_rn(__uint_as_float(o[8 * c + 2 * e]) * mul,
                                                            __uint_as_float(o[8 * c + 2 * e + 1]) * mul);
                            w[e] = *reinterpret_cast<uint32_t*>(&t);
                        }
                        *reinterpret_cast<uint4*>(dst + ((c ^ (row & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
                    }
                    ptx::fence_proxy_async();
                    __syncwarp();
                    if (lane == 0 && n0 + q4 * 32 < p.Sk) {
                        int c[4];
                        c[0] = g * 64;
                        coords(which == 0 ? p.pos_dv : p.pos_dk, n0 + q4 * 32, hkv, b, c);
                        tma_store_4d(which == 0 ? &tmap_dv : &tmap_dk, stage + g * (kBN * 128) + q4 * 32 * 128, c[0], c[1],
                                     c[2], c[3]);
                        ptx::bulk_commit();
                    }
                }
            }
        }
        if (lane == 0) ptx::bulk_wait<0>();
    }
    ptx::tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<1>(tmem_base, 512);
}

// Delta = rowsum(O * dO) and a padded copy of the LSE ([B*Hq, Sq_pad], +inf / 0 in the padding) so that the main
// kernel can fetch 64-row slices with 256-byte bulk copies and get masking of out-of-range query rows for free.
// kD / 8 lanes own a row (16-byte vectors), a warp has 4 x (32 / lanes-per-row) rows in flight: the kernel streams
// 2 x [B, Hq, Sq, D] bf16 once and is bound by HBM, not by load latency.
template <int kD>
__global__ void __launch_bounds__(256) fa_bwd_tc_prep_kernel(const bf16* __restrict__ o, const bf16* __restrict__ dout,
                                                             const float* __restrict__ lse, float* __restrict__ lse_pad,
                                                             float* __restrict__ delta_pad, long long o_sb,
                                                             long long o_sh, long long o_ss, long long do_sb,
                                                             long long do_sh, long long do_ss, int B, int Hq, int Sq,
                                                             int Sq_pad) {
    constexpr int L = kD / 8, R = 32 / L, U = 4;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int sub = lane / L, l = lane % L;
    const long long total = (long long)B * Hq * Sq_pad;
    const long long r0 = ((long long)blockIdx.x * 8 + warp) * (R * U) + sub;
    bf16x8 a[U], g[U];
    bool valid[U];
    long long lse_idx[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const long long r = r0 + u * R;
        const int i = (int)(r % Sq_pad);
        const int h = (int)((r / Sq_pad) % Hq);
        const long long b = r / ((long long)Sq_pad * Hq);
        valid[u] = r < total && i < Sq;
        lse_idx[u] = (b * Hq + h) * Sq + i;
        if (valid[u]) {
            a[u] = reinterpret_cast<const bf16x8*>(o + b * o_sb + h * o_sh + (long long)i * o_ss)[l];
            g[u] = reinterpret_cast<const bf16x8*>(dout + b * do_sb + h * do_sh + (long long)i * do_ss)[l];
        }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
        float acc = 0.f;
        if (valid[u]) {
            float fa[8], fg[8];
            unpack8(a[u], fa);
            unpack8(g[u], fg);
#pragma unroll
            for (int e = 0; e < 8; ++e) acc = fmaf(fa[e], fg[e], acc);
        }
#pragma unroll
        for (int off = L / 2; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
        const long long r = r0 + u * R;
        if (l == 0 && r < total) {
            delta_pad[r] = valid[u] ? acc : 0.f;
            lse_pad[r] = valid[u] ? lse[lse_idx[u]] : INFINITY;
        }
    }
}

// fp32 [B, Hq, Sq, D] accumulator -> strided bf16 dq
__global__ void fa_bwd_tc_dq_convert_kernel(const float* __restrict__ acc, bf16* __restrict__ dq, long long sb,
                                            long long sh, long long ss, int B, int Hq, int Sq, int kD) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // one thread = 8 elements
    const long long total = (long long)B * Hq * Sq * (kD / 8);
    if (idx >= total) return;
    const int c = (int)(idx % (kD / 8));
    const long long r = idx / (kD / 8);
    const int i = (int)(r % Sq);
    const int h = (int)((r / Sq) % Hq);
    const int b = (int)(r / ((long long)Sq * Hq));
    const float4 a0 = reinterpret_cast<const float4*>(acc + r * kD)[2 * c];
    const float4 a1 = reinterpret_cast<const float4*>(acc + r * kD)[2 * c + 1];
    float f[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
    reinterpret_cast<bf16x8*>(dq + b * sb + h * sh + (long long)i * ss)[c] = pack8(f);
}

// d-major fp32 accumulator [B, Hq, D, Sq_pad] -> strided bf16 dq [.., Sq, D]: 32 query rows per block through a padded
// shared tile.  Reads: 128-byte rows along q, all D / 8 of a thread in flight; writes: two adjacent lanes write the two
// 16-byte halves of a 32-byte sector, 16 query rows per warp (tile reads 2-way conflicted at most).
template <int kD>
__global__ void __launch_bounds__(256)
fa_bwd_tc_dq_transpose_kernel(const float* __restrict__ acc, bf16* __restrict__ dq, long long sb, long long sh,
                              long long ss, int Hq, int Sq, int Sq_pad) {
    __shared__ float tile[kD][33];
    const int q0 = blockIdx.x * 32, h = blockIdx.y, b = blockIdx.z;
    const float* src = acc + ((long long)b * Hq + h) * kD * Sq_pad;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    float v[kD / 8];
    const bool in = q0 + lane < Sq_pad;
#pragma unroll
    for (int j = 0; j < kD / 8; ++j) v[j] = in ? src[(long long)(w + 8 * j) * Sq_pad + q0 + lane] : 0.f;
#pragma unroll
    for (int j = 0; j < kD / 8; ++j) tile[w + 8 * j][lane] = v[j];
    __syncthreads();
    const int r = (w & 1) * 16 + (lane >> 1);
    if (q0 + r >= Sq) return;
    bf16* dst = dq + b * sb + h * sh + (long long)(q0 + r) * ss;
#pragma unroll
    for (int c = (w >> 1) * 2 + (lane & 1); c < kD / 8; c += 8) {
        float f[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = tile[c * 8 + e][r];
        reinterpret_cast<bf16x8*>(dst)[c] = pack8(f);
    }
}

template <int kD, bool CAUSAL, bool DROP>
int launch_bwd_tc(dim3 grid, cudaStream_t st, const CUtensorMap& tq, const CUtensorMap& tdo, const CUtensorMap& tk,
                  const CUtensorMap& tv, const CUtensorMap& tdk, const CUtensorMap& tdv, const CUtensorMap& tdq,
                  const CUtensorMap& tdka, const CUtensorMap& tdva, const BwdParams& p) {
    static cudaError_t once = cudaFuncSetAttribute(fa_bwd_tc_kernel<kD, CAUSAL, DROP>,
                                                   cudaFuncAttributeMaxDynamicSharedMemorySize, BwdSmem<kD>::kBytes);
    NXD_CHECK_CUDA(once);
    fa_bwd_tc_kernel<kD, CAUSAL, DROP><<<grid, kThreadsBwd, BwdSmem<kD>::kBytes, st>>>(tq, tdo, tk, tv, tdk, tdv, tdq, tdka, tdva, p);
    NXD_LAUNCH_CHECK();
    return 0;
}

}  // namespace

// Number of CTAs that should share one (K/V block, kv head) so that the grid fills the GPU: 1 = every CTA owns its
// dK / dV tile (no accumulators needed), > 1 = pass `dkv_accum` to nxd_flash_attn_bwd_tc.
NXD_API int nxd_flash_attn_bwd_tc_num_splits(int B, int Hq, int Hkv, int Sq, int Sk) {
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const long long ctas = (long long)((Sk + kBN - 1) / kBN) * Hkv * B;
    if (ctas >= 2LL * sms) return 1;
    const int group = Hq / (Hkv > 0 ? Hkv : 1);
    const long long iters = (long long)((Sq + kBM - 1) / kBM) * group;     // longest (q head, q tile) list of a CTA
    long long n = (4LL * sms + ctas - 1) / ctas;
    if (n > iters / 8) n = iters / 8;                                      // keep >= 8 tiles per CTA (prologue / epilogue)
    if (n > 16) n = 16;
    return n < 1 ? 1 : (int)n;
}

// Head dim 128 or 64.  scratch: lse_pad and delta_pad are [B*Hq*Sq_pad] floats (Sq_pad = Sq rounded up to 64);
// dq_accum is a ZERO-FILLED [B, Hq, D, Sq_pad] fp32 buffer (d-major).  dkv_accum: NULL, or a ZERO-FILLED [2, B, Hkv, Sk, D] fp32
// buffer -- then nxd_flash_attn_bwd_tc_num_splits() CTAs share every K/V block (small batch x heads: tensor
// parallel shards, long sequences) and add their partial dK / dV with TMA reduce-add.  strides as nxd_flash_attn_bwd
// (24 values).
static int bwd_tc_impl(const void* q, const void* k, const void* v, const void* o, const void* dout, const float* lse,
                       float* lse_pad, float* delta_pad, float* dq_accum, float* dkv_accum, void* dq, void* dk, void* dv,
                       const long long* strides, int B, int Hq, int Hkv, int Sq, int Sk, int D, float scale, int causal,
                       float dropout_p, unsigned long long seed, void* stream) {
    NXD_REQUIRE(D == 128 || D == 64, "flash_attn_bwd_tc: head dim %d not supported (64 or 128)", D);
    NXD_REQUIRE(B > 0 && Hq > 0 && Hkv > 0 && Sq > 0 && Sk > 0 && Hq % Hkv == 0, "flash_attn_bwd_tc: bad problem");
    for (int i = 0; i < 24; ++i)
        NXD_REQUIRE(strides[i] % 8 == 0, "flash_attn_bwd_tc: stride %d (=%lld) must be a multiple of 8 elements", i,
                    strides[i]);
    cudaStream_t st = (cudaStream_t)stream;
    BwdParams p{};
    p.B = B; p.Hq = Hq; p.Hkv = Hkv; p.Sq = Sq; p.Sk = Sk;
    p.Sq_pad = (Sq + kBM - 1) / kBM * kBM;
    p.lse_pad = lse_pad; p.delta_pad = delta_pad;
    p.scale = scale; p.scale_log2 = scale * 1.4426950408889634f;
    NXD_REQUIRE(dropout_p >= 0.f && dropout_p < 1.f, "flash_attn_bwd_tc: dropout_p %f not in [0, 1)", dropout_p);
    p.drop.threshold = dropout_p > 0.f ? (uint32_t)fmin(4294967295.0, (double)dropout_p * 4294967296.0) : 0u;
    p.drop.inv_keep = 1.f / (1.f - dropout_p);
    p.drop.seed = (uint32_t)(seed ^ (seed >> 32));
    p.nsplit = dkv_accum ? nxd_flash_attn_bwd_tc_num_splits(B, Hq, Hkv, Sq, Sk) : 1;
    {
        const long long rows = (long long)B * Hq * p.Sq_pad;
        if (D == 128) {
            fa_bwd_tc_prep_kernel<128><<<(unsigned)((rows + 63) / 64), 256, 0, st>>>(
                (const bf16*)o, (const bf16*)dout, lse, lse_pad, delta_pad, strides[9], strides[10], strides[11],
                strides[12], strides[13], strides[14], B, Hq, Sq, p.Sq_pad);
        } else {
            fa_bwd_tc_prep_kernel<64><<<(unsigned)((rows + 127) / 128), 256, 0, st>>>(
                (const bf16*)o, (const bf16*)dout, lse, lse_pad, delta_pad, strides[9], strides[10], strides[11],
                strides[12], strides[13], strides[14], B, Hq, Sq, p.Sq_pad);
        }
        NXD_LAUNCH_CHECK();
    }
    CUtensorMap tq, tdo, tk, tv, tdk, tdv, tdq, tdka, tdva;
    int rc;
    if ((rc = make_tmap_bhsd(&tq, q, D, Sq, Hq, B, strides[2], strides[1], strides[0], kBM, p.pos_q))) return rc;
    if ((rc = make_tmap_bhsd(&tk, k, D, Sk, Hkv, B, strides[5], strides[4], strides[3], kBN, p.pos_k))) return rc;
    if ((rc = make_tmap_bhsd(&tv, v, D, Sk, Hkv, B, strides[8], strides[7], strides[6], kBN, p.pos_v))) return rc;
    if ((rc = make_tmap_bhsd(&tdo, dout, D, Sq, Hq, B, strides[14], strides[13], strides[12], kBM, p.pos_do))) return rc;
    if ((rc = make_tmap_bhsd(&tdk, dk, D, Sk, Hkv, B, strides[20], strides[19], strides[18], 32, p.pos_dk))) return rc;
    if ((rc = make_tmap_bhsd(&tdv, dv, D, Sk, Hkv, B, strides[23], strides[22], strides[21], 32, p.pos_dv))) return rc;
    // dQ accumulator: d-major [B, Hq, D, Sq_pad] fp32, so that a warp's 32 (d) x 32 (q) block of dQ^T leaves TMEM
    // and shared memory without a transpose; box = 32 d rows x 32 q columns (128 bytes)
    if ((rc = make_tmap_bhsd(&tdq, dq_accum, p.Sq_pad, D, Hq, B, p.Sq_pad, (long long)D * p.Sq_pad,
                             (long long)Hq * D * p.Sq_pad, 32, p.pos_dq, 4)))
        return rc;
    const long long kv_elems = (long long)B * Hkv * Sk * D;
    float* dk_acc = p.nsplit > 1 ? dkv_accum : dq_accum;   // unused maps still have to be valid
    float* dv_acc = p.nsplit > 1 ? dkv_accum + kv_elems : dq_accum;
    if (p.nsplit > 1) {
        if ((rc = make_tmap_bhsd(&tdka, dk_acc, D, Sk, Hkv, B, D, (long long)Sk * D, (long long)Hkv * Sk * D, 32, p.pos_acc, 4)))
            return rc;
        if ((rc = make_tmap_bhsd(&tdva, dv_acc, D, Sk, Hkv, B, D, (long long)Sk * D, (long long)Hkv * Sk * D, 32, p.pos_acc, 4)))
            return rc;
    } else {
        tdka = tdq;
        tdva = tdq;
    }
    dim3 grid(Hkv * p.nsplit, B, (Sk + kBN - 1) / kBN);
#define NXD_BWD_ARGS grid, st, tq, tdo, tk, tv, tdk, tdv, tdq, tdka, tdva, p
    const bool drop = p.drop.threshold != 0;
    if (D == 128) {
        if (drop) rc = causal ? launch_bwd_tc<128, true, true>(NXD_BWD_ARGS) : launch_bwd_tc<128, false, true>(NXD_BWD_ARGS);
        else rc = causal ? launch_bwd_tc<128, true, false>(NXD_BWD_ARGS) : launch_bwd_tc<128, false, false>(NXD_BWD_ARGS);
    } else {
        if (drop) rc = causal ? launch_bwd_tc<64, true, true>(NXD_BWD_ARGS) : launch_bwd_tc<64, false, true>(NXD_BWD_ARGS);
        else rc = causal ? launch_bwd_tc<64, true, false>(NXD_BWD_ARGS) : launch_bwd_tc<64, false, false>(NXD_BWD_ARGS);
    }
#undef NXD_BWD_ARGS
    if (rc) return rc;
    if (p.nsplit > 1) {
        const long long n = kv_elems / 8;
        fa_bwd_tc_dq_convert_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(dk_acc, (bf16*)dk, strides[18], strides[19],
                                                                                strides[20], B, Hkv, Sk, D);
        fa_bwd_tc_dq_convert_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(dv_acc, (bf16*)dv, strides[21], strides[22],
                                                                                strides[23], B, Hkv, Sk, D);
        NXD_LAUNCH_CHECK_N(2);
    }
    {
        dim3 g((unsigned)((Sq + 31) / 32), (unsigned)Hq, (unsigned)B);
        if (D == 128)
            fa_bwd_tc_dq_transpose_kernel<128><<<g, 256, 0, st>>>(dq_accum, (bf16*)dq, strides[15], strides[16],
                                                                  strides[17], Hq, Sq, p.Sq_pad);
        else
            fa_bwd_tc_dq_transpose_kernel<64><<<g, 256, 0, st>>>(dq_accum, (bf16*)dq, strides[15], strides[16],
                                                                 strides[17], Hq, Sq, p.Sq_pad);
        NXD_LAUNCH_CHECK();
    }
    return 0;
}

NXD_API int nxd_flash_attn_bwd_tc(const void* q, const void* k, const void* v, const void* o, const void* dout,
                                  const float* lse, float* lse_pad, float* delta_pad, float* dq_accum, float* dkv_accum,
                                  void* dq, void* dk, void* dv, const long long* strides, int B, int Hq, int Hkv, int Sq,
                                  int Sk, int D, float scale, int causal, void* stream) {
    return bwd_tc_impl(q, k, v, o, dout, lse, lse_pad, delta_pad, dq_accum, dkv_accum, dq, dk, dv, strides, B, Hq, Hkv, Sq,
                       Sk, D, scale, causal, 0.f, 0ull, stream);
}

// Backward of nxd_flash_attn_fwd_tc_dropout: same dropout_p and seed as the forward call.
NXD_API int nxd_flash_attn_bwd_tc_dropout(const void* q, const void* k, const void* v, const void* o, const void* dout,
                                          const float* lse, float* lse_pad, float* delta_pad, float* dq_accum,
                                          float* dkv_accum, void* dq, void* dk, void* dv, const long long* strides, int B,
                                          int Hq, int Hkv, int Sq, int Sk, int D, float scale, int causal,
                                          float dropout_p, unsigned long long seed, void* stream) {
    return bwd_tc_impl(q, k, v, o, dout, lse, lse_pad, delta_pad, dq_accum, dkv_accum, dq, dk, dv, strides, B, Hq, Hkv, Sq,
                       Sk, D, scale, causal, dropout_p, seed, stream);
}
