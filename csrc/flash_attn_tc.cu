// Flash attention FORWARD on 5th-gen tensor cores (tcgen05 + TMEM + TMA), sm_100a  (SURVEY G4/G6, C23).
//
// Same contract as nxd_flash_attn_fwd in flash_attn.cu (reference: the external NKI `flash_fwd` kernel behind
// src/neuronx_distributed/kernels/flash_attn.py:151-191 and the inference prefill kernel,
// examples/inference/modules/attention/attention_base.py:98-139); this is the fast path, the mma.sync kernel
// stays as the portable fallback and as a second oracle in the tests.
//
// One CTA = 128 query rows of one (batch, head); 192 threads, TWO CTAs per SM so that one CTA's softmax overlaps
// the other CTA's MMAs on the shared tensor core:
//   warp 0    TMA producer: Q once (4-D tensor map over the strided [B,H,S,D] view), then K/V tiles of 64 rows into a
//             3-slot ring in CONSUMPTION order  K0, K1, V0, K2, V1, ...  (so a slot's next load never waits on a
//             late MMA).  K and V tiles have the same smem image (64-wide column groups of 128-byte swizzled rows);
//             only the UMMA descriptor differs: K is the K-major B operand of S = Q K^T, V the MN-major B of O += P V.
//   warp 1    TMEM allocation (O: D columns, S: 64 columns) and MMA issue.  S(j+1) is issued BEFORE waiting for
//             P(j), so it runs while the softmax warps work on tile j.
//   warps 2-5 softmax: one thread per query row (TMEM lane).  tcgen05.ld S -> scale, mask, running max, exp2 ->
//             bf16 P written to tensor memory (tcgen05.st), the A operand of the PV MMA.  O stays in TMEM across the
//             whole loop; it is rescaled (tcgen05.ld / st) only when the running max moved by more than 2^8
//             ("lazy rescale"), which is rare after the first tiles.  Epilogue: O / l -> bf16 -> smem -> TMA store.
#include "common.cuh"
#include "gemm_tcgen05.cuh"
#include "attn_tc_common.cuh"
#include "ptx.cuh"

#include <type_traits>

namespace {
using namespace nxd_attn;

constexpr int kBM = 128;
constexpr int kBN = 64;
constexpr int kSlots = 3;
constexpr int kThreadsTc = 192;
constexpr float kRescaleThreshold = 8.0f;  // log2 units

struct TcParams {
    float* lse;
    int B, Hq, Hkv, Sq, Sk;
    int pos_q[3], pos_k[3], pos_v[3], pos_o[3];  // coordinate slot (1..3) of (row, head, batch) in each tensor map
    float scale_log2;
    DropParams drop;
};

template <int D>
struct TcSmem {
    static constexpr int kQBytes = kBM * D * 2;
    static constexpr int kSlotBytes = kBN * D * 2;
    static constexpr int kPBytes = kBM * kBN * 2;
    static constexpr int kQOff = 0;
    static constexpr int kRingOff = kQBytes;
    static constexpr int kPOff = kRingOff + kSlots * kSlotBytes;
    static constexpr int kBarOff = kPOff + kPBytes;
    static constexpr int kNumBars = 1 + 2 * kSlots + 4;
    static constexpr int kBytes = kBarOff + kNumBars * 8 + 16 + 1024;  // + manual 1024-byte alignment slack
    static constexpr int kTmemCols = 256;   // O (D) + S (64) + bf16 P as the tensor-memory A operand of P V (32)
};

template <int D, bool CAUSAL, bool DROP>
__global__ void __launch_bounds__(kThreadsTc, 2)
fa_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                 const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ CUtensorMap tmap_o,
                 const TcParams p) {
    using L = TcSmem<D>;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = align_smem_1024(smem_raw);
    uint8_t* sQ = smem + L::kQOff;
    uint8_t* sRing = smem + L::kRingOff;
    uint8_t* sP = smem + L::kPOff;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kBarOff);
    uint64_t* q_full = bars;
    uint64_t* slot_full = bars + 1;
    uint64_t* slot_empty = slot_full + kSlots;
    uint64_t* s_full = slot_empty + kSlots;
    uint64_t* s_empty = s_full + 1;
    uint64_t* p_full = s_empty + 1;
    uint64_t* pv_done = p_full + 1;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(pv_done + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // grid = (q heads, batch, q blocks): the q block is the slowest launch dimension and counted from the end, so the
    // long causal rows of EVERY head are scheduled before any short one (longest-first list scheduling)
    const int mb = gridDim.z - 1 - blockIdx.z;
    const int hq = blockIdx.x, b = blockIdx.y;
    const int hkv = hq / (p.Hq / p.Hkv);
    const int m0 = mb * kBM;
    const int offset = p.Sk - p.Sq;
    int n_tiles = (p.Sk + kBN - 1) / kBN;
    if (CAUSAL) n_tiles = min(n_tiles, max((m0 + kBM + offset + kBN - 1) / kBN, 0));

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tensormap(&tmap_q);
        ptx::prefetch_tensormap(&tmap_k);
        ptx::prefetch_tensormap(&tmap_v);
        ptx::prefetch_tensormap(&tmap_o);
        ptx::mbar_init(q_full, 1);
        for (int i = 0; i < kSlots; ++i) {
            ptx::mbar_init(&slot_full[i], 1);
            ptx::mbar_init(&slot_empty[i], 1);
        }
        ptx::mbar_init(s_full, 1);
        ptx::mbar_init(s_empty, 128);
        ptx::mbar_init(p_full, 128);
        ptx::mbar_init(pv_done, 1);
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<1>(tmem_ptr_smem, L::kTmemCols);
    ptx::tcgen05_fence_before();
    __syncthreads();
    ptx::tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;
    const uint32_t tmem_o = tmem_base;
    const uint32_t tmem_s = tmem_base + D;
    const uint32_t tmem_p = tmem_base + D + kBN;   // 128 lanes x 64 keys bf16 = 32 packed columns

    if (warp == 0) {
        // ============================ TMA producer ============================
        if (lane == 0 && n_tiles > 0) {
            int c[4];
            ptx::mbar_arrive_expect_tx(q_full, L::kQBytes);
#pragma unroll
            for (int g = 0; g < D / 64; ++g) {
                c[0] = g * 64;
                coords(p.pos_q, m0, hq, b, c);
                tma_load_4d(&tmap_q, q_full, sQ + g * (kBM * 128), c[0], c[1], c[2], c[3]);
            }
            int it = 0;
            auto load = [&](const CUtensorMap* map, const int* pos, int row0) {
                const int slot = it % kSlots;
                const uint32_t ph = (uint32_t)(it / kSlots) & 1u;
                ptx::mbar_wait(&slot_empty[slot], ph ^ 1u);
                ptx::mbar_arrive_expect_tx(&slot_full[slot], L::kSlotBytes);
#pragma unroll
                for (int g = 0; g < D / 64; ++g) {
                    c[0] = g * 64;
                    coords(pos, row0, hkv, b, c);
                    tma_load_4d(map, &slot_full[slot], sRing + slot * L::kSlotBytes + g * (kBN * 128), c[0], c[1], c[2],
                                c[3]);
                }
                ++it;
            };
            load(&tmap_k, p.pos_k, 0);
            for (int j = 0; j < n_tiles; ++j) {
                if (j + 1 < n_tiles) load(&tmap_k, p.pos_k, (j + 1) * kBN);
                load(&tmap_v, p.pos_v, j * kBN);
            }
        }
    } else if (warp == 1) {
        // ============================ MMA issuer ============================
        if (lane == 0 && n_tiles > 0) {
            constexpr uint32_t idesc_s = ptx::make_idesc_bf16(kBM, kBN, 0, 0);
            constexpr uint32_t idesc_o = ptx::make_idesc_bf16(kBM, D, 0, 1);
            const uint32_t q_addr = ptx::smem_u32(sQ), ring_addr = ptx::smem_u32(sRing);
            int it = 0;
            auto issue_s = [&](int j) {
                const int slot = it % kSlots;
                const uint32_t ph = (uint32_t)(it / kSlots) & 1u;
                ++it;
                ptx::mbar_wait(&slot_full[slot], ph);
                ptx::mbar_wait(s_empty, ((uint32_t)j & 1u) ^ 1u);  // softmax(j-1) has read S out of TMEM
                ptx::tcgen05_fence_after();
                const uint32_t kb = ring_addr + slot * L::kSlotBytes;
#pragma unroll
                for (int kk = 0; kk < D / 16; ++kk) {
                    const uint64_t da = ptx::make_smem_desc(q_addr + (kk >> 2) * (kBM * 128) + (kk & 3) * 32, 16, 1024);
                    const uint64_t db = ptx::make_smem_desc(kb + (kk >> 2) * (kBN * 128) + (kk & 3) * 32, 16, 1024);
                    ptx::umma_bf16<1>(tmem_s, da, db, idesc_s, kk != 0 ? 1u : 0u);
                }
                ptx::umma_commit(&slot_empty[slot]);
                ptx::umma_commit(s_full);
            };
            ptx::mbar_wait(q_full, 0);
            issue_s(0);
            for (int j = 0; j < n_tiles; ++j) {
                if (j + 1 < n_tiles) issue_s(j + 1);
                const int slot = it % kSlots;
                const uint32_t ph = (uint32_t)(it / kSlots) & 1u;
                ++it;
                ptx::mbar_wait(&slot_full[slot], ph);
                ptx::mbar_wait(p_full, (uint32_t)j & 1u);
                ptx::tcgen05_fence_after();
                const uint32_t vb = ring_addr + slot * L::kSlotBytes;
#pragma unroll
                for (int kk = 0; kk < kBN / 16; ++kk) {
                    const uint64_t db = ptx::make_smem_desc(vb + kk * 2048, kBN * 128, 1024);
                    ptx::umma_bf16_ts(tmem_o, tmem_p + kk * 8, db, idesc_o, (j | kk) != 0 ? 1u : 0u);
                }
                ptx::umma_commit(&slot_empty[slot]);
                ptx::umma_commit(pv_done);
            }
        }
    } else {
        // ============================ softmax / epilogue ============================
        const int q = warp & 3;                 // TMEM lane quadrant of this warp
        const int row = q * 32 + lane;          // row inside the q tile == TMEM lane
        const int grow = m0 + row;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        // Running max in RAW score units (the scale is positive, so max commutes with it); the exponent argument is one
        // FFMA per element: s * scale_log2 - m * scale_log2.
        float m_run = -INFINITY, l_run = 0.f;
        // one tile of the online softmax; MASK is instantiated only for tiles on the causal diagonal / past Sk
        auto tile_softmax = [&](auto mask_tag, int j) {
            constexpr bool MASK = decltype(mask_tag)::value;
            uint32_t sr[kBN];
            ptx::tmem_ld_32x32b_x32(tmem_s + lane_off, sr);
            ptx::tmem_ld_32x32b_x32(tmem_s + lane_off + 32, sr + 32);
            ptx::tmem_ld_wait();
            ptx::tcgen05_fence_before();
            ptx::mbar_arrive(s_empty);
            const int n0 = j * kBN;
            if (MASK) {
                // columns c >= first_bad are masked: beyond the causal diagonal of this row, or beyond the sequence end
                int first_bad = p.Sk - n0;
                if (CAUSAL) first_bad = min(first_bad, grow + offset + 1 - n0);
#pragma unroll
                for (int c = 0; c < kBN; ++c)
                    if (c >= first_bad) sr[c] = 0xff800000u;   // -inf
            }
            float mx0 = __uint_as_float(sr[0]), mx1 = __uint_as_float(sr[1]), mx2 = __uint_as_float(sr[2]),
                  mx3 = __uint_as_float(sr[3]);
#pragma unroll
            for (int c = 4; c < kBN; c += 4) {
                mx0 = fmaxf(mx0, __uint_as_float(sr[c]));
                mx1 = fmaxf(mx1, __uint_as_float(sr[c + 1]));
                mx2 = fmaxf(mx2, __uint_as_float(sr[c + 2]));
                mx3 = fmaxf(mx3, __uint_as_float(sr[c + 3]));
            }
            const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
            // lazy rescale: keep the stale max unless it moved by more than 2^8 (in scaled log2 units)
            float alpha = 1.f;
            bool rescale = false;
            if ((mx - m_run) * p.scale_log2 > kRescaleThreshold || m_run == -INFINITY) {
                const float m_new = fmaxf(m_run, mx);
                if (m_new != -INFINITY) {
                    alpha = ex2_approx((m_run - m_new) * p.scale_log2);   // 0 when m_run was -inf
                    rescale = (j > 0) && (m_run != -INFINITY);
                    m_run = m_new;
                }
            }
            const float neg_m = (m_run == -INFINITY) ? 0.f : -m_run * p.scale_log2;
            float rs0 = 0.f, rs1 = 0.f;
            uint32_t pk[kBN / 2];
#pragma unroll
            for (int c = 0; c < kBN; c += 2) {
                const float p0 = ex2_approx(fmaf(__uint_as_float(sr[c]), p.scale_log2, neg_m));
                const float p1 = ex2_approx(fmaf(__uint_as_float(sr[c + 1]), p.scale_log2, neg_m));
                rs0 += p0;
                rs1 += p1;
                if (DROP) {   // the row sum (softmax denominator) is taken before dropout, the PV operand after it
                    const uint32_t hs = drop_head_seed(p.drop.seed, (uint32_t)(b * p.Hq + hq));
                    const bool k0 = drop_keep(hs, (uint32_t)grow, (uint32_t)(n0 + c), p.drop.threshold);
                    const bool k1 = drop_keep(hs, (uint32_t)grow, (uint32_t)(n0 + c + 1), p.drop.threshold);
                    pk[c >> 1] = pack_bf16x2(k0 ? p0 * p.drop.inv_keep : 0.f, k1 ? p1 * p.drop.inv_keep : 0.f);
                } else {
                    pk[c >> 1] = pack_bf16x2(p0, p1);
                }
            }
            l_run = l_run * alpha + (rs0 + rs1);
            if (j > 0) {
                ptx::mbar_wait(pv_done, ((uint32_t)(j - 1)) & 1u);  // P buffer free, O up to date
                ptx::tcgen05_fence_after();
                if (__any_sync(0xffffffffu, rescale)) {
#pragma unroll 1
                    for (int c = 0; c < D / 32; ++c) {
                        uint32_t o[32];
                        ptx::tmem_ld_32x32b_x32(tmem_o + lane_off + c * 32, o);
                        ptx::tmem_ld_wait();
                        const float a_o = rescale ? alpha : 1.f;
#pragma unroll
                        for (int e = 0; e < 32; ++e) o[e] = __float_as_uint(__uint_as_float(o[e]) * a_o);
                        tmem_st_32x32b_x32(tmem_o + lane_off + c * 32, o);
                    }
                    tmem_st_wait();
                }
            }
            // P goes to TENSOR MEMORY (this thread's row, 64 keys = 32 packed words): the A operand of the P V MMA never
            // touches shared memory (no 16 KB store + re-read per tile, no async-proxy fence)
            tmem_st_32x32b_x16(tmem_p + lane_off, pk);
            tmem_st_32x32b_x16(tmem_p + lane_off + 16, pk + 16);
            tmem_st_wait();
            ptx::tcgen05_fence_before();
            ptx::mbar_arrive(p_full);
        };
        for (int j = 0; j < n_tiles; ++j) {
            ptx::mbar_wait(s_full, (uint32_t)j & 1u);
            ptx::tcgen05_fence_after();
            const int n0 = j * kBN;
            // warp-uniform: the tile crosses the sequence end, or the causal diagonal of one of this warp's 32 rows
            const bool need_mask = (n0 + kBN > p.Sk) || (CAUSAL && (n0 + kBN - 1 > m0 + q * 32 + offset));
            if (need_mask) tile_softmax(std::true_type{}, j);
            else tile_softmax(std::false_type{}, j);
        }
        // ---- epilogue ----
        if (n_tiles > 0) {
            ptx::mbar_wait(pv_done, ((uint32_t)(n_tiles - 1)) & 1u);
            ptx::tcgen05_fence_after();
        }
        const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
        if (p.lse && grow < p.Sq)
            p.lse[((long long)b * p.Hq + hq) * p.Sq + grow] = l_run > 0.f ? m_run * p.scale_log2 + log2f(l_run) : INFINITY;
        // every MMA that read Q has completed: reuse the Q tile as staging (warp q owns rows [32q, 32q+32) of
        // each 64-column group: 32 rows x 128 bytes = one TMA store box)
#pragma unroll 1
        for (int g = 0; g < D / 64; ++g) {
            uint32_t o[64];
            if (n_tiles > 0) {
                ptx::tmem_ld_32x32b_x32(tmem_o + lane_off + g * 64, o);
                ptx::tmem_ld_32x32b_x32(tmem_o + lane_off + g * 64 + 32, o + 32);
                ptx::tmem_ld_wait();
            } else {
#pragma unroll
                for (int e = 0; e < 64; ++e) o[e] = 0u;
            }
            uint8_t* dst = sQ + g * (kBM * 128) + row * 128;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                uint32_t w[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    bf162 t = __floats2bfloat162_rn(__uint_as_float(o[8 * c + 2 * e]) * inv,
                                                    __uint_as_float(o[8 * c + 2 * e + 1]) * inv);
                    w[e] = *reinterpret_cast<uint32_t*>(&t);
                }
                st_shared_v4(dst + ((c ^ (row & 7)) << 4), w[0], w[1], w[2], w[3]);
            }
            ptx::fence_proxy_async();
            __syncwarp();
            if (lane == 0 && m0 + q * 32 < p.Sq) {
                int c[4];
                c[0] = g * 64;
                coords(p.pos_o, m0 + q * 32, hq, b, c);
                tma_store_4d(&tmap_o, sQ + g * (kBM * 128) + q * 32 * 128, c[0], c[1], c[2], c[3]);
                ptx::bulk_commit();
            }
        }
        if (lane == 0) ptx::bulk_wait<0>();
    }
    ptx::tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<1>(tmem_base, L::kTmemCols);
}

template <int D, bool CAUSAL, bool DROP>
int launch_fwd_tc_one(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const CUtensorMap& to,
                      const TcParams& p, cudaStream_t st) {
    using L = TcSmem<D>;
    dim3 grid(p.Hq, p.B, (p.Sq + kBM - 1) / kBM);
    static cudaError_t once = cudaFuncSetAttribute(fa_fwd_tc_kernel<D, CAUSAL, DROP>,
                                                   cudaFuncAttributeMaxDynamicSharedMemorySize, L::kBytes);
    NXD_CHECK_CUDA(once);
    fa_fwd_tc_kernel<D, CAUSAL, DROP><<<grid, kThreadsTc, L::kBytes, st>>>(tq, tk, tv, to, p);
    NXD_LAUNCH_CHECK();
    return 0;
}

template <int D>
int launch_fwd_tc(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const CUtensorMap& to,
                  const TcParams& p, int causal, cudaStream_t st) {
    const bool drop = p.drop.threshold != 0;
    if (causal) return drop ? launch_fwd_tc_one<D, true, true>(tq, tk, tv, to, p, st)
                            : launch_fwd_tc_one<D, true, false>(tq, tk, tv, to, p, st);
    return drop ? launch_fwd_tc_one<D, false, true>(tq, tk, tv, to, p, st)
                : launch_fwd_tc_one<D, false, false>(tq, tk, tv, to, p, st);
}

}  // namespace

// Same arguments as nxd_flash_attn_fwd.  Extra requirement: 16-byte aligned base pointers.
static int fwd_tc_impl(const void* q, const void* k, const void* v, void* o, float* lse, const long long* strides, int B,
                       int Hq, int Hkv, int Sq, int Sk, int D, float scale, int causal, float dropout_p,
                       unsigned long long seed, void* stream) {
    NXD_REQUIRE(D == 64 || D == 128, "flash_attn_tc: head dim %d not supported (64 or 128)", D);
    NXD_REQUIRE(B > 0 && Hq > 0 && Hkv > 0 && Sq > 0 && Sk > 0 && Hq % Hkv == 0, "flash_attn_tc: bad problem");
    for (int i = 0; i < 12; ++i)
        NXD_REQUIRE(strides[i] % 8 == 0, "flash_attn_tc: stride %d (=%lld) must be a multiple of 8 elements", i,
                    strides[i]);
    NXD_REQUIRE((((uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)o) & 15) == 0,
                "flash_attn_tc: tensors must be 16-byte aligned");
    TcParams p{};
    p.lse = lse;
    p.B = B; p.Hq = Hq; p.Hkv = Hkv; p.Sq = Sq; p.Sk = Sk;
    p.scale_log2 = scale * 1.4426950408889634f;
    NXD_REQUIRE(dropout_p >= 0.f && dropout_p < 1.f, "flash_attn_tc: dropout_p %f not in [0, 1)", dropout_p);
    p.drop.threshold = dropout_p > 0.f ? (uint32_t)fmin(4294967295.0, (double)dropout_p * 4294967296.0) : 0u;
    p.drop.inv_keep = 1.f / (1.f - dropout_p);
    p.drop.seed = (uint32_t)(seed ^ (seed >> 32));
    CUtensorMap tq, tk, tv, to;
    int rc;
    if ((rc = make_tmap_bhsd(&tq, q, D, Sq, Hq, B, strides[2], strides[1], strides[0], kBM, p.pos_q))) return rc;
    if ((rc = make_tmap_bhsd(&tk, k, D, Sk, Hkv, B, strides[5], strides[4], strides[3], kBN, p.pos_k))) return rc;
    if ((rc = make_tmap_bhsd(&tv, v, D, Sk, Hkv, B, strides[8], strides[7], strides[6], kBN, p.pos_v))) return rc;
    if ((rc = make_tmap_bhsd(&to, o, D, Sq, Hq, B, strides[11], strides[10], strides[9], 32, p.pos_o))) return rc;
    return D == 64 ? launch_fwd_tc<64>(tq, tk, tv, to, p, causal, (cudaStream_t)stream)
                   : launch_fwd_tc<128>(tq, tk, tv, to, p, causal, (cudaStream_t)stream);
}

NXD_API int nxd_flash_attn_fwd_tc(const void* q, const void* k, const void* v, void* o, float* lse,
                                  const long long* strides, int B, int Hq, int Hkv, int Sq, int Sk, int D, float scale,
                                  int causal, void* stream) {
    return fwd_tc_impl(q, k, v, o, lse, strides, B, Hq, Hkv, Sq, Sk, D, scale, causal, 0.f, 0ull, stream);
}

// Same, with attention dropout: P is multiplied by a keep mask / (1 - dropout_p) before the PV product; the mask is a
// function of (seed, batch, head, query, key) that nxd_flash_attn_bwd_tc_dropout regenerates.
NXD_API int nxd_flash_attn_fwd_tc_dropout(const void* q, const void* k, const void* v, void* o, float* lse,
                                          const long long* strides, int B, int Hq, int Hkv, int Sq, int Sk, int D,
                                          float scale, int causal, float dropout_p, unsigned long long seed,
                                          void* stream) {
    return fwd_tc_impl(q, k, v, o, lse, strides, B, Hq, Hkv, Sq, Sk, D, scale, causal, dropout_p, seed, stream);
}
