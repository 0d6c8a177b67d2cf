// Persistent, warp-specialised bf16 GEMM on 5th-gen tensor cores (tcgen05 + TMEM + TMA), sm_100a.
//
//   D[M, N] (+)= A_op[M, K] * B_op[N, K]^T       (fp32 accumulate in TMEM; D bf16 or fp32)
//
// This is the compute core behind ColumnParallelLinear / RowParallelLinear forward, dgrad and
// wgrad (reference src/neuronx_distributed/parallel_layers/layers.py:325,348,409 -- plain
// torch.matmul lowered by XLA).  The three operand layouts are handled with the operand-major
// bits of the UMMA descriptors, never with transposed copies:
//     "nt"  A = X[M,K]   (K-major)    B = W[N,K]   (K-major)      forward
//     "nn"  A = dY[M,K]  (K-major)    B = W[K,N]   (MN-major)     dgrad
//     "tn"  A = dY[K,M]  (MN-major)   B = X[K,N]   (MN-major)     wgrad
//
// Structure (one CTA pair = one cluster of 2 CTAs = one 256 x BLOCK_N output tile):
//   warp 0   TMA producer: each CTA loads its 128 rows of A and its BLOCK_N/2 rows of B per
//            64-wide K block into a kStages-deep smem ring (SWIZZLE_128B); transaction bytes of
//            both CTAs are credited to the LEADER CTA's "full" mbarrier.
//   warp 1   MMA issuer (leader CTA only, one elected lane): tcgen05.mma.cta_group::2 with
//            M=256, N=BLOCK_N, K=16; tcgen05.commit (multicast to both CTAs) releases smem slots
//            and publishes finished accumulators.
//   warp 2   TMEM allocator (2 accumulator stages x BLOCK_N fp32 columns).
//   warps 4-7 epilogue: tcgen05.ld the 128 x BLOCK_N accumulator (lane = row), convert, stage 32 x 128 B
//            boxes in swizzled smem and hand them to the TMA unit (bulk tensor store; fp32
//            accumulate = TMA reduce-add).  The accumulator is double buffered so the epilogue of
//            tile i overlaps the main loop of tile i+1.  (The first version stored 16 B per thread
//            straight from registers; at K = 3072 that epilogue (~50k cycles / tile) was longer than
//            the main loop (~25k) and capped the kernel at 0.66 of the measured peak.)
// Tiles are visited persistently (grid = #SM/2 clusters) in an L2-friendly grouped order.
#include "common.cuh"
#include "gemm_tcgen05.cuh"
#include "ptx.cuh"

namespace nxd_gemm {

// GROUPED: `batches` independent problems of the same shape (the experts of an MoE layer) in ONE launch.  The tensor
// maps are 3-D, the outermost coordinate is the problem index; a tile never mixes two problems because rows past M
// are out of bounds of dimension 1 (zero-filled on load, clipped on store), so M needs no alignment.
template <int BN, bool A_MN, bool B_MN, int OUT_MODE, bool GROUPED = false>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm_pair_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const __grid_constant__ CUtensorMap tmap_d, int M, int N, int K, int batches) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    using L = SmemLayout<BN>;
    uint8_t* smem = ptx::align_smem_1024(smem_raw);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarrierOffset);
    constexpr int kStages = L::kStages;
    uint64_t* empty_bar = full_bar + kStages;
    uint64_t* tmem_full_bar = empty_bar + kStages;
    uint64_t* tmem_empty_bar = tmem_full_bar + kAccStages;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty_bar + kAccStages);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t cta_rank = ptx::cluster_ctarank();
    const bool is_leader = cta_rank == 0;

    const int num_m_tiles = (M + kTileM - 1) / kTileM;
    const int num_n_tiles = (N + BN - 1) / BN;
    const int tiles_per_batch = num_m_tiles * num_n_tiles;
    const int num_tiles = tiles_per_batch * (GROUPED ? batches : 1);
    const int num_k_blocks = (K + kBlockK - 1) / kBlockK;
    const int cluster_id = blockIdx.x >> 1;
    const int num_clusters = gridDim.x >> 1;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tensormap(&tmap_a);
        ptx::prefetch_tensormap(&tmap_b);
        ptx::prefetch_tensormap(&tmap_d);
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < kStages; ++i) {
            ptx::mbar_init(&full_bar[i], 1);
            ptx::mbar_init(&empty_bar[i], 1);
        }
        for (int i = 0; i < kAccStages; ++i) {
            ptx::mbar_init(&tmem_full_bar[i], 1);
            ptx::mbar_init(&tmem_empty_bar[i], 2 * kEpilogueWarps);  // both CTAs' epilogue warps
        }
        ptx::fence_barrier_init();
    }
    if (warp == 2) ptx::tmem_alloc<2>(tmem_ptr_smem, L::kTmemCols);
    ptx::tcgen05_fence_before();
    ptx::cluster_sync();
    ptx::tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    if (warp == 0) {
        // ======================= TMA producer =======================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
                int tm, tn;
                const int e = GROUPED ? tile / tiles_per_batch : 0;
                tile_coords(GROUPED ? tile - e * tiles_per_batch : tile, num_m_tiles, num_n_tiles, tm, tn);
                const int m0 = tm * kTileM + (int)cta_rank * kBlockM;
                const int n0 = tn * BN + (int)cta_rank * (BN / 2);
                for (int kb = 0; kb < num_k_blocks; ++kb) {
                    ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * L::kStageBytes;
                    uint8_t* sb = sa + L::kABytes;
                    const int k0 = kb * kBlockK;
                    if (is_leader) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * L::kStageBytes);
                    auto load = [&](const CUtensorMap* map, void* dst, int c0, int c1) {
                        if constexpr (GROUPED) ptx::tma_load_3d_pair(map, &full_bar[stage], dst, c0, c1, e);
                        else ptx::tma_load_2d_pair(map, &full_bar[stage], dst, c0, c1);
                    };
                    if constexpr (!A_MN) {
                        load(&tmap_a, sa, k0, m0);
                    } else {
#pragma unroll
                        for (int g = 0; g < kBlockM / 64; ++g) load(&tmap_a, sa + g * (kBlockK * 128), m0 + g * 64, k0);
                    }
                    if constexpr (!B_MN) {
                        load(&tmap_b, sb, k0, n0);
                    } else {
#pragma unroll
                        for (int g = 0; g < (BN / 2) / 64; ++g) load(&tmap_b, sb + g * (kBlockK * 128), n0 + g * 64, k0);
                    }
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ======================= MMA issuer (leader CTA) =======================
        if (is_leader && lane == 0) {
            constexpr uint32_t idesc = ptx::make_idesc_bf16(kTileM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
            // K-major: 8-row groups 1024 B apart; step 32 B per UMMA_K inside the 128 B swizzle row.
            // MN-major: 64-element MN groups kBlockK*128 B apart (LBO), 8-k-row groups 1024 B apart
            // (SBO); step 16 k-rows = 2048 B per UMMA_K.
            constexpr uint32_t a_lbo = A_MN ? kBlockK * 128 : 16, b_lbo = B_MN ? kBlockK * 128 : 16;
            constexpr uint32_t a_kstep = A_MN ? 2048 : 32, b_kstep = B_MN ? 2048 : 32;
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
                ptx::mbar_wait_cluster(&tmem_empty_bar[acc], acc_phase ^ 1);
                ptx::tcgen05_fence_after();
                const uint32_t tmem_d = tmem_base + acc * BN;
                for (int kb = 0; kb < num_k_blocks; ++kb) {
                    ptx::mbar_wait_cluster(&full_bar[stage], phase);
                    ptx::tcgen05_fence_after();
                    const uint32_t sa = ptx::smem_u32(smem + stage * L::kStageBytes);
                    const uint32_t sb = sa + L::kABytes;
#pragma unroll
                    for (int k = 0; k < kBlockK / 16; ++k) {
                        const uint64_t da = ptx::make_smem_desc(sa + k * a_kstep, a_lbo, 1024);
                        const uint64_t db = ptx::make_smem_desc(sb + k * b_kstep, b_lbo, 1024);
                        ptx::umma_bf16<2>(tmem_d, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    ptx::umma_commit_pair(&empty_bar[stage], 0b11);  // frees the slot in both CTAs
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
                ptx::umma_commit_pair(&tmem_full_bar[acc], 0b11);  // accumulator ready (both CTAs)
                if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else if (warp >= kFirstEpilogueWarp) {
        // ======================= epilogue =======================
        // TMEM -> registers -> swizzled smem staging -> TMA store (or TMA fp32 reduce-add): every global write is
        // a full 128-byte row segment issued by the TMA unit, nothing goes through the LSU store path.
        const int q = warp & 3;  // TMEM lane quadrant this warp may access
        constexpr int kCols = (OUT_MODE == kOutBF16) ? 64 : 32;  // accumulator columns per 128-byte staging row
        uint8_t* stg = smem + L::kStagingOffset + q * L::kStagingBytesPerWarp;
        int buf = 0;
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
            int tm, tn;
            const int e = GROUPED ? tile / tiles_per_batch : 0;
            tile_coords(GROUPED ? tile - e * tiles_per_batch : tile, num_m_tiles, num_n_tiles, tm, tn);
            const int row0 = tm * kTileM + (int)cta_rank * kBlockM + q * 32;
            const int n0 = tn * BN;
            ptx::mbar_wait(&tmem_full_bar[acc], acc_phase);
            ptx::tcgen05_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN;
#pragma unroll 1
            for (int c = 0; c < BN / kCols; ++c) {
                if (lane == 0) ptx::bulk_wait_read<1>();  // the store that used this buffer two chunks ago has read it
                __syncwarp();
                uint32_t r[kCols];
                ptx::tmem_ld_32x32b_x32(taddr + c * kCols, r);
                if constexpr (kCols == 64) ptx::tmem_ld_32x32b_x32(taddr + c * kCols + 32, r + 32);
                ptx::tmem_ld_wait();
                if (c == BN / kCols - 1) {  // accumulator fully read: hand it back to the MMA warp early
                    ptx::tcgen05_fence_before();
                    __syncwarp();
                    if (lane == 0) ptx::mbar_arrive_cluster(&tmem_empty_bar[acc], 0);
                }
                uint8_t* dst = stg + buf * (32 * 128) + lane * 128;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    uint4 v;
                    if constexpr (OUT_MODE == kOutBF16) {
                        bf162 p0 = __floats2bfloat162_rn(__uint_as_float(r[8 * j]), __uint_as_float(r[8 * j + 1]));
                        bf162 p1 = __floats2bfloat162_rn(__uint_as_float(r[8 * j + 2]), __uint_as_float(r[8 * j + 3]));
                        bf162 p2 = __floats2bfloat162_rn(__uint_as_float(r[8 * j + 4]), __uint_as_float(r[8 * j + 5]));
                        bf162 p3 = __floats2bfloat162_rn(__uint_as_float(r[8 * j + 6]), __uint_as_float(r[8 * j + 7]));
                        v = make_uint4(*reinterpret_cast<uint32_t*>(&p0), *reinterpret_cast<uint32_t*>(&p1),
                                       *reinterpret_cast<uint32_t*>(&p2), *reinterpret_cast<uint32_t*>(&p3));
                    } else {
                        v = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
                    }
                    *reinterpret_cast<uint4*>(dst + ((j ^ (lane & 7)) << 4)) = v;
                }
                ptx::fence_proxy_async();
                __syncwarp();
                if (lane == 0) {
                    const int col0 = n0 + c * kCols;
                    if (col0 < N && row0 < M) {
                        if constexpr (GROUPED) {
                            if constexpr (OUT_MODE == kOutF32Acc)
                                ptx::tma_reduce_add_3d(&tmap_d, stg + buf * (32 * 128), col0, row0, e);
                            else
                                ptx::tma_store_3d(&tmap_d, stg + buf * (32 * 128), col0, row0, e);
                        } else if constexpr (OUT_MODE == kOutF32Acc) {
                            ptx::tma_reduce_add_2d(&tmap_d, stg + buf * (32 * 128), col0, row0);
                        } else {
                            ptx::tma_store_2d(&tmap_d, stg + buf * (32 * 128), col0, row0);
                        }
                    }
                    ptx::bulk_commit();
                }
                buf ^= 1;
            }
            if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
        }
        if (lane == 0) ptx::bulk_wait<0>();
    }

    // ---- teardown: both CTAs must be done with TMEM (and with remote barrier traffic) ----
    ptx::tcgen05_fence_before();
    ptx::cluster_sync();
    if (warp == 2) ptx::tmem_dealloc<2>(tmem_base, L::kTmemCols);
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------

template <int BN, bool A_MN, bool B_MN, int OUT_MODE>
static int launch(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& td, int M, int N, int K,
                  cudaStream_t stream) {
    using L = SmemLayout<BN>;
    auto kern = gemm_pair_kernel<BN, A_MN, B_MN, OUT_MODE>;
    static bool configured = false;
    if (!configured) {
        NXD_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kSmemBytes));
        configured = true;
    }
    const int num_tiles = ((M + kTileM - 1) / kTileM) * ((N + BN - 1) / BN);
    int clusters = nxd_num_sms() / 2;
    if (clusters > num_tiles) clusters = num_tiles;
    kern<<<dim3(clusters * 2), dim3(kThreads), L::kSmemBytes, stream>>>(ta, tb, td, M, N, K, 1);
    NXD_LAUNCH_CHECK();
    return 0;
}

template <int BN, int OUT_MODE>
static int dispatch_layout(int layout, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& td, int M,
                           int N, int K, cudaStream_t s) {
    switch (layout) {
        case kLayoutNT: return launch<BN, false, false, OUT_MODE>(ta, tb, td, M, N, K, s);
        case kLayoutNN: return launch<BN, false, true, OUT_MODE>(ta, tb, td, M, N, K, s);
        default: return launch<BN, true, true, OUT_MODE>(ta, tb, td, M, N, K, s);
    }
}

int pick_block_n(int M, int N) {
    // Fewer, fatter tiles amortise operand traffic (256 x 256 per CTA pair is the sweet spot).  Measured on every
    // Llama-3.2 training / prefill shape (tools/bench_kernels.py, NXD_GEMM_CASES=wgrad,dgrad,prefill): a wave of
    // 256 x 128 tiles takes ~0.95 of the time of a wave of 256 x 256 tiles (the A operand traffic is the same), so the
    // narrow tile only pays when it does not add a wave.
    const int clusters = nxd_num_sms() / 2;
    const int mt = (M + kTileM - 1) / kTileM;
    const long t256 = (long)mt * ((N + 255) / 256), t128 = (long)mt * ((N + 127) / 128);
    const long w256 = (t256 + clusters - 1) / clusters, w128 = (t128 + clusters - 1) / clusters;
    return (w256 * 100 <= w128 * 95) ? 256 : 128;
}

int run(int layout, const void* a, const void* b, long long lda, long long ldb, const EpilogueArgs& ep, int M, int N,
        int K, int block_n, cudaStream_t stream) {
    NXD_REQUIRE(layout >= 0 && layout <= 2, "gemm: bad layout %d", layout);
    NXD_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: empty problem %dx%dx%d", M, N, K);
    NXD_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && ep.ldd % 8 == 0, "gemm: leading dims must be multiples of 8 elements");
    NXD_REQUIRE(N % 8 == 0 && (ep.mode == kOutBF16 || N % 4 == 0), "gemm: N must be a multiple of 8");
    NXD_REQUIRE(((uintptr_t)a & 15) == 0 && ((uintptr_t)b & 15) == 0 && ((uintptr_t)ep.d & 15) == 0,
                "gemm: operands must be 16-byte aligned");
    const bool a_mn = layout == kLayoutTN;
    const bool b_mn = layout != kLayoutNT;
    const int bn = block_n > 0 ? block_n : pick_block_n(M, N);
    NXD_REQUIRE(bn == 128 || bn == 256, "gemm: block_n must be 128 or 256");
    CUtensorMap ta, tb;
    int rc;
    // A: K-major -> global [M rows, K cols], box {64 k, 128 rows}; MN-major -> global [K rows, M cols], box {64 m, 64 k}
    if (!a_mn) rc = make_tmap_2d(&ta, a, (uint64_t)K, (uint64_t)M, (uint64_t)lda, 64, kBlockM);
    else rc = make_tmap_2d(&ta, a, (uint64_t)M, (uint64_t)K, (uint64_t)lda, 64, kBlockK);
    if (rc) return rc;
    if (!b_mn) rc = make_tmap_2d(&tb, b, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, 64, bn / 2);
    else rc = make_tmap_2d(&tb, b, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, 64, kBlockK);
    if (rc) return rc;
    // D: staged per warp as 32 rows x 128 bytes (64 bf16 or 32 fp32 columns)
    CUtensorMap td;
    if (ep.mode == kOutBF16) rc = make_tmap_2d(&td, ep.d, (uint64_t)N, (uint64_t)M, (uint64_t)ep.ldd, 64, 32, 2);
    else rc = make_tmap_2d(&td, ep.d, (uint64_t)N, (uint64_t)M, (uint64_t)ep.ldd, 32, 32, 4);
    if (rc) return rc;
#define NXD_GEMM_DISPATCH(BN_)                                                                           \
    switch (ep.mode) {                                                                                   \
        case kOutBF16: return dispatch_layout<BN_, kOutBF16>(layout, ta, tb, td, M, N, K, stream);       \
        case kOutF32: return dispatch_layout<BN_, kOutF32>(layout, ta, tb, td, M, N, K, stream);         \
        case kOutF32Acc: return dispatch_layout<BN_, kOutF32Acc>(layout, ta, tb, td, M, N, K, stream);   \
        default: nxd_set_error("gemm: bad output mode %d", ep.mode); return -1;                          \
    }
    if (bn == 256) { NXD_GEMM_DISPATCH(256) }
    NXD_GEMM_DISPATCH(128)
#undef NXD_GEMM_DISPATCH
}

template <int BN, bool A_MN, bool B_MN, int OUT_MODE>
static int launch_grouped(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& td, int E, int M, int N,
                          int K, cudaStream_t stream) {
    using L = SmemLayout<BN>;
    auto kern = gemm_pair_kernel<BN, A_MN, B_MN, OUT_MODE, true>;
    static bool configured = false;
    if (!configured) {
        NXD_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kSmemBytes));
        configured = true;
    }
    const long num_tiles = (long)E * ((M + kTileM - 1) / kTileM) * ((N + BN - 1) / BN);
    long clusters = nxd_num_sms() / 2;
    if (clusters > num_tiles) clusters = num_tiles;
    kern<<<dim3((unsigned)clusters * 2), dim3(kThreads), L::kSmemBytes, stream>>>(ta, tb, td, M, N, K, E);
    NXD_LAUNCH_CHECK();
    return 0;
}

// E problems D_e[M, N] (+)= A_e op B_e of one shape, operands stacked with strides sa / sb / sd (elements) between
// problems.  The tile shape comes from the total tile count (E * tiles per problem), M needs no alignment.
int run_grouped(int layout, const void* a, const void* b, long long lda, long long ldb, long long sa, long long sb,
                const EpilogueArgs& ep, long long sd, int E, int M, int N, int K, cudaStream_t stream) {
    NXD_REQUIRE(layout >= 0 && layout <= 2, "gemm_grouped: bad layout %d", layout);
    NXD_REQUIRE(E > 0 && M > 0 && N > 0 && K > 0, "gemm_grouped: empty problem %d x %dx%dx%d", E, M, N, K);
    NXD_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && ep.ldd % 8 == 0 && sa % 8 == 0 && sb % 8 == 0 && sd % 8 == 0,
                "gemm_grouped: leading dims and problem strides must be multiples of 8 elements");
    NXD_REQUIRE(N % 8 == 0, "gemm_grouped: N must be a multiple of 8");
    NXD_REQUIRE(((uintptr_t)a & 15) == 0 && ((uintptr_t)b & 15) == 0 && ((uintptr_t)ep.d & 15) == 0,
                "gemm_grouped: operands must be 16-byte aligned");
    const bool a_mn = layout == kLayoutTN;
    const bool b_mn = layout != kLayoutNT;
    const int bn = pick_block_n(M * E, N);
    CUtensorMap ta, tb, td;
    int rc;
    if (!a_mn) rc = make_tmap_3d(&ta, a, (uint64_t)K, (uint64_t)M, (uint64_t)E, (uint64_t)lda, (uint64_t)sa, 64, kBlockM);
    else rc = make_tmap_3d(&ta, a, (uint64_t)M, (uint64_t)K, (uint64_t)E, (uint64_t)lda, (uint64_t)sa, 64, kBlockK);
    if (rc) return rc;
    if (!b_mn) rc = make_tmap_3d(&tb, b, (uint64_t)K, (uint64_t)N, (uint64_t)E, (uint64_t)ldb, (uint64_t)sb, 64, bn / 2);
    else rc = make_tmap_3d(&tb, b, (uint64_t)N, (uint64_t)K, (uint64_t)E, (uint64_t)ldb, (uint64_t)sb, 64, kBlockK);
    if (rc) return rc;
    if (ep.mode == kOutBF16)
        rc = make_tmap_3d(&td, ep.d, (uint64_t)N, (uint64_t)M, (uint64_t)E, (uint64_t)ep.ldd, (uint64_t)sd, 64, 32, 2);
    else
        rc = make_tmap_3d(&td, ep.d, (uint64_t)N, (uint64_t)M, (uint64_t)E, (uint64_t)ep.ldd, (uint64_t)sd, 32, 32, 4);
    if (rc) return rc;
#define NXD_GG_LAYOUT(BN_, MODE_)                                                                              \
    switch (layout) {                                                                                          \
        case kLayoutNT: return launch_grouped<BN_, false, false, MODE_>(ta, tb, td, E, M, N, K, stream);       \
        case kLayoutNN: return launch_grouped<BN_, false, true, MODE_>(ta, tb, td, E, M, N, K, stream);        \
        default: return launch_grouped<BN_, true, true, MODE_>(ta, tb, td, E, M, N, K, stream);                \
    }
#define NXD_GG_MODE(BN_)                                                          \
    switch (ep.mode) {                                                            \
        case kOutBF16: NXD_GG_LAYOUT(BN_, kOutBF16)                               \
        case kOutF32: NXD_GG_LAYOUT(BN_, kOutF32)                                 \
        case kOutF32Acc: NXD_GG_LAYOUT(BN_, kOutF32Acc)                           \
        default: nxd_set_error("gemm_grouped: bad output mode %d", ep.mode); return -1; \
    }
    if (bn == 256) { NXD_GG_MODE(256) }
    NXD_GG_MODE(128)
#undef NXD_GG_MODE
#undef NXD_GG_LAYOUT
}

}  // namespace nxd_gemm

// Grouped (batched over experts) GEMM: E problems of one shape in ONE launch.  a / b / d point at problem 0, stride_*
// are the element distances between consecutive problems; layouts, pitches and out_mode as nxd_gemm_bf16.
NXD_API int nxd_gemm_grouped_bf16(int layout, const void* a, const void* b, void* d, int E, int M, int N, int K,
                                  long long lda, long long ldb, long long ldd, long long stride_a, long long stride_b,
                                  long long stride_d, int out_mode, void* stream) {
    nxd_gemm::EpilogueArgs ep{};
    ep.d = d;
    ep.ldd = ldd;
    ep.mode = out_mode;
    return nxd_gemm::run_grouped(layout, a, b, lda, ldb, stride_a, stride_b, ep, stride_d, E, M, N, K,
                                 (cudaStream_t)stream);
}

// C-ABI: layout 0 = nt, 1 = nn, 2 = tn (see header).  out_mode 0 = bf16, 1 = fp32, 2 = fp32 accumulate.
// lda / ldb / ldd are row pitches in elements of the row-major operands as they sit in memory.
NXD_API int nxd_gemm_bf16(int layout, const void* a, const void* b, void* d, int M, int N, int K, long long lda,
                          long long ldb, long long ldd, int out_mode, int block_n, void* stream) {
    nxd_gemm::EpilogueArgs ep{};
    ep.d = d;
    ep.ldd = ldd;
    ep.mode = out_mode;
    return nxd_gemm::run(layout, a, b, lda, ldb, ep, M, N, K, block_n, (cudaStream_t)stream);
}
