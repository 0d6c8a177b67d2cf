// Fused GEMM -> REDUCE-SCATTER (sequence parallel) in ONE kernel over NVLink peer memory.
//
// Reference: RowParallelLinear forward computes the partial product and then calls
// `reduce_scatter_to_sequence_parallel_region` (src/neuronx_distributed/parallel_layers/layers.py:785-795,
// mappings.py:118-147); ColumnParallel backward does the same for the input gradient (layers.py:348,394-407).
//
//   Y_d[R, N] = sum_r ( X_r[tp*R, Kloc] * W_r_op )[rows d*R .. (d+1)*R)          rank d keeps Y_d
//
// One persistent tcgen05 kernel per rank:
//   * tiles are visited destination by destination in ring order (rank+1, rank+2, ..., own slab LAST), so the
//     partial sums that have to cross NVLink leave first and the transfer overlaps the remaining math;
//   * the epilogue does not write a local C: each 32 x 128 B box goes straight from shared memory to the DESTINATION
//     rank's staging buffer (slot = source rank) with a TMA store on a peer-mapped tensor map; when all boxes of a
//     256-row tile have landed, the warp bumps a counter on the destination's signal pad (red.release.sys);
//   * after its last tile every CTA turns into a reducer: it waits (ld.acquire.sys) until all tp sources have
//     delivered a row tile of the own slab, sums the tp staged partials in fp32 and writes the bf16 result.
// Staging and counters are double-buffered / epoch-stamped exactly like the all-gather kernel, so no barrier
// separates consecutive calls.
#include "common.cuh"
#include "gemm_tcgen05.cuh"
#include "ptx.cuh"

namespace nxd_gemm {

constexpr int kRsMaxTp = 8;

struct RsTmaps {
    CUtensorMap d[kRsMaxTp];   // d[r]: rank r's staging slot for partials coming from THIS rank ([R, N] bf16)
};

struct RsArgs {
    const bf16* stage_local;        // local staging [tp][R][N] (this call's buffer)
    bf16* out;                      // [R, N] result (pitch ldo)
    long long ldo;
    uint32_t* sig_local;
    uint32_t* sig_peer[kRsMaxTp];
    int slot_cnt;                   // counters [src][m tile]  (tp * m_tiles_per_slab entries)
    int tp, rank, R;
    uint32_t epoch;
    unsigned long long* trace;      // optional: globaltimer stamps (debug)
    // expert-parallel COMBINE (grouped GEMM -> all-to-all): row block `dest` of x holds the tokens that came
    // from rank `dest`; its result tiles go back to that rank and NOTHING is summed (the staging buffer [src][R][N] of
    // the destination IS its output [E, C, N]).  expert e of a slab uses weight rows [e * w_expert_stride, ...).
    int expert_rows;                // 0 = reduce-scatter mode
    int w_expert_stride;
};

__device__ __forceinline__ unsigned long long rs_gtime() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// increment after an explicit __threadfence_system(): fence + relaxed atomic = release, but the fence is paid once
__device__ __forceinline__ void red_relaxed_sys_add(uint32_t* p, uint32_t v) {
    asm volatile("red.relaxed.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

template <int BN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm_rs_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const __grid_constant__ RsTmaps tmaps_d, const RsArgs rs, int M, int N, int K) {
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
    const int m_tiles_per_slab = rs.R / kTileM;
    const int num_n_tiles = (N + BN - 1) / BN;
    const int num_tiles = rs.tp * m_tiles_per_slab * num_n_tiles;
    const int num_k_blocks = (K + kBlockK - 1) / kBlockK;
    const int cluster_id = blockIdx.x >> 1;
    const int num_clusters = gridDim.x >> 1;
    const bool tr = rs.trace != nullptr && blockIdx.x == 0;
    if (tr && threadIdx.x == 0) rs.trace[0] = rs_gtime();

    // tile -> (destination rank in ring order, own slab last; m tile inside the slab; n tile)
    auto tile_coords_rs = [&](int tile, int& dest, int& mt, int& tn) {
        tn = tile % num_n_tiles;
        const int mg = tile / num_n_tiles;
        mt = mg % m_tiles_per_slab;
        dest = (rs.rank + 1 + mg / m_tiles_per_slab) % rs.tp;
    };

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tensormap(&tmap_a);
        ptx::prefetch_tensormap(&tmap_b);
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < kStages; ++i) {
            ptx::mbar_init(&full_bar[i], 1);
            ptx::mbar_init(&empty_bar[i], 1);
        }
        for (int i = 0; i < kAccStages; ++i) {
            ptx::mbar_init(&tmem_full_bar[i], 1);
            ptx::mbar_init(&tmem_empty_bar[i], 2 * kEpilogueWarps);
        }
        ptx::fence_barrier_init();
    }
    if (warp == 2) ptx::tmem_alloc<2>(tmem_ptr_smem, L::kTmemCols);
    ptx::tcgen05_fence_before();
    ptx::cluster_sync();
    ptx::tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
                int dest, mt, tn;
                tile_coords_rs(tile, dest, mt, tn);
                const int m0 = dest * rs.R + mt * kTileM + (int)cta_rank * kBlockM;
                const int n0 = tn * BN + (int)cta_rank * (BN / 2);
                const int w_off = rs.expert_rows > 0 ? (mt * kTileM / rs.expert_rows) * rs.w_expert_stride : 0;
                for (int kb = 0; kb < num_k_blocks; ++kb) {
                    ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * L::kStageBytes;
                    uint8_t* sb = sa + L::kABytes;
                    const int k0 = kb * kBlockK;
                    if (is_leader) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * L::kStageBytes);
                    ptx::tma_load_2d_pair(&tmap_a, &full_bar[stage], sa, k0, m0);
                    if constexpr (!B_MN) {
                        ptx::tma_load_2d_pair(&tmap_b, &full_bar[stage], sb, k0, n0 + w_off);
                    } else {
#pragma unroll
                        for (int g = 0; g < (BN / 2) / 64; ++g)
                            ptx::tma_load_2d_pair(&tmap_b, &full_bar[stage], sb + g * (kBlockK * 128), n0 + g * 64,
                                                  k0 + w_off);
                    }
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (is_leader && lane == 0) {
            constexpr uint32_t idesc = ptx::make_idesc_bf16(kTileM, BN, 0, B_MN ? 1 : 0);
            constexpr uint32_t b_lbo = B_MN ? kBlockK * 128 : 16;
            constexpr uint32_t b_kstep = B_MN ? 2048 : 32;
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
                        const uint64_t da = ptx::make_smem_desc(sa + k * 32, 16, 1024);
                        const uint64_t db = ptx::make_smem_desc(sb + k * b_kstep, b_lbo, 1024);
                        ptx::umma_bf16<2>(tmem_d, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    ptx::umma_commit_pair(&empty_bar[stage], 0b11);
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
                ptx::umma_commit_pair(&tmem_full_bar[acc], 0b11);
                if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else if (warp >= kFirstEpilogueWarp) {
        const int q = warp & 3;
        uint8_t* stg = smem + L::kStagingOffset + q * L::kStagingBytesPerWarp;
        int buf = 0;
        int acc = 0;
        uint32_t acc_phase = 0;
        constexpr int kGroupsPerTile = BN / 64;
        for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
            int dest, mt, tn;
            tile_coords_rs(tile, dest, mt, tn);
            const int row_in_slab = mt * kTileM + (int)cta_rank * kBlockM + q * 32;
            const int n0 = tn * BN;
            ptx::mbar_wait(&tmem_full_bar[acc], acc_phase);
            ptx::tcgen05_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN;
#pragma unroll 1
            for (int c = 0; c < kGroupsPerTile; ++c) {
                if (lane == 0) ptx::bulk_wait_read<1>();
                __syncwarp();
                uint32_t r[64];
                ptx::tmem_ld_32x32b_x32(taddr + c * 64, r);
                ptx::tmem_ld_32x32b_x32(taddr + c * 64 + 32, r + 32);
                ptx::tmem_ld_wait();
                if (c == kGroupsPerTile - 1) {
                    ptx::tcgen05_fence_before();
                    __syncwarp();
                    if (lane == 0) ptx::mbar_arrive_cluster(&tmem_empty_bar[acc], 0);
                }
                uint8_t* dst = stg + buf * (32 * 128) + lane * 128;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    bf162 p0 = __floats2bfloat162_rn(__uint_as_float(r[8 * j]), __uint_as_float(r[8 * j + 1]));
                    bf162 p1 = __floats2bfloat162_rn(__uint_as_float(r[8 * j + 2]), __uint_as_float(r[8 * j + 3]));
                    bf162 p2 = __floats2bfloat162_rn(__uint_as_float(r[8 * j + 4]), __uint_as_float(r[8 * j + 5]));
                    bf162 p3 = __floats2bfloat162_rn(__uint_as_float(r[8 * j + 6]), __uint_as_float(r[8 * j + 7]));
                    *reinterpret_cast<uint4*>(dst + ((j ^ (lane & 7)) << 4)) =
                        make_uint4(*reinterpret_cast<uint32_t*>(&p0), *reinterpret_cast<uint32_t*>(&p1),
                                   *reinterpret_cast<uint32_t*>(&p2), *reinterpret_cast<uint32_t*>(&p3));
                }
                ptx::fence_proxy_async();
                __syncwarp();
                if (lane == 0) {
                    const int col0 = n0 + c * 64;
                    // box into the destination's staging slot for THIS source rank (peer HBM over NVLink, or local)
                    if (col0 < N) ptx::tma_store_2d(&tmaps_d.d[dest], stg + buf * (32 * 128), col0, row_in_slab);
                    ptx::bulk_commit();
                }
                buf ^= 1;
            }
            if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
        }
        if (tr && warp == kFirstEpilogueWarp && lane == 0) rs.trace[1] = rs_gtime();   // last tile's boxes issued
        if (lane == 0) {
            // ONE system fence per warp for the whole GEMM phase, then a relaxed increment per tile: a fence per tile
            // drains the in-flight NVLink stores every time (tools/bench_p2p.py: 4x lower push bandwidth), and the
            // consumers -- the reduction below, on CTAs that are busy with GEMM tiles until now anyway -- gain nothing
            // from earlier signals.
            ptx::bulk_wait<0>();
            if (tr && warp == kFirstEpilogueWarp) rs.trace[2] = rs_gtime();            // ... and landed
            asm volatile("fence.proxy.async;" ::: "memory");
            __threadfence_system();
            for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
                int dest, mt, tn;
                tile_coords_rs(tile, dest, mt, tn);
                red_relaxed_sys_add(rs.sig_peer[dest] + rs.slot_cnt + rs.rank * m_tiles_per_slab + mt, 1u);
            }
        }
    }
    ptx::tcgen05_fence_before();
    ptx::cluster_sync();
    if (warp == 2) ptx::tmem_dealloc<2>(tmem_base, L::kTmemCols);

    // =========================== reduction of the own slab ===========================
    // work item = (m tile of the own slab, 64-column group); every source contributes num_n_tiles tiles x 8 epilogue
    // warps to the counter of an m tile
    const int col_groups = (N + 63) / 64;
    const int items = m_tiles_per_slab * col_groups;
    const uint32_t want = rs.epoch * (uint32_t)(num_n_tiles * 2 * kEpilogueWarps);
    if (tr && threadIdx.x == 0) rs.trace[3] = rs_gtime();   // GEMM phase of this CTA over
    if (rs.expert_rows > 0) {
        // combine mode: nothing to add -- the kernel only has to outlive the arrival of every source's tiles, so that
        // its completion means "the local output is complete"
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < rs.tp * m_tiles_per_slab; i += gridDim.x * blockDim.x) {
            const uint32_t* cnt = rs.sig_local + rs.slot_cnt + i;
            while ((int32_t)(ld_acquire_sys_u32(cnt) - want) < 0) { __nanosleep(64); }
        }
        return;
    }
    for (int item = blockIdx.x; item < items; item += gridDim.x) {
        const int mt = item / col_groups, cg = item % col_groups;
        if (threadIdx.x < rs.tp) {
            const uint32_t* cnt = rs.sig_local + rs.slot_cnt + threadIdx.x * m_tiles_per_slab + mt;
            while ((int32_t)(ld_acquire_sys_u32(cnt) - want) < 0) { __nanosleep(64); }
        }
        __syncthreads();
        if (tr && threadIdx.x == 0 && item == (int)blockIdx.x) rs.trace[4] = rs_gtime();   // first item's inputs complete
        // 256 rows x 64 columns: 8 consecutive threads cover one 128-byte row segment, 32 rows per pass.  All 8 passes
        // of one source are loaded back to back (8 independent 16-byte loads in flight per thread) before they are
        // accumulated: the loop is latency bound otherwise.
        const int col = cg * 64 + (threadIdx.x & 7) * 8;
        if (col < N) {
            constexpr int kPasses = kTileM / 32;
            float acc8[kPasses][8];
#pragma unroll
            for (int pass = 0; pass < kPasses; ++pass)
#pragma unroll
                for (int e = 0; e < 8; ++e) acc8[pass][e] = 0.f;
            const int row_base = mt * kTileM + (threadIdx.x >> 3);
            // two sources per round: 16 independent 16-byte loads in flight per thread
            for (int s = 0; s < rs.tp; s += 2) {
                bf16x8 pk[2][kPasses];
                const int ns = min(2, rs.tp - s);
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    if (u < ns) {
#pragma unroll
                        for (int pass = 0; pass < kPasses; ++pass) {
                            const bf16* src = rs.stage_local + ((size_t)(s + u) * rs.R + row_base + pass * 32) * N + col;
                            asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];"
                                         : "=r"(reinterpret_cast<uint32_t*>(&pk[u][pass])[0]),
                                           "=r"(reinterpret_cast<uint32_t*>(&pk[u][pass])[1]),
                                           "=r"(reinterpret_cast<uint32_t*>(&pk[u][pass])[2]),
                                           "=r"(reinterpret_cast<uint32_t*>(&pk[u][pass])[3])
                                         : "l"(src));
                        }
                    }
                }
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    if (u < ns) {
#pragma unroll
                        for (int pass = 0; pass < kPasses; ++pass) {
                            float f[8];
                            unpack8(pk[u][pass], f);
#pragma unroll
                            for (int e = 0; e < 8; ++e) acc8[pass][e] += f[e];
                        }
                    }
                }
            }
#pragma unroll
            for (int pass = 0; pass < kPasses; ++pass)
                *reinterpret_cast<bf16x8*>(rs.out + (size_t)(row_base + pass * 32) * rs.ldo + col) = pack8(acc8[pass]);
        }
        __syncthreads();
    }
    if (tr && threadIdx.x == 0) rs.trace[5] = rs_gtime();
}

}  // namespace nxd_gemm

// layout 0 "nt": W = [N, K];  1 "nn": W = [K, N].  x [tp*R, K] (pitch ldx), out [R, N] (pitch ldo).
// stage_ptrs / sig_ptrs: HOST arrays of tp device pointers: every rank's staging buffer for THIS call
// ([tp][R][N] bf16, slot index = source rank) and signal pad.  R % 256 == 0, N % 8 == 0.
static int gemm_rs_impl(int layout, const void* x, const void* w, void* out, int R, int N, int K, long long ldx,
                        long long ldw, long long ldo, void* const* stage_ptrs, void* const* sig_ptrs, int tp, int rank,
                        int slot_cnt, unsigned epoch, int block_n, void* trace, int expert_rows, void* stream) {
    using namespace nxd_gemm;
    NXD_REQUIRE(expert_rows == 0 || (expert_rows % kTileM == 0 && R % expert_rows == 0),
                "gemm_a2a: rows per expert (%d) must be a multiple of %d and divide the slab (%d)", expert_rows, kTileM, R);
    const int experts = expert_rows > 0 ? R / expert_rows : 1;
    NXD_REQUIRE(layout == 0 || layout == 1, "gemm_rs: layout must be 0 (nt) or 1 (nn)");
    NXD_REQUIRE(tp >= 2 && tp <= kRsMaxTp && rank >= 0 && rank < tp, "gemm_rs: bad tp/rank %d/%d", tp, rank);
    NXD_REQUIRE(R > 0 && R % kTileM == 0, "gemm_rs: rows per rank (%d) must be a multiple of %d", R, kTileM);
    NXD_REQUIRE(N % 8 == 0 && ldx % 8 == 0 && ldw % 8 == 0 && ldo % 8 == 0, "gemm_rs: N, pitches must be multiples of 8");
    const int M = tp * R;
    const int bn = block_n > 0 ? block_n : 256;
    NXD_REQUIRE(bn == 128 || bn == 256, "gemm_rs: block_n must be 128 or 256");
    RsArgs rs{};
    rs.stage_local = (const bf16*)stage_ptrs[rank];
    rs.out = (bf16*)out;
    rs.ldo = ldo;
    rs.sig_local = (uint32_t*)sig_ptrs[rank];
    for (int r = 0; r < tp; ++r) rs.sig_peer[r] = (uint32_t*)sig_ptrs[r];
    rs.slot_cnt = slot_cnt; rs.tp = tp; rs.rank = rank; rs.R = R; rs.epoch = epoch;
    rs.trace = (unsigned long long*)trace;
    rs.expert_rows = expert_rows;
    rs.w_expert_stride = layout == 0 ? N : K;
    CUtensorMap ta, tb;
    RsTmaps td;
    int rc;
    if ((rc = make_tmap_2d(&ta, x, (uint64_t)K, (uint64_t)M, (uint64_t)ldx, 64, kBlockM))) return rc;
    if (layout == 0) rc = make_tmap_2d(&tb, w, (uint64_t)K, (uint64_t)N * experts, (uint64_t)ldw, 64, bn / 2);
    else rc = make_tmap_2d(&tb, w, (uint64_t)N, (uint64_t)K * experts, (uint64_t)ldw, 64, kBlockK);
    if (rc) return rc;
    for (int r = 0; r < tp; ++r) {
        const bf16* base = (const bf16*)stage_ptrs[r] + (size_t)rank * R * N;   // my slot on rank r
        if ((rc = make_tmap_2d(&td.d[r], base, (uint64_t)N, (uint64_t)R, (uint64_t)N, 64, 32, 2))) return rc;
    }
    for (int r = tp; r < kRsMaxTp; ++r) td.d[r] = td.d[0];
    const int clusters = nxd_num_sms() / 2;
    cudaStream_t st = (cudaStream_t)stream;
#define NXD_RS_LAUNCH(BN_, BMN_)                                                                              \
    {                                                                                                         \
        using L = SmemLayout<BN_>;                                                                            \
        static cudaError_t once = cudaFuncSetAttribute(gemm_rs_kernel<BN_, BMN_>,                             \
                                                       cudaFuncAttributeMaxDynamicSharedMemorySize, L::kSmemBytes); \
        NXD_CHECK_CUDA(once);                                                                                 \
        gemm_rs_kernel<BN_, BMN_><<<dim3(clusters * 2), dim3(kThreads), L::kSmemBytes, st>>>(ta, tb, td, rs, M, N, K); \
    }
    if (bn == 256) {
        if (layout == 0) NXD_RS_LAUNCH(256, false) else NXD_RS_LAUNCH(256, true)
    } else {
        if (layout == 0) NXD_RS_LAUNCH(128, false) else NXD_RS_LAUNCH(128, true)
    }
#undef NXD_RS_LAUNCH
    NXD_LAUNCH_CHECK();
    return 0;
}

NXD_API int nxd_gemm_rs_bf16(int layout, const void* x, const void* w, void* out, int R, int N, int K, long long ldx,
                             long long ldw, long long ldo, void* const* stage_ptrs, void* const* sig_ptrs, int tp, int rank,
                             int slot_cnt, unsigned epoch, int block_n, void* trace, void* stream) {
    return gemm_rs_impl(layout, x, w, out, R, N, K, ldx, ldw, ldo, stage_ptrs, sig_ptrs, tp, rank, slot_cnt, epoch,
                        block_n, trace, 0, stream);
}

// Second expert GEMM fused with the expert-parallel COMBINE (reference: expert matmul then
// `exit_expert_parallel_region` all-to-all, modules/moe/experts.py:145-152): x is [ep(source)][E_loc, C, K], slab s
// holds the tokens that rank s dispatched here; tile results are TMA-stored straight into rank s's output
// [ep(owner)][E_loc, C, N] = [E, C, N] (out_ptrs: every rank's symmetric output buffer of THIS call).  When the kernel
// completes, out_ptrs[rank] is complete.  W: stacked expert weights as in nxd_a2a_gemm_bf16.  C % 256 == 0.
NXD_API int nxd_gemm_a2a_bf16(int layout, const void* x, const void* w, int R, int C, int N, int K, long long ldx,
                              long long ldw, void* const* out_ptrs, void* const* sig_ptrs, int ep, int rank, int slot_cnt,
                              unsigned epoch, int block_n, void* trace, void* stream) {
    NXD_REQUIRE(C > 0, "gemm_a2a: capacity must be positive");
    return gemm_rs_impl(layout, x, w, out_ptrs[rank], R, N, K, ldx, ldw, (long long)N, out_ptrs, sig_ptrs, ep, rank, slot_cnt,
                        epoch, block_n, trace, C, stream);
}
