// Fused sequence-parallel ALL-GATHER -> GEMM in ONE kernel over NVLink peer memory.
//
// Reference: ColumnParallelLinear forward under sequence parallelism does `xm.all_gather(input)` and then
// `matmul(total_input, W^T)` as two dependent ops (src/neuronx_distributed/parallel_layers/layers.py:312-325); the
// RowParallel backward does the same with dY (mappings.py:307-308 + layers.py:348).  Here the gather is performed BY
// THE GEMM KERNEL while its tensor-core tiles run:
//
//   Y[tp*R, N] = concat_r(X_r)[tp*R, K] * W_op        X_r = rank r's [R, K] shard (bf16), W local
//
//   * the gathered buffer X_total lives in symmetric (peer-mapped, CUDA-IPC) memory on every rank;
//   * a few "communication clusters" of the grid (default 6 of 74 CTA pairs) run no MMA at all: one thread per CTA
//     drives the TMA engine -- cp.async.bulk of a 16 KB piece of the LOCAL shard into shared memory, then one
//     cp.async.bulk store of that piece into EVERY peer's X_total over NVLink (posted writes: bandwidth bound, no
//     round trip; an earlier pull version was limited by bytes-in-flight x NVLink latency) -- and, once the stores of
//     a 256-row chunk have landed, bumps that chunk's counter on every peer's signal pad (red.release.sys);
//   * the remaining clusters are the persistent tcgen05 GEMM of gemm_tcgen05.cu.  Their tile order follows the
//     arrival order (own slab first, read straight from the local shard), and the TMA producer polls the chunk
//     counter (ld.acquire.gpu) before it loads A tiles of a remote slab.  Math on slab d overlaps the transfer of
//     slab d+1; the gathered X_total is a by-product kept for the weight-gradient GEMM.
//
// Buffer reuse needs no extra barrier: X_total is double-buffered, counters are epoch-stamped, and rank A cannot
// finish call n+1 before rank B entered call n+1 (A needs B's rows of that call), i.e. before B finished with the
// buffer of call n, which is the one A's call n+2 will write.
#include "common.cuh"
#include "gemm_tcgen05.cuh"
#include "ptx.cuh"

namespace nxd_gemm {

constexpr int kMaxTp = 8;
constexpr int kChunkRows = kTileM;       // readiness granularity = one M tile (256 rows)
constexpr int kCommSub = 16384;          // bytes per TMA bulk op of the copy engine
constexpr int kCommBufs = 12;            // smem ring of the copy engine (192 KB in flight per CTA)

struct AgArgs {
    const bf16* x_local;          // [R, K] this rank's shard, dense
    bf16* xt_peer[kMaxTp];        // every rank's X_total [tp*R, K] for this call (peer-mapped; [rank] = local)
    uint32_t* sig_local;          // local signal pad
    uint32_t* sig_peer[kMaxTp];   // peers' signal pads
    int slot_chunks;              // flag index: chunk counters [src][chunk]   (tp * chunks_per_slab entries)
    int tp, rank, R;
    int comm_ctas;
    uint32_t epoch;
    unsigned long long* trace;    // optional: globaltimer stamps of a few events (debug)
    // expert-parallel dispatch (all-to-all -> grouped GEMM): x_local is [tp][R, K] (slab d goes to rank d
    // only), a slab holds R / expert_rows experts and expert e multiplies weight rows [e * w_expert_stride, ...)
    int expert_rows;              // 0 = all-gather mode
    int w_expert_stride;
    // NVLS (run at 2 and 8 GPUs): multicast address of X_total for this call.  One multimem.st reaches
    // every rank's copy through the switch, so 1/tp of the bytes leave the GPU compared with the per-peer pushes.
    bf16* mc_total;               // nullptr = per-peer TMA pushes
};

__device__ __forceinline__ unsigned long long gtime() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void red_release_sys_add_u32(uint32_t* p, uint32_t v) {
    asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_relaxed_sys_add_u32(uint32_t* p, uint32_t v) {
    asm volatile("red.relaxed.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t atom_add_release_gpu(uint32_t* p, uint32_t v) {
    uint32_t old;
    asm volatile("atom.add.release.gpu.global.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
    return old;
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void multimem_st_16(void* mc_addr, const uint4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr), "f"(__uint_as_float(v.x)),
                 "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
                 : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     ptx::smem_u32(dst)),
                 "l"(reinterpret_cast<uint64_t>(src)), "r"(bytes), "r"(ptx::smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(reinterpret_cast<uint64_t>(dst)),
                 "r"(ptx::smem_u32(src)), "r"(bytes)
                 : "memory");
}

// One continuous stream of <= 16 KB TMA copies (global -> smem ring -> global), driven by ONE thread.  src(j, ptr,
// bytes) describes item j; store(j, smem, bytes) issues its bulk store(s); done(j) is called for every item with
// (j + 1) % sig_every == 0 once its stores (and all older ones) have fully landed -- kLag store groups later, so the
// pipe never drains between signals.  Loads run (kCommBufs - kReadLag) items ahead of the stores.
template <typename SrcFn, typename StoreFn, typename DoneFn>
__device__ __forceinline__ void tma_stream(uint8_t* ring, uint64_t* bars, uint32_t& phases, int n_items, int sig_every,
                                           SrcFn src, StoreFn store, DoneFn done) {
    constexpr int kLag = 4;        // store groups allowed to be incomplete before a chunk is signalled
    constexpr int kReadLag = 4;    // store groups whose smem READ may still be pending when a buffer is refilled
    if (n_items <= 0) return;
    auto issue_load = [&](int j) {
        const uint8_t* p; uint32_t bytes;
        src(j, p, bytes);
        const int b = j % kCommBufs;
        ptx::mbar_arrive_expect_tx(&bars[b], bytes);
        bulk_g2s(ring + b * kCommSub, p, bytes, &bars[b]);
    };
    const int ahead = kCommBufs - kReadLag;
    for (int j = 0; j < min(n_items, ahead); ++j) issue_load(j);
    int next_signal = sig_every > 0 ? sig_every - 1 : n_items;
    for (int j = 0; j < n_items; ++j) {
        const int b = j % kCommBufs;
        ptx::mbar_wait(&bars[b], (phases >> b) & 1u);
        phases ^= (1u << b);
        const uint8_t* p; uint32_t bytes;
        src(j, p, bytes);
        store(j, ring + b * kCommSub, bytes);
        ptx::bulk_commit();
        if (j + ahead < n_items) {
            ptx::bulk_wait_read<kReadLag>();   // the buffer about to be refilled was stored >= kReadLag groups ago
            issue_load(j + ahead);
        }
        while (next_signal <= j - kLag) {
            ptx::bulk_wait<kLag>();
            done(next_signal);
            next_signal += sig_every;
        }
    }
    ptx::bulk_wait<0>();   // every byte is in global memory (local and peers)
    while (next_signal < n_items) {
        done(next_signal);
        next_signal += sig_every;
    }
}

template <int BN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
ag_gemm_kernel(const __grid_constant__ CUtensorMap tmap_a_local, const __grid_constant__ CUtensorMap tmap_a_total,
               const __grid_constant__ CUtensorMap tmap_b, const __grid_constant__ CUtensorMap tmap_d,
               const AgArgs ag, int M, int N, int K) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    using L = SmemLayout<BN>;
    uint8_t* smem = ptx::align_smem_1024(smem_raw);
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t cta_rank = ptx::cluster_ctarank();
    const int num_clusters_all = gridDim.x >> 1;
    const int comm_clusters = ag.comm_ctas >> 1;
    const int num_clusters = num_clusters_all - comm_clusters;   // GEMM clusters
    // the copy engine takes the LOWEST block indices so that it is scheduled first
    const int cluster_id = (int)(blockIdx.x >> 1) - comm_clusters;
    const int chunks_per_slab = ag.R / kChunkRows;
    const size_t slab_bytes = (size_t)ag.R * K * 2;

    if (cluster_id < 0 && ag.mc_total != nullptr && ag.expert_rows == 0) {
        // ============== communication CTAs, NVLS variant: all threads, 16-byte loads -> multicast stores ==============
        const int cid = (int)blockIdx.x;
        const size_t chunk_bytes = (size_t)kChunkRows * K * 2;
        const size_t piece = ((chunk_bytes / ag.comm_ctas) + 15) & ~size_t(15);
        const size_t p0 = min((size_t)cid * piece, chunk_bytes);
        const size_t p1 = min(p0 + piece, chunk_bytes);
        const int n16 = (int)((p1 - p0) >> 4);
        for (int c = 0; c < chunks_per_slab; ++c) {
            const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const uint8_t*>(ag.x_local) +
                                                              (size_t)c * chunk_bytes + p0);
            uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(ag.mc_total) + (size_t)ag.rank * slab_bytes +
                                                  (size_t)c * chunk_bytes + p0);
            for (int i = threadIdx.x; i < n16; i += blockDim.x) multimem_st_16(dst + i, __ldg(src + i));
            __syncthreads();
            if (threadIdx.x == 0) {
                __threadfence_system();
                for (int d = 1; d < ag.tp; ++d) {
                    const int peer = (ag.rank + d) % ag.tp;
                    red_relaxed_sys_add_u32(ag.sig_peer[peer] + ag.slot_chunks + ag.rank * chunks_per_slab + c, 1u);
                }
            }
        }
        return;
    }
    if (cluster_id < 0) {
        // =========================== communication CTAs (copy engine, no MMA) ===========================
        if (threadIdx.x != 0) return;
        const int cid = (int)blockIdx.x;                           // 0 .. comm_ctas-1
        uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kCommBufs * kCommSub);
        for (int i = 0; i < kCommBufs; ++i) ptx::mbar_init(&bars[i], 1);
        ptx::fence_barrier_init();
        uint32_t phases = 0;
        // per-chunk piece of this CTA: bytes [p0, p1) of every 256-row chunk of the LOCAL shard
        const size_t chunk_bytes = (size_t)kChunkRows * K * 2;
        const size_t piece = ((chunk_bytes / ag.comm_ctas) + 15) & ~size_t(15);
        const size_t p0 = min((size_t)cid * piece, chunk_bytes);
        const size_t p1 = min(p0 + piece, chunk_bytes);
        const int ipp = p1 > p0 ? (int)((p1 - p0 + kCommSub - 1) / kCommSub) : 0;      // items per piece
        if (ag.trace && cid == 0) ag.trace[0] = gtime();
        auto bump = [&](int c) {   // this CTA's piece of chunk c of MY shard has landed everywhere
            // ONE system-scope fence, then relaxed increments: a release per peer would pay the fence tp - 1 times
            fence_proxy_async_all();
            __threadfence_system();
            for (int d = 1; d < ag.tp; ++d) {
                const int peer = (ag.rank + d) % ag.tp;
                red_relaxed_sys_add_u32(ag.sig_peer[peer] + ag.slot_chunks + ag.rank * chunks_per_slab + c, 1u);
            }
        };
        if (ipp == 0) {  // more copy CTAs than 16-byte pieces: nothing to move, but the counters expect every CTA
            for (int c = 0; c < chunks_per_slab; ++c) bump(c);
        }
        if (ag.expert_rows == 0) {
            tma_stream(smem, bars, phases, ipp * chunks_per_slab, ipp,
                       [&](int j, const uint8_t*& src, uint32_t& bytes) {
                           const int c = j / ipp, i = j % ipp;
                           bytes = (uint32_t)min((size_t)kCommSub, p1 - p0 - (size_t)i * kCommSub);
                           src = reinterpret_cast<const uint8_t*>(ag.x_local) + (size_t)c * chunk_bytes + p0 + (size_t)i * kCommSub;
                       },
                       [&](int j, const uint8_t* sm, uint32_t bytes) {
                           const int c = j / ipp, i = j % ipp;
                           const size_t off = (size_t)ag.rank * slab_bytes + (size_t)c * chunk_bytes + p0 + (size_t)i * kCommSub;
                           // ring order: every rank starts with a different peer, so no link is a hot spot
                           for (int d = 1; d <= ag.tp; ++d) {
                               const int peer = (ag.rank + d) % ag.tp;   // d == tp: own X_total (by-product for wgrad)
                               bulk_s2g(reinterpret_cast<uint8_t*>(ag.xt_peer[peer]) + off, sm, bytes);
                           }
                       },
                       [&](int j_last) { bump(j_last / ipp); });
        } else {
            // all-to-all: item = (chunk, destination, piece); slab d of x_local goes to rank d only (own slab too: the
            // gathered buffer is kept for the weight gradient)
            const int per_chunk = ipp * ag.tp;
            tma_stream(smem, bars, phases, per_chunk * chunks_per_slab, per_chunk,
                       [&](int j, const uint8_t*& src, uint32_t& bytes) {
                           const int c = j / per_chunk, r = j % per_chunk, d = r / ipp, i = r % ipp;
                           const int peer = (ag.rank + 1 + d) % ag.tp;
                           bytes = (uint32_t)min((size_t)kCommSub, p1 - p0 - (size_t)i * kCommSub);
                           src = reinterpret_cast<const uint8_t*>(ag.x_local) + (size_t)peer * slab_bytes +
                                 (size_t)c * chunk_bytes + p0 + (size_t)i * kCommSub;
                       },
                       [&](int j, const uint8_t* sm, uint32_t bytes) {
                           const int c = j / per_chunk, r = j % per_chunk, d = r / ipp, i = r % ipp;
                           const int peer = (ag.rank + 1 + d) % ag.tp;
                           const size_t off = (size_t)ag.rank * slab_bytes + (size_t)c * chunk_bytes + p0 + (size_t)i * kCommSub;
                           bulk_s2g(reinterpret_cast<uint8_t*>(ag.xt_peer[peer]) + off, sm, bytes);
                       },
                       [&](int j_last) { bump(j_last / per_chunk); });
        }
        if (ag.trace && cid == 0) ag.trace[1] = gtime();
        return;
    }

    // =========================== GEMM CTAs (same pipeline as gemm_pair_kernel) ===========================
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarrierOffset);
    constexpr int kStages = L::kStages;
    uint64_t* empty_bar = full_bar + kStages;
    uint64_t* tmem_full_bar = empty_bar + kStages;
    uint64_t* tmem_empty_bar = tmem_full_bar + kAccStages;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty_bar + kAccStages);
    const bool is_leader = cta_rank == 0;
    const int m_tiles_per_slab = chunks_per_slab;            // kChunkRows == kTileM
    const int num_n_tiles = (N + BN - 1) / BN;
    const int num_tiles = ag.tp * m_tiles_per_slab * num_n_tiles;
    const int num_k_blocks = (K + kBlockK - 1) / kBlockK;

    // tile -> (slab, m tile inside the slab, n tile) in ARRIVAL order; n fastest so that A rows are shared in L2.
    // Own slab first (no wait).  Every peer pushes chunk c of its shard to all ranks at the same time, so the remote
    // rows arrive chunk-major: chunk 0 of every slab, then chunk 1 of every slab, ...
    auto tile_coords_ag = [&](int tile, int& slab, int& mt, int& tn) {
        tn = tile % num_n_tiles;
        const int mg = tile / num_n_tiles;
        if (mg < m_tiles_per_slab) {
            mt = mg;
            slab = ag.rank;
        } else {
            const int idx = mg - m_tiles_per_slab;
            mt = idx / (ag.tp - 1);
            slab = (ag.rank + 1 + idx % (ag.tp - 1)) % ag.tp;
        }
    };

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tensormap(&tmap_a_local);
        ptx::prefetch_tensormap(&tmap_a_total);
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
                int slab, mt, tn;
                tile_coords_ag(tile, slab, mt, tn);
                const bool own = slab == ag.rank;
                if (!own) {
                    // wait until every copy CTA of the SOURCE rank delivered its piece of this 256-row chunk
                    const uint32_t* cnt = ag.sig_local + ag.slot_chunks + slab * chunks_per_slab + mt;
                    const uint32_t want = ag.epoch * (uint32_t)ag.comm_ctas;
                    const bool tr = ag.trace && cluster_id == 0 && is_leader;
                    if (tr && ag.trace[8] == 0) ag.trace[8] = gtime();
                    while ((int32_t)(ld_acquire_sys(cnt) - want) < 0) { __nanosleep(32); }
                    if (tr && ag.trace[9] == 0) ag.trace[9] = gtime();
                    if (tr) ag.trace[10] = gtime();
                    fence_proxy_async_all();
                }
                const CUtensorMap* ma = own ? &tmap_a_local : &tmap_a_total;
                // dispatch mode: x_local holds one slab per destination, the own one at row rank * R
                const int m0 = ((own && ag.expert_rows == 0) ? 0 : slab * ag.R) + mt * kTileM + (int)cta_rank * kBlockM;
                const int n0 = tn * BN + (int)cta_rank * (BN / 2);
                const int w_off = ag.expert_rows > 0 ? (mt * kTileM / ag.expert_rows) * ag.w_expert_stride : 0;
                for (int kb = 0; kb < num_k_blocks; ++kb) {
                    ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * L::kStageBytes;
                    uint8_t* sb = sa + L::kABytes;
                    const int k0 = kb * kBlockK;
                    if (is_leader) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * L::kStageBytes);
                    ptx::tma_load_2d_pair(ma, &full_bar[stage], sa, k0, m0);
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
        for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
            int slab, mt, tn;
            tile_coords_ag(tile, slab, mt, tn);
            const int row0 = slab * ag.R + mt * kTileM + (int)cta_rank * kBlockM + q * 32;
            const int n0 = tn * BN;
            ptx::mbar_wait(&tmem_full_bar[acc], acc_phase);
            ptx::tcgen05_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN;
#pragma unroll 1
            for (int c = 0; c < BN / 64; ++c) {
                if (lane == 0) ptx::bulk_wait_read<1>();
                __syncwarp();
                uint32_t r[64];
                ptx::tmem_ld_32x32b_x32(taddr + c * 64, r);
                ptx::tmem_ld_32x32b_x32(taddr + c * 64 + 32, r + 32);
                ptx::tmem_ld_wait();
                if (c == BN / 64 - 1) {
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
                    if (col0 < N && row0 < M) ptx::tma_store_2d(&tmap_d, stg + buf * (32 * 128), col0, row0);
                    ptx::bulk_commit();
                }
                buf ^= 1;
            }
            if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
        }
        if (lane == 0) ptx::bulk_wait<0>();
    }
    ptx::tcgen05_fence_before();
    ptx::cluster_sync();
    if (warp == 2) ptx::tmem_dealloc<2>(tmem_base, L::kTmemCols);
    if (ag.trace && cluster_id == 0 && is_leader && threadIdx.x == 0) ag.trace[11] = gtime();
}

}  // namespace nxd_gemm

// layout 0 "nt": W = [N, K];  1 "nn": W = [K, N].  x_local [R, K] dense, y [tp*R, N] (pitch ldy).
// xt_ptrs / sig_ptrs: HOST arrays of tp device pointers (own entry included): every rank's symmetric X_total buffer
// [tp*R, K] for this call, and signal pad.  R % 256 == 0, K % 64 == 0, N % 8 == 0.  trace: optional 16 x u64 (device).
static int ag_gemm_impl(int layout, const void* x_local, const void* w, void* y, int R, int N, int K, long long ldw,
                        long long ldy, void* const* xt_ptrs, void* const* sig_ptrs, int tp, int rank, int slot_chunks,
                        unsigned epoch, int comm_ctas, int block_n, void* trace, int expert_rows, void* stream,
                        void* mc_total = nullptr) {
    using namespace nxd_gemm;
    NXD_REQUIRE(expert_rows == 0 || (expert_rows % kChunkRows == 0 && R % expert_rows == 0),
                "a2a_gemm: rows per expert (%d) must be a multiple of %d and divide the slab (%d)", expert_rows, kChunkRows, R);
    const int experts = expert_rows > 0 ? R / expert_rows : 1;
    NXD_REQUIRE(layout == 0 || layout == 1, "ag_gemm: layout must be 0 (nt) or 1 (nn)");
    NXD_REQUIRE(tp >= 2 && tp <= kMaxTp && rank >= 0 && rank < tp, "ag_gemm: bad tp/rank %d/%d", tp, rank);
    NXD_REQUIRE(R > 0 && R % kChunkRows == 0, "ag_gemm: rows per rank (%d) must be a multiple of %d", R, kChunkRows);
    NXD_REQUIRE(K % 64 == 0 && N % 8 == 0 && ldw % 8 == 0 && ldy % 8 == 0, "ag_gemm: K %% 64, N %% 8 required");
    NXD_REQUIRE(comm_ctas >= 2 && comm_ctas % 2 == 0, "ag_gemm: comm_ctas must be even and >= 2");
    const int M = tp * R;
    const int bn = block_n > 0 ? block_n : 256;
    NXD_REQUIRE(bn == 128 || bn == 256, "ag_gemm: block_n must be 128 or 256");
    AgArgs ag{};
    ag.x_local = (const bf16*)x_local;
    for (int r = 0; r < tp; ++r) {
        ag.xt_peer[r] = (bf16*)xt_ptrs[r];
        ag.sig_peer[r] = (uint32_t*)sig_ptrs[r];
    }
    ag.sig_local = (uint32_t*)sig_ptrs[rank];
    ag.slot_chunks = slot_chunks;
    ag.tp = tp; ag.rank = rank; ag.R = R; ag.comm_ctas = comm_ctas; ag.epoch = epoch;
    ag.trace = (unsigned long long*)trace;
    ag.expert_rows = expert_rows;
    ag.mc_total = (bf16*)mc_total;
    ag.w_expert_stride = layout == 0 ? N : K;     // stacked expert weights: [E * N, K] (nt) or [E * K, N] (nn)
    CUtensorMap ta_l, ta_t, tb, td;
    int rc;
    if ((rc = make_tmap_2d(&ta_l, x_local, (uint64_t)K, (uint64_t)(expert_rows > 0 ? M : R), (uint64_t)K, 64, kBlockM)))
        return rc;
    if ((rc = make_tmap_2d(&ta_t, xt_ptrs[rank], (uint64_t)K, (uint64_t)M, (uint64_t)K, 64, kBlockM))) return rc;
    if (layout == 0) rc = make_tmap_2d(&tb, w, (uint64_t)K, (uint64_t)N * experts, (uint64_t)ldw, 64, bn / 2);
    else rc = make_tmap_2d(&tb, w, (uint64_t)N, (uint64_t)K * experts, (uint64_t)ldw, 64, kBlockK);
    if (rc) return rc;
    if ((rc = make_tmap_2d(&td, y, (uint64_t)N, (uint64_t)M, (uint64_t)ldy, 64, 32, 2))) return rc;
    const int clusters = nxd_num_sms() / 2;
    NXD_REQUIRE(comm_ctas / 2 < clusters, "ag_gemm: too many communication CTAs");
    cudaStream_t st = (cudaStream_t)stream;
#define NXD_AG_LAUNCH(BN_, BMN_)                                                                              \
    {                                                                                                         \
        using L = SmemLayout<BN_>;                                                                            \
        static cudaError_t once = cudaFuncSetAttribute(ag_gemm_kernel<BN_, BMN_>,                             \
                                                       cudaFuncAttributeMaxDynamicSharedMemorySize, L::kSmemBytes); \
        NXD_CHECK_CUDA(once);                                                                                 \
        ag_gemm_kernel<BN_, BMN_><<<dim3(clusters * 2), dim3(kThreads), L::kSmemBytes, st>>>(ta_l, ta_t, tb, td, ag, M, \
                                                                                             N, K);           \
    }
    if (bn == 256) {
        if (layout == 0) NXD_AG_LAUNCH(256, false) else NXD_AG_LAUNCH(256, true)
    } else {
        if (layout == 0) NXD_AG_LAUNCH(128, false) else NXD_AG_LAUNCH(128, true)
    }
#undef NXD_AG_LAUNCH
    NXD_LAUNCH_CHECK();
    return 0;
}

NXD_API int nxd_ag_gemm_bf16(int layout, const void* x_local, const void* w, void* y, int R, int N, int K,
                             long long ldw, long long ldy, void* const* xt_ptrs, void* const* sig_ptrs, int tp,
                             int rank, int slot_chunks, unsigned epoch, int comm_ctas, int block_n, void* trace,
                             void* stream) {
    return ag_gemm_impl(layout, x_local, w, y, R, N, K, ldw, ldy, xt_ptrs, sig_ptrs, tp, rank, slot_chunks, epoch,
                        comm_ctas, block_n, trace, 0, stream);
}

// Expert-parallel DISPATCH fused with the first expert GEMM (reference: modules/moe/experts.py:121-128
// `enter_expert_parallel_region` all-to-all, then the expert matmul): x_local is [ep][E_loc, C, K] (slab d = the
// tokens this rank routed to the experts of rank d), every rank ends up with X_total = [ep(source)][E_loc, C, K] in
// its symmetric buffer and y[(s, e, c), :] = X_total[s, e, c, :] * W_e.  W: stacked expert weights, layout 0 = [E_loc * N, K]
// (each expert [N, K]), 1 = [E_loc * K, N].  R = E_loc * C rows per slab, C % 256 == 0.  Other arguments as
// nxd_ag_gemm_bf16 (tp = ep).
NXD_API int nxd_a2a_gemm_bf16(int layout, const void* x_local, const void* w, void* y, int R, int C, int N, int K,
                              long long ldw, long long ldy, void* const* xt_ptrs, void* const* sig_ptrs, int ep,
                              int rank, int slot_chunks, unsigned epoch, int comm_ctas, int block_n, void* trace,
                              void* stream) {
    NXD_REQUIRE(C > 0, "a2a_gemm: capacity must be positive");
    return ag_gemm_impl(layout, x_local, w, y, R, N, K, ldw, ldy, xt_ptrs, sig_ptrs, ep, rank, slot_chunks, epoch,
                        comm_ctas, block_n, trace, C, stream);
}

// EXPERIMENTAL (compiled, not yet run on hardware): nxd_ag_gemm_bf16 whose copy CTAs write the local shard ONCE to the
// NVLS multicast address `mc_total` of the gathered buffer (the same buffer xt_ptrs point to, mapped through
// cuMulticast* / torch.distributed._symmetric_memory) instead of pushing it to every peer.
NXD_API int nxd_ag_gemm_bf16_nvls(int layout, const void* x_local, const void* w, void* y, int R, int N, int K,
                                  long long ldw, long long ldy, void* const* xt_ptrs, void* const* sig_ptrs, int tp,
                                  int rank, int slot_chunks, unsigned epoch, int comm_ctas, int block_n, void* mc_total,
                                  void* stream) {
    NXD_REQUIRE(mc_total != nullptr, "ag_gemm_nvls: multicast pointer required");
    return ag_gemm_impl(layout, x_local, w, y, R, N, K, ldw, ldy, xt_ptrs, sig_ptrs, tp, rank, slot_chunks, epoch,
                        comm_ctas, block_n, nullptr, 0, stream, mc_total);
}
