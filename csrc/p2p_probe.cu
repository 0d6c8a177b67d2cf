// NVLink push-bandwidth probe (tooling, no counterpart in the reference): measures what a kernel on THIS GPU can
// sustain when it copies a local buffer into `n_dst` peers' symmetric memory, for the three ways a fused
// compute+collective kernel can issue the writes.  tools/bench_p2p.py sweeps it; the numbers pick the copy-engine
// geometry of fused_ag_gemm.cu / fused_gemm_rs.cu.
//   mode 0: one thread per CTA, cp.async.bulk global->smem ring->peer (what fused_ag_gemm.cu's copy engine does)
//   mode 1: all threads, 16-byte ld.global / st.global to every destination (classic SM copy)
//   mode 2: like 0 but the local read is skipped after the first fill (isolates the store side)
#include "common.cuh"
#include "ptx.cuh"

namespace {
constexpr int kMaxDst = 8;
struct ProbeArgs {
    const uint8_t* src;
    uint8_t* dst[kMaxDst];
    size_t bytes;       // bytes of src copied to EVERY destination
    int n_dst;
    int chunk;          // bytes per bulk op (mode 0/2), multiple of 16, <= 32768
    int bufs;           // ring depth (mode 0/2)
    int read_lag;
    int sig_mode;       // 0 none; 1 fence + release red per peer; 2 fence + relaxed reds; 3 relaxed reds only; 4 fence only
    int sig_every;      // items per signal
    size_t flag_off;    // byte offset of this writer's flag inside every destination buffer
};

__device__ __forceinline__ void p_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     ptx::smem_u32(dst)),
                 "l"(reinterpret_cast<uint64_t>(src)), "r"(bytes), "r"(ptx::smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void p_bulk_s2g(void* dst, const void* src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(reinterpret_cast<uint64_t>(dst)),
                 "r"(ptx::smem_u32(src)), "r"(bytes)
                 : "memory");
}

template <int kMode>
__global__ void __launch_bounds__(512, 1) p2p_push_kernel(const ProbeArgs a) {
    extern __shared__ __align__(128) uint8_t psm[];
    if constexpr (kMode == 1) {
        const size_t n16 = a.bytes / 16;
        const uint4* s = reinterpret_cast<const uint4*>(a.src);
        const size_t stride = (size_t)gridDim.x * blockDim.x;
        size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
        for (; i + 3 * stride < n16; i += 4 * stride) {
            uint4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) v[u] = __ldg(s + i + u * stride);
            for (int d = 0; d < a.n_dst; ++d) {
                uint4* o = reinterpret_cast<uint4*>(a.dst[d]);
#pragma unroll
                for (int u = 0; u < 4; ++u) o[i + u * stride] = v[u];
            }
        }
        for (; i < n16; i += stride) {
            const uint4 v = __ldg(s + i);
            for (int d = 0; d < a.n_dst; ++d) reinterpret_cast<uint4*>(a.dst[d])[i] = v;
        }
        return;
    } else {
        if (threadIdx.x != 0) return;
        uint64_t* bars = reinterpret_cast<uint64_t*>(psm + (size_t)a.bufs * a.chunk);
        for (int i = 0; i < a.bufs; ++i) ptx::mbar_init(&bars[i], 1);
        ptx::fence_barrier_init();
        // this CTA's contiguous piece
        const size_t per = ((a.bytes / gridDim.x) + 15) & ~size_t(15);
        const size_t p0 = min((size_t)blockIdx.x * per, a.bytes), p1 = min(p0 + per, a.bytes);
        const int n_items = (int)((p1 - p0 + a.chunk - 1) / a.chunk);
        uint32_t phases = 0;
        const int ahead = a.bufs - a.read_lag;
        auto issue_load = [&](int j) {
            const int b = j % a.bufs;
            const uint32_t bytes = (uint32_t)min((size_t)a.chunk, p1 - p0 - (size_t)j * a.chunk);
            ptx::mbar_arrive_expect_tx(&bars[b], bytes);
            p_bulk_g2s(psm + (size_t)b * a.chunk, a.src + p0 + (size_t)j * a.chunk, bytes, &bars[b]);
        };
        const int preload = min(n_items, kMode == 2 ? a.bufs : ahead);
        for (int j = 0; j < preload; ++j) issue_load(j);
        for (int j = 0; j < n_items; ++j) {
            const int b = j % a.bufs;
            if (kMode == 0 || j < a.bufs) {
                ptx::mbar_wait(&bars[b], (phases >> b) & 1u);
                phases ^= (1u << b);
            }
            const uint32_t bytes = (uint32_t)min((size_t)a.chunk, p1 - p0 - (size_t)j * a.chunk);
            for (int d = 0; d < a.n_dst; ++d)
                p_bulk_s2g(a.dst[d] + p0 + (size_t)j * a.chunk, psm + (size_t)b * a.chunk, bytes);
            ptx::bulk_commit();
            if (a.sig_mode != 0 && (j + 1) % a.sig_every == 0 && j >= 4) {
                ptx::bulk_wait<4>();     // everything but the 4 newest groups has landed
                if (a.sig_mode == 1 || a.sig_mode == 2 || a.sig_mode == 4) __threadfence_system();
                if (a.sig_mode != 4) {
                    for (int d = 0; d < a.n_dst; ++d) {
                        uint32_t* f = reinterpret_cast<uint32_t*>(a.dst[d] + a.flag_off);
                        if (a.sig_mode == 1) asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(f), "r"(1u) : "memory");
                        else asm volatile("red.relaxed.sys.global.add.u32 [%0], %1;" ::"l"(f), "r"(1u) : "memory");
                    }
                }
            }
            if (kMode == 0 && j + ahead < n_items) {
                // wait until the buffer about to be refilled has been read (read_lag groups may still be reading)
                switch (a.read_lag) {
                    case 0: ptx::bulk_wait_read<0>(); break;
                    case 1: ptx::bulk_wait_read<1>(); break;
                    case 2: ptx::bulk_wait_read<2>(); break;
                    case 4: ptx::bulk_wait_read<4>(); break;
                    default: ptx::bulk_wait_read<8>(); break;
                }
                issue_load(j + ahead);
            }
        }
        ptx::bulk_wait<0>();
    }
}
}  // namespace

// dst_ptrs: host array of n_dst device pointers (peer-mapped).  Returns 0 or a CUDA error code.
NXD_API int nxd_p2p_push_probe(int mode, const void* src, const void* const* dst_ptrs, int n_dst, size_t bytes,
                                  int ctas, int chunk, int bufs, int read_lag, int sig_mode, int sig_every, size_t flag_off,
                                  void* stream) {
    NXD_REQUIRE(n_dst >= 1 && n_dst <= kMaxDst && bytes % 16 == 0 && chunk % 16 == 0 && chunk <= 32768 && bufs >= 1 &&
                    bufs <= 32 && (size_t)bufs * chunk + 512 <= 220 * 1024,
                "nxd_p2p_push_probe: bad arguments");
    ProbeArgs a{};
    a.src = (const uint8_t*)src;
    for (int i = 0; i < n_dst; ++i) a.dst[i] = (uint8_t*)dst_ptrs[i];
    a.bytes = bytes; a.n_dst = n_dst; a.chunk = chunk; a.bufs = bufs; a.read_lag = read_lag;
    a.sig_mode = sig_mode; a.sig_every = sig_every > 0 ? sig_every : 1; a.flag_off = flag_off;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t smem = mode == 1 ? 0 : (size_t)bufs * chunk + 512;
    if (mode == 0) {
        NXD_CHECK_CUDA(cudaFuncSetAttribute(p2p_push_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        p2p_push_kernel<0><<<ctas, 32, smem, st>>>(a);
    } else if (mode == 1) {
        p2p_push_kernel<1><<<ctas, 512, 0, st>>>(a);
    } else {
        NXD_CHECK_CUDA(cudaFuncSetAttribute(p2p_push_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        p2p_push_kernel<2><<<ctas, 32, smem, st>>>(a);
    }
    NXD_LAUNCH_CHECK();
    return 0;
}
