// Causal / full flash attention forward + backward with native GQA (C23).
//
// Provides what the reference gets from the external NKI kernels `flash_fwd` / `flash_attn_bwd`
// (src/neuronx_distributed/kernels/flash_attn.py:19-28,151-191) and from `attention_isa_kernel`
// for inference prefill (examples/inference/modules/attention/attention_base.py:98-139):
//   O = softmax(scale * Q K^T + causal mask) V     q [B,Hq,Sq,D], k/v [B,Hkv,Sk,D], D in {64,128}
// without the reference's S % 2048 restriction and without materialising repeat_kv
// (examples/training/llama/modeling_llama_nxd.py:409-410): q-head h reads kv-head h / (Hq/Hkv).
//
// Tensors are addressed through (batch, head, row) element strides with unit stride on D, so the
// [S, B, heads, D] activations of the sequence-parallel transformer block are consumed in place.
//
// Structure (per kernel: 256 threads = 8 warps, each warp owns 16 rows of the resident tile):
//   fwd     grid (q-blocks of 128, Hq, B): K/V tiles of 64 rows stream through a 2-deep cp.async
//           ring (XOR-swizzled 16 B chunks, conflict-free ldmatrix); S and O live in registers;
//           online softmax in the exp2 domain; LSE (log2 units) saved for the backward.
//   delta   Delta[b,h,i] = sum_d O*dO.
//   dkdv    grid (kv-blocks of 128, Hkv, B): loops over the q-heads of the GQA group and over q
//           tiles of 64 rows; computes S^T and dP^T so that P^T / dS^T come out in A-fragment
//           layout for dV += P^T dO and dK += dS^T Q.  The GQA reduction of dK/dV over the group
//           happens in registers (the reference needs the kv-shared all-reduce only across ranks).
//   dq      grid (q-blocks of 128, Hq, B): recomputes S, dP per K/V tile; dQ += dS K.
// Deterministic: no atomics anywhere.
#include "common.cuh"

namespace {

constexpr int kThreads = 256;

struct FaParams {
    const bf16* q; const bf16* k; const bf16* v; bf16* o;
    const bf16* dout; bf16* dq; bf16* dk; bf16* dv;
    float* lse;    // [B, Hq, Sq] log2-domain logsumexp of the scaled scores
    float* delta;  // [B, Hq, Sq]
    long long q_sb, q_sh, q_ss, k_sb, k_sh, k_ss, v_sb, v_sh, v_ss, o_sb, o_sh, o_ss;
    long long do_sb, do_sh, do_ss, dq_sb, dq_sh, dq_ss, dk_sb, dk_sh, dk_ss, dv_sb, dv_sh, dv_ss;
    int B, Hq, Hkv, Sq, Sk;
    float scale, scale_log2;
};

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
    const int sz = valid ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src, bool valid) {
    const int sz = valid ? 4 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(addr));
}
// c[16x8] += a[16x16] * b[16x8]   (bf16 inputs, fp32 accumulate)
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
    bf162 t = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float quad_max(float v) {
    v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 1));
    return fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 2));
}
__device__ __forceinline__ float quad_sum(float v) {
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    return v + __shfl_xor_sync(0xffffffffu, v, 2);
}

// smem tile [rows][D] bf16, 16-byte chunks XOR-swizzled with the row index.
template <int D>
__device__ __forceinline__ uint32_t tile_addr(uint32_t base, int row, int chunk) {
    return base + row * (D * 2) + ((chunk ^ (row & 7)) << 4);
}

template <int D, int ROWS>
__device__ __forceinline__ void load_tile_async(uint32_t sbase, const bf16* g, long long stride_row, int row0,
                                                int rows_total) {
    constexpr int CH = D / 8;
    for (int i = threadIdx.x; i < ROWS * CH; i += kThreads) {
        const int r = i / CH, c = i % CH;
        const bool ok = row0 + r < rows_total;
        const bf16* src = g + (long long)(ok ? row0 + r : 0) * stride_row + c * 8;
        cp_async16(tile_addr<D>(sbase, r, c), src, ok);
    }
}

// A fragment (16 rows x 16 k) of a [rows][D] tile.
template <int D>
__device__ __forceinline__ void load_a(uint32_t (&a)[4], uint32_t sbase, int r0, int kk, int lane) {
    const int row = r0 + (lane & 7) + ((lane >> 3) & 1) * 8;
    ldsm_x4(a, tile_addr<D>(sbase, row, kk * 2 + (lane >> 4)));
}
// B fragments for two adjacent 8-wide n blocks from an [n][D] tile (contraction over D).
template <int D>
__device__ __forceinline__ void load_b(uint32_t (&b)[4], uint32_t sbase, int n0, int kk, int lane) {
    const int row = n0 + (lane & 7) + (lane >> 4) * 8;
    ldsm_x4(b, tile_addr<D>(sbase, row, kk * 2 + ((lane >> 3) & 1)));
}
// B fragments for two adjacent 8-wide d blocks from a [k][D] tile (contraction over rows).
template <int D>
__device__ __forceinline__ void load_bt(uint32_t (&b)[4], uint32_t sbase, int k0, int dj, int lane) {
    const int row = k0 + (lane & 7) + ((lane >> 3) & 1) * 8;
    ldsm_x4_t(b, tile_addr<D>(sbase, row, dj * 2 + (lane >> 4)));
}

// Write this warp's 16 x D fp32 accumulator (C-fragment layout) as bf16 rows through its own
// 16 rows of a staging tile, then 16-byte coalesced stores.
template <int D>
__device__ __forceinline__ void store_rows(float (&acc)[D / 8][4], float s0, float s1, uint32_t sbase,
                                           uint8_t* sptr, int r0, bf16* g, long long stride_row, int grow0,
                                           int rows_total, int lane) {
    const int g4 = lane >> 2, t4 = lane & 3;
#pragma unroll
    for (int j = 0; j < D / 8; ++j) {
        // chunk j holds cols j*8..j*8+7; this thread's pair sits at byte offset t4*4 inside it
        const uint32_t a0 = tile_addr<D>(sbase, r0 + g4, j) + t4 * 4;
        const uint32_t a1 = tile_addr<D>(sbase, r0 + g4 + 8, j) + t4 * 4;
        const uint32_t v0 = pack2(acc[j][0] * s0, acc[j][1] * s0);
        const uint32_t v1 = pack2(acc[j][2] * s1, acc[j][3] * s1);
        asm volatile("st.shared.b32 [%0], %1;" ::"r"(a0), "r"(v0) : "memory");
        asm volatile("st.shared.b32 [%0], %1;" ::"r"(a1), "r"(v1) : "memory");
    }
    __syncwarp();
    constexpr int CH = D / 8;
    for (int i = lane; i < 16 * CH; i += 32) {
        const int r = i / CH, c = i % CH;
        if (grow0 + r < rows_total) {
            const uint32_t off = tile_addr<D>(0, r0 + r, c);
            const uint4 val = *reinterpret_cast<const uint4*>(sptr + off);
            *reinterpret_cast<uint4*>(g + (long long)(grow0 + r) * stride_row + c * 8) = val;
        }
    }
}

// =============================================================================================
// forward
// =============================================================================================
template <int D, bool CAUSAL>
__global__ void __launch_bounds__(kThreads, 1) fa_fwd_kernel(const FaParams p) {
    constexpr int BM = 128, BN = 64;
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t sQ = (uint32_t)__cvta_generic_to_shared(smem);
    const uint32_t sK = sQ + BM * D * 2;
    const uint32_t sV = sK + 2 * BN * D * 2;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g4 = lane >> 2, t4 = lane & 3;
    const int mb = gridDim.x - 1 - blockIdx.x;  // heavy (late) causal blocks first
    const int hq = blockIdx.y, b = blockIdx.z;
    const int hkv = hq / (p.Hq / p.Hkv);
    const int m0 = mb * BM;
    const int offset = p.Sk - p.Sq;
    const bf16* qg = p.q + b * p.q_sb + hq * p.q_sh;
    const bf16* kg = p.k + b * p.k_sb + hkv * p.k_sh;
    const bf16* vg = p.v + b * p.v_sb + hkv * p.v_sh;

    int n_blocks = (p.Sk + BN - 1) / BN;
    if (CAUSAL) {
        const int lim = (m0 + BM + offset + BN - 1) / BN;
        n_blocks = min(n_blocks, max(lim, 0));
    }

    load_tile_async<D, BM>(sQ, qg, p.q_ss, m0, p.Sq);
    if (n_blocks > 0) {
        load_tile_async<D, BN>(sK, kg, p.k_ss, 0, p.Sk);
        load_tile_async<D, BN>(sV, vg, p.v_ss, 0, p.Sk);
    }
    cp_async_commit();

    float o[D / 8][4];
#pragma unroll
    for (int j = 0; j < D / 8; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
    float m_i[2] = {-INFINITY, -INFINITY}, l_i[2] = {0.f, 0.f};
    uint32_t qf[D / 16][4];
    const int r0 = warp * 16;
    const int row_a = m0 + r0 + g4, row_b = row_a + 8;

    for (int nb = 0; nb < n_blocks; ++nb) {
        const int buf = nb & 1;
        if (nb + 1 < n_blocks) {
            load_tile_async<D, BN>(sK + (buf ^ 1) * BN * D * 2, kg, p.k_ss, (nb + 1) * BN, p.Sk);
            load_tile_async<D, BN>(sV + (buf ^ 1) * BN * D * 2, vg, p.v_ss, (nb + 1) * BN, p.Sk);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        if (nb == 0) {
#pragma unroll
            for (int kk = 0; kk < D / 16; ++kk) load_a<D>(qf[kk], sQ, r0, kk, lane);
        }
        const uint32_t sKb = sK + buf * BN * D * 2, sVb = sV + buf * BN * D * 2;
        float s[BN / 8][4];
#pragma unroll
        for (int j = 0; j < BN / 8; ++j) s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
#pragma unroll
            for (int nj = 0; nj < BN / 16; ++nj) {
                uint32_t bfr[4];
                load_b<D>(bfr, sKb, nj * 16, kk, lane);
                mma16816(s[2 * nj], qf[kk], bfr[0], bfr[1]);
                mma16816(s[2 * nj + 1], qf[kk], bfr[2], bfr[3]);
            }
        }
        const int n0 = nb * BN;
        const bool need_mask = (n0 + BN > p.Sk) || (CAUSAL && (n0 + BN - 1 > m0 + r0 + offset));
        float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
        for (int j = 0; j < BN / 8; ++j) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                float val = s[j][e] * p.scale_log2;
                if (need_mask) {
                    const int col = n0 + j * 8 + t4 * 2 + (e & 1);
                    const int row = (e < 2) ? row_a : row_b;
                    if (col >= p.Sk || (CAUSAL && col > row + offset)) val = -INFINITY;
                }
                s[j][e] = val;
            }
            mx0 = fmaxf(mx0, fmaxf(s[j][0], s[j][1]));
            mx1 = fmaxf(mx1, fmaxf(s[j][2], s[j][3]));
        }
        mx0 = quad_max(mx0);
        mx1 = quad_max(mx1);
        const float mn0 = fmaxf(m_i[0], mx0), mn1 = fmaxf(m_i[1], mx1);
        const float ms0 = (mn0 == -INFINITY) ? 0.f : mn0, ms1 = (mn1 == -INFINITY) ? 0.f : mn1;
        const float al0 = exp2f(m_i[0] - ms0), al1 = exp2f(m_i[1] - ms1);
        m_i[0] = mn0;
        m_i[1] = mn1;
        float rs0 = 0.f, rs1 = 0.f;
        uint32_t pa[BN / 16][4];
#pragma unroll
        for (int j = 0; j < BN / 8; ++j) {
            const float p0 = exp2f(s[j][0] - ms0), p1 = exp2f(s[j][1] - ms0);
            const float p2 = exp2f(s[j][2] - ms1), p3 = exp2f(s[j][3] - ms1);
            rs0 += p0 + p1;
            rs1 += p2 + p3;
            pa[j >> 1][(j & 1) * 2] = pack2(p0, p1);
            pa[j >> 1][(j & 1) * 2 + 1] = pack2(p2, p3);
        }
        l_i[0] = l_i[0] * al0 + rs0;
        l_i[1] = l_i[1] * al1 + rs1;
#pragma unroll
        for (int j = 0; j < D / 8; ++j) {
            o[j][0] *= al0; o[j][1] *= al0; o[j][2] *= al1; o[j][3] *= al1;
        }
#pragma unroll
        for (int kk = 0; kk < BN / 16; ++kk) {
#pragma unroll
            for (int dj = 0; dj < D / 16; ++dj) {
                uint32_t bfr[4];
                load_bt<D>(bfr, sVb, kk * 16, dj, lane);
                mma16816(o[2 * dj], pa[kk], bfr[0], bfr[1]);
                mma16816(o[2 * dj + 1], pa[kk], bfr[2], bfr[3]);
            }
        }
        __syncthreads();
    }
    if (n_blocks == 0) {
        cp_async_wait<0>();
        __syncthreads();
    }
    const float l0 = quad_sum(l_i[0]), l1 = quad_sum(l_i[1]);
    const float inv0 = l0 > 0.f ? 1.f / l0 : 0.f, inv1 = l1 > 0.f ? 1.f / l1 : 0.f;
    if (p.lse && t4 == 0) {
        float* lg = p.lse + ((long long)b * p.Hq + hq) * p.Sq;
        if (row_a < p.Sq) lg[row_a] = l0 > 0.f ? m_i[0] + log2f(l0) : INFINITY;
        if (row_b < p.Sq) lg[row_b] = l1 > 0.f ? m_i[1] + log2f(l1) : INFINITY;
    }
    // every warp only touches its own 16 rows of the (now dead) Q tile
    store_rows<D>(o, inv0, inv1, sQ, smem, r0, p.o + b * p.o_sb + hq * p.o_sh, p.o_ss, m0 + r0, p.Sq, lane);
}

// =============================================================================================
// backward: Delta
// =============================================================================================
template <int D>
__global__ void fa_delta_kernel(const FaParams p) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + warp;
    const long long total = (long long)p.B * p.Hq * p.Sq;
    if (row >= total) return;
    const int i = (int)(row % p.Sq);
    const int h = (int)((row / p.Sq) % p.Hq);
    const int b = (int)(row / ((long long)p.Sq * p.Hq));
    const bf16* o = p.o + b * p.o_sb + h * p.o_sh + (long long)i * p.o_ss;
    const bf16* d = p.dout + b * p.do_sb + h * p.do_sh + (long long)i * p.do_ss;
    float acc = 0.f;
#pragma unroll
    for (int c = 0; c < D / 64; ++c) {
        const bf162 a = reinterpret_cast<const bf162*>(o)[lane + c * 32];
        const bf162 g = reinterpret_cast<const bf162*>(d)[lane + c * 32];
        const float2 af = __bfloat1622float2(a), gf = __bfloat1622float2(g);
        acc += af.x * gf.x + af.y * gf.y;
    }
    acc = warp_sum(acc);
    if (lane == 0) p.delta[row] = acc;
}

// =============================================================================================
// backward: dK, dV  (one CTA = 128 kv rows of one kv head; loops over the GQA group's q heads)
// =============================================================================================
template <int D, bool CAUSAL>
__global__ void __launch_bounds__(kThreads, 1) fa_bwd_dkdv_kernel(const FaParams p) {
    constexpr int BN = 128, BM = 64;
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t sK = (uint32_t)__cvta_generic_to_shared(smem);
    const uint32_t sV = sK + BN * D * 2;
    const uint32_t sQ = sV + BN * D * 2;            // 2 buffers
    const uint32_t sDO = sQ + 2 * BM * D * 2;       // 2 buffers
    const uint32_t sL = sDO + 2 * BM * D * 2;       // 2 x BM floats
    const uint32_t sDl = sL + 2 * BM * 4;           // 2 x BM floats
    const float* sLp = reinterpret_cast<const float*>(smem + (sL - sK));
    const float* sDlp = reinterpret_cast<const float*>(smem + (sDl - sK));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g4 = lane >> 2, t4 = lane & 3;
    const int nbk = blockIdx.x, hkv = blockIdx.y, b = blockIdx.z;
    const int group = p.Hq / p.Hkv;
    const int n0 = nbk * BN;
    const int offset = p.Sk - p.Sq;
    const int r0 = warp * 16;
    const int krow_a = n0 + r0 + g4, krow_b = krow_a + 8;

    int mb_start = 0;
    if (CAUSAL) mb_start = max(0, (n0 - offset) / BM);  // first q tile that can see key n0
    const int mb_end = (p.Sq + BM - 1) / BM;
    const int tiles_per_head = max(mb_end - mb_start, 0);
    const int total = tiles_per_head * group;

    auto issue = [&](int it, int buf) {
        const int hq = hkv * group + it / tiles_per_head;
        const int mb = mb_start + it % tiles_per_head;
        load_tile_async<D, BM>(sQ + buf * BM * D * 2, p.q + b * p.q_sb + hq * p.q_sh, p.q_ss, mb * BM, p.Sq);
        load_tile_async<D, BM>(sDO + buf * BM * D * 2, p.dout + b * p.do_sb + hq * p.do_sh, p.do_ss, mb * BM, p.Sq);
        if (threadIdx.x < BM) {
            const int i = mb * BM + threadIdx.x;
            const bool ok = i < p.Sq;
            const long long base = ((long long)b * p.Hq + hq) * p.Sq + (ok ? i : 0);
            cp_async4(sL + (buf * BM + threadIdx.x) * 4, p.lse + base, ok);
            cp_async4(sDl + (buf * BM + threadIdx.x) * 4, p.delta + base, ok);
        }
    };

    load_tile_async<D, BN>(sK, p.k + b * p.k_sb + hkv * p.k_sh, p.k_ss, n0, p.Sk);
    load_tile_async<D, BN>(sV, p.v + b * p.v_sb + hkv * p.v_sh, p.v_ss, n0, p.Sk);
    if (total > 0) issue(0, 0);
    cp_async_commit();

    float dk[D / 8][4], dv[D / 8][4];
#pragma unroll
    for (int j = 0; j < D / 8; ++j) {
        dk[j][0] = dk[j][1] = dk[j][2] = dk[j][3] = 0.f;
        dv[j][0] = dv[j][1] = dv[j][2] = dv[j][3] = 0.f;
    }

    for (int it = 0; it < total; ++it) {
        const int buf = it & 1;
        if (it + 1 < total) {
            issue(it + 1, buf ^ 1);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        const int mb = mb_start + it % tiles_per_head;
        const int mrow0 = mb * BM;
        const uint32_t sQb = sQ + buf * BM * D * 2, sDOb = sDO + buf * BM * D * 2;
        // S^T[n][m] = K Q^T ; dP^T[n][m] = V dO^T
        float st[BM / 8][4], dpt[BM / 8][4];
#pragma unroll
        for (int j = 0; j < BM / 8; ++j) {
            st[j][0] = st[j][1] = st[j][2] = st[j][3] = 0.f;
            dpt[j][0] = dpt[j][1] = dpt[j][2] = dpt[j][3] = 0.f;
        }
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
            uint32_t ka[4], va[4];
            load_a<D>(ka, sK, r0, kk, lane);
            load_a<D>(va, sV, r0, kk, lane);
#pragma unroll
            for (int mj = 0; mj < BM / 16; ++mj) {
                uint32_t bq[4], bd[4];
                load_b<D>(bq, sQb, mj * 16, kk, lane);
                mma16816(st[2 * mj], ka, bq[0], bq[1]);
                mma16816(st[2 * mj + 1], ka, bq[2], bq[3]);
                load_b<D>(bd, sDOb, mj * 16, kk, lane);
                mma16816(dpt[2 * mj], va, bd[0], bd[1]);
                mma16816(dpt[2 * mj + 1], va, bd[2], bd[3]);
            }
        }
        // P^T = exp2(S^T * c - L[m]);  dS^T = P^T * (dP^T - Delta[m])
        uint32_t pa[BM / 16][4], dsa[BM / 16][4];
#pragma unroll
        for (int j = 0; j < BM / 8; ++j) {
            const int c0 = j * 8 + t4 * 2;
            const float L0 = sLp[buf * BM + c0], L1 = sLp[buf * BM + c0 + 1];
            const float D0 = sDlp[buf * BM + c0], D1 = sDlp[buf * BM + c0 + 1];
            float pv[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int mcol = mrow0 + c0 + (e & 1);
                const int krow = (e < 2) ? krow_a : krow_b;
                const float L = (e & 1) ? L1 : L0;
                float val = exp2f(st[j][e] * p.scale_log2 - L);
                if (mcol >= p.Sq || krow >= p.Sk || (CAUSAL && krow > mcol + offset)) val = 0.f;
                pv[e] = val;
            }
            const float d0 = pv[0] * (dpt[j][0] - D0), d1 = pv[1] * (dpt[j][1] - D1);
            const float d2 = pv[2] * (dpt[j][2] - D0), d3 = pv[3] * (dpt[j][3] - D1);
            pa[j >> 1][(j & 1) * 2] = pack2(pv[0], pv[1]);
            pa[j >> 1][(j & 1) * 2 + 1] = pack2(pv[2], pv[3]);
            dsa[j >> 1][(j & 1) * 2] = pack2(d0, d1);
            dsa[j >> 1][(j & 1) * 2 + 1] = pack2(d2, d3);
        }
        // dV += P^T dO ; dK += dS^T Q   (contraction over the q rows of the tile)
#pragma unroll
        for (int kk = 0; kk < BM / 16; ++kk) {
#pragma unroll
            for (int dj = 0; dj < D / 16; ++dj) {
                uint32_t bd[4], bq[4];
                load_bt<D>(bd, sDOb, kk * 16, dj, lane);
                mma16816(dv[2 * dj], pa[kk], bd[0], bd[1]);
                mma16816(dv[2 * dj + 1], pa[kk], bd[2], bd[3]);
                load_bt<D>(bq, sQb, kk * 16, dj, lane);
                mma16816(dk[2 * dj], dsa[kk], bq[0], bq[1]);
                mma16816(dk[2 * dj + 1], dsa[kk], bq[2], bq[3]);
            }
        }
        __syncthreads();
    }
    if (total == 0) {
        cp_async_wait<0>();
    }
    __syncthreads();
    // K / V tiles are dead: stage the results through them (each warp its own 16 rows)
    store_rows<D>(dv, 1.f, 1.f, sV, smem + (sV - sK), r0, p.dv + b * p.dv_sb + hkv * p.dv_sh, p.dv_ss, n0 + r0, p.Sk,
                  lane);
    store_rows<D>(dk, p.scale, p.scale, sK, smem, r0, p.dk + b * p.dk_sb + hkv * p.dk_sh, p.dk_ss, n0 + r0, p.Sk,
                  lane);
}

// =============================================================================================
// backward: dQ
// =============================================================================================
template <int D, bool CAUSAL>
__global__ void __launch_bounds__(kThreads, 1) fa_bwd_dq_kernel(const FaParams p) {
    constexpr int BM = 128, BN = 64;
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t sQ = (uint32_t)__cvta_generic_to_shared(smem);
    const uint32_t sDO = sQ + BM * D * 2;
    const uint32_t sK = sDO + BM * D * 2;
    const uint32_t sV = sK + 2 * BN * D * 2;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g4 = lane >> 2, t4 = lane & 3;
    const int mb = gridDim.x - 1 - blockIdx.x;
    const int hq = blockIdx.y, b = blockIdx.z;
    const int hkv = hq / (p.Hq / p.Hkv);
    const int m0 = mb * BM;
    const int offset = p.Sk - p.Sq;
    const bf16* kg = p.k + b * p.k_sb + hkv * p.k_sh;
    const bf16* vg = p.v + b * p.v_sb + hkv * p.v_sh;

    int n_blocks = (p.Sk + BN - 1) / BN;
    if (CAUSAL) {
        const int lim = (m0 + BM + offset + BN - 1) / BN;
        n_blocks = min(n_blocks, max(lim, 0));
    }
    load_tile_async<D, BM>(sQ, p.q + b * p.q_sb + hq * p.q_sh, p.q_ss, m0, p.Sq);
    load_tile_async<D, BM>(sDO, p.dout + b * p.do_sb + hq * p.do_sh, p.do_ss, m0, p.Sq);
    if (n_blocks > 0) {
        load_tile_async<D, BN>(sK, kg, p.k_ss, 0, p.Sk);
        load_tile_async<D, BN>(sV, vg, p.v_ss, 0, p.Sk);
    }
    cp_async_commit();

    const int r0 = warp * 16;
    const int row_a = m0 + r0 + g4, row_b = row_a + 8;
    const long long sbase = ((long long)b * p.Hq + hq) * p.Sq;
    const float La = row_a < p.Sq ? p.lse[sbase + row_a] : INFINITY;
    const float Lb = row_b < p.Sq ? p.lse[sbase + row_b] : INFINITY;
    const float Da = row_a < p.Sq ? p.delta[sbase + row_a] : 0.f;
    const float Db = row_b < p.Sq ? p.delta[sbase + row_b] : 0.f;

    float dq[D / 8][4];
#pragma unroll
    for (int j = 0; j < D / 8; ++j) dq[j][0] = dq[j][1] = dq[j][2] = dq[j][3] = 0.f;

    for (int nb = 0; nb < n_blocks; ++nb) {
        const int buf = nb & 1;
        if (nb + 1 < n_blocks) {
            load_tile_async<D, BN>(sK + (buf ^ 1) * BN * D * 2, kg, p.k_ss, (nb + 1) * BN, p.Sk);
            load_tile_async<D, BN>(sV + (buf ^ 1) * BN * D * 2, vg, p.v_ss, (nb + 1) * BN, p.Sk);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        const uint32_t sKb = sK + buf * BN * D * 2, sVb = sV + buf * BN * D * 2;
        float s[BN / 8][4], dp[BN / 8][4];
#pragma unroll
        for (int j = 0; j < BN / 8; ++j) {
            s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
            dp[j][0] = dp[j][1] = dp[j][2] = dp[j][3] = 0.f;
        }
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
            uint32_t qa[4], da[4];
            load_a<D>(qa, sQ, r0, kk, lane);
            load_a<D>(da, sDO, r0, kk, lane);
#pragma unroll
            for (int nj = 0; nj < BN / 16; ++nj) {
                uint32_t bk[4], bv[4];
                load_b<D>(bk, sKb, nj * 16, kk, lane);
                mma16816(s[2 * nj], qa, bk[0], bk[1]);
                mma16816(s[2 * nj + 1], qa, bk[2], bk[3]);
                load_b<D>(bv, sVb, nj * 16, kk, lane);
                mma16816(dp[2 * nj], da, bv[0], bv[1]);
                mma16816(dp[2 * nj + 1], da, bv[2], bv[3]);
            }
        }
        const int n0 = nb * BN;
        uint32_t dsa[BN / 16][4];
#pragma unroll
        for (int j = 0; j < BN / 8; ++j) {
            float dsv[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int col = n0 + j * 8 + t4 * 2 + (e & 1);
                const int row = (e < 2) ? row_a : row_b;
                const float L = (e < 2) ? La : Lb;
                const float Dl = (e < 2) ? Da : Db;
                float pv = exp2f(s[j][e] * p.scale_log2 - L);
                if (col >= p.Sk || (CAUSAL && col > row + offset)) pv = 0.f;
                dsv[e] = pv * (dp[j][e] - Dl);
            }
            dsa[j >> 1][(j & 1) * 2] = pack2(dsv[0], dsv[1]);
            dsa[j >> 1][(j & 1) * 2 + 1] = pack2(dsv[2], dsv[3]);
        }
#pragma unroll
        for (int kk = 0; kk < BN / 16; ++kk) {
#pragma unroll
            for (int dj = 0; dj < D / 16; ++dj) {
                uint32_t bk[4];
                load_bt<D>(bk, sKb, kk * 16, dj, lane);
                mma16816(dq[2 * dj], dsa[kk], bk[0], bk[1]);
                mma16816(dq[2 * dj + 1], dsa[kk], bk[2], bk[3]);
            }
        }
        __syncthreads();
    }
    if (n_blocks == 0) cp_async_wait<0>();
    __syncthreads();
    store_rows<D>(dq, p.scale, p.scale, sQ, smem, r0, p.dq + b * p.dq_sb + hq * p.dq_sh, p.dq_ss, m0 + r0, p.Sq,
                  lane);
}

template <typename K>
int set_smem(K kern, int bytes) {
    NXD_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    return 0;
}

template <int D>
int fwd_launch(const FaParams& p, int causal, cudaStream_t st) {
    constexpr int smem = (128 + 4 * 64) * D * 2;
    dim3 grid((p.Sq + 127) / 128, p.Hq, p.B);
    if (causal) {
        static int once = set_smem(fa_fwd_kernel<D, true>, smem);
        if (once) return once;
        fa_fwd_kernel<D, true><<<grid, kThreads, smem, st>>>(p);
    } else {
        static int once = set_smem(fa_fwd_kernel<D, false>, smem);
        if (once) return once;
        fa_fwd_kernel<D, false><<<grid, kThreads, smem, st>>>(p);
    }
    NXD_LAUNCH_CHECK();
    return 0;
}

template <int D>
int bwd_launch(const FaParams& p, int causal, cudaStream_t st) {
    {
        const long long rows = (long long)p.B * p.Hq * p.Sq;
        fa_delta_kernel<D><<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(p);
        NXD_LAUNCH_CHECK();
    }
    constexpr int smem_kv = (2 * 128 + 4 * 64) * D * 2 + 4 * 64 * 4;
    constexpr int smem_q = (2 * 128 + 4 * 64) * D * 2;
    dim3 grid_kv((p.Sk + 127) / 128, p.Hkv, p.B), grid_q((p.Sq + 127) / 128, p.Hq, p.B);
    if (causal) {
        static int once = set_smem(fa_bwd_dkdv_kernel<D, true>, smem_kv) | set_smem(fa_bwd_dq_kernel<D, true>, smem_q);
        if (once) return once;
        fa_bwd_dkdv_kernel<D, true><<<grid_kv, kThreads, smem_kv, st>>>(p);
        NXD_LAUNCH_CHECK();
        fa_bwd_dq_kernel<D, true><<<grid_q, kThreads, smem_q, st>>>(p);
    } else {
        static int once =
            set_smem(fa_bwd_dkdv_kernel<D, false>, smem_kv) | set_smem(fa_bwd_dq_kernel<D, false>, smem_q);
        if (once) return once;
        fa_bwd_dkdv_kernel<D, false><<<grid_kv, kThreads, smem_kv, st>>>(p);
        NXD_LAUNCH_CHECK();
        fa_bwd_dq_kernel<D, false><<<grid_q, kThreads, smem_q, st>>>(p);
    }
    NXD_LAUNCH_CHECK();
    return 0;
}

int check_common(const FaParams& p, int D) {
    NXD_REQUIRE(D == 64 || D == 128, "flash_attn: head dim %d not supported (64 or 128)", D);
    NXD_REQUIRE(p.B > 0 && p.Hq > 0 && p.Hkv > 0 && p.Sq > 0 && p.Sk > 0, "flash_attn: empty problem");
    NXD_REQUIRE(p.Hq % p.Hkv == 0, "flash_attn: Hq %d not a multiple of Hkv %d", p.Hq, p.Hkv);
    NXD_REQUIRE(p.B <= 65535 && p.Hq <= 65535, "flash_attn: grid limits exceeded");
    return 0;
}

}  // namespace

// strides: 12 (fwd) / 24 (bwd) element strides in the order (batch, head, row) for q, k, v, o[, do, dq, dk, dv].
NXD_API int nxd_flash_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse,
                               const long long* strides, int B, int Hq, int Hkv, int Sq, int Sk, int D, float scale,
                               int causal, void* stream) {
    FaParams p{};
    p.q = (const bf16*)q; p.k = (const bf16*)k; p.v = (const bf16*)v; p.o = (bf16*)o; p.lse = lse;
    p.q_sb = strides[0]; p.q_sh = strides[1]; p.q_ss = strides[2];
    p.k_sb = strides[3]; p.k_sh = strides[4]; p.k_ss = strides[5];
    p.v_sb = strides[6]; p.v_sh = strides[7]; p.v_ss = strides[8];
    p.o_sb = strides[9]; p.o_sh = strides[10]; p.o_ss = strides[11];
    p.B = B; p.Hq = Hq; p.Hkv = Hkv; p.Sq = Sq; p.Sk = Sk;
    p.scale = scale; p.scale_log2 = scale * 1.4426950408889634f;
    if (int rc = check_common(p, D)) return rc;
    NXD_REQUIRE(p.q_ss % 8 == 0 && p.k_ss % 8 == 0 && p.v_ss % 8 == 0 && p.o_ss % 8 == 0 && p.q_sh % 8 == 0 &&
                    p.k_sh % 8 == 0 && p.v_sh % 8 == 0 && p.o_sh % 8 == 0 && p.q_sb % 8 == 0 && p.k_sb % 8 == 0 &&
                    p.v_sb % 8 == 0 && p.o_sb % 8 == 0,
                "flash_attn: strides must be multiples of 8 elements (16-byte rows)");
    return D == 64 ? fwd_launch<64>(p, causal, (cudaStream_t)stream) : fwd_launch<128>(p, causal, (cudaStream_t)stream);
}

NXD_API int nxd_flash_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* dout,
                               const float* lse, float* delta, void* dq, void* dk, void* dv,
                               const long long* strides, int B, int Hq, int Hkv, int Sq, int Sk, int D, float scale,
                               int causal, void* stream) {
    FaParams p{};
    p.q = (const bf16*)q; p.k = (const bf16*)k; p.v = (const bf16*)v; p.o = (bf16*)const_cast<void*>(o);
    p.dout = (const bf16*)dout; p.lse = const_cast<float*>(lse); p.delta = delta;
    p.dq = (bf16*)dq; p.dk = (bf16*)dk; p.dv = (bf16*)dv;
    p.q_sb = strides[0]; p.q_sh = strides[1]; p.q_ss = strides[2];
    p.k_sb = strides[3]; p.k_sh = strides[4]; p.k_ss = strides[5];
    p.v_sb = strides[6]; p.v_sh = strides[7]; p.v_ss = strides[8];
    p.o_sb = strides[9]; p.o_sh = strides[10]; p.o_ss = strides[11];
    p.do_sb = strides[12]; p.do_sh = strides[13]; p.do_ss = strides[14];
    p.dq_sb = strides[15]; p.dq_sh = strides[16]; p.dq_ss = strides[17];
    p.dk_sb = strides[18]; p.dk_sh = strides[19]; p.dk_ss = strides[20];
    p.dv_sb = strides[21]; p.dv_sh = strides[22]; p.dv_ss = strides[23];
    p.B = B; p.Hq = Hq; p.Hkv = Hkv; p.Sq = Sq; p.Sk = Sk;
    p.scale = scale; p.scale_log2 = scale * 1.4426950408889634f;
    if (int rc = check_common(p, D)) return rc;
    for (int i = 0; i < 24; ++i)
        NXD_REQUIRE(strides[i] % 8 == 0, "flash_attn_bwd: stride %d (=%lld) must be a multiple of 8 elements", i,
                    strides[i]);
    return D == 64 ? bwd_launch<64>(p, causal, (cudaStream_t)stream) : bwd_launch<128>(p, causal, (cudaStream_t)stream);
}
