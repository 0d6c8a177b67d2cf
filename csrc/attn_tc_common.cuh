// Helpers shared by the tcgen05 attention kernels (flash_attn_tc.cu forward, flash_attn_bwd_tc.cu backward).
#pragma once
#include "common.cuh"
#include "ptx.cuh"

namespace nxd_attn {

// 4-D map over a strided [B, H, S, D] view, inner box = 128 bytes (64 bf16 / 32 fp32), SWIZZLE_128B.  The three outer
// dims are ordered by ascending stride; pos[] returns the coordinate slot (1..3) of (row, head, batch).
int make_tmap_bhsd(CUtensorMap* out, const void* base, int D, long long S, long long H, long long B, long long s_s,
                   long long s_h, long long s_b, int box_rows, int* pos, int elem_bytes = 2);

__device__ __forceinline__ void tma_load_4d(const CUtensorMap* map, uint64_t* bar, void* dst, int c0, int c1, int c2,
                                            int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(ptx::smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(ptx::smem_u32(bar)), "r"(c0), "r"(c1),
        "r"(c2), "r"(c3)
        : "memory");
}
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* map, const void* src, int c0, int c1, int c2, int c3) {
    asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
                     reinterpret_cast<uint64_t>(map)),
                 "r"(ptx::smem_u32(src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
                 : "memory");
}
// Coordinates 1..3 of a 4-D map whose (row, head, batch) dims sit in slots pos[0..2].  Written as selects on purpose:
// indexing a local array with pos[] would put the array in local memory (STL / LDL per TMA issue).
__device__ __forceinline__ void coords(const int* pos, int row, int h, int b, int* c) {
    const int pr = pos[0], ph = pos[1];
    c[1] = pr == 1 ? row : (ph == 1 ? h : b);
    c[2] = pr == 2 ? row : (ph == 2 ? h : b);
    c[3] = pr == 3 ? row : (ph == 3 ? h : b);
}

using ptx::align_smem_1024;

__device__ __forceinline__ void st_shared_v4(void* dst, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(ptx::smem_u32(dst)), "r"(a), "r"(b), "r"(c), "r"(d)
                 : "memory");
}
__device__ __forceinline__ float4 ld_shared_f4(const void* src) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "r"(ptx::smem_u32(src)));
    return v;
}
__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}

__device__ __forceinline__ void tmem_st_32x32b_x32(uint32_t taddr, const uint32_t* r) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
        "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
        "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
        "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
        : "memory");
}
__device__ __forceinline__ void tmem_st_32x32b_x16(uint32_t taddr, const uint32_t* r) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
        "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }


// 1-D bulk copy global -> shared, completion counted on an mbarrier (bytes % 16 == 0, both sides 16-byte aligned)
__device__ __forceinline__ void bulk_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     ptx::smem_u32(dst)),
                 "l"(reinterpret_cast<uint64_t>(src)), "r"(bytes), "r"(ptx::smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void tma_reduce_add_4d(const CUtensorMap* map, const void* src, int c0, int c1, int c2,
                                                  int c3) {
    asm volatile(
        "cp.reduce.async.bulk.tensor.4d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
            reinterpret_cast<uint64_t>(map)),
        "r"(ptx::smem_u32(src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
// Attention dropout (reference: nki_flash_attn_func(..., dropout_p, ..., seed), kernels/flash_attn.py:122-127,151-160).
// The keep decision of element (batch b, q head h, query i, key j) is a pure function of (seed, b, h, i, j), so the
// forward pass (thread = query row) and the backward pass (thread = key row) regenerate the same mask without storing
// it: a 32-bit integer hash (two multiply-xorshift rounds) compared with dropout_p * 2^32.
struct DropParams {
    uint32_t threshold;   // drop iff hash < threshold; 0 = no dropout
    float inv_keep;       // 1 / (1 - dropout_p)
    uint32_t seed;
};
__host__ __device__ __forceinline__ uint32_t drop_head_seed(uint32_t seed, uint32_t bh) { return seed ^ (bh * 0x9E3779B1u); }
__host__ __device__ __forceinline__ bool drop_keep(uint32_t head_seed, uint32_t i, uint32_t j, uint32_t threshold) {
    uint32_t x = (i * 0x85EBCA77u) ^ (j * 0xC2B2AE3Du) ^ head_seed;
    x ^= x >> 16;
    x *= 0x7FEB352Du;
    x ^= x >> 15;
    x *= 0x846CA68Bu;
    x ^= x >> 16;
    return x >= threshold;
}

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

}  // namespace nxd_attn
