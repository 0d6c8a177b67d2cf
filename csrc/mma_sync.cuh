// Warp-level tensor-core helpers (mma.sync m16n8k16 bf16, ldmatrix, cp.async) for kernels whose tiles are too small for
// tcgen05 (M = 16 query heads of one KV head in token generation).
#pragma once
#include "common.cuh"

namespace mma {

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
    const int sz = valid ? 16 : 0;      // src-size 0: the 16 bytes are zero-filled
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
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

// smem tile [rows][D] bf16, 16-byte chunks XOR-swizzled with the row index (conflict-free ldmatrix).
template <int D>
__device__ __forceinline__ uint32_t tile_addr(uint32_t base, int row, int chunk) {
    return base + row * (D * 2) + ((chunk ^ (row & 7)) << 4);
}
// B fragments for two adjacent 8-wide n blocks from an [n][D] tile (contraction over D): {b0, b1} of block n0, {b0, b1}
// of block n0 + 8.
template <int D>
__device__ __forceinline__ void load_b(uint32_t (&b)[4], uint32_t sbase, int n0, int kk, int lane) {
    const int row = n0 + (lane & 7) + (lane >> 4) * 8;
    ldsm_x4(b, tile_addr<D>(sbase, row, kk * 2 + ((lane >> 3) & 1)));
}
// B fragments for two adjacent 8-wide d blocks from a [k][D] tile (contraction over the 16 rows k0 .. k0 + 15).
template <int D>
__device__ __forceinline__ void load_bt(uint32_t (&b)[4], uint32_t sbase, int k0, int dj, int lane) {
    const int row = k0 + (lane & 7) + ((lane >> 3) & 1) * 8;
    ldsm_x4_t(b, tile_addr<D>(sbase, row, dj * 2 + (lane >> 4)));
}

}  // namespace mma
