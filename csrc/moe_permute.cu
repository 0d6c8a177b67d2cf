// MoE token permutation kernels (SURVEY C18, §2.5 "MoE routing/permutation"): the capacity-factor dispatch and combine of
// reference src/neuronx_distributed/modules/moe/expert_mlps.py:183-265 (gather of token rows into the expert-aligned
// [E, C, H] buffer keyed by token_assignments, affinity-weighted scatter-combine back to [T, H]) as two bandwidth-bound
// kernels and their gradients, instead of chains of indexing ops:
//
//   nxd_moe_gather      out[r]  = slot_token[r] >= 0 ? scale[r] * x[slot_token[r]] : 0        r over E*C slots
//   nxd_moe_combine     out[t]  = sum_j w[t, j] * y[slot_of[t, j]]                              (w = 0 for dropped pairs)
//   nxd_moe_combine_dw  dw[t,j] = <dout[t], y[slot_of[t, j]]>
//
// Forward dispatch = gather (scale NULL); its backward = combine with w = kept (1 / 0).  Forward combine = combine; its
// backward w.r.t. y = gather of dout rows with scale = affinity of the slot, w.r.t. the affinities = combine_dw.
#include "common.cuh"

namespace {

// one warp per output row, 16-byte vectors
__global__ void moe_gather_kernel(const bf16* __restrict__ x, const int* __restrict__ slot_token,
                                  const float* __restrict__ scale, bf16* __restrict__ out, long long slots, int H) {
    const int lane = threadIdx.x & 31;
    const long long r = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= slots) return;
    const int t = slot_token[r];
    const int nvec = H >> 3;
    bf16x8* dst = reinterpret_cast<bf16x8*>(out + r * H);
    if (t < 0) {
        bf16x8 z;
#pragma unroll
        for (int i = 0; i < 4; ++i) z.v[i] = __floats2bfloat162_rn(0.f, 0.f);
        for (int c = lane; c < nvec; c += 32) dst[c] = z;
        return;
    }
    const bf16x8* src = reinterpret_cast<const bf16x8*>(x + (long long)t * H);
    if (scale == nullptr) {
        for (int c = lane; c < nvec; c += 32) dst[c] = src[c];
        return;
    }
    const float s = scale[r];
    for (int c = lane; c < nvec; c += 32) {
        float f[8];
        unpack8(src[c], f);
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] *= s;
        dst[c] = pack8(f);
    }
}

// one warp per token
__global__ void moe_combine_kernel(const bf16* __restrict__ y, const int* __restrict__ slot_of,
                                   const float* __restrict__ w, bf16* __restrict__ out, long long T, int k, int H) {
    const int lane = threadIdx.x & 31;
    const long long t = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (t >= T) return;
    const int nvec = H >> 3;
    for (int c = lane; c < nvec; c += 32) {
        float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int j = 0; j < k; ++j) {
            const float wj = w[t * k + j];
            if (wj == 0.f) continue;
            float f[8];
            unpack8(reinterpret_cast<const bf16x8*>(y + (long long)slot_of[t * k + j] * H)[c], f);
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e] = fmaf(wj, f[e], acc[e]);
        }
        reinterpret_cast<bf16x8*>(out + t * H)[c] = pack8(acc);
    }
}

// one warp per (token, choice)
__global__ void moe_combine_dw_kernel(const bf16* __restrict__ y, const bf16* __restrict__ dout,
                                      const int* __restrict__ slot_of, const unsigned char* __restrict__ kept,
                                      float* __restrict__ dw, long long pairs, int k, int H) {
    const int lane = threadIdx.x & 31;
    const long long p = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (p >= pairs) return;
    if (!kept[p]) {
        if (lane == 0) dw[p] = 0.f;
        return;
    }
    const long long t = p / k;
    const bf16x8* a = reinterpret_cast<const bf16x8*>(y + (long long)slot_of[p] * H);
    const bf16x8* b = reinterpret_cast<const bf16x8*>(dout + t * H);
    float acc = 0.f;
    for (int c = lane; c < (H >> 3); c += 32) {
        float fa[8], fb[8];
        unpack8(a[c], fa);
        unpack8(b[c], fb);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc = fmaf(fa[e], fb[e], acc);
    }
    acc = warp_sum(acc);
    if (lane == 0) dw[p] = acc;
}

inline unsigned blocks_for(long long rows) { return (unsigned)((rows + 7) / 8); }   // 8 warps per block

}  // namespace

// x [T, H] bf16; slot_token [slots] int32 (token held by each expert slot, -1 = empty); scale [slots] fp32 or NULL;
// out [slots, H].  H % 8 == 0.
NXD_API int nxd_moe_gather(const void* x, const int* slot_token, const float* scale, void* out, long long slots, int H,
                           void* stream) {
    NXD_REQUIRE(H % 8 == 0, "moe_gather: H=%d must be a multiple of 8", H);
    if (slots == 0) return 0;
    moe_gather_kernel<<<blocks_for(slots), 256, 0, (cudaStream_t)stream>>>((const bf16*)x, slot_token, scale, (bf16*)out,
                                                                           slots, H);
    NXD_LAUNCH_CHECK();
    return 0;
}

// y [slots, H] bf16; slot_of [T, k] int32; w [T, k] fp32 (0 for dropped pairs: their slot index is not read); out [T, H].
NXD_API int nxd_moe_combine(const void* y, const int* slot_of, const float* w, void* out, long long T, int k, int H,
                            void* stream) {
    NXD_REQUIRE(H % 8 == 0 && k >= 1, "moe_combine: H=%d must be a multiple of 8, k=%d >= 1", H, k);
    if (T == 0) return 0;
    moe_combine_kernel<<<blocks_for(T), 256, 0, (cudaStream_t)stream>>>((const bf16*)y, slot_of, w, (bf16*)out, T, k, H);
    NXD_LAUNCH_CHECK();
    return 0;
}

// dw [T, k] fp32 = <dout[t], y[slot_of[t, j]]> for kept pairs, 0 otherwise.  kept [T, k] uint8.
NXD_API int nxd_moe_combine_dw(const void* y, const void* dout, const int* slot_of, const unsigned char* kept, float* dw,
                               long long T, int k, int H, void* stream) {
    NXD_REQUIRE(H % 8 == 0 && k >= 1, "moe_combine_dw: H=%d must be a multiple of 8, k=%d >= 1", H, k);
    if (T == 0) return 0;
    moe_combine_dw_kernel<<<blocks_for(T * k), 256, 0, (cudaStream_t)stream>>>((const bf16*)y, (const bf16*)dout, slot_of,
                                                                               kept, dw, T * k, k, H);
    NXD_LAUNCH_CHECK();
    return 0;
}
