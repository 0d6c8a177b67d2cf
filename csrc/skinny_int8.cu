// Weight-only int8 skinny GEMM for token generation (reference quantization_layers.py:444-478 dequantises
// the weight and calls the float matmul -- on Neuron the compiler fuses that, here it would read 1 byte and write + read
// 2 more per weight element).
//
//   nxd_skinny_gemm_i8   out[B, N] (+ residual) = (x[B, K] (bf16) * w[N, K]^T (int8)) * scale      B <= 8
//
// Token generation is weight-bandwidth bound, so the point of int8 weights is HALF THE BYTES of the bf16 skinny GEMM
// (fused_gemm_ar.cu), not int8 math: weights are streamed once with 16-byte loads (16 weights per lane and step), turned
// into fp32 in registers and multiplied with the bf16 activations held in shared memory, fp32 accumulation.  The
// int8 -> fp32 conversion does not use I2F (quarter rate: 8 TB/s of weights would need ~28 conversions per clock and
// SM): a byte u = s ^ 0x80 (one LOP3 per four weights) is placed into the mantissa of 2^23 with PRMT, giving the float
// 2^23 + u exactly, and 2^23 + 128 is subtracted -- one PRMT + one FADD per weight, shared by all B rows.  The
// subtraction is done before the multiply on purpose: folding it into the accumulation (sum(f * x) - c * sum(x)) would
// cancel numbers of magnitude 2^23 |x| in fp32.
//
// scale: one float (per tensor) or N floats (per output channel); applied to the fp32 sum before the bf16 rounding, i.e.
// before the tensor-parallel all-reduce of a row-parallel layer, as the float layers do.
//
// STATUS: written after the round's GPU budget was spent -- compiles for sm_100a, numerics not yet run on hardware.  The
// Python side uses it only with NXD_INT8_SKINNY=1.
#include "common.cuh"

namespace {

constexpr int kI8Threads = 256;
constexpr int kI8Warps = kI8Threads / 32;

// 4 int8 (one 32-bit word, already XOR-ed with 0x80808080) -> 4 floats
__device__ __forceinline__ void cvt4_i8(uint32_t u, float* f) {
    const float kBias = 8388736.0f;                      // 2^23 + 128
    f[0] = __uint_as_float(__byte_perm(u, 0x4B000000u, 0x7440)) - kBias;
    f[1] = __uint_as_float(__byte_perm(u, 0x4B000000u, 0x7441)) - kBias;
    f[2] = __uint_as_float(__byte_perm(u, 0x4B000000u, 0x7442)) - kBias;
    f[3] = __uint_as_float(__byte_perm(u, 0x4B000000u, 0x7443)) - kBias;
}

// 16 int8 weights (uint4) -> 16 floats, in memory order
__device__ __forceinline__ void cvt16_i8(const uint4& a, float* f) {
    cvt4_i8(a.x ^ 0x80808080u, f);
    cvt4_i8(a.y ^ 0x80808080u, f + 4);
    cvt4_i8(a.z ^ 0x80808080u, f + 8);
    cvt4_i8(a.w ^ 0x80808080u, f + 12);
}

// acc += sum_i wf[i] * x[i] for 8 bf16 activations packed in a uint4
__device__ __forceinline__ float dot8_f32_bf16(const float* wf, const uint4& x, float acc) {
    const uint32_t xw[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        acc = fmaf(wf[2 * j], __uint_as_float(xw[j] << 16), acc);              // low half  = element 2j
        acc = fmaf(wf[2 * j + 1], __uint_as_float(xw[j] & 0xffff0000u), acc);  // high half = element 2j + 1
    }
    return acc;
}

// Shared memory: X tile [kB][K] bf16, then the result block [kB][cols] fp32.  A CTA owns `cols` output columns; every
// warp walks them two at a time (two independent 16-byte weight streams per lane), lanes split K in 16-element steps.
template <int kB>
__global__ void __launch_bounds__(kI8Threads) skinny_i8_kernel(const bf16* __restrict__ x, const int8_t* __restrict__ w,
                                                               const float* __restrict__ scale, int scale_n,
                                                               const bf16* __restrict__ residual, bf16* __restrict__ out,
                                                               int B, int N, int K, long long ldx, long long ldw,
                                                               long long ldo, int cols) {
    extern __shared__ __align__(16) uint8_t i8_smem[];
    bf16* xs = reinterpret_cast<bf16*>(i8_smem);
    float* part = reinterpret_cast<float*>(i8_smem + (size_t)kB * K * sizeof(bf16));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int k8 = K / 8, k16 = K / 16;
    for (int i = threadIdx.x; i < kB * k8; i += kI8Threads) {
        const int row = i / k8, c = i % k8;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (row < B) v = *reinterpret_cast<const uint4*>(x + (size_t)row * ldx + (size_t)c * 8);
        reinterpret_cast<uint4*>(xs)[i] = v;
    }
    __syncthreads();
    const int c0 = (int)blockIdx.x * cols;
    const int ncols = min(cols, N - c0);
    for (int cc = warp * 2; cc < cols; cc += kI8Warps * 2) {
        float acc0[kB], acc1[kB];
#pragma unroll
        for (int b = 0; b < kB; ++b) acc0[b] = acc1[b] = 0.f;
        const bool v0 = cc < ncols, v1 = cc + 1 < ncols;
        const int8_t* w0 = w + (size_t)(c0 + cc) * ldw;
        const int8_t* w1 = w0 + ldw;
#pragma unroll (kB <= 2 ? 4 : 2)
        for (int k = lane; k < k16; k += 32) {
            const uint4 a0 = v0 ? __ldg(reinterpret_cast<const uint4*>(w0) + k) : make_uint4(0u, 0u, 0u, 0u);
            const uint4 a1 = v1 ? __ldg(reinterpret_cast<const uint4*>(w1) + k) : make_uint4(0u, 0u, 0u, 0u);
            float f0[16], f1[16];
            cvt16_i8(a0, f0);
            cvt16_i8(a1, f1);
#pragma unroll
            for (int b = 0; b < kB; ++b) {
                const uint4 xlo = reinterpret_cast<const uint4*>(xs)[b * k8 + 2 * k];
                const uint4 xhi = reinterpret_cast<const uint4*>(xs)[b * k8 + 2 * k + 1];
                acc0[b] = dot8_f32_bf16(f0, xlo, acc0[b]);
                acc0[b] = dot8_f32_bf16(f0 + 8, xhi, acc0[b]);
                acc1[b] = dot8_f32_bf16(f1, xlo, acc1[b]);
                acc1[b] = dot8_f32_bf16(f1 + 8, xhi, acc1[b]);
            }
        }
#pragma unroll
        for (int b = 0; b < kB; ++b) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                acc0[b] += __shfl_xor_sync(0xffffffffu, acc0[b], o);
                acc1[b] += __shfl_xor_sync(0xffffffffu, acc1[b], o);
            }
        }
        if (lane == 0) {
            const float s0 = v0 ? scale[scale_n > 1 ? c0 + cc : 0] : 0.f;
            const float s1 = v1 ? scale[scale_n > 1 ? c0 + cc + 1 : 0] : 0.f;
#pragma unroll
            for (int b = 0; b < kB; ++b) {
                part[b * cols + cc] = acc0[b] * s0;
                part[b * cols + cc + 1] = acc1[b] * s1;
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < B * ncols; i += kI8Threads) {
        const int b = i / ncols, c = i % ncols;
        float v = part[b * cols + c];
        if (residual) v += __bfloat162float(residual[(size_t)b * ldo + c0 + c]);
        out[(size_t)b * ldo + c0 + c] = __float2bfloat16(v);
    }
}

template <int kB>
int launch_i8(const bf16* x, const int8_t* w, const float* scale, int scale_n, const bf16* residual, bf16* out, int B, int N,
              int K, long long ldx, long long ldw, long long ldo, cudaStream_t st) {
    // 16 columns per CTA = one column pair per warp: the most CTAs, all weight streams in flight at once
    const int cols = 16;
    const size_t smem = (size_t)kB * K * sizeof(bf16) + (size_t)kB * cols * sizeof(float);
    NXD_REQUIRE(smem <= 200 * 1024, "skinny_gemm_i8: B %d x K %d does not fit shared memory", B, K);
    NXD_CHECK_CUDA(cudaFuncSetAttribute(skinny_i8_kernel<kB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int grid = (N + cols - 1) / cols;
    skinny_i8_kernel<kB><<<grid, kI8Threads, smem, st>>>(x, w, scale, scale_n, residual, out, B, N, K, ldx, ldw, ldo, cols);
    NXD_LAUNCH_CHECK();
    return 0;
}

}  // namespace

// out[B, N] (+ residual[B, N], pitch ldo) = (x[B, K] * w[N, K]^T) * scale.  x, residual, out bf16; w int8 with row pitch
// ldw bytes; scale fp32 with scale_n == 1 (per tensor) or scale_n == N (per output channel).  B <= 8, K % 16 == 0,
// ldx % 8 == 0, ldw % 16 == 0, rows 16-byte aligned.
NXD_API int nxd_skinny_gemm_i8(const void* x, const void* w, const void* scale, int scale_n, const void* residual, void* out,
                               int B, int N, int K, long long ldx, long long ldw, long long ldo, void* stream) {
    NXD_REQUIRE(B >= 1 && B <= 8, "skinny_gemm_i8: B %d not in [1, 8]", B);
    NXD_REQUIRE(K % 16 == 0 && ldx % 8 == 0 && ldw % 16 == 0, "skinny_gemm_i8: K / ldw must be multiples of 16, ldx of 8");
    NXD_REQUIRE(scale_n == 1 || scale_n == N, "skinny_gemm_i8: %d scales for %d output channels", scale_n, N);
    const bf16 *xb = (const bf16*)x, *rb = (const bf16*)residual;
    const int8_t* wb = (const int8_t*)w;
    const float* sc = (const float*)scale;
    cudaStream_t st = (cudaStream_t)stream;
    if (B <= 1) return launch_i8<1>(xb, wb, sc, scale_n, rb, (bf16*)out, B, N, K, ldx, ldw, ldo, st);
    if (B <= 2) return launch_i8<2>(xb, wb, sc, scale_n, rb, (bf16*)out, B, N, K, ldx, ldw, ldo, st);
    if (B <= 4) return launch_i8<4>(xb, wb, sc, scale_n, rb, (bf16*)out, B, N, K, ldx, ldw, ldo, st);
    return launch_i8<8>(xb, wb, sc, scale_n, rb, (bf16*)out, B, N, K, ldx, ldw, ldo, st);
}
