// SwiGLU and RoPE for sm_100a.
//
// SwiGLU: reference examples/training/llama/modeling_llama_nxd.py:197-208 -- the fused
// gate_up projection (ColumnParallel, stride=2) yields [T, 2*I] with the local gate
// chunk first and the local up chunk second; out = silu(gate) * up.
//
// RoPE: reference modeling_llama_nxd.py:398-399 (HF rotate-half) and the inference
// variant examples/inference/modules/attention/utils.py:27-41.  Applied IN PLACE
// on the fused QKV GEMM output viewed as [T, heads_total, D]: the first
// `n_rot_heads` heads (q heads then k heads) are rotated, v heads are untouched.
// Backward is the same kernel with the sine negated (inverse rotation).
#include "common.cuh"

namespace {

__global__ void swiglu_fwd_kernel(const bf16* __restrict__ gu, bf16* __restrict__ out, size_t rows, int I) {
    const int nvec = I >> 3;
    const size_t total = rows * (size_t)nvec;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (size_t)gridDim.x * blockDim.x) {
        const size_t r = idx / nvec;
        const int c = (int)(idx % nvec);
        const bf16x8* g = reinterpret_cast<const bf16x8*>(gu + r * 2 * (size_t)I);
        const bf16x8* u = reinterpret_cast<const bf16x8*>(gu + r * 2 * (size_t)I + I);
        float gv[8], uv[8], o[8];
        bf16x8 pg = g[c], pu = u[c];
        unpack8(pg, gv);
        unpack8(pu, uv);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float sg = 1.f / (1.f + __expf(-gv[k]));
            o[k] = gv[k] * sg * uv[k];
        }
        reinterpret_cast<bf16x8*>(out + r * (size_t)I)[c] = pack8(o);
    }
}

__global__ void swiglu_bwd_kernel(const bf16* __restrict__ gu, const bf16* __restrict__ dy,
                                  bf16* __restrict__ dgu, size_t rows, int I) {
    const int nvec = I >> 3;
    const size_t total = rows * (size_t)nvec;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (size_t)gridDim.x * blockDim.x) {
        const size_t r = idx / nvec;
        const int c = (int)(idx % nvec);
        const bf16x8* g = reinterpret_cast<const bf16x8*>(gu + r * 2 * (size_t)I);
        const bf16x8* u = reinterpret_cast<const bf16x8*>(gu + r * 2 * (size_t)I + I);
        const bf16x8* d = reinterpret_cast<const bf16x8*>(dy + r * (size_t)I);
        float gv[8], uv[8], dv[8], dg[8], du[8];
        bf16x8 pg = g[c], pu = u[c], pd = d[c];
        unpack8(pg, gv);
        unpack8(pu, uv);
        unpack8(pd, dv);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float sg = 1.f / (1.f + __expf(-gv[k]));
            const float silu = gv[k] * sg;
            dg[k] = dv[k] * uv[k] * sg * (1.f + gv[k] * (1.f - sg));
            du[k] = dv[k] * silu;
        }
        reinterpret_cast<bf16x8*>(dgu + r * 2 * (size_t)I)[c] = pack8(dg);
        reinterpret_cast<bf16x8*>(dgu + r * 2 * (size_t)I + I)[c] = pack8(du);
    }
}

// One thread rotates 8 (x1, x2) pairs: columns [8c, 8c+8) and [D/2 + 8c, ...).
__global__ void rope_kernel(bf16* __restrict__ x, const float* __restrict__ cos_t,
                            const float* __restrict__ sin_t, const int* __restrict__ positions, size_t T,
                            int heads_total, int n_rot_heads, int D, int pos_div, int pos_mod, float sin_sign) {
    const int half_vec = D >> 4;  // 16-byte vectors in half a head
    const size_t total = T * (size_t)n_rot_heads * half_vec;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (size_t)gridDim.x * blockDim.x) {
        const int c = (int)(idx % half_vec);
        const size_t th = idx / half_vec;
        const int h = (int)(th % n_rot_heads);
        const size_t t = th / n_rot_heads;
        const int pos = positions ? positions[t] : (int)((t / pos_div) % pos_mod);
        bf16* base = x + (t * heads_total + h) * (size_t)D;
        bf16x8* p1 = reinterpret_cast<bf16x8*>(base) + c;
        bf16x8* p2 = reinterpret_cast<bf16x8*>(base + (D >> 1)) + c;
        float a[8], b[8], oa[8], ob[8];
        bf16x8 va = *p1, vb = *p2;
        unpack8(va, a);
        unpack8(vb, b);
        const float4* cp = reinterpret_cast<const float4*>(cos_t + (size_t)pos * (D >> 1) + c * 8);
        const float4* sp = reinterpret_cast<const float4*>(sin_t + (size_t)pos * (D >> 1) + c * 8);
        const float4 c0 = cp[0], c1 = cp[1], s0 = sp[0], s1 = sp[1];
        const float cs[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
        const float sn[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float s = sn[k] * sin_sign;
            oa[k] = a[k] * cs[k] - b[k] * s;
            ob[k] = b[k] * cs[k] + a[k] * s;
        }
        *p1 = pack8(oa);
        *p2 = pack8(ob);
    }
}

inline int grid_for(size_t work_items, int threads) {
    size_t blocks = (work_items + threads - 1) / threads;
    const size_t cap = (size_t)nxd_num_sms() * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

}  // namespace

NXD_API int nxd_swiglu_fwd(const void* gate_up, void* out, long long rows, int I, void* stream) {
    NXD_REQUIRE(I % 8 == 0, "swiglu: I=%d must be a multiple of 8", I);
    if (rows == 0) return 0;
    const size_t work = (size_t)rows * (I / 8);
    swiglu_fwd_kernel<<<grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)gate_up, (bf16*)out,
                                                                             (size_t)rows, I);
    NXD_LAUNCH_CHECK();
    return 0;
}

NXD_API int nxd_swiglu_bwd(const void* gate_up, const void* dy, void* dgate_up, long long rows, int I,
                           void* stream) {
    NXD_REQUIRE(I % 8 == 0, "swiglu_bwd: I=%d must be a multiple of 8", I);
    if (rows == 0) return 0;
    const size_t work = (size_t)rows * (I / 8);
    swiglu_bwd_kernel<<<grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>(
        (const bf16*)gate_up, (const bf16*)dy, (bf16*)dgate_up, (size_t)rows, I);
    NXD_LAUNCH_CHECK();
    return 0;
}

// x: [T, heads_total, D] bf16 (in place).  cos/sin: [max_pos, D/2] fp32.
// position of token t = positions[t] if given, else (t / pos_div) % pos_mod.
// inverse != 0 applies the transpose rotation (backward).
NXD_API int nxd_rope_inplace(void* x, const float* cos_t, const float* sin_t, const int* positions, long long T,
                             int heads_total, int n_rot_heads, int D, int pos_div, int pos_mod, int inverse,
                             void* stream) {
    NXD_REQUIRE(D % 16 == 0, "rope: head dim %d must be a multiple of 16", D);
    NXD_REQUIRE(n_rot_heads <= heads_total, "rope: n_rot_heads > heads_total");
    if (T == 0 || n_rot_heads == 0) return 0;
    const size_t work = (size_t)T * n_rot_heads * (D / 16);
    rope_kernel<<<grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>(
        (bf16*)x, cos_t, sin_t, positions, (size_t)T, heads_total, n_rot_heads, D, pos_div < 1 ? 1 : pos_div,
        pos_mod < 1 ? 1 : pos_mod, inverse ? -1.f : 1.f);
    NXD_LAUNCH_CHECK();
    return 0;
}
