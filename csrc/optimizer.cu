// Flat-buffer optimizer kernels for the ZeRO-1 step.
//
// The reference composes torch_xla's ZeroRedundancyOptimizer with a Python AdamW
// (src/neuronx_distributed/optimizer/zero_redundancy_optimizer.py:29-156,
//  src/neuronx_distributed/utils/adamw_fp32_optim_params.py:88-153) and computes the grad norm with
// per-parameter torch.norm calls (parallel_layers/grads.py:33-177).  Here every rank owns
// one contiguous fp32 master shard + exp_avg + exp_avg_sq and the whole local step is:
//   sumsq (two-stage, deterministic)  ->  [scalar all-reduce on host stream]  ->
//   adamw_step which reads the clip coefficient FROM DEVICE MEMORY (no host sync), updates
//   the fp32 state and writes the bf16 parameter shard that is then all-gathered.
#include "common.cuh"

namespace {

template <typename T>
__device__ __forceinline__ float to_f(T v);
template <>
__device__ __forceinline__ float to_f<float>(float v) { return v; }
template <>
__device__ __forceinline__ float to_f<bf16>(bf16 v) { return __bfloat162float(v); }

template <typename T>
__global__ void sumsq_partial_kernel(const T* __restrict__ x, size_t n, float* __restrict__ partial) {
    __shared__ float red[32];
    float acc = 0.f;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float v = to_f<T>(x[i]);
        acc += v * v;
    }
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) partial[blockIdx.x] = acc;
}

__global__ void sumsq_final_kernel(const float* __restrict__ partial, int P, float* __restrict__ out,
                                   int accumulate) {
    __shared__ float red[32];
    float acc = 0.f;
    for (int i = threadIdx.x; i < P; i += blockDim.x) acc += partial[i];
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) out[0] = accumulate ? out[0] + acc : acc;
}

// clip_coef[0] = min(1, max_norm / (sqrt(total_sumsq) + 1e-6));  norm_out[0] = sqrt(total_sumsq)
__global__ void clip_coef_kernel(const float* __restrict__ total_sumsq, float max_norm, float* __restrict__ coef,
                                 float* __restrict__ norm_out) {
    const float norm = sqrtf(total_sumsq[0]);
    if (norm_out) norm_out[0] = norm;
    float c = 1.f;
    if (max_norm > 0.f) c = fminf(1.f, max_norm / (norm + 1e-6f));
    coef[0] = c;
}

template <typename G>
__global__ void adamw_kernel(float* __restrict__ master, float* __restrict__ m, float* __restrict__ v,
                             const G* __restrict__ grad, bf16* __restrict__ param_out, size_t n, float lr,
                             float beta1, float beta2, float eps, float wd, float bc1, float bc2,
                             float grad_scale, const float* __restrict__ clip_coef) {
    const float scale = grad_scale * (clip_coef ? clip_coef[0] : 1.f);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float g = to_f<G>(grad[i]) * scale;
        float mi = m[i], vi = v[i], p = master[i];
        mi = beta1 * mi + (1.f - beta1) * g;
        vi = beta2 * vi + (1.f - beta2) * g * g;
        const float mhat = mi / bc1;
        const float vhat = vi / bc2;
        p = p * (1.f - lr * wd) - lr * mhat / (sqrtf(vhat) + eps);
        m[i] = mi;
        v[i] = vi;
        master[i] = p;
        if (param_out) param_out[i] = __float2bfloat16(p);
    }
}

template <typename T>
__global__ void scale_kernel(T* __restrict__ x, size_t n, float s, const float* __restrict__ dev_scale) {
    const float k = s * (dev_scale ? dev_scale[0] : 1.f);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        if constexpr (sizeof(T) == 4) x[i] = x[i] * k;
        else x[i] = __float2bfloat16(__bfloat162float(x[i]) * k);
    }
}

inline int blocks_for(size_t n) {
    size_t b = (n + 1023) / 1024;
    const size_t cap = (size_t)nxd_num_sms() * 8;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (int)b;
}

}  // namespace

NXD_API int nxd_sumsq_num_partials() { return nxd_num_sms() * 8; }

// out[0] (+)= sum(x^2).  dtype: 0 = fp32, 1 = bf16.  scratch: nxd_sumsq_num_partials() floats.
NXD_API int nxd_sumsq(const void* x, long long n, int dtype, float* scratch, float* out, int accumulate,
                      void* stream) {
    NXD_REQUIRE(scratch != nullptr && out != nullptr, "sumsq: scratch/out required");
    cudaStream_t s = (cudaStream_t)stream;
    const int P = blocks_for((size_t)n);
    if (dtype == 0) sumsq_partial_kernel<float><<<P, 256, 0, s>>>((const float*)x, (size_t)n, scratch);
    else sumsq_partial_kernel<bf16><<<P, 256, 0, s>>>((const bf16*)x, (size_t)n, scratch);
    NXD_LAUNCH_CHECK();
    sumsq_final_kernel<<<1, 256, 0, s>>>(scratch, P, out, accumulate);
    NXD_LAUNCH_CHECK();
    return 0;
}

NXD_API int nxd_clip_coef(const float* total_sumsq, float max_norm, float* coef, float* norm_out, void* stream) {
    clip_coef_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(total_sumsq, max_norm, coef, norm_out);
    NXD_LAUNCH_CHECK();
    return 0;
}

// grad_dtype: 0 = fp32, 1 = bf16.  clip_coef may be null.  param_out (bf16) may be null.
NXD_API int nxd_adamw_step(float* master, float* exp_avg, float* exp_avg_sq, const void* grad, int grad_dtype,
                           void* param_out, long long n, float lr, float beta1, float beta2, float eps,
                           float weight_decay, int step, float grad_scale, const float* clip_coef,
                           void* stream) {
    if (n == 0) return 0;
    const float bc1 = 1.f - powf(beta1, (float)step);
    const float bc2 = 1.f - powf(beta2, (float)step);
    cudaStream_t s = (cudaStream_t)stream;
    const int B = blocks_for((size_t)n);
    if (grad_dtype == 0)
        adamw_kernel<float><<<B, 256, 0, s>>>(master, exp_avg, exp_avg_sq, (const float*)grad, (bf16*)param_out,
                                              (size_t)n, lr, beta1, beta2, eps, weight_decay, bc1, bc2,
                                              grad_scale, clip_coef);
    else
        adamw_kernel<bf16><<<B, 256, 0, s>>>(master, exp_avg, exp_avg_sq, (const bf16*)grad, (bf16*)param_out,
                                             (size_t)n, lr, beta1, beta2, eps, weight_decay, bc1, bc2,
                                             grad_scale, clip_coef);
    NXD_LAUNCH_CHECK();
    return 0;
}

NXD_API int nxd_scale_inplace(void* x, long long n, int dtype, float scale, const float* dev_scale, void* stream) {
    if (n == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    const int B = blocks_for((size_t)n);
    if (dtype == 0) scale_kernel<float><<<B, 256, 0, s>>>((float*)x, (size_t)n, scale, dev_scale);
    else scale_kernel<bf16><<<B, 256, 0, s>>>((bf16*)x, (size_t)n, scale, dev_scale);
    NXD_LAUNCH_CHECK();
    return 0;
}
