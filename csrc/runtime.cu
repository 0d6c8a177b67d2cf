// Library-wide runtime helpers: error text, version, device properties.
#include <stdarg.h>
#include <string.h>

#include <stdlib.h>

#include "common.cuh"

static thread_local char g_err[1024] = {0};

void nxd_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

NXD_API const char* nxd_last_error() { return g_err; }

static unsigned long long g_launches = 0;   // relaxed counter: exactness across threads is not needed
int nxd_pdl_enabled() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("NXD_PDL");
        v = (e != nullptr && e[0] == '0') ? 0 : 1;
    }
    return v;
}

void nxd_count_launches(int n) { __atomic_fetch_add(&g_launches, (unsigned long long)n, __ATOMIC_RELAXED); }
NXD_API unsigned long long nxd_kernel_launches() { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }

NXD_API int nxd_version() { return 100; }

NXD_API int nxd_device_sm_count() { return nxd_num_sms(); }

// Writes `bytes` of zeros/pattern through a buffer larger than L2 (bench L2 flush).
__global__ void l2_flush_kernel(uint4* p, size_t n, uint32_t v) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) p[i] = make_uint4(v, v, v, v);
}

NXD_API int nxd_l2_flush(void* buf, size_t bytes, void* stream) {
    if (bytes < 16) return 0;
    l2_flush_kernel<<<nxd_num_sms() * 8, 256, 0, (cudaStream_t)stream>>>((uint4*)buf, bytes / 16, 0u);
    NXD_LAUNCH_CHECK();
    return 0;
}
