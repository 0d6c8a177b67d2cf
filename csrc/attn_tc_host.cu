// Host-side 4-D TMA descriptors for the tcgen05 attention kernels.
#include <cudaTypedefs.h>

#include "attn_tc_common.cuh"

namespace nxd_attn {

static PFN_cuTensorMapEncodeTiled_v12000 get_encode_fn() {
    // The encode call is a DRIVER entry point and needs a current context on the calling thread.  A thread that has
    // made no runtime call yet (an autograd worker whose first CUDA work is one of our launches) has none: bind the
    // primary context of the thread's device once per thread.
    static thread_local bool ctx_bound = false;
    if (!ctx_bound) {
        cudaFree(nullptr);
        ctx_bound = true;
    }
    static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
    if (fn) return fn;
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
        return nullptr;
    fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(ptr);
    return fn;
}

// 4-D bf16 (or fp32) map over a strided [B, H, S, D] view.  The three outer dims are ordered by ascending stride (TMA wants
// a conventional layout); pos[] returns the coordinate slot (1..3) of (row, head, batch).
int make_tmap_bhsd(CUtensorMap* out, const void* base, int D, long long S, long long H, long long B, long long s_s,
                   long long s_h, long long s_b, int box_rows, int* pos, int elem_bytes) {
    auto encode = get_encode_fn();
    NXD_REQUIRE(encode != nullptr, "cuTensorMapEncodeTiled could not be resolved from the driver");
    struct Dim { long long size, stride; int box, id; };
    Dim d[3] = {{S, s_s, box_rows, 0}, {H, s_h, 1, 1}, {B, s_b, 1, 2}};
    for (int i = 0; i < 3; ++i)
        for (int j = i + 1; j < 3; ++j) {
            // size-1 dims carry meaningless strides: sort them last
            const bool swap = (d[j].size > 1 && d[i].size == 1) ||
                              (d[j].size > 1 && d[i].size > 1 && d[j].stride < d[i].stride);
            if (swap) { Dim t = d[i]; d[i] = d[j]; d[j] = t; }
        }
    cuuint64_t dims[4] = {(cuuint64_t)D, 0, 0, 0};
    cuuint64_t strides[3];
    cuuint32_t box[4] = {(cuuint32_t)(128 / elem_bytes), 1, 1, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    long long prev_extent = (long long)D * elem_bytes;
    for (int i = 0; i < 3; ++i) {
        dims[i + 1] = (cuuint64_t)d[i].size;
        long long st = d[i].stride * elem_bytes;
        if (d[i].size == 1 && (st < prev_extent || st % 16)) st = prev_extent;  // any legal value
        strides[i] = (cuuint64_t)st;
        prev_extent = st * d[i].size;
        box[i + 1] = (cuuint32_t)d[i].box;
        pos[d[i].id] = i + 1;
    }
    CUresult r = encode(out, elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    NXD_REQUIRE(r == CUDA_SUCCESS,
                "cuTensorMapEncodeTiled(4d) failed (%d): dims=[%d,%llu,%llu,%llu] strides=[%llu,%llu,%llu] box_rows=%d",
                (int)r, D, (unsigned long long)dims[1], (unsigned long long)dims[2], (unsigned long long)dims[3],
                (unsigned long long)strides[0], (unsigned long long)strides[1], (unsigned long long)strides[2],
                box_rows);
    return 0;
}


}  // namespace nxd_attn
