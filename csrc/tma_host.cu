// Host-side TMA descriptor (CUtensorMap) construction.  cuTensorMapEncodeTiled lives in the
// driver; it is resolved at run time through cudaGetDriverEntryPoint so the library links
// against cudart only (no libcuda at build time on the GPU-less build box).
#include <cudaTypedefs.h>

#include "gemm_tcgen05.cuh"

namespace nxd_gemm {

static PFN_cuTensorMapEncodeTiled_v12000 get_encode() {
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
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || p == nullptr) return nullptr;
    fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
    return fn;
}

int make_tmap_2d(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t pitch_elems,
                 uint32_t box_inner, uint32_t box_outer, int elem_bytes) {
    auto encode = get_encode();
    NXD_REQUIRE(encode != nullptr, "cuTensorMapEncodeTiled could not be resolved from the driver");
    cuuint64_t dims[2] = {inner, outer};
    cuuint64_t strides[1] = {pitch_elems * (cuuint64_t)elem_bytes};  // bytes, for dim 1
    cuuint32_t box[2] = {box_inner, box_outer};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = encode(out, elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    NXD_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d): inner=%llu outer=%llu pitch=%llu box=%ux%u",
                (int)r, (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)pitch_elems,
                box_inner, box_outer);
    return 0;
}

int make_tmap_3d(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t batch, uint64_t pitch_elems,
                 uint64_t batch_stride_elems, uint32_t box_inner, uint32_t box_outer, int elem_bytes) {
    auto encode = get_encode();
    NXD_REQUIRE(encode != nullptr, "cuTensorMapEncodeTiled could not be resolved from the driver");
    cuuint64_t dims[3] = {inner, outer, batch};
    cuuint64_t strides[2] = {pitch_elems * (cuuint64_t)elem_bytes, batch_stride_elems * (cuuint64_t)elem_bytes};
    if (batch == 1 && strides[1] < strides[0] * outer) strides[1] = strides[0] * outer;   // any legal value
    cuuint32_t box[3] = {box_inner, box_outer, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = encode(out, elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3,
                        const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    NXD_REQUIRE(r == CUDA_SUCCESS,
                "cuTensorMapEncodeTiled(3d) failed (%d): inner=%llu outer=%llu batch=%llu pitch=%llu stride=%llu box=%ux%u",
                (int)r, (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)batch,
                (unsigned long long)pitch_elems, (unsigned long long)batch_stride_elems, box_inner, box_outer);
    return 0;
}

}  // namespace nxd_gemm
