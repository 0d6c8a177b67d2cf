// KV-cache addressing shared by the token-generation kernels (decode_fused.cu, gemm_smallm.cu).
#pragma once
#include "common.cuh"

namespace nxd_decode {

struct CacheDst {
    bf16* k;
    bf16* v;
    const int* block_table;  // null -> contiguous cache [lines, n_kv, S_max, D]
    const int* seq_ids;      // null -> cache line = batch row
    long long stride_b, stride_h;
    int page_size, max_pages;
};

__device__ __forceinline__ size_t cache_row(const CacheDst& c, int b, int h, int pos, int D) {
    const int line = c.seq_ids ? c.seq_ids[b] : b;
    if (c.block_table) {
        const int page = c.block_table[(size_t)line * c.max_pages + pos / c.page_size];
        return (size_t)page * c.stride_b + (size_t)h * c.stride_h + (size_t)(pos % c.page_size) * D;
    }
    return (size_t)line * c.stride_b + (size_t)h * c.stride_h + (size_t)pos * D;
}

}  // namespace nxd_decode
