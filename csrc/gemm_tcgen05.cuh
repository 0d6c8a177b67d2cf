// Shared definitions for the tcgen05 GEMM family (plain GEMM and the fused comm+GEMM kernels).
#pragma once
#include "common.cuh"

namespace nxd_gemm {

constexpr int kBlockM = 128;          // rows of A (and of D) per CTA
constexpr int kTileM = 2 * kBlockM;   // rows per CTA pair (UMMA M = 256, cta_group::2)
constexpr int kBlockK = 64;           // 64 bf16 = 128 B = one SWIZZLE_128B row
constexpr int kAccStages = 2;         // TMEM accumulator double buffering
constexpr int kThreads = 256;         // 8 warps: TMA, MMA, TMEM alloc, spare, 4 x epilogue
constexpr int kFirstEpilogueWarp = 4;
constexpr int kEpilogueWarps = 4;
constexpr int kGroupM = 8;            // m-tiles per rasterisation group (L2 reuse of B)

enum { kLayoutNT = 0, kLayoutNN = 1, kLayoutTN = 2 };
enum { kOutBF16 = 0, kOutF32 = 1, kOutF32Acc = 2 };

template <int BN>
struct SmemLayout {
    static constexpr int kABytes = kBlockM * kBlockK * 2;        // 16 KB
    static constexpr int kBBytes = (BN / 2) * kBlockK * 2;       // this CTA's half of B
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kStages = (BN == 256) ? 6 : 8;
    // epilogue staging: per epilogue warp two 32-row x 128-byte buffers (SWIZZLE_128B boxes for the TMA store)
    static constexpr int kStagingOffset = kStages * kStageBytes;
    static constexpr int kStagingBytesPerWarp = 2 * 32 * 128;
    static constexpr int kBarrierOffset = kStagingOffset + kEpilogueWarps * kStagingBytesPerWarp;
    static constexpr int kBarrierBytes = (2 * kStages + 2 * kAccStages) * 8 + 16;
    static constexpr int kSmemBytes = kBarrierOffset + kBarrierBytes + 1024;  // +1024: manual alignment slack
    static constexpr int kTmemCols = kAccStages * BN;            // 512 or 256 (power of two)
    static_assert(kSmemBytes <= 232448, "shared memory budget exceeded");
};

struct EpilogueArgs {
    void* d;          // output, row-major, pitch ldd elements
    long long ldd;
    int mode;         // kOutBF16 / kOutF32 / kOutF32Acc
};

// Grouped rasterisation: tiles are numbered so that kGroupM consecutive m-tiles share each n-tile
// back to back -> concurrently resident clusters re-use the same B panel out of L2.
__device__ __forceinline__ void tile_coords(int tile, int num_m_tiles, int num_n_tiles, int& tm, int& tn) {
    const int per_group = kGroupM * num_n_tiles;
    const int group = tile / per_group;
    const int first_m = group * kGroupM;
    const int rows_in_group = min(kGroupM, num_m_tiles - first_m);
    const int in_group = tile - group * per_group;
    tm = first_m + in_group % rows_in_group;
    tn = in_group / rows_in_group;
}

// 2-D tensor map: `inner` contiguous elements, `outer` rows with pitch `pitch_elems`;
// box = {box_inner, box_outer}; SWIZZLE_128B (box_inner * elem_bytes must be 128 B).  elem_bytes 2 = bf16, 4 = fp32.
int make_tmap_2d(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t pitch_elems,
                 uint32_t box_inner, uint32_t box_outer, int elem_bytes = 2);

// 3-D variant: `batch` stacked 2-D problems `batch_stride_elems` apart (coordinate 2 = problem index).
int make_tmap_3d(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t batch, uint64_t pitch_elems,
                 uint64_t batch_stride_elems, uint32_t box_inner, uint32_t box_outer, int elem_bytes = 2);

int pick_block_n(int M, int N);

int run(int layout, const void* a, const void* b, long long lda, long long ldb, const EpilogueArgs& ep, int M, int N,
        int K, int block_n, cudaStream_t stream);

}  // namespace nxd_gemm
