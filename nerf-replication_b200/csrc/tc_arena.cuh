// tc_arena.cuh -- layout of the training arenas shared by the tensor-core forward (save), dgrad and wgrad kernels.
//
// Everything a training step keeps between kernels is stored as 16 KB blocks of 128 rows x 64 bf16 columns in
// the un-swizzled tcgen05 "core matrix" order, split in two 64-row halves:
//     byte offset of (row r, column c) = (r / 64) * 8192 + (c / 8) * 1024 + (r % 64) * 16 + (c % 8) * 2
// i.e. 8 rows x 8 columns form one contiguous 128-byte core matrix.  Two properties make this the layout:
//   * the producer kernels hold one row per thread (TMEM lane) and 8 consecutive columns per 16-byte register
//     quad, so a warp's st.global.v4 covers 32 rows x 16 B = 512 contiguous bytes: the images are written
//     straight from registers, fully coalesced, without a second pass through shared memory;
//   * wgrad (dW = dZ^T X, reduction over rows) bulk-copies half-block pieces (8 KB, contiguous) into shared
//     memory and uses them as MN-major operands: descriptor LBO = 128 (next 8 rows), SBO = 1024 (next 8
//     columns), one K = 16 step = +256 bytes -- no kernel ever transposes or re-tiles.
// The ReLU masks the dgrad chain needs are kept separately as bits (32 B per row and layer instead of 512 B).
//
// Tile T = ((pair_super_tile * 2 + cta_rank) * 2 + xy) covers rows [T*128, T*128+128) of the MLP row space
// (row g = ray * S + sample), identical in all three kernels.
#pragma once
#include <stdint.h>

namespace nerf {
namespace arena {

constexpr int kBlockBytes = 16384;
// ---- activation arena (written by the forward) : block slots per tile
constexpr int kActPE = 0;        // PE(pts): 63 values, column 63 = 1.0 (the "ones" column that yields bias gradients)
constexpr int kActH = 1;         // h_l (post-ReLU), l = 0..7 : slots 1 + 4 l + b
constexpr int kActFeat = 33;     // feature_linear output (linear), 4 blocks
constexpr int kActPEDir = 37;    // PE(dir): 27 values, column 63 = 1.0
constexpr int kActHV = 38;       // views hidden (post-ReLU), 2 blocks
constexpr int kActBlocks = 40;
// ---- ReLU bit masks, stored behind the tile's activation blocks: [mask layer][row 0..127][word 0..7] uint32 -- one 32-byte
//      record per row and layer, written / read with ONE 256-bit st.global / ld.global per row (a warp covers 1 KB of full
//      sectors; round 2 first kept [layer][word][row] with eight 4-byte accesses per row); word w covers columns [32 w, 32 w + 32):
//      column pair i (columns 32 w + 2 i, + 1) has its two flags at bits 8 (i & 1) + 7 - (i >> 1) and that + 16
//      (tc_common.cuh pos_bits32 / mask2_from_bits: the consumer expands a pair with one shift-shared PRMT)
constexpr int kMaskH = 0;        // h_l > 0, l = 0..7
constexpr int kMaskHV = 8;       // views hidden > 0 (words 0..3)
constexpr int kMaskLayers = 9;
constexpr int kMaskRowBytes = 32;              // eight words of one row, contiguous
constexpr int kMaskLayerBytes = 128 * kMaskRowBytes;
// ---- gradient arena (written by the dgrad chain) : block slots per tile
constexpr int kGradZ = 0;        // dZ_l (grad wrt pre-activation of trunk layer l), l = 0..7 : slots 4 l + b
constexpr int kGradFeat = 32;    // grad wrt feature_linear output, 4 blocks
constexpr int kGradHV = 36;      // grad wrt views pre-activation, 2 blocks
constexpr int kGradBlocks = 38;

__host__ __device__ constexpr size_t act_mask_offset() { return (size_t)kActBlocks * kBlockBytes; }
__host__ __device__ constexpr size_t act_tile_bytes() { return act_mask_offset() + (size_t)kMaskLayers * kMaskLayerBytes; }
// offset of a row's first 16-byte chunk inside a block; chunk c8 (8 columns) adds c8 * 1024
__host__ __device__ constexpr uint32_t cm_row_offset(int row) { return (uint32_t)((row >> 6) * 8192 + (row & 63) * 16); }
__host__ __device__ constexpr size_t grad_tile_bytes() { return (size_t)kGradBlocks * kBlockBytes; }
// number of 128-row tiles the pair-tiled kernels touch for M rows
__host__ __device__ constexpr int64_t num_tiles(int64_t M) { return ((M + 511) / 512) * 4; }

}  // namespace arena
}  // namespace nerf
