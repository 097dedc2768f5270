// tc_arena.cuh -- layout of the training arenas shared by the tensor-core forward (save), dgrad and wgrad kernels.
//
// Everything a training step keeps between kernels is stored as "tile images": one 16 KB block = 128 rows x 64
// bf16 columns, K-major with the 128-byte swizzle -- byte for byte what a tcgen05 shared-memory operand looks
// like.  The same image is a K-major operand [rows x cols] (forward / dgrad: A = activations) and an MN-major
// operand [cols x rows] (wgrad: dW = dZ^T X reduces over rows), so no kernel ever transposes or re-tiles.
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
// ---- gradient arena (written by the dgrad chain) : block slots per tile
constexpr int kGradZ = 0;        // dZ_l (grad wrt pre-activation of trunk layer l), l = 0..7 : slots 4 l + b
constexpr int kGradFeat = 32;    // grad wrt feature_linear output, 4 blocks
constexpr int kGradHV = 36;      // grad wrt views pre-activation, 2 blocks
constexpr int kGradBlocks = 38;

__host__ __device__ constexpr size_t act_tile_bytes() { return (size_t)kActBlocks * kBlockBytes; }
__host__ __device__ constexpr size_t grad_tile_bytes() { return (size_t)kGradBlocks * kBlockBytes; }
// number of 128-row tiles the pair-tiled kernels touch for M rows
__host__ __device__ constexpr int64_t num_tiles(int64_t M) { return ((M + 511) / 512) * 4; }

}  // namespace arena
}  // namespace nerf
