// pack.cuh -- fp32 master weights -> the bf16 operand images of the tcgen05 kernels, as per-chunk device functions shared
// by the stand-alone pack kernels (nerf_set_params) and the fused clip+Adam+re-pack kernel (adam.cu):
//   forward images (mlp_tc2.cu):  per net 39 pre-swizzled 32 KB stage images (L0..L7, feature, views; the skip / view
//                                 concatenations of network.py:54-67 are baked in as extra K-blocks) + the bias slabs
//   dgrad images (mlp_tc2_bwd.cu): per net 42 transposed stage images B[n][k] = W_layer[kb*64 + k][col0 + n]
// The source weights are read with ld.global.cg (L2): the fused Adam kernel re-packs values other CTAs wrote earlier in
// the same launch, which the non-coherent path (__ldg) must not be used for.
#pragma once
#include "engine.h"
#include "tc_common.cuh"
#include "tc_arena.cuh"

namespace nerf {

// ---- forward schedule ------------------------------------------------------------------------------
namespace fwdpack {
using namespace tc;
constexpr int kStageBytes = 32768;
constexpr int kNumSched = 10;                      // L0..L7, FEAT, VIEWS
constexpr int kNumStages = 39;

__host__ __device__ constexpr int sched_nkb(int li) { return li == 0 ? 1 : (li == 5 || li == 9) ? 5 : 4; }
__host__ __device__ constexpr int sched_start(int li) {
  int s = 0;
  for (int i = 0; i < li; ++i) s += sched_nkb(i);
  return s;
}
static_assert(sched_start(kNumSched) == kNumStages, "stage count");
__host__ __device__ constexpr int sched_n(int li) { return li == 9 ? 128 : 256; }
// schedule index -> reference layer id (state_dict order): 0..7 trunk, 8 -> feature(9), 9 -> views(8)
__host__ __device__ constexpr int sched_layer(int li) { return li < 8 ? li : (li == 8 ? L_FEAT : L_VIEWS); }

// after the stage images of both nets: per (net, schedule entry) one 8 KB bias slab (csrc/mlp_tc2.cu folds the bias into the GEMM)
constexpr size_t kBiasSlabBytes = 8192;
inline size_t wpack_bias_offset() { return (size_t)2 * kNumStages * kStageBytes; }
inline size_t wpack_bytes() { return wpack_bias_offset() + (size_t)2 * kNumSched * kBiasSlabBytes; }

constexpr int64_t kWeightChunks = (int64_t)2 * kNumStages * 256 * 8;
// rep / n_rep: optional replicas of the stage images (engine.h wrep): the chunk is also written to each of them
__device__ __forceinline__ void pack_weight_chunk(int64_t e, const float* __restrict__ flat, __nv_bfloat16* __restrict__ wpack,
                                                  __nv_bfloat16* __restrict__ rep = nullptr, int n_rep = 0) {
  {
    const int kc = (int)(e & 7), n = (int)((e >> 3) & 255);
    const int s = (int)((e >> 11) % kNumStages), net = (int)((e >> 11) / kNumStages);
    int li = 0;
    while (li + 1 < kNumSched && sched_start(li + 1) <= s) ++li;
    const int kb = s - sched_start(li);
    if (n >= sched_n(li)) return;
    const int layer = sched_layer(li);
    const LayerDim d = layer_dim(layer);
    const float* wrow = flat + (int64_t)net * NERF_PARAMS_PER_NET + w_off(layer) + (int64_t)n * d.in;
    uint32_t pk[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float w2[2];
#pragma unroll
      for (int h2 = 0; h2 < 2; ++h2) {
        const int k = kc * 8 + 2 * j + h2;
        int col;   // input column of the reference weight, -1 = zero padding
        if (li == 0) col = k < 63 ? k : -1;
        else if (li == 5) col = kb == 0 ? (k < 63 ? k : -1) : 63 + (kb - 1) * 64 + k;     // [pts63 | h256]
        else if (li == 9) col = kb < 4 ? kb * 64 + k : (k < 27 ? 256 + k : -1);           // [feature256 | dir27]
        else col = kb * 64 + k;
        w2[h2] = col < 0 ? 0.f : __ldcg(wrow + col);
      }
      const __nv_bfloat162 b2 = __floats2bfloat162_rn(w2[0], w2[1]);
      pk[j] = *reinterpret_cast<const uint32_t*>(&b2);
    }
    const int64_t off = ((int64_t)net * kNumStages + s) * kStageBytes + sw128_offset(n, kc);
    *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(wpack) + off) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
    for (int r = 0; r < n_rep; ++r)
      *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(rep) + (int64_t)r * (2 * kNumStages * kStageBytes) + off) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
  }
}

// Bias slabs for the CTA-pair kernel: the bias enters the accumulator through one extra K = 16 MMA per layer, A = a constant
// slab whose columns 0..2 are 1.0, B = this slab: row n holds (hi, mid, lo, 0 ...) with hi + mid + lo = bias[n] split into three
// bf16 terms (24 mantissa bits, products by 1.0 are exact, fp32 accumulation).  Layout: un-swizzled K-major core matrices
// (8 rows x 16 bytes), element (r, k) of a CTA's half at (r / 8) * 256 + (k / 8) * 128 + (r % 8) * 16 + (k % 8) * 2; each CTA
// of the pair owns N / 2 consecutive rows, its half starts at rank * 4096.  One thread per row.
constexpr int kBiasRows = 2 * kNumSched * 256;
__device__ __forceinline__ void pack_bias_row(int e, const float* __restrict__ flat, uint8_t* __restrict__ slabs) {
  const int n = e & 255, li = (e >> 8) % kNumSched, net = (e >> 8) / kNumSched;
  const int N = sched_n(li);
  if (n >= N) return;
  const float b = __ldcg(flat + (int64_t)net * NERF_PARAMS_PER_NET + b_off(sched_layer(li)) + n);
  const __nv_bfloat16 hi = __float2bfloat16_rn(b);
  const float r1 = b - __bfloat162float(hi);
  const __nv_bfloat16 mid = __float2bfloat16_rn(r1);
  const __nv_bfloat16 lo = __float2bfloat16_rn(r1 - __bfloat162float(mid));
  const int rank = n / (N / 2), r = n % (N / 2);
  uint8_t* dst = slabs + ((size_t)net * kNumSched + li) * kBiasSlabBytes + (size_t)rank * 4096 + (r >> 3) * 256 + (r & 7) * 16;
  const uint32_t w0 = (uint32_t)__bfloat16_as_ushort(hi) | ((uint32_t)__bfloat16_as_ushort(mid) << 16);
  *reinterpret_cast<uint4*>(dst) = make_uint4(w0, (uint32_t)__bfloat16_as_ushort(lo), 0u, 0u);
  *reinterpret_cast<uint4*>(dst + 128) = make_uint4(0u, 0u, 0u, 0u);
}

}  // namespace fwdpack

// ---- dgrad schedule (steps of the fused input-gradient chain; the kernel itself is in mlp_tc2_bwd.cu) --------
namespace bw {
using namespace tc;
constexpr int kTile = 128;
constexpr int kRowsPerPair = 512;
constexpr int kStageImage = 32768;
constexpr int kHalfBytes = 16384;
constexpr int kRing = 6;        // weight stages in flight
// weight producers: warp 0 and warps 10.. (one warp issues one bulk copy per ~300-450 clk whatever its size: see mlp_tc2.cu)
#ifndef NERF_NUM_PROD_BWD
#define NERF_NUM_PROD_BWD 3
#endif
constexpr int kNumProd = NERF_NUM_PROD_BWD;
constexpr int kFirstExtraProd = 10;
constexpr int kThreads = 320 + 32 * (kNumProd - 1);      // warps: 0 producer 0, 1 issuer / forwarder (+ TMEM alloc), 2-9 tile warps, 10.. producers 1..
constexpr int kNumSteps = 11;      // V1, F, L7, L6, PE5, L5, L4, L3, L2, L1, P0
constexpr int kNumStages = 42;
enum Step { S_V1 = 0, S_F = 1, S_L7 = 2, S_L6 = 3, S_PE5 = 4, S_L5 = 5, S_L4 = 6, S_L3 = 7, S_L2 = 8, S_L1 = 9, S_P0 = 10 };

__host__ __device__ constexpr int step_nkb(int s) { return s == S_V1 ? 2 : 4; }
__host__ __device__ constexpr int step_n(int s) { return (s == S_PE5 || s == S_P0) ? 64 : 256; }
__host__ __device__ constexpr int step_stage0(int s) {
  int a = 0;
  for (int i = 0; i < s; ++i) a += step_nkb(i);
  return a;
}
static_assert(step_stage0(kNumSteps) == kNumStages, "dgrad stage count");
// reference layer whose weight the step multiplies by, and the first input column of that weight it produces
__host__ __device__ constexpr int step_layer(int s) {
  return s == S_V1 ? L_VIEWS : s == S_F ? L_FEAT : s == S_L7 ? 7 : s == S_L6 ? 6 : (s == S_PE5 || s == S_L5) ? 5
       : s == S_L4 ? 4 : s == S_L3 ? 3 : s == S_L2 ? 2 : s == S_L1 ? 1 : 0;
}
__host__ __device__ constexpr int step_col0(int s) { return s == S_L5 ? 63 : 0; }
__host__ __device__ constexpr int step_ncols(int s) { return (s == S_PE5 || s == S_P0) ? 63 : 256; }
// trunk layer whose ReLU bit mask the step's epilogue applies (-1: none) and gradient-arena slot it writes
__host__ __device__ constexpr int step_mask_layer(int s) {
  return s == S_F ? 7 : s == S_L7 ? 6 : s == S_L6 ? 5 : s == S_L5 ? 4 : s == S_L4 ? 3 : s == S_L3 ? 2 : s == S_L2 ? 1
       : s == S_L1 ? 0 : -1;
}
__host__ __device__ constexpr int step_out_slot(int s) {
  return s == S_V1 ? arena::kGradFeat : s == S_F ? arena::kGradZ + 28 : s == S_L7 ? arena::kGradZ + 24
       : s == S_L6 ? arena::kGradZ + 20 : s == S_L5 ? arena::kGradZ + 16 : s == S_L4 ? arena::kGradZ + 12
       : s == S_L3 ? arena::kGradZ + 8 : s == S_L2 ? arena::kGradZ + 4 : s == S_L1 ? arena::kGradZ + 0 : -1;
}

constexpr int64_t kDgradChunks = (int64_t)2 * kNumStages * 256 * 8;     // one 16-byte chunk each: (net, stage, n, 8 consecutive k)
__device__ __forceinline__ void pack_dgrad_chunk(int64_t e, const float* __restrict__ flat, __nv_bfloat16* __restrict__ wpack) {
  {
    // n fastest: neighbouring threads read neighbouring columns of the same weight rows
    const int n = (int)(e & 255), kc = (int)((e >> 8) & 7);
    const int stg = (int)((e >> 11) % kNumStages), net = (int)((e >> 11) / kNumStages);
    int s = 0;
    while (s + 1 < kNumSteps && step_stage0(s + 1) <= stg) ++s;
    const int kb = stg - step_stage0(s);
    if (n >= step_n(s)) return;
    const int layer = step_layer(s);
    const LayerDim d = layer_dim(layer);
    const bool col_ok = n < step_ncols(s);
    const float* wcol = flat + (int64_t)net * NERF_PARAMS_PER_NET + w_off(layer) + step_col0(s) + n;
    uint32_t pk[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float w2[2];
#pragma unroll
      for (int h2 = 0; h2 < 2; ++h2) {
        const int out = kb * 64 + kc * 8 + 2 * j + h2;     // output feature of the layer = reduction index of dgrad
        w2[h2] = (col_ok && out < d.out) ? __ldcg(wcol + (int64_t)out * d.in) : 0.f;
      }
      const __nv_bfloat162 b2 = __floats2bfloat162_rn(w2[0], w2[1]);
      pk[j] = *reinterpret_cast<const uint32_t*>(&b2);
    }
    uint8_t* dst = reinterpret_cast<uint8_t*>(wpack) + ((int64_t)net * kNumStages + stg) * kStageImage + sw128_offset(n, kc);
    *reinterpret_cast<uint4*>(dst) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
  }
}

}  // namespace bw

}  // namespace nerf
