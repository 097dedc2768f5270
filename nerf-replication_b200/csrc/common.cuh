// common.cuh -- shared definitions of the nerf_b200 engine (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include "../../include/nerf_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "nerf_b200 targets sm_100a (B200) only"
#endif

namespace nerf {

// ---- network geometry (reference network.py:9-47; lego.yaml:31-43) ---------------------------------
constexpr int kW = 256;          // trunk width
constexpr int kPE = 63;          // 3 + 3*2*10
constexpr int kPEDir = 27;       // 3 + 3*2*4
constexpr int kLayers = NERF_NUM_LAYERS;
enum Layer { L_PTS0 = 0, L_PTS5 = 5, L_PTS7 = 7, L_VIEWS = 8, L_FEAT = 9, L_ALPHA = 10, L_RGB = 11 };

struct LayerDim { int out, in; };
__host__ __device__ constexpr LayerDim layer_dim(int l) {
  return l == 0 ? LayerDim{256, 63}
       : l == 5 ? LayerDim{256, 319}
       : l <= 7 ? LayerDim{256, 256}
       : l == 8 ? LayerDim{128, 283}
       : l == 9 ? LayerDim{256, 256}
       : l == 10 ? LayerDim{1, 256}
                 : LayerDim{3, 128};
}
// offset (in floats) of layer l's weight inside one net's flat block; bias follows the weight.
__host__ __device__ constexpr int w_off(int l) {
  int o = 0;
  for (int i = 0; i < l; ++i) o += layer_dim(i).out * layer_dim(i).in + layer_dim(i).out;
  return o;
}
__host__ __device__ constexpr int b_off(int l) { return w_off(l) + layer_dim(l).out * layer_dim(l).in; }
static_assert(w_off(kLayers) == NERF_PARAMS_PER_NET, "flat layout mismatch");

// ---- error plumbing --------------------------------------------------------------------------------
struct Status {
  int code = NERF_OK;
  std::string msg;
};

#define NERF_CUDA_TRY(h, expr)                                                                   \
  do {                                                                                           \
    cudaError_t _e = (expr);                                                                     \
    if (_e != cudaSuccess) {                                                                     \
      return (h)->fail(NERF_E_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));          \
    }                                                                                            \
  } while (0)

// ---- small device helpers --------------------------------------------------------------------------
// torch.linspace in fp32 (CPU and CUDA kernels contract start + step*i into an FMA; verified bit-equal
// against torch 2.11 for (2,6,64), (0,1,128), (2,6,128), (0,1,256)).
__device__ __forceinline__ float linspace_at(float start, float end, int steps, int i) {
  const float step = __fdiv_rn(__fsub_rn(end, start), (float)(steps - 1));
  return (i < steps / 2) ? __fmaf_rn(step, (float)i, start) : __fmaf_rn(-step, (float)(steps - 1 - i), end);
}

// pts = o + d * t exactly as torch evaluates it: separately rounded mul, then add (no FMA).
__device__ __forceinline__ float ray_point(float o, float d, float t) { return __fadd_rn(o, __fmul_rn(d, t)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// inclusive warp scans
__device__ __forceinline__ float warp_scan_add(float v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    float n = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += n;
  }
  return v;
}
__device__ __forceinline__ float warp_scan_mul(float v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    float n = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v *= n;
  }
  return v;
}

// Philox-4x32-10 (counter-based) -> one uniform in [0,1) per (seed, stream, index).
__device__ __forceinline__ float philox_uniform(uint64_t seed, uint32_t stream_id, uint64_t idx) {
  uint32_t c0 = (uint32_t)idx, c1 = (uint32_t)(idx >> 32), c2 = stream_id, c3 = 0x9E3779B9u;
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    c0 = hi1 ^ c1 ^ k0; c1 = lo1; c2 = hi0 ^ c3 ^ k1; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return (float)(c0 >> 8) * (1.0f / 16777216.0f);
}

}  // namespace nerf
