// mlp_tc.cu -- the hot kernel: fused NeRF MLP forward on tcgen05 / TMEM (sm_100a), bf16 operands, fp32
// accumulate.  Replaces Network.forward + batchify + NeRF.forward (network.py:49-74,163-171,199-258):
// positional encoding, the 8x256 skip-connected trunk, feature/alpha heads, the view branch and the rgb
// head run inside ONE persistent kernel; sample features never leave the SM between layers.
//
// Per CTA (one per SM, persistent over "super-tiles" of 256 sample rows = two 128-row MMA tiles X,Y):
//   warp 0      weight producer: streams the layer weights as pre-swizzled 32 KB stage images with
//               cp.async.bulk (TMA engine) into a shared-memory ring, completion on mbarriers
//   warp 1      MMA issuer: one lane issues tcgen05.mma (M=128, N=256|128, K=16) for X then Y on every
//               weight stage (each stage is read from L2 once per 256 rows), commits to mbarriers
//   warp 2      TMEM allocator (512 columns: X accumulators [0,256), Y accumulators [256,512))
//   warps 4-7   tile X: positional encoding -> bf16 A operand; per layer epilogue TMEM -> +bias, ReLU ->
//               bf16 -> A operand of the next layer (in place, 128B-swizzled K-major); sigma / rgb heads
//   warps 8-11  same for tile Y
// Shared memory (bytes): A_X 64K | A_Y 64K | PE_X 16K | PE_Y 16K | W ring 2 x 32K | barriers+heads 3K.
#include "engine.h"
#include "tc_common.cuh"
#include "rows.cuh"

namespace nerf {
using namespace tc;

constexpr int kTile = 128;
constexpr int kSuper = 2 * kTile;
constexpr int kStageBytes = 32768;
constexpr int kRing = 2;
constexpr int kNumSched = 10;                      // L0..L7, FEAT, VIEWS
constexpr int kNumStages = 39;
constexpr uint32_t kOffA = 0, kOffPE = 131072, kOffW = 163840, kOffBar = 229376;
constexpr uint32_t kOffHeads = kOffBar + 128;      // w_alpha[256] | w_rgb[384]
constexpr uint32_t kSmemBytes = kOffHeads + (256 + 384) * 4;   // 232064 <= 232448
constexpr int kThreads = 384;

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
size_t tc_wpack_bias_offset() { return (size_t)2 * kNumStages * kStageBytes; }
size_t tc_wpack_bytes() { return tc_wpack_bias_offset() + (size_t)2 * kNumSched * kBiasSlabBytes; }

// Packs fp32 [out][in] weights into the stage images the kernel streams: stage = [N rows x 64 k] bf16,
// K-major, 128B swizzle.  One thread per 16-byte chunk (net, stage, n, 8 consecutive k): the schedule lookup is done once
// per chunk, the eight source weights are consecutive floats, the destination is one 16-byte store.
__global__ void pack_weights_kernel(const float* __restrict__ flat, __nv_bfloat16* __restrict__ wpack) {
  const int64_t total = (int64_t)2 * kNumStages * 256 * 8;
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int kc = (int)(e & 7), n = (int)((e >> 3) & 255);
    const int s = (int)((e >> 11) % kNumStages), net = (int)((e >> 11) / kNumStages);
    int li = 0;
    while (li + 1 < kNumSched && sched_start(li + 1) <= s) ++li;
    const int kb = s - sched_start(li);
    if (n >= sched_n(li)) continue;
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
        w2[h2] = col < 0 ? 0.f : __ldg(wrow + col);
      }
      const __nv_bfloat162 b2 = __floats2bfloat162_rn(w2[0], w2[1]);
      pk[j] = *reinterpret_cast<const uint32_t*>(&b2);
    }
    uint8_t* dst = reinterpret_cast<uint8_t*>(wpack) + ((int64_t)net * kNumStages + s) * kStageBytes + sw128_offset(n, kc);
    *reinterpret_cast<uint4*>(dst) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
  }
}

// Bias slabs for the CTA-pair kernel: the bias enters the accumulator through one extra K = 16 MMA per layer, A = a constant
// slab whose columns 0..2 are 1.0, B = this slab: row n holds (hi, mid, lo, 0 ...) with hi + mid + lo = bias[n] split into three
// bf16 terms (24 mantissa bits, products by 1.0 are exact, fp32 accumulation).  Layout: un-swizzled K-major core matrices
// (8 rows x 16 bytes), element (r, k) of a CTA's half at (r / 8) * 256 + (k / 8) * 128 + (r % 8) * 16 + (k % 8) * 2; each CTA
// of the pair owns N / 2 consecutive rows, its half starts at rank * 4096.  One thread per row.
__global__ void pack_bias_slabs_kernel(const float* __restrict__ flat, uint8_t* __restrict__ slabs) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= 2 * kNumSched * 256) return;
  const int n = e & 255, li = (e >> 8) % kNumSched, net = (e >> 8) / kNumSched;
  const int N = sched_n(li);
  if (n >= N) return;
  const float b = __ldg(flat + (int64_t)net * NERF_PARAMS_PER_NET + b_off(sched_layer(li)) + n);
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

void tc_pack_weights(NerfEngine* h, const float* flat, __nv_bfloat16* wpack, cudaStream_t st) {
  pack_weights_kernel<<<h->num_sms * 4, 256, 0, st>>>(flat, wpack);
  pack_bias_slabs_kernel<<<(2 * kNumSched * 256 + 255) / 256, 256, 0, st>>>(flat, reinterpret_cast<uint8_t*>(wpack) + tc_wpack_bias_offset());
  h->launches += 2;
}

// ---- the fused forward kernel ---------------------------------------------------------------------
struct TcSmem {
  uint64_t w_full[kRing], w_empty[kRing], acc_full[2], act_ready[2];
  uint32_t tmem_slot;
};

// write 8 consecutive bf16 (one 16-byte chunk) of row `row` into a [128 x 64] swizzled K-block
__device__ __forceinline__ void store_chunk(uint8_t* blk, uint32_t row_base, uint32_t row_xor, int chunk, uint32_t a,
                                            uint32_t b, uint32_t c, uint32_t d) {
  *reinterpret_cast<uint4*>(blk + row_base + (((uint32_t)chunk ^ row_xor) << 4)) = make_uint4(a, b, c, d);
}

__global__ void __launch_bounds__(kThreads, 1)
mlp_tc_forward_kernel(RowSrc src, float4* __restrict__ raw, const __nv_bfloat16* __restrict__ wpack,
                      const float* __restrict__ P, int64_t M, int64_t n_super) {
  extern __shared__ __align__(1024) uint8_t smem[];
  TcSmem& sb = *reinterpret_cast<TcSmem*>(smem + kOffBar);
  float* s_walpha = reinterpret_cast<float*>(smem + kOffHeads);
  float* s_wrgb = s_walpha + 256;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if ((smem_u32(smem) & 1023u) != 0) mbar_timeout(0xA11, smem_u32(smem));   // swizzle-128B needs 1024B alignment
  if (tid == 0) {
    for (int i = 0; i < kRing; ++i) { mbar_init(&sb.w_full[i], 1); mbar_init(&sb.w_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&sb.acc_full[i], 1); mbar_init(&sb.act_ready[i], kTile); }
    fence_barrier_init();
  }
  if (warp == 2) { tmem_alloc(&sb.tmem_slot, 512); tmem_relinquish(); }
  for (int i = tid; i < 256; i += kThreads) s_walpha[i] = __ldg(P + w_off(L_ALPHA) + i);
  for (int i = tid; i < 384; i += kThreads) s_wrgb[i] = __ldg(P + w_off(L_RGB) + i);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sb.tmem_slot;

  if (warp == 0) {
    // ===== weight producer =====
    if (lane == 0) {
      uint32_t it = 0;
      for (int64_t st = blockIdx.x; st < n_super; st += gridDim.x) {
        for (int li = 0; li < kNumSched; ++li) {
          const uint32_t bytes = (uint32_t)sched_n(li) * 128u;
          for (int kb = 0; kb < sched_nkb(li); ++kb, ++it) {
            const uint32_t slot = it % kRing, par = (it / kRing) & 1u;
            mbar_wait(&sb.w_empty[slot], par ^ 1u, 0x10 + slot);
            mbar_arrive_expect_tx(&sb.w_full[slot], bytes);
            bulk_g2s(smem + kOffW + slot * kStageBytes,
                     reinterpret_cast<const uint8_t*>(wpack) + (size_t)(sched_start(li) + kb) * kStageBytes, bytes,
                     &sb.w_full[slot]);
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      uint32_t it = 0, lc = 0;
      const uint32_t sA = smem_u32(smem + kOffA), sPE = smem_u32(smem + kOffPE), sW = smem_u32(smem + kOffW);
      for (int64_t st = blockIdx.x; st < n_super; st += gridDim.x) {
        for (int li = 0; li < kNumSched; ++li, ++lc) {
          const int nkb = sched_nkb(li);
          const uint32_t idesc = make_idesc_bf16(kTile, sched_n(li));
          for (int kb = 0; kb < nkb; ++kb, ++it) {
            const uint32_t slot = it % kRing, par = (it / kRing) & 1u;
            mbar_wait(&sb.w_full[slot], par, 0x20 + slot);
            tc_fence_after();
            // A source of this k-block: PE buffer or hidden-activation block
            int ablk;   // -1 = PE buffer
            if (li == 0) ablk = -1;
            else if (li == 5) ablk = kb - 1;
            else if (li == 9) ablk = kb < 4 ? kb : -1;
            else ablk = kb;
#pragma unroll
            for (int tile = 0; tile < 2; ++tile) {
              if (kb == 0) { mbar_wait(&sb.act_ready[tile], lc & 1u, 0x30 + tile); tc_fence_after(); }
              const uint32_t a_base = ablk < 0 ? sPE + tile * 16384u : sA + tile * 65536u + (uint32_t)ablk * 16384u;
              const uint32_t b_base = sW + slot * kStageBytes;
#pragma unroll
              for (int kk = 0; kk < 4; ++kk) {
                umma_ss(tmem + tile * 256u, make_smem_desc(a_base + kk * 32, 16, 1024, kLayoutSW128),
                        make_smem_desc(b_base + kk * 32, 16, 1024, kLayoutSW128), idesc, (kb | kk) != 0);
              }
              if (kb == nkb - 1) umma_commit(&sb.acc_full[tile]);
            }
            umma_commit(&sb.w_empty[slot]);
          }
        }
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ===== PE + epilogue warps =====
    const int tile = (warp - 4) >> 2, q = warp & 3;
    const int row = q * 32 + lane;
    const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16) + tile * 256u;
    uint8_t* sAct = smem + kOffA + tile * 65536;
    uint8_t* sPe = smem + kOffPE + tile * 16384;
    const uint32_t row_base = (uint32_t)((row >> 3) * 1024 + (row & 7) * 128), row_xor = (uint32_t)(row & 7);
    uint32_t lc = 0;
    for (int64_t st = blockIdx.x; st < n_super; st += gridDim.x) {
      const int64_t g = st * kSuper + tile * kTile + row;
      const bool valid = g < M;
      float p[3] = {0.f, 0.f, 0.f}, vd[3] = {0.f, 0.f, 0.f};
      if (valid) row_point(src, g, p, vd);
      {  // positional encoding of the point: 63 values + zero pad -> one swizzled K-block
        float f = 1.0f;
        float prev[3] = {p[0], p[1], p[2]};   // columns 0..2 of the current chunk stream
        // stream the 64 columns in order: [x y z | sin(f x..z) cos(f x..z)]_k ; pack 8 at a time
        float colbuf[8];
        int nbuf = 0, chunk = 0;
        auto push = [&](float v) {
          colbuf[nbuf++] = v;
          if (nbuf == 8) {
            store_chunk(sPe, row_base, row_xor, chunk++, pack_bf16x2(colbuf[0], colbuf[1]), pack_bf16x2(colbuf[2], colbuf[3]),
                        pack_bf16x2(colbuf[4], colbuf[5]), pack_bf16x2(colbuf[6], colbuf[7]));
            nbuf = 0;
          }
        };
        push(prev[0]); push(prev[1]); push(prev[2]);
#pragma unroll
        for (int k = 0; k < 10; ++k, f *= 2.0f) {
          float sn[3], cs[3];
#pragma unroll
          for (int c = 0; c < 3; ++c) sincosf(p[c] * f, &sn[c], &cs[c]);
          push(sn[0]); push(sn[1]); push(sn[2]); push(cs[0]); push(cs[1]); push(cs[2]);
        }
        push(0.f);
      }
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(&sb.act_ready[tile]);

      float sigma = 0.f;
      for (int li = 0; li < kNumSched; ++li, ++lc) {
        mbar_wait(&sb.acc_full[tile], lc & 1u, 0x40 + tile);
        tc_fence_after();
        const float* bias = P + b_off(sched_layer(li));
        if (li < 9) {
#pragma unroll 1
          for (int c0 = 0; c0 < 256; c0 += 32) {
            uint32_t v[32];
            tmem_ld32(taddr + c0, v);
            tmem_wait_ld();
            uint32_t pk[16];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 bb = __ldg(reinterpret_cast<const float4*>(bias + c0) + j);
              float f0 = __uint_as_float(v[4 * j]) + bb.x, f1 = __uint_as_float(v[4 * j + 1]) + bb.y;
              float f2 = __uint_as_float(v[4 * j + 2]) + bb.z, f3 = __uint_as_float(v[4 * j + 3]) + bb.w;
              if (li == 8) {   // feature_linear: no activation
                pk[2 * j] = pack_bf16x2(f0, f1); pk[2 * j + 1] = pack_bf16x2(f2, f3);
              } else {
                if (li == 7) {   // alpha head on the fp32 post-ReLU activations
                  const float4 wa = *reinterpret_cast<const float4*>(s_walpha + c0 + 4 * j);
                  sigma = fmaf(fmaxf(f0, 0.f), wa.x, sigma); sigma = fmaf(fmaxf(f1, 0.f), wa.y, sigma);
                  sigma = fmaf(fmaxf(f2, 0.f), wa.z, sigma); sigma = fmaf(fmaxf(f3, 0.f), wa.w, sigma);
                }
                pk[2 * j] = pack_bf16x2_relu(f0, f1); pk[2 * j + 1] = pack_bf16x2_relu(f2, f3);
              }
            }
            uint8_t* blk = sAct + (c0 >> 6) * 16384;
            const int ch0 = (c0 & 63) >> 3;
#pragma unroll
            for (int j = 0; j < 4; ++j) store_chunk(blk, row_base, row_xor, ch0 + j, pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
          }
          if (li == 5) {
            // layer 5 has consumed PE(pts): the buffer now carries PE(dir) (27 values + zero pad) for the view branch
            float colbuf[8];
            int nbuf = 0, chunk = 0;
            auto push = [&](float x) {
              colbuf[nbuf++] = x;
              if (nbuf == 8) {
                store_chunk(sPe, row_base, row_xor, chunk++, pack_bf16x2(colbuf[0], colbuf[1]), pack_bf16x2(colbuf[2], colbuf[3]),
                            pack_bf16x2(colbuf[4], colbuf[5]), pack_bf16x2(colbuf[6], colbuf[7]));
                nbuf = 0;
              }
            };
            push(vd[0]); push(vd[1]); push(vd[2]);
            float f = 1.0f;
#pragma unroll
            for (int k = 0; k < 4; ++k, f *= 2.0f) {
              float sn[3], cs[3];
#pragma unroll
              for (int c = 0; c < 3; ++c) sincosf(vd[c] * f, &sn[c], &cs[c]);
              push(sn[0]); push(sn[1]); push(sn[2]); push(cs[0]); push(cs[1]); push(cs[2]);
            }
#pragma unroll
            for (int z = 27; z < 64; ++z) push(0.f);
          }
          fence_proxy_async_smem();
          tc_fence_before();
          mbar_arrive(&sb.act_ready[tile]);
        } else {
          // view branch output (128 wide, ReLU) -> rgb head on CUDA cores; sigma joined from layer 7
          float r0 = 0.f, r1 = 0.f, r2 = 0.f;
#pragma unroll 1
          for (int c0 = 0; c0 < 128; c0 += 32) {
            uint32_t v[32];
            tmem_ld32(taddr + c0, v);
            tmem_wait_ld();
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 bb = __ldg(reinterpret_cast<const float4*>(bias + c0) + j);
              const float hv[4] = {fmaxf(__uint_as_float(v[4 * j]) + bb.x, 0.f), fmaxf(__uint_as_float(v[4 * j + 1]) + bb.y, 0.f),
                                   fmaxf(__uint_as_float(v[4 * j + 2]) + bb.z, 0.f), fmaxf(__uint_as_float(v[4 * j + 3]) + bb.w, 0.f)};
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const int c = c0 + 4 * j + e;
                r0 = fmaf(hv[e], s_wrgb[c], r0); r1 = fmaf(hv[e], s_wrgb[128 + c], r1); r2 = fmaf(hv[e], s_wrgb[256 + c], r2);
              }
            }
          }
          tc_fence_before();
          if (valid) {
            float4 o = make_float4(r0 + __ldg(P + b_off(L_RGB)), r1 + __ldg(P + b_off(L_RGB) + 1), r2 + __ldg(P + b_off(L_RGB) + 2),
                                   sigma + __ldg(P + b_off(L_ALPHA)));
            if (src.mask && !src.mask[g]) o = make_float4(0.f, 0.f, 0.f, 0.f);
            raw[g] = o;
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem, 512);
}

static unsigned int g_debug_host[8];

int mlp_tc_forward(NerfEngine* h, int net, const RowSrc& src, float* raw, cudaStream_t st) {
  const int64_t M = src.R * src.S;
  if (M <= 0) return NERF_OK;
  static bool attr_set_dev[64] = {};          // the opt-in shared-memory size is a per-device function attribute
  bool& attr_set = attr_set_dev[h->device & 63];
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(mlp_tc_forward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
    if (e != cudaSuccess) return h->fail(NERF_E_CUDA, std::string("mlp_tc smem attribute: ") + cudaGetErrorString(e));
    attr_set = true;
  }
  const int64_t n_super = (M + kSuper - 1) / kSuper;
  const int grid = (int)std::min<int64_t>(n_super, h->num_sms);
  const float* P = h->params + (net == NERF_NET_FINE ? NERF_PARAMS_PER_NET : 0);
  const __nv_bfloat16* wp = h->wpack + (size_t)(net == NERF_NET_FINE ? 1 : 0) * kNumStages * (kStageBytes / 2);
  mlp_tc_forward_kernel<<<grid, kThreads, kSmemBytes, st>>>(src, reinterpret_cast<float4*>(raw), wp, P, M, n_super);
  h->launches++;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return h->fail(NERF_E_CUDA, std::string("mlp_tc_forward launch: ") + cudaGetErrorString(e));
  return NERF_OK;
}

// reads the watchdog record of this translation unit (valid after a trapped launch, before device reset)
int tc_debug_read(unsigned int out[8]) {
  cudaError_t e = cudaMemcpyFromSymbol(g_debug_host, g_tc_debug, sizeof(g_debug_host));
  for (int i = 0; i < 8; ++i) out[i] = g_debug_host[i];
  return e == cudaSuccess ? 0 : -1;
}

NERF_TC_DEFINE_DEBUG_SINK(tc1_set_debug_sink)

}  // namespace nerf
