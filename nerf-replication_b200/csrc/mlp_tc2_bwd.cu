// mlp_tc2_bwd.cu -- tensor-core backward of the NeRF MLP (bf16 operands, fp32 accumulate), two kernels:
//
//   dgrad chain  (this file, part 1): per 128-row tile the whole chain  d raw -> d hv -> d feature -> dZ7 ... dZ0
//                (-> d PE -> d t for the fine net) runs inside one persistent CTA-pair kernel; the gradient tile
//                stays in shared memory as the A operand of the next tcgen05.mma (M=256, cta_group::2), transposed
//                weight stage images stream through a 6-deep ring, the ReLU masks are the bit masks the forward
//                saved (32 bytes per row and layer, loaded into registers ahead of the accumulator wait), and every
//                dZ_l tile is written to the gradient arena straight from the epilogue registers.
//   wgrad        (part 2): dW_l = dZ_l^T X_{l-1} summed over all rows -- the saved tile images are used directly as
//                MN-major operands (reduction dimension = rows), accumulators [256 x 384] fp32 in TMEM across a CTA
//                pair, one red.global.add flush per (layer, tile range) work unit.
//
// Reference graph: autograd of network.py:49-74 as exercised by trainers/nerf.py:27-33: the fine
// net needs all weight gradients and the input gradient down to t; the coarse net only receives d sigma.
#include <algorithm>
#include <vector>
#include "engine.h"
#include "tc_common.cuh"
#include "rows.cuh"
#include "tc_arena.cuh"
#include "pack.cuh"

namespace nerf {
using namespace tc;

namespace bw {

// shared memory map
constexpr uint32_t kOffA = 0, kOffW = 131072;
constexpr uint32_t kOffHeads = kOffW + kRing * kHalfBytes;                                  // 229376: w_alpha[256] | w_rgb[384]
constexpr uint32_t kOffBar = kOffHeads + (256 + 384) * 4;                                   // 231936
constexpr uint32_t kSmemBytes = kOffBar + 256;                                              // 232192 <= 232448

struct Bars {
  uint64_t w_full[kRing], w_empty[kRing], acc_full[2], act_ready[2], acc_free[2];
  uint32_t tmem_slot;
};
static_assert(sizeof(Bars) <= 256, "barrier block");

}  // namespace bw

namespace bw {

__device__ __forceinline__ void store_chunk(uint8_t* blk, uint32_t row_base, uint32_t row_xor, int chunk, uint32_t a,
                                            uint32_t b, uint32_t c, uint32_t d) {
  if (NERF_EXP & 1024) return;     // timing experiment: the epilogues write no A operand
  *reinterpret_cast<uint4*>(blk + row_base + (((uint32_t)chunk ^ row_xor) << 4)) = make_uint4(a, b, c, d);
}
constexpr float kInv2PiHi = 0.15915493667125702f, kInv2PiLo = 6.4206382432985265e-09f;
__device__ __forceinline__ void sincos_turns(float x, float c_hi, float c_lo, float& s, float& c) {
  const float r = x * c_hi;
  const float k = rintf(r);
  float fr = fmaf(x, c_hi, -k);
  fr = fmaf(x, c_lo, fr);
  const float ang = fr * 6.283185307179586f;
  s = __sinf(ang);
  c = __cosf(ang);
}

// NERF_DEFER_STORES (default 0): experiment -- gradient-arena copy after the hand-over, re-read from shared memory; measured
// slower (see mlp_tc2.cu).
#ifndef NERF_DEFER_STORES
#define NERF_DEFER_STORES 0
#endif
// 16 packed words (32 columns) -> shared-memory A operand (swizzled) [+ this row of the gradient arena block]
// PASS 0: both; 1: A operand only (first pass of the two-phase epilogue); 2: arena only (second pass over the same accumulator)
#ifndef NERF_TWO_PHASE
#define NERF_TWO_PHASE 0      // measured: forward 0.37 -> 0.43 ms, dgrad 0.34 -> 0.43 ms (see the comment in mlp_tc2.cu / profiles/exp_train_r2.md)
#endif
template <int PASS = 0>
__device__ __forceinline__ void put_chunk(const uint32_t (&pk)[16], int c0, uint8_t* sA, uint32_t row_base, uint32_t row_xor, uint8_t* grow) {
  uint8_t* blk = sA + (c0 >> 6) * 16384;
  const int ch0 = (c0 & 63) >> 3;
  if (PASS != 2) {
#pragma unroll
    for (int j = 0; j < 4; ++j) store_chunk(blk, row_base, row_xor, ch0 + j, pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
  }
  if (!NERF_DEFER_STORES && PASS != 1) {
    uint8_t* gb = grow + (c0 >> 6) * 16384 + ch0 * 1024;
#pragma unroll
    for (int j = 0; j < 4; ++j) st_global_cs(gb + j * 1024, pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
  }
}
// re-read this row's n_blocks x 128 bytes from the (swizzled) A operand and stream them to the arena blocks at grow
__device__ __forceinline__ void flush_row_to_arena(const uint8_t* sA, uint32_t row_base, uint32_t row_xor, uint8_t* grow, int n_blocks) {
  if (!NERF_DEFER_STORES) return;
#pragma unroll
  for (int b = 0; b < 4; ++b) {
    if (b < n_blocks) {
#pragma unroll
      for (int ch = 0; ch < 8; ++ch) {
        const uint4 v = *reinterpret_cast<const uint4*>(sA + b * 16384 + row_base + (((uint32_t)ch ^ row_xor) << 4));
        st_global_cs(grow + b * 16384 + ch * 1024, v.x, v.y, v.z, v.w);
      }
    }
  }
}

// Epilogue of a masked step (F, L7..L1): 256 accumulator columns -> (+ d_sigma * w_alpha for F) -> bf16 -> AND
// with the saved ReLU bit mask (mw: 8 words of this row) -> next A operand (in place) + gradient arena.
template <bool ADD_ALPHA, int PASS = 0>
__device__ __forceinline__ void epi_masked(uint32_t taddr, const uint32_t (&mw)[8], float dsigma, const float* __restrict__ s_walpha,
                                           uint8_t* sA, uint32_t row_base, uint32_t row_xor, uint8_t* grow) {
  uint32_t va[32], vb[32];
  auto chunk = [&](const uint32_t (&v)[32], int c0) {
    const uint32_t bits = mw[c0 >> 5];
    uint32_t pk[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      float f0 = __uint_as_float(v[2 * i]), f1 = __uint_as_float(v[2 * i + 1]);
      if (ADD_ALPHA) {
        const float2 wa = *reinterpret_cast<const float2*>(s_walpha + c0 + 2 * i);
        f0 = fmaf(dsigma, wa.x, f0);
        f1 = fmaf(dsigma, wa.y, f1);
      }
      pk[i] = pack_bf16x2(f0, f1) & mask2_from_bits(bits, i);
    }
    put_chunk<PASS>(pk, c0, sA, row_base, row_xor, grow);
  };
  tmem_ld32(taddr, va);
#pragma unroll
  for (int c = 0; c < 8; c += 2) {
    tmem_wait_ld();
    tmem_ld32(taddr + (c + 1) * 32, vb);
    chunk(va, c * 32);
    tmem_wait_ld();
    if (c + 2 < 8) tmem_ld32(taddr + (c + 2) * 32, va);
    chunk(vb, (c + 1) * 32);
  }
}

// Epilogue of the linear step V1 (d feature): 256 columns -> bf16 -> A operand + gradient arena.
template <int PASS = 0>
__device__ __forceinline__ void epi_linear(uint32_t taddr, uint8_t* sA, uint32_t row_base, uint32_t row_xor, uint8_t* grow) {
  uint32_t va[32], vb[32];
  auto chunk = [&](const uint32_t (&v)[32], int c0) {
    uint32_t pk[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) pk[j] = pack_bf16x2(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));
    put_chunk<PASS>(pk, c0, sA, row_base, row_xor, grow);
  };
  tmem_ld32(taddr, va);
#pragma unroll
  for (int c = 0; c < 8; c += 2) {
    tmem_wait_ld();
    tmem_ld32(taddr + (c + 1) * 32, vb);
    chunk(va, c * 32);
    tmem_wait_ld();
    if (c + 2 < 8) tmem_ld32(taddr + (c + 2) * 32, va);
    chunk(vb, (c + 1) * 32);
  }
}

struct DgradArgs {
  RowSrc src;                       // rays + depths (FINE: needed for d PE -> d t); R, S define the row space
  const float4* d_raw;              // [M] (d r, d g, d b, d sigma) w.r.t. the pre-activation outputs
  const uint8_t* act_arena;         // saved forward activations (tc_arena.cuh)
  uint8_t* grad_arena;              // out: dZ tile images
  float4* pe_scratch;               // FINE: [M, 16] float4 -- d PE contribution of layer 5 (skip connection)
  float* d_t;                       // FINE: [M] += d loss / d t
  const __nv_bfloat16* wpack;       // dgrad stage images of this net
  const float* P;                   // fp32 parameters of this net (heads)
  int64_t M, n_pair;
};

constexpr bool kTwoPhase = NERF_TWO_PHASE && !NERF_DEFER_STORES;
template <bool FINE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1) mlp_tc2_dgrad_kernel(DgradArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  Bars& sb = *reinterpret_cast<Bars*>(smem + kOffBar);
  float* s_walpha = reinterpret_cast<float*>(smem + kOffHeads);
  float* s_wrgb = s_walpha + 256;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t rank = cluster_ctarank();
  const int64_t pair0 = blockIdx.x >> 1, pair_stride = gridDim.x >> 1;
  // step list of this net: the coarse net only receives d sigma: trunk only, no input gradient
  constexpr int kSteps = FINE ? 11 : 7;
  // fine: all 11 steps in order; coarse: L7, L6, L5, L4, L3, L2, L1 (arithmetic, not a table: a local array would live in
  // local memory, and with all of L1 configured as shared memory every lookup would be an L2 round trip)
  auto step_at = [&](int i) { return FINE ? i : (i < 2 ? i + 2 : i + 3); };
  static_assert(S_L7 == 2 && S_L6 == 3 && S_L5 == 5 && S_L1 == 9 && S_P0 == 10, "step numbering");
  constexpr int kInitBlocks = FINE ? 2 : 4;                                   // hv (2 blocks) / h7 (4 blocks)
  constexpr int kInitMask = FINE ? arena::kMaskHV : arena::kMaskH + 7;

  if ((smem_u32(smem) & 1023u) != 0) mbar_timeout(0xA13, smem_u32(smem));
  if (tid == 0) {
    for (int i = 0; i < kRing; ++i) { mbar_init(&sb.w_full[i], rank == 0 ? 2 : 1); mbar_init(&sb.w_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&sb.acc_full[i], 1); mbar_init(&sb.act_ready[i], 8); mbar_init(&sb.acc_free[i], 8); }
    fence_barrier_init();
  }
  for (int i = tid; i < 256; i += kThreads) s_walpha[i] = __ldg(a.P + w_off(L_ALPHA) + i);
  for (int i = tid; i < 384; i += kThreads) s_wrgb[i] = __ldg(a.P + w_off(L_RGB) + i);
  cluster_sync_all();
  if (warp == 1) { tmem_alloc2(&sb.tmem_slot, 512); tmem_relinquish2(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sb.tmem_slot;

  const int prod = warp == 0 ? 0 : (warp >= kFirstExtraProd ? warp - kFirstExtraProd + 1 : -1);
  if (prod >= 0) {
    // ===== weight producers: producer p issues the stages it % kNumProd == p =====
    // The whole warp runs the loop converged (all lanes poll the barrier, lane 0 issues): a lone lane-0 loop inside a diverged
    // warp sees its barrier complete ~2 000 clk late (event trace of the forward, profiles/exp_train_r2.md).
    {
      uint32_t it = 0;
      for (int64_t st = pair0; st < a.n_pair; st += pair_stride)
        for (int i = 0; i < kSteps; ++i) {
          const int s = step_at(i);
          const uint32_t bytes = (uint32_t)step_n(s) * 64u;
          for (int rep = 0; rep < 2; ++rep)            // once per tile: see the issuer
          for (int kb = 0; kb < step_nkb(s); ++kb, ++it) {
            if ((int)(it % kNumProd) != prod) continue;
            const uint32_t slot = it % kRing, par = (it / kRing) & 1u;
            mbar_wait(&sb.w_empty[slot], par ^ 1u, 0x10 + slot);
            if (lane == 0) {
              mbar_arrive_expect_tx(&sb.w_full[slot], bytes);
              bulk_g2s(smem + kOffW + slot * kHalfBytes,
                       reinterpret_cast<const uint8_t*>(a.wpack) + (size_t)(step_stage0(s) + kb) * kStageImage + (size_t)rank * bytes,
                       bytes, &sb.w_full[slot]);
            }
            __syncwarp();
          }
        }
    }
    __syncwarp();
  } else if (warp == 1 && rank == 1) {
    // ===== forwarder (converged warp, see the producer) =====
    {
      uint32_t it = 0;
      for (int64_t st = pair0; st < a.n_pair; st += pair_stride)
        for (int i = 0; i < kSteps; ++i)
          for (int kb = 0; kb < 2 * step_nkb(step_at(i)); ++kb, ++it) {
            const uint32_t slot = it % kRing, par = (it / kRing) & 1u;
            mbar_wait(&sb.w_full[slot], par, 0x50 + slot);
            if (lane == 0) mbar_arrive_remote(mapa_u32(smem_u32(&sb.w_full[slot]), 0));
            __syncwarp();
          }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===== MMA issuer (leader) =====
    {                                   // the whole warp, converged: see umma_ss2_w (tc_common.cuh)
      uint32_t it = 0, lc = 0;
      const uint32_t sA = smem_u32(smem + kOffA), sW = smem_u32(smem + kOffW);
      for (int64_t st = pair0; st < a.n_pair; st += pair_stride)
        for (int i = 0; i < kSteps; ++i, ++lc) {
          const int s = step_at(i), nkb = step_nkb(s);
          const uint32_t idesc = make_idesc_bf16(256, step_n(s));
          // one tile's whole step, then the other's (the step's weights are streamed twice, L2 hits): tile Y's MMAs run
          // under tile X's epilogue instead of both accumulators finishing -- and being drained -- together
#pragma unroll
          for (int tile = 0; tile < 2; ++tile) {
            mbar_wait(&sb.act_ready[tile], lc & 1u, 0x30 + tile);
            if (kTwoPhase) mbar_wait(&sb.acc_free[tile], lc & 1u, 0x38 + tile);      // the accumulator has been read twice (two-phase epilogue)
            tc_fence_after();
            for (int kb = 0; kb < nkb; ++kb, ++it) {
              const uint32_t slot = it % kRing, par = (it / kRing) & 1u;
              if (!(NERF_EXP & 2)) mbar_wait(&sb.w_full[slot], par, 0x20 + slot);
              tc_fence_after();
              const uint32_t a_base = sA + tile * 65536u + (uint32_t)kb * 16384u;
              const uint32_t b_base = sW + slot * kHalfBytes;
              umma_ss2_w_x4(tmem + tile * 256u, make_smem_desc(a_base, 16, 1024, kLayoutSW128), make_smem_desc(b_base, 16, 1024, kLayoutSW128), idesc, kb != 0);
              umma_commit2_w(&sb.w_empty[slot], 0x3);
            }
            umma_commit2_w(&sb.acc_full[tile], 0x3);
          }
        }
    }
    __syncwarp();
  } else if (warp >= 2) {
    // ===== gradient warps: init step + per-step epilogues =====
    const int tile = (warp - 2) >> 2, q = warp & 3;
    const int row = q * 32 + lane;
    const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16) + tile * 256u;
    uint8_t* sA = smem + kOffA + tile * 65536;
    const uint32_t row_base = (uint32_t)((row >> 3) * 1024 + (row & 7) * 128), row_xor = (uint32_t)(row & 7);
    const uint32_t ready_addr = mapa_u32(smem_u32(&sb.act_ready[tile]), 0);
    auto signal_ready = [&]() {
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_remote(ready_addr);
    };
    const uint32_t free_addr = mapa_u32(smem_u32(&sb.acc_free[tile]), 0);
    auto signal_free = [&]() {              // two-phase epilogue: every tcgen05.ld of this warp on the accumulator has completed
      if (!kTwoPhase) return;
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_remote(free_addr);
    };
    uint32_t lc = 0;
    // ReLU bit masks of this row (saved by the forward; tc_arena.cuh).  The epilogues are the critical path of this kernel,
    // so a mask is fetched one step AHEAD of its use (mw_pre: the next step's words, or the next super-tile's init mask): a
    // load issued just before the accumulator wait would expose its full HBM latency every step.
    auto load_mask_at = [&](int64_t st_, int layer, uint32_t (&mw)[8]) {
      const int64_t T_ = (st_ * 2 + rank) * 2 + tile;
      const uint8_t* mp = a.act_arena + (size_t)T_ * arena::act_tile_bytes() + arena::act_mask_offset() + row * arena::kMaskRowBytes + layer * arena::kMaskLayerBytes;
      if (NERF_EXP & 4) {
#pragma unroll
        for (int w = 0; w < 8; ++w) mw[w] = 0xFFFFFFFFu;
      } else {
        ld_global_mask8(mp, mw);
      }
    };
    uint32_t mw_pre[8];
    if (pair0 < a.n_pair) load_mask_at(pair0, kInitMask, mw_pre);
    for (int64_t st = pair0; st < a.n_pair; st += pair_stride) {
      const int64_t T = (st * 2 + rank) * 2 + tile;
      const int64_t g = T * kTile + row;
      const bool valid = g < a.M;
      // this row's slice of the tile's gradient-arena blocks
      uint8_t* grow0 = a.grad_arena + (size_t)T * arena::grad_tile_bytes() + arena::cm_row_offset(row);
      const float4 dr = valid ? __ldg(a.d_raw + g) : make_float4(0.f, 0.f, 0.f, 0.f);
      float p[3] = {0.f, 0.f, 0.f}, vd[3];
      if (FINE && valid) row_point(a.src, g, p, vd);

      // ---- init: the first A operand from d raw ----
      {
        uint32_t mw[8];
#pragma unroll
        for (int w = 0; w < 8; ++w) mw[w] = mw_pre[w];
        if (step_mask_layer(step_at(0)) >= 0) load_mask_at(st, step_mask_layer(step_at(0)), mw_pre);     // for the first step
        uint8_t* grow = grow0 + (size_t)(FINE ? arena::kGradHV : arena::kGradZ + 28) * arena::kBlockBytes;
#pragma unroll
        for (int c32 = 0; c32 < 2 * kInitBlocks; ++c32) {
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int c = c32 * 32 + 2 * i;
            float f0, f1;
            if (FINE) {      // d hv_pre = W_rgb^T d rgb, masked by hv > 0   (rgb_linear.weight [3,128])
              f0 = dr.x * s_wrgb[c] + dr.y * s_wrgb[128 + c] + dr.z * s_wrgb[256 + c];
              f1 = dr.x * s_wrgb[c + 1] + dr.y * s_wrgb[128 + c + 1] + dr.z * s_wrgb[256 + c + 1];
            } else {         // d h7 = w_alpha d sigma, masked by h7 > 0
              f0 = dr.w * s_walpha[c]; f1 = dr.w * s_walpha[c + 1];
            }
            pk[i] = pack_bf16x2(f0, f1) & mask2_from_bits(mw[c32], i);
          }
          put_chunk(pk, c32 * 32, sA, row_base, row_xor, grow);
        }
      }
      signal_ready();
      signal_free();                          // the previous super-tile's last step has drained the accumulator
      flush_row_to_arena(sA, row_base, row_xor, grow0 + (size_t)(FINE ? arena::kGradHV : arena::kGradZ + 28) * arena::kBlockBytes, kInitBlocks);

      for (int i = 0; i < kSteps; ++i, ++lc) {
        const int s = step_at(i);
        uint32_t mw_next[8];                      // this step's mask words (fetched during the previous step)
#pragma unroll
        for (int w = 0; w < 8; ++w) mw_next[w] = mw_pre[w];
        if (i + 1 < kSteps) {
          if (step_mask_layer(step_at(i + 1)) >= 0) load_mask_at(st, step_mask_layer(step_at(i + 1)), mw_pre);
        } else if (st + pair_stride < a.n_pair) {
          load_mask_at(st + pair_stride, kInitMask, mw_pre);
        }
        mbar_wait(&sb.acc_full[tile], lc & 1u, 0x40 + tile);
        tc_fence_after();
        exp_bulk_wait(lane);
        if (s == S_PE5) {
          // d PE from the skip connection: park 64 fp32 columns in HBM until the end of the chain
          uint32_t v[32];
#pragma unroll
          for (int hlf = 0; hlf < 2; ++hlf) {
            tmem_ld32(taddr + hlf * 32, v);
            tmem_wait_ld();
            if (valid) {
#pragma unroll
              for (int j = 0; j < 8; ++j)
                a.pe_scratch[g * 16 + hlf * 8 + j] = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]),
                                                                __uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3]));
            }
          }
          signal_ready();
          signal_free();
        } else if (s == S_P0) {
          // d PE = dZ0 W0 + (layer-5 part) ; PE backward ; d t += d . d pts
          float dpe[64];
          uint32_t v[32];
#pragma unroll
          for (int hlf = 0; hlf < 2; ++hlf) {
            tmem_ld32(taddr + hlf * 32, v);
            tmem_wait_ld();
#pragma unroll
            for (int j = 0; j < 32; ++j) dpe[hlf * 32 + j] = __uint_as_float(v[j]);
          }
          tc_fence_before();
          if (valid) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float4 sc = a.pe_scratch[g * 16 + j];
              dpe[4 * j] += sc.x; dpe[4 * j + 1] += sc.y; dpe[4 * j + 2] += sc.z; dpe[4 * j + 3] += sc.w;
            }
            const int64_t ray = g / a.src.S;
            float acc = 0.f;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
              float dp = dpe[c];
              float chi = kInv2PiHi, clo = kInv2PiLo, f = 1.0f;
#pragma unroll
              for (int k = 0; k < 10; ++k, chi *= 2.0f, clo *= 2.0f, f *= 2.0f) {
                float sn, cs;
                sincos_turns(p[c], chi, clo, sn, cs);
                dp += f * (cs * dpe[3 + 6 * k + c] - sn * dpe[3 + 6 * k + 3 + c]);
              }
              acc += dp * __ldg(a.src.rays_d + 3 * ray + c);
            }
            a.d_t[g] += acc;
          }
        } else {
          uint8_t* grow = grow0 + (size_t)step_out_slot(s) * arena::kBlockBytes;
          constexpr int kP1 = kTwoPhase ? 1 : 0;      // two-phase: A operand + hand-over first, the arena copy from a second read
          if (s == S_V1) epi_linear<kP1>(taddr, sA, row_base, row_xor, grow);
          else if (s == S_F) epi_masked<true, kP1>(taddr, mw_next, dr.w, s_walpha, sA, row_base, row_xor, grow);
          else epi_masked<false, kP1>(taddr, mw_next, 0.f, s_walpha, sA, row_base, row_xor, grow);
          if (i + 1 < kSteps) signal_ready(); else tc_fence_before();
          if (kTwoPhase) {
            if (s == S_V1) epi_linear<2>(taddr, sA, row_base, row_xor, grow);
            else if (s == S_F) epi_masked<true, 2>(taddr, mw_next, dr.w, s_walpha, sA, row_base, row_xor, grow);
            else epi_masked<false, 2>(taddr, mw_next, 0.f, s_walpha, sA, row_base, row_xor, grow);
            if (i + 1 < kSteps) signal_free(); else tc_fence_before();
          }
          exp_bulk_rows_to_arena(sA, q, lane, a.grad_arena + (size_t)T * arena::grad_tile_bytes() + (size_t)step_out_slot(s) * arena::kBlockBytes, 4);
          flush_row_to_arena(sA, row_base, row_xor, grow, 4);
        }
      }
    }
  }
  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) tmem_dealloc2(tmem, 512);
}

}  // namespace bw

// d_raw [M,4] -> gradient arena (+ d_t for the fine net).  Row space and tiling identical to the saving forward.
// full: all layers receive gradients and d_t is produced (fine net; coarse net when N_importance == 0);
// !full: only d sigma arrives (coarse net of the hierarchical path): trunk + alpha only.
int mlp_tc2_dgrad(NerfEngine* h, int net, bool full, const RowSrc& src, const float* d_raw, const uint8_t* act_arena,
                  uint8_t* grad_arena, float* pe_scratch, float* d_t, cudaStream_t st) {
  using namespace bw;
  const int64_t M = src.R * src.S;
  if (M <= 0) return NERF_OK;
  static bool attr_set_dev[64] = {};          // the opt-in shared-memory size is a per-device function attribute
  bool& attr_set = attr_set_dev[h->device & 63];
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(mlp_tc2_dgrad_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(mlp_tc2_dgrad_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
    if (e != cudaSuccess) return h->fail(NERF_E_CUDA, std::string("mlp_tc2_dgrad smem attribute: ") + cudaGetErrorString(e));
    attr_set = true;
  }
  DgradArgs a;
  a.src = src; a.d_raw = reinterpret_cast<const float4*>(d_raw); a.act_arena = act_arena; a.grad_arena = grad_arena;
  a.pe_scratch = reinterpret_cast<float4*>(pe_scratch); a.d_t = d_t;
  a.wpack = h->wpack_bwd + (size_t)(net == NERF_NET_FINE ? 1 : 0) * kNumStages * (kStageImage / 2);
  a.P = h->params + (net == NERF_NET_FINE ? NERF_PARAMS_PER_NET : 0);
  a.M = M; a.n_pair = (M + kRowsPerPair - 1) / kRowsPerPair;
  // the fewest CTA pairs that keep the number of waves (1 024 training rays: 384 pair tiles = 6 waves on 74 pairs as on 64):
  // every pair then runs the same number of tiles and the idle SMs give their power budget to the busy ones
  const int64_t max_pairs = h->num_sms / 2, waves = (a.n_pair + max_pairs - 1) / max_pairs;
#ifdef NERF_FULL_GRID      // A/B builds: every pair the device has
  const int grid = 2 * (int)std::min<int64_t>(a.n_pair, max_pairs);
#else
  const int grid = 2 * (int)((a.n_pair + waves - 1) / waves);
#endif
  if (full) mlp_tc2_dgrad_kernel<true><<<grid, kThreads, kSmemBytes, st>>>(a);
  else mlp_tc2_dgrad_kernel<false><<<grid, kThreads, kSmemBytes, st>>>(a);
  h->launches++;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return h->fail(NERF_E_CUDA, std::string(full ? "mlp_tc2_dgrad<full> launch: " : "mlp_tc2_dgrad<sigma> launch: ") + cudaGetErrorString(e));
  return NERF_OK;
}


// =====================================================================================================
// part 2: weight gradients
// =====================================================================================================
namespace wg {

constexpr int kFlushWarps = 8;         // two warps per TMEM lane quarter, each flushing half of the columns
constexpr int kThreads = 64 + 32 * kFlushWarps;   // warp 0 producer, warp 1 MMA issuer / forwarder, warps 2-9 flush
constexpr int kHalfRows = 64;          // one ring stage = 64 of a tile's 128 rows (the first / second 8 KB of every block)
constexpr int kPiece = 8192;
constexpr int kStageBytes = 5 * kPiece;   // A' 2 pieces | B' main 2 pieces | B' extra 1 piece
constexpr int kStages = 4;
constexpr uint32_t kOffStage = 0, kOffXpose = kStages * kStageBytes;          // 163840
constexpr uint32_t kOffOnes = kOffXpose + kFlushWarps * 32 * 33 * 4;          // 197632: constant B' extra piece, column 63 = 1
constexpr uint32_t kOffBar = kOffOnes + kPiece;                               // 205824
constexpr uint32_t kSmemBytes = kOffBar + 128;
constexpr int kNumJobsFine = 10, kNumJobsCoarse = 8;

struct Job { int layer, a_slot, main_slot, extra_slot, main_col0, extra_col0, extra_ncols, out_dim; };
__device__ __forceinline__ Job job_of(int j) {
  Job b;
  if (j < 8) {
    b.layer = j; b.a_slot = arena::kGradZ + 4 * j;
    b.main_slot = j == 0 ? -1 : (j == 5 ? arena::kActH + 16 : arena::kActH + 4 * (j - 1));
    b.extra_slot = arena::kActPE; b.main_col0 = j == 5 ? 63 : 0; b.extra_col0 = 0;
    b.extra_ncols = (j == 0 || j == 5) ? 63 : 0; b.out_dim = 256;
  } else if (j == 8) {
    b.layer = L_FEAT; b.a_slot = arena::kGradFeat; b.main_slot = arena::kActH + 28; b.extra_slot = arena::kActPE;
    b.main_col0 = 0; b.extra_col0 = 0; b.extra_ncols = 0; b.out_dim = 256;
  } else {
    b.layer = L_VIEWS; b.a_slot = arena::kGradHV; b.main_slot = arena::kActFeat; b.extra_slot = arena::kActPEDir;
    b.main_col0 = 0; b.extra_col0 = 256; b.extra_ncols = 27; b.out_dim = 128;
  }
  return b;
}
// relative cost of one tile of a job (bytes a pair streams), used by the host to size the tile ranges
inline double job_weight(int j) { return j == 0 ? 0.6 : j == 9 ? 0.7 : 1.0; }

struct Bars { uint64_t full[kStages], empty[kStages], acc_done, tmem_free; uint32_t tmem_slot; };

struct WgradArgs {
  const uint8_t* act_arena; const uint8_t* grad_arena; float* G;     // G: flat gradient of this net (+=)
  int n_tiles, chunk, n_jobs, n_units;
};

// unit u = (job u % n_jobs, tile range u / n_jobs): pairs that flush at the same time hit different layers' dW
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1) mlp_tc2_wgrad_kernel(WgradArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  Bars& sb = *reinterpret_cast<Bars*>(smem + kOffBar);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t rank = cluster_ctarank();
  const int pair0 = blockIdx.x >> 1, pair_stride = gridDim.x >> 1;

  if ((smem_u32(smem) & 1023u) != 0) mbar_timeout(0xA14, smem_u32(smem));
  if (tid == 0) {
    for (int i = 0; i < kStages; ++i) { mbar_init(&sb.full[i], rank == 0 ? 2 : 1); mbar_init(&sb.empty[i], 1); }
    mbar_init(&sb.acc_done, 1); mbar_init(&sb.tmem_free, 2 * kFlushWarps);
    fence_barrier_init();
  }
  // jobs whose layer has no PE input columns still need the bias gradient: their B' extra operand is this constant
  // piece (64 rows x 64 columns, all zero except column 63 = 1.0) instead of a PE block fetched from the arena
  for (int i = tid; i < kPiece / 16; i += kThreads) {
    const bool ones = i >= 7 * 64;                                    // 16-byte chunk i = (column group i / 64, row i % 64)
    reinterpret_cast<uint4*>(smem + kOffOnes)[i] = make_uint4(0u, 0u, 0u, ones ? 0x3F800000u : 0u);   // bf16 1.0 in column 8*7+7
  }
  fence_proxy_async_smem();
  cluster_sync_all();
  if (warp == 1) { tmem_alloc2(&sb.tmem_slot, 512); tmem_relinquish2(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sb.tmem_slot;

  if (warp == 0) {
    // ===== producer: per half tile, my halves of dZ^T (A') and X^T (B' main) + the extra (PE) piece where the job has one =====
    if (lane == 0) {
      uint32_t it = 0;
      for (int u = pair0; u < a.n_units; u += pair_stride) {
        const Job jb = job_of(u % a.n_jobs);
        const int t0 = (u / a.n_jobs) * a.chunk, t1 = min(t0 + a.chunk, a.n_tiles);
        const int a_blk = jb.out_dim == 256 ? 2 * (int)rank : 0;       // views (128 outputs): both CTAs take blocks 0,1
        for (int t = t0; t < t1; ++t) {
          const uint8_t* gt = a.grad_arena + (size_t)t * arena::grad_tile_bytes() + (size_t)(jb.a_slot + a_blk) * 16384;
          const uint8_t* at = a.act_arena + (size_t)t * arena::act_tile_bytes();
          const uint8_t* mt = at + (size_t)(jb.main_slot + 2 * (int)rank) * 16384;
          const uint8_t* xt = at + (size_t)jb.extra_slot * 16384;
          for (int hf = 0; hf < 2; ++hf, ++it) {
            const uint32_t slot = it % kStages, par = (it / kStages) & 1u;
            mbar_wait(&sb.empty[slot], par ^ 1u, 0x810 + slot);
            uint8_t* dst = smem + kOffStage + slot * kStageBytes;
            const uint32_t o = hf * kPiece;
            mbar_arrive_expect_tx(&sb.full[slot], (2u + (jb.main_slot >= 0 ? 2u : 0u) + (jb.extra_ncols > 0 ? 1u : 0u)) * kPiece);
            bulk_g2s(dst, gt + o, kPiece, &sb.full[slot]);
            bulk_g2s(dst + kPiece, gt + 16384 + o, kPiece, &sb.full[slot]);
            if (jb.main_slot >= 0) {
              bulk_g2s(dst + 2 * kPiece, mt + o, kPiece, &sb.full[slot]);
              bulk_g2s(dst + 3 * kPiece, mt + 16384 + o, kPiece, &sb.full[slot]);
            }
            if (jb.extra_ncols > 0) bulk_g2s(dst + 4 * kPiece, xt + o, kPiece, &sb.full[slot]);
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1 && rank == 1) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int u = pair0; u < a.n_units; u += pair_stride) {
        const int t0 = (u / a.n_jobs) * a.chunk, t1 = min(t0 + a.chunk, a.n_tiles);
        for (int k = 2 * (t1 - t0); k > 0; --k, ++it) {
          const uint32_t slot = it % kStages, par = (it / kStages) & 1u;
          mbar_wait(&sb.full[slot], par, 0x850 + slot);
          mbar_arrive_remote(mapa_u32(smem_u32(&sb.full[slot]), 0));
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===== MMA issuer: D[256 x (256 | 128)] += dZ^T X, 4 K-steps of 16 rows per half tile =====
    {                                   // the whole warp, converged: see umma_ss2_w (tc_common.cuh)
      uint32_t it = 0, uc = 0;
      const uint32_t idesc_main = make_idesc_bf16(256, 256, 1, 1), idesc_extra = make_idesc_bf16(256, 128, 1, 1);
      for (int u = pair0; u < a.n_units; u += pair_stride, ++uc) {
        const Job jb = job_of(u % a.n_jobs);
        const int t0 = (u / a.n_jobs) * a.chunk, t1 = min(t0 + a.chunk, a.n_tiles);
        if (uc > 0) { mbar_wait(&sb.tmem_free, (uc - 1) & 1u, 0x830); tc_fence_after(); }
        const uint32_t it0 = it;
        for (int k = 2 * (t1 - t0); k > 0; --k, ++it) {
          const uint32_t slot = it % kStages, par = (it / kStages) & 1u;
          mbar_wait(&sb.full[slot], par, 0x820 + slot);
          tc_fence_after();
          const uint32_t base = smem_u32(smem + kOffStage + slot * kStageBytes);
          const uint32_t xbase = jb.extra_ncols > 0 ? base + 4 * kPiece : smem_u32(smem + kOffOnes);
#pragma unroll
          // core-matrix pieces (tc_arena.cuh) as MN-major operands: LBO = 128 (next 8 rows), SBO = 1024 (next 8 columns;
          // consecutive pieces continue the column index: 8 KB = 8 x SBO), one K = 16 step = 256 bytes
          for (int ks = 0; ks < kHalfRows / 16; ++ks) {
            const uint32_t acc = (it > it0 || ks > 0) ? 1u : 0u;
            const uint64_t ad = make_smem_desc(base + ks * 256, 128, 1024, kLayoutNone);
            if (jb.main_slot >= 0)
              umma_ss2_w(tmem, ad, make_smem_desc(base + 2 * kPiece + ks * 256, 128, 1024, kLayoutNone), idesc_main, acc);
            umma_ss2_w(tmem + 256, ad, make_smem_desc(xbase + ks * 256, 128, 1024, kLayoutNone), idesc_extra, acc);
          }
          umma_commit2_w(&sb.empty[slot], 0x3);
        }
        umma_commit2_w(&sb.acc_done, 0x3);
      }
    }
    __syncwarp();
  } else {
    // ===== flush warps: TMEM -> (transpose through shared memory) -> coalesced red.global.add =====
    const int fw = warp - 2, q = warp & 3, hsel = fw >> 2;   // TMEM lane quarter this warp may access; column half
    float* xp = reinterpret_cast<float*>(smem + kOffXpose) + fw * 32 * 33;
    const uint32_t free_addr = mapa_u32(smem_u32(&sb.tmem_free), 0);
    uint32_t uc = 0;
    for (int u = pair0; u < a.n_units; u += pair_stride, ++uc) {
      const Job jb = job_of(u % a.n_jobs);
      mbar_wait(&sb.acc_done, uc & 1u, 0x840);
      tc_fence_after();
      const LayerDim d = layer_dim(jb.layer);
      const int n_base = 128 * (int)rank + q * 32;              // first output row of this warp
      const bool rows_valid = n_base < jb.out_dim;
      float* Wg = a.G + w_off(jb.layer);
      auto flush_cols = [&](int tcol0, int c_begin, int c_end, int ncols, int gcol0) {   // D[:, tcol0 + c] -> dW[n][gcol0 + c]
        for (int c0 = c_begin; c0 < c_end && c0 < ncols; c0 += 32) {
          uint32_t v[32];
          tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + tcol0 + c0, v);
          tmem_wait_ld();
#pragma unroll
          for (int j = 0; j < 32; ++j) xp[lane * 33 + j] = __uint_as_float(v[j]);
          __syncwarp();
          if (rows_valid && c0 + lane < ncols) {
#pragma unroll 8
            for (int r = 0; r < 32; ++r) atomicAdd(Wg + (size_t)(n_base + r) * d.in + gcol0 + c0 + lane, xp[r * 33 + lane]);
          }
          __syncwarp();
        }
      };
      if (jb.main_slot >= 0) flush_cols(0, 128 * hsel, 128 * hsel + 128, 256, jb.main_col0);
      if (jb.extra_ncols > 0) flush_cols(256, 32 * hsel, 32 * hsel + 32, jb.extra_ncols, jb.extra_col0);
      if (hsel == 1) {  // bias gradient: the "ones" column (63) of the extra block
        uint32_t v[32];
        tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + 256 + 32, v);
        tmem_wait_ld();
        if (rows_valid) atomicAdd(a.G + b_off(jb.layer) + n_base + lane, __uint_as_float(v[31]));
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_remote(free_addr);
    }
  }
  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) tmem_dealloc2(tmem, 512);
}

// Heads on CUDA cores: alpha_linear (d sigma^T h7) and rgb_linear (d rgb^T hv) + their biases, straight from the
// saved arena blocks.  One CTA per SM, 12 warps: warp w streams head block w % 6 (0..3: h7, 4..5: hv) of every second
// tile of its CTA; a lane owns one 16-byte chunk (8 columns) of every fourth row, so a warp instruction reads 8 runs
// of 64 contiguous bytes.  Partial sums are folded inside the CTA first: the 644 outputs are shared by every CTA, and
// contended red.global.add is what this kernel would otherwise spend its time on.
constexpr int kHeadThreads = 384;
__global__ void __launch_bounds__(kHeadThreads) wgrad_heads_kernel(const uint8_t* __restrict__ act_arena, const float4* __restrict__ d_raw,
                                                                   int64_t M, int n_tiles, int fine, float* __restrict__ G) {
  __shared__ float s_red[6][8][28];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int hb = warp % 6, tp = warp / 6;
  const int j = lane & 7, rs = lane >> 3;
  const bool is_rgb = hb >= 4;
  const bool active = !(is_rgb && !fine);
  const int slot = is_rgb ? arena::kActHV + (hb - 4) : arena::kActH + 28 + hb;
  float acc[3][8], bs[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int k = 0; k < 3; ++k)
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[k][e] = 0.f;
  if (active) {
    for (int t = blockIdx.x + gridDim.x * tp; t < n_tiles; t += 2 * gridDim.x) {
      const uint8_t* blk = act_arena + (size_t)t * arena::act_tile_bytes() + (size_t)slot * 16384;
      const int64_t g0 = (int64_t)t * 128;
#pragma unroll 8
      for (int i = 0; i < 32; ++i) {
        const int r = 4 * i + rs;
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(blk + arena::cm_row_offset(r) + j * 1024));
        const float4 dr = g0 + r < M ? __ldg(d_raw + g0 + r) : make_float4(0.f, 0.f, 0.f, 0.f);
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float x0 = __uint_as_float(w[e] << 16), x1 = __uint_as_float(w[e] & 0xFFFF0000u);
          if (is_rgb) {
            acc[0][2 * e] = fmaf(dr.x, x0, acc[0][2 * e]); acc[0][2 * e + 1] = fmaf(dr.x, x1, acc[0][2 * e + 1]);
            acc[1][2 * e] = fmaf(dr.y, x0, acc[1][2 * e]); acc[1][2 * e + 1] = fmaf(dr.y, x1, acc[1][2 * e + 1]);
            acc[2][2 * e] = fmaf(dr.z, x0, acc[2][2 * e]); acc[2][2 * e + 1] = fmaf(dr.z, x1, acc[2][2 * e + 1]);
          } else {
            acc[0][2 * e] = fmaf(dr.w, x0, acc[0][2 * e]); acc[0][2 * e + 1] = fmaf(dr.w, x1, acc[0][2 * e + 1]);
          }
        }
        if (j == 0) { bs[0] += dr.x; bs[1] += dr.y; bs[2] += dr.z; bs[3] += dr.w; }
      }
    }
  }
  // fold the four row groups (lanes l, l^8, l^16, l^24), then the two warps of a head block
#pragma unroll
  for (int k = 0; k < 3; ++k)
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      acc[k][e] += __shfl_xor_sync(0xffffffffu, acc[k][e], 8);
      acc[k][e] += __shfl_xor_sync(0xffffffffu, acc[k][e], 16);
    }
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    bs[k] += __shfl_xor_sync(0xffffffffu, bs[k], 8);
    bs[k] += __shfl_xor_sync(0xffffffffu, bs[k], 16);
  }
  if (tp == 1 && rs == 0) {
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
      for (int e = 0; e < 8; ++e) s_red[hb][j][k * 8 + e] = acc[k][e];
#pragma unroll
    for (int k = 0; k < 4; ++k) s_red[hb][j][24 + k] = bs[k];
  }
  __syncthreads();
  if (tp == 0 && rs == 0 && active) {
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[k][e] += s_red[hb][j][k * 8 + e];
#pragma unroll
    for (int k = 0; k < 4; ++k) bs[k] += s_red[hb][j][24 + k];
    if (is_rgb) {
      const int c0 = (hb - 4) * 64 + 8 * j;
#pragma unroll
      for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int e = 0; e < 8; ++e) atomicAdd(G + w_off(L_RGB) + k * 128 + c0 + e, acc[k][e]);
      if (hb == 4 && j == 0) for (int k = 0; k < 3; ++k) atomicAdd(G + b_off(L_RGB) + k, bs[k]);
    } else {
      const int c0 = hb * 64 + 8 * j;
#pragma unroll
      for (int e = 0; e < 8; ++e) atomicAdd(G + w_off(L_ALPHA) + c0 + e, acc[0][e]);
      if (hb == 0 && j == 0) atomicAdd(G + b_off(L_ALPHA), bs[3]);
    }
  }
}

// Tile-range length: units (job, range) are dealt round-robin to the CTA pairs; pick the range length that minimises the
// busiest pair's work, one flush costing about kFlushTiles tile-loads.
static int pick_chunk(int n_tiles, int n_jobs, int pairs) {
  constexpr double kFlushTiles = 10.0;
  double best = 1e300; int best_chunk = n_tiles;
  std::vector<double> load(pairs);
  for (int chunk = std::min(n_tiles, 8); chunk <= n_tiles; ++chunk) {
    const int ranges = (n_tiles + chunk - 1) / chunk, units = n_jobs * ranges;
    std::fill(load.begin(), load.end(), 0.0);
    for (int u = 0; u < units; ++u) {
      const int t0 = (u / n_jobs) * chunk, len = std::min(chunk, n_tiles - t0);
      load[u % pairs] += job_weight(u % n_jobs) * len + kFlushTiles;
    }
    const double mk = *std::max_element(load.begin(), load.end());
    if (mk < best) { best = mk; best_chunk = chunk; }
  }
  return best_chunk;
}

}  // namespace wg

// flat_grad_net += weight / bias gradients of one net from the arenas written by the saving forward and the dgrad chain.
int mlp_tc2_wgrad(NerfEngine* h, bool full, int64_t M, const float* d_raw, const uint8_t* act_arena, const uint8_t* grad_arena,
                  float* flat_grad_net, cudaStream_t st) {
  using namespace wg;
  if (M <= 0) return NERF_OK;
  static bool attr_set_dev[64] = {};          // the opt-in shared-memory size is a per-device function attribute
  bool& attr_set = attr_set_dev[h->device & 63];
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(mlp_tc2_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
    if (e != cudaSuccess) return h->fail(NERF_E_CUDA, std::string("mlp_tc2_wgrad smem attribute: ") + cudaGetErrorString(e));
    attr_set = true;
  }
  WgradArgs a;
  a.act_arena = act_arena; a.grad_arena = grad_arena; a.G = flat_grad_net;
  a.n_tiles = (int)arena::num_tiles(M);
  a.n_jobs = full ? kNumJobsFine : kNumJobsCoarse;
  const int pairs = h->num_sms / 2;
  static int memo_tiles[2] = {-1, -1}, memo_chunk[2] = {0, 0};
  if (memo_tiles[full] != a.n_tiles) { memo_tiles[full] = a.n_tiles; memo_chunk[full] = pick_chunk(a.n_tiles, a.n_jobs, pairs); }
  a.chunk = memo_chunk[full];
  a.n_units = a.n_jobs * ((a.n_tiles + a.chunk - 1) / a.chunk);
  const int grid = 2 * std::min(a.n_units, pairs);
  // The heads kernel is independent of the main kernel (disjoint outputs) and small: it runs on a forked stream next
  // to the persistent main kernel (5 KB of shared memory and 384 threads fit beside it on every SM).
  if (!h->side) {
    if (cudaStreamCreateWithFlags(&h->side, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&h->side_fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&h->side_join, cudaEventDisableTiming) != cudaSuccess)
      return h->fail(NERF_E_CUDA, "mlp_tc2_wgrad: side stream");
  }
  cudaEventRecord(h->side_fork, st);
  cudaStreamWaitEvent(h->side, h->side_fork, 0);
  prefer_max_shared(reinterpret_cast<const void*>(wgrad_heads_kernel), h->device);
  wgrad_heads_kernel<<<std::min(a.n_tiles, h->num_sms), kHeadThreads, 0, h->side>>>(act_arena, reinterpret_cast<const float4*>(d_raw), M,
                                                                                    a.n_tiles, full ? 1 : 0, flat_grad_net);
  cudaEventRecord(h->side_join, h->side);
  mlp_tc2_wgrad_kernel<<<grid, kThreads, kSmemBytes, st>>>(a);
  cudaStreamWaitEvent(st, h->side_join, 0);
  h->launches += 2;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return h->fail(NERF_E_CUDA, std::string("mlp_tc2_wgrad launch: ") + cudaGetErrorString(e));
  return NERF_OK;
}

NERF_TC_DEFINE_DEBUG_SINK(tc2b_set_debug_sink)

}  // namespace nerf
