// mlp_tc2.cu -- fused NeRF MLP forward on CTA pairs (tcgen05 cta_group::2), the default bf16 fast path.
// Replaces Network.forward + batchify + NeRF.forward (network.py:49-74,163-171,199-258).
//
// Why pairs: with one CTA per SM (mlp_tc.cu) every SM streams the full 1.19 MB of layer weights per 256 rows,
// the 64 KB weight ring leaves no shared memory for biases (they were fetched from L2 in the epilogue: the
// critical path of round-1a, profiles/mlp_r1a_hot_sass.txt) and the B operand costs 64 B/clk of shared-memory
// bandwidth.  A CTA pair issues M=256 MMAs: each CTA supplies its own 128 rows of A and HALF of the weight tile,
// so per SM the weight traffic, the weight ring and the B-operand bandwidth are halved.
//
// Per CTA (cluster = 2 CTAs on one TPC, persistent over "pair super-tiles" of 512 rows; a CTA owns 256 of them
// as two 128-row tiles X, Y):
//   warp 0      weight producer: this CTA's half of every stage image (16 KB: 128 weight rows x 64 k, bf16,
//               128B-swizzled by the packer) by 2-SM TMA tile loads into a 3-deep ring; both halves complete on the
//               LEADER's mbarrier (.cta_group::2); bias slabs (4 KB) by cp.async.bulk into their own 2-slot ring
//   warp 1      rank 0: MMA issuer -- the warp runs converged, one elected lane issues tcgen05.mma.cta_group::2
//               (M=256, N=256|128, K=16); tcgen05.commit multicast to both CTAs' barriers.  Render: every stage feeds X
//               then Y, X one operation ahead; training: one tile's whole layer, then the other's
//               rank 1: forwarder -- relays "my half of the bias slab has landed" to the leader's barrier
//               both: TMEM allocation (cta_group::2, 512 columns: X accumulators [0,256), Y [256,512))
//   warps 2-5   tile X: positional encoding (computed ahead, under the MMAs) -> bf16 A operand; per layer TMEM -> ReLU ->
//               bf16 -> A operand of the next layer (in place); sigma / rgb heads on CUDA cores
//   warps 6-9   same for tile Y
// Shared memory (bytes): A_X 64K | A_Y 64K | PE_X 16K | PE_Y 16K | W ring 3 x 16K | ones slab 4K | bias ring 2 x 4K | heads 2.5K | barriers
// Biases: folded into the GEMM -- every layer starts with one K = 16 MMA (A = constant slab whose columns 0..2 are 1.0,
// B = the layer's bias slab: bias split into three bf16 terms) that also initialises the accumulator, so the epilogue is
// TMEM -> ReLU/bf16 -> st.shared only (the 64 LDS.128 + 64 FADD2 per row and layer it replaces cost 11 % of the kernel).
#include <cuda.h>            // CUtensorMap (types only: the encoder is fetched through cudaGetDriverEntryPoint)
#include "engine.h"
#include "tc_common.cuh"
#include "rows.cuh"
#include "tc_arena.cuh"

namespace nerf {
using namespace tc;

namespace v2 {

constexpr int kTile = 128;
constexpr int kRowsPerCta = 2 * kTile;          // X + Y
constexpr int kRowsPerPair = 2 * kRowsPerCta;   // 512
// (a packed stage image is 256 rows x 128 B = 32 KB, both halves: mlp_tc.cu pack_weights_kernel; the tensor maps address it by rows)
constexpr int kHalfBytes = 16384;               // this CTA's half of a 256-row stage (views layer: 8192)
constexpr int kRingReal = 3;
// diagnostic builds (NERF_EXP bits 3-4): 1 or 2 extra ring slots ALIASED onto the PE buffers -- wrong results, but it times a
// deeper weight ring before shared memory is found for one
constexpr int kRing = kRingReal + ((NERF_EXP >> 3) & 3);
constexpr int kNumSched = 10;                   // L0..L7, FEAT, VIEWS
constexpr int kNumStages = 39;
// One warp can issue one bulk / TMA copy (expect_tx + cp.async.bulk.tensor) per ~300-450 clk whatever its size, several warps
// issue in parallel (profiles/ubench/handshake.cu: 307 -> 154 -> 102 -> 77 clk per 1 KB stage with 1 / 2 / 3 / 4 issuing warps;
// 16 KB stages then reach 86 B/clk/SM).  A single producer warp paced the per-tile (training) order at ~870 clk per stage
// against 512 clk of MMAs, so the weight ring is fed by kNumProd warps, warp p taking the stages it % kNumProd == p.
#ifndef NERF_NUM_PROD
#define NERF_NUM_PROD 3
#endif
constexpr int kNumProd = NERF_NUM_PROD;
constexpr int kFirstExtraProd = 10;   // warps: 0 producer 0, 1 issuer / forwarder (+ TMEM alloc), 2-9 tile warps (TMEM lane quarter = warp % 4), 10.. producers 1..
constexpr int kThreads = 320 + 32 * (kNumProd - 1);      // 384 threads -> still 168 registers per thread
constexpr int kOnesBytes = 4096;                // constant A slab of the bias MMA: 128 rows x 16 k, columns 0..2 = 1.0
constexpr int kBiasSlabBytes = 8192;            // packed bias slab of one schedule entry (both halves), mlp_tc.cu pack_bias_slabs_kernel
constexpr uint32_t kOffA = 0, kOffPE = 131072, kOffW = 163840;
constexpr uint32_t kOffBias = kOffW + kRingReal * kHalfBytes;     // 212992
constexpr int kBiasRing = 2;                                      // bias slabs have their own small ring (4 KB slots)
constexpr uint32_t kOffBiasRing = kOffBias + kOnesBytes;          // 217088
constexpr uint32_t kOffHeads = kOffBiasRing + kBiasRing * 4096;   // 225280: w_alpha[256] | w_rgb[384]
constexpr uint32_t kOffBar = kOffHeads + (256 + 384) * 4;         // 227840
constexpr uint32_t kSmemBytes = kOffBar + 256;                    // 228096 <= 232448

__host__ __device__ constexpr int sched_nkb(int li) { return li == 0 ? 1 : (li == 5 || li == 9) ? 5 : 4; }
__host__ __device__ constexpr int sched_start(int li) {
  int s = 0;
  for (int i = 0; i < li; ++i) s += sched_nkb(i);
  return s;
}
static_assert(sched_start(kNumSched) == kNumStages, "stage count");
__host__ __device__ constexpr int sched_n(int li) { return li == 9 ? 128 : 256; }
__host__ __device__ constexpr int sched_layer(int li) { return li < 8 ? li : (li == 8 ? L_FEAT : L_VIEWS); }

__host__ __device__ constexpr uint32_t ring_slot_off(uint32_t slot) { return slot < kRingReal ? kOffW + slot * kHalfBytes : kOffPE + (slot - kRingReal) * kHalfBytes; }

struct Bars {
  uint64_t w_full[kRing], w_empty[kRing], b_full[kBiasRing], b_empty[kBiasRing], acc_full[2], act_ready[2], acc_free[2];
  uint32_t tmem_slot;
};

__device__ __forceinline__ void store_chunk(uint8_t* blk, uint32_t row_base, uint32_t row_xor, int chunk, uint32_t a,
                                            uint32_t b, uint32_t c, uint32_t d) {
  if (NERF_EXP & 1024) return;     // timing experiment: the epilogues write no A operand
  *reinterpret_cast<uint4*>(blk + row_base + (((uint32_t)chunk ^ row_xor) << 4)) = make_uint4(a, b, c, d);
}

// sin / cos of (x * 2^k) for the bf16 A operand: x * 2^k / (2 pi) is reduced to [-0.5, 0.5] turns with a two-term
// constant (error ~1e-7 turns, far below bf16 resolution), then MUFU.SIN / MUFU.COS on [-pi, pi].
// c_hi/c_lo = 2^k / (2 pi) split in two floats.
__device__ __forceinline__ void sincos_turns(float x, float c_hi, float c_lo, float& s, float& c) {
  const float r = x * c_hi;
  const float k = rintf(r);
  float fr = fmaf(x, c_hi, -k);
  fr = fmaf(x, c_lo, fr);
  const float ang = fr * 6.283185307179586f;
  s = __sinf(ang);
  c = __cosf(ang);
}
constexpr float kInv2PiHi = 0.15915493667125702f;          // fl(1/(2 pi))
constexpr float kInv2PiLo = 6.4206382432985265e-09f;          // 1/(2 pi) - kInv2PiHi

// One 64-column bf16 K-block row, [v0 v1 v2 | sin(2^k v) cos(2^k v)]_{k<L} | zero pad | 1.0, packed into 32 registers (8 chunks
// of 8 columns).  The encoding is computed ahead of time, while the tile's MMAs run and its epilogue warps would only spin
// on the accumulator barrier (60 + 24 MUFU pairs per row are ~4 % of the kernel when they sit between two layers).
template <int L>
__device__ __forceinline__ void compute_pe_block(const float v[3], uint32_t (&out)[32]) {
  float col[8];
  int n = 0, chunk = 0;
  auto push = [&](float x) {
    col[n++] = x;
    if (n == 8) {
      out[4 * chunk + 0] = pack_bf16x2(col[0], col[1]); out[4 * chunk + 1] = pack_bf16x2(col[2], col[3]);
      out[4 * chunk + 2] = pack_bf16x2(col[4], col[5]); out[4 * chunk + 3] = pack_bf16x2(col[6], col[7]);
      ++chunk;
      n = 0;
    }
  };
  push(v[0]); push(v[1]); push(v[2]);
  float chi = kInv2PiHi, clo = kInv2PiLo;
#pragma unroll
  for (int k = 0; k < L; ++k, chi *= 2.0f, clo *= 2.0f) {
    float sn[3], cs[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) sincos_turns(v[c], chi, clo, sn[c], cs[c]);
    push(sn[0]); push(sn[1]); push(sn[2]); push(cs[0]); push(cs[1]); push(cs[2]);
  }
#pragma unroll
  for (int z = 3 + 6 * L; z < 63; ++z) push(0.f);
  push(1.0f);     // column 63 = 1: its weight column is zero in the forward; in wgrad it yields the bias gradient
}
// Stores the packed row into the swizzled K-block; grow != nullptr (training forward): the same chunks also go to this row of
// the block's arena image (tc_arena.cuh).
__device__ __forceinline__ void store_pe_block(uint8_t* blk, uint32_t row_base, uint32_t row_xor, const uint32_t (&pe)[32], uint8_t* grow) {
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    if (grow) st_global_cs(grow + c * 1024, pe[4 * c], pe[4 * c + 1], pe[4 * c + 2], pe[4 * c + 3]);
    if (blk) store_chunk(blk, row_base, row_xor, c, pe[4 * c], pe[4 * c + 1], pe[4 * c + 2], pe[4 * c + 3]);
  }
}

// Experiment kept behind NERF_DEFER_STORES (default 0 = off): write the arena copy of a finished layer AFTER the tile has
// been handed back to the MMA issuer, by re-reading the row's 4 x 128 bytes from the swizzled A operand (untouched until
// the same threads run the tile's next epilogue), to take the 32 st.global.v4 per row and the proxy fence behind them off
// the MMA -> epilogue -> MMA chain (without any arena store the forward is 10 % and the dgrad chain 21 % faster).
// Measured on B200: SLOWER (forward 0.400 -> 0.426 ms, dgrad 0.348 -> 0.368 ms per 1024-ray step): the extra 64 KB of
// shared-memory reads per tile and layer compete with the tensor core's operand fetch (SS-mode MMAs read 64 B/clk of the
// 128 B/clk the SM has) -- the same reason the bulk shared->global store variant of round 1 lost.
#ifndef NERF_DEFER_STORES
#define NERF_DEFER_STORES 0
#endif
__device__ __forceinline__ void flush_row_to_arena(const uint8_t* sAct, uint32_t row_base, uint32_t row_xor, uint8_t* grow, int n_blocks) {
#pragma unroll
  for (int b = 0; b < 4; ++b) {
    if (b < n_blocks) {
#pragma unroll
      for (int ch = 0; ch < 8; ++ch) {
        const uint4 v = *reinterpret_cast<const uint4*>(sAct + b * 16384 + row_base + (((uint32_t)ch ^ row_xor) << 4));
        st_global_cs(grow + b * 16384 + ch * 1024, v.x, v.y, v.z, v.w);
      }
    }
  }
}

// One 32-column chunk of a trunk-layer epilogue: acc + bias -> (ReLU) -> bf16 -> 4 x 16-byte chunks of the next
// layer's A operand.  MODE 0: ReLU; 1: ReLU + sigma head accumulation on the fp32 values; 2: linear (feature).
// grow != nullptr (training forward): the chunk is also stored to this row of the layer's arena blocks, and mbits
// receives the ReLU bit mask of the 32 columns.
template <int MODE>
__device__ __forceinline__ void epi_chunk(const uint32_t (&v)[32], int c0,
                                          const float* __restrict__ s_walpha, float& sigma, uint8_t* sAct,
                                          uint32_t row_base, uint32_t row_xor, uint8_t* grow, uint8_t* gmask, uint32_t (&mbits)[8]) {
  uint32_t pk[16];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float f0 = __uint_as_float(v[4 * j]), f1 = __uint_as_float(v[4 * j + 1]);      // bias already in the accumulator
    const float f2 = __uint_as_float(v[4 * j + 2]), f3 = __uint_as_float(v[4 * j + 3]);
    if (MODE == 2) {
      pk[2 * j] = pack_bf16x2(f0, f1); pk[2 * j + 1] = pack_bf16x2(f2, f3);
    } else {
      if (MODE == 1) {
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
  if (grow) {
    if (MODE != 2 && !(NERF_EXP & 64)) mbits[c0 >> 5] = pos_bits32(pk);      // the row's mask record is stored once per layer
    if (!NERF_DEFER_STORES) {         // (deferred flavour: arena copy after the hand-over, flush_row_to_arena)
      uint8_t* gb = grow + (c0 >> 6) * 16384 + ch0 * 1024;
#pragma unroll
      for (int j = 0; j < 4; ++j) st_global_cs(gb + j * 1024, pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
    }
  }
}

// Whole 256-column epilogue of one tile row, TMEM loads software-pipelined against the math of the previous chunk.
// grow / gmask (training forward): this row of the layer's arena blocks / of its bit-mask record.
template <int MODE>
__device__ __forceinline__ void epi_layer(uint32_t taddr, const float* __restrict__ s_walpha,
                                          float& sigma, uint8_t* sAct, uint32_t row_base, uint32_t row_xor, uint8_t* grow,
                                          uint8_t* gmask, uint32_t (&mbits)[8]) {
  uint32_t va[32], vb[32];
  tmem_ld32(taddr, va);
#pragma unroll
  for (int c = 0; c < 8; c += 2) {
    tmem_wait_ld();
    tmem_ld32(taddr + (c + 1) * 32, vb);
    epi_chunk<MODE>(va, c * 32, s_walpha, sigma, sAct, row_base, row_xor, grow, gmask, mbits);
    tmem_wait_ld();
    if (c + 2 < 8) tmem_ld32(taddr + (c + 2) * 32, va);
    epi_chunk<MODE>(vb, (c + 1) * 32, s_walpha, sigma, sAct, row_base, row_xor, grow, gmask, mbits);
  }
  if (gmask && MODE != 2 && !NERF_DEFER_STORES) st_global_mask8(gmask, mbits);
}

// View-branch epilogue: 128 columns, ReLU, then the 128 -> 3 rgb head on CUDA cores.
template <bool SAVE>
__device__ __forceinline__ void epi_views(uint32_t taddr, const float* __restrict__ s_wrgb,
                                          float& r0, float& r1, float& r2, uint8_t* grow, uint8_t* gmask) {
  uint32_t va[32], vb[32];
  uint32_t mv[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};      // 128 columns: words 0..3 of the row's mask record
  auto chunk = [&](const uint32_t (&v)[32], int c0) {
    uint32_t pk[16];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float f0 = __uint_as_float(v[4 * j]), f1 = __uint_as_float(v[4 * j + 1]);
      const float f2 = __uint_as_float(v[4 * j + 2]), f3 = __uint_as_float(v[4 * j + 3]);
      if (SAVE) { pk[2 * j] = pack_bf16x2_relu(f0, f1); pk[2 * j + 1] = pack_bf16x2_relu(f2, f3); }
      const float hv[4] = {fmaxf(f0, 0.f), fmaxf(f1, 0.f), fmaxf(f2, 0.f), fmaxf(f3, 0.f)};
      const float4 w0 = *reinterpret_cast<const float4*>(s_wrgb + c0 + 4 * j);
      const float4 w1 = *reinterpret_cast<const float4*>(s_wrgb + 128 + c0 + 4 * j);
      const float4 w2 = *reinterpret_cast<const float4*>(s_wrgb + 256 + c0 + 4 * j);
      r0 = fmaf(hv[0], w0.x, r0); r0 = fmaf(hv[1], w0.y, r0); r0 = fmaf(hv[2], w0.z, r0); r0 = fmaf(hv[3], w0.w, r0);
      r1 = fmaf(hv[0], w1.x, r1); r1 = fmaf(hv[1], w1.y, r1); r1 = fmaf(hv[2], w1.z, r1); r1 = fmaf(hv[3], w1.w, r1);
      r2 = fmaf(hv[0], w2.x, r2); r2 = fmaf(hv[1], w2.y, r2); r2 = fmaf(hv[2], w2.z, r2); r2 = fmaf(hv[3], w2.w, r2);
    }
    if (SAVE) {   // keep the hidden view activations (arena image + bit mask) for the backward pass
      uint8_t* gb = grow + (c0 >> 6) * 16384 + ((c0 & 63) >> 3) * 1024;
#pragma unroll
      for (int j = 0; j < 4; ++j) st_global_cs(gb + j * 1024, pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
      if (!(NERF_EXP & 64)) mv[c0 >> 5] = pos_bits32(pk);
    }
  };
  tmem_ld32(taddr, va);
#pragma unroll
  for (int c = 0; c < 4; c += 2) {
    tmem_wait_ld();
    tmem_ld32(taddr + (c + 1) * 32, vb);
    chunk(va, c * 32);
    tmem_wait_ld();
    if (c + 2 < 4) tmem_ld32(taddr + (c + 2) * 32, va);
    chunk(vb, (c + 1) * 32);
  }
  if (SAVE) st_global_mask8(gmask, mv);
}

// TMA tile load of this CTA's half of a weight stage (rows [row0, row0 + box rows) of the packed image seen as a
// [rows x 64] bf16 matrix) that completes on an mbarrier of EITHER CTA of the pair (.cta_group::2): both halves signal the
// leader's w_full directly.  A plain cp.async.bulk can only signal a barrier of its destination CTA, so the peer's arrival
// had to be relayed by a forwarder warp (peer barrier -> poll -> remote arrive): that hop sat on the ring's round trip.
// The maps see the packed image as rows of (128 << kTmaRowShift) bytes.  A CTA's half of a stage is one contiguous run, so
// any 2-D factorisation copies the same bytes.  Measured A/B on the C2 render (profiles/exp_train_r2.md): 128-byte rows
// 154.2-154.4 ms, 512-byte rows 154.9-155.5 ms, training step unchanged -> 128-byte rows.
constexpr int kTmaRowShift = 0;
__device__ __forceinline__ void tma_stage_cg2(void* smem_dst, const CUtensorMap* tmap, int row0, uint32_t mbar_cluster_addr) {
  asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(tmap), "r"(0), "r"(row0 >> kTmaRowShift), "r"(mbar_cluster_addr)
               : "memory");
}

// SAVE: training forward -- every A-operand block (PE, h_0..h_7, feature, PE(dir), view hidden) is also written to the
// activation arena straight from the epilogue registers (coalesced st.global.v4 in the arena's core-matrix order,
// tc_arena.cuh), together with the ReLU bit masks the dgrad chain needs.  (Bulk shared->global stores of the same
// blocks cost 40 % of the kernel: their shared-memory reads compete with the MMA operand fetch.)
// NERF_TWO_PHASE (training forward): the epilogue first does what the render epilogue does -- accumulator -> ReLU -> bf16 ->
// next layer's A operand -> hand-over -- and only then reads the SAME accumulator a second time to write the arena copy and
// the ReLU bit masks.  In the per-tile order the issuer is busy with the other tile's MMAs for ~2 000 clk after a hand-over,
// so the second pass (~1 300 clk of stores) was meant to leave the MMA -> epilogue -> MMA chain; a second barrier (acc_free)
// keeps the issuer from overwriting an accumulator that is still being read.  Parity green, but SLOWER (forward 0.37 -> 0.43
// ms, dgrad 0.34 -> 0.43 ms): the chain of ONE tile is  MMAs -> epilogue -> accumulator free -> MMAs, and with a single
// accumulator per tile the second pass sits on it just like the stores did -- only now after a second TMEM read and a second
// pack.  Kept as an option (default off).
#ifndef NERF_TWO_PHASE
#define NERF_TWO_PHASE 0      // measured: forward 0.37 -> 0.43 ms, dgrad 0.34 -> 0.43 ms (see the comment in mlp_tc2.cu / profiles/exp_train_r2.md)
#endif
// second pass of a trunk / feature layer: accumulator (bias included) -> (ReLU) -> bf16 -> this row of the arena blocks + mask record
template <int MODE>
__device__ __forceinline__ void epi_store_layer(uint32_t taddr, uint8_t* grow, uint8_t* gmask) {
  uint32_t va[32], vb[32], mbits[8];
  auto chunk = [&](const uint32_t (&v)[32], int c0) {
    uint32_t pk[16];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float f0 = __uint_as_float(v[4 * j]), f1 = __uint_as_float(v[4 * j + 1]);
      const float f2 = __uint_as_float(v[4 * j + 2]), f3 = __uint_as_float(v[4 * j + 3]);
      if (MODE == 2) { pk[2 * j] = pack_bf16x2(f0, f1); pk[2 * j + 1] = pack_bf16x2(f2, f3); }
      else { pk[2 * j] = pack_bf16x2_relu(f0, f1); pk[2 * j + 1] = pack_bf16x2_relu(f2, f3); }
    }
    uint8_t* gb = grow + (c0 >> 6) * 16384 + ((c0 & 63) >> 3) * 1024;
#pragma unroll
    for (int j = 0; j < 4; ++j) st_global_cs(gb + j * 1024, pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
    if (MODE != 2) mbits[c0 >> 5] = pos_bits32(pk);
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
  if (MODE != 2) st_global_mask8(gmask, mbits);
}

template <bool SAVE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
mlp_tc2_forward_kernel(const __grid_constant__ CUtensorMap tm128, const __grid_constant__ CUtensorMap tm64, int stage_row0, int n_rep,
                       RowSrc src, float4* __restrict__ raw, const uint8_t* __restrict__ wbias, const float* __restrict__ P, int64_t M, int64_t n_pair,
                       uint8_t* __restrict__ act_arena) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if (!SAVE && src.row_count) {      // compacted row list (fast_sampling work skipping): the row count lives on the device
    M = min((int64_t)__ldg(src.row_count), M);
    n_pair = (M + kRowsPerPair - 1) / kRowsPerPair;
  }
  constexpr bool kPerTile = SAVE && !(NERF_EXP & 32);   // NERF_EXP bit 5: the saving forward in the render (shared-stage) order
  constexpr int kReps = kPerTile ? 2 : 1;            // weight passes per layer (training: once per tile, see the issuer)
  constexpr bool kTwoPhase = SAVE && NERF_TWO_PHASE && !NERF_DEFER_STORES;
  Bars& sb = *reinterpret_cast<Bars*>(smem + kOffBar);
  float* s_walpha = reinterpret_cast<float*>(smem + kOffHeads);
  float* s_wrgb = s_walpha + 256;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t rank = cluster_ctarank();
  const int64_t pair0 = blockIdx.x >> 1, pair_stride = gridDim.x >> 1;
  stage_row0 += (int)(pair0 % n_rep) * (2 * kNumStages * 256);              // this pair's replica of the stage images

  if ((smem_u32(smem) & 1023u) != 0) mbar_timeout(0xA12, smem_u32(smem));   // swizzle-128B needs 1024B alignment
  if (tid == 0) {
    for (int i = 0; i < kRing; ++i) { mbar_init(&sb.w_full[i], 1); mbar_init(&sb.w_empty[i], 1); }   // w_full: used in the leader only
    for (int i = 0; i < kBiasRing; ++i) { mbar_init(&sb.b_full[i], rank == 0 ? 2 : 1); mbar_init(&sb.b_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&sb.acc_full[i], 1); mbar_init(&sb.act_ready[i], 8); mbar_init(&sb.acc_free[i], 8); }
    fence_barrier_init();
  }
  // constant A slab of the bias MMAs: 128 rows x 16 k bf16, un-swizzled K-major core matrices, columns 0..2 = 1.0
  for (int i = tid; i < kOnesBytes / 16; i += kThreads)
    reinterpret_cast<uint4*>(smem + kOffBias)[i] = (i & 8) ? make_uint4(0u, 0u, 0u, 0u) : make_uint4(0x3F803F80u, 0x00003F80u, 0u, 0u);
  fence_proxy_async_smem();
  for (int i = tid; i < 256; i += kThreads) s_walpha[i] = __ldg(P + w_off(L_ALPHA) + i);
  for (int i = tid; i < 384; i += kThreads) s_wrgb[i] = __ldg(P + w_off(L_RGB) + i);
  cluster_sync_all();                      // barrier inits visible cluster-wide before any remote arrive / multicast commit
  if (warp == 1) { tmem_alloc2(&sb.tmem_slot, 512); tmem_relinquish2(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sb.tmem_slot;

  const int prod = warp == 0 ? 0 : (warp >= kFirstExtraProd ? warp - kFirstExtraProd + 1 : -1);
  if (prod >= 0) {
    // ===== weight producers (both CTAs): my half of every stage; producer p issues the stages it % kNumProd == p, the last
    // one also the bias slabs =====
    // The whole warp runs the loop converged (all lanes poll the barrier, lane 0 issues): a lone lane-0 loop inside a diverged
    // warp sees its barrier complete ~2 000 clk late (event trace, profiles/exp_train_r2.md).
    {
      uint32_t it = 0, itb = 0, tn_p = 0;
      const uint32_t w_full_leader = mapa_u32(smem_u32(&sb.w_full[0]), 0);
      for (int64_t st = pair0; st < n_pair; st += pair_stride) {
        for (int li = 0; li < kNumSched; ++li) {
          // half of the stage: N/2 rows x 128 B (NERF_EXP bit 11, timing experiment: 8 rows = 1 KB only, same handshakes)
          const uint32_t bytes = (NERF_EXP & 2048) ? 1024u : (uint32_t)sched_n(li) * 64u;
          const int row_li = stage_row0 + sched_start(li) * 256 + (int)rank * (sched_n(li) / 2);
          for (int rep = 0; rep < kReps; ++rep) {                             // training: once per tile (see the issuer)
            if (prod == kNumProd - 1) {                                       // the layer's bias slab (N/2 rows x 32 B), own ring
              const uint32_t slot = itb % kBiasRing, par = (itb / kBiasRing) & 1u;
              mbar_wait(&sb.b_empty[slot], par ^ 1u, 0x18 + slot);
              if (lane == 0) {
                mbar_arrive_expect_tx(&sb.b_full[slot], bytes / 4);
                bulk_g2s(smem + kOffBiasRing + slot * 4096, wbias + (size_t)li * kBiasSlabBytes + (size_t)rank * 4096, bytes / 4, &sb.b_full[slot]);
              }
              __syncwarp();
            }
            ++itb;
            for (int kb = 0; kb < sched_nkb(li); ++kb, ++it) {
              if ((int)(it % kNumProd) != prod) continue;
              const uint32_t slot = it % kRing, par = (it / kRing) & 1u;
              mbar_wait(&sb.w_empty[slot], par ^ 1u, 0x10 + slot);
              if (lane == 0) {
                if (prod == 0) trace_event(0x3000u | (rep << 8) | (li << 4) | (4u + kb), 1024, tn_p, 3000, 5000);
                // both halves complete on the leader's barrier, which expects the bytes of the pair
                if (rank == 0) mbar_arrive_expect_tx(&sb.w_full[slot], 2 * bytes);
                tma_stage_cg2(smem + ring_slot_off(slot), sched_n(li) == 256 ? &tm128 : &tm64,
                              row_li + kb * 256, w_full_leader + slot * 8u);
                if (prod == 0) trace_event(0x3000u | (rep << 8) | (li << 4) | (9u + kb), 1024, tn_p, 3000, 5000);      // load issued
              }
              __syncwarp();
            }
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1 && rank == 1) {
    // ===== forwarder (peer CTA): relay the arrival of my half of a bias slab to the leader's b_full (converged warp) =====
    {
      uint32_t itb = 0;
      for (int64_t st = pair0; st < n_pair; st += pair_stride)
        for (int li = 0; li < kNumSched; ++li)
          for (int rep = 0; rep < kReps; ++rep) {
            const uint32_t bslot = itb % kBiasRing, bpar = (itb / kBiasRing) & 1u;
            ++itb;
            mbar_wait(&sb.b_full[bslot], bpar, 0x58 + bslot);
            if (lane == 0) mbar_arrive_remote(mapa_u32(smem_u32(&sb.b_full[bslot]), 0));
            __syncwarp();
          }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===== MMA issuer (leader CTA) =====
    {                                   // the whole warp, converged: see umma_ss2_w
      uint32_t it = 0, itb = 0, lc = 0, tn = 0;
      const uint32_t sA = smem_u32(smem + kOffA), sPE = smem_u32(smem + kOffPE);
      for (int64_t st = pair0; st < n_pair; st += pair_stride) {
        for (int li = 0; li < kNumSched; ++li, ++lc) {
          const int nkb = sched_nkb(li);
          const uint32_t idesc = make_idesc_bf16(256, sched_n(li));
          // A source of k-block kb: -1 = PE buffer, else hidden-activation block
          auto a_block = [&](int kb) { return li == 0 ? -1 : li == 5 ? kb - 1 : li == 9 ? (kb < 4 ? kb : -1) : kb; };
          // bias slab: one K = 16 MMA (un-swizzled K-major operands) that also initialises the accumulator
          auto issue_bias = [&](int tile, uint32_t bslot) {
            umma_ss2_w(tmem + tile * 256u, make_smem_desc(smem_u32(smem + kOffBias), 128, 256, kLayoutNone),
                     make_smem_desc(smem_u32(smem + kOffBiasRing) + bslot * 4096u, 128, 256, kLayoutNone), idesc, 0u);
          };
          auto issue = [&](int tile, int kb, uint32_t slot) {
            const int ablk = a_block(kb);
            const uint32_t a_base = ablk < 0 ? sPE + tile * 16384u : sA + tile * 65536u + (uint32_t)ablk * 16384u;
            const uint32_t b_base = smem_u32(smem) + ring_slot_off(slot);
            umma_ss2_w_x4(tmem + tile * 256u, make_smem_desc(a_base, 16, 1024, kLayoutSW128), make_smem_desc(b_base, 16, 1024, kLayoutSW128), idesc, 1u);
          };
          if (kPerTile) {
            // Training forward: one tile's whole layer, then the other's -- the layer's weights are streamed twice (L2
            // hits) so that tile Y's MMAs run under tile X's long (arena-storing) epilogue.
#pragma unroll
            for (int tile = 0; tile < 2; ++tile) {
              mbar_wait(&sb.act_ready[tile], lc & 1u, 0x30 + tile);
              if (kTwoPhase) mbar_wait(&sb.acc_free[tile], lc & 1u, 0x38 + tile);      // ... and its accumulator has been read twice
              tc_fence_after();
              if (lane == 0) trace_event(0x1000u | (tile << 8) | (li << 4) | 1u, 1024, tn, 0, 3000);
              {
                const uint32_t bslot = itb % kBiasRing, bpar = (itb / kBiasRing) & 1u;
                ++itb;
                if (!(NERF_EXP & 128)) mbar_wait(&sb.b_full[bslot], bpar, 0x28 + bslot);
                tc_fence_after();
                if (lane == 0) trace_event(0x1000u | (tile << 8) | (li << 4) | 2u, 1024, tn, 0, 3000);
                issue_bias(tile, bslot);
                umma_commit2_w(&sb.b_empty[bslot], 0x3);
              }
              for (int kb = 0; kb < nkb; ++kb, ++it) {
                const uint32_t slot = it % kRing, par = (it / kRing) & 1u;
                if (!(NERF_EXP & 2)) mbar_wait(&sb.w_full[slot], par, 0x20 + slot);
                tc_fence_after();
                if (lane == 0) trace_event(0x1000u | (tile << 8) | (li << 4) | (4u + kb), 1024, tn, 0, 3000);
                issue(tile, kb, slot);
                umma_commit2_w(&sb.w_empty[slot], 0x3);
              }
              umma_commit2_w(&sb.acc_full[tile], 0x3);
              if (lane == 0) trace_event(0x1000u | (tile << 8) | (li << 4) | 3u, 1024, tn, 0, 3000);
            }
          } else {
            // Render: every weight stage feeds both tiles before it is released (half the weight traffic of the per-tile
            // order, which starves on the ring's round trip).  Tile X runs one operation ahead of tile Y (bias MMA, then the
            // k-blocks): X's accumulator is complete -- and its epilogue starts -- one k-block earlier, and its first MMAs
            // of the next layer run under the tail of Y's epilogue instead of waiting for it.
            const uint32_t bslot = itb % kBiasRing, bpar = (itb / kBiasRing) & 1u;
            ++itb;
            mbar_wait(&sb.b_full[bslot], bpar, 0x28 + bslot);
            tc_fence_after();
            mbar_wait(&sb.act_ready[0], lc & 1u, 0x30);
            if (kTwoPhase) mbar_wait(&sb.acc_free[0], lc & 1u, 0x38);
            tc_fence_after();
            issue_bias(0, bslot);
            for (int kb = 0; kb < nkb; ++kb, ++it) {
              const uint32_t slot = it % kRing, par = (it / kRing) & 1u;
              mbar_wait(&sb.w_full[slot], par, 0x20 + slot);
              tc_fence_after();
              issue(0, kb, slot);
              if (kb == nkb - 1) umma_commit2_w(&sb.acc_full[0], 0x3);
              if (kb == 0) {
                mbar_wait(&sb.act_ready[1], lc & 1u, 0x31);
                if (kTwoPhase) mbar_wait(&sb.acc_free[1], lc & 1u, 0x39);
                tc_fence_after();
                issue_bias(1, bslot);
                umma_commit2_w(&sb.b_empty[bslot], 0x3);
              } else {
                const uint32_t pslot = (it - 1) % kRing;
                issue(1, kb - 1, pslot);
                umma_commit2_w(&sb.w_empty[pslot], 0x3);
              }
            }
            {
              const uint32_t pslot = (it - 1) % kRing;
              issue(1, nkb - 1, pslot);
              umma_commit2_w(&sb.acc_full[1], 0x3);
              umma_commit2_w(&sb.w_empty[pslot], 0x3);
            }
          }
        }
      }
    }
    __syncwarp();
  } else if (warp >= 2) {
    // ===== PE + epilogue warps =====
    const int tile = (warp - 2) >> 2, q = warp & 3;
    const int row = q * 32 + lane;
    const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16) + tile * 256u;
    uint8_t* sAct = smem + kOffA + tile * 65536;
    uint8_t* sPe = smem + kOffPE + tile * 16384;
    const uint32_t row_base = (uint32_t)((row >> 3) * 1024 + (row & 7) * 128), row_xor = (uint32_t)(row & 7);
    const uint32_t ready_addr = mapa_u32(smem_u32(&sb.act_ready[tile]), 0);
    auto signal_ready = [&]() {
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_remote(ready_addr);
    };
    const uint32_t free_addr = mapa_u32(smem_u32(&sb.acc_free[tile]), 0);
    auto signal_free = [&]() {              // every tcgen05.ld of this warp on the tile's accumulator has completed
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_remote(free_addr);
    };
    uint32_t lc = 0, tn_e = 0;
    // this row of pair super-tile st_: validity, global row (through the compacted row list when rows are skipped), point, view direction
    auto row_setup = [&](int64_t st_, bool& valid_, int64_t& g_, float (&p_)[3], float (&vd_)[3]) {
      const int64_t gl = st_ * kRowsPerPair + (int64_t)rank * kRowsPerCta + tile * kTile + row;
      valid_ = gl < M;
      g_ = (!SAVE && src.row_list && valid_) ? (int64_t)__ldg(src.row_list + gl) : gl;
      p_[0] = p_[1] = p_[2] = 0.f; vd_[0] = vd_[1] = vd_[2] = 0.f;
      if (valid_) row_point(src, g_, p_, vd_);
    };
    uint32_t pe[32];                      // the next PE block of this row, packed (PE(pts) of the next super-tile or PE(dir) of this one)
    bool valid = false;
    int64_t g = 0;
    float vd[3] = {0.f, 0.f, 0.f};
    if (pair0 < n_pair) {
      float p[3];
      row_setup(pair0, valid, g, p, vd);
      compute_pe_block<10>(p, pe);
    }
    for (int64_t st = pair0; st < n_pair; st += pair_stride) {
      // training forward: this row's slice of the tile's arena blocks / bit-mask records (tc_arena.cuh)
      uint8_t* tile_base = SAVE ? act_arena + (size_t)((st * 2 + rank) * 2 + tile) * arena::act_tile_bytes() : nullptr;
      uint8_t* grow0 = SAVE ? tile_base + arena::cm_row_offset(row) : nullptr;
      uint8_t* gmask0 = SAVE ? tile_base + arena::act_mask_offset() + row * arena::kMaskRowBytes : nullptr;
      auto gblk = [&](int slot) { return SAVE ? grow0 + (size_t)slot * arena::kBlockBytes : nullptr; };
      auto gmsk = [&](int layer) { return SAVE ? gmask0 + layer * arena::kMaskLayerBytes : nullptr; };
      constexpr bool kDefer = SAVE && NERF_DEFER_STORES;
      constexpr bool kLate = kDefer || kTwoPhase;       // arena copies of the PE blocks leave after the hand-over, from the registers
      store_pe_block(sPe, row_base, row_xor, pe, kLate ? nullptr : gblk(arena::kActPE));
      signal_ready();
      if (kTwoPhase) signal_free();                      // the previous super-tile's last epilogue has drained the accumulator
      if (kLate) store_pe_block(nullptr, row_base, row_xor, pe, gblk(arena::kActPE));

      float sigma = 0.f;
      bool nvalid = false;
      int64_t ng = 0;
      float nvd[3] = {0.f, 0.f, 0.f};
      for (int li = 0; li < kNumSched; ++li, ++lc) {
        // encodings computed under this tile's MMAs: PE(dir) before layer 5's epilogue, the next super-tile's PE(pts) before the last
        if (li == 5) compute_pe_block<4>(vd, pe);
        if (li == 9 && st + pair_stride < n_pair) {
          float np[3];
          row_setup(st + pair_stride, nvalid, ng, np, nvd);
          compute_pe_block<10>(np, pe);
        }
        mbar_wait(&sb.acc_full[tile], lc & 1u, 0x40 + tile);
        tc_fence_after();
        if (SAVE) exp_bulk_wait(lane);
        const bool tr = SAVE && q == 0 && lane == 0;
        if (tr) trace_event(0x2000u | (tile << 8) | (li << 4) | 1u, 1024, tn_e, 5000u + 1000u * tile, 6000u + 1000u * tile);
        if (li < 9) {
          uint8_t* grow = gblk(li < 8 ? arena::kActH + 4 * li : arena::kActFeat);
          uint8_t* grow1 = kTwoPhase ? nullptr : grow;                                        // first pass: no arena stores
          uint8_t* gmask1 = kTwoPhase ? nullptr : gmsk(arena::kMaskH + (li < 8 ? li : 0));
          uint32_t mbits[8];
          if (li == 7) epi_layer<1>(taddr, s_walpha, sigma, sAct, row_base, row_xor, grow1, gmask1, mbits);
          else if (li == 8) epi_layer<2>(taddr, s_walpha, sigma, sAct, row_base, row_xor, grow1, nullptr, mbits);
          else epi_layer<0>(taddr, s_walpha, sigma, sAct, row_base, row_xor, grow1, gmask1, mbits);
          // layer 5 has consumed PE(pts): the buffer now carries PE(dir) for the view branch
          if (li == 5) store_pe_block(sPe, row_base, row_xor, pe, kLate ? nullptr : gblk(arena::kActPEDir));
          if (tr) trace_event(0x2000u | (tile << 8) | (li << 4) | 2u, 1024, tn_e, 5000u + 1000u * tile, 6000u + 1000u * tile);
          signal_ready();
          if (tr) trace_event(0x2000u | (tile << 8) | (li << 4) | 3u, 1024, tn_e, 5000u + 1000u * tile, 6000u + 1000u * tile);
          if (SAVE) exp_bulk_rows_to_arena(sAct, q, lane, tile_base + (size_t)(li < 8 ? arena::kActH + 4 * li : arena::kActFeat) * arena::kBlockBytes, 4);
          if (kTwoPhase) {     // second pass over the same accumulator: arena copy + ReLU bit masks, then release it
            if (li == 8) epi_store_layer<2>(taddr, grow, nullptr);
            else epi_store_layer<0>(taddr, grow, gmsk(arena::kMaskH + li));
            signal_free();
            if (li == 5) store_pe_block(nullptr, row_base, row_xor, pe, gblk(arena::kActPEDir));
          }
          if (kDefer) {        // arena copy of this layer's output, its ReLU bit masks and (layer 5) PE(dir), off the critical path
            flush_row_to_arena(sAct, row_base, row_xor, grow, 4);
            if (li != 8) st_global_mask8(gmsk(arena::kMaskH + li), mbits);
            if (li == 5) store_pe_block(nullptr, row_base, row_xor, pe, gblk(arena::kActPEDir));
          }
        } else {
          float r0 = 0.f, r1 = 0.f, r2 = 0.f;
          epi_views<SAVE>(taddr, s_wrgb, r0, r1, r2, gblk(arena::kActHV), gmsk(arena::kMaskHV));
          tc_fence_before();
          if (valid) {
            float4 o = make_float4(r0 + __ldg(P + b_off(L_RGB)), r1 + __ldg(P + b_off(L_RGB) + 1), r2 + __ldg(P + b_off(L_RGB) + 2),
                                   sigma + __ldg(P + b_off(L_ALPHA)));
            if (src.mask && !src.mask[g]) o = make_float4(0.f, 0.f, 0.f, 0.f);
            raw[g] = o;
          }
        }
      }
      valid = nvalid; g = ng; vd[0] = nvd[0]; vd[1] = nvd[1]; vd[2] = nvd[2];
    }
  }
  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) tmem_dealloc2(tmem, 512);
}

}  // namespace v2

// act_arena != NULL: training forward, activations saved as tile images (needs arena::num_tiles(M) * act_tile_bytes()).
int mlp_tc2_forward(NerfEngine* h, int net, const RowSrc& src, float* raw, uint8_t* act_arena, cudaStream_t st) {
  using namespace v2;
  const int64_t M = src.R * src.S;
  if (M <= 0) return NERF_OK;
  static bool attr_set_dev[64] = {};          // the opt-in shared-memory size is a per-device function attribute
  bool& attr_set = attr_set_dev[h->device & 63];
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(mlp_tc2_forward_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(mlp_tc2_forward_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
    if (e != cudaSuccess) return h->fail(NERF_E_CUDA, std::string("mlp_tc2 smem attribute: ") + cudaGetErrorString(e));
    attr_set = true;
  }
  const int64_t n_pair = (M + kRowsPerPair - 1) / kRowsPerPair;
  // the fewest CTA pairs that keep the number of waves (1 024 training rays: 384 pair tiles = 6 waves on 74 pairs as on 64):
  // every pair then runs the same number of tiles and the idle SMs give their power budget to the busy ones
  const int64_t max_pairs = h->num_sms / 2, waves = (n_pair + max_pairs - 1) / max_pairs;
#ifdef NERF_FULL_GRID      // A/B builds: every pair the device has
  const int grid = 2 * (int)std::min<int64_t>(n_pair, max_pairs);
#else
  const int grid = 2 * (int)((n_pair + waves - 1) / waves);
#endif
  const float* P = h->params + (net == NERF_NET_FINE ? NERF_PARAMS_PER_NET : 0);
  // tensor maps over the packed stage images of both nets: [2 * 39 * 256 rows x 64 bf16] (rows of 128 bytes, already
  // swizzled by the packer, so the map copies them verbatim); boxes of 128 rows (N = 256 layers) and 64 rows (view layer)
  static_assert(sizeof(CUtensorMap) == sizeof(h->tmaps[0]), "CUtensorMap size");
  CUtensorMap& tm128 = *reinterpret_cast<CUtensorMap*>(h->tmaps[0]);
  CUtensorMap& tm64 = *reinterpret_cast<CUtensorMap*>(h->tmaps[1]);
  const int n_rep = (h->w_replicas > 1 && h->wrep) ? h->w_replicas : 1;
  const void* stage_base = n_rep > 1 ? (const void*)h->wrep : (const void*)h->wpack;
  if (h->tmap_base != stage_base) {
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                 const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess || !fn)
      return h->fail(NERF_E_CUDA, "mlp_tc2: cuTensorMapEncodeTiled not available");
    // rows of (128 << kTmaRowShift) bytes (see tma_stage_cg2); box = a CTA's half of a stage: 16 KB (8 KB for the view layer)
    const cuuint64_t gdim[2] = {64u << kTmaRowShift, ((cuuint64_t)n_rep * 2 * kNumStages * 256) >> kTmaRowShift}, gstride[1] = {128u << kTmaRowShift};
    const cuuint32_t estr[2] = {1, 1};
    for (int v = 0; v < 2; ++v) {
      const cuuint32_t box[2] = {64u << kTmaRowShift, (NERF_EXP & 2048) ? 8u : ((v == 0 ? 128u : 64u) >> kTmaRowShift)};
      CUresult r = reinterpret_cast<EncodeFn>(fn)(v == 0 ? &tm128 : &tm64, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(stage_base), gdim, gstride, box, estr,
                                                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) return h->fail(NERF_E_CUDA, "mlp_tc2: cuTensorMapEncodeTiled failed");
    }
    h->tmap_base = stage_base;
  }
  const int stage_row0 = (net == NERF_NET_FINE ? 1 : 0) * kNumStages * 256;
  const uint8_t* wb = reinterpret_cast<const uint8_t*>(h->wpack) + tc_wpack_bias_offset() + (size_t)(net == NERF_NET_FINE ? 1 : 0) * kNumSched * kBiasSlabBytes;
  if (act_arena) mlp_tc2_forward_kernel<true><<<grid, kThreads, kSmemBytes, st>>>(tm128, tm64, stage_row0, n_rep, src, reinterpret_cast<float4*>(raw), wb, P, M, n_pair, act_arena);
  else mlp_tc2_forward_kernel<false><<<grid, kThreads, kSmemBytes, st>>>(tm128, tm64, stage_row0, n_rep, src, reinterpret_cast<float4*>(raw), wb, P, M, n_pair, nullptr);
  h->launches++;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return h->fail(NERF_E_CUDA, std::string("mlp_tc2_forward launch: ") + cudaGetErrorString(e));
  return NERF_OK;
}

int tc2_debug_read(unsigned int out[8]) {
  unsigned int tmp[8];
  cudaError_t e = cudaMemcpyFromSymbol(tmp, g_tc_debug, sizeof(tmp));
  for (int i = 0; i < 8; ++i) out[i] = tmp[i];
  return e == cudaSuccess ? 0 : -1;
}

NERF_TC_DEFINE_DEBUG_SINK(tc2_set_debug_sink)

}  // namespace nerf
