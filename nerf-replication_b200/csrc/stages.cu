// stages.cu -- the non-MLP stages of Renderer.render as warp-per-ray CUDA kernels (sm_100a).
//   get_rays          src/datasets/nerf/blender.py:95-127            (a-B)
//   stratified_t      volume_renderer.py:27-65                        (a-C)
//   freq_encode       src/models/encoding/freq.py:7-32                (a-D, standalone tap)
//   composite         volume_renderer.py:67-96, 414-430               (a-G, a-J)
//   sample_pdf+merge  volume_renderer.py:98-272, 349-356              (a-H, a-I)
//   backward of composite / sample_pdf / MSE                          (a-K)
// All are HBM-bound streaming kernels: one warp owns one ray, lanes own samples i = lane + 32*j so that
// every global access is a fully coalesced 128/512-byte request; scans/reductions are warp shuffles.
#include "engine.h"

namespace nerf {

constexpr int kWarpsPerBlock = 4;
constexpr float kEps = 1e-5f;
// unroll factors of the per-ray loops over 32-sample blocks (S = 192 -> 6 blocks): the blocks' loads and double-precision exp
// chains are independent, only the scan carry is serial -- but unrolling costs occupancy in the render kernels and buys
// nothing in the 1024-ray training kernels (measured, see the defaults), so both stay at 1; the knobs remain for A/B builds.
#ifndef NERF_STAGE_UNROLL_RENDER
#define NERF_STAGE_UNROLL_RENDER 1      // measured on the 65 536-ray chunk: composite_kernel 88.6 us (1) / 137 us (3) / 120 us (6)
#endif
#ifndef NERF_STAGE_UNROLL_TRAIN
#define NERF_STAGE_UNROLL_TRAIN 1       // 1024-ray step: composite_loss_bwd 16.0 us (1) / 17.3 us (6): no gain from unrolling either
#endif
constexpr int kUnrollRender = NERF_STAGE_UNROLL_RENDER, kUnrollTrain = NERF_STAGE_UNROLL_TRAIN;
#ifndef NERF_PDF_MINBLOCKS
#define NERF_PDF_MINBLOCKS 8            // sample_pdf_kernel, 65 536 rays: 184.6 us (8 CTAs/SM, 64 registers) / 201 us (6) / 197 us (10)
#endif

static inline int grid_for(int64_t n_rays, int num_sms) {
  int64_t blocks = (n_rays + kWarpsPerBlock - 1) / kWarpsPerBlock;
  // cap at a multiple of the SM count (16 resident 128-thread CTAs per SM); kernels grid-stride over rays
  int64_t cap = (int64_t)num_sms * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

// ------------------------------------------------------------------------------------------------
// get_rays: float64 math like the NumPy original, rounded to fp32 at the end (collate, blender.py:163).
__global__ void get_rays_kernel(Camera cam, const int64_t* __restrict__ pixel_ids, int64_t pix0, int64_t n,
                                float* __restrict__ rays_o, float* __restrict__ rays_d) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t pid = pixel_ids ? pixel_ids[i] : pix0 + i;
    const double u = (double)(pid % cam.W), v = (double)(pid / cam.W);
    const double dx = (u - cam.cx) / cam.focal, dy = -(v - cam.cy) / cam.focal, dz = -1.0;
    double wx = cam.r64[0] * dx + cam.r64[1] * dy + cam.r64[2] * dz;
    double wy = cam.r64[3] * dx + cam.r64[4] * dy + cam.r64[5] * dz;
    double wz = cam.r64[6] * dx + cam.r64[7] * dy + cam.r64[8] * dz;
    const double nrm = sqrt(wx * wx + wy * wy + wz * wz);   // numpy divides by the norm (blender.py:126)
    rays_d[3 * i + 0] = (float)(wx / nrm);
    rays_d[3 * i + 1] = (float)(wy / nrm);
    rays_d[3 * i + 2] = (float)(wz / nrm);
    rays_o[3 * i + 0] = (float)cam.o64[0];
    rays_o[3 * i + 1] = (float)cam.o64[1];
    rays_o[3 * i + 2] = (float)cam.o64[2];
  }
}

void launch_get_rays(NerfEngine* h, const Camera& cam, const int64_t* pixel_ids, int64_t pix0, int64_t n,
                     float* rays_o, float* rays_d, cudaStream_t st) {
  if (n <= 0) return;
  int blocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)h->num_sms * 8);
  get_rays_kernel<<<blocks, 256, 0, st>>>(cam, pixel_ids, pix0, n, rays_o, rays_d);
  h->launches++;
}

// ------------------------------------------------------------------------------------------------
__global__ void stratified_t_kernel(int64_t n, int S, float t_near, float t_far, const float* __restrict__ jitter,
                                    int philox, uint64_t seed, uint64_t call_id, float* __restrict__ t) {
  const int64_t total = n * S;
  for (int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; g < total; g += (int64_t)gridDim.x * blockDim.x) {
    const int i = (int)(g % S);
    float ti = linspace_at(t_near, t_far, S, i);
    if (jitter || philox) {
      // lower/upper = segment bounds built from mid-points (volume_renderer.py:50-56)
      const float tp = i > 0 ? linspace_at(t_near, t_far, S, i - 1) : ti;
      const float tn = i < S - 1 ? linspace_at(t_near, t_far, S, i + 1) : ti;
      const float lower = i > 0 ? 0.5f * (ti + tp) : ti;
      const float upper = i < S - 1 ? 0.5f * (tn + ti) : ti;
      const float r = jitter ? jitter[g] : philox_uniform(seed, 0x51u + (uint32_t)call_id * 2u, (uint64_t)g);
      ti = __fadd_rn(lower, __fmul_rn(__fsub_rn(upper, lower), r));
    }
    t[g] = ti;
  }
}

void launch_stratified_t(NerfEngine* h, int64_t n, int S, const float* jitter, bool philox, uint64_t call_id,
                         float* t, cudaStream_t st) {
  if (n <= 0) return;
  int blocks = (int)std::min<int64_t>((n * S + 255) / 256, (int64_t)h->num_sms * 8);
  prefer_max_shared(reinterpret_cast<const void*>(stratified_t_kernel), h->device);
  stratified_t_kernel<<<blocks, 256, 0, st>>>(n, S, h->cfg.t_near, h->cfg.t_far, jitter, philox ? 1 : 0,
                                              h->cfg.seed, call_id, t);
  h->launches++;
}

// ------------------------------------------------------------------------------------------------
// Standalone positional encoding tap: out[m, 3 + 6L] = [x, sin(2^0 x), cos(2^0 x), ..., cos(2^(L-1) x)].
__global__ void freq_encode_kernel(const float* __restrict__ x, int64_t m, int L, float* __restrict__ out) {
  const int C = 3 + 6 * L;
  const int64_t total = m * C;
  for (int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; g < total; g += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = g / C;
    const int c = (int)(g % C);
    float v;
    if (c < 3) {
      v = x[row * 3 + c];
    } else {
      const int k = (c - 3) / 6, r = (c - 3) % 6;
      const float a = x[row * 3 + (r % 3)] * exp2f((float)k);   // exact: power-of-two scaling
      v = r < 3 ? sinf(a) : cosf(a);
    }
    out[g] = v;
  }
}

void launch_freq_encode(NerfEngine* h, const float* x, int64_t m, int n_freqs, float* out, cudaStream_t st) {
  if (m <= 0) return;
  int blocks = (int)std::min<int64_t>((m * (3 + 6 * n_freqs) + 255) / 256, (int64_t)h->num_sms * 8);
  freq_encode_kernel<<<blocks, 256, 0, st>>>(x, m, n_freqs, out);
  h->launches++;
}

// ------------------------------------------------------------------------------------------------
// Per-sample transmittance pieces, evaluated exactly like weights_computation (volume_renderer.py:80-94):
//   delta_i = t_{i+1}-t_i (last 1e10);  e = exp(-sigma*delta);  alpha = 1-e;  q = clamp(1-alpha, 1e-10, 1)
// exp rounded from a double-precision evaluation: within 0.5 ulp(fp32), which is what the CPU
// libm of the reference delivers in ~99% of calls (CUDA's expf is 2 ulp).  alpha = 1 - exp(-sigma delta)
// cancels catastrophically at sigma delta ~ 1e-3, so 1 ulp of exp is 1e-4 relative in the sampling pdf.
__device__ __forceinline__ float exp_cr(float x) { return (float)exp((double)x); }
// colours: sigmoid is well conditioned (no cancellation), MUFU.EX2 accuracy (|err| < 4e-7 for |x| < 20) is far inside the
// 1e-5 gate; three of the four double-precision exp calls per sample of the forward composite go away
__device__ __forceinline__ float sigmoid_fast(float x) { return __fdividef(1.0f, 1.0f + __expf(-x)); }
struct AlphaQ { float alpha, q, e, delta; };
__device__ __forceinline__ AlphaQ alpha_q(float sigma, float t_i, float t_next, bool last) {
  AlphaQ r;
  r.delta = last ? 1e10f : __fsub_rn(t_next, t_i);
  r.e = exp_cr(-__fmul_rn(sigma, r.delta));
  r.alpha = __fsub_rn(1.0f, r.e);
  r.q = fminf(fmaxf(__fsub_rn(1.0f, r.alpha), 1e-10f), 1.0f);
  return r;
}

// composite forward: one warp per ray.  T is an exclusive *product* scan of q over the samples (the same
// arithmetic as torch.cumprod up to association; equivalent to the log-space form
// T_i = exp(-sum_{j<i} min(sigma_j delta_j, ln 1e10)) without the extra exp/log round trip).
__global__ void composite_kernel(const float4* __restrict__ raw, const float* __restrict__ t, int64_t n, int S,
                                 int white, float* __restrict__ rgb, float* __restrict__ depth,
                                 float* __restrict__ acc, float* __restrict__ weights) {
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = blockIdx.x * (int64_t)kWarpsPerBlock + (threadIdx.x >> 5);
  const int nblk = (S + 31) / 32;
  for (int64_t ray = warp0; ray < n; ray += (int64_t)gridDim.x * kWarpsPerBlock) {
    const float4* rr = raw + ray * S;
    const float* tt = t + ray * S;
    float carry = 1.0f, sr = 0.f, sg = 0.f, sb = 0.f, sd = 0.f, sa = 0.f;
#pragma unroll kUnrollRender      // independent loads / exp chains of consecutive 32-sample blocks overlap; same operations, same order
    for (int j = 0; j < nblk; ++j) {
      const int i = j * 32 + lane;
      const bool valid = i < S;
      float4 v = valid ? __ldg(rr + i) : make_float4(0.f, 0.f, 0.f, 0.f);
      const float ti = valid ? __ldg(tt + i) : 0.f;
      const float tn = (i + 1 < S) ? __ldg(tt + i + 1) : 0.f;
      const float sigma = fmaxf(v.w, 0.f);
      AlphaQ a = alpha_q(sigma, ti, tn, i == S - 1);
      const float q = valid ? a.q : 1.0f;
      const float incl = warp_scan_mul(q, lane);
      float excl = __shfl_up_sync(0xffffffffu, incl, 1);
      if (lane == 0) excl = 1.0f;
      const float T = carry * excl;
      carry *= __shfl_sync(0xffffffffu, incl, 31);
      const float w = valid ? T * a.alpha : 0.f;
      if (weights && valid) weights[ray * S + i] = w;
      sr += w * sigmoid_fast(v.x);
      sg += w * sigmoid_fast(v.y);
      sb += w * sigmoid_fast(v.z);
      sd += w * ti;
      sa += w;
    }
    sr = warp_sum(sr); sg = warp_sum(sg); sb = warp_sum(sb); sd = warp_sum(sd); sa = warp_sum(sa);
    if (lane == 0) {
      if (white) { const float bg = 1.0f - sa; sr += bg; sg += bg; sb += bg; }
      if (rgb) { rgb[ray * 3 + 0] = sr; rgb[ray * 3 + 1] = sg; rgb[ray * 3 + 2] = sb; }
      if (depth) depth[ray] = sd;
      if (acc) acc[ray] = sa;
    }
  }
}

void launch_composite(NerfEngine* h, const float* raw, const float* t, int64_t n, int S, int white,
                      float* rgb, float* depth, float* acc, float* weights, cudaStream_t st) {
  if (n <= 0) return;
  prefer_max_shared(reinterpret_cast<const void*>(composite_kernel), h->device);
  composite_kernel<<<grid_for(n, h->num_sms), kWarpsPerBlock * 32, 0, st>>>(
      reinterpret_cast<const float4*>(raw), t, n, S, white, rgb, depth, acc, weights);
  h->launches++;
}

// ------------------------------------------------------------------------------------------------
// Hierarchical sampling.  Per-warp shared memory: t_c[Nc] | cdf[Nc] | keys[Npad] | kidx[Npad] | bin flags[Nc] | fine valid[Nf]
// (the last two only serve the ESS/ERT pruning mask of fast_sampling).
struct PdfSmem {
  float* tc; float* cdf; float* keys; int* kidx; int* flags; int* fvalid;
};
__host__ __device__ __forceinline__ size_t pdf_smem_words(int Nc, int Nf, int Npad) { return (size_t)3 * Nc + 2 * Npad + Nf; }
__device__ __forceinline__ PdfSmem pdf_smem(float* base, int warp, int Nc, int Nf, int Npad) {
  float* p = base + (size_t)warp * pdf_smem_words(Nc, Nf, Npad);
  PdfSmem s; s.tc = p; s.cdf = p + Nc; s.keys = p + 2 * Nc; s.kidx = reinterpret_cast<int*>(p + 2 * Nc + Npad);
  s.flags = s.kidx + Npad; s.fvalid = s.flags + Nc;
  return s;
}
static inline int next_pow2(int v) { int p = 32; while (p < v) p <<= 1; return p; }

// Builds t_c and cdf[0..Nc-2] in shared memory (volume_renderer.py:127-140); returns sum of (w+eps).
// `wtmp` (>= Nc floats, may alias keys) receives w'_j = w_{j+1}+eps for j in [0,Nc-2).
// prune != nullptr (a-M, volume_renderer.py:116-133,158-165): flags[k] bit 0 = coarse bin k is "empty" (w < w_thr),
// bit 1 = the transmittance fell below ert_thr at or before bin k (prefix OR = the reference's cummax); ray_stats
// receives (sum sigma, max sigma).
struct PruneArgs { int on; float w_thr, ert_thr; uint8_t* valid_fine; uint8_t* mask_sorted; int32_t* row_list; int32_t* row_count; };
__device__ __forceinline__ float build_cdf(const float4* __restrict__ raw_c, const float* __restrict__ t_c,
                                           const float* __restrict__ cdf_in, int Nc, int lane,
                                           float* tc, float* cdf, float* wtmp, const PruneArgs* prune = nullptr,
                                           int* flags = nullptr, float* ray_stats = nullptr) {
  const int nblk = (Nc + 31) / 32;
  float carry = 1.0f, ssum = 0.f, sig_sum = 0.f, sig_max = 0.f;
  bool ert_seen = false;
#pragma unroll 2
  for (int j = 0; j < nblk; ++j) {
    const int i = j * 32 + lane;
    const bool valid = i < Nc;
    const float ti = valid ? __ldg(t_c + i) : 0.f;
    const float tn = (i + 1 < Nc) ? __ldg(t_c + i + 1) : 0.f;
    const float sigma = valid ? fmaxf(__ldg(raw_c + i).w, 0.f) : 0.f;
    AlphaQ a = alpha_q(sigma, ti, tn, i == Nc - 1);
    const float q = valid ? a.q : 1.0f;
    const float incl = warp_scan_mul(q, lane);
    float excl = __shfl_up_sync(0xffffffffu, incl, 1);
    if (lane == 0) excl = 1.0f;
    const float T = carry * excl;
    carry *= __shfl_sync(0xffffffffu, incl, 31);
    if (valid) tc[i] = ti;
    const bool interior = i >= 1 && i <= Nc - 2;
    const float w = __fmul_rn(T, a.alpha);
    if (interior) {                          // weights[..., 1:-1] + eps
      const float wp = __fadd_rn(w, kEps);
      wtmp[i - 1] = wp;
      ssum += wp;
    }
    if (prune) {
      sig_sum += sigma; sig_max = fmaxf(sig_max, sigma);
      const unsigned below_thr = __ballot_sync(0xffffffffu, interior && T < prune->ert_thr);
      const bool ert = ert_seen || (below_thr & (0xffffffffu >> (31 - lane))) != 0u;
      ert_seen = ert_seen || below_thr != 0u;
      if (interior) flags[i - 1] = (w < prune->w_thr ? 1 : 0) | (ert ? 2 : 0);
    }
  }
  ssum = warp_sum(ssum);
  if (prune) {
    sig_sum = warp_sum(sig_sum);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sig_max = fmaxf(sig_max, __shfl_xor_sync(0xffffffffu, sig_max, o));
    ray_stats[0] = sig_sum; ray_stats[1] = sig_max;
  }
  __syncwarp();
  const int L = Nc - 1;                       // cdf length
  if (cdf_in) {
    for (int k = lane; k < L; k += 32) cdf[k] = __ldg(cdf_in + k);
  } else {
    float run = 0.f;
    const int pblk = (Nc - 2 + 31) / 32;
    if (lane == 0) cdf[0] = 0.f;
    for (int j = 0; j < pblk; ++j) {
      const int k = j * 32 + lane;
      const float pdf = (k < Nc - 2) ? __fdiv_rn(wtmp[k], ssum) : 0.f;
      const float incl = warp_scan_add(pdf, lane);
      if (k < Nc - 2) cdf[k + 1] = run + incl;
      run += __shfl_sync(0xffffffffu, incl, 31);
    }
  }
  __syncwarp();
  return ssum;
}

struct FineSample { float t, u, c0, c1, b0, b1; int inds, below, above; };
// lo = searchsorted(cdf, u, right=True): first index with cdf > u
__device__ __forceinline__ FineSample inverse_cdf_finish(const float* tc, const float* cdf, int Nc, float u, int lo) {
  FineSample f;
  f.inds = lo; f.u = u;
  f.below = min(max(lo - 1, 0), Nc - 3);      // the reference clamps at Nc-3 (volume_renderer.py:153-154)
  f.above = min(max(lo, 0), Nc - 3);
  f.c0 = cdf[f.below]; f.c1 = cdf[f.above];
  f.b0 = 0.5f * __fadd_rn(tc[f.below + 1], tc[f.below]);
  f.b1 = 0.5f * __fadd_rn(tc[f.above + 1], tc[f.above]);
  float denom = __fsub_rn(f.c1, f.c0);
  if (denom < kEps) denom = 1.0f;
  const float tt = __fdiv_rn(__fsub_rn(u, f.c0), denom);
  f.t = __fadd_rn(f.b0, __fmul_rn(tt, __fsub_rn(f.b1, f.b0)));
  return f;
}
__device__ __forceinline__ FineSample inverse_cdf(const float* tc, const float* cdf, int Nc, float u) {
  const int L = Nc - 1;
  int lo = 0, hi = L;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (cdf[mid] <= u) lo = mid + 1; else hi = mid;
  }
  return inverse_cdf_finish(tc, cdf, Nc, u, lo);
}

// Nc == 64: cdf has 63 = 2^6 - 1 entries, so the search of inverse_cdf takes exactly six probes whatever u is; four samples
// of a lane run their (identical) probe sequences interleaved instead of as four dependent shared-memory chains.
__device__ __forceinline__ void search4_cdf63(const float* cdf, const float (&uu)[4], int (&lo)[4]) {
  int hi[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) { lo[j] = 0; hi[j] = 63; }
#pragma unroll
  for (int it = 0; it < 6; ++it) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int mid = (lo[j] + hi[j]) >> 1;
      if (cdf[mid] <= uu[j]) lo[j] = mid + 1; else hi[j] = mid;
    }
  }
}

__device__ __forceinline__ float fine_u(const float* __restrict__ u_row, int philox, uint64_t seed, uint64_t call_id,
                                        int64_t ray, int Nf, int k) {
  if (u_row) return __ldg(u_row + k);
  if (philox) return philox_uniform(seed, 0x52u + (uint32_t)call_id * 2u, (uint64_t)(ray * Nf + k));
  return linspace_at(0.f, 1.f, Nf, k);
}

// warp-level bitonic sort of (key, idx) pairs in shared memory, ascending, ties by idx (== stable sort).
__device__ __forceinline__ void warp_bitonic_sort(float* keys, int* kidx, int Npad, int lane) {
  for (int k = 2; k <= Npad; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int e = lane; e < (Npad >> 1); e += 32) {
        const int i = 2 * e - (e & (j - 1));
        const int p = i + j;
        const bool up = (i & k) == 0;
        const float ka = keys[i], kb = keys[p];
        const int ia = kidx[i], ib = kidx[p];
        const bool a_gt_b = (ka > kb) || (ka == kb && ia > ib);
        if (a_gt_b == up) { keys[i] = kb; keys[p] = ka; kidx[i] = ib; kidx[p] = ia; }
      }
      __syncwarp();
    }
  }
}

// Merge of two sorted lists already in shared memory -- coarse depths tc[0..Nc) and fine depths keys[Nc..Nc+Nf) -- into
// keys/kidx[0..Nt) in exactly the order the stable sort above produces (coarse before fine on equal depths): every element
// finds its output slot by one binary search in the other list (merge path), ~45 probes per lane at 64+128 instead of the
// 36 compare-exchange rounds of the bitonic network.  Returns false (nothing written) when either list is not monotone
// (random u in training, or an inverse-CDF rounding that steps back by an ulp) or the per-lane slots do not fit.
constexpr int kMergeSlots = 16;   // elements per lane held in registers: Nc + Nf <= 512
__device__ __forceinline__ bool warp_merge_sorted(const float* tc, float* keys, int* kidx, int Nc, int Nf, int lane) {
  const float* tf = keys + Nc;
  const int per_c = (Nc + 31) >> 5, per_f = (Nf + 31) >> 5;
  if (per_c + per_f > kMergeSlots) return false;
  bool ok = true;
  for (int i = lane; i + 1 < Nc; i += 32) ok = ok && tc[i] <= tc[i + 1];
  for (int k = lane; k + 1 < Nf; k += 32) ok = ok && tf[k] <= tf[k + 1];
  if (!__all_sync(0xffffffffu, ok)) return false;
  if (Nc == 64 && Nf == 128) {
    // the default sampling: 2 coarse + 4 fine elements per lane.  The same probe sequences as the loops below, but with a fixed
    // trip count (8 probes cover 128 entries, a converged search idles), so that the six searches of a lane interleave
    // instead of running one dependent shared-memory chain after the other.
    float v[6];
    int lo[6], hi[6];
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      v[j] = j < 2 ? tc[lane + 32 * j] : tf[lane + 32 * (j - 2)];
      lo[j] = 0; hi[j] = j < 2 ? 128 : 64;
    }
#pragma unroll
    for (int it = 0; it < 8; ++it) {
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        if (lo[j] < hi[j]) {
          const int mid = (lo[j] + hi[j]) >> 1;
          const bool right = j < 2 ? (tf[mid] < v[j]) : (tc[mid] <= v[j]);   // coarse: fine depths strictly below; fine: coarse at or below
          if (right) lo[j] = mid + 1; else hi[j] = mid;
        }
      }
    }
    __syncwarp();                             // every probe of keys[Nc..] is done before the slots are overwritten
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      const int src = j < 2 ? lane + 32 * j : Nc + lane + 32 * (j - 2);
      const int dst = (j < 2 ? lane + 32 * j : lane + 32 * (j - 2)) + lo[j];
      keys[dst] = v[j];
      kidx[dst] = src;
    }
    __syncwarp();
    return true;
  }
  float val[kMergeSlots];
  int pos[kMergeSlots];
#pragma unroll
  for (int j = 0; j < kMergeSlots; ++j) {
    pos[j] = -1;
    if (j < per_c) {
      const int i = lane + 32 * j;
      if (i < Nc) {
        const float v = tc[i];
        int lo = 0, hi = Nf;                  // number of fine depths strictly below v
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (tf[mid] < v) lo = mid + 1; else hi = mid; }
        val[j] = v; pos[j] = i + lo;
      }
    } else if (j < per_c + per_f) {
      const int k = lane + 32 * (j - per_c);
      if (k < Nf) {
        const float v = tf[k];
        int lo = 0, hi = Nc;                  // number of coarse depths at or below v
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (tc[mid] <= v) lo = mid + 1; else hi = mid; }
        val[j] = v; pos[j] = k + lo;
      }
    }
  }
  __syncwarp();                               // every probe of keys[Nc..] is done before the slots are overwritten
#pragma unroll
  for (int j = 0; j < kMergeSlots; ++j) {
    if (pos[j] >= 0) {
      keys[pos[j]] = val[j];
      kidx[pos[j]] = j < per_c ? lane + 32 * j : Nc + lane + 32 * (j - per_c);
    }
  }
  __syncwarp();
  return true;
}

__global__ void __launch_bounds__(kWarpsPerBlock * 32, NERF_PDF_MINBLOCKS) sample_pdf_kernel(const float4* __restrict__ raw_c, const float* __restrict__ t_c, int64_t n, int Nc,
                                  int Nf, int Npad, const float* __restrict__ u, int philox, uint64_t seed,
                                  uint64_t call_id, const float* __restrict__ cdf_in, float* __restrict__ t_fine,
                                  int64_t* __restrict__ inds, int64_t* __restrict__ below,
                                  int64_t* __restrict__ above, float* __restrict__ cdf_out,
                                  float* __restrict__ t_sorted, int32_t* __restrict__ sort_idx, PruneArgs prune) {
  extern __shared__ float smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  PdfSmem s = pdf_smem(smem, warp, Nc, Nf, Npad);
  const int Nt = Nc + Nf;
  for (int64_t ray = blockIdx.x * (int64_t)kWarpsPerBlock + warp; ray < n; ray += (int64_t)gridDim.x * kWarpsPerBlock) {
    float stats[2] = {0.f, 0.f};
    build_cdf(raw_c + ray * Nc, t_c + ray * Nc, cdf_in ? cdf_in + ray * (Nc - 1) : nullptr, Nc, lane, s.tc, s.cdf, s.keys,
              prune.on ? &prune : nullptr, s.flags, stats);
    const bool empty_ray = stats[0] < 1e-3f, object_ray = stats[1] > 0.5f;      // volume_renderer.py:116-122
    if (cdf_out) for (int k = lane; k < Nc - 1; k += 32) cdf_out[ray * (Nc - 1) + k] = s.cdf[k];
    __syncwarp();
    for (int k = lane; k < Nc; k += 32) { s.keys[k] = s.tc[k]; s.kidx[k] = k; }
    auto emit = [&](int k, const FineSample& f) {
      s.keys[Nc + k] = f.t; s.kidx[Nc + k] = Nc + k;
      const int64_t o = ray * Nf + k;
      if (t_fine) t_fine[o] = f.t;
      if (inds) inds[o] = f.inds;
      if (below) below[o] = f.below;
      if (above) above[o] = f.above;
      if (prune.on) {      // ESS (:170-186): object rays need both neighbour bins empty, background rays either; ERT (:158-168)
        const int fb = s.flags[f.below], fa = s.flags[f.above];
        const bool ess = object_ray ? ((fb & fa) & 1) != 0 : ((fb | fa) & 1) != 0;
        const bool valid = !(ess || (fb & 2) != 0) && !empty_ray;
        s.fvalid[k] = valid ? 1 : 0;
        if (prune.valid_fine) prune.valid_fine[o] = valid ? 1 : 0;
      }
    };
    const float* u_row = u ? u + ray * Nf : nullptr;
    if (Nc == 64) {
      for (int k0 = lane; k0 < Nf; k0 += 128) {
        float uu[4];
        int lo[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int k = k0 + 32 * j;
          uu[j] = k < Nf ? fine_u(u_row, philox, seed, call_id, ray, Nf, k) : 0.f;
        }
        search4_cdf63(s.cdf, uu, lo);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int k = k0 + 32 * j;
          if (k < Nf) emit(k, inverse_cdf_finish(s.tc, s.cdf, Nc, uu[j], lo[j]));
        }
      }
    } else {
      for (int k = lane; k < Nf; k += 32) emit(k, inverse_cdf(s.tc, s.cdf, Nc, fine_u(u_row, philox, seed, call_id, ray, Nf, k)));
    }
    for (int k = Nt + lane; k < Npad; k += 32) { s.keys[k] = __int_as_float(0x7f800000); s.kidx[k] = 0x7fffffff; }
    __syncwarp();
    if (t_sorted || sort_idx || (prune.on && (prune.mask_sorted || prune.row_list))) {
      if (!warp_merge_sorted(s.tc, s.keys, s.kidx, Nc, Nf, lane)) warp_bitonic_sort(s.keys, s.kidx, Npad, lane);
      for (int k = lane; k < Nt; k += 32) {
        if (t_sorted) t_sorted[ray * Nt + k] = s.keys[k];
        if (sort_idx) sort_idx[ray * Nt + k] = s.kidx[k];
        // coarse samples stay valid; fine samples carry their flag through the same permutation (:359-370)
        if (prune.on && prune.mask_sorted) prune.mask_sorted[ray * Nt + k] = s.kidx[k] < Nc ? 1 : (uint8_t)s.fvalid[s.kidx[k] - Nc];
      }
      if (prune.on && prune.row_list) {
        // work skipping: append this ray's valid rows (sorted order) to the compacted row list the fine MLP pass walks;
        // one global atomic per ray reserves the segment
        int total = 0;
        for (int k0 = 0; k0 < Nt; k0 += 32) {
          const int k = k0 + lane;
          const bool v = k < Nt && (s.kidx[k] < Nc || s.fvalid[s.kidx[k] - Nc] != 0);
          total += __popc(__ballot_sync(0xffffffffu, v));
        }
        int base = 0;
        if (lane == 0) base = atomicAdd(prune.row_count, total);
        base = __shfl_sync(0xffffffffu, base, 0);
        for (int k0 = 0; k0 < Nt; k0 += 32) {
          const int k = k0 + lane;
          const bool v = k < Nt && (s.kidx[k] < Nc || s.fvalid[s.kidx[k] - Nc] != 0);
          const unsigned b = __ballot_sync(0xffffffffu, v);
          if (v) prune.row_list[base + __popc(b & ((1u << lane) - 1u))] = (int32_t)(ray * Nt + k);
          base += __popc(b);
        }
      }
    }
    __syncwarp();
  }
}

int launch_sample_pdf(NerfEngine* h, const float* raw_c, const float* t_c, int64_t n, int Nc, int Nf,
                      const float* u, bool philox_u, uint64_t call_id, const float* cdf_in, float* t_fine,
                      int64_t* inds, int64_t* below, int64_t* above, float* cdf_out, float* t_sorted,
                      int32_t* sort_idx, cudaStream_t st, float prune_w_thr, uint8_t* valid_fine, uint8_t* mask_sorted,
                      int32_t* row_list, int32_t* row_count) {
  if (Nc < 4 || Nf < 1 || Nc + Nf > 4096) return h->fail(NERF_E_INVALID, "sample_pdf: need Nc>=4, Nf>=1, Nc+Nf<=4096");
  if (n <= 0) return NERF_OK;
  const int Npad = next_pow2(Nc + Nf);
  PruneArgs prune;
  prune.on = (valid_fine || mask_sorted) ? 1 : 0; prune.w_thr = prune_w_thr; prune.ert_thr = 0.45f;   // ert_threshold default (:98)
  prune.valid_fine = valid_fine; prune.mask_sorted = mask_sorted; prune.row_list = row_list; prune.row_count = row_count;
  if (row_list && (int64_t)n * (Nc + Nf) >= (int64_t)1 << 31) return h->fail(NERF_E_INVALID, "sample_pdf: row list needs n*(Nc+Nf) < 2^31");
  const size_t smem = (size_t)kWarpsPerBlock * pdf_smem_words(Nc, Nf, Npad) * sizeof(float);
  if (smem > 48 * 1024) cudaFuncSetAttribute(sample_pdf_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  prefer_max_shared(reinterpret_cast<const void*>(sample_pdf_kernel), h->device);
  sample_pdf_kernel<<<grid_for(n, h->num_sms), kWarpsPerBlock * 32, smem, st>>>(
      reinterpret_cast<const float4*>(raw_c), t_c, n, Nc, Nf, Npad, u, philox_u ? 1 : 0, h->cfg.seed, call_id, cdf_in,
      t_fine, inds, below, above, cdf_out, t_sorted, sort_idx, prune);
  h->launches++;
  return NERF_OK;
}

// ------------------------------------------------------------------------------------------------
// MSE loss (trainers/nerf.py:17,33): loss = mean over n*3; d_rgb = 2 (rgb-gt) / (3n) * scale.
// One block: deterministic summation order.
__global__ void mse_grad_kernel(const float* __restrict__ rgb, const float* __restrict__ gt, int64_t n3, float scale,
                                float* __restrict__ d_rgb, float* __restrict__ loss) {
  __shared__ float red[32];
  float acc = 0.f;
  const float gscale = 2.0f * scale / (float)n3;
  for (int64_t i = threadIdx.x; i < n3; i += blockDim.x) {
    const float d = rgb[i] - gt[i];
    acc += d * d;
    if (d_rgb) d_rgb[i] = d * gscale;
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0 && loss) *loss = v / (float)n3;
  }
}

void launch_mse_grad(NerfEngine* h, const float* rgb, const float* gt, int64_t n, float scale, float* d_rgb,
                     float* loss, cudaStream_t st) {
  mse_grad_kernel<<<1, 1024, 0, st>>>(rgb, gt, n * 3, scale, d_rgb, loss);
  h->launches++;
}

// ------------------------------------------------------------------------------------------------
// composite backward (autograd of volume_renderer.py:67-96,414-430), one warp per ray.
// Per-warp shared memory: T[S] | alpha[S] | ddelta[S].
// composite_fwd_pass: T_i, alpha_i into shared memory (forward order); with `maps` it also folds the forward reductions
// rgb / depth / acc in exactly the order composite_kernel uses (so a fused caller produces the same map bits).
struct RayMaps { float r, g, b, depth, acc; };
__device__ __forceinline__ RayMaps composite_fwd_pass(const float4* __restrict__ rr, const float* __restrict__ tt, int S, int lane,
                                                      float* sT, float* sA, bool maps) {
  const int nblk = (S + 31) / 32;
  float carry = 1.0f, sr = 0.f, sg = 0.f, sb = 0.f, sd = 0.f, sa = 0.f;
#pragma unroll kUnrollTrain
  for (int j = 0; j < nblk; ++j) {
    const int i = j * 32 + lane;
    const bool valid = i < S;
    const float4 v = valid ? __ldg(rr + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    const float sigma = fmaxf(v.w, 0.f);
    const float ti = valid ? __ldg(tt + i) : 0.f;
    const float tn = (i + 1 < S) ? __ldg(tt + i + 1) : 0.f;
    AlphaQ a = alpha_q(sigma, ti, tn, i == S - 1);
    const float q = valid ? a.q : 1.0f;
    const float incl = warp_scan_mul(q, lane);
    float excl = __shfl_up_sync(0xffffffffu, incl, 1);
    if (lane == 0) excl = 1.0f;
    const float T = carry * excl;
    if (valid) { sT[i] = T; sA[i] = a.alpha; }
    carry *= __shfl_sync(0xffffffffu, incl, 31);
    if (maps) {
      const float w = valid ? T * a.alpha : 0.f;
      sr += w * sigmoid_fast(v.x); sg += w * sigmoid_fast(v.y); sb += w * sigmoid_fast(v.z);
      sd += w * ti; sa += w;
    }
  }
  RayMaps m{0.f, 0.f, 0.f, 0.f, 0.f};
  if (maps) { m.r = warp_sum(sr); m.g = warp_sum(sg); m.b = warp_sum(sb); m.depth = warp_sum(sd); m.acc = warp_sum(sa); }
  __syncwarp();
  return m;
}

// Reverse pass of one ray: (d rgb_map = g0..g2, d depth_map = gd, direct d w from d_w_in -- a shared- or global-memory row
// written earlier by this warp or by an earlier kernel; plain loads) -> d_raw row, d_t row (may be NULL).
__device__ __forceinline__ void composite_bwd_ray(const float4* __restrict__ rr, const float* __restrict__ tt, int S, int white, int lane,
                                                  float g0, float g1, float g2, float gd, const float* d_w_in,
                                                  float4* __restrict__ d_raw_row, float* __restrict__ d_t_row,
                                                  const float* sT, const float* sA, float* sD) {
  const int nblk = (S + 31) / 32;
  const float wb = white ? 1.0f : 0.0f;
  // no colour gradient arrives at the coarse pass of the hierarchical step (only d w through the inverse CDF): skip the
  // three sigmoids per sample there (per-ray condition, uniform across the warp)
  const bool need_rgb = (g0 != 0.f) || (g1 != 0.f) || (g2 != 0.f);
  // suffix sums S_k = sum_{j>k} dw_j w_j
  float suffix_carry = 0.f;
#pragma unroll kUnrollTrain
  for (int j = nblk - 1; j >= 0; --j) {
    const int i = j * 32 + lane;
    const bool valid = i < S;
    float4 v = valid ? __ldg(rr + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    const float T = valid ? sT[i] : 0.f, alpha = valid ? sA[i] : 0.f;
    const float w = T * alpha;
    // the same sigmoid the forward composite evaluates (MUFU.EX2, |err| < 4e-7): the backward differentiates what was rendered
    const float c0 = need_rgb ? sigmoid_fast(v.x) : 0.f, c1 = need_rgb ? sigmoid_fast(v.y) : 0.f, c2 = need_rgb ? sigmoid_fast(v.z) : 0.f;
    float dw = g0 * (c0 - wb) + g1 * (c1 - wb) + g2 * (c2 - wb) + gd * (valid ? __ldg(tt + i) : 0.f);
    if (d_w_in && valid) dw += d_w_in[i];
    // inclusive suffix scan of dw*w over the lanes of this block
    float sfx = valid ? dw * w : 0.f;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const float nv = __shfl_down_sync(0xffffffffu, sfx, o);
      if (lane + o < 32) sfx += nv;
    }
    const float own = valid ? dw * w : 0.f;
    const float Sk = suffix_carry + (sfx - own);            // strictly-after sum
    suffix_carry += __shfl_sync(0xffffffffu, sfx, 0);
    if (valid) {
      const float one_m_a = __fsub_rn(1.0f, alpha);
      const float q = fminf(fmaxf(one_m_a, 1e-10f), 1.0f);
      const float pass = (one_m_a >= 1e-10f && one_m_a <= 1.0f) ? 1.0f : 0.0f;   // clamp backward (inclusive)
      const float dalpha = T * dw - pass * Sk / q;
      const float sigma = fmaxf(v.w, 0.f);
      const float ti = __ldg(tt + i);
      const float delta = (i == S - 1) ? 1e10f : __fsub_rn(__ldg(tt + i + 1), ti);
      const float e = exp_cr(-__fmul_rn(sigma, delta));      // d alpha / d x = exp(-x)
      const float dx = dalpha * e;
      const float dsigma = dx * delta;
      sD[i] = (i == S - 1) ? 0.f : dx * sigma;               // d delta (last delta is a constant)
      float4 o;
      o.x = g0 * w * c0 * (1.0f - c0);
      o.y = g1 * w * c1 * (1.0f - c1);
      o.z = g2 * w * c2 * (1.0f - c2);
      o.w = v.w > 0.f ? dsigma : 0.f;                        // relu backward
      d_raw_row[i] = o;
    }
  }
  __syncwarp();
  if (d_t_row) {
    for (int i = lane; i < S; i += 32) {
      const float prev = i > 0 ? sD[i - 1] : 0.f;
      // delta_{i-1} = t_i - t_{i-1}; delta_i = t_{i+1} - t_i; depth_map = sum w t adds gd * w_i
      d_t_row[i] = prev - sD[i] + gd * sT[i] * sA[i];
    }
  }
  __syncwarp();
}

__global__ void composite_bwd_kernel(const float4* __restrict__ raw, const float* __restrict__ t, int64_t n, int S,
                                     int white, const float* __restrict__ d_rgb_map, const float* __restrict__ d_depth_map,
                                     const float* __restrict__ d_w_in, float4* __restrict__ d_raw, float* __restrict__ d_t) {
  extern __shared__ float smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* sT = smem + (size_t)warp * 3 * S;
  float* sA = sT + S;
  float* sD = sA + S;
  for (int64_t ray = blockIdx.x * (int64_t)kWarpsPerBlock + warp; ray < n; ray += (int64_t)gridDim.x * kWarpsPerBlock) {
    const float4* rr = raw + ray * S;
    const float* tt = t + ray * S;
    composite_fwd_pass(rr, tt, S, lane, sT, sA, false);
    float g0 = 0.f, g1 = 0.f, g2 = 0.f;
    if (d_rgb_map) { g0 = __ldg(d_rgb_map + ray * 3); g1 = __ldg(d_rgb_map + ray * 3 + 1); g2 = __ldg(d_rgb_map + ray * 3 + 2); }
    const float gd = d_depth_map ? __ldg(d_depth_map + ray) : 0.f;
    composite_bwd_ray(rr, tt, S, white, lane, g0, g1, g2, gd, d_w_in ? d_w_in + ray * S : nullptr, d_raw + ray * S,
                      d_t ? d_t + ray * S : nullptr, sT, sA, sD);
  }
}

void launch_composite_bwd(NerfEngine* h, const float* raw, const float* t, int64_t n, int S, int white,
                          const float* d_rgb_map, const float* d_depth_map, const float* d_w, float* d_raw, float* d_t,
                          cudaStream_t st) {
  if (n <= 0) return;
  const size_t smem = (size_t)kWarpsPerBlock * 3 * S * sizeof(float);
  if (smem > 48 * 1024) cudaFuncSetAttribute(composite_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  prefer_max_shared(reinterpret_cast<const void*>(composite_bwd_kernel), h->device);
  composite_bwd_kernel<<<grid_for(n, h->num_sms), kWarpsPerBlock * 32, smem, st>>>(
      reinterpret_cast<const float4*>(raw), t, n, S, white, d_rgb_map, d_depth_map, d_w, reinterpret_cast<float4*>(d_raw), d_t);
  h->launches++;
}

// Fused tail of the training forward + head of the backward (one launch instead of composite + mse_grad + composite_bwd):
// per ray  rgb_map = composite(raw, t)  (volume_renderer.py:414-430; same arithmetic and reduction order as composite_kernel),
// err = rgb_map - gt, d rgb_map = 2 err / (3 n) * scale (MSELoss, trainers/nerf.py:33), then the composite backward.
// Loss: every warp leaves its ray's squared error in sq[ray]; the last block to finish folds sq[0..n) in a fixed order
// (deterministic, like the one-block mse_grad_kernel) and resets the counter.
__global__ void composite_loss_bwd_kernel(const float4* __restrict__ raw, const float* __restrict__ t, int64_t n, int S, int white,
                                          const float* __restrict__ gt, float gscale, float* __restrict__ rgb_out,
                                          float4* __restrict__ d_raw, float* __restrict__ d_t, float* __restrict__ sq,
                                          unsigned int* __restrict__ done, float* __restrict__ loss) {
  extern __shared__ float smem[];
  __shared__ float red[kWarpsPerBlock];
  __shared__ bool is_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* sT = smem + (size_t)warp * 3 * S;
  float* sA = sT + S;
  float* sD = sA + S;
  for (int64_t ray = blockIdx.x * (int64_t)kWarpsPerBlock + warp; ray < n; ray += (int64_t)gridDim.x * kWarpsPerBlock) {
    const float4* rr = raw + ray * S;
    const float* tt = t + ray * S;
    RayMaps m = composite_fwd_pass(rr, tt, S, lane, sT, sA, true);
    if (white) { const float bg = 1.0f - m.acc; m.r += bg; m.g += bg; m.b += bg; }
    const float e0 = m.r - __ldg(gt + ray * 3), e1 = m.g - __ldg(gt + ray * 3 + 1), e2 = m.b - __ldg(gt + ray * 3 + 2);
    if (lane == 0) {
      if (rgb_out) { rgb_out[ray * 3] = m.r; rgb_out[ray * 3 + 1] = m.g; rgb_out[ray * 3 + 2] = m.b; }
      sq[ray] = (e0 * e0 + e1 * e1) + e2 * e2;
    }
    composite_bwd_ray(rr, tt, S, white, lane, e0 * gscale, e1 * gscale, e2 * gscale, 0.f, nullptr, d_raw + ray * S,
                      d_t ? d_t + ray * S : nullptr, sT, sA, sD);
  }
  // last block: deterministic fold of the per-ray squared errors
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(done, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  float acc = 0.f;
  for (int64_t i = threadIdx.x; i < n; i += blockDim.x) acc += __ldcg(sq + i);
  acc = warp_sum(acc);
  if (lane == 0) red[warp] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float v = 0.f;
    for (int w = 0; w < kWarpsPerBlock; ++w) v += red[w];
    *loss = v / (3.0f * (float)n);
    *done = 0u;
  }
}

void launch_composite_loss_bwd(NerfEngine* h, const float* raw, const float* t, int64_t n, int S, int white, const float* gt,
                               float scale, float* rgb_out, float* d_raw, float* d_t, float* sq, unsigned int* done, float* loss,
                               cudaStream_t st) {
  if (n <= 0) return;
  const size_t smem = (size_t)kWarpsPerBlock * 3 * S * sizeof(float);
  if (smem > 48 * 1024) cudaFuncSetAttribute(composite_loss_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const float gscale = 2.0f * scale / (3.0f * (float)n);
  prefer_max_shared(reinterpret_cast<const void*>(composite_loss_bwd_kernel), h->device);
  composite_loss_bwd_kernel<<<grid_for(n, h->num_sms), kWarpsPerBlock * 32, smem, st>>>(
      reinterpret_cast<const float4*>(raw), t, n, S, white, gt, gscale, rgb_out, reinterpret_cast<float4*>(d_raw), d_t, sq, done, loss);
  h->launches++;
}

// ------------------------------------------------------------------------------------------------
// sample_pdf backward: d t_sorted -> (un-sort) d t_fine -> d cdf (scatter) -> d pdf (suffix sum) ->
// d w_coarse.  Forward quantities are recomputed.  Per-warp smem: forward layout + dcdf[Nc] + dtf[Nf].
// d_raw_c != NULL (fused training step): the coarse composite backward runs in the same warp on the d w row it has just
// produced (kept in shared memory; + 3 Nc floats of T / alpha / d delta) and writes d_raw_c [n,Nc,4] -- d_w_coarse may
// then be NULL.  Only d sigma is non-zero there: the coarse maps are not in the loss.
__global__ void sample_pdf_bwd_kernel(const float4* __restrict__ raw_c, const float* __restrict__ t_c, int64_t n, int Nc,
                                      int Nf, int Npad, const float* __restrict__ u, int philox, uint64_t seed,
                                      uint64_t call_id, const float* __restrict__ d_t_sorted,
                                      const int32_t* __restrict__ sort_idx, float* __restrict__ d_w_coarse,
                                      int white, float4* __restrict__ d_raw_c) {
  extern __shared__ float smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int per_warp = 2 * Nc + 2 * Npad + Nc + Nf + (d_raw_c ? 3 * Nc : 0);
  float* base = smem + (size_t)warp * per_warp;
  PdfSmem s; s.tc = base; s.cdf = base + Nc; s.keys = base + 2 * Nc; s.kidx = reinterpret_cast<int*>(base + 2 * Nc + Npad);
  float* dcdf = base + 2 * Nc + 2 * Npad;
  float* dtf = dcdf + Nc;
  const int Nt = Nc + Nf;
  for (int64_t ray = blockIdx.x * (int64_t)kWarpsPerBlock + warp; ray < n; ray += (int64_t)gridDim.x * kWarpsPerBlock) {
    const float ssum = build_cdf(raw_c + ray * Nc, t_c + ray * Nc, nullptr, Nc, lane, s.tc, s.cdf, s.keys);
    // s.keys[0..Nc-3] now holds w'_j = w_{j+1} + eps
    for (int k = lane; k < Nc; k += 32) dcdf[k] = 0.f;
    for (int k = lane; k < Nf; k += 32) dtf[k] = 0.f;
    __syncwarp();
    for (int p = lane; p < Nt; p += 32) {
      const int idx = __ldg(sort_idx + ray * Nt + p);
      if (idx >= Nc) dtf[idx - Nc] = __ldg(d_t_sorted + ray * Nt + p);
    }
    __syncwarp();
    auto scatter = [&](int k, const FineSample& f) {
      const float g = dtf[k];
      const float db = __fsub_rn(f.b1, f.b0);
      const float den_raw = __fsub_rn(f.c1, f.c0);
      const float d_tt = g * db;
      float d_c0, d_c1;
      if (den_raw < kEps) { d_c0 = -d_tt; d_c1 = 0.f; }        // where(denom<eps, 1, denom): no grad via denom
      else {
        const float tt = __fdiv_rn(__fsub_rn(f.u, f.c0), den_raw);
        const float d_num = d_tt / den_raw, d_den = -d_tt * tt / den_raw;
        d_c0 = -d_num - d_den; d_c1 = d_den;
      }
      atomicAdd(&dcdf[f.below], d_c0);
      atomicAdd(&dcdf[f.above], d_c1);
    };
    const float* u_row = u ? u + ray * Nf : nullptr;
    if (Nc == 64) {
      for (int k0 = lane; k0 < Nf; k0 += 128) {
        float uu[4];
        int lo[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int k = k0 + 32 * j;
          uu[j] = k < Nf ? fine_u(u_row, philox, seed, call_id, ray, Nf, k) : 0.f;
        }
        search4_cdf63(s.cdf, uu, lo);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int k = k0 + 32 * j;
          if (k < Nf) scatter(k, inverse_cdf_finish(s.tc, s.cdf, Nc, uu[j], lo[j]));
        }
      }
    } else {
      for (int k = lane; k < Nf; k += 32) scatter(k, inverse_cdf(s.tc, s.cdf, Nc, fine_u(u_row, philox, seed, call_id, ray, Nf, k)));
    }
    __syncwarp();
    // dpdf_j = sum_{k>j} dcdf_k, j in [0, Nc-2);  dot = sum_j dpdf_j pdf_j
    const int P = Nc - 2;
    const int pblk = (P + 31) / 32;
    float carry = 0.f, dot = 0.f;
    for (int j = pblk - 1; j >= 0; --j) {
      const int k = j * 32 + lane;
      // element k of the suffix scan is dcdf[k+1]
      float v = (k < P) ? dcdf[k + 1] : 0.f;
      float sfx = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const float nv = __shfl_down_sync(0xffffffffu, sfx, o);
        if (lane + o < 32) sfx += nv;
      }
      const float dp = carry + sfx;                            // inclusive over indices >= k+1 of dcdf
      carry += __shfl_sync(0xffffffffu, sfx, 0);
      __syncwarp();
      if (k < P) { dcdf[k + 1] = dp; dot += dp * __fdiv_rn(s.keys[k], ssum); }
    }
    dot = warp_sum(dot);
    __syncwarp();
    for (int i = lane; i < Nc; i += 32) {
      float o = 0.f;
      if (i >= 1 && i <= Nc - 2) o = (dcdf[i] - dot) / ssum;   // dcdf[i] holds dpdf_{i-1}
      if (d_w_coarse) d_w_coarse[ray * Nc + i] = o;
      if (d_raw_c) dcdf[i] = o;                                // same index read above: in place
    }
    __syncwarp();
    if (d_raw_c) {
      float* sT = dtf + Nf;
      composite_fwd_pass(raw_c + ray * Nc, t_c + ray * Nc, Nc, lane, sT, sT + Nc, false);
      composite_bwd_ray(raw_c + ray * Nc, t_c + ray * Nc, Nc, white, lane, 0.f, 0.f, 0.f, 0.f, dcdf, d_raw_c + ray * Nc, nullptr,
                        sT, sT + Nc, sT + 2 * Nc);
    }
  }
}

int launch_sample_pdf_bwd(NerfEngine* h, const float* raw_c, const float* t_c, int64_t n, int Nc, int Nf,
                          const float* u, bool philox_u, uint64_t call_id, const float* d_t_sorted,
                          const int32_t* sort_idx, float* d_w_coarse, cudaStream_t st, int white, float* d_raw_c) {
  if (n <= 0) return NERF_OK;
  const int Npad = next_pow2(Nc + Nf);
  const size_t smem = (size_t)kWarpsPerBlock * (2 * Nc + 2 * Npad + Nc + Nf + (d_raw_c ? 3 * Nc : 0)) * sizeof(float);
  if (smem > 48 * 1024) cudaFuncSetAttribute(sample_pdf_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  prefer_max_shared(reinterpret_cast<const void*>(sample_pdf_bwd_kernel), h->device);
  sample_pdf_bwd_kernel<<<grid_for(n, h->num_sms), kWarpsPerBlock * 32, smem, st>>>(
      reinterpret_cast<const float4*>(raw_c), t_c, n, Nc, Nf, Npad, u, philox_u ? 1 : 0, h->cfg.seed, call_id,
      d_t_sorted, sort_idx, d_w_coarse, white, reinterpret_cast<float4*>(d_raw_c));
  h->launches++;
  return NERF_OK;
}

}  // namespace nerf
