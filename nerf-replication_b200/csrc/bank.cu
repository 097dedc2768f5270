// bank.cu -- the steps either side of the hot path, kept on the device:
//   ray bank     Dataset.__init__/__getitem__, src/datasets/nerf/blender.py:8-170: every view's image and pose stay
//                resident in HBM; one kernel turns pixel ids into (rays_o, rays_d, colors) exactly as __getitem__ does
//                (float64 camera math rounded to fp32, colors = image / 255), so a training step needs no DataLoader
//                worker and no host->device copy.  Pixel ids are the caller's (np.random.choice in the reference,
//                blender.py:108-110), a row-major range (the val/test branch, :112-115), or drawn here without
//                replacement by a keyed bijection of [0, H*W) (a cycle-walking Feistel network: O(1) per ray, no
//                permutation array).
//   evaluator    Evaluator.psnr_metric / evaluate, src/evaluators/nerf.py:23-30, 80-103: clipped fp32 MSE and the
//                uint8-quantised PSNR sum *with numpy's uint8 wrap-around* ((a-b) mod 256, squared mod 256), as
//                exact integers.
// Both are HBM-bound gather / streaming kernels.
#include <mutex>
#include <vector>
#include "engine.h"

namespace nerf {

// ---- keyed bijection of [0, n) -----------------------------------------------------------------------
// Balanced Feistel network over 2*hb bits (2^(2 hb) >= n), 6 rounds, cycle-walked back into [0, n).
struct PermKey { uint32_t k[6]; int hb; };

__host__ __device__ inline uint32_t perm_mix(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}
__host__ __device__ inline uint64_t perm_at(const PermKey& key, uint64_t i, uint64_t n) {
  const uint32_t mask = (1u << key.hb) - 1u;
  uint64_t x = i;
  do {
    uint32_t l = (uint32_t)(x >> key.hb) & mask, r = (uint32_t)x & mask;
#pragma unroll
    for (int round = 0; round < 6; ++round) {
      const uint32_t f = perm_mix(r + key.k[round]) & mask;
      const uint32_t nl = r;
      r = l ^ f;
      l = nl;
    }
    x = ((uint64_t)l << key.hb) | r;
  } while (x >= n);
  return x;
}
static PermKey make_perm_key(uint64_t seed, uint64_t draw, uint64_t n) {
  PermKey key{};
  int bits = 2;
  while (bits < 64 && (1ull << bits) < n) ++bits;
  key.hb = (bits + 1) / 2;
  for (int r = 0; r < 6; ++r) {   // splitmix64 of (seed, draw, round)
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * (draw * 8ull + (uint64_t)r + 1ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    key.k[r] = (uint32_t)z;
  }
  return key;
}

// ---- pixel ids -> rays + colours ------------------------------------------------------------------------
struct BankArgs {
  const void* image;      // [H,W,3] of this view, fp32 (0..255) or uint8
  const double* pose;     // 12 doubles: c2w rows 0..2 (3x4 row-major)
  double focal, cx, cy;
  int W;
  int64_t n_pix;
};

template <typename PIX>
__global__ void bank_sample_kernel(BankArgs a, int mode, const int64_t* __restrict__ pixel_ids, int64_t pix0, PermKey key,
                                   int64_t n, float* __restrict__ rays_o, float* __restrict__ rays_d,
                                   float* __restrict__ colors, int64_t* __restrict__ ids_out) {
  double c2w[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) c2w[k] = __ldg(a.pose + k);
  const PIX* img = reinterpret_cast<const PIX*>(a.image);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t pid = mode == 0 ? pixel_ids[i] : mode == 1 ? pix0 + i : (int64_t)perm_at(key, (uint64_t)i, (uint64_t)a.n_pix);
    // blender.py:119-127 -- float64 like the NumPy original, rounded to fp32 at the end (:150-152 .float())
    const double u = (double)(pid % a.W), v = (double)(pid / a.W);
    const double dx = (u - a.cx) / a.focal, dy = -(v - a.cy) / a.focal, dz = -1.0;
    const double wx = c2w[0] * dx + c2w[1] * dy + c2w[2] * dz;
    const double wy = c2w[4] * dx + c2w[5] * dy + c2w[6] * dz;
    const double wz = c2w[8] * dx + c2w[9] * dy + c2w[10] * dz;
    const double nrm = sqrt(wx * wx + wy * wy + wz * wz);
    if (rays_d) { rays_d[3 * i + 0] = (float)(wx / nrm); rays_d[3 * i + 1] = (float)(wy / nrm); rays_d[3 * i + 2] = (float)(wz / nrm); }
    if (rays_o) { rays_o[3 * i + 0] = (float)c2w[3]; rays_o[3 * i + 1] = (float)c2w[7]; rays_o[3 * i + 2] = (float)c2w[11]; }
    if (colors) {   // blender.py:117: image[v, u] / 255.0 in fp32
#pragma unroll
      for (int c = 0; c < 3; ++c) colors[3 * i + c] = __fdiv_rn((float)img[3 * pid + c], 255.0f);
    }
    if (ids_out) ids_out[i] = pid;
  }
}

// ---- evaluator sums ---------------------------------------------------------------------------------------
constexpr int kEvalThreads = 256;

// Deterministic grid reduction of one double + one uint64 per thread: warp shuffle, fixed-order fold per block, then the
// last block to finish folds the per-block partials in a fixed lane-strided order (same grid => same bits).
__device__ __forceinline__ void grid_fold(double a, unsigned long long b, double* part_a, unsigned long long* part_b,
                                          unsigned int* done, double* out_a, unsigned long long* out_b) {
  __shared__ double s_a[kEvalThreads / 32];
  __shared__ unsigned long long s_b[kEvalThreads / 32];
  __shared__ bool s_last;
  const int tid = threadIdx.y * blockDim.x + threadIdx.x, nblk = gridDim.x * gridDim.y, blk = blockIdx.y * gridDim.x + blockIdx.x;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_down_sync(0xffffffffu, a, o);
    b += __shfl_down_sync(0xffffffffu, b, o);
  }
  if ((tid & 31) == 0) { s_a[tid >> 5] = a; s_b[tid >> 5] = b; }
  __syncthreads();
  if (tid == 0) {
    double fa = 0.0; unsigned long long fb = 0;
    for (int w = 0; w < kEvalThreads / 32; ++w) { fa += s_a[w]; fb += s_b[w]; }
    part_a[blk] = fa; part_b[blk] = fb;
    __threadfence();
    s_last = atomicAdd(done, 1u) == (unsigned int)nblk - 1u;
  }
  __syncthreads();
  if (s_last && tid < 32) {      // the last block's first warp folds the per-block partials: lane-strided, then a fixed tree
    __threadfence();
    double fa = 0.0; unsigned long long fb = 0;
    for (int k = tid; k < nblk; k += 32) { fa += __ldcg(part_a + k); fb += __ldcg(part_b + k); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      fa += __shfl_down_sync(0xffffffffu, fa, o);
      fb += __shfl_down_sync(0xffffffffu, fb, o);
    }
    if (tid == 0) {
      *out_a = fa;
      if (out_b) *out_b = fb;
      *done = 0;
    }
  }
}

__global__ void __launch_bounds__(kEvalThreads)
eval_image_kernel(const float* __restrict__ pred, const float* __restrict__ gt, int64_t n, double* __restrict__ part_sse,
                  unsigned long long* __restrict__ part_sq8, unsigned int* __restrict__ done, double* __restrict__ sse_out,
                  unsigned long long* __restrict__ sq8_out) {
  double sse = 0.0;
  unsigned long long sq8 = 0;
  auto term = [&](float pv, float gv) {
    const float p = fminf(fmaxf(pv, 0.f), 1.f), g = fminf(fmaxf(gv, 0.f), 1.f);            // np.clip(x, 0, 1), nerf.py:98-99
    const float d = __fsub_rn(p, g);
    sse += (double)__fmul_rn(d, d);                                                          // (pred - gt) ** 2 in fp32, :101
    // psnr_metric, nerf.py:24-26: (x * 255).astype(np.uint8) truncates; uint8 - uint8 and ** 2 wrap modulo 256
    const uint32_t pq = (uint32_t)__float2int_rz(__fmul_rn(p, 255.0f)), gq = (uint32_t)__float2int_rz(__fmul_rn(g, 255.0f));
    const uint32_t dq = (pq - gq) & 255u;
    sq8 += (dq * dq) & 255u;
  };
  const int64_t tid0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
  const bool vec = ((reinterpret_cast<uintptr_t>(pred) | reinterpret_cast<uintptr_t>(gt)) & 15) == 0;
  const int64_t n4 = vec ? n / 4 : 0;
  for (int64_t i = tid0; i < n4; i += nthr) {          // 16-byte loads; the per-thread order of terms is fixed by the grid
    const float4 p = __ldg(reinterpret_cast<const float4*>(pred) + i), g = __ldg(reinterpret_cast<const float4*>(gt) + i);
    term(p.x, g.x); term(p.y, g.y); term(p.z, g.z); term(p.w, g.w);
  }
  for (int64_t i = n4 * 4 + tid0; i < n; i += nthr) term(pred[i], gt[i]);
  grid_fold(sse, sq8, part_sse, part_sq8, done, sse_out, sq8_out);
}

// ssim_metric, nerf.py:52-77: both images quantised with (x * 255).astype(np.uint8) (:66-67; x already clipped, :98-99)
__global__ void quantise_u8_kernel(const float* __restrict__ x, int64_t n, uint8_t* __restrict__ q) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    q[i] = (uint8_t)__float2int_rz(__fmul_rn(fminf(fmaxf(x[i], 0.f), 1.f), 255.0f));
}

// skimage.metrics.structural_similarity(uint8, uint8, win_size=win, channel_axis=2) (uniform window, K1 .01, K2 .03, sample
// covariance, data_range 255): one thread per (pixel of the border-cropped map, channel); the window sums are exact
// integers, the SSIM value is evaluated in float64 and summed over the map.
__global__ void __launch_bounds__(kEvalThreads)
ssim_kernel(const uint8_t* __restrict__ p8, const uint8_t* __restrict__ g8, int H, int W, int win, double* __restrict__ part,
            unsigned long long* __restrict__ part_n, unsigned int* __restrict__ done, double* __restrict__ out) {
  // the block's 32 x 8 outputs read a (32+win-1) x (8+win-1) pixel tile of both images: staged once in shared memory
  extern __shared__ uint8_t s_tile[];
  const int tw = 32 + win - 1, th = 8 + win - 1, row_bytes = tw * 3;
  uint8_t* sp = s_tile;
  uint8_t* sg = s_tile + th * row_bytes;
  const int bx = blockIdx.x * 32, by = blockIdx.y * 8, tid = threadIdx.y * 32 + threadIdx.x;
  for (int e = tid; e < th * row_bytes; e += kEvalThreads) {
    const int ty = e / row_bytes, tb = e % row_bytes;
    const int gy = by + ty, gxb = bx * 3 + tb;
    const bool in = gy < H && gxb < W * 3;
    sp[e] = in ? __ldg(p8 + (size_t)gy * W * 3 + gxb) : 0;
    sg[e] = in ? __ldg(g8 + (size_t)gy * W * 3 + gxb) : 0;
  }
  __syncthreads();
  const int x0 = bx + threadIdx.x, y0 = by + threadIdx.y;
  const int ow = W - win + 1, oh = H - win + 1;
  double acc = 0.0;
  if (x0 < ow && y0 < oh) {
    uint32_t sx[3] = {0, 0, 0}, sy[3] = {0, 0, 0}, sxx[3] = {0, 0, 0}, syy[3] = {0, 0, 0}, sxy[3] = {0, 0, 0};
    for (int dy = 0; dy < win; ++dy) {
      const uint8_t* pr = sp + (threadIdx.y + dy) * row_bytes + threadIdx.x * 3;
      const uint8_t* gr = sg + (threadIdx.y + dy) * row_bytes + threadIdx.x * 3;
      for (int dx = 0; dx < win; ++dx) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const uint32_t a = pr[dx * 3 + c], b = gr[dx * 3 + c];
          sx[c] += a; sy[c] += b; sxx[c] += a * a; syy[c] += b * b; sxy[c] += a * b;
        }
      }
    }
    const double np_ = (double)(win * win), cov_norm = np_ / (np_ - 1.0);
    const double C1 = (0.01 * 255.0) * (0.01 * 255.0), C2 = (0.03 * 255.0) * (0.03 * 255.0);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const double ux = sx[c] / np_, uy = sy[c] / np_, uxx = sxx[c] / np_, uyy = syy[c] / np_, uxy = sxy[c] / np_;
      const double vx = cov_norm * (uxx - ux * ux), vy = cov_norm * (uyy - uy * uy), vxy = cov_norm * (uxy - ux * uy);
      acc += ((2.0 * ux * uy + C1) * (2.0 * vxy + C2)) / ((ux * ux + uy * uy + C1) * (vx + vy + C2));
    }
  }
  grid_fold(acc, 0ull, part, part_n, done, out, nullptr);
}

}  // namespace nerf

using namespace nerf;

static NerfEngine::Bank* bank_of(NerfEngine* h, int bank) {
  if (!h || bank < 0 || bank >= (int)h->banks.size() || !h->banks[bank].images) return nullptr;
  return &h->banks[bank];
}

extern "C" int nerf_bank_create(NerfEngine* h, int n_views, int H, int W, double focal, double cx, double cy, int storage,
                                int* bank_out) {
  if (!h || !bank_out) return NERF_E_INVALID;
  if (n_views <= 0 || H <= 0 || W <= 0 || !(focal > 0.0) || (storage != 0 && storage != 1))
    return h->fail(NERF_E_INVALID, "nerf_bank_create: bad argument");
  if ((int64_t)H * W > (1ll << 32)) return h->fail(NERF_E_INVALID, "nerf_bank_create: views larger than 2^32 pixels are not supported");
  cudaSetDevice(h->device);
  NerfEngine::Bank b{};
  b.V = n_views; b.H = H; b.W = W; b.focal = focal; b.cx = cx; b.cy = cy; b.storage = storage;
  const size_t bytes = (size_t)n_views * H * W * 3 * (storage ? 1 : 4);
  cudaError_t e = cudaMalloc(&b.images, bytes);
  if (e != cudaSuccess) { cudaGetLastError(); return h->fail(NERF_E_NOMEM, std::string("nerf_bank_create: cudaMalloc of ") + std::to_string(bytes) + " B: " + cudaGetErrorString(e)); }
  e = cudaMalloc(&b.poses, (size_t)n_views * 12 * sizeof(double));
  if (e != cudaSuccess) { cudaGetLastError(); cudaFree(b.images); return h->fail(NERF_E_NOMEM, "nerf_bank_create: cudaMalloc of the pose table failed"); }
  b.loaded.assign(n_views, 0);
  h->banks.push_back(b);
  *bank_out = (int)h->banks.size() - 1;
  return NERF_OK;
}

extern "C" int nerf_bank_destroy(NerfEngine* h, int bank) {
  NerfEngine::Bank* b = bank_of(h, bank);
  if (!b) return h ? h->fail(NERF_E_INVALID, "nerf_bank_destroy: no such bank") : NERF_E_INVALID;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  cudaFree(b->images); cudaFree(b->poses);
  b->images = nullptr; b->poses = nullptr;
  return NERF_OK;
}

extern "C" int nerf_bank_set_view(NerfEngine* h, int bank, int view, const void* image_host, const double* c2w_host, void* stream) {
  NerfEngine::Bank* b = bank_of(h, bank);
  if (!b) return h ? h->fail(NERF_E_INVALID, "nerf_bank_set_view: no such bank") : NERF_E_INVALID;
  if (view < 0 || view >= b->V || !image_host || !c2w_host) return h->fail(NERF_E_INVALID, "nerf_bank_set_view: bad argument");
  cudaSetDevice(h->device);
  const size_t view_bytes = (size_t)b->H * b->W * 3 * (b->storage ? 1 : 4);
  NERF_CUDA_TRY(h, cudaMemcpyAsync((uint8_t*)b->images + view_bytes * view, image_host, view_bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
  NERF_CUDA_TRY(h, cudaMemcpyAsync(b->poses + 12 * (size_t)view, c2w_host, 12 * sizeof(double), cudaMemcpyHostToDevice, (cudaStream_t)stream));
  NERF_CUDA_TRY(h, cudaStreamSynchronize((cudaStream_t)stream));   // the host buffers may be reused on return
  b->loaded[view] = 1;
  return NERF_OK;
}

extern "C" int nerf_bank_sample(NerfEngine* h, int bank, int view, int mode, const int64_t* pixel_ids, int64_t pix0, uint64_t draw,
                                int64_t n, float* rays_o, float* rays_d, float* colors, int64_t* ids_out, void* stream) {
  NerfEngine::Bank* b = bank_of(h, bank);
  if (!b) return h ? h->fail(NERF_E_INVALID, "nerf_bank_sample: no such bank") : NERF_E_INVALID;
  if (n == 0) return NERF_OK;
  const int64_t n_pix = (int64_t)b->H * b->W;
  if (view < 0 || view >= b->V || n < 0 || mode < 0 || mode > 2) return h->fail(NERF_E_INVALID, "nerf_bank_sample: bad argument");
  if (!b->loaded[view]) return h->fail(NERF_E_STATE, "nerf_bank_sample: view " + std::to_string(view) + " was never uploaded (nerf_bank_set_view)");
  if (mode == 0 && !pixel_ids) return h->fail(NERF_E_INVALID, "nerf_bank_sample: mode 0 needs pixel_ids");
  if (mode == 1 && (pix0 < 0 || pix0 + n > n_pix)) return h->fail(NERF_E_INVALID, "nerf_bank_sample: pixel range outside the image");
  if (mode == 2 && n > n_pix) return h->fail(NERF_E_INVALID, "nerf_bank_sample: cannot draw more pixels than the view has without replacement");
  cudaSetDevice(h->device);
  BankArgs a{};
  a.image = (const uint8_t*)b->images + (size_t)n_pix * 3 * (b->storage ? 1 : 4) * view;
  a.pose = b->poses + 12 * (size_t)view;
  a.focal = b->focal; a.cx = b->cx; a.cy = b->cy; a.W = b->W; a.n_pix = n_pix;
  const PermKey key = make_perm_key(h->cfg.seed, draw, (uint64_t)n_pix);
  const int blocks = (int)std::min<int64_t>((n + 127) / 128, (int64_t)h->num_sms * 16);
  if (b->storage) bank_sample_kernel<uint8_t><<<blocks, 128, 0, (cudaStream_t)stream>>>(a, mode, pixel_ids, pix0, key, n, rays_o, rays_d, colors, ids_out);
  else bank_sample_kernel<float><<<blocks, 128, 0, (cudaStream_t)stream>>>(a, mode, pixel_ids, pix0, key, n, rays_o, rays_d, colors, ids_out);
  h->launches++;
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? NERF_OK : h->fail(NERF_E_CUDA, std::string("nerf_bank_sample: ") + cudaGetErrorString(e));
}

// host-side copy of the keyed bijection (the CPU oracle and the multi-rank tests check the device ids against it)
extern "C" int nerf_bank_permutation_host(uint64_t seed, uint64_t draw, int64_t n_pix, int64_t n, int64_t* ids_out) {
  if (n_pix <= 0 || n < 0 || n > n_pix || !ids_out || n_pix > (1ll << 32)) return NERF_E_INVALID;
  const PermKey key = make_perm_key(seed, draw, (uint64_t)n_pix);
  for (int64_t i = 0; i < n; ++i) ids_out[i] = (int64_t)perm_at(key, (uint64_t)i, (uint64_t)n_pix);
  return NERF_OK;
}

extern "C" int nerf_bank_train_step(NerfEngine* h, int bank, int view, int mode, const int64_t* pixel_ids, int64_t pix0, uint64_t draw,
                                    int64_t n, const float* jitter, const float* u, float grad_scale, float* loss, float* flat_grad,
                                    float* rgb_out, void* stream) {
  if (!h) return NERF_E_INVALID;
  if (n <= 0) return h->fail(NERF_E_INVALID, "nerf_bank_train_step: bad argument");
  float *o, *d, *c;
  int rc;
  if ((rc = h->buf("bank.o", (size_t)n * 3, &o)) || (rc = h->buf("bank.d", (size_t)n * 3, &d)) || (rc = h->buf("bank.c", (size_t)n * 3, &c))) return rc;
  if ((rc = nerf_bank_sample(h, bank, view, mode, pixel_ids, pix0, draw, n, o, d, c, nullptr, stream))) return rc;
  return nerf_forward_backward(h, o, d, c, n, jitter, u, grad_scale, loss, flat_grad, rgb_out, stream);
}

extern "C" int nerf_eval_image(NerfEngine* h, const float* rgb_pred, const float* rgb_gt, int64_t n_elems, double* sse_out,
                               unsigned long long* sq8_out, void* stream) {
  if (!h) return NERF_E_INVALID;
  if (!rgb_pred || !rgb_gt || n_elems <= 0 || !sse_out || !sq8_out) return h->fail(NERF_E_INVALID, "nerf_eval_image: bad argument");
  cudaSetDevice(h->device);
  const int blocks = (int)std::min<int64_t>((n_elems + kEvalThreads - 1) / kEvalThreads, (int64_t)h->num_sms * 4);
  double* part_sse; unsigned long long* part_sq8; unsigned int* done;
  int rc;
  const bool fresh = h->ws.find("eval.done") == h->ws.end();
  if ((rc = h->buf("eval.sse", (size_t)h->num_sms * 8, &part_sse)) || (rc = h->buf("eval.sq8", (size_t)h->num_sms * 8, &part_sq8)) ||
      (rc = h->buf("eval.done", 1, &done))) return rc;
  if (fresh) NERF_CUDA_TRY(h, cudaMemsetAsync(done, 0, sizeof(unsigned int), (cudaStream_t)stream));
  eval_image_kernel<<<blocks, kEvalThreads, 0, (cudaStream_t)stream>>>(rgb_pred, rgb_gt, n_elems, part_sse, part_sq8, done, sse_out, sq8_out);
  h->launches++;
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? NERF_OK : h->fail(NERF_E_CUDA, std::string("nerf_eval_image: ") + cudaGetErrorString(e));
}

// ---- allocation-free variant of nerf_eval_image ------------------------------------------------------
// (no counterpart in the reference; for callers that cannot call cudaMalloc: graph capture, serving loops.)  The per-block
// partials and the "blocks done" counter of the grid fold live in a file-scope device array instead of the engine's
// workspace (h->buf = cudaMalloc on first use).  Every GPU that loads the library holds these NERF_EVAL_STATIC_SCRATCH_BYTES
// until exit; every call on a device shares them, so two calls that overlap on the GPU race.
struct EvalStaticScratch {
  double part_sse[NERF_EVAL_STATIC_MAX_BLOCKS];
  unsigned long long part_sq8[NERF_EVAL_STATIC_MAX_BLOCKS];
  unsigned int done;
  unsigned int pad[63];
};
static_assert(sizeof(EvalStaticScratch) == NERF_EVAL_STATIC_SCRATCH_BYTES, "header constant");
static __device__ __align__(256) unsigned char g_eval_static_scratch[sizeof(EvalStaticScratch)];

// In host code `g_eval_static_scratch` is only the symbol's host-side stand-in: the device address is looked up once per
// device, with that device current, and that pointer is what the kernel gets.
static int eval_static_scratch(NerfEngine* h, EvalStaticScratch** out) {
  constexpr int kMaxDev = 64;
  static std::mutex mu;
  static EvalStaticScratch* ptr[kMaxDev] = {};
  if (h->device < 0 || h->device >= kMaxDev) return h->fail(NERF_E_INVALID, "nerf_eval_image_static: device index out of range");
  std::lock_guard<std::mutex> lk(mu);
  if (!ptr[h->device]) {
    void* p = nullptr;
    cudaError_t e = cudaSetDevice(h->device);
    if (e == cudaSuccess) e = cudaGetSymbolAddress(&p, g_eval_static_scratch);
    if (e != cudaSuccess || !p) return h->fail(NERF_E_CUDA, std::string("nerf_eval_image_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
    ptr[h->device] = static_cast<EvalStaticScratch*>(p);
  }
  *out = ptr[h->device];
  return NERF_OK;
}

extern "C" int nerf_eval_image_static(NerfEngine* h, const float* rgb_pred, const float* rgb_gt, int64_t n_elems, double* sse_out,
                                      unsigned long long* sq8_out, void* stream) {
  if (!h) return NERF_E_INVALID;
  if (!rgb_pred || !rgb_gt || n_elems <= 0 || !sse_out || !sq8_out) return h->fail(NERF_E_INVALID, "nerf_eval_image_static: bad argument");
  if (n_elems > (int64_t)NERF_EVAL_STATIC_MAX_ELEMS)
    return h->fail(NERF_E_INVALID, "nerf_eval_image_static: n_elems " + std::to_string((long long)n_elems) + " exceeds NERF_EVAL_STATIC_MAX_ELEMS (" +
                                       std::to_string((long long)NERF_EVAL_STATIC_MAX_ELEMS) + "); use nerf_eval_image");
  // the same grid as nerf_eval_image => the same per-thread order of terms and the same fold => the same bits
  const int64_t blocks = std::min<int64_t>((n_elems + kEvalThreads - 1) / kEvalThreads, (int64_t)h->num_sms * 4);
  if (blocks > NERF_EVAL_STATIC_MAX_BLOCKS)
    return h->fail(NERF_E_INVALID, "nerf_eval_image_static: the grid of this device (" + std::to_string((long long)blocks) +
                                       " blocks) exceeds the static scratch (NERF_EVAL_STATIC_MAX_BLOCKS); use nerf_eval_image");
  EvalStaticScratch* s = nullptr;
  if (int rc = eval_static_scratch(h, &s)) return rc;
  cudaSetDevice(h->device);
  cudaStream_t st = (cudaStream_t)stream;
  // the array keeps the last call's contents: the counter must start at zero (the partials are overwritten before they are read)
  NERF_CUDA_TRY(h, cudaMemsetAsync(&s->done, 0, sizeof(unsigned int), st));
  eval_image_kernel<<<(int)blocks, kEvalThreads, 0, st>>>(rgb_pred, rgb_gt, n_elems, s->part_sse, s->part_sq8, &s->done, sse_out, sq8_out);
  h->launches++;
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? NERF_OK : h->fail(NERF_E_CUDA, std::string("nerf_eval_image_static: ") + cudaGetErrorString(e));
}

extern "C" int nerf_eval_ssim(NerfEngine* h, const float* rgb_pred, const float* rgb_gt, int H, int W, int win_size, double* ssim_sum_out,
                              void* stream) {
  if (!h) return NERF_E_INVALID;
  if (!rgb_pred || !rgb_gt || !ssim_sum_out || win_size < 3 || !(win_size & 1) || win_size > 31 || H < win_size || W < win_size)
    return h->fail(NERF_E_INVALID, "nerf_eval_ssim: bad argument (odd win_size in [3,31], image at least win_size on a side)");
  cudaSetDevice(h->device);
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t n = (int64_t)H * W * 3;
  const dim3 block(32, kEvalThreads / 32), grid((W - win_size + 1 + 31) / 32, (H - win_size + 1 + 7) / 8);
  uint8_t *p8, *g8; double* part; unsigned long long* part_n; unsigned int* done;
  int rc;
  const bool fresh = h->ws.find("ssim.done") == h->ws.end();
  if ((rc = h->buf("ssim.p8", (size_t)n, &p8)) || (rc = h->buf("ssim.g8", (size_t)n, &g8)) ||
      (rc = h->buf("ssim.part", (size_t)grid.x * grid.y, &part)) || (rc = h->buf("ssim.partn", (size_t)grid.x * grid.y, &part_n)) ||
      (rc = h->buf("ssim.done", 1, &done))) return rc;
  if (fresh) NERF_CUDA_TRY(h, cudaMemsetAsync(done, 0, sizeof(unsigned int), st));
  const int qb = (int)std::min<int64_t>((n + 255) / 256, (int64_t)h->num_sms * 8);
  quantise_u8_kernel<<<qb, 256, 0, st>>>(rgb_pred, n, p8);
  quantise_u8_kernel<<<qb, 256, 0, st>>>(rgb_gt, n, g8);
  ssim_kernel<<<grid, block, 2 * (8 + win_size - 1) * (32 + win_size - 1) * 3, st>>>(p8, g8, H, W, win_size, part, part_n, done, ssim_sum_out);
  h->launches += 3;
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? NERF_OK : h->fail(NERF_E_CUDA, std::string("nerf_eval_ssim: ") + cudaGetErrorString(e));
}
