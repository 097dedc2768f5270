// api.cu -- the extern "C" boundary (include/nerf_b200.h) and the render / train-step orchestration.
// Renderer.render (volume_renderer.py:290-432) becomes: stratified t -> coarse MLP -> sample_pdf+merge ->
// fine MLP -> composite, each a kernel of this library launched on the caller's stream.
#include <algorithm>
#include <cstring>
#include <cstdlib>
#include <map>
#include "engine.h"
#include "tc_arena.cuh"

namespace nerf {
int tc2_debug_read(unsigned int out[8]);
void tc2_set_debug_sink(unsigned int*);
void tc2b_set_debug_sink(unsigned int*);
void tc2_set_debug_sink_waitprof(unsigned long long*);
void tc2b_set_debug_sink_waitprof(unsigned long long*);
}
using namespace nerf;

static std::string g_create_error;
static void drop_step_graph(NerfEngine* h);
static unsigned int* g_watchdog_host = nullptr;   // host-mapped, survives a trapped kernel

static int check_cfg(const NerfConfig& c, std::string* why) {
  if (c.n_samples < 4 || c.n_samples > 1024) { *why = "n_samples must be in [4,1024]"; return NERF_E_INVALID; }
  if (c.n_importance < 0 || c.n_samples + c.n_importance > 4096) { *why = "n_importance out of range"; return NERF_E_INVALID; }
  if (c.precision != NERF_PRECISION_BF16 && c.precision != NERF_PRECISION_FP32) { *why = "unknown precision"; return NERF_E_INVALID; }
  if (!(c.t_far > c.t_near)) { *why = "t_far must exceed t_near"; return NERF_E_INVALID; }
  return NERF_OK;
}

extern "C" int nerf_version(void) { return 100; }

extern "C" int nerf_create(const NerfConfig* cfg, int device, NerfEngine** out) {
  if (!cfg || !out) { g_create_error = "null argument"; return NERF_E_INVALID; }
  std::string why;
  if (check_cfg(*cfg, &why)) { g_create_error = why; return NERF_E_INVALID; }
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || device < 0 || device >= count) {
    g_create_error = std::string("no such CUDA device (") + (e == cudaSuccess ? "index out of range" : cudaGetErrorString(e)) + "); the engine has no CPU fallback";
    cudaGetLastError();
    return NERF_E_CUDA;
  }
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, device);
  if (prop.major != 10) {
    g_create_error = "device is sm_" + std::to_string(prop.major * 10 + prop.minor) + "; this library is built for sm_100a (B200) only";
    return NERF_E_INVALID;
  }
  cudaSetDevice(device);
  if (!g_watchdog_host) {
    unsigned int* hp = nullptr;
    if (cudaHostAlloc(&hp, 256, cudaHostAllocMapped) == cudaSuccess) {
      memset(hp, 0, 256);
      unsigned int* dp = nullptr;
      if (cudaHostGetDevicePointer(&dp, hp, 0) == cudaSuccess) {
        g_watchdog_host = hp;
        tc2_set_debug_sink(dp); tc2b_set_debug_sink(dp);
      }
    }
    cudaGetLastError();
  }
  NerfEngine* h = new NerfEngine();
  h->cfg = *cfg;
  h->device = device;
  h->num_sms = prop.multiProcessorCount;
  if (const char* v = getenv("NERF_B200_WREPLICAS")) h->w_replicas = std::max(1, std::min(74, atoi(v)));
  *out = h;
  return NERF_OK;
}

extern "C" int nerf_destroy(NerfEngine* h) {
  if (!h) return NERF_OK;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  drop_step_graph(h);
  for (auto& kv : h->ws) if (kv.second.p) cudaFree(kv.second.p);
  for (auto& b : h->banks) { if (b.images) cudaFree(b.images); if (b.poses) cudaFree(b.poses); }
  if (h->wpack) cudaFree(h->wpack);
  if (h->wpack_bwd) cudaFree(h->wpack_bwd);
  if (h->wrep) cudaFree(h->wrep);
  for (cudaEvent_t e : h->prof_pool) cudaEventDestroy(e);
  if (h->side) cudaStreamDestroy(h->side);
  if (h->cap_stream) cudaStreamDestroy(h->cap_stream);
  if (h->pack_ev) cudaEventDestroy(h->pack_ev);
  if (h->fine_grad_ev) cudaEventDestroy(h->fine_grad_ev);
  if (h->side_fork) cudaEventDestroy(h->side_fork);
  if (h->side_join) cudaEventDestroy(h->side_join);
  for (int i = 0; i < 2; ++i) {
    if (h->hs[i]) cudaStreamDestroy(h->hs[i]);
    if (h->hev[i]) cudaEventDestroy(h->hev[i]);
    if (h->pinned[i]) cudaFreeHost(h->pinned[i]);
  }
  cudaGetLastError();
  delete h;
  return NERF_OK;
}

extern "C" const char* nerf_last_error(const NerfEngine* h) { return h ? h->err.c_str() : g_create_error.c_str(); }
extern "C" int64_t nerf_launch_count(const NerfEngine* h) { return h ? h->launches : 0; }

extern "C" int nerf_set_config(NerfEngine* h, const NerfConfig* cfg) {
  if (!h || !cfg) return NERF_E_INVALID;
  std::string why;
  if (check_cfg(*cfg, &why)) return h->fail(NERF_E_INVALID, why);
  h->cfg = *cfg;
  return NERF_OK;
}

static int post(NerfEngine* h, const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return h->fail(NERF_E_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
  return NERF_OK;
}

// The packed bf16 images are written on the caller's stream; the host-buffer render path runs on the engine's own
// streams and must not read them before the (re)pack has finished.
static int record_pack(NerfEngine* h, cudaStream_t st) {
  if (!h->pack_ev) NERF_CUDA_TRY(h, cudaEventCreateWithFlags(&h->pack_ev, cudaEventDisableTiming));
  NERF_CUDA_TRY(h, cudaEventRecord(h->pack_ev, st));
  return NERF_OK;
}

extern "C" int nerf_set_params(NerfEngine* h, const float* flat, void* stream) {
  if (!h || !flat) return NERF_E_INVALID;
  if (reinterpret_cast<uintptr_t>(flat) & 15) return h->fail(NERF_E_INVALID, "flat parameter buffer must be 16-byte aligned");
  cudaSetDevice(h->device);
  h->params = flat;
  if (!h->wpack) {
    h->wpack_bytes = tc_wpack_bytes();
    NERF_CUDA_TRY(h, cudaMalloc(&h->wpack, h->wpack_bytes));
    NERF_CUDA_TRY(h, cudaMemsetAsync(h->wpack, 0, h->wpack_bytes, (cudaStream_t)stream));
  }
  if (!h->wpack_bwd) NERF_CUDA_TRY(h, cudaMalloc(&h->wpack_bwd, tc_bwd_wpack_bytes()));
  if (h->w_replicas > 1 && !h->wrep) NERF_CUDA_TRY(h, cudaMalloc(&h->wrep, tc_wpack_bias_offset() * h->w_replicas));   // replicas of the stage images
  tc_pack_all(h, flat, h->wpack, h->wpack_bwd, (cudaStream_t)stream);
  h->params_set = true;
  if (int rc = record_pack(h, (cudaStream_t)stream)) return rc;
  return post(h, "nerf_set_params");
}

// ---- stage entry points ---------------------------------------------------------------------------
static Camera make_camera(const double* c2w, int H, int W, double focal, double cx, double cy) {
  Camera cam{};
  for (int r = 0; r < 3; ++r) {
    for (int c = 0; c < 3; ++c) { cam.r64[3 * r + c] = c2w[4 * r + c]; cam.r[3 * r + c] = (float)c2w[4 * r + c]; }
    cam.o64[r] = c2w[4 * r + 3]; cam.o[r] = (float)c2w[4 * r + 3];
  }
  cam.focal = focal; cam.cx = cx; cam.cy = cy; cam.H = H; cam.W = W;
  return cam;
}

extern "C" int nerf_get_rays(NerfEngine* h, const double* c2w_host, int H, int W, double focal, double cx, double cy,
                             const int64_t* pixel_ids, int64_t pix0, int64_t n, float* rays_o, float* rays_d, void* stream) {
  if (h && n == 0) return NERF_OK;
  if (!h || !c2w_host || !rays_o || !rays_d || n < 0 || H <= 0 || W <= 0) return h ? h->fail(NERF_E_INVALID, "nerf_get_rays: bad argument") : NERF_E_INVALID;
  if (!pixel_ids && (pix0 < 0 || pix0 + n > (int64_t)H * W)) return h->fail(NERF_E_INVALID, "nerf_get_rays: pixel range outside the image");
  cudaSetDevice(h->device);
  launch_get_rays(h, make_camera(c2w_host, H, W, focal, cx, cy), pixel_ids, pix0, n, rays_o, rays_d, (cudaStream_t)stream);
  return post(h, "nerf_get_rays");
}

extern "C" int nerf_stratified_t(NerfEngine* h, int64_t n, int S, const float* jitter, float* t, void* stream) {
  if (h && n == 0) return NERF_OK;
  if (!h || !t || n < 0 || S < 2) return h ? h->fail(NERF_E_INVALID, "nerf_stratified_t: bad argument") : NERF_E_INVALID;
  cudaSetDevice(h->device);
  launch_stratified_t(h, n, S, jitter, false, 0, t, (cudaStream_t)stream);
  return post(h, "nerf_stratified_t");
}

extern "C" int nerf_freq_encode(NerfEngine* h, const float* x, int64_t m, int n_freqs, float* out, void* stream) {
  if (h && m == 0) return NERF_OK;
  if (!h || !x || !out || m < 0 || n_freqs < 0 || n_freqs > 16) return h ? h->fail(NERF_E_INVALID, "nerf_freq_encode: bad argument") : NERF_E_INVALID;
  cudaSetDevice(h->device);
  launch_freq_encode(h, x, m, n_freqs, out, (cudaStream_t)stream);
  return post(h, "nerf_freq_encode");
}

static const char* act_arena_name(int net) { return net == NERF_NET_FINE ? "tc.act_f" : "tc.act_c"; }

static int run_mlp(NerfEngine* h, int net, const RowSrc& src, float* raw, bool keep, cudaStream_t st) {
  if (h->cfg.precision == NERF_PRECISION_FP32) return mlp_fp32_forward(h, net, src, raw, keep, st);
  uint8_t* act = nullptr;
  if (keep) {   // training forward: activations saved as tile images for the tensor-core backward
    int rc0 = h->buf(act_arena_name(net), (size_t)arena::num_tiles(src.R * src.S) * arena::act_tile_bytes(), &act);
    if (rc0) return rc0;
  }
  if (h->profiling) cudaEventRecord(h->prof_event(), st);
  const int rc = mlp_tc2_forward(h, net, src, raw, act, st);
  if (h->profiling) { cudaEventRecord(h->prof_event(), st); h->prof_rows += src.R * src.S; h->prof_class.push_back(0); h->prof_class_rows.push_back(src.R * src.S); }
  return rc;
}

extern "C" int nerf_profile_enable(NerfEngine* h, int on) {
  if (!h) return NERF_E_INVALID;
  h->profiling = on != 0;
  return NERF_OK;
}

extern "C" int nerf_profile_read_classes(NerfEngine* h, double* ms3, int64_t* launches3, int64_t* rows3) {
  if (!h) return NERF_E_INVALID;
  cudaSetDevice(h->device);
  NERF_CUDA_TRY(h, cudaDeviceSynchronize());
  double ms[3] = {0, 0, 0};
  int64_t n[3] = {0, 0, 0}, rows[3] = {0, 0, 0};
  for (size_t i = 0; i + 1 < h->prof_used; i += 2) {
    float t = 0.f;
    NERF_CUDA_TRY(h, cudaEventElapsedTime(&t, h->prof_pool[i], h->prof_pool[i + 1]));
    const int c = i / 2 < h->prof_class.size() ? h->prof_class[i / 2] : 0;
    ms[c] += t; n[c]++; rows[c] += i / 2 < h->prof_class_rows.size() ? h->prof_class_rows[i / 2] : 0;
  }
  for (int c = 0; c < 3; ++c) {
    if (ms3) ms3[c] = ms[c];
    if (launches3) launches3[c] = n[c];
    if (rows3) rows3[c] = rows[c];
  }
  h->prof_used = 0; h->prof_rows = 0; h->prof_class.clear(); h->prof_class_rows.clear();
  return NERF_OK;
}

extern "C" int nerf_profile_read(NerfEngine* h, double* mlp_ms, int64_t* mlp_launches, int64_t* mlp_rows) {
  double ms[3]; int64_t n[3], rows[3];
  const int rc = nerf_profile_read_classes(h, ms, n, rows);
  if (rc) return rc;
  if (mlp_ms) *mlp_ms = ms[0];
  if (mlp_launches) *mlp_launches = n[0];
  if (mlp_rows) *mlp_rows = rows[0];
  return NERF_OK;
}

extern "C" int nerf_mlp_forward(NerfEngine* h, int which_net, const float* pts, const float* viewdirs, const uint8_t* mask,
                                int64_t R, int S, float* raw, void* stream) {
  if (!h) return NERF_E_INVALID;
  if (!h->params_set) return h->fail(NERF_E_STATE, "nerf_mlp_forward before nerf_set_params");
  if (R == 0) return NERF_OK;
  if (!pts || !viewdirs || !raw || R < 0 || S < 1 || (which_net != NERF_NET_COARSE && which_net != NERF_NET_FINE))
    return h->fail(NERF_E_INVALID, "nerf_mlp_forward: bad argument");
  cudaSetDevice(h->device);
  RowSrc src;
  src.pts = pts; src.viewdirs = viewdirs; src.mask = mask; src.R = R; src.S = S;
  int rc = run_mlp(h, which_net, src, raw, false, (cudaStream_t)stream);
  return rc ? rc : post(h, "nerf_mlp_forward");
}

extern "C" int nerf_composite(NerfEngine* h, const float* raw, const float* t, int64_t n, int S, int white_bkgd, float* rgb,
                              float* depth, float* acc, float* weights, void* stream) {
  if (h && n == 0) return NERF_OK;
  if (!h || !raw || !t || n < 0 || S < 1) return h ? h->fail(NERF_E_INVALID, "nerf_composite: bad argument") : NERF_E_INVALID;
  cudaSetDevice(h->device);
  launch_composite(h, raw, t, n, S, white_bkgd, rgb, depth, acc, weights, (cudaStream_t)stream);
  return post(h, "nerf_composite");
}

extern "C" int nerf_sample_pdf(NerfEngine* h, const float* raw_c, const float* t_c, int64_t n, int Nc, int Nf, const float* u,
                               const float* cdf_in, float* t_fine, int64_t* inds, int64_t* below, int64_t* above,
                               float* cdf_out, float* t_sorted, int32_t* sort_idx, void* stream) {
  if (h && n == 0) return NERF_OK;
  if (!h || !raw_c || !t_c || n < 0) return h ? h->fail(NERF_E_INVALID, "nerf_sample_pdf: bad argument") : NERF_E_INVALID;
  cudaSetDevice(h->device);
  int rc = launch_sample_pdf(h, raw_c, t_c, n, Nc, Nf, u, false, 0, cdf_in, t_fine, inds, below, above, cdf_out, t_sorted,
                             sort_idx, (cudaStream_t)stream);
  return rc ? rc : post(h, "nerf_sample_pdf");
}

extern "C" int nerf_sample_pdf_prune(NerfEngine* h, const float* raw_c, const float* t_c, int64_t n, int Nc, int Nf, const float* u,
                                     const float* cdf_in, float weights_threshold, float* t_fine, uint8_t* valid_fine,
                                     uint8_t* mask_sorted, float* t_sorted, int32_t* sort_idx, void* stream) {
  if (h && n == 0) return NERF_OK;
  if (!h || !raw_c || !t_c || n < 0 || (!valid_fine && !mask_sorted))
    return h ? h->fail(NERF_E_INVALID, "nerf_sample_pdf_prune: bad argument") : NERF_E_INVALID;
  cudaSetDevice(h->device);
  int rc = launch_sample_pdf(h, raw_c, t_c, n, Nc, Nf, u, false, 0, cdf_in, t_fine, nullptr, nullptr, nullptr, nullptr, t_sorted,
                             sort_idx, (cudaStream_t)stream, weights_threshold, valid_fine, mask_sorted);
  return rc ? rc : post(h, "nerf_sample_pdf_prune");
}

// ---- Renderer.render ----------------------------------------------------------------------------
struct RenderTaps {   // device workspace of one chunk; kept for the backward pass
  float *t_c = nullptr, *raw_c = nullptr, *t_s = nullptr, *raw_f = nullptr;
  int32_t* sort_idx = nullptr;
  uint64_t call_id = 0;
};

static uint64_t g_call_counter = 0;

// One chunk of rays through the whole pipeline.  `slot` selects a private workspace (host path ping-pong).
static int render_chunk(NerfEngine* h, const float* o, const float* d, int64_t n, const float* jitter, const float* u,
                        const NerfRenderOut& out, bool keep, int slot, RenderTaps* taps, cudaStream_t st, bool final_composite = true) {
  const NerfConfig& c = h->cfg;
  const int Nc = c.n_samples, Nf = c.n_importance, Nt = Nc + Nf;
  const uint64_t call_id = ++g_call_counter;
  char nm[4][32];
  snprintf(nm[0], 32, "r%d.t_c", slot); snprintf(nm[1], 32, "r%d.raw_c", slot);
  snprintf(nm[2], 32, "r%d.t_s", slot); snprintf(nm[3], 32, "r%d.raw_f", slot);
  RenderTaps tp;
  tp.call_id = call_id;
  int rc;
  if ((rc = h->buf(nm[0], (size_t)n * Nc, &tp.t_c))) return rc;
  if ((rc = h->buf(nm[1], (size_t)n * Nc * 4, &tp.raw_c))) return rc;
  const bool perturb = c.perturb != 0;
  launch_stratified_t(h, n, Nc, perturb ? jitter : nullptr, perturb && !jitter, call_id, tp.t_c, st);
  RowSrc src;
  src.rays_o = o; src.rays_d = d; src.t = tp.t_c; src.R = n; src.S = Nc;
  if ((rc = run_mlp(h, NERF_NET_COARSE, src, tp.raw_c, keep, st))) return rc;
  if (Nf > 0) {
    if ((rc = h->buf(nm[2], (size_t)n * Nt, &tp.t_s))) return rc;
    if ((rc = h->buf(nm[3], (size_t)n * Nt * 4, &tp.raw_f))) return rc;
    if (keep) {
      char nb[32]; snprintf(nb, 32, "r%d.sort_idx", slot);
      if ((rc = h->buf(nb, (size_t)n * Nt, &tp.sort_idx))) return rc;
    }
    const bool tu = c.train_u != 0;
    uint8_t* prune_mask = nullptr;       // a-M: fine-pass validity mask (fast_sampling); masked samples get raw = 0
    if (c.fast_sampling) {
      char nb[32]; snprintf(nb, 32, "r%d.prune", slot);
      if ((rc = h->buf(nb, (size_t)n * Nt, &prune_mask))) return rc;
    }
    // fast_sampling == 2 (bf16 tensor-core path): masked rows are not evaluated at all -- sample_pdf also emits the
    // compacted list of valid rows, raw_f is zero-filled and the fine MLP pass walks the list (same results as the mask).
    // The training forward (keep) always masks: its tile images are indexed by row for the backward kernels.  A masked
    // sample has raw = 0, hence weight 0 and relu'(0) = 0: composite_bwd hands the MLP backward an exactly zero d_raw row,
    // which is the reference's autograd through `out[m] = flat` (network.py:238-253) -- no gradient from pruned samples.
    int32_t *row_list = nullptr, *row_count = nullptr;
    if (c.fast_sampling == 2 && !keep && c.precision == NERF_PRECISION_BF16) {
      char nb[2][32]; snprintf(nb[0], 32, "r%d.rows", slot); snprintf(nb[1], 32, "r%d.nrows", slot);
      if ((rc = h->buf(nb[0], (size_t)n * Nt, &row_list)) || (rc = h->buf(nb[1], 1, &row_count))) return rc;
      NERF_CUDA_TRY(h, cudaMemsetAsync(row_count, 0, sizeof(int32_t), st));
      NERF_CUDA_TRY(h, cudaMemsetAsync(tp.raw_f, 0, (size_t)n * Nt * 4 * sizeof(float), st));
    }
    if ((rc = launch_sample_pdf(h, tp.raw_c, tp.t_c, n, Nc, Nf, tu ? u : nullptr, tu && !u, call_id, nullptr, nullptr, nullptr,
                                nullptr, nullptr, nullptr, tp.t_s, tp.sort_idx, st, c.weights_threshold, nullptr, prune_mask,
                                row_list, row_count))) return rc;
    src.t = tp.t_s; src.S = Nt; src.mask = row_list ? nullptr : prune_mask; src.row_list = row_list; src.row_count = row_count;
    if ((rc = run_mlp(h, NERF_NET_FINE, src, tp.raw_f, keep, st))) return rc;
    if (final_composite) launch_composite(h, tp.raw_f, tp.t_s, n, Nt, c.white_bkgd, out.rgb, out.depth, out.acc, nullptr, st);
    if (out.rgb_coarse || out.depth_coarse || out.acc_coarse)
      launch_composite(h, tp.raw_c, tp.t_c, n, Nc, c.white_bkgd, out.rgb_coarse, out.depth_coarse, out.acc_coarse, nullptr, st);
  } else {
    if (final_composite) launch_composite(h, tp.raw_c, tp.t_c, n, Nc, c.white_bkgd, out.rgb, out.depth, out.acc, nullptr, st);
    if (out.rgb_coarse || out.depth_coarse || out.acc_coarse)
      launch_composite(h, tp.raw_c, tp.t_c, n, Nc, c.white_bkgd, out.rgb_coarse, out.depth_coarse, out.acc_coarse, nullptr, st);
  }
  if (taps) *taps = tp;
  return post(h, "render");
}

static NerfRenderOut offset_out(const NerfRenderOut& o, int64_t r) {
  NerfRenderOut q = o;
  if (q.rgb) q.rgb += 3 * r;
  if (q.depth) q.depth += r;
  if (q.acc) q.acc += r;
  if (q.rgb_coarse) q.rgb_coarse += 3 * r;
  if (q.depth_coarse) q.depth_coarse += r;
  if (q.acc_coarse) q.acc_coarse += r;
  return q;
}

constexpr int64_t kRenderChunk = 65536;   // rays per pass: bounds the HBM workspace (~5 KB/ray at 64+128)

extern "C" int nerf_render_rays(NerfEngine* h, const float* rays_o, const float* rays_d, int64_t n, const float* jitter,
                                const float* u, const NerfRenderOut* out, void* stream) {
  if (!h) return NERF_E_INVALID;
  if (!h->params_set) return h->fail(NERF_E_STATE, "nerf_render_rays before nerf_set_params");
  if (n == 0) return NERF_OK;
  if (!rays_o || !rays_d || !out || n < 0) return h->fail(NERF_E_INVALID, "nerf_render_rays: bad argument");
  cudaSetDevice(h->device);
  const int Nc = h->cfg.n_samples, Nf = h->cfg.n_importance;
  for (int64_t r = 0; r < n; r += kRenderChunk) {
    const int64_t m = std::min(kRenderChunk, n - r);
    int rc = render_chunk(h, rays_o + 3 * r, rays_d + 3 * r, m, jitter ? jitter + r * Nc : nullptr, u ? u + r * Nf : nullptr,
                          offset_out(*out, r), false, 0, nullptr, (cudaStream_t)stream);
    if (rc) return rc;
  }
  return NERF_OK;
}

extern "C" int nerf_render_camera(NerfEngine* h, const double* c2w_host, int H, int W, double focal, double cx, double cy,
                                  int64_t pix0, int64_t n, const NerfRenderOut* out, void* stream) {
  if (!h) return NERF_E_INVALID;
  if (!h->params_set) return h->fail(NERF_E_STATE, "nerf_render_camera before nerf_set_params");
  if (n == 0) return NERF_OK;
  if (!c2w_host || !out || n < 0 || pix0 < 0 || pix0 + n > (int64_t)H * W) return h->fail(NERF_E_INVALID, "nerf_render_camera: bad argument");
  cudaSetDevice(h->device);
  const Camera cam = make_camera(c2w_host, H, W, focal, cx, cy);
  for (int64_t r = 0; r < n; r += kRenderChunk) {
    const int64_t m = std::min(kRenderChunk, n - r);
    float *o, *d;
    int rc;
    if ((rc = h->buf("cam.o", (size_t)m * 3, &o))) return rc;
    if ((rc = h->buf("cam.d", (size_t)m * 3, &d))) return rc;
    launch_get_rays(h, cam, nullptr, pix0 + r, m, o, d, (cudaStream_t)stream);
    if ((rc = render_chunk(h, o, d, m, nullptr, nullptr, offset_out(*out, r), false, 0, nullptr, (cudaStream_t)stream))) return rc;
  }
  return NERF_OK;
}

// Host-buffer path: chunks ping-pong over two internal streams so the H2D of chunk i+1 and the D2H of
// chunk i-1 overlap the kernels of chunk i.  (fp32 validation mode shares one MLP workspace -> one stream.)
extern "C" int nerf_render_rays_host(NerfEngine* h, const float* rays_o_host, const float* rays_d_host, int64_t n,
                                     float* rgb_host, float* depth_host) {
  if (!h) return NERF_E_INVALID;
  if (!h->params_set) return h->fail(NERF_E_STATE, "nerf_render_rays_host before nerf_set_params");
  if (n == 0) return NERF_OK;
  if (!rays_o_host || !rays_d_host || !rgb_host || !depth_host || n < 0) return h->fail(NERF_E_INVALID, "nerf_render_rays_host: bad argument");
  cudaSetDevice(h->device);
  const int nslots = h->cfg.precision == NERF_PRECISION_FP32 ? 1 : 2;
  const size_t per_ray = 10 * sizeof(float);   // o3 d3 rgb3 depth1
  const size_t need = (size_t)kRenderChunk * per_ray;
  for (int i = 0; i < 2; ++i) {
    if (!h->hs[i]) NERF_CUDA_TRY(h, cudaStreamCreateWithFlags(&h->hs[i], cudaStreamNonBlocking));
    if (!h->hev[i]) NERF_CUDA_TRY(h, cudaEventCreateWithFlags(&h->hev[i], cudaEventDisableTiming));
    if (h->pinned_bytes < need) {
      if (h->pinned[i]) cudaFreeHost(h->pinned[i]);
      NERF_CUDA_TRY(h, cudaMallocHost(&h->pinned[i], need));
    }
  }
  h->pinned_bytes = std::max(h->pinned_bytes, need);
  // the internal streams start after the last weight (re)pack issued on any caller stream (nerf_set_params / nerf_clip_adam)
  if (h->pack_ev) for (int i = 0; i < 2; ++i) NERF_CUDA_TRY(h, cudaStreamWaitEvent(h->hs[i], h->pack_ev, 0));
  struct Pending { int64_t r = -1, m = 0; } pend[2];
  auto drain = [&](int s) {   // wait for slot s and copy its maps out of pinned memory
    if (pend[s].r < 0) return;
    cudaEventSynchronize(h->hev[s]);
    const float* pin = reinterpret_cast<const float*>(h->pinned[s]);
    memcpy(rgb_host + 3 * pend[s].r, pin + 6 * kRenderChunk, (size_t)pend[s].m * 3 * sizeof(float));
    memcpy(depth_host + pend[s].r, pin + 9 * kRenderChunk, (size_t)pend[s].m * sizeof(float));
    pend[s].r = -1;
  };
  int64_t ci = 0;
  for (int64_t r = 0; r < n; r += kRenderChunk, ++ci) {
    const int s = (int)(ci % nslots);
    drain(s);
    const int64_t m = std::min(kRenderChunk, n - r);
    float* pin = reinterpret_cast<float*>(h->pinned[s]);
    memcpy(pin, rays_o_host + 3 * r, (size_t)m * 3 * sizeof(float));
    memcpy(pin + 3 * kRenderChunk, rays_d_host + 3 * r, (size_t)m * 3 * sizeof(float));
    float *o, *d, *rgb, *dep;
    char nm[32];
    int rc;
    snprintf(nm, 32, "h%d.o", s); if ((rc = h->buf(nm, (size_t)kRenderChunk * 3, &o))) return rc;
    snprintf(nm, 32, "h%d.d", s); if ((rc = h->buf(nm, (size_t)kRenderChunk * 3, &d))) return rc;
    snprintf(nm, 32, "h%d.rgb", s); if ((rc = h->buf(nm, (size_t)kRenderChunk * 3, &rgb))) return rc;
    snprintf(nm, 32, "h%d.dep", s); if ((rc = h->buf(nm, (size_t)kRenderChunk, &dep))) return rc;
    cudaStream_t st = h->hs[s];
    NERF_CUDA_TRY(h, cudaMemcpyAsync(o, pin, (size_t)m * 3 * sizeof(float), cudaMemcpyHostToDevice, st));
    NERF_CUDA_TRY(h, cudaMemcpyAsync(d, pin + 3 * kRenderChunk, (size_t)m * 3 * sizeof(float), cudaMemcpyHostToDevice, st));
    NerfRenderOut out{};
    out.rgb = rgb; out.depth = dep;
    if ((rc = render_chunk(h, o, d, m, nullptr, nullptr, out, false, 1 + s, nullptr, st))) return rc;
    NERF_CUDA_TRY(h, cudaMemcpyAsync(pin + 6 * kRenderChunk, rgb, (size_t)m * 3 * sizeof(float), cudaMemcpyDeviceToHost, st));
    NERF_CUDA_TRY(h, cudaMemcpyAsync(pin + 9 * kRenderChunk, dep, (size_t)m * sizeof(float), cudaMemcpyDeviceToHost, st));
    NERF_CUDA_TRY(h, cudaEventRecord(h->hev[s], st));
    pend[s].r = r; pend[s].m = m;
  }
  for (int s = 0; s < 2; ++s) drain(s);
  return post(h, "nerf_render_rays_host");
}

// ---- training step ------------------------------------------------------------------------------
// loss = MSE(render(rays), rgb_gt) (trainers/nerf.py:27-33); backward follows the reference autograd
// graph: fine MLP gets rgb+sigma gradients and an input gradient that flows
// through PE -> t_fine -> inverse CDF -> coarse weights -> coarse sigma -> coarse trunk.
static int train_forward(NerfEngine* h, const float* rays_o, const float* rays_d, int64_t n, const float* jitter,
                         const float* u, float* rgb, float* depth, bool final_composite, cudaStream_t st) {
  NerfRenderOut out{};
  out.rgb = rgb; out.depth = depth;
  RenderTaps tp;
  int rc = render_chunk(h, rays_o, rays_d, n, jitter, u, out, true, 9, &tp, st, final_composite);
  if (rc) { h->taps.n = -1; return rc; }
  h->taps.t_c = tp.t_c; h->taps.raw_c = tp.raw_c; h->taps.t_s = tp.t_s; h->taps.raw_f = tp.raw_f;
  h->taps.sort_idx = tp.sort_idx; h->taps.call_id = tp.call_id; h->taps.n = n; h->taps.jitter = jitter; h->taps.u = u;
  return NERF_OK;
}

// MLP backward of one net: fp32 validation kernels, or the tensor-core dgrad chain + wgrad over the saved tile images.
static int mlp_backward(NerfEngine* h, int net, const RowSrc& src, const float* d_raw, bool sigma_only, float* G, float* d_t,
                        cudaStream_t st) {
  if (h->cfg.precision == NERF_PRECISION_FP32) return mlp_fp32_backward(h, net, src, d_raw, sigma_only, G, d_t, st);
  const int64_t M = src.R * src.S;
  const bool full = !sigma_only;
  uint8_t *act, *grad;
  float* scratch = nullptr;
  int rc;
  if ((rc = h->buf(act_arena_name(net), (size_t)arena::num_tiles(M) * arena::act_tile_bytes(), &act))) return rc;
  if ((rc = h->buf("tc.grad", (size_t)arena::num_tiles(M) * arena::grad_tile_bytes(), &grad))) return rc;
  if (full && (rc = h->buf("tc.pe_scratch", (size_t)M * 64, &scratch))) return rc;
  if (full && !d_t) {   // coarse-only rendering: the input gradient is not needed; give the kernel a sink
    if ((rc = h->buf("tc.dt_sink", (size_t)M, &d_t))) return rc;
  }
  auto mark = [&](int cls) {
    if (!h->profiling) return;
    cudaEventRecord(h->prof_event(), st);
    if (cls >= 0) { h->prof_class.push_back(cls); h->prof_class_rows.push_back(M); }
  };
  mark(-1);
  if ((rc = mlp_tc2_dgrad(h, net, full, src, d_raw, act, grad, scratch, d_t, st))) return rc;
  mark(1); mark(-1);
  rc = mlp_tc2_wgrad(h, full, M, d_raw, act, grad, G, st);
  mark(2);
  return rc;
}

// The fine net's half of flat_grad is final here (its wgrad and heads kernels are the last writers): data-parallel callers
// may start reducing it while the coarse net's backward runs (nerf_stream_wait_fine_grads).
static int mark_fine_grads(NerfEngine* h, cudaStream_t st) {
  if (!h->fine_grad_ev) NERF_CUDA_TRY(h, cudaEventCreateWithFlags(&h->fine_grad_ev, cudaEventDisableTiming));
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(st, &cs);
  NERF_CUDA_TRY(h, cudaEventRecordWithFlags(h->fine_grad_ev, st, cs == cudaStreamCaptureStatusActive ? cudaEventRecordExternal : cudaEventRecordDefault));
  h->fine_grad_recorded = true;
  return NERF_OK;
}

// Backward of the last train_forward.  fused_loss != NULL (nerf_forward_backward): the final pass's composite, the MSE loss
// and its composite backward are one kernel (d rgb = 2 (rgb - gt) grad_scale / 3n computed per ray), and the coarse composite
// backward runs inside the sample_pdf backward kernel; otherwise d_rgb / d_depth are the caller's map gradients.
struct FusedLoss { const float* gt; float scale; float* loss; float* rgb_out; };
static int train_backward(NerfEngine* h, const float* rays_o, const float* rays_d, int64_t n, const float* d_rgb,
                          const float* d_depth, const FusedLoss* fl, float* flat_grad, cudaStream_t st) {
  if (h->taps.n != n) return h->fail(NERF_E_STATE, "train_backward: no matching nerf_train_forward");
  const NerfConfig& c = h->cfg;
  const int Nc = c.n_samples, Nf = c.n_importance, Nt = Nc + Nf;
  const NerfEngine::Taps& tp = h->taps;
  int rc;
  RowSrc src;
  src.rays_o = rays_o; src.rays_d = rays_d; src.R = n;
  float* sq = nullptr; unsigned int* done = nullptr;
  if (fl) {
    const bool fresh = h->ws.find("tr.done") == h->ws.end();
    if ((rc = h->buf("tr.sq", (size_t)n, &sq)) || (rc = h->buf("tr.done", 1, &done))) return rc;
    if (fresh) NERF_CUDA_TRY(h, cudaMemsetAsync(done, 0, sizeof(unsigned int), st));
  }
  if (Nf > 0) {
    float *d_raw_f, *d_t_s, *d_raw_c;
    if ((rc = h->buf("tr.d_raw_f", (size_t)n * Nt * 4, &d_raw_f))) return rc;
    if ((rc = h->buf("tr.d_t_s", (size_t)n * Nt, &d_t_s))) return rc;
    if ((rc = h->buf("tr.d_raw_c", (size_t)n * Nc * 4, &d_raw_c))) return rc;
    if (fl) launch_composite_loss_bwd(h, tp.raw_f, tp.t_s, n, Nt, c.white_bkgd, fl->gt, fl->scale, fl->rgb_out, d_raw_f, d_t_s, sq, done, fl->loss, st);
    else launch_composite_bwd(h, tp.raw_f, tp.t_s, n, Nt, c.white_bkgd, d_rgb, d_depth, nullptr, d_raw_f, d_t_s, st);
    src.t = tp.t_s; src.S = Nt;
    if ((rc = mlp_backward(h, NERF_NET_FINE, src, d_raw_f, false, flat_grad + NERF_PARAMS_PER_NET, d_t_s, st))) return rc;
    if ((rc = mark_fine_grads(h, st))) return rc;
    const bool tu = c.train_u != 0;
    if ((rc = launch_sample_pdf_bwd(h, tp.raw_c, tp.t_c, n, Nc, Nf, tu ? tp.u : nullptr, tu && !tp.u, tp.call_id, d_t_s,
                                    tp.sort_idx, nullptr, st, c.white_bkgd, d_raw_c))) return rc;
    src.t = tp.t_c; src.S = Nc;
    if ((rc = mlp_backward(h, NERF_NET_COARSE, src, d_raw_c, true, flat_grad, nullptr, st))) return rc;
  } else {
    float* d_raw_c;
    if ((rc = h->buf("tr.d_raw_c", (size_t)n * Nc * 4, &d_raw_c))) return rc;
    if (fl) launch_composite_loss_bwd(h, tp.raw_c, tp.t_c, n, Nc, c.white_bkgd, fl->gt, fl->scale, fl->rgb_out, d_raw_c, nullptr, sq, done, fl->loss, st);
    else launch_composite_bwd(h, tp.raw_c, tp.t_c, n, Nc, c.white_bkgd, d_rgb, d_depth, nullptr, d_raw_c, nullptr, st);
    if ((rc = mark_fine_grads(h, st))) return rc;     // the fine net is unused: its (zero) half is final from the start
    src.t = tp.t_c; src.S = Nc;
    if ((rc = mlp_backward(h, NERF_NET_COARSE, src, d_raw_c, false, flat_grad, nullptr, st))) return rc;
  }
  return post(h, "train_backward");
}

extern "C" int nerf_stream_wait_fine_grads(NerfEngine* h, void* stream) {
  if (!h) return NERF_E_INVALID;
  if (!h->fine_grad_recorded) return h->fail(NERF_E_STATE, "nerf_stream_wait_fine_grads: no backward pass has been issued on this handle");
  cudaSetDevice(h->device);
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing((cudaStream_t)stream, &cs);
  NERF_CUDA_TRY(h, cudaStreamWaitEvent((cudaStream_t)stream, h->fine_grad_ev, cs == cudaStreamCaptureStatusActive ? cudaEventWaitExternal : cudaEventWaitDefault));
  return NERF_OK;
}

extern "C" int nerf_train_forward(NerfEngine* h, const float* rays_o, const float* rays_d, int64_t n, const float* jitter,
                                  const float* u, float* rgb, float* depth, uint64_t* forward_id, void* stream) {
  if (!h) return NERF_E_INVALID;
  if (!h->params_set) return h->fail(NERF_E_STATE, "nerf_train_forward before nerf_set_params");
  if (!rays_o || !rays_d || !rgb || n <= 0) return h->fail(NERF_E_INVALID, "nerf_train_forward: bad argument");
  cudaSetDevice(h->device);
  const int rc = train_forward(h, rays_o, rays_d, n, jitter, u, rgb, depth, true, (cudaStream_t)stream);
  if (!rc && forward_id) *forward_id = h->taps.call_id;
  return rc;
}

extern "C" int nerf_train_backward(NerfEngine* h, const float* rays_o, const float* rays_d, int64_t n, const float* d_rgb,
                                   const float* d_depth, float* flat_grad, int zero_grad, uint64_t forward_id, void* stream) {
  if (!h) return NERF_E_INVALID;
  if (!rays_o || !rays_d || !flat_grad || n <= 0 || (!d_rgb && !d_depth)) return h->fail(NERF_E_INVALID, "nerf_train_backward: bad argument");
  // the handle keeps the activations of ONE forward: a backward for an older forward would silently use the newer one's
  if (forward_id != 0 && forward_id != h->taps.call_id)
    return h->fail(NERF_E_STATE, "nerf_train_backward: the activations of this forward were overwritten by a later nerf_train_forward "
                                 "(one forward, then its backward; the handle keeps one set of activations)");
  cudaSetDevice(h->device);
  if (zero_grad) NERF_CUDA_TRY(h, cudaMemsetAsync(flat_grad, 0, (size_t)NERF_PARAMS_TOTAL * sizeof(float), (cudaStream_t)stream));
  return train_backward(h, rays_o, rays_d, n, d_rgb, d_depth, nullptr, flat_grad, (cudaStream_t)stream);
}

// The whole step body on one stream (also what a CUDA graph of the step captures).
static int forward_backward_body(NerfEngine* h, const float* rays_o, const float* rays_d, const float* rgb_gt, int64_t n,
                                 const float* jitter, const float* u, float grad_scale, float* loss, float* flat_grad,
                                 float* rgb_out, cudaStream_t st) {
  int rc;
  NERF_CUDA_TRY(h, cudaMemsetAsync(flat_grad, 0, (size_t)NERF_PARAMS_TOTAL * sizeof(float), st));
  if ((rc = train_forward(h, rays_o, rays_d, n, jitter, u, nullptr, nullptr, false, st))) return rc;
  FusedLoss fl{rgb_gt, grad_scale, loss, rgb_out};
  return train_backward(h, rays_o, rays_d, n, nullptr, nullptr, &fl, flat_grad, st);
}

// CUDA graph of the step.  Everything the step launches is a pure function of (device pointers, n, config), so the second
// identical call is captured and later ones replay the graph: ~14 launches become one.  The caller's rays / colours are
// first copied into engine-owned staging buffers (three small device copies outside the graph) so that callers that
// rotate their batch tensors still hit the graph.  Falls back to plain launches whenever the step is not replayable:
// in-kernel Philox randomness (keyed by a per-call counter), profiling brackets, graph mode off.
struct StepKey {
  int64_t n; const void *jitter, *u, *loss, *grad, *rgb_out, *params; float scale; NerfConfig cfg;
  bool operator==(const StepKey& o) const {
    return n == o.n && jitter == o.jitter && u == o.u && loss == o.loss && grad == o.grad && rgb_out == o.rgb_out && params == o.params &&
           scale == o.scale && memcmp(&cfg, &o.cfg, sizeof(cfg)) == 0;
  }
};
struct StepGraph { StepKey key{}; bool warm = false; cudaGraphExec_t exec = nullptr; int64_t launches = 0; };
static std::map<NerfEngine*, StepGraph> g_step_graphs;

static void drop_step_graph(NerfEngine* h) {
  auto it = g_step_graphs.find(h);
  if (it == g_step_graphs.end()) return;
  if (it->second.exec) cudaGraphExecDestroy(it->second.exec);
  g_step_graphs.erase(it);
}

extern "C" int nerf_set_graph_mode(NerfEngine* h, int on) {
  if (!h) return NERF_E_INVALID;
  h->graph_mode = on != 0;
  if (!on) drop_step_graph(h);
  return NERF_OK;
}

extern "C" int nerf_forward_backward(NerfEngine* h, const float* rays_o, const float* rays_d, const float* rgb_gt, int64_t n,
                                     const float* jitter, const float* u, float grad_scale, float* loss, float* flat_grad,
                                     float* rgb_out, void* stream) {
  if (!h) return NERF_E_INVALID;
  if (!h->params_set) return h->fail(NERF_E_STATE, "nerf_forward_backward before nerf_set_params");
  if (!rays_o || !rays_d || !rgb_gt || !loss || !flat_grad || n <= 0) return h->fail(NERF_E_INVALID, "nerf_forward_backward: bad argument");
  cudaSetDevice(h->device);
  cudaStream_t st = (cudaStream_t)stream;
  const NerfConfig& c = h->cfg;
  const bool philox = (c.perturb && !jitter) || (c.train_u && !u && c.n_importance > 0);
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(st, &cs);
  // Stage the inputs in engine-owned buffers (no-op when the caller already uses them, e.g. nerf_bank_train_step).  The
  // copies take the direction from the pointers (unified addressing), so rays_o / rays_d / rgb_gt may also be HOST arrays
  // (pinned for an asynchronous copy): the end-to-end path then needs no device tensors of its own.
  float *so, *sd, *sg;
  int rc;
  if ((rc = h->buf("bank.o", (size_t)n * 3, &so)) || (rc = h->buf("bank.d", (size_t)n * 3, &sd)) || (rc = h->buf("bank.c", (size_t)n * 3, &sg))) return rc;
  if (cs == cudaStreamCaptureStatusNone) {
    const size_t bytes = (size_t)n * 3 * sizeof(float);
    if (rays_o != so) NERF_CUDA_TRY(h, cudaMemcpyAsync(so, rays_o, bytes, cudaMemcpyDefault, st));
    if (rays_d != sd) NERF_CUDA_TRY(h, cudaMemcpyAsync(sd, rays_d, bytes, cudaMemcpyDefault, st));
    if (rgb_gt != sg) NERF_CUDA_TRY(h, cudaMemcpyAsync(sg, rgb_gt, bytes, cudaMemcpyDefault, st));
  } else {
    so = const_cast<float*>(rays_o); sd = const_cast<float*>(rays_d); sg = const_cast<float*>(rgb_gt);   // caller captures: device pointers as given
  }
  if (!h->graph_mode || philox || h->profiling || cs != cudaStreamCaptureStatusNone) {
    return forward_backward_body(h, so, sd, sg, n, jitter, u, grad_scale, loss, flat_grad, rgb_out, st);
  }
  StepKey key{n, jitter, u, loss, flat_grad, rgb_out, h->params, grad_scale, c};
  StepGraph& sg_ = g_step_graphs[h];
  if (sg_.exec && sg_.key == key) {
    NERF_CUDA_TRY(h, cudaGraphLaunch(sg_.exec, st));
    h->launches += sg_.launches;
    h->taps.n = -1;                                   // the split-form backward cannot follow a replayed step
    return NERF_OK;
  }
  if (sg_.exec) { cudaGraphExecDestroy(sg_.exec); sg_.exec = nullptr; }
  if (!(sg_.warm && sg_.key == key)) {
    // first call with this signature: plain launches (sizes every workspace buffer, sets the kernel attributes)
    sg_.key = key; sg_.warm = true;
    return forward_backward_body(h, so, sd, sg, n, jitter, u, grad_scale, loss, flat_grad, rgb_out, st);
  }
  // second identical call: capture, instantiate, launch.  The capture runs on a private stream (the caller's may be the
  // legacy default stream, which cannot be captured); the graph itself is stream-agnostic and is launched on the caller's.
  const int64_t l0 = h->launches;
  if (!h->cap_stream && cudaStreamCreateWithFlags(&h->cap_stream, cudaStreamNonBlocking) != cudaSuccess) {
    cudaGetLastError();
    h->graph_mode = false;
    return forward_backward_body(h, so, sd, sg, n, jitter, u, grad_scale, loss, flat_grad, rgb_out, st);
  }
  cudaError_t e = cudaStreamBeginCapture(h->cap_stream, cudaStreamCaptureModeThreadLocal);
  if (e != cudaSuccess) {
    cudaGetLastError();
    h->graph_mode = false;
    return forward_backward_body(h, so, sd, sg, n, jitter, u, grad_scale, loss, flat_grad, rgb_out, st);
  }
  rc = forward_backward_body(h, so, sd, sg, n, jitter, u, grad_scale, loss, flat_grad, rgb_out, h->cap_stream);
  cudaGraph_t graph = nullptr;
  e = cudaStreamEndCapture(h->cap_stream, &graph);
  if (rc || e != cudaSuccess || !graph) {
    if (graph) cudaGraphDestroy(graph);
    cudaGetLastError();
    sg_.warm = false;
    h->graph_mode = false;                            // do not try again on this handle
    if (rc) return rc;
    return forward_backward_body(h, so, sd, sg, n, jitter, u, grad_scale, loss, flat_grad, rgb_out, st);
  }
  sg_.launches = h->launches - l0;
  e = cudaGraphInstantiate(&sg_.exec, graph, 0);
  cudaGraphDestroy(graph);
  if (e != cudaSuccess) { sg_.exec = nullptr; return h->fail(NERF_E_CUDA, std::string("cudaGraphInstantiate: ") + cudaGetErrorString(e)); }
  NERF_CUDA_TRY(h, cudaGraphLaunch(sg_.exec, st));
  h->taps.n = -1;
  return NERF_OK;
}

extern "C" int nerf_clip_adam(NerfEngine* h, float* flat_param, const float* flat_grad, float* m, float* v, float lr, float beta1,
                              float beta2, float eps, float clip, int step, void* stream) {
  if (!h || !flat_param || !flat_grad || !m || !v || step < 1) return h ? h->fail(NERF_E_INVALID, "nerf_clip_adam: bad argument") : NERF_E_INVALID;
  if ((reinterpret_cast<uintptr_t>(flat_param) | reinterpret_cast<uintptr_t>(flat_grad) | reinterpret_cast<uintptr_t>(m) |
       reinterpret_cast<uintptr_t>(v)) & 15)
    return h->fail(NERF_E_INVALID, "nerf_clip_adam: buffers must be 16-byte aligned");
  cudaSetDevice(h->device);
  if (h->params_set && flat_param == h->params && h->wpack) {
    // the engine's packed bf16 images follow the masters: update + re-pack in one cooperative launch
    if (int rc = launch_clip_adam_pack(h, flat_param, flat_grad, m, v, NERF_PARAMS_TOTAL, lr, beta1, beta2, eps, clip, step, h->wpack,
                                       reinterpret_cast<uint8_t*>(h->wpack) + tc_wpack_bias_offset(), h->wpack_bwd, (cudaStream_t)stream))
      return rc;
    if (int rc = record_pack(h, (cudaStream_t)stream)) return rc;
  } else {
    launch_clip_adam(h, flat_param, flat_grad, m, v, NERF_PARAMS_TOTAL, lr, beta1, beta2, eps, clip, step, (cudaStream_t)stream);
  }
  return post(h, "nerf_clip_adam");
}

// Wait profile of the CTA-pair kernels (tc_common.cuh waitprof_add).  on != 0: (re)start with zeroed counters; on == 0: stop.
// out (may be NULL): 2 x 512 uint64 read before the state change -- [0,512) forward TU, [512,1024) backward TU; each
// [b] = cycles in slow-path waits of tag bucket b = (tag >> 4) & 255, [256 + b] = number of such waits.
static unsigned long long* g_waitprof_buf = nullptr;
// Event trace of the diagnostic build -DNERF_TC_TRACE (tc_common.cuh trace_event): out[0] = number of records, out[1..n) =
// records (clock64 << 16 | tag) of CTA 0 since the last nerf_debug_waitprof(1, ...).  Zeros in any other build.
extern "C" int nerf_debug_trace(unsigned long long* out, int n) {
  if (!out || n < 1 || n > 8192) return NERF_E_INVALID;
  cudaDeviceSynchronize();
  if (!g_waitprof_buf) { memset(out, 0, (size_t)n * sizeof(unsigned long long)); return NERF_OK; }
  return cudaMemcpy(out, g_waitprof_buf + 1024, (size_t)n * sizeof(unsigned long long), cudaMemcpyDeviceToHost) == cudaSuccess ? NERF_OK : NERF_E_CUDA;
}

extern "C" int nerf_debug_waitprof(int on, unsigned long long* out) {
  static unsigned long long* buf = nullptr;
  cudaDeviceSynchronize();
  if (out) {
    if (buf) cudaMemcpy(out, buf, 1024 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    else memset(out, 0, 1024 * sizeof(unsigned long long));
  }
  if (on) {
    if (!buf && cudaMalloc(&buf, (1024 + 8192) * sizeof(unsigned long long)) != cudaSuccess) return NERF_E_CUDA;
    cudaMemset(buf, 0, (1024 + 8192) * sizeof(unsigned long long));
    g_waitprof_buf = buf;
    nerf::tc2_set_debug_sink_waitprof(buf);
    nerf::tc2b_set_debug_sink_waitprof(buf + 512);
  } else {
    nerf::tc2_set_debug_sink_waitprof(nullptr);
    nerf::tc2b_set_debug_sink_waitprof(nullptr);
  }
  return cudaDeviceSynchronize() == cudaSuccess ? NERF_OK : NERF_E_CUDA;
}

// watchdog record of the tcgen05 kernels (valid after a trapped launch); out: 8 uint32
extern "C" int nerf_debug_tc_watchdog(unsigned int* out8) {
  if (g_watchdog_host && g_watchdog_host[0]) { for (int i = 0; i < 64; ++i) out8[i] = g_watchdog_host[i]; return 0; }
  unsigned int b[8];
  const int rc = nerf::tc2_debug_read(b);
  for (int i = 0; i < 8; ++i) out8[i] = b[i];
  return rc;
}
