// engine.h -- the opaque handle behind the C ABI and the launchers each .cu file provides.
#pragma once
#include <cstdlib>
#include <map>
#include <utility>
#include <string>
#include <vector>
#include "common.cuh"

struct NerfEngine {
  NerfConfig cfg{};
  int device = 0;
  int num_sms = 148;
  std::string err;
  int64_t launches = 0;

  const float* params = nullptr;   // caller-owned fp32 masters (flat layout)
  __nv_bfloat16* wpack = nullptr;  // packed bf16 weight images for the tcgen05 kernels
  size_t wpack_bytes = 0;
  __nv_bfloat16* wpack_bwd = nullptr;   // transposed stage images for the dgrad chain (mlp_tc2_bwd.cu)
  // Replicas of the forward stage images (NERF_B200_WREPLICAS = R > 1): CTA pair p streams replica p % R, so that the 74
  // pairs, which walk the layers in lock step, do not all pull the same 256 L2 lines of the current stage at the same time
  __nv_bfloat16* wrep = nullptr;
  int w_replicas = 1;
  // tensor maps (CUtensorMap, 128 bytes each) over the packed stage images for the 2-SM TMA weight loads of mlp_tc2.cu:
  // boxes of 128 / 64 rows; encoded for tmap_base == wpack
  alignas(64) unsigned char tmaps[2][128] = {};
  const void* tmap_base = nullptr;
  bool params_set = false;

  // grow-only named workspace buffers
  struct Buf { void* p = nullptr; size_t bytes = 0; };
  std::map<std::string, Buf> ws;

  // ray banks (bank.cu): every view's image and pose resident in HBM
  struct Bank { void* images = nullptr; double* poses = nullptr; int V = 0, H = 0, W = 0; double focal = 0, cx = 0, cy = 0;
                int storage = 0; std::vector<uint8_t> loaded; };
  std::vector<Bank> banks;

  // taps of the last training forward (nerf_train_forward), consumed by nerf_train_backward
  struct Taps { float *t_c = nullptr, *raw_c = nullptr, *t_s = nullptr, *raw_f = nullptr; int32_t* sort_idx = nullptr;
                uint64_t call_id = 0; int64_t n = -1; const float *jitter = nullptr, *u = nullptr; } taps;

  // Ordering of the packed weights against the engine's private streams: recorded on the caller's stream after every
  // (re)pack (nerf_set_params, nerf_clip_adam); nerf_render_rays_host makes its internal streams wait for it.
  cudaEvent_t pack_ev = nullptr;
  // Data-parallel hook: recorded on the caller's stream when the fine net's half of flat_grad is final, i.e. before the
  // coarse net's backward starts (nerf_stream_wait_fine_grads)
  cudaEvent_t fine_grad_ev = nullptr;
  bool fine_grad_recorded = false;
  cudaStream_t cap_stream = nullptr;        // private stream the step graph is captured on
  bool graph_mode = true;                   // nerf_forward_backward may capture / replay the step as a CUDA graph

  // host-path resources (nerf_render_rays_host)
  cudaStream_t side = nullptr;              // forked work that overlaps a persistent kernel (wgrad heads)
  cudaEvent_t side_fork = nullptr, side_join = nullptr;
  cudaStream_t hs[2] = {nullptr, nullptr};
  cudaEvent_t hev[2] = {nullptr, nullptr};
  void* pinned[2] = {nullptr, nullptr};
  size_t pinned_bytes = 0;

  // optional event bracketing of the fused-MLP launches (nerf_profile_enable / nerf_profile_read)
  bool profiling = false;
  std::vector<cudaEvent_t> prof_pool;      // all events ever created (reused)
  size_t prof_used = 0;                    // events handed out since the last read (pairs: begin, end)
  int64_t prof_rows = 0;
  std::vector<int> prof_class;             // class of each bracketed launch: 0 MLP forward, 1 dgrad chain, 2 wgrad
  std::vector<int64_t> prof_class_rows;
  cudaEvent_t prof_event() {
    if (prof_used == prof_pool.size()) { cudaEvent_t e; cudaEventCreate(&e); prof_pool.push_back(e); }
    return prof_pool[prof_used++];
  }

  int fail(int code, const std::string& m) { err = m; return code; }

  template <typename T>
  int buf(const char* name, size_t count, T** out) {
    Buf& b = ws[name];
    size_t need = count * sizeof(T);
    if (need == 0) need = sizeof(T);
    if (b.bytes < need) {
      if (b.p) {
        // other streams may still use the old block: full sync before freeing (rare, grow-only)
        cudaDeviceSynchronize();
        cudaFree(b.p);
        b.p = nullptr; b.bytes = 0;
      }
      size_t grow = need + need / 8;
      cudaError_t e = cudaMalloc(&b.p, grow);
      if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(NERF_E_NOMEM, std::string("cudaMalloc(") + name + ", " + std::to_string(grow) + " B): " + cudaGetErrorString(e));
      }
      b.bytes = grow;
    }
    *out = reinterpret_cast<T*>(b.p);
    return NERF_OK;
  }
};

namespace nerf {

// Between two persistent tcgen05 kernels (228 KB of shared memory per CTA) a small stage kernel with the default carve-out
// would make every SM switch its L1 / shared-memory split and back; these kernels stream through L2 anyway, so they ask for
// the same (maximum shared) split.  NERF_B200_CARVEOUT=0 keeps the driver default (A/B measurements).
inline void prefer_max_shared(const void* func, int device) {
  static std::map<std::pair<const void*, int>, bool> done;
  static int enabled = -1;
  if (enabled < 0) { const char* v = getenv("NERF_B200_CARVEOUT"); enabled = (v && v[0] == '0') ? 0 : 1; }
  if (!enabled) return;
  bool& d = done[std::make_pair(func, device)];
  if (d) return;
  d = true;
  cudaFuncSetAttribute(func, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
  cudaGetLastError();
}

// ---- stages.cu ---------------------------------------------------------------------------------
struct Camera { float r[9]; float o[3]; double r64[9]; double o64[3]; double focal, cx, cy; int H, W; };
void launch_get_rays(NerfEngine* h, const Camera& cam, const int64_t* pixel_ids, int64_t pix0, int64_t n,
                     float* rays_o, float* rays_d, cudaStream_t st);
void launch_stratified_t(NerfEngine* h, int64_t n, int S, const float* jitter, bool philox, uint64_t call_id,
                         float* t, cudaStream_t st);
void launch_freq_encode(NerfEngine* h, const float* x, int64_t m, int n_freqs, float* out, cudaStream_t st);
void launch_composite(NerfEngine* h, const float* raw, const float* t, int64_t n, int S, int white,
                      float* rgb, float* depth, float* acc, float* weights, cudaStream_t st);
int launch_sample_pdf(NerfEngine* h, const float* raw_c, const float* t_c, int64_t n, int Nc, int Nf,
                      const float* u, bool philox_u, uint64_t call_id, const float* cdf_in, float* t_fine,
                      int64_t* inds, int64_t* below, int64_t* above, float* cdf_out, float* t_sorted,
                      int32_t* sort_idx, cudaStream_t st, float prune_w_thr = 0.25f, uint8_t* valid_fine = nullptr,
                      uint8_t* mask_sorted = nullptr,      // a-M: ESS/ERT mask outputs (either non-NULL turns it on)
                      int32_t* row_list = nullptr, int32_t* row_count = nullptr);   // + compacted list of the valid sorted rows
// backward stages
void launch_mse_grad(NerfEngine* h, const float* rgb, const float* gt, int64_t n, float scale,
                     float* d_rgb, float* loss, cudaStream_t st);
// composite backward: d_rgb_map [n,3] / d_depth_map [n] (or NULL) and/or direct d_w [n,S] (or NULL) -> d_raw [n,S,4] (rgb
// channels zero when d_rgb_map NULL), d_t [n,S] (or NULL; gradient w.r.t. the sample depths).
void launch_composite_bwd(NerfEngine* h, const float* raw, const float* t, int64_t n, int S, int white,
                          const float* d_rgb_map, const float* d_depth_map, const float* d_w, float* d_raw, float* d_t,
                          cudaStream_t st);
// sample_pdf backward: d_t_sorted [n,Nt] (total gradient at sorted positions) -> d_w_coarse [n,Nc] (may be NULL when
// d_raw_c is given); d_raw_c != NULL fuses the coarse composite backward behind it (d w -> d_raw_c [n,Nc,4])
int launch_sample_pdf_bwd(NerfEngine* h, const float* raw_c, const float* t_c, int64_t n, int Nc, int Nf,
                          const float* u, bool philox_u, uint64_t call_id, const float* d_t_sorted,
                          const int32_t* sort_idx, float* d_w_coarse, cudaStream_t st, int white = 0, float* d_raw_c = nullptr);
// fused composite + MSE loss + composite backward of the final pass (training step): see stages.cu
void launch_composite_loss_bwd(NerfEngine* h, const float* raw, const float* t, int64_t n, int S, int white, const float* gt,
                               float scale, float* rgb_out, float* d_raw, float* d_t, float* sq, unsigned int* done, float* loss,
                               cudaStream_t st);

// ---- mlp_fp32.cu -------------------------------------------------------------------------------
// Row source for the MLP kernels: either explicit points or rays + depths.
struct RowSrc {
  const float* pts = nullptr;       // [R*S,3] explicit points (mode B) or NULL
  const float* viewdirs = nullptr;  // [R,3] unit view dirs (mode B) or NULL -> normalise rays_d
  const float* rays_o = nullptr;    // [R,3]
  const float* rays_d = nullptr;    // [R,3]
  const float* t = nullptr;         // [R,S]
  const uint8_t* mask = nullptr;    // [R,S] or NULL
  // optional compacted row list (fast_sampling work skipping, CTA-pair forward only): the kernel evaluates rows
  // row_list[0 .. *row_count) instead of all R*S rows; both live in device memory (no host round trip)
  const int32_t* row_list = nullptr;
  const int32_t* row_count = nullptr;
  int64_t R = 0;
  int S = 0;
};
int mlp_fp32_forward(NerfEngine* h, int net, const RowSrc& src, float* raw, bool keep_acts, cudaStream_t st);
// backward of the last keep_acts forward of `net`: d_raw [M,4] -> flat_grad (+=), d_t [R,S] (+=, may be NULL)
int mlp_fp32_backward(NerfEngine* h, int net, const RowSrc& src, const float* d_raw, bool coarse_only_sigma,
                      float* flat_grad_net, float* d_t, cudaStream_t st);

// ---- pack_weights.cu (bf16 operand images of the tcgen05 forward) ------------------------------------
size_t tc_wpack_bytes();
size_t tc_wpack_bias_offset();   // byte offset of the bias slabs ([net][schedule entry] x 8 KB) inside the packed buffer
size_t tc_bwd_wpack_bytes();
// forward stage images + bias slabs (wpack) and the transposed dgrad stage images (wpack_bwd) in one launch
void tc_pack_all(NerfEngine* h, const float* flat, __nv_bfloat16* wpack, __nv_bfloat16* wpack_bwd, cudaStream_t st);

// ---- mlp_tc2.cu (tcgen05 cta_group::2, CTA pairs; default) ----------------------------------------
int mlp_tc2_forward(NerfEngine* h, int net, const RowSrc& src, float* raw, uint8_t* act_arena, cudaStream_t st);

// ---- mlp_tc2_bwd.cu (tensor-core backward) -----------------------------------------------------------
int mlp_tc2_dgrad(NerfEngine* h, int net, bool full, const RowSrc& src, const float* d_raw, const uint8_t* act_arena,
                  uint8_t* grad_arena, float* pe_scratch, float* d_t, cudaStream_t st);

int mlp_tc2_wgrad(NerfEngine* h, bool full, int64_t M, const float* d_raw, const uint8_t* act_arena, const uint8_t* grad_arena,
                  float* flat_grad_net, cudaStream_t st);

// ---- adam.cu -----------------------------------------------------------------------------------
void launch_clip_adam(NerfEngine* h, float* p, const float* g, float* m, float* v, int64_t n, float lr,
                      float b1, float b2, float eps, float clip, int step, cudaStream_t st);
// the same + the re-pack of both bf16 weight images in one cooperative launch (slabs = wpack + tc_wpack_bias_offset())
int launch_clip_adam_pack(NerfEngine* h, float* p, const float* g, float* m, float* v, int64_t n, float lr, float b1, float b2,
                          float eps, float clip, int step, __nv_bfloat16* wpack, uint8_t* slabs, __nv_bfloat16* wpack_bwd, cudaStream_t st);
void launch_fill(NerfEngine* h, float* p, int64_t n, float v, cudaStream_t st);

}  // namespace nerf
