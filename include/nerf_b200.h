/*
 * nerf_b200.h -- C ABI of libnerf_b200.so, the B200-native (sm_100a) NeRF ray-render / train engine.
 *
 * Drop-in boundary for rkin100g/Nerf-Replication's hot path.  The reference has no native ABI on this
 * path (its only native code is the unused pybind hash encoder,
 * src/models/encoding/hashencoder/src/bindings.cpp:5-8), so every entry point below cites the Python
 * interface it replaces.  File:line citations are relative to the reference root.
 *
 * Conventions
 *   - plain C: opaque handle, raw pointers, sizes; no torch / C++ types.
 *   - every pointer is a DEVICE pointer on the handle's device unless the name ends in `_host`.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); every launch of a call
 *     goes to that stream; calls are asynchronous unless documented otherwise (the *_host calls sync).
 *   - the caller owns inputs/outputs; the engine owns its handle, workspace and packed weights.
 *   - return value: 0 = NERF_OK, negative = NERF_E_*; nerf_last_error() has the message.  Unsupported
 *     configurations are errors -- there is no CPU or library fallback anywhere.
 *   - a handle is bound to one device, is not thread-safe, and is meant for one process per GPU.
 *
 * Flat parameter layout ("flat", NERF_PARAMS_TOTAL floats): the reference's Network.state_dict() order
 * (network.py:22-47,144-161): for net in {model, model_fine}: pts_linears.0..7, views_linears.0,
 * feature_linear, alpha_linear, rgb_linear; each as weight[out][in] (row-major, PyTorch layout) followed by
 * bias[out].  Gradients, Adam moments use the same layout.
 */
#ifndef NERF_B200_H_
#define NERF_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NERF_OK 0
#define NERF_E_INVALID (-1)     /* bad argument / unsupported configuration            */
#define NERF_E_CUDA (-2)        /* a CUDA runtime call failed                          */
#define NERF_E_NOMEM (-3)       /* workspace allocation failed                         */
#define NERF_E_STATE (-4)       /* call order (e.g. render before nerf_set_params)     */

#define NERF_PARAMS_PER_NET 595844
#define NERF_PARAMS_TOTAL 1191688
#define NERF_NUM_LAYERS 12      /* per net, in state_dict order                         */

#define NERF_PRECISION_BF16 0   /* tcgen05 bf16 operands, fp32 accumulate (fast path)  */
#define NERF_PRECISION_FP32 1   /* fp32 FFMA validation mode                           */

#define NERF_NET_COARSE 0       /* Network.model       (network.py:144)                */
#define NERF_NET_FINE 1         /* Network.model_fine  (network.py:154)                */

typedef struct NerfEngine NerfEngine;

/* Knobs of Renderer.__init__ (volume_renderer.py:9-24) + near/far (:27). */
typedef struct NerfConfig {
  int32_t n_samples;     /* cfg.N_samples      (64)                                   */
  int32_t n_importance;  /* cfg.N_importance   (128); 0 = coarse only                 */
  int32_t white_bkgd;    /* cfg.white_bkgd     (1)                                    */
  int32_t perturb;       /* cfg.perturb && task=="train": jittered coarse t           */
  int32_t train_u;       /* task=="train": u ~ U[0,1) instead of linspace (:143-147)  */
  int32_t precision;     /* NERF_PRECISION_*                                          */
  float t_near, t_far;   /* 2.0, 6.0                                                  */
  uint64_t seed;         /* Philox seed for in-kernel jitter/u when no arrays given   */
  int32_t fast_sampling; /* cfg.fast_sampling  (0): ESS/ERT pruning mask on the fine pass (:116-244, :359-374);
                          * 1 = mask like the reference (masked samples evaluated, raw zeroed), 2 = same results but the
                          * masked samples are skipped (compacted row list; bf16 render path -- fp32 mode and the training calls mask)   */
  float weights_threshold; /* cfg.weights_threshold (0.25): ESS "empty bin" weight threshold (:132-133)        */
} NerfConfig;

/* Optional outputs / taps of a render call; any pointer may be NULL. */
typedef struct NerfRenderOut {
  float* rgb;            /* [n,3]  final-pass rgb_map   (volume_renderer.py:425,432)  */
  float* depth;          /* [n]    final-pass depth_map (:426)                        */
  float* acc;            /* [n]    sum of weights (:430)                              */
  float* rgb_coarse;     /* [n,3]  same maps for the coarse pass (extension)          */
  float* depth_coarse;   /* [n]                                                       */
  float* acc_coarse;     /* [n]                                                       */
} NerfRenderOut;

/* ---- lifecycle ------------------------------------------------------------------------------- */
int nerf_create(const NerfConfig* cfg, int device, NerfEngine** out);
int nerf_destroy(NerfEngine* h);
const char* nerf_last_error(const NerfEngine* h);        /* h may be NULL: last create() error   */
int nerf_set_config(NerfEngine* h, const NerfConfig* cfg);
/* Network parameters (replaces nn.Module ownership, network.py:126-161 / net_utils.py:346-379).
 * `flat` stays owned by the caller and must stay alive: the engine reads fp32 masters from it and keeps a
 * packed bf16 copy that nerf_set_params / nerf_clip_adam refresh. */
int nerf_set_params(NerfEngine* h, const float* flat, void* stream);
int nerf_version(void);
/* Number of engine kernels launched through this handle since creation (bench.py's gpu_launches). */
int64_t nerf_launch_count(const NerfEngine* h);

/* Kernel timing for bench.py's roofline line: while enabled, CUDA events bracket every fused-MLP launch on
 * the stream it is launched on.  nerf_profile_read synchronises the device, returns the summed duration
 * (ms), the number of bracketed launches and the MLP rows (sample evaluations) they processed since the
 * last read, and resets the counters. */
int nerf_profile_enable(NerfEngine* h, int on);
int nerf_profile_read(NerfEngine* h, double* mlp_ms, int64_t* mlp_launches, int64_t* mlp_rows);
/* same, per kernel class (arrays of 3): 0 = fused MLP forward, 1 = dgrad chain, 2 = wgrad (+ heads). */
int nerf_profile_read_classes(NerfEngine* h, double* ms3, int64_t* launches3, int64_t* rows3);

/* ---- stage kernels (parity taps; each is also used by the fused calls) ------------------------- */
/* a-B get_rays, src/datasets/nerf/blender.py:95-127.  c2w_host: 12 doubles (3x4 row-major).  pixel_ids
 * (device int64[n]) or NULL for the pixel range [pix0, pix0+n) in row-major order. */
int nerf_get_rays(NerfEngine* h, const double* c2w_host, int H, int W, double focal, double cx, double cy,
                  const int64_t* pixel_ids, int64_t pix0, int64_t n, float* rays_o, float* rays_d, void* stream);
/* a-C stratified t, volume_renderer.py:27-65.  jitter [n,S] or NULL (perturb=0 -> linspace). */
int nerf_stratified_t(NerfEngine* h, int64_t n, int S, const float* jitter, float* t, void* stream);
/* a-D Encoder.embed, src/models/encoding/freq.py:7-32.  x [m,3] -> out [m, 3+6*n_freqs] fp32. */
int nerf_freq_encode(NerfEngine* h, const float* x, int64_t m, int n_freqs, float* out, void* stream);
/* a-E/a-F Network.forward, network.py:199-258.  pts [R,S,3], viewdirs [R,3], mask [R,S] uint8 or NULL,
 * raw [R,S,4] = (r,g,b,sigma) pre-activation; masked-out points get raw = 0 (:235). */
int nerf_mlp_forward(NerfEngine* h, int which_net, const float* pts, const float* viewdirs,
                     const uint8_t* mask, int64_t R, int S, float* raw, void* stream);
/* a-G + a-J weights_computation + composite, volume_renderer.py:67-96,414-430.
 * raw [n,S,4], t [n,S] -> maps (any NULL skipped), weights [n,S] or NULL. */
int nerf_composite(NerfEngine* h, const float* raw, const float* t, int64_t n, int S, int white_bkgd,
                   float* rgb, float* depth, float* acc, float* weights, void* stream);
/* a-H + a-I fine_sample_points + merge, volume_renderer.py:98-272,349-356.
 * raw_c [n,Nc,4] (sigma = relu(raw[...,3])), t_c [n,Nc]; u [n,Nf] or NULL (linspace);
 * cdf_in [n,Nc-1] or NULL (inject the oracle's cdf for bit-exact index parity).
 * Outputs (any may be NULL): t_fine [n,Nf], inds/below/above int64 [n,Nf], cdf_out [n,Nc-1],
 * t_sorted [n,Nc+Nf], sort_idx int32 [n,Nc+Nf] (index into cat[t_c,t_fine]). */
int nerf_sample_pdf(NerfEngine* h, const float* raw_c, const float* t_c, int64_t n, int Nc, int Nf,
                    const float* u, const float* cdf_in, float* t_fine, int64_t* inds, int64_t* below,
                    int64_t* above, float* cdf_out, float* t_sorted, int32_t* sort_idx, void* stream);
/* a-M the same with the reference's fast_sampling=True mask (volume_renderer.py:116-133,158-192,359-370):
 * valid_fine [n,Nf] uint8 = fine_sample_points' valid_mask; mask_sorted [n,Nc+Nf] uint8 = that mask merged with
 * the all-valid coarse samples and permuted like t_sorted (what Network.forward receives).  Either may be NULL. */
int nerf_sample_pdf_prune(NerfEngine* h, const float* raw_c, const float* t_c, int64_t n, int Nc, int Nf, const float* u,
                          const float* cdf_in, float weights_threshold, float* t_fine, uint8_t* valid_fine,
                          uint8_t* mask_sorted, float* t_sorted, int32_t* sort_idx, void* stream);

/* ---- the hot call: Renderer.render, volume_renderer.py:290-432 ------------------------------------ */
int nerf_render_rays(NerfEngine* h, const float* rays_o, const float* rays_d, int64_t n,
                     const float* jitter, const float* u, const NerfRenderOut* out, void* stream);
/* camera mode: get_rays fused in front (pixels [pix0, pix0+n) of an HxW view). */
int nerf_render_camera(NerfEngine* h, const double* c2w_host, int H, int W, double focal, double cx,
                       double cy, int64_t pix0, int64_t n, const NerfRenderOut* out, void* stream);
/* Host-buffer variant (the e2e path): pageable or pinned host arrays in, host maps out; H2D/D2H are
 * chunked and overlapped with compute on internal streams; synchronous on return. */
int nerf_render_rays_host(NerfEngine* h, const float* rays_o_host, const float* rays_d_host, int64_t n,
                          float* rgb_host, float* depth_host);

/* ---- training: trainers/nerf.py:27-33 (MSE), trainer.py:56-60, train/optimizer.py:8-28 ------------- */
/* loss = mean((render(rays)-rgb_gt)^2); writes d loss/d params * grad_scale into flat_grad (overwritten,
 * not accumulated), loss into *loss (device float).  Follows the reference autograd graph incl. the
 * un-detached t_fine path.  rays_o / rays_d / rgb_gt [n,3] may be device OR host arrays (they are
 * copied into engine-owned staging buffers first; pinned host memory keeps the copy asynchronous); all other pointers
 * are device pointers. */
int nerf_forward_backward(NerfEngine* h, const float* rays_o, const float* rays_d, const float* rgb_gt,
                          int64_t n, const float* jitter, const float* u, float grad_scale,
                          float* loss, float* flat_grad, float* rgb_out, void* stream);
/* Split form used by the autograd drop-in (Renderer.render under torch.enable_grad()): the forward keeps
 * its activations inside the handle until the matching backward.  d_rgb [n,3] / d_depth [n] are the incoming
 * gradients of the maps (either may be NULL); gradients are accumulated into flat_grad (zeroed first when
 * zero_grad != 0).  The handle keeps the activations of ONE forward: *forward_id (host, may be NULL) identifies it, and
 * nerf_train_backward fails with NERF_E_STATE when its forward_id (0 = do not check) is no longer the latest forward. */
int nerf_train_forward(NerfEngine* h, const float* rays_o, const float* rays_d, int64_t n, const float* jitter,
                       const float* u, float* rgb, float* depth, uint64_t* forward_id, void* stream);
int nerf_train_backward(NerfEngine* h, const float* rays_o, const float* rays_d, int64_t n, const float* d_rgb,
                        const float* d_depth, float* flat_grad, int zero_grad, uint64_t forward_id, void* stream);
/* Data-parallel hook (DistributedDataParallel's overlapped gradient reduction, src/train/trainers/trainer.py:16-21): the
 * fine net's half of flat_grad ([NERF_PARAMS_PER_NET, NERF_PARAMS_TOTAL)) is final before the coarse net's backward starts.
 * Makes `stream` wait for that point of the last nerf_forward_backward / nerf_train_backward / nerf_bank_train_step issued
 * on this handle, so that a collective on that half can overlap the rest of the backward pass. */
int nerf_stream_wait_fine_grads(NerfEngine* h, void* stream);
/* nerf_forward_backward / nerf_bank_train_step replay their launches as one CUDA graph from the third identical call on
 * (default on; off = plain stream launches, e.g. under a per-kernel profiler). */
int nerf_set_graph_mode(NerfEngine* h, int on);
/* g = clamp(g,-clip,clip) (clip<=0: off); torch.optim.Adam step `step` (1-based), weight_decay 0;
 * updates flat_param/m/v in place and refreshes the engine's packed bf16 weights. */
int nerf_clip_adam(NerfEngine* h, float* flat_param, const float* flat_grad, float* m, float* v,
                   float lr, float beta1, float beta2, float eps, float clip, int step, void* stream);

/* ---- either side of the path, on the device -------------------------------- */
/* Ray bank = Dataset of src/datasets/nerf/blender.py:8-170 resident in HBM: all views' images (fp32 0..255 as the
 * reference keeps them after cv2.resize, :50-52; storage 1 = uint8 when the pixel values are integral) and poses
 * (transform_matrix rows 0..2, 12 doubles).  focal/cx/cy as blender.py:102-103.  Several banks (train / val / test
 * splits) may live in one handle; nerf_destroy frees them all. */
int nerf_bank_create(NerfEngine* h, int n_views, int H, int W, double focal, double cx, double cy, int storage,
                     int* bank_out);
int nerf_bank_destroy(NerfEngine* h, int bank);
/* image_host: [H,W,3] fp32 (storage 0) or uint8 (storage 1); c2w_host: 12 doubles.  Synchronous on return. */
int nerf_bank_set_view(NerfEngine* h, int bank, int view, const void* image_host, const double* c2w_host, void* stream);
/* Dataset.__getitem__ (blender.py:84-170) for n pixels of `view`: rays_o/rays_d [n,3] (float64 camera math rounded
 * to fp32), colors [n,3] = image[v,u]/255, ids_out [n] int64 (any output may be NULL).  mode 0: pixel_ids (device
 * int64[n], e.g. the caller's np.random.choice, :108-110); mode 1: the row-major range [pix0, pix0+n) (val/test
 * branch, :112-115); mode 2: n distinct pixels drawn on the device -- ids[i] = P(i) for a bijection P of [0,H*W) keyed
 * by (NerfConfig.seed, draw). */
int nerf_bank_sample(NerfEngine* h, int bank, int view, int mode, const int64_t* pixel_ids, int64_t pix0, uint64_t draw,
                     int64_t n, float* rays_o, float* rays_d, float* colors, int64_t* ids_out, void* stream);
/* the same bijection on the host (checker for mode 2; no CUDA call) */
int nerf_bank_permutation_host(uint64_t seed, uint64_t draw, int64_t n_pix, int64_t n, int64_t* ids_out);
/* nerf_bank_sample + nerf_forward_backward in one call: a training step whose only host input is (view, draw). */
int nerf_bank_train_step(NerfEngine* h, int bank, int view, int mode, const int64_t* pixel_ids, int64_t pix0, uint64_t draw,
                         int64_t n, const float* jitter, const float* u, float grad_scale, float* loss, float* flat_grad,
                         float* rgb_out, void* stream);
/* Evaluator sums, src/evaluators/nerf.py:23-30,96-103, over n_elems floats of rgb_pred / rgb_gt:
 * *sse_out (device double) = sum of fp32 (clip(pred,0,1)-clip(gt,0,1))^2 -> mse = sse/n_elems (:101);
 * *sq8_out (device uint64) = sum of psnr_metric's uint8 terms, ((p8-g8) mod 256)^2 mod 256 with x8 = trunc(clip(x)*255)
 * (numpy uint8 arithmetic wraps, :24-26) -> mse8 = sq8/n_elems, psnr = 100 if mse8 < 1e-10 else 10 log10(255^2/mse8). */
int nerf_eval_image(NerfEngine* h, const float* rgb_pred, const float* rgb_gt, int64_t n_elems, double* sse_out,
                    unsigned long long* sq8_out, void* stream);
/* nerf_eval_image without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same
 * arguments, same kernel, same grid and therefore the same bits as nerf_eval_image; the scratch of the grid reduction
 * (per-block partial sums + a counter) is a file-scope device array of NERF_EVAL_STATIC_SCRATCH_BYTES that every GPU loading
 * the library holds until exit, instead of the engine's cudaMalloc'ed workspace.  The call performs no cudaMalloc / cudaFree
 * (the first call on a device looks the array's address up with cudaGetSymbolAddress: make that call outside a capture).
 * Limits: n_elems <= NERF_EVAL_STATIC_MAX_ELEMS and a grid of at most NERF_EVAL_STATIC_MAX_BLOCKS blocks (4 per SM: devices
 * up to 256 SMs); beyond either the call returns NERF_E_INVALID and nerf_last_error() says which.  All calls on a device
 * share the array (it keeps the last call's contents; the counter is cleared on `stream` before each launch): calls that
 * overlap on the GPU -- different streams, or different engines on one device -- race; serialise them. */
#define NERF_EVAL_STATIC_MAX_ELEMS (1 << 27)      /* floats per image: 134 217 728 (e.g. 6688 x 6688 x 3)      */
#define NERF_EVAL_STATIC_MAX_BLOCKS 1024
#define NERF_EVAL_STATIC_SCRATCH_BYTES 16640      /* 1024 x (8 + 8) B of partials + a 256 B counter line      */
int nerf_eval_image_static(NerfEngine* h, const float* rgb_pred, const float* rgb_gt, int64_t n_elems, double* sse_out,
                           unsigned long long* sq8_out, void* stream);
/* ssim_metric, src/evaluators/nerf.py:52-77 = skimage.metrics.structural_similarity(uint8 pred, uint8 gt, win_size,
 * channel_axis=2) on x8 = trunc(clip(x,0,1)*255) of [H,W,3] images: *ssim_sum_out (device double) = sum of the SSIM map
 * over its border-cropped (H-win+1)x(W-win+1) region and the 3 channels; ssim = sum / (3 (H-win+1)(W-win+1)). */
int nerf_eval_ssim(NerfEngine* h, const float* rgb_pred, const float* rgb_gt, int H, int W, int win_size, double* ssim_sum_out,
                   void* stream);

/* ---- debug ------------------------------------------------------------------------------------------
 * (the tcgen05 descriptor probe battery of the bring-up lives in its own library: include/nerf_b200_probe.h) */
/* watchdog record of the tcgen05 kernels: a barrier that never completes traps the kernel instead of
 * hanging the GPU and leaves {tag, blockIdx, threadIdx, parity, ...} here.  out8: 64 uint32 (host): [0,8) first record,
 * [8,64) one word per stuck waiter (tag | warp << 12 | block << 16 | parity << 31). */
int nerf_debug_tc_watchdog(unsigned int* out8);
/* wait profile of the CTA-pair kernels (diagnostic; off by default).  on != 0 starts it with zeroed counters, 0 stops it.
 * out (host, may be NULL) receives the counters accumulated so far: 2 x 512 uint64, [0,512) forward kernel,
 * [512,1024) dgrad/wgrad kernels; within each, [b] = cycles warps spent waiting on barriers tagged b, [256+b] = waits. */
int nerf_debug_waitprof(int on, unsigned long long* out);
/* event trace of the diagnostic build -DNERF_TC_TRACE (`build.py --trace`): out[0] = number of records, out[1..n) =
 * (clock64 << 16 | tag) records of CTA 0 since the last nerf_debug_waitprof(1, ...); all zero in any other build.  n <= 8192. */
int nerf_debug_trace(unsigned long long* out, int n);

#ifdef __cplusplus
}
#endif
#endif /* NERF_B200_H_ */
