// adam.cu -- fused gradient clip + Adam step over the flat fp32 parameter buffer.
// Replaces torch.nn.utils.clip_grad_value_(params, 40) (trainer.py:59) followed by torch.optim.Adam.step
// with one param-group per tensor, lr 5e-4, eps 1e-8, weight_decay 0 (train/optimizer.py:8-28): 48 tensors
// x several foreach kernels become one streaming pass: 16 B read + 12 B written per parameter.
#include <cooperative_groups.h>
#include "pack.cuh"

namespace nerf {

// clip_grad_value_ is torch.clamp: a NaN gradient stays NaN (fminf / fmaxf would return the finite bound and hide a
// diverged step in a quietly wrong update)
__device__ __forceinline__ float clamp_keep_nan(float g, float clip) { return g != g ? g : fminf(fmaxf(g, -clip), clip); }

__global__ void clip_adam_kernel(float4* __restrict__ p, const float4* __restrict__ g, float4* __restrict__ m,
                                 float4* __restrict__ v, int64_t n4, float beta1, float beta2, float step_size,
                                 float inv_bc2_sqrt, float eps, float clip) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float4 pp = p[i], gg = __ldg(g + i), mm = m[i], vv = v[i];
    float* pa = &pp.x; float* ga = &gg.x; float* ma = &mm.x; float* va = &vv.x;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float gr = ga[c];
      if (clip > 0.f) gr = clamp_keep_nan(gr, clip);
      ma[c] = ma[c] + (gr - ma[c]) * (1.0f - beta1);              // exp_avg.lerp_(grad, 1-beta1)
      va[c] = va[c] * beta2 + (1.0f - beta2) * gr * gr;           // mul_(beta2).addcmul_(g, g, 1-beta2)
      const float denom = sqrtf(va[c]) * inv_bc2_sqrt + eps;      // sqrt(v)/sqrt(bc2) + eps
      pa[c] = pa[c] - step_size * (ma[c] / denom);                // addcdiv_(m, denom, -lr/bc1)
    }
    p[i] = pp; m[i] = mm; v[i] = vv;
  }
}

// clip + Adam + re-pack in ONE cooperative launch (training step): phase 1 = the update above, grid-wide barrier, phase 2 =
// the bf16 operand images of both tensor-core kernels + the bias slabs from the fresh fp32 masters (pack.cuh).  Replaces
// four launches (clip_adam, pack_weights, pack_bias_slabs, pack_dgrad_weights) whose run time is mostly launch latency.
__global__ void __launch_bounds__(256) clip_adam_pack_kernel(float4* __restrict__ p, const float4* __restrict__ g, float4* __restrict__ m,
                                                             float4* __restrict__ v, int64_t n4, float beta1, float beta2, float step_size,
                                                             float inv_bc2_sqrt, float eps, float clip, __nv_bfloat16* __restrict__ wpack,
                                                             uint8_t* __restrict__ slabs, __nv_bfloat16* __restrict__ wpack_bwd,
                                                             __nv_bfloat16* __restrict__ rep, int n_rep) {
  const int64_t tid0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x, stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = tid0; i < n4; i += stride) {
    float4 pp = p[i], gg = __ldg(g + i), mm = m[i], vv = v[i];
    float* pa = &pp.x; float* ga = &gg.x; float* ma = &mm.x; float* va = &vv.x;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float gr = ga[c];
      if (clip > 0.f) gr = clamp_keep_nan(gr, clip);
      ma[c] = ma[c] + (gr - ma[c]) * (1.0f - beta1);
      va[c] = va[c] * beta2 + (1.0f - beta2) * gr * gr;
      const float denom = sqrtf(va[c]) * inv_bc2_sqrt + eps;
      pa[c] = pa[c] - step_size * (ma[c] / denom);
    }
    p[i] = pp; m[i] = mm; v[i] = vv;
  }
  __threadfence();
  cooperative_groups::this_grid().sync();
  const float* flat = reinterpret_cast<const float*>(p);
  const int64_t total = fwdpack::kWeightChunks + bw::kDgradChunks + fwdpack::kBiasRows;
  for (int64_t e = tid0; e < total; e += stride) {
    if (e < fwdpack::kWeightChunks) fwdpack::pack_weight_chunk(e, flat, wpack, rep, n_rep);
    else if (e < fwdpack::kWeightChunks + bw::kDgradChunks) bw::pack_dgrad_chunk(e - fwdpack::kWeightChunks, flat, wpack_bwd);
    else fwdpack::pack_bias_row((int)(e - fwdpack::kWeightChunks - bw::kDgradChunks), flat, slabs);
  }
}

struct AdamScalars { float step_size, inv_bc2_sqrt; };
static AdamScalars adam_scalars(float lr, float b1, float b2, int step) {
  const double bc1 = 1.0 - pow((double)b1, (double)step);
  const double bc2 = 1.0 - pow((double)b2, (double)step);
  return {(float)((double)lr / bc1), (float)(1.0 / sqrt(bc2))};
}

int launch_clip_adam_pack(NerfEngine* h, float* p, const float* g, float* m, float* v, int64_t n, float lr, float b1, float b2,
                          float eps, float clip, int step, __nv_bfloat16* wpack, uint8_t* slabs, __nv_bfloat16* wpack_bwd, cudaStream_t st) {
  static int blocks_per_sm_dev[64] = {};
  int& bps = blocks_per_sm_dev[h->device & 63];
  if (bps == 0) {
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, clip_adam_pack_kernel, 256, 0) != cudaSuccess || bps < 1)
      return h->fail(NERF_E_CUDA, "clip_adam_pack: occupancy query failed");
  }
  const AdamScalars sc = adam_scalars(lr, b1, b2, step);
  float4* p4 = reinterpret_cast<float4*>(p); const float4* g4 = reinterpret_cast<const float4*>(g);
  float4* m4 = reinterpret_cast<float4*>(m); float4* v4 = reinterpret_cast<float4*>(v);
  int64_t n4 = n / 4;
  float step_size = sc.step_size, inv = sc.inv_bc2_sqrt;
  __nv_bfloat16* rep = h->wrep;
  int n_rep = h->wrep ? h->w_replicas : 0;
  void* args[] = {&p4, &g4, &m4, &v4, &n4, &b1, &b2, &step_size, &inv, &eps, &clip, &wpack, &slabs, &wpack_bwd, &rep, &n_rep};
  const int grid = h->num_sms * std::min(bps, 8);     // every CTA resident: the grid barrier needs it
  prefer_max_shared(reinterpret_cast<const void*>(clip_adam_pack_kernel), h->device);
  cudaError_t e = cudaLaunchCooperativeKernel(reinterpret_cast<void*>(clip_adam_pack_kernel), dim3(grid), dim3(256), args, 0, st);
  if (e != cudaSuccess) return h->fail(NERF_E_CUDA, std::string("clip_adam_pack launch: ") + cudaGetErrorString(e));
  h->launches++;
  return NERF_OK;
}

void launch_clip_adam(NerfEngine* h, float* p, const float* g, float* m, float* v, int64_t n, float lr, float b1,
                      float b2, float eps, float clip, int step, cudaStream_t st) {
  const AdamScalars sc = adam_scalars(lr, b1, b2, step);
  const float step_size = sc.step_size, inv_bc2_sqrt = sc.inv_bc2_sqrt;
  const int64_t n4 = n / 4;   // NERF_PARAMS_TOTAL is a multiple of 4
  int blocks = (int)std::min<int64_t>((n4 + 255) / 256, (int64_t)h->num_sms * 8);
  clip_adam_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<float4*>(p), reinterpret_cast<const float4*>(g),
                                           reinterpret_cast<float4*>(m), reinterpret_cast<float4*>(v), n4, b1, b2,
                                           step_size, inv_bc2_sqrt, eps, clip);
  h->launches++;
}

__global__ void fill_kernel(float* __restrict__ p, int64_t n, float v) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}
void launch_fill(NerfEngine* h, float* p, int64_t n, float v, cudaStream_t st) {
  if (n <= 0) return;
  int blocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)h->num_sms * 8);
  fill_kernel<<<blocks, 256, 0, st>>>(p, n, v);
  h->launches++;
}

}  // namespace nerf
