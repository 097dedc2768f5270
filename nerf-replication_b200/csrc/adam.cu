// adam.cu -- fused gradient clip + Adam step over the flat fp32 parameter buffer.
// Replaces torch.nn.utils.clip_grad_value_(params, 40) (trainer.py:59) followed by torch.optim.Adam.step
// with one param-group per tensor, lr 5e-4, eps 1e-8, weight_decay 0 (train/optimizer.py:8-28): 48 tensors
// x several foreach kernels become one streaming pass: 16 B read + 12 B written per parameter.
#include "engine.h"

namespace nerf {

__global__ void clip_adam_kernel(float4* __restrict__ p, const float4* __restrict__ g, float4* __restrict__ m,
                                 float4* __restrict__ v, int64_t n4, float beta1, float beta2, float step_size,
                                 float inv_bc2_sqrt, float eps, float clip) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float4 pp = p[i], gg = __ldg(g + i), mm = m[i], vv = v[i];
    float* pa = &pp.x; float* ga = &gg.x; float* ma = &mm.x; float* va = &vv.x;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float gr = ga[c];
      if (clip > 0.f) gr = fminf(fmaxf(gr, -clip), clip);
      ma[c] = ma[c] + (gr - ma[c]) * (1.0f - beta1);              // exp_avg.lerp_(grad, 1-beta1)
      va[c] = va[c] * beta2 + (1.0f - beta2) * gr * gr;           // mul_(beta2).addcmul_(g, g, 1-beta2)
      const float denom = sqrtf(va[c]) * inv_bc2_sqrt + eps;      // sqrt(v)/sqrt(bc2) + eps
      pa[c] = pa[c] - step_size * (ma[c] / denom);                // addcdiv_(m, denom, -lr/bc1)
    }
    p[i] = pp; m[i] = mm; v[i] = vv;
  }
}

void launch_clip_adam(NerfEngine* h, float* p, const float* g, float* m, float* v, int64_t n, float lr, float b1,
                      float b2, float eps, float clip, int step, cudaStream_t st) {
  const double bc1 = 1.0 - pow((double)b1, (double)step);
  const double bc2 = 1.0 - pow((double)b2, (double)step);
  const float step_size = (float)((double)lr / bc1);
  const float inv_bc2_sqrt = (float)(1.0 / sqrt(bc2));
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
