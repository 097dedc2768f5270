// mlp_fp32.cu -- fp32 FFMA "validation mode" of the NeRF MLP (network.py:9-74,199-258) forward and
// backward.  Layer-by-layer SIMT GEMMs with activations in HBM: this is the accuracy reference of the
// engine (north_star: 1e-5 gate) and the only mode whose backward keeps fp32 activations; the fast path
// is the fused tcgen05 kernel in mlp_tc.cu.
#include "engine.h"
#include "rows.cuh"

namespace nerf {

// writes PE(pts) into s5[:, 0:63] (ld 320) and PE(dir) into v[:, 256:283] (ld 288)
__global__ void pe_rows_kernel(RowSrc src, int64_t row0, int64_t m, float* __restrict__ s5, float* __restrict__ v) {
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < m; r += (int64_t)gridDim.x * blockDim.x) {
    float p[3], vd[3];
    row_point(src, row0 + r, p, vd);
    float* a = s5 + r * 320;
    float* b = v + r * 288 + 256;
#pragma unroll
    for (int c = 0; c < 3; ++c) { a[c] = p[c]; b[c] = vd[c]; }
    float f = 1.0f;
    for (int k = 0; k < 10; ++k, f *= 2.0f) {
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        float sn, cs;
        sincosf(p[c] * f, &sn, &cs);
        a[3 + 6 * k + c] = sn; a[3 + 6 * k + 3 + c] = cs;
      }
    }
    f = 1.0f;
    for (int k = 0; k < 4; ++k, f *= 2.0f) {
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        float sn, cs;
        sincosf(vd[c] * f, &sn, &cs);
        b[3 + 6 * k + c] = sn; b[3 + 6 * k + 3 + c] = cs;
      }
    }
  }
}

// d t[g] += d . dpts, dpts_c = dPE[c] + sum_k 2^k (cos(2^k x_c) dPE_sin[k,c] - sin(2^k x_c) dPE_cos[k,c])
__global__ void pe_bwd_kernel(RowSrc src, int64_t m, const float* __restrict__ dpe /*[m,64]*/, float* __restrict__ d_t) {
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < m; r += (int64_t)gridDim.x * blockDim.x) {
    float p[3], vd[3];
    row_point(src, r, p, vd);
    const float* g = dpe + r * 64;
    const int64_t ray = r / src.S;
    float acc = 0.f;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float dp = g[c];
      float f = 1.0f;
      for (int k = 0; k < 10; ++k, f *= 2.0f) {
        float sn, cs;
        sincosf(p[c] * f, &sn, &cs);
        dp += f * (cs * g[3 + 6 * k + c] - sn * g[3 + 6 * k + 3 + c]);
      }
      acc += dp * __ldg(src.rays_d + 3 * ray + c);
    }
    d_t[r] += acc;
  }
}

__global__ void zero_masked_rows_kernel(const uint8_t* __restrict__ mask, int64_t m, float4* __restrict__ raw) {
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < m; r += (int64_t)gridDim.x * blockDim.x)
    if (!mask[r]) raw[r] = make_float4(0.f, 0.f, 0.f, 0.f);
}

// ---- SIMT GEMM: C[M,N] = epi(A[M,K] * op(B)) -----------------------------------------------------
// TRANS_B = true : B is W[N,K] row-major (y = x W^T, forward)       B(k,n) = W[n*ldb + k]
// TRANS_B = false: B is W[K,N] row-major (dx = dy W, dgrad)         B(k,n) = W[k*ldb + n]
struct GemmEpi {
  const float* bias = nullptr;   // [N]
  int relu = 0;
  int accumulate = 0;            // C += result (before mask)
  const float* mask_src = nullptr;  // multiply by (mask_src[m*ld_mask + n] > 0)
  int ld_mask = 0;
};

constexpr int BM = 64, BN = 64, BK = 16;

template <bool TRANS_B>
__global__ void __launch_bounds__(256) gemm_kernel(const float* __restrict__ A, int lda, const float* __restrict__ B,
                                                   int ldb, float* __restrict__ C, int ldc, int64_t M, int N, int K,
                                                   GemmEpi epi) {
  __shared__ __align__(16) float As[BK][BM + 4];
  __shared__ __align__(16) float Bs[BK][BN + 4];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int64_t m0 = (int64_t)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += BK) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {          // A tile: 64 rows x 16 k; consecutive threads along k
      const int e = tid + i * 256;
      const int r = e >> 4, kk = e & 15;
      const int64_t gm = m0 + r;
      const int gk = k0 + kk;
      As[kk][r] = (gm < M && gk < K) ? __ldg(A + gm * lda + gk) : 0.f;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int e = tid + i * 256;
      if (TRANS_B) {                       // W[n][k]: consecutive threads along k
        const int c = e >> 4, kk = e & 15;
        const int gn = n0 + c, gk = k0 + kk;
        Bs[kk][c] = (gn < N && gk < K) ? __ldg(B + (int64_t)gn * ldb + gk) : 0.f;
      } else {                             // W[k][n]: consecutive threads along n
        const int kk = e >> 6, c = e & 63;
        const int gn = n0 + c, gk = k0 + kk;
        Bs[kk][c] = (gn < N && gk < K) ? __ldg(B + (int64_t)gk * ldb + gn) : 0.f;
      }
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      const float4 a = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t gm = m0 + ty * 4 + i;
    if (gm >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gn = n0 + tx * 4 + j;
      if (gn >= N) continue;
      float v = acc[i][j];
      if (epi.bias) v += __ldg(epi.bias + gn);
      if (epi.accumulate) v += C[gm * ldc + gn];
      if (epi.relu) v = fmaxf(v, 0.f);
      if (epi.mask_src) v = (__ldg(epi.mask_src + gm * epi.ld_mask + gn) > 0.f) ? v : 0.f;
      C[gm * ldc + gn] = v;
    }
  }
}

// wgrad: dW[Nout,Kin] += dY[M,Nout]^T X[M,Kin];  db[Nout] += column sums of dY.
// grid = (ceil(Kin/64), ceil(Nout/64), splits); every CTA reduces rows [split*rows_per, ...) and
// atomically adds its 64x64 partial.
__global__ void __launch_bounds__(256) wgrad_kernel(const float* __restrict__ dY, int ldy, const float* __restrict__ X,
                                                    int ldx, int64_t M, int Nout, int Kin, int64_t rows_per,
                                                    float* __restrict__ dW, float* __restrict__ db) {
  __shared__ __align__(16) float Ys[BK][BM + 4];
  __shared__ __align__(16) float Xs[BK][BN + 4];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int k0 = blockIdx.x * BN;     // Kin tile
  const int n0 = blockIdx.y * BM;     // Nout tile
  const int64_t r_begin = (int64_t)blockIdx.z * rows_per;
  const int64_t r_end = min(M, r_begin + rows_per);
  float acc[4][4] = {};
  float bsum = 0.f;
  for (int64_t r0 = r_begin; r0 < r_end; r0 += BK) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int e = tid + i * 256;
      const int rr = e >> 6, c = e & 63;
      const int64_t gr = r0 + rr;
      const bool ok = gr < r_end;
      Ys[rr][c] = (ok && n0 + c < Nout) ? __ldg(dY + gr * ldy + n0 + c) : 0.f;
      Xs[rr][c] = (ok && k0 + c < Kin) ? __ldg(X + gr * ldx + k0 + c) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      const float4 a = *reinterpret_cast<const float4*>(&Ys[kk][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Xs[kk][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    if (db && blockIdx.x == 0 && tid < BM) {
#pragma unroll
      for (int kk = 0; kk < BK; ++kk) bsum += Ys[kk][tid];
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int gn = n0 + ty * 4 + i;
    if (gn >= Nout) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gk = k0 + tx * 4 + j;
      if (gk < Kin) atomicAdd(dW + (int64_t)gn * Kin + gk, acc[i][j]);
    }
  }
  if (db && blockIdx.x == 0 && tid < BM && n0 + tid < Nout) atomicAdd(db + n0 + tid, bsum);
}

static void gemm(NerfEngine* h, bool trans_b, const float* A, int lda, const float* B, int ldb, float* C, int ldc,
                 int64_t M, int N, int K, const GemmEpi& epi, cudaStream_t st) {
  if (M <= 0) return;
  dim3 grid((unsigned)((M + BM - 1) / BM), (unsigned)((N + BN - 1) / BN));
  if (trans_b) gemm_kernel<true><<<grid, 256, 0, st>>>(A, lda, B, ldb, C, ldc, M, N, K, epi);
  else gemm_kernel<false><<<grid, 256, 0, st>>>(A, lda, B, ldb, C, ldc, M, N, K, epi);
  h->launches++;
}

static void wgrad(NerfEngine* h, const float* dY, int ldy, const float* X, int ldx, int64_t M, int Nout, int Kin,
                  float* dW, float* db, cudaStream_t st) {
  if (M <= 0) return;
  const int tiles = ((Kin + BN - 1) / BN) * ((Nout + BM - 1) / BM);
  int64_t splits = std::max<int64_t>(1, std::min<int64_t>((M + 511) / 512, (int64_t)(h->num_sms * 4 + tiles - 1) / tiles));
  int64_t rows_per = ((M + splits - 1) / splits + BK - 1) / BK * BK;
  splits = (M + rows_per - 1) / rows_per;
  dim3 grid((unsigned)((Kin + BN - 1) / BN), (unsigned)((Nout + BM - 1) / BM), (unsigned)splits);
  wgrad_kernel<<<grid, 256, 0, st>>>(dY, ldy, X, ldx, M, Nout, Kin, rows_per, dW, db);
  h->launches++;
}

static inline int ew_blocks(NerfEngine* h, int64_t m) { return (int)std::min<int64_t>((m + 255) / 256, (int64_t)h->num_sms * 8); }

// ---- activation workspace ---------------------------------------------------------------------
struct Acts {
  float *s5, *v, *hv, *h[8];   // h[4] aliases s5+63 (ld 320)
  float *ga, *gb, *g128, *dpe;
};
static int ld_of(int l) { return l == 4 ? 320 : 256; }

static int get_acts(NerfEngine* h, int net, int64_t rows, bool keep, bool bwd, Acts* a) {
  char name[64];
  const char* tag = keep ? (net == NERF_NET_FINE ? "f32f" : "f32c") : "f32i";
  int rc;
  snprintf(name, sizeof name, "%s.s5", tag);
  if ((rc = h->buf(name, (size_t)rows * 320, &a->s5))) return rc;
  snprintf(name, sizeof name, "%s.v", tag);
  if ((rc = h->buf(name, (size_t)rows * 288, &a->v))) return rc;
  snprintf(name, sizeof name, "%s.hv", tag);
  if ((rc = h->buf(name, (size_t)rows * 128, &a->hv))) return rc;
  if (keep) {
    for (int l = 0; l < 8; ++l) {
      if (l == 4) { a->h[4] = a->s5 + 63; continue; }
      snprintf(name, sizeof name, "%s.h%d", tag, l);
      if ((rc = h->buf(name, (size_t)rows * 256, &a->h[l]))) return rc;
    }
  } else {
    float *p0, *p1;
    if ((rc = h->buf("f32i.p0", (size_t)rows * 256, &p0))) return rc;
    if ((rc = h->buf("f32i.p1", (size_t)rows * 256, &p1))) return rc;
    for (int l = 0; l < 8; ++l) a->h[l] = (l & 1) ? p1 : p0;
    a->h[4] = a->s5 + 63;
    a->h[5] = p1; a->h[6] = p0; a->h[7] = p1;
  }
  a->ga = a->gb = a->g128 = a->dpe = nullptr;
  if (bwd) {
    if ((rc = h->buf("f32b.ga", (size_t)rows * 256, &a->ga))) return rc;
    if ((rc = h->buf("f32b.gb", (size_t)rows * 256, &a->gb))) return rc;
    if ((rc = h->buf("f32b.g128", (size_t)rows * 128, &a->g128))) return rc;
    if ((rc = h->buf("f32b.dpe", (size_t)rows * 64, &a->dpe))) return rc;
  }
  return NERF_OK;
}

static void forward_rows(NerfEngine* h, const float* P, const RowSrc& src, int64_t row0, int64_t m, const Acts& a,
                         float* raw /* rows [row0,row0+m) */, cudaStream_t st) {
  pe_rows_kernel<<<ew_blocks(h, m), 256, 0, st>>>(src, row0, m, a.s5, a.v);
  h->launches++;
  GemmEpi e;
  e.relu = 1;
  const float* x = a.s5; int ldx = 320;
  for (int l = 0; l < 8; ++l) {
    const LayerDim d = layer_dim(l);
    e.bias = P + b_off(l);
    gemm(h, true, x, ldx, P + w_off(l), d.in, a.h[l], ld_of(l), m, d.out, d.in, e, st);
    if (l == 4) { x = a.s5; ldx = 320; } else { x = a.h[l]; ldx = 256; }
  }
  GemmEpi lin;
  lin.bias = P + b_off(L_ALPHA);
  gemm(h, true, a.h[7], 256, P + w_off(L_ALPHA), 256, raw + 3, 4, m, 1, 256, lin, st);
  lin.bias = P + b_off(L_FEAT);
  gemm(h, true, a.h[7], 256, P + w_off(L_FEAT), 256, a.v, 288, m, 256, 256, lin, st);
  e.bias = P + b_off(L_VIEWS);
  gemm(h, true, a.v, 288, P + w_off(L_VIEWS), 283, a.hv, 128, m, 128, 283, e, st);
  lin.bias = P + b_off(L_RGB);
  gemm(h, true, a.hv, 128, P + w_off(L_RGB), 128, raw, 4, m, 3, 128, lin, st);
}

int mlp_fp32_forward(NerfEngine* h, int net, const RowSrc& src, float* raw, bool keep_acts, cudaStream_t st) {
  const int64_t M = src.R * src.S;
  if (M <= 0) return NERF_OK;
  const float* P = h->params + (net == NERF_NET_FINE ? NERF_PARAMS_PER_NET : 0);
  Acts a;
  int rc;
  if (keep_acts) {
    if ((rc = get_acts(h, net, M, true, false, &a))) return rc;
    forward_rows(h, P, src, 0, M, a, raw, st);
  } else {
    const int64_t chunk = std::min<int64_t>(M, 1 << 18);
    if ((rc = get_acts(h, net, chunk, false, false, &a))) return rc;
    for (int64_t r0 = 0; r0 < M; r0 += chunk)
      forward_rows(h, P, src, r0, std::min<int64_t>(chunk, M - r0), a, raw + 4 * r0, st);
  }
  if (src.mask) {
    zero_masked_rows_kernel<<<ew_blocks(h, M), 256, 0, st>>>(src.mask, M, reinterpret_cast<float4*>(raw));
    h->launches++;
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return h->fail(NERF_E_CUDA, std::string("mlp_fp32_forward: ") + cudaGetErrorString(e));
  return NERF_OK;
}

// Backward through the activations kept by the last keep_acts forward of `net` over the same rows.
// sigma_only: the coarse net of the reference graph -- only d_raw[:,3] is non-zero, rgb/views/feature
// receive exactly-zero grads and no input gradient is needed.
int mlp_fp32_backward(NerfEngine* h, int net, const RowSrc& src, const float* d_raw, bool sigma_only,
                      float* G /* flat grad of this net */, float* d_t, cudaStream_t st) {
  const int64_t M = src.R * src.S;
  if (M <= 0) return NERF_OK;
  const float* P = h->params + (net == NERF_NET_FINE ? NERF_PARAMS_PER_NET : 0);
  Acts a;
  int rc;
  if ((rc = get_acts(h, net, M, true, true, &a))) return rc;
  GemmEpi plain;
  float* cur;   // dZ of the layer being processed (pre-activation gradient)
  {
    // d h7 = d_alpha (x) w_alpha [+ dFeat W_feat]; then relu mask of h7
    GemmEpi e0;
    if (sigma_only) { e0.mask_src = a.h[7]; e0.ld_mask = 256; }
    gemm(h, false, d_raw + 3, 4, P + w_off(L_ALPHA), 256, a.gb, 256, M, 256, 1, e0, st);
    wgrad(h, d_raw + 3, 4, a.h[7], 256, M, 1, 256, G + w_off(L_ALPHA), G + b_off(L_ALPHA), st);
    if (!sigma_only) {
      wgrad(h, d_raw, 4, a.hv, 128, M, 3, 128, G + w_off(L_RGB), G + b_off(L_RGB), st);
      GemmEpi e1; e1.mask_src = a.hv; e1.ld_mask = 128;
      gemm(h, false, d_raw, 4, P + w_off(L_RGB), 128, a.g128, 128, M, 128, 3, e1, st);
      wgrad(h, a.g128, 128, a.v, 288, M, 128, 283, G + w_off(L_VIEWS), G + b_off(L_VIEWS), st);
      gemm(h, false, a.g128, 128, P + w_off(L_VIEWS), 283, a.ga, 256, M, 256, 128, plain, st);   // dFeat
      wgrad(h, a.ga, 256, a.h[7], 256, M, 256, 256, G + w_off(L_FEAT), G + b_off(L_FEAT), st);
      GemmEpi e2; e2.accumulate = 1; e2.mask_src = a.h[7]; e2.ld_mask = 256;
      gemm(h, false, a.ga, 256, P + w_off(L_FEAT), 256, a.gb, 256, M, 256, 256, e2, st);
    }
    cur = a.gb;
  }
  float* other = a.ga;
  for (int l = 7; l >= 0; --l) {
    const LayerDim d = layer_dim(l);
    const float* x = (l == 0 || l == 5) ? a.s5 : a.h[l - 1];
    const int ldx = (l == 0 || l == 5) ? 320 : ld_of(l - 1);
    wgrad(h, cur, 256, x, ldx, M, 256, d.in, G + w_off(l), G + b_off(l), st);
    if (l == 0) {
      if (!sigma_only) {   // d PE += dZ0 W0
        GemmEpi e; e.accumulate = 1;
        gemm(h, false, cur, 256, P + w_off(0), 63, a.dpe, 64, M, 63, 256, e, st);
      }
      break;
    }
    if (l == 5) {
      if (!sigma_only) gemm(h, false, cur, 256, P + w_off(5), 319, a.dpe, 64, M, 63, 256, plain, st);
      GemmEpi e; e.mask_src = a.s5 + 63; e.ld_mask = 320;
      gemm(h, false, cur, 256, P + w_off(5) + 63, 319, other, 256, M, 256, 256, e, st);
    } else {
      GemmEpi e; e.mask_src = x; e.ld_mask = ldx;
      gemm(h, false, cur, 256, P + w_off(l), d.in, other, 256, M, 256, 256, e, st);
    }
    std::swap(cur, other);
  }
  if (!sigma_only && d_t) {
    pe_bwd_kernel<<<ew_blocks(h, M), 256, 0, st>>>(src, M, a.dpe, d_t);
    h->launches++;
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return h->fail(NERF_E_CUDA, std::string("mlp_fp32_backward: ") + cudaGetErrorString(e));
  return NERF_OK;
}

}  // namespace nerf
