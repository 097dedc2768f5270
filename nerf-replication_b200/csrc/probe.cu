// probe.cu -- bring-up battery for the hand-written tcgen05 path: one CTA, one 128 x N x K bf16 GEMM,
// operands staged through shared memory (or TMEM for A) in each layout the engine uses.  Lets a GPU run
// tell descriptor/layout bugs apart from pipeline bugs in the fused kernels.
//   variant 0: A,B in smem, K-major, 128B swizzle        (the K-major operands of mlp_tc2.cu)
//   variant 1: A,B in smem, K-major, no swizzle ("interleaved" 8x16B core matrices)
//   variant 2: A in TMEM (bf16 pairs per 32-bit column), B in smem 128B swizzle
#include "engine.h"
#include "tc_common.cuh"

namespace nerf {
using namespace tc;

__global__ void __launch_bounds__(128, 1)
umma_probe_kernel(int variant, const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ D, int N, int K) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar_done;
  __shared__ uint32_t tmem_base_slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int kblocks = K / 64;
  uint8_t* sA = smem;                                   // kblocks x 16 KB
  uint8_t* sB = smem + (size_t)kblocks * 16384;          // kblocks x (N*128) B

  if (warp == 0) { tmem_alloc(&tmem_base_slot, 512); tmem_relinquish(); }
  if (tid == 32) { mbar_init(&bar_done, 1); fence_barrier_init(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_slot;

  // stage operands (fp32 -> bf16) into shared memory
  for (int e = tid; e < 128 * K; e += 128) {
    const int r = e / K, k = e % K;
    const __nv_bfloat16 v = __float2bfloat16(A[e]);
    uint32_t off;
    if (variant == 1) off = (uint32_t)((k / 8) * (128 * 16) + r * 16 + (k % 8) * 2);
    else off = (uint32_t)((k / 64) * 16384) + sw128_offset(r, (k % 64) / 8) + (k % 8) * 2;
    *reinterpret_cast<__nv_bfloat16*>(sA + off) = v;
  }
  for (int e = tid; e < N * K; e += 128) {
    const int r = e / K, k = e % K;
    const __nv_bfloat16 v = __float2bfloat16(B[e]);
    uint32_t off;
    if (variant == 1) off = (uint32_t)((k / 8) * (N * 16) + r * 16 + (k % 8) * 2);
    else off = (uint32_t)((k / 64) * (N * 128)) + sw128_offset(r, (k % 64) / 8) + (k % 8) * 2;
    *reinterpret_cast<__nv_bfloat16*>(sB + off) = v;
  }
  if (variant == 2) {
    // A into TMEM columns [256, 256 + K/2): thread = row, 2 bf16 per column (low half = even k)
    const int row = warp * 32 + lane;
    for (int c0 = 0; c0 < K / 2; c0 += 16) {
      uint32_t v[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] = pack_bf16x2(A[row * K + 2 * (c0 + j)], A[row * K + 2 * (c0 + j) + 1]);
      tmem_st16(tmem + ((uint32_t)(warp * 32) << 16) + 256 + c0, v);
    }
    tmem_wait_st();
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  if (tid == 0) {
    const uint32_t idesc = make_idesc_bf16(128, N);
    for (int ks = 0; ks < K / 16; ++ks) {
      uint64_t ad, bd;
      if (variant == 1) {
        ad = make_smem_desc(smem_u32(sA) + ks * 2 * (128 * 16), 128 * 16, 128, kLayoutNone);
        bd = make_smem_desc(smem_u32(sB) + ks * 2 * (N * 16), N * 16, 128, kLayoutNone);
      } else {
        const int kb = ks / 4, kk = ks % 4;
        ad = make_smem_desc(smem_u32(sA) + kb * 16384 + kk * 32, 16, 1024, kLayoutSW128);
        bd = make_smem_desc(smem_u32(sB) + kb * (N * 128) + kk * 32, 16, 1024, kLayoutSW128);
      }
      if (variant == 2) umma_ts(tmem, tmem + 256 + ks * 8, bd, idesc, ks > 0);
      else umma_ss(tmem, ad, bd, idesc, ks > 0);
    }
    umma_commit(&bar_done);
  }
  mbar_wait(&bar_done, 0, 0x100 + variant);
  tc_fence_after();
  {
    const int row = warp * 32 + lane;
    for (int c0 = 0; c0 < N; c0 += 32) {
      uint32_t v[32];
      tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
      tmem_wait_ld();
#pragma unroll
      for (int j = 0; j < 32; ++j) D[row * N + c0 + j] = __uint_as_float(v[j]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

// variant 3: CTA pair (cluster of 2, tcgen05 cta_group::2): D[256 x N] = A[256 x K] * B[N x K]^T.
// CTA r stages A rows [128r, 128r+128) and B rows [N/2 r, N/2 r + N/2) (both K-major, 128B swizzle) at the SAME
// shared-memory offsets; the leader (rank 0) issues M=256 MMAs; the commit is multicast to both CTAs; each CTA
// reads its 128 accumulator rows from its own TMEM.
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
umma_probe2_kernel(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ D, int N, int K) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar_done, bar_ready;
  __shared__ uint32_t tmem_base_slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t rank = cluster_ctarank();
  const int kblocks = K / 64, Nh = N / 2;
  uint8_t* sA = smem;
  uint8_t* sB = smem + (size_t)kblocks * 16384;

  if (tid == 32) { mbar_init(&bar_done, 1); mbar_init(&bar_ready, 256); fence_barrier_init(); }
  cluster_sync_all();
  if (warp == 0) { tmem_alloc2(&tmem_base_slot, 512); tmem_relinquish2(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_slot;

  for (int e = tid; e < 128 * K; e += 128) {
    const int r = e / K, k = e % K;
    const uint32_t off = (uint32_t)((k / 64) * 16384) + sw128_offset(r, (k % 64) / 8) + (k % 8) * 2;
    *reinterpret_cast<__nv_bfloat16*>(sA + off) = __float2bfloat16(A[(size_t)(rank * 128 + r) * K + k]);
  }
  for (int e = tid; e < Nh * K; e += 128) {
    const int r = e / K, k = e % K;
    const uint32_t off = (uint32_t)((k / 64) * (Nh * 128)) + sw128_offset(r, (k % 64) / 8) + (k % 8) * 2;
    *reinterpret_cast<__nv_bfloat16*>(sB + off) = __float2bfloat16(B[(size_t)(rank * Nh + r) * K + k]);
  }
  fence_proxy_async_smem();
  mbar_arrive_cluster(mapa_u32(smem_u32(&bar_ready), 0));      // 128 threads x 2 CTAs arrive on the leader's barrier

  if (rank == 0 && tid == 0) {
    mbar_wait_cluster(&bar_ready, 0, 0x200);
    tc_fence_after();
    const uint32_t idesc = make_idesc_bf16(256, N);
    for (int ks = 0; ks < K / 16; ++ks) {
      const int kb = ks / 4, kk = ks % 4;
      const uint64_t ad = make_smem_desc(smem_u32(sA) + kb * 16384 + kk * 32, 16, 1024, kLayoutSW128);
      const uint64_t bd = make_smem_desc(smem_u32(sB) + kb * (Nh * 128) + kk * 32, 16, 1024, kLayoutSW128);
      umma_ss2(tmem, ad, bd, idesc, ks > 0);
    }
    umma_commit2(&bar_done, 0x3);
  }
  mbar_wait(&bar_done, 0, 0x103);
  tc_fence_after();
  {
    const int row = warp * 32 + lane;
    for (int c0 = 0; c0 < N; c0 += 32) {
      uint32_t v[32];
      tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
      tmem_wait_ld();
#pragma unroll
      for (int j = 0; j < 32; ++j) D[(size_t)(rank * 128 + row) * N + c0 + j] = __uint_as_float(v[j]);
    }
  }
  tc_fence_before();
  cluster_sync_all();
  if (warp == 0) tmem_dealloc2(tmem, 512);
}

// variant 4: both operands MN-major (the wgrad form): D[128 x N] = At^T * Bt with At [K x 128], Bt [K x N]
// (K = reduction dim = "rows").  The operands are staged as [rows x 64-column] 128B-swizzled blocks -- the very
// images the forward saves -- and described as MN-major: LBO = distance between 64-column blocks, SBO = 1024
// (8-row groups), one K=16 step = 16 rows = +2048 bytes.
// layout 0: the 128B-swizzled tile images above.  layout 1/2: un-swizzled "core matrix" images (tc_arena.cuh, half-tile
// piece): element (row r < 64, column c) at (c/8)*1024 + r*16 + (c%8)*2 -- 8 rows x 8 columns = one contiguous 128-byte
// core matrix; layout 1 describes it with LBO = 128 (next 8 rows), SBO = 1024 (next 8 columns), layout 2 swaps the two.
__global__ void __launch_bounds__(128, 1)
umma_probe_mn_kernel(const float* __restrict__ At, const float* __restrict__ Bt, float* __restrict__ D, int N, int K, int layout) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar_done;
  __shared__ uint32_t tmem_base_slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  uint8_t* sA = smem;                 // 2 blocks x 16 KB
  uint8_t* sB = smem + 2 * 16384;     // N/64 blocks x 16 KB
  if (warp == 0) { tmem_alloc(&tmem_base_slot, 512); tmem_relinquish(); }
  if (tid == 32) { mbar_init(&bar_done, 1); fence_barrier_init(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_slot;
  for (int e = tid; e < K * 128; e += 128) {
    const int r = e / 128, c = e % 128;
    const uint32_t off = layout == 0 ? (c / 64) * 16384 + sw128_offset(r, (c % 64) / 8) + (c % 8) * 2 : (c / 8) * 1024 + r * 16 + (c % 8) * 2;
    *reinterpret_cast<__nv_bfloat16*>(sA + off) = __float2bfloat16(At[e]);
  }
  for (int e = tid; e < K * N; e += 128) {
    const int r = e / N, c = e % N;
    const uint32_t off = layout == 0 ? (c / 64) * 16384 + sw128_offset(r, (c % 64) / 8) + (c % 8) * 2 : (c / 8) * 1024 + r * 16 + (c % 8) * 2;
    *reinterpret_cast<__nv_bfloat16*>(sB + off) = __float2bfloat16(Bt[e]);
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (tid == 0) {
    const uint32_t idesc = make_idesc_bf16(128, N, 1, 1);
    for (int ks = 0; ks < K / 16; ++ks) {
      uint64_t ad, bd;
      if (layout == 0) {
        ad = make_smem_desc(smem_u32(sA) + ks * 2048, 16384, 1024, kLayoutSW128);
        bd = make_smem_desc(smem_u32(sB) + ks * 2048, 16384, 1024, kLayoutSW128);
      } else {
        const uint32_t lbo = layout == 1 ? 128 : 1024, sbo = layout == 1 ? 1024 : 128;
        ad = make_smem_desc(smem_u32(sA) + ks * 256, lbo, sbo, kLayoutNone);
        bd = make_smem_desc(smem_u32(sB) + ks * 256, lbo, sbo, kLayoutNone);
      }
      umma_ss(tmem, ad, bd, idesc, ks > 0);
    }
    umma_commit(&bar_done);
  }
  mbar_wait(&bar_done, 0, 0x104);
  tc_fence_after();
  {
    const int row = warp * 32 + lane;
    for (int c0 = 0; c0 < N; c0 += 32) {
      uint32_t v[32];
      tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
      tmem_wait_ld();
#pragma unroll
      for (int j = 0; j < 32; ++j) D[row * N + c0 + j] = __uint_as_float(v[j]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

}  // namespace nerf

namespace nerf { NERF_TC_DEFINE_DEBUG_SINK(tcp_set_debug_sink) }

extern "C" int nerf_debug_umma_probe(int variant, const float* a, const float* b, float* d, int M, int N, int K, void* stream) {
  using namespace nerf;
  if (variant >= 4 && variant <= 6) {   // a = At [K x 128], b = Bt [K x N]; 5/6: core-matrix layout, K <= 64
    if (M != 128 || N % 64 || N < 64 || N > 256 || K % 16 || K < 16 || K > (variant == 4 ? 128 : 64)) return NERF_E_INVALID;
    const size_t smem4 = (size_t)(2 + N / 64) * 16384 + 1024;
    if (cudaFuncSetAttribute(umma_probe_mn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem4) != cudaSuccess) return NERF_E_CUDA;
    umma_probe_mn_kernel<<<1, 128, smem4, (cudaStream_t)stream>>>(a, b, d, N, K, variant - 4);
    return cudaGetLastError() == cudaSuccess ? NERF_OK : NERF_E_CUDA;
  }
  if (variant == 3) {
    if (M != 256 || N % 32 || N < 32 || N > 256 || K % 64 || K < 64 || K > 256) return NERF_E_INVALID;
    const size_t smem2 = (size_t)(K / 64) * 16384 + (size_t)K * (N / 2) * 2 + 1024;
    if (cudaFuncSetAttribute(umma_probe2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2) != cudaSuccess) return NERF_E_CUDA;
    umma_probe2_kernel<<<2, 128, smem2, (cudaStream_t)stream>>>(a, b, d, N, K);
    return cudaGetLastError() == cudaSuccess ? NERF_OK : NERF_E_CUDA;
  }
  if (M != 128 || N % 16 || N < 16 || N > 256 || K % 64 || K < 64 || K > 256 || variant < 0 || variant > 2) return NERF_E_INVALID;
  const size_t smem = (size_t)(K / 64) * 16384 + (size_t)K * N * 2 + 1024;
  cudaError_t e = cudaFuncSetAttribute(umma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return NERF_E_CUDA;
  umma_probe_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(variant, a, b, d, N, K);
  return cudaGetLastError() == cudaSuccess ? NERF_OK : NERF_E_CUDA;
}
