// handshake.cu -- micro-benchmark of the weight-ring HANDSHAKE of the CTA-pair MLP kernels, without MMAs, epilogues or stores:
// how many clocks does one ring stage cost when the only work is
//   producer (both CTAs, one warp): wait w_empty[slot] -> expect_tx (leader) -> TMA half stage to the LEADER's w_full[slot]
//   issuer   (leader, one warp)   : wait w_full[slot]  -> release the slot in both CTAs
// with the release done by tcgen05.commit (multicast, as in the product) or by two remote mbarrier arrives, the loads as
// cp.async.bulk.tensor.2d.cta_group::2 (product) and 16 KB or 1 KB per half stage, the loops run by a converged warp or by
// lane 0 alone.   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I nerf-replication_b200/csrc -o hs handshake.cu -lcudart
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "tc_common.cuh"

using namespace nerf::tc;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

constexpr int kMaxRing = 8;
constexpr int kStages = 39;

__device__ __forceinline__ void tma_cg2(void* smem_dst, const CUtensorMap* tmap, int row0, uint32_t mbar_cluster_addr) {
  asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(tmap), "r"(0), "r"(row0), "r"(mbar_cluster_addr)
               : "memory");
}

__device__ __forceinline__ void tma_local(void* smem_dst, const CUtensorMap* tmap, int row0, uint64_t* bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(tmap), "r"(0), "r"(row0), "r"(smem_u32(bar))
               : "memory");
}

struct Bars { uint64_t w_full[kMaxRing], w_empty[kMaxRing]; uint32_t tmem_slot; };

// release: 0 = tcgen05.commit multicast, 1 = two remote arrives.  lone: 1 = loops run by lane 0 only (diverged warp)
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
handshake_kernel(const __grid_constant__ CUtensorMap map, int ring, int iters, int rows, int release, int lone, int local, unsigned long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ Bars sb;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t rank = cluster_ctarank();
  if (tid == 0) {
    for (int i = 0; i < ring; ++i) { mbar_init(&sb.w_full[i], local == 2 ? (1u << 19) : 1u); mbar_init(&sb.w_empty[i], 1); }   // issue-only: phases never complete
    fence_barrier_init();
  }
  cluster_sync_all();
  if (warp == 1) { tmem_alloc2(&sb.tmem_slot, 32); tmem_relinquish2(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t bytes = (uint32_t)rows * 128u;
  const long long t0 = clock64();
  if (local == 2) {
    // pure issue rate: `release + 1` producer warps, warp w takes the stages it % nprod == w; nobody waits for anything
    const int nprod = release + 1;
    if (warp < nprod) {
      for (int it = warp; it < iters; it += nprod) {
        const uint32_t slot = it % ring;
        if (lane == 0) {
          mbar_arrive_expect_tx(&sb.w_full[slot], bytes);
          tma_local(smem + slot * 16384, &map, (it % kStages) * 256 + (int)rank * 128, &sb.w_full[slot]);
        }
        __syncwarp();
      }
      if (lane == 0 && rank == 0) atomicMax(&out[blockIdx.x >> 1], (unsigned long long)(clock64() - t0));
    }
  } else if (warp == 0) {
    const uint32_t w_full_leader = mapa_u32(smem_u32(&sb.w_full[0]), 0);
    if (!lone || lane == 0) {
      for (int it = 0; it < iters; ++it) {
        const uint32_t slot = it % ring, par = (it / ring) & 1u;
        mbar_wait(&sb.w_empty[slot], par ^ 1u, 0x10 + slot);
        if (lane == 0) {
          if (local) {          // no cross-CTA traffic at all: each CTA loads to its own barrier and has its own consumer
            mbar_arrive_expect_tx(&sb.w_full[slot], bytes);
            tma_local(smem + slot * 16384, &map, (it % kStages) * 256 + (int)rank * 128, &sb.w_full[slot]);
          } else {
            if (rank == 0) mbar_arrive_expect_tx(&sb.w_full[slot], 2 * bytes);
            tma_cg2(smem + slot * 16384, &map, (it % kStages) * 256 + (int)rank * 128, w_full_leader + slot * 8u);
          }
        }
        if (!lone) __syncwarp();
      }
    }
    __syncwarp();
  } else if (warp == 1 && local) {
    for (int it = 0; it < iters; ++it) {
      const uint32_t slot = it % ring, par = (it / ring) & 1u;
      mbar_wait(&sb.w_full[slot], par, 0x20 + slot);
      if (lane == 0) mbar_arrive(&sb.w_empty[slot]);
      __syncwarp();
    }
    if (lane == 0 && rank == 0) out[blockIdx.x >> 1] = (unsigned long long)(clock64() - t0);
  } else if (warp == 1 && rank == 0) {
    const uint32_t e0 = mapa_u32(smem_u32(&sb.w_empty[0]), 0), e1 = mapa_u32(smem_u32(&sb.w_empty[0]), 1);
    if (!lone || lane == 0) {
      for (int it = 0; it < iters; ++it) {
        const uint32_t slot = it % ring, par = (it / ring) & 1u;
        mbar_wait(&sb.w_full[slot], par, 0x20 + slot);
        tc_fence_after();
        if (release == 0) {
          if (lone) umma_commit2(&sb.w_empty[slot], 0x3); else umma_commit2_w(&sb.w_empty[slot], 0x3);
        } else if (lane == 0) {
          mbar_arrive_remote(e0 + slot * 8u);
          mbar_arrive_remote(e1 + slot * 8u);
        }
        if (!lone) __syncwarp();
      }
    }
    __syncwarp();
    if (lane == 0) out[blockIdx.x >> 1] = (unsigned long long)(clock64() - t0);
  }
  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) tmem_dealloc2(sb.tmem_slot, 32);
}

int main() {
  CK(cudaSetDevice(0));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  const int sms = prop.multiProcessorCount;
  const size_t img_bytes = (size_t)kStages * 32768;
  uint8_t* img;
  CK(cudaMalloc(&img, img_bytes));
  CK(cudaMemset(img, 1, img_bytes));
  unsigned long long* out;
  CK(cudaMalloc(&out, sizeof(unsigned long long) * sms));
  CK(cudaFuncSetAttribute(handshake_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxRing * 16384));
  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                               const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
  const int iters = 39 * 40;
  std::vector<unsigned long long> h(sms);
  printf("# %d SMs = %d CTA pairs, %d ring stages per pair; clk/stage = slowest pair's cycles / stages\n", sms, sms / 2, iters);
  printf("%-8s %-10s %-10s %-5s %10s\n", "rows", "release", "warp", "ring", "clk/stage");
  for (int rows : {128, 8})
    for (int local = 0; local < 3; ++local)
    for (int release = 0; release < (local == 2 ? 4 : local ? 1 : 2); ++release)
      for (int lone = 0; lone < (local ? 1 : 2); ++lone)
        for (int ring : {1, 2, 3, 5, 8}) {
          CUtensorMap map;
          const cuuint64_t gdim[2] = {64, (cuuint64_t)(img_bytes / 128)}, gstride[1] = {128};
          const cuuint32_t box[2] = {64, (cuuint32_t)rows}, estr[2] = {1, 1};
          CUresult r = reinterpret_cast<EncodeFn>(fn)(&map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, img, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                                      CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
          if (r != CUDA_SUCCESS) { printf("tensor map encode failed %d\n", (int)r); return 1; }
          for (int rep = 0; rep < 2; ++rep) {
            CK(cudaMemset(out, 0, sizeof(unsigned long long) * sms));
            handshake_kernel<<<sms & ~1, 128, ring * 16384>>>(map, ring, iters, rows, release, lone, local, out);
            CK(cudaGetLastError());
            CK(cudaDeviceSynchronize());
          }
          CK(cudaMemcpy(h.data(), out, sizeof(unsigned long long) * (sms / 2), cudaMemcpyDeviceToHost));
          unsigned long long mx = 0;
          for (int i = 0; i < sms / 2; ++i) if (h[i] > mx) mx = h[i];
          printf("%-8d %-10s %-10s %-5d %10.0f\n", rows, local == 2 ? (release == 0 ? "issue x1" : release == 1 ? "issue x2" : release == 2 ? "issue x3" : "issue x4") : local ? "local" : (release ? "arrive x2" : "commit"), lone ? "lane 0" : "converged", ring, (double)mx / iters);
        }
  return 0;
}
