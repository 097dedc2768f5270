// wstream.cu -- micro-benchmark of the weight-stage stream of the CTA-pair MLP kernels (round 2): what do
// L2 -> shared-memory bulk loads of 16 KB deliver per SM when every SM streams THE SAME 1.25 MB image through a ring of
// R slots?  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o wstream wstream.cu -lcudart ; ./wstream
// Modes: 1-D cp.async.bulk or 2-D tensor-map loads (128-byte rows, box 64 x 128 bf16 = the product's maps); all CTAs read the
// same image or each its own copy; the consumer releases a slot at once or after `delay` clocks (512 = the MMAs of a stage).
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(c)); }
__device__ __forceinline__ void mbar_arrive(uint64_t* b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
  uint32_t ok = 0;
  while (!ok) {
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(smem_u32(b)), "r"(parity) : "memory");
  }
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes),
               "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tma2d_g2s(void* dst, const CUtensorMap* map, int x, int y, uint64_t* bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(smem_u32(dst)), "l"(map),
               "r"(x), "r"(y), "r"(smem_u32(bar))
               : "memory");
}

constexpr int kChunk = 16384;        // a CTA's half of a 32 KB weight stage
constexpr int kStages = 39;          // stages per net
constexpr int kMaxRing = 12;

// out[cta] = {cycles, sum of (full seen - issue) over the chunks}
__global__ void __launch_bounds__(64, 1)
stream_kernel(const __grid_constant__ CUtensorMap map, const uint8_t* img, size_t cta_stride, int ring, int iters, int delay, int use_map,
              unsigned long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t full[kMaxRing], empty[kMaxRing];
  __shared__ long long t_issue[kMaxRing];
  if (threadIdx.x == 0) {
    for (int i = 0; i < ring; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const uint8_t* base = img + (size_t)blockIdx.x * cta_stride;
  const int half = blockIdx.x & 1;
  const long long t0 = clock64();
  if (threadIdx.x == 0) {                               // producer
    for (int i = 0; i < iters; ++i) {
      const int slot = i % ring;
      const uint32_t par = (i / ring) & 1u;
      mbar_wait(&empty[slot], par ^ 1u);
      const int stage = i % kStages;
      mbar_expect(&full[slot], kChunk);
      t_issue[slot] = clock64();
      if (use_map) tma2d_g2s(smem + slot * kChunk, &map, 0, (int)((cta_stride * blockIdx.x) / 128) + stage * 256 + half * 128, &full[slot]);
      else bulk_g2s(smem + slot * kChunk, base + (size_t)stage * 2 * kChunk + (size_t)half * kChunk, kChunk, &full[slot]);
    }
  } else if (threadIdx.x == 32) {                       // consumer
    long long lat = 0;
    for (int i = 0; i < iters; ++i) {
      const int slot = i % ring;
      const uint32_t par = (i / ring) & 1u;
      mbar_wait(&full[slot], par);
      const long long t = clock64();
      lat += t - *(volatile long long*)&t_issue[slot];
      if (delay) { while (clock64() - t < delay) {} }
      mbar_arrive(&empty[slot]);
    }
    out[2 * blockIdx.x] = (unsigned long long)(clock64() - t0);
    out[2 * blockIdx.x + 1] = (unsigned long long)lat;
  }
}

int main(int argc, char** argv) {
  int dev = 0;
  CK(cudaSetDevice(dev));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, dev));
  const int sms = prop.multiProcessorCount;
  const size_t img_bytes = (size_t)kStages * 2 * kChunk;               // 1.25 MB
  uint8_t* img;
  CK(cudaMalloc(&img, img_bytes * sms));
  CK(cudaMemset(img, 1, img_bytes * sms));
  unsigned long long* out;
  CK(cudaMalloc(&out, sizeof(unsigned long long) * 2 * sms));
  CK(cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxRing * kChunk));

  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                               const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
  CUtensorMap map;
  {
    const cuuint64_t gdim[2] = {64, (cuuint64_t)(img_bytes * sms / 128)}, gstride[1] = {128};
    const cuuint32_t box[2] = {64, 128}, estr[2] = {1, 1};
    CUresult r = reinterpret_cast<EncodeFn>(fn)(&map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, img, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                                CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("tensor map encode failed %d\n", (int)r); return 1; }
  }
  const int iters = 39 * 40;
  std::vector<unsigned long long> h(2 * sms);
  printf("# sms %d, chunk %d B, %d chunks per CTA; B/clk/SM = bytes / slowest CTA's cycles; lat = mean issue->full clocks\n", sms, kChunk, iters);
  printf("%-6s %-9s %-5s %-5s %-6s %10s %10s %8s\n", "grid", "source", "path", "ring", "delay", "B/clk/SM", "chip B/clk", "lat");
  const int grids[2] = {sms, 1};
  for (int gi = 0; gi < 2; ++gi)
    for (int distinct = 0; distinct < 2; ++distinct)
      for (int use_map = 0; use_map < 2; ++use_map)
        for (int delay : {0, 512})
          for (int ring : {1, 2, 3, 4, 6, 8, 12}) {
            const int grid = grids[gi];
            if (grid == 1 && (distinct || delay)) continue;
            const size_t stride = distinct ? img_bytes : 0;
            for (int rep = 0; rep < 2; ++rep) {          // first run warms L2
              stream_kernel<<<grid, 64, ring * kChunk>>>(map, img, stride, ring, iters, delay, use_map, out);
              CK(cudaGetLastError());
              CK(cudaDeviceSynchronize());
            }
            CK(cudaMemcpy(h.data(), out, sizeof(unsigned long long) * 2 * grid, cudaMemcpyDeviceToHost));
            unsigned long long mx = 0, lat = 0;
            for (int i = 0; i < grid; ++i) { if (h[2 * i] > mx) mx = h[2 * i]; lat += h[2 * i + 1]; }
            const double bpc = (double)iters * kChunk / (double)mx;
            printf("%-6d %-9s %-5s %-5d %-6d %10.2f %10.0f %8.0f\n", grid, distinct ? "distinct" : "same", use_map ? "tma2d" : "bulk", ring, delay, bpc, bpc * grid,
                   (double)lat / ((double)iters * grid));
          }
  return 0;
}
