// pack_weights.cu -- stand-alone pack kernels (after nerf_set_params; the training step re-packs inside the fused
// clip+Adam kernel, adam.cu).  The per-chunk bodies and the schedules live in pack.cuh.
#include "pack.cuh"

namespace nerf {

size_t tc_wpack_bias_offset() { return fwdpack::wpack_bias_offset(); }
size_t tc_wpack_bytes() { return fwdpack::wpack_bytes(); }
size_t tc_bwd_wpack_bytes() { return (size_t)2 * bw::kNumStages * bw::kStageImage; }

// one thread per 16-byte chunk of the forward images, then of the dgrad images, then per bias-slab row
__global__ void pack_all_kernel(const float* __restrict__ flat, __nv_bfloat16* __restrict__ wpack, uint8_t* __restrict__ slabs,
                                __nv_bfloat16* __restrict__ wpack_bwd, __nv_bfloat16* __restrict__ rep, int n_rep) {
  const int64_t total = fwdpack::kWeightChunks + bw::kDgradChunks + fwdpack::kBiasRows;
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    if (e < fwdpack::kWeightChunks) fwdpack::pack_weight_chunk(e, flat, wpack, rep, n_rep);
    else if (e < fwdpack::kWeightChunks + bw::kDgradChunks) bw::pack_dgrad_chunk(e - fwdpack::kWeightChunks, flat, wpack_bwd);
    else fwdpack::pack_bias_row((int)(e - fwdpack::kWeightChunks - bw::kDgradChunks), flat, slabs);
  }
}

void tc_pack_all(NerfEngine* h, const float* flat, __nv_bfloat16* wpack, __nv_bfloat16* wpack_bwd, cudaStream_t st) {
  pack_all_kernel<<<h->num_sms * 8, 256, 0, st>>>(flat, wpack, reinterpret_cast<uint8_t*>(wpack) + tc_wpack_bias_offset(), wpack_bwd,
                                                  h->wrep, h->wrep ? h->w_replicas : 0);
  h->launches++;
}

}  // namespace nerf
