/* nerf_b200_probe.h -- C ABI of nerf-replication_b200/libnerf_b200_probe.so: the tcgen05 descriptor probe battery
 * used to bring up the operand layouts of the fused kernels (csrc/probe.cu).  Test infrastructure: it is NOT part
 * of the product library libnerf_b200.so and nothing on the render / train path links or loads it.  Nothing in the
 * reference corresponds to it (its only native ABI is the unused hash encoder, hashencoder/src/bindings.cpp:5-8). */
#ifndef NERF_B200_PROBE_H_
#define NERF_B200_PROBE_H_
#ifdef __cplusplus
extern "C" {
#endif

/* One [M x N x K] bf16 GEMM (fp32 accumulate, device pointers a, b, d fp32) through one operand layout:
 *   variant 0: A, B in shared memory, K-major, 128B swizzle (M = 128)         1: K-major, un-swizzled core matrices
 *   variant 2: A in TMEM, B in shared memory 128B swizzle                     3: cta_group::2, M = 256 over a CTA pair
 *   variant 4: both operands MN-major (a = At [K x 128], b = Bt [K x N])       5/6: MN-major core-matrix layouts (K <= 64)
 * Returns 0, or a negative NERF_E_* code for unsupported shapes / launch failures. */
int nerf_debug_umma_probe(int variant, const float* a, const float* b, float* d, int M, int N, int K,
                          void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NERF_B200_PROBE_H_ */
