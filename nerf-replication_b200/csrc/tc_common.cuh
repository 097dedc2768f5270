// tc_common.cuh -- hand-written sm_100a primitives: mbarrier, bulk-copy TMA, tcgen05 (UMMA) + TMEM.
// Inline PTX only; descriptor bit layouts follow the PTX ISA "tcgen05 matrix/instruction descriptor"
// tables (same fields CUTLASS's cute/arch/mma_sm100_desc.hpp encodes).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace nerf {
namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- debug watchdog: a deadlocked barrier traps instead of hanging the GPU -------------------------
// g_tc_debug[0] = first failing code (barrier tag), [1] = blockIdx, [2] = extra
static __device__ unsigned int g_tc_debug[8];   // one copy per translation unit
// optional host-mapped copy (survives the trap that kills the context); set per translation unit by tc_set_debug_sink
static __device__ unsigned int* g_tc_debug_host = nullptr;
constexpr long long kWatchdogCycles = 6000000000LL;   // ~3-4 s at B200 clocks
// wait profile (diagnostic build -DNERF_TC_WAITPROF + nerf_debug_waitprof): [b] = cycles warps spent in barrier waits
// tagged b = (tag >> 4) & 255, [256 + b] = number of such waits.  Compiled out of the product library.
static __device__ unsigned long long* g_tc_waitprof = nullptr;
#ifdef NERF_TC_WAITPROF
__device__ __forceinline__ long long waitprof_now() { return clock64(); }
__device__ __forceinline__ void waitprof_add(uint32_t tag, long long t0) {
  unsigned long long* p = g_tc_waitprof;
  if (p != nullptr && (__ffs(__activemask()) - 1) == (int)(threadIdx.x & 31)) {
    atomicAdd(p + ((tag >> 4) & 255u), (unsigned long long)(clock64() - t0));
    atomicAdd(p + 256 + ((tag >> 4) & 255u), 1ull);
  }
}
#else
__device__ __forceinline__ long long waitprof_now() { return 0; }
__device__ __forceinline__ void waitprof_add(uint32_t, long long) {}
#endif
// event trace (diagnostic build -DNERF_TC_TRACE, read with nerf_debug_trace): CTA 0 appends (clock64 << 16 | tag) records
// behind the wait-profile counters; `base` = offset of the trace area relative to this translation unit's counter block
#ifdef NERF_TC_TRACE
// plain stores into a per-role region [lo, hi) of the trace area (no atomics: a returning atomic costs ~800 clk of the
// thread being traced); `n` is the role's own running count (a register)
__device__ __forceinline__ void trace_event(uint32_t tag, int base, uint32_t& n, uint32_t lo, uint32_t hi) {
  unsigned long long* p = g_tc_waitprof;
  if (p != nullptr && blockIdx.x == 0 && lo + n < hi) {
    p[base + 1 + lo + n] = ((unsigned long long)clock64() << 16) | tag;
    ++n;
  }
}
#else
__device__ __forceinline__ void trace_event(uint32_t, int, uint32_t&, uint32_t, uint32_t) {}
#endif

// ---- mbarrier ---------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// NERF_MBAR_POLL (A/B builds): 0 = mbarrier.try_wait (may suspend the thread for a system-dependent time), 1 = mbarrier.test_wait
// (never suspends: pure polling), N > 1 = try_wait with suspendTimeHint = N ns
#ifndef NERF_MBAR_POLL
#define NERF_MBAR_POLL 0
#endif
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
#if NERF_MBAR_POLL == 1
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
#elif NERF_MBAR_POLL > 1
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity), "r"((uint32_t)NERF_MBAR_POLL)
      : "memory");
#else
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
#endif
  return ok != 0;
}
static __device__ __noinline__ void mbar_timeout(uint32_t tag, uint32_t parity) {
  // every waiter that times out logs itself (host-mapped: [8..24) = tag | warp << 12 | block << 16 | parity << 31),
  // then lingers so that the other stuck waiters can log too before the trap kills the grid
  // one log line per warp: the lowest active lane speaks
  const bool speaker = (__ffs(__activemask()) - 1) == (int)(threadIdx.x & 31);
  const unsigned int n = speaker ? atomicAdd(&g_tc_debug[4], 1u) : 1000u;
  if (n == 0) {
    g_tc_debug[0] = tag; g_tc_debug[1] = blockIdx.x; g_tc_debug[2] = threadIdx.x; g_tc_debug[3] = parity;
  }
  if (g_tc_debug_host) {
    volatile unsigned int* hp = g_tc_debug_host;
    if (n == 0) { hp[1] = blockIdx.x; hp[2] = threadIdx.x; hp[3] = parity; hp[0] = tag; }
    if (n < 56) hp[8 + n] = tag | ((threadIdx.x >> 5) << 12) | (blockIdx.x << 16) | (parity << 31);
  }
  __threadfence_system();
  const long long t0 = clock64();
  while (clock64() - t0 < 1500000000LL) { }
  __trap();
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity, uint32_t tag) {
  const long long tp = waitprof_now();
  if (mbar_try_wait(bar, parity)) { waitprof_add(tag, tp); return; }
  const long long t0 = clock64();
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++spins & 0x3FFu) == 0 && clock64() - t0 > kWatchdogCycles) mbar_timeout(tag, parity);
  }
  waitprof_add(tag, tp);
}

// host side: point this translation unit's watchdog at a host-mapped buffer
#define NERF_TC_DEFINE_DEBUG_SINK(fn)                                                              \
  void fn(unsigned int* mapped) { cudaMemcpyToSymbol(nerf::tc::g_tc_debug_host, &mapped, sizeof(mapped)); } \
  void fn##_waitprof(unsigned long long* buf) { cudaMemcpyToSymbol(nerf::tc::g_tc_waitprof, &buf, sizeof(buf)); }

// generic-proxy writes (st.shared) -> visible to the async proxy (UMMA / TMA reads)
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- bulk copy global -> shared (TMA engine, no tensor map): completes on an mbarrier ---------------
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// ---- TMEM -----------------------------------------------------------------------------------
// whole warp, .sync.aligned.  Writes the base address (lane 0, column base) to *smem_slot.
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_slot, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// 32 lanes x 32 columns of 32-bit: thread t of the warp gets lane (base_lane + t), columns c..c+31.
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
// 32 lanes x 16 columns store (thread t writes lane base+t, columns c..c+15)
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}

// ---- UMMA descriptors -------------------------------------------------------------------------
// Shared-memory matrix descriptor (64-bit):
//   [0,14)  start address >> 4        [16,30) leading-dim byte offset >> 4
//   [32,46) stride-dim byte offset>>4 [46,48) version = 1 on sm_100
//   [49,52) base offset               [61,64) layout: 0 none, 2 = 128B swizzle, 4 = 64B, 6 = 32B
constexpr uint64_t kLayoutNone = 0, kLayoutSW128 = 2;
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint64_t layout) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  d |= layout << 61;
  return d;
}
// Instruction descriptor for kind::f16, bf16 x bf16 -> f32, both operands K-major:
//   [4,6) D fmt (1 = f32)  [7,10) A fmt (1 = bf16)  [10,13) B fmt (1 = bf16)
//   [15] A major (0 = K)   [16] B major (0 = K)     [17,23) N>>3   [24,29) M>>4
__host__ __device__ constexpr uint32_t make_idesc_bf16(int M, int N, int a_mn_major = 0, int b_mn_major = 0) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// D[tmem] (+)= A[smem] * B[smem]^T   (one thread issues)
__device__ __forceinline__ void umma_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem]^T
__device__ __forceinline__ void umma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// all previously issued MMAs of this thread arrive on `bar` when complete (implies fence::before_thread_sync)
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// ---- 128B-swizzled K-major tile addressing ------------------------------------------------------
// A tile of R rows x 64 bf16 (128 B per row): 8-row groups of 1024 B; the 16-byte chunk index is XORed
// with (row & 7).  Matches UMMA layout type SWIZZLE_128B with SBO = 1024 (and TMA's 128B swizzle).
__host__ __device__ __forceinline__ uint32_t sw128_offset(int row, int chunk16 /*0..7*/) {
  return (uint32_t)((row >> 3) * 1024 + (row & 7) * 128 + ((chunk16 ^ (row & 7)) << 4));
}

// per-half "x > 0" mask of a bf16x2 word: 0xFFFF where the half is positive (HSET2.BF16)
__device__ __forceinline__ uint32_t pos_mask2(uint32_t h) {
  uint32_t d, z = 0;
  asm("set.gt.u32.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(h), "r"(z));
  return d;
}
// 16 bf16x2 words (32 columns) -> one mask word.  Bit layout chosen for the consumer (mask2_from_bits): with k = i >> 1 and
// b = i & 1, the flag of word i's low half is bit 8 b + 7 - k and the flag of its high half bit 8 (b + 2) + 7 - k, so that
// (bits << k) carries both flags in the sign bits of bytes b and b + 2.
__device__ __forceinline__ uint32_t pos_bits32(const uint32_t (&pk)[16]) {
  uint32_t m = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) m |= pos_mask2(pk[i]) & (0x00010001u << (8 * (i & 1) + 7 - (i >> 1)));
  return m;
}
// inverse for one word: 0xFFFF in the halves whose flags are set -- one shift per word pair and one PRMT in sign-replicate
// mode (selector nibble 8 | byte) instead of shift + and + multiply
__device__ __forceinline__ uint32_t mask2_from_bits(uint32_t bits, int i) {
  uint32_t d;
  const uint32_t m = bits << (i >> 1);
  if (i & 1) asm("prmt.b32 %0, %1, %1, 0xBB99;" : "=r"(d) : "r"(m));
  else asm("prmt.b32 %0, %1, %1, 0xAA88;" : "=r"(d) : "r"(m));
  return d;
}
// streaming 16-byte global store (written once, read by a later kernel after far more than L2-size traffic)
// NERF_EXP (diagnostic builds only, build.py --exp N; never defined in the product library): timing experiments that break
// results on purpose -- bit 0: no arena stores, bit 1: MMA issuers do not wait for weight stages, bit 2: dgrad loads no masks
#ifndef NERF_EXP
#define NERF_EXP 0
#endif
__device__ __forceinline__ void st_global_cs(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  if (NERF_EXP & 1) return;
#ifndef NERF_ARENA_ST
#define NERF_ARENA_ST "st.global.cs.v4.b32"      // A/B builds: .cg / .wt / default write-back measured no faster
#endif
  if (NERF_EXP & 4096) {      // timing experiment: the same number of store instructions, a quarter of the bytes
    asm volatile("st.global.cs.b32 [%0], %1;" ::"l"(p), "r"(a ^ b ^ c ^ d) : "memory");
    return;
  }
  asm volatile(NERF_ARENA_ST " [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// one row's 32-byte bit-mask record (tc_arena.cuh): 256-bit global store / read-only load (sm_100: STG.256 / LDG.256)
__device__ __forceinline__ void st_global_mask8(void* p, const uint32_t (&m)[8]) {
  if (NERF_EXP & 64) return;
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(m[0]), "r"(m[1]), "r"(m[2]), "r"(m[3]), "r"(m[4]),
               "r"(m[5]), "r"(m[6]), "r"(m[7])
               : "memory");
}
__device__ __forceinline__ void ld_global_mask8(const void* p, uint32_t (&m)[8]) {
  asm volatile("ld.global.nc.v8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(m[0]), "=r"(m[1]), "=r"(m[2]), "=r"(m[3]), "=r"(m[4]), "=r"(m[5]), "=r"(m[6]), "=r"(m[7])
               : "l"(p));
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ uint32_t pack_bf16x2_relu(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

}  // namespace tc
}  // namespace nerf

// ---- CTA-pair (cta_group::2) primitives ---------------------------------------------------------------
namespace nerf {
namespace tc {

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cta address of this CTA -> shared::cluster address of the same offset in CTA `rank`
__device__ __forceinline__ uint32_t mapa_u32(uint32_t saddr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
// arrive on an mbarrier of another CTA of the cluster (address from mapa_u32); releases this thread's prior
// writes at cluster scope
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// same, default semantics (release at CTA scope): what CUTLASS's ClusterBarrier::arrive(cta_id) emits.  The
// .release.cluster form above costs MEMBAR.ALL.GPU + ERRBAR per arrive (profiles/mlp_r1b_hot_sass.txt); the
// CTA-scope form is enough when the data guarded is this CTA's own shared memory read by this SM's tensor core.
__device__ __forceinline__ void mbar_arrive_remote(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_cluster(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
#if NERF_MBAR_POLL == 1
      "mbarrier.test_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
#else
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
#endif
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// wait on a local barrier that CTAs of the whole cluster arrive on
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity, uint32_t tag) {
  if (mbar_try_wait_cluster(bar, parity)) return;
  const long long t0 = clock64();
  uint32_t spins = 0;
  while (!mbar_try_wait_cluster(bar, parity)) {
    if ((++spins & 0x3FFu) == 0 && clock64() - t0 > kWatchdogCycles) mbar_timeout(tag, parity);
  }
}

__device__ __forceinline__ void tmem_alloc2(uint32_t* smem_slot, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish2() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// D[tmem of both CTAs] (+)= A[smem: 128 rows per CTA] * B[smem: N/2 rows per CTA]^T; issued by one thread of
// the leader CTA (cluster rank 0); the descriptors name the same shared-memory offsets in both CTAs.
__device__ __forceinline__ void umma_ss2(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// completion of all prior MMAs of this thread -> arrive on the mbarrier at this offset in every CTA of `mask`
__device__ __forceinline__ void umma_commit2(uint64_t* bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)),
               "h"(mask)
               : "memory");
}

}  // namespace tc
}  // namespace nerf

// ---- bulk copy shared -> global (TMA engine) + named barriers ------------------------------------------
namespace nerf {
namespace tc {
__device__ __forceinline__ void bulk_s2g(void* gmem_dst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// all committed bulk stores of this thread have finished READING shared memory (it may be overwritten)
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// ... have completed entirely (required before the CTA exits)
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// NERF_EXP bit 8 (timing experiment, build with bits 0 + 8 = 257; the arena then holds the swizzled operand image, which wgrad
// does not understand): the warp's 32 rows of n_blocks A-operand blocks go to the arena as bulk async copies issued by one
// lane after the hand-over, instead of 32 st.global.v4 per row in the epilogue -- what would a bulk-store arena buy?
__device__ __forceinline__ void exp_bulk_rows_to_arena(const uint8_t* s_blk0, int q, int lane, uint8_t* g_blk0, int n_blocks) {
  if (!(NERF_EXP & 256)) return;
  if (lane == 0) {
    for (int b = 0; b < n_blocks; ++b) bulk_s2g(g_blk0 + b * 16384 + q * 4096, s_blk0 + b * 16384 + q * 4096, 4096);
    bulk_commit();
  }
}
__device__ __forceinline__ void exp_bulk_wait(int lane) {
  if (!(NERF_EXP & 256)) return;
  if (lane == 0) bulk_wait_read0();
  __syncwarp();
}
__device__ __forceinline__ void named_bar_sync(uint32_t id, uint32_t nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
// Warp-collective forms for the issuer warp: all 32 lanes run the (uniform) issue loop and one elected lane issues.  With the
// loop inside `if (lane == 0)` every operand lives in a per-thread register and each tcgen05 instruction costs an
// ELECT + R2UR.BROADCAST + BRA.U.ANY sequence (~22 SASS instructions, ~100 clk of the issuing thread per 128-clk MMA).
__device__ __forceinline__ void umma_ss2_w(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  if (NERF_EXP & 512) return;      // timing experiment: no MMAs at all (the commits then complete at once)
  asm volatile(
      "{\n\t.reg .pred p, pe;\n\t"
      "elect.sync _|pe, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@pe tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// The four K = 16 MMAs of one 64-wide k-block in ONE election: the operand descriptors advance by 32 bytes (2 address units)
// per MMA.  Issued one by one through umma_ss2_w, every MMA costs an ELECT, a VOTEU and six R2UR.BROADCAST (the descriptors
// are recomputed in vector registers: ~20 SASS instructions, ~600-850 clk of the issuing warp per 512-clk stage, which paced
// the per-tile order of the training kernels).
__device__ __forceinline__ void umma_ss2_w_x4(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate_first) {
  if (NERF_EXP & 512) return;
  asm volatile(
      "{\n\t.reg .pred p, pe;\n\t.reg .b64 da, db;\n\t"
      "elect.sync _|pe, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@pe tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "setp.eq.b32 p, 0, 0;\n\t"
      "add.u64 da, %1, 2;\n\tadd.u64 db, %2, 2;\n\t"
      "@pe tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %3, p;\n\t"
      "add.u64 da, %1, 4;\n\tadd.u64 db, %2, 4;\n\t"
      "@pe tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %3, p;\n\t"
      "add.u64 da, %1, 6;\n\tadd.u64 db, %2, 6;\n\t"
      "@pe tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate_first)
      : "memory");
}
__device__ __forceinline__ void umma_commit2_w(uint64_t* bar, uint16_t mask) {
  asm volatile(
      "{\n\t.reg .pred pe;\n\t"
      "elect.sync _|pe, 0xffffffff;\n\t"
      "@pe tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n\t}" ::"r"(
          smem_u32(bar)),
      "h"(mask)
      : "memory");
}

}  // namespace tc
}  // namespace nerf
