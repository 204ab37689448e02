// w4a16_common.cuh -- sm_100a building blocks shared by the W4A16 kernels:
// in-register INT4 -> fp16 dequant, mbarrier / bulk-copy / TMA / tcgen05 PTX wrappers.
//
// The dequant arithmetic is bit-exact with the reference
// (csrc/quantization/awq/dequantize.cuh:17-84 + gemm_kernels.cu:153-160):
//   W = fma.rn.f16( (q - z), s, +0 ).
// The zero point is folded into the magic-number subtraction, which stays exact:
//   low  nibbles: (1024 + q) - (1024 + z)                = q - z
//   high nibbles: fma(1024 + 16 q, 1/16, -(64 + z))      = q - z
// so a word costs 1 shift + 4 lop3 + 2 sub + 2 fma + 4 fma(scale) = 13 instructions
// instead of the reference's 17 (dequant 9 + 4 sub + 4 fma).
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace b200 {

// ----------------------------------------------------------------------------- dequant

// Nibble index of logical element j inside a packed word: {0,4,1,5,2,6,3,7}[j].
__host__ __device__ __forceinline__ int awq_nibble(int j) { return ((j & 1) << 2) | (j >> 1); }

__device__ __forceinline__ uint32_t lop3_and_or(uint32_t a, uint32_t mask, uint32_t orv) {
  uint32_t r;
  // (a & mask) | orv : LUT = (0xf0 & 0xcc) | 0xaa = 0xea
  asm("lop3.b32 %0, %1, %2, %3, 0xea;" : "=r"(r) : "r"(a), "r"(mask), "r"(orv));
  return r;
}
__device__ __forceinline__ uint32_t hsub2_u32(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("sub.rn.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ uint32_t hfma2_u32(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}

// Per-(group, channel-octet) constants for the AWQ-layout path: eight zeros and eight scales.
struct ZS8 {
  uint32_t zlo[2];  // half2(1024+z) for element pairs (0,1) and (4,5)
  uint32_t zhi[2];  // half2(-(64+z)) for element pairs (2,3) and (6,7)
  uint32_t s[4];    // half2 scales for pairs (0,1) (2,3) (4,5) (6,7)
};

// zeros word (eight zeros for eight consecutive channels) -> folded constants.
__device__ __forceinline__ void make_zs8(uint32_t zw, uint4 sc, ZS8& o) {
  const uint32_t t = zw >> 8;
  // pair (0,1): nibbles 0,4 ; pair (2,3): nibbles 1,5 ; pair (4,5): nibbles 2,6 ; (6,7): 3,7
  o.zlo[0] = (zw & 0x000f000fu) | 0x64006400u;            // 1024 + z
  o.zlo[1] = (t & 0x000f000fu) | 0x64006400u;
  o.zhi[0] = (zw & 0x00f000f0u) | 0xd400d400u;            // -(64 + z): 0xd400 + (z << 4)
  o.zhi[1] = (t & 0x00f000f0u) | 0xd400d400u;
  o.s[0] = sc.x; o.s[1] = sc.y; o.s[2] = sc.z; o.s[3] = sc.w;
}

// One packed word of eight channels (AWQ layout) -> eight fp16 (as 4 x half2) in logical order.
__device__ __forceinline__ uint4 dequant_word_zs8(uint32_t w, const ZS8& c) {
  const uint32_t t = w >> 8;
  uint32_t a0 = lop3_and_or(w, 0x000f000fu, 0x64006400u);
  uint32_t a1 = lop3_and_or(w, 0x00f000f0u, 0x64006400u);
  uint32_t a2 = lop3_and_or(t, 0x000f000fu, 0x64006400u);
  uint32_t a3 = lop3_and_or(t, 0x00f000f0u, 0x64006400u);
  a0 = hsub2_u32(a0, c.zlo[0]);
  a1 = hfma2_u32(a1, 0x2c002c00u, c.zhi[0]);
  a2 = hsub2_u32(a2, c.zlo[1]);
  a3 = hfma2_u32(a3, 0x2c002c00u, c.zhi[1]);
  uint4 r;
  r.x = hfma2_u32(a0, c.s[0], 0u);
  r.y = hfma2_u32(a1, c.s[1], 0u);
  r.z = hfma2_u32(a2, c.s[2], 0u);
  r.w = hfma2_u32(a3, c.s[3], 0u);
  return r;
}

// B200 layout: a word holds eight K-rows of ONE channel, so zero and scale are scalars.
//   zlo = half2(1024+z), zhi = half2(-(64+z)), s2 = half2(s,s)
__device__ __forceinline__ void make_zs1(uint32_t z, uint16_t s, uint32_t& zlo, uint32_t& zhi,
                                         uint32_t& s2) {
  zlo = 0x64006400u + z * 0x00010001u;
  zhi = 0xd400d400u + (z << 4) * 0x00010001u;
  s2 = (uint32_t)s * 0x00010001u;
}
__device__ __forceinline__ void dequant_word_zs1(uint32_t w, uint32_t zlo, uint32_t zhi,
                                                 uint32_t s2, uint32_t& o0, uint32_t& o1,
                                                 uint32_t& o2, uint32_t& o3) {
  const uint32_t t = w >> 8;
  uint32_t a0 = lop3_and_or(w, 0x000f000fu, 0x64006400u);
  uint32_t a1 = lop3_and_or(w, 0x00f000f0u, 0x64006400u);
  uint32_t a2 = lop3_and_or(t, 0x000f000fu, 0x64006400u);
  uint32_t a3 = lop3_and_or(t, 0x00f000f0u, 0x64006400u);
  a0 = hsub2_u32(a0, zlo);
  a1 = hfma2_u32(a1, 0x2c002c00u, zhi);
  a2 = hsub2_u32(a2, zlo);
  a3 = hfma2_u32(a3, 0x2c002c00u, zhi);
  o0 = hfma2_u32(a0, s2, 0u);  // rows (0,1)
  o1 = hfma2_u32(a1, s2, 0u);  // rows (2,3)
  o2 = hfma2_u32(a2, s2, 0u);  // rows (4,5)
  o3 = hfma2_u32(a3, s2, 0u);  // rows (6,7)
}

// One packed word -> (q - z) for its eight K rows as four half2 (exact; NO scale): the decode
// variants apply the group scale to the fp32 accumulator instead (9 instead of 13 operations).
__device__ __forceinline__ void unpack_sub_zero(uint32_t w, uint32_t zlo, uint32_t zhi,
                                                uint32_t& o0, uint32_t& o1, uint32_t& o2,
                                                uint32_t& o3) {
  const uint32_t t = w >> 8;
  o0 = hsub2_u32(lop3_and_or(w, 0x000f000fu, 0x64006400u), zlo);               // rows (0,1)
  o1 = hfma2_u32(lop3_and_or(w, 0x00f000f0u, 0x64006400u), 0x2c002c00u, zhi);  // rows (2,3)
  o2 = hsub2_u32(lop3_and_or(t, 0x000f000fu, 0x64006400u), zlo);               // rows (4,5)
  o3 = hfma2_u32(lop3_and_or(t, 0x00f000f0u, 0x64006400u), 0x2c002c00u, zhi);  // rows (6,7)
}

// Zero point of channel c (0..127 inside a tile) from a row of 16 packed zero words.
__device__ __forceinline__ uint32_t zero_of_channel(const uint32_t* zrow, int c) {
  return (zrow[c >> 3] >> (4 * awq_nibble(c & 7))) & 0xFu;
}

// ----------------------------------------------------------------------------- smem / mbarrier

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}

// Bounded waits.  Inside one GPU a pipeline bug must surface as a trapped launch, never as a
// hung device: B200_WATCHDOG_CYCLES (~2 s) bounds every wait on this GPU's own CTAs.  Waits on
// ANOTHER rank are different: seconds of skew are ordinary in serving (uneven checkpoint
// load, a GC pause, a debugger), the reference's kernels and NCCL simply wait -- so every
// cross-rank spin shares B200_PEER_WATCHDOG_CYCLES, ~10 minutes by default (0 = wait forever),
// and traps only as a last resort against a dead peer.
#ifndef B200_WATCHDOG_CYCLES
#define B200_WATCHDOG_CYCLES (4000000000ll)  // ~2 s at 1.9 GHz
#endif
#ifndef B200_PEER_WATCHDOG_CYCLES
#define B200_PEER_WATCHDOG_CYCLES (1200000000000ll)  // ~10 min at 1.9 GHz
#endif
__device__ __forceinline__ bool peer_wait_expired(long long t0) {
  return B200_PEER_WATCHDOG_CYCLES > 0 && clock64() - t0 > B200_PEER_WATCHDOG_CYCLES;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > B200_WATCHDOG_CYCLES) {
      printf("b200_w4a16: mbarrier timeout block %d thread %d bar 0x%x parity %u\n", blockIdx.x,
             threadIdx.x, bar, parity);
      __trap();
    }
  }
}

// ----------------------------------------------------------------------------- async copies

// 1-D bulk copy global -> shared, completion counted on an mbarrier (bytes % 16 == 0).
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes,
                                         uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
      : "memory");
}
// Same with an L2 eviction-priority hint (weights are streamed once: evict_first).
__device__ __forceinline__ void bulk_g2s_hint(uint32_t dst, const void* src, uint32_t bytes,
                                              uint32_t bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1], %2, [%3], %4;"
      ::"r"(dst), "l"(src), "r"(bytes), "r"(bar), "l"(policy)
      : "memory");
}
__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t policy_evict_normal() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
// 2-D tiled TMA load (tensor map in kernel param space).
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const void* tmap, int c0, int c1,
                                            uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%2, %3}], [%4];"
      ::"r"(dst), "l"(tmap), "r"(c0), "r"(c1), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}

// ----------------------------------------------------------------------------- tcgen05 / TMEM

template <int COLS>
__device__ __forceinline__ void tmem_alloc(uint32_t smem_slot) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_slot),
               "n"(COLS)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS)
               : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_wait_st() {
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tc_wait_ld() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// MMA-completion -> mbarrier arrive (implies tcgen05.fence::before_thread_sync).
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar)
               : "memory");
}
// D[tmem] (+)= A[tmem] . B[smem desc]   (kind::f16, fp32 accumulate)
__device__ __forceinline__ void tc_mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc,
                                          uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// D[tmem] (+)= A[smem desc] . B[smem desc]
__device__ __forceinline__ void tc_mma_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                          uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// Instruction descriptor, kind::f16: fp16 A/B (K-major), fp32 D, M=128, N=n.
// Bit layout per cute/arch/mma_sm100_desc.hpp (InstrDescriptor): c_format[4,6) a_format[7,10)
// b_format[10,13) a_major[15] b_major[16] n>>3 [17,23) m>>4 [24,29).
__host__ __device__ constexpr uint32_t make_idesc_f16_m128(int n) {
  return (1u << 4) | (0u << 7) | (0u << 10) | (0u << 15) | (0u << 16) |
         ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}
// Shared-memory operand descriptor, K-major, 128-byte swizzle, 8-row groups 1024 B apart
// (SmemDescriptor: start>>4 [0,14) LBO>>4 [16,30) SBO>>4 [32,46) version=1 [46,48) layout=2 [61,64)).
__device__ __forceinline__ uint64_t make_desc_k_sw128(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) |
         (1ull << 46) | (2ull << 61);
}

// TMEM -> registers, 32 lanes x 32 bit, N consecutive columns per thread.
__device__ __forceinline__ void tmem_ld_x8(uint32_t taddr, uint32_t (&v)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]),
                 "=r"(v[6]), "=r"(v[7])
               : "r"(taddr)
               : "memory");
}
// TMEM -> registers, 16 consecutive columns per thread.
__device__ __forceinline__ void tmem_ld_x16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
        "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),
        "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
// registers -> TMEM, 16 columns per thread.
__device__ __forceinline__ void tmem_st_x16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]),
      "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]),
      "r"(v[14]), "r"(v[15])
      : "memory");
}
// registers -> TMEM, 32 columns per thread.
__device__ __forceinline__ void tmem_st_x32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
      "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]),
      "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]),
      "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]),
      "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]),
      "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
      : "memory");
}

// ----------------------------------------------------------------------------- CTA pairs (cta_group::2)
// Two CTAs of a cluster on the SMs of one TPC issue ONE tcgen05.mma of M = 256: each holds 128
// rows of A (its own TMEM) and of D, and HALF of the B operand (N / 2 rows) in its own shared
// memory.  Only the leader (cluster rank 0) issues MMAs and owns the barriers the MMA thread
// waits on; the peer signals them remotely.  Forms as in CUTLASS (cute/arch/mma_sm100_umma.hpp
// SM100_MMA_F16BF16_2x1SM_TS, copy_sm100_tma.hpp SM100_TMA_2SM_LOAD_2D, cutlass/arch/barrier.h
// umma_arrive_multicast_2x1SM, tmem_allocator_sm100.hpp Allocator2Sm).
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cta address -> the same offset in CTA `rank` of the cluster (shared::cluster address)
__device__ __forceinline__ uint32_t map_to_cta(uint32_t saddr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
// Remote arrive with the default (CTA-scope release) semantics, as CUTLASS's ClusterBarrier does:
// an explicit .release.cluster costs a full drain of the SM's memory pipeline per arrive -- every
// dequant group of the peer CTA then finished ~2000 cycles late, all at the same instant
// (profiles/r2u_trace_prefill_two_release_cluster.txt).  The TMEM data it publishes is ordered by
// tcgen05.wait::st + tcgen05.fence::before_thread_sync on this side and fence::after_thread_sync
// on the MMA issuer's.
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_alloc_pair(uint32_t smem_slot) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_slot),
               "n"(COLS)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS)
               : "memory");
}
// MMA completion -> arrive on the barrier at this offset in BOTH CTAs of the pair.
__device__ __forceinline__ void tc_commit_pair(uint32_t bar) {
  const uint16_t mask = 3;
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      ::"r"(bar), "h"(mask)
      : "memory");
}
// D[tmem, both CTAs] (+)= A[tmem, 128 rows per CTA] . B[smem desc, N / 2 rows per CTA]
__device__ __forceinline__ void tc_mma_ts_pair(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc,
                                               uint32_t idesc, uint32_t accumulate) {
  const uint32_t z = 0;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, {%5, %5, %5, %5, %5, %5, %5, %5}, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate), "r"(z)
      : "memory");
}
// 2-D tiled TMA load whose bytes are counted on the LEADER's barrier (peer bit cleared).
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const void* tmap, int c0, int c1,
                                                 uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%2, %3}], [%4];"
      ::"r"(dst), "l"(tmap), "r"(c0), "r"(c1), "r"(bar & 0xFEFFFFFFu)
      : "memory");
}
// Instruction descriptor, kind::f16, fp16 A/B (K-major), fp32 D, M = m (128 or 256 for a pair), N = n.
__host__ __device__ constexpr uint32_t make_idesc_f16(int m, int n) {
  return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// One thread of a converged warp.  Code under `if (elect_one())` is known by the compiler to
// run on a single thread, so operands of UBLKCP / UTMALDG / UTCHMMA move to uniform registers
// directly instead of through a per-instruction waterfall loop (which `lane == 0` triggers).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "elect.sync _|P, 0xffffffff;\n\t"
      "selp.b32 %0, 1, 0, P;\n\t}"
      : "=r"(pred));
  return pred != 0;
}

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// Device-scope acquire-release counter bump: publishes everything this CTA wrote before the
// preceding barrier and makes the other arrivers' earlier writes visible after the next one.
__device__ __forceinline__ int atom_add_acq_rel_gpu(int* addr, int v) {
  int old;
  asm volatile("atom.add.acq_rel.gpu.global.s32 %0, [%1], %2;" : "=r"(old) : "l"(addr), "r"(v)
               : "memory");
  return old;
}

// Fire-and-forget publish / polling read used by the deferred stream-K fix-up.
__device__ __forceinline__ void red_add_release_gpu(int* addr, int v) {
  asm volatile("red.release.gpu.global.add.s32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_gpu(const int* addr) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
  return v;
}

// 16-byte accesses for data another CTA (or another GPU) writes while we poll it.
//  * `volatile` = strong, system scope (SASS LDG/STG.E.128.STRONG.SYS): always performed at
//    the coherence point; what the NVLink packets and every RE-poll use.  Strong loads of one
//    thread complete one after the other, though -- a chain of L2 round trips as long as the
//    number of contributors (measured: profiles/r2c_trace_decode_4096.log, 2.2 us).
//  * `weak_na` = weak, L1 no-allocate (LDG.E.NA.128): tuning builds only (flag 2).  It was
//    tried for the first look at a slot and REJECTED: it neither shortened the fix-up nor stayed
//    correct (stale words made back-to-back launches differ, profiles/r2e notes in DESIGN.md).
__device__ __forceinline__ uint4 ld_volatile_u4(const void* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_volatile_u4(void* p, uint4 v) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_weak_na_u4(const void* p) {
  uint4 v;
  asm volatile("ld.weak.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
// First look at a polled slot: strong like every other poll (flag 2 of a tuning build: weak).
constexpr int kFlagNoEvictFirst = 1, kFlagWeakFirstPoll = 2;
__device__ __forceinline__ uint4 ld_first_poll_u4(const void* p, int flags) {
  return (flags & kFlagWeakFirstPoll) ? ld_weak_na_u4(p) : ld_volatile_u4(p);
}

// Stream-K partial tiles travel through L2 as SELF-FLAGGING fp32 words: a slot that has not been
// written holds kPartEmpty (an fp32 NaN bit pattern no arithmetic produces: hardware NaNs are
// 0x7fffffff) in every word; the contributor stores its sums (a real value can never alias the
// marker: part_word clamps), the owner polls the words themselves and puts the marker back.
// One L2 hop, no fence, no counter, no flag -- and the workspace is left as it was found.
constexpr uint32_t kPartEmpty = 0xFFFFFFFFu;
__device__ __forceinline__ uint32_t part_word(float v) { return min(__float_as_uint(v), 0xFFFFFFFEu); }
__device__ __forceinline__ uint4 part_pack(float a, float b, float c, float d) {
  return make_uint4(part_word(a), part_word(b), part_word(c), part_word(d));
}
__device__ __forceinline__ bool part_ready(const uint4& v) {
  return v.x != kPartEmpty && v.y != kPartEmpty && v.z != kPartEmpty && v.w != kPartEmpty;
}
// Wait for one 16-byte group of a contributor's partial tile, take it, restore the marker.
__device__ __forceinline__ float4 part_take(uint4* slot, uint4 first) {
  uint4 v = first;
  if (!part_ready(v)) {
    const long long t0 = clock64();
    do {
      v = ld_volatile_u4(slot);
      if (clock64() - t0 > B200_WATCHDOG_CYCLES) {
        printf("b200_w4a16: stream-K partial timeout block %d thread %d\n", blockIdx.x, threadIdx.x);
        __trap();
      }
    } while (!part_ready(v));
  }
  st_volatile_u4(slot, make_uint4(kPartEmpty, kPartEmpty, kPartEmpty, kPartEmpty));
  return make_float4(__uint_as_float(v.x), __uint_as_float(v.y), __uint_as_float(v.z),
                     __uint_as_float(v.w));
}

// silu(gate) * up on two GEMM outputs, exactly as the unfused pair computes it: the GEMM rounds
// both to fp16, then csrc/activation_kernels.cu:10-27 of the reference --
// (half)(float(x) / (1 + expf(-float(x)))) * y, a half multiply.  Shared by the standalone
// kernel and the decode GEMM's epilogue so the two agree bit for bit.
__device__ __forceinline__ uint16_t silu_mul_f16(float gate, float up) {
  const __half gh = __float2half_rn(gate), uh = __float2half_rn(up);
  const float gf = __half2float(gh);
  const __half sh = __float2half_rn(gf / (1.0f + expf(-gf)));
  return __half_as_ushort(__hmul(sh, uh));
}

}  // namespace b200
