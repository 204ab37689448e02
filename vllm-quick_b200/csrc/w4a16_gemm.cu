// w4a16_gemm.cu -- the W4A16 group-quantized GEMM for sm_100a.
//
//   y[M,N] (fp16) = x[M,K] (fp16) . dequant(qweight, scales, qzeros)[K,N]     fp32 accumulate
//
// Replaces the reference's gemm_forward_4bit_cuda_m16nXk32 + Tensor.sum(0)
// (csrc/quantization/awq/gemm_kernels.cu:30-272, :389-446) and its dequantize + cuBLAS branch
// (vllm/model_executor/layers/quantization/awq.py:154-159).  Design:
//
//  * swap-AB: the tcgen05 "M" dimension (128 TMEM lanes) is OUTPUT CHANNELS, the tcgen05 "N"
//    dimension is TOKENS (16..128), so a decode batch of 1..16 tokens still fills the MMA.
//  * weights arrive as contiguous 8 KiB tiles (128 channels x 128 rows, B200 layout) by 1-D
//    bulk async copy; activations by 2-D TMA with the 128-byte swizzle tcgen05 expects.
//  * four dequant warpgroups take tiles round-robin.  A thread reads one 16-byte word group
//    (bank-conflict free), expands it in registers with the reference's exact lop3 / sub / fma
//    sequence, and writes the fp16 pairs straight into TENSOR MEMORY with tcgen05.st:
//    thread = TMEM lane = channel, so the registers already are the MMA A-operand -- no
//    shared-memory round trip for the dequantized tile (the reference's :168 store +
//    :198-203 ldmatrix).  A ring of 4-6 operand buffers decouples them from the MMA.
//  * one thread issues tcgen05.mma (A from TMEM, B from shared memory); two fp32 accumulator
//    buffers in TMEM let a dedicated epilogue warpgroup drain tile i while tile i+1 runs.
//  * stream-K: the (channel tile, token tile, K block) space is cut into equal contiguous
//    ranges, one per SM.  A CTA does its two partial output tiles FIRST, so the cross-CTA
//    reduction (fp32 partials through L2, fixed CTA order, done by whichever CTA arrives
//    last -- no second kernel, deterministic) overlaps the rest of the weight stream.
//  * programmatic dependent launch: the weight pipeline is filled BEFORE griddepcontrol.wait
//    (weights never depend on the previous kernel), so launch ramp overlaps the previous op.
#include "w4a16_common.cuh"
#include "w4a16_internal.h"
#include "w4a16_streamk.cuh"

#include <stdlib.h>

namespace b200 {

namespace {

constexpr int kProducerWarp = 0;
constexpr int kMmaWarp = 1;
constexpr int kAllocWarp = 2;
constexpr int kXProducerWarp = 3;
constexpr int kFirstEpiWarp = 4;                        // warps 4..7 (warp % 4 == lane quarter)
constexpr int kEpiThreads = 128;
constexpr int kFirstDequantWarp = 8;
constexpr int kGroups = 4;                              // dequant warpgroups, tiles round-robin
constexpr int kDequantThreads = kGroups * 128;
constexpr int kThreads = kFirstDequantWarp * 32 + kDequantThreads;  // 768
constexpr int kScaleRowsMax = 4;                        // G = 32 -> four (scale, zero) rows / tile
constexpr int kTmemACols = 64;                          // one dequantized tile: 128 fp16 per lane
constexpr int kTmemCols = 512;
constexpr int kMaxSmem = 227 * 1024;

template <int MTOK>
struct Cfg {
  static constexpr int kABufs = MTOK <= 64 ? 6 : 4;    // dequantized-operand ring in TMEM
  static constexpr int kDCol = kABufs * kTmemACols;    // accumulators start here
  static constexpr int kDStride = MTOK <= 64 ? 64 : 128;  // two accumulator buffers
  static constexpr int kXBytes = MTOK * 256;           // two swizzle atoms of [MTOK x 64] fp16
  static constexpr int kWOff = kXBytes;
  static constexpr int kSOff = kWOff + kTileWords * 4;
  static constexpr int kZOff = kSOff + kScaleRowsMax * 256;
  static constexpr int kStageBytes = ((kZOff + kScaleRowsMax * 64 + 1023) / 1024) * 1024;
  static constexpr int kBarBytes = 1024;
  static constexpr int kStagesFit = (kMaxSmem - kBarBytes - 1024) / kStageBytes;
  static constexpr int kStages = kStagesFit > 16 ? 16 : kStagesFit;
  static constexpr int kSmemBytes = kBarBytes + kStages * kStageBytes + 1024;  // + align slack
  // A tcgen05.commit costs the tensor pipe ~110 cycles (scripts/probes/mma_probe.cu), so MMA
  // completion is signalled once per kCommit tiles on a ring of kDoneBars barriers; both the
  // producer (stage reuse) and the dequant groups (operand-buffer reuse) wait on that ring.
  static constexpr int kCommit = 2;
  static constexpr int kRingMax = kStages > kABufs ? kStages : kABufs;
  static constexpr int kDoneBars = (kRingMax + kCommit - 1) / kCommit + 1;
  static_assert(kStages > kGroups, "stage ring must be deeper than the dequant group count");
  static_assert(kABufs <= kStages, "operand-ring wait must imply the stage's previous phase landed");
  static_assert(kDCol + 2 * kDStride <= kTmemCols, "TMEM budget");
};

struct GemmParams {
  const uint32_t* qw;
  const uint16_t* scales;
  const uint32_t* qz;
  uint16_t* y;
  float* partials;
  int* counters;
  int M, K, N, G;
  int KB, NT, MT;
  int T;        // total (output tile, K block) work items
  int P;        // CTAs
  int R;        // (scale, zero) rows per K block: max(1, 128 / G)
  int Gdiv;     // K blocks per group: max(1, G / 128)
  int flags;    // bit 0: no L2 evict-first hint on the weight stream
  uint32_t* trace;  // B200_TRACE builds only: per-CTA event timestamps
};

#ifdef B200_TRACE
// event types: 0 producer issues tile, 1 dequant sees tile landed, 2 dequant starts (operand
// buffer free), 3 dequant done, 4 MMA starts tile, 5 MMA issued, 6 epilogue seg (2*idx: dfull
// seen, 2*idx+1: segment finished incl. fix-up), 7 misc (0 setup done, 1 exit)
#define TR(type, i)                                                                          \
  do {                                                                                       \
    if (p.trace && (i) < 64) p.trace[((size_t)blockIdx.x * 8 + (type)) * 64 + (i)] = (uint32_t)clock64(); \
  } while (0)
#define TRG(i)                                                                               \
  do {                                                                                       \
    if (p.trace) {                                                                           \
      unsigned long long gt_;                                                                \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt_));                                \
      p.trace[((size_t)blockIdx.x * 8 + 7) * 64 + (i)] = (uint32_t)gt_;                      \
    }                                                                                        \
  } while (0)
#else
#define TR(type, i) do { } while (0)
#define TRG(i) do { } while (0)
#endif

template <int MTOK>
__global__ void __launch_bounds__(kThreads, 1)
w4a16_gemm_kernel(const GemmParams p, const __grid_constant__ CUtensorMap tmap_x) {
  using C = Cfg<MTOK>;
  if (threadIdx.x == 0) { TRG(2); TR(7, 5); }
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));

  // barrier block
  const uint32_t bar_full = smem_base;                          // [kStages]   TMA landed
  const uint32_t bar_done = bar_full + 8 * C::kStages;          // [kDoneBars] MMA done, per group
  const uint32_t bar_afull = bar_done + 8 * C::kDoneBars;       // [kABufs]    operand dequantized
  const uint32_t bar_dfull = bar_afull + 8 * C::kABufs;         // [2] accumulator complete
  const uint32_t bar_dempty = bar_dfull + 16;                   // [2] accumulator drained
  constexpr int kNumBars = C::kStages + C::kDoneBars + C::kABufs + 4;
  const uint32_t tmem_slot = smem_base + 8 * kNumBars;
  static_assert(8 * kNumBars + 8 <= C::kBarBytes, "barrier block");
  volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(
      smem_gen + (tmem_slot - smem_base));
  const uint32_t stage0 = smem_base + C::kBarBytes;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // This CTA's contiguous slice of the work space, and the order it is walked in.
  const int t_begin = range_begin(blockIdx.x, p.T, p.P);
  const int t_end = range_begin(blockIdx.x + 1, p.T, p.P);
  const Walk walk(t_begin, t_end, p.KB);
  const int n_total = t_end - t_begin;
  const int npre = n_total < C::kStages ? n_total : C::kStages;

  // Weight stream of this CTA (elected thread of warp 0): tiles in walk order with addresses
  // advanced incrementally -- consecutive K blocks of a channel tile are contiguous in the
  // B200 layout, so the steady state is an add, an mbarrier wait and three bulk copies.
  struct WStream {
    const GemmParams& p;
    const Walk& walk;
    int idx, left, nt, kb;
    uint32_t tx;
    const uint32_t* wsrc;
    const uint16_t* s0;
    const uint32_t* z0;
    uint32_t nvalid;
    __device__ __forceinline__ WStream(const GemmParams& pp, const Walk& w)
        : p(pp), walk(w), idx(-1), left(0), nt(0), kb(0), tx(0), wsrc(nullptr), s0(nullptr),
          z0(nullptr), nvalid(0) {}
    __device__ __forceinline__ void open() {
      if (++idx >= walk.nseg) return;
      int t0, n;
      walk.seg(idx, t0, n);
      const int o = t0 / p.KB;
      kb = t0 - o * p.KB;
      nt = o / p.MT;
      left = n;
      nvalid = (uint32_t)min(kTileN, p.N - nt * kTileN);
      tx = (uint32_t)Cfg<MTOK>::kXBytes + kTileWords * 4 + (uint32_t)p.R * (nvalid * 2 + nvalid / 2);
      wsrc = p.qw + ((size_t)nt * p.KB + kb) * kTileWords;
      s0 = p.scales + nt * kTileN;
      z0 = p.qz + nt * 16;
    }
    __device__ __forceinline__ void next() {
      ++kb;
      wsrc += kTileWords;
      if (--left == 0) open();
    }
  };
  const uint64_t pol_w = (p.flags & 1) ? policy_evict_normal() : policy_evict_first();
  auto issue_w = [&](const WStream& c, int stage) {
    const uint32_t sb = stage0 + stage * C::kStageBytes;
    const uint32_t bar = bar_full + 8 * stage;
    mbar_arrive_expect_tx(bar, c.tx);
    bulk_g2s_hint(sb + C::kWOff, c.wsrc, kTileWords * 4, bar, pol_w);
    const uint32_t g0 = p.R > 1 ? (uint32_t)(c.kb * p.R)
                                : (p.Gdiv == 1 ? (uint32_t)c.kb : (uint32_t)c.kb / (uint32_t)p.Gdiv);
    for (int r = 0; r < p.R; ++r) {
      bulk_g2s(sb + C::kSOff + r * 256, c.s0 + (size_t)(g0 + r) * p.N, c.nvalid * 2, bar);
      bulk_g2s(sb + C::kZOff + r * 64, c.z0 + (size_t)(g0 + r) * (p.N >> 3), c.nvalid / 2, bar);
    }
  };

  WStream cw(p, walk);                                      // opened by the producer thread only
  if (warp == kProducerWarp) {
    // Lane-parallel barrier init, then the weight side of the first npre stages goes out
    // BEFORE the CTA-wide setup barrier and without ever waiting for the previous kernel:
    // weights never depend on it, so launch ramp and TMEM allocation overlap the first fetch.
    for (int b = lane; b < kNumBars; b += 32) {
      const uint32_t count = b < C::kStages + C::kDoneBars ? 1u
                             : b < C::kStages + C::kDoneBars + C::kABufs ? 128u
                             : b < kNumBars - 2 ? 1u : (uint32_t)kEpiThreads;
      mbar_init(smem_base + 8 * b, count);
    }
    fence_barrier_init();
    __syncwarp();
    if (elect_one()) {
      // Let the next kernel in the stream start its own weight prefetch as soon as an SM frees.
      asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
      cw.open();
      for (int i = 0; i < npre; ++i) { TR(0, i); issue_w(cw, i); cw.next(); }
    }
    __syncwarp();
  } else if (warp == kAllocWarp) {
    tmem_alloc<kTmemCols>(tmem_slot);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot_gen;
  if (threadIdx.x == 0) { TR(7, 0); TRG(3); }

  if (warp == kProducerWarp) {
    // -- steady state: refill stage it % kStages once the MMAs of tile it - kStages are done.
    // elect.sync picks the same lane as before the setup barrier, so `cw` carries over.
    __syncwarp();
    if (elect_one()) {
      int stage = npre % C::kStages;
      for (int it = npre; it < n_total; ++it) {
        const int g = (it - C::kStages) / C::kCommit;
        mbar_wait(bar_done + 8 * (g % C::kDoneBars), (g / C::kDoneBars) & 1);
        TR(0, it);
        issue_w(cw, stage);
        cw.next();
        if (++stage == C::kStages) stage = 0;
      }
    }
  } else if (warp == kXProducerWarp) {
    // Activation stream (elected thread of warp 3): two 64-column TMA boxes per tile, issued
    // once the previous kernel's output is visible.
    __syncwarp();
    if (elect_one()) {
      tma_prefetch_desc(&tmap_x);
      asm volatile("griddepcontrol.wait;" ::: "memory");      // x is the previous kernel's output
      int it = 0;
      for (int idx = 0; idx < walk.nseg; ++idx) {
        int t0, n;
        walk.seg(idx, t0, n);
        const int o = t0 / p.KB;
        int kb = t0 - o * p.KB;
        const int m0 = (o % p.MT) * MTOK;
        for (int i = 0; i < n; ++i, ++it, ++kb) {
          const int stage = it % C::kStages;
          if (it >= C::kStages) {
            const int g = (it - C::kStages) / C::kCommit;
            mbar_wait(bar_done + 8 * (g % C::kDoneBars), (g / C::kDoneBars) & 1);
          }
          const uint32_t sb = stage0 + stage * C::kStageBytes;
          const uint32_t bar = bar_full + 8 * stage;
          tma_load_2d(sb, &tmap_x, kb * kTileK, m0, bar);
          tma_load_2d(sb + MTOK * 128, &tmap_x, kb * kTileK + 64, m0, bar);
        }
      }
    }
  } else if (warp == kMmaWarp) {
    __syncwarp();
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_f16_m128(MTOK);
      int stage = 0, buf = 0, itg = 0;
      uint32_t aphase = 0;
      for (int idx = 0; idx < walk.nseg; ++idx) {
        int t0, n;
        walk.seg(idx, t0, n);
        const int d = idx & 1;
        mbar_wait(bar_dempty + 8 * d, ((idx >> 1) & 1) ^ 1);  // drained by the epilogue warps
        tc_fence_after();
        const uint32_t dcol = tmem + C::kDCol + d * C::kDStride;
        for (int i = 0; i < n; ++i) {
          // One wait per tile: the dequant group arrives on bar_afull only after it has seen
          // this stage's bar_full complete (weights, scales, zeros AND the activation TMA
          // share that barrier), so the activations are covered transitively -- a second
          // ~100-cycle try_wait here made the issuing thread, not the tensor pipe, the
          // bottleneck at small token tiles.
          mbar_wait(bar_afull + 8 * buf, aphase);
          tc_fence_after();
          TR(4, itg);
          const uint32_t xb = stage0 + stage * C::kStageBytes;
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            const uint64_t bdesc = make_desc_k_sw128(xb + (kk >> 2) * (MTOK * 128) + (kk & 3) * 32);
            tc_mma_ts(dcol, tmem + buf * kTmemACols + kk * 8, bdesc, idesc, (i | kk) != 0);
          }
          if ((itg + 1) % C::kCommit == 0 || itg == n_total - 1) {
            const int g = itg / C::kCommit;
            tc_commit(bar_done + 8 * (g % C::kDoneBars));
          }
          TR(5, itg);
          ++itg;
          if (++stage == C::kStages) stage = 0;
          if (++buf == C::kABufs) { buf = 0; aphase ^= 1u; }
        }
        tc_commit(bar_dfull + 8 * d);
      }
    }
  } else if (warp >= kFirstDequantWarp) {
    // ---- dequant warpgroups: tile it -> group it % 4, stage it % kStages, operand it % kABufs
    const int wi = warp - kFirstDequantWarp;
    const int grp = wi >> 2, q = wi & 3;
    const int c = q * 32 + lane;                              // channel inside the tile
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    const int zsh = 4 * awq_nibble(c & 7);
    int stage = grp % C::kStages, buf = grp % C::kABufs;
    uint32_t fphase = 0;
    for (int it = grp; it < n_total; it += kGroups) {
      const uint8_t* sb = smem_gen + C::kBarBytes + stage * C::kStageBytes;
      // Operand buffer it % kABufs was last read by the MMAs of tile it - kABufs.  This wait
      // comes FIRST on purpose: the stage this tile sits in was last used by tile it - kStages
      // (< it - kABufs + 1), and a parity wait on its full barrier is only meaningful once
      // that earlier phase has completed -- bulk copies of different stages can land out of
      // order, and a group that runs ahead would otherwise see the stale phase as "landed".
      if (it >= C::kABufs) {
        const int g = (it - C::kABufs) / C::kCommit;
        mbar_wait(bar_done + 8 * (g % C::kDoneBars), (g / C::kDoneBars) & 1);
      }
      mbar_wait(bar_full + 8 * stage, fphase);
      if ((threadIdx.x & 127) == 0) TR(1, it);
      uint4 wv[4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        wv[j] = *reinterpret_cast<const uint4*>(sb + C::kWOff + j * 2048 + c * 16);
      const uint16_t* sS = reinterpret_cast<const uint16_t*>(sb + C::kSOff);
      const uint32_t* sZ = reinterpret_cast<const uint32_t*>(sb + C::kZOff);
      uint32_t zlo = 0, zhi = 0, s2 = 0;
      make_zs1((sZ[c >> 3] >> zsh) & 0xFu, sS[c], zlo, zhi, s2);
      tc_fence_after();
      if ((threadIdx.x & 127) == 0) TR(2, it);
      const uint32_t abase = tmem + lane_base + buf * kTmemACols;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (p.R > 1) {
          const int r = (j * 32) / p.G;
          if (j == 0 || r != ((j - 1) * 32) / p.G)
            make_zs1((sZ[r * 16 + (c >> 3)] >> zsh) & 0xFu, sS[r * 128 + c], zlo, zhi, s2);
        }
        uint32_t v[16];
        dequant_word_zs1(wv[j].x, zlo, zhi, s2, v[0], v[1], v[2], v[3]);
        dequant_word_zs1(wv[j].y, zlo, zhi, s2, v[4], v[5], v[6], v[7]);
        dequant_word_zs1(wv[j].z, zlo, zhi, s2, v[8], v[9], v[10], v[11]);
        dequant_word_zs1(wv[j].w, zlo, zhi, s2, v[12], v[13], v[14], v[15]);
        tmem_st_x16(abase + j * 16, v);
      }
      tc_wait_st();
      tc_fence_before();
      mbar_arrive(bar_afull + 8 * buf);
      if ((threadIdx.x & 127) == 0) TR(3, it);
      stage += kGroups;
      if (stage >= C::kStages) { stage -= C::kStages; fphase ^= 1u; }
      buf += kGroups;
      if (buf >= C::kABufs) buf -= C::kABufs;
    }
  } else if (warp >= kFirstEpiWarp) {
    // ---- epilogue warpgroup: drains accumulators; a whole tile goes to y, a partial one is
    // published as fp32 and finished after the main loop by the CTA that owns the tile's first
    // K block (deferred stream-K fix-up: nothing here waits on another CTA mid-stream)
    const int q = warp - kFirstEpiWarp;
    const int c = q * 32 + lane;                              // channel inside the tile
    const int et = threadIdx.x - kFirstEpiWarp * 32;          // 0..127
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    constexpr int kPartVec = MTOK * kTileN / 4;               // float4 per partial tile
    int own_o = -1, own_slot = 0, own_last = 0;               // tile this CTA finishes, if any
    asm volatile("griddepcontrol.wait;" ::: "memory");        // y / workspace may alias earlier work
    for (int idx = 0; idx < walk.nseg; ++idx) {
      int t0, n;
      walk.seg(idx, t0, n);
      const int o = t0 / p.KB;
      const int nt = o / p.MT, mt = o - nt * p.MT;
      const int d = idx & 1;
      const bool whole = (n == p.KB);
      const int n_glob = nt * kTileN + c;
      const int m0 = mt * MTOK;
      const uint32_t dbase = tmem + lane_base + C::kDCol + d * C::kDStride;
      const int slot = idx == 0 ? 0 : 1;
      // partial layout: float4 v of channel c holds tokens 4 v .. 4 v + 3 -> [v][c], coalesced
      float4* part = reinterpret_cast<float4*>(p.partials) +
                     ((size_t)blockIdx.x * 2 + slot) * kPartVec + c;
      mbar_wait(bar_dfull + 8 * d, (idx >> 1) & 1);
      tc_fence_after();
      if (et == 0) TR(6, 2 * idx);
#pragma unroll 1
      for (int cc = 0; cc < MTOK; cc += 16) {
        uint32_t v[16];
        tmem_ld_x16(dbase + cc, v);
        tc_wait_ld();
        if (whole) {
          if (n_glob < p.N) {
#pragma unroll
            for (int e = 0; e < 16; ++e) {
              const int m = m0 + cc + e;
              if (m < p.M)
                p.y[(size_t)m * p.N + n_glob] =
                    __half_as_ushort(__float2half_rn(__uint_as_float(v[e])));
            }
          }
        } else {
#pragma unroll
          for (int e = 0; e < 16; e += 4)
            __stcg(part + ((cc + e) >> 2) * kTileN,
                   make_float4(__uint_as_float(v[e]), __uint_as_float(v[e + 1]),
                               __uint_as_float(v[e + 2]), __uint_as_float(v[e + 3])));
        }
      }
      tc_fence_before();
      mbar_arrive(bar_dempty + 8 * d);
      if (!whole) {
        const int o_lo = o * p.KB;
        const int p_first = owner_of(o_lo, p.P, p.T);
        named_bar_sync(1, kEpiThreads);                       // every partial store is issued
        if (et == 0) red_add_release_gpu(p.counters + p_first, 1);   // publish, do not wait
        if (p_first == (int)blockIdx.x) {                     // at most one such tile per CTA
          own_o = o;
          own_slot = slot;
          own_last = owner_of(o_lo + p.KB - 1, p.P, p.T);
        }
      }
      if (et == 0) TR(6, 2 * idx + 1);
    }

    // ---- deferred fix-up: sum the contributors of the owned tile in CTA order (bit-identical
    // run to run).  The others handled this tile FIRST in their walk, so their partials are
    // normally long there; every CTA of the grid is resident (grid <= SM count), so the wait
    // cannot deadlock.
    if (own_o >= 0) {
      if (et == 0) {
        const int need = own_last - (int)blockIdx.x + 1;
        const long long w0 = clock64();
        while (ld_acquire_gpu(p.counters + blockIdx.x) < need) {
          if (clock64() - w0 > B200_WATCHDOG_CYCLES) {
            printf("b200_w4a16: stream-K fix-up timeout block %d\n", blockIdx.x);
            __trap();
          }
        }
        p.counters[blockIdx.x] = 0;                           // leave the workspace zeroed
      }
      named_bar_sync(1, kEpiThreads);
      const int nt = own_o / p.MT, mt = own_o - nt * p.MT;
      const int n_glob = nt * kTileN + c;
      const int m0 = mt * MTOK;
      const float4* mine = reinterpret_cast<const float4*>(p.partials) +
                           ((size_t)blockIdx.x * 2 + own_slot) * kPartVec + c;
      // later contributors start inside the tile: their partial sits in slot 0
      const float4* others = reinterpret_cast<const float4*>(p.partials) +
                             ((size_t)(blockIdx.x + 1) * 2) * kPartVec + c;
      const int nother = own_last - (int)blockIdx.x;
      if (n_glob < p.N) {
#pragma unroll 1
        for (int cc = 0; cc < MTOK; cc += 16) {
          float4 acc[4], f[2][4];
#pragma unroll
          for (int e = 0; e < 4; ++e) acc[e] = __ldcg(mine + ((cc >> 2) + e) * kTileN);
          if (nother > 0) {
#pragma unroll
            for (int e = 0; e < 4; ++e) f[0][e] = __ldcg(others + ((cc >> 2) + e) * kTileN);
          }
          // two contributors in flight: the loads of j + 1 overlap the adds of j
#pragma unroll 1
          for (int j = 0; j < nother; j += 2) {
            const float4* s1 = others + (size_t)(j + 1) * 2 * kPartVec;
            if (j + 1 < nother) {
#pragma unroll
              for (int e = 0; e < 4; ++e) f[1][e] = __ldcg(s1 + ((cc >> 2) + e) * kTileN);
            }
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              acc[e].x += f[0][e].x; acc[e].y += f[0][e].y; acc[e].z += f[0][e].z; acc[e].w += f[0][e].w;
            }
            if (j + 2 < nother) {
#pragma unroll
              for (int e = 0; e < 4; ++e)
                f[0][e] = __ldcg(s1 + (size_t)2 * kPartVec + ((cc >> 2) + e) * kTileN);
            }
            if (j + 1 < nother) {
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                acc[e].x += f[1][e].x; acc[e].y += f[1][e].y; acc[e].z += f[1][e].z; acc[e].w += f[1][e].w;
              }
            }
          }
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float r[4] = {acc[e].x, acc[e].y, acc[e].z, acc[e].w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int m = m0 + cc + 4 * e + u;
              if (m < p.M) p.y[(size_t)m * p.N + n_glob] = __half_as_ushort(__float2half_rn(r[u]));
            }
          }
        }
      }
    }
    if (et == 0) { TRG(4); TR(7, 6); }
  }

  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) TR(7, 1);
  if (warp == kAllocWarp) {
    tc_fence_after();
    tmem_dealloc<kTmemCols>(tmem);
  }
}

// ------------------------------------------------------------------------------------------

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(sym);
  }
  return fn;
}

template <int MTOK>
int launch_impl(const GemmParams& prm, const CUtensorMap& tmap, const GemmPlan& plan, bool pdl,
                cudaStream_t stream) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(w4a16_gemm_kernel<MTOK>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Cfg<MTOK>::kSmemBytes);
    if (e != cudaSuccess) return (int)e;
    attr_set = true;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(plan.ctas);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = Cfg<MTOK>::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  return (int)cudaLaunchKernelEx(&cfg, w4a16_gemm_kernel<MTOK>, prm, tmap);
}

int gemm_flags() {
  static const int f = [] {
    const char* e = getenv("B200_W4A16_FLAGS");
    return e ? atoi(e) : 0;
  }();
  return f;
}

uint32_t* g_trace = nullptr;

template <int MTOK>
void fill_plan(GemmPlan* plan) {
  plan->stages = Cfg<MTOK>::kStages;
  plan->smem_bytes = Cfg<MTOK>::kSmemBytes;
}

}  // namespace

void set_trace_buffer(uint32_t* dev) { g_trace = dev; }
uint32_t* trace_buffer() { return g_trace; }

// B200_W4A16_KERNEL=tcgen05|decode pins the variant (tuning / tests); default: by shape.
static int forced_kind() {
  static const int k = [] {
    const char* e = getenv("B200_W4A16_KERNEL");
    if (!e) return -1;
    if (e[0] == 't') return (int)kKindTcgen05;
    if (e[0] == 'd') return (int)kKindDecode;
    return -1;
  }();
  return k;
}

int make_gemm_plan(int M, int K, int N, int G, int split_k_hint, int sm_count, GemmPlan* plan) {
  (void)split_k_hint;
  if (sm_count <= 0) sm_count = kMaxCtas;
  plan->n_tiles = (N + kTileN - 1) / kTileN;
  plan->k_blocks = K / kTileK;
  const bool decode_ok = M <= kDecodeMaxTokens && G % kTileK == 0;
  const int fk = forced_kind();
  const bool decode = decode_ok && fk != (int)kKindTcgen05;
  if (decode) {
    // Decode variant: stream-K over (channel tile, K block), two CTAs per SM.
    plan->kind = kKindDecode;
    plan->m_tiles = 1;
    const long long T = (long long)plan->n_tiles * plan->k_blocks;
    if (T >= (1LL << 31)) return 1;
    long long ctas = (long long)sm_count * kDecodeCtasPerSm;
    if (T < ctas) ctas = T;
    if (ctas < 1) ctas = 1;
    plan->ctas = (int)ctas;
    const long long per = T / ctas;
    plan->max_split = (plan->k_blocks == 0 || (T % ctas == 0 && per % plan->k_blocks == 0))
                          ? 1
                          : (int)((plan->k_blocks + per - 1) / (per > 0 ? per : 1) + 1);
    fill_decode_plan(M, plan);
    return 0;
  }
  plan->kind = kKindTcgen05;
  const int tok = M <= 16 ? 16 : M <= 32 ? 32 : M <= 64 ? 64 : 128;
  plan->tok_tile = tok;
  plan->m_tiles = (M + tok - 1) / tok;
  const long long outs = (long long)plan->n_tiles * plan->m_tiles;
  const long long T = outs * plan->k_blocks;
  if (T >= (1LL << 31)) return 1;
  // Persistent stream-K: one CTA per SM, equal contiguous ranges of (output tile, K block).
  long long ctas = T < sm_count ? T : sm_count;
  if (ctas < 1) ctas = 1;
  plan->ctas = (int)ctas;
  const long long per = T / ctas;  // shortest range
  plan->max_split = (plan->k_blocks == 0 || (T % ctas == 0 && per % plan->k_blocks == 0))
                        ? 1
                        : (int)((plan->k_blocks + per - 1) / (per > 0 ? per : 1) + 1);
  plan->ws_partial_bytes = (size_t)ctas * 2 * tok * kTileN * sizeof(float);
  plan->ws_counter_bytes = (size_t)ctas * sizeof(int);
  switch (tok) {
    case 16: fill_plan<16>(plan); break;
    case 32: fill_plan<32>(plan); break;
    case 64: fill_plan<64>(plan); break;
    default: fill_plan<128>(plan); break;
  }
  return 0;
}

int launch_gemm(const uint16_t* x, const int32_t* qw_b200, const uint16_t* scales,
                const int32_t* qzeros, uint16_t* y, int M, int K, int N, int G,
                const GemmPlan& plan, void* ws_partials, int* ws_counters, bool pdl,
                cudaStream_t stream) {
  EncodeTiledFn encode = get_encode_fn();
  if (!encode) return (int)cudaErrorNotSupported;
  CUtensorMap tmap;
  const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)M};
  const cuuint64_t strides[1] = {(cuuint64_t)K * 2};
  const cuuint32_t box[2] = {64u, (cuuint32_t)plan.tok_tile};
  const cuuint32_t estr[2] = {1u, 1u};
  CUresult r = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<uint16_t*>(x), dims,
                      strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return (int)cudaErrorInvalidValue;

  GemmParams prm;
  prm.qw = reinterpret_cast<const uint32_t*>(qw_b200);
  prm.scales = scales;
  prm.qz = reinterpret_cast<const uint32_t*>(qzeros);
  prm.y = y;
  prm.partials = reinterpret_cast<float*>(ws_partials);
  prm.counters = ws_counters;
  prm.M = M; prm.K = K; prm.N = N; prm.G = G;
  prm.KB = plan.k_blocks; prm.NT = plan.n_tiles; prm.MT = plan.m_tiles;
  prm.T = plan.n_tiles * plan.m_tiles * plan.k_blocks;
  prm.P = plan.ctas;
  prm.R = G >= kTileK ? 1 : kTileK / G;
  prm.Gdiv = G >= kTileK ? G / kTileK : 1;
  prm.flags = gemm_flags();
  prm.trace = g_trace;
  switch (plan.tok_tile) {
    case 16: return launch_impl<16>(prm, tmap, plan, pdl, stream);
    case 32: return launch_impl<32>(prm, tmap, plan, pdl, stream);
    case 64: return launch_impl<64>(prm, tmap, plan, pdl, stream);
    default: return launch_impl<128>(prm, tmap, plan, pdl, stream);
  }
}

}  // namespace b200
