// w4a16_gemm.cu -- the tcgen05 W4A16 group-quantized GEMM for sm_100a (batches above 16 tokens).
//
//   y[M,N] (fp16) = x[M,K] (fp16) . dequant(qweight, scales, qzeros)[K,N]     fp32 accumulate
//
// Replaces the reference's gemm_forward_4bit_cuda_m16nXk32 + Tensor.sum(0)
// (csrc/quantization/awq/gemm_kernels.cu:30-272, :389-446) and its dequantize + cuBLAS branch
// (vllm/model_executor/layers/quantization/awq.py:154-159).  Design:
//
//  * swap-AB: the tcgen05 "M" dimension (128 TMEM lanes) is OUTPUT CHANNELS, the tcgen05 "N"
//    dimension is TOKENS (MTOK = 16..128).
//  * TWO independent rings in shared memory.  Weights: contiguous 8 KiB tiles (128 channels x
//    128 rows, B200 layout) + their scale / zero rows by 1-D bulk copy, 8-16 stages deep (HBM
//    latency), recycled as soon as the dequant warps have READ them.  Activations: [MTOK x 128]
//    fp16 tiles by 2-D TMA with the 128-byte swizzle tcgen05 expects, 4-8 stages (L2 latency),
//    recycled when the MMAs that read them have completed.  (One shared ring of 5 x 41 KiB
//    stages starved the tensor pipe at 128 tokens: 852 cycles per tile where the MMAs need 512,
//    profiles/r2j_trace_prefill_m2048_5120x15360.txt.)
//  * PAIR = 2 (128-token tiles, long K): a CTA works on TWO adjacent channel tiles per activation
//    tile -- 256 channels x 128 tokens per K block -- so an activation tile is fetched from L2 once
//    per two weight tiles (24 instead of 40 KiB of L2 -> SM traffic per 128 x 128 x 128 MMA block).
//  * TWO (128-token tiles): CTA pairs.  The two CTAs of a cluster (the SMs of one TPC) walk the same
//    (token tile, K block) steps on different channel tiles; the leader's MMA thread issues
//    tcgen05.mma.cta_group::2 (M = 256) for both, each CTA holds HALF of the activation tile in its
//    shared memory (2-SM TMA loads counted on the leader's barrier), dequantizes its own weight tiles
//    into its own TMEM and signals the leader's operand barrier remotely; commits are multicast to
//    both CTAs' barriers.  With short K each CTA of a pair takes ONE channel tile (two accumulator
//    sets), with long K two (one set, drained half by half).
//  * one tcgen05.commit per weight tile at 128 tokens: the commit frees the tile's TMEM operand
//    buffer, and the operand ring -- dequant latency + MMAs up to the commit -- not memory, paced
//    the kernel at 650-670 cycles per tile with one commit per two tiles; 592 now (MMAs need 512).
//  * TP = 2 (CTA pairs, >= 32 token tiles): one dequantized weight tile in TMEM feeds the MMAs of
//    two 128-token tiles -- half the dequant work and weight bytes per flop; steady state = the
//    tensor pipe (513 cycles per 128 x 128 x 128 block).
//  * four dequant warpgroups take weight tiles round-robin.  A thread reads one 16-byte word
//    group (bank-conflict free), expands it in registers with the reference's exact lop3 / sub /
//    fma sequence, and writes the fp16 pairs straight into TENSOR MEMORY with tcgen05.st:
//    thread = TMEM lane = channel, so the registers already are the MMA A-operand -- no
//    shared-memory round trip for the dequantized tile (the reference's :168 store +
//    :198-203 ldmatrix).  A ring of 4-6 operand buffers decouples them from the MMA.
//  * one thread issues tcgen05.mma (A from TMEM, B from shared memory); fp32 accumulators in
//    TMEM, drained by a dedicated epilogue warpgroup while the next output tile's MMAs run.
//  * stream-K: the (channel-tile group, token tile, K block) space is cut into equal contiguous
//    ranges, one per SM.  A CTA does its two partial output tiles FIRST and publishes them as
//    self-flagging fp32 words; after its main loop the CTA that owns a split tile's first K
//    block sums the contributors in CTA order with ALL its dequant + epilogue threads, every
//    poll of a thread in flight at once (no second kernel, deterministic).
//  * programmatic dependent launch: the weight ring is filled BEFORE griddepcontrol.wait
//    (weights never depend on the previous kernel), so launch ramp overlaps the previous op.
//
// A decode-mode variant of this kernel (unscaled (q - z) operands, group scale applied to
// per-tile accumulators by the epilogue) was built and measured in round 2 and REMOVED: at
// <= 64 tokens a tcgen05.mma of K = 16 costs the tensor pipe 41-56 cycles whatever N is, eight
// of them per 8 KiB tile are level with the HBM tile period, and the per-tile accumulator drain
// made it slower than this kernel at 24..64 tokens and no faster than the mma.sync decode
// kernel at <= 16 (profiles/r2j_decode_variants.md).
#include "w4a16_common.cuh"
#include "w4a16_internal.h"
#include "w4a16_streamk.cuh"

#include <stdlib.h>

namespace b200 {

namespace {

// Warp roles.  The issue arbiter of a sub-partition favours the highest warp id
// (B300_MICROARCH.md, multi-warp arbiter), so the single-thread roles whose latency paces the
// whole pipeline -- above all the MMA issuer -- take the TOP warp ids, one per sub-partition;
// the epilogue sits above the dequant warps.
constexpr int kGroups = 4;                              // dequant warpgroups, weight tiles round-robin
constexpr int kDequantThreads = kGroups * 128;          // warps 0..15 (warp % 4 == TMEM lane quarter)
constexpr int kFirstEpiWarp = 16;                       // warps 16..19 (warp % 4 == lane quarter)
constexpr int kEpiThreads = 128;
constexpr int kFixThreads = kDequantThreads + kEpiThreads;   // everyone who takes part in the fix-up
constexpr int kAllocWarp = 20;                          // TMEM allocation + second weight producer
constexpr int kXProducerWarp = 21;
constexpr int kProducerWarp = 22;
constexpr int kMmaWarp = 23;
constexpr int kThreads = 24 * 32;                       // 768
constexpr int kScaleRowsMax = 4;                        // G = 32 -> four (scale, zero) rows / tile
constexpr int kTmemACols = 64;                          // one dequantized tile: 128 fp16 per lane
constexpr int kTmemCols = 512;
constexpr int kMaxSmem = 227 * 1024;

template <int V>
struct IntC { static constexpr int value = V; };

// CTA pairs (cta_group::2) for 128-token tiles: measured +2-10 % on 10 of 12 Llama-2-13B prefill points
// (profiles/r2x_kb_ab_two{0,1}.jsonl); B200_W4A16_TWO=0 in a tuning build switches them off
#ifndef B200_GEMM_TWO_CTA_DEFAULT
#define B200_GEMM_TWO_CTA_DEFAULT 1
#endif
// ring depths below 128 tokens (tuning builds may override: make variant)
#ifndef B200_XST64
#define B200_XST64 6
#endif
#ifndef B200_WST64
#define B200_WST64 12
#endif
#ifndef B200_XST32
#define B200_XST32 8
#endif
#ifndef B200_WST32
#define B200_WST32 16
#endif

// TWO: a pair of CTAs (cluster of 2 on one TPC) issues tcgen05.mma.cta_group::2 -- each CTA holds its
// own channel tiles (A, D) and HALF of the activation tile (MTOK / 2 tokens) in shared memory.
// TP: token tiles per work unit (1 or 2): with 2, ONE dequantized weight tile in TMEM feeds the MMAs
// of two 128-token tiles -- half the dequant work and half the weight bytes per flop.
template <int MTOK, int PAIR, bool TWO = false, int TP = 1>
struct Cfg {
  // accumulators: PAIR per buffer set; two sets (drain overlaps the next tile's MMAs) unless
  // 2 x 128 columns per set leave no room (MTOK = 128, PAIR = 2: one set, drained half by half)
  static constexpr int kAccSlots = PAIR * TP;          // accumulators per set: (channel tile, token tile)
  static constexpr int kAccBufs = (MTOK * kAccSlots > 128) ? 1 : 2;
  static constexpr int kAccCols = kAccBufs * kAccSlots * MTOK;
  // dequantized-operand ring in TMEM: what the accumulators leave, at most 6, even
  static constexpr int kABufsFit = (kTmemCols - kAccCols) / kTmemACols;
  static constexpr int kABufs = kABufsFit >= 6 ? 6 : 4;
  static constexpr int kDCol = kABufs * kTmemACols;    // accumulators start here
  // activation ring
  static constexpr int kXRows = TWO ? MTOK / 2 : MTOK; // token rows of the activation tile in THIS CTA
  static constexpr int kXTileBytes = kXRows * 256;     // two swizzle atoms of [kXRows x 64] fp16
  static constexpr int kXBytes = kXTileBytes * TP;      // one stage: the unit's TP token tiles
  static constexpr int kXStages = TWO ? (TP == 2 ? 4 : 6) : MTOK >= 128 ? 4 : MTOK == 64 ? B200_XST64 : B200_XST32;
  // weight ring: [8 KiB tile | 4 x 256 B scale rows | 4 x 64 B zero rows]
  static constexpr int kSOff = kTileWords * 4;
  static constexpr int kZOff = kSOff + kScaleRowsMax * 256;
  static constexpr int kWStageBytes = kZOff + kScaleRowsMax * 64;            // 9472
  static constexpr int kWStages = TWO ? (TP == 2 ? 8 : 12) : MTOK >= 128 ? 8 : MTOK == 64 ? B200_WST64 : B200_WST32;
  static constexpr int kBarBytes = 1024;
  static constexpr int kXOff = kBarBytes;                                    // 1024-aligned
  static constexpr int kWOff = kXOff + kXStages * kXBytes;
  static constexpr int kSmemBytes = kWOff + kWStages * kWStageBytes + 1024;  // + align slack
  // Weight tiles per tcgen05.commit.  The commit frees the tile's operand buffer: with 128-token tiles
  // (a tile's MMAs take 512 cycles) the operand ring, not memory, paced the kernel -- a buffer's cycle
  // is dequant (~900) + wait + the MMAs up to the commit + wake-up, and with one commit per TWO
  // tiles four buffers gave 650-670 cycles per tile in every configuration measured, CTA pairs
  // included (profiles/r2v_trace_prefill_two.txt).  Below 128 tokens two tiles share a commit.
#ifndef B200_COMMIT128
#define B200_COMMIT128 1
#endif
  static constexpr int kCommit = MTOK >= 128 ? B200_COMMIT128 : 2;
  // MMA completion is signalled once per kCommit weight tiles (a tcgen05.commit costs the tensor
  // pipe ~110 cycles, scripts/probes/mma_probe.cu) on a ring of kDoneBars barriers; the
  // activation producer (stage reuse) and the dequant groups (operand-buffer reuse) wait on it.
  static constexpr int kRingTiles = kXStages * PAIR > kABufs ? kXStages * PAIR : kABufs;
  static constexpr int kDoneBars = kRingTiles / kCommit + 2;
  static constexpr int kNumBars = 2 * kWStages + kXStages + kDoneBars + kABufs + kAccBufs + kAccBufs * kAccSlots;
  // A dequant group waits for the MMAs of tile w - kABufs before it looks at tile w's stage:
  // with kWStages >= kABufs the stage's previous tenant (tile w - kWStages) has then been
  // dequantized, i.e. had landed -- the parity wait cannot pass on a stale phase.
  static_assert(kWStages >= kABufs, "stage reuse must be covered by the operand-ring wait");
  static_assert(kWStages % 2 == 0 && kABufs % kCommit == 0, "producer / commit parity");
  static_assert(kWStageBytes % 128 == 0 && kXBytes % 1024 == 0, "ring alignment");
  static_assert(kDCol + kAccCols <= kTmemCols, "TMEM budget");
  static_assert(kSmemBytes <= kMaxSmem, "shared memory budget");
  static_assert(8 * kNumBars + 8 <= kBarBytes, "barrier block");
};

struct GemmParams {
  const uint32_t* qw;
  const uint16_t* scales;
  const uint32_t* qz;
  uint16_t* y;
  uint4* partials;   // [CTA][2 slots][PAIR][MTOK x 128 / 4] self-flagging fp32 quads (kPartEmpty = unwritten)
  int M, K, N, G;
  int KB, NT, MT;
  int NTP;      // channel-tile groups: ceil(NT / PAIR)
  int T;        // total (output unit, K block) work items
  int P;        // CTAs
  int R;        // (scale, zero) rows per K block: max(1, 128 / G)
  int Gdiv;     // K blocks per group: max(1, G / 128)
  int flags;    // bit 0: no L2 evict-first hint on the weight stream
  uint32_t* trace;  // B200_TRACE builds only: per-CTA event timestamps
  // DYN instantiations (b200_w4a16_gemm_rows): the number of valid rows of x / y is read from
  // the device (<= M, the buffers' rows); token tiles, work items and ranges follow it.
  const int* m_dev;
  int split_cap, align_pct;   // DYN: the planner's grid policy (aligned_workers), applied to the live problem
};

#ifdef B200_TRACE
// event types: 0 producer issues weight tile, 1 dequant sees tile landed, 2 dequant starts (operand
// buffer free), 3 dequant done, 4 MMA starts weight tile, 5 MMA issued, 6 epilogue seg (2*idx: dfull
// seen, 2*idx+1: segment finished), 7 misc (0 setup done, 1 exit, 2..4 globaltimer: entry /
// setup / epilogue done, 5..6 clock at entry / epilogue done)
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

// The output unit this CTA finishes (it holds the unit's first K block but not all of them), if
// any: only the first walk segment or the tail segment can be partial, at most one of them starts
// at a unit boundary.  slot = which of the CTA's two partial slots its own piece sits in.
struct Owned {
  int o, slot, last;   // o < 0: none; last = highest CTA contributing to the unit
};
__device__ __forceinline__ Owned find_owned(const Walk& walk, int KB, int P, int T, int cta) {
  (void)cta;
  Owned r{-1, 0, 0};
  const int nlook = walk.nseg < 2 ? walk.nseg : 2;
  for (int idx = 0; idx < nlook; ++idx) {
    int t0, n;
    walk.seg(idx, t0, n);
    if (n != KB && t0 % KB == 0) {                       // t0 lies in this CTA's own range: it owns the unit
      r.o = t0 / KB;
      r.slot = idx == 0 ? 0 : 1;
      r.last = owner_of(t0 + KB - 1, P, T);
    }
  }
  return r;
}

// DYN: the token count comes from device memory (the routed rows of a mixture-of-experts block:
// how many tokens chose an expert is only known on the device).  The grid is the one planned
// for the buffers' M rows; every CTA derives the token tiles, the work-item count and its range
// from the live count, so an expert with few tokens streams its weights once per LIVE token
// tile.  Rows beyond the count are never stored (whatever the gather buffer holds there only
// reaches accumulator columns of its own).
template <int MTOK, int PAIR, bool TWO, int TP, bool DYN>
__global__ void __launch_bounds__(kThreads, 1)
w4a16_gemm_kernel(const GemmParams p, const __grid_constant__ CUtensorMap tmap_x) {
  using C = Cfg<MTOK, PAIR, TWO, TP>;
  constexpr int kAS = C::kAccSlots;
  // CTA pair: rank in the cluster (0 = leader: issues the MMAs, owns the barriers their issuer waits
  // on); the work space is cut per PAIR of CTAs, both walk the same (token tile, K block) steps
  const uint32_t crank = TWO ? cluster_ctarank() : 0u;
  const int vblock = TWO ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  int VP = TWO ? p.P >> 1 : p.P;
  if (threadIdx.x == 0) { TRG(2); TR(7, 5); }
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));

  // barrier block
  const uint32_t bar_wfull = smem_base;                          // [kWStages]  weight tile landed
  const uint32_t bar_wempty = bar_wfull + 8 * C::kWStages;       // [kWStages]  weight tile read by its dequant group
  const uint32_t bar_xfull = bar_wempty + 8 * C::kWStages;       // [kXStages]  activation tile landed
  const uint32_t bar_done = bar_xfull + 8 * C::kXStages;         // [kDoneBars] MMAs of kCommit weight tiles done
  const uint32_t bar_afull = bar_done + 8 * C::kDoneBars;        // [kABufs]    operand dequantized
  const uint32_t bar_dfull = bar_afull + 8 * C::kABufs;          // [kAccBufs]  accumulator set complete
  const uint32_t bar_dempty = bar_dfull + 8 * C::kAccBufs;       // [kAccBufs * kAccSlots] accumulator drained
  const uint32_t tmem_slot = smem_base + 8 * C::kNumBars;
  volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(
      smem_gen + (tmem_slot - smem_base));
  const uint32_t xring = smem_base + C::kXOff;
  const uint32_t wring = smem_base + C::kWOff;
  const uint8_t* wring_gen = smem_gen + C::kWOff;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  int M = p.M, MT = p.MT, T = p.T;
  if constexpr (DYN) {
    static_assert(!TWO && TP == 1, "dynamic row counts: single-CTA variants only");
    asm volatile("griddepcontrol.wait;" ::: "memory");        // the count is the previous kernel's output
    const int m = *reinterpret_cast<const volatile int*>(p.m_dev);
    M = m < p.M ? m : p.M;
    if (M <= 0) return;                                       // nobody chose this expert
    MT = (M + MTOK - 1) / MTOK;
    T = p.NTP * MT * p.KB;
    // The planner's grid policy on the live problem (aligned_workers): a whole number of CTAs per
    // output unit when the live units are fewer than the CTAs.
    const int units = p.NTP * MT;
    if (units <= VP && p.align_pct > 0) {
      int S = VP / units;
      bool capped = false;
      if (p.split_cap > 0 && S > p.split_cap) { S = p.split_cap; capped = true; }
      if (S > p.KB) S = p.KB;
      if (capped || units * S * 100 >= VP * p.align_pct) VP = units * S;
    }
    // Fewer work items than CTAs: the first T CTAs take one each.  (Equal ranges over all P would
    // leave idle CTAs BETWEEN the contributors of a unit; the fix-up takes them to be consecutive.)
    if (T < VP) VP = T;
    if (vblock >= VP) return;
  }

  // This CTA's contiguous slice of the work space, and the order it is walked in.  A work item
  // ("unit step") is one K block of one output unit = PAIR channel tiles x one token tile.
  const int t_begin = range_begin(vblock, T, VP);
  const int t_end = range_begin(vblock + 1, T, VP);
  const Walk walk(t_begin, t_end, p.KB);
  const int n_units = t_end - t_begin;
  const int n_w = n_units * PAIR;                               // weight tiles in stream order
  const int npre = n_w < C::kWStages ? n_w : C::kWStages;
  if (threadIdx.x == 0) TR(7, 8);                               // ranges and walk derived

  // Weight stream of this CTA: for every unit step of the walk its PAIR weight tiles, every
  // address advanced incrementally -- consecutive K blocks of a channel tile are contiguous in
  // the B200 layout, so a step is a few adds, an mbarrier wait and three bulk copies.  It has
  // to be that lean: a producer is ONE thread running a dependent chain per tile.  Two producer
  // threads on different sub-partitions (warps kProducerWarp / kAllocWarp) take the tiles
  // alternately.  A channel tile beyond the matrix (odd NT, PAIR = 2) re-reads the last one:
  // its MMAs run, the epilogue drops the result.
  struct WStream {
    const GemmParams& p;
    const Walk& walk;
    int crank_, MT_;
    int idx, left, h, gcnt;
    uint32_t sb0, zb0, sb1, zb1;                        // scale / zero row bytes of tile h = 0 / 1
    const uint32_t *w0, *w1, *z0, *z1;                  // weight tile, zero row
    const uint16_t *s0, *s1;                            // scale row
    __device__ __forceinline__ WStream(const GemmParams& pp, const Walk& w, int cr, int mt)
        : p(pp), walk(w), crank_(cr), MT_(mt), idx(-1), left(0), h(0), gcnt(0), sb0(0), zb0(0), sb1(0), zb1(0),
          w0(nullptr), w1(nullptr), z0(nullptr), z1(nullptr), s0(nullptr), s1(nullptr) {}
    __device__ __forceinline__ void open() {
      h = 0;
      if (++idx >= walk.nseg) return;
      int t0, n;
      walk.seg(idx, t0, n);
      const int o = t0 / p.KB;
      const int kb = t0 - o * p.KB;
      const int ntp = o / MT_;
      left = n;
      const int g0 = p.R > 1 ? kb * p.R : kb / p.Gdiv;
      gcnt = p.R > 1 ? 0 : kb - g0 * p.Gdiv;                  // K blocks into the current group
      // channel tiles of this CTA in the group (a tile beyond the matrix re-reads the last one)
      int nta = (TWO ? ntp * 2 + crank_ : ntp) * PAIR;
      if (nta >= p.NT) nta = p.NT - 1;
      const int ntb = (PAIR == 2 && nta + 1 < p.NT) ? nta + 1 : nta;
      const uint32_t va = (uint32_t)min(kTileN, p.N - nta * kTileN);
      const uint32_t vb = (uint32_t)min(kTileN, p.N - ntb * kTileN);
      sb0 = va * 2; zb0 = va / 2; sb1 = vb * 2; zb1 = vb / 2;
      w0 = p.qw + ((size_t)nta * p.KB + kb) * kTileWords;
      w1 = p.qw + ((size_t)ntb * p.KB + kb) * kTileWords;
      s0 = p.scales + nta * kTileN + (size_t)g0 * p.N;
      s1 = p.scales + ntb * kTileN + (size_t)g0 * p.N;
      z0 = p.qz + nta * 16 + (size_t)g0 * (p.N >> 3);
      z1 = p.qz + ntb * 16 + (size_t)g0 * (p.N >> 3);
    }
    __device__ __forceinline__ void next() {                  // to the next weight tile of the stream
      if (PAIR == 2 && h == 0) { h = 1; return; }
      h = 0;
      w0 += kTileWords; w1 += kTileWords;
      if (p.R > 1) {
        const size_t ds = (size_t)p.R * p.N, dz = (size_t)p.R * (p.N >> 3);
        s0 += ds; s1 += ds; z0 += dz; z1 += dz;
      } else if (++gcnt == p.Gdiv) {
        gcnt = 0;
        s0 += p.N; s1 += p.N; z0 += p.N >> 3; z1 += p.N >> 3;
      }
      if (--left == 0) open();
    }
  };
  const uint64_t pol_w = (p.flags & 1) ? policy_evict_normal() : policy_evict_first();
  auto issue_w = [&](const WStream& c, int stage) {
    const uint32_t sb = wring + stage * C::kWStageBytes;
    const uint32_t bar = bar_wfull + 8 * stage;
    const bool hb = PAIR == 2 && c.h == 1;
    const uint32_t sbytes = hb ? c.sb1 : c.sb0, zbytes = hb ? c.zb1 : c.zb0;
    const uint32_t* wsrc = hb ? c.w1 : c.w0;
    const uint16_t* sptr = hb ? c.s1 : c.s0;
    const uint32_t* zptr = hb ? c.z1 : c.z0;
    mbar_arrive_expect_tx(bar, kTileWords * 4 + (uint32_t)p.R * (sbytes + zbytes));
    bulk_g2s_hint(sb, wsrc, kTileWords * 4, bar, pol_w);
    bulk_g2s(sb + C::kSOff, sptr, sbytes, bar);
    bulk_g2s(sb + C::kZOff, zptr, zbytes, bar);
    for (int r = 1; r < p.R; ++r) {                           // G < 128: several rows per K block
      bulk_g2s(sb + C::kSOff + r * 256, sptr + (size_t)r * p.N, sbytes, bar);
      bulk_g2s(sb + C::kZOff + r * 64, zptr + (size_t)r * (p.N >> 3), zbytes, bar);
    }
  };

  WStream cw(p, walk, (int)crank, MT);                          // opened by a producer thread only
  if (warp == kProducerWarp) {
    // Lane-parallel barrier init, then the first npre weight tiles go out BEFORE the CTA-wide
    // setup barrier and without ever waiting for the previous kernel: weights never depend on
    // it, so launch ramp and TMEM allocation overlap the first fetch.
    for (int b = lane; b < C::kNumBars; b += 32) {
      uint32_t count = 1u;                                   // wfull / xfull / done / dfull: one arriver
      if (b >= C::kWStages && b < 2 * C::kWStages) count = 4u;                    // wempty: one lane per warp
      const int a0 = 2 * C::kWStages + C::kXStages + C::kDoneBars;
      if (b >= a0 && b < a0 + C::kABufs) count = TWO ? 256u : 128u;               // afull (pair: both CTAs' dequant groups)
      if (b >= a0 + C::kABufs + C::kAccBufs) count = (uint32_t)kEpiThreads * (TWO ? 2u : 1u);   // dempty
      mbar_init(smem_base + 8 * b, count);
    }
    fence_barrier_init();
    __syncwarp();
    if (elect_one()) {
      TR(7, 9);                                                 // barriers initialised
      // Let the next kernel in the stream start its own weight prefetch as soon as an SM frees.
      asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
      cw.open();
      TR(7, 10);                                                // weight cursor opened
      for (int i = 0; i < npre; ++i) { TR(0, i); issue_w(cw, i); cw.next(); }
      TR(7, 11);                                                // first ring issued
    }
    __syncwarp();
  } else if (warp == kAllocWarp) {
    if (TWO) tmem_alloc_pair<kTmemCols>(tmem_slot); else tmem_alloc<kTmemCols>(tmem_slot);
    if (lane == 0) TR(7, 12);                                   // TMEM allocated
  }
  tc_fence_before();
  __syncthreads();
  if (TWO) cluster_sync_all();          // both CTAs' barriers exist before anything signals across
  tc_fence_after();
  const uint32_t tmem = *tmem_slot_gen;
  if (threadIdx.x == 0) { TR(7, 0); TRG(3); }

  if (warp == kProducerWarp || warp == kAllocWarp) {
    // -- steady state: refill weight stage w % kWStages once its dequant group has read tile
    // w - kWStages.  Producer r takes the tiles with w % 2 == r (kWStages is even: always the
    // same stages); both walk the same cursor.  elect.sync picks the same lane as before the
    // setup barrier, so warp 22's `cw` carries over.
    __syncwarp();
    if (elect_one() && npre < n_w) {
      const int r = warp == kProducerWarp ? 0 : 1;
      if (r == 1) {
        cw.open();
        for (int i = 0; i < npre; ++i) cw.next();
      }
      int w = npre;
      if ((w & 1) != r) { cw.next(); ++w; }
      int stage = w % C::kWStages;
      uint32_t par = (uint32_t)(w / C::kWStages - 1) & 1u;     // phase of wempty[stage] to wait for
      for (; w < n_w; w += 2) {
        mbar_wait(bar_wempty + 8 * stage, par);
        TR(0, w);
        issue_w(cw, stage);
        cw.next();
        if (w + 1 < n_w) cw.next();
        stage += 2;
        if (stage >= C::kWStages) { stage -= C::kWStages; par ^= 1u; }
      }
    }
  } else if (warp == kXProducerWarp) {
    // Activation stream (elected thread of warp 21): two 64-column TMA boxes per unit step,
    // issued once the previous kernel's output is visible; stage u % kXStages is refilled when
    // the MMAs of step u - kXStages have completed, i.e. when the commit group of that step's
    // LAST weight tile has: group ((u - kXStages) * PAIR + PAIR - 1) / kCommit.
    __syncwarp();
    if (elect_one()) {
      tma_prefetch_desc(&tmap_x);
      asm volatile("griddepcontrol.wait;" ::: "memory");      // x is the previous kernel's output
      TR(7, 13);                                                // dependency resolved (activation producer)
      int u = 0;
      uint32_t sb = xring, bar = bar_xfull;
      for (int idx = 0; idx < walk.nseg; ++idx) {
        int t0, n;
        walk.seg(idx, t0, n);
        const int o = t0 / p.KB;
        int k0 = (t0 - o * p.KB) * kTileK;
        const int m0 = (o % MT) * (TP * MTOK) + (TWO ? (int)crank * C::kXRows : 0);   // pair: this CTA's token half
        for (int i = 0; i < n; ++i, ++u, k0 += kTileK) {
          if (u >= C::kXStages) {
            const int g = ((u - C::kXStages) * PAIR + PAIR - 1) / C::kCommit;
            mbar_wait(bar_done + 8 * (g % C::kDoneBars), (uint32_t)(g / C::kDoneBars) & 1u);
          }
          if (TWO) {                                          // both halves are counted on the leader's barrier
            if (crank == 0) mbar_arrive_expect_tx(bar, 2u * (uint32_t)C::kXBytes);
#pragma unroll
            for (int t = 0; t < TP; ++t) {
              tma_load_2d_pair(sb + t * C::kXTileBytes, &tmap_x, k0, m0 + t * MTOK, bar);
              tma_load_2d_pair(sb + t * C::kXTileBytes + C::kXRows * 128, &tmap_x, k0 + 64, m0 + t * MTOK, bar);
            }
          } else {
            mbar_arrive_expect_tx(bar, (uint32_t)C::kXBytes);
#pragma unroll
            for (int t = 0; t < TP; ++t) {
              tma_load_2d(sb + t * C::kXTileBytes, &tmap_x, k0, m0 + t * MTOK, bar);
              tma_load_2d(sb + t * C::kXTileBytes + C::kXRows * 128, &tmap_x, k0 + 64, m0 + t * MTOK, bar);
            }
          }
          sb += C::kXBytes;
          bar += 8;
          if (bar == bar_xfull + 8 * C::kXStages) { sb = xring; bar = bar_xfull; }
        }
      }
    }
  } else if (warp == kMmaWarp) {
    __syncwarp();
    if (elect_one() && crank == 0) {                          // pair: the leader issues for both CTAs
      // ONE thread issues every tcgen05.mma of the CTA, so its instruction count per tile is a
      // floor for the tile period: descriptors are a base plus compile-time constants, ring
      // positions advance incrementally (no modulo), and the operand barrier of the NEXT weight
      // tile is tested before this tile's MMAs so its latency hides behind them.
      constexpr uint32_t idesc = make_idesc_f16(TWO ? 256 : 128, MTOK);
      const uint64_t bdesc0 = make_desc_k_sw128(xring);           // activation stage 0, K slice 0
      constexpr uint32_t kXStep = C::kXBytes >> 4;                // descriptor address units (16 B)
      uint32_t x_off = 0;                                         // stage * kXStep
      uint32_t xbar = bar_xfull, xphase = 0;
      int buf = 0, w = 0;                                         // operand buffer, weight tile index
      uint32_t aphase = 0;
      int done_slot = 0;                                          // (w / kCommit) % kDoneBars
      bool a_ready = n_w > 0 && mbar_try_wait(bar_afull, 0);
      for (int idx = 0; idx < walk.nseg; ++idx) {
        int t0, n;
        walk.seg(idx, t0, n);
        const int d = idx % C::kAccBufs;                          // accumulator set of this output unit
        const uint32_t dpar = ((uint32_t)(idx / C::kAccBufs) & 1u) ^ 1u;
        for (int i = 0; i < n; ++i) {
          mbar_wait(xbar, xphase);                                // activation tile of this step
          const uint64_t bd = bdesc0 + x_off;
#pragma unroll
          for (int h = 0; h < PAIR; ++h) {
            if (i == 0) {                                         // first touch of accumulators (d, h, *):
#pragma unroll
              for (int t = 0; t < TP; ++t) mbar_wait(bar_dempty + 8 * ((d * PAIR + h) * TP + t), dpar);   // drained by the epilogue warps
              tc_fence_after();
            }
            if (!a_ready) mbar_wait(bar_afull + 8 * buf, aphase);
            tc_fence_after();
            TR(4, w);
            const uint32_t acol = tmem + buf * kTmemACols;
            const uint32_t dcol0 = tmem + C::kDCol + (d * PAIR + h) * TP * MTOK;
            // ring positions of the next weight tile, and an early look at its operand barrier
            int nbuf = buf + 1;
            uint32_t nphase = aphase;
            if (nbuf == C::kABufs) { nbuf = 0; nphase ^= 1u; }
            a_ready = w + 1 < n_w && mbar_try_wait(bar_afull + 8 * nbuf, nphase);
#pragma unroll
            for (int t = 0; t < TP; ++t)                          // the operand buffer feeds every token tile of the unit
#pragma unroll
              for (int kk = 0; kk < 8; ++kk) {
                const uint64_t bk = bd + (uint32_t)(t * (C::kXTileBytes / 16) + (kk >> 2) * (C::kXRows * 128 / 16) + (kk & 3) * 2);
                const uint32_t dcol = dcol0 + t * MTOK;
                if (TWO) tc_mma_ts_pair(dcol, acol + kk * 8, bk, idesc, (uint32_t)((i | kk) != 0));
                else tc_mma_ts(dcol, acol + kk * 8, bk, idesc, (uint32_t)((i | kk) != 0));
              }
            ++w;
            if ((w & (C::kCommit - 1)) == 0 || w == n_w) {
              if (TWO) tc_commit_pair(bar_done + 8 * done_slot); else tc_commit(bar_done + 8 * done_slot);
              if (++done_slot == C::kDoneBars) done_slot = 0;
            }
            TR(5, w - 1);
            buf = nbuf;
            aphase = nphase;
          }
          x_off += kXStep;
          xbar += 8;
          if (xbar == bar_xfull + 8 * C::kXStages) { x_off = 0; xbar = bar_xfull; xphase ^= 1u; }
        }
        if (TWO) tc_commit_pair(bar_dfull + 8 * d); else tc_commit(bar_dfull + 8 * d);
      }
    }
  } else if (warp < kFirstEpiWarp) {
    // ---- dequant warpgroups: weight tile w -> group w % 4, stage w % kWStages, operand w % kABufs
    const int grp = warp >> 2, q = warp & 3;
    const int c = q * 32 + lane;                              // channel inside the tile
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    const int zsh = 4 * awq_nibble(c & 7);
    int stage = grp % C::kWStages, buf = grp % C::kABufs;
    uint32_t fphase = 0;
    for (int w = grp; w < n_w; w += kGroups) {
      const uint8_t* sb = wring_gen + stage * C::kWStageBytes;
      // Operand buffer w % kABufs was last read by the MMAs of weight tile w - kABufs.
      if (w >= C::kABufs) {
        const int g = (w - C::kABufs) / C::kCommit;
        mbar_wait(bar_done + 8 * (g % C::kDoneBars), (uint32_t)(g / C::kDoneBars) & 1u);
      }
      // (that wait also covers the stage's previous phase, see Cfg: no stale pass here)
      mbar_wait(bar_wfull + 8 * stage, fphase);
      if ((threadIdx.x & 127) == 0) TR(1, w);
      uint4 wv[4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        wv[j] = *reinterpret_cast<const uint4*>(sb + j * 2048 + c * 16);
      const uint16_t* sS = reinterpret_cast<const uint16_t*>(sb + C::kSOff);
      const uint32_t* sZ = reinterpret_cast<const uint32_t*>(sb + C::kZOff);
      uint32_t zlo[kScaleRowsMax], zhi[kScaleRowsMax], s2[kScaleRowsMax];
      make_zs1((sZ[c >> 3] >> zsh) & 0xFu, sS[c], zlo[0], zhi[0], s2[0]);
      if (p.R > 1) {
#pragma unroll
        for (int r = 1; r < kScaleRowsMax; ++r)
          if (r < p.R) make_zs1((sZ[r * 16 + (c >> 3)] >> zsh) & 0xFu, sS[r * 128 + c], zlo[r], zhi[r], s2[r]);
      }
      // everything this thread needs from the stage is in registers: hand it back to the producer
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_wempty + 8 * stage);
      tc_fence_after();
      if ((threadIdx.x & 127) == 0) TR(2, w);
      const uint32_t abase = tmem + lane_base + buf * kTmemACols;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        // rows 32 j .. 32 j + 31 of the tile: (scale, zero) row (32 j) / G when G < 128
        uint32_t zl = zlo[0], zh = zhi[0], sc = s2[0];
        if (p.R > 1) {
          const int r = (j * 32) / p.G;
#pragma unroll
          for (int rr = 1; rr < kScaleRowsMax; ++rr)
            if (r == rr) { zl = zlo[rr]; zh = zhi[rr]; sc = s2[rr]; }
        }
        uint32_t v[16];
        dequant_word_zs1(wv[j].x, zl, zh, sc, v[0], v[1], v[2], v[3]);
        dequant_word_zs1(wv[j].y, zl, zh, sc, v[4], v[5], v[6], v[7]);
        dequant_word_zs1(wv[j].z, zl, zh, sc, v[8], v[9], v[10], v[11]);
        dequant_word_zs1(wv[j].w, zl, zh, sc, v[12], v[13], v[14], v[15]);
        tmem_st_x16(abase + j * 16, v);
      }
      tc_wait_st();
      tc_fence_before();
      if (TWO && crank != 0) mbar_arrive_cluster(map_to_cta(bar_afull + 8 * buf, 0));   // the leader's barrier
      else mbar_arrive(bar_afull + 8 * buf);
      if ((threadIdx.x & 127) == 0) TR(3, w);
      stage += kGroups;
      if (stage >= C::kWStages) { stage -= C::kWStages; fphase ^= 1u; }
      buf += kGroups;
      if (buf >= C::kABufs) buf -= C::kABufs;
    }
  } else if (warp < kFirstEpiWarp + kEpiThreads / 32) {
    // ---- epilogue warpgroup: drains accumulators; a whole unit goes to y, a partial one is
    // published as self-flagging fp32 quads and finished after the main loop (fix-up below):
    // nothing here waits on another CTA mid-stream
    const int q = warp - kFirstEpiWarp;
    const int c = q * 32 + lane;                              // channel inside the tile
    const int et = threadIdx.x - kFirstEpiWarp * 32;          // 0..127
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    constexpr int kPartVec = MTOK * kTileN / 4;               // quads per partial tile
    asm volatile("griddepcontrol.wait;" ::: "memory");        // y / workspace may alias earlier work
    for (int idx = 0; idx < walk.nseg; ++idx) {
      int t0, n;
      walk.seg(idx, t0, n);
      const int o = t0 / p.KB;
      const int ntp = o / MT, mt = o - ntp * MT;
      const int d = idx % C::kAccBufs;
      const bool whole = (n == p.KB);
      const int slot = idx == 0 ? 0 : 1;
      mbar_wait(bar_dfull + 8 * d, (uint32_t)(idx / C::kAccBufs) & 1u);
      tc_fence_after();
      if (et == 0) TR(6, 2 * idx);
#pragma unroll 1
      for (int a = 0; a < kAS; ++a) {                         // accumulator a = (channel tile h, token tile t)
        const int h = a / TP, t = a % TP;
        const int m0 = (mt * TP + t) * MTOK;
        const int n_glob = ((TWO ? ntp * 2 + (int)crank : ntp) * PAIR + h) * kTileN + c;
        // partial layout: quad v of channel c holds tokens 4 v .. 4 v + 3 -> [v][c], coalesced
        uint4* part = p.partials + (((size_t)blockIdx.x * 2 + slot) * kAS + a) * kPartVec + c;
        const uint32_t dbase = tmem + lane_base + C::kDCol + (d * kAS + a) * MTOK;
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
                if (m < M)
                  p.y[(size_t)m * p.N + n_glob] =
                      __half_as_ushort(__float2half_rn(__uint_as_float(v[e])));
              }
            }
          } else {
#pragma unroll
            for (int e = 0; e < 16; e += 4)
              st_volatile_u4(part + ((cc + e) >> 2) * kTileN,
                             part_pack(__uint_as_float(v[e]), __uint_as_float(v[e + 1]),
                                       __uint_as_float(v[e + 2]), __uint_as_float(v[e + 3])));
          }
        }
        tc_fence_before();
        if (TWO && crank != 0) mbar_arrive_cluster(map_to_cta(bar_dempty + 8 * (d * kAS + a), 0));
        else mbar_arrive(bar_dempty + 8 * (d * kAS + a));
      }
      if (et == 0) TR(6, 2 * idx + 1);
    }
    if (et == 0) { TRG(4); TR(7, 6); }
  }

  // ---- deferred stream-K fix-up, by every dequant and epilogue thread of the CTA that owns a
  // split unit's first K block: sum the contributors' partial tiles in CTA order (bit-identical
  // run to run) and write y.  The quads are polled where they land (self-flagging words, one L2
  // hop) and handed back empty; ALL polls of a thread's pass are issued before the first is
  // looked at (one round trip per pass instead of one per contributor).  The others handled
  // this unit FIRST in their walk, so their partials are normally long there; every CTA of the
  // grid is resident (grid <= SM count), so the wait cannot deadlock.
  if (warp < kFirstEpiWarp + kEpiThreads / 32) {
    const Owned own = find_owned(walk, p.KB, VP, T, vblock);
    if (own.o >= 0) {
      asm volatile("griddepcontrol.wait;" ::: "memory");      // the dequant warps write y here
      constexpr int kPartVec = MTOK * kTileN / 4;
      constexpr int kQuads = kAS * kPartVec;
      const int ntp = own.o / MT, mt = own.o - ntp * MT;
      const int nother = own.last - vblock;
      constexpr int kStep = TWO ? 2 : 1;                      // pair: the same-rank CTA of the next pair
      uint4* const mine = p.partials + ((size_t)blockIdx.x * 2 + own.slot) * kQuads;
      uint4* const others = p.partials + ((size_t)(blockIdx.x + kStep) * 2) * kQuads;   // their slot 0
      const size_t cstride = (size_t)kStep * 2 * kQuads;
      const int ft = threadIdx.x;                               // 0 .. kFixThreads - 1
      auto pass = [&](auto kq_, auto kb_) {
        constexpr int KQ = decltype(kq_)::value, KB = decltype(kb_)::value;
#pragma unroll 1
        for (int q0 = ft; q0 < kQuads; q0 += kFixThreads * KQ) {
          uint4 f0[KQ], f[KQ][KB];
#pragma unroll
          for (int i = 0; i < KQ; ++i) {
            const int qd = q0 + i * kFixThreads;
            if (qd < kQuads) f0[i] = ld_first_poll_u4(mine + qd, p.flags);
          }
#pragma unroll
          for (int i = 0; i < KQ; ++i)
#pragma unroll
            for (int u = 0; u < KB; ++u) {
              const int qd = q0 + i * kFixThreads;
              if (qd < kQuads && u < nother) f[i][u] = ld_first_poll_u4(others + u * cstride + qd, p.flags);
            }
          float4 acc[KQ];
#pragma unroll
          for (int i = 0; i < KQ; ++i) {
            const int qd = q0 + i * kFixThreads;
            if (qd < kQuads) acc[i] = part_take(mine + qd, f0[i]);
          }
          for (int base = 0; base < nother; base += KB) {
            if (base > 0) {
#pragma unroll
              for (int i = 0; i < KQ; ++i)
#pragma unroll
                for (int u = 0; u < KB; ++u) {
                  const int qd = q0 + i * kFixThreads;
                  if (qd < kQuads && base + u < nother)
                    f[i][u] = ld_first_poll_u4(others + (size_t)(base + u) * cstride + qd, p.flags);
                }
            }
#pragma unroll
            for (int i = 0; i < KQ; ++i)
#pragma unroll
              for (int u = 0; u < KB; ++u) {                    // CTA order: deterministic
                const int qd = q0 + i * kFixThreads;
                if (qd < kQuads && base + u < nother) {
                  const float4 g4 = part_take(others + (size_t)(base + u) * cstride + qd, f[i][u]);
                  acc[i].x += g4.x; acc[i].y += g4.y; acc[i].z += g4.z; acc[i].w += g4.w;
                }
              }
          }
#pragma unroll
          for (int i = 0; i < KQ; ++i) {
            const int qd = q0 + i * kFixThreads;
            if (qd < kQuads) {
              const int a = qd / kPartVec, r = qd % kPartVec;
              const int h = a / TP, t = a % TP;
              const int n_glob = ((TWO ? ntp * 2 + (int)crank : ntp) * PAIR + h) * kTileN + (r % kTileN);
              const int m = (mt * TP + t) * MTOK + 4 * (r / kTileN);
              if (n_glob < p.N) {
                const float rr[4] = {acc[i].x, acc[i].y, acc[i].z, acc[i].w};
#pragma unroll
                for (int u = 0; u < 4; ++u)
                  if (m + u < M) p.y[(size_t)(m + u) * p.N + n_glob] = __half_as_ushort(__float2half_rn(rr[u]));
              }
            }
          }
        }
      };
      // few contributors (prefill: a unit is split between two or three CTAs): more quads per
      // pass; many (small decode-like problems): more contributors per pass
      if (nother <= 3) pass(IntC<2>{}, IntC<3>{});
      else pass(IntC<1>{}, IntC<7>{});
    }
  }

  tc_fence_before();
  __syncthreads();
  if (TWO) cluster_sync_all();          // the peer may still read our shared memory / signal our barriers
  if (threadIdx.x == 0) TR(7, 1);
  if (warp == kAllocWarp) {
    tc_fence_after();
    if (TWO) tmem_dealloc_pair<kTmemCols>(tmem); else tmem_dealloc<kTmemCols>(tmem);
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

template <int MTOK, int PAIR, bool TWO = false, int TP = 1, bool DYN = false>
int launch_impl(const GemmParams& prm, const CUtensorMap& tmap, const GemmPlan& plan, bool pdl,
                cudaStream_t stream) {
  // the opt-in is per device: a process may drive several GPUs
  static bool attr_set[kMaxDevices] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= kMaxDevices || !attr_set[dev]) {
    cudaError_t e = cudaFuncSetAttribute(w4a16_gemm_kernel<MTOK, PAIR, TWO, TP, DYN>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Cfg<MTOK, PAIR, TWO, TP>::kSmemBytes);
    if (e != cudaSuccess) return (int)e;
    if (dev >= 0 && dev < kMaxDevices) attr_set[dev] = true;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(plan.ctas);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = Cfg<MTOK, PAIR, TWO, TP>::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (TWO) {                                              // CTA pairs: clusters of two on one TPC
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = 2;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
  }
  if (pdl) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  return (int)cudaLaunchKernelEx(&cfg, w4a16_gemm_kernel<MTOK, PAIR, TWO, TP, DYN>, prm, tmap);
}

int gemm_flags() {
#ifdef B200_TUNING   // `make tune` only: the product library reads no environment variables
  static const int f = [] {
    const char* e = getenv("B200_W4A16_FLAGS");
    return e ? atoi(e) : 0;
  }();
  return f;
#else
  return 0;
#endif
}

uint32_t* g_trace = nullptr;

template <int MTOK, int PAIR, bool TWO = false, int TP = 1>
void fill_plan(GemmPlan* plan) {
  plan->stages = Cfg<MTOK, PAIR, TWO, TP>::kWStages;
  plan->smem_bytes = Cfg<MTOK, PAIR, TWO, TP>::kSmemBytes;
}

// Tuning builds (`make tune`, -DB200_TUNING) can pin the variant with
// B200_W4A16_KERNEL=tcgen05|decode and the channel-tile grouping with B200_W4A16_PAIR=1|2; the
// product library dispatches by shape only.
int forced_kind() {
#ifdef B200_TUNING
  static const int k = [] {
    const char* e = getenv("B200_W4A16_KERNEL");
    if (!e) return -1;
    if (e[0] == 't') return (int)kKindTcgen05;
    if (e[0] == 'd') return (int)kKindDecode;
    return -1;
  }();
  return k;
#else
  return -1;
#endif
}
int forced_two() {
#ifdef B200_TUNING
  static const int k = [] {
    const char* e = getenv("B200_W4A16_TWO");
    return e ? atoi(e) : -1;
  }();
  return k;
#else
  return -1;
#endif
}
// Grid sizing when the output units are fewer than the SMs (see aligned_workers).
#ifndef B200_GEMM_SPLIT_CAP
#define B200_GEMM_SPLIT_CAP 4        // most workers sharing one output unit
#endif
#ifndef B200_GEMM_SPLIT_MIN_TOK
#define B200_GEMM_SPLIT_MIN_TOK 32   // ... from this token tile up (partial tiles of 16 KiB and more)
#endif
#ifndef B200_GEMM_ALIGN_PCT
#define B200_GEMM_ALIGN_PCT 64       // unit-aligned ranges when they keep at least this share of the workers busy
#endif
#ifndef B200_DECODE_ALIGN_PCT
#define B200_DECODE_ALIGN_PCT 80     // the decode kernel: same idea, ...
#endif
#ifndef B200_DECODE_ALIGN_MAX_ITEMS
#define B200_DECODE_ALIGN_MAX_ITEMS 30   // ... for short walks only (tiles per CTA); longer ones balance better over all SMs
#endif
int tuning_int(const char* name, int dflt) {
#ifdef B200_TUNING
  const char* e = getenv(name);
  return e ? atoi(e) : dflt;
#else
  (void)name;
  return dflt;
#endif
}
int split_cap() { static const int k = tuning_int("B200_W4A16_SPLITCAP", B200_GEMM_SPLIT_CAP); return k; }
int split_min_tok() { static const int k = tuning_int("B200_W4A16_SPLITMINTOK", B200_GEMM_SPLIT_MIN_TOK); return k; }
int align_pct() { static const int k = tuning_int("B200_W4A16_ALIGNPCT", B200_GEMM_ALIGN_PCT); return k; }
int decode_align_pct() { static const int k = tuning_int("B200_W4A16_DALIGNPCT", B200_DECODE_ALIGN_PCT); return k; }

// Stream-K cuts units x KB work items into equal ranges over all workers; when the units are fewer
// than the workers every unit is shared by workers / units of them, ranges straddle unit borders
// (two segments per worker) and the unit's owner sums the others' partial tiles through L2 ON ITS
// OWN while everybody else has left: at 128-token tiles 64 KiB per contributor and channel tile
// at ~36 B/clk -- 25 us of a 41 us launch on 8192 -> 2560 at 256 tokens
// (profiles/r4e_trace_m256_8192x2560.txt).  Two remedies, both a smaller grid:
//   * a cap on the workers per unit (cap > 0): few units with a long K;
//   * S = floor(workers / units) workers per unit, so that ranges END at unit borders (one
//     segment per worker, S - 1 contributors, none at S = 1), when that keeps `pct` percent of the
//     workers busy.
// Measured, three interleaved repetitions (profiles/r4g_*): 7-25 % on the Llama-2-7B / 13B / 70B-TP8
// layers at 32-1024 tokens wherever the grid changes, 2.1-2.3x on 8192 -> 1280 at 64-128 tokens,
// never slower beyond the run-to-run spread; the decode kernel gains 3-11 % on walks of up to 28
// tiles per CTA (4096 -> 4096 at 16 tokens: 6.44 -> 5.89 us) and loses 1-10 % on longer ones,
// hence B200_DECODE_ALIGN_MAX_ITEMS.
// Returns the worker count to use, or 0 to keep them all.
long long aligned_workers(long long units, int k_blocks, long long workers, int cap, int pct) {
  if (units <= 0 || units > workers || pct <= 0) return 0;
  long long S = workers / units;
  bool capped = false;
  if (cap > 0 && S > cap) { S = cap; capped = true; }
  if (S > k_blocks) S = k_blocks;
  if (S < 1) S = 1;
  const long long aligned = units * S;
  if (capped || aligned * 100 >= workers * pct) return aligned;
  return 0;
}
int forced_tp() {
#ifdef B200_TUNING
  static const int k = [] {
    const char* e = getenv("B200_W4A16_TP");
    return e ? atoi(e) : 0;
  }();
  return k;
#else
  return 0;
#endif
}
int forced_pair() {
#ifdef B200_TUNING
  static const int k = [] {
    const char* e = getenv("B200_W4A16_PAIR");
    return e ? atoi(e) : 0;
  }();
  return k;
#else
  return 0;
#endif
}

}  // namespace

void set_trace_buffer(uint32_t* dev) { g_trace = dev; }
uint32_t* trace_buffer() { return g_trace; }

int make_gemm_plan(int M, int K, int N, int G, int split_k_hint, int sm_count, GemmPlan* plan) {
  (void)split_k_hint;
  if (sm_count <= 0) sm_count = kMaxCtas;
  if (sm_count > kMaxCtas) sm_count = kMaxCtas;
  plan->n_tiles = (N + kTileN - 1) / kTileN;
  plan->k_blocks = K / kTileK;
  plan->pair = 1;
  plan->two_cta = 0;
  plan->tok_pair = 1;
  const int fk = forced_kind();
  const bool group_ok = G % kTileK == 0;                  // one (scale, zero) row per 128-row tile
  bool decode = group_ok && M <= kDecodeMaxTokens;
  if (fk == (int)kKindTcgen05) decode = false;
  if (decode) {
    // Decode variant: stream-K over (channel tile, K block), one persistent CTA per SM.
    plan->kind = kKindDecode;
    plan->m_tiles = 1;
    const long long T = (long long)plan->n_tiles * plan->k_blocks;
    if (T >= (1LL << 31)) return 1;
    long long ctas = (long long)sm_count * kDecodeCtasPerSm;
    if (T < ctas) ctas = T;
    {   // unit-aligned ranges on short walks: one segment per CTA, fewer partial tiles to fix up
      const long long a = aligned_workers(plan->n_tiles, plan->k_blocks, ctas, 0, decode_align_pct());
      if (a > 0 && T <= a * B200_DECODE_ALIGN_MAX_ITEMS) ctas = a;
    }
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
  // Channel tiles in groups of two for 128-token tiles (prefill: the activation tile is 4x the
  // weight tile and L2 -> SM bytes bound the kernel) when that still leaves every SM a long walk
  // (>= 8 unit steps): one activation tile then feeds two weight tiles.  Below 128 tokens single
  // tiles measured faster on every shape but one (profiles/r2q_ring_pair_variants.md), and small
  // problems balance better with the finer work items.
  {
    const long long units2 = (long long)((plan->n_tiles + 1) / 2) * plan->m_tiles * plan->k_blocks;
    if (tok == 128 && plan->n_tiles >= 2 && units2 >= 8LL * sm_count) plan->pair = 2;
    const int fp = forced_pair();
    if (fp == 1) plan->pair = 1;
    if (fp == 2 && plan->n_tiles >= 2) plan->pair = 2;
  }
  // CTA pairs (cta_group::2): with channel-tile pairs, groups of FOUR channel tiles per work unit,
  // two per CTA, the activation tile split between the two SMs -- 16 instead of 24 KiB into an SM
  // per 128 x 128 x 128 MMA block.  Needs an even grid and a long walk per pair.
  {
    const long long units4 = (long long)((plan->n_tiles + 3) / 4) * plan->m_tiles * plan->k_blocks;
    const int pairs = sm_count / 2;
    if (plan->pair == 2 && plan->n_tiles >= 4 && pairs >= 1 && units4 >= 8LL * pairs) plan->two_cta = B200_GEMM_TWO_CTA_DEFAULT;
    // Short K and not too many token tiles: ONE channel tile per CTA of the pair -- two accumulator
    // sets again (a pair's drain stall is paid once per output unit, i.e. per K / 128 steps), finer
    // work units; +8-27 % at 512-2048 tokens on the K = 5120 layers, a loss at K = 13824 and at
    // 8192 tokens (profiles/r3c_kb_pair{1,2}_two.jsonl).
    if (plan->two_cta && plan->k_blocks < 64 && plan->m_tiles <= 32) plan->pair = 1;
    const int ft = forced_two();
    if (ft == 0) plan->two_cta = 0;
    if (ft == 1 && plan->pair == 2 && plan->n_tiles >= 4 && pairs >= 1 && units4 >= pairs) plan->two_cta = 1;
    // tuning: CTA pairs with ONE channel tile per CTA (two accumulator sets again)
    if (ft == 1 && tok == 128 && plan->pair == 1 && plan->n_tiles >= 2 && pairs >= 1) plan->two_cta = 1;
  }
  // Token-tile pairs on CTA pairs with one channel tile per CTA: one dequantized weight tile in TMEM
  // feeds the MMAs of TWO 128-token tiles (half the dequant work and weight bytes per flop: the
  // steady state is then the tensor pipe itself, 513 cycles per 128 x 128 x 128 block,
  // profiles/r3h_trace_tp2.txt).  Units are twice as long and the accumulators single-buffered,
  // which loses on short walks: from 32 token tiles up, on wide or deep layers (+5-14 % at
  // 4096-8192 tokens, up to 1.37 PFLOP/s; profiles/r3h_kb_{policy,tp2}.jsonl).
  if (plan->two_cta && plan->m_tiles >= 32 && (plan->n_tiles >= 64 || plan->k_blocks >= 64)) {
    plan->pair = 1;
    plan->tok_pair = 2;
  }
  if (forced_tp() == 1) plan->tok_pair = 1;
  if (forced_tp() == 2 && plan->two_cta && plan->m_tiles >= 2) { plan->pair = 1; plan->tok_pair = 2; }
  const int per_unit = plan->pair * (plan->two_cta ? 2 : 1);   // channel tiles per work unit
  const int ntp = (plan->n_tiles + per_unit - 1) / per_unit;
  const int mtp = (plan->m_tiles + plan->tok_pair - 1) / plan->tok_pair;
  const long long T = (long long)ntp * mtp * plan->k_blocks;
  if (T >= (1LL << 31)) return 1;
  // Persistent stream-K: one CTA (or CTA pair) per SM (pair), equal contiguous ranges of (output
  // unit, K block).
  long long ctas = T < sm_count ? T : sm_count;
  if (plan->two_cta) {
    const long long workers = T < sm_count / 2 ? T : sm_count / 2;
    ctas = 2 * workers;
  }
  // Few output units with a long K (8192 -> 1280: 10 units of 64 K blocks): equal ranges over all
  // SMs would cut every unit among ~15 CTAs, and its owner would sum that many 32-64 KiB partial
  // tiles through L2 on its own -- 22 us at 64 tokens for 2 us of bytes.  At these tile sizes the
  // contributors per unit are capped; the idle SMs cost less than the serial reduction.
  {
    const long long units = (long long)ntp * mtp;
    long long workers = plan->two_cta ? ctas / 2 : ctas;
    const long long aligned = aligned_workers(units, plan->k_blocks, workers, tok >= split_min_tok() ? split_cap() : 0,
                                              align_pct());
    if (aligned > 0) ctas = plan->two_cta ? 2 * aligned : aligned;
  }
  if (ctas < 1) ctas = 1;
  plan->ctas = (int)ctas;
  const long long workers_ = plan->two_cta ? ctas / 2 : ctas;
  const long long per = T / workers_;  // shortest range
  plan->max_split = (plan->k_blocks == 0 || (T % workers_ == 0 && per % plan->k_blocks == 0))
                        ? 1
                        : (int)((plan->k_blocks + per - 1) / (per > 0 ? per : 1) + 1);
  plan->ws_partial_bytes = (size_t)ctas * 2 * plan->pair * plan->tok_pair * tok * kTileN * sizeof(float);
  if (plan->two_cta) {
    if (plan->tok_pair == 2) fill_plan<128, 1, true, 2>(plan);
    else if (plan->pair == 2) fill_plan<128, 2, true>(plan); else fill_plan<128, 1, true>(plan);
    return 0;
  }
  if (plan->pair == 2) {
    switch (tok) {
      case 16: fill_plan<16, 2>(plan); break;
      case 32: fill_plan<32, 2>(plan); break;
      case 64: fill_plan<64, 2>(plan); break;
      default: fill_plan<128, 2>(plan); break;
    }
    return 0;
  }
  switch (tok) {
    case 16: fill_plan<16, 1>(plan); break;
    case 32: fill_plan<32, 1>(plan); break;
    case 64: fill_plan<64, 1>(plan); break;
    default: fill_plan<128, 1>(plan); break;
  }
  return 0;
}

// Plan for a launch whose live row count is read on the device (<= M_max, the buffers' rows):
// single-CTA, single-tile variant, token tile by M_max, the grid of the M_max problem.
int make_gemm_plan_rows(int M_max, int K, int N, int G, int sm_count, GemmPlan* plan) {
  (void)G;
  if (sm_count <= 0) sm_count = kMaxCtas;
  if (sm_count > kMaxCtas) sm_count = kMaxCtas;
  plan->kind = kKindTcgen05;
  plan->n_tiles = (N + kTileN - 1) / kTileN;
  plan->k_blocks = K / kTileK;
  plan->pair = 1;
  plan->two_cta = 0;
  plan->tok_pair = 1;
  const int tok = M_max <= 32 ? 32 : M_max <= 64 ? 64 : 128;
  plan->tok_tile = tok;
  plan->m_tiles = (M_max + tok - 1) / tok;
  const long long T = (long long)plan->n_tiles * plan->m_tiles * plan->k_blocks;
  if (T >= (1LL << 31)) return 1;
  long long ctas = T < sm_count ? T : sm_count;
  if (ctas < 1) ctas = 1;
  plan->ctas = (int)ctas;
  plan->max_split = plan->k_blocks + 1;                     // the live problem may be much smaller
  plan->ws_partial_bytes = (size_t)ctas * 2 * tok * kTileN * sizeof(float);
  switch (tok) {
    case 32: fill_plan<32, 1>(plan); break;
    case 64: fill_plan<64, 1>(plan); break;
    default: fill_plan<128, 1>(plan); break;
  }
  return 0;
}

int launch_gemm(const uint16_t* x, const int32_t* qw_b200, const uint16_t* scales,
                const int32_t* qzeros, uint16_t* y, int M, int K, int N, int G,
                const GemmPlan& plan, void* ws_partials, bool pdl, cudaStream_t stream,
                const int* m_dev) {
  EncodeTiledFn encode = get_encode_fn();
  if (!encode) return (int)cudaErrorNotSupported;
  CUtensorMap tmap;
  const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)M};
  const cuuint64_t strides[1] = {(cuuint64_t)K * 2};
  const cuuint32_t box[2] = {64u, (cuuint32_t)(plan.two_cta ? plan.tok_tile / 2 : plan.tok_tile)};
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
  prm.partials = reinterpret_cast<uint4*>(ws_partials);
  prm.M = M; prm.K = K; prm.N = N; prm.G = G;
  prm.KB = plan.k_blocks; prm.NT = plan.n_tiles;
  prm.MT = (plan.m_tiles + plan.tok_pair - 1) / plan.tok_pair;      // token-tile groups
  const int per_unit = plan.pair * (plan.two_cta ? 2 : 1);
  prm.NTP = (plan.n_tiles + per_unit - 1) / per_unit;
  prm.T = prm.NTP * prm.MT * plan.k_blocks;
  prm.P = plan.ctas;
  prm.R = G >= kTileK ? 1 : kTileK / G;
  prm.Gdiv = G >= kTileK ? G / kTileK : 1;
  prm.flags = gemm_flags();
  // (The weight stream keeps its L2 evict-first hint at every batch size.  Keeping weights
  // resident across token tiles was measured -- profiles/r2t_kb_prefill_evict_normal.jsonl --
  // and lost: the operand with the long reuse distance is the ACTIVATION matrix, re-read once
  // per channel-tile group; 13824 -> 5120 at 2048 tokens fell from 1144 to 932 TFLOP/s.)
  prm.trace = g_trace;
  prm.m_dev = m_dev;
  prm.split_cap = plan.tok_tile >= split_min_tok() ? split_cap() : 0;
  prm.align_pct = align_pct();
  if (m_dev) {
    if (plan.two_cta || plan.pair != 1 || plan.tok_pair != 1) return (int)cudaErrorInvalidValue;
    switch (plan.tok_tile) {
      case 32: return launch_impl<32, 1, false, 1, true>(prm, tmap, plan, pdl, stream);
      case 64: return launch_impl<64, 1, false, 1, true>(prm, tmap, plan, pdl, stream);
      case 128: return launch_impl<128, 1, false, 1, true>(prm, tmap, plan, pdl, stream);
      default: return (int)cudaErrorInvalidValue;
    }
  }
  if (plan.two_cta) {
    if (plan.tok_pair == 2) return launch_impl<128, 1, true, 2>(prm, tmap, plan, pdl, stream);
    if (plan.pair == 2) return launch_impl<128, 2, true>(prm, tmap, plan, pdl, stream);
    return launch_impl<128, 1, true>(prm, tmap, plan, pdl, stream);
  }
  if (plan.pair == 2) {
    switch (plan.tok_tile) {
      case 16: return launch_impl<16, 2>(prm, tmap, plan, pdl, stream);
      case 32: return launch_impl<32, 2>(prm, tmap, plan, pdl, stream);
      case 64: return launch_impl<64, 2>(prm, tmap, plan, pdl, stream);
      default: return launch_impl<128, 2>(prm, tmap, plan, pdl, stream);
    }
  }
  switch (plan.tok_tile) {
    case 16: return launch_impl<16, 1>(prm, tmap, plan, pdl, stream);
    case 32: return launch_impl<32, 1>(prm, tmap, plan, pdl, stream);
    case 64: return launch_impl<64, 1>(prm, tmap, plan, pdl, stream);
    default: return launch_impl<128, 1>(prm, tmap, plan, pdl, stream);
  }
}

}  // namespace b200
