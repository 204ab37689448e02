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
//  * dequant warps read one 16-byte word group per thread (bank-conflict free), expand it in
//    registers with the reference's exact lop3 / sub / fma sequence, and write the fp16 pairs
//    straight into TENSOR MEMORY with tcgen05.st: thread = TMEM lane = channel, so the
//    registers already are the MMA A-operand -- no shared-memory round trip for the
//    dequantized tile (the reference's :168 store + :198-203 ldmatrix).
//  * one thread issues tcgen05.mma (A from TMEM, B from shared memory), accumulators in TMEM.
//  * stream-K: the (channel tile, token tile, K block) space is cut into equal contiguous
//    ranges, one per SM; an output tile touched by several CTAs is reduced in fixed CTA order
//    by whichever CTA arrives last (fp32 partials through L2, no second kernel, deterministic).
//  * programmatic dependent launch: the weight pipeline is filled BEFORE griddepcontrol.wait
//    (weights never depend on the previous kernel), so launch ramp overlaps the previous op.
#include "w4a16_common.cuh"
#include "w4a16_internal.h"

namespace b200 {

namespace {

constexpr int kProducerWarp = 0;
constexpr int kMmaWarp = 1;
constexpr int kAllocWarp = 2;
constexpr int kFirstDequantWarp = 4;
constexpr int kDequantGroups = 2;                       // warpgroups alternating tiles
constexpr int kDequantThreads = kDequantGroups * 128;
constexpr int kThreads = kFirstDequantWarp * 32 + kDequantThreads;  // 384
constexpr int kScaleRowsMax = 4;                        // G = 32 -> four (scale, zero) rows / tile
constexpr int kTmemACols = 64;                          // one dequantized tile: 128 fp16 per lane
constexpr int kTmemDCol = kDequantGroups * kTmemACols;  // accumulators start here
constexpr int kTmemCols = 256;

template <int MTOK>
struct Cfg {
  static constexpr int kXBytes = MTOK * 256;           // two swizzle atoms of [MTOK x 64] fp16
  static constexpr int kWOff = kXBytes;
  static constexpr int kSOff = kWOff + kTileWords * 4;
  static constexpr int kZOff = kSOff + kScaleRowsMax * 256;
  static constexpr int kStageBytes = ((kZOff + kScaleRowsMax * 64 + 1023) / 1024) * 1024;
  static constexpr int kStages = MTOK <= 16 ? 8 : MTOK <= 32 ? 6 : 4;
  static constexpr int kBarBytes = 1024;
  static constexpr int kSmemBytes = kBarBytes + kStages * kStageBytes + 1024;  // + align slack
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
  long long T;  // total (output tile, K block) work items
  int P;        // CTAs
  int R;        // (scale, zero) rows per K block: max(1, 128 / G)
  int Gdiv;     // K blocks per group: max(1, G / 128)
};

__device__ __forceinline__ int owner_of(long long t, int P, long long T) {
  return (int)(((t + 1) * P - 1) / T);
}

template <int MTOK>
__global__ void __launch_bounds__(kThreads, 1)
w4a16_gemm_kernel(const GemmParams p, const __grid_constant__ CUtensorMap tmap_x) {
  using C = Cfg<MTOK>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));

  // barrier block
  const uint32_t bar_full = smem_base;                       // [kStages]
  const uint32_t bar_empty = smem_base + 8 * C::kStages;     // [kStages]
  const uint32_t bar_afull = smem_base + 16 * C::kStages;    // [kDequantGroups]
  const uint32_t bar_aempty = bar_afull + 8 * kDequantGroups;
  const uint32_t bar_dfull = bar_aempty + 8 * kDequantGroups;
  const uint32_t bar_dempty = bar_dfull + 8;
  const uint32_t tmem_slot = bar_dempty + 8;
  const uint32_t flag_slot = tmem_slot + 4;
  volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(
      smem_gen + (tmem_slot - smem_base));
  volatile uint32_t* flag_gen = reinterpret_cast<volatile uint32_t*>(
      smem_gen + (flag_slot - smem_base));
  const uint32_t stage0 = smem_base + C::kBarBytes;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == kProducerWarp && lane == 0) {
    tma_prefetch_desc(&tmap_x);
    for (int s = 0; s < C::kStages; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 1);
    }
    for (int g = 0; g < kDequantGroups; ++g) {
      mbar_init(bar_afull + 8 * g, 128);
      mbar_init(bar_aempty + 8 * g, 1);
    }
    mbar_init(bar_dfull, 1);
    mbar_init(bar_dempty, kDequantThreads);
    fence_barrier_init();
  } else if (warp == kAllocWarp) {
    tmem_alloc<kTmemCols>(tmem_slot);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot_gen;

  // This CTA's contiguous slice of the work space.
  const long long t_begin = (long long)blockIdx.x * p.T / p.P;
  const long long t_end = (long long)(blockIdx.x + 1) * p.T / p.P;

  if (warp == kProducerWarp) {
    if (lane == 0) {
      // Let the next kernel in the stream start its own weight prefetch right away.
      asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
      const uint64_t pol_w = policy_evict_first();
      const int npre = (int)((t_end - t_begin) < C::kStages ? (t_end - t_begin) : C::kStages);
      // -- phase 1: weight side of the first npre stages (independent of the previous kernel)
      // Tile coordinates advance incrementally: no divisions on the single producer thread.
      struct Coord { int kb, nt, mt; };
      auto advance = [&](Coord& c) {
        if (++c.kb == p.KB) {
          c.kb = 0;
          if (++c.mt == p.MT) { c.mt = 0; ++c.nt; }
        }
      };
      Coord cw, cx;
      {
        const int o = (int)(t_begin / p.KB);
        cw.kb = (int)(t_begin - (long long)o * p.KB);
        cw.nt = o / p.MT;
        cw.mt = o - cw.nt * p.MT;
        cx = cw;
      }
      const int n_total = (int)(t_end - t_begin);
      const uint32_t zs_row_bytes_full = kTileN * 2 + kTileN / 2;
      auto issue_w = [&](const Coord& c, int stage) {
        const int nvalid = min(kTileN, p.N - c.nt * kTileN);
        const uint32_t sb = stage0 + stage * C::kStageBytes;
        const uint32_t bar = bar_full + 8 * stage;
        const uint32_t zs_row = nvalid == kTileN ? zs_row_bytes_full : (uint32_t)(nvalid * 2 + nvalid / 2);
        mbar_arrive_expect_tx(bar, C::kXBytes + kTileWords * 4 + p.R * zs_row);
        bulk_g2s_hint(sb + C::kWOff, p.qw + ((size_t)c.nt * p.KB + c.kb) * kTileWords,
                      kTileWords * 4, bar, pol_w);
        const int g0 = p.R > 1 ? c.kb * p.R : (p.Gdiv == 1 ? c.kb : c.kb / p.Gdiv);
        for (int r = 0; r < p.R; ++r) {
          bulk_g2s(sb + C::kSOff + r * 256, p.scales + (size_t)(g0 + r) * p.N + c.nt * kTileN,
                   nvalid * 2, bar);
          bulk_g2s(sb + C::kZOff + r * 64, p.qz + (size_t)(g0 + r) * (p.N >> 3) + c.nt * 16,
                   nvalid / 2, bar);
        }
      };
      auto issue_x = [&](const Coord& c, int stage) {
        const uint32_t sb = stage0 + stage * C::kStageBytes;
        const uint32_t bar = bar_full + 8 * stage;
        tma_load_2d(sb, &tmap_x, c.kb * kTileK, c.mt * MTOK, bar);
        tma_load_2d(sb + MTOK * 128, &tmap_x, c.kb * kTileK + 64, c.mt * MTOK, bar);
      };
      for (int i = 0; i < npre; ++i) { issue_w(cw, i); advance(cw); }
      asm volatile("griddepcontrol.wait;" ::: "memory");
      for (int i = 0; i < npre; ++i) { issue_x(cx, i); advance(cx); }
      // -- phase 2: steady state
      int stage = npre % C::kStages;
      uint32_t ephase = npre == C::kStages ? 0u : 1u;  // parity to wait for on empty[stage]
      for (int it = npre; it < n_total; ++it) {
        mbar_wait(bar_empty + 8 * stage, ephase);
        issue_w(cw, stage);
        issue_x(cw, stage);
        advance(cw);
        if (++stage == C::kStages) { stage = 0; ephase ^= 1u; }
      }
    }
  } else if (warp == kMmaWarp) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_f16_m128(MTOK);
      int it = 0, seg = 0;
      int kb0 = (int)(t_begin % p.KB);
      int remaining = (int)(t_end - t_begin);
      int stage = 0;
      uint32_t fphase = 0;
      while (remaining > 0) {
        const int n = min(p.KB - kb0, remaining);
        mbar_wait(bar_dempty, (seg & 1) ^ 1);  // accumulators drained by the previous epilogue
        tc_fence_after();
        for (int i = 0; i < n; ++i, ++it) {
          const int a = it % kDequantGroups;
          mbar_wait(bar_full + 8 * stage, fphase);
          mbar_wait(bar_afull + 8 * a, (it / kDequantGroups) & 1);
          tc_fence_after();
          const uint32_t xb = stage0 + stage * C::kStageBytes;
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            const uint64_t bdesc = make_desc_k_sw128(xb + (kk >> 2) * (MTOK * 128) + (kk & 3) * 32);
            tc_mma_ts(tmem + kTmemDCol, tmem + a * kTmemACols + kk * 8, bdesc, idesc,
                      (i | kk) != 0);
          }
          tc_commit(bar_empty + 8 * stage);
          tc_commit(bar_aempty + 8 * a);
          if (++stage == C::kStages) { stage = 0; fphase ^= 1u; }
        }
        tc_commit(bar_dfull);
        remaining -= n;
        kb0 = 0;
        ++seg;
      }
    }
  } else if (warp >= kFirstDequantWarp) {
    const int wi = warp - kFirstDequantWarp;
    const int grp = wi >> 2, q = wi & 3;
    const int c = q * 32 + lane;                              // channel inside the tile
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    const int zsh = 4 * awq_nibble(c & 7);
    asm volatile("griddepcontrol.wait;" ::: "memory");        // y / workspace may alias earlier work
    int it = 0, seg = 0;
    long long t = t_begin;
    while (t < t_end) {
      const int o = (int)(t / p.KB), kb0 = (int)(t - (long long)o * p.KB);
      const int n = (int)min((long long)(p.KB - kb0), t_end - t);
      const int nt = o / p.MT, mt = o % p.MT;
      // ---- dequantize this group's tiles of the segment into tensor memory
      for (int i = 0; i < n; ++i, ++it) {
        if ((it % kDequantGroups) != grp) continue;
        const int stage = it % C::kStages;
        const uint8_t* sb = smem_gen + C::kBarBytes + stage * C::kStageBytes;
        mbar_wait(bar_full + 8 * stage, (it / C::kStages) & 1);
        uint4 wv[4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
          wv[j] = *reinterpret_cast<const uint4*>(sb + C::kWOff + j * 2048 + c * 16);
        const uint16_t* sS = reinterpret_cast<const uint16_t*>(sb + C::kSOff);
        const uint32_t* sZ = reinterpret_cast<const uint32_t*>(sb + C::kZOff);
        mbar_wait(bar_aempty + 8 * grp, ((it / kDequantGroups) & 1) ^ 1);
        tc_fence_after();
        uint32_t zlo = 0, zhi = 0, s2 = 0;
        int rprev = -1;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int r = (p.R == 1) ? 0 : (j * 32) / p.G;
          if (r != rprev) {
            make_zs1((sZ[r * 16 + (c >> 3)] >> zsh) & 0xFu, sS[r * 128 + c], zlo, zhi, s2);
            rprev = r;
          }
          uint32_t v[16];
          dequant_word_zs1(wv[j].x, zlo, zhi, s2, v[0], v[1], v[2], v[3]);
          dequant_word_zs1(wv[j].y, zlo, zhi, s2, v[4], v[5], v[6], v[7]);
          dequant_word_zs1(wv[j].z, zlo, zhi, s2, v[8], v[9], v[10], v[11]);
          dequant_word_zs1(wv[j].w, zlo, zhi, s2, v[12], v[13], v[14], v[15]);
          tmem_st_x16(tmem + lane_base + grp * kTmemACols + j * 16, v);
        }
        tc_wait_st();
        tc_fence_before();
        mbar_arrive(bar_afull + 8 * grp);
      }
      // ---- epilogue of the segment: this group owns half of the token columns
      mbar_wait(bar_dfull, seg & 1);
      tc_fence_after();
      const bool whole = (kb0 == 0 && n == p.KB);
      const int n_glob = nt * kTileN + c;
      constexpr int kColsPerGroup = MTOK / kDequantGroups;
      const int col0 = grp * kColsPerGroup;
      float* part = p.partials +
                    ((size_t)blockIdx.x * 2 + (t == t_begin ? 0 : 1)) * (size_t)(MTOK * kTileN);
#pragma unroll 1
      for (int cc = 0; cc < kColsPerGroup; cc += 8) {
        uint32_t v[8];
        tmem_ld_x8(tmem + lane_base + kTmemDCol + col0 + cc, v);
        tc_wait_ld();
        if (whole) {
          if (n_glob < p.N) {
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const int m = mt * MTOK + col0 + cc + e;
              if (m < p.M)
                p.y[(size_t)m * p.N + n_glob] = __half_as_ushort(__float2half_rn(__uint_as_float(v[e])));
            }
          }
        } else {
#pragma unroll
          for (int e = 0; e < 8; ++e) part[(col0 + cc + e) * kTileN + c] = __uint_as_float(v[e]);
        }
      }
      tc_fence_before();
      mbar_arrive(bar_dempty);
      if (!whole) {
        // stream-K fix-up: the last CTA to arrive sums all contributors in CTA order.
        const long long o_lo = (long long)o * p.KB;
        const int p_first = owner_of(o_lo, p.P, p.T);
        const int p_last = owner_of(o_lo + p.KB - 1, p.P, p.T);
        __threadfence();
        named_bar_sync(1, kDequantThreads);
        if (wi == 0 && lane == 0) {
          const int old = atomicAdd(p.counters + o, 1);
          *flag_gen = (old == p_last - p_first) ? 1u : 0u;
        }
        named_bar_sync(1, kDequantThreads);
        const bool last = (*flag_gen != 0u);
        named_bar_sync(1, kDequantThreads);  // flag consumed before anyone can overwrite it
        if (last) {
          __threadfence();
          if (wi == 0 && lane == 0) p.counters[o] = 0;
          if (n_glob < p.N) {
            for (int cc = 0; cc < kColsPerGroup; ++cc) {
              const int col = col0 + cc;
              const int m = mt * MTOK + col;
              if (m >= p.M) break;
              float acc = 0.f;
              for (int pc = p_first; pc <= p_last; ++pc) {
                const long long pc_begin = (long long)pc * p.T / p.P;
                const float* src = p.partials +
                                   ((size_t)pc * 2 + (pc_begin >= o_lo ? 0 : 1)) * (size_t)(MTOK * kTileN);
                acc += __ldcg(src + col * kTileN + c);
              }
              p.y[(size_t)m * p.N + n_glob] = __half_as_ushort(__float2half_rn(acc));
            }
          }
        }
      }
      t += n;
      ++seg;
    }
  }

  tc_fence_before();
  __syncthreads();
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

template <int MTOK>
void fill_plan(GemmPlan* plan) {
  plan->stages = Cfg<MTOK>::kStages;
  plan->smem_bytes = Cfg<MTOK>::kSmemBytes;
}

}  // namespace

int make_gemm_plan(int M, int K, int N, int G, int split_k_hint, int sm_count, GemmPlan* plan) {
  (void)G;
  (void)split_k_hint;
  const int tok = M <= 16 ? 16 : M <= 32 ? 32 : M <= 64 ? 64 : 128;
  plan->tok_tile = tok;
  plan->n_tiles = (N + kTileN - 1) / kTileN;
  plan->m_tiles = (M + tok - 1) / tok;
  plan->k_blocks = K / kTileK;
  const long long outs = (long long)plan->n_tiles * plan->m_tiles;
  const long long T = outs * plan->k_blocks;
  if (sm_count <= 0) sm_count = kMaxCtas;
  // Few output tiles: stream-K over one wave of SMs.  Many: one CTA per output tile.
  long long ctas = outs >= 2LL * sm_count ? outs : (T < sm_count ? T : sm_count);
  if (ctas < 1) ctas = 1;
  plan->ctas = (int)ctas;
  const long long per = (T + ctas - 1) / ctas;
  plan->max_split = (ctas == outs) ? 1 : (int)((plan->k_blocks + per - 1) / per + 1);
  plan->ws_partial_bytes = (ctas == outs) ? 0 : (size_t)ctas * 2 * tok * kTileN * sizeof(float);
  plan->ws_counter_bytes = (size_t)outs * sizeof(int);
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
  prm.T = (long long)plan.n_tiles * plan.m_tiles * plan.k_blocks;
  prm.P = plan.ctas;
  prm.R = G >= kTileK ? 1 : kTileK / G;
  prm.Gdiv = G >= kTileK ? G / kTileK : 1;
  switch (plan.tok_tile) {
    case 16: return launch_impl<16>(prm, tmap, plan, pdl, stream);
    case 32: return launch_impl<32>(prm, tmap, plan, pdl, stream);
    case 64: return launch_impl<64>(prm, tmap, plan, pdl, stream);
    default: return launch_impl<128>(prm, tmap, plan, pdl, stream);
  }
}

}  // namespace b200
