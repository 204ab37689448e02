// w4a16_norm.cu -- residual add + RMSNorm, standalone and fused with the one-shot NVLink
// all-reduce that follows every row-parallel W4A16 GEMM.
//
// Replaces:
//   rms_norm / fused_add_rms_norm          csrc/layernorm_kernels.cu:12-76, 78-129
//   RMSNorm.forward                        vllm/model_executor/layers/layernorm.py:46-66
//   and, fused, the sequence of a Llama decoder layer
//       o_proj / down_proj all-reduce  ->  fused_add_rms_norm
//   (vllm/model_executor/models/llama.py:205-222, layers/linear.py:541-542).
//
// Arithmetic (same as the reference, fp16 storage):
//   x   = float(in) [+ float(residual)]          residual_out = half(x)
//   var = sum(x * x) / hidden                    (x before rounding, fp32)
//   out = half(float(half(x)) * rsqrt(var + eps)) * weight        (half multiply, RNE)
// Without a residual half(x) == in.  The sum of squares is a fixed-order tree (lane shuffle,
// warps, cluster ranks), so results are deterministic; it is not the reference's
// blockReduceSum order (floating point: tolerance parity, see tests/test_gpu_norm.py).
//
// Layout: a token row of `hidden` halfs is cut into 16-byte chunks; a thread-block CLUSTER of
// C = ceil(hidden / 2048) CTAs x 256 threads owns a row, CTA j holding chunks [256 j, 256 j +
// 256): every value stays in registers between the statistics pass and the scaling pass, and
// the C partial sums of squares meet through distributed shared memory -- no second read of
// the row, no global scratch.  Row clusters stride over the rows.
//
// Fused variant: the all-reduce protocol of w4a16_allreduce.cu (same region, flags and epochs,
// so both kernels may be mixed on one stream): block b stages ITS chunks of every row it owns
// into the rank's slot, crosses the per-block NVLink barrier, then sums the peers' copies in
// rank order, rounds to fp16 (== the unfused all-reduce result, bit for bit) and continues
// into the add + norm above without the [rows, hidden] round trip through HBM and without a
// second launch.
#include <cooperative_groups.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "w4a16_common.cuh"
#include "w4a16_internal.h"

namespace cg = cooperative_groups;

namespace b200 {

namespace {

constexpr int kNormThreads = 256;
constexpr int kNormMaxCluster = 8;     // hidden <= 8 * 256 * 8 = 16384

struct NormParams {
  uint8_t* region[kArMaxRanks];   // fused variant: peer-mapped regions, index = rank
  const uint4* in;                // [rows, hidden] half
  const uint4* residual_in;       // [rows, hidden] half or nullptr
  const uint4* weight;            // [hidden] half
  uint4* out;                     // [rows, hidden] half (may alias in)
  uint4* residual_out;            // [rows, hidden] half or nullptr (may alias residual_in)
  float eps;
  float inv_hidden;
  uint32_t rows;
  uint32_t chunks_per_row;        // hidden / 8
  uint32_t slot_bytes;
  int rank;
  int barrier;
  // LL variant: flag-in-data packets pushed into the peers' regions (see add_rmsnorm_kernel)
  uint32_t ll_off;                // byte offset of the packet area inside a region
  uint32_t ll_slot_bytes;         // packet bytes of one source rank and parity (2 x message)
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_volatile_v4(const uint4* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}

__device__ __forceinline__ void unpack8(float (&f)[8], const uint4& v) {
  const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __half22float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}

// WORLD == 0: plain (x read from `in`); WORLD >= 1: x = all-reduce over WORLD ranks of `in`.
// CPT: 16-byte chunks per thread (chunk j of a thread = cta_base + j * 256 + tid: coalesced).
// CPT = 1 spreads a row over the most CTAs (decode: few rows, latency); CPT = 4 puts up to
// 8192 halfs in ONE CTA, so the per-row cluster barrier (a membar + barrier pair that was 40 %
// of the stall samples, profiles/r1e_ncu_full_add_rmsnorm.txt) shrinks to a __syncthreads.
//
// LL ("flag in data", WORLD >= 2, CPT == 1): instead of stage -> system fence -> flag -> barrier
// -> remote read, every 16-byte chunk is PUSHED to each peer as two 16-byte packets
// {data0, epoch, data1, epoch} (one vector store each, no fence); the receiver spins on the
// packets themselves until both epoch words match -- one NVLink one-way trip instead of two
// plus a round trip.  The packet area [2 parities][WORLD sources][2 x message bytes] lives behind
// the other slots of the region and only ever holds packets, so a stale word can never look
// like the current epoch (epochs only grow).  Same epochs, same parity rule as the other kernels.
__device__ __forceinline__ void st_volatile_v4(uint4* p, uint4 v) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}

template <int WORLD, int CPT, bool LL = false>
__global__ void __launch_bounds__(kNormThreads)
add_rmsnorm_kernel(const NormParams p) {
  cg::cluster_group cluster = cg::this_cluster();
  const uint32_t csize = cluster.num_blocks();
  const uint32_t crank = cluster.block_rank();
  const uint32_t row0 = blockIdx.x / csize;
  const uint32_t row_step = gridDim.x / csize;
  uint32_t ch[CPT];                                            // chunks of the row this thread owns
  bool active[CPT];
#pragma unroll
  for (int j = 0; j < CPT; ++j) {
    ch[j] = (crank * CPT + j) * kNormThreads + threadIdx.x;
    active[j] = ch[j] < p.chunks_per_row;
  }

  __shared__ float s_warp[2 * CPT * (kNormThreads / 32)];       // [row parity][group][warp]
  __shared__ float s_part[2 * CPT];                            // [row parity][group]: per-CTA partials
  __shared__ uint32_t s_epoch;

  asm volatile("griddepcontrol.wait;" ::: "memory");

  uint32_t slot_off = 0;
  if constexpr (WORLD > 0) {
    static_assert(WORLD == 0 || CPT == 1, "the fused variant maps one chunk per thread");
    uint8_t* mine = p.region[p.rank];
    uint32_t* my_flags = reinterpret_cast<uint32_t*>(mine) + blockIdx.x * kArMaxRanks;
    uint32_t* epochs = reinterpret_cast<uint32_t*>(mine + kArFlagBytes);
    if (threadIdx.x == 0) s_epoch = epochs[blockIdx.x] + 1;
    __syncthreads();
    const uint32_t epoch = s_epoch;
    slot_off = kArHeaderBytes + (epoch & 1u) * p.slot_bytes;
    if constexpr (LL) {
      slot_off = p.ll_off + (epoch & 1u) * (uint32_t)WORLD * p.ll_slot_bytes;
      if (active[0]) {
        for (uint32_t r = row0; r < p.rows; r += row_step) {
          const size_t c = (size_t)r * p.chunks_per_row + ch[0];
          const uint4 v = p.in[c];
          const uint4 pk0 = make_uint4(v.x, epoch, v.y, epoch), pk1 = make_uint4(v.z, epoch, v.w, epoch);
#pragma unroll
          for (int q = 0; q < WORLD; ++q) {
            if (q == p.rank) continue;
            uint4* dst = reinterpret_cast<uint4*>(p.region[q] + slot_off + (uint32_t)p.rank * p.ll_slot_bytes) + 2 * c;
            st_volatile_v4(dst, pk0);
            st_volatile_v4(dst + 1, pk1);
          }
        }
      }
    } else {
    // 1. stage this thread's chunk of every row of this cluster into the own slot
    if (active[0]) {
      uint4* my_slot = reinterpret_cast<uint4*>(mine + slot_off);
      for (uint32_t r = row0; r < p.rows; r += row_step) {
        const size_t c = (size_t)r * p.chunks_per_row + ch[0];
        my_slot[c] = p.in[c];
      }
    }
    __threadfence_system();
    __syncthreads();
    // 2. cross-GPU barrier of block b (peers' block b staged exactly the chunks read below)
    if (p.barrier) {
      if (threadIdx.x < WORLD) {
        const int q = threadIdx.x;
        uint32_t* peer_flag =
            reinterpret_cast<uint32_t*>(p.region[q]) + blockIdx.x * kArMaxRanks + p.rank;
        st_release_sys(peer_flag, epoch);
        const long long t0 = clock64();
        while ((int32_t)(ld_acquire_sys(my_flags + q) - epoch) < 0) {     // a peer may already be a call ahead
          if (peer_wait_expired(t0)) {
            printf("b200_w4a16: all-reduce+norm barrier timeout rank %d block %d peer %d\n",
                   p.rank, blockIdx.x, q);
            __trap();
          }
        }
      }
      __syncthreads();
    }
    }  // !LL
    // every block's epoch advances on every call (block 0 covers the idle ones): the slot
    // parity is then a property of the call, whatever grid the previous calls used.
    if (threadIdx.x == 0) epochs[blockIdx.x] = epoch;
    if (blockIdx.x == 0 && threadIdx.x >= gridDim.x && threadIdx.x < kArMaxBlocks)
      epochs[threadIdx.x] = epoch;
  }

  uint4 wv[CPT];
#pragma unroll
  for (int j = 0; j < CPT; ++j) wv[j] = active[j] ? p.weight[ch[j]] : make_uint4(0, 0, 0, 0);

  // Raw operands of one row: the next row's are requested before the current row's statistics
  // meet at the barrier (twice the bytes in flight per thread).
  constexpr int NV = WORLD > 0 ? WORLD : 1;
  uint4 cur[CPT][NV], nxt[CPT][NV];
  uint4 cur_res[CPT], nxt_res[CPT];
#pragma unroll
  for (int j = 0; j < CPT; ++j) cur_res[j] = nxt_res[j] = make_uint4(0, 0, 0, 0);
  auto fetch = [&](uint32_t r, uint4 (&v)[CPT][NV], uint4 (&res)[CPT]) {
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
      if (!active[j]) continue;
      const size_t c = (size_t)r * p.chunks_per_row + ch[j];
      if constexpr (WORLD > 0 && LL) {
        const uint32_t epoch = s_epoch;
#pragma unroll
        for (int q = 0; q < WORLD; ++q) {
          if (q == p.rank) { v[j][q] = p.in[c]; continue; }
          const uint4* src = reinterpret_cast<const uint4*>(p.region[p.rank] + slot_off + (uint32_t)q * p.ll_slot_bytes) + 2 * c;
          uint4 a, b;
          const long long t0 = clock64();
          for (;;) {
            a = ld_volatile_v4(src);
            b = ld_volatile_v4(src + 1);
            if (a.y == epoch && a.w == epoch && b.y == epoch && b.w == epoch) break;
            if (!p.barrier || peer_wait_expired(t0)) {
              printf("b200_w4a16: all-reduce+norm packet timeout rank %d block %d source %d\n", p.rank,
                     blockIdx.x, q);
              __trap();
            }
          }
          v[j][q] = make_uint4(a.x, a.z, b.x, b.z);
        }
      } else if constexpr (WORLD > 0) {
#pragma unroll
        for (int q = 0; q < WORLD; ++q)
          v[j][q] = ld_volatile_v4(reinterpret_cast<const uint4*>(p.region[q] + slot_off) + c);
      } else {
        v[j][0] = p.in[c];
      }
      if (p.residual_in) res[j] = p.residual_in[c];
    }
  };
  if (row0 < p.rows) fetch(row0, cur, cur_res);

  uint32_t it = 0;
  for (uint32_t r = row0; r < p.rows; r += row_step, ++it) {
    if (r + row_step < p.rows) fetch(r + row_step, nxt, nxt_res);
    uint4 xr[CPT];                          // half(x): the residual stream
    float ssq[CPT];
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
      xr[j] = make_uint4(0, 0, 0, 0);
      ssq[j] = 0.f;
      if (!active[j]) continue;
      float x[8];
      uint4 xin;
      if constexpr (WORLD > 0) {
        float acc[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = 0.f;
#pragma unroll
        for (int q = 0; q < WORLD; ++q) {
          float f[8];
          unpack8(f, cur[j][q]);
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[i] += f[i];
        }
        __half2* xh = reinterpret_cast<__half2*>(&xin);
#pragma unroll
        for (int i = 0; i < 4; ++i) xh[i] = __floats2half2_rn(acc[2 * i], acc[2 * i + 1]);
      } else {
        xin = cur[j][0];
      }
      unpack8(x, xin);
      if (p.residual_in) {
        float res[8];
        unpack8(res, cur_res[j]);
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] += res[i];
        __half2* h = reinterpret_cast<__half2*>(&xr[j]);
#pragma unroll
        for (int i = 0; i < 4; ++i) h[i] = __floats2half2_rn(x[2 * i], x[2 * i + 1]);
      } else {
        xr[j] = xin;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) ssq[j] = fmaf(x[i], x[i], ssq[j]);
      if (p.residual_out) p.residual_out[(size_t)r * p.chunks_per_row + ch[j]] = xr[j];
    }
    // ONE summation order for every shape of this kernel (so the fused and the standalone
    // variants agree bit for bit whatever CPT / cluster size they run with): the row is cut
    // into groups of 256 chunks (group = crank * CPT + j); inside a group the 8 values of a
    // thread, then the lane tree, then the 8 warps in order; then the groups in order.
    float* sw = s_warp + (it & 1) * (CPT * (kNormThreads / 32));
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) ssq[j] += __shfl_xor_sync(0xffffffffu, ssq[j], o);
      if ((threadIdx.x & 31) == 0) sw[j * (kNormThreads / 32) + (threadIdx.x >> 5)] = ssq[j];
    }
    __syncthreads();
    float total = 0.f;
    if (csize == 1) {                       // the whole row lives in this CTA
#pragma unroll
      for (int j = 0; j < CPT; ++j) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < kNormThreads / 32; ++w) t += sw[j * (kNormThreads / 32) + w];
        total += t;
      }
    } else {
      float* sp = s_part + (it & 1) * CPT;
      if (threadIdx.x < CPT) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < kNormThreads / 32; ++w) t += sw[threadIdx.x * (kNormThreads / 32) + w];
        sp[threadIdx.x] = t;
      }
      cluster.sync();
      for (uint32_t k = 0; k < csize; ++k) {
        const float* rp = cluster.map_shared_rank(sp, k);
#pragma unroll
        for (int j = 0; j < CPT; ++j) total += rp[j];
      }
    }
    const float inv = rsqrtf(total * p.inv_hidden + p.eps);
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
      if (!active[j]) continue;
      float xf[8];
      unpack8(xf, xr[j]);
      uint4 o;
      __half2* oh = reinterpret_cast<__half2*>(&o);
      const __half2* wh = reinterpret_cast<const __half2*>(&wv[j]);
#pragma unroll
      for (int i = 0; i < 4; ++i)
        oh[i] = __hmul2(__floats2half2_rn(xf[2 * i] * inv, xf[2 * i + 1] * inv), wh[i]);
      p.out[(size_t)r * p.chunks_per_row + ch[j]] = o;
    }
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
#pragma unroll
      for (int q = 0; q < NV; ++q) cur[j][q] = nxt[j][q];
      cur_res[j] = nxt_res[j];
    }
  }
  if (csize > 1) cluster.sync();   // nobody leaves while a peer CTA may still read its partial
}

template <int WORLD, int CPT = 1, bool LL = false>
int launch_norm(const NormParams& prm, int cluster, int row_clusters, cudaStream_t stream) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(cluster * row_clusters));
  cfg.blockDim = dim3(kNormThreads);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)cluster;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return (int)cudaLaunchKernelEx(&cfg, add_rmsnorm_kernel<WORLD, CPT, LL>, prm);
}

int fill_params(NormParams* prm, const uint16_t* in, const uint16_t* residual_in,
                const uint16_t* weight, uint16_t* out, uint16_t* residual_out, float eps,
                size_t rows, int hidden) {
  for (int q = 0; q < kArMaxRanks; ++q) prm->region[q] = nullptr;
  prm->in = reinterpret_cast<const uint4*>(in);
  prm->residual_in = reinterpret_cast<const uint4*>(residual_in);
  prm->weight = reinterpret_cast<const uint4*>(weight);
  prm->out = reinterpret_cast<uint4*>(out);
  prm->residual_out = reinterpret_cast<uint4*>(residual_out);
  prm->eps = eps;
  prm->inv_hidden = 1.0f / (float)hidden;
  prm->rows = (uint32_t)rows;
  prm->chunks_per_row = (uint32_t)(hidden / 8);
  prm->slot_bytes = 0;
  prm->ll_off = 0;
  prm->ll_slot_bytes = 0;
  prm->rank = 0;
  prm->barrier = 0;
  return (int)((prm->chunks_per_row + kNormThreads - 1) / kNormThreads);   // cluster size
}

}  // namespace

namespace {

// out[r, n..n+7] = silu(gate) * up, 16 bytes out per thread.  Plain: gate = in[r, n], up =
// in[r, d + n] (csrc/activation_kernels.cu:17-28).  Paired: gate = in[r, 2 n'..], the 16-byte
// chunk 2 j of the row, up = chunk 2 j + 1 -- the order the gate_up GEMM produces when its
// weight was channel-paired for the fused epilogue (used for batches above the decode kernel).
__global__ void __launch_bounds__(256)
silu_and_mul_kernel(const uint4* __restrict__ in, uint4* __restrict__ out, size_t total_chunks,
                    uint32_t chunks_per_row, int paired, const int32_t* __restrict__ rows_dev) {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (rows_dev) {                                             // live rows known on the device only
    const size_t live = (size_t)max(*rows_dev, 0) * chunks_per_row;
    if (live < total_chunks) total_chunks = live;
  }
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total_chunks;
       i += (size_t)gridDim.x * blockDim.x) {
    const size_t r = i / chunks_per_row;
    const uint32_t j = (uint32_t)(i - r * chunks_per_row);
    const uint4* row = in + r * 2 * chunks_per_row;
    const uint4 gv = paired ? row[2 * j] : row[j];
    const uint4 uv = paired ? row[2 * j + 1] : row[chunks_per_row + j];
    float gf[8], uf[8];
    unpack8(gf, gv);
    unpack8(uf, uv);
    uint4 o;
    uint16_t* oh = reinterpret_cast<uint16_t*>(&o);
#pragma unroll
    for (int k = 0; k < 8; ++k) oh[k] = silu_mul_f16(gf[k], uf[k]);
    out[i] = o;
  }
}

}  // namespace

int launch_silu_and_mul(const uint16_t* in, uint16_t* out, size_t rows, int d, int paired,
                        cudaStream_t stream, const int32_t* rows_dev) {
  const size_t total = rows * (size_t)(d / 8);
  size_t blocks = (total + 255) / 256;
  if (blocks > (size_t)kMaxCtas * 8) blocks = (size_t)kMaxCtas * 8;
  silu_and_mul_kernel<<<(unsigned)blocks, 256, 0, stream>>>(
      reinterpret_cast<const uint4*>(in), reinterpret_cast<uint4*>(out), total, (uint32_t)(d / 8), paired, rows_dev);
  return (int)cudaGetLastError();
}

int norm_max_hidden() { return kNormMaxCluster * kNormThreads * 8; }

int launch_add_rmsnorm(const uint16_t* in, const uint16_t* residual_in, const uint16_t* weight,
                       uint16_t* out, uint16_t* residual_out, float eps, size_t rows, int hidden,
                       cudaStream_t stream) {
  NormParams prm;
  int cluster = fill_params(&prm, in, residual_in, weight, out, residual_out, eps, rows, hidden);
  static const size_t cpt4_min_rows = [] {
    const char* e = getenv("B200_NORM_CPT4_MIN_ROWS");      // tuning override
    return e ? (size_t)atol(e) : (size_t)17;
  }();
  if (rows >= cpt4_min_rows) {
    // above decode batches: whole-row CTAs, four chunks per thread, (almost) no cluster
    // barrier -- faster from 32 rows up (2.4 vs 2.8 us) and 0.91-0.97 of the copy bandwidth
    // at prefill sizes (profiles/r1e_norm_bench.jsonl)
    cluster = (cluster + 3) / 4;
    size_t row_clusters = (size_t)(kMaxCtas * 4 / cluster);
    if (row_clusters > rows) row_clusters = rows;
    return launch_norm<0, 4>(prm, cluster, (int)row_clusters, stream);
  }
  // few rows (decode): one chunk per thread, a row spread over a cluster
  size_t row_clusters = (size_t)(kMaxCtas * 8 / cluster);
  if (row_clusters > rows) row_clusters = rows;
  return launch_norm<0>(prm, cluster, (int)row_clusters, stream);
}

int launch_allreduce_add_rmsnorm(void* const* regions, int rank, int world, const uint16_t* in,
                                 const uint16_t* residual_in, const uint16_t* weight,
                                 uint16_t* out, uint16_t* residual_out, float eps, size_t rows,
                                 int hidden, size_t max_message_bytes, int barrier,
                                 cudaStream_t stream, size_t ll_offset, size_t ll_message_bytes) {
  NormParams prm;
  const int cluster = fill_params(&prm, in, residual_in, weight, out, residual_out, eps, rows, hidden);
  for (int q = 0; q < world; ++q) prm.region[q] = static_cast<uint8_t*>(regions[q]);
  prm.slot_bytes = (uint32_t)((max_message_bytes + 255) / 256 * 256);
  prm.rank = rank;
  prm.barrier = barrier;
  size_t row_clusters = (size_t)(kArMaxBlocks / cluster);      // one barrier flag set per block
  if (row_clusters > rows) row_clusters = rows;
  if (ll_message_bytes > 0) {
    prm.ll_off = (uint32_t)ll_offset;
    prm.ll_slot_bytes = (uint32_t)(2 * ll_message_bytes);
    switch (world) {
      case 2: return launch_norm<2, 1, true>(prm, cluster, (int)row_clusters, stream);
      case 3: return launch_norm<3, 1, true>(prm, cluster, (int)row_clusters, stream);
      case 4: return launch_norm<4, 1, true>(prm, cluster, (int)row_clusters, stream);
      case 5: return launch_norm<5, 1, true>(prm, cluster, (int)row_clusters, stream);
      case 6: return launch_norm<6, 1, true>(prm, cluster, (int)row_clusters, stream);
      case 7: return launch_norm<7, 1, true>(prm, cluster, (int)row_clusters, stream);
      case 8: return launch_norm<8, 1, true>(prm, cluster, (int)row_clusters, stream);
      default: return (int)cudaErrorInvalidValue;
    }
  }
  switch (world) {
    case 1: return launch_norm<1>(prm, cluster, (int)row_clusters, stream);
    case 2: return launch_norm<2>(prm, cluster, (int)row_clusters, stream);
    case 3: return launch_norm<3>(prm, cluster, (int)row_clusters, stream);
    case 4: return launch_norm<4>(prm, cluster, (int)row_clusters, stream);
    case 5: return launch_norm<5>(prm, cluster, (int)row_clusters, stream);
    case 6: return launch_norm<6>(prm, cluster, (int)row_clusters, stream);
    case 7: return launch_norm<7>(prm, cluster, (int)row_clusters, stream);
    case 8: return launch_norm<8>(prm, cluster, (int)row_clusters, stream);
    default: return (int)cudaErrorInvalidValue;
  }
}

}  // namespace b200
