// w4a16_moe.cu -- token gather / weighted scatter-add around the routed expert GEMMs of a
// mixture-of-experts block.
//
// Replaces the dense "every local expert on every token, then multiply by a mostly-zero
// routing weight" loop of MixtralMoE.forward (vllm/model_executor/models/mixtral_quant.py:
// 143-157): tokens routed to an expert are gathered into a dense [M_e, hidden] batch, the
// expert's three W4A16 GEMMs run on M_e rows instead of T, and the result goes back as
//     acc[idx[i], :] = half(float(acc[idx[i], :]) + float(half(float(y[i, :]) * w[i])))
// -- the reference's `current.mul_(expert_weights)` then `final.add_(current)` arithmetic
// (fp32 product rounded to fp16, fp16 sum rounded once), expert after expert in index order.
// A token selects an expert at most once, so rows never collide inside one call: no atomics,
// deterministic.  HBM-bound byte movement: 16-byte chunks, one row per block-row of threads.
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "w4a16_internal.h"

namespace b200 {

namespace {

__global__ void __launch_bounds__(256)
moe_gather_kernel(const uint4* __restrict__ x, const int32_t* __restrict__ idx,
                  uint4* __restrict__ out, size_t total_chunks, uint32_t chunks_per_row,
                  const int32_t* __restrict__ rows_dev) {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (rows_dev) {                                             // live rows known on the device only
    const size_t live = (size_t)max(*rows_dev, 0) * chunks_per_row;
    if (live < total_chunks) total_chunks = live;
  }
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total_chunks;
       i += (size_t)gridDim.x * blockDim.x) {
    const size_t r = i / chunks_per_row;
    const uint32_t c = (uint32_t)(i - r * chunks_per_row);
    out[i] = x[(size_t)idx[r] * chunks_per_row + c];
  }
}

__global__ void __launch_bounds__(256)
moe_scatter_add_kernel(const uint4* __restrict__ y, const int32_t* __restrict__ idx,
                       const float* __restrict__ w, uint4* __restrict__ acc, size_t total_chunks,
                       uint32_t chunks_per_row, const int32_t* __restrict__ rows_dev) {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (rows_dev) {
    const size_t live = (size_t)max(*rows_dev, 0) * chunks_per_row;
    if (live < total_chunks) total_chunks = live;
  }
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total_chunks;
       i += (size_t)gridDim.x * blockDim.x) {
    const size_t r = i / chunks_per_row;
    const uint32_t c = (uint32_t)(i - r * chunks_per_row);
    const float wr = w[r];
    uint4* dst = acc + (size_t)idx[r] * chunks_per_row + c;
    const uint4 yv = y[i];
    uint4 av = *dst;
    const __half* yh = reinterpret_cast<const __half*>(&yv);
    __half* ah = reinterpret_cast<__half*>(&av);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const __half scaled = __float2half_rn(__half2float(yh[k]) * wr);
      ah[k] = __float2half_rn(__half2float(ah[k]) + __half2float(scaled));
    }
    *dst = av;
  }
}

// The grouped expert launch (w4a16_decode.cu, GROUPED) leaves y[e] = expert e's output for every
// local expert with tokens; this is the loop over experts of the reference (:152-157) as one
// pass: out[t] = (((0 + half(y[0][t] w[t][0])) + half(y[1][t] w[t][1])) + ...), experts without
// tokens skipped (their y is unwritten; a zero weight times a finite y would add +0).
__global__ void __launch_bounds__(256)
moe_combine_kernel(const uint4* __restrict__ y, const float* __restrict__ w,
                   const int32_t* __restrict__ counts, uint4* __restrict__ out, int n_experts,
                   int w_stride, size_t total_chunks, uint32_t chunks_per_row) {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total_chunks;
       i += (size_t)gridDim.x * blockDim.x) {
    const size_t r = i / chunks_per_row;
    uint4 av = make_uint4(0u, 0u, 0u, 0u);
    __half* ah = reinterpret_cast<__half*>(&av);
    for (int e = 0; e < n_experts; ++e) {
      if (counts[e] <= 0) continue;
      const float wr = w[r * (size_t)w_stride + e];
      const uint4 yv = y[(size_t)e * total_chunks + i];
      const __half* yh = reinterpret_cast<const __half*>(&yv);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const __half scaled = __float2half_rn(__half2float(yh[k]) * wr);
        ah[k] = __float2half_rn(__half2float(ah[k]) + __half2float(scaled));
      }
    }
    out[i] = av;
  }
}

// The router of a mixture-of-experts block as one kernel (mixtral_quant.py:133-142: gate GEMM ->
// softmax in fp32 -> top-k -> renormalise): a CTA per token.  The logits are rounded to fp16
// as the reference's fp16 gate GEMM leaves them, the softmax runs over those in fp32.
// Writes the top-k weights / expert ids (descending, as torch.topk), the dense [T, E] weight
// matrix (zero where an expert was not chosen) and bumps counts[e] (zeroed by the launcher).
constexpr int kRouteWarps = 8;
template <int EMAX>
__global__ void __launch_bounds__(kRouteWarps * 32)
moe_route_kernel(const uint4* __restrict__ x, const uint4* __restrict__ gate_w, float* __restrict__ topk_w,
                 long long* __restrict__ topk_idx, float* __restrict__ dense_w, int32_t* __restrict__ counts,
                 int T, int E, int top_k, int chunks_per_row) {
  __shared__ float part[kRouteWarps][EMAX];
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  (void)T;
  const int t = blockIdx.x;                                  // a CTA per token: the row is read once,
  float acc[EMAX];                                           // every thread's loads are in flight together
#pragma unroll
  for (int e = 0; e < EMAX; ++e) acc[e] = 0.f;
  for (int c = threadIdx.x; c < chunks_per_row; c += kRouteWarps * 32) {
    const uint4 xv = x[(size_t)t * chunks_per_row + c];
    const __half2* xh = reinterpret_cast<const __half2*>(&xv);
#pragma unroll
    for (int e = 0; e < EMAX; ++e) {
      if (e < E) {
        const uint4 gv = gate_w[(size_t)e * chunks_per_row + c];
        const __half2* gh = reinterpret_cast<const __half2*>(&gv);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float2 a = __half22float2(xh[k]), b = __half22float2(gh[k]);
          acc[e] = fmaf(a.x, b.x, acc[e]);
          acc[e] = fmaf(a.y, b.y, acc[e]);
        }
      }
    }
  }
#pragma unroll
  for (int e = 0; e < EMAX; ++e) {
    float v = acc[e];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) part[warp][e] = v;
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  float logit[EMAX];
  float mx = -INFINITY;
#pragma unroll
  for (int e = 0; e < EMAX; ++e) {
    float v = 0.f;
#pragma unroll
    for (int w = 0; w < kRouteWarps; ++w) v += part[w][e];   // fixed order: deterministic
    logit[e] = e < E ? __half2float(__float2half_rn(v)) : -INFINITY;
    mx = fmaxf(mx, logit[e]);
  }
  float prob[EMAX], sum = 0.f;
#pragma unroll
  for (int e = 0; e < EMAX; ++e) {
    prob[e] = e < E ? expf(logit[e] - mx) : 0.f;
    sum += prob[e];
  }
#pragma unroll
  for (int e = 0; e < EMAX; ++e) prob[e] = prob[e] / sum;
  // selection twice: the first pass finds the renormalisation sum, the second writes the
  // renormalised weights (nothing is read back from global memory)
  uint32_t taken = 0;
  float ksum = 0.f;
#pragma unroll 1
  for (int pass = 0; pass < 2; ++pass) {
    taken = 0;
    for (int j = 0; j < top_k; ++j) {
      int best = -1;
      float bv = -1.f;
#pragma unroll
      for (int e = 0; e < EMAX; ++e)
        if (e < E && !((taken >> e) & 1u) && (best < 0 || prob[e] > bv)) { bv = prob[e]; best = e; }   // NaN-safe
      taken |= 1u << best;
      if (pass == 0) {
        ksum += bv;
        topk_idx[(size_t)t * top_k + j] = best;
        atomicAdd(counts + best, 1);
      } else {
        topk_w[(size_t)t * top_k + j] = bv / ksum;
      }
    }
  }
#pragma unroll
  for (int e = 0; e < EMAX; ++e)
    if (e < E) dense_w[(size_t)t * E + e] = ((taken >> e) & 1u) ? prob[e] / ksum : 0.f;
}

// Per local expert, the tokens that chose it (ascending) and their routing weights, from the
// router's top-k lists: CTA e scans the (token, choice) pairs in order, a block-wide prefix sum
// gives every hit its row.  tok / w: [n_local][tokens].  A token names an expert at most once.
constexpr int kCompactThreads = 256;
__global__ void __launch_bounds__(kCompactThreads)
moe_compact_kernel(const long long* __restrict__ topk_idx, const float* __restrict__ topk_w,
                   int32_t* __restrict__ tok, float* __restrict__ w, int tokens, int top_k, int first) {
  __shared__ int warp_sum[kCompactThreads / 32];
  __shared__ int base_s;
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const int e = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int total = tokens * top_k;
  if (threadIdx.x == 0) base_s = 0;
  __syncthreads();
  for (int i0 = 0; i0 < total; i0 += kCompactThreads) {
    const int i = i0 + threadIdx.x;
    const bool hit = i < total && topk_idx[i] == (long long)(first + e);
    const unsigned m = __ballot_sync(0xffffffffu, hit);
    if (lane == 0) warp_sum[warp] = __popc(m);
    __syncthreads();
    int before = base_s;
    for (int q = 0; q < warp; ++q) before += warp_sum[q];
    if (hit) {
      const int pos = before + __popc(m & ((1u << lane) - 1u));
      tok[(size_t)e * tokens + pos] = i / top_k;
      w[(size_t)e * tokens + pos] = topk_w[i];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      int t = 0;
      for (int q = 0; q < kCompactThreads / 32; ++q) t += warp_sum[q];
      base_s += t;
    }
    __syncthreads();
  }
}

// The small kernels start with griddepcontrol.wait: launched programmatically their launch
// latency hides behind the tail of the kernel before them.
template <typename... KArgs, typename... Args>
int launch_pdl(void (*kernel)(KArgs...), unsigned grid, unsigned block, cudaStream_t stream, Args... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(block);
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return (int)cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

unsigned grid_for(size_t total) {
  size_t blocks = (total + 255) / 256;
  if (blocks > (size_t)kMaxCtas * 8) blocks = (size_t)kMaxCtas * 8;
  return (unsigned)(blocks ? blocks : 1);
}

}  // namespace

int launch_moe_gather(const uint16_t* x, const int32_t* idx, uint16_t* out, size_t rows, int hidden,
                      cudaStream_t stream, const int32_t* rows_dev) {
  const size_t total = rows * (size_t)(hidden / 8);
  return launch_pdl(moe_gather_kernel, grid_for(total), 256, stream, reinterpret_cast<const uint4*>(x), idx,
                    reinterpret_cast<uint4*>(out), total, (uint32_t)(hidden / 8), rows_dev);
}

int launch_moe_scatter_add(const uint16_t* y, const int32_t* idx, const float* w, uint16_t* acc,
                           size_t rows, int hidden, cudaStream_t stream, const int32_t* rows_dev) {
  const size_t total = rows * (size_t)(hidden / 8);
  return launch_pdl(moe_scatter_add_kernel, grid_for(total), 256, stream, reinterpret_cast<const uint4*>(y),
                    idx, w, reinterpret_cast<uint4*>(acc), total, (uint32_t)(hidden / 8), rows_dev);
}

int launch_moe_compact(const long long* topk_idx, const float* topk_w, int32_t* tok, float* w,
                       size_t tokens, int top_k, int first, int n_local, cudaStream_t stream) {
  return launch_pdl(moe_compact_kernel, (unsigned)n_local, kCompactThreads, stream, topk_idx, topk_w, tok, w,
                    (int)tokens, top_k, first);
}

int launch_moe_route(const uint16_t* x, const uint16_t* gate_w, float* topk_w, long long* topk_idx,
                     float* dense_w, int32_t* counts, size_t tokens, int hidden, int n_experts,
                     int top_k, cudaStream_t stream) {
  cudaError_t e = cudaMemsetAsync(counts, 0, (size_t)n_experts * sizeof(int32_t), stream);
  if (e != cudaSuccess) return (int)e;
  const unsigned grid = (unsigned)tokens;
  const uint4* xv = reinterpret_cast<const uint4*>(x);
  const uint4* gv = reinterpret_cast<const uint4*>(gate_w);
  if (n_experts <= 8)
    return launch_pdl(moe_route_kernel<8>, grid, kRouteWarps * 32, stream, xv, gv, topk_w, topk_idx, dense_w,
                      counts, (int)tokens, n_experts, top_k, hidden / 8);
  return launch_pdl(moe_route_kernel<kMaxRouteExperts>, grid, kRouteWarps * 32, stream, xv, gv, topk_w, topk_idx,
                    dense_w, counts, (int)tokens, n_experts, top_k, hidden / 8);
}

int launch_moe_combine(const uint16_t* y, const float* w, const int32_t* counts, uint16_t* out,
                       int n_experts, int w_stride, size_t rows, int hidden, cudaStream_t stream) {
  const size_t total = rows * (size_t)(hidden / 8);
  return launch_pdl(moe_combine_kernel, grid_for(total), 256, stream, reinterpret_cast<const uint4*>(y), w,
                    counts, reinterpret_cast<uint4*>(out), n_experts, w_stride, total, (uint32_t)(hidden / 8));
}

}  // namespace b200
