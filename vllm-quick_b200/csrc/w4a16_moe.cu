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
                  uint4* __restrict__ out, size_t total_chunks, uint32_t chunks_per_row) {
  asm volatile("griddepcontrol.wait;" ::: "memory");
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
                       uint32_t chunks_per_row) {
  asm volatile("griddepcontrol.wait;" ::: "memory");
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

unsigned grid_for(size_t total) {
  size_t blocks = (total + 255) / 256;
  if (blocks > (size_t)kMaxCtas * 8) blocks = (size_t)kMaxCtas * 8;
  return (unsigned)(blocks ? blocks : 1);
}

}  // namespace

int launch_moe_gather(const uint16_t* x, const int32_t* idx, uint16_t* out, size_t rows, int hidden,
                      cudaStream_t stream) {
  const size_t total = rows * (size_t)(hidden / 8);
  moe_gather_kernel<<<grid_for(total), 256, 0, stream>>>(
      reinterpret_cast<const uint4*>(x), idx, reinterpret_cast<uint4*>(out), total, (uint32_t)(hidden / 8));
  return (int)cudaGetLastError();
}

int launch_moe_scatter_add(const uint16_t* y, const int32_t* idx, const float* w, uint16_t* acc,
                           size_t rows, int hidden, cudaStream_t stream) {
  const size_t total = rows * (size_t)(hidden / 8);
  moe_scatter_add_kernel<<<grid_for(total), 256, 0, stream>>>(
      reinterpret_cast<const uint4*>(y), idx, w, reinterpret_cast<uint4*>(acc), total, (uint32_t)(hidden / 8));
  return (int)cudaGetLastError();
}

}  // namespace b200
