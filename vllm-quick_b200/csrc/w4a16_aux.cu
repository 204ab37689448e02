// w4a16_aux.cu -- the HBM-bound helpers of the W4A16 path: standalone dequantize and the
// AWQ <-> B200 layout conversion.  Integer/byte work: no tensor cores, coalesced 16-byte
// accesses, grid sized from the problem.
//
// Reference behaviour restated here:
//   awq_dequantize  csrc/quantization/awq/gemm_kernels.cu:274-381 (kernel + host)
//   nibble order    csrc/quantization/awq/dequantize.cuh:17-84
#include "w4a16_common.cuh"
#include "w4a16_internal.h"

namespace b200 {

// ------------------------------------------------------------------------------------------
// Dequantize, AWQ (checkpoint) layout.  One thread = one packed column (8 channels) x 4 rows:
// 4 coalesced 4-byte loads, 4 x 16-byte stores, zeros/scales fetched once per 4 rows.
// Algorithmic bytes per word of qweight: 4 read + 16 written (+3.9 % for zeros/scales).
__global__ void __launch_bounds__(256)
dequant_awq_kernel(const uint32_t* __restrict__ qw, const uint4* __restrict__ scales,
                   const uint32_t* __restrict__ qz, uint4* __restrict__ W, int K, int NW, int G) {
  const int nw = blockIdx.x * blockDim.x + threadIdx.x;
  const int k0 = (blockIdx.y * blockDim.y + threadIdx.y) * 4;
  if (nw >= NW || k0 >= K) return;
  const int g = k0 / G;  // G % 4 == 0: the four rows share one group
  ZS8 c;
  make_zs8(__ldg(qz + (size_t)g * NW + nw), __ldg(scales + (size_t)g * NW + nw), c);
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) w[i] = (k0 + i < K) ? __ldg(qw + (size_t)(k0 + i) * NW + nw) : 0u;
#pragma unroll
  for (int i = 0; i < 4; ++i)
    if (k0 + i < K) W[(size_t)(k0 + i) * NW + nw] = dequant_word_zs8(w[i], c);
}

// ------------------------------------------------------------------------------------------
// Dequantize, B200 layout.  One thread = one channel x 32 rows (one 16-byte load); a warp
// covers 32 consecutive channels so each of its 2-byte stores lands in one 64-byte segment.
__global__ void __launch_bounds__(128)
dequant_b200_kernel(const uint4* __restrict__ qw, const uint16_t* __restrict__ scales,
                    const uint32_t* __restrict__ qz, uint16_t* __restrict__ W, int K, int N,
                    int G, int KB) {
  const int c = threadIdx.x;                  // channel inside the tile
  const int j = blockIdx.x & 3;               // 32-row slice of the tile
  const int kb = (blockIdx.x >> 2) % KB;
  const int nt = (blockIdx.x >> 2) / KB;
  const int n = nt * kTileN + c;
  if (n >= N) return;
  const int k0 = kb * kTileK + j * 32;
  const int g = k0 / G;                       // G % 32 == 0: 32 rows share one group
  const uint32_t z = (__ldg(qz + (size_t)g * (N >> 3) + (n >> 3)) >> (4 * awq_nibble(n & 7))) & 0xFu;
  uint32_t zlo, zhi, s2;
  make_zs1(z, __ldg(scales + (size_t)g * N + n), zlo, zhi, s2);
  const uint4 v = __ldg(qw + ((size_t)(nt * KB + kb) * kTileWords >> 2) + j * 128 + c);
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    uint32_t o[4];
    dequant_word_zs1(w[i], zlo, zhi, s2, o[0], o[1], o[2], o[3]);
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      const size_t k = k0 + i * 8 + p * 2;
      W[k * N + n] = (uint16_t)(o[p] & 0xFFFFu);
      W[(k + 1) * N + n] = (uint16_t)(o[p] >> 16);
    }
  }
}

// ------------------------------------------------------------------------------------------
// AWQ -> B200 layout.  One CTA = one 128x128 tile staged through shared memory: coalesced
// 64-byte row reads in, fully coalesced 8 KiB tile out.  Channels >= N are written as zero.
__global__ void __launch_bounds__(256)
repack_awq_to_b200_kernel(const uint32_t* __restrict__ src, uint32_t* __restrict__ dst, int K,
                          int N, int KB) {
  __shared__ uint32_t tile[kTileK][17];
  const int kb = blockIdx.x % KB, nt = blockIdx.x / KB;
  const int NW = N >> 3;
  for (int i = threadIdx.x; i < kTileK * 16; i += blockDim.x) {
    const int r = i >> 4, cw = i & 15;
    const int nw = nt * 16 + cw;
    tile[r][cw] = (nw < NW) ? __ldg(src + (size_t)(kb * kTileK + r) * NW + nw) : 0u;
  }
  __syncthreads();
  uint32_t* out = dst + (size_t)(nt * KB + kb) * kTileWords;
  for (int idx = threadIdx.x; idx < kTileWords; idx += blockDim.x) {
    const int w = idx & 3, c = (idx >> 2) & 127, j = idx >> 9;
    const int r0 = (j * 4 + w) * 8;
    const int sh = 4 * awq_nibble(c & 7);
    uint32_t word = 0;
#pragma unroll
    for (int e = 0; e < 8; ++e)
      word |= ((tile[r0 + e][c >> 3] >> sh) & 0xFu) << (4 * awq_nibble(e));
    out[idx] = word;
  }
  __threadfence();  // a dependent GEMM may prefetch these words as soon as this CTA retires
}

// B200 -> AWQ layout (inverse permutation; used by tests and checkpoint export).
__global__ void __launch_bounds__(256)
repack_b200_to_awq_kernel(const uint32_t* __restrict__ src, uint32_t* __restrict__ dst, int K,
                          int N, int KB) {
  __shared__ uint32_t tile[kTileWords];
  const int kb = blockIdx.x % KB, nt = blockIdx.x / KB;
  const int NW = N >> 3;
  const uint32_t* in = src + (size_t)(nt * KB + kb) * kTileWords;
  for (int i = threadIdx.x; i < kTileWords; i += blockDim.x) tile[i] = __ldg(in + i);
  __syncthreads();
  for (int i = threadIdx.x; i < kTileK * 16; i += blockDim.x) {
    const int r = i >> 4, cw = i & 15;
    const int nw = nt * 16 + cw;
    if (nw >= NW) continue;
    const int o = r >> 3, e = r & 7;
    const int sh = 4 * awq_nibble(e);
    uint32_t word = 0;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int c = cw * 8 + q;
      word |= ((tile[(o >> 2) * 512 + c * 4 + (o & 3)] >> sh) & 0xFu) << (4 * awq_nibble(q));
    }
    dst[(size_t)(kb * kTileK + r) * NW + nw] = word;
  }
  __threadfence();
}

// ------------------------------------------------------------------------------------------

int launch_dequant(const int32_t* qweight, const uint16_t* scales, const int32_t* qzeros,
                   uint16_t* W, int K, int N, int G, int layout, cudaStream_t stream) {
  if (K == 0 || N == 0) return cudaSuccess;
  if (layout == 0) {
    const int NW = N / 8;
    dim3 block(64, 4);
    dim3 grid((NW + 63) / 64, (K / 4 + 3) / 4);
    dequant_awq_kernel<<<grid, block, 0, stream>>>(
        reinterpret_cast<const uint32_t*>(qweight), reinterpret_cast<const uint4*>(scales),
        reinterpret_cast<const uint32_t*>(qzeros), reinterpret_cast<uint4*>(W), K, NW, G);
  } else {
    const int KB = K / kTileK, NT = (N + kTileN - 1) / kTileN;
    dequant_b200_kernel<<<NT * KB * 4, 128, 0, stream>>>(
        reinterpret_cast<const uint4*>(qweight), scales,
        reinterpret_cast<const uint32_t*>(qzeros), W, K, N, G, KB);
  }
  return (int)cudaGetLastError();
}

int launch_repack(const int32_t* src, int32_t* dst, int K, int N, int direction,
                  cudaStream_t stream) {
  if (K == 0 || N == 0) return cudaSuccess;
  const int KB = K / kTileK, NT = (N + kTileN - 1) / kTileN;
  if (direction == 0)
    repack_awq_to_b200_kernel<<<NT * KB, 256, 0, stream>>>(
        reinterpret_cast<const uint32_t*>(src), reinterpret_cast<uint32_t*>(dst), K, N, KB);
  else
    repack_b200_to_awq_kernel<<<NT * KB, 256, 0, stream>>>(
        reinterpret_cast<const uint32_t*>(src), reinterpret_cast<uint32_t*>(dst), K, N, KB);
  return (int)cudaGetLastError();
}

}  // namespace b200
