// hmma_probe.cu -- tuning probe (not product code): legacy mma.sync.m16n8k16 (HMMA.16816.F32)
// issue rate per SM on sm_100a as a function of resident warps, four independent accumulators
// per warp.  Decides whether the decode kernel (w4a16_decode.cu) can afford M/8 HMMAs per
// packed word.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 hmma_probe.cu -o hmma_probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void probe(int iters, long long* out, float* sink) {
  float d[4][4] = {};
  uint32_t a0 = threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 * 11, b1 = a0 * 13;
  __syncthreads();
  const long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int q = 0; q < 4; ++q)
      asm volatile(
          "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
          "{%8,%9}, {%0,%1,%2,%3};"
          : "+f"(d[q][0]), "+f"(d[q][1]), "+f"(d[q][2]), "+f"(d[q][3])
          : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = t1 - t0;
  float s = 0;
  for (int q = 0; q < 4; ++q) for (int e = 0; e < 4; ++e) s += d[q][e];
  if (s == 123.456f) sink[0] = s;
}

int main() {
  long long* dout; float* sink;
  cudaMalloc(&dout, 8); cudaMalloc(&sink, 4);
  const int iters = 4000;
  for (int warps = 1; warps <= 32; warps *= 2) {
    probe<<<148, warps * 32>>>(iters, dout, sink);
    probe<<<148, warps * 32>>>(iters, dout, sink);
    cudaError_t e = cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, dout, 8, cudaMemcpyDeviceToHost);
    const double per_sm = (double)h / (iters * 4.0 * warps);
    printf("warps/SM=%2d: %.2f cycles per HMMA.16816 per SM (%.0f dense fp16 flop/cycle/SM) %s\n", warps,
           per_sm, 4096.0 / per_sm, e == cudaSuccess ? "" : cudaGetErrorString(e));
  }
  return 0;
}
