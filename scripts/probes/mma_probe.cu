// mma_probe.cu -- tuning probe (not product code): cycles per tcgen05.mma kind::f16 for
// M=128, K=16 and N in {16,32,64,128,256}, A from TMEM (ts) or shared memory (ss), with and
// without concurrent tcgen05.st traffic from four other warps.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../../vllm-quick_b200/csrc mma_probe.cu -o mma_probe
#include <cstdio>
#include <cstdlib>
#include "w4a16_common.cuh"
using namespace b200;

template <int N, int MODE>  // MODE 0: ts, 1: ss
__global__ void __launch_bounds__(192, 1) probe(int tiles, int st_on, int commit_every, int nacc, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar = base, slot = base + 64;
  const uint32_t bsm = base + 1024;            // B tile: N x 128 K fp16, two swizzle atoms
  const uint32_t asm_ = bsm + N * 256;         // A tile (ss mode): 128 x 128 K fp16
  volatile uint32_t* slot_gen = reinterpret_cast<volatile uint32_t*>(smem_raw + (slot - smem_u32(smem_raw)));
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (N * 256 + 128 * 256) / 4; i += blockDim.x)
    reinterpret_cast<uint32_t*>(smem_raw + (bsm - smem_u32(smem_raw)))[i] = 0;
  if (threadIdx.x == 0) { mbar_init(bar, 1); mbar_init(bar + 8, 1); fence_barrier_init(); }
  if (warp == 1) tmem_alloc<512>(slot);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *slot_gen;
  __shared__ volatile int stop;
  if (threadIdx.x == 0) stop = 0;
  __syncthreads();
  if (warp == 0) {
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_f16_m128(N);
      const long long t0 = clock64();
      int nc = 0, nw = 0;  // commits issued / waited; commit i uses barrier i & 1, parity (i >> 1) & 1
      for (int t = 0; t < tiles; ++t) {
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          const uint64_t bdesc = make_desc_k_sw128(bsm + (kk >> 2) * (N * 128) + (kk & 3) * 32);
          if (MODE == 0) {
            tc_mma_ts(tmem + 256 + (kk % nacc) * N, tmem + (t & 3) * 64 + kk * 8, bdesc, idesc, 1);
          } else {
            const uint64_t adesc = make_desc_k_sw128(asm_ + (kk >> 2) * (128 * 128) + (kk & 3) * 32);
            tc_mma_ss(tmem + 256, adesc, bdesc, idesc, 1);
          }
        }
        if ((t + 1) % commit_every == 0) {
          if (nc - nw == 2) { mbar_wait(bar + 8 * (nw & 1), (nw >> 1) & 1); ++nw; }
          tc_commit(bar + 8 * (nc & 1));
          ++nc;
        }
      }
      if (nc - nw == 2) { mbar_wait(bar + 8 * (nw & 1), (nw >> 1) & 1); ++nw; }
      tc_commit(bar + 8 * (nc & 1));
      ++nc;
      while (nw < nc) { mbar_wait(bar + 8 * (nw & 1), (nw >> 1) & 1); ++nw; }
      const long long t1 = clock64();
      stop = 1;
      if (blockIdx.x == 0) out[0] = t1 - t0;
    }
  } else if (warp >= 2 && st_on) {
    const int q = warp & 3;  // warps 2..5 -> quarters 2,3,0,1
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    uint32_t v[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = threadIdx.x + i;
    int n = 0;
    while (!stop) {
      tmem_st_x16(tmem + lane_base + (n & 15) * 16, v);
      if ((n & 3) == 3) tc_wait_st();
      ++n;
    }
    tc_wait_st();
    if (blockIdx.x == 0 && threadIdx.x == 64) out[1] = n;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc<512>(tmem); }
}

template <int N, int MODE>
void run(int tiles, int st_on, int commit_every, long long* dout, int nacc = 1) {
  const int smem = 2048 + N * 256 + 128 * 256 + 1024;
  cudaFuncSetAttribute(probe<N, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaMemset(dout, 0, 16);
  probe<N, MODE><<<148, 192, smem>>>(tiles, st_on, commit_every, nacc, dout);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[2];
  cudaMemcpy(h, dout, 16, cudaMemcpyDeviceToHost);
  printf("N=%3d %s st=%d nacc=%d commit_every=%d: %.1f cycles/mma  (st x16 issued %lld, %.1f B/cycle)  %s\n", N,
         MODE ? "ss" : "ts", st_on, nacc, commit_every, (double)h[0] / (tiles * 8.0), h[1],
         h[0] ? (double)h[1] * 2048.0 * 4 / h[0] : 0.0, e == cudaSuccess ? "" : cudaGetErrorString(e));
}

int main() {
  long long* dout;
  cudaMalloc(&dout, 16);
  const int tiles = 2000;
  for (int nacc = 1; nacc <= 4; nacc *= 2)
    for (int ce = 1; ce <= 8; ce *= 2) {
      run<16, 0>(tiles, 0, ce, dout, nacc);
      run<32, 0>(tiles, 0, ce, dout, nacc);
      if (nacc <= 2) run<64, 0>(tiles, 0, ce, dout, nacc);
    }
  run<16, 0>(tiles, 1, 4, dout, 2);
  run<16, 0>(tiles, 0, 1000000, dout, 2);
  run<16, 0>(tiles, 0, 1000000, dout, 4);
  run<128, 0>(tiles, 0, 4, dout, 1);
  return 0;
}
