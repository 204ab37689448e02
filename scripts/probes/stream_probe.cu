// stream_probe.cu -- tuning probe (not product code): how fast can 148 persistent CTAs stream a
// large buffer from HBM into shared memory, (a) with cp.async.bulk (UBLKCP) 8 KiB copies on an
// mbarrier ring, (b) with 16-byte cp.async (LDGSTS) issued by W producer warps, D groups deep.
// Answers whether the 5.3 TB/s ceiling of the decode kernel's weight stream is the bulk-copy
// engine or HBM.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../../vllm-quick_b200/csrc stream_probe.cu -o stream_probe
#include <cstdio>
#include <cstdlib>
#include "w4a16_common.cuh"
using namespace b200;

constexpr int kTile = 8192;

template <int STAGES>
__global__ void __launch_bounds__(64, 1) bulk_kernel(const uint8_t* src, size_t tiles_per_cta, unsigned* sink) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 127u) & ~127u;
  const uint32_t bar_full = base, bar_empty = base + 8 * STAGES, data = base + 1024;
  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(bar_full + 8 * i, 1); mbar_init(bar_empty + 8 * i, 1); }
    fence_barrier_init();
  }
  __syncthreads();
  const uint8_t* mine = src + (size_t)blockIdx.x * tiles_per_cta * kTile;
  if (threadIdx.x == 0) {
    const uint64_t pol = policy_evict_first();
    for (size_t it = 0; it < tiles_per_cta; ++it) {
      const int s = it % STAGES;
      if (it >= STAGES) mbar_wait(bar_empty + 8 * s, ((it / STAGES) - 1) & 1);
      mbar_arrive_expect_tx(bar_full + 8 * s, kTile);
      bulk_g2s_hint(data + s * kTile, mine + it * kTile, kTile, bar_full + 8 * s, pol);
    }
  } else if (threadIdx.x == 32) {
    unsigned acc = 0;
    for (size_t it = 0; it < tiles_per_cta; ++it) {
      const int s = it % STAGES;
      mbar_wait(bar_full + 8 * s, (it / STAGES) & 1);
      acc += *reinterpret_cast<volatile unsigned*>(smem_raw + (data - smem_u32(smem_raw)) + s * kTile);
      mbar_arrive(bar_empty + 8 * s);
    }
    if (acc == 0x12345678u) sink[0] = acc;
  }
}

// W warps; warp w streams its own contiguous slice with 16-byte cp.async, DEPTH groups of
// GROUP_INSTR x 512 B in flight per warp.
template <int WARPS, int DEPTH, int GROUP_INSTR>
__global__ void __launch_bounds__(WARPS * 32, 1) ldgsts_kernel(const uint8_t* src, size_t bytes_per_cta, unsigned* sink) {
  extern __shared__ uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = ((smem_u32(smem_raw) + 127u) & ~127u) + warp * (DEPTH * GROUP_INSTR * 512);
  const size_t per_warp = bytes_per_cta / WARPS;
  const uint8_t* mine = src + (size_t)blockIdx.x * bytes_per_cta + (size_t)warp * per_warp + lane * 16;
  const size_t groups = per_warp / (GROUP_INSTR * 512);
  unsigned acc = 0;
  for (size_t g = 0; g < groups + DEPTH; ++g) {
    if (g < groups) {
      const uint32_t dst = base + (g % DEPTH) * (GROUP_INSTR * 512) + lane * 16;
#pragma unroll
      for (int i = 0; i < GROUP_INSTR; ++i)
        asm volatile("cp.async.cg.shared.global.L2::128B [%0], [%1], 16;" ::"r"(dst + i * 512), "l"(mine + g * (GROUP_INSTR * 512) + i * 512) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    if (g >= DEPTH - 1) {
      asm volatile("cp.async.wait_group %0;" ::"n"(DEPTH - 1) : "memory");
      acc += *reinterpret_cast<volatile unsigned*>(smem_raw + (base - smem_u32(smem_raw)) + ((g + 1) % DEPTH) * (GROUP_INSTR * 512) + lane * 16);
    }
  }
  if (acc == 0x12345678u) sink[0] = acc;
}

template <typename F>
void time_it(const char* name, size_t bytes, F launch) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  launch(); launch();
  cudaEventRecord(e0);
  const int reps = 5;
  for (int i = 0; i < reps; ++i) launch();
  cudaEventRecord(e1);
  cudaError_t e = cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  printf("%-46s %8.1f GB/s  %s\n", name, bytes * reps / (ms * 1e-3) / 1e9, e == cudaSuccess ? "" : cudaGetErrorString(e));
}

int main() {
  const size_t tiles_per_cta = 1024;                       // 8 MiB per CTA, 1.16 GiB total
  const size_t bytes = (size_t)148 * tiles_per_cta * kTile;
  uint8_t* src; unsigned* sink;
  cudaMalloc(&src, bytes); cudaMalloc(&sink, 4);
  cudaMemset(src, 1, bytes);
#define BULK(S) { cudaFuncSetAttribute(bulk_kernel<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2048 + S * kTile); \
    time_it("bulk 8 KiB, " #S " stages", bytes, [&] { bulk_kernel<S><<<148, 64, 2048 + S * kTile>>>(src, tiles_per_cta, sink); }); }
  BULK(4) BULK(8) BULK(16) BULK(24)
#define LDG(W, D, G) { const int sm = 256 + W * D * G * 512; cudaFuncSetAttribute(ldgsts_kernel<W, D, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm); \
    time_it("cp.async 16 B, " #W " warps x " #D " groups x " #G " instr", bytes, [&] { ldgsts_kernel<W, D, G><<<148, W * 32, sm>>>(src, tiles_per_cta * kTile, sink); }); }
  LDG(4, 4, 4) LDG(4, 8, 4) LDG(8, 4, 4) LDG(8, 8, 4) LDG(12, 8, 4) LDG(16, 4, 4) LDG(16, 8, 2) LDG(8, 8, 8)
  return 0;
}
