// w4a16_allreduce.cu -- one-shot fp16 all-reduce over NVLink peer memory for the row-parallel
// linears (o_proj, down_proj) of the W4A16 path.
//
// Replaces, for messages up to a few MiB inside one NVSwitch domain:
//   tensor_model_parallel_all_reduce  vllm/model_executor/parallel_utils/communication_op.py:17-41
//   and is the B200 counterpart of the reference's own (disabled for world_size > 1,
//   vllm/config.py:403-410) one-stage kernel  csrc/custom_all_reduce.cuh:214-232, 144-202.
//
// Every rank owns one REGION (cudaMalloc'd, exported with CUDA IPC, mapped by all peers):
//     [ flags  kArMaxBlocks x kArMaxRanks x u32 ][ per-block epoch counters ][ slot 0 ][ slot 1 ]
// A call on rank r with grid B:
//   1. block b copies its chunk set of `in` into slot (epoch & 1) of r's own region;
//   2. block b publishes: one thread per peer p stores epoch into flags[b][r] of p's region
//      (st.release.sys through NVLink) and spins (ld.acquire.sys) until flags[b][q] of its OWN
//      region carries the epoch for every q -- peers' block b have then written exactly the
//      chunk set this block is about to read;
//   3. block b sums its chunks over all ranks in rank order with fp32 accumulation (every
//      rank computes bit-identical results) and writes `out`.
// Two data slots alternate by epoch, so a rank may run ahead into call e+1 while a peer still
// reads call e; slot reuse at call e+2 is safe because every rank passed barrier e+1 only
// after it finished reading call e.  The epoch lives in device memory (one counter per
// block), so the launch is CUDA-graph capturable and replayable.  Nothing here allocates,
// synchronises or touches NCCL.
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "w4a16_common.cuh"
#include "w4a16_internal.h"

namespace b200 {

namespace {

constexpr int kArThreads = 256;

struct ArParams {
  uint8_t* region[kArMaxRanks];   // peer-mapped base pointers, index = rank
  const uint4* in;
  uint4* out;
  uint32_t nchunks;               // 16-byte chunks
  uint32_t slot_bytes;
  int rank, world;
  int barrier;                    // 0 only in the single-GPU emulation test
  uint32_t ll_off, ll_slot_bytes; // packet variant: area offset, bytes per source rank and parity
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// Peer data changes between calls at the same address: never serve it from L1.
__device__ __forceinline__ uint4 ld_volatile_v4(const uint4* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}

__device__ __forceinline__ void add8(float (&acc)[8], uint4 v) {
  const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = __half22float2(h[i]);
    acc[2 * i] += f.x;
    acc[2 * i + 1] += f.y;
  }
}

template <int WORLD>
__global__ void __launch_bounds__(kArThreads)
allreduce_oneshot_kernel(const ArParams p) {
  uint8_t* mine = p.region[p.rank];
  uint32_t* my_flags = reinterpret_cast<uint32_t*>(mine) + blockIdx.x * kArMaxRanks;
  uint32_t* my_epoch = reinterpret_cast<uint32_t*>(mine + kArFlagBytes) + blockIdx.x;
  __shared__ uint32_t s_epoch;
  // The input is the previous kernel's output (programmatic dependent launch aware).
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (threadIdx.x == 0) s_epoch = *my_epoch + 1;
  __syncthreads();
  const uint32_t epoch = s_epoch;
  const uint32_t slot_off = kArHeaderBytes + (epoch & 1u) * p.slot_bytes;

  // 1. stage this block's chunk set into the own slot
  uint4* my_slot = reinterpret_cast<uint4*>(mine + slot_off);
  for (uint32_t c = blockIdx.x * kArThreads + threadIdx.x; c < p.nchunks; c += gridDim.x * kArThreads)
    my_slot[c] = p.in[c];
  __threadfence_system();
  __syncthreads();

  // 2. cross-GPU barrier of block b
  if (p.barrier) {
    if (threadIdx.x < WORLD) {
      const int q = threadIdx.x;
      uint32_t* peer_flag = reinterpret_cast<uint32_t*>(p.region[q]) + blockIdx.x * kArMaxRanks + p.rank;
      st_release_sys(peer_flag, epoch);
      const long long t0 = clock64();
      while ((int32_t)(ld_acquire_sys(my_flags + q) - epoch) < 0) {     // a peer may already be a call ahead
        if (peer_wait_expired(t0)) {
          printf("b200_w4a16: all-reduce barrier timeout rank %d block %d peer %d\n", p.rank,
                 blockIdx.x, q);
          __trap();
        }
      }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) *my_epoch = epoch;
  // Every block's epoch advances on every call (block 0 covers the blocks this grid does not
  // have): the slot parity is then a property of the CALL, so calls of different sizes -- and
  // the all-reduce + RMSNorm kernel of w4a16_norm.cu, which maps chunks to blocks differently
  // -- may be mixed on one stream: rank r reaches call e+2 only after a peer's block 0 arrived
  // at call e+1, i.e. after that peer's call e (all of its blocks) had finished reading.
  if (blockIdx.x == 0 && threadIdx.x >= gridDim.x && threadIdx.x < kArMaxBlocks)
    reinterpret_cast<uint32_t*>(mine + kArFlagBytes)[threadIdx.x] = epoch;

  // 3. reduce in rank order (identical on every rank), fp32 accumulate
  for (uint32_t c = blockIdx.x * kArThreads + threadIdx.x; c < p.nchunks; c += gridDim.x * kArThreads) {
    uint4 v[WORLD];
#pragma unroll
    for (int q = 0; q < WORLD; ++q)
      v[q] = ld_volatile_v4(reinterpret_cast<const uint4*>(p.region[q] + slot_off) + c);
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
#pragma unroll
    for (int q = 0; q < WORLD; ++q) add8(acc, v[q]);
    uint4 o;
    __half2* oh = reinterpret_cast<__half2*>(&o);
#pragma unroll
    for (int i = 0; i < 4; ++i) oh[i] = __floats2half2_rn(acc[2 * i], acc[2 * i + 1]);
    p.out[c] = o;
  }
}

// Packet ("flag in data") variant: see add_rmsnorm_kernel<WORLD, 1, true> in w4a16_norm.cu for
// the protocol.  Every 16-byte chunk goes to each peer as two stores {data, epoch, data, epoch};
// the receiver spins on the packets: one NVLink one-way trip, no fence, no flag, no barrier.
template <int WORLD>
__global__ void __launch_bounds__(kArThreads)
allreduce_ll_kernel(const ArParams p) {
  uint8_t* mine = p.region[p.rank];
  uint32_t* epochs = reinterpret_cast<uint32_t*>(mine + kArFlagBytes);
  __shared__ uint32_t s_epoch;
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (threadIdx.x == 0) s_epoch = epochs[blockIdx.x] + 1;
  __syncthreads();
  const uint32_t epoch = s_epoch;
  const uint32_t base = p.ll_off + (epoch & 1u) * (uint32_t)WORLD * p.ll_slot_bytes;
  for (uint32_t c = blockIdx.x * kArThreads + threadIdx.x; c < p.nchunks; c += gridDim.x * kArThreads) {
    const uint4 v = p.in[c];
    const uint4 pk0 = make_uint4(v.x, epoch, v.y, epoch), pk1 = make_uint4(v.z, epoch, v.w, epoch);
#pragma unroll
    for (int q = 0; q < WORLD; ++q) {
      if (q == p.rank) continue;
      uint4* dst = reinterpret_cast<uint4*>(p.region[q] + base + (uint32_t)p.rank * p.ll_slot_bytes) + 2 * (size_t)c;
      asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(dst), "r"(pk0.x), "r"(pk0.y), "r"(pk0.z), "r"(pk0.w) : "memory");
      asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(dst + 1), "r"(pk1.x), "r"(pk1.y), "r"(pk1.z), "r"(pk1.w) : "memory");
    }
  }
  if (threadIdx.x == 0) epochs[blockIdx.x] = epoch;
  if (blockIdx.x == 0 && threadIdx.x >= gridDim.x && threadIdx.x < kArMaxBlocks) epochs[threadIdx.x] = epoch;
  for (uint32_t c = blockIdx.x * kArThreads + threadIdx.x; c < p.nchunks; c += gridDim.x * kArThreads) {
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
#pragma unroll
    for (int q = 0; q < WORLD; ++q) {
      if (q == p.rank) { add8(acc, p.in[c]); continue; }
      const uint4* src = reinterpret_cast<const uint4*>(mine + base + (uint32_t)q * p.ll_slot_bytes) + 2 * (size_t)c;
      uint4 a, b;
      const long long t0 = clock64();
      for (;;) {
        a = ld_volatile_v4(src);
        b = ld_volatile_v4(src + 1);
        if (a.y == epoch && a.w == epoch && b.y == epoch && b.w == epoch) break;
        if (!p.barrier || peer_wait_expired(t0)) {
          printf("b200_w4a16: all-reduce packet timeout rank %d block %d source %d\n", p.rank, blockIdx.x, q);
          __trap();
        }
      }
      add8(acc, make_uint4(a.x, a.z, b.x, b.z));
    }
    uint4 o;
    __half2* oh = reinterpret_cast<__half2*>(&o);
#pragma unroll
    for (int i = 0; i < 4; ++i) oh[i] = __floats2half2_rn(acc[2 * i], acc[2 * i + 1]);
    p.out[c] = o;
  }
}

// Two-shot variant for messages the one-shot pull handles badly (every rank reading every peer's
// WHOLE message: (world - 1) x n bytes into each GPU): reduce-scatter + all-gather, the counterpart
// of the reference's two-stage kernel (csrc/custom_all_reduce.cuh:239-290, dispatch :520-531).
//   1. block b stages its chunk set of `in` into the own slot (as the one-shot kernel);
//   2. per-block NVLink barrier A (flags[b][rank] = epoch);
//   3. rank r sums chunks [r n / W, (r + 1) n / W) over all ranks (rank order, fp32) and PUSHES
//      the fp16 result over that chunk in EVERY rank's slot -- the only reader of peer q's copy of
//      such a chunk is rank r itself, which has read it before it overwrites it;
//   4. system fence, per-block barrier B (flags2[b][rank] = epoch);
//   5. the own slot now holds the whole reduced message: block b copies its chunk set to `out`.
// Each GPU moves 2 (W - 1) / W x n bytes over NVLink instead of (W - 1) x n.  Chunk -> block is
// the same map in every step and on every rank, so the per-block barriers cover exactly the
// chunks a block touches.  Epochs, slots and the parity argument are shared with the other
// kernels (a rank cannot reach call e + 2 before every peer has finished call e: barrier B).
// `mode` (single-GPU emulation only, see tests/test_gpu_allreduce.py): 1 = all steps with barriers;
// 2 = step 3 alone (inputs staged by the test); 3 = step 5 alone, on the epoch mode 2 left.
template <int WORLD>
__global__ void __launch_bounds__(2 * kArThreads)
allreduce_twoshot_kernel(const ArParams p) {
  constexpr int kT = 2 * kArThreads;
  uint8_t* mine = p.region[p.rank];
  uint32_t* my_flags = reinterpret_cast<uint32_t*>(mine) + blockIdx.x * kArMaxRanks;
  uint32_t* my_flags2 = reinterpret_cast<uint32_t*>(mine + kArFlags2Off) + blockIdx.x * kArMaxRanks;
  uint32_t* my_epoch = reinterpret_cast<uint32_t*>(mine + kArFlagBytes) + blockIdx.x;
  __shared__ uint32_t s_epoch;
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const int mode = p.barrier;
  if (threadIdx.x == 0) s_epoch = *my_epoch + (mode == 3 ? 0u : 1u);
  __syncthreads();
  const uint32_t epoch = s_epoch;
  const uint32_t slot_off = kArHeaderBytes + (epoch & 1u) * p.slot_bytes;
  uint4* my_slot = reinterpret_cast<uint4*>(mine + slot_off);
  const uint32_t first = blockIdx.x * kT + threadIdx.x, stride = gridDim.x * kT;

  auto wait_peers = [&](uint32_t* flags_base_off_own, uint32_t off) {
    if (threadIdx.x < WORLD) {
      const int q = threadIdx.x;
      uint32_t* peer_flag = reinterpret_cast<uint32_t*>(p.region[q] + off) + blockIdx.x * kArMaxRanks + p.rank;
      st_release_sys(peer_flag, epoch);
      const long long t0 = clock64();
      while ((int32_t)(ld_acquire_sys(flags_base_off_own + q) - epoch) < 0) {
        if (peer_wait_expired(t0)) {
          printf("b200_w4a16: two-shot all-reduce barrier timeout rank %d block %d peer %d\n", p.rank,
                 blockIdx.x, q);
          __trap();
        }
      }
    }
    __syncthreads();
  };

  if (mode != 3) {
    // 1. stage (the emulation stages every rank's input by hand: with the ranks run one after
    // the other, a later rank's staging would overwrite what an earlier one pushed)
    if (mode == 1) {
      for (uint32_t c = first; c < p.nchunks; c += stride) my_slot[c] = p.in[c];
      __threadfence_system();
    }
    __syncthreads();
    // 2. barrier A
    if (mode == 1) wait_peers(my_flags, 0u);
    if (threadIdx.x == 0) *my_epoch = epoch;
    if (blockIdx.x == 0 && threadIdx.x >= gridDim.x && threadIdx.x < kArMaxBlocks)
      reinterpret_cast<uint32_t*>(mine + kArFlagBytes)[threadIdx.x] = epoch;
    // 3. reduce the own range, push it everywhere
    const uint32_t lo = (uint32_t)(((uint64_t)p.nchunks * (uint32_t)p.rank) / WORLD);
    const uint32_t hi = (uint32_t)(((uint64_t)p.nchunks * (uint32_t)(p.rank + 1)) / WORLD);
    for (uint32_t c = first; c < p.nchunks; c += stride) {
      if (c < lo || c >= hi) continue;
      uint4 v[WORLD];
#pragma unroll
      for (int q = 0; q < WORLD; ++q)
        v[q] = ld_volatile_v4(reinterpret_cast<const uint4*>(p.region[q] + slot_off) + c);
      float acc[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] = 0.f;
#pragma unroll
      for (int q = 0; q < WORLD; ++q) add8(acc, v[q]);
      uint4 o;
      __half2* oh = reinterpret_cast<__half2*>(&o);
#pragma unroll
      for (int i = 0; i < 4; ++i) oh[i] = __floats2half2_rn(acc[2 * i], acc[2 * i + 1]);
#pragma unroll
      for (int q = 0; q < WORLD; ++q)
        reinterpret_cast<uint4*>(p.region[q] + slot_off)[c] = o;
    }
    __threadfence_system();
    __syncthreads();
    // 4. barrier B
    if (mode == 1) wait_peers(my_flags2, (uint32_t)kArFlags2Off);
  }
  // 5. the own slot holds the reduced message
  if (mode != 2)
    for (uint32_t c = first; c < p.nchunks; c += stride) p.out[c] = ld_volatile_v4(my_slot + c);
}

template <int WORLD>
int launch_ar_twoshot(const ArParams& prm, int blocks, cudaStream_t stream) {
  allreduce_twoshot_kernel<WORLD><<<blocks, 2 * kArThreads, 0, stream>>>(prm);
  return (int)cudaGetLastError();
}

template <int WORLD>
int launch_ar_ll(const ArParams& prm, int blocks, cudaStream_t stream) {
  allreduce_ll_kernel<WORLD><<<blocks, kArThreads, 0, stream>>>(prm);
  return (int)cudaGetLastError();
}

template <int WORLD>
int launch_ar(const ArParams& prm, int blocks, cudaStream_t stream) {
  allreduce_oneshot_kernel<WORLD><<<blocks, kArThreads, 0, stream>>>(prm);
  return (int)cudaGetLastError();
}

}  // namespace

size_t ar_region_bytes(size_t max_message_bytes) {
  const size_t slot = (max_message_bytes + 255) / 256 * 256;
  return kArHeaderBytes + 2 * slot;
}

int launch_allreduce_twoshot(void* const* regions, int rank, int world, const uint16_t* in,
                             uint16_t* out, size_t n_halfs, size_t max_message_bytes, int mode,
                             cudaStream_t stream) {
  ArParams prm;
  prm.ll_off = 0;
  prm.ll_slot_bytes = 0;
  for (int q = 0; q < kArMaxRanks; ++q) prm.region[q] = q < world ? static_cast<uint8_t*>(regions[q]) : nullptr;
  prm.in = reinterpret_cast<const uint4*>(in);
  prm.out = reinterpret_cast<uint4*>(out);
  prm.nchunks = (uint32_t)(n_halfs / 8);
  prm.slot_bytes = (uint32_t)((max_message_bytes + 255) / 256 * 256);
  prm.rank = rank;
  prm.world = world;
  prm.barrier = mode;
  int blocks = (int)((prm.nchunks + 2 * kArThreads - 1) / (2 * kArThreads));
  if (blocks < 1) blocks = 1;
  if (blocks > kArMaxBlocks) blocks = kArMaxBlocks;
  switch (world) {
    case 2: return launch_ar_twoshot<2>(prm, blocks, stream);
    case 3: return launch_ar_twoshot<3>(prm, blocks, stream);
    case 4: return launch_ar_twoshot<4>(prm, blocks, stream);
    case 5: return launch_ar_twoshot<5>(prm, blocks, stream);
    case 6: return launch_ar_twoshot<6>(prm, blocks, stream);
    case 7: return launch_ar_twoshot<7>(prm, blocks, stream);
    case 8: return launch_ar_twoshot<8>(prm, blocks, stream);
    default: return (int)cudaErrorInvalidValue;
  }
}

int launch_allreduce(void* const* regions, int rank, int world, const uint16_t* in, uint16_t* out,
                     size_t n_halfs, size_t max_message_bytes, int barrier, cudaStream_t stream,
                     size_t ll_offset, size_t ll_message_bytes) {
  ArParams prm;
  prm.ll_off = (uint32_t)ll_offset;
  prm.ll_slot_bytes = (uint32_t)(2 * ll_message_bytes);
  for (int q = 0; q < kArMaxRanks; ++q) prm.region[q] = q < world ? static_cast<uint8_t*>(regions[q]) : nullptr;
  prm.in = reinterpret_cast<const uint4*>(in);
  prm.out = reinterpret_cast<uint4*>(out);
  prm.nchunks = (uint32_t)(n_halfs / 8);
  prm.slot_bytes = (uint32_t)((max_message_bytes + 255) / 256 * 256);
  prm.rank = rank;
  prm.world = world;
  prm.barrier = barrier;
  int blocks = (int)((prm.nchunks + kArThreads - 1) / kArThreads);
  if (blocks < 1) blocks = 1;
  if (blocks > kArMaxBlocks) blocks = kArMaxBlocks;
  if (ll_message_bytes > 0) {
    switch (world) {
      case 2: return launch_ar_ll<2>(prm, blocks, stream);
      case 3: return launch_ar_ll<3>(prm, blocks, stream);
      case 4: return launch_ar_ll<4>(prm, blocks, stream);
      case 5: return launch_ar_ll<5>(prm, blocks, stream);
      case 6: return launch_ar_ll<6>(prm, blocks, stream);
      case 7: return launch_ar_ll<7>(prm, blocks, stream);
      case 8: return launch_ar_ll<8>(prm, blocks, stream);
      default: return (int)cudaErrorInvalidValue;
    }
  }
  switch (world) {
    case 1: return launch_ar<1>(prm, blocks, stream);
    case 2: return launch_ar<2>(prm, blocks, stream);
    case 3: return launch_ar<3>(prm, blocks, stream);
    case 4: return launch_ar<4>(prm, blocks, stream);
    case 5: return launch_ar<5>(prm, blocks, stream);
    case 6: return launch_ar<6>(prm, blocks, stream);
    case 7: return launch_ar<7>(prm, blocks, stream);
    case 8: return launch_ar<8>(prm, blocks, stream);
    default: return (int)cudaErrorInvalidValue;
  }
}

}  // namespace b200
