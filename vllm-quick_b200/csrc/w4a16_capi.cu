// w4a16_capi.cu -- the C ABI of libb200_w4a16.so (declared in include/b200_w4a16.h).
// Argument validation mirrors the reference host functions
// (csrc/quantization/awq/gemm_kernels.cu:332-446); nothing here allocates, synchronises
// or throws.  There is no CPU path: without an sm_100 device every compute entry point
// returns B200_ERR_NO_DEVICE.
#include <stdlib.h>
#include <string.h>

#include "../../include/b200_w4a16.h"
#include "w4a16_internal.h"

namespace {

thread_local int g_last_cuda_error = 0;

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Workspace layout: [0, kMaxPartialBytes) self-flagging partial-tile slots (whatever the plan,
// a launch uses a prefix of it and hands it back all-0xFF) | from kMaxPartialBytes on: scratch
// for the on-the-fly interleave of checkpoint-layout weights.  The scratch must never overlap
// ANY plan's slots: leftovers of a repack would read as arrived partial sums.
constexpr size_t kMaxPartialBytes = (size_t)b200::kMaxCtas * 2 * 2 * 128 * b200::kTileN * sizeof(float);

struct DeviceInfo {
  int status;    // b200_status_t of the probe
  int sm_count;
};

// Probe of the CURRENT device (a process may drive several): compute capability and the SM
// count the planner may rely on -- the stream-K kernels need every CTA of the grid resident,
// so on a device that exposes fewer SMs (MIG slice, MPS share) the grid shrinks with it.
DeviceInfo device_info() {
  static DeviceInfo cache[b200::kMaxDevices];
  static bool known[b200::kMaxDevices] = {};
  DeviceInfo d{B200_ERR_NO_DEVICE, b200::kMaxCtas};
  int dev = 0, major = 0, sms = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return d; }
  if (dev >= 0 && dev < b200::kMaxDevices && known[dev]) return cache[dev];
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess ||
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) {
    cudaGetLastError();
    return d;
  }
  if (major == 10) {
    d.status = B200_OK;
    d.sm_count = sms < b200::kMaxCtas ? sms : b200::kMaxCtas;
  }
  if (dev >= 0 && dev < b200::kMaxDevices) { cache[dev] = d; known[dev] = true; }
  return d;
}

int sm_count_or_default() {
  // workspace sizing and plan queries must work before a device is touched: one full B200
  // (an upper bound for the scratch; launches plan with the probed count).
  return b200::kMaxCtas;
}

inline int cuda_fail(int e) {
  g_last_cuda_error = e;
  return B200_ERR_CUDA;
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

bool gemm_shape_ok(int M, int K, int N, int G) {
  if (M < 0 || K <= 0 || N <= 0 || G <= 0) return false;
  if (N % 64 != 0 || G % 32 != 0 || K % G != 0 || K % b200::kTileK != 0) return false;
  if (!(G == 32 || G == 64 || G % 128 == 0)) return false;
  return true;
}

// one rank's fused GEMM output as 16-byte packets {data, epoch, data, epoch}: 16 tokens x N fp16
// = 32 B of data per channel, 64 B of packets; N rounded up to whole tiles
size_t fused_slot_bytes(int out_features) {
  return (size_t)((out_features + b200::kTileN - 1) / b200::kTileN) * b200::kTileN * 64;
}

bool use_pdl() {
#ifdef B200_TUNING   // `make tune` only: the product library reads no environment variables
  static const bool on = [] {
    const char* e = getenv("B200_W4A16_NO_PDL");
    return !(e && e[0] == '1');
  }();
  return on;
#else
  return true;
#endif
}

}  // namespace

extern "C" {

int b200_w4a16_version(void) { return 200; }

const char* b200_w4a16_strerror(int status) {
  switch (status) {
    case B200_OK: return "ok";
    case B200_ERR_BAD_SHAPE: return "shape precondition violated";
    case B200_ERR_BAD_ARG: return "null or misaligned pointer, or unknown layout";
    case B200_ERR_WORKSPACE: return "workspace missing, too small or misaligned (initialise it with b200_w4a16_workspace_init)";
    case B200_ERR_CUDA: return "CUDA call failed";
    case B200_ERR_NO_DEVICE: return "no sm_100 device (this library has no fallback path)";
    case B200_ERR_STATIC_LIMIT:
      return "b200_w4a16_gemm_noalloc: more than 16 tokens need more scratch than the library's built-in "
             "array holds (use b200_w4a16_gemm with a workspace)";
    default: return "unknown status";
  }
}

int b200_w4a16_last_cuda_error(void) { return g_last_cuda_error; }

/* tuning aid, not part of the public ABI: per-CTA pipeline timestamps in B200_TRACE builds */
void b200_w4a16_debug_set_trace(void* dev_buffer) {
  b200::set_trace_buffer(static_cast<uint32_t*>(dev_buffer));
}

size_t b200_w4a16_repacked_words(int K, int N) {
  if (K <= 0 || N <= 0) return 0;
  return (size_t)((N + b200::kTileN - 1) / b200::kTileN) * (size_t)(K / b200::kTileK) *
         b200::kTileWords;
}

size_t b200_w4a16_workspace_bytes(int M, int K, int N, int layout) {
  if (M <= 0 || K <= 0 || N <= 0) return 256;
  b200::GemmPlan plan;
  b200::make_gemm_plan(M, K, N, 128, 0, sm_count_or_default(), &plan);
  size_t bytes = align_up(plan.ws_partial_bytes, 256);
  if (layout == B200_LAYOUT_AWQ) bytes = kMaxPartialBytes + align_up(b200_w4a16_repacked_words(K, N) * 4, 256);
  return bytes < 256 ? 256 : bytes;
}

int b200_w4a16_workspace_init(void* workspace, size_t workspace_bytes, b200_stream_t stream) {
  if (!workspace || workspace_bytes == 0 || !aligned16(workspace)) return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  // every word of an unwritten partial-tile slot reads as kPartEmpty (w4a16_common.cuh)
  cudaError_t e = cudaMemsetAsync(workspace, 0xFF, workspace_bytes, reinterpret_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? B200_OK : cuda_fail((int)e);
}

int b200_w4a16_gemm_plan(int M, int K, int N, int G, int split_k_hint, int* tok_tile,
                         int* split_k, int* grid, int* smem_bytes) {
  if (!gemm_shape_ok(M, K, N, G)) return B200_ERR_BAD_SHAPE;
  b200::GemmPlan plan;
  if (b200::make_gemm_plan(M > 0 ? M : 1, K, N, G, split_k_hint, sm_count_or_default(), &plan))
    return B200_ERR_BAD_SHAPE;
  if (tok_tile) *tok_tile = plan.tok_tile;
  if (split_k) *split_k = plan.max_split;
  if (grid) *grid = plan.ctas;
  if (smem_bytes) *smem_bytes = plan.smem_bytes;
  return B200_OK;
}

int b200_w4a16_gemm_variant(int M, int K, int N, int G) {
  if (!gemm_shape_ok(M, K, N, G)) return -B200_ERR_BAD_SHAPE;
  b200::GemmPlan plan;
  if (b200::make_gemm_plan(M > 0 ? M : 1, K, N, G, 0, sm_count_or_default(), &plan))
    return -B200_ERR_BAD_SHAPE;
  return plan.kind == b200::kKindDecode ? B200_VARIANT_DECODE : B200_VARIANT_TCGEN05;
}

int b200_w4a16_gemm_unit_tiles(int M, int K, int N, int G) {
  if (!gemm_shape_ok(M, K, N, G)) return -B200_ERR_BAD_SHAPE;
  b200::GemmPlan plan;
  if (b200::make_gemm_plan(M > 0 ? M : 1, K, N, G, 0, sm_count_or_default(), &plan))
    return -B200_ERR_BAD_SHAPE;
  return plan.pair * (plan.two_cta ? 2 : 1);
}

int b200_w4a16_gemm_cta_pair(int M, int K, int N, int G) {
  if (!gemm_shape_ok(M, K, N, G)) return -B200_ERR_BAD_SHAPE;
  b200::GemmPlan plan;
  if (b200::make_gemm_plan(M > 0 ? M : 1, K, N, G, 0, sm_count_or_default(), &plan))
    return -B200_ERR_BAD_SHAPE;
  return plan.two_cta ? plan.tok_pair : 0;
}

int b200_w4a16_gemm(const uint16_t* x, const int32_t* qweight, const uint16_t* scales,
                    const int32_t* qzeros, uint16_t* y, int M, int K, int N, int G, int layout,
                    int split_k_hint, void* workspace, size_t workspace_bytes,
                    b200_stream_t stream) {
  return b200_w4a16_gemm_gated(x, qweight, scales, qzeros, y, M, K, N, G, layout, split_k_hint,
                               nullptr, workspace, workspace_bytes, stream);
}

static int gemm_impl(const uint16_t* x, const int32_t* qweight, const uint16_t* scales,
                     const int32_t* qzeros, uint16_t* y, int M, int K, int N, int G, int layout,
                     int split_k_hint, const int32_t* gate, int act, void* workspace,
                     size_t workspace_bytes, b200_stream_t stream);

int b200_w4a16_gemm_gated(const uint16_t* x, const int32_t* qweight, const uint16_t* scales,
                          const int32_t* qzeros, uint16_t* y, int M, int K, int N, int G,
                          int layout, int split_k_hint, const int32_t* gate, void* workspace,
                          size_t workspace_bytes, b200_stream_t stream) {
  return gemm_impl(x, qweight, scales, qzeros, y, M, K, N, G, layout, split_k_hint, gate, 0,
                   workspace, workspace_bytes, stream);
}

int b200_w4a16_gemm_silu_mul(const uint16_t* x, const int32_t* qweight, const uint16_t* scales,
                             const int32_t* qzeros, uint16_t* y, int M, int K, int N, int G,
                             int split_k_hint, void* workspace, size_t workspace_bytes,
                             b200_stream_t stream) {
  return b200_w4a16_gemm_silu_mul_gated(x, qweight, scales, qzeros, y, M, K, N, G, split_k_hint,
                                        nullptr, workspace, workspace_bytes, stream);
}

int b200_w4a16_gemm_silu_mul_gated(const uint16_t* x, const int32_t* qweight, const uint16_t* scales,
                                   const int32_t* qzeros, uint16_t* y, int M, int K, int N, int G,
                                   int split_k_hint, const int32_t* gate, void* workspace,
                                   size_t workspace_bytes, b200_stream_t stream) {
  if (!gemm_shape_ok(M, K, N, G)) return B200_ERR_BAD_SHAPE;
  // y is [M, N / 2] written with 16-byte stores when the gate is closed: N / 2 % 8 == 0
  if (M > b200::kDecodeMaxTokens || G % b200::kTileK != 0 || N % 16 != 0) return B200_ERR_BAD_SHAPE;
  return gemm_impl(x, qweight, scales, qzeros, y, M, K, N, G, B200_LAYOUT_B200, split_k_hint,
                   gate, 1, workspace, workspace_bytes, stream);
}

int b200_w4a16_silu_and_mul(const uint16_t* in, uint16_t* out, size_t rows, int d, int paired,
                            b200_stream_t stream) {
  if (d <= 0 || d % 8 != 0 || (paired != 0 && paired != 1)) return B200_ERR_BAD_SHAPE;
  if (rows == 0) return B200_OK;
  if (!in || !out || !aligned16(in) || !aligned16(out)) return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  int e = b200::launch_silu_and_mul(in, out, rows, d, paired, reinterpret_cast<cudaStream_t>(stream));
  return e == 0 ? B200_OK : cuda_fail(e);
}

static int gemm_impl(const uint16_t* x, const int32_t* qweight, const uint16_t* scales,
                     const int32_t* qzeros, uint16_t* y, int M, int K, int N, int G, int layout,
                     int split_k_hint, const int32_t* gate, int act, void* workspace,
                     size_t workspace_bytes, b200_stream_t stream) {
  if (!gemm_shape_ok(M, K, N, G)) return B200_ERR_BAD_SHAPE;
  if (gate && (reinterpret_cast<uintptr_t>(gate) & 3u)) return B200_ERR_BAD_ARG;
  if (layout != B200_LAYOUT_AWQ && layout != B200_LAYOUT_B200) return B200_ERR_BAD_ARG;
  if (M == 0) return B200_OK;
  if (!x || !qweight || !scales || !qzeros || !y) return B200_ERR_BAD_ARG;
  if (!aligned16(x) || !aligned16(qweight) || !aligned16(scales) || !aligned16(qzeros) ||
      !aligned16(y))
    return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;

  b200::GemmPlan plan;
  if (b200::make_gemm_plan(M, K, N, G, split_k_hint, dev.sm_count, &plan))
    return B200_ERR_BAD_SHAPE;
  const size_t part_bytes = align_up(plan.ws_partial_bytes, 256);
  size_t need = part_bytes;
  if (layout == B200_LAYOUT_AWQ) need = kMaxPartialBytes + align_up(b200_w4a16_repacked_words(K, N) * 4, 256);
  if (part_bytes > kMaxPartialBytes || !workspace || workspace_bytes < need || !aligned16(workspace))
    return B200_ERR_WORKSPACE;
  if (act && plan.kind != b200::kKindDecode) return B200_ERR_BAD_SHAPE;

  uint8_t* ws = static_cast<uint8_t*>(workspace);
  void* partials = ws;
  const int32_t* qw = qweight;
  bool pdl = use_pdl();
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (layout == B200_LAYOUT_AWQ) {
    // Checkpoint-layout weights: interleave into scratch first.  The weight prefetch of the
    // GEMM must then wait for the repack, so programmatic launch is off for this call.
    int32_t* scratch = reinterpret_cast<int32_t*>(ws + kMaxPartialBytes);
    int e = b200::launch_repack(qweight, scratch, K, N, B200_REPACK_AWQ_TO_B200, s);
    if (e != 0) return cuda_fail(e);
    qw = scratch;
    pdl = false;
  }
  int e = plan.kind == b200::kKindDecode
              ? b200::launch_decode(x, qw, scales, qzeros, y, M, K, N, G, plan, partials, pdl, s,
                                    nullptr, gate, act)
              : b200::launch_gemm(x, qw, scales, qzeros, y, M, K, N, G, plan, partials, pdl, s);
  if (e != 0) return cuda_fail(e);
  return B200_OK;
}

int b200_w4a16_dequant(const int32_t* qweight, const uint16_t* scales, const int32_t* qzeros,
                       uint16_t* W, int K, int N, int G, int layout, b200_stream_t stream) {
  if (K < 0 || N < 0 || G <= 0) return B200_ERR_BAD_SHAPE;
  if (layout == B200_LAYOUT_AWQ) {
    if (N % 8 != 0 || K % G != 0 || G % 4 != 0) return B200_ERR_BAD_SHAPE;
  } else if (layout == B200_LAYOUT_B200) {
    if (N % 64 != 0 || K % G != 0 || K % b200::kTileK != 0 || G % 32 != 0) return B200_ERR_BAD_SHAPE;
  } else {
    return B200_ERR_BAD_ARG;
  }
  if (K == 0 || N == 0) return B200_OK;
  if (!qweight || !scales || !qzeros || !W) return B200_ERR_BAD_ARG;
  if (!aligned16(qweight) || !aligned16(scales) || !aligned16(W) ||
      (reinterpret_cast<uintptr_t>(qzeros) & 3u))
    return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  int e = b200::launch_dequant(qweight, scales, qzeros, W, K, N, G, layout,
                               reinterpret_cast<cudaStream_t>(stream));
  return e == 0 ? B200_OK : cuda_fail(e);
}

int b200_w4a16_repack(const int32_t* src, int32_t* dst, int K, int N, int direction,
                      b200_stream_t stream) {
  if (K < 0 || N < 0 || K % b200::kTileK != 0 || N % 64 != 0) return B200_ERR_BAD_SHAPE;
  if (direction != B200_REPACK_AWQ_TO_B200 && direction != B200_REPACK_B200_TO_AWQ)
    return B200_ERR_BAD_ARG;
  if (K == 0 || N == 0) return B200_OK;
  if (!src || !dst || src == dst || (reinterpret_cast<uintptr_t>(src) & 3u) ||
      (reinterpret_cast<uintptr_t>(dst) & 3u))
    return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  int e = b200::launch_repack(src, dst, K, N, direction, reinterpret_cast<cudaStream_t>(stream));
  return e == 0 ? B200_OK : cuda_fail(e);
}

size_t b200_w4a16_ar_region_bytes(size_t max_message_bytes, int world, int fused_out_features) {
  if (world < 1 || world > b200::kArMaxRanks || fused_out_features < 0) return 0;
  return b200::ar_region_bytes(max_message_bytes) +
         2 * (size_t)world * fused_slot_bytes(fused_out_features);
}

int b200_w4a16_gemm_allreduce(const uint16_t* x, const int32_t* qweight, const uint16_t* scales,
                              const int32_t* qzeros, uint16_t* y, int M, int K, int N, int G,
                              int split_k_hint, void* workspace, size_t workspace_bytes,
                              void* const* regions, int rank, int world, size_t max_message_bytes,
                              int fused_out_features, b200_stream_t stream) {
  if (!gemm_shape_ok(M, K, N, G)) return B200_ERR_BAD_SHAPE;
  if (M > b200::kDecodeMaxTokens || G % b200::kTileK != 0 || N > fused_out_features ||
      (N + b200::kTileN - 1) / b200::kTileN > b200::kArMaxTiles)
    return B200_ERR_BAD_SHAPE;
  if (world < 2 || world > b200::kArMaxRanks || rank < 0 || rank >= world) return B200_ERR_BAD_ARG;
  if (M == 0) return B200_OK;
  if (!x || !qweight || !scales || !qzeros || !y || !regions) return B200_ERR_BAD_ARG;
  if (!aligned16(x) || !aligned16(qweight) || !aligned16(scales) || !aligned16(qzeros) || !aligned16(y))
    return B200_ERR_BAD_ARG;
  for (int q = 0; q < world; ++q)
    if (!regions[q] || !aligned16(regions[q])) return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  b200::GemmPlan plan;
  if (b200::make_gemm_plan(M, K, N, G, split_k_hint, dev.sm_count, &plan)) return B200_ERR_BAD_SHAPE;
  if (plan.kind != b200::kKindDecode) return B200_ERR_BAD_SHAPE;   // pinned to another variant (tuning builds)
  const size_t part_bytes = align_up(plan.ws_partial_bytes, 256);
  if (!workspace || workspace_bytes < part_bytes || !aligned16(workspace)) return B200_ERR_WORKSPACE;
  b200::FusedArParams ar;
  for (int q = 0; q < b200::kArMaxRanks; ++q)
    ar.region[q] = q < world ? static_cast<uint8_t*>(regions[q]) : nullptr;
  ar.rank = rank;
  ar.world = world;
  ar.ctr_off = b200::kArFusedCtrOff;
  ar.data_off = (uint32_t)b200::ar_region_bytes(max_message_bytes);
  ar.slot_bytes = (uint32_t)fused_slot_bytes(fused_out_features);
  uint8_t* ws = static_cast<uint8_t*>(workspace);
  int e = b200::launch_decode(x, qweight, scales, qzeros, y, M, K, N, G, plan, ws, use_pdl(),
                              reinterpret_cast<cudaStream_t>(stream), &ar);
  return e == 0 ? B200_OK : cuda_fail(e);
}

int b200_w4a16_ar_alloc(size_t bytes, void** dev_ptr) {
  if (!dev_ptr || bytes == 0) return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  cudaError_t e = cudaMalloc(dev_ptr, bytes);
  if (e != cudaSuccess) return cuda_fail((int)e);
  e = cudaMemset(*dev_ptr, 0, bytes);
  if (e != cudaSuccess) return cuda_fail((int)e);
  e = cudaDeviceSynchronize();
  return e == cudaSuccess ? B200_OK : cuda_fail((int)e);
}

int b200_w4a16_ar_free(void* dev_ptr) {
  if (!dev_ptr) return B200_OK;
  cudaError_t e = cudaFree(dev_ptr);
  return e == cudaSuccess ? B200_OK : cuda_fail((int)e);
}

int b200_w4a16_ar_export(const void* dev_ptr, void* handle64) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  if (!dev_ptr || !handle64) return B200_ERR_BAD_ARG;
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, const_cast<void*>(dev_ptr));
  if (e != cudaSuccess) return cuda_fail((int)e);
  memcpy(handle64, &h, sizeof(h));
  return B200_OK;
}

int b200_w4a16_ar_import(const void* handle64, void** dev_ptr) {
  if (!handle64 || !dev_ptr) return B200_ERR_BAD_ARG;
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  cudaError_t e = cudaIpcOpenMemHandle(dev_ptr, h, cudaIpcMemLazyEnablePeerAccess);
  return e == cudaSuccess ? B200_OK : cuda_fail((int)e);
}

int b200_w4a16_ar_unimport(void* dev_ptr) {
  if (!dev_ptr) return B200_OK;
  cudaError_t e = cudaIpcCloseMemHandle(dev_ptr);
  return e == cudaSuccess ? B200_OK : cuda_fail((int)e);
}

int b200_w4a16_allreduce(void* const* regions, int rank, int world, const uint16_t* in,
                         uint16_t* out, size_t n_halfs, size_t max_message_bytes, int barrier,
                         b200_stream_t stream) {
  if (world < 1 || world > b200::kArMaxRanks || rank < 0 || rank >= world) return B200_ERR_BAD_ARG;
  if (n_halfs % 8 != 0 || n_halfs * 2 > max_message_bytes || n_halfs / 8 > 0xFFFFFFFFull)
    return B200_ERR_BAD_SHAPE;
  if (n_halfs == 0) return B200_OK;
  if (!regions || !in || !out || !aligned16(in) || !aligned16(out)) return B200_ERR_BAD_ARG;
  for (int q = 0; q < world; ++q)
    if (!regions[q] || !aligned16(regions[q])) return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  int e = b200::launch_allreduce(regions, rank, world, in, out, n_halfs, max_message_bytes,
                                 barrier, reinterpret_cast<cudaStream_t>(stream));
  return e == 0 ? B200_OK : cuda_fail(e);
}

int b200_w4a16_allreduce_twoshot(void* const* regions, int rank, int world, const uint16_t* in,
                                 uint16_t* out, size_t n_halfs, size_t max_message_bytes, int mode,
                                 b200_stream_t stream) {
  if (world < 2 || world > b200::kArMaxRanks || rank < 0 || rank >= world) return B200_ERR_BAD_ARG;
  if (mode < 1 || mode > 3) return B200_ERR_BAD_ARG;
  if (n_halfs % 8 != 0 || n_halfs * 2 > max_message_bytes || n_halfs / 8 > 0xFFFFFFFFull)
    return B200_ERR_BAD_SHAPE;
  if (n_halfs == 0) return B200_OK;
  if (!regions || !in || !out || !aligned16(in) || !aligned16(out)) return B200_ERR_BAD_ARG;
  for (int q = 0; q < world; ++q)
    if (!regions[q] || !aligned16(regions[q])) return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  int e = b200::launch_allreduce_twoshot(regions, rank, world, in, out, n_halfs, max_message_bytes,
                                         mode, reinterpret_cast<cudaStream_t>(stream));
  return e == 0 ? B200_OK : cuda_fail(e);
}

int b200_w4a16_moe_gather(const uint16_t* x, const int32_t* idx, uint16_t* out, size_t rows,
                          int hidden, b200_stream_t stream) {
  if (hidden <= 0 || hidden % 8 != 0) return B200_ERR_BAD_SHAPE;
  if (rows == 0) return B200_OK;
  if (!x || !idx || !out || !aligned16(x) || !aligned16(out) || (reinterpret_cast<uintptr_t>(idx) & 3u))
    return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  int e = b200::launch_moe_gather(x, idx, out, rows, hidden, reinterpret_cast<cudaStream_t>(stream));
  return e == 0 ? B200_OK : cuda_fail(e);
}

int b200_w4a16_moe_scatter_add(const uint16_t* y, const int32_t* idx, const float* w, uint16_t* acc,
                               size_t rows, int hidden, b200_stream_t stream) {
  if (hidden <= 0 || hidden % 8 != 0) return B200_ERR_BAD_SHAPE;
  if (rows == 0) return B200_OK;
  if (!y || !idx || !w || !acc || !aligned16(y) || !aligned16(acc) ||
      (reinterpret_cast<uintptr_t>(idx) & 3u) || (reinterpret_cast<uintptr_t>(w) & 3u))
    return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  int e = b200::launch_moe_scatter_add(y, idx, w, acc, rows, hidden, reinterpret_cast<cudaStream_t>(stream));
  return e == 0 ? B200_OK : cuda_fail(e);
}

int b200_w4a16_moe_grouped_gemm(const uint16_t* x, const int32_t* const* qweight,
                                const uint16_t* const* scales, const int32_t* const* qzeros,
                                uint16_t* y, int M, int K, int N, int G, int n_experts,
                                const int32_t* counts, size_t x_expert_stride, int act,
                                void* workspace, size_t workspace_bytes, b200_stream_t stream) {
  if (!gemm_shape_ok(M, K, N, G)) return B200_ERR_BAD_SHAPE;
  if (M > b200::kDecodeMaxTokens || G % b200::kTileK != 0) return B200_ERR_BAD_SHAPE;
  if (n_experts < 1 || n_experts > b200::kMaxGroupExperts) return B200_ERR_BAD_SHAPE;
  if ((act != 0 && act != 1) || (act && N % 16 != 0)) return B200_ERR_BAD_SHAPE;
  if (x_expert_stride != 0 && x_expert_stride != (size_t)M * K) return B200_ERR_BAD_SHAPE;
  if (M == 0) return B200_OK;
  if (!x || !qweight || !scales || !qzeros || !y || !counts) return B200_ERR_BAD_ARG;
  if (!aligned16(x) || !aligned16(y) || (reinterpret_cast<uintptr_t>(counts) & 3u)) return B200_ERR_BAD_ARG;
  b200::ExpertGroup grp{};
  for (int e = 0; e < n_experts; ++e) {
    if (!qweight[e] || !scales[e] || !qzeros[e] || !aligned16(qweight[e]) || !aligned16(scales[e]) ||
        !aligned16(qzeros[e]))
      return B200_ERR_BAD_ARG;
    grp.qw[e] = qweight[e];
    grp.scales[e] = scales[e];
    grp.qz[e] = qzeros[e];
  }
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  b200::GemmPlan plan;                                      // one expert's plan: grid and tile counts
  if (b200::make_gemm_plan(M, K, N, G, 0, dev.sm_count, &plan) || plan.kind != b200::kKindDecode)
    return B200_ERR_BAD_SHAPE;
  if ((long long)plan.n_tiles * plan.k_blocks * n_experts >= (1LL << 31)) return B200_ERR_BAD_SHAPE;
  if (n_experts > 1) {
    // The single-expert plan may have sized the grid to end ranges at tile borders (short walks);
    // over several experts' tiles the walk is long and border alignment is moot: all SMs.
    const long long t_one = (long long)plan.n_tiles * plan.k_blocks;
    plan.ctas = (int)(t_one < dev.sm_count ? t_one : dev.sm_count);
    b200::fill_decode_plan(M, &plan);                        // workspace bytes follow the grid
  }
  const size_t part_bytes = align_up(plan.ws_partial_bytes, 256);
  if (part_bytes > kMaxPartialBytes || !workspace || workspace_bytes < part_bytes || !aligned16(workspace))
    return B200_ERR_WORKSPACE;
  grp.counts = counts;
  grp.n_experts = n_experts;
  grp.x_stride = (long long)x_expert_stride;
  grp.y_stride = (long long)M * (act ? N / 2 : N);
  int e = b200::launch_decode_grouped(x, grp, y, M, K, N, G, plan, workspace, use_pdl(),
                                      reinterpret_cast<cudaStream_t>(stream), act);
  return e == 0 ? B200_OK : cuda_fail(e);
}

int b200_w4a16_gemm_rows(const uint16_t* x, const int32_t* qweight, const uint16_t* scales,
                         const int32_t* qzeros, uint16_t* y, int max_rows, const int32_t* rows_dev,
                         int K, int N, int G, void* workspace, size_t workspace_bytes,
                         b200_stream_t stream) {
  if (!gemm_shape_ok(max_rows, K, N, G)) return B200_ERR_BAD_SHAPE;
  if (max_rows == 0) return B200_OK;
  if (!rows_dev || (reinterpret_cast<uintptr_t>(rows_dev) & 3u)) return B200_ERR_BAD_ARG;
  if (max_rows <= b200::kDecodeMaxTokens && G % b200::kTileK == 0)   // decode kernel: all rows
    return gemm_impl(x, qweight, scales, qzeros, y, max_rows, K, N, G, B200_LAYOUT_B200, 0, nullptr, 0,
                     workspace, workspace_bytes, stream);
  if (!x || !qweight || !scales || !qzeros || !y) return B200_ERR_BAD_ARG;
  if (!aligned16(x) || !aligned16(qweight) || !aligned16(scales) || !aligned16(qzeros) || !aligned16(y))
    return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  b200::GemmPlan plan;
  if (b200::make_gemm_plan_rows(max_rows, K, N, G, dev.sm_count, &plan)) return B200_ERR_BAD_SHAPE;
  const size_t part_bytes = align_up(plan.ws_partial_bytes, 256);
  if (part_bytes > kMaxPartialBytes || !workspace || workspace_bytes < part_bytes || !aligned16(workspace))
    return B200_ERR_WORKSPACE;
  int e = b200::launch_gemm(x, qweight, scales, qzeros, y, max_rows, K, N, G, plan, workspace, use_pdl(),
                            reinterpret_cast<cudaStream_t>(stream), rows_dev);
  return e == 0 ? B200_OK : cuda_fail(e);
}

int b200_w4a16_moe_compact(const int64_t* topk_idx, const float* topk_w, int32_t* tok, float* w,
                           size_t tokens, int top_k, int first_expert, int n_local,
                           b200_stream_t stream) {
  if (top_k < 1 || first_expert < 0 || n_local < 1 || tokens * (size_t)top_k > (size_t)0x7fffffff)
    return B200_ERR_BAD_SHAPE;
  if (tokens == 0) return B200_OK;
  if (!topk_idx || !topk_w || !tok || !w || (reinterpret_cast<uintptr_t>(topk_idx) & 7u) ||
      (reinterpret_cast<uintptr_t>(topk_w) & 3u) || (reinterpret_cast<uintptr_t>(tok) & 3u) ||
      (reinterpret_cast<uintptr_t>(w) & 3u))
    return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  int e = b200::launch_moe_compact(reinterpret_cast<const long long*>(topk_idx), topk_w, tok, w, tokens,
                                   top_k, first_expert, n_local, reinterpret_cast<cudaStream_t>(stream));
  return e == 0 ? B200_OK : cuda_fail(e);
}

int b200_w4a16_moe_gather_rows(const uint16_t* x, const int32_t* idx, const int32_t* rows_dev,
                               uint16_t* out, size_t max_rows, int hidden, b200_stream_t stream) {
  if (hidden <= 0 || hidden % 8 != 0) return B200_ERR_BAD_SHAPE;
  if (max_rows == 0) return B200_OK;
  if (!x || !idx || !out || !rows_dev || !aligned16(x) || !aligned16(out) ||
      (reinterpret_cast<uintptr_t>(idx) & 3u) || (reinterpret_cast<uintptr_t>(rows_dev) & 3u))
    return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  int e = b200::launch_moe_gather(x, idx, out, max_rows, hidden, reinterpret_cast<cudaStream_t>(stream), rows_dev);
  return e == 0 ? B200_OK : cuda_fail(e);
}

int b200_w4a16_moe_scatter_add_rows(const uint16_t* y, const int32_t* idx, const float* w,
                                    const int32_t* rows_dev, uint16_t* acc, size_t max_rows,
                                    int hidden, b200_stream_t stream) {
  if (hidden <= 0 || hidden % 8 != 0) return B200_ERR_BAD_SHAPE;
  if (max_rows == 0) return B200_OK;
  if (!y || !idx || !w || !acc || !rows_dev || !aligned16(y) || !aligned16(acc) ||
      (reinterpret_cast<uintptr_t>(idx) & 3u) || (reinterpret_cast<uintptr_t>(w) & 3u) ||
      (reinterpret_cast<uintptr_t>(rows_dev) & 3u))
    return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  int e = b200::launch_moe_scatter_add(y, idx, w, acc, max_rows, hidden,
                                       reinterpret_cast<cudaStream_t>(stream), rows_dev);
  return e == 0 ? B200_OK : cuda_fail(e);
}

int b200_w4a16_silu_and_mul_rows(const uint16_t* in, uint16_t* out, size_t max_rows,
                                 const int32_t* rows_dev, int d, int paired, b200_stream_t stream) {
  if (d <= 0 || d % 8 != 0 || (paired != 0 && paired != 1)) return B200_ERR_BAD_SHAPE;
  if (max_rows == 0) return B200_OK;
  if (!in || !out || !rows_dev || !aligned16(in) || !aligned16(out) ||
      (reinterpret_cast<uintptr_t>(rows_dev) & 3u))
    return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  int e = b200::launch_silu_and_mul(in, out, max_rows, d, paired, reinterpret_cast<cudaStream_t>(stream), rows_dev);
  return e == 0 ? B200_OK : cuda_fail(e);
}

int b200_w4a16_moe_route(const uint16_t* x, const uint16_t* gate_w, float* topk_w, int64_t* topk_idx,
                         float* dense_w, int32_t* counts, size_t tokens, int hidden, int n_experts,
                         int top_k, b200_stream_t stream) {
  if (hidden <= 0 || hidden % 8 != 0 || n_experts < 1 || n_experts > b200::kMaxRouteExperts ||
      top_k < 1 || top_k > n_experts || tokens > (size_t)0x7fffffff)
    return B200_ERR_BAD_SHAPE;
  if (!counts || (reinterpret_cast<uintptr_t>(counts) & 3u)) return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (tokens == 0) {                                        // nobody chose anybody
    cudaError_t e = cudaMemsetAsync(counts, 0, (size_t)n_experts * sizeof(int32_t), s);
    return e == cudaSuccess ? B200_OK : cuda_fail((int)e);
  }
  if (!x || !gate_w || !topk_w || !topk_idx || !dense_w || !aligned16(x) || !aligned16(gate_w) ||
      (reinterpret_cast<uintptr_t>(topk_w) & 3u) || (reinterpret_cast<uintptr_t>(topk_idx) & 7u) ||
      (reinterpret_cast<uintptr_t>(dense_w) & 3u))
    return B200_ERR_BAD_ARG;
  int e = b200::launch_moe_route(x, gate_w, topk_w, reinterpret_cast<long long*>(topk_idx), dense_w,
                                 counts, tokens, hidden, n_experts, top_k, s);
  return e == 0 ? B200_OK : cuda_fail(e);
}

int b200_w4a16_moe_combine(const uint16_t* y, const float* w, const int32_t* counts, uint16_t* out,
                           int n_experts, int w_stride, size_t rows, int hidden,
                           b200_stream_t stream) {
  if (hidden <= 0 || hidden % 8 != 0 || n_experts < 1 || w_stride < n_experts) return B200_ERR_BAD_SHAPE;
  if (rows == 0) return B200_OK;
  if (!y || !w || !counts || !out || !aligned16(y) || !aligned16(out) ||
      (reinterpret_cast<uintptr_t>(w) & 3u) || (reinterpret_cast<uintptr_t>(counts) & 3u))
    return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  int e = b200::launch_moe_combine(y, w, counts, out, n_experts, w_stride, rows, hidden,
                                   reinterpret_cast<cudaStream_t>(stream));
  return e == 0 ? B200_OK : cuda_fail(e);
}

int b200_w4a16_add_rms_norm(const uint16_t* in, const uint16_t* residual_in, const uint16_t* weight,
                      uint16_t* out, uint16_t* residual_out, float eps, size_t rows, int hidden,
                      b200_stream_t stream) {
  if (hidden <= 0 || hidden % 8 != 0 || hidden > b200::norm_max_hidden() ||
      rows > 0xFFFFFFFFull / (size_t)(hidden / 8))
    return B200_ERR_BAD_SHAPE;
  if (!(eps >= 0.f)) return B200_ERR_BAD_ARG;
  if (rows == 0) return B200_OK;
  if (!in || !weight || !out || !aligned16(in) || !aligned16(weight) || !aligned16(out) ||
      !aligned16(residual_in) || !aligned16(residual_out))
    return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  int e = b200::launch_add_rmsnorm(in, residual_in, weight, out, residual_out, eps, rows, hidden,
                                   reinterpret_cast<cudaStream_t>(stream));
  return e == 0 ? B200_OK : cuda_fail(e);
}

int b200_w4a16_allreduce_add_rms_norm(void* const* regions, int rank, int world, const uint16_t* in,
                                const uint16_t* residual_in, const uint16_t* weight,
                                uint16_t* out, uint16_t* residual_out, float eps, size_t rows,
                                int hidden, size_t max_message_bytes, int barrier,
                                b200_stream_t stream) {
  if (world < 1 || world > b200::kArMaxRanks || rank < 0 || rank >= world) return B200_ERR_BAD_ARG;
  if (hidden <= 0 || hidden % 8 != 0 || hidden > b200::norm_max_hidden() ||
      rows > max_message_bytes / (2 * (size_t)hidden))
    return B200_ERR_BAD_SHAPE;
  if (!(eps >= 0.f)) return B200_ERR_BAD_ARG;
  if (rows == 0) return B200_OK;
  if (!regions || !in || !weight || !out || !aligned16(in) || !aligned16(weight) ||
      !aligned16(out) || !aligned16(residual_in) || !aligned16(residual_out))
    return B200_ERR_BAD_ARG;
  for (int q = 0; q < world; ++q)
    if (!regions[q] || !aligned16(regions[q])) return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  int e = b200::launch_allreduce_add_rmsnorm(regions, rank, world, in, residual_in, weight, out,
                                             residual_out, eps, rows, hidden, max_message_bytes,
                                             barrier, reinterpret_cast<cudaStream_t>(stream));
  return e == 0 ? B200_OK : cuda_fail(e);
}

int b200_w4a16_allreduce_ll(void* const* regions, int rank, int world, const uint16_t* in,
                            uint16_t* out, size_t n_halfs, size_t ll_offset, size_t ll_message_bytes,
                            int barrier, b200_stream_t stream) {
  if (world < 2 || world > b200::kArMaxRanks || rank < 0 || rank >= world) return B200_ERR_BAD_ARG;
  if (n_halfs % 8 != 0 || ll_message_bytes == 0 || n_halfs * 2 > ll_message_bytes) return B200_ERR_BAD_SHAPE;
  if (ll_offset % 256 != 0 || ll_offset < (size_t)b200::kArHeaderBytes ||
      ll_offset + 4 * (size_t)world * ll_message_bytes > 0xFFFFFFFFull)
    return B200_ERR_BAD_ARG;
  if (n_halfs == 0) return B200_OK;
  if (!regions || !in || !out || !aligned16(in) || !aligned16(out)) return B200_ERR_BAD_ARG;
  for (int q = 0; q < world; ++q)
    if (!regions[q] || !aligned16(regions[q])) return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  int e = b200::launch_allreduce(regions, rank, world, in, out, n_halfs, 0, barrier,
                                 reinterpret_cast<cudaStream_t>(stream), ll_offset, ll_message_bytes);
  return e == 0 ? B200_OK : cuda_fail(e);
}

int b200_w4a16_allreduce_add_rms_norm_ll(void* const* regions, int rank, int world,
                                         const uint16_t* in, const uint16_t* residual_in,
                                         const uint16_t* weight, uint16_t* out,
                                         uint16_t* residual_out, float eps, size_t rows, int hidden,
                                         size_t ll_offset, size_t ll_message_bytes, int barrier,
                                         b200_stream_t stream) {
  if (world < 2 || world > b200::kArMaxRanks || rank < 0 || rank >= world) return B200_ERR_BAD_ARG;
  if (hidden <= 0 || hidden % 8 != 0 || hidden > b200::norm_max_hidden() || ll_message_bytes == 0 ||
      rows > ll_message_bytes / (2 * (size_t)hidden))
    return B200_ERR_BAD_SHAPE;
  if (!(eps >= 0.f) || ll_offset % 256 != 0 || ll_offset < (size_t)b200::kArHeaderBytes ||
      ll_offset + 4 * (size_t)world * ll_message_bytes > 0xFFFFFFFFull)
    return B200_ERR_BAD_ARG;
  if (rows == 0) return B200_OK;
  if (!regions || !in || !weight || !out || !aligned16(in) || !aligned16(weight) ||
      !aligned16(out) || !aligned16(residual_in) || !aligned16(residual_out))
    return B200_ERR_BAD_ARG;
  for (int q = 0; q < world; ++q)
    if (!regions[q] || !aligned16(regions[q])) return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  int e = b200::launch_allreduce_add_rmsnorm(regions, rank, world, in, residual_in, weight, out,
                                             residual_out, eps, rows, hidden, 0, barrier,
                                             reinterpret_cast<cudaStream_t>(stream), ll_offset,
                                             ll_message_bytes);
  return e == 0 ? B200_OK : cuda_fail(e);
}

}  // extern "C"

// ---- b200_w4a16_gemm_noalloc: the decode-batch GEMM on scratch built into the library ---------
// (no counterpart in the reference; see include/b200_w4a16.h).  The array is the decode kernel's
// partial-tile workspace for the largest plan it is used with: [148 CTAs][2 slots][16 tokens x 128
// channels] fp32.  Every GPU that loads the library holds its own copy until the process exits.
namespace {

static __device__ __align__(256) uint8_t g_noalloc_scratch[B200_NOALLOC_SCRATCH_BYTES];

// In host code `g_noalloc_scratch` is a host-side stand-in for the symbol, not a device address:
// the address a kernel can use comes from cudaGetSymbolAddress, once per device, with that
// device current.  (A benign race between threads: every thread stores the same value.)
int noalloc_scratch(int dev, uint8_t** out) {
  static uint8_t* addr[b200::kMaxDevices] = {};
  if (dev >= 0 && dev < b200::kMaxDevices && addr[dev]) { *out = addr[dev]; return 0; }
  void* p = nullptr;
  cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
  if (e != cudaSuccess || !p) return e != cudaSuccess ? (int)e : (int)cudaErrorInvalidSymbol;
  if (dev >= 0 && dev < b200::kMaxDevices) addr[dev] = static_cast<uint8_t*>(p);
  *out = static_cast<uint8_t*>(p);
  return 0;
}

}  // namespace

extern "C" int b200_w4a16_gemm_noalloc(const uint16_t* x, const int32_t* qweight, const uint16_t* scales,
                                       const int32_t* qzeros, uint16_t* y, int M, int K, int N, int G,
                                       b200_stream_t stream) {
  if (!gemm_shape_ok(M, K, N, G)) return B200_ERR_BAD_SHAPE;
  if (M > B200_NOALLOC_MAX_TOKENS) return B200_ERR_STATIC_LIMIT;
  if (G % b200::kTileK != 0) return B200_ERR_BAD_SHAPE;      // the decode kernel's group sizes
  if (M == 0) return B200_OK;
  if (!x || !qweight || !scales || !qzeros || !y) return B200_ERR_BAD_ARG;
  if (!aligned16(x) || !aligned16(qweight) || !aligned16(scales) || !aligned16(qzeros) || !aligned16(y))
    return B200_ERR_BAD_ARG;
  const DeviceInfo dev = device_info();
  if (dev.status != B200_OK) return dev.status;
  b200::GemmPlan plan;
  if (b200::make_gemm_plan(M, K, N, G, 0, dev.sm_count, &plan)) return B200_ERR_BAD_SHAPE;
  if (plan.kind != b200::kKindDecode) return B200_ERR_STATIC_LIMIT;
  const size_t part_bytes = align_up(plan.ws_partial_bytes, 256);
  if (part_bytes > (size_t)B200_NOALLOC_SCRATCH_BYTES) return B200_ERR_STATIC_LIMIT;
  int cur = 0;
  cudaError_t ce = cudaGetDevice(&cur);
  if (ce != cudaSuccess) return cuda_fail((int)ce);
  uint8_t* scratch = nullptr;
  int e = noalloc_scratch(cur, &scratch);
  if (e != 0) return cuda_fail(e);
  // The array is shared by every call and keeps what the last one left (a launch that faulted
  // mid-way, another process state after a device reset: zeros): the partial-tile slots must read
  // "unwritten" (all 0xFF) when the kernel starts, so the part this launch uses is cleared first.
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  ce = cudaMemsetAsync(scratch, 0xFF, part_bytes, s);
  if (ce != cudaSuccess) return cuda_fail((int)ce);
  // no programmatic launch: the kernel must not start before the clear has finished
  e = b200::launch_decode(x, qweight, scales, qzeros, y, M, K, N, G, plan, scratch, /*pdl=*/false, s,
                          nullptr, nullptr, 0);
  return e == 0 ? B200_OK : cuda_fail(e);
}

