// w4a16_internal.h -- launcher prototypes shared by the translation units of libb200_w4a16.so.
// Not part of the public ABI (that is include/b200_w4a16.h).
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace b200 {

constexpr int kTileN = 128;        // output channels per weight tile  (tcgen05 M)
constexpr int kTileK = 128;        // input rows per weight tile       (8 x UMMA_K)
constexpr int kTileWords = 2048;   // int32 words per packed tile (8 KiB)
constexpr int kMaxCtas = 148;      // one CTA per SM on B200
constexpr int kDecodeCtasPerSm = 1;    // decode kernel residency (w4a16_decode.cu)
constexpr int kDecodeMaxTokens = 16;   // mma.sync decode kernel (w4a16_decode.cu) up to here
constexpr int kMaxDevices = 64;        // per-device launch state (shared-memory opt-in)

enum GemmKind { kKindTcgen05 = 0, kKindDecode = 1 };

// How a GEMM is cut into CTAs (stream-K over the (channel tile, token tile, K block) space).
struct GemmPlan {
  int kind;          // GemmKind: tcgen05 kernel or the mma.sync decode kernel
  int tok_tile;      // tokens per MMA tile (tcgen05 N): 16 / 32 / 64 / 128
  int n_tiles;       // ceil(N / 128)
  int pair;          // tcgen05 kernel: channel tiles per CTA and work unit (1 or 2, sharing one activation tile)
  int two_cta;       // tcgen05 kernel: CTA pairs (cta_group::2), the activation tile split between two SMs
  int tok_pair;      // tcgen05 kernel: token tiles per work unit (2: one dequantized weight tile feeds two)
  int m_tiles;       // ceil(M / tok_tile)
  int k_blocks;      // K / 128
  int ctas;          // grid size
  int stages;        // smem pipeline depth
  int smem_bytes;    // dynamic shared memory per CTA
  int max_split;     // most CTAs contributing to one output tile
  size_t ws_partial_bytes;  // fp32 partial-tile scratch (self-flagging words: all 0xFF on entry and exit)
};

int make_gemm_plan(int M, int K, int N, int G, int split_k_hint, int sm_count, GemmPlan* plan);

// y[M,N] = x[M,K] . dequant(W) with W in the B200 layout.  Returns cudaError_t.
int launch_gemm(const uint16_t* x, const int32_t* qw_b200, const uint16_t* scales,
                const int32_t* qzeros, uint16_t* y, int M, int K, int N, int G,
                const GemmPlan& plan, void* ws_partials, bool pdl, cudaStream_t stream,
                const int* m_dev = nullptr);
// The same with the live row count (<= M, the buffers' rows) read on the device: plan with
// make_gemm_plan_rows, pass the count's address as m_dev (tcgen05 kernel, single-tile variants).
int make_gemm_plan_rows(int M_max, int K, int N, int G, int sm_count, GemmPlan* plan);

// Row-parallel fusion of the decode kernel with the all-reduce (w4a16_decode.cu): where in the
// ranks' peer-mapped regions the fused path keeps its epoch counter, per-tile flags and slots.
constexpr int kArMaxRanks = 8;
constexpr int kArMaxTiles = 128;       // channel tiles of a fused GEMM: N <= 16384
struct FusedArParams {
  uint8_t* region[kArMaxRanks];        // index = rank; own region at [rank]
  int rank, world;                     // world <= 1: not fused
  uint32_t ctr_off;                    // two u32: epoch of the last fused call, CTA done counter
  uint32_t data_off;                   // packet slots [2 parities][world source ranks][slot_bytes]
  uint32_t slot_bytes;                 // one rank's output as {data, epoch, data, epoch} packets: 64 B / channel
};

// Decode variant (M <= 16, G % 128 == 0): same layout, workspace protocol and semantics.
void fill_decode_plan(int M, GemmPlan* plan);
int launch_decode(const uint16_t* x, const int32_t* qw_b200, const uint16_t* scales,
                  const int32_t* qzeros, uint16_t* y, int M, int K, int N, int G,
                  const GemmPlan& plan, void* ws_partials, bool pdl, cudaStream_t stream,
                  const FusedArParams* ar = nullptr, const int* gate = nullptr, int act = 0);

// The local experts of a mixture-of-experts block as ONE decode-kernel launch (w4a16_decode.cu,
// GROUPED instantiation): same shapes, one set of weight pointers per expert.  counts[e] (device)
// = tokens routed to local expert e; the (expert, channel tile, K block) space is cut over the
// experts WITH tokens only, so every SM streams an equal share whatever the routing.
constexpr int kMaxGroupExperts = 8;
struct ExpertGroup {
  const int32_t* qw[kMaxGroupExperts];      // B200 layout
  const uint16_t* scales[kMaxGroupExperts];
  const int32_t* qz[kMaxGroupExperts];
  const int32_t* counts;                    // [n_experts], device
  int n_experts;
  long long x_stride;                       // halfs between the experts' activations (0: shared x)
  long long y_stride;                       // halfs between the experts' outputs
};
int launch_decode_grouped(const uint16_t* x, const ExpertGroup& grp, uint16_t* y, int M, int K,
                          int N, int G, const GemmPlan& plan, void* ws_partials, bool pdl,
                          cudaStream_t stream, int act);
// final[t] = sum over local experts e with counts[e] > 0, in index order, of
// half(final[t] + half(y[e][t] * w[t][first + e])) -- the reference's mul_ / add_ arithmetic.
int launch_moe_combine(const uint16_t* y, const float* w, const int32_t* counts, uint16_t* out,
                       int n_experts, int w_stride, size_t rows, int hidden, cudaStream_t stream);

// Router: fp16 gate GEMM -> fp32 softmax -> top-k -> renormalise, a warp per token.
constexpr int kMaxRouteExperts = 32;
int launch_moe_route(const uint16_t* x, const uint16_t* gate_w, float* topk_w, long long* topk_idx,
                     float* dense_w, int32_t* counts, size_t tokens, int hidden, int n_experts,
                     int top_k, cudaStream_t stream);

// silu(x[:, :d]) * x[:, d:] (paired == 0: the reference's [gate | up] halves) or, paired == 1,
// the channel-paired order the fused gate_up GEMM uses (every 16 channels = 8 gate + 8 up).
int launch_silu_and_mul(const uint16_t* in, uint16_t* out, size_t rows, int d, int paired,
                        cudaStream_t stream, const int32_t* rows_dev = nullptr);

// NVLink peer-memory all-reduce (w4a16_allreduce.cu).  Region = header + two data slots.
// Region layout: [0, 2 KiB) flags[block][rank] | [2 KiB, +256) per-block epochs |
// [3 KiB, +8) fused epoch + done counter | [4 KiB, +2 KiB) flags2[block][rank] |
// [16 KiB ...) two all-reduce slots, then
// (fused_bytes > 0) 2 x world fused packet slots.
constexpr int kArMaxBlocks = 64;
constexpr int kArFlagBytes = kArMaxBlocks * kArMaxRanks * 4;   // flags[block][rank]
constexpr int kArFusedCtrOff = 3072;
constexpr int kArFlags2Off = 4096;                             // second barrier of the two-shot kernel: flags2[block][rank]
constexpr int kArHeaderBytes = 16384;
static_assert(kArFlagBytes + kArMaxBlocks * 4 <= kArFusedCtrOff, "all-reduce header layout");
static_assert(kArFusedCtrOff + 8 <= kArFlags2Off && kArFlags2Off + kArFlagBytes <= kArHeaderBytes, "all-reduce header layout");
size_t ar_region_bytes(size_t max_message_bytes);
int launch_allreduce(void* const* regions, int rank, int world, const uint16_t* in, uint16_t* out,
                     size_t n_halfs, size_t max_message_bytes, int barrier, cudaStream_t stream,
                     size_t ll_offset = 0, size_t ll_message_bytes = 0);

// Reduce-scatter + all-gather over the same regions (messages above the one-shot sizes).
// mode 1 = collective; 2 / 3 = the two halves without barriers (single-GPU emulation).
int launch_allreduce_twoshot(void* const* regions, int rank, int world, const uint16_t* in,
                             uint16_t* out, size_t n_halfs, size_t max_message_bytes, int mode,
                             cudaStream_t stream);

// Residual add + RMSNorm, standalone and fused behind the one-shot all-reduce (w4a16_norm.cu).
// hidden % 8 == 0, hidden <= norm_max_hidden(); residual_in / residual_out may be nullptr.
int norm_max_hidden();
int launch_add_rmsnorm(const uint16_t* in, const uint16_t* residual_in, const uint16_t* weight,
                       uint16_t* out, uint16_t* residual_out, float eps, size_t rows, int hidden,
                       cudaStream_t stream);
int launch_allreduce_add_rmsnorm(void* const* regions, int rank, int world, const uint16_t* in,
                                 const uint16_t* residual_in, const uint16_t* weight,
                                 uint16_t* out, uint16_t* residual_out, float eps, size_t rows,
                                 int hidden, size_t max_message_bytes, int barrier,
                                 cudaStream_t stream, size_t ll_offset = 0,
                                 size_t ll_message_bytes = 0);

// Routed mixture-of-experts plumbing (w4a16_moe.cu): out[i] = x[idx[i]];
// acc[idx[i]] = half(acc[idx[i]] + half(y[i] * w[i])).  hidden % 8 == 0, idx rows distinct.
// rows_dev (optional, device): the live row count, <= rows -- for row sets only the device knows.
int launch_moe_gather(const uint16_t* x, const int32_t* idx, uint16_t* out, size_t rows, int hidden,
                      cudaStream_t stream, const int32_t* rows_dev = nullptr);
int launch_moe_scatter_add(const uint16_t* y, const int32_t* idx, const float* w, uint16_t* acc,
                           size_t rows, int hidden, cudaStream_t stream, const int32_t* rows_dev = nullptr);
// tok[e][0 .. counts[first + e]) = tokens that chose local expert e (ascending), w = their weights.
int launch_moe_compact(const long long* topk_idx, const float* topk_w, int32_t* tok, float* w,
                       size_t tokens, int top_k, int first, int n_local, cudaStream_t stream);

// B200_TRACE builds: device buffer of 148 x 8 x 64 uint32 timestamps (nullptr = off).
void set_trace_buffer(uint32_t* dev);
uint32_t* trace_buffer();

int launch_dequant(const int32_t* qweight, const uint16_t* scales, const int32_t* qzeros,
                   uint16_t* W, int K, int N, int G, int layout, cudaStream_t stream);

int launch_repack(const int32_t* src, int32_t* dst, int K, int N, int direction,
                  cudaStream_t stream);

}  // namespace b200
