/*
 * b200_w4a16.h -- C ABI of the B200-native W4A16 group-quantized GEMM library
 * (libb200_w4a16.so, sm_100a only, no CPU fallback).
 *
 * This is the drop-in boundary for the QUICK/AWQ hot path of SqueezeBits/vllm-quick.
 * The reference binds C++ functions taking torch::Tensor directly through pybind11
 * (csrc/pybind.cpp:51-55, csrc/ops.h:68-83); it has no C ABI.  Each entry point below
 * names the reference interface it replaces.  All pointers are DEVICE pointers on the
 * current CUDA device unless stated otherwise; tensors are dense row-major; `half` is
 * IEEE binary16 passed as uint16_t storage.  Nothing here allocates, synchronises or
 * throws: every call validates its arguments, enqueues work on `stream` and returns a
 * b200_status_t.  All entry points are CUDA-graph capturable.
 *
 * Weight layouts (`layout` argument):
 *   B200_LAYOUT_AWQ  (0): qweight[K, N/8] int32, eight 4-bit values per word with logical
 *                         element j in nibble {0,4,1,5,2,6,3,7}[j]  -- the checkpoint
 *                         format the reference consumes (csrc/quantization/awq/dequantize.cuh:17-84).
 *   B200_LAYOUT_B200 (1): the tile-major, K-packed interleave produced by
 *                         b200_w4a16_repack: 128x128 tiles of 2048
 *                         words, word j*512+c*4+w of tile (nt,kb) holding channel nt*128+c,
 *                         rows kb*128+(j*4+w)*8+e, e=0..7 in nibble {0,4,1,5,2,6,3,7}[e].
 * scales[K/G, N] half and qzeros[K/G, N/8] int32 stay in checkpoint format for both.
 */
#ifndef B200_W4A16_H_
#define B200_W4A16_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st *b200_stream_t; /* == cudaStream_t */

typedef enum {
  B200_OK = 0,
  B200_ERR_BAD_SHAPE = 1,  /* shape precondition violated (reference: std::invalid_argument,
                              csrc/quantization/awq/gemm_kernels.cu:412-419) */
  B200_ERR_BAD_ARG = 2,    /* null / misaligned pointer, unknown layout */
  B200_ERR_WORKSPACE = 3,  /* workspace missing, misaligned or smaller than b200_w4a16_workspace_bytes */
  B200_ERR_CUDA = 4,       /* a CUDA runtime / driver call failed (see b200_w4a16_last_cuda_error) */
  B200_ERR_NO_DEVICE = 5,  /* no sm_100 device: the library has no fallback path */
  B200_ERR_STATIC_LIMIT = 6 /* b200_w4a16_gemm_noalloc: the problem needs more scratch than the
                               library's built-in array holds (M > B200_NOALLOC_MAX_TOKENS) */
} b200_status_t;

enum { B200_LAYOUT_AWQ = 0, B200_LAYOUT_B200 = 1 };
enum { B200_REPACK_AWQ_TO_B200 = 0, B200_REPACK_B200_TO_AWQ = 1 };

/* ABI version (major*100 + minor). */
int b200_w4a16_version(void);

/* Static string for a status code. */
const char *b200_w4a16_strerror(int status);

/* cudaError_t of the last failing CUDA call seen by this thread (0 if none). */
int b200_w4a16_last_cuda_error(void);

/* Number of int32 words of the B200-layout tensor for a [K, N/8] AWQ qweight
 * (== K*N/8 when N % 128 == 0; N % 128 == 64 pads the last tile column). */
size_t b200_w4a16_repacked_words(int K, int N);

/* Bytes of scratch b200_w4a16_gemm needs for this problem.  The caller initialises it ONCE
 * with b200_w4a16_workspace_init (outside any CUDA-graph capture); every launch leaves it in
 * that state again.  One workspace may be shared by all GEMMs issued on one stream (size it
 * for the largest), and by every CUDA graph captured on it.  `layout` B200_LAYOUT_AWQ adds
 * room for an on-the-fly repack.
 * Replaces: the internal `torch::empty({split_k, M, N})` scratch of awq_gemm
 * (csrc/quantization/awq/gemm_kernels.cu:401). */
size_t b200_w4a16_workspace_bytes(int M, int K, int N, int layout);

/* Puts a workspace into its idle state: split-K partial tiles cross CTAs as self-flagging fp32
 * words, an unwritten slot reads 0xFFFFFFFF in every word (asynchronous fill on `stream`).
 * A workspace that was never initialised makes the first split GEMM take garbage for
 * partial sums -- hence a call of its own instead of a "zero it yourself" rule. */
int b200_w4a16_workspace_init(void *workspace, size_t workspace_bytes, b200_stream_t stream);

/* y[M,N] = x[M,K] . dequant(qweight, scales, qzeros), fp16 in / fp32 accumulate / fp16 out.
 * Replaces: awq_gemm(in_feats, kernel, scaling_factors, zeros, split_k_iters)
 *           (csrc/ops.h:69-74, csrc/quantization/awq/gemm_kernels.cu:389-446), including
 *           its trailing `.sum(0)` reduction kernel (:445), and the dequantize + torch.matmul
 *           branch of AWQLinearMethod.apply_weights for >= 256 tokens
 *           (vllm/model_executor/layers/quantization/awq.py:154-159).
 * Preconditions (B200_ERR_BAD_SHAPE otherwise): M >= 0, N % 64 == 0, G % 32 == 0,
 * K % G == 0, K % 128 == 0, G in {32, 64} or G % 128 == 0.  M == 0 is a no-op.
 * `split_k_hint` is the reference's split_k_iters: accepted, used only as a tuning hint
 * (<= 0 lets the library choose).
 * Cost of `layout` = B200_LAYOUT_AWQ: the whole weight is interleaved into the workspace on
 * EVERY call (a second kernel that reads and writes K * N / 2 bytes, and no programmatic overlap
 * with the previous launch) -- at decode sizes more than twice the time of a B200_LAYOUT_B200
 * call.  It exists so that a literal awq_gemm binding works; a linear method should repack once
 * after loading (b200_w4a16_repack) and pass layout 1 from then on. */
int b200_w4a16_gemm(const uint16_t *x, const int32_t *qweight, const uint16_t *scales,
                    const int32_t *qzeros, uint16_t *y, int M, int K, int N, int G,
                    int layout, int split_k_hint, void *workspace, size_t workspace_bytes,
                    b200_stream_t stream);

/* b200_w4a16_gemm for ONE routed expert of a mixture-of-experts block.  `gate` (device
 * pointer, may be NULL = always compute) holds the number of tokens of this batch routed to
 * the expert, written by the router earlier on the same stream.  Decode batches (M <= 16):
 * when *gate == 0 the launch streams no weights and writes y = 0 -- the caller multiplies the
 * expert's output by a zero routing weight anyway -- which removes the weight traffic of
 * every expert no token selected (6 of 8 at one token, top-2).  Larger batches ignore the
 * gate and compute.  CUDA-graph capturable (the flag is read on the device).
 * Replaces: the dense "every local expert on every token" loop of MixtralMoE.forward
 *           (vllm/model_executor/models/mixtral_quant.py:143-154), expert by expert. */
int b200_w4a16_gemm_gated(const uint16_t *x, const int32_t *qweight, const uint16_t *scales,
                          const int32_t *qzeros, uint16_t *y, int M, int K, int N, int G,
                          int layout, int split_k_hint, const int32_t *gate, void *workspace,
                          size_t workspace_bytes, b200_stream_t stream);

/* W[K,N] half = fp16((q - z) * s), bit-exact with the reference.
 * Replaces: awq_dequantize(kernel, scaling_factors, zeros, split_k_iters, thx, thy)
 *           (csrc/ops.h:76-82, csrc/quantization/awq/gemm_kernels.cu:274-381); the three
 *           trailing launch-shape arguments of the reference have no meaning here.
 * Preconditions: N % 8 == 0 (AWQ layout) / N % 64 == 0 (B200 layout), K % G == 0,
 * K % 128 == 0 for the B200 layout. */
int b200_w4a16_dequant(const int32_t *qweight, const uint16_t *scales, const int32_t *qzeros,
                       uint16_t *W, int K, int N, int G, int layout, b200_stream_t stream);

/* Layout conversion of qweight between the checkpoint format and the B200 interleave
 * (a bijection on the nibbles; `dst` has b200_w4a16_repacked_words(K,N) words for
 * AWQ_TO_B200 and K*N/8 words for B200_TO_AWQ).  src and dst must not overlap.
 * Replaces: nothing in the mounted checkout (QUICK's offline interleave is not vendored,
 * SURVEY.md F1); precedent for a post-load shuffle is gptq_shuffle
 * (vllm/model_executor/layers/quantization/gptq.py:195-204).
 * Preconditions: K % 128 == 0, N % 64 == 0. */
int b200_w4a16_repack(const int32_t *src, int32_t *dst, int K, int N, int direction,
                      b200_stream_t stream);

/* Introspection for tests and benchmarks: how the library would launch the GEMM.
 * Fills tok_tile (tokens per CTA tile), split_k, grid (CTAs), smem_bytes; returns status. */
int b200_w4a16_gemm_plan(int M, int K, int N, int G, int split_k_hint, int *tok_tile,
                         int *split_k, int *grid, int *smem_bytes);

/* Which kernel variant b200_w4a16_gemm launches for this problem: B200_VARIANT_DECODE
 * (M <= 16 and G % 128 == 0: warp-level mma.sync with the dequantized registers as the A
 * fragment, bandwidth bound) or B200_VARIANT_TCGEN05 (tcgen05 MMA with TMEM operands and
 * accumulators; compute bound from a few hundred tokens up).  Negative: -b200_status_t.  The
 * reference picks between its GEMM kernel and dequantize + cuBLAS on `num_tokens >= 256`
 * (vllm/model_executor/layers/quantization/awq.py:154). */
enum { B200_VARIANT_TCGEN05 = 0, B200_VARIANT_DECODE = 1 };
int b200_w4a16_gemm_variant(int M, int K, int N, int G);

/* Introspection: channel tiles per work unit of the tcgen05 kernel for this problem -- 1; 2 when
 * one activation tile feeds two adjacent 128-channel weight tiles of a CTA; 4 when a CTA pair
 * (tcgen05 cta_group::2) shares the activation tile, two channel tiles per CTA (large problems
 * with 128-token tiles).  Always 1 for B200_VARIANT_DECODE.  Negative: -b200_status_t. */
int b200_w4a16_gemm_unit_tiles(int M, int K, int N, int G);

/* Introspection: 0 = single CTAs; 1 = the tcgen05 kernel runs this problem on CTA pairs (clusters
 * of two CTAs issuing tcgen05.mma.cta_group::2, the activation tile split between the two SMs);
 * 2 = CTA pairs whose work units span two 128-token tiles (one dequantized weight tile feeds both). */
int b200_w4a16_gemm_cta_pair(int M, int K, int N, int G);

/* ---- NVLink peer-memory all-reduce for the row-parallel linears --------------------------
 * Replaces: tensor_model_parallel_all_reduce after RowParallelLinear's GEMM
 * (vllm/model_executor/layers/linear.py:541-542,
 *  vllm/model_executor/parallel_utils/communication_op.py:17-41) for fp16 messages that fit a
 * registered region; the reference's own peer-memory kernel is csrc/custom_all_reduce.cuh:214-232
 * (init / registration: csrc/custom_all_reduce.cu:12-148), disabled in that version for
 * world_size > 1 (vllm/config.py:403-410).  One process per GPU; every rank allocates one
 * region, exports it, imports the peers' and passes the `world` base pointers (index = rank)
 * to every call.  Larger messages go through NCCL in the Python layer. */

/* Bytes of one rank's region: all-reduce messages up to max_message_bytes, plus (when
 * fused_out_features > 0) the packet slots of the fused row-parallel GEMM + all-reduce for
 * outputs of up to fused_out_features channels x 16 tokens from `world` ranks (two parities x
 * world x 64 bytes per channel).  0 on bad arguments. */
size_t b200_w4a16_ar_region_bytes(size_t max_message_bytes, int world, int fused_out_features);
/* cudaMalloc + zero-fill a region on the current device / free it. */
int b200_w4a16_ar_alloc(size_t bytes, void **dev_ptr);
int b200_w4a16_ar_free(void *dev_ptr);
/* CUDA IPC: 64-byte handle of a region from b200_w4a16_ar_alloc / map a peer's handle /
 * unmap it. */
int b200_w4a16_ar_export(const void *dev_ptr, void *handle64);
int b200_w4a16_ar_import(const void *handle64, void **dev_ptr);
int b200_w4a16_ar_unimport(void *dev_ptr);
/* out[n_halfs] = sum over ranks of in[n_halfs] (fp16, fp32 accumulate, rank order: every rank
 * gets bit-identical results).  `regions` is a HOST array of `world` device pointers (own
 * region at index `rank`, peers' mappings elsewhere).  n_halfs % 8 == 0,
 * 2 * n_halfs <= max_message_bytes, in/out 16-byte aligned, in != out allowed to alias
 * nothing else.  Collective: every rank of the group calls it with the same n_halfs, in the
 * same order.  CUDA-graph capturable.  `barrier` must be 1 (0 is the single-GPU emulation used
 * by the tests: peers' slots are pre-filled and no cross-GPU wait happens). */
int b200_w4a16_allreduce(void *const *regions, int rank, int world, const uint16_t *in,
                         uint16_t *out, size_t n_halfs, size_t max_message_bytes, int barrier,
                         b200_stream_t stream);

/* The same all-reduce as reduce-scatter + all-gather over the same regions, for messages above
 * the one-shot sizes (every GPU moves 2 (world - 1) / world x the message over NVLink instead of
 * (world - 1) x): rank r sums chunk range r over all ranks and pushes the result into every
 * rank's slot, a second NVLink barrier, then a local copy.  Same result bits as
 * b200_w4a16_allreduce (rank-order fp32 sums), same regions, epochs and mixing rules, CUDA-graph
 * capturable.  mode: 1 = collective call; 2 / 3 = reduce-scatter / final copy alone, without
 * staging and cross-GPU waits (single-GPU emulation in the tests only).  world >= 2.
 * Replaces: the reference's two-stage kernel (csrc/custom_all_reduce.cuh:239-290, dispatch
 * :520-531) behind tensor_model_parallel_all_reduce (communication_op.py:17-41). */
int b200_w4a16_allreduce_twoshot(void *const *regions, int rank, int world, const uint16_t *in,
                                 uint16_t *out, size_t n_halfs, size_t max_message_bytes,
                                 int mode, b200_stream_t stream);

/* Row-parallel GEMM fused with its all-reduce, decode batches (M <= 16, G % 128 == 0):
 *   y[M,N] = sum over ranks of  x_rank[M,K_rank] . dequant(qweight_rank, ...)      (fp16 out)
 * ONE kernel per rank: a channel tile that finishes its K reduction is rounded to fp16 and
 * pushed over NVLink into every rank's region as 16-byte packets {data, epoch, data, epoch}
 * (peer stores; no fence, no flag: the receiver polls the packets themselves), and each rank
 * sums the arrived copies in rank order with fp32 accumulation -- so the transfer of early
 * tiles overlaps the weight stream of later ones, and the result is bit-identical on every
 * rank and to b200_w4a16_gemm followed by b200_w4a16_allreduce.
 * Replaces: RowParallelLinear.forward's apply_weights + tensor_model_parallel_all_reduce pair
 * (vllm/model_executor/layers/linear.py:538-542).  qweight is in the B200 layout.  `regions`,
 * `max_message_bytes`, `fused_out_features` as given to b200_w4a16_ar_region_bytes when the
 * regions were sized (N <= fused_out_features).  Collective: every rank calls it for the same
 * (M, N), in the same order.  CUDA-graph capturable. */
int b200_w4a16_gemm_allreduce(const uint16_t *x, const int32_t *qweight, const uint16_t *scales,
                              const int32_t *qzeros, uint16_t *y, int M, int K, int N, int G,
                              int split_k_hint, void *workspace, size_t workspace_bytes,
                              void *const *regions, int rank, int world,
                              size_t max_message_bytes, int fused_out_features,
                              b200_stream_t stream);

/* ---- routed mixture-of-experts: token gather / weighted scatter-add --------
 * out[i, :] = x[idx[i], :] for i < rows (idx: int32 device array of token rows).
 * acc[idx[i], :] = half(float(acc[idx[i], :]) + float(half(float(y[i, :]) * w[i]))), w fp32 --
 * the arithmetic of `current_hidden_states.mul_(expert_weights)` followed by
 * `final_hidden_states.add_(current_hidden_states)`.  The rows named by idx must be distinct
 * (a token selects an expert at most once).  hidden % 8 == 0.
 * Replaces: the dense all-tokens expert loop of MixtralMoE.forward
 *           (vllm/model_executor/models/mixtral_quant.py:143-157) together with
 *           b200_w4a16_gemm on the gathered [rows, hidden] batch. */
int b200_w4a16_moe_gather(const uint16_t *x, const int32_t *idx, uint16_t *out, size_t rows,
                          int hidden, b200_stream_t stream);
int b200_w4a16_moe_scatter_add(const uint16_t *y, const int32_t *idx, const float *w,
                               uint16_t *acc, size_t rows, int hidden, b200_stream_t stream);

/* The local experts of a mixture-of-experts block as ONE launch at decode batches (M <= 16,
 * G % 128 == 0, n_experts <= 8, all experts the same K, N, G; weights in the B200 layout):
 *   y[e][M, N'] = x_e[M, K] . dequant(qweight[e], scales[e], qzeros[e])   for every local expert e
 *                 with counts[e] > 0;   y[e] of the others is NOT written.
 * `qweight` / `scales` / `qzeros` are HOST arrays of n_experts device pointers; `counts` is a
 * device int32[n_experts] = tokens routed to each local expert (the router's output: read on
 * the device, no host synchronisation, CUDA-graph capturable).  x_expert_stride = halfs between
 * the experts' activations (0: all experts read the same x -- the merged w1 | w3 projection;
 * M * K: expert e reads x + e * M * K -- the w2 projection behind it); y[e] = y + e * M * N'.
 * act == 0: N' = N.  act == 1: SiLU-and-multiply in the epilogue over channel-paired gate / up
 * weights (see b200_w4a16_gemm_silu_mul), N' = N / 2.  The (expert, channel tile, K block)
 * space is cut into equal ranges over the experts WITH tokens only: every SM streams an equal
 * share of the live weights whatever the routing.  Per expert within the GEMM tolerance of
 * b200_w4a16_gemm / b200_w4a16_gemm_silu_mul (other stream-K cut, other fp32 summation order; the
 * same bits when one expert is live and the grids coincide).
 * Replaces: the per-expert launches of the dense loop in MixtralMoE.forward
 *           (vllm/model_executor/models/mixtral_quant.py:143-157; expert body :84-93). */
int b200_w4a16_moe_grouped_gemm(const uint16_t *x, const int32_t *const *qweight,
                                const uint16_t *const *scales, const int32_t *const *qzeros,
                                uint16_t *y, int M, int K, int N, int G, int n_experts,
                                const int32_t *counts, size_t x_expert_stride, int act,
                                void *workspace, size_t workspace_bytes, b200_stream_t stream);

/* ---- row sets whose size only the device knows (routed experts under CUDA graphs) ----------
 * The tokens routed to an expert are counted on the device; reading the count back would
 * synchronise the host and break graph capture.  These entries take the buffers' row capacity
 * (`max_rows`) from the host and the LIVE row count from `rows_dev` (device int32, clamped to
 * [0, max_rows]); rows beyond it are neither read as indices nor written.
 *
 * b200_w4a16_gemm_rows: y[r, :] = x[r, :] . dequant(W) for r < *rows_dev (B200 layout; x and y
 * have max_rows rows; y rows beyond the count are unspecified).  Above 16 rows the tcgen05
 * kernel derives its token tiles, work items and per-CTA ranges from the live count, so an
 * expert with few tokens streams its weights once per LIVE token tile; up to 16 rows it is
 * b200_w4a16_gemm on all max_rows rows.  0 live rows: nothing is streamed.
 * b200_w4a16_moe_compact: from the router's top-k lists, per local expert e < n_local the tokens
 * that chose expert first_expert + e, ascending, and their routing weights:
 * tok[e * tokens + i], w[e * tokens + i] for i < counts[first_expert + e].
 * Replaces: the argsort + host read of the counts of the routed path above, i.e. the dense loop
 *           of MixtralMoE.forward (vllm/model_executor/models/mixtral_quant.py:143-157). */
int b200_w4a16_gemm_rows(const uint16_t *x, const int32_t *qweight, const uint16_t *scales,
                         const int32_t *qzeros, uint16_t *y, int max_rows, const int32_t *rows_dev,
                         int K, int N, int G, void *workspace, size_t workspace_bytes,
                         b200_stream_t stream);
int b200_w4a16_moe_compact(const int64_t *topk_idx, const float *topk_w, int32_t *tok, float *w,
                           size_t tokens, int top_k, int first_expert, int n_local,
                           b200_stream_t stream);
int b200_w4a16_moe_gather_rows(const uint16_t *x, const int32_t *idx, const int32_t *rows_dev,
                               uint16_t *out, size_t max_rows, int hidden, b200_stream_t stream);
int b200_w4a16_moe_scatter_add_rows(const uint16_t *y, const int32_t *idx, const float *w,
                                    const int32_t *rows_dev, uint16_t *acc, size_t max_rows,
                                    int hidden, b200_stream_t stream);
int b200_w4a16_silu_and_mul_rows(const uint16_t *in, uint16_t *out, size_t max_rows,
                                 const int32_t *rows_dev, int d, int paired, b200_stream_t stream);

/* The router of a mixture-of-experts block as one launch (plus a memset node for `counts`):
 *   logits = half(x[tokens, hidden] . gate_w[n_experts, hidden]^T)     fp32 accumulate, fp16 result
 *   p = softmax(float(logits));  (topk_w, topk_idx) = top_k largest of p, descending;
 *   topk_w /= sum(topk_w);  dense_w[t, e] = topk_w of e for token t, 0 where e was not chosen;
 *   counts[e] = tokens that chose e.
 * topk_w fp32 [tokens, top_k], topk_idx int64 [tokens, top_k], dense_w fp32 [tokens, n_experts],
 * counts int32 [n_experts]; hidden % 8 == 0, n_experts <= 32, top_k <= n_experts.
 * Replaces: self.gate(hidden_states), F.softmax(..., dtype=torch.float), torch.topk and the
 *           renormalisation of MixtralMoE.forward (vllm/model_executor/models/mixtral_quant.py:133-142)
 *           plus the token counting the expert gates need. */
int b200_w4a16_moe_route(const uint16_t *x, const uint16_t *gate_w, float *topk_w,
                         int64_t *topk_idx, float *dense_w, int32_t *counts, size_t tokens,
                         int hidden, int n_experts, int top_k, b200_stream_t stream);

/* out[t, :] = sum over e < n_experts with counts[e] > 0, in index order, of
 *   half(float(out[t, :]) + float(half(float(y[e][t, :]) * w[t * w_stride + e])))   (out starts 0)
 * -- the `mul_(expert_weights)` / `add_` pair of mixtral_quant.py:152-157 for all local experts
 * in one pass over y[n_experts][rows, hidden] (the grouped launch's output).  w fp32. */
int b200_w4a16_moe_combine(const uint16_t *y, const float *w, const int32_t *counts,
                           uint16_t *out, int n_experts, int w_stride, size_t rows, int hidden,
                           b200_stream_t stream);

/* ---- SiLU-and-multiply behind the gate_up projection ----------------------
 * out[rows, d] = half(silu(gate)) * up, silu(x) = half(float(x) / (1 + expf(-float(x)))), the
 * reference's arithmetic.  paired == 0: in[rows, 2 d] = [gate | up] halves, the reference's
 * layout.  paired == 1: every 16 input channels are 8 gate channels followed by the 8 up
 * channels of the same indices (channel-paired gate_up weight, see below).  d % 8 == 0.
 * Replaces: silu_and_mul(out, input)   (csrc/activation_kernels.cu:17-52; SiluAndMul.forward,
 *           vllm/model_executor/layers/activation.py:14-38; llama.py:76-80). */
int b200_w4a16_silu_and_mul(const uint16_t *in, uint16_t *out, size_t rows, int d, int paired,
                            b200_stream_t stream);

/* gate_up GEMM with the SiLU-and-multiply in its epilogue, decode batches (M <= 16,
 * G % 128 == 0):  y[M, N/2] = silu_and_mul(x[M,K] . dequant(W[K,N])) without the [M, N]
 * round trip and the second launch.  W (B200 layout), scales and qzeros hold the merged
 * gate_up projection with its output channels PAIRED: channel block 16 j..16 j + 15 = gate
 * channels 8 j..8 j + 7, then up channels 8 j..8 j + 7 -- a permutation of whole packed words
 * done once at load time (MergedColumnParallelLinear.pair_gate_up, before the repack).  The
 * two then sit in one thread's accumulator fragment.  Bit-identical to b200_w4a16_gemm on the
 * same weights followed by b200_w4a16_silu_and_mul(paired = 1).  N % 16 == 0.
 * Replaces: gate_up_proj + act_fn of LlamaMLP.forward (vllm/model_executor/models/llama.py:76-80). */
int b200_w4a16_gemm_silu_mul(const uint16_t *x, const int32_t *qweight, const uint16_t *scales,
                             const int32_t *qzeros, uint16_t *y, int M, int K, int N, int G,
                             int split_k_hint, void *workspace, size_t workspace_bytes,
                             b200_stream_t stream);

/* The same behind a routed-expert gate (see b200_w4a16_gemm_gated): `gate` points at an int32
 * on the device; 0 = no token of the batch chose this expert -- nothing is streamed and
 * y[M, N/2] is zero-filled.  NULL = no gate.  Used for the merged w1 / w3 projection of a
 * Mixtral expert (vllm/model_executor/models/mixtral_quant.py:87-93). */
int b200_w4a16_gemm_silu_mul_gated(const uint16_t *x, const int32_t *qweight,
                                   const uint16_t *scales, const int32_t *qzeros, uint16_t *y,
                                   int M, int K, int N, int G, int split_k_hint,
                                   const int32_t *gate, void *workspace, size_t workspace_bytes,
                                   b200_stream_t stream);

/* ---- residual add + RMSNorm around the row-parallel linears ---------------
 * out[rows, hidden] = half(float(half(x)) * rsqrt(mean(x^2) + eps)) * weight,
 *   x = float(in) + float(residual_in)  (residual_in == NULL: x = in),
 *   residual_out = half(x)              (residual_out == NULL: not written),
 * fp16 storage, fp32 statistics taken from x BEFORE it is rounded -- the reference's
 * arithmetic.  out may alias in, residual_out may alias residual_in (the reference works in
 * place).  hidden % 8 == 0, hidden <= 16384; rows == 0 is a no-op.
 * Replaces: rms_norm(out, input, weight, eps)            (csrc/ops.h, csrc/layernorm_kernels.cu:78-102)
 *           fused_add_rms_norm(input, residual, weight, eps)      (csrc/layernorm_kernels.cu:104-129)
 *           as called by RMSNorm.forward (vllm/model_executor/layers/layernorm.py:46-66). */
int b200_w4a16_add_rms_norm(const uint16_t *in, const uint16_t *residual_in, const uint16_t *weight,
                      uint16_t *out, uint16_t *residual_out, float eps, size_t rows, int hidden,
                      b200_stream_t stream);

/* The same with x = all-reduce(in) over the ranks (rounded to fp16 exactly as
 * b200_w4a16_allreduce rounds it): ONE kernel for the "row-parallel all-reduce, then
 * fused_add_rms_norm" pair of a decoder layer -- the reduced rows never travel through HBM
 * between the two and the second launch disappears.  Result is bit-identical on every rank
 * and to b200_w4a16_allreduce followed by b200_w4a16_add_rms_norm.  `in` is this rank's partial sum
 * (RowParallelLinear with reduce_results=False); rows * hidden * 2 <= max_message_bytes.
 * Collective; shares regions, flags and epochs with b200_w4a16_allreduce (calls of both kinds
 * may be mixed on one stream).  CUDA-graph capturable.  `barrier` as in b200_w4a16_allreduce.
 * Replaces: tensor_model_parallel_all_reduce (vllm/model_executor/layers/linear.py:541-542)
 *           + post_attention_layernorm / next layer's input_layernorm
 *           (vllm/model_executor/models/llama.py:205-222). */
int b200_w4a16_allreduce_add_rms_norm(void *const *regions, int rank, int world, const uint16_t *in,
                                const uint16_t *residual_in, const uint16_t *weight,
                                uint16_t *out, uint16_t *residual_out, float eps, size_t rows,
                                int hidden, size_t max_message_bytes, int barrier,
                                b200_stream_t stream);

/* ---- flag-in-data ("packet") variants of the two collectives above ---------------------------
 * Instead of stage / system fence / flag / barrier / remote read, every 16-byte chunk is
 * PUSHED to each peer as two 16-byte packets {data, epoch, data, epoch} and the receiver spins
 * on the packets themselves -- one NVLink one-way trip on the critical path, no fence, no
 * barrier.  The packet area is `4 * world * ll_message_bytes` bytes at byte `ll_offset` (a
 * multiple of 256, behind everything b200_w4a16_ar_region_bytes covers) of every rank's
 * region: allocate `ll_offset + 4 * world * ll_message_bytes` with b200_w4a16_ar_alloc.  The
 * message must fit ll_message_bytes.  Results are bit-identical to the barrier variants; the
 * epochs are shared with them (calls of all kinds may be mixed on one stream); CUDA-graph
 * capturable.  `barrier` 0 = single-GPU emulation (a missing packet traps instead of being
 * waited for).  world >= 2.
 * Replaces: the same reference sites as b200_w4a16_allreduce (communication_op.py:17-41) and
 * b200_w4a16_allreduce_add_rms_norm (linear.py:541-542 + llama.py:205-222). */

/* out[n_halfs] = sum over ranks of in[n_halfs]; n_halfs % 8 == 0, 2 * n_halfs <= ll_message_bytes. */
int b200_w4a16_allreduce_ll(void *const *regions, int rank, int world, const uint16_t *in,
                            uint16_t *out, size_t n_halfs, size_t ll_offset,
                            size_t ll_message_bytes, int barrier, b200_stream_t stream);

/* all-reduce + residual add + RMSNorm; rows * hidden * 2 <= ll_message_bytes. */
int b200_w4a16_allreduce_add_rms_norm_ll(void *const *regions, int rank, int world,
                                         const uint16_t *in, const uint16_t *residual_in,
                                         const uint16_t *weight, uint16_t *out,
                                         uint16_t *residual_out, float eps, size_t rows,
                                         int hidden, size_t ll_offset, size_t ll_message_bytes,
                                         int barrier, b200_stream_t stream);

/* ---- GEMM without any caller- or runtime-provided scratch ------------------------------------
 * No counterpart in the reference: added for callers that
 * cannot allocate (code under CUDA-graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).
 *   y[M,N] = x[M,K] . dequant(qweight, scales, qzeros)       qweight in the B200 layout
 * is b200_w4a16_gemm's decode-batch path (M <= B200_NOALLOC_MAX_TOKENS, G % 128 == 0; same
 * kernel, same plan, bit-identical result) with the stream-K partial-tile scratch in an array built
 * into the library: B200_NOALLOC_SCRATCH_BYTES of device memory per GPU that loads the library,
 * held until the process exits.  The call allocates nothing (no cudaMalloc / cudaFree, no
 * allocating library routine): it clears the part of the array the launch uses (asynchronously,
 * on `stream`) and launches the kernel.
 * M > B200_NOALLOC_MAX_TOKENS returns B200_ERR_STATIC_LIMIT (use b200_w4a16_gemm with a
 * workspace); G % 128 != 0 returns B200_ERR_BAD_SHAPE; other preconditions as b200_w4a16_gemm.
 * EVERY call on a GPU shares the one array: calls that may overlap on the GPU (different
 * streams, or graph branches, without an ordering between them) race on it and give wrong
 * results; calls ordered on one stream are safe. */
#define B200_NOALLOC_MAX_TOKENS 16
#define B200_NOALLOC_SCRATCH_BYTES (148u * 2u * 2048u * 4u) /* 2 424 832: [CTA][2 slots][16 x 128 fp32] */
int b200_w4a16_gemm_noalloc(const uint16_t *x, const int32_t *qweight, const uint16_t *scales,
                            const int32_t *qzeros, uint16_t *y, int M, int K, int N, int G,
                            b200_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_W4A16_H_ */
