/*
 * oracle/awq_oracle.c -- CPU restatement of the reference's W4A16 (AWQ) hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (vllm-quick_b200/) may
 * import, link or execute this file; only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs use it, and only as the checker.
 *
 * PARITY PINNING: the reference (stock vLLM v0.3.0 checkout at /root/reference) has
 * no test, golden vector or fixture for this path and no CPU
 * code path (its arithmetic is __global__ kernels with inline PTX).  This oracle is
 * pinned two ways instead:
 *   (1) a bit-level emulation of the reference's PTX sequence (lop3/sub/fma on fp16
 *       pairs) is compared against the arithmetic restatement for all 2^16 low/high
 *       half-word patterns (tests/test_oracle.py);
 *   (2) outputs of the reference's own kernels, compiled unmodified for sm_100a into
 *       oracle/_ref/ and run on a B200, are committed under tests/golden/ and both
 *       this oracle and the CUDA path are checked against them.
 * Until (2) has been generated the header of tests/golden/README says "parity unpinned".
 *
 * Every function cites the reference file:line it follows (paths relative to
 * /root/reference).
 *
 * Build: see oracle/Makefile (gcc -O2 -fopenmp -shared -fPIC).
 */
#include <stdint.h>
#include <stddef.h>
#include <string.h>

/* ---------------------------------------------------------------- fp16 helpers */

static inline float h2f(uint16_t b)
{
    _Float16 h;
    memcpy(&h, &b, 2);
    return (float)h;
}

static inline uint16_t f2h(float f)
{
    _Float16 h = (_Float16)f; /* round-to-nearest-even, subnormals kept */
    uint16_t b;
    memcpy(&b, &h, 2);
    return b;
}

/* fp16 fused multiply-add with a single rounding (PTX fma.rn.f16).  For fp16
 * inputs the product has <= 22 significant bits and is exact in double; the sum
 * with a third fp16 is exact in double too (exponent span < 53 bits for all
 * finite fp16), so one double->fp16 rounding reproduces the hardware result. */
static inline uint16_t hfma(uint16_t a, uint16_t b, uint16_t c)
{
    double r = (double)h2f(a) * (double)h2f(b) + (double)h2f(c);
    _Float16 h = (_Float16)r;
    uint16_t o;
    memcpy(&o, &h, 2);
    return o;
}

static inline uint16_t hsub(uint16_t a, uint16_t b)
{
    double r = (double)h2f(a) - (double)h2f(b);
    _Float16 h = (_Float16)r;
    uint16_t o;
    memcpy(&o, &h, 2);
    return o;
}

/* ---------------------------------------------------------------- a1: unpack */

/* Logical element j of a packed word lives in nibble AWQ_ORDER[j].
 * Follows csrc/quantization/awq/dequantize.cuh:17-84: the four lop3 extractions
 * take nibbles (0,4), (1,5), (2,6), (3,7) and the uint4 result is read back as
 * eight consecutive halves. */
static const int AWQ_ORDER[8] = {0, 4, 1, 5, 2, 6, 3, 7};

void oracle_unpack_word(uint32_t w, uint8_t out[8])
{
    for (int j = 0; j < 8; ++j)
        out[j] = (uint8_t)((w >> (4 * AWQ_ORDER[j])) & 0xF);
}

uint32_t oracle_pack_word(const uint8_t in[8])
{
    uint32_t w = 0;
    for (int j = 0; j < 8; ++j)
        w |= ((uint32_t)(in[j] & 0xF)) << (4 * AWQ_ORDER[j]);
    return w;
}

/* Bit-level emulation of dequantize_s4_to_fp16x2 (dequantize.cuh:17-84):
 *   lop3 (w & 0x000f000f) | 0x64006400      :42-44   -> halves 1024+q
 *   lop3 (w & 0x00f000f0) | 0x64006400      :46-48   -> halves 1024+16q
 *   same two on (w >> 8)                    :40,50-56
 *   sub.f16x2  h, 0x64006400                :74,78
 *   fma.rn.f16x2 h, 0x2c00(1/16), 0xd400(-64) :76,80
 * out[] holds the eight fp16 bit patterns in the order the uint4 is stored. */
void oracle_dequant_word_lop3(uint32_t w, uint16_t out[8])
{
    const uint32_t BOTTOM = 0x000f000fu, TOP = 0x00f000f0u, MAGIC = 0x64006400u;
    uint32_t top = w >> 8;
    uint32_t h[4];
    h[0] = (w & BOTTOM) | MAGIC;
    h[1] = (w & TOP) | MAGIC;
    h[2] = (top & BOTTOM) | MAGIC;
    h[3] = (top & TOP) | MAGIC;
    for (int i = 0; i < 4; ++i) {
        uint16_t lo = (uint16_t)(h[i] & 0xFFFF), hi = (uint16_t)(h[i] >> 16);
        if ((i & 1) == 0) {
            lo = hsub(lo, 0x6400);
            hi = hsub(hi, 0x6400);
        } else {
            lo = hfma(lo, 0x2c00, 0xd400);
            hi = hfma(hi, 0x2c00, 0xd400);
        }
        out[2 * i] = lo;
        out[2 * i + 1] = hi;
    }
}

/* ---------------------------------------------------------------- a2/a5: dequant */

/* W[k,n] = fp16_rn( (q - z) * s + 0 ), one rounding.
 * Element pairing follows gemm_kernels.cu:128-130,137,146 and :296-311:
 *   q from qweight[k, n/8], z from qzeros[k/G, n/8], s from scales[k/G, n].
 * Arithmetic follows gemm_kernels.cu:153-160 / :313-320: sub.f16x2 (exact, both
 * operands are small integers) then fma.rn.f16x2 with a +0 addend; the +0 addend
 * matters only for the sign of an exact zero product (-0 + +0 = +0). */
void oracle_awq_dequantize(const int32_t *qweight, const uint16_t *scales,
                           const int32_t *qzeros, uint16_t *W, int K, int N, int G)
{
    const int NW = N / 8;
#pragma omp parallel for schedule(static)
    for (int k = 0; k < K; ++k) {
        const int g = k / G;
        for (int nw = 0; nw < NW; ++nw) {
            uint8_t q[8], z[8];
            oracle_unpack_word((uint32_t)qweight[(size_t)k * NW + nw], q);
            oracle_unpack_word((uint32_t)qzeros[(size_t)g * NW + nw], z);
            for (int j = 0; j < 8; ++j) {
                float d = (float)((int)q[j] - (int)z[j]);
                float s = h2f(scales[(size_t)g * N + nw * 8 + j]);
                float p = d * s; /* exact: 5-bit int x 11-bit mantissa */
                p = p + 0.0f;    /* fma addend +0: turns -0 into +0 */
                W[(size_t)k * N + nw * 8 + j] = f2h(p);
            }
        }
    }
}

/* Same result, computed through the PTX emulation end to end (slow; small cases). */
void oracle_awq_dequantize_lop3(const int32_t *qweight, const uint16_t *scales,
                                const int32_t *qzeros, uint16_t *W, int K, int N, int G)
{
    const int NW = N / 8;
    for (int k = 0; k < K; ++k) {
        const int g = k / G;
        for (int nw = 0; nw < NW; ++nw) {
            uint16_t q[8], z[8];
            oracle_dequant_word_lop3((uint32_t)qweight[(size_t)k * NW + nw], q);
            oracle_dequant_word_lop3((uint32_t)qzeros[(size_t)g * NW + nw], z);
            for (int j = 0; j < 8; ++j) {
                uint16_t d = hsub(q[j], z[j]);
                W[(size_t)k * N + nw * 8 + j] =
                    hfma(d, scales[(size_t)g * N + nw * 8 + j], 0x0000);
            }
        }
    }
}

/* ---------------------------------------------------------------- a3/a4: GEMM */

/* Y[m,n] = sum_k fp32(x[m,k]) * fp32(W[k,n]) in fp32, k ascending.  This is the
 * "fp32 dequant-then-matmul oracle" of BASELINE.json north_star; Yf keeps fp32,
 * Yh (optional) is its fp16 rounding. */
void oracle_gemm_f32(const uint16_t *x, const uint16_t *W, float *Yf, uint16_t *Yh,
                     int M, int K, int N)
{
#pragma omp parallel for schedule(static)
    for (int m = 0; m < M; ++m) {
        float *acc = Yf + (size_t)m * N;
        for (int n = 0; n < N; ++n) acc[n] = 0.0f;
        for (int k = 0; k < K; ++k) {
            const float xv = h2f(x[(size_t)m * K + k]);
            const uint16_t *wr = W + (size_t)k * N;
            for (int n = 0; n < N; ++n) acc[n] += xv * h2f(wr[n]);
        }
        if (Yh)
            for (int n = 0; n < N; ++n) Yh[(size_t)m * N + n] = f2h(acc[n]);
    }
}

/* Emulates the reference's split-K numerics (gemm_kernels.cu:112-115 chunk
 * interleave k = (i*split + z)*32, fp32 accumulate per slice :242-253, fp16
 * partial store :262-270, then Tensor.sum(0) :445 which accumulates in fp32 and
 * rounds once).  Used to report how far the reference itself is from the fp32
 * oracle; accumulation order inside a slice is k ascending (the tensor-core
 * order is not specified, so this is a model, not a bit-exact replay). */
void oracle_gemm_ref_splitk(const uint16_t *x, const uint16_t *W, uint16_t *Yh,
                            int M, int K, int N, int split_k)
{
    const int chunks = K / 32;
#pragma omp parallel for schedule(static)
    for (int m = 0; m < M; ++m) {
        for (int n = 0; n < N; ++n) {
            float total = 0.0f;
            for (int z = 0; z < split_k; ++z) {
                float acc = 0.0f;
                for (int c = z; c < chunks; c += split_k)
                    for (int kk = 0; kk < 32; ++kk) {
                        int k = c * 32 + kk;
                        acc += h2f(x[(size_t)m * K + k]) * h2f(W[(size_t)k * N + n]);
                    }
                total += h2f(f2h(acc));
            }
            Yh[(size_t)m * N + n] = f2h(total);
        }
    }
}

/* ---------------------------------------------------------------- packing */

/* q[K,N] (one nibble value per byte) -> qweight[K,N/8] in the AWQ order.  The
 * quantiser side lives in AutoAWQ (not vendored); the reference only holds the
 * inverse (dequantize.cuh), which fixes the order uniquely. */
void oracle_awq_pack(const uint8_t *q, int32_t *qweight, int K, int N)
{
    const int NW = N / 8;
    for (size_t i = 0; i < (size_t)K * NW; ++i)
        qweight[i] = (int32_t)oracle_pack_word(q + i * 8);
}

void oracle_awq_unpack(const int32_t *qweight, uint8_t *q, int K, int N)
{
    const int NW = N / 8;
    for (size_t i = 0; i < (size_t)K * NW; ++i)
        oracle_unpack_word((uint32_t)qweight[i], q + i * 8);
}

/* ---------------------------------------------------------------- B200 ("quick") layout
 *
 * The mounted reference does not contain QUICK's layout; the build
 * defines its own tile-major, K-packed interleave:
 *   tile (nt, kb) = 128 output channels x 128 input rows = 2048 words, stored
 *   contiguously at word offset ((nt * KB) + kb) * 2048, KB = K/128;
 *   inside a tile, word index = j*512 + c*4 + w  (j 0..3, c 0..127, w 0..3) holds
 *   channel n = nt*128 + c and the eight rows k = kb*128 + (j*4+w)*8 + e, e = 0..7,
 *   with row e in nibble AWQ_ORDER[e] (so the same lop3 sequence yields the
 *   (k, k+1) half2 pairs tcgen05 wants in one TMEM column).
 * Channels >= N (when N % 128 == 64) are stored as zero nibbles.
 */
size_t oracle_b200_words(int K, int N)
{
    return (size_t)((N + 127) / 128) * (size_t)(K / 128) * 2048u;
}

void oracle_b200_repack(const int32_t *qweight, int32_t *out, int K, int N)
{
    const int NW = N / 8, KB = K / 128, NT = (N + 127) / 128;
    for (int nt = 0; nt < NT; ++nt)
        for (int kb = 0; kb < KB; ++kb) {
            int32_t *tile = out + ((size_t)nt * KB + kb) * 2048;
            for (int j = 0; j < 4; ++j)
                for (int c = 0; c < 128; ++c)
                    for (int w = 0; w < 4; ++w) {
                        const int n = nt * 128 + c;
                        uint32_t word = 0;
                        if (n < N)
                            for (int e = 0; e < 8; ++e) {
                                const int k = kb * 128 + (j * 4 + w) * 8 + e;
                                uint32_t src = (uint32_t)qweight[(size_t)k * NW + n / 8];
                                uint32_t nib = (src >> (4 * AWQ_ORDER[n % 8])) & 0xF;
                                word |= nib << (4 * AWQ_ORDER[e]);
                            }
                        tile[j * 512 + c * 4 + w] = (int32_t)word;
                    }
        }
}

void oracle_b200_unrepack(const int32_t *in, int32_t *qweight, int K, int N)
{
    const int NW = N / 8, KB = K / 128;
    memset(qweight, 0, (size_t)K * NW * sizeof(int32_t));
    for (int k = 0; k < K; ++k)
        for (int n = 0; n < N; ++n) {
            const int nt = n / 128, c = n % 128, kb = k / 128, o = (k % 128) / 8, e = k % 8;
            const int32_t *tile = in + ((size_t)nt * KB + kb) * 2048;
            uint32_t word = (uint32_t)tile[(o / 4) * 512 + c * 4 + (o % 4)];
            uint32_t nib = (word >> (4 * AWQ_ORDER[e])) & 0xF;
            uint32_t *dst = (uint32_t *)&qweight[(size_t)k * NW + n / 8];
            *dst |= nib << (4 * AWQ_ORDER[n % 8]);
        }
}
