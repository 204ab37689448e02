// w4a16_streamk.cuh -- stream-K work decomposition shared by the W4A16 GEMM kernels.
// The (channel tile, token tile, K block) space is cut into P equal contiguous ranges; a CTA
// walks its range partial segments first so the cross-CTA reduction overlaps the remaining
// weight stream.
#pragma once

namespace b200 {

#ifndef B200_HD
#ifdef __CUDACC__
#define B200_HD __host__ __device__ __forceinline__
#else
#define B200_HD inline      // tests/streamk_check.cpp compiles this header for the host
#endif
#endif

// floor(n / P) for 0 <= n < 2^22, 1 <= P < 2^11 through one float multiply: (n + 0.5) / P is at
// least 0.5 / P away from an integer, far more than the rounding of the product (2^-22 relative).
B200_HD unsigned small_div(unsigned n, float inv_p) {
  return (unsigned)(((float)n + 0.5f) * inv_p);
}

// Range p of P equal contiguous ranges of [0, T): begins at floor(p T / P).  Every thread of every
// CTA evaluates this at kernel entry, on the critical path to the first weight fetch, so no 64-bit
// division (a ~100-instruction dependent chain): with T = q P + r,
//   floor(p T / P) = p q + floor(p r / P),   p r < P^2 < 2^22  (P <= 2 x 148).
B200_HD int range_begin(int p, int T, int P) {
  const unsigned q = (unsigned)T / (unsigned)P, r = (unsigned)T - q * (unsigned)P;
  return (int)((unsigned)p * q + small_div((unsigned)p * r, 1.0f / (float)P));
}
// The range that holds work item t: floor(((t + 1) P - 1) / T) = the largest p with
// range_begin(p) <= t.  A float estimate (relative error 2^-21: off by at most one) settled with
// range_begin.
B200_HD int owner_of(int t, int P, int T) {
  int p = (int)(((float)t + 0.5f) * ((float)P / (float)T));
  if (p > P - 1) p = P - 1;
  while (p > 0 && range_begin(p, T, P) > t) --p;
  while (p + 1 < P && range_begin(p + 1, T, P) <= t) ++p;
  return p;
}

// The order in which a CTA walks its contiguous work range [t_begin, t_end): the segment
// that finishes someone else's output tile first, then the segment that starts a tile someone
// else finishes, then its whole tiles.  Every warp role derives the same sequence.
struct Walk {
  int t_begin, t_end, KB;
  int first_end;   // end of the first segment
  int tail_start;  // start of a distinct, partial last segment, or -1
  int nseg;
  __device__ __forceinline__ Walk(int tb, int te, int kb) : t_begin(tb), t_end(te), KB(kb) {
    if (te <= tb) { first_end = tb; tail_start = -1; nseg = 0; return; }
    const int fe = (tb / kb + 1) * kb;
    first_end = fe < te ? fe : te;
    const int ls = ((te - 1) / kb) * kb;
    tail_start = (ls >= first_end && te - ls < kb) ? ls : -1;
    const int mid_end = tail_start >= 0 ? tail_start : te;
    nseg = 1 + (tail_start >= 0 ? 1 : 0) + (mid_end - first_end) / kb;
  }
  __device__ __forceinline__ void seg(int idx, int& t0, int& n) const {
    if (idx == 0) { t0 = t_begin; n = first_end - t_begin; return; }
    if (tail_start >= 0) {
      if (idx == 1) { t0 = tail_start; n = t_end - tail_start; return; }
      --idx;
    }
    t0 = first_end + (idx - 1) * KB;
    n = KB;
  }
};

// Tile-by-tile cursor over a Walk (used by the single producer thread).
struct TileCursor {
  const Walk& w;
  int MT;
  int idx, left;
  int kb, nt, mt;
  __device__ __forceinline__ TileCursor(const Walk& walk, int mt_count)
      : w(walk), MT(mt_count), idx(-1), left(0), kb(0), nt(0), mt(0) { load_next(); }
  __device__ __forceinline__ void load_next() {
    if (++idx >= w.nseg) return;
    int t0, n;
    w.seg(idx, t0, n);
    const int o = t0 / w.KB;
    kb = t0 - o * w.KB;
    nt = o / MT;
    mt = o - nt * MT;
    left = n;
  }
  __device__ __forceinline__ void next() {
    ++kb;
    if (--left == 0) load_next();
  }
};

}  // namespace b200
