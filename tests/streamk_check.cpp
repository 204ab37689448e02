// Host check of the division-free stream-K range arithmetic (csrc/w4a16_streamk.cuh) against the
// 64-bit formulas it replaces.  Built and run by tests/test_streamk.py.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#define __device__
#define __forceinline__ inline
#include "../vllm-quick_b200/csrc/w4a16_streamk.cuh"

static int ref_begin(int p, int T, int P) { return (int)((long long)p * T / P); }
static int ref_owner(int t, int P, int T) { return (int)((((long long)t + 1) * P - 1) / T); }

static long long checks = 0;
static int check(int T, int P) {
  for (int p = 0; p <= P; ++p) {
    ++checks;
    if (b200::range_begin(p, T, P) != ref_begin(p, T, P)) {
      printf("range_begin(%d, %d, %d) = %d, want %d\n", p, T, P, b200::range_begin(p, T, P), ref_begin(p, T, P));
      return 1;
    }
  }
  // every range boundary and its neighbours, plus a stride through the interior
  for (int p = 0; p < P; ++p) {
    const int b = ref_begin(p, T, P);
    for (int d = -1; d <= 1; ++d) {
      const int t = b + d;
      if (t < 0 || t >= T) continue;
      ++checks;
      if (b200::owner_of(t, P, T) != ref_owner(t, P, T)) {
        printf("owner_of(%d, %d, %d) = %d, want %d\n", t, P, T, b200::owner_of(t, P, T), ref_owner(t, P, T));
        return 1;
      }
    }
  }
  const int step = T / 997 + 1;
  for (long long tt = 0; tt < T; tt += step) {
    const int t = (int)tt;
    ++checks;
    if (b200::owner_of(t, P, T) != ref_owner(t, P, T)) {
      printf("owner_of(%d, %d, %d) = %d, want %d\n", t, P, T, b200::owner_of(t, P, T), ref_owner(t, P, T));
      return 1;
    }
  }
  if (b200::owner_of(T - 1, P, T) != P - 1) { printf("last item not in the last range (%d, %d)\n", T, P); return 1; }
  return 0;
}

int main() {
  // every grid size up to two per SM (CTA pairs halve it), every small T, and large ones
  for (int P = 1; P <= 296; ++P) {
    for (int T = P; T <= P + 2100; ++T)
      if (check(T, P)) return 1;
    uint64_t x = 88172645463325252ull + (uint64_t)P;
    for (int i = 0; i < 300; ++i) {
      x ^= x << 13; x ^= x >> 7; x ^= x << 17;
      const int shift = 33 + (int)(x % 20);             // T from ~2^11 up to 2^31 - 1
      int T = (int)((x >> shift) & 0x7fffffffu);
      if (T < P) T = P;
      if (check(T, P)) return 1;
    }
    const int big[] = {0x7fffffff, 0x7ffffffe, 1 << 30, (1 << 30) + 1, 1 << 24, (1 << 24) - 1, 57344 / 128 * 64 * 8};
    for (int T : big)
      if (T >= P && check(T, P)) return 1;
  }
  printf("ok %lld checks\n", checks);
  return 0;
}
