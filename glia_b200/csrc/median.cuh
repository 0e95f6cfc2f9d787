// Sorted value stores and the selection primitive behind the median features (median.cu).
#pragma once
#include <vector>
#include "handle.cuh"

namespace glia {

// keys: order-preserving u32 image of the float values (f2ord), sorted inside each base set;
// set s of a store occupies [start[s], start[s + 1]).  All images of a store share the starts.
struct MedStore {
  const uint32_t* rkeys[GLIA_MAX_IMG]; const long long* rstart;  // per base region, per region image
  const uint32_t* ekeys[GLIA_MAX_IMG]; const long long* estart;  // per directed base boundary, per boundary image
};

// number of keys <= v in the sorted segment [lo, hi)
__device__ __forceinline__ long long med_count_le(const uint32_t* __restrict__ k, long long lo, long long hi, uint32_t v) {
  long long a = lo, b = hi;
  while (a < b) { const long long m = (a + b) >> 1; if (k[m] <= v) a = m + 1; else b = m; }
  return a - lo;
}

// The n/2-th smallest value (0-based; util/stats.hxx:83-91) of the union of nseg sorted segments, by one warp.
// seg(i, lo, hi) yields segment i or returns false to skip it; every lane must call with the same arguments.
// Returns 0.0 for an empty union (ImageRealFeats::generate returns before touching its zero-initialised fields).
template <class SegFn>
__device__ double warp_median(const uint32_t* __restrict__ keys, int nseg, SegFn seg, int lane) {
  long long cnt = 0;
  for (int i = lane; i < nseg; i += 32) { long long lo, hi; if (seg(i, lo, hi)) cnt += hi - lo; }
  for (int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  if (cnt == 0) return 0.0;
  const long long k = cnt / 2;
  uint32_t lo = 0, hi = 0xFFFFFFFFu;  // smallest v with count(<= v) > k
  while (lo < hi) {
    const uint32_t mid = lo + ((hi - lo) >> 1);
    long long c = 0;
    for (int i = lane; i < nseg; i += 32) { long long a, b; if (seg(i, a, b)) c += med_count_le(keys, a, b, mid); }
    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if (c > k) hi = mid; else lo = mid + 1;
  }
  return (double)ord2f(lo);
}

}  // namespace glia

glia::MedStore glia_median_view(const glia_handle* h);
int glia_median_build(glia_handle* h);
int glia_median_init_candidates(glia_handle* h);
int glia_median_queries(glia_handle* h, const std::vector<int>& kind, const std::vector<long long>& off,
                        const std::vector<int>& len, const std::vector<int>& seg, double* out, double* dOutKeep);
