// Medians (the reference's GLIA_USE_MEDIAN_AS_FEATS build): order statistics of voxel sets that are unions of base
// sets.
//
// Replaces (reference, /root/reference/code):
//   ImageRealFeats::generate, median branch        type/feat.hxx:711-724
//   stats::amedian (n/2-th order statistic)        util/stats.hxx:83-91
//   ImageRealDiffFeats::generate (medianDiff)      type/feat.hxx:801-810
// The reference copies the values of every pixel of a region (or boundary) into a vector and runs nth_element, for
// every region of every candidate.  A median is not a mergeable statistic, so next to the records this keeps the VALUES
// once: per region image the voxels' values sorted inside each base region, per boundary image the boundary voxels'
// values sorted inside each directed base boundary (one 64-bit radix sort of (set, value key) per image).  Every set
// the path asks about - a region of the merge tree, its boundary entries, a pair boundary - is a union of such
// sorted segments; its n/2-th smallest value is found by a binary search over the 32-bit value keys, where one probe
// counts "how many values <= v" with one binary search per segment (a warp or a CTA per set).
#include <cub/cub.cuh>
#include "median.cuh"

namespace glia {

// One 64-bit key per voxel: (region index << 32 | value key); masked voxels sort behind the last region.
__global__ void med_region_keys_kernel(const uint32_t* __restrict__ labels, const uint32_t* __restrict__ mask,
                                       const float* __restrict__ img, const uint32_t* __restrict__ rank, size_t n,
                                       uint32_t R, unsigned long long* __restrict__ keys, int contourOnly) {
  size_t p = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (p >= n) return;
  if (mask && mask[p] == 0 && !contourOnly) { keys[p] = (unsigned long long)R << 32; return; }
  keys[p] = ((unsigned long long)rank[labels[p]] << 32) | f2ord(img[p]);
}

// One key per voxel: (directed boundary index << 32 | value key) for a boundary voxel - the first differing valid
// neighbour in the order x-, x+, y-, y+, z-, z+ names the boundary (type/neighbor.hxx:72-126, util/struct.hxx:95-108) -
// and E << 32 for every other voxel.
__global__ void med_edge_keys_kernel(const uint32_t* __restrict__ labels, const uint32_t* __restrict__ mask,
                                     const float* __restrict__ img, const uint32_t* __restrict__ rank,
                                     const int* __restrict__ row, const int* __restrict__ eb, int D, int X, int Y, int Z,
                                     uint32_t E, unsigned long long* __restrict__ keys, int contourOnly) {
  const size_t n = (size_t)X * Y * Z;
  size_t p = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (p >= n) return;
  unsigned long long key = (unsigned long long)E << 32;
  if (!mask || mask[p] != 0 || contourOnly) {
    const int x = (int)(p % X), y = (int)((p / X) % Y), z = (int)(p / ((size_t)X * Y));
    const int c[3] = {x, y, z}, ext[3] = {X, Y, Z};
    const size_t stride[3] = {1, (size_t)X, (size_t)X * Y};
    const uint32_t me = labels[p];
    bool found = false; uint32_t nb = 0;
    for (int d = 0; d < D && !found; ++d)
      for (int sgn = -1; sgn <= 1 && !found; sgn += 2) {
        const int cc = c[d] + sgn;
        if (cc < 0 || cc >= ext[d]) continue;
        const size_t q = sgn < 0 ? p - stride[d] : p + stride[d];
        if (mask && mask[q] == 0) continue;
        if (labels[q] != me) { found = true; nb = labels[q]; }
      }
    if (found) {
      const int a = (int)rank[me], b = (int)rank[nb];
      int lo = row[a], hi = row[a + 1];  // boundaries of a are sorted by b
      while (lo < hi) { const int m = (lo + hi) >> 1; if (eb[m] < b) lo = m + 1; else hi = m; }
      key = ((unsigned long long)lo << 32) | f2ord(img[p]);
    }
  }
  keys[p] = key;
}

__global__ void med_strip_kernel(const unsigned long long* __restrict__ keys, size_t n, uint32_t* __restrict__ out) {
  size_t p = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (p < n) out[p] = (uint32_t)keys[p];
}

// start[s] = first sorted key of set s (s = 0 .. nSets; start[nSets] = number of valid keys)
__global__ void med_starts_kernel(const unsigned long long* __restrict__ keys, size_t n, long long nSets,
                                  long long* __restrict__ start) {
  long long s = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (s > nSets) return;
  const unsigned long long want = (unsigned long long)s << 32;
  size_t lo = 0, hi = n;
  while (lo < hi) { const size_t m = (lo + hi) >> 1; if (keys[m] < want) lo = m + 1; else hi = m; }
  start[s] = (long long)lo;
}

// ---- K6: medians of the initial candidates -------------------------------------------------------------------------
// per base region r: [region image i] the middle element of its own segment; [boundary image j] over all its
// directed boundaries (every entry of a base region is live).  One warp per region.
__global__ void __launch_bounds__(256)
med_regions_kernel(const __grid_constant__ MedStore S, int nR, int nB, long long R, const int* __restrict__ row,
                   double* __restrict__ out) {
  const long long r = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (r >= R) return;
  double* o = out + (size_t)r * (nR + nB);
  const long long a = S.rstart[r], n = S.rstart[r + 1] - a;
  if (lane < nR) o[lane] = n ? (double)ord2f(S.rkeys[lane][a + n / 2]) : 0.0;
  const int e0 = row[r], deg = row[r + 1] - e0;
  for (int j = 0; j < nB; ++j) {
    auto seg = [&](int i, long long& lo, long long& hi) { lo = S.estart[e0 + i]; hi = S.estart[e0 + i + 1]; return true; };
    const double m = warp_median(S.ekeys[j], deg, seg, lane);
    if (lane == 0) o[nR + j] = m;
  }
}

// per candidate c = (a, b): [Z region per region image][boundary-of Z per boundary image][pair boundary per boundary
// image]; entries(Z) = the boundaries of a and of b except the mutual pair (type/region.hxx:66-75), pair boundary =
// (a -> b) u (b -> a) (util/struct.hxx:10-16).  One warp per candidate.
__global__ void __launch_bounds__(256)
med_candidates_kernel(const __grid_constant__ MedStore S, int nR, int nB, long long C, const int* __restrict__ cand,
                      const int* __restrict__ ea, const int* __restrict__ eb, const int* __restrict__ rev,
                      const int* __restrict__ row, double* __restrict__ out) {
  const long long c = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (c >= C) return;
  const int e = cand[c], rv = rev[e], a = ea[e], b = eb[e];
  double* o = out + (size_t)c * (nR + 2 * nB);
  for (int i = 0; i < nR; ++i) {
    auto seg = [&](int k, long long& lo, long long& hi) { const int r = k ? b : a; lo = S.rstart[r]; hi = S.rstart[r + 1]; return true; };
    const double m = warp_median(S.rkeys[i], 2, seg, lane);
    if (lane == 0) o[i] = m;
  }
  const int ra = row[a], da = row[a + 1] - ra, rb = row[b], db = row[b + 1] - rb;
  for (int j = 0; j < nB; ++j) {
    auto segZ = [&](int k, long long& lo, long long& hi) {
      const int x = k < da ? ra + k : rb + (k - da);
      if (x == e || x == rv) return false;
      lo = S.estart[x]; hi = S.estart[x + 1];
      return true;
    };
    const double mz = warp_median(S.ekeys[j], da + db, segZ, lane);
    auto segP = [&](int k, long long& lo, long long& hi) { const int x = k ? rv : e; lo = S.estart[x]; hi = S.estart[x + 1]; return true; };
    const double mp = warp_median(S.ekeys[j], 2, segP, lane);
    if (lane == 0) { o[nR + j] = mz; o[nR + nB + j] = mp; }
  }
}

// ---- generic queries (K5 / K7): a set = an explicit list of base segments ------------------------------------------
// query q: kind[q] = store (0 region, 1 boundary) * GLIA_MAX_IMG + image, segments seg[off[q] .. off[q] + len[q]).
// GROUP = 32: one warp per query; GROUP = 256: one CTA per query (long lists).
template <int GROUP>
__global__ void __launch_bounds__(256)
med_query_kernel(const __grid_constant__ MedStore S, const int* __restrict__ qid, int nq, const int* __restrict__ kind,
                 const long long* __restrict__ off, const int* __restrict__ len, const int* __restrict__ seg,
                 double* __restrict__ out) {
  __shared__ long long red[8];
  const int w = GROUP == 32 ? (int)((blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5) : (int)blockIdx.x;
  if (w >= nq) return;  // uniform per group
  const int q = qid[w], lane = GROUP == 32 ? (threadIdx.x & 31) : (int)threadIdx.x;
  const int kd = kind[q], store = kd / GLIA_MAX_IMG, img = kd % GLIA_MAX_IMG;
  const uint32_t* keys = store ? S.ekeys[img] : S.rkeys[img];
  const long long* start = store ? S.estart : S.rstart;
  const int* sg = seg + off[q];
  const int n = len[q];
  auto total = [&](long long v) -> long long {  // group-wide sum
    if (GROUP == 32) { for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o); return v; }
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    long long t = 0;
    for (int i = 0; i < 8; ++i) t += red[i];
    return t;
  };
  long long cnt = 0;
  for (int i = lane; i < n; i += GROUP) cnt += start[sg[i] + 1] - start[sg[i]];
  cnt = total(cnt);
  double m = 0.0;
  if (cnt > 0) {
    const long long k = cnt / 2;
    uint32_t lo = 0, hi = 0xFFFFFFFFu;
    while (lo < hi) {
      const uint32_t mid = lo + ((hi - lo) >> 1);
      long long c = 0;
      for (int i = lane; i < n; i += GROUP) c += med_count_le(keys, start[sg[i]], start[sg[i] + 1], mid);
      c = total(c);
      if (c > k) hi = mid; else lo = mid + 1;
    }
    m = (double)ord2f(lo);
  }
  if (lane == 0) out[q] = m;
}

}  // namespace glia

using namespace glia;
#define CK(expr) GLIA_CUDA_CHECK(h, expr)
static inline unsigned nblk(long long n, int per) { return (unsigned)((n + per - 1) / per); }

MedStore glia_median_view(const glia_handle* h) {
  MedStore S;
  for (int i = 0; i < GLIA_MAX_IMG; ++i) { S.rkeys[i] = h->medRKeys[i].as<uint32_t>(); S.ekeys[i] = h->medEKeys[i].as<uint32_t>(); }
  S.rstart = h->medRStart.as<long long>(); S.estart = h->medEStart.as<long long>();
  return S;
}

// Sorted value stores of the graph just finalised (whole volume in this handle).
int glia_median_build(glia_handle* h) {
  const glia_params& prm = h->prm;
  const VoxParams& P = h->vp;
  cudaStream_t st = h->stream;
  h->hasMedian = false;
  const int D = prm.ndim, X = prm.dims[0], Y = prm.dims[1], Z = D == 3 ? prm.dims[2] : 1;
  const size_t n = (size_t)X * Y * Z;
  const long long R = h->R, E = h->E;
  if (n == 0 || R == 0) { h->hasMedian = true; return GLIA_OK; }
  CK(h->medTmpA.ensure(8 * n)); CK(h->medTmpB.ensure(8 * n));
  CK(h->medRStart.ensure(8 * (size_t)(R + 2))); CK(h->medEStart.ensure(8 * (size_t)(E + 2)));
  unsigned long long* ka = h->medTmpA.as<unsigned long long>();
  unsigned long long* kb = h->medTmpB.as<unsigned long long>();
  auto bits_for = [](long long v) { int b = 1; while ((1ll << b) <= v) ++b; return b; };
  auto sort_into = [&](long long nSets, DevBuf& outKeys, long long* start, bool wantStart) -> int {
    cub::DoubleBuffer<unsigned long long> db(ka, kb);
    size_t tmp = 0;
    const int endBit = 32 + bits_for(nSets);
    CK(cub::DeviceRadixSort::SortKeys(nullptr, tmp, db, (long long)n, 0, endBit, st));
    CK(h->cubTmp.ensure(tmp));
    CK(cub::DeviceRadixSort::SortKeys(h->cubTmp.p, tmp, db, (long long)n, 0, endBit, st));
    CK(outKeys.ensure(4 * n));
    med_strip_kernel<<<nblk((long long)n, 256), 256, 0, st>>>(db.Current(), n, outKeys.as<uint32_t>());
    if (wantStart) med_starts_kernel<<<nblk(nSets + 1, 256), 256, 0, st>>>(db.Current(), n, nSets, start);
    h->launches += 3;
    CK(cudaGetLastError());
    return GLIA_OK;
  };
  int s;
  for (int i = 0; i < prm.n_rimg; ++i) {
    med_region_keys_kernel<<<nblk((long long)n, 256), 256, 0, st>>>(P.labels, P.mask, h->vc.reg[i].data, h->rank.as<uint32_t>(), n, (uint32_t)R, ka, P.contourOnly);
    h->launches++;
    if ((s = sort_into(R, h->medRKeys[i], h->medRStart.as<long long>(), i == 0))) return s;
  }
  for (int j = 0; j < prm.n_bimg && E > 0; ++j) {
    med_edge_keys_kernel<<<nblk((long long)n, 256), 256, 0, st>>>(P.labels, P.mask, h->vc.edge[j].data, h->rank.as<uint32_t>(),
                                                                   h->row.as<int>(), h->eb.as<int>(), D, X, Y, Z, (uint32_t)E, ka, P.contourOnly);
    h->launches++;
    if ((s = sort_into(E, h->medEKeys[j], h->medEStart.as<long long>(), j == 0))) return s;
  }
  if (E == 0 || prm.n_bimg == 0) CK(cudaMemsetAsync(h->medEStart.p, 0, 8 * (size_t)(E + 2), st));
  if (prm.n_rimg == 0) {  // starts are still needed by nobody; keep them defined
    CK(cudaMemsetAsync(h->medRStart.p, 0, 8 * (size_t)(R + 2), st));
  }
  h->hasMedian = true;
  return GLIA_OK;
}

// per-region and per-candidate medians of the initial candidates (featurize.cu reads them)
int glia_median_init_candidates(glia_handle* h) {
  if (!h->hasMedian) { h->err = "median stores were not built"; return GLIA_E_STATE; }
  const glia_params& prm = h->prm;
  const int nR = prm.n_rimg, nB = prm.n_bimg;
  cudaStream_t st = h->stream;
  MedStore S = glia_median_view(h);
  CK(h->medReg.ensure(8 * (size_t)(h->R + 1) * (nR + nB + 1)));
  CK(h->medCand.ensure(8 * (size_t)(h->C + 1) * (nR + 2 * nB + 1)));
  if (h->R) med_regions_kernel<<<nblk(h->R * 32, 256), 256, 0, st>>>(S, nR, nB, h->R, h->row.as<int>(), h->medReg.as<double>());
  if (h->C) med_candidates_kernel<<<nblk(h->C * 32, 256), 256, 0, st>>>(S, nR, nB, h->C, h->cand.as<int>(), h->ea.as<int>(), h->eb.as<int>(),
                                                                         h->rev.as<int>(), h->row.as<int>(), h->medCand.as<double>());
  h->launches += 2;
  CK(cudaGetLastError());
  return GLIA_OK;
}

// Runs nq queries (host arrays) and returns their medians (host array out[nq]).  Lists of >= 256 segments get a CTA.
int glia_median_queries(glia_handle* h, const std::vector<int>& kind, const std::vector<long long>& off,
                        const std::vector<int>& len, const std::vector<int>& seg, double* out, double* dOutKeep) {
  const int nq = (int)kind.size();
  if (!nq) return GLIA_OK;
  if (!h->hasMedian) { h->err = "median stores were not built"; return GLIA_E_STATE; }
  cudaStream_t st = h->stream;
  std::vector<int> small, big;
  for (int q = 0; q < nq; ++q) (len[q] >= 256 ? big : small).push_back(q);
  std::vector<int> qid(small);
  qid.insert(qid.end(), big.begin(), big.end());
  DevBuf &bK = h->medQ[0], &bO = h->medQ[1], &bL = h->medQ[2], &bS = h->medQ[3], &bI = h->medQ[4], &bR = h->medQ[5];
  CK(bK.ensure(4 * (size_t)nq)); CK(bO.ensure(8 * (size_t)nq)); CK(bL.ensure(4 * (size_t)nq));
  CK(bS.ensure(4 * (seg.size() + 1))); CK(bI.ensure(4 * (size_t)nq)); CK(bR.ensure(8 * (size_t)nq));
  CK(cudaMemcpyAsync(bK.p, kind.data(), 4 * (size_t)nq, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(bO.p, off.data(), 8 * (size_t)nq, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(bL.p, len.data(), 4 * (size_t)nq, cudaMemcpyHostToDevice, st));
  if (!seg.empty()) CK(cudaMemcpyAsync(bS.p, seg.data(), 4 * seg.size(), cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(bI.p, qid.data(), 4 * (size_t)nq, cudaMemcpyHostToDevice, st));
  MedStore S = glia_median_view(h);
  double* dOut = dOutKeep ? dOutKeep : bR.as<double>();
  if (!small.empty()) {
    med_query_kernel<32><<<nblk((long long)small.size() * 32, 256), 256, 0, st>>>(S, bI.as<int>(), (int)small.size(), bK.as<int>(), bO.as<long long>(),
                                                                                  bL.as<int>(), bS.as<int>(), dOut);
    h->launches++;
  }
  if (!big.empty()) {
    med_query_kernel<256><<<(unsigned)big.size(), 256, 0, st>>>(S, bI.as<int>() + small.size(), (int)big.size(), bK.as<int>(), bO.as<long long>(),
                                                                bL.as<int>(), bS.as<int>(), dOut);
    h->launches++;
  }
  CK(cudaGetLastError());
  if (out) {
    CK(cudaMemcpyAsync(out, dOut, 8 * (size_t)nq, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
  }
  return GLIA_OK;
}
