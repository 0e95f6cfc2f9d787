// K6: boundary-classifier feature row from mergeable records (device code, header-only).
//
// Replaces (reference, /root/reference/code):
//   RegionFeats::generate / serialize               hmt/bc_feat.hxx:69-127
//   BoundaryFeats::generate / serialize             hmt/bc_feat.hxx:156-214
//   BoundaryClassificationFeats::serialize          hmt/bc_feat.hxx:232-238
//   selectFeatures (--simpf)                        hmt/bc_feat.hxx:247-279
//   RegionShapeFeats / ...DiffFeats / ...IntraDiff  type/feat.hxx:71-90, :124-132, :176-186
//   ImageRegionShape(IntraDiff)Feats                type/feat.hxx:485-502, :567-589
//   ImageLabelFeats / ImageRealFeats (+Diff)        type/feat.hxx:632-669, :706-810
//   stats::entropy, distL1, distX2                  util/stats.hxx:145-185
//   sdivide / slog / ssqrt                          glia_base.hxx:77-84
// A record accessor exposes add(i), mx(i), f64(i) of one statistics record (common.cuh); the
// row is a pure function of seven records: region X, Y, Z; boundary-of X, Y, Z; pair boundary.
#pragma once
#include <float.h>
#include <math.h>
#include "handle.cuh"

namespace glia {

#define GLIA_FEPS 2.22e-16
__device__ __forceinline__ double sdivide(double a, double b, double d) { return fabs(b) >= GLIA_FEPS ? a / b : d; }
__device__ __forceinline__ double slog(double x, double d) { return x > 0.0 ? log(x) : d; }
__device__ __forceinline__ double ssqrt(double x, double d) { return x >= 0.0 ? sqrt(x) : d; }

struct RecSingle {
  const uint32_t* a; const uint32_t* m; const double* f;
  __device__ uint32_t add(int i) const { return a[i]; }
  __device__ uint32_t mx(int i) const { return m[i]; }
  __device__ double f64(int i) const { return f[i]; }
};
struct RecSum2 {  // union of two disjoint records
  RecSingle p, q;
  __device__ uint32_t add(int i) const { return p.a[i] + q.a[i]; }
  __device__ uint32_t mx(int i) const { return max(p.m[i], q.m[i]); }
  __device__ double f64(int i) const { return p.f[i] + q.f[i]; }
};
// boundary-of Z = X u Y for base regions: all entries of X and Y minus the mutual pair;
// max / f64 words are re-reduced by the caller (not subtractable), add words by subtraction.
struct RecBoundaryZ {
  const uint32_t *pa, *qa, *ea, *ra;
  uint32_t mxw[2 * GLIA_MAX_IMG]; double f[2 * GLIA_MAX_IMG];
  __device__ uint32_t add(int i) const { return pa[i] + qa[i] - ea[i] - ra[i]; }
  __device__ uint32_t mx(int i) const { return mxw[i]; }
  __device__ double f64(int i) const { return f[i]; }
};

struct ImgFeat { double entropy, mean, stddev, mn, mx, median; };

// util/image_stats.hxx:39-53 (normalise by n), util/stats.hxx:145-152, type/feat.hxx:706-738
// log2 of a voxel count: table (host log2, FeatCfg::log2tab) for counts below GLIA_LOG2_TAB, else computed
#define GLIA_LOG2_TAB 4096
__device__ __forceinline__ double log2_count(const double* __restrict__ tab, uint32_t c) {
  return c < GLIA_LOG2_TAB ? __ldg(tab + c) : log2((double)c);
}

template <class Rec>
__device__ void img_feat(const double* __restrict__ l2tab, const Rec& r, uint32_t n, int bins, int addOff, int maxOff,
                         int f64Off, ImgFeat& o, double* histOut = nullptr) {
  // histOut: the normalised histogram itself (GLIA_USE_HISTOGRAM_AS_FEATS, type/feat.hxx:608-626; util/stats.hxx:125-131)
  if (histOut) for (int i = 0; i < bins; ++i) histOut[i] = n ? r.add(addOff + i) / (double)n : 0.0;
  o.median = 0.0;
  // stats::entropy (util/stats.hxx:145-152): - sum p log2 p over bins with |p| >= FEPS, p = c / n.  log2 p is taken as
  // log2 c - log2 n from the count table (the FP64 log2 and the division per bin were ~60 % of the featurise kernels);
  // a bin holding all n voxels contributes exactly 0 as in the reference.  Agreement with the reference's
  // log2(c / n): ~1e-15 relative (the entropy columns were never bit-identical: libm vs CUDA log2).
  o.entropy = 0.0;
  if (n) {
    const double ln = log2_count(l2tab, n), rn = 1.0 / (double)n;
    for (int i = 0; i < bins; ++i) {
      const uint32_t c = r.add(addOff + i);
      if (c) o.entropy -= ((double)c * rn) * (log2_count(l2tab, c) - ln);  // c >= 1 => p >= 1/n > FEPS for n < 2^52
    }
  }
  if (n == 0 || maxOff < 0) { o.mean = o.stddev = o.mn = o.mx = 0.0; return; }
  int ni = (int)n;
  o.mean = r.f64(f64Off) / ni;
  o.stddev = ssqrt(r.f64(f64Off + 1) / ni - o.mean * o.mean, 0.0);
  uint32_t kmin = ~r.mx(maxOff), kmax = r.mx(maxOff + 1);
  // all-NaN sets leave the words at 0: the reference then keeps its +-DBL_MAX initialisers
  o.mn = r.mx(maxOff) ? (double)ord2f(kmin) : DBL_MAX;
  o.mx = kmax ? (double)ord2f(kmax) : -DBL_MAX;
}

template <class R1, class R2>
__device__ void hist_dists(const R1& a, uint32_t na, const R2& b, uint32_t nb, int bins, int addOff,
                           double& l1, double& x2) {
  l1 = 0.0; x2 = 0.0;
  for (int i = 0; i < bins; ++i) {
    double p = na ? (double)a.add(addOff + i) / (double)na : 0.0;
    double q = nb ? (double)b.add(addOff + i) / (double)nb : 0.0;
    double d = p - q;
    l1 += fabs(d);
    x2 += d * d / (p + q + GLIA_FEPS);
  }
}

struct ShapeFeat {
  double area, perim, compactness, bboxArea, bboxSize[3];
  double validPerims[GLIA_MAX_THR], rValidPerims[GLIA_MAX_THR];
  uint32_t nReg, nBnd;  // voxel counts of the region and of its boundary entries
};

// RegionShapeFeats::generate + ImageRegionShapeFeats::generate
template <class RR, class RB>
__device__ void shape_feat(const FeatCfg& F, const RR& reg, const RB& bnd, ShapeFeat& s) {
  uint32_t area = reg.add(0), border = reg.add(1), nbnd = bnd.add(0);
  s.nReg = area; s.nBnd = nbnd;
  s.area = (double)area;
  s.perim = (double)((unsigned long long)nbnd + border);
  s.compactness = sdivide(pow(s.perim, (double)F.D / (F.D - 1)), s.area, 0.0);
  s.area = sdivide(s.area, F.normArea, 0.0);
  s.perim = sdivide(s.perim, F.normLen, 0.0);
  s.bboxArea = 1.0;
  for (int d = 0; d < F.D; ++d) {
    int ax = F.axis[d];
    uint32_t wmin = reg.mx(ax), wmax = reg.mx(3 + ax);
    double bb = (wmin && wmax) ? (double)((long long)(wmax - 1) - (long long)(~wmin)) : 0.0;
    s.bboxSize[d] = sdivide(bb, F.normLen, 0.0);
    s.bboxArea *= bb;
  }
  s.bboxArea = sdivide(s.bboxArea, F.normArea, 0.0);
  for (int t = 0; t < F.nThr; ++t) {
    double vp = (double)bnd.add(1 + t);
    s.validPerims[t] = sdivide(vp, F.normLen, 0.0);
    s.rValidPerims[t] = sdivide(vp, (double)nbnd, 0.0);
  }
}

// Serialises ImageFeats = ImageLabelFeats + ImageRealFeats (type/feat.hxx:826-840): [bins] entropy [median] mean stddev
// min max; `real` false = ImageLabelFeats alone.
__device__ __forceinline__ double* push_img(const FeatCfg& F, const ImgFeat& g, bool real, double* p) {
  *p++ = g.entropy;
  if (!real) return p;
  if (F.median) *p++ = g.median;
  *p++ = g.mean; *p++ = g.stddev; *p++ = g.mn; *p++ = g.mx;
  return p;
}

// rm / bm: medians of the region images / boundary images of THIS region (element i at rm[i * rmStride]); read only
// when F.median.
template <class RR, class RB>
__device__ double* push_region(const FeatCfg& F, const ShapeFeat& s, const RR& reg, const RB& bnd,
                               const double* sal, double* p, const double* rm = nullptr, int rmStride = 0,
                               const double* bm = nullptr, int bmStride = 0) {
  const int lg = F.useLog;
  *p++ = lg ? slog(s.area, 0) : s.area;
  *p++ = lg ? slog(s.perim, 0) : s.perim;
  *p++ = s.compactness;
  *p++ = lg ? slog(s.bboxArea, 0) : s.bboxArea;
  for (int d = 0; d < F.D; ++d) *p++ = lg ? slog(s.bboxSize[d], 0) : s.bboxSize[d];
  for (int t = 0; t < F.nThr; ++t) *p++ = lg ? slog(s.validPerims[t], 0) : s.validPerims[t];
  for (int t = 0; t < F.nThr; ++t) *p++ = s.rValidPerims[t];
  ImgFeat g;
  for (int i = 0; i < F.nR; ++i) {
    img_feat(F.log2tab, reg, s.nReg, F.rBins[i], F.rAdd[i], F.rMax[i], F.rF64[i], g, F.hist ? p : nullptr);
    if (F.hist) p += F.rBins[i];
    if (F.median && rm) g.median = rm[i * rmStride];
    p = push_img(F, g, true, p);
  }
  for (int i = 0; i < F.nRL; ++i) {
    img_feat(F.log2tab, reg, s.nReg, F.rlBins[i], F.rlAdd[i], -1, 0, g, F.hist ? p : nullptr);
    if (F.hist) p += F.rlBins[i];
    p = push_img(F, g, false, p);
  }
  for (int i = 0; i < F.nB; ++i) {
    img_feat(F.log2tab, bnd, s.nBnd, F.bBins[i], F.bAdd[i], F.bMax[i], F.bF64[i], g, F.hist ? p : nullptr);
    if (F.hist) p += F.bBins[i];
    if (F.median && bm) g.median = bm[i * bmStride];
    p = push_img(F, g, true, p);
  }
  if (sal) *p++ = *sal;
  return p;
}

// Full candidate row.  (X, Y) in call order; x1 = the one with the smaller normalised area
// (strict >, hmt/main_merge_order_bc.cxx:77-80).  sal: null or {salX, salY, salZ}.
template <class RX, class RY, class RZ, class BX, class BY, class BZ, class PB>
__device__ void featurize_row(const FeatCfg& F, const RX& rx, const RY& ry, const RZ& rz,
                              const BX& bx, const BY& by, const BZ& bz, const PB& pb,
                              const double* sal, double* out, int part = -1, bool tieSwap = false,
                              const double* med = nullptr) {
  // med (F.median): medians of this row's seven voxel sets, [3 * i + w] for region image i of w = X, Y, Z and
  // [3 * nR + 4 * j + w] for boundary image j of w = boundary-of X, Y, Z and the pair boundary (median.cu).
  // tieSwap: for EQUAL areas the reference keeps the call order; the caller says whether the reference's
  // call order was (Y, X) (orient.cu) - only the x1 / x2 blocks depend on it.
  // part: -1 = whole row; 0 = x0 (pair terms), 1 / 2 = first / second region block, 3 = merged block.
  // The four parts write disjoint ranges, so four threads can build one row (merge_order.cu).
  ShapeFeat sx, sy, sz;
  shape_feat(F, rx, bx, sx);
  shape_feat(F, ry, by, sy);
  if (part < 0 || part == 3) shape_feat(F, rz, bz, sz);
  if (part > 0) {
    if (F.simple) return;
    const bool sw = sx.area > sy.area || (tieSwap && sx.area == sy.area);
    const int xi = F.xiLen() + (sal ? 1 : 0);
    const int x0len = F.dim + (sal ? 5 : 0) - 3 * xi;
    double* q = out + x0len + (size_t)(part - 1) * xi;
    const double* mb = med ? med + 3 * F.nR : nullptr;
    const bool first = (part == 1) != sw;  // X is written by part 1 unless swapped
    if (part == 3) push_region(F, sz, rz, bz, sal ? sal + 2 : nullptr, q, med ? med + 2 : nullptr, 3, mb ? mb + 2 : nullptr, 4);
    else if (first) push_region(F, sx, rx, bx, sal ? sal + 0 : nullptr, q, med, 3, mb, 4);
    else push_region(F, sy, ry, by, sal ? sal + 1 : nullptr, q, med ? med + 1 : nullptr, 3, mb ? mb + 1 : nullptr, 4);
    return;
  }
  const bool swp = sx.area > sy.area || (tieSwap && sx.area == sy.area);
  const ShapeFeat& s1 = swp ? sy : sx;
  const ShapeFeat& s2 = swp ? sx : sy;
  const int T = F.nThr, lg = F.useLog;
  const uint32_t nb = pb.add(0);
  double* p = out;
  const double areaDiff = fabs(s1.area - s2.area), perimDiff = fabs(s1.perim - s2.perim);
  const double bl = sdivide(ceil(nb / 2.0), F.normLen, 0.0);
  ImgFeat g, g1, g2;
  if (F.simple) {  // bc_feat.hxx:247-279 (logs were applied before selection)
    *p++ = lg ? slog(s1.area, 0) : s1.area;
    *p++ = lg ? slog(s2.area, 0) : s2.area;
    *p++ = lg ? slog(s1.perim, 0) : s1.perim;
    *p++ = lg ? slog(s2.perim, 0) : s2.perim;
    *p++ = lg ? slog(bl, 0) : bl;
    for (int k = 0; k < F.nB; ++k) {
      img_feat(F.log2tab, pb, nb, F.bBins[k], F.bAdd[k], F.bMax[k], F.bF64[k], g);
      *p++ = g.mean;
      if (F.median) *p++ = med ? med[3 * F.nR + 4 * k + 3] : 0.0;  // bc_feat.hxx:265-267
    }
    for (int i = 0; i < F.nR; ++i) {
      double l1, x2;
      if (swp) { img_feat(F.log2tab, ry, sy.nReg, F.rBins[i], F.rAdd[i], F.rMax[i], F.rF64[i], g1); img_feat(F.log2tab, rx, sx.nReg, F.rBins[i], F.rAdd[i], F.rMax[i], F.rF64[i], g2); hist_dists(ry, sy.nReg, rx, sx.nReg, F.rBins[i], F.rAdd[i], l1, x2); }
      else     { img_feat(F.log2tab, rx, sx.nReg, F.rBins[i], F.rAdd[i], F.rMax[i], F.rF64[i], g1); img_feat(F.log2tab, ry, sy.nReg, F.rBins[i], F.rAdd[i], F.rMax[i], F.rF64[i], g2); hist_dists(rx, sx.nReg, ry, sy.nReg, F.rBins[i], F.rAdd[i], l1, x2); }
      *p++ = fabs(g1.mean - g2.mean); *p++ = l1; *p++ = x2; *p++ = fabs(g1.entropy - g2.entropy);
    }
    for (int i = 0; i < F.nRL; ++i) {
      double l1, x2;
      if (swp) hist_dists(ry, sy.nReg, rx, sx.nReg, F.rlBins[i], F.rlAdd[i], l1, x2);
      else hist_dists(rx, sx.nReg, ry, sy.nReg, F.rlBins[i], F.rlAdd[i], l1, x2);
      *p++ = l1; *p++ = x2;
    }
    return;
  }
  *p++ = lg ? slog(areaDiff, 0) : areaDiff;
  *p++ = sdivide(areaDiff, s1.area, 0.0);
  *p++ = sdivide(areaDiff, s2.area, 0.0);
  *p++ = lg ? slog(perimDiff, 0) : perimDiff;
  *p++ = sdivide(perimDiff, s1.perim, 0.0);
  *p++ = sdivide(perimDiff, s2.perim, 0.0);
  *p++ = lg ? slog(bl, 0) : bl;
  *p++ = sdivide(bl, s1.area, 0.0);
  *p++ = sdivide(bl, s2.area, 0.0);
  *p++ = sdivide(bl, s1.perim, 0.0);
  *p++ = sdivide(bl, s2.perim, 0.0);
  double vbl[GLIA_MAX_THR];
  for (int t = 0; t < T; ++t) vbl[t] = sdivide(ceil(pb.add(1 + t) / 2.0), F.normLen, 0.0);
  for (int t = 0; t < T; ++t) *p++ = lg ? slog(vbl[t], 0) : vbl[t];
  for (int t = 0; t < T; ++t) *p++ = sdivide(vbl[t], bl, 0.0);
  for (int t = 0; t < T; ++t) *p++ = sdivide(vbl[t], s1.perim, 0.0);
  for (int t = 0; t < T; ++t) *p++ = sdivide(vbl[t], s2.perim, 0.0);
  for (int i = 0; i < F.nR; ++i) {
    double l1, x2;
    img_feat(F.log2tab, rx, sx.nReg, F.rBins[i], F.rAdd[i], F.rMax[i], F.rF64[i], g1);
    img_feat(F.log2tab, ry, sy.nReg, F.rBins[i], F.rAdd[i], F.rMax[i], F.rF64[i], g2);
    hist_dists(rx, sx.nReg, ry, sy.nReg, F.rBins[i], F.rAdd[i], l1, x2);  // symmetric in (X,Y)
    *p++ = l1; *p++ = x2;
    *p++ = fabs(g1.entropy - g2.entropy);
    if (F.median) *p++ = med ? fabs(med[3 * i] - med[3 * i + 1]) : 0.0;  // type/feat.hxx:803-805
    *p++ = fabs(g1.mean - g2.mean);
    *p++ = fabs(g1.stddev - g2.stddev); *p++ = fabs(g1.mn - g2.mn); *p++ = fabs(g1.mx - g2.mx);
  }
  for (int i = 0; i < F.nRL; ++i) {
    double l1, x2;
    img_feat(F.log2tab, rx, sx.nReg, F.rlBins[i], F.rlAdd[i], -1, 0, g1);
    img_feat(F.log2tab, ry, sy.nReg, F.rlBins[i], F.rlAdd[i], -1, 0, g2);
    hist_dists(rx, sx.nReg, ry, sy.nReg, F.rlBins[i], F.rlAdd[i], l1, x2);
    *p++ = l1; *p++ = x2; *p++ = fabs(g1.entropy - g2.entropy);
  }
  for (int k = 0; k < F.nB; ++k) {
    img_feat(F.log2tab, pb, nb, F.bBins[k], F.bAdd[k], F.bMax[k], F.bF64[k], g, F.hist ? p : nullptr);
    if (F.hist) p += F.bBins[k];
    if (F.median && med) g.median = med[3 * F.nR + 4 * k + 3];
    p = push_img(F, g, true, p);
  }
  if (sal) {  // bc_feat.hxx:208-213; sal = {X, Y, Z}
    double s1v = swp ? sal[1] : sal[0], s2v = swp ? sal[0] : sal[1];
    double d02 = fabs(s1v - sal[2]), d12 = fabs(s2v - sal[2]);
    *p++ = d02 < d12 ? d02 : d12;
    *p++ = d02 < d12 ? d12 : d02;
  }
  if (part == 0) return;
  const double* mb = med ? med + 3 * F.nR : nullptr;
  if (swp) {
    p = push_region(F, sy, ry, by, sal ? sal + 1 : nullptr, p, med ? med + 1 : nullptr, 3, mb ? mb + 1 : nullptr, 4);
    p = push_region(F, sx, rx, bx, sal ? sal + 0 : nullptr, p, med, 3, mb, 4);
  } else {
    p = push_region(F, sx, rx, bx, sal ? sal + 0 : nullptr, p, med, 3, mb, 4);
    p = push_region(F, sy, ry, by, sal ? sal + 1 : nullptr, p, med ? med + 1 : nullptr, 3, mb ? mb + 1 : nullptr, 4);
  }
  push_region(F, sz, rz, bz, sal ? sal + 2 : nullptr, p, med ? med + 2 : nullptr, 3, mb ? mb + 2 : nullptr, 4);
}

}  // namespace glia
