// section-classifier features of region pairs across two 2D sections (LINK3D).
//
// Replaces (reference, /root/reference/code):
//   sc_feat operation()                               hmt/main_sc_feat.cxx:27-117
//   RegionFeatsWithLocation / RegionPairFeats / SectionClassificationFeats   hmt/sc_feat.hxx:10-170
//   RegionLocationFeats, RegionLocationDiffFeats      type/feat.hxx:290-370
//   RegionAdvShapeFeats2D (+Diff): central moments, scale-invariant and Hu moments, eccentricity
//                                                     type/feat.hxx:191-287, alg/geometry.hxx:44-108
//   RegionSetDiffFeats                                type/feat.hxx:373-440
//   RegionShapeDiffFeats, ImageDiffFeats              type/feat.hxx:124-132, :859-880
//   stats::getOverlap (image against image)           util/image_stats.hxx:308-335
// The reference walks every pixel of every region once more for the centroid, once more for the moments, and both
// images for the overlaps.  What touches pixels here are two small kernels: per-region power sums of bbox-relative
// coordinates up to third order as exact 64-bit integers (the "shape moments" of north_star: from them the moments
// about ANY point follow in closed form), and the overlap counts of the listed pairs.  The RegionFeats part of a row is
// the per-region block the bc path already builds (featurize.cu), under the caller's normalisation.  The per-pair
// assembly is a few hundred flops per row and is done on the host with the reference's own operation order.
#include <algorithm>
#include <math.h>
#include <vector>
#include "handle.cuh"

int glia_region_blocks_cfg(glia_handle* h, const glia::FeatCfg& F, double* dBlk);

namespace glia {

#define SC_NSUM 10  // n, Sx, Sy, Sxx, Sxy, Syy, Sxxx, Sxxy, Sxyy, Syyy (coordinates relative to the region's bbox minimum)

// 2D volumes run as X x 1 x rows: reference y = kernel z.  Max words of a region record: ~xmin at [0], ~ymin at [2].
__global__ void sc_moment_sums_kernel(const uint32_t* __restrict__ labels, const uint32_t* __restrict__ mask,
                                      const uint32_t* __restrict__ rank, const uint32_t* __restrict__ flags,
                                      uint32_t labelBound, const uint32_t* __restrict__ cMax, int nMax, int X, int Y,
                                      unsigned long long* __restrict__ sums) {
  const size_t p = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (p >= (size_t)X * Y) return;
  if (mask && mask[p] == 0) return;
  const uint32_t l = labels[p];
  if (l >= labelBound || !flags[l]) return;
  const uint32_t r = rank[l];
  const unsigned long long x = (uint32_t)(p % X) - (~cMax[(size_t)r * nMax + 0]);
  const unsigned long long y = (uint32_t)(p / X) - (~cMax[(size_t)r * nMax + 2]);
  unsigned long long* s = sums + (size_t)r * SC_NSUM;
  atomicAdd(s + 0, 1ull); atomicAdd(s + 1, x); atomicAdd(s + 2, y);
  atomicAdd(s + 3, x * x); atomicAdd(s + 4, x * y); atomicAdd(s + 5, y * y);
  atomicAdd(s + 6, x * x * x); atomicAdd(s + 7, x * x * y); atomicAdd(s + 8, x * y * y); atomicAdd(s + 9, y * y * y);
}

// overlap counts of the listed pairs: keys = sorted unique (region index in section 0) << 32 | (region index in section 1)
__global__ void sc_overlap_kernel(const uint32_t* __restrict__ lab0, const uint32_t* __restrict__ mask0,
                                  const uint32_t* __restrict__ rank0, const uint32_t* __restrict__ flags0, uint32_t lb0,
                                  const uint32_t* __restrict__ lab1, const uint32_t* __restrict__ mask1,
                                  const uint32_t* __restrict__ rank1, const uint32_t* __restrict__ flags1, uint32_t lb1,
                                  size_t n, const unsigned long long* __restrict__ keys, int nkeys,
                                  unsigned long long* __restrict__ count) {
  const size_t p = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (p >= n) return;
  if ((mask0 && mask0[p] == 0) || (mask1 && mask1[p] == 0)) return;
  const uint32_t a = lab0[p], b = lab1[p];
  if (a == 0 || b == 0 || a >= lb0 || b >= lb1 || !flags0[a] || !flags1[b]) return;  // BG_VAL excluded on both sides
  const unsigned long long k = ((unsigned long long)rank0[a] << 32) | rank1[b];
  int lo = 0, hi = nkeys;
  while (lo < hi) { const int m = (lo + hi) >> 1; if (keys[m] < k) lo = m + 1; else hi = m; }
  if (lo < nkeys && keys[lo] == k) atomicAdd(count + lo, 1ull);
}

}  // namespace glia

using namespace glia;
#define CK(expr) GLIA_CUDA_CHECK(h, expr)

namespace {

const double FEPS_ = 2.22e-16;
inline double sdivide(double a, double b, double d) { return fabs(b) >= FEPS_ ? a / b : d; }
inline double slog(double x, double d) { return x > 0.0 ? log(x) : d; }
inline double ssqrt(double x, double d) { return x >= 0.0 ? sqrt(x) : d; }

struct AdvShape { double centroid[2]; double cm[7], hu[7], ecc; };

// Moments about the point the reference uses - the centroid divided by the normalising length (type/feat.hxx:336-344,
// hmt/sc_feat.hxx:58-60) - from the integer power sums about the bbox minimum (x0, y0): with a = cx - x0, b = cy - y0,
// sum (x' - a)^i (y' - b)^j expands into the sums; long double keeps the cancellation below the reference's own rounding.
void adv_shape(const unsigned long long* s, uint32_t x0, uint32_t y0, double normLen, AdvShape& A) {
  typedef long double ld;
  const ld n = (ld)s[0];
  const double sx = (double)((ld)s[1] + n * (ld)x0), sy = (double)((ld)s[2] + n * (ld)y0);  // exact: integers below 2^53
  const double siz = (double)s[0] * normLen;
  A.centroid[0] = sx / siz; A.centroid[1] = sy / siz;
  const ld a = (ld)A.centroid[0] - (ld)x0, b = (ld)A.centroid[1] - (ld)y0;
  const ld S10 = s[1], S01 = s[2], S20 = s[3], S11 = s[4], S02 = s[5], S30 = s[6], S21 = s[7], S12 = s[8], S03 = s[9];
  double ms[7];
  ms[0] = (double)(S02 - 2 * b * S01 + n * b * b);                                            // m02
  ms[1] = (double)(S03 - 3 * b * S02 + 3 * b * b * S01 - n * b * b * b);                      // m03
  ms[2] = (double)(S11 - a * S01 - b * S10 + n * a * b);                                      // m11
  ms[3] = (double)(S12 - 2 * b * S11 + b * b * S10 - a * S02 + 2 * a * b * S01 - n * a * b * b);  // m12
  ms[4] = (double)(S20 - 2 * a * S10 + n * a * a);                                            // m20
  ms[5] = (double)(S21 - 2 * a * S11 + a * a * S01 - b * S20 + 2 * a * b * S10 - n * a * a * b);  // m21
  ms[6] = (double)(S30 - 3 * a * S20 + 3 * a * a * S10 - n * a * a * a);                      // m30
  const double m00 = (double)s[0], m002 = m00 * m00, m003 = pow(m00, 2.5);
  double q[7];  // alg/geometry.hxx:64-77
  q[0] = sdivide(ms[0], m002, 0.0); q[1] = sdivide(ms[1], m003, 0.0); q[2] = sdivide(ms[2], m002, 0.0);
  q[3] = sdivide(ms[3], m003, 0.0); q[4] = sdivide(ms[4], m002, 0.0); q[5] = sdivide(ms[5], m003, 0.0);
  q[6] = sdivide(ms[6], m003, 0.0);
  if (normLen > 0.0) {  // type/feat.hxx:231-240 (m21 is divided by L^2 there)
    const double l2 = normLen * normLen, l3 = l2 * normLen;
    ms[0] /= l2; ms[1] /= l3; ms[2] /= l2; ms[3] /= l3; ms[4] /= l2; ms[5] /= l2; ms[6] /= l3;
  }
  for (int i = 0; i < 7; ++i) A.cm[i] = ms[i];
  const double m02 = q[0], m03 = q[1], m11 = q[2], m12 = q[3], m20 = q[4], m21 = q[5], m30 = q[6];
  auto sq = [](double v) { return v * v; };  // std::pow(v, 2)
  A.hu[0] = m20 + m02;
  A.hu[1] = sq(m20 - m02) + 4.0 * m11 * m11;
  A.hu[2] = sq(m30 - 3.0 * m12) + sq(3.0 * m21 - m03);
  A.hu[3] = sq(m30 + m12) + sq(m21 + m03);
  A.hu[4] = (m30 - 3.0 * m12) * (m30 + m12) * (sq(m30 + m12) - 3.0 * sq(m21 + m03)) +
            (3.0 * m21 - m03) * (m21 + m03) * (3.0 * sq(m30 + m12) - sq(m21 + m03));
  A.hu[5] = (m20 - m02) * (sq(m30 + m12) - sq(m21 + m03)) + 4.0 * m11 * (m30 + m12) * (m03 + m21);
  A.hu[6] = (3.0 * m21 - m03) * (m12 + m30) * (sq(m30 + m12) - 3.0 * sq(m21 + m03)) -
            (m30 - 3.0 * m12) * (m12 + m03) * (3.0 * sq(m30 + m12) - sq(m21 + m03));
  const double ea = A.cm[4] + A.cm[0], eb = ssqrt(sq(A.cm[4] - A.cm[0]) + 4.0 * A.cm[2] * A.cm[2], 0.0);
  A.ecc = sdivide(ea + eb, ea - eb, 0.0);  // alg/geometry.hxx:100-106
}

struct Section {  // what a row needs of one section, on the host
  glia_handle* h;
  std::vector<uint32_t> lab, cAdd, cMax;
  std::vector<unsigned long long> sums;
  std::vector<double> blk;
};

}  // namespace

extern "C" int glia_b200_sc_row_dim(const glia_handle* h) {
  if (!h) return -1;
  const FeatCfg& F = h->fc;
  return 6 + 1 + 8 + 15 + (3 + F.lenReal()) * (F.nR + F.nRL) + 2 * (F.xiLen() + 15);
}

extern "C" int glia_b200_sc_feat(glia_handle* h0, glia_handle* h1, const uint32_t* pairs, int64_t n64, double norm_area,
                                 double norm_length, int use_log, double* rows) {
  glia_handle* h = h0;
  if (!h0 || !h1 || n64 < 0 || n64 > 0x3fffffff || (n64 > 0 && (!pairs || !rows))) return GLIA_E_ARG;
  if (!h0->built || !h1->built) return GLIA_E_STATE;
  const FeatCfg &F0 = h0->fc, &F1 = h1->fc;
  bool same = h0->prm.ndim == 2 && h1->prm.ndim == 2 && h0->prm.dims[0] == h1->prm.dims[0] && h0->prm.dims[1] == h1->prm.dims[1] &&
              h0->prm.device == h1->prm.device && F0.nR == F1.nR && F0.nRL == F1.nRL && F0.nB == F1.nB && F0.nThr == F1.nThr &&
              F0.hist == F1.hist && !F0.median && !F1.median && !F0.simple && !F1.simple && F0.nRL <= F0.nR &&
              !h0->prm.z_end && !h1->prm.z_end;
  for (int i = 0; same && i < F0.nR; ++i) same = F0.rBins[i] == F1.rBins[i];
  for (int i = 0; same && i < F0.nRL; ++i) same = F0.rlBins[i] == F1.rlBins[i];
  for (int i = 0; same && i < F0.nB; ++i) same = F0.bBins[i] == F1.bBins[i];
  if (!same) { h->err = "sc_feat: two 2D handles of equal extent and feature configuration on one device are needed (no median / simple layout)"; return GLIA_E_ARG; }
  if (n64 == 0) return GLIA_OK;
  CK(cudaSetDevice(h0->prm.device));
  const int n = (int)n64, X = h0->prm.dims[0], Y = h0->prm.dims[1];
  // main_sc_feat.cxx:48-56: the larger volume / diagonal of the two images (equal here) unless given
  if (norm_area <= 0.0) norm_area = (double)X * Y;
  if (norm_length <= 0.0) norm_length = sqrt((double)X * X + (double)Y * Y);
  const int xi = F0.xiLen(), ld = xi + 2;
  Section S[2];
  S[0].h = h0; S[1].h = h1;
  for (int k = 0; k < 2; ++k) {
    glia_handle* g = S[k].h;
    cudaStream_t st = g->stream;
    int s = glia_ensure_full_regions(g);
    if (s != GLIA_OK) { if (g != h) h->err = g->err; return s; }
    const size_t R = (size_t)g->R;
    const RecLayout LR = g->vp.reg;
    FeatCfg F = g->fc;
    F.normArea = norm_area; F.normLen = norm_length; F.useLog = use_log != 0;
    DevBuf &dBlk = g->k5[5], &dSum = g->k5[6];
    CK(dBlk.ensure(8 * (R + 1) * ld)); CK(dSum.ensure(8 * (R + 1) * SC_NSUM));
    s = glia_region_blocks_cfg(g, F, dBlk.as<double>());
    if (s != GLIA_OK) { if (g != h) h->err = g->err; return s; }
    CK(cudaMemsetAsync(dSum.p, 0, 8 * (R + 1) * SC_NSUM, st));
    sc_moment_sums_kernel<<<(unsigned)(((size_t)X * Y + 255) / 256), 256, 0, st>>>(
        g->vp.labels, g->vp.mask, g->rank.as<uint32_t>(), g->flags.as<uint32_t>(), g->vp.labelBound, g->cMax.as<uint32_t>(),
        LR.nMax, X, Y, dSum.as<unsigned long long>());
    g->launches += 2;
    CK(cudaGetLastError());
    S[k].lab.resize(R); S[k].cAdd.resize(R * LR.nAdd); S[k].cMax.resize(R * LR.nMax); S[k].sums.resize(R * SC_NSUM); S[k].blk.resize(R * ld);
    if (R) {
      CK(cudaMemcpyAsync(S[k].lab.data(), g->regLabel.p, 4 * R, cudaMemcpyDeviceToHost, st));
      CK(cudaMemcpyAsync(S[k].cAdd.data(), g->cAdd.p, 4 * R * LR.nAdd, cudaMemcpyDeviceToHost, st));
      CK(cudaMemcpyAsync(S[k].cMax.data(), g->cMax.p, 4 * R * LR.nMax, cudaMemcpyDeviceToHost, st));
      CK(cudaMemcpyAsync(S[k].sums.data(), dSum.p, 8 * R * SC_NSUM, cudaMemcpyDeviceToHost, st));
      CK(cudaMemcpyAsync(S[k].blk.data(), dBlk.p, 8 * R * ld, cudaMemcpyDeviceToHost, st));
    }
    CK(cudaStreamSynchronize(st));
  }
  // ---- the pairs: labels -> region indices; overlap counts of the distinct pairs
  std::vector<int> ia(n), ib(n);
  std::vector<unsigned long long> keys(n);
  for (int i = 0; i < n; ++i) {
    auto f0 = std::lower_bound(S[0].lab.begin(), S[0].lab.end(), pairs[2 * i]);
    auto f1 = std::lower_bound(S[1].lab.begin(), S[1].lab.end(), pairs[2 * i + 1]);
    if (f0 == S[0].lab.end() || *f0 != pairs[2 * i] || f1 == S[1].lab.end() || *f1 != pairs[2 * i + 1]) {
      h->err = "sc_feat: pair " + std::to_string(i) + " names a region that is not in its section";
      return GLIA_E_ORDER;
    }
    ia[i] = (int)(f0 - S[0].lab.begin()); ib[i] = (int)(f1 - S[1].lab.begin());
    keys[i] = ((unsigned long long)ia[i] << 32) | (unsigned)ib[i];
  }
  std::vector<unsigned long long> uniq(keys);
  std::sort(uniq.begin(), uniq.end());
  uniq.erase(std::unique(uniq.begin(), uniq.end()), uniq.end());
  std::vector<unsigned long long> ov(uniq.size());
  {
    cudaStream_t st = h0->stream;
    DevBuf &dK = h0->k5[7], &dC = h0->k5[8];
    CK(dK.ensure(8 * uniq.size())); CK(dC.ensure(8 * uniq.size()));
    CK(cudaMemcpyAsync(dK.p, uniq.data(), 8 * uniq.size(), cudaMemcpyHostToDevice, st));
    CK(cudaMemsetAsync(dC.p, 0, 8 * uniq.size(), st));
    sc_overlap_kernel<<<(unsigned)(((size_t)X * Y + 255) / 256), 256, 0, st>>>(
        h0->vp.labels, h0->vp.mask, h0->rank.as<uint32_t>(), h0->flags.as<uint32_t>(), h0->vp.labelBound,
        h1->vp.labels, h1->vp.mask, h1->rank.as<uint32_t>(), h1->flags.as<uint32_t>(), h1->vp.labelBound,
        (size_t)X * Y, dK.as<unsigned long long>(), (int)uniq.size(), dC.as<unsigned long long>());
    h0->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(ov.data(), dC.p, 8 * uniq.size(), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
  }
  // ---- per-region advanced shape (only the regions that occur)
  std::vector<AdvShape> adv[2];
  std::vector<char> have[2];
  for (int k = 0; k < 2; ++k) { adv[k].resize(S[k].lab.size()); have[k].assign(S[k].lab.size(), 0); }
  auto shape_of = [&](int k, int r) -> const AdvShape& {
    if (!have[k][r]) {
      const int nMax = S[k].h->vp.reg.nMax;
      adv_shape(&S[k].sums[(size_t)r * SC_NSUM], ~S[k].cMax[(size_t)r * nMax + 0], ~S[k].cMax[(size_t)r * nMax + 2], norm_length, adv[k][r]);
      have[k][r] = 1;
    }
    return adv[k][r];
  };
  // ---- rows (hmt/sc_feat.hxx:119-147, :160-166; logs as RegionPairFeats::log)
  const int dim = glia_b200_sc_row_dim(h0);
  std::vector<double> out((size_t)n * dim);
  const bool lg = use_log != 0;
  const int nAdd = h0->vp.reg.nAdd, T = F0.nThr, D = 2;
  for (int i = 0; i < n; ++i) {
    const int r[2] = {ia[i], ib[i]};
    const double* B[2] = {&S[0].blk[(size_t)r[0] * ld], &S[1].blk[(size_t)r[1] * ld]};
    const uint32_t* A[2] = {&S[0].cAdd[(size_t)r[0] * nAdd], &S[1].cAdd[(size_t)r[1] * nAdd]};
    const AdvShape& a0 = shape_of(0, r[0]); const AdvShape& a1 = shape_of(1, r[1]);
    double* p = &out[(size_t)i * dim];
    const double area0 = B[0][xi], area1 = B[1][xi], per0 = B[0][xi + 1], per1 = B[1][xi + 1];
    const double areaDiff = fabs(area0 - area1), perimDiff = fabs(per0 - per1);
    *p++ = lg ? slog(areaDiff, 0.0) : areaDiff; *p++ = sdivide(areaDiff, area0, 0.0); *p++ = sdivide(areaDiff, area1, 0.0);
    *p++ = lg ? slog(perimDiff, 0.0) : perimDiff; *p++ = sdivide(perimDiff, per0, 0.0); *p++ = sdivide(perimDiff, per1, 0.0);
    const double d0 = a0.centroid[0] - a1.centroid[0], d1 = a0.centroid[1] - a1.centroid[1];
    double dist = ssqrt(d0 * d0 + d1 * d1, 0.0);
    if (lg) dist = std::max(0.0, slog(dist, 0.0));
    *p++ = dist;
    const size_t ki = std::lower_bound(uniq.begin(), uniq.end(), keys[i]) - uniq.begin();
    const double o = (double)ov[ki], sd0 = area0 - o, sd1 = area1 - o, sym = sd0 + sd1;  // areas: the normalised ones, as there
    *p++ = lg ? slog(o, 0.0) : o; *p++ = lg ? slog(sd0, 0.0) : sd0; *p++ = lg ? slog(sd1, 0.0) : sd1; *p++ = lg ? slog(sym, 0.0) : sym;
    *p++ = o / area0; *p++ = o / area1; *p++ = sd0 / area0; *p++ = sd1 / area1;
    for (int k = 0; k < 7; ++k) { const double d = fabs(a0.cm[k] - a1.cm[k]); *p++ = lg ? slog(d, 0.0) : d; }
    for (int k = 0; k < 7; ++k) *p++ = fabs(a0.hu[k] - a1.hu[k]);
    *p++ = fabs(a0.ecc - a1.ecc);
    for (int pass = 0; pass < 2; ++pass) {  // imageRegionDiff, then labelImageRegionDiff - which reads region[i] as well (:136-146)
      const int cnt = pass ? F0.nRL : F0.nR;
      int off = 4 + D + 2 * T;
      for (int k = 0; k < cnt; ++k) {
        const int bins = F0.rBins[k];
        const uint32_t n0 = A[0][0], n1 = A[1][0];
        double l1 = 0.0, x2 = 0.0;  // util/stats.hxx:155-185 on the normalised histograms
        for (int b = 0; b < bins; ++b) {
          const double pa = n0 ? A[0][F0.rAdd[k] + b] / (double)n0 : 0.0, pb = n1 ? A[1][F1.rAdd[k] + b] / (double)n1 : 0.0;
          const double d = pa - pb;
          l1 += fabs(d); x2 += d * d / (pa + pb + FEPS_);
        }
        const double* g0 = B[0] + off + (F0.hist ? bins : 0);  // entropy, mean, stddev, min, max
        const double* g1 = B[1] + off + (F0.hist ? bins : 0);
        *p++ = l1; *p++ = x2;
        for (int j = 0; j < 5; ++j) *p++ = fabs(g0[j] - g1[j]);
        off += F0.lenR(k);
      }
    }
    const int first = area0 > area1 ? 1 : 0;  // keep region 0 area <= region 1 area (main_sc_feat.cxx:98-101)
    for (int k = 0; k < 2; ++k) {
      const int w = k ? 1 - first : first;
      const AdvShape& a = w ? a1 : a0;
      for (int j = 0; j < xi; ++j) *p++ = B[w][j];
      for (int j = 0; j < 7; ++j) *p++ = a.cm[j];
      for (int j = 0; j < 7; ++j) *p++ = a.hu[j];
      *p++ = a.ecc;
    }
  }
  if (h0->prm.device_pointers) {
    CK(cudaMemcpyAsync(rows, out.data(), 8 * out.size(), cudaMemcpyHostToDevice, h0->stream));
    CK(cudaStreamSynchronize(h0->stream));
  } else memcpy(rows, out.data(), 8 * out.size());
  return GLIA_OK;
}
