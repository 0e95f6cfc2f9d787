// Initial-candidate rows from per-region blocks (device code shared by featurize.cu and score_fused.cu).
#pragma once
#include "featurize.cuh"

namespace glia {

struct GraphView {
  long long R, E, C;
  const int* cand; const int* ea; const int* eb; const int* rev; const int* row;
  const uint32_t* regLabel;
  const uint32_t* cAdd; const uint32_t* cMax; const double* cF64;        // regions
  const uint32_t* sAdd; const uint32_t* sMax; const double* sF64;        // directed boundaries
  const uint32_t* aAdd; const uint32_t* aMax; const double* aF64;        // per-region boundary aggregate
  const uint32_t* aMax2; const int* aOwn;                                // second best / holder of the best
};


// Row of initial candidate c written to out[0 .. F.dim): the x1 / x2 blocks are copied from the per-region blocks
// (region_blocks_kernel), x0 (pair terms) and x3 (merged region) are computed here.  parts: bit 0 = x0 and the
// x1 block, bit 1 = the x2 block and x3 - two threads can build one row (disjoint ranges).
__device__ __forceinline__ void candidate_row_from_blocks(const FeatCfg& F, const GraphView& G, int xi,
                                                          const double* __restrict__ blk,
                                                          const unsigned char* __restrict__ orient, long long c,
                                                          double* out, int parts, const double* __restrict__ medC = nullptr) {
  // medC (F.median): this candidate's medians [Z region per region image][boundary-of Z per boundary image][pair
  // boundary per boundary image] (median.cu); those of X and Y are already inside their blocks.
  const int e = G.cand[c], rv = G.rev[e];
  const int a = G.ea[e], b = G.eb[e];
  const RecLayout LR = F.reg, LE = F.edge;
  RecSingle rx{G.cAdd + (size_t)a * LR.nAdd, G.cMax + (size_t)a * LR.nMax, G.cF64 + (size_t)a * LR.nF64};
  RecSingle ry{G.cAdd + (size_t)b * LR.nAdd, G.cMax + (size_t)b * LR.nMax, G.cF64 + (size_t)b * LR.nF64};
  RecSum2 rz{rx, ry};
  RecSingle pe{G.sAdd + (size_t)e * LE.nAdd, G.sMax + (size_t)e * LE.nMax, G.sF64 + (size_t)e * LE.nF64};
  RecSingle pr{G.sAdd + (size_t)rv * LE.nAdd, G.sMax + (size_t)rv * LE.nMax, G.sF64 + (size_t)rv * LE.nF64};
  RecSum2 pb{pe, pr};  // pair boundary b(X,Y) = (a->b) u (b->a)   (util/struct.hxx:10-16)
  const int ld = xi + 2;
  const double* A = blk + (size_t)a * ld;
  const double* B = blk + (size_t)b * ld;
  // strict >, hmt/main_merge_order_bc.cxx:77-80; equal areas keep the reference's call order (orient.cu)
  const bool swp = A[xi] > B[xi] || (orient && orient[c] && A[xi] == B[xi]);
  const double* B1 = swp ? B : A;
  const double* B2 = swp ? A : B;
  const double a1 = B1[xi], a2 = B2[xi], p1 = B1[xi + 1], p2 = B2[xi + 1];
  const int x0len = F.dim - 3 * xi;
  if (parts & 1) {
    const int T = F.nThr, lg = F.useLog;
    const uint32_t nb = pb.add(0), nX = rx.add(0), nY = ry.add(0);
    double* p = out;
    const double areaDiff = fabs(a1 - a2), perimDiff = fabs(p1 - p2);
    const double bl = sdivide(ceil(nb / 2.0), F.normLen, 0.0);
    *p++ = lg ? slog(areaDiff, 0) : areaDiff;
    *p++ = sdivide(areaDiff, a1, 0.0);
    *p++ = sdivide(areaDiff, a2, 0.0);
    *p++ = lg ? slog(perimDiff, 0) : perimDiff;
    *p++ = sdivide(perimDiff, p1, 0.0);
    *p++ = sdivide(perimDiff, p2, 0.0);
    *p++ = lg ? slog(bl, 0) : bl;
    *p++ = sdivide(bl, a1, 0.0);
    *p++ = sdivide(bl, a2, 0.0);
    *p++ = sdivide(bl, p1, 0.0);
    *p++ = sdivide(bl, p2, 0.0);
    double vbl[GLIA_MAX_THR];
    for (int t = 0; t < T; ++t) vbl[t] = sdivide(ceil(pb.add(1 + t) / 2.0), F.normLen, 0.0);
    for (int t = 0; t < T; ++t) *p++ = lg ? slog(vbl[t], 0) : vbl[t];
    for (int t = 0; t < T; ++t) *p++ = sdivide(vbl[t], bl, 0.0);
    for (int t = 0; t < T; ++t) *p++ = sdivide(vbl[t], p1, 0.0);
    for (int t = 0; t < T; ++t) *p++ = sdivide(vbl[t], p2, 0.0);
    int off = 4 + F.D + 2 * T;  // first image feature inside a block
    const int nreal = 1 + F.lenReal();  // entropy, [median], mean, stddev, min, max
    for (int i = 0; i < F.nR; ++i) {
      double l1, x2;
      const double* g1 = A + off + (F.hist ? F.rBins[i] : 0);
      const double* g2 = B + off + (F.hist ? F.rBins[i] : 0);
      hist_dists(rx, nX, ry, nY, F.rBins[i], F.rAdd[i], l1, x2);
      *p++ = l1; *p++ = x2;
      for (int k = 0; k < nreal; ++k) *p++ = fabs(g1[k] - g2[k]);
      off += F.lenR(i);
    }
    for (int i = 0; i < F.nRL; ++i) {
      double l1, x2;
      hist_dists(rx, nX, ry, nY, F.rlBins[i], F.rlAdd[i], l1, x2);
      const int eo = off + (F.hist ? F.rlBins[i] : 0);
      *p++ = l1; *p++ = x2; *p++ = fabs(A[eo] - B[eo]);
      off += F.lenRL(i);
    }
    ImgFeat g;
    for (int k = 0; k < F.nB; ++k) {
      img_feat(F.log2tab, pb, nb, F.bBins[k], F.bAdd[k], F.bMax[k], F.bF64[k], g, F.hist ? p : nullptr);
      if (F.hist) p += F.bBins[k];
      if (F.median && medC) g.median = medC[F.nR + F.nB + k];
      p = push_img(F, g, true, p);
    }
    for (int k = 0; k < xi; ++k) out[x0len + k] = B1[k];
  }
  if (parts & 2) {
    for (int k = 0; k < xi; ++k) out[x0len + xi + k] = B2[k];
    RecBoundaryZ bz;
    bz.pa = G.aAdd + (size_t)a * LE.nAdd; bz.qa = G.aAdd + (size_t)b * LE.nAdd; bz.ea = pe.a; bz.ra = pr.a;
    // entries(Z) = all boundaries of a and of b except the mutual pair (region.hxx:66-75).  Instead of
    // walking the ~2 deg boundary records per candidate: max words from the per-region best / second best
    // (the excluded boundary may hold the best), sums as aggregate - pair (exact for dyadic values, ~1e-13
    // relative otherwise; non-finite totals take the walk so that inf - inf cannot appear).
    bool walk = false;
    for (int i = 0; i < LE.nF64; ++i) {
      const double ta = G.aF64[(size_t)a * LE.nF64 + i], tb = G.aF64[(size_t)b * LE.nF64 + i];
      walk = walk || !isfinite(ta) || !isfinite(tb);
      bz.f[i] = (ta - pe.f[i]) + (tb - pr.f[i]);
    }
    for (int i = 0; i < LE.nMax; ++i) {
      const size_t ia = (size_t)a * LE.nMax + i, ib = (size_t)b * LE.nMax + i;
      const uint32_t ma = G.aOwn[ia] == e ? G.aMax2[ia] : G.aMax[ia];
      const uint32_t mb = G.aOwn[ib] == rv ? G.aMax2[ib] : G.aMax[ib];
      bz.mxw[i] = max(ma, mb);
    }
    if (walk) {
      for (int i = 0; i < LE.nF64; ++i) bz.f[i] = 0.0;
      for (int side = 0; side < 2; ++side) {
        int r = side ? b : a;
        for (int k = G.row[r]; k < G.row[r + 1]; ++k) {
          if (k == e || k == rv) continue;
          for (int i = 0; i < LE.nF64; ++i) bz.f[i] += G.sF64[(size_t)k * LE.nF64 + i];
        }
      }
    }
    ShapeFeat sz;
    shape_feat(F, rz, bz, sz);
    push_region(F, sz, rz, bz, nullptr, out + x0len + 2 * xi, medC, 1, medC ? medC + F.nR : nullptr, 1);
  }
}

}  // namespace glia
