// K6 driver: featurise the initial merge candidates (mutual base pairs) on the device.
//
// Replaces TBoundaryTable::init's initFb calls (type/boundary_table.hxx:91-114,
// util/struct_merge_bc.hxx:18-23) and fBcFeat (hmt/main_merge_order_bc.cxx:54-95): where the
// reference builds a temporary merged region and re-walks every pixel of r0, r1 and r0 u r1,
// a candidate row here is a pure function of the records of a, b, their boundary aggregates
// and the two directed boundaries a->b, b->a.
#include "featurize.cuh"
#include "featurize_init.cuh"
#include "median.cuh"

namespace glia {

__global__ void __launch_bounds__(128)
init_candidates_kernel(const __grid_constant__ FeatCfg F, const __grid_constant__ GraphView G,
                       const unsigned char* __restrict__ orient, uint32_t* __restrict__ keys, double* __restrict__ rows,
                       const double* __restrict__ medReg, const double* __restrict__ medCand) {
  long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (c >= G.C) return;
  const int e = G.cand[c], rv = G.rev[e];
  const int a = G.ea[e], b = G.eb[e];
  if (keys) { keys[2 * c] = G.regLabel[a]; keys[2 * c + 1] = G.regLabel[b]; }
  if (!rows) return;
  const RecLayout LR = F.reg, LE = F.edge;
  RecSingle rx{G.cAdd + (size_t)a * LR.nAdd, G.cMax + (size_t)a * LR.nMax, G.cF64 + (size_t)a * LR.nF64};
  RecSingle ry{G.cAdd + (size_t)b * LR.nAdd, G.cMax + (size_t)b * LR.nMax, G.cF64 + (size_t)b * LR.nF64};
  RecSum2 rz{rx, ry};
  RecSingle bx{G.aAdd + (size_t)a * LE.nAdd, G.aMax + (size_t)a * LE.nMax, G.aF64 + (size_t)a * LE.nF64};
  RecSingle by{G.aAdd + (size_t)b * LE.nAdd, G.aMax + (size_t)b * LE.nMax, G.aF64 + (size_t)b * LE.nF64};
  RecSingle pe{G.sAdd + (size_t)e * LE.nAdd, G.sMax + (size_t)e * LE.nMax, G.sF64 + (size_t)e * LE.nF64};
  RecSingle pr{G.sAdd + (size_t)rv * LE.nAdd, G.sMax + (size_t)rv * LE.nMax, G.sF64 + (size_t)rv * LE.nF64};
  RecSum2 pb{pe, pr};  // pair boundary b(X,Y) = (a->b) u (b->a)   (util/struct.hxx:10-16)
  RecBoundaryZ bz;
  bz.pa = bx.a; bz.qa = by.a; bz.ea = pe.a; bz.ra = pr.a;
  for (int i = 0; i < LE.nMax; ++i) bz.mxw[i] = 0;
  for (int i = 0; i < LE.nF64; ++i) bz.f[i] = 0.0;
  // entries(Z): all directed boundaries of a then of b except the mutual pair (region.hxx:66-75)
  for (int side = 0; side < 2; ++side) {
    int r = side ? b : a;
    for (int k = G.row[r]; k < G.row[r + 1]; ++k) {
      if (k == e || k == rv) continue;
      for (int i = 0; i < LE.nMax; ++i) bz.mxw[i] = max(bz.mxw[i], G.sMax[(size_t)k * LE.nMax + i]);
      for (int i = 0; i < LE.nF64; ++i) bz.f[i] += G.sF64[(size_t)k * LE.nF64 + i];
    }
  }
  double med[7 * GLIA_MAX_IMG];  // featurize_row's layout, from the per-region and per-candidate medians (median.cu)
  if (F.median && medReg && medCand) {
    const double* ma = medReg + (size_t)a * (F.nR + F.nB);
    const double* mb = medReg + (size_t)b * (F.nR + F.nB);
    const double* mc = medCand + (size_t)c * (F.nR + 2 * F.nB);
    for (int i = 0; i < F.nR; ++i) { med[3 * i] = ma[i]; med[3 * i + 1] = mb[i]; med[3 * i + 2] = mc[i]; }
    for (int j = 0; j < F.nB; ++j) {
      double* m = med + 3 * F.nR + 4 * j;
      m[0] = ma[F.nR + j]; m[1] = mb[F.nR + j]; m[2] = mc[F.nR + j]; m[3] = mc[F.nR + F.nB + j];
    }
  }
  featurize_row(F, rx, ry, rz, bx, by, bz, pb, nullptr, rows + (size_t)c * F.dim, -1, orient && orient[c],
                (F.median && medReg && medCand) ? med : nullptr);
}

// ---- v2: the x1 / x2 blocks of a row depend on ONE base region only, and each region takes part in
// ~2 deg candidates, so they are computed once per region (region_blocks_kernel) and copied; a
// candidate thread is left with x0 (differences, pair boundary) and x3 (merged region).  Same
// device functions, same operation order: rows are bit-identical to init_candidates_kernel's
// (3 instead of 9 entropy evaluations per candidate; log2 was ~60 % of the v1 kernel,
// profiles/r1_v3_ncu_c4s.txt).
// block layout: [0, xi) = push_region output, [xi] = area, [xi+1] = perim (normalised, before --logs)
__global__ void __launch_bounds__(128)
region_blocks_kernel(const __grid_constant__ FeatCfg F, const __grid_constant__ GraphView G, int xi,
                     double* __restrict__ blk, const double* __restrict__ medReg) {
  long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (r >= G.R) return;
  const RecLayout LR = F.reg, LE = F.edge;
  RecSingle rx{G.cAdd + (size_t)r * LR.nAdd, G.cMax + (size_t)r * LR.nMax, G.cF64 + (size_t)r * LR.nF64};
  RecSingle bx{G.aAdd + (size_t)r * LE.nAdd, G.aMax + (size_t)r * LE.nMax, G.aF64 + (size_t)r * LE.nF64};
  ShapeFeat s;
  shape_feat(F, rx, bx, s);
  double* o = blk + (size_t)r * (xi + 2);
  const double* m = medReg ? medReg + (size_t)r * (F.nR + F.nB) : nullptr;
  push_region(F, s, rx, bx, nullptr, o, m, 1, m ? m + F.nR : nullptr, 1);
  o[xi] = s.area; o[xi + 1] = s.perim;
}

// One thread per candidate.  The kernel is bound by the latency of its scattered record loads (long-scoreboard stalls, 16 %
// issue utilisation at 24 warps / SM: profiles/r2_ncu_featurize_c4s.txt); splitting a row over two threads (more loads in
// flight) was measured SLOWER (C4 7.6 -> 9.0 ms: both halves re-read the shared records and the row writes interleave).
__global__ void __launch_bounds__(128)
init_candidates_blocks_kernel(const __grid_constant__ FeatCfg F, const __grid_constant__ GraphView G, int xi,
                              const double* __restrict__ blk, const unsigned char* __restrict__ orient,
                              uint32_t* __restrict__ keys, double* __restrict__ rows, const double* __restrict__ medCand) {
  long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (c >= G.C) return;
  if (keys) { const int e = G.cand[c]; keys[2 * c] = G.regLabel[G.ea[e]]; keys[2 * c + 1] = G.regLabel[G.eb[e]]; }
  if (!rows) return;
  candidate_row_from_blocks(F, G, xi, blk, orient, c, rows + (size_t)c * F.dim, 3,
                            medCand ? medCand + (size_t)c * (F.nR + 2 * F.nB) : nullptr);
}

}  // namespace glia

using namespace glia;

GraphView glia_graph_view(const glia_handle* h) {
  GraphView G;
  G.R = h->R; G.E = h->E; G.C = h->C;
  G.cand = h->cand.as<int>(); G.ea = h->ea.as<int>(); G.eb = h->eb.as<int>();
  G.rev = h->rev.as<int>(); G.row = h->row.as<int>();
  G.regLabel = h->regLabel.as<uint32_t>();
  G.cAdd = h->cAdd.as<uint32_t>(); G.cMax = h->cMax.as<uint32_t>(); G.cF64 = h->cF64.as<double>();
  G.sAdd = h->sAdd.as<uint32_t>(); G.sMax = h->sMax.as<uint32_t>(); G.sF64 = h->sF64.as<double>();
  G.aAdd = h->bAggAdd.as<uint32_t>(); G.aMax = h->bAggMax.as<uint32_t>(); G.aF64 = h->bAggF64.as<double>();
  G.aMax2 = h->bAggMax2.as<uint32_t>(); G.aOwn = h->bAggOwn.as<int>();
  return G;
}

// per-region x1 / x2 blocks (once per build; the candidates of a region copy them)
int glia_region_blocks_device(glia_handle* h, int* xiOut) {
  const glia_params& c = h->prm;
  const int xi = h->fc.xiLen();
  if (xiOut) *xiOut = xi;
  if (h->R == 0) return GLIA_OK;
  GraphView G = glia_graph_view(h);
  GLIA_CUDA_CHECK(h, h->regBlocks.ensure(8 * (size_t)(h->R + 1) * (xi + 2)));
  region_blocks_kernel<<<(unsigned)((h->R + 127) / 128), 128, 0, h->stream>>>(h->fc, G, xi, h->regBlocks.as<double>(),
                                                                              h->fc.median ? h->medReg.as<double>() : nullptr);
  h->launches++;
  GLIA_CUDA_CHECK(h, cudaGetLastError());
  return GLIA_OK;
}

// The same blocks under another feature configuration (normalisation / --logs of the caller: sc_feat.cu), into dBlk
// [R][F.xiLen() + 2]; the handle's own blocks are left alone.
int glia_region_blocks_cfg(glia_handle* h, const glia::FeatCfg& F, double* dBlk) {
  if (h->R == 0) return GLIA_OK;
  GraphView G = glia_graph_view(h);
  region_blocks_kernel<<<(unsigned)((h->R + 127) / 128), 128, 0, h->stream>>>(F, G, F.xiLen(), dBlk, nullptr);
  h->launches++;
  GLIA_CUDA_CHECK(h, cudaGetLastError());
  return GLIA_OK;
}

// dKeys [C][2] or null, dRows [C][dim] or null: device buffers
int glia_featurize_init_device(glia_handle* h, uint32_t* dKeys, double* dRows) {
  if (h->C == 0) return GLIA_OK;
  GraphView G = glia_graph_view(h);
  unsigned blocks = (unsigned)((h->C + 127) / 128);
  const glia_params& c = h->prm;
  // orientation bits of equal-area pairs are only valid for the build they were installed after
  const unsigned char* orient = (h->hasOrient && h->orientC == h->C) ? h->orient.as<unsigned char>() : nullptr;
  const int xi = h->fc.xiLen();
  const bool med = h->fc.median && dRows;
  if (med) { int st = glia_median_init_candidates(h); if (st != GLIA_OK) return st; }
  if (h->fc.simple || !dRows) {
    init_candidates_kernel<<<blocks, 128, 0, h->stream>>>(h->fc, G, orient, dKeys, dRows,
                                                          med ? h->medReg.as<double>() : nullptr, med ? h->medCand.as<double>() : nullptr);
  } else {
    int st = glia_region_blocks_device(h, nullptr);
    if (st != GLIA_OK) return st;
    init_candidates_blocks_kernel<<<blocks, 128, 0, h->stream>>>(h->fc, G, xi, h->regBlocks.as<double>(), orient, dKeys, dRows,
                                                                 med ? h->medCand.as<double>() : nullptr);
  }
  h->launches++;
  GLIA_CUDA_CHECK(h, cudaGetLastError());
  return GLIA_OK;
}
