// C ABI of glia_b200 (include/glia_b200.h): handle life cycle, the build pipeline
// (voxel pass -> graph), dumps for parity tests, candidate featurisation and scoring.
#include <math.h>
#include <string.h>
#include <algorithm>
#include <vector>
#include "handle.cuh"

namespace glia {
cudaError_t launch_label_max(const uint32_t* labels, size_t n, uint32_t* out, cudaStream_t st);
}  // namespace glia
int glia_finalize_graph(glia_handle* h);
int glia_featurize_init_device(glia_handle* h, uint32_t* dKeys, double* dRows);
int glia_rf_label_index(const glia_rf* m, int label);
int glia_rf_score_device(const glia_rf* m, const double* dX, size_t N, int D, int labelIdx,
                         int* dVotes, double* dProb, cudaStream_t st);

using namespace glia;

#define CK(expr) GLIA_CUDA_CHECK(h, expr)

static float f32_ceil(double b) {  // smallest float >= b
  float f = (float)b;
  if ((double)f < b) f = nextafterf(f, INFINITY);
  return f;
}
static float f32_floor(double b) {  // largest float <= b
  float f = (float)b;
  if ((double)f > b) f = nextafterf(f, -INFINITY);
  return f;
}
static uint64_t pow2ceil(uint64_t v) { uint64_t p = 1; while (p < v) p <<= 1; return p; }
static int pow2floor(long long v) { int p = 1; while ((long long)p * 2 <= v) p <<= 1; return p; }

static void fill_channel(Channel& c, const glia_image_spec& s, int addOff, int maxOff, int f64Off) {
  memset(&c, 0, sizeof(c));
  c.data = nullptr;
  c.bins = s.bins; c.addOff = addOff; c.maxOff = maxOff; c.f64Off = f64Off;
  c.lo_dn = f32_floor(s.lo); c.hi_up = f32_ceil(s.hi);
  // util/image_stats.hxx:17-22: bounds[0] = interval, bounds[i] = bounds[i-1] + interval (no lo offset)
  double interval = (s.hi - s.lo) / s.bins, b = interval;
  for (int i = 0; i < s.bins; ++i) { c.bound_up[i] = f32_ceil(b); b += interval; }
}

extern "C" int glia_b200_create(const glia_params* p, glia_handle** out) {
  if (!p || !out) return GLIA_E_ARG;
  *out = nullptr;
  if (p->ndim != 2 && p->ndim != 3) return GLIA_E_ARG;
  if (p->n_rimg < 0 || p->n_rimg > GLIA_MAX_IMG || p->n_rlimg < 0 || p->n_rlimg > GLIA_MAX_IMG ||
      p->n_bimg < 0 || p->n_bimg > GLIA_MAX_IMG || p->n_thr < 0 || p->n_thr > GLIA_MAX_THR)
    return GLIA_E_ARG;
  for (int d = 0; d < p->ndim; ++d) if (p->dims[d] <= 0) return GLIA_E_ARG;
  for (int i = 0; i < p->n_rimg; ++i) if (p->rimg[i].bins < 1 || p->rimg[i].bins > GLIA_MAX_BINS) return GLIA_E_ARG;
  for (int i = 0; i < p->n_rlimg; ++i) if (p->rlimg[i].bins < 1 || p->rlimg[i].bins > GLIA_MAX_BINS) return GLIA_E_ARG;
  for (int i = 0; i < p->n_bimg; ++i) if (p->bimg[i].bins < 1 || p->bimg[i].bins > GLIA_MAX_BINS) return GLIA_E_ARG;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || p->device < 0 || p->device >= ndev)
    return GLIA_E_CUDA;  // no CPU fallback
  if (cudaSetDevice(p->device) != cudaSuccess) return GLIA_E_CUDA;
  glia_handle* h = new glia_handle();
  h->prm = *p;
  if (p->ndim == 2) h->prm.dims[2] = 1;
  if (p->stream) h->stream = (cudaStream_t)p->stream;
  else { if (cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess) { delete h; return GLIA_E_CUDA; } h->ownStream = true; }
  for (int i = 0; i < 8; ++i) cudaEventCreate(&h->ev[i]);
  // kernel axes: 2D images run as X x 1 x rows so that one kernel serves both dimensions
  if (p->ndim == 2) { h->KX = p->dims[0]; h->KY = 1; h->KZ = p->dims[1]; }
  else { h->KX = p->dims[0]; h->KY = p->dims[1]; h->KZ = p->dims[2]; }
  int zb = 0, ze = h->KZ, hl = 0, hh = 0;
  if (p->z_end > 0 && p->use_median_feat) { glia_b200_destroy(h); return GLIA_E_ARG; }  // medians need the whole volume (median.cu)
  if (p->z_end > 0) {
    if (p->ndim != 3 || p->z_begin < 0 || p->z_end > h->KZ || p->z_begin >= p->z_end ||
        p->halo_lo < 0 || p->halo_hi < 0 || (p->z_begin > 0 && p->halo_lo < 1) ||
        (p->z_end < h->KZ && p->halo_hi < 1) || p->z_begin - p->halo_lo < 0 ||
        p->z_end + p->halo_hi > h->KZ) {
      glia_b200_destroy(h); return GLIA_E_ARG;
    }
    zb = p->z_begin; ze = p->z_end; hl = p->halo_lo; hh = p->halo_hi;
  }
  VoxParams& P = h->vp;
  memset(&P, 0, sizeof(P));
  memset(&h->vc, 0, sizeof(h->vc));
  P.contourOnly = p->contour_only != 0;
  P.ndim = p->ndim; P.X = h->KX; P.Y = h->KY; P.Z = h->KZ;
  P.zBegin = zb; P.zEnd = ze; P.zBuf0 = zb - hl; P.zBufN = (ze + hh) - (zb - hl);
  h->nvoxBuf = (size_t)P.X * P.Y * P.zBufN;
  // 2D rows of 128 + 2 halo voxels stay inside the 256-element box limit of a TMA tensor copy (voxel_pass.cu)
  if (p->ndim == 2) { P.TX = std::min(128, ((P.X + 31) / 32) * 32); P.TY = 1; P.TZ = 16; }
  else { P.TX = 32; P.TY = 8; P.TZ = 16; }
  P.SXp = P.TX + 8;  // 4 voxels of padding on either side: 16-byte aligned tile rows (voxel_pass.cu)
  P.tilesX = (P.X + P.TX - 1) / P.TX; P.tilesY = (P.Y + P.TY - 1) / P.TY;
  P.tilesZ = (ze - zb + P.TZ - 1) / P.TZ;
  P.nThr = p->n_thr;
  for (int t = 0; t < p->n_thr; ++t) P.thr_up[t] = f32_ceil(p->thr[t]);
  // record layouts
  FeatCfg& F = h->fc;
  memset(&F, 0, sizeof(F));
  int add = 2, mx = 6, f64 = 0, ch = 0;
  for (int i = 0; i < p->n_rimg; ++i, ++ch) {
    fill_channel(h->vc.reg[ch], p->rimg[i], add, mx, f64);
    F.rBins[i] = p->rimg[i].bins; F.rAdd[i] = add; F.rMax[i] = mx; F.rF64[i] = f64;
    add += p->rimg[i].bins; mx += 2; f64 += 2;
  }
  for (int i = 0; i < p->n_rlimg; ++i, ++ch) {
    fill_channel(h->vc.reg[ch], p->rlimg[i], add, mx, f64);
    F.rlBins[i] = p->rlimg[i].bins; F.rlAdd[i] = add;
    add += p->rlimg[i].bins; mx += 2; f64 += 2;
  }
  P.nRegCh = ch; P.reg.nAdd = add; P.reg.nMax = mx; P.reg.nF64 = f64;
  add = 1 + p->n_thr; mx = 0; f64 = 0;
  for (int i = 0; i < p->n_bimg; ++i) {
    fill_channel(h->vc.edge[i], p->bimg[i], add, mx, f64);
    F.bBins[i] = p->bimg[i].bins; F.bAdd[i] = add; F.bMax[i] = mx; F.bF64[i] = f64;
    add += p->bimg[i].bins; mx += 2; f64 += 2;
  }
  P.nEdgeCh = p->n_bimg; P.edge.nAdd = add; P.edge.nMax = mx; P.edge.nF64 = f64;
  // shared-memory tables: ~10 KB for regions, ~30 KB for directed boundaries
  long long regBytes = 4 + 4LL * (P.reg.nAdd + P.reg.nMax) + 8LL * P.reg.nF64;
  long long edgeBytes = 8 + 4LL * (P.edge.nAdd + P.edge.nMax) + 8LL * P.edge.nF64;
  P.regSlots = std::max(8, std::min(128, pow2floor(10240 / regBytes)));
  P.edgeSlots = std::max(16, std::min(1024, pow2floor(30720 / edgeBytes)));
  // feature configuration
  F.D = p->ndim;
  F.axis[0] = 0; F.axis[1] = (p->ndim == 2) ? 2 : 1; F.axis[2] = 2;
  F.nThr = p->n_thr; F.nR = p->n_rimg; F.nRL = p->n_rlimg; F.nB = p->n_bimg;
  F.reg = P.reg; F.edge = P.edge;
  double vol = 1.0, diag = 0.0;  // util/image.hxx:173-192
  for (int d = 0; d < p->ndim; ++d) { vol *= (double)p->dims[d]; diag += (double)p->dims[d] * p->dims[d]; }
  F.normArea = p->normalize ? vol : 1.0;
  F.normLen = p->normalize ? sqrt(diag) : 1.0;
  F.useLog = p->use_log; F.simple = p->simple;
  F.hist = p->use_hist_feat != 0; F.median = p->use_median_feat != 0;
  {  // log2 of small counts (entropy terms, featurize.cuh)
    std::vector<double> tab(4096);
    tab[0] = 0.0;
    for (int c = 1; c < 4096; ++c) tab[c] = log2((double)c);
    if (h->log2tab.ensure(8 * tab.size()) != cudaSuccess ||
        cudaMemcpy(h->log2tab.p, tab.data(), 8 * tab.size(), cudaMemcpyHostToDevice) != cudaSuccess) { glia_b200_destroy(h); return GLIA_E_CUDA; }
    F.log2tab = h->log2tab.as<double>();
  }
  *out = h;
  F.dim = glia_b200_row_dim(h, 0);
  return GLIA_OK;
}

extern "C" void glia_b200_destroy(glia_handle* h) {
  if (!h) return;
  cudaSetDevice(h->prm.device);
  glia::DevBuf* bufs[] = {&h->inLabels, &h->inMask, &h->inPb, &h->regAdd, &h->regMax, &h->regF64,
                          &h->edgeKeys, &h->edgeAdd, &h->edgeMax, &h->edgeF64, &h->flags, &h->rank,
                          &h->regLabel, &h->cAdd, &h->cMax, &h->cF64, &h->eSlot, &h->eKey,
                          &h->eSlotSorted, &h->eKeySorted, &h->sAdd, &h->sMax, &h->sF64, &h->ea,
                          &h->eb, &h->row, &h->rev, &h->cand, &h->bAggAdd, &h->bAggMax, &h->bAggF64,
                          &h->cubTmp, &h->scalars, &h->rows, &h->sal, &h->keysOut, &h->interest,
                          &h->exportBuf, &h->exportCursor, &h->regBlocks, &h->bAggMax2, &h->bAggOwn,
                          &h->orient, &h->tieBuf, &h->firstVox, &h->log2tab, &h->medRStart, &h->medEStart,
                          &h->medTmpA, &h->medTmpB, &h->medReg, &h->medCand};
  for (auto b : bufs) b->release();
  for (auto& b : h->inImg) b.release();
  for (auto& b : h->medRKeys) b.release();
  for (auto& b : h->medEKeys) b.release();
  for (auto& b : h->medQ) b.release();
  for (auto& b : h->k5) b.release();
  for (int i = 0; i < 8; ++i) cudaEventDestroy(h->ev[i]);
  if (h->copyStream) { for (int i = 0; i < 4; ++i) cudaEventDestroy(h->evCopy[i]); cudaStreamDestroy(h->copyStream); }
  for (int i = 0; i < 2; ++i) { if (h->pin[i]) cudaFreeHost(h->pin[i]); if (h->evPin[i]) cudaEventDestroy(h->evPin[i]); }
  if (h->ownStream) cudaStreamDestroy(h->stream);
  delete h;
}

// A cudaMemcpy into pageable memory goes through the driver's small staging buffer (~3 GB/s measured for 81 MB of
// bc_feat rows); two pinned 8 MB buffers filled asynchronously while the host drains the other one reach PCIe speed.
int glia_copy_to_host(glia_handle* h, void* dst, const void* dsrc, size_t bytes) {
  const size_t CH = 8u << 20;
  cudaStream_t st = h->stream;
  if (bytes <= CH) {
    GLIA_CUDA_CHECK(h, cudaMemcpyAsync(dst, dsrc, bytes, cudaMemcpyDeviceToHost, st));
    GLIA_CUDA_CHECK(h, cudaStreamSynchronize(st));
    return GLIA_OK;
  }
  for (int i = 0; i < 2; ++i) {
    if (!h->pin[i]) GLIA_CUDA_CHECK(h, cudaMallocHost(&h->pin[i], CH));
    if (!h->evPin[i]) GLIA_CUDA_CHECK(h, cudaEventCreateWithFlags(&h->evPin[i], cudaEventDisableTiming));
  }
  const size_t nch = (bytes + CH - 1) / CH;
  auto len = [&](size_t k) { return std::min(CH, bytes - k * CH); };
  GLIA_CUDA_CHECK(h, cudaMemcpyAsync(h->pin[0], dsrc, len(0), cudaMemcpyDeviceToHost, st));
  GLIA_CUDA_CHECK(h, cudaEventRecord(h->evPin[0], st));
  for (size_t k = 0; k < nch; ++k) {
    if (k + 1 < nch) {
      GLIA_CUDA_CHECK(h, cudaMemcpyAsync(h->pin[(k + 1) & 1], (const char*)dsrc + (k + 1) * CH, len(k + 1), cudaMemcpyDeviceToHost, st));
      GLIA_CUDA_CHECK(h, cudaEventRecord(h->evPin[(k + 1) & 1], st));
    }
    GLIA_CUDA_CHECK(h, cudaEventSynchronize(h->evPin[k & 1]));
    memcpy((char*)dst + k * CH, h->pin[k & 1], len(k));
  }
  return GLIA_OK;
}

// Re-points one configured image at a new buffer of the same extent (a handle is reused for a
// stream of slices/volumes: only the data changes, the histogram specs stay).
extern "C" int glia_b200_set_image(glia_handle* h, int kind, int i, const float* data) {
  if (!h || !data || i < 0) return GLIA_E_ARG;
  glia_params& p = h->prm;
  if (kind == 0 && i < p.n_rimg) p.rimg[i].data = data;
  else if (kind == 1 && i < p.n_rlimg) p.rlimg[i].data = data;
  else if (kind == 2 && i < p.n_bimg) p.bimg[i].data = data;
  else return GLIA_E_ARG;
  return GLIA_OK;
}

// Tuning / test hooks of a handle (no environment variables are read on the data path).
extern "C" int glia_b200_tune(glia_handle* h, const char* key, int value) {
  if (!h || !key) return GLIA_E_ARG;
  if (!strcmp(key, "generic")) h->forceGeneric = value != 0;
  else if (!strcmp(key, "notma")) h->noTma = value != 0;
  else if (!strcmp(key, "used_tma")) return h->usedTma ? GLIA_OK : GLIA_E_STATE;  // query: did the last generic pass stage by TMA
  else if (!strcmp(key, "tz")) { if (value < 0 || value > 31) return GLIA_E_ARG; h->tuneTZ = value; }
  else if (!strcmp(key, "slots")) { if (value < 0 || value > 2048) return GLIA_E_ARG; h->tuneSlots = value; }
  else return GLIA_E_ARG;
  return GLIA_OK;
}

extern "C" const char* glia_b200_last_error(const glia_handle* h) { return h ? h->err.c_str() : "null handle"; }

extern "C" int glia_b200_row_dim(const glia_handle* h, int with_sal) {
  if (!h) return -1;
  return h->fc.rowLen(with_sal);
}

extern "C" int glia_b200_counts(const glia_handle* h, int64_t* R, int64_t* E, int64_t* C) {
  if (!h) return GLIA_E_ARG;
  if (!h->built) return GLIA_E_STATE;
  if (R) *R = h->R;
  if (E) *E = h->E;
  if (C) *C = h->C;
  return GLIA_OK;
}

extern "C" double glia_b200_last_ms(const glia_handle* h, int which) {
  if (!h || which < 0 || which > 4) return -1.0;
  return h->ms[which];
}
extern "C" int64_t glia_b200_launch_count(const glia_handle* h) { return h ? h->launches : 0; }
