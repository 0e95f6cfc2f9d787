// C ABI, part 2: glia_b200_build and the candidate entry points.
#include <math.h>
#include <string.h>
#include <algorithm>
#include <vector>
#include <stdlib.h>
#include <cuda.h>
#include "handle.cuh"
#include "voxel_fast.cuh"

namespace glia {
cudaError_t launch_voxel_pass_fast(const FastParams& P, cudaStream_t st);
cudaError_t launch_voxel_pass_generic(const VoxParams& P, const VoxChannels& C, const CUtensorMap* tm, cudaStream_t st);
cudaError_t launch_label_max(const uint32_t* labels, size_t n, uint32_t* out, cudaStream_t st);
}  // namespace glia
int glia_finalize_graph(glia_handle* h);
int glia_median_build(glia_handle* h);
int glia_featurize_init_device(glia_handle* h, uint32_t* dKeys, double* dRows);
int glia_rf_label_index(const glia_rf* m, int label);
int glia_rf_score_device(const glia_rf* m, const double* dX, size_t N, int D, int labelIdx,
                         int* dVotes, double* dProb, cudaStream_t st);

using namespace glia;
#define CK(expr) GLIA_CUDA_CHECK(h, expr)
static uint64_t pow2ceil(uint64_t v) { uint64_t p = 1; while (p < v) p <<= 1; return p; }


// ---- value classes of the fast voxel kernel (voxel_fast.cuh) ---------------------------------
// Reference comparisons, already turned into float compares by fill_channel (api.cu):
//   util/image_stats.hxx:17-36  lo < v  <=> v >  lo_dn      v < hi <=> v < hi_up     v < bounds[i] <=> v < bound_up[i]
//   type/feat.hxx:498           v >= t  <=> v >= thr_up
// Each is a step function of v with one breakpoint c in the form "v >= c" (v > c <=> v >= nextup(c)); the
// sorted union of the breakpoints partitions the floats into classes on which every statistic is constant.
static int host_bin_of(const Channel& c, float v) {  // same as bin_of (voxel_pass.cu)
  if (v > c.lo_dn && v < c.hi_up) {
    for (int i = 0; i < c.bins; ++i) if (v < c.bound_up[i]) return i;
    return -1;
  }
  return (v <= c.lo_dn) ? 0 : c.bins - 1;
}
static bool build_value_classes(const Channel& rc, const Channel& ec, int nThr, const float* thr_up, ValueClasses& V) {
  memset(&V, 0, sizeof(V));
  std::vector<float> bp;
  const Channel* chs[2] = {&rc, &ec};
  for (const Channel* c : chs) {
    bp.push_back(nextafterf(c->lo_dn, INFINITY));
    bp.push_back(c->hi_up);
    for (int i = 0; i < c->bins; ++i) bp.push_back(c->bound_up[i]);
  }
  for (int t = 0; t < nThr; ++t) bp.push_back(thr_up[t]);
  std::vector<float> u;
  for (float b : bp) if (b == b && b > 0.0f && b <= 1.0f) u.push_back(b);  // the kernel only sees values in [0, 1]
  std::sort(u.begin(), u.end());
  u.erase(std::unique(u.begin(), u.end()), u.end());
  if ((int)u.size() + 1 > GLIA_FAST_MAX_CLS || nThr > 3) return false;
  V.nCls = (int)u.size() + 1;
  for (size_t i = 0; i < u.size(); ++i) V.bound[i] = u[i];
  for (int j = 0; j < V.nCls; ++j) {
    const float rep = j == 0 ? 0.0f : u[j - 1];  // smallest member of the class
    V.rBin[j] = (signed char)host_bin_of(rc, rep);
    V.eBin[j] = (signed char)host_bin_of(ec, rep);
    unsigned m = 0;
    for (int t = 0; t < nThr; ++t) if (rep >= thr_up[t]) m |= 1u << t;
    V.thr[j] = (unsigned char)m;
  }
  // lookup table: the kernel reads entry max(0, cell(v) - 1), which must never exceed the true class:
  // entry q counts the breakpoints <= lower edge of cell q (entry 0: none, the value may lie below lutLo).
  V.lutLo = u.empty() ? 0.0f : u.front();
  const float span = u.empty() ? 0.0f : u.back() - u.front();
  V.lutScale = span > 0.0f ? 256.0f / span : 0.0f;
  for (int q = 0; q < 256; ++q) {
    int n = 0;
    if (q > 0 && span > 0.0f) {
      const double edge = (double)V.lutLo + (double)q * ((double)span / 256.0);
      for (float b : u) if ((double)b <= edge) ++n;
    }
    V.lut[q] = (unsigned char)n;
  }
  // the kernel starts from lut[max(0, cell - 1)] and takes at most two steps up: a value of cell c (or, by float
  // rounding of the cell index, c + 1) must not lie more than two breakpoints above that entry
  for (int c = 0; c < 256 && span > 0.0f; ++c) {
    const double hi = (double)V.lutLo + (double)(c + 2) * ((double)span / 256.0);
    int n = 0;
    for (float b : u) if ((double)b <= hi) ++n;
    if (n - (int)V.lut[c > 0 ? c - 1 : 0] > 2) return false;
  }
  return true;
}


// ---- TMA tensor maps of the generic voxel kernel (voxel_pass.cu): the label / mask volume as a rank-3 u32 tensor,
// box = tile + halo.  The encoder is a driver entry point fetched at run time (no libcuda link dependency).
// Returns false when the volume does not meet the copy engine's rules (16-byte aligned base and strides, box <= 256).
static bool make_tile_map(CUtensorMap* out, const void* base, const VoxParams& P) {
  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static EncodeFn encode = nullptr;
  static bool looked = false;
  if (!looked) {
    looked = true;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      encode = (EncodeFn)fn;
  }
  const int hy = P.ndim == 2 ? 0 : 1;
  if (!encode || !base || ((uintptr_t)base & 15) || (P.X & 3) || P.SXp > 256 || P.TZ + 2 > 256) return false;
  cuuint64_t dims[3] = {(cuuint64_t)P.X, (cuuint64_t)P.Y, (cuuint64_t)P.zBufN};
  cuuint64_t strides[2] = {4ull * P.X, 4ull * P.X * P.Y};
  cuuint32_t box[3] = {(cuuint32_t)P.SXp, (cuuint32_t)(P.TY + 2 * hy), (cuuint32_t)(P.TZ + 2)};
  cuuint32_t estr[3] = {1, 1, 1};
  return encode(out, CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, const_cast<void*>(base), dims, strides, box, estr,
                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// inputs -> device, table initialisation, voxel pass (raw tables).  Leaves the stream running.
int glia_build_raw(glia_handle* h, const uint32_t* labels, const uint32_t* mask, const float* pb) {
  if (!h || !labels || !pb) return GLIA_E_ARG;
  const glia_params& prm = h->prm;
  for (int i = 0; i < prm.n_rimg; ++i) if (!prm.rimg[i].data) { h->err = "rimg data null"; return GLIA_E_ARG; }
  for (int i = 0; i < prm.n_rlimg; ++i) if (!prm.rlimg[i].data) { h->err = "rlimg data null"; return GLIA_E_ARG; }
  for (int i = 0; i < prm.n_bimg; ++i) if (!prm.bimg[i].data) { h->err = "bimg data null"; return GLIA_E_ARG; }
  CK(cudaSetDevice(prm.device));
  cudaStream_t st = h->stream;
  h->built = false;
  VoxParams& P = h->vp;
  const size_t n = h->nvoxBuf;
  // ---- inputs on the device (host buffers are staged; identical host pointers once)
  std::vector<std::pair<const void*, const void*>> staged;
  int nextImg = 0;
  auto dev_of = [&](const void* src, size_t bytes, DevBuf& buf, const void** out) -> int {
    if (prm.device_pointers) { *out = src; return GLIA_OK; }
    for (auto& s : staged) if (s.first == src) { *out = s.second; return GLIA_OK; }
    CK(buf.ensure(bytes));
    staged.push_back({src, buf.p});  // copied below, in z-chunks that overlap the voxel pass
    *out = buf.p;
    return GLIA_OK;
  };
  const void* d = nullptr;
  int s;
  if ((s = dev_of(labels, 4 * n, h->inLabels, &d))) return s; P.labels = (const uint32_t*)d;
  P.mask = nullptr;
  if (mask) { if ((s = dev_of(mask, 4 * n, h->inMask, &d))) return s; P.mask = (const uint32_t*)d; }
  if ((s = dev_of(pb, 4 * n, h->inPb, &d))) return s; P.pb = (const float*)d;
  int ch = 0;
  for (int i = 0; i < prm.n_rimg; ++i, ++ch) { if ((s = dev_of(prm.rimg[i].data, 4 * n, h->inImg[nextImg++], &d))) return s; h->vc.reg[ch].data = (const float*)d; }
  for (int i = 0; i < prm.n_rlimg; ++i, ++ch) { if ((s = dev_of(prm.rlimg[i].data, 4 * n, h->inImg[nextImg++], &d))) return s; h->vc.reg[ch].data = (const float*)d; }
  for (int i = 0; i < prm.n_bimg; ++i) { if ((s = dev_of(prm.bimg[i].data, 4 * n, h->inImg[nextImg++], &d))) return s; h->vc.edge[i].data = (const float*)d; }

  // Host inputs: the upload is cut into z-chunks on a copy stream; the voxel pass of chunk k starts
  // when chunks <= k+1 have arrived (it reads one plane beyond its own), so PCIe and SMs overlap.
  const size_t plane = (size_t)P.X * P.Y;
  int nChunks = 1;
  if (!staged.empty() && prm.label_bound != 0 && 4 * n >= (16u << 20) && P.zEnd - P.zBegin >= 64) nChunks = 4;
  int zCut[5];
  for (int k = 0; k <= nChunks; ++k) zCut[k] = P.zBegin + (int)(((long long)(P.zEnd - P.zBegin) * k) / nChunks);
  if (!staged.empty()) {
    if (nChunks == 1) {
      for (auto& c : staged) CK(cudaMemcpyAsync((void*)c.second, c.first, 4 * n, cudaMemcpyHostToDevice, st));
    } else {
      if (!h->copyStream) {
        CK(cudaStreamCreateWithFlags(&h->copyStream, cudaStreamNonBlocking));
        for (int k = 0; k < 4; ++k) CK(cudaEventCreateWithFlags(&h->evCopy[k], cudaEventDisableTiming));
      }
      for (int k = 0; k < nChunks; ++k) {
        const size_t p0 = k == 0 ? 0 : (size_t)(zCut[k] - P.zBuf0);
        const size_t p1 = k == nChunks - 1 ? (size_t)P.zBufN : (size_t)(zCut[k + 1] - P.zBuf0);
        for (auto& c : staged)
          CK(cudaMemcpyAsync((char*)c.second + 4 * plane * p0, (const char*)c.first + 4 * plane * p0,
                             4 * plane * (p1 - p0), cudaMemcpyHostToDevice, h->copyStream));
        CK(cudaEventRecord(h->evCopy[k], h->copyStream));
      }
    }
  }
  CK(cudaEventRecord(h->ev[0], st));
  CK(h->scalars.ensure(64));
  // ---- label bound
  uint32_t LB = prm.label_bound;
  if (LB == 0) {
    CK(cudaMemsetAsync(h->scalars.p, 0, 64, st));
    CK(launch_label_max(P.labels, n, h->scalars.as<uint32_t>(), st));
    h->launches++;
    uint32_t mxl = 0;
    CK(cudaMemcpyAsync(&mxl, h->scalars.p, 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (mxl == 0xFFFFFFFFu) { h->err = "label 0xFFFFFFFF is reserved"; return GLIA_E_ARG; }
    LB = mxl + 1;
  }
  P.labelBound = LB;
  // ---- raw tables (zero == identity for every word kind)
  size_t owned = (size_t)P.X * P.Y * (size_t)(P.zEnd - P.zBegin);
  uint64_t cap = prm.edge_capacity;
  if (cap == 0) {
    // directed boundaries per region: ~5.7 in 2D, ~13.8 in 3D (synthetic EM); the table is
    // grown x4 and the pass re-run if this guess is too small
    uint64_t est = (P.ndim == 2 ? 8ull : 16ull) * LB;
    // a z-slab sees (and, after the exchange, owns) its share of the boundaries plus those near its faces
    if (P.zEnd - P.zBegin < P.Z) est = 2 * est * (uint64_t)(P.zEnd - P.zBegin) / (uint64_t)P.Z + 4096;
    est = std::min<uint64_t>(owned, est);
    cap = pow2ceil(est + est / 2 + 1024);
    if (cap < h->capHint) cap = h->capHint;
  } else cap = pow2ceil(cap);
  const uint64_t capMax = pow2ceil(2 * (uint64_t)owned + 1024);
  if (cap > capMax) cap = capMax;
  for (;;) {  // grows the directed-boundary table and re-runs the pass if it overflowed
  P.edgeCap = cap;
  CK(h->regAdd.ensure(4 * (size_t)LB * P.reg.nAdd));
  CK(h->regMax.ensure(4 * (size_t)LB * P.reg.nMax));
  CK(h->regF64.ensure(8 * (size_t)LB * std::max(1, P.reg.nF64)));
  CK(h->edgeKeys.ensure(8 * (size_t)cap));
  CK(h->edgeAdd.ensure(4 * (size_t)cap * P.edge.nAdd));
  CK(h->edgeMax.ensure(4 * (size_t)cap * std::max(1, P.edge.nMax)));
  CK(h->edgeF64.ensure(8 * (size_t)cap * std::max(1, P.edge.nF64)));
  CK(cudaMemsetAsync(h->regAdd.p, 0, 4 * (size_t)LB * P.reg.nAdd, st));
  CK(cudaMemsetAsync(h->regMax.p, 0, 4 * (size_t)LB * P.reg.nMax, st));
  CK(cudaMemsetAsync(h->regF64.p, 0, 8 * (size_t)LB * P.reg.nF64, st));
  CK(cudaMemsetAsync(h->edgeKeys.p, 0xFF, 8 * (size_t)cap, st));
  CK(cudaMemsetAsync(h->edgeAdd.p, 0, 4 * (size_t)cap * P.edge.nAdd, st));
  CK(cudaMemsetAsync(h->edgeMax.p, 0, 4 * (size_t)cap * P.edge.nMax, st));
  CK(cudaMemsetAsync(h->edgeF64.p, 0, 8 * (size_t)cap * P.edge.nF64, st));
  CK(cudaMemsetAsync(h->scalars.p, 0, 64, st));
  h->launches += 8;
  P.regAdd = h->regAdd.as<uint32_t>(); P.regMax = h->regMax.as<uint32_t>(); P.regF64 = h->regF64.as<double>();
  P.edgeKeys = (unsigned long long*)h->edgeKeys.p;
  P.edgeAdd = h->edgeAdd.as<uint32_t>(); P.edgeMax = h->edgeMax.as<uint32_t>(); P.edgeF64 = h->edgeF64.as<double>();
  P.overflow = (int*)h->scalars.p + 12;
  // ---- voxel pass
  CK(cudaEventRecord(h->ev[6], st));
  {
    ValueClasses vcls;
    const bool fast = !h->forceGeneric && !h->fastDomainFailed && !P.mask && prm.n_rimg == 1 && prm.n_rlimg == 0 &&
                      prm.n_bimg == 1 && prm.n_thr <= 3 && h->vc.reg[0].data == P.pb &&
                      h->vc.edge[0].data == P.pb && h->vc.reg[0].bins <= 16 && h->vc.edge[0].bins <= 16 &&
                      P.Y <= 8 * 65535 &&  // grid y extent
                      build_value_classes(h->vc.reg[0], h->vc.edge[0], P.nThr, P.thr_up, vcls);
    h->usedFast = fast;
    if (fast) {
      FastParams F;
      memset(&F, 0, sizeof(F));
      F.labels = P.labels; F.pb = P.pb; F.ndim = P.ndim; F.X = P.X; F.Y = P.Y; F.Z = P.Z;
      F.zBegin = P.zBegin; F.zEnd = P.zEnd; F.zBuf0 = P.zBuf0;
      // CTA = 8 warps: 256 x 1 voxels per row step in 2D, 32 x 8 in 3D, marching TZ rows / planes.
      // 512 table slots keep a CTA at ~51 KB (3D) / ~66 KB (2D) of shared memory: 4 / 3 CTAs per SM.
      if (P.ndim == 2) { F.warpsX = 8; F.warpsY = 1; } else { F.warpsX = 1; F.warpsY = 8; }
      F.TZ = h->tuneTZ > 0 ? h->tuneTZ : (P.ndim == 3 ? 24 : 16);  // measured: profiles/r2_voxel_notes.txt
      if (h->tuneTZ <= 0 && P.ndim == 3) {  // even out the tiles of a short slab: 128 planes -> 6 tiles of 22, not 5 x 24 + 8
        const int planes = P.zEnd - P.zBegin, nt = (planes + F.TZ - 1) / F.TZ;
        if (nt > 0) F.TZ = (planes + nt - 1) / nt;
      }
      if (F.TZ + F.warpsY > 32) F.TZ = 32 - F.warpsY;  // y and z occupancy masks share one word
      while (F.TZ > 1 && (long long)P.X * P.Y * (F.TZ + 3) >= (1ll << 31)) F.TZ >>= 1;
      F.tilesX = (P.X + 32 * F.warpsX - 1) / (32 * F.warpsX);
      F.tilesY = (P.Y + F.warpsY - 1) / F.warpsY;
      F.tilesZ = (P.zEnd - P.zBegin + F.TZ - 1) / F.TZ;
      F.nThr = P.nThr;
      F.rBins = h->vc.reg[0].bins; F.eBins = h->vc.edge[0].bins;
      F.vc = vcls;
      F.reg = P.reg; F.edge = P.edge; F.labelBound = P.labelBound;
      F.regAdd = P.regAdd; F.regMax = P.regMax; F.regF64 = P.regF64;
      F.edgeCap = P.edgeCap; F.edgeKeys = P.edgeKeys; F.edgeAdd = P.edgeAdd; F.edgeMax = P.edgeMax; F.edgeF64 = P.edgeF64;
      F.overflow = P.overflow;
      F.slots = h->tuneSlots > 0 ? (int)pow2ceil((uint64_t)h->tuneSlots) : 512;
      for (int k = 0; k < nChunks; ++k) {
        if (nChunks > 1) CK(cudaStreamWaitEvent(st, h->evCopy[k + 1 < nChunks ? k + 1 : k], 0));
        F.zBegin = zCut[k]; F.zEnd = zCut[k + 1];
        F.tilesZ = (F.zEnd - F.zBegin + F.TZ - 1) / F.TZ;
        CK(launch_voxel_pass_fast(F, st));
      }
    } else {
      // label (and mask) tiles are staged by TMA when the volume meets the copy engine's alignment rules
      CUtensorMap tm[2];
      memset(tm, 0, sizeof(tm));
      const bool tma = !h->noTma && make_tile_map(&tm[0], P.labels, P) && (!P.mask || make_tile_map(&tm[1], P.mask, P));
      h->usedTma = tma;
      for (int k = 0; k < nChunks; ++k) {
        if (nChunks > 1) CK(cudaStreamWaitEvent(st, h->evCopy[k + 1 < nChunks ? k + 1 : k], 0));
        VoxParams Pk = P;
        Pk.zBegin = zCut[k]; Pk.zEnd = zCut[k + 1];
        Pk.tilesZ = (Pk.zEnd - Pk.zBegin + Pk.TZ - 1) / Pk.TZ;
        CK(launch_voxel_pass_generic(Pk, h->vc, tma ? tm : nullptr, st));
      }
    }
  }
  h->launches += nChunks;
  CK(cudaEventRecord(h->ev[1], st));
  int ovf = 0;
  CK(cudaMemcpyAsync(&ovf, P.overflow, 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  if (ovf & 2) { h->err = "label value >= label_bound (0xFFFFFFFF is reserved)"; return GLIA_E_ARG; }
  if (ovf & 4) {  // a value outside [0, 1] (or NaN): the fast kernel's fixed-point sums do not cover it
    h->fastDomainFailed = true;  // sticky: later volumes of this handle go to the generic kernel directly
    nChunks = 1; zCut[0] = P.zBegin; zCut[1] = P.zEnd;  // the inputs are resident by now
    continue;
  }
  if (ovf & 1) {
    if (prm.edge_capacity != 0 || cap >= capMax) { h->err = "directed-boundary table full: raise edge_capacity"; return GLIA_E_CAPACITY; }
    cap *= 4;
    if (cap > capMax) cap = capMax;
    h->capHint = cap;
    continue;
  }
  break;
  }
  h->partial = true;
  h->hasOrient = false;  // orientation bits belong to one build
  return GLIA_OK;
}

extern "C" int glia_b200_build(glia_handle* h, const uint32_t* labels, const uint32_t* mask,
                               const float* pb) {
  int s = glia_build_raw(h, labels, mask, pb);
  if (s != GLIA_OK) return s;
  cudaStream_t st = h->stream;
  // ---- graph (a handle created for a z-slab sees only its own planes: multi-GPU callers use
  // glia_b200_build_partial + the exchange of shard.cu instead)
  s = glia_finalize_graph(h);
  if (s != GLIA_OK) return s;
  if (h->prm.use_median_feat && (s = glia_median_build(h)) != GLIA_OK) return s;
  CK(cudaEventRecord(h->ev[2], st));
  CK(cudaStreamSynchronize(st));
  float ms = 0;
  cudaEventElapsedTime(&ms, h->ev[6], h->ev[1]); h->ms[0] = ms;
  cudaEventElapsedTime(&ms, h->ev[1], h->ev[2]); h->ms[1] = ms;
  cudaEventElapsedTime(&ms, h->ev[0], h->ev[6]); h->ms[4] = ms;
  h->built = true;
  h->err.clear();
  return GLIA_OK;
}

// ---- dumps ---------------------------------------------------------------------------------
template <class T>
static int d2h(glia_handle* h, std::vector<T>& v, const void* src, size_t count) {
  v.resize(count);
  if (!count) return GLIA_OK;
  CK(cudaMemcpyAsync(v.data(), src, sizeof(T) * count, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  return GLIA_OK;
}

extern "C" int glia_b200_get_regions(glia_handle* h, uint32_t* label, uint32_t* count,
                                     uint32_t* border, int32_t* bbmin, int32_t* bbmax) {
  if (!h) return GLIA_E_ARG;
  if (!h->built) return GLIA_E_STATE;
  const RecLayout L = h->vp.reg;
  std::vector<uint32_t> lab, add, mx;
  int s;
  if ((s = glia_ensure_full_regions(h))) return s;
  if ((s = d2h(h, lab, h->regLabel.p, (size_t)h->R))) return s;
  if ((s = d2h(h, add, h->cAdd.p, (size_t)h->R * L.nAdd))) return s;
  if ((s = d2h(h, mx, h->cMax.p, (size_t)h->R * L.nMax))) return s;
  for (long long r = 0; r < h->R; ++r) {
    if (label) label[r] = lab[r];
    if (count) count[r] = add[r * L.nAdd];
    if (border) border[r] = add[r * L.nAdd + 1];
    for (int d = 0; d < 3; ++d) {
      int ax = h->fc.axis[d];
      bool used = d < h->prm.ndim;
      uint32_t wmin = mx[r * L.nMax + ax], wmax = mx[r * L.nMax + 3 + ax];
      if (bbmin) bbmin[3 * r + d] = (used && wmin) ? (int32_t)(~wmin) : INT32_MAX;
      if (bbmax) bbmax[3 * r + d] = (used && wmax) ? (int32_t)(wmax - 1) : INT32_MIN;
    }
  }
  return GLIA_OK;
}

static int dump_stats(glia_handle* h, long long n, const RecLayout& L, int bins, int addOff,
                      int maxOff, int f64Off, const void* dAdd, const void* dMax, const void* dF64,
                      uint32_t* hist, double* sum, double* sum2, float* mn, float* mx) {
  std::vector<uint32_t> add, m;
  std::vector<double> f;
  int s;
  if ((s = d2h(h, add, dAdd, (size_t)n * L.nAdd))) return s;
  if ((s = d2h(h, m, dMax, (size_t)n * L.nMax))) return s;
  if ((s = d2h(h, f, dF64, (size_t)n * L.nF64))) return s;
  for (long long i = 0; i < n; ++i) {
    if (hist) for (int b = 0; b < bins; ++b) hist[i * bins + b] = add[i * L.nAdd + addOff + b];
    if (sum) sum[i] = f[i * L.nF64 + f64Off];
    if (sum2) sum2[i] = f[i * L.nF64 + f64Off + 1];
    uint32_t wmin = m[i * L.nMax + maxOff], wmax = m[i * L.nMax + maxOff + 1];
    if (mn) mn[i] = wmin ? ord2f(~wmin) : 0.0f;
    if (mx) mx[i] = wmax ? ord2f(wmax) : 0.0f;
  }
  return GLIA_OK;
}

extern "C" int glia_b200_get_region_stats(glia_handle* h, int kind, int i, uint32_t* hist,
                                          double* sum, double* sum2, float* mn, float* mx) {
  if (!h) return GLIA_E_ARG;
  if (!h->built) return GLIA_E_STATE;
  int nk = kind == 0 ? h->prm.n_rimg : h->prm.n_rlimg;
  if (kind < 0 || kind > 1 || i < 0 || i >= nk) return GLIA_E_ARG;
  const Channel& c = h->vc.reg[kind == 0 ? i : h->prm.n_rimg + i];
  { int s = glia_ensure_full_regions(h); if (s != GLIA_OK) return s; }
  return dump_stats(h, h->R, h->vp.reg, c.bins, c.addOff, c.maxOff, c.f64Off, h->cAdd.p, h->cMax.p,
                    h->cF64.p, hist, sum, sum2, mn, mx);
}

extern "C" int glia_b200_get_edges(glia_handle* h, uint32_t* a, uint32_t* b, uint32_t* count,
                                   uint8_t* mutual, uint32_t* thr_count) {
  if (!h) return GLIA_E_ARG;
  if (!h->built) return GLIA_E_STATE;
  const RecLayout L = h->vp.edge;
  const int T = h->prm.n_thr;
  std::vector<uint32_t> lab, add;
  std::vector<int> ea, eb, rev;
  int s;
  if ((s = d2h(h, lab, h->regLabel.p, (size_t)h->R))) return s;
  if ((s = d2h(h, ea, h->ea.p, (size_t)h->E))) return s;
  if ((s = d2h(h, eb, h->eb.p, (size_t)h->E))) return s;
  if ((s = d2h(h, rev, h->rev.p, (size_t)h->E))) return s;
  if ((s = d2h(h, add, h->sAdd.p, (size_t)h->E * L.nAdd))) return s;
  for (long long e = 0; e < h->E; ++e) {
    if (a) a[e] = lab[ea[e]];
    if (b) b[e] = lab[eb[e]];
    if (count) count[e] = add[e * L.nAdd];
    if (mutual) mutual[e] = rev[e] >= 0;
    if (thr_count) for (int t = 0; t < T; ++t) thr_count[e * T + t] = add[e * L.nAdd + 1 + t];
  }
  return GLIA_OK;
}

extern "C" int glia_b200_get_edge_stats(glia_handle* h, int i, uint32_t* hist, double* sum,
                                        double* sum2, float* mn, float* mx) {
  if (!h) return GLIA_E_ARG;
  if (!h->built) return GLIA_E_STATE;
  if (i < 0 || i >= h->prm.n_bimg) return GLIA_E_ARG;
  const Channel& c = h->vc.edge[i];
  return dump_stats(h, h->E, h->vp.edge, c.bins, c.addOff, c.maxOff, c.f64Off, h->sAdd.p, h->sMax.p,
                    h->sF64.p, hist, sum, sum2, mn, mx);
}

// ---- candidates ------------------------------------------------------------------------------
extern "C" int glia_b200_init_candidates(glia_handle* h, uint32_t* keys, double* rows) {
  if (!h) return GLIA_E_ARG;
  if (!h->built) return GLIA_E_STATE;
  CK(cudaSetDevice(h->prm.device));
  const size_t C = (size_t)h->C, dim = (size_t)h->fc.dim;
  if (!C) return GLIA_OK;
  uint32_t* dKeys = nullptr; double* dRows = nullptr;
  if (h->prm.device_pointers) { dKeys = keys; dRows = rows; }
  else {
    if (keys) { CK(h->keysOut.ensure(8 * C)); dKeys = h->keysOut.as<uint32_t>(); }
    if (rows) { CK(h->rows.ensure(8 * C * dim)); dRows = h->rows.as<double>(); }
  }
  CK(cudaEventRecord(h->ev[3], h->stream));
  int s = glia_featurize_init_device(h, dKeys, dRows);
  if (s) return s;
  CK(cudaEventRecord(h->ev[4], h->stream));
  if (!h->prm.device_pointers) {
    if (keys) CK(cudaMemcpyAsync(keys, dKeys, 8 * C, cudaMemcpyDeviceToHost, h->stream));
    if (rows) { int s2 = glia_copy_to_host(h, rows, dRows, 8 * C * dim); if (s2 != GLIA_OK) return s2; }
  }
  CK(cudaStreamSynchronize(h->stream));
  float ms = 0; cudaEventElapsedTime(&ms, h->ev[3], h->ev[4]); h->ms[2] = ms;
  return GLIA_OK;
}

extern "C" int glia_b200_score_candidates(glia_handle* h, const glia_rf* rf, double* sal) {
  if (!h || !rf || !sal) return GLIA_E_ARG;
  if (!h->built) return GLIA_E_STATE;
  if (rf->device != h->prm.device) { h->err = "forest lives on another device"; return GLIA_E_ARG; }
  if (rf->mdim > h->fc.dim) { h->err = "forest uses more features than the row has"; return GLIA_E_ARG; }
  int li = glia_rf_label_index(rf, -1);  // BC_LABEL_MERGE, hmt/bc_label.hxx:11
  if (li < 0) { h->err = "forest has no class -1"; return GLIA_E_ARG; }
  CK(cudaSetDevice(h->prm.device));
  const size_t C = (size_t)h->C, dim = (size_t)h->fc.dim;
  if (!C) return GLIA_OK;
  double* dSal = sal;
  if (!h->prm.device_pointers) { CK(h->sal.ensure(8 * C)); dSal = h->sal.as<double>(); }
  CK(cudaEventRecord(h->ev[3], h->stream));
  int s;
  CK(h->rows.ensure(8 * C * dim));
  s = glia_featurize_init_device(h, nullptr, h->rows.as<double>());
  if (s) return s;
  CK(cudaEventRecord(h->ev[4], h->stream));
  s = glia_rf_score_device(rf, h->rows.as<double>(), C, (int)dim, li, nullptr, dSal, h->stream);
  h->launches++;
  if (s) { h->err = "rf kernel launch failed"; return s; }
  CK(cudaEventRecord(h->ev[5], h->stream));
  if (!h->prm.device_pointers) CK(cudaMemcpyAsync(sal, dSal, 8 * C, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  float ms = 0;
  cudaEventElapsedTime(&ms, h->ev[3], h->ev[4]); h->ms[2] = ms;
  cudaEventElapsedTime(&ms, h->ev[4], h->ev[5]); h->ms[3] = ms;
  return GLIA_OK;
}
