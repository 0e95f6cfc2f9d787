// glia_b200 — the handle behind the C ABI (include/glia_b200.h).
#pragma once
#include <chrono>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#include "common.cuh"

namespace glia {

// Grow-only device buffer (reused across calls on the same handle).
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <class T> T* as() const { return (T*)p; }
};

// Featurisation constants (device copy of the feature configuration).
struct FeatCfg {
  int D;             // DIMENSION
  int axis[3];       // kernel axis of reference dimension d (2D volumes run as X x 1 x Y)
  int nThr, nR, nRL, nB;
  int rBins[GLIA_MAX_IMG], rlBins[GLIA_MAX_IMG], bBins[GLIA_MAX_IMG];
  int rAdd[GLIA_MAX_IMG], rMax[GLIA_MAX_IMG], rF64[GLIA_MAX_IMG];     // offsets in region records
  int rlAdd[GLIA_MAX_IMG];
  int bAdd[GLIA_MAX_IMG], bMax[GLIA_MAX_IMG], bF64[GLIA_MAX_IMG];     // offsets in edge records
  RecLayout reg, edge;
  double normArea, normLen;
  int useLog, simple;
  int hist, median;  // GLIA_USE_HISTOGRAM_AS_FEATS / GLIA_USE_MEDIAN_AS_FEATS of the reference build, as run-time fields
  int dim;           // row length without saliency terms
  const double* log2tab;  // device table: log2 of the counts 0 .. 4095 (featurize.cuh)
  // serialised lengths (type/feat.hxx:608-626, :677-700, :826-840): ImageLabelFeats = [bins] + entropy,
  // ImageRealFeats = [median] + mean, stddev, min, max
  __host__ __device__ int lenReal() const { return median ? 5 : 4; }
  __host__ __device__ int lenLabel(int bins) const { return (hist ? bins : 0) + 1; }
  __host__ __device__ int lenR(int i) const { return lenLabel(rBins[i]) + lenReal(); }
  __host__ __device__ int lenRL(int i) const { return lenLabel(rlBins[i]); }
  __host__ __device__ int lenB(int i) const { return lenLabel(bBins[i]) + lenReal(); }
  __host__ __device__ int xiLen() const {  // one region block (x1 / x2 / x3) without its saliency
    int n = D + 4 + 2 * nThr;
    for (int i = 0; i < nR; ++i) n += lenR(i);
    for (int i = 0; i < nRL; ++i) n += lenRL(i);
    for (int i = 0; i < nB; ++i) n += lenB(i);
    return n;
  }
  __host__ __device__ int x0Len() const {  // pair terms without saliency
    int n = 11 + 4 * nThr + nR * (3 + lenReal()) + 3 * nRL;
    for (int i = 0; i < nB; ++i) n += lenB(i);
    return n;
  }
  __host__ __device__ int rowLen(int withSal) const {
    if (simple) return 5 + nB * (median ? 2 : 1) + 4 * nR + 2 * nRL;
    return x0Len() + 3 * xiLen() + (withSal ? 5 : 0);
  }
  __host__ __device__ int nMed() const { return median ? 3 * nR + 4 * nB : 0; }  // medians of one row (featurize.cuh)
};

// GLIA_B200_TRACE=1: per-phase wall times on stderr (each mark synchronises the stream; diagnostics only).
struct PhaseTrace {
  bool on; cudaStream_t st; std::chrono::steady_clock::time_point t0;
  PhaseTrace(cudaStream_t s) : on(getenv("GLIA_B200_TRACE") != nullptr), st(s) { if (on) { cudaStreamSynchronize(st); t0 = std::chrono::steady_clock::now(); } }
  void mark(const char* what) {
    if (!on) return;
    cudaStreamSynchronize(st);
    auto t1 = std::chrono::steady_clock::now();
    fprintf(stderr, "[glia trace] %-28s %8.1f us\n", what, std::chrono::duration<double, std::micro>(t1 - t0).count());
    t0 = t1;
  }
};

// Completes per-node tables that every rank of a sharded build holds a partial copy of (merge.cu asks, shard.cu
// answers with ncclAllReduce); in place, on the given stream.
struct TableReducer {
  enum Type { U32, I32, F64 };
  enum Op { SUM, MAX, MIN };
  virtual int all_reduce(void* buf, size_t count, Type t, Op op, cudaStream_t st) = 0;
  virtual ~TableReducer() {}
};

}  // namespace glia

// alg::EnsembleRandomForest + opt::ThresholdModelDistributor (alg/rf.hxx:63-98, type/function.hxx:71-85):
// a row picks model 0 if x[dim1] < thr, else 1 if x[dim0] < thr, else 2; the trees of all models
// are concatenated on the device and a candidate walks the tree range [tb[m], te[m]) of its model.
struct RfDist {
  int n;            // 1 = plain forest
  int dim0, dim1;
  double thr;
  int tb[3], te[3];
  __host__ __device__ int begin(int m) const { return m == 0 ? tb[0] : (m == 1 ? tb[1] : tb[2]); }
  __host__ __device__ int end(int m) const { return m == 0 ? te[0] : (m == 1 ? te[1] : te[2]); }
  __host__ __device__ int pick(const double* x) const {
    if (n <= 1) return 0;
    if (x[dim1] < thr) return 0;
    if (x[dim0] < thr) return 1;
    return 2;
  }
};

struct glia_rf {
  int device = 0;
  int ntree = 0, nrnodes = 0, nclass = 0, mdim = 0;
  std::vector<int> treemap, nodestatus, bestvar, nodeclass, ndbigtree, orig_labels;
  std::vector<double> xbestsplit;
  // device: compact breadth-first forest (rf.cu)
  int totalNodes = 0;
  int* d_treeOff = nullptr;      // [ntree+1]
  void* d_nodes = nullptr;       // packed nodes
  void* d_nodesW = nullptr;      // walk encoding for rf_score_smem_kernel: leaves are self-loops (rf.cu)
  void* d_blocks = nullptr;      // two-level 16-byte blocks (rf_walk.cuh, two-class forests)
  double* d_blockThr = nullptr;  // exact thresholds [3 * block + j] for the rare undecided compare
  int* d_blockRoot = nullptr;    // [ntree] root block of every tree
  long long nBlocks = 0;
  int maxDepth = 0;
  RfDist dist = {1, 0, 0, 0.0, {0, 0, 0}, {0, 0, 0}};  // dist.te[0] = ntree for a plain forest
};

struct glia_handle {
  glia_params prm;
  std::string err;
  cudaStream_t stream = nullptr;
  bool ownStream = false;
  cudaEvent_t ev[8];
  cudaStream_t copyStream = nullptr;  // chunked upload of host inputs (api_build.cu)
  cudaEvent_t evCopy[4];
  double ms[5] = {-1, -1, -1, -1, -1};
  long long launches = 0;
  bool built = false;
  bool partial = false;  // raw tables valid (glia_b200_build_partial / glia_b200_build)
  bool usedFast = false;
  bool noTma = false, usedTma = false;  // generic voxel kernel: tile staged by plain loads instead of TMA ("notma")
  bool forceGeneric = false;      // glia_b200_tune(h, "generic", 1): test hook, run the generic voxel kernel
  bool fastDomainFailed = false;  // a value outside [0, 1] was met: the fast voxel kernel is skipped from then on
  int tuneTZ = 0, tuneSlots = 0;  // tile depth / table slots of the fast voxel kernel (0 = default)
  long long tableOccupied = 0;  // occupied slots of the boundary table at the last export (shard.cu)
  uint64_t capHint = 0;  // boundary-table capacity that turned out to be needed (grown tables are kept)

  // kernel-axis geometry (2D runs as X x 1 x rows)
  int KX = 0, KY = 0, KZ = 0;
  size_t nvoxBuf = 0;

  glia::VoxParams vp;
  glia::VoxChannels vc;
  glia::FeatCfg fc;

  // staging of host inputs
  glia::DevBuf inLabels, inMask, inPb, inImg[3 * GLIA_MAX_IMG];
  // raw tables of the voxel pass
  glia::DevBuf regAdd, regMax, regF64, edgeKeys, edgeAdd, edgeMax, edgeF64, flags;
  // finalised graph
  long long R = 0, E = 0, C = 0;
  long long ownLo = 0, ownHi = 0;                          // region indices owned by this handle (all of them unless sharded)
  glia::DevBuf rank, regLabel, cAdd, cMax, cF64;          // regions ascending by label
  glia::DevBuf eSlot, eKey, eSlotSorted, eKeySorted;      // compaction + sort
  glia::DevBuf sAdd, sMax, sF64, ea, eb, row, rev, cand;  // directed boundaries, CSR by a
  glia::DevBuf bAggAdd, bAggMax, bAggF64;                 // per-region sum over its boundaries
  glia::DevBuf bAggMax2, bAggOwn;                         // second-best max words and the boundary holding the best
  glia::DevBuf cubTmp, scalars, rows, sal, keysOut;
  glia::DevBuf regBlocks;                                 // per-region x1/x2 row blocks (featurize.cu)
  glia::DevBuf log2tab;
  glia::DevBuf orient, tieBuf, firstVox;                  // equal-area orientation bits (orient.cu)
  bool hasOrient = false;
  long long orientC = 0;
  glia::DevBuf interest, exportBuf, exportCursor;         // multi-GPU exchange (shard.cu)
  // median features (median.cu): sorted value keys per base region / directed boundary, per image
  glia::DevBuf medRKeys[GLIA_MAX_IMG], medEKeys[GLIA_MAX_IMG], medRStart, medEStart, medTmpA, medTmpB;
  glia::DevBuf medReg, medCand, medQ[6];
  glia::DevBuf k5[40];                                    // scratch of the bc_feat stage (merge.cu), kept between calls
  bool hasMedian = false;
  int shardWorld = 0, shardMe = 0;  // set by glia_b200_build_sharded around its graph finalisation (interest mask valid)
  bool regionsPartial = false;  // compact region records hold the regions of interest only (sharded build)
  // large results to pageable host memory: two pinned staging buffers, D2H of chunk k+1 overlaps the host copy of chunk k
  void* pin[2] = {nullptr, nullptr};
  cudaEvent_t evPin[2] = {nullptr, nullptr};
};

int glia_ensure_full_regions(glia_handle* h);  // graph.cu
// device -> (pageable) host copy of a large result on the handle's stream; returns when dst is complete (api.cu)
int glia_copy_to_host(glia_handle* h, void* dst, const void* dsrc, size_t bytes);
