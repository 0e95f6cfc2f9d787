// K5: hmt feature stage for a GIVEN merge order (the "per-level update of merged-region
// features": a segmented reduction over child-region records, tree level by tree level).
//
// Replaces (reference, /root/reference/code):
//   bc_feat operation()                         hmt/main_bc_feat.cxx:27-112
//   TRegionMap::set(order) / merge              type/region_map.hxx:67-68, :113-118
//   TRegion::merge (mutual entries cancel)      type/region.hxx:66-75
//   getBoundary / TRegion::boundaryWith         util/struct.hxx:10-16, type/region.hxx:42-51
//   genSaliencyMap                              hmt/bc_feat.hxx:12-26
// The reference applies all merges and then re-walks the pixel lists of every tree node.  Here
// a node's record is the combination of its two children's records, and the bookkeeping of
// which directed boundaries are still entries of a node follows from the tree alone:
//   a mutual boundary (a->b) stops being an entry exactly at LCA(a,b); a non-mutual one never.
// So  B(Z) = B(X) + B(Y) - C(Z)   with C(Z) = sum of mutual boundaries whose LCA is Z,
// the pair boundary of merge Z is C(Z) plus the non-mutual boundaries between X and Y whose
// target base region still has a live entry, and min/max words (not subtractable) are pushed
// up the tree from each boundary's source leaf to just below its LCA.
#include <algorithm>
#include <unordered_map>
#include <vector>
#include "featurize.cuh"
#include "median.cuh"

namespace glia {

struct TreeView {
  int N, R, K;            // nodes, leaves, lifting levels
  const int* up;          // [K][N], up[0] = parent (root -> itself)
  const int* depth;       // [N]
};

__global__ void lift_kernel(const int* __restrict__ upPrev, const int* __restrict__ cntPrev,
                            int* __restrict__ upNext, int* __restrict__ cntNext, int N) {
  int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= N) return;
  int u = upPrev[v];
  upNext[v] = upPrev[u];
  cntNext[v] = cntPrev[v] + cntPrev[u];
}

__device__ __forceinline__ int lca_of(const TreeView& T, int a, int b) {
  int da = T.depth[a], db = T.depth[b];
  if (da < db) { int t = a; a = b; b = t; t = da; da = db; db = t; }
  int diff = da - db;
  for (int k = 0; diff; ++k, diff >>= 1)
    if (diff & 1) a = T.up[(size_t)k * T.N + a];
  if (a == b) return a;
  for (int k = T.K - 1; k >= 0; --k) {
    int ua = T.up[(size_t)k * T.N + a], ub = T.up[(size_t)k * T.N + b];
    if (ua != ub) { a = ua; b = ub; }
  }
  int pa = T.up[a], pb = T.up[b];
  return (pa == pb && pa != a) ? pa : -1;  // different trees of the forest: never joined
}

// per directed boundary: LCA of its end points; per base region: depth of the highest LCA
// among its mutual boundaries (topd) and whether it owns a non-mutual boundary (hasNM)
// A sharded build holds the boundaries of its owned regions AND of their neighbours; a boundary is handled by the
// rank that owns its source region, [ownLo, ownHi) (every directed boundary exactly once across the ranks).
__global__ void edge_lca_kernel(TreeView T, const int* __restrict__ ea, const int* __restrict__ eb,
                                const int* __restrict__ rev, int E, int* __restrict__ lca,
                                int* __restrict__ topd, int* __restrict__ hasNM, int ownLo, int ownHi) {
  int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  int a = ea[e], b = eb[e];
  if (a < ownLo || a >= ownHi) { lca[e] = -1; return; }
  int l = lca_of(T, a, b);
  lca[e] = l;
  if (rev[e] >= 0) atomicMin(topd + a, l >= 0 ? T.depth[l] : -1);  // never cancelled: always live
  else hasNM[a] = 1;
}

// accumulate boundary records by LCA: mutual -> C[lca]; non-mutual with an active target -> NMA[lca];
// push min/max words up from the source leaf to just below the node where the entry dies
__global__ void edge_scatter_kernel(TreeView T, const int* __restrict__ ea, const int* __restrict__ eb,
                                    const int* __restrict__ rev, const int* __restrict__ lca,
                                    const int* __restrict__ topd, const int* __restrict__ hasNM, int E,
                                    RecLayout L, const uint32_t* __restrict__ sAdd,
                                    const uint32_t* __restrict__ sMax, const double* __restrict__ sF64,
                                    uint32_t* cAdd, uint32_t* cMax, double* cF64, uint32_t* nAdd,
                                    uint32_t* nMax, double* nF64, uint32_t* bMax, int ownLo, int ownHi) {
  long long w = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  int lane = threadIdx.x & 31;
  if (w >= E) return;
  int e = (int)w, a = ea[e], b = eb[e], l = lca[e];
  if (a < ownLo || a >= ownHi) return;
  bool mutual = rev[e] >= 0;
  if (l >= 0) {
    uint32_t* dA; uint32_t* dM; double* dF; bool go = true;
    if (mutual) { dA = cAdd; dM = cMax; dF = cF64; }
    else {
      // util/struct.hxx:10-16 / region.hxx:42-51: counted only if the target base region still
      // has a live entry inside its side of the merge
      go = hasNM[b] || topd[b] <= T.depth[l];
      dA = nAdd; dM = nMax; dF = nF64;
    }
    if (go) {
      for (int i = lane; i < L.nAdd; i += 32) { uint32_t v = sAdd[(size_t)e * L.nAdd + i]; if (v) atomicAdd(dA + (size_t)l * L.nAdd + i, v); }
      for (int i = lane; i < L.nMax; i += 32) { uint32_t v = sMax[(size_t)e * L.nMax + i]; if (v) atomicMax(dM + (size_t)l * L.nMax + i, v); }
      for (int i = lane; i < L.nF64; i += 32) { double v = sF64[(size_t)e * L.nF64 + i]; if (v != 0.0) atomicAdd(dF + (size_t)l * L.nF64 + i, v); }
    }
  }
  // min/max words of the boundary-of records: entry (a->b) is live in every ancestor-or-self
  // of leaf a strictly below the LCA (mutual) or in all of them (non-mutual / never joined)
  if (lane < L.nMax) {
    uint32_t v = sMax[(size_t)e * L.nMax + lane];
    if (v) {
      int stop = (mutual && l >= 0) ? l : -1;
      int node = a;
      for (;;) {
        if (node == stop) break;
        atomicMax(bMax + (size_t)node * L.nMax + lane, v);
        int p = T.up[node];
        if (p == node) break;
        node = p;
      }
    }
  }
}

// one warp per node of the current level: combine the two children
__global__ void level_combine_kernel(const int* __restrict__ nodes, int count,
                                     const int* __restrict__ child0, const int* __restrict__ child1,
                                     RecLayout LR, RecLayout LE, uint32_t* rAdd, uint32_t* rMax,
                                     double* rF64, uint32_t* bAdd, double* bF64,
                                     const uint32_t* __restrict__ cAdd, const double* __restrict__ cF64) {
  int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (w >= count) return;
  size_t z = nodes[w], x = child0[z], y = child1[z];
  for (int i = lane; i < LR.nAdd; i += 32) rAdd[z * LR.nAdd + i] = rAdd[x * LR.nAdd + i] + rAdd[y * LR.nAdd + i];
  for (int i = lane; i < LR.nMax; i += 32) rMax[z * LR.nMax + i] = max(rMax[x * LR.nMax + i], rMax[y * LR.nMax + i]);
  for (int i = lane; i < LR.nF64; i += 32) rF64[z * LR.nF64 + i] = rF64[x * LR.nF64 + i] + rF64[y * LR.nF64 + i];
  for (int i = lane; i < LE.nAdd; i += 32) bAdd[z * LE.nAdd + i] = bAdd[x * LE.nAdd + i] + bAdd[y * LE.nAdd + i] - cAdd[z * LE.nAdd + i];
  for (int i = lane; i < LE.nF64; i += 32) bF64[z * LE.nF64 + i] = bF64[x * LE.nF64 + i] + bF64[y * LE.nF64 + i] - cF64[z * LE.nF64 + i];
}

// The narrow top of a deep tree (every remaining level holds <= 256 nodes; a greedy or random merge order on 10^5
// regions has ~10^4 such levels): ONE CTA walks those levels in order, a warp per node, a barrier per level, instead of
// one launch per level.  trip[i] = (node, child0, child1) in level order.
__global__ void __launch_bounds__(1024)
level_tail_kernel(const int4* __restrict__ trip, const int* __restrict__ levelOff, int l0, int l1,
                  RecLayout LR, RecLayout LE, uint32_t* rAdd, uint32_t* rMax, double* rF64, uint32_t* bAdd, double* bF64,
                  const uint32_t* __restrict__ cAdd, const double* __restrict__ cF64) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int l = l0; l <= l1; ++l) {
    const int beg = levelOff[l], cnt = levelOff[l + 1] - beg;
    for (int w = warp; w < cnt; w += 32) {
      const int4 t = trip[beg + w];
      const size_t z = t.x, x = t.y, y = t.z;
      for (int i = lane; i < LR.nAdd; i += 32) rAdd[z * LR.nAdd + i] = rAdd[x * LR.nAdd + i] + rAdd[y * LR.nAdd + i];
      for (int i = lane; i < LR.nMax; i += 32) rMax[z * LR.nMax + i] = max(rMax[x * LR.nMax + i], rMax[y * LR.nMax + i]);
      for (int i = lane; i < LR.nF64; i += 32) rF64[z * LR.nF64 + i] = rF64[x * LR.nF64 + i] + rF64[y * LR.nF64 + i];
      for (int i = lane; i < LE.nAdd; i += 32) bAdd[z * LE.nAdd + i] = bAdd[x * LE.nAdd + i] + bAdd[y * LE.nAdd + i] - cAdd[z * LE.nAdd + i];
      for (int i = lane; i < LE.nF64; i += 32) bF64[z * LE.nF64 + i] = bF64[x * LE.nF64 + i] + bF64[y * LE.nF64 + i] - cF64[z * LE.nF64 + i];
    }
    __syncthreads();  // the next level reads what this one wrote (same CTA: the barrier orders global memory too)
  }
}

struct BcFeatView {
  int n, R, withSal, dim;
  int first, count;                    // rows [first, first + count) are written (a rank's share, shard.cu)
  const int* nodeX; const int* nodeY;  // [n] nodes merged at step i (node of the step = R + i)
  const double* nodeSal;               // [N] or null
  const uint32_t* rAdd; const uint32_t* rMax; const double* rF64;
  const uint32_t* bAdd; const uint32_t* bMax; const double* bF64;
  const uint32_t* cAdd; const uint32_t* cMax; const double* cF64;
  const uint32_t* nAdd; const uint32_t* nMax; const double* nF64;
  const double* med; int nMed;         // [n][nMed] medians of the rows' voxel sets (featurize.cuh), or null
};

__global__ void __launch_bounds__(128)
bc_feat_rows_kernel(const __grid_constant__ FeatCfg F, const __grid_constant__ BcFeatView V,
                    double* __restrict__ rows) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= V.count) return;
  rows += (size_t)i * V.dim;
  i += V.first;
  size_t x = V.nodeX[i], y = V.nodeY[i], z = (size_t)V.R + i;
  const RecLayout LR = F.reg, LE = F.edge;
  auto R = [&](size_t v) { return RecSingle{V.rAdd + v * LR.nAdd, V.rMax + v * LR.nMax, V.rF64 + v * LR.nF64}; };
  auto B = [&](size_t v) { return RecSingle{V.bAdd + v * LE.nAdd, V.bMax + v * LE.nMax, V.bF64 + v * LE.nF64}; };
  RecSum2 pb{RecSingle{V.cAdd + z * LE.nAdd, V.cMax + z * LE.nMax, V.cF64 + z * LE.nF64},
             RecSingle{V.nAdd + z * LE.nAdd, V.nMax + z * LE.nMax, V.nF64 + z * LE.nF64}};
  double sal[3];
  if (V.withSal) { sal[0] = V.nodeSal[x]; sal[1] = V.nodeSal[y]; sal[2] = V.nodeSal[z]; }
  featurize_row(F, R(x), R(y), R(z), B(x), B(y), B(z), pb, V.withSal ? sal : nullptr,
                rows, -1, false, V.med ? V.med + (size_t)i * V.nMed : nullptr);
}

}  // namespace glia

using namespace glia;
#define CK(expr) GLIA_CUDA_CHECK(h, expr)
static inline unsigned nblk(long long n, int per) { return (unsigned)((n + per - 1) / per); }

namespace {
struct Scratch {  // the i-th request of a call reuses the handle's i-th grow-only buffer: no cudaMalloc / cudaFree after the first call
  glia_handle* h; int next = 0;
  explicit Scratch(glia_handle* hh) : h(hh) {}
  template <class T> cudaError_t get(T** out, size_t count, bool zero, cudaStream_t st) {
    if (next >= (int)(sizeof(h->k5) / sizeof(h->k5[0]))) return cudaErrorMemoryAllocation;
    glia::DevBuf& b = h->k5[next++];
    const size_t bytes = sizeof(T) * (count ? count : 1);
    cudaError_t e = b.ensure(bytes);
    if (e != cudaSuccess) return e;
    *out = (T*)b.p;
    if (zero) e = cudaMemsetAsync(b.p, 0, bytes, st);
    return e;
  }
};
}  // namespace

// K5 for one handle, or - with a reducer - for one rank of a z-slab sharded build (shard.cu): every directed boundary
// lives on exactly one rank, so the per-node sums / maxima scattered from the local boundaries are partial and one
// all-reduce per array completes them; the tree, the level-wise combination (cheap) and the region records (already
// global) are replicated, and each rank featurises its contiguous share of the merges.
int glia_bc_feat_impl(glia_handle* h, const uint32_t* order, int64_t n64, const double* sal_in, double init_sal,
                      double sal_bias, double* rows, glia::TableReducer* red, int world, int rank, int64_t* firstOut,
                      int64_t* countOut) {
  if (!h || (n64 > 0 && (!order || !rows)) || n64 < 0 || n64 > 0x3fffffff) return GLIA_E_ARG;
  if (!h->built) return GLIA_E_STATE;
  if (firstOut) *firstOut = 0;
  if (countOut) *countOut = 0;
  if (red && h->fc.nMed()) { h->err = "median features need the whole volume in one handle"; return GLIA_E_STATE; }
  if (n64 == 0) return GLIA_OK;
  CK(cudaSetDevice(h->prm.device));
  cudaStream_t st = h->stream;
  const int n = (int)n64, R = (int)h->R, E = (int)h->E, N = R + n;
  const int withSal = sal_in != nullptr;
  const RecLayout LR = h->vp.reg, LE = h->vp.edge;
  PhaseTrace tr(st);
  // ---- host: labels -> nodes, tree, levels
  std::vector<uint32_t> lab(R);
  if (R) CK(cudaMemcpyAsync(lab.data(), h->regLabel.p, 4 * (size_t)R, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  std::vector<int> parent(N), c0(N, -1), c1(N, -1), level(N, 0), nodeX(n), nodeY(n);
  std::vector<double> nodeSal(N, init_sal);
  for (int v = 0; v < N; ++v) parent[v] = v;
  // keys of merged regions -> nodes: a direct table when they are dense (the reference numbers them upwards from the
  // largest label + 1, util/struct_merge.hxx:18-21), a hash map otherwise
  uint32_t kMin = 0xFFFFFFFFu, kMax = 0;
  for (int i = 0; i < n; ++i) { kMin = std::min(kMin, order[3 * i + 2]); kMax = std::max(kMax, order[3 * i + 2]); }
  const bool dense = (uint64_t)kMax - kMin < 8ull * (uint64_t)n + 1024;
  std::vector<int> direct;
  std::unordered_map<uint32_t, int> merged;
  if (dense) direct.assign((size_t)(kMax - kMin) + 1, -1); else merged.reserve((size_t)n * 2);
  auto set_merged = [&](uint32_t key, int node) { if (dense) direct[key - kMin] = node; else merged[key] = node; };
  auto resolve = [&](uint32_t key) -> int {
    if (dense) { if (key >= kMin && key <= kMax && direct[key - kMin] >= 0) return direct[key - kMin]; }
    else { auto it = merged.find(key); if (it != merged.end()) return it->second; }
    auto lb = std::lower_bound(lab.begin(), lab.end(), key);
    return (lb != lab.end() && *lb == key) ? (int)(lb - lab.begin()) : -1;
  };
  int maxLevel = 0;
  for (int i = 0; i < n; ++i) {
    int x = resolve(order[3 * i]), y = resolve(order[3 * i + 1]), z = R + i;
    if (x < 0 || y < 0 || x == y || parent[x] != x || parent[y] != y) {
      h->err = "merge order: step " + std::to_string(i) + " references an unknown or already merged region";
      return GLIA_E_ORDER;
    }
    parent[x] = parent[y] = z; c0[z] = x; c1[z] = y; nodeX[i] = x; nodeY[i] = y;
    level[z] = 1 + std::max(level[x], level[y]);
    maxLevel = std::max(maxLevel, level[z]);
    set_merged(order[3 * i + 2], z);
    nodeSal[z] = withSal ? sal_in[i] + sal_bias : 0.0;  // hmt/bc_feat.hxx:12-26
  }
  std::vector<int> levelOff(maxLevel + 2, 0), levelNodes(n);
  for (int i = 0; i < n; ++i) levelOff[level[R + i] + 1]++;
  for (int l = 0; l <= maxLevel; ++l) levelOff[l + 1] += levelOff[l];
  { std::vector<int> pos(levelOff.begin(), levelOff.end() - 1);
    for (int i = 0; i < n; ++i) levelNodes[pos[level[R + i]]++] = R + i; }
  int K = 1; while ((1ll << K) < N) ++K;
  ++K;
  tr.mark("k5: host tree");
  // ---- device buffers
  Scratch S(h);
  int *dUp, *dCntA, *dCntB, *dC0, *dC1, *dX, *dY, *dLevelNodes, *dLca, *dTopd, *dHasNM;
  double* dSal = nullptr; double* dRows;
  uint32_t *rAdd, *rMax, *bAdd, *bMax, *cAdd, *cMax, *nAdd, *nMax;
  double *rF64, *bF64, *cF64, *nF64;
  CK(S.get(&dUp, (size_t)K * N, false, st)); CK(S.get(&dCntA, N, false, st)); CK(S.get(&dCntB, N, false, st));
  CK(S.get(&dC0, N, false, st)); CK(S.get(&dC1, N, false, st)); CK(S.get(&dX, n, false, st)); CK(S.get(&dY, n, false, st));
  CK(S.get(&dLevelNodes, n, false, st)); CK(S.get(&dLca, E, false, st)); CK(S.get(&dTopd, R, false, st));
  CK(S.get(&dHasNM, R, true, st));
  CK(S.get(&rAdd, (size_t)N * LR.nAdd, false, st)); CK(S.get(&rMax, (size_t)N * LR.nMax, false, st));
  CK(S.get(&rF64, (size_t)N * LR.nF64, false, st));
  CK(S.get(&bAdd, (size_t)N * LE.nAdd, false, st)); CK(S.get(&bMax, (size_t)N * LE.nMax, true, st));
  CK(S.get(&bF64, (size_t)N * LE.nF64, false, st));
  CK(S.get(&cAdd, (size_t)N * LE.nAdd, true, st)); CK(S.get(&cMax, (size_t)N * LE.nMax, true, st));
  CK(S.get(&cF64, (size_t)N * LE.nF64, true, st));
  CK(S.get(&nAdd, (size_t)N * LE.nAdd, true, st)); CK(S.get(&nMax, (size_t)N * LE.nMax, true, st));
  CK(S.get(&nF64, (size_t)N * LE.nF64, true, st));
  const int dim = glia_b200_row_dim(h, withSal);
  const int share = (n + world - 1) / world;
  const int first = std::min(n, rank * share), count = std::min(n, first + share) - first;
  if (firstOut) *firstOut = first;
  if (countOut) *countOut = count;
  if (h->prm.device_pointers) dRows = rows; else CK(S.get(&dRows, (size_t)std::max(count, 1) * dim, false, st));
  if (withSal) { CK(S.get(&dSal, N, false, st)); CK(cudaMemcpyAsync(dSal, nodeSal.data(), 8 * (size_t)N, cudaMemcpyHostToDevice, st)); }
  std::vector<int> cnt0(N);
  for (int v = 0; v < N; ++v) cnt0[v] = parent[v] != v;
  CK(cudaMemcpyAsync(dUp, parent.data(), 4 * (size_t)N, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(dCntA, cnt0.data(), 4 * (size_t)N, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(dC0, c0.data(), 4 * (size_t)N, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(dC1, c1.data(), 4 * (size_t)N, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(dX, nodeX.data(), 4 * (size_t)n, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(dY, nodeY.data(), 4 * (size_t)n, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(dLevelNodes, levelNodes.data(), 4 * (size_t)n, cudaMemcpyHostToDevice, st));
  CK(cudaMemsetAsync(dTopd, 0x7f, 4 * (size_t)R, st));
  // leaves: region records and boundary-of aggregates of the base regions
  { int s = glia_ensure_full_regions(h); if (s != GLIA_OK) return s; }
  CK(cudaMemcpyAsync(rAdd, h->cAdd.p, 4 * (size_t)R * LR.nAdd, cudaMemcpyDeviceToDevice, st));
  CK(cudaMemcpyAsync(rMax, h->cMax.p, 4 * (size_t)R * LR.nMax, cudaMemcpyDeviceToDevice, st));
  CK(cudaMemcpyAsync(rF64, h->cF64.p, 8 * (size_t)R * LR.nF64, cudaMemcpyDeviceToDevice, st));
  CK(cudaMemcpyAsync(bAdd, h->bAggAdd.p, 4 * (size_t)R * LE.nAdd, cudaMemcpyDeviceToDevice, st));
  CK(cudaMemcpyAsync(bF64, h->bAggF64.p, 8 * (size_t)R * LE.nF64, cudaMemcpyDeviceToDevice, st));
  tr.mark("k5: alloc + upload");
  // ---- binary lifting + depth
  int* cntPrev = dCntA; int* cntNext = dCntB;
  for (int k = 0; k + 1 < K; ++k) {
    lift_kernel<<<nblk(N, 256), 256, 0, st>>>(dUp + (size_t)k * N, cntPrev, dUp + (size_t)(k + 1) * N, cntNext, N);
    std::swap(cntPrev, cntNext);
    h->launches++;
  }
  TreeView T{N, R, K, dUp, cntPrev};
  tr.mark("k5: lifting");
  if (E > 0) {
    edge_lca_kernel<<<nblk(E, 256), 256, 0, st>>>(T, h->ea.as<int>(), h->eb.as<int>(), h->rev.as<int>(), E, dLca, dTopd, dHasNM,
                                                  (int)h->ownLo, (int)h->ownHi);
  }
  if (red) {  // which base regions still own live entries is a property of ALL boundaries
    int s;
    if ((s = red->all_reduce(dTopd, (size_t)R, glia::TableReducer::I32, glia::TableReducer::MIN, st))) return s;
    if ((s = red->all_reduce(dHasNM, (size_t)R, glia::TableReducer::I32, glia::TableReducer::MAX, st))) return s;
  }
  if (E > 0) {
    edge_scatter_kernel<<<nblk((long long)E * 32, 256), 256, 0, st>>>(
        T, h->ea.as<int>(), h->eb.as<int>(), h->rev.as<int>(), dLca, dTopd, dHasNM, E, LE, h->sAdd.as<uint32_t>(),
        h->sMax.as<uint32_t>(), h->sF64.as<double>(), cAdd, cMax, cF64, nAdd, nMax, nF64, bMax, (int)h->ownLo, (int)h->ownHi);
    h->launches += 2;
  }
  if (red) {
    typedef glia::TableReducer TR;
    int s;
    // leaves: boundary-of aggregates, each taken from the rank that owns the region
    if (h->ownLo > 0) { CK(cudaMemsetAsync(bAdd, 0, 4 * (size_t)h->ownLo * LE.nAdd, st)); CK(cudaMemsetAsync(bF64, 0, 8 * (size_t)h->ownLo * LE.nF64, st)); }
    if (h->ownHi < R) { CK(cudaMemsetAsync(bAdd + (size_t)h->ownHi * LE.nAdd, 0, 4 * (size_t)(R - h->ownHi) * LE.nAdd, st));
                        CK(cudaMemsetAsync(bF64 + (size_t)h->ownHi * LE.nF64, 0, 8 * (size_t)(R - h->ownHi) * LE.nF64, st)); }
    if ((s = red->all_reduce(bAdd, (size_t)R * LE.nAdd, TR::U32, TR::SUM, st))) return s;
    if ((s = red->all_reduce(bF64, (size_t)R * LE.nF64, TR::F64, TR::SUM, st))) return s;
    if ((s = red->all_reduce(bMax, (size_t)N * LE.nMax, TR::U32, TR::MAX, st))) return s;
    if ((s = red->all_reduce(cAdd, (size_t)N * LE.nAdd, TR::U32, TR::SUM, st))) return s;
    if ((s = red->all_reduce(cMax, (size_t)N * LE.nMax, TR::U32, TR::MAX, st))) return s;
    if ((s = red->all_reduce(cF64, (size_t)N * LE.nF64, TR::F64, TR::SUM, st))) return s;
    if ((s = red->all_reduce(nAdd, (size_t)N * LE.nAdd, TR::U32, TR::SUM, st))) return s;
    if ((s = red->all_reduce(nMax, (size_t)N * LE.nMax, TR::U32, TR::MAX, st))) return s;
    if ((s = red->all_reduce(nF64, (size_t)N * LE.nF64, TR::F64, TR::SUM, st))) return s;
  }
  tr.mark("k5: boundary lca + scatter");
  // ---- per-level update of merged-region records
  int tail0 = maxLevel + 1;  // first level of the narrow top: every level from here on holds <= 256 nodes
  while (tail0 > 1 && levelOff[tail0] - levelOff[tail0 - 1] <= 256) --tail0;
  for (int l = 1; l < tail0; ++l) {
    int cnt = levelOff[l + 1] - levelOff[l];
    if (!cnt) continue;
    level_combine_kernel<<<nblk((long long)cnt * 32, 256), 256, 0, st>>>(
        dLevelNodes + levelOff[l], cnt, dC0, dC1, LR, LE, rAdd, rMax, rF64, bAdd, bF64, cAdd, cF64);
    h->launches++;
  }
  if (tail0 <= maxLevel) {
    std::vector<int4> trip(n);
    for (int i = 0; i < n; ++i) { const int z = levelNodes[i]; trip[i] = make_int4(z, c0[z], c1[z], 0); }
    int4* dTrip; int* dLevelOff;
    CK(S.get(&dTrip, (size_t)n, false, st)); CK(S.get(&dLevelOff, levelOff.size(), false, st));
    CK(cudaMemcpyAsync(dTrip, trip.data(), sizeof(int4) * (size_t)n, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(dLevelOff, levelOff.data(), 4 * levelOff.size(), cudaMemcpyHostToDevice, st));
    level_tail_kernel<<<1, 1024, 0, st>>>(dTrip, dLevelOff, tail0, maxLevel, LR, LE, rAdd, rMax, rF64, bAdd, bF64, cAdd, cF64);
    h->launches++;
    CK(cudaStreamSynchronize(st));  // trip is a local
  }
  tr.mark("k5: level combine");
  // ---- medians (GLIA_USE_MEDIAN_AS_FEATS): every set of a row is a union of base segments named by the tree
  double* dMed = nullptr;
  const int nMed = h->fc.nMed();
  if (nMed) {
    const int nR = h->prm.n_rimg, nB = h->prm.n_bimg;
    std::vector<int> ea(E), eb(E), rev(E), row(R + 1);
    if (E) { CK(cudaMemcpyAsync(ea.data(), h->ea.p, 4 * (size_t)E, cudaMemcpyDeviceToHost, st));
             CK(cudaMemcpyAsync(eb.data(), h->eb.p, 4 * (size_t)E, cudaMemcpyDeviceToHost, st));
             CK(cudaMemcpyAsync(rev.data(), h->rev.p, 4 * (size_t)E, cudaMemcpyDeviceToHost, st)); }
    CK(cudaMemcpyAsync(row.data(), h->row.p, 4 * (size_t)(R + 1), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    // leaves in depth-first order: the base regions of a node are one range [lo, hi) of `leaf`
    std::vector<int> leaf, pos(R, -1), lo(N, 0), hi(N, 0);
    leaf.reserve(R);
    {
      std::vector<std::pair<int, int>> stack;  // (node, state)
      for (int root = 0; root < N; ++root) {
        if (parent[root] != root) continue;
        stack.push_back({root, 0});
        while (!stack.empty()) {
          auto [v, state] = stack.back();
          if (v < R) { lo[v] = (int)leaf.size(); pos[v] = lo[v]; leaf.push_back(v); hi[v] = lo[v] + 1; stack.pop_back(); continue; }
          if (state == 0) { lo[v] = (int)leaf.size(); stack.back().second = 1; stack.push_back({c0[v], 0}); }
          else if (state == 1) { stack.back().second = 2; stack.push_back({c1[v], 0}); }
          else { hi[v] = (int)leaf.size(); stack.pop_back(); }
        }
      }
    }
    auto inside = [&](int b, int v) { return pos[b] >= lo[v] && pos[b] < hi[v]; };
    auto live = [&](int e, int v) { return rev[e] < 0 || !inside(eb[e], v); };  // entry of node v (type/region.hxx:66-75)
    std::vector<double> nodeMed((size_t)N * (nR + nB), 0.0), pairMed((size_t)n * (nB ? nB : 1), 0.0);
    std::vector<int> qKind, qLen, seg; std::vector<long long> qOff; std::vector<double*> qDst;
    auto flush = [&]() -> int {
      if (qKind.empty()) return GLIA_OK;
      std::vector<double> res(qKind.size());
      int s = glia_median_queries(h, qKind, qOff, qLen, seg, res.data(), nullptr);
      if (s != GLIA_OK) return s;
      for (size_t q = 0; q < res.size(); ++q) *qDst[q] = res[q];
      qKind.clear(); qLen.clear(); qOff.clear(); qDst.clear(); seg.clear();
      return GLIA_OK;
    };
    auto add_queries = [&](int store, int nImg, long long off, int len, double* dst) {
      for (int i = 0; i < nImg; ++i) { qKind.push_back(store * GLIA_MAX_IMG + i); qOff.push_back(off); qLen.push_back(len); qDst.push_back(dst + i); }
    };
    std::vector<char> used(N, 0);
    for (int i = 0; i < n; ++i) used[nodeX[i]] = used[nodeY[i]] = used[R + i] = 1;
    for (int v = 0; v < N; ++v) {
      if (!used[v]) continue;
      long long off = (long long)seg.size();
      for (int k = lo[v]; k < hi[v]; ++k) seg.push_back(leaf[k]);
      add_queries(0, nR, off, hi[v] - lo[v], &nodeMed[(size_t)v * (nR + nB)]);
      if (nB) {
        off = (long long)seg.size();
        for (int k = lo[v]; k < hi[v]; ++k)
          for (int e = row[leaf[k]]; e < row[leaf[k] + 1]; ++e) if (live(e, v)) seg.push_back(e);
        add_queries(1, nB, off, (int)((long long)seg.size() - off), &nodeMed[(size_t)v * (nR + nB) + nR]);
      }
      if (v >= R && nB) {  // pair boundary of the merge (util/struct.hxx:10-16, type/region.hxx:42-51)
        off = (long long)seg.size();
        const int sides[2][2] = {{c0[v], c1[v]}, {c1[v], c0[v]}};
        for (auto& sd : sides) {
          const int A = sd[0], B = sd[1];
          for (int k = lo[A]; k < hi[A]; ++k)
            for (int e = row[leaf[k]]; e < row[leaf[k] + 1]; ++e) {
              if (!live(e, A) || !inside(eb[e], B)) continue;
              bool act = false;  // the target must still own an entry inside its side
              for (int e2 = row[eb[e]]; e2 < row[eb[e] + 1] && !act; ++e2) act = live(e2, B);
              if (act) seg.push_back(e);
            }
        }
        add_queries(1, nB, off, (int)((long long)seg.size() - off), &pairMed[(size_t)(v - R) * nB]);
      }
      if (seg.size() > (32u << 20)) { int s = flush(); if (s != GLIA_OK) return s; }
    }
    { int s = flush(); if (s != GLIA_OK) return s; }
    std::vector<double> med((size_t)n * nMed);
    for (int i = 0; i < n; ++i) {
      const int x = nodeX[i], y = nodeY[i], z = R + i;
      double* m = &med[(size_t)i * nMed];
      for (int k = 0; k < nR; ++k) {
        m[3 * k] = nodeMed[(size_t)x * (nR + nB) + k]; m[3 * k + 1] = nodeMed[(size_t)y * (nR + nB) + k];
        m[3 * k + 2] = nodeMed[(size_t)z * (nR + nB) + k];
      }
      for (int k = 0; k < nB; ++k) {
        double* q = m + 3 * nR + 4 * k;
        q[0] = nodeMed[(size_t)x * (nR + nB) + nR + k]; q[1] = nodeMed[(size_t)y * (nR + nB) + nR + k];
        q[2] = nodeMed[(size_t)z * (nR + nB) + nR + k]; q[3] = pairMed[(size_t)i * nB + k];
      }
    }
    CK(S.get(&dMed, med.size(), false, st));
    CK(cudaMemcpyAsync(dMed, med.data(), 8 * med.size(), cudaMemcpyHostToDevice, st));
    CK(cudaStreamSynchronize(st));  // med is a local
  }
  BcFeatView V{n, R, withSal, dim, first, count, dX, dY, dSal, rAdd, rMax, rF64, bAdd, bMax, bF64, cAdd, cMax, cF64, nAdd, nMax, nF64, dMed, nMed};
  if (count) bc_feat_rows_kernel<<<nblk(count, 128), 128, 0, st>>>(h->fc, V, dRows);
  h->launches++;
  CK(cudaGetLastError());
  tr.mark("k5: rows kernel");
  if (!h->prm.device_pointers && count) { int s = glia_copy_to_host(h, rows, dRows, 8 * (size_t)count * dim); if (s != GLIA_OK) return s; }
  CK(cudaStreamSynchronize(st));
  tr.mark("k5: rows to host");
  if (tr.on) fprintf(stderr, "[glia trace] k5: %d merges, %d levels\n", n, maxLevel);
  return GLIA_OK;
}

extern "C" int glia_b200_bc_feat(glia_handle* h, const uint32_t* order, int64_t n64,
                                 const double* sal_in, double init_sal, double sal_bias,
                                 double* rows) {
  return glia_bc_feat_impl(h, order, n64, sal_in, init_sal, sal_bias, rows, nullptr, 1, 0, nullptr, nullptr);
}
