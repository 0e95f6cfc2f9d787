// K7: random-forest scoring of merge candidates + the rf_old model-file reader.
//
// Replaces (reference, /root/reference/code):
//   alg::RandomForest::operator()              alg/rf.hxx:35-36
//   ml::rf::Model::predict / helperPredict     ml/rf/rf.hxx:362-439
//   extern classForest (third party)           ml/rf/rf.hxx:8-14  -> predictClassTree semantics:
//     k = 0; while nodestatus[k] != -1: k = (x[bestvar[k]-1] <= xbestsplit[k]) ? left : right;
//     votes[nodeclass[k]-1] += 1            (source absent from the reference tree)
//   rf_old::readModelFromBinaryFile, readArray ml/rf/ml_rf_model.cxx:459-563, :48-70
//
// Device layout: every tree is compacted to its used nodes (ndbigtree[t], not the padded
// nrnodes stride) and re-ordered breadth-first with sibling pairs adjacent, so a node is one
// 16-byte record {threshold, left-child (right = left+1) | class, feature | -1}.  One thread
// scores one candidate; all lanes of a warp walk the same tree at the same time, so the upper
// levels are broadcast fetches.  Trees whose nodes fit are staged in shared memory.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <math.h>
#include <queue>
#include "handle.cuh"
#include "rf_walk.cuh"

namespace glia {

struct __align__(16) RfNode {
  double thr;
  int left;  // internal: absolute index of the left child; leaf: class index
  int feat;  // internal: 0-based feature; leaf: -1
};

// votesOut: [N][nclass] or null.  Dynamic smem: blockDim.x * nclass ints when votesOut.
template <bool WANT_VOTES>
__global__ void __launch_bounds__(128)
rf_score_kernel(const RfNode* __restrict__ nodes, const int* __restrict__ treeOff, const RfDist dist,
                int nclass, int labelIdx, const double* __restrict__ X, size_t N, int D,
                int* __restrict__ votesOut, double* __restrict__ prob) {
  extern __shared__ int svotes[];
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const bool active = i < N;
  const double* x = X + (active ? i : 0) * (size_t)D;
  if (WANT_VOTES)
    for (int c = 0; c < nclass; ++c) svotes[c * blockDim.x + threadIdx.x] = 0;
  int hits = 0;
  if (active) {
    const int mdl = dist.pick(x), tb = dist.begin(mdl), ntree = dist.end(mdl) - tb;
    for (int t = tb; t < tb + ntree; ++t) {
      int k = __ldg(treeOff + t);
      RfNode nd = nodes[k];
      while (nd.feat >= 0) {
        double v = __ldg(x + nd.feat);
        k = nd.left + (v <= nd.thr ? 0 : 1);
        nd = nodes[k];
      }
      if (WANT_VOTES) svotes[nd.left * blockDim.x + threadIdx.x] += 1;
      hits += (nd.left == labelIdx);
    }
    if (prob) prob[i] = (double)hits / (double)ntree;  // ml/rf/rf.hxx:366-369
    if (WANT_VOTES)
      for (int c = 0; c < nclass; ++c) votesOut[i * nclass + c] = svotes[c * blockDim.x + threadIdx.x];
  }
}


// node as one 16-byte load: {thr.lo, thr.hi, left, feat}
__device__ __forceinline__ int4 rf_node(const int4* np, int k) { return __ldg(np + k); }
// alg/rf.hxx:35-36 -> classForest: x <= threshold goes left, anything else (NaN too) goes right.
// Nodes are in the walk encoding (rf_build_device): leaf = sign bit of the feature word, class in the low word.
// A walk that has reached its leaf issues no loads.  Measured on B200 (profiles/r1_rf_v4_notes.txt): stepping
// finished walks unconditionally over their self-loop is 7.5 instructions per level instead of 9 but 28 % slower
// on C5 - the kernel is bound by L1 wavefronts of the divergent 16-byte node loads, not by issue slots.
#define RF_STEP_P(n)                                                                             \
  if (n.w >= 0) {                                                                                \
    const double xv_ = rf_lds(xs + 8u * (uint32_t)n.w);                                          \
    n = rf_node(np, n.z + (xv_ <= __hiloint2double(n.y, n.x) ? 0 : 1)); }

// v2: a CTA scores RF_CPB candidates whose rows are staged ONCE in shared memory (one coalesced
// copy; the v1 kernel re-streamed them from DRAM ~85 times, profiles/r1_v1_ncu_rf.txt); four
// threads share a candidate and split the trees, two trees in flight per thread; sibling pairs
// are 32-byte aligned so that one L2 sector serves either outcome of a split.
template <bool ENS, int NS>  // NS: tree slices per candidate (CTA = 64 * NS threads); ENS: model ensemble, the tree range depends on the row (alg/rf.hxx:97-98)
__global__ void __launch_bounds__(64 * NS, NS == 8 ? 3 : 6)
rf_score_smem_kernel(const RfNode* __restrict__ nodes, const int* __restrict__ treeOff, const RfDist dist,
                     int labelIdx, const double* __restrict__ X, size_t N, int D, int ldRow,
                     double* __restrict__ prob) {
  extern __shared__ __align__(16) double srow[];
  const size_t c0 = (size_t)blockIdx.x * RF_CPB;
  const int ncand = (int)min((size_t)RF_CPB, N - c0);
  // coalesced copy of the CTA's rows (contiguous in X) into padded shared rows
  for (int i = threadIdx.x; i < ncand * D; i += blockDim.x) {
    int c = i / D, f = i - c * D;
    srow[c * ldRow + f] = X[c0 * D + i];
  }
  __syncthreads();
  // v3 mapping: a warp = 32 candidates walking the SAME trees (lane = candidate), the 8 warps of
  // the CTA = 2 candidate halves x 4 tree slices.  The upper levels of a tree are then one or two
  // 32-byte sectors per warp-level fetch instead of 4 trees x 8 candidates worth of sectors
  // (v2 mapped 4 lanes to one candidate; L1 wavefronts were the limiter, profiles/r1_v3_ncu_c4s.txt).
  __shared__ int svote[RF_CPB];
  __shared__ int sntree[RF_CPB];
  __shared__ volatile uint32_t spin[1];
  __shared__ volatile unsigned long long spin64[1];
  if (threadIdx.x < RF_CPB) svote[threadIdx.x] = 0;
  if (threadIdx.x == 0) { spin[0] = (uint32_t)__cvta_generic_to_shared(srow); spin64[0] = (unsigned long long)nodes; }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cand = (warp & 1) * 32 + lane, slice = warp >> 1;
  int hits = 0;
  if (cand < ncand) {
    const double* x = srow + cand * ldRow;
    const int mdl = ENS ? dist.pick(x) : 0;
    const int tbeg = ENS ? dist.begin(mdl) : 0, ntree = ENS ? dist.end(mdl) : dist.te[0];  // uniform when !ENS
    if (slice == 0) sntree[cand] = ntree - tbeg;
    // The row base (a 32-bit shared-window address) and the node base are read back from shared memory
    // once: a value ptxas can recompute from a kernel parameter is rematerialised inside every step
    // (S2R + LDC + 5 integer ops; 20 instructions per tree level where 8 do the work).
    const uint32_t xs = spin[0] + 8u * (uint32_t)(cand * ldRow);
    const int4* np = reinterpret_cast<const int4*>(spin64[0]);
    int t = tbeg + slice;
    for (; t + 3 * NS < ntree; t += 4 * NS) {  // four trees in flight
      int4 na = rf_node(np, __ldg(treeOff + t)), nb = rf_node(np, __ldg(treeOff + t + NS));
      int4 nc = rf_node(np, __ldg(treeOff + t + 2 * NS)), nd = rf_node(np, __ldg(treeOff + t + 3 * NS));
      while ((na.w & nb.w & nc.w & nd.w) >= 0) {  // some walk still on an internal node (leaf: sign bit)
        RF_STEP_P(na) RF_STEP_P(nb) RF_STEP_P(nc) RF_STEP_P(nd)
      }
      hits += (na.x == labelIdx) + (nb.x == labelIdx) + (nc.x == labelIdx) + (nd.x == labelIdx);
    }
    for (; t < ntree; t += NS) {
      int4 nd = rf_node(np, __ldg(treeOff + t));
      while (nd.w >= 0) RF_STEP_P(nd)
      hits += (nd.x == labelIdx);
    }
    atomicAdd(&svote[cand], hits);
  }
  __syncthreads();
  if (threadIdx.x < ncand) prob[c0 + threadIdx.x] = (double)svote[threadIdx.x] / (double)sntree[threadIdx.x];
}


// One walk per thread: measured on B200 (profiles/r2_rf_notes.txt) more resident warps beat more walks in flight
// per thread (28 registers -> 4 CTAs of 512 threads = 64 warps per SM).
template <bool ENS, int NS>  // NS tree slices per candidate (CTA = 64 * NS threads)
__global__ void __launch_bounds__(64 * NS, NS == 8 ? 4 : 8)
rf_score_blk_kernel(const uint4* __restrict__ blocks, const double* __restrict__ thr, const int* __restrict__ treeRoot, const RfDist dist,
                    int labelIdx, const double* __restrict__ X, size_t N, int D, int ldRow,
                    double* __restrict__ prob) {
  extern __shared__ __align__(16) uint32_t frow[];  // the CTA's rows as 25-bit order keys, odd stride: one bank per lane for equal features
  const size_t c0 = (size_t)blockIdx.x * RF_CPB;
  const int ncand = (int)min((size_t)RF_CPB, N - c0);
  for (int i = threadIdx.x; i < ncand * D; i += blockDim.x) {
    int c = i / D, f = i - c * D;
    frow[c * ldRow + f] = rf_xkey(X[c0 * D + i]);
  }
  __shared__ int svote[RF_CPB];
  __shared__ int sntree[RF_CPB];
  if (threadIdx.x < RF_CPB) svote[threadIdx.x] = 0;
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cand = (warp & 1) * 32 + lane, slice = warp >> 1;
  if (cand < ncand) {
    const double* xg = X + (c0 + cand) * (size_t)D;
    const int mdl = ENS ? dist.pick(xg) : 0;
    const int tbeg = ENS ? dist.begin(mdl) : 0, ntree = ENS ? dist.end(mdl) : dist.te[0];
    if (slice == 0) sntree[cand] = ntree - tbeg;
    const uint32_t xs = (uint32_t)__cvta_generic_to_shared(frow + cand * ldRow);
    int hits = 0;
    for (int t = tbeg + slice; t < ntree; t += NS)
      hits += (int)rf_walk_blocks(blocks, thr, (uint32_t)__ldg(treeRoot + t), xs, xg, (uint32_t)labelIdx);
    atomicAdd(&svote[cand], hits);
  }
  __syncthreads();
  if (threadIdx.x < ncand) prob[c0 + threadIdx.x] = (double)svote[threadIdx.x] / (double)sntree[threadIdx.x];
}

}  // namespace glia

using namespace glia;

static int rf_build_device(glia_rf* m) {
  // breadth-first compaction, siblings adjacent
  std::vector<RfNode> nodes;
  std::vector<int> treeOff(m->ntree + 1);
  const int nr = m->nrnodes;
  for (int t = 0; t < m->ntree; ++t) {
    const size_t o = (size_t)t * nr;
    const int* tm = m->treemap.data() + 2 * o;
    if (nodes.size() & 1) nodes.push_back(RfNode{0.0, 0, -1});
    treeOff[t] = (int)nodes.size();  // even: sibling pairs stay 32-byte aligned
    std::queue<std::pair<int, int>> q;  // (original node, new index)
    nodes.push_back(RfNode{0.0, 0, -1});
    nodes.push_back(RfNode{0.0, 0, -1});  // pad
    q.push({0, treeOff[t]});
    int visited = 0;
    while (!q.empty()) {
      auto cur = q.front(); q.pop();
      int k = cur.first;
      if (k < 0 || k >= nr || ++visited > nr) return GLIA_E_IO;
      if (m->nodestatus[o + k] == -1) {
        int cls = m->nodeclass[o + k] - 1;
        if (cls < 0 || cls >= m->nclass) return GLIA_E_IO;
        nodes[cur.second] = RfNode{0.0, cls, -1};
      } else {
        int f = m->bestvar[o + k] - 1;
        if (f < 0) return GLIA_E_IO;
        if (f + 1 > m->mdim) m->mdim = f + 1;
        int left = (int)nodes.size();
        nodes.push_back(RfNode{0.0, 0, -1});
        nodes.push_back(RfNode{0.0, 0, -1});
        nodes[cur.second] = RfNode{m->xbestsplit[o + k], left, f};
        q.push({tm[2 * k] - 1, left});
        q.push({tm[2 * k + 1] - 1, left + 1});
      }
    }
  }
  treeOff[m->ntree] = (int)nodes.size();
  m->totalNodes = (int)nodes.size();
  if (m->dist.n <= 1) { m->dist.n = 1; m->dist.tb[0] = 0; m->dist.te[0] = m->ntree; }
  if (cudaSetDevice(m->device) != cudaSuccess) return GLIA_E_CUDA;
  if (cudaMalloc(&m->d_nodes, sizeof(RfNode) * nodes.size()) != cudaSuccess) return GLIA_E_CUDA;
  if (cudaMalloc((void**)&m->d_treeOff, 4 * treeOff.size()) != cudaSuccess) return GLIA_E_CUDA;
  if (cudaMemcpy(m->d_nodes, nodes.data(), sizeof(RfNode) * nodes.size(), cudaMemcpyHostToDevice) != cudaSuccess) return GLIA_E_CUDA;
  if (cudaMemcpy(m->d_treeOff, treeOff.data(), 4 * treeOff.size(), cudaMemcpyHostToDevice) != cudaSuccess) return GLIA_E_CUDA;
  {  // walk encoding: a leaf is {NaN with the class in the low word, self - 1, feature 0 | sign bit}.  "x <= NaN" is
     // false for every x, so the step "left + !(x <= thr)" of a leaf lands on itself: walks need no leaf test.
    std::vector<RfNode> w(nodes);
    for (size_t i = 0; i < w.size(); ++i)
      if (w[i].feat < 0) {
        const uint64_t bits = 0x7FF8000000000000ull | (uint32_t)w[i].left;
        memcpy(&w[i].thr, &bits, 8);
        w[i].left = (int)i - 1;
        w[i].feat = (int)0x80000000u;
      }
    if (cudaMalloc(&m->d_nodesW, sizeof(RfNode) * w.size()) != cudaSuccess) return GLIA_E_CUDA;
    if (cudaMemcpy(m->d_nodesW, w.data(), sizeof(RfNode) * w.size(), cudaMemcpyHostToDevice) != cudaSuccess) return GLIA_E_CUDA;
  }
  // two-level blocks (rf_score_blk_kernel, rf_walk.cuh): two classes, features < 128, block ids < 2^20
  if (m->nclass == 2 && m->mdim <= 128) {
    struct HB { double t[3]; int f[3]; uint32_t leaf, cls, base; };
    std::vector<HB> blk;
    std::vector<int> root(m->ntree);
    bool tooBig = false;
    for (int t = 0; t < m->ntree && !tooBig; ++t) {
      std::queue<std::pair<int, uint32_t>> q;  // (node index in `nodes`, block index)
      root[t] = (int)blk.size();
      blk.push_back(HB());
      q.push({treeOff[t], (uint32_t)root[t]});
      while (!q.empty()) {
        auto cur = q.front(); q.pop();
        const RfNode& r = nodes[cur.first];
        HB B = HB();
        int next[4] = {-1, -1, -1, -1};
        if (r.feat < 0) {  // a tree that is a single leaf: every outcome ends here
          B.leaf = 0xF; B.cls = r.left ? 0xFu : 0u;
        } else {
          B.t[0] = r.thr; B.f[0] = r.feat;
          for (int c = 0; c < 2; ++c) {
            const RfNode& ch = nodes[r.left + c];
            if (ch.feat < 0) { B.leaf |= 3u << (2 * c); if (ch.left) B.cls |= 3u << (2 * c); continue; }
            B.t[1 + c] = ch.thr; B.f[1 + c] = ch.feat;
            for (int g = 0; g < 2; ++g) {
              const RfNode& gc = nodes[ch.left + g];
              const int o = 2 * c + g;
              if (gc.feat < 0) { B.leaf |= 1u << o; if (gc.left) B.cls |= 1u << o; }
              else next[o] = ch.left + g;
            }
          }
        }
        // the four child blocks sit at base + outcome; slots beyond the last continuing outcome are not allocated
        int last = -1;
        for (int o = 0; o < 4; ++o) if (next[o] >= 0) last = o;
        if (blk.size() + 4 >= (1u << 20)) { tooBig = true; break; }
        B.base = (uint32_t)blk.size();
        for (int o = 0; o <= last; ++o) {
          if (next[o] >= 0) q.push({next[o], (uint32_t)blk.size()});
          blk.push_back(HB());
        }
        blk[cur.second] = B;
      }
    }
    if (!tooBig) {
      std::vector<uint4> packed(blk.size());
      std::vector<double> thr(3 * blk.size());
      for (size_t i = 0; i < blk.size(); ++i) {
        uint32_t w[3];
        for (int j = 0; j < 3; ++j) {
          float lo = (float)blk[i].t[j];  // floor_f32(t): largest float <= t
          if ((double)lo > blk[i].t[j]) lo = nextafterf(lo, -INFINITY);
          w[j] = (f2ord(lo) >> 7) | (uint32_t)blk[i].f[j] << RFB_KBITS;
          thr[3 * i + j] = blk[i].t[j];
        }
        packed[i] = make_uint4(w[0], w[1], w[2], blk[i].base | blk[i].leaf << 24 | blk[i].cls << 28);
      }
      if (cudaMalloc(&m->d_blocks, sizeof(uint4) * packed.size()) != cudaSuccess) return GLIA_E_CUDA;
      if (cudaMalloc((void**)&m->d_blockThr, 8 * thr.size()) != cudaSuccess) return GLIA_E_CUDA;
      if (cudaMalloc((void**)&m->d_blockRoot, 4 * root.size()) != cudaSuccess) return GLIA_E_CUDA;
      if (cudaMemcpy(m->d_blocks, packed.data(), sizeof(uint4) * packed.size(), cudaMemcpyHostToDevice) != cudaSuccess) return GLIA_E_CUDA;
      if (cudaMemcpy(m->d_blockThr, thr.data(), 8 * thr.size(), cudaMemcpyHostToDevice) != cudaSuccess) return GLIA_E_CUDA;
      if (cudaMemcpy(m->d_blockRoot, root.data(), 4 * root.size(), cudaMemcpyHostToDevice) != cudaSuccess) return GLIA_E_CUDA;
      m->nBlocks = (long long)packed.size();
    }
  }
  return GLIA_OK;
}

extern "C" int glia_b200_rf_from_arrays(int ntree, int nrnodes, int nclass, int mdim,
                                        const int* treemap, const int* nodestatus,
                                        const double* xbestsplit, const int* bestvar,
                                        const int* nodeclass, const int* ndbigtree,
                                        const int* orig_labels, int device, glia_rf** out) {
  if (!out || ntree <= 0 || nrnodes <= 0 || nclass <= 0 || nclass > 64 || !treemap || !nodestatus ||
      !xbestsplit || !bestvar || !nodeclass || !orig_labels)
    return GLIA_E_ARG;
  glia_rf* m = new glia_rf();
  size_t n = (size_t)ntree * nrnodes;
  m->device = device; m->ntree = ntree; m->nrnodes = nrnodes; m->nclass = nclass; m->mdim = mdim;
  m->treemap.assign(treemap, treemap + 2 * n);
  m->nodestatus.assign(nodestatus, nodestatus + n);
  m->xbestsplit.assign(xbestsplit, xbestsplit + n);
  m->bestvar.assign(bestvar, bestvar + n);
  m->nodeclass.assign(nodeclass, nodeclass + n);
  if (ndbigtree) m->ndbigtree.assign(ndbigtree, ndbigtree + ntree);
  m->orig_labels.assign(orig_labels, orig_labels + nclass);
  int st = rf_build_device(m);
  if (st != GLIA_OK) { glia_b200_rf_destroy(m); return st; }
  *out = m;
  return GLIA_OK;
}

// alg::EnsembleRandomForest (alg/rf.hxx:63-98) with opt::ThresholdModelDistributor
// (type/function.hxx:71-85; --bcm m0 m1 m2 --bcmd dim0 dim1 thr, hmt/main_merge_order_bc.cxx:103-109):
// the models' trees are concatenated into one device forest, a row walks the range of its model.
extern "C" int glia_b200_rf_ensemble(const glia_rf* const* models, int n, int dim0, int dim1,
                                     double threshold, glia_rf** out) {
  if (!models || !out || n < 1 || n > 3 || dim0 < 0 || dim1 < 0) return GLIA_E_ARG;
  for (int i = 0; i < n; ++i) {
    if (!models[i] || models[i]->dist.n != 1 || models[i]->device != models[0]->device ||
        models[i]->nclass != models[0]->nclass || models[i]->orig_labels != models[0]->orig_labels)
      return GLIA_E_ARG;  // one label table for all models
  }
  glia_rf* m = new glia_rf();
  m->device = models[0]->device; m->nclass = models[0]->nclass; m->orig_labels = models[0]->orig_labels;
  int nr = 0, nt = 0;
  for (int i = 0; i < n; ++i) { nr = std::max(nr, models[i]->nrnodes); nt += models[i]->ntree; m->mdim = std::max(m->mdim, models[i]->mdim); }
  m->mdim = std::max(m->mdim, std::max(dim0, dim1) + 1);
  m->nrnodes = nr; m->ntree = nt;
  const size_t tot = (size_t)nt * nr;
  m->treemap.assign(2 * tot, 0); m->nodestatus.assign(tot, 0); m->xbestsplit.assign(tot, 0.0);
  m->bestvar.assign(tot, 0); m->nodeclass.assign(tot, 0);
  m->dist.n = n; m->dist.dim0 = dim0; m->dist.dim1 = dim1; m->dist.thr = threshold;
  int t0 = 0;
  for (int i = 0; i < n; ++i) {
    const glia_rf* s = models[i];
    for (int t = 0; t < s->ntree; ++t) {
      const size_t so = (size_t)t * s->nrnodes, d = (size_t)(t0 + t) * nr;
      for (int k = 0; k < s->nrnodes; ++k) {
        m->treemap[2 * (d + k)] = s->treemap[2 * (so + k)]; m->treemap[2 * (d + k) + 1] = s->treemap[2 * (so + k) + 1];
        m->nodestatus[d + k] = s->nodestatus[so + k]; m->xbestsplit[d + k] = s->xbestsplit[so + k];
        m->bestvar[d + k] = s->bestvar[so + k]; m->nodeclass[d + k] = s->nodeclass[so + k];
      }
      // a tree shorter than the padded stride: its unused nodes are never reached (node 0 is the root)
    }
    m->dist.tb[i] = t0; t0 += s->ntree; m->dist.te[i] = t0;
  }
  for (int i = n; i < 3; ++i) { m->dist.tb[i] = m->dist.tb[n - 1]; m->dist.te[i] = m->dist.te[n - 1]; }
  int st = rf_build_device(m);
  if (st != GLIA_OK) { glia_b200_rf_destroy(m); return st; }
  *out = m;
  return GLIA_OK;
}

// ---- rf_old binary model file -------------------------------------------------------------
// The file starts with a raw dump of rf_old::Model (ml/rf/ml_rf.h:97-157; 520 bytes on
// x86-64/libstdc++): only the {rows, cols} pairs of each array header are meaningful on read
// (the pointers are stale).  Offsets of those pairs, in declaration order:
namespace {
struct ArrDim { int rows, cols; };
// 4 std::vector headers (96 bytes), then array headers {ptr(8), int[2]} = 16 bytes each.
const size_t OFF_NCAT = 96, OFF_CATF = 112, OFF_XBEST = 136, OFF_CLASSWT = 152, OFF_CUTOFF = 168,
             OFF_TREEMAP = 184, OFF_NODESTATUS = 200, OFF_NODECLASS = 216, OFF_BESTVAR = 232,
             OFF_NDBIG = 248, OFF_ORIG = 272, OFF_NEWL = 288;
ArrDim dim_at(const unsigned char* hdr, size_t off) { ArrDim d; memcpy(&d, hdr + off + 8, 8); return d; }

template <class T>
bool read_array(FILE* f, std::vector<T>& out, long long size) {  // ml_rf_model.cxx:48-70
  out.assign((size_t)(size > 0 ? size : 0), T(0));
  if (size <= 0) return true;
  if (size > 128) {
    unsigned char sparse;
    if (fread(&sparse, 1, 1, f) != 1) return false;
    if (sparse) {
      int num;
      if (fread(&num, 4, 1, f) != 1 || num < 0) return false;
      for (int i = 0; i < num; ++i) {
        int idx; T v;
        if (fread(&idx, 4, 1, f) != 1 || fread(&v, sizeof(T), 1, f) != 1) return false;
        if (idx < 0 || idx >= size) return false;
        out[idx] = v;
      }
      return true;
    }
  }
  return fread(out.data(), sizeof(T), (size_t)size, f) == (size_t)size;
}
// file matrix is rows x cols row-major; native layout is its transpose (ml_rf_model.cxx:542-557)
template <class T>
std::vector<T> transposed(const std::vector<T>& x, int rows, int cols) {
  std::vector<T> r(x.size());
  for (int i = 0; i < cols; ++i)
    for (int j = 0; j < rows; ++j) r[(size_t)i * rows + j] = x[(size_t)j * cols + i];
  return r;
}
}  // namespace

extern "C" int glia_b200_rf_load(const char* path, int device, glia_rf** out) {
  if (!path || !out) return GLIA_E_ARG;
  FILE* f = fopen(path, "rb");
  if (!f) return GLIA_E_IO;
  unsigned char hdr[520];
  if (fread(hdr, 520, 1, f) != 1) { fclose(f); return GLIA_E_IO; }
  for (int i = 0; i < 4; ++i) {  // categorical-feature vectors must be empty (begin == end)
    if (memcmp(hdr + 24 * i, hdr + 24 * i + 8, 8) != 0) { fclose(f); return GLIA_E_IO; }
  }
  auto nel = [&](size_t off) { ArrDim d = dim_at(hdr, off); return (long long)d.rows * d.cols; };
  std::vector<int> ncat, catf, tm, ns, nc, bv, nb, ol, nl;
  std::vector<double> xb, cw, co;
  int nrnodes = 0, ntree = 0, mtry = 0, nclass = 0;
  bool ok = true;
  ok = ok && read_array(f, ncat, nel(OFF_NCAT));
  ok = ok && read_array(f, catf, nel(OFF_CATF));
  ok = ok && fread(&nrnodes, 4, 1, f) == 1 && fread(&ntree, 4, 1, f) == 1;
  ok = ok && read_array(f, xb, nel(OFF_XBEST));
  ok = ok && read_array(f, cw, nel(OFF_CLASSWT));
  ok = ok && read_array(f, co, nel(OFF_CUTOFF));
  ok = ok && read_array(f, tm, nel(OFF_TREEMAP));
  ok = ok && read_array(f, ns, nel(OFF_NODESTATUS));
  ok = ok && read_array(f, nc, nel(OFF_NODECLASS));
  ok = ok && read_array(f, bv, nel(OFF_BESTVAR));
  ok = ok && read_array(f, nb, nel(OFF_NDBIG));
  ok = ok && fread(&mtry, 4, 1, f) == 1;
  ok = ok && read_array(f, ol, nel(OFF_ORIG));
  ok = ok && read_array(f, nl, nel(OFF_NEWL));
  ok = ok && fread(&nclass, 4, 1, f) == 1;
  fclose(f);  // the remaining arrays are out-of-bag statistics, unused for scoring
  for (int v : catf) if (v) ok = false;
  long long n = (long long)nrnodes * ntree;
  if (!ok || ntree <= 0 || nrnodes <= 0 || nclass <= 0 || (long long)xb.size() != n ||
      (long long)tm.size() != 2 * n || (long long)ns.size() != n || (long long)nc.size() != n ||
      (long long)bv.size() != n || (long long)ol.size() < nclass)
    return GLIA_E_IO;
  ArrDim dx = dim_at(hdr, OFF_XBEST), dt = dim_at(hdr, OFF_TREEMAP), ds = dim_at(hdr, OFF_NODESTATUS),
         dc = dim_at(hdr, OFF_NODECLASS), db = dim_at(hdr, OFF_BESTVAR);
  std::vector<double> xbT = transposed(xb, dx.rows, dx.cols);
  std::vector<int> tmT = transposed(tm, dt.rows, dt.cols), nsT = transposed(ns, ds.rows, ds.cols),
                   ncT = transposed(nc, dc.rows, dc.cols), bvT = transposed(bv, db.rows, db.cols);
  return glia_b200_rf_from_arrays(ntree, nrnodes, nclass, 0, tmT.data(), nsT.data(), xbT.data(),
                                  bvT.data(), ncT.data(), nb.empty() ? nullptr : nb.data(),
                                  ol.data(), device, out);
}

extern "C" void glia_b200_rf_destroy(glia_rf* m) {
  if (!m) return;
  if (m->d_nodes) cudaFree(m->d_nodes);
  if (m->d_treeOff) cudaFree(m->d_treeOff);
  if (m->d_nodesW) cudaFree(m->d_nodesW);
  if (m->d_blocks) cudaFree(m->d_blocks);
  if (m->d_blockThr) cudaFree(m->d_blockThr);
  if (m->d_blockRoot) cudaFree(m->d_blockRoot);
  delete m;
}

extern "C" int glia_b200_rf_info(const glia_rf* m, int* ntree, int* nrnodes, int* nclass, int* mdim) {
  if (!m) return GLIA_E_ARG;
  if (ntree) *ntree = m->ntree;
  if (nrnodes) *nrnodes = m->nrnodes;
  if (nclass) *nclass = m->nclass;
  if (mdim) *mdim = m->mdim;
  return GLIA_OK;
}

int glia_rf_label_index(const glia_rf* m, int label) {
  for (int i = 0; i < m->nclass; ++i) if (m->orig_labels[i] == label) return i;
  return -1;
}

// device-side scoring of N rows already on the device
int glia_rf_score_device(const glia_rf* m, const double* dX, size_t N, int D, int labelIdx,
                         int* dVotes, double* dProb, cudaStream_t st) {
  if (!N) return GLIA_OK;
  unsigned blocks = (unsigned)((N + 127) / 128);
  if (dVotes)
    rf_score_kernel<true><<<blocks, 128, 128 * m->nclass * sizeof(int), st>>>(
        (const RfNode*)m->d_nodes, m->d_treeOff, m->dist, m->nclass, labelIdx, dX, N, D, dVotes, dProb);
  else {
    const int ldRow = D | 1;  // odd stride (in doubles): rows start on different banks
    const size_t smem = sizeof(double) * (size_t)RF_CPB * ldRow;
    const unsigned nb = (unsigned)((N + RF_CPB - 1) / RF_CPB);
    if (smem > 200 * 1024) {
      rf_score_kernel<false><<<blocks, 128, 0, st>>>((const RfNode*)m->d_nodes, m->d_treeOff, m->dist,
                                                     m->nclass, labelIdx, dX, N, D, nullptr, dProb);
    } else if (m->d_blocks) {
      // wide rows leave room for only a few CTAs of staged rows per SM: spread a candidate's trees over 8 warps
      // instead of 4 so that the SM still holds enough warps (profiles/r1_rf_v4_notes.txt)
      const bool wide = smem > 28 * 1024;
      const size_t smemF = sizeof(float) * (size_t)RF_CPB * ldRow;  // float rows
      const uint4* bl = (const uint4*)m->d_blocks;
#define RF_LAUNCH(ENS_, NS_)                                                                                    \
      { cudaFuncSetAttribute(rf_score_blk_kernel<ENS_, NS_>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024); \
        rf_score_blk_kernel<ENS_, NS_><<<nb, 64 * NS_, smemF, st>>>(bl, m->d_blockThr, m->d_blockRoot, m->dist, labelIdx, dX, N, D, ldRow, dProb); }
      if (m->dist.n > 1) { if (wide) RF_LAUNCH(true, 8) else RF_LAUNCH(true, 4) }
      else { if (wide) RF_LAUNCH(false, 8) else RF_LAUNCH(false, 4) }
#undef RF_LAUNCH
    } else {  // more than two classes: one 16-byte node per level
      cudaFuncSetAttribute(rf_score_smem_kernel<false, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
      cudaFuncSetAttribute(rf_score_smem_kernel<true, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
      if (m->dist.n > 1)
        rf_score_smem_kernel<true, 4><<<nb, 256, smem, st>>>((const RfNode*)m->d_nodesW, m->d_treeOff, m->dist, labelIdx, dX, N, D, ldRow, dProb);
      else
        rf_score_smem_kernel<false, 4><<<nb, 256, smem, st>>>((const RfNode*)m->d_nodesW, m->d_treeOff, m->dist, labelIdx, dX, N, D, ldRow, dProb);
    }
  }
  return cudaGetLastError() == cudaSuccess ? GLIA_OK : GLIA_E_CUDA;
}

extern "C" int glia_b200_rf_score(const glia_rf* m, const double* X, size_t N, int D, int label,
                                  int* votes, double* prob, int device_pointers, void* stream) {
  if (!m || !X || D < m->mdim || (!votes && !prob)) return GLIA_E_ARG;
  int li = glia_rf_label_index(m, label);
  if (li < 0) return GLIA_E_ARG;  // ml/rf/rf.hxx:370 perr
  if (cudaSetDevice(m->device) != cudaSuccess) return GLIA_E_CUDA;
  cudaStream_t st = (cudaStream_t)stream;
  if (device_pointers) {
    int s = glia_rf_score_device(m, X, N, D, li, votes, prob, st);
    return s;
  }
  // host buffers: stream the rows through the device in chunks
  const size_t chunk = 1 << 20;
  double* dX = nullptr; int* dV = nullptr; double* dP = nullptr;
  size_t c = N < chunk ? (N ? N : 1) : chunk;
  int status = GLIA_OK;
  if (cudaMalloc(&dX, 8 * c * D) != cudaSuccess || cudaMalloc(&dP, 8 * c) != cudaSuccess ||
      (votes && cudaMalloc(&dV, 4 * c * m->nclass) != cudaSuccess))
    status = GLIA_E_CUDA;
  for (size_t o = 0; o < N && status == GLIA_OK; o += c) {
    size_t n = N - o < c ? N - o : c;
    if (cudaMemcpyAsync(dX, X + o * D, 8 * n * D, cudaMemcpyHostToDevice, st) != cudaSuccess) { status = GLIA_E_CUDA; break; }
    status = glia_rf_score_device(m, dX, n, D, li, votes ? dV : nullptr, dP, st);
    if (status != GLIA_OK) break;
    if (prob && cudaMemcpyAsync(prob + o, dP, 8 * n, cudaMemcpyDeviceToHost, st) != cudaSuccess) status = GLIA_E_CUDA;
    if (votes && cudaMemcpyAsync(votes + o * m->nclass, dV, 4 * n * m->nclass, cudaMemcpyDeviceToHost, st) != cudaSuccess) status = GLIA_E_CUDA;
    if (cudaStreamSynchronize(st) != cudaSuccess) status = GLIA_E_CUDA;
  }
  cudaFree(dX); cudaFree(dV); cudaFree(dP);
  return status;
}
