// K8: greedy merge order with the boundary classifier (the merge-order stage).
//
// Replaces (reference, /root/reference/code):
//   genMergeOrderGreedyUsingBoundaryClassifier     util/struct_merge_bc.hxx:8-58
//   genMergeOrderGreedy                            util/struct_merge.hxx:10-33
//   TBoundaryTable::{init,top,update}              type/boundary_table.hxx:47-52, :91-167
//   TRegionMap::merge / TRegion::merge             type/region_map.hxx:113-118, type/region.hxx:66-75
//   fBcFeat / fBcPred                              hmt/main_merge_order_bc.cxx:54-139
// The loop is a strict sequential chain (each pop depends on the scores of the previous
// merge's new candidates), so the host keeps the TOPOLOGY (which regions touch, the ordered
// table and the priority queue, exactly as the reference does) and combines the mergeable
// records of the merged regions; the device featurises and RF-scores every batch of new
// candidates (one launch per merge).  No pixel is ever re-walked: where the reference's
// fBcFeat costs O(|r0| + |r1| + |r0 u r1|) voxels per candidate, a candidate here costs
// O(deg) record combinations.
#include <algorithm>
#include <map>
#include <set>
#include <queue>
#include <unordered_map>
#include <vector>
#include <chrono>
#include "featurize.cuh"
#include "median.cuh"

int glia_featurize_init_device(glia_handle* h, uint32_t* dKeys, double* dRows);
int glia_rf_label_index(const glia_rf* m, int label);
int glia_rf_score_device(const glia_rf* m, const double* dX, size_t N, int D, int labelIdx,
                         int* dVotes, double* dProb, cudaStream_t st);

namespace glia {

struct __align__(16) RfNode16 { double thr; int left; int feat; };  // == RfNode (rf.cu)

// One CTA per candidate.  Four threads (one per warp) build the four parts of the feature row from the
// six staged records (region X, Y; boundary-of X, Y, Z; pair boundary); the row is then staged in shared memory and every thread walks
// its own trees of the forest: the chain of a merge is latency-bound (m is tens of candidates, a walk is
// ~20 dependent L2 loads), so the trees of one candidate are spread over 256 threads, not 32.
#define MERGE_TPB 256
__global__ void __launch_bounds__(MERGE_TPB)
merge_batch_kernel(const __grid_constant__ FeatCfg F, int m, const uint32_t* __restrict__ stage,
                   int stageWords, const RfNode16* __restrict__ nodes,
                   const int* __restrict__ treeOff, const RfDist dist, int labelIdx,
                   double* __restrict__ rows, double* __restrict__ sal, const double* __restrict__ med, int nMed,
                   double* __restrict__ rowsHost) {
  // stage / sal / med / rowsHost are pinned HOST buffers read and written in place (unified addressing): a merge is
  // one launch and one synchronisation, no separate copies (58 -> 45 us per merge on the chain's critical path);
  // rows is device scratch (the row is re-read for the forest walk).
  extern __shared__ __align__(16) double srow[];
  __shared__ int shits;
  const int w = blockIdx.x;
  if (w >= m) return;
  const RecLayout LR = F.reg, LE = F.edge;
  const int regW = LR.nAdd + LR.nMax + ((LR.nAdd + LR.nMax) & 1) + 2 * LR.nF64;
  const int edgeW = LE.nAdd + LE.nMax + ((LE.nAdd + LE.nMax) & 1) + 2 * LE.nF64;
  // the candidate's six records arrive in ONE coalesced read of the host-mapped staging buffer (word-by-word reads
  // over PCIe by the featurising threads cost more than the copy they replaced)
  uint32_t* sstage = (uint32_t*)(srow + ((F.dim + 1) & ~1));
  for (int i = threadIdx.x; i < stageWords; i += MERGE_TPB) sstage[i] = stage[(size_t)w * stageWords + i];
  const uint32_t* base = sstage;
  double* row = rows + (size_t)w * F.dim;
  if (threadIdx.x == 0) shits = 0;
  __syncthreads();
  if ((threadIdx.x & 31) == 0 && threadIdx.x < 128) {  // four warps, one part of the row each
    auto regRec = [&](int k) {
      const uint32_t* p = base + (size_t)k * regW;
      return RecSingle{p, p + LR.nAdd, (const double*)(p + LR.nAdd + LR.nMax + ((LR.nAdd + LR.nMax) & 1))};
    };
    auto edgeRec = [&](int k) {
      const uint32_t* p = base + 2 * (size_t)regW + (size_t)k * edgeW;
      return RecSingle{p, p + LE.nAdd, (const double*)(p + LE.nAdd + LE.nMax + ((LE.nAdd + LE.nMax) & 1))};
    };
    RecSingle rx = regRec(0), ry = regRec(1);
    RecSum2 rz{rx, ry};
    featurize_row(F, rx, ry, rz, edgeRec(0), edgeRec(1), edgeRec(2), edgeRec(3), nullptr, row,
                  F.simple ? (threadIdx.x == 0 ? -1 : 4) : (int)(threadIdx.x >> 5), false,
                  med ? med + (size_t)w * nMed : nullptr);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < F.dim; i += MERGE_TPB) { const double v = row[i]; srow[i] = v; if (rowsHost) rowsHost[(size_t)w * F.dim + i] = v; }
  __syncthreads();
  int hits = 0;
  const int mdl = dist.pick(srow), tb = dist.begin(mdl), ntree = dist.end(mdl) - tb;  // alg/rf.hxx:97-98
  for (int t = tb + (int)threadIdx.x; t < tb + ntree; t += MERGE_TPB) {
    int k = treeOff[t];
    RfNode16 nd = nodes[k];
    while (nd.feat >= 0) {
      k = nd.left + (srow[nd.feat] <= nd.thr ? 0 : 1);
      nd = nodes[k];
    }
    hits += (nd.left == labelIdx);
  }
  for (int o = 16; o; o >>= 1) hits += __shfl_xor_sync(0xffffffffu, hits, o);
  if ((threadIdx.x & 31) == 0 && hits) atomicAdd(&shits, hits);
  __syncthreads();
  if (threadIdx.x == 0) sal[w] = (double)shits / (double)ntree;
}

}  // namespace glia

using namespace glia;
#define CK(expr) GLIA_CUDA_CHECK(h, expr)

namespace {

// host-side mergeable record (common.cuh word kinds)
struct Rec {
  std::vector<uint32_t> add, mx;
  std::vector<double> f;
  void init(const RecLayout& L) { add.assign(L.nAdd, 0); mx.assign(L.nMax, 0); f.assign(L.nF64, 0.0); }
  void merge(const Rec& o) {
    for (size_t i = 0; i < add.size(); ++i) add[i] += o.add[i];
    for (size_t i = 0; i < mx.size(); ++i) mx[i] = std::max(mx[i], o.mx[i]);
    for (size_t i = 0; i < f.size(); ++i) f[i] += o.f[i];
  }
};

struct Item { double sal; unsigned long long seq; std::vector<double> row; };
typedef std::pair<int, int> Key;

struct Greedy {
  glia_handle* h;
  RecLayout LR, LE;
  int R, E;
  std::vector<int> ea, eb, rev, row;
  std::vector<uint32_t> sAdd, sMax; std::vector<double> sF64;   // base directed boundaries
  std::vector<Rec> reg;                                          // region record per id
  std::vector<Rec> btot;                                         // boundary-of totals (add, f64 only)
  std::vector<std::vector<int>> nmOut;                           // non-mutual boundaries by source region
  std::vector<std::unordered_map<int, Rec>> dm;                  // dm[x][w] = mutual boundaries x -> w
  // per region and max word: the largest value over its boundary entries, the neighbour owning it (-1: a non-mutual
  // entry, -2: none) and the largest over all OTHER entries.  Kept up to date across merges instead of re-scanned
  // for every new candidate (the scans were most of the staging time, profiles/r2_merge_order_wall.txt).
  struct Top2 { std::vector<uint32_t> b1, b2; std::vector<int> o1; bool valid = false; };
  std::vector<Top2> t2;
  // median features only (median.cu): the base sets behind the records
  std::vector<std::map<int, std::vector<int>>> de;               // de[x][w] = base directed boundaries of dm[x][w]
  std::vector<std::vector<int>> mem;                             // base regions of a region id
  std::vector<double> medOf;                                     // [id][nR + nB]: region / boundary-of medians
  std::vector<int> ufParent, idOfRoot, repBase;  // union-find over base regions

  void ufUnion(int r0, int r1, int r2) {
    int a = find(repBase[r0]), b = find(repBase[r1]);
    ufParent[b] = a; idOfRoot[a] = r2; repBase[r2] = a;
  }

  int find(int b) { while (ufParent[b] != b) { ufParent[b] = ufParent[ufParent[b]]; b = ufParent[b]; } return b; }
  int cur(int base) { return idOfRoot[find(base)]; }
  Rec edgeRec(int e) const {
    Rec r; r.add.assign(sAdd.begin() + (size_t)e * LE.nAdd, sAdd.begin() + (size_t)(e + 1) * LE.nAdd);
    r.mx.assign(sMax.begin() + (size_t)e * LE.nMax, sMax.begin() + (size_t)(e + 1) * LE.nMax);
    r.f.assign(sF64.begin() + (size_t)e * LE.nF64, sF64.begin() + (size_t)(e + 1) * LE.nF64);
    return r;
  }
  // does base region b still own a live entry when it sits inside current region `inside`?
  bool active(int b, int inside) {
    for (int e = row[b]; e < row[b + 1]; ++e)
      if (rev[e] < 0 || cur(eb[e]) != inside) return true;
    return false;
  }
};

void put_rec(uint32_t*& p, const Rec& r) {
  size_t na = r.add.size(), nm = r.mx.size();
  if (na) memcpy(p, r.add.data(), 4 * na);
  if (nm) memcpy(p + na, r.mx.data(), 4 * nm);
  size_t off = na + nm + ((na + nm) & 1);
  if (!r.f.empty()) memcpy(p + off, r.f.data(), 8 * r.f.size());
  p += off + 2 * r.f.size();
}

}  // namespace

extern "C" int glia_b200_merge_order_bc(glia_handle* h, const glia_rf* rf, uint32_t* order,
                                        double* sal, double* rows, int64_t* n_merges,
                                        int64_t* n_featurized) {
  if (!h || !rf || !order || !sal) return GLIA_E_ARG;
  if (!h->built) return GLIA_E_STATE;
  // order / sal / rows are always HOST buffers (the chain itself runs on the host)
  if (rf->device != h->prm.device) { h->err = "forest lives on another device"; return GLIA_E_ARG; }
  if (rf->mdim > h->fc.dim) { h->err = "forest uses more features than the row has"; return GLIA_E_ARG; }
  const int li = glia_rf_label_index(rf, -1);  // BC_LABEL_MERGE, hmt/bc_label.hxx:11
  if (li < 0) { h->err = "forest has no class -1"; return GLIA_E_ARG; }
  CK(cudaSetDevice(h->prm.device));
  cudaStream_t st = h->stream;
  const int R = (int)h->R, E = (int)h->E, C = (int)h->C, dim = h->fc.dim;
  if (n_merges) *n_merges = 0;
  if (n_featurized) *n_featurized = 0;
  if (R == 0) return GLIA_OK;
  Greedy G;
  G.h = h; G.LR = h->vp.reg; G.LE = h->vp.edge; G.R = R; G.E = E;
  const RecLayout LR = G.LR, LE = G.LE;
  // ---- download the finalised graph
  std::vector<uint32_t> lab(R), cAdd((size_t)R * LR.nAdd), cMax((size_t)R * LR.nMax), aAdd((size_t)R * LE.nAdd);
  std::vector<double> cF64((size_t)R * LR.nF64), aF64((size_t)R * LE.nF64);
  std::vector<int> cand(C);
  G.ea.resize(E); G.eb.resize(E); G.rev.resize(E); G.row.resize(R + 1);
  G.sAdd.resize((size_t)E * LE.nAdd); G.sMax.resize((size_t)E * LE.nMax); G.sF64.resize((size_t)E * LE.nF64);
  auto dl = [&](void* dst, const void* src, size_t bytes) -> cudaError_t {
    return bytes ? cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, st) : cudaSuccess;
  };
  { int s = glia_ensure_full_regions(h); if (s != GLIA_OK) return s; }
  CK(dl(lab.data(), h->regLabel.p, 4 * (size_t)R));
  CK(dl(cAdd.data(), h->cAdd.p, 4 * cAdd.size())); CK(dl(cMax.data(), h->cMax.p, 4 * cMax.size()));
  CK(dl(cF64.data(), h->cF64.p, 8 * cF64.size()));
  CK(dl(aAdd.data(), h->bAggAdd.p, 4 * aAdd.size())); CK(dl(aF64.data(), h->bAggF64.p, 8 * aF64.size()));
  CK(dl(cand.data(), h->cand.p, 4 * (size_t)C));
  CK(dl(G.ea.data(), h->ea.p, 4 * (size_t)E)); CK(dl(G.eb.data(), h->eb.p, 4 * (size_t)E));
  CK(dl(G.rev.data(), h->rev.p, 4 * (size_t)E)); CK(dl(G.row.data(), h->row.p, 4 * (size_t)(R + 1)));
  CK(dl(G.sAdd.data(), h->sAdd.p, 4 * G.sAdd.size())); CK(dl(G.sMax.data(), h->sMax.p, 4 * G.sMax.size()));
  CK(dl(G.sF64.data(), h->sF64.p, 8 * G.sF64.size()));
  // ---- initial candidates: featurise + score in bulk on the device
  std::vector<double> sal0(C), rows0;
  if (C) {
    CK(h->rows.ensure(8 * (size_t)C * dim));
    CK(h->sal.ensure(8 * (size_t)C));
    int s = glia_featurize_init_device(h, nullptr, h->rows.as<double>());
    if (s) return s;
    s = glia_rf_score_device(rf, h->rows.as<double>(), C, dim, li, nullptr, h->sal.as<double>(), st);
    h->launches++;
    if (s) return s;
    CK(dl(sal0.data(), h->sal.p, 8 * (size_t)C));
    if (rows) { rows0.resize((size_t)C * dim); CK(dl(rows0.data(), h->rows.p, 8 * rows0.size())); }
  }
  CK(cudaStreamSynchronize(st));
  // ---- host state
  const int maxIds = 2 * R;
  G.reg.resize(maxIds); G.btot.resize(maxIds); G.nmOut.resize(maxIds); G.dm.resize(maxIds); G.t2.resize(maxIds);
  G.ufParent.resize(R); G.idOfRoot.resize(R); G.repBase.assign(maxIds, 0);
  for (int a = 0; a < R; ++a) {
    G.ufParent[a] = a; G.idOfRoot[a] = a; G.repBase[a] = a;
    Rec& r = G.reg[a]; r.init(LR);
    std::copy(cAdd.begin() + (size_t)a * LR.nAdd, cAdd.begin() + (size_t)(a + 1) * LR.nAdd, r.add.begin());
    std::copy(cMax.begin() + (size_t)a * LR.nMax, cMax.begin() + (size_t)(a + 1) * LR.nMax, r.mx.begin());
    std::copy(cF64.begin() + (size_t)a * LR.nF64, cF64.begin() + (size_t)(a + 1) * LR.nF64, r.f.begin());
    Rec& b = G.btot[a]; b.init(LE); b.mx.clear();
    std::copy(aAdd.begin() + (size_t)a * LE.nAdd, aAdd.begin() + (size_t)(a + 1) * LE.nAdd, b.add.begin());
    std::copy(aF64.begin() + (size_t)a * LE.nF64, aF64.begin() + (size_t)(a + 1) * LE.nF64, b.f.begin());
    for (int e = G.row[a]; e < G.row[a + 1]; ++e) {
      if (G.rev[e] < 0) G.nmOut[a].push_back(e);
      else G.dm[a].emplace(G.eb[e], G.edgeRec(e));
    }
  }
  const int nMed = h->fc.nMed(), nRi = h->prm.n_rimg, nBi = h->prm.n_bimg, nMedReg = nRi + nBi;
  if (nMed) {  // GLIA_USE_MEDIAN_AS_FEATS: keep the base sets behind every record (median.cu)
    G.de.resize(maxIds); G.mem.resize(maxIds); G.medOf.assign((size_t)maxIds * nMedReg, 0.0);
    if (C) {  // medians of the base regions were computed for the initial candidates
      CK(dl(G.medOf.data(), h->medReg.p, 8 * (size_t)R * nMedReg));
      CK(cudaStreamSynchronize(st));
    } else {
      int s = glia_median_init_candidates(h);
      if (s) return s;
      CK(dl(G.medOf.data(), h->medReg.p, 8 * (size_t)R * nMedReg));
      CK(cudaStreamSynchronize(st));
    }
    for (int a = 0; a < R; ++a) {
      G.mem[a].assign(1, a);
      for (int e = G.row[a]; e < G.row[a + 1]; ++e) if (G.rev[e] >= 0) G.de[a][G.eb[e]].assign(1, e);
    }
  }
  // The reference keeps an ordered table and a multimap queue (type/boundary_table.hxx:19-33).  Nothing here walks the
  // table in key order (the keys touching a merge come from the adjacency), so it is a hash map; the queue is a binary
  // heap with lazy deletion: an entry is live while the table still holds its key with the same insertion sequence.
  auto tkey = [](const Key& k) { return ((unsigned long long)(unsigned)k.first << 32) | (unsigned)k.second; };
  std::unordered_map<unsigned long long, Item> table;
  table.reserve((size_t)C * 2 + 16);
  typedef std::tuple<double, unsigned long long, Key> QKey;  // largest saliency first, among equals the latest insertion
  std::priority_queue<QKey> queue;
  unsigned long long seq = 0;
  for (int c = 0; c < C; ++c) {  // ascending (min,max) key order == candidate order
    int e = cand[c];
    Item it; it.sal = sal0[c]; it.seq = seq++;
    if (rows) it.row.assign(rows0.begin() + (size_t)c * dim, rows0.begin() + (size_t)(c + 1) * dim);
    Key k(G.ea[e], G.eb[e]);
    queue.push(QKey(it.sal, it.seq, k));
    table.emplace(tkey(k), std::move(it));
  }
  rows0.clear(); rows0.shrink_to_fit();
  long long nfeat = C;
  const uint32_t maxLabel = lab[R - 1];
  auto label_of = [&](int id) -> uint32_t { return id < R ? lab[id] : maxLabel + 1 + (uint32_t)(id - R); };
  // staging (pinned) + device scratch for the per-merge batches
  const int regW = LR.nAdd + LR.nMax + ((LR.nAdd + LR.nMax) & 1) + 2 * LR.nF64;
  const int edgeW = LE.nAdd + LE.nMax + ((LE.nAdd + LE.nMax) & 1) + 2 * LE.nF64;
  const int stageWords = 2 * regW + 4 * edgeW;
  size_t capBatch = 0;
  uint32_t* hStage = nullptr; double* hSal = nullptr; double* hRows = nullptr;
  double* dRows = nullptr; double* hMed = nullptr;
  auto release = [&]() {
    if (hStage) cudaFreeHost(hStage); if (hSal) cudaFreeHost(hSal); if (hRows) cudaFreeHost(hRows); if (hMed) cudaFreeHost(hMed);
    if (dRows) cudaFree(dRows);
    hStage = nullptr; hSal = nullptr; hRows = nullptr; hMed = nullptr; dRows = nullptr;
  };
  auto ensure_batch = [&](size_t m) -> cudaError_t {
    if (m <= capBatch) return cudaSuccess;
    release();
    capBatch = std::max<size_t>(2 * m, 256);
    cudaError_t e;
    if ((e = cudaMallocHost((void**)&hStage, 4 * capBatch * stageWords)) != cudaSuccess) return e;
    if ((e = cudaMallocHost((void**)&hSal, 8 * capBatch)) != cudaSuccess) return e;
    if ((e = cudaMallocHost((void**)&hRows, 8 * capBatch * dim)) != cudaSuccess) return e;
    if (nMed && (e = cudaMallocHost((void**)&hMed, 8 * capBatch * nMed)) != cudaSuccess) return e;
    return cudaMalloc((void**)&dRows, 8 * capBatch * dim);
  };
  struct Guard { decltype(release)& r; ~Guard() { r(); } } guard{release};

  int M = 0;
  int nextId = R;
  const bool trace = getenv("GLIA_B200_TRACE") != nullptr;
  double tBook = 0.0, tStage = 0.0, tDev = 0.0, tIns = 0.0;
  auto now = []() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  auto top2 = [&](int X) -> const Greedy::Top2& {  // per max word: best, its owning neighbour, best of the others
    Greedy::Top2& t = G.t2[X];
    if (t.valid) return t;
    t.b1.assign(LE.nMax, 0); t.b2.assign(LE.nMax, 0); t.o1.assign(LE.nMax, -2);
    auto feed = [&](const uint32_t* mx, int own) {
      for (int i = 0; i < LE.nMax; ++i) {
        uint32_t v = mx[i];
        if (v > t.b1[i]) { t.b2[i] = t.b1[i]; t.b1[i] = v; t.o1[i] = own; }
        else if (v > t.b2[i]) t.b2[i] = v;
      }
    };
    for (auto& kv : G.dm[X]) feed(kv.second.mx.data(), kv.first);
    for (int e : G.nmOut[X]) feed(G.sMax.data() + (size_t)e * LE.nMax, -1);
    t.valid = true;
    return t;
  };
  std::vector<int> seenStamp(maxIds, -1);
  while (!table.empty()) {
    const double tA = trace ? now() : 0.0;
    // top: largest saliency, among equals the most recently inserted (boundary_table.hxx:47-52)
    auto it01 = table.end();
    Key k01;
    for (;;) {  // skip heap entries whose table item is gone or was re-created since
      const QKey& q = queue.top();
      k01 = std::get<2>(q);
      it01 = table.find(tkey(k01));
      const bool live = it01 != table.end() && it01->second.seq == std::get<1>(q);
      queue.pop();
      if (live) break;
    }
    const int r0 = k01.first, r1 = k01.second, r2 = nextId++;
    order[3 * (size_t)M] = label_of(r0); order[3 * (size_t)M + 1] = label_of(r1); order[3 * (size_t)M + 2] = label_of(r2);
    sal[M] = it01->second.sal;
    if (rows) memcpy(rows + (size_t)M * dim, it01->second.row.data(), 8 * (size_t)dim);
    ++M;
    table.erase(it01);
    // ---- rmap.merge(r0, r1, r2) on the records
    G.reg[r2] = G.reg[r0]; G.reg[r2].merge(G.reg[r1]);
    Rec& bz = G.btot[r2]; bz = G.btot[r0];
    {
      const Rec& b1 = G.btot[r1];
      const Rec& d01 = G.dm[r0][r1]; const Rec& d10 = G.dm[r1][r0];
      for (int i = 0; i < LE.nAdd; ++i) bz.add[i] = bz.add[i] + b1.add[i] - d01.add[i] - d10.add[i];
      for (int i = 0; i < LE.nF64; ++i) bz.f[i] = bz.f[i] + b1.f[i] - d01.f[i] - d10.f[i];
    }
    G.nmOut[r2] = std::move(G.nmOut[r0]);
    G.nmOut[r2].insert(G.nmOut[r2].end(), G.nmOut[r1].begin(), G.nmOut[r1].end());
    G.nmOut[r1].clear();
    // keys touching r0 or r1, ascending; first encounter of each neighbour (boundary_table.hxx:127-166)
    std::vector<std::pair<Key, int>> keys;
    for (int r : {r0, r1})
      for (auto& kv : G.dm[r]) {
        int W = kv.first;
        if (W == r0 || W == r1) continue;
        keys.push_back({Key(std::min(W, r), std::max(W, r)), W});
      }
    std::sort(keys.begin(), keys.end());
    std::vector<int> Ws;
    for (auto& kw : keys) if (seenStamp[kw.second] != M) { seenStamp[kw.second] = M; Ws.push_back(kw.second); }
    // adjacency of r2 and of its neighbours (the records of r0 / r1 are moved, not copied: their maps are dropped below)
    G.dm[r2].reserve(Ws.size() * 2 + 2);
    for (int W : Ws) {
      Rec out, in; bool haveOut = false, haveIn = false;
      const uint32_t* vin[2] = {nullptr, nullptr};  // max words of W -> r0, W -> r1 before they are merged
      Rec inParts[2];
      int k = 0;
      for (int r : {r0, r1}) {
        auto f = G.dm[r].find(W);
        if (f != G.dm[r].end()) { if (!haveOut) { out = std::move(f->second); haveOut = true; } else out.merge(f->second); }
        auto g = G.dm[W].find(r);
        if (g != G.dm[W].end()) { inParts[k] = std::move(g->second); vin[k] = inParts[k].mx.data(); G.dm[W].erase(g); }
        ++k;
      }
      // W's cached maxima: its entries towards r0 / r1 become ONE entry towards r2 holding their maximum
      Greedy::Top2& tw = G.t2[W];
      if (tw.valid)
        for (int i = 0; i < LE.nMax && tw.valid; ++i) {
          if (tw.o1[i] != r0 && tw.o1[i] != r1) continue;       // the best entry is elsewhere: nothing changes
          const uint32_t* other = tw.o1[i] == r0 ? vin[1] : vin[0];
          const uint32_t vo = other ? other[i] : 0u;
          if (vo != 0 && vo == tw.b2[i]) tw.valid = false;      // the runner-up may have been the other merged entry: rescan
          else tw.o1[i] = r2;
        }
      for (int q = 0; q < 2; ++q)
        if (vin[q]) { if (!haveIn) { in = std::move(inParts[q]); haveIn = true; } else in.merge(inParts[q]); }
      if (!haveOut) out.init(LE);
      if (!haveIn) in.init(LE);
      G.dm[r2].emplace(W, std::move(out));
      G.dm[W].emplace(r2, std::move(in));
      if (nMed) {
        std::vector<int> lo, li;
        for (int r : {r0, r1}) {
          auto f = G.de[r].find(W);
          if (f != G.de[r].end()) lo.insert(lo.end(), f->second.begin(), f->second.end());
          auto g = G.de[W].find(r);
          if (g != G.de[W].end()) { li.insert(li.end(), g->second.begin(), g->second.end()); G.de[W].erase(g); }
        }
        G.de[r2].emplace(W, std::move(lo));
        G.de[W].emplace(r2, std::move(li));
      }
    }
    if (nMed) {
      G.de[r0].clear(); G.de[r1].clear();
      G.mem[r2] = std::move(G.mem[r0]);
      G.mem[r2].insert(G.mem[r2].end(), G.mem[r1].begin(), G.mem[r1].end());
      G.mem[r1].clear(); G.mem[r1].shrink_to_fit();
    }
    G.ufUnion(r0, r1, r2);
    // Clean-up that neither the staging below nor the batch kernel reads - the emptied maps of r0 / r1, the stale table
    // items, the heap rebuild - runs AFTER the launch, while the GPU featurises and scores the new candidates; it only
    // has to be done before the next pop.
    auto finish_merge = [&]() {
      { std::unordered_map<int, Rec>().swap(G.dm[r0]); std::unordered_map<int, Rec>().swap(G.dm[r1]); }
      G.t2[r0] = Greedy::Top2(); G.t2[r1] = Greedy::Top2();
      for (auto& kw : keys) table.erase(tkey(kw.first));  // their heap entries die with them
      if (queue.size() > 4 * table.size() + 4096) {       // mostly dead entries: rebuild the heap from the table
        std::vector<QKey> live;
        live.reserve(table.size());
        for (auto& kv : table) live.push_back(QKey(kv.second.sal, kv.second.seq, Key((int)(kv.first >> 32), (int)(unsigned)kv.first)));
        queue = std::priority_queue<QKey>(std::less<QKey>(), std::move(live));
      }
    };
    const size_t m = Ws.size();
    if (!m) { finish_merge(); continue; }
    {
      cudaError_t e = ensure_batch(m);
      if (e != cudaSuccess) { h->err = std::string("merge_order_bc batch buffers: ") + cudaGetErrorString(e); return GLIA_E_CUDA; }
    }
    const double tB = trace ? now() : 0.0;
    // ---- stage the six records of every new candidate (X = W, Y = r2)
    std::unordered_map<int, std::vector<int>> nmByTarget;
    for (int e : G.nmOut[r2]) { int t = G.cur(G.eb[e]); if (t != r2) nmByTarget[t].push_back(e); }
    const Greedy::Top2& ty = top2(r2);
    const std::vector<uint32_t>&yb1 = ty.b1, &yb2 = ty.b2; const std::vector<int>& yo1 = ty.o1;
    uint32_t* p = hStage;
    // records are written straight into the staging buffer; the two scratch records below are reused for
    // every candidate (no allocation per candidate)
    auto put_parts = [&](const uint32_t* add, size_t na, const uint32_t* mx, size_t nm, const double* f, size_t nf) {
      if (na) memcpy(p, add, 4 * na);
      if (nm) memcpy(p + na, mx, 4 * nm);
      const size_t off = na + nm + ((na + nm) & 1);
      if (nf) memcpy(p + off, f, 8 * nf);
      p += off + 2 * nf;
    };
    Rec bzz, pb;
    bzz.init(LE); pb.init(LE);
    // median queries of this batch: r2's own sets once, then per candidate the merged region, its entries and the
    // pair boundary; each set is a list of base segments (median.cu)
    std::vector<int> qKind, qLen, qSeg; std::vector<long long> qOff;
    auto add_q = [&](int store, int nImg, long long off) {
      for (int i = 0; i < nImg; ++i) { qKind.push_back(store * GLIA_MAX_IMG + i); qOff.push_back(off); qLen.push_back((int)((long long)qSeg.size() - off)); }
    };
    if (nMed) {
      long long off = 0;
      qSeg.insert(qSeg.end(), G.mem[r2].begin(), G.mem[r2].end());
      add_q(0, nRi, off);
      off = (long long)qSeg.size();
      for (auto& kv : G.de[r2]) qSeg.insert(qSeg.end(), kv.second.begin(), kv.second.end());
      qSeg.insert(qSeg.end(), G.nmOut[r2].begin(), G.nmOut[r2].end());
      add_q(1, nBi, off);
    }
    for (int W : Ws) {
      const Rec& dWY = G.dm[W][r2]; const Rec& dYW = G.dm[r2][W];
      put_rec(p, G.reg[W]);
      put_rec(p, G.reg[r2]);
      const Greedy::Top2& tw = top2(W);
      const std::vector<uint32_t>&best1 = tw.b1, &best2 = tw.b2; const std::vector<int>& owner1 = tw.o1;
      const Rec& bx = G.btot[W];   // boundary-of totals: add + f64; the max words are best1 / yb1
      const Rec& by = G.btot[r2];
      for (int i = 0; i < LE.nAdd; ++i) bzz.add[i] = bx.add[i] + by.add[i] - dWY.add[i] - dYW.add[i];
      for (int i = 0; i < LE.nF64; ++i) bzz.f[i] = bx.f[i] + by.f[i] - dWY.f[i] - dYW.f[i];
      for (int i = 0; i < LE.nMax; ++i)
        bzz.mx[i] = std::max(owner1[i] == r2 ? best2[i] : best1[i], yo1[i] == W ? yb2[i] : yb1[i]);
      pb.add = dWY.add; pb.mx = dWY.mx; pb.f = dWY.f;  // pair boundary (util/struct.hxx:10-16)
      pb.merge(dYW);
      auto merge_base = [&](int e) {
        const uint32_t* a = G.sAdd.data() + (size_t)e * LE.nAdd;
        const uint32_t* m = G.sMax.data() + (size_t)e * LE.nMax;
        const double* f = G.sF64.data() + (size_t)e * LE.nF64;
        for (int i = 0; i < LE.nAdd; ++i) pb.add[i] += a[i];
        for (int i = 0; i < LE.nMax; ++i) pb.mx[i] = std::max(pb.mx[i], m[i]);
        for (int i = 0; i < LE.nF64; ++i) pb.f[i] += f[i];
      };
      std::vector<int> pairNM;  // non-mutual boundaries counted into the pair boundary
      for (int e : G.nmOut[W]) if (G.cur(G.eb[e]) == r2 && G.active(G.eb[e], r2)) { merge_base(e); if (nMed) pairNM.push_back(e); }
      auto nt = nmByTarget.find(W);
      if (nt != nmByTarget.end()) for (int e : nt->second) if (G.active(G.eb[e], W)) { merge_base(e); if (nMed) pairNM.push_back(e); }
      if (nMed) {
        long long off = (long long)qSeg.size();  // region of Z = W u r2
        qSeg.insert(qSeg.end(), G.mem[W].begin(), G.mem[W].end());
        qSeg.insert(qSeg.end(), G.mem[r2].begin(), G.mem[r2].end());
        add_q(0, nRi, off);
        off = (long long)qSeg.size();            // entries of Z: those of W and r2 except the mutual pair (region.hxx:66-75)
        for (auto& kv : G.de[W]) if (kv.first != r2) qSeg.insert(qSeg.end(), kv.second.begin(), kv.second.end());
        for (auto& kv : G.de[r2]) if (kv.first != W) qSeg.insert(qSeg.end(), kv.second.begin(), kv.second.end());
        qSeg.insert(qSeg.end(), G.nmOut[W].begin(), G.nmOut[W].end());
        qSeg.insert(qSeg.end(), G.nmOut[r2].begin(), G.nmOut[r2].end());
        add_q(1, nBi, off);
        off = (long long)qSeg.size();            // pair boundary
        { auto f = G.de[W].find(r2); if (f != G.de[W].end()) qSeg.insert(qSeg.end(), f->second.begin(), f->second.end()); }
        { auto f = G.de[r2].find(W); if (f != G.de[r2].end()) qSeg.insert(qSeg.end(), f->second.begin(), f->second.end()); }
        qSeg.insert(qSeg.end(), pairNM.begin(), pairNM.end());
        add_q(1, nBi, off);
      }
      put_parts(bx.add.data(), bx.add.size(), best1.data(), best1.size(), bx.f.data(), bx.f.size());
      put_parts(by.add.data(), by.add.size(), yb1.data(), yb1.size(), by.f.data(), by.f.size());
      put_rec(p, bzz); put_rec(p, pb);
    }
    if (nMed) {
      std::vector<double> res(qKind.size());
      int s = glia_median_queries(h, qKind, qOff, qLen, qSeg, res.data(), nullptr);
      if (s) return s;
      double* my = &G.medOf[(size_t)r2 * nMedReg];
      for (int i = 0; i < nMedReg; ++i) my[i] = res[i];
      std::vector<double> med(m * nMed);
      for (size_t j = 0; j < m; ++j) {
        const double* mw = &G.medOf[(size_t)Ws[j] * nMedReg];
        const double* rz = &res[nMedReg + j * (nRi + 2 * nBi)];
        double* o = &med[j * nMed];
        for (int i = 0; i < nRi; ++i) { o[3 * i] = mw[i]; o[3 * i + 1] = my[i]; o[3 * i + 2] = rz[i]; }
        for (int i = 0; i < nBi; ++i) {
          double* q = o + 3 * nRi + 4 * i;
          q[0] = mw[nRi + i]; q[1] = my[nRi + i]; q[2] = rz[nRi + i]; q[3] = rz[nRi + nBi + i];
        }
      }
      memcpy(hMed, med.data(), 8 * med.size());
    }
    const double tC = trace ? now() : 0.0;
    merge_batch_kernel<<<(unsigned)m, MERGE_TPB, 8 * (size_t)((dim + 1) & ~1) + 4 * (size_t)stageWords, st>>>(
        h->fc, (int)m, hStage, stageWords, (const RfNode16*)rf->d_nodes, rf->d_treeOff, rf->dist, li, dRows, hSal,
        nMed ? hMed : nullptr, nMed, rows ? hRows : nullptr);
    h->launches++;
    finish_merge();  // overlapped with the kernel
    CK(cudaStreamSynchronize(st));
    const double tD = trace ? now() : 0.0;
    nfeat += (long long)m;
    for (size_t j = 0; j < m; ++j) {
      Item it; it.sal = hSal[j]; it.seq = seq++;
      if (rows) it.row.assign(hRows + j * dim, hRows + (j + 1) * dim);
      Key k(Ws[j], r2);
      queue.push(QKey(it.sal, it.seq, k));
      table.emplace(tkey(k), std::move(it));
    }
    if (trace) { const double tE = now(); tBook += tB - tA; tStage += tC - tB; tDev += tD - tC; tIns += tE - tD; }
  }
  if (trace)
    fprintf(stderr, "[glia_b200] merge_order_bc: %d merges, %lld candidates | bookkeeping %.3f s, staging %.3f s, device round trips %.3f s, queue inserts %.3f s\n",
            M, nfeat, tBook, tStage, tDev, tIns);
  if (n_merges) *n_merges = M;
  if (n_featurized) *n_featurized = nfeat;
  return GLIA_OK;
}
