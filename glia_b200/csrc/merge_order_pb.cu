// merge order from the boundary probability alone, and pre_merge.
//
// Replaces (reference, /root/reference/code):
//   merge_order_pb operation()                       hmt/main_merge_order_pb.cxx:14-40
//   genMergeOrderGreedyUsingPbMean / ...ApproxMedian util/struct_merge.hxx:36-136
//   genMergeOrderGreedy, TBoundaryTable::{init,top,update}  util/struct_merge.hxx:10-33, type/boundary_table.hxx:47-52, :91-167
//   pre_merge operation() + its condition            gadget/main_pre_merge.cxx:7-87
//   transformKeys / transformImage                   util/struct_merge.hxx:188-209, util/image.hxx:246-259
// The reference reads the image once - the pixels of every mutual pair boundary - and then only combines item data:
// (mean, n) by a weighted mean, or the value lists by concatenation + nth_element.  Here the voxel pass has already
// reduced every directed boundary to (n, sum) and, for the median variant, to a sorted value segment (median.cu), so
// the initial table is read off the graph; the chain itself is the reference's ordered table + priority queue on the
// host (an item update is O(1) for the mean), and the median of a concatenation is a selection over the base segments
// behind it on the device (one launch per merge).  pre_merge's renamed label image is one LUT gather over the volume.
#include <algorithm>
#include <map>
#include <set>
#include <vector>
#include "featurize.cuh"
#include "median.cuh"

namespace glia {

// out[p] = final key of the region of labels[p]; masked-out voxels keep their label (they are in no region).
__global__ void relabel_kernel(const uint32_t* __restrict__ labels, const uint32_t* __restrict__ mask,
                               const uint32_t* __restrict__ rank, const uint32_t* __restrict__ lut, size_t n,
                               uint32_t labelBound, uint32_t* __restrict__ out) {
  size_t p = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (p >= n) return;
  const uint32_t l = labels[p];
  out[p] = ((mask && mask[p] == 0) || l >= labelBound) ? l : lut[rank[l]];
}

}  // namespace glia

using namespace glia;
#define CK(expr) GLIA_CUDA_CHECK(h, expr)

namespace {

typedef std::pair<int, int> Key;
struct PbItem { double sal; unsigned long long seq; double mean; long long n; std::vector<int> segs; };
typedef std::tuple<double, unsigned long long, Key> QKey;

struct PreMerge {  // gadget/main_pre_merge.cxx:26-76
  std::vector<long long> size; std::vector<double> sum, rpb; std::vector<char> hasRpb;
  std::vector<int> thr; double rpbThr;
  bool pass(int k0, int k1) {
    int id[2] = {k0, k1};
    if (size[id[0]] > size[id[1]]) std::swap(id[0], id[1]);
    if (size[id[0]] < thr[0]) return true;
    if (thr.size() > 1)
      for (int k = 0; k < 2; ++k)
        if (size[id[k]] < thr[1]) {
          if (!hasRpb[id[k]]) { rpb[id[k]] = fabs((double)size[id[k]]) >= GLIA_FEPS ? sum[id[k]] / (double)size[id[k]] : 0.0; hasRpb[id[k]] = 1; }
          if (rpb[id[k]] > rpbThr) return true;
        }
    return false;
  }
};

// type 1 = approximate median, 2 = mean.  pm != null: pre_merge's condition and region updates.
int pb_chain(glia_handle* h, int type, PreMerge* pm, uint32_t* order, double* sal, int64_t* nMerges,
             std::vector<int>* parentOut) {
  const int R = (int)h->R, E = (int)h->E, C = (int)h->C;
  cudaStream_t st = h->stream;
  const RecLayout LE = h->vp.edge;
  if (nMerges) *nMerges = 0;
  if (R == 0) return GLIA_OK;
  if (h->prm.n_bimg < 1) { h->err = "merge order from pb needs the pb map as boundary image 0"; return GLIA_E_STATE; }
  if (type == 1 && !h->hasMedian) { h->err = "the median variant needs use_median_feat"; return GLIA_E_STATE; }
  std::vector<uint32_t> lab(R), sAdd((size_t)E * LE.nAdd);
  std::vector<double> sF64((size_t)E * LE.nF64);
  std::vector<int> ea(E), eb(E), rev(E), cand(C);
  auto dl = [&](void* dst, const void* src, size_t bytes) -> cudaError_t {
    return bytes ? cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, st) : cudaSuccess;
  };
  CK(dl(lab.data(), h->regLabel.p, 4 * (size_t)R));
  CK(dl(sAdd.data(), h->sAdd.p, 4 * sAdd.size())); CK(dl(sF64.data(), h->sF64.p, 8 * sF64.size()));
  CK(dl(ea.data(), h->ea.p, 4 * (size_t)E)); CK(dl(eb.data(), h->eb.p, 4 * (size_t)E));
  CK(dl(rev.data(), h->rev.p, 4 * (size_t)E)); CK(dl(cand.data(), h->cand.p, 4 * (size_t)C));
  CK(cudaStreamSynchronize(st));
  const int sumOff = h->fc.bF64[0];
  const int maxIds = 2 * R;
  std::vector<std::set<int>> nbr(maxIds);
  std::map<Key, PbItem> table;
  std::set<QKey> queue;
  unsigned long long seq = 0;
  // ---- init: mutual base pairs in key order; pair boundary = (a -> b) u (b -> a)   (util/struct.hxx:10-16)
  std::vector<int> qKind, qLen, qSeg; std::vector<long long> qOff;
  for (int c = 0; c < C; ++c) {
    const int e = cand[c], rv = rev[e];
    PbItem it;
    it.n = (long long)sAdd[(size_t)e * LE.nAdd] + (long long)sAdd[(size_t)rv * LE.nAdd];
    const double s = sF64[(size_t)e * LE.nF64 + sumOff] + sF64[(size_t)rv * LE.nF64 + sumOff];
    it.mean = fabs((double)it.n) >= GLIA_FEPS ? s / (double)it.n : 0.0;  // struct_merge.hxx:52-55
    it.sal = -it.mean; it.seq = seq++;
    if (type == 1) {
      it.segs = {e, rv};
      qKind.push_back(1 * GLIA_MAX_IMG + 0); qOff.push_back((long long)qSeg.size()); qLen.push_back(2);
      qSeg.push_back(e); qSeg.push_back(rv);
    }
    table.emplace(Key(ea[e], eb[e]), std::move(it));
    nbr[ea[e]].insert(eb[e]); nbr[eb[e]].insert(ea[e]);
  }
  if (type == 1 && C) {
    std::vector<double> med(C);
    int s = glia_median_queries(h, qKind, qOff, qLen, qSeg, med.data(), nullptr);
    if (s != GLIA_OK) return s;
    int c = 0;
    for (auto& kv : table) kv.second.sal = -med[c++];  // table order == candidate order (ascending key)
  }
  for (auto& kv : table) queue.insert(QKey(kv.second.sal, kv.second.seq, kv.first));
  const uint32_t maxLabel = lab[R - 1];
  auto label_of = [&](int id) -> uint32_t { return id < R ? lab[id] : maxLabel + 1 + (uint32_t)(id - R); };
  int M = 0, nextId = R;
  while (!table.empty()) {
    // top(fcond): largest saliency, among equals the most recently inserted; the first one passing the condition
    auto qit = queue.rbegin();
    if (pm) { while (qit != queue.rend() && !pm->pass(std::get<2>(*qit).first, std::get<2>(*qit).second)) ++qit; }
    if (qit == queue.rend()) break;
    const QKey q = *qit;
    const Key k01 = std::get<2>(q);
    const int r0 = k01.first, r1 = k01.second, r2 = nextId++;
    order[3 * (size_t)M] = label_of(r0); order[3 * (size_t)M + 1] = label_of(r1); order[3 * (size_t)M + 2] = label_of(r2);
    sal[M] = std::get<0>(q);
    ++M;
    if (parentOut) { (*parentOut)[r0] = r2; (*parentOut)[r1] = r2; }
    if (pm) { pm->size[r2] = pm->size[r0] + pm->size[r1]; pm->sum[r2] = pm->sum[r0] + pm->sum[r1]; }
    queue.erase(q);
    table.erase(k01);
    nbr[r0].erase(r1); nbr[r1].erase(r0);
    // keys touching r0 or r1, ascending; first encounter of each neighbour (boundary_table.hxx:127-166)
    std::vector<std::pair<Key, int>> keys;
    for (int r : {r0, r1}) for (int W : nbr[r]) keys.push_back({Key(std::min(W, r), std::max(W, r)), W});
    std::sort(keys.begin(), keys.end());
    std::vector<int> Ws;
    { std::set<int> seen; for (auto& kw : keys) if (seen.insert(kw.second).second) Ws.push_back(kw.second); }
    std::vector<PbItem> fresh(Ws.size());
    qKind.clear(); qLen.clear(); qSeg.clear(); qOff.clear();
    for (size_t j = 0; j < Ws.size(); ++j) {
      const int W = Ws[j];
      PbItem& it = fresh[j];
      double acc = 0.0; long long n = 0;  // struct_merge.hxx:62-76: the (rs, r0) item first, then (rs, r1)
      for (int r : {r0, r1}) {
        auto f = table.find(Key(std::min(W, r), std::max(W, r)));
        if (f == table.end()) continue;
        acc += f->second.mean * (double)f->second.n; n += f->second.n;
        if (type == 1) it.segs.insert(it.segs.end(), f->second.segs.begin(), f->second.segs.end());
        queue.erase(QKey(f->second.sal, f->second.seq, f->first));
        table.erase(f);
        nbr[W].erase(r);
      }
      it.n = n; it.mean = fabs((double)n) >= GLIA_FEPS ? acc / (double)n : 0.0;
      it.sal = -it.mean; it.seq = seq++;
      if (type == 1) {
        qKind.push_back(1 * GLIA_MAX_IMG + 0); qOff.push_back((long long)qSeg.size()); qLen.push_back((int)it.segs.size());
        qSeg.insert(qSeg.end(), it.segs.begin(), it.segs.end());
      }
      nbr[W].insert(r2); nbr[r2].insert(W);
    }
    nbr[r0].clear(); nbr[r1].clear();
    if (type == 1 && !Ws.empty()) {
      std::vector<double> med(Ws.size());
      int s = glia_median_queries(h, qKind, qOff, qLen, qSeg, med.data(), nullptr);
      if (s != GLIA_OK) return s;
      for (size_t j = 0; j < Ws.size(); ++j) fresh[j].sal = -med[j];
    }
    for (size_t j = 0; j < Ws.size(); ++j) {
      const Key k(Ws[j], r2);
      queue.insert(QKey(fresh[j].sal, fresh[j].seq, k));
      table.emplace(k, std::move(fresh[j]));
    }
  }
  if (nMerges) *nMerges = M;
  return GLIA_OK;
}

}  // namespace

// labels_out[p] = lut[region index of labels[p]] over the volume the handle was built from (masked-out voxels keep their
// label): one gather, 8 bytes per voxel.  labels_out is a host or device buffer as the handle's inputs are.
int glia_relabel_by_region(glia_handle* h, const std::vector<uint32_t>& lut, uint32_t* labels_out) {
  cudaStream_t st = h->stream;
  const size_t n = h->nvoxBuf, R = lut.size();
  DevBuf &dLut = h->k5[0], &dOut = h->k5[1];
  CK(dLut.ensure(4 * (R + 1)));
  CK(cudaMemcpyAsync(dLut.p, lut.data(), 4 * R, cudaMemcpyHostToDevice, st));
  uint32_t* out = labels_out;
  if (!h->prm.device_pointers) { CK(dOut.ensure(4 * n)); out = dOut.as<uint32_t>(); }
  relabel_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(h->vp.labels, h->vp.mask, h->rank.as<uint32_t>(), dLut.as<uint32_t>(), n,
                                                             h->vp.labelBound, out);
  h->launches++;
  CK(cudaGetLastError());
  if (!h->prm.device_pointers) { int s = glia_copy_to_host(h, labels_out, out, 4 * n); if (s != GLIA_OK) return s; }
  CK(cudaStreamSynchronize(st));
  return GLIA_OK;
}

extern "C" int glia_b200_merge_order_pb(glia_handle* h, int type, uint32_t* order, double* sal, int64_t* n_merges) {
  if (!h || !order || !sal || (type != 1 && type != 2)) return GLIA_E_ARG;
  if (!h->built) return GLIA_E_STATE;
  CK(cudaSetDevice(h->prm.device));
  return pb_chain(h, type, nullptr, order, sal, n_merges, nullptr);
}

extern "C" int glia_b200_pre_merge(glia_handle* h, const int* size_thresholds, int n_thresholds, double rpb_threshold,
                                   uint32_t* order, double* sal, int64_t* n_merges, uint32_t* labels_out) {
  if (!h || !size_thresholds || n_thresholds < 1 || !order || !sal) return GLIA_E_ARG;
  if (!h->built) return GLIA_E_STATE;
  if (h->prm.n_rimg < 1) { h->err = "pre_merge needs the pb map as region image 0"; return GLIA_E_STATE; }
  CK(cudaSetDevice(h->prm.device));
  cudaStream_t st = h->stream;
  const int R = (int)h->R;
  if (n_merges) *n_merges = 0;
  const RecLayout LR = h->vp.reg;
  PreMerge pm;
  pm.thr.assign(size_thresholds, size_thresholds + n_thresholds); pm.rpbThr = rpb_threshold;
  pm.size.assign(2 * (size_t)R + 1, 0); pm.sum.assign(2 * (size_t)R + 1, 0.0); pm.rpb.assign(2 * (size_t)R + 1, 0.0);
  pm.hasRpb.assign(2 * (size_t)R + 1, 0);
  std::vector<uint32_t> cAdd((size_t)R * LR.nAdd);
  std::vector<double> cF64((size_t)R * LR.nF64);
  { int s = glia_ensure_full_regions(h); if (s != GLIA_OK) return s; }
  if (R) {
    CK(cudaMemcpyAsync(cAdd.data(), h->cAdd.p, 4 * cAdd.size(), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(cF64.data(), h->cF64.p, 8 * cF64.size(), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
  }
  for (int r = 0; r < R; ++r) { pm.size[r] = cAdd[(size_t)r * LR.nAdd]; pm.sum[r] = cF64[(size_t)r * LR.nF64 + h->fc.rF64[0]]; }
  std::vector<int> parent(2 * (size_t)R + 1);
  for (size_t i = 0; i < parent.size(); ++i) parent[i] = (int)i;
  int64_t M = 0;
  int s = pb_chain(h, 2, &pm, order, sal, &M, &parent);
  if (s != GLIA_OK) return s;
  if (n_merges) *n_merges = M;
  if (!labels_out || R == 0) return GLIA_OK;
  // transformKeys + transformImage: every base region takes the key of the root of its merges
  std::vector<uint32_t> lab(R), lut(R);
  CK(cudaMemcpyAsync(lab.data(), h->regLabel.p, 4 * (size_t)R, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  const uint32_t maxLabel = lab[R - 1];
  for (int r = 0; r < R; ++r) {
    int v = r;
    while (parent[v] != v) v = parent[v];
    lut[r] = v < R ? lab[v] : maxLabel + 1 + (uint32_t)(v - R);
  }
  return glia_relabel_by_region(h, lut, labels_out);
}
