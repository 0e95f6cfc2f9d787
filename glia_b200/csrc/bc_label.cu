// boundary-classifier training labels from a ground-truth segmentation.
//
// Replaces (reference, /root/reference/code):
//   bc_label_ri operation()                              hmt/main_bc_label_ri.cxx:31-155
//   genBoundaryClassificationLabelF1 / ...RI             hmt/bc_label.hxx:44-122
//   stats::pairStats over regions (truth BG excluded)    util/image_stats.hxx:173-193, util/stats.hxx:186-230
//   stats::pairF1, precision, recall, f1, randIndex      util/image_stats.hxx:222-245, util/stats.hxx:232-265
//   genTree                                              hmt/tree_build.hxx:12-38
// For every merge the reference re-walks the pixels of r0, r1 and r0 u r1, builds a (region, truth label) contingency
// map of 512-bit integers and counts pixel pairs.  Here the volume is read ONCE: one 64-bit key
// (region index << 32 | truth label) per unmasked, non-background voxel, one radix sort, one run-length encode - the
// contingency table of the BASE regions (the same sort-by-key machinery as the median stores).  Everything the label
// rules need of a tree node v then follows from two integers,
//     n_v = #voxels of v with a truth label,        T_v = sum over truth labels t of C(c_vt, 2),
// because for a set S of disjoint nodes whose union is a node u (the two children, or any "best split" of u)
//     TP = sum_S T_v,   pairs sharing a region = sum_S C(n_v, 2),   pairs sharing a truth label = T_u,   all pairs = C(n_u, 2)
// (util/stats.hxx:194-229 with keys0 = S).  T_u = T_x + T_y + sum_t c_xt c_yt is accumulated while the children's
// sparse count vectors are merged small-into-large up the tree on the host (O(total log) map updates).
#include <algorithm>
#include <unordered_map>
#include <vector>
#include <cub/cub.cuh>
#include "handle.cuh"

namespace glia {

__global__ void bcl_keys_kernel(const uint32_t* __restrict__ labels, const uint32_t* __restrict__ mask,
                                const uint32_t* __restrict__ truth, const uint32_t* __restrict__ rank, size_t n,
                                uint32_t labelBound, unsigned long long* __restrict__ keys) {
  size_t p = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (p >= n) return;
  const uint32_t l = labels[p], t = truth[p];
  // masked-out voxels are in no region (TRegionMap(seg, mask, ...)); truth BG_VAL = 0 is excluded from every count
  keys[p] = ((mask && mask[p] == 0) || t == 0 || l >= labelBound) ? ~0ull : (((unsigned long long)rank[l] << 32) | t);
}

__global__ void bcl_count_heads_kernel(const unsigned long long* __restrict__ keys, size_t n, unsigned long long* count) {
  size_t p = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const bool head = p < n && (p == 0 || keys[p] != keys[p - 1]);
  const int c = __syncthreads_count(head);
  if (threadIdx.x == 0 && c) atomicAdd(count, (unsigned long long)c);
}

}  // namespace glia

using namespace glia;
#define CK(expr) GLIA_CUDA_CHECK(h, expr)

namespace {

typedef unsigned __int128 u128;
const int LABEL_SPLIT = 1, LABEL_MERGE = -1;  // hmt/bc_label.hxx:9-11
const double FEPS_ = 2.22e-16;
inline bool isfeq(double a, double b) { return fabs(a - b) < FEPS_; }  // glia_base.hxx:71-72
inline u128 choose2(u128 n) { return n * (n - 1) / 2; }                 // n == 0 -> 0 * (2^128 - 1) / 2 == 0

struct PairCounts { u128 tp, tn, fp, fn; };
// S: disjoint nodes with union u.  tpS = sum T_v, sameRegion = sum C(n_v, 2).
inline PairCounts pair_counts(u128 tpS, u128 sameRegion, u128 Tu, u128 nu) {
  PairCounts c;
  c.tp = tpS;
  c.tn = choose2(nu) - Tu + tpS - sameRegion;  // util/stats.hxx:226-227
  c.fp = sameRegion - tpS;
  c.fn = Tu - tpS;
  return c;
}
inline double ratio(u128 num, u128 den) {  // util/stats.hxx:243-256
  const double d = (double)den;
  return (double)num / (!isfeq(d, 0.0) ? d : (d + FEPS_));
}
struct F1 { double f, prec, rec; };
inline F1 pair_f1(const PairCounts& c) {
  F1 r;
  r.prec = ratio(c.tp, c.tp + c.fp);
  r.rec = ratio(c.tp, c.tp + c.fn);
  r.f = 2.0 * r.prec * r.rec / (r.prec + r.rec);  // util/stats.hxx:259-261 (0 / 0 = NaN compares false, as there)
  return r;
}
inline double rand_index(const PairCounts& c) {  // util/stats.hxx:232-240
  const double num = (double)(c.tp + c.tn);
  double den = (double)(c.fp + c.fn);
  den += num;
  return num / (!isfeq(den, 0.0) ? den : (den + FEPS_));
}

}  // namespace

extern "C" int glia_b200_bc_label(glia_handle* h, const uint32_t* truth, const uint32_t* order, int64_t n64,
                                  int use_pair_f1, int opt_split, int global_opt, int tweak, double max_prec_drop,
                                  int32_t* bc_labels, double* merge_score, double* split_score) {
  if (!h || !truth || (n64 > 0 && (!order || !bc_labels)) || n64 < 0 || n64 > 0x3fffffff || global_opt < 0 || global_opt > 2)
    return GLIA_E_ARG;
  if (!h->built) return GLIA_E_STATE;
  if (n64 == 0) return GLIA_OK;
  CK(cudaSetDevice(h->prm.device));
  cudaStream_t st = h->stream;
  const int n = (int)n64, R = (int)h->R, N = R + n;
  const size_t nv = h->nvoxBuf;
  // ---- device: contingency table of the base regions
  const uint32_t* dTruth = truth;
  if (!h->prm.device_pointers) {
    CK(h->k5[2].ensure(4 * nv));
    CK(cudaMemcpyAsync(h->k5[2].p, truth, 4 * nv, cudaMemcpyHostToDevice, st));
    dTruth = h->k5[2].as<uint32_t>();
  }
  CK(h->medTmpA.ensure(8 * nv)); CK(h->medTmpB.ensure(8 * nv));
  CK(h->scalars.ensure(64));
  bcl_keys_kernel<<<(unsigned)((nv + 255) / 256), 256, 0, st>>>(h->vp.labels, h->vp.mask, dTruth, h->rank.as<uint32_t>(), nv,
                                                                 h->vp.labelBound, h->medTmpA.as<unsigned long long>());
  cub::DoubleBuffer<unsigned long long> db(h->medTmpA.as<unsigned long long>(), h->medTmpB.as<unsigned long long>());
  size_t tmp = 0;
  CK(cub::DeviceRadixSort::SortKeys(nullptr, tmp, db, (long long)nv, 0, 64, st));
  CK(h->cubTmp.ensure(tmp));
  CK(cub::DeviceRadixSort::SortKeys(h->cubTmp.p, tmp, db, (long long)nv, 0, 64, st));
  unsigned long long* dCount = h->scalars.as<unsigned long long>();
  CK(cudaMemsetAsync(dCount, 0, 8, st));
  bcl_count_heads_kernel<<<(unsigned)((nv + 255) / 256), 256, 0, st>>>(db.Current(), nv, dCount);
  unsigned long long runs = 0;
  CK(cudaMemcpyAsync(&runs, dCount, 8, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  h->launches += 4;
  std::vector<unsigned long long> keys((size_t)runs);
  std::vector<unsigned int> cnt((size_t)runs);
  if (runs) {
    CK(h->k5[3].ensure(8 * (size_t)runs)); CK(h->k5[4].ensure(4 * (size_t)runs + 8));
    int* dRuns = (int*)(h->k5[4].as<unsigned int>() + runs);
    tmp = 0;
    CK(cub::DeviceRunLengthEncode::Encode(nullptr, tmp, db.Current(), h->k5[3].as<unsigned long long>(), h->k5[4].as<unsigned int>(), dRuns, (long long)nv, st));
    CK(h->cubTmp.ensure(tmp));
    CK(cub::DeviceRunLengthEncode::Encode(h->cubTmp.p, tmp, db.Current(), h->k5[3].as<unsigned long long>(), h->k5[4].as<unsigned int>(), dRuns, (long long)nv, st));
    h->launches++;
    CK(cudaMemcpyAsync(keys.data(), h->k5[3].p, 8 * (size_t)runs, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(cnt.data(), h->k5[4].p, 4 * (size_t)runs, cudaMemcpyDeviceToHost, st));
  }
  std::vector<uint32_t> lab(R);
  if (R) CK(cudaMemcpyAsync(lab.data(), h->regLabel.p, 4 * (size_t)R, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  // ---- host: tree (genTree semantics: a key names the node that carries it), n_v and T_v bottom-up
  typedef std::unordered_map<uint32_t, unsigned long long> Sparse;
  std::vector<Sparse*> vec(N, nullptr);
  std::vector<u128> nn(N, 0), T(N, 0);
  std::vector<Sparse> store(R);
  for (size_t i = 0; i < (size_t)runs; ++i) {
    if (keys[i] == ~0ull) continue;  // masked / background voxels
    const int r = (int)(keys[i] >> 32);
    if (r >= R) continue;
    store[r][(uint32_t)keys[i]] += cnt[i];
    nn[r] += cnt[i];
    T[r] += choose2(cnt[i]);
  }
  for (int r = 0; r < R; ++r) vec[r] = &store[r];
  std::unordered_map<uint32_t, int> merged;
  merged.reserve((size_t)n * 2);
  std::vector<int> c0(N, -1), c1(N, -1), parent(N, -1);
  auto resolve = [&](uint32_t key) -> int {
    auto it = merged.find(key);
    if (it != merged.end()) return it->second;
    auto lb = std::lower_bound(lab.begin(), lab.end(), key);
    return (lb != lab.end() && *lb == key) ? (int)(lb - lab.begin()) : -1;
  };
  for (int i = 0; i < n; ++i) {
    const int x = resolve(order[3 * i]), y = resolve(order[3 * i + 1]), z = R + i;
    if (x < 0 || y < 0 || x == y || parent[x] >= 0 || parent[y] >= 0) {
      h->err = "merge order: step " + std::to_string(i) + " references an unknown or already merged region";
      return GLIA_E_ORDER;
    }
    parent[x] = parent[y] = z; c0[z] = x; c1[z] = y;
    merged.emplace(order[3 * i + 2], z);
    Sparse* big = vec[x]; Sparse* small = vec[y];
    if (small->size() > big->size()) std::swap(big, small);
    u128 t = T[x] + T[y];
    for (auto& kv : *small) {  // T_z = T_x + T_y + sum_t c_xt c_yt
      unsigned long long& c = (*big)[kv.first];
      t += (u128)c * kv.second;
      c += kv.second;
    }
    small->clear();
    vec[z] = big; vec[x] = vec[y] = nullptr;
    nn[z] = nn[x] + nn[y];
    T[z] = t;
  }
  auto merge_counts = [&](int z) { return pair_counts(T[z], choose2(nn[z]), T[z], nn[z]); };
  // split set of z given as (sum T, sum C(n, 2)) of its parts
  std::vector<u128> sT(N), sP(N);
  if (global_opt == 0) {
    std::vector<double> mS(n), sS(n);
    for (int i = 0; i < n; ++i) {
      const int z = R + i, x = c0[z], y = c1[z];
      const PairCounts cm = merge_counts(z), cs = pair_counts(T[x] + T[y], choose2(nn[x]) + choose2(nn[y]), T[z], nn[z]);
      if (use_pair_f1) {  // hmt/bc_label.hxx:44-70
        const F1 s = pair_f1(cs), m = pair_f1(cm);
        mS[i] = m.f; sS[i] = s.f;
        if (max_prec_drop < 1.0 && s.prec - m.prec > max_prec_drop) bc_labels[i] = LABEL_SPLIT;
        else if (tweak)
          bc_labels[i] = (m.f > s.f || (s.prec < FEPS_ && s.rec < FEPS_ && m.prec < FEPS_ && m.rec < FEPS_) ||
                          (s.f == m.f && s.prec > 0.9 && m.prec > 0.9)) ? LABEL_MERGE : LABEL_SPLIT;
        else bc_labels[i] = m.f > s.f ? LABEL_MERGE : LABEL_SPLIT;
      } else {            // hmt/bc_label.hxx:92-108
        mS[i] = rand_index(cm); sS[i] = rand_index(cs);
        bc_labels[i] = mS[i] > sS[i] ? LABEL_MERGE : LABEL_SPLIT;
      }
    }
    if (use_pair_f1 && opt_split) {  // main_bc_label_ri.cxx:58-82: compare the merge with the best split found so far
      for (int v = 0; v < R; ++v) { sT[v] = T[v]; sP[v] = choose2(nn[v]); }
      for (int i = 0; i < n; ++i) {
        const int z = R + i, x = c0[z], y = c1[z];
        const u128 st_ = sT[x] + sT[y], sp_ = sP[x] + sP[y];
        bool split = bc_labels[i] == LABEL_SPLIT;
        if (!split) {
          const F1 s = pair_f1(pair_counts(st_, sp_, T[z], nn[z]));
          if (!(mS[i] > s.f)) { bc_labels[i] = LABEL_SPLIT; split = true; }
        }
        if (split) { sT[z] = st_; sP[z] = sp_; } else { sT[z] = T[z]; sP[z] = choose2(nn[z]); }
      }
    }
    if (merge_score) std::copy(mS.begin(), mS.end(), merge_score);
    if (split_score) std::copy(sS.begin(), sS.end(), split_score);
    return GLIA_OK;
  }
  // ---- global optimum (main_bc_label_ri.cxx:84-146): best splits bottom-up, then path consistency
  std::vector<int> lbl(N, LABEL_MERGE);
  for (int v = 0; v < R; ++v) { sT[v] = T[v]; sP[v] = choose2(nn[v]); }
  for (int i = 0; i < n; ++i) {
    const int z = R + i, x = c0[z], y = c1[z];
    const u128 st_ = sT[x] + sT[y], sp_ = sP[x] + sP[y];
    const F1 m = pair_f1(merge_counts(z)), s = pair_f1(pair_counts(st_, sp_, T[z], nn[z]));
    if (m.f > s.f) { lbl[z] = LABEL_MERGE; sT[z] = T[z]; sP[z] = choose2(nn[z]); }
    else { lbl[z] = LABEL_SPLIT; sT[z] = st_; sP[z] = sp_; }
  }
  if (global_opt == 1) {  // below a merge node everything is a merge (top-down: parents have larger indices); the
    // reference starts at tree.root() = the LAST node, so only that node's tree is made consistent (:125-137)
    std::vector<char> inMain(N, 0);
    inMain[N - 1] = 1;
    for (int z = N - 2; z >= R; --z) {
      if (parent[z] < 0 || !inMain[parent[z]]) continue;
      inMain[z] = 1;
      if (lbl[parent[z]] == LABEL_MERGE) lbl[z] = LABEL_MERGE;
    }
  } else {                // above a split node everything is a split
    for (int z = R; z < N; ++z)
      if (lbl[z] == LABEL_SPLIT && parent[z] >= 0) lbl[parent[z]] = LABEL_SPLIT;
  }
  for (int i = 0; i < n; ++i) bc_labels[i] = lbl[R + i];
  if (merge_score) std::fill(merge_score, merge_score + n, 0.0);
  if (split_score) std::fill(split_score, split_score + n, 0.0);
  return GLIA_OK;
}
