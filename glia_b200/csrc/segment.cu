// resolve merge trees into a final segmentation.
//
// Replaces (reference, /root/reference/code):
//   segment_greedy operation()                 hmt/main_segment_greedy.cxx:30-95
//   genTree / genTreeWithNodePotentials        hmt/tree_build.hxx:12-62
//   resolveTreeGreedy (trees, picks)           hmt/tree_greedy.hxx:80-152
//   genLabelTransform / genFinalSegmentation   hmt/tree_segment.hxx:24-65
//   transformImage (lmap, mask, fillMissing)   util/image.hxx:211-242
// The tree logic is a few passes over 2 R nodes on the host; the part that touches the volume - renaming every voxel
// to the key of the picked node above its superpixel - is one LUT gather on the device (relabel_kernel,
// merge_order_pb.cu).  The reference finds every pick by a linear argmax over all nodes of all trees (O(n^2)); nodes
// only ever become invalid, so one descending sort by (potential, tree, node) and a cursor give the same picks.
#include <algorithm>
#include <unordered_map>
#include <vector>
#include "handle.cuh"

int glia_relabel_by_region(glia_handle* h, const std::vector<uint32_t>& lut, uint32_t* labels_out);

#define CK(expr) GLIA_CUDA_CHECK(h, expr)

namespace {
struct Node { uint32_t label; int parent, c0, c1; double potential; };
struct Tree {
  std::vector<Node> nodes;
  bool leaf(int i) const { return nodes[i].c0 < 0; }
  template <class F> void descendants(int i, F f) const {  // breadth first, child by child (type/tree.hxx:120-124)
    std::vector<int> q;
    for (int c : {nodes[i].c0, nodes[i].c1}) {
      if (c < 0) continue;
      q.assign(1, c);
      for (size_t k = 0; k < q.size(); ++k) {
        const int v = q[k];
        f(v);
        if (nodes[v].c0 >= 0) { q.push_back(nodes[v].c0); q.push_back(nodes[v].c1); }
      }
    }
  }
  template <class F> void leaves(int i, F f) const {  // type/tree.hxx:154-162
    if (leaf(i)) { f(i); return; }
    descendants(i, [&](int v) { if (leaf(v)) f(v); });
  }
};

// hmt/tree_build.hxx:12-62: nodes in order of first appearance in the merge order
void build_tree(Tree& t, const uint32_t* order, int64_t n, const double* mergeProbs) {
  std::unordered_map<uint32_t, int> nmap;
  nmap.reserve((size_t)(2 * n + 1) * 2);
  t.nodes.reserve((size_t)(2 * n + 1));
  auto add = [&](uint32_t label, int c0, int c1) {
    t.nodes.push_back(Node{label, -1, c0, c1, 1.0});
    return (int)t.nodes.size() - 1;
  };
  for (int64_t i = 0; i < n; ++i) {
    const uint32_t x0 = order[3 * i], x1 = order[3 * i + 1], x2 = order[3 * i + 2];
    auto it0 = nmap.find(x0);
    const int n0 = it0 != nmap.end() ? it0->second : (nmap[x0] = add(x0, -1, -1));
    auto it1 = nmap.find(x1);
    const int n1 = it1 != nmap.end() ? it1->second : (nmap[x1] = add(x1, -1, -1));
    const int z = add(x2, n0, n1);
    t.nodes[n0].parent = z; t.nodes[n1].parent = z;
    if (mergeProbs) {  // tree_build.hxx:47-59: children take the split probability of this merge
      t.nodes[z].potential = mergeProbs[i];
      const double pSplit = 1.0 - mergeProbs[i];
      for (int c : {n0, n1}) {
        if (t.leaf(c)) t.nodes[c].potential = pSplit * pSplit;
        else t.nodes[c].potential *= pSplit;
      }
    }
    nmap.emplace(x2, z);
  }
  if (mergeProbs && !t.nodes.empty()) t.nodes.back().potential *= t.nodes.back().potential;  // root, :60
}
}  // namespace

extern "C" int glia_b200_segment_greedy(glia_handle* h, int n_trees, const uint32_t* const* orders, const int64_t* n_merges,
                                        const double* const* merge_probs, const double* const* region_probs,
                                        int ignore_missing, uint32_t* labels_out, int32_t* picks, int64_t* n_picks) {
  if (!h || n_trees < 1 || !orders || !n_merges || !labels_out) return GLIA_E_ARG;
  if (!h->built) return GLIA_E_STATE;
  CK(cudaSetDevice(h->prm.device));
  std::vector<Tree> trees(n_trees);
  size_t total = 0;
  for (int i = 0; i < n_trees; ++i) {
    if (n_merges[i] < 0 || (n_merges[i] > 0 && !orders[i])) return GLIA_E_ARG;
    build_tree(trees[i], orders[i], n_merges[i], merge_probs ? merge_probs[i] : nullptr);
    if (region_probs && region_probs[i])  // main_segment_greedy.cxx:52-58
      for (size_t k = 0; k < trees[i].nodes.size(); ++k) trees[i].nodes[k].potential *= std::max(region_probs[i][k], 2.22e-16);
    total += trees[i].nodes.size();
  }
  // ---- resolveTreeGreedy (hmt/tree_greedy.hxx:107-152)
  std::vector<std::vector<char>> valid(n_trees);
  std::vector<std::unordered_map<uint32_t, int>> lnmap(n_trees);
  struct Cand { double p; int t, v; };
  std::vector<Cand> cand;
  cand.reserve(total);
  for (int i = 0; i < n_trees; ++i) {
    const Tree& t = trees[i];
    valid[i].assign(t.nodes.size(), 1);
    if (!t.nodes.empty()) t.leaves((int)t.nodes.size() - 1, [&](int v) { lnmap[i][t.nodes[v].label] = v; });  // leaves below the root
    for (size_t v = 0; v < t.nodes.size(); ++v) cand.push_back(Cand{t.nodes[v].potential, i, (int)v});
  }
  // the reference's argmax replaces its pick only on a strictly larger potential: first maximum in (tree, node) order
  std::stable_sort(cand.begin(), cand.end(), [](const Cand& a, const Cand& b) { return a.p > b.p; });
  std::vector<std::pair<int, int>> pk;
  std::vector<uint32_t> llabels;
  for (const Cand& c : cand) {
    if (!valid[c.t][c.v]) continue;
    pk.push_back({c.t, c.v});
    const Tree& t = trees[c.t];
    valid[c.t][c.v] = 0;
    for (int a = t.nodes[c.v].parent; a >= 0; a = t.nodes[a].parent) valid[c.t][a] = 0;
    llabels.clear();
    t.descendants(c.v, [&](int v) { valid[c.t][v] = 0; if (t.leaf(v)) llabels.push_back(t.nodes[v].label); });
    for (uint32_t ll : llabels)
      for (int i = 0; i < n_trees; ++i) {
        if (i == c.t) continue;
        auto it = lnmap[i].find(ll);
        if (it == lnmap[i].end()) continue;
        valid[i][it->second] = 0;
        for (int a = trees[i].nodes[it->second].parent; a >= 0; a = trees[i].nodes[a].parent) valid[i][a] = 0;
      }
  }
  // NaN potentials never win a comparison in the reference either way; they sort last here and are picked last there.
  // ---- genLabelTransform (hmt/tree_segment.hxx:24-35) + transformImage (util/image.hxx:211-242)
  std::unordered_map<uint32_t, uint32_t> lmap;
  uint32_t key = 1;
  for (auto& p : pk) { trees[p.first].leaves(p.second, [&](int v) { lmap[trees[p.first].nodes[v].label] = key; }); ++key; }
  const size_t R = (size_t)h->R;
  std::vector<uint32_t> lab(R), lut(R);
  if (R) { CK(cudaMemcpyAsync(lab.data(), h->regLabel.p, 4 * R, cudaMemcpyDeviceToHost, h->stream)); CK(cudaStreamSynchronize(h->stream)); }
  for (size_t r = 0; r < R; ++r) {
    auto it = lmap.find(lab[r]);
    lut[r] = it != lmap.end() ? it->second : (ignore_missing ? 0u : lab[r]);
  }
  if (n_picks) *n_picks = (int64_t)pk.size();
  if (picks) for (size_t i = 0; i < pk.size(); ++i) { picks[2 * i] = pk[i].first; picks[2 * i + 1] = pk[i].second; }
  return glia_relabel_by_region(h, lut, labels_out);
}
