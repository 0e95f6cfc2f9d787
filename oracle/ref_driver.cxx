// TEST INFRASTRUCTURE — not part of the product.
//
// Thin C-ABI harness around the UNMODIFIED reference headers, which are
// included from /root/reference/code at build time (never copied into this
// repo).  It exists to (1) pin oracle/glia_oracle.c against the reference's own
// code, (2) generate the golden fixtures under tests/golden/, and (3) serve as
// the "reference" CPU baseline in bench.py.  Built by oracle/Makefile into
// oracle/_ref/libglia_ref{2d,3d}.so (git-ignored).
//
// The harness itself contains no algorithm: it only marshals raw buffers into
// the reference's image type, calls the reference entry points in the same way
// hmt/main_merge_order_bc.cxx:33-159 and hmt/main_bc_feat.cxx:27-112 do, and
// copies the results out.  The random-forest vote is NOT available from the
// reference tree (third-party classForest), so the predictor is
// a caller-supplied callback.
#include <cmath>
#include <cstring>
#include <cstdint>
#include <chrono>
#include "hmt/bc_feat.hxx"
#include "type/region_map.hxx"
#include "util/struct_merge_bc.hxx"
#include "util/struct_merge.hxx"
#include "hmt/tree_build.hxx"
#include "hmt/tree_greedy.hxx"
#include "hmt/bc_label.hxx"
#ifndef GLIA_3D
#include "hmt/sc_feat.hxx"
#endif
#include <queue>
#include "util/mp.hxx"

using namespace glia;
using namespace glia::hmt;

typedef double (*ref_pred_fn)(const double* row, int dim, void* ctx);
typedef TRegionMap<Label, Point<DIMENSION>> RegionMap;
typedef LabelImage<DIMENSION> LImage;
typedef RealImage<DIMENSION> RImage;

namespace {

struct Results {
  // RAG dump
  std::vector<uint32_t> reg_label, reg_count, reg_border;
  std::vector<uint32_t> edge_a, edge_b, edge_count;
  // candidates / rows
  std::vector<uint32_t> cand_keys;   // 3 per featurised candidate (r0, r1, r2-as-called)
  std::vector<double> cand_rows;     // dim per candidate, call sequence
  std::vector<double> cand_scores;   // one per candidate (NaN if no predictor)
  int dim = 0;
  std::vector<uint32_t> order;       // 3 per merge
  std::vector<double> sal;
  std::vector<double> merge_rows;    // -b output: one row per merge
  std::vector<uint32_t> out_labels;  // pre_merge: the label image after transformImage
  double t_rmap = 0, t_feat = 0, t_pred = 0, t_total = 0;
  int nthreads = 1;
};
Results G;

double now() {
  return std::chrono::duration<double>(
      std::chrono::steady_clock::now().time_since_epoch()).count();
}

template <typename TImg, typename T>
typename TImg::Pointer wrap(const T* buf, const int* dims) {
  auto img = TImg::New();
  typename TImg::RegionType reg;
  for (int d = 0; d < DIMENSION; ++d) {
    reg.GetModifiableIndex()[d] = 0;
    reg.GetModifiableSize()[d] = dims[d];
  }
  img->SetRegions(reg);
  img->Allocate();
  std::memcpy(img->GetBufferPointer(), buf, sizeof(T) * img->GetBufferSize());
  return img;
}

// util/image.hxx:173-192 pulls in ~25 ITK filters; these two helpers are
// restated here (product of sizes; sqrt of sum of squared sizes).
double imageVolume(const int* dims) {
  unsigned long v = 1;
  for (int d = 0; d < DIMENSION; ++d) v *= dims[d];
  return (double)v;
}
double imageDiagonal(const int* dims) {
  double r = 0.0;
  for (int d = 0; d < DIMENSION; ++d) { double l = dims[d]; r += l * l; }
  return std::sqrt(r);
}

typedef std::vector<ImageHistPair<RImage::Pointer>> ImgList;

void makeList(ImgList& out, int n, const float* const* imgs, const int* bins,
              const double* lo, const double* hi, const int* dims) {
  for (int i = 0; i < n; ++i) {
    out.emplace_back(wrap<RImage>(imgs[i], dims), (glia::uint)bins[i],
                     std::make_pair(lo[i], hi[i]));
  }
}

}  // namespace

extern "C" {

int glia_ref_dimension() { return DIMENSION; }

// mode: 0 RAG dump only; 1 initial candidates (TBoundaryTable::init only);
//       2 full greedy merge order; 3 bc_feat stage for a given order.
int glia_ref_run(
    int mode, const int* dims, const uint32_t* labels, const uint32_t* maskBuf,
    const float* pb,
    int nR, const float* const* rImgs, const int* rBins, const double* rLo, const double* rHi,
    int nRL, const float* const* rlImgs, const int* rlBins, const double* rlLo, const double* rlHi,
    int nB, const float* const* bImgs, const int* bBins, const double* bLo, const double* bHi,
    int nThr, const double* thr, int normalizeShape, int useLogShape,
    int useSimpleFeatures, ref_pred_fn pred, void* ctx,
    const uint32_t* orderIn, int nOrder, const double* salIn, double initSal,
    double salBias, int maxThreads)
{
  G = Results();
  double tStart = now();
  auto segImage = wrap<LImage>(labels, dims);
  LImage::Pointer mask = maskBuf ? wrap<LImage>(maskBuf, dims) : LImage::Pointer(nullptr);
  auto pbImage = wrap<RImage>(pb, dims);
  ImgList rImages, rlImages, bImages;
  makeList(rImages, nR, rImgs, rBins, rLo, rHi, dims);
  makeList(rlImages, nRL, rlImgs, rlBins, rlLo, rlHi, dims);
  makeList(bImages, nB, bImgs, bBins, bLo, bHi, dims);
  std::vector<double> boundaryThresholds(thr, thr + nThr);
  double normalizingArea = normalizeShape ? imageVolume(dims) : 1.0;
  double normalizingLength = normalizeShape ? imageDiagonal(dims) : 1.0;
#ifdef GLIA_MT
  G.nthreads = maxThreads > 0 ? maxThreads : omp_get_max_threads();
#endif

  if (mode == 3) {
    // ---- hmt/main_bc_feat.cxx:27-112 call sequence ----
    std::vector<TTriple<Label>> order;
    for (int i = 0; i < nOrder; ++i)
      order.emplace_back(orderIn[3 * i], orderIn[3 * i + 1], orderIn[3 * i + 2]);
    std::unordered_map<Label, double> saliencyMap;
    if (salIn) {
      std::vector<double> saliencies(salIn, salIn + nOrder);
      genSaliencyMap(saliencyMap, order, saliencies, initSal, salBias);
    }
    double t0 = now();
    RegionMap rmap(segImage, mask, order, false);
    G.t_rmap = now() - t0;
    t0 = now();
    int rn = rmap.size();
    std::vector<std::pair<Label, std::shared_ptr<RegionFeats>>> rfeats(rn);
    parfor(rmap, true, [&](RegionMap::const_iterator rit, int i) {
      rfeats[i].first = rit->first;
      rfeats[i].second = std::make_shared<RegionFeats>();
      rfeats[i].second->generate(
          rit->second, normalizingArea, normalizingLength, pbImage,
          boundaryThresholds, rImages, rlImages, bImages,
          ccpointer(saliencyMap, rit->first));
    }, maxThreads);
    std::unordered_map<Label, std::shared_ptr<RegionFeats>> rfmap;
    for (auto const& rfp : rfeats) { rfmap[rfp.first] = rfp.second; }
    int bn = order.size();
    std::vector<BoundaryClassificationFeats> bfeats(bn);
    parfor(0, bn, true, [&](int i) {
      Label r0 = order[i].x0, r1 = order[i].x1, r2 = order[i].x2;
      bfeats[i].x1 = rfmap.find(r0)->second.get();
      bfeats[i].x2 = rfmap.find(r1)->second.get();
      bfeats[i].x3 = rfmap.find(r2)->second.get();
      if (bfeats[i].x1->shape->area > bfeats[i].x2->shape->area) {
        std::swap(r0, r1);
        std::swap(bfeats[i].x1, bfeats[i].x2);
      }
      RegionMap::Region::Boundary b;
      getBoundary(b, rmap.find(r0)->second, rmap.find(r1)->second);
      bfeats[i].x0.generate(
          b, normalizingLength, *bfeats[i].x1, *bfeats[i].x2, *bfeats[i].x3,
          pbImage, boundaryThresholds, bImages);
    }, maxThreads);
    if (useLogShape) {
      parfor(0, rn, false, [&rfeats](int i) { rfeats[i].second->log(); }, maxThreads);
      parfor(0, bn, false, [&bfeats](int i) { bfeats[i].x0.log(); }, maxThreads);
    }
    G.t_feat = now() - t0;
    for (int i = 0; i < bn; ++i) {
      std::vector<FVal> row;
      if (useSimpleFeatures) { selectFeatures(row, bfeats[i]); }
      else { bfeats[i].serialize(row); }
      G.dim = row.size();
      G.merge_rows.insert(G.merge_rows.end(), row.begin(), row.end());
    }
    G.order.assign(orderIn, orderIn + 3 * (size_t)nOrder);
    G.t_total = now() - tStart;
    return 0;
  }

  double t0 = now();
  RegionMap rmap(segImage, mask, false);
  G.t_rmap = now() - t0;
  // RAG dump (sorted for stable comparison)
  {
    std::map<Label, std::pair<uint32_t, uint32_t>> regs;
    std::map<std::pair<Label, Label>, uint32_t> edges;
    for (auto const& rp : rmap) {
      regs[rp.first] = std::make_pair((uint32_t)rp.second.size(),
                                      (uint32_t)rp.second.border.size());
      for (auto const& bp : rp.second.boundary)
        edges[bp.first] = (uint32_t)bp.second->size();
    }
    for (auto const& r : regs) {
      G.reg_label.push_back(r.first);
      G.reg_count.push_back(r.second.first);
      G.reg_border.push_back(r.second.second);
    }
    for (auto const& e : edges) {
      G.edge_a.push_back(e.first.first);
      G.edge_b.push_back(e.first.second);
      G.edge_count.push_back(e.second);
    }
  }
  if (mode == 0) { G.t_total = now() - tStart; return 0; }

  // ---- hmt/main_merge_order_bc.cxx:50-146 call sequence ----
  rmap.reserve(4 * rmap.size() + 16);  // keep init() iteration rehash-free
  std::unordered_map<std::pair<Label, Label>, std::vector<FVal>> bcfmap;
  double tFeat = 0.0, tPred = 0.0;
  auto fBcFeat = [&](std::vector<FVal>& data, RegionMap::Region const& reg0,
                     RegionMap::Region const& reg1, RegionMap::Region const& reg2,
                     Label r0, Label r1, Label r2) {
    double tf = now();
    Label c0 = r0, c1 = r1;
    auto rf0 = std::make_shared<RegionFeats>();
    auto rf1 = std::make_shared<RegionFeats>();
    auto rf2 = std::make_shared<RegionFeats>();
    rf0->generate(reg0, normalizingArea, normalizingLength, pbImage,
                  boundaryThresholds, rImages, rlImages, bImages, nullptr);
    rf1->generate(reg1, normalizingArea, normalizingLength, pbImage,
                  boundaryThresholds, rImages, rlImages, bImages, nullptr);
    rf2->generate(reg2, normalizingArea, normalizingLength, pbImage,
                  boundaryThresholds, rImages, rlImages, bImages, nullptr);
    BoundaryClassificationFeats bcf;
    bcf.x1 = rf0.get();
    bcf.x2 = rf1.get();
    bcf.x3 = rf2.get();
    if (bcf.x1->shape->area > bcf.x2->shape->area) {
      std::swap(r0, r1);
      std::swap(bcf.x1, bcf.x2);
    }
    RegionMap::Region::Boundary b;
    getBoundary(b, reg0, reg1);
    bcf.x0.generate(b, normalizingLength, *bcf.x1, *bcf.x2, *bcf.x3, pbImage,
                    boundaryThresholds, bImages);
    if (useLogShape) {
      bcf.x0.log();
      rf0->log();
      rf1->log();
      rf2->log();
    }
    if (useSimpleFeatures) { selectFeatures(data, bcf); }
    else { bcf.serialize(data); }
    bcfmap[std::make_pair(std::min(r0, r1), std::max(r0, r1))] = data;
    tFeat += now() - tf;
    G.dim = data.size();
    G.cand_keys.push_back(c0);
    G.cand_keys.push_back(c1);
    G.cand_keys.push_back(r2);
    G.cand_rows.insert(G.cand_rows.end(), data.begin(), data.end());
  };
  auto fBcPred = [&](std::vector<FVal> const& data) -> double {
    double tp = now();
    double s = pred ? pred(data.data(), (int)data.size(), ctx) : 0.0;
    tPred += now() - tp;
    G.cand_scores.push_back(s);
    return s;
  };
  typedef TBoundaryTable<std::vector<FVal>, RegionMap> BT;
  if (mode == 1) {
    // Same lambdas genMergeOrderGreedyUsingBoundaryClassifier builds
    // (util/struct_merge_bc.hxx:18-27), table constructed but loop not run.
    auto initFb = [&rmap, &fBcFeat](std::vector<FVal>& data, Label r0, Label r1) {
      rmap.erase(BG_VAL);
      auto rit2 = rmap.merge(r0, r1, BG_VAL);
      fBcFeat(data, rmap.find(r0)->second, rmap.find(r1)->second,
              rit2->second, r0, r1, BG_VAL);
    };
    auto initFsal = [&fBcPred](std::vector<FVal> const& data, Label, Label) -> double {
      return fBcPred(data);
    };
    BT bt(rmap, initFb, initFsal);
    // re-emit in table (sorted key) order: rows were recorded in call order,
    // scores in table order; reorder rows to table order.
    std::map<std::pair<Label, Label>, size_t> callIdx;
    size_t n = G.cand_keys.size() / 3;
    for (size_t i = 0; i < n; ++i) {
      Label a = G.cand_keys[3 * i], b = G.cand_keys[3 * i + 1];
      callIdx[std::make_pair(std::min(a, b), std::max(a, b))] = i;
    }
    std::vector<uint32_t> keys;
    std::vector<double> rows;
    for (auto const& it : bt.table()) {
      size_t i = callIdx[it.first];
      keys.push_back(G.cand_keys[3 * i]);      // as passed to fBcFeat: the direction
      keys.push_back(G.cand_keys[3 * i + 1]);  // met first in unordered_map iteration
      keys.push_back(0);
      rows.insert(rows.end(), G.cand_rows.begin() + i * G.dim,
                  G.cand_rows.begin() + (i + 1) * G.dim);
    }
    G.cand_keys.swap(keys);
    G.cand_rows.swap(rows);
  } else {
    std::vector<TTriple<Label>> order;
    std::vector<double> saliencies;
    genMergeOrderGreedyUsingBoundaryClassifier<std::vector<FVal>>(
        order, saliencies, rmap, fBcFeat, fBcPred, f_true<BT&, BT::iterator>);
    for (auto const& m : order) {
      G.order.push_back(m.x0);
      G.order.push_back(m.x1);
      G.order.push_back(m.x2);
      auto const& row = bcfmap.find(std::make_pair(m.x0, m.x1))->second;
      G.merge_rows.insert(G.merge_rows.end(), row.begin(), row.end());
    }
    G.sal = saliencies;
  }
  G.t_feat = tFeat;
  G.t_pred = tPred;
  G.t_total = now() - tStart;
  return 0;
}

// ---- hmt/main_merge_order_pb.cxx:14-40 and gadget/main_pre_merge.cxx:7-87 ----
// mode 4: merge_order_pb (type 1 = approximate median, 2 = mean; rmap with contours only, regions not updated).
// mode 5: pre_merge: genMergeOrderGreedyUsingPbMean with the size / region-pb condition of main_pre_merge.cxx:26-76
//         (restated here because it lives in a main file), transformKeys + transformImage (util/image.hxx:246-259 is
//         behind the ITK filter includes: its 8-line loop is restated), no relabel (an ITK filter).
int glia_ref_run_pb(int mode, const int* dims, const uint32_t* labels, const uint32_t* maskBuf, const float* pb,
                    int type, const int* sizeThr, int nSizeThr, double rpbThreshold)
{
  G = Results();
  double tStart = now();
  auto segImage = wrap<LImage>(labels, dims);
  LImage::Pointer mask = maskBuf ? wrap<LImage>(maskBuf, dims) : LImage::Pointer(nullptr);
  auto pbImage = wrap<RImage>(pb, dims);
  std::vector<TTriple<Label>> order;
  std::vector<double> saliencies;
  if (mode == 4) {
    RegionMap rmap(segImage, mask, true);  // Only use contours
    if (type == 1) {
      genMergeOrderGreedyUsingPbApproxMedian(
          order, saliencies, rmap, false, pbImage,
          f_true<TBoundaryTable<std::vector<double>, RegionMap>&, TBoundaryTable<std::vector<double>, RegionMap>::iterator>,
          f_null<std::vector<double>&, Label, Label>);
    } else {
      genMergeOrderGreedyUsingPbMean(
          order, saliencies, rmap, false, pbImage,
          f_true<TBoundaryTable<std::pair<double, int>, RegionMap>&, TBoundaryTable<std::pair<double, int>, RegionMap>::iterator>);
    }
  } else {
    std::vector<int> sizeThresholds(sizeThr, sizeThr + nSizeThr);
    RegionMap rmap(segImage, mask, false);
    std::unordered_map<Label, double> rpbs;
    genMergeOrderGreedyUsingPbMean(
        order, saliencies, rmap, true, pbImage,
        [&pbImage, &sizeThresholds, &rmap, rpbThreshold, &rpbs](
            TBoundaryTable<std::pair<double, int>, RegionMap> const& bt,
            TBoundaryTable<std::pair<double, int>, RegionMap>::iterator btit) {
          Label key0 = btit->first.first;
          Label key1 = btit->first.second;
          auto const* pr0 = &rmap.find(key0)->second;
          auto const* pr1 = &rmap.find(key1)->second;
          auto sz0 = pr0->size();
          auto sz1 = pr1->size();
          if (sz0 > sz1) { std::swap(key0, key1); std::swap(pr0, pr1); std::swap(sz0, sz1); }
          if (sz0 < sizeThresholds[0]) { return true; }
          if (sizeThresholds.size() > 1) {
            const RegionMap::Region* prs[2] = {pr0, pr1};
            Label keys[2] = {key0, key1};
            decltype(sz0) szs[2] = {sz0, sz1};
            for (int k = 0; k < 2; ++k) {
              if (szs[k] < sizeThresholds[1]) {
                auto rpit = rpbs.find(keys[k]);
                if (rpit != rpbs.end()) { if (rpit->second > rpbThreshold) { return true; } }
                else {
                  double rpb = 0.0;
                  prs[k]->traverse([&pbImage, &rpb](RegionMap::Region::Point const& p) { rpb += pbImage->GetPixel(p); });
                  rpb = sdivide(rpb, szs[k], 0.0);
                  rpbs[keys[k]] = rpb;
                  if (rpb > rpbThreshold) { return true; }
                }
              }
            }
          }
          return false;
        });
    std::unordered_map<Label, Label> lmap;
    transformKeys(lmap, order);
    for (auto const& lp : lmap) {  // transformImage(segImage, rmap, lmap)
      auto rit = rmap.find(lp.first);
      if (rit != rmap.end()) {
        rit->second.traverse([&segImage, &lp](RegionMap::Region::Point const& p) { segImage->SetPixel(p, lp.second); });
      }
    }
    size_t n = segImage->GetBufferSize();
    G.out_labels.assign(segImage->GetBufferPointer(), segImage->GetBufferPointer() + n);
  }
  for (auto const& m : order) { G.order.push_back(m.x0); G.order.push_back(m.x1); G.order.push_back(m.x2); }
  G.sal = saliencies;
  G.t_total = now() - tStart;
  return 0;
}

// ---- hmt/main_segment_greedy.cxx:30-95 (tree resolution + final segmentation) ----
// Trees with node potentials from merge probabilities (hmt/tree_build.hxx:41-62), optional region probabilities,
// resolveTreeGreedy over all trees (hmt/tree_greedy.hxx:107-152), then genLabelTransform + transformImage
// (hmt/tree_segment.hxx:24-35, util/image.hxx:211-242 - both behind the ITK filter includes, their loops are restated).
struct SegNodeData { Label label; double potential; };
int glia_ref_segment_greedy(const int* dims, const uint32_t* labels, const uint32_t* maskBuf, int nTree,
                            const uint32_t* const* ordersIn, const int* nMerges, const double* const* mergeProbsIn,
                            const double* const* regionProbsIn, int ignore)
{
  G = Results();
  typedef TTree<SegNodeData> Tree;
  std::vector<std::vector<TTriple<Label>>> orders(nTree);
  std::vector<Tree> trees(nTree);
  for (int i = 0; i < nTree; ++i) {
    for (int k = 0; k < nMerges[i]; ++k)
      orders[i].emplace_back(ordersIn[i][3 * k], ordersIn[i][3 * k + 1], ordersIn[i][3 * k + 2]);
    if (mergeProbsIn && mergeProbsIn[i]) {
      std::vector<double> mergeProbs(mergeProbsIn[i], mergeProbsIn[i] + nMerges[i]);
      genTreeWithNodePotentials(trees[i], orders[i], mergeProbs.begin());
    } else {
      genTree(trees[i], orders[i], [](Tree::Node& node, Label r) { node.data.label = r; node.data.potential = 1.0; });
    }
    if (regionProbsIn && regionProbsIn[i]) {
      const double* rp = regionProbsIn[i];
      for (auto& tn : trees[i]) { tn.data.potential *= std::max(*rp++, FEPS); }
    }
  }
  auto segImage = wrap<LImage>(labels, dims);
  LImage::Pointer mask = maskBuf ? wrap<LImage>(maskBuf, dims) : LImage::Pointer(nullptr);
  std::vector<std::pair<int, int>> picks;
  resolveTreeGreedy(picks, trees, [](Tree::Node const& node0, Tree::Node const& node1) -> bool {
    return node0.data.potential < node1.data.potential; });
  std::unordered_map<Label, Label> lmap;   // genLabelTransform(lmap, trees, picks, 1u)
  Label keyToAssign = 1u;
  for (auto const& pick : picks) {
    trees[pick.first].traverseLeaves(pick.second, [keyToAssign, &lmap](Tree::Node const& node) { lmap[node.data.label] = keyToAssign; });
    ++keyToAssign;
  }
  size_t n = segImage->GetBufferSize();
  Label* px = segImage->GetBufferPointer();
  const Label* mk = mask.IsNull() ? nullptr : mask->GetBufferPointer();
  for (size_t p = 0; p < n; ++p) {   // transformImage(image, lmap, mask[, fillMissing = ignore])
    if (mk && mk[p] == MASK_OUT_VAL) continue;
    auto lit = lmap.find(px[p]);
    if (lit != lmap.end()) px[p] = lit->second;
    else if (ignore) px[p] = BG_VAL;
  }
  G.out_labels.assign(px, px + n);
  for (auto const& pk : picks) { G.order.push_back((uint32_t)pk.first); G.order.push_back((uint32_t)pk.second); G.order.push_back(0); }
  for (int i = 0; i < nTree; ++i) for (auto const& tn : trees[i]) G.sal.push_back(tn.data.potential);
  return 0;
}

// ---- hmt/main_bc_label_ri.cxx:31-155 (boundary-classifier training labels from a truth image) ----
// The label functions are the reference's own templates (hmt/bc_label.hxx:44-122, util/image_stats.hxx:173-245,
// util/stats.hxx:186-265), instantiated with __int128 where the reference's main uses boost's int512_t (boost is not
// in this image; the pair counts of these inputs are far below 2^63, so both integer types hold the same values).
// The option logic around them (optSplit, globalOpt 1 / 2) lives in the main file and is restated here.
typedef __int128 RefInt;
struct LabelNodeData { Label label; std::vector<int> bestSplits; int bcLabel = BC_LABEL_MERGE; };
int glia_ref_bc_label(const int* dims, const uint32_t* labels, const uint32_t* maskBuf, const uint32_t* truth,
                      const uint32_t* orderIn, int nOrder, int usePairF1, int optSplit, int globalOpt, int tweak,
                      double maxPrecDrop)
{
  G = Results();
  auto segImage = wrap<LImage>(labels, dims);
  auto truthImage = wrap<LImage>(truth, dims);
  LImage::Pointer mask = maskBuf ? wrap<LImage>(maskBuf, dims) : LImage::Pointer(nullptr);
  std::vector<TTriple<Label>> order;
  for (int i = 0; i < nOrder; ++i) order.emplace_back(orderIn[3 * i], orderIn[3 * i + 1], orderIn[3 * i + 2]);
  RegionMap rmap(segImage, mask, order, false);
  int n = order.size();
  std::vector<int> bcLabels;
  std::vector<double> mergeF1s(n, 0.0), splitF1s(n, 0.0);
  if (globalOpt == 0) {
    bcLabels.resize(n);
    if (usePairF1) {
      for (int i = 0; i < n; ++i)
        bcLabels[i] = genBoundaryClassificationLabelF1<RefInt>(mergeF1s[i], splitF1s[i], rmap, order[i].x0, order[i].x1,
                                                               order[i].x2, truthImage, tweak != 0, maxPrecDrop);
      if (optSplit) {
        std::unordered_map<Label, std::vector<RegionMap::Region const*>> smap;
        for (int i = 0; i < n; ++i) {
          auto sit0 = citerator(smap, order[i].x0, 1, &rmap.find(order[i].x0)->second);
          auto sit1 = citerator(smap, order[i].x1, 1, &rmap.find(order[i].x1)->second);
          std::vector<RegionMap::Region const*> pSplitRegions, pMergeRegions{&rmap.find(order[i].x2)->second};
          append(pSplitRegions, sit0->second, sit1->second);
          if (bcLabels[i] == BC_LABEL_SPLIT) { splice(smap[order[i].x2], pSplitRegions); }
          else {
            double optSplitF1, optSplitPrec, optSplitRec;
            stats::pairF1<RefInt>(optSplitF1, optSplitPrec, optSplitRec, pSplitRegions, truthImage, {BG_VAL});
            if (mergeF1s[i] > optSplitF1) { splice(smap[order[i].x2], pMergeRegions); }
            else { bcLabels[i] = BC_LABEL_SPLIT; splice(smap[order[i].x2], pSplitRegions); }
          }
        }
      }
    } else {
      for (int i = 0; i < n; ++i)
        bcLabels[i] = genBoundaryClassificationLabelRI<RefInt>(mergeF1s[i], splitF1s[i], rmap, order[i].x0, order[i].x1,
                                                               order[i].x2, truthImage);
    }
  } else {
    typedef TTree<LabelNodeData> Tree;
    Tree tree;
    genTree(tree, order, [](Tree::Node& node, Label r) { node.data.label = r; });
    for (auto& node : tree) {
      if (node.isLeaf()) { node.data.bcLabel = BC_LABEL_MERGE; node.data.bestSplits.push_back(node.self); }
      else {
        std::vector<RegionMap::Region const*> pMergeRegions{&rmap.find(node.data.label)->second};
        std::vector<RegionMap::Region const*> pSplitRegions;
        for (auto c : node.children)
          for (auto j : tree[c].data.bestSplits) pSplitRegions.push_back(&rmap.find(tree[j].data.label)->second);
        double mergeF1 = stats::pairF1<RefInt>(pMergeRegions, truthImage, {BG_VAL});
        double splitF1 = stats::pairF1<RefInt>(pSplitRegions, truthImage, {BG_VAL});
        if (mergeF1 > splitF1) { node.data.bcLabel = BC_LABEL_MERGE; node.data.bestSplits.push_back(node.self); }
        else {
          node.data.bcLabel = BC_LABEL_SPLIT;
          for (auto c : node.children) for (auto j : tree[c].data.bestSplits) node.data.bestSplits.push_back(j);
        }
      }
    }
    if (globalOpt == 1) {
      std::queue<int> q;
      q.push(tree.root());
      while (!q.empty()) {
        int ni = q.front(); q.pop();
        if (tree[ni].data.bcLabel == BC_LABEL_MERGE)
          tree.traverseDescendants(ni, [](Tree::Node& node) { node.data.bcLabel = BC_LABEL_MERGE; });
        else for (auto c : tree[ni].children) q.push(c);
      }
    } else {
      for (auto const& node : tree)
        if (node.data.bcLabel == BC_LABEL_SPLIT)
          tree.traverseAncestors(node.self, [](Tree::Node& tn) { tn.data.bcLabel = BC_LABEL_SPLIT; });
    }
    for (auto const& node : tree) if (!node.isLeaf()) bcLabels.push_back(node.data.bcLabel);
  }
  for (int v : bcLabels) G.order.push_back((uint32_t)(int32_t)v);
  G.sal = mergeF1s;
  G.cand_scores = splitF1s;
  return 0;
}

#ifndef GLIA_3D
// ---- hmt/main_sc_feat.cxx:27-117 (section-classifier features of region pairs across two 2D
// sections: RegionFeatsWithLocation incl. central / Hu moments and eccentricity, RegionPairFeats incl. overlaps).
// The feature classes are the reference's own (hmt/sc_feat.hxx, type/feat.hxx, alg/geometry.hxx); the call sequence of
// the main file is restated.  pairs: [n][2] = (label in section 0, label in section 1).
int glia_ref_sc_feat(const int* dims, const uint32_t* labels0, const uint32_t* mask0Buf, const float* pb0,
                     const uint32_t* labels1, const uint32_t* mask1Buf, const float* pb1, int bins, double lo, double hi,
                     int nThr, const double* thr, double normalizingArea, double normalizingLength, int useLogShape,
                     const uint32_t* pairs, int nPairs)
{
  G = Results();
  auto segImage0 = wrap<LImage>(labels0, dims), segImage1 = wrap<LImage>(labels1, dims);
  LImage::Pointer mask0 = mask0Buf ? wrap<LImage>(mask0Buf, dims) : LImage::Pointer(nullptr);
  LImage::Pointer mask1 = mask1Buf ? wrap<LImage>(mask1Buf, dims) : LImage::Pointer(nullptr);
  auto pbImage0 = wrap<RImage>(pb0, dims), pbImage1 = wrap<RImage>(pb1, dims);
  ImgList rImages0, rlImages0, bImages0, rImages1, rlImages1, bImages1;   // --rbi pb (prepareImages, hmt_util.hxx:17-56)
  rImages0.emplace_back(pbImage0, (glia::uint)bins, std::make_pair(lo, hi)); bImages0.push_back(rImages0.back());
  rImages1.emplace_back(pbImage1, (glia::uint)bins, std::make_pair(lo, hi)); bImages1.push_back(rImages1.back());
  std::vector<double> boundaryThresholds(thr, thr + nThr);
  if (normalizingArea <= 0.0) normalizingArea = imageVolume(dims);       // max over two images of equal extent
  if (normalizingLength <= 0.0) normalizingLength = imageDiagonal(dims);
  RegionMap rmap0(segImage0, mask0, false), rmap1(segImage1, mask1, false);
  std::unordered_map<Label, std::shared_ptr<RegionFeatsWithLocation>> rfmap0, rfmap1;
  for (auto const& rp : rmap0) {
    auto f = std::make_shared<RegionFeatsWithLocation>();
    f->generate(rp.second, normalizingArea, normalizingLength, pbImage0, boundaryThresholds, rImages0, rlImages0, bImages0, nullptr);
    rfmap0[rp.first] = f;
  }
  for (auto const& rp : rmap1) {
    auto f = std::make_shared<RegionFeatsWithLocation>();
    f->generate(rp.second, normalizingArea, normalizingLength, pbImage1, boundaryThresholds, rImages1, rlImages1, bImages1, nullptr);
    rfmap1[rp.first] = f;
  }
  std::unordered_map<std::pair<Label, Label>, int> overlaps;
  stats::getOverlap(overlaps, segImage0, mask0, {BG_VAL}, segImage1, mask1, {BG_VAL});
  std::vector<SectionClassificationFeats> sfeats(nPairs);
  for (int i = 0; i < nPairs; ++i) {
    Label r0 = pairs[2 * i], r1 = pairs[2 * i + 1];
    auto rf0 = rfmap0.find(r0)->second.get();
    auto rf1 = rfmap1.find(r1)->second.get();
    sfeats[i].x0.generate(*rf0, *rf1, (double)clookup(overlaps, std::make_pair(r0, r1), 0));
    if (rf0->shape->area > rf1->shape->area) { std::swap(r0, r1); std::swap(rf0, rf1); }
    sfeats[i].x1 = rf0;
    sfeats[i].x2 = rf1;
  }
  if (useLogShape) {
    for (auto& f : rfmap0) f.second->log();
    for (auto& f : rfmap1) f.second->log();
    for (auto& f : sfeats) f.x0.log();
  }
  for (int i = 0; i < nPairs; ++i) {
    std::vector<FVal> row;
    sfeats[i].serialize(row);
    G.dim = row.size();
    G.merge_rows.insert(G.merge_rows.end(), row.begin(), row.end());
  }
  return 0;
}
#endif

// Leaf RegionFeats under the reference's own parfor (the GLIA_MT path,
// hmt/main_bc_feat.cxx:59-69) — timing only, for the CPU baseline.
double glia_ref_time_leaf_feats(
    const int* dims, const uint32_t* labels, const float* pb, int bins,
    double lo, double hi, int nThr, const double* thr, int maxThreads,
    double* tRmapOut)
{
  auto segImage = wrap<LImage>(labels, dims);
  LImage::Pointer mask(nullptr);
  auto pbImage = wrap<RImage>(pb, dims);
  ImgList rImages, rlImages, bImages;
  rImages.emplace_back(pbImage, (glia::uint)bins, std::make_pair(lo, hi));
  bImages.push_back(rImages.back());
  std::vector<double> boundaryThresholds(thr, thr + nThr);
  double t0 = now();
  RegionMap rmap(segImage, mask, false);
  if (tRmapOut) *tRmapOut = now() - t0;
  int rn = rmap.size();
  std::vector<std::shared_ptr<RegionFeats>> rfeats(rn);
  t0 = now();
  parfor(rmap, true, [&](RegionMap::const_iterator rit, int i) {
    rfeats[i] = std::make_shared<RegionFeats>();
    rfeats[i]->generate(rit->second, 1.0, 1.0, pbImage, boundaryThresholds,
                        rImages, rlImages, bImages, nullptr);
  }, maxThreads);
  return now() - t0;
}

#define GETTER(name, type, vec) \
  size_t glia_ref_n_##name() { return G.vec.size(); } \
  const type* glia_ref_##name() { return G.vec.data(); }
GETTER(reg_label, uint32_t, reg_label)
GETTER(reg_count, uint32_t, reg_count)
GETTER(reg_border, uint32_t, reg_border)
GETTER(edge_a, uint32_t, edge_a)
GETTER(edge_b, uint32_t, edge_b)
GETTER(edge_count, uint32_t, edge_count)
GETTER(cand_keys, uint32_t, cand_keys)
GETTER(cand_rows, double, cand_rows)
GETTER(cand_scores, double, cand_scores)
GETTER(order, uint32_t, order)
GETTER(sal, double, sal)
GETTER(merge_rows, double, merge_rows)
GETTER(out_labels, uint32_t, out_labels)
int glia_ref_dim() { return G.dim; }
int glia_ref_nthreads() { return G.nthreads; }
double glia_ref_time(int which) {
  switch (which) {
    case 0: return G.t_rmap;
    case 1: return G.t_feat;
    case 2: return G.t_pred;
    default: return G.t_total;
  }
}

}  // extern "C"
