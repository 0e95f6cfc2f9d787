// segment_greedy — drop-in for hmt/main_segment_greedy.cxx: resolve merge tree(s) into the final segmentation.
// Flags as the reference's (:101-127); -b (boundary confidence image) and -r (ITK relabel) are refused, -z ignored.
#include "hmt_setup.hpp"

int main(int argc, char** argv) {
  gh::Args a;
  a.add_alias("s", "segImage"); a.add_alias("o", "mergeOrders"); a.add_alias("p", "mergeProbs"); a.add_alias("n", "regionProbs");
  a.add_alias("m", "maskImage"); a.add_alias("i", "ignore"); a.add_alias("r", "relabel"); a.add_alias("u", "write16");
  a.add_alias("z", "compress"); a.add_alias("f", "finalSegImage"); a.add_alias("b", "bcImage");
  if (!a.parse(argc, argv) || a.has("help")) {
    std::cerr << "Usage: segment_greedy -s SEG -o ORDER [ORDER...] [-p PROBS...] [-n REGIONPROBS...] [-m MASK] [-i 0|1] [-u 0|1] -f OUT [--device N]\n";
    return a.has("help") ? EXIT_SUCCESS : EXIT_FAILURE;
  }
  a.require("segImage"); a.require("mergeOrders");
  if (a.flag("relabel")) gh::perr("Error: -r (ITK relabel) is not supported by this build");
  if (a.has("bcImage") && !a.str("bcImage").empty()) gh::perr("Error: -b (boundary confidence image) is not supported by this build");
  if (!a.has("finalSegImage") || a.str("finalSegImage").empty()) return EXIT_SUCCESS;  // nothing to write
  const std::vector<std::string> of = a.strs("mergeOrders"), pf = a.strs("mergeProbs"), rf = a.strs("regionProbs");
  const int nTree = (int)of.size();
  std::vector<std::vector<uint32_t>> orders(nTree);
  std::vector<std::vector<double>> probs(nTree), rprobs(nTree);
  std::vector<const uint32_t*> op(nTree); std::vector<int64_t> nm(nTree);
  std::vector<const double*> pp(nTree, nullptr), rp(nTree, nullptr);
  for (int i = 0; i < nTree; ++i) {
    orders[i] = gh::read_order(of[i]);
    op[i] = orders[i].data(); nm[i] = (int64_t)(orders[i].size() / 3);
    if ((int)pf.size() > i) {
      probs[i] = gh::read_values(pf[i]);
      if ((int64_t)probs[i].size() < nm[i]) gh::perr("Error: merge probability file shorter than merge order");
      pp[i] = probs[i].data();
    }
    if (!rf.empty()) {
      if ((int)rf.size() <= i) gh::perr("Error: one region probability file per merge order is needed");
      rprobs[i] = gh::read_values(rf[i]);
      if ((int64_t)rprobs[i].size() < 2 * nm[i] + 1 && nm[i] > 0) gh::perr("Error: region probability file shorter than the tree");
      rp[i] = rprobs[i].data();
    }
  }
  gh::HmtInputs in;
  in.seg = gh::read_label_image(a.str("segImage"));
  glia_params& p = in.prm;
  memset(&p, 0, sizeof(p));
  p.ndim = in.seg.ndim;
  for (int d = 0; d < 3; ++d) p.dims[d] = in.seg.dims[d];
  p.device = a.has("device") ? atoi(a.str("device").c_str()) : 0;
  std::vector<float> zeros(in.seg.size(), 0.0f);  // no image statistics are needed: the graph build only indexes the labels
  in.pb = zeros.data();
  if (a.has("maskImage") && !a.str("maskImage").empty()) {
    in.mask = gh::read_label_image(a.str("maskImage"));
    if (in.mask.size() != in.seg.size()) gh::perr("Error: mask size mismatch");
    in.hasMask = true;
  }
  gh::Handle H;
  gh::build_graph(in, H);
  std::vector<uint32_t> out(in.seg.size());
  const bool anyProbs = !pf.empty();
  H.check(glia_b200_segment_greedy(H.h, nTree, op.data(), nm.data(), anyProbs ? pp.data() : nullptr, rf.empty() ? nullptr : rp.data(),
                                   a.flag("ignore", true) ? 1 : 0, out.data(), nullptr, nullptr), "glia_b200_segment_greedy");
  gh::write_label_image(a.str("finalSegImage"), in.seg, out.data(), a.flag("write16"));
  return EXIT_SUCCESS;
}
