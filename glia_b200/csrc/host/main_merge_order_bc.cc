// merge_order_bc — drop-in for hmt/main_merge_order_bc.cxx (same flags, same output files):
// greedy merge order + saliencies (+ per-merge boundary-classifier features) from a superpixel
// image, a boundary map and a random-forest boundary classifier.  The work runs on the GPU
// through libglia_b200; there is no CPU path.
#include "hmt_setup.hpp"

int main(int argc, char** argv) {
  gh::Args a;
  gh::add_alias_common(a);
  if (!a.parse(argc, argv) || a.has("help")) {
    std::cerr << "Usage: merge_order_bc --bct 1 --bcm MODEL -s SEG --pb PB [--rbi IMG --rbb BINS --rbl LO --rbu HI]...\n"
                 "       [--rli/--rlb/--rll/--rlu] [--ri/--rb/--rl/--ru] [--bi/--bb/--bl/--bu] [--bt T...] [-m MASK]\n"
                 "       [--bcm M1 --bcm M2 --bcmd DIM0 DIM1 THR] [-n 0|1] [-l 0|1] [--simpf 0|1] [--histf 0|1] [--medf 0|1] -o ORDER [--sal SAL] [-b BCFEAT] [--device N]\n"
                 "       [--ref-order N]  equal-area pairs follow the unordered_map call order of the reference (N = its rmap.reserve, 0 = none)\n"
                 "Images: .npy, .mha, .mhd (raw).\n";
    return a.has("help") ? EXIT_SUCCESS : EXIT_FAILURE;
  }
  a.require("bct"); a.require("bcm"); a.require("mergeOrder");
  if (atoi(a.str("bct").c_str()) != 1) gh::perr("Error: only --bct 1 (random forest) is supported by this build");
  const std::vector<std::string> bcm = a.strs("bcm");
  const std::vector<double> bcmd = a.nums("bcmd");
  if (bcm.size() > 3) gh::perr("Error: at most three --bcm models");
  if (bcm.size() > 1 && bcmd.size() != 3) gh::perr("Error: incorrect boundary classifier model distributor args...");  // main_merge_order_bc.cxx:103-104
  gh::HmtInputs in;
  gh::setup_inputs(a, in);
  gh::Handle H;
  gh::build_graph(in, H);
  glia_rf* rf = nullptr;
  int st = GLIA_OK;
  if (bcm.size() == 1) {
    st = glia_b200_rf_load(bcm[0].c_str(), in.prm.device, &rf);
    if (st != GLIA_OK) gh::perr("Error: cannot read model file " + bcm[0]);
  } else {  // alg::EnsembleRandomForest, main_merge_order_bc.cxx:102-109
    glia_rf* sub[3] = {nullptr, nullptr, nullptr};
    for (size_t i = 0; i < bcm.size(); ++i)
      if (glia_b200_rf_load(bcm[i].c_str(), in.prm.device, &sub[i]) != GLIA_OK) gh::perr("Error: cannot read model file " + bcm[i]);
    st = glia_b200_rf_ensemble(sub, (int)bcm.size(), (int)bcmd[0], (int)bcmd[1], bcmd[2], &rf);
    for (size_t i = 0; i < bcm.size(); ++i) glia_b200_rf_destroy(sub[i]);
    if (st != GLIA_OK) gh::perr("Error: cannot combine the boundary classifier models");
  }
  int64_t R = 0, E = 0, C = 0;
  H.check(glia_b200_counts(H.h, &R, &E, &C), "glia_b200_counts");
  const int dim = glia_b200_row_dim(H.h, 0);
  const bool wantRows = a.has("bfeat") && !a.str("bfeat").empty();
  std::vector<uint32_t> order(3 * (size_t)(R > 0 ? R : 1));
  std::vector<double> sal((size_t)(R > 0 ? R : 1)), rows(wantRows ? (size_t)R * dim : 0);
  if (a.has("ref-order")) {  // include/glia_b200.h: equal-area candidates keep the reference binary's call order
    std::vector<uint8_t> secondFirst((size_t)(C > 0 ? C : 1));
    H.check(glia_b200_reference_orientation(H.h, (int64_t)atoll(a.str("ref-order").c_str()), secondFirst.data()), "glia_b200_reference_orientation");
    H.check(glia_b200_set_tie_orientation(H.h, secondFirst.data()), "glia_b200_set_tie_orientation");
  }
  int64_t nm = 0, nf = 0;
  H.check(glia_b200_merge_order_bc(H.h, rf, order.data(), sal.data(), wantRows ? rows.data() : nullptr, &nm, &nf),
          "glia_b200_merge_order_bc");
  gh::write_order(a.str("mergeOrder"), order.data(), (size_t)nm);
  if (a.has("sal") && !a.str("sal").empty()) gh::write_values(a.str("sal"), sal.data(), (size_t)nm, -1);
  if (wantRows) gh::write_rows(a.str("bfeat"), rows.data(), (size_t)nm, (size_t)dim, 8);  // FLT_PREC
  glia_b200_rf_destroy(rf);
  return EXIT_SUCCESS;
}
