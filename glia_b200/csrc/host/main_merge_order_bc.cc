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
                 "       [-n 0|1] [-l 0|1] [--simpf 0|1] -o ORDER [--sal SAL] [-b BCFEAT] [--device N]\n"
                 "Images: .npy, .mha, .mhd (raw).\n";
    return a.has("help") ? EXIT_SUCCESS : EXIT_FAILURE;
  }
  a.require("bct"); a.require("bcm"); a.require("mergeOrder");
  if (atoi(a.str("bct").c_str()) != 1) gh::perr("Error: only --bct 1 (random forest) is supported by this build");
  if (a.strs("bcm").size() != 1) gh::perr("Error: ensemble models (--bcmd) are not supported by this build");
  gh::HmtInputs in;
  gh::setup_inputs(a, in);
  gh::Handle H;
  gh::build_graph(in, H);
  glia_rf* rf = nullptr;
  int st = glia_b200_rf_load(a.str("bcm").c_str(), in.prm.device, &rf);
  if (st != GLIA_OK) gh::perr("Error: cannot read model file " + a.str("bcm"));
  int64_t R = 0, E = 0, C = 0;
  H.check(glia_b200_counts(H.h, &R, &E, &C), "glia_b200_counts");
  const int dim = glia_b200_row_dim(H.h, 0);
  const bool wantRows = a.has("bfeat") && !a.str("bfeat").empty();
  std::vector<uint32_t> order(3 * (size_t)(R > 0 ? R : 1));
  std::vector<double> sal((size_t)(R > 0 ? R : 1)), rows(wantRows ? (size_t)R * dim : 0);
  int64_t nm = 0, nf = 0;
  H.check(glia_b200_merge_order_bc(H.h, rf, order.data(), sal.data(), wantRows ? rows.data() : nullptr, &nm, &nf),
          "glia_b200_merge_order_bc");
  gh::write_order(a.str("mergeOrder"), order.data(), (size_t)nm);
  if (a.has("sal") && !a.str("sal").empty()) gh::write_values(a.str("sal"), sal.data(), (size_t)nm, -1);
  if (wantRows) gh::write_rows(a.str("bfeat"), rows.data(), (size_t)nm, (size_t)dim, 8);  // FLT_PREC
  glia_b200_rf_destroy(rf);
  return EXIT_SUCCESS;
}
