// bc_feat — drop-in for hmt/main_bc_feat.cxx: boundary-classifier feature rows for every merge of
// a given merge order (optionally with saliency features: -y SAL --s0 INIT --sb BIAS).
// Rows are written in the column-parseable `serialize` layout (value + " " each);
// the reference's non-simple operator<< output, which fuses two columns, is not reproduced.
#include "hmt_setup.hpp"

int main(int argc, char** argv) {
  gh::Args a;
  gh::add_alias_common(a);
  if (!a.parse(argc, argv) || a.has("help")) {
    std::cerr << "Usage: bc_feat -s SEG -o ORDER [-y SAL --s0 INIT --sb BIAS] --pb PB [--rbi/--rbb/--rbl/--rbu]...\n"
                 "       [--bt T...] [-m MASK] [-n 0|1] [-l 0|1] [--simpf 0|1] [--histf 0|1] [--medf 0|1] -b BCFEAT [--device N]\n";
    return a.has("help") ? EXIT_SUCCESS : EXIT_FAILURE;
  }
  a.require("mergeOrder"); a.require("bfeat");
  gh::HmtInputs in;
  gh::setup_inputs(a, in);
  std::vector<uint32_t> order = gh::read_order(a.str("mergeOrder"));
  const size_t n = order.size() / 3;
  std::vector<double> sal;
  const bool withSal = a.has("saliency") && !a.str("saliency").empty();
  if (withSal) {
    sal = gh::read_values(a.str("saliency"));
    if (sal.size() < n) gh::perr("Error: saliency file shorter than merge order");
  }
  const double initSal = a.has("s0") ? atof(a.str("s0").c_str()) : 1.0;
  const double salBias = a.has("sb") ? atof(a.str("sb").c_str()) : 0.0;
  gh::Handle H;
  gh::build_graph(in, H);
  const int dim = glia_b200_row_dim(H.h, withSal ? 1 : 0);
  std::vector<double> rows(n * (size_t)dim);
  H.check(glia_b200_bc_feat(H.h, order.data(), (int64_t)n, withSal ? sal.data() : nullptr, initSal, salBias, rows.data()),
          "glia_b200_bc_feat");
  gh::write_rows(a.str("bfeat"), rows.data(), n, (size_t)dim, 8);
  return EXIT_SUCCESS;
}
