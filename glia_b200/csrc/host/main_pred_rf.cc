// pred_rf — drop-in for ml/rf/main_pred_rf.cxx: batch random-forest scoring of feature files.
//   pred_rf --m MODEL --f FEATS... --l LABEL --p PREDS...
#include "../../../include/glia_b200.h"
#include "host_util.hpp"

int main(int argc, char** argv) {
  gh::Args a;
  if (!a.parse(argc, argv) || a.has("help")) {
    std::cerr << "Usage: pred_rf --m MODEL --f FEAT_FILE... --l LABEL --p PRED_FILE... [--device N]\n";
    return a.has("help") ? EXIT_SUCCESS : EXIT_FAILURE;
  }
  a.require("m"); a.require("f"); a.require("l"); a.require("p");
  if (a.strs("m").size() != 1) gh::perr("Error: ensemble models (--md) are not supported by this build");
  auto feats = a.strs("f"), preds = a.strs("p");
  if (feats.size() != preds.size()) gh::perr("Error: need one --p per --f");
  const int device = a.has("device") ? atoi(a.str("device").c_str()) : 0;
  glia_rf* rf = nullptr;
  if (glia_b200_rf_load(a.str("m").c_str(), device, &rf) != GLIA_OK) gh::perr("Error: cannot read model file " + a.str("m"));
  const int label = atoi(a.str("l").c_str());
  for (size_t i = 0; i < feats.size(); ++i) {
    size_t n = 0, d = 0;
    std::vector<double> X = gh::read_rows(feats[i], n, d);
    std::vector<double> p(n);
    int st = glia_b200_rf_score(rf, X.data(), n, (int)d, label, nullptr, p.data(), 0, nullptr);
    if (st != GLIA_OK) gh::perr("Error: glia_b200_rf_score failed (" + std::to_string(st) + ")");
    gh::write_values(preds[i], p.data(), n, 8);  // FLT_PREC
  }
  glia_b200_rf_destroy(rf);
  return EXIT_SUCCESS;
}
