// bc_label_ri — drop-in for hmt/main_bc_label_ri.cxx: boundary-classifier training labels (-1 merge / +1 split) for
// every merge of an order, from a ground-truth segmentation.  Flags as the reference's (:162-185).
#include "hmt_setup.hpp"

int main(int argc, char** argv) {
  gh::Args a;
  a.add_alias("s", "segImage"); a.add_alias("o", "mergeOrder"); a.add_alias("t", "truthImage"); a.add_alias("n", "maskImage");
  a.add_alias("g", "opt"); a.add_alias("p", "optSplit"); a.add_alias("w", "tweak"); a.add_alias("d", "mpd"); a.add_alias("l", "bclabel");
  if (!a.parse(argc, argv) || a.has("help")) {
    std::cerr << "Usage: bc_label_ri -s SEG -o ORDER -t TRUTH [-n MASK] [--f1 0|1] [-g 0|1|2] [-p 0|1] [-w 0|1] [-d MPD] [-l BCLABEL] [--device N]\n";
    return a.has("help") ? EXIT_SUCCESS : EXIT_FAILURE;
  }
  a.require("segImage"); a.require("mergeOrder"); a.require("truthImage");
  const int globalOpt = a.has("opt") ? atoi(a.str("opt").c_str()) : 0;
  if (globalOpt < 0 || globalOpt > 2) gh::perr("Error: unsupported globalOpt type...");
  gh::HmtInputs in;
  in.seg = gh::read_label_image(a.str("segImage"));
  gh::Image truth = gh::read_label_image(a.str("truthImage"));
  if (truth.size() != in.seg.size()) gh::perr("Error: truth image size mismatch");
  glia_params& p = in.prm;
  memset(&p, 0, sizeof(p));
  p.ndim = in.seg.ndim;
  for (int d = 0; d < 3; ++d) p.dims[d] = in.seg.dims[d];
  p.device = a.has("device") ? atoi(a.str("device").c_str()) : 0;
  std::vector<float> zeros(in.seg.size(), 0.0f);  // no image statistics are needed
  in.pb = zeros.data();
  if (a.has("maskImage") && !a.str("maskImage").empty()) {
    in.mask = gh::read_label_image(a.str("maskImage"));
    if (in.mask.size() != in.seg.size()) gh::perr("Error: mask size mismatch");
    in.hasMask = true;
  }
  std::vector<uint32_t> order = gh::read_order(a.str("mergeOrder"));
  const size_t n = order.size() / 3;
  gh::Handle H;
  gh::build_graph(in, H);
  std::vector<int32_t> labels(n ? n : 1);
  H.check(glia_b200_bc_label(H.h, truth.u32.data(), order.data(), (int64_t)n, a.flag("f1", true) ? 1 : 0, a.flag("optSplit") ? 1 : 0,
                             globalOpt, a.flag("tweak") ? 1 : 0, a.has("mpd") ? atof(a.str("mpd").c_str()) : 1.0, labels.data(),
                             nullptr, nullptr), "glia_b200_bc_label");
  if (a.has("bclabel") && !a.str("bclabel").empty()) {
    std::ofstream os(a.str("bclabel"));
    if (!os) gh::perr("Error: cannot write file " + a.str("bclabel"));
    for (size_t i = 0; i < n; ++i) os << labels[i] << "\n";  // writeData(file, labels, "\n")
  }
  return EXIT_SUCCESS;
}
