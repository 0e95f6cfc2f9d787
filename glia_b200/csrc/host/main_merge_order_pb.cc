// merge_order_pb — drop-in for hmt/main_merge_order_pb.cxx: greedy merge order from the boundary probability alone
// (-t 1: approximate median [default], -t 2: mean), flags as the reference's (:44-60).
#include "hmt_setup.hpp"

int main(int argc, char** argv) {
  gh::Args a;
  a.add_alias("s", "segImage"); a.add_alias("p", "pbImage"); a.add_alias("m", "maskImage"); a.add_alias("t", "type");
  a.add_alias("o", "mergeOrder"); a.add_alias("y", "saliency");
  if (!a.parse(argc, argv) || a.has("help")) {
    std::cerr << "Usage: merge_order_pb -s SEG -p PB [-m MASK] [-t 1|2] [-o ORDER] [-y SAL] [--device N]\n";
    return a.has("help") ? EXIT_SUCCESS : EXIT_FAILURE;
  }
  a.require("segImage"); a.require("pbImage");
  const int type = a.has("type") ? atoi(a.str("type").c_str()) : 1;
  if (type != 1 && type != 2) gh::perr("Error: unsupported boundary stats type...");
  gh::HmtInputs in;
  in.seg = gh::read_label_image(a.str("segImage"));
  glia_params& p = in.prm;
  memset(&p, 0, sizeof(p));
  p.ndim = in.seg.ndim;
  for (int d = 0; d < 3; ++d) p.dims[d] = in.seg.dims[d];
  in.pb = gh::load_real(in, a.str("pbImage"), in.seg);
  p.n_bimg = 1; p.bimg[0].data = in.pb; p.bimg[0].bins = 1; p.bimg[0].lo = 0.0; p.bimg[0].hi = 1.0;
  p.contour_only = 1;                 // RegionMap rmap(segImage, mask, true), main_merge_order_pb.cxx:23
  p.use_median_feat = type == 1;
  p.device = a.has("device") ? atoi(a.str("device").c_str()) : 0;
  if (a.has("maskImage") && !a.str("maskImage").empty()) {
    in.mask = gh::read_label_image(a.str("maskImage"));
    if (in.mask.size() != in.seg.size()) gh::perr("Error: mask size mismatch");
    in.hasMask = true;
  }
  gh::Handle H;
  gh::build_graph(in, H);
  int64_t R = 0, n = 0;
  H.check(glia_b200_counts(H.h, &R, nullptr, nullptr), "glia_b200_counts");
  std::vector<uint32_t> order(3 * (size_t)std::max<int64_t>(R, 1));
  std::vector<double> sal((size_t)std::max<int64_t>(R, 1));
  H.check(glia_b200_merge_order_pb(H.h, type, order.data(), sal.data(), &n), "glia_b200_merge_order_pb");
  if (a.has("mergeOrder") && !a.str("mergeOrder").empty()) gh::write_order(a.str("mergeOrder"), order.data(), (size_t)n);
  if (a.has("saliency") && !a.str("saliency").empty()) gh::write_values(a.str("saliency"), sal.data(), (size_t)n, 6);
  return EXIT_SUCCESS;
}
