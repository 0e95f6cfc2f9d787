// pre_merge — drop-in for gadget/main_pre_merge.cxx: merge small superpixels (and small ones lying on strong
// boundary probability) into their best neighbour, write the renamed label image.  Flags as the reference's
// (:95-118); -r (ITK RelabelComponentImageFilter) is not supported, -z (compression) is ignored.
#include "hmt_setup.hpp"

int main(int argc, char** argv) {
  gh::Args a;
  a.add_alias("s", "segImage"); a.add_alias("p", "pbImage"); a.add_alias("m", "maskImage"); a.add_alias("t", "sizeThreshold");
  a.add_alias("b", "rpbThreshold"); a.add_alias("r", "relabel"); a.add_alias("u", "write16"); a.add_alias("z", "compress");
  a.add_alias("o", "outputImage");
  if (!a.parse(argc, argv) || a.has("help")) {
    std::cerr << "Usage: pre_merge -s SEG -p PB [-m MASK] -t SIZE [SIZE2] [-b RPB] [-u 0|1] -o OUT [--device N]\n";
    return a.has("help") ? EXIT_SUCCESS : EXIT_FAILURE;
  }
  a.require("segImage"); a.require("pbImage"); a.require("sizeThreshold"); a.require("outputImage");
  if (a.flag("relabel")) gh::perr("Error: -r (ITK relabel) is not supported by this build");
  std::vector<int> thr;
  for (double v : a.nums("sizeThreshold")) thr.push_back((int)v);
  const double rpb = a.has("rpbThreshold") ? atof(a.str("rpbThreshold").c_str()) : 0.0;
  gh::HmtInputs in;
  in.seg = gh::read_label_image(a.str("segImage"));
  glia_params& p = in.prm;
  memset(&p, 0, sizeof(p));
  p.ndim = in.seg.ndim;
  for (int d = 0; d < 3; ++d) p.dims[d] = in.seg.dims[d];
  in.pb = gh::load_real(in, a.str("pbImage"), in.seg);
  p.n_rimg = 1; p.rimg[0].data = in.pb; p.rimg[0].bins = 1; p.rimg[0].lo = 0.0; p.rimg[0].hi = 1.0;
  p.n_bimg = 1; p.bimg[0] = p.rimg[0];
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
  std::vector<uint32_t> order(3 * (size_t)std::max<int64_t>(R, 1)), out(in.seg.size());
  std::vector<double> sal((size_t)std::max<int64_t>(R, 1));
  H.check(glia_b200_pre_merge(H.h, thr.data(), (int)thr.size(), rpb, order.data(), sal.data(), &n, out.data()), "glia_b200_pre_merge");
  gh::write_label_image(a.str("outputImage"), in.seg, out.data(), a.flag("write16"));
  return EXIT_SUCCESS;
}
