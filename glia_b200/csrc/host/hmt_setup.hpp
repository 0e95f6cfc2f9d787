// Shared set-up of the hmt programs: the image options of hmt/main_merge_order_bc.cxx:172-242 and
// hmt/main_bc_feat.cxx:125-185, prepareImages (hmt/hmt_util.hxx:17-56), handle creation + build.
#pragma once
#include <memory>
#include "../../../include/glia_b200.h"
#include "host_util.hpp"

namespace gh {

struct HmtInputs {
  Image seg, mask;
  bool hasMask = false;
  std::map<std::string, std::shared_ptr<Image>> images;  // by file name: identical files are loaded once
  const float* pb = nullptr;
  glia_params prm;
};

inline const float* load_real(HmtInputs& in, const std::string& file, const Image& seg) {
  auto it = in.images.find(file);
  if (it == in.images.end()) {
    auto im = std::make_shared<Image>(read_real_image(file));
    if (im->ndim != seg.ndim || im->size() != seg.size()) perr("Error: image size mismatch: " + file);
    it = in.images.emplace(file, im).first;
  }
  return it->second->f32.data();
}

inline void add_alias_common(Args& a) {
  a.add_alias("s", "segImage"); a.add_alias("m", "maskImage"); a.add_alias("n", "ns"); a.add_alias("l", "logs");
  a.add_alias("o", "mergeOrder"); a.add_alias("b", "bfeat"); a.add_alias("y", "saliency");
}

// fills in.prm (images, histogram specs, thresholds, flags) exactly as prepareImages orders them:
// rImages = rbi then ri; bImages = rbi then bi; rlImages = rli.
inline void setup_inputs(const Args& a, HmtInputs& in) {
  a.require("segImage"); a.require("pb");
  in.seg = read_label_image(a.str("segImage"));
  glia_params& p = in.prm;
  memset(&p, 0, sizeof(p));
  p.ndim = in.seg.ndim;
  for (int d = 0; d < 3; ++d) p.dims[d] = in.seg.dims[d];
  struct Group { const char *files, *bins, *lo, *hi; };
  const Group rb{"rbi", "rbb", "rbl", "rbu"}, rl{"rli", "rlb", "rll", "rlu"}, r{"ri", "rb", "rl", "ru"}, b{"bi", "bb", "bl", "bu"};
  auto add = [&](const Group& g, glia_image_spec* dst, int& n, glia_image_spec* dst2, int* n2) {
    auto files = a.strs(g.files); auto bins = a.nums(g.bins); auto lo = a.nums(g.lo); auto hi = a.nums(g.hi);
    if (bins.size() < files.size() || lo.size() < files.size() || hi.size() < files.size())
      perr(std::string("Error: --") + g.files + " needs one --" + g.bins + "/--" + g.lo + "/--" + g.hi + " per image");
    for (size_t i = 0; i < files.size(); ++i) {
      if (n >= GLIA_MAX_IMG || (dst2 && *n2 >= GLIA_MAX_IMG)) perr("Error: too many images");
      glia_image_spec s; s.data = load_real(in, files[i], in.seg); s.bins = (int)bins[i]; s.lo = lo[i]; s.hi = hi[i];
      dst[n++] = s;
      if (dst2) dst2[(*n2)++] = s;
    }
  };
  add(rb, p.rimg, p.n_rimg, p.bimg, &p.n_bimg);
  add(r, p.rimg, p.n_rimg, nullptr, nullptr);
  add(b, p.bimg, p.n_bimg, nullptr, nullptr);
  add(rl, p.rlimg, p.n_rlimg, nullptr, nullptr);
  in.pb = load_real(in, a.str("pb"), in.seg);
  if (a.has("maskImage") && !a.str("maskImage").empty()) {
    in.mask = read_label_image(a.str("maskImage"));
    if (in.mask.size() != in.seg.size()) perr("Error: mask size mismatch");
    in.hasMask = true;
  }
  auto bt = a.nums("bt");
  if (bt.size() > GLIA_MAX_THR) perr("Error: too many --bt thresholds");
  p.n_thr = (int)bt.size();
  for (size_t i = 0; i < bt.size(); ++i) p.thr[i] = bt[i];
  p.normalize = a.flag("ns"); p.use_log = a.flag("logs"); p.simple = a.flag("simpf");
  // the reference's CMake options GLIA_HMT_HIST_FEAT / GLIA_HMT_MEDIAN_FEAT (CMakeLists.txt:53-64) as run-time flags
  p.use_hist_feat = a.flag("histf"); p.use_median_feat = a.flag("medf");
  p.device = a.has("device") ? atoi(a.str("device").c_str()) : 0;
}

struct Handle {
  glia_handle* h = nullptr;
  ~Handle() { if (h) glia_b200_destroy(h); }
  void check(int st, const char* what) {
    if (st != GLIA_OK) perr(std::string("Error: ") + what + " failed (" + std::to_string(st) + "): " + (h ? glia_b200_last_error(h) : "no CUDA device / bad arguments"));
  }
};

inline void build_graph(HmtInputs& in, Handle& H) {
  H.check(glia_b200_create(&in.prm, &H.h), "glia_b200_create");
  H.check(glia_b200_build(H.h, in.seg.u32.data(), in.hasMask ? in.mask.u32.data() : nullptr, in.pb), "glia_b200_build");
}

}  // namespace gh
