// Replay of the reference's region-map iteration order (host-only C++; used by orient.cu and by the CPU tests).
//
// TBoundaryTable::init (type/boundary_table.hxx:92-107) walks a std::unordered_map<Label, Region> that was filled
// in the iteration order of the point map (type/region_map.hxx:79-95), itself filled in the iteration order of
// genPointMap's count map (util/struct.hxx:61-92), whose keys arrive in raster order of first appearance.  The
// resulting order belongs to the C++ library; replaying the same containers with the same keys reproduces it for
// the libstdc++ this file is compiled with.  The mapped types do not influence the order.
#pragma once
#include <stddef.h>
#include <stdint.h>
#include <unordered_map>
#include <vector>

namespace glia_host {

// labels_by_first_appearance: the R region labels in raster order of their first unmasked voxel.
// has_mutual[i]: region i (same indexing) has at least one mutual boundary, i.e. its visit creates a table entry.
// reserve_hint > 0: rmap.reserve(reserve_hint) before the loop (harnesses that keep the loop rehash-free).
// Returns visit position per region (same indexing); 0xFFFFFFFF = the loop never reached it.
inline std::vector<uint32_t> reference_visit_order(const std::vector<uint32_t>& labels_by_first_appearance,
                                                   const std::vector<unsigned char>& has_mutual, long long reserve_hint) {
  const size_t R = labels_by_first_appearance.size();
  std::unordered_map<uint32_t, uint32_t> cmap;  // genCountMap: cmap[key] = 1 at first appearance
  for (uint32_t l : labels_by_first_appearance) cmap[l] = 1;
  std::unordered_map<uint32_t, uint32_t> pmap;  // genPointMap: pmap[cp.first] in cmap order
  for (auto const& cp : cmap) pmap[cp.first];
  std::unordered_map<uint32_t, uint32_t> index;
  index.reserve(R);
  for (size_t i = 0; i < R; ++i) index[labels_by_first_appearance[i]] = (uint32_t)i;
  std::unordered_map<uint32_t, uint32_t> rmap;  // TRegionMap::init: emplace in pmap order
  for (auto const& pp : pmap) rmap.emplace(pp.first, index[pp.first]);
  if (reserve_hint > 0) rmap.reserve((size_t)reserve_hint);
  std::vector<uint32_t> pos(R, 0xFFFFFFFFu);
  uint32_t visit = 0;
  bool scratch = false;  // initFb inserts the scratch region BG_VAL = 0 at the first table entry (struct_merge_bc.hxx:19-20)
  for (auto const& rp : rmap) {
    if (rp.first == 0 && scratch) continue;
    const uint32_t r = rp.second;
    if (pos[r] == 0xFFFFFFFFu) pos[r] = visit++;
    if (!scratch && has_mutual[r]) { rmap.erase(0u); rmap.emplace(0u, 0u); scratch = true; }
  }
  return pos;
}

}  // namespace glia_host
