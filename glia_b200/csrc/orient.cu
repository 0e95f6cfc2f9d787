// Equal-area candidates: which region becomes x1.
//
// fBcFeat puts the smaller region first and swaps on a STRICT ">" only (hmt/main_merge_order_bc.cxx:76-80), so
// for two regions of equal area x1 is whichever region TBoundaryTable::init passed first: the one its
// range-for over the region map reaches first (type/boundary_table.hxx:92-107).  That map is a
// std::unordered_map<Label, Region> filled in the iteration order of a second unordered_map (the point map,
// type/region_map.hxx:79-95) which was filled in the iteration order of a third (genPointMap's count map,
// util/struct.hxx:61-92, keyed in raster order of first appearance).  The order is a property of the C++
// library, not of the algorithm, so the device path uses the deterministic rule "x1 = smaller label" unless the
// caller installs orientation bits (glia_b200_set_tie_orientation).  glia_b200_reference_orientation computes
// the bits the reference built with THIS toolchain would produce, by running the same three containers with
// the same keys in the same order - including the insertion of the scratch key BG_VAL = 0 during the loop
// (util/struct_merge_bc.hxx:19-20), which can rehash the map in mid-iteration.
#include <algorithm>
#include <vector>
#include "handle.cuh"
#include "host/ref_order.hpp"

namespace glia {

__global__ void first_voxel_kernel(const uint32_t* __restrict__ labels, const uint32_t* __restrict__ mask,
                                   size_t n, uint32_t labelBound, uint32_t* __restrict__ first) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    if (mask && mask[i] == 0) continue;
    const uint32_t l = labels[i];
    if (l < labelBound && first[l] > (uint32_t)i) atomicMin(first + l, (uint32_t)i);
  }
}

// tie[c] = 1 when both ends of candidate c have the same voxel count
__global__ void candidate_ties_kernel(const int* __restrict__ cand, const int* __restrict__ ea, const int* __restrict__ eb,
                                      long long C, const uint32_t* __restrict__ cAdd, int nAdd,
                                      unsigned char* __restrict__ tie, unsigned long long* __restrict__ count) {
  long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (c >= C) return;
  const int e = cand[c];
  const bool t = cAdd[(size_t)ea[e] * nAdd] == cAdd[(size_t)eb[e] * nAdd];
  tie[c] = t ? 1 : 0;
  if (t) atomicAdd(count, 1ull);
}

}  // namespace glia

using namespace glia;
#define CK(expr) GLIA_CUDA_CHECK(h, expr)

extern "C" int glia_b200_candidate_ties(glia_handle* h, uint8_t* tie, int64_t* n_ties) {
  if (!h) return GLIA_E_ARG;
  if (!h->built) return GLIA_E_STATE;
  CK(cudaSetDevice(h->prm.device));
  const size_t C = (size_t)h->C;
  if (n_ties) *n_ties = 0;
  if (!C) return GLIA_OK;
  CK(h->tieBuf.ensure(C + 16));
  CK(h->scalars.ensure(64));
  CK(cudaMemsetAsync(h->scalars.p, 0, 8, h->stream));
  candidate_ties_kernel<<<(unsigned)((C + 255) / 256), 256, 0, h->stream>>>(
      h->cand.as<int>(), h->ea.as<int>(), h->eb.as<int>(), (long long)C, h->cAdd.as<uint32_t>(), h->vp.reg.nAdd,
      h->tieBuf.as<unsigned char>(), h->scalars.as<unsigned long long>());
  h->launches++;
  CK(cudaGetLastError());
  unsigned long long n = 0;
  CK(cudaMemcpyAsync(&n, h->scalars.p, 8, cudaMemcpyDeviceToHost, h->stream));
  if (tie) CK(cudaMemcpyAsync(tie, h->tieBuf.p, C, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  if (n_ties) *n_ties = (int64_t)n;
  return GLIA_OK;
}

extern "C" int glia_b200_set_tie_orientation(glia_handle* h, const uint8_t* second_first) {
  if (!h) return GLIA_E_ARG;
  if (!second_first) { h->hasOrient = false; return GLIA_OK; }
  if (!h->built) return GLIA_E_STATE;
  CK(cudaSetDevice(h->prm.device));
  const size_t C = (size_t)h->C;
  CK(h->orient.ensure(C + 16));
  if (C) CK(cudaMemcpyAsync(h->orient.p, second_first, C, cudaMemcpyHostToDevice, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  h->hasOrient = true;
  h->orientC = (long long)C;
  return GLIA_OK;
}

extern "C" int glia_b200_reference_orientation(glia_handle* h, int64_t reserve_hint, uint8_t* second_first) {
  if (!h || !second_first) return GLIA_E_ARG;
  if (!h->built || h->vp.zEnd - h->vp.zBegin != h->vp.Z) return GLIA_E_STATE;  // whole volumes only
  CK(cudaSetDevice(h->prm.device));
  const VoxParams& P = h->vp;
  const size_t n = h->nvoxBuf;
  if (n >= 0xFFFFFFFFull) { h->err = "reference_orientation: more than 2^32 voxels"; return GLIA_E_ARG; }
  const uint32_t LB = P.labelBound;
  const size_t R = (size_t)h->R, E = (size_t)h->E, C = (size_t)h->C;
  // first voxel of every label in raster order (x fastest): the key order of genCountMap's map
  CK(h->firstVox.ensure(4 * (size_t)LB + 16));
  CK(cudaMemsetAsync(h->firstVox.p, 0xFF, 4 * (size_t)LB, h->stream));
  first_voxel_kernel<<<148 * 8, 256, 0, h->stream>>>(P.labels, P.mask, n, LB, h->firstVox.as<uint32_t>());
  h->launches++;
  CK(cudaGetLastError());
  std::vector<uint32_t> first(LB), label(R);
  std::vector<int> row(R + 1), ea(E), eb(E), rev(E), cand(C);
  cudaStream_t st = h->stream;
  CK(cudaMemcpyAsync(first.data(), h->firstVox.p, 4 * (size_t)LB, cudaMemcpyDeviceToHost, st));
  if (R) CK(cudaMemcpyAsync(label.data(), h->regLabel.p, 4 * R, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(row.data(), h->row.p, 4 * (R + 1), cudaMemcpyDeviceToHost, st));
  if (E) {
    CK(cudaMemcpyAsync(ea.data(), h->ea.p, 4 * E, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(eb.data(), h->eb.p, 4 * E, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(rev.data(), h->rev.p, 4 * E, cudaMemcpyDeviceToHost, st));
  }
  if (C) CK(cudaMemcpyAsync(cand.data(), h->cand.p, 4 * C, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  // regions in order of first appearance
  std::vector<uint32_t> seq(R);
  for (size_t r = 0; r < R; ++r) seq[r] = (uint32_t)r;
  std::sort(seq.begin(), seq.end(), [&](uint32_t a, uint32_t b) { return first[label[a]] < first[label[b]]; });
  std::vector<uint32_t> seqLabel(R);
  std::vector<unsigned char> hasMutual(R, 0);
  for (size_t i = 0; i < R; ++i) {
    const uint32_t r = seq[i];
    seqLabel[i] = label[r];
    for (int e = row[r]; e < row[r + 1] && !hasMutual[i]; ++e) hasMutual[i] = rev[e] >= 0;
  }
  const std::vector<uint32_t> visit = glia_host::reference_visit_order(seqLabel, hasMutual, reserve_hint);
  std::vector<uint32_t> pos(R);
  for (size_t i = 0; i < R; ++i) pos[seq[i]] = visit[i];
  for (size_t c = 0; c < C; ++c) {
    const int e = cand[c];
    const uint32_t pa = pos[ea[e]], pb = pos[eb[e]];
    second_first[c] = pb < pa ? 1 : 0;  // regions the loop never reached (rehash) keep the default
  }
  return GLIA_OK;
}
