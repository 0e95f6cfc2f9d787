// K2: from the raw tables of the voxel pass to the finalised region adjacency graph.
//
// Replaces (reference, /root/reference/code):
//   boundaryMap[(a,b)].push_back + TPointPairMap::prepare   util/struct.hxx:104, type/point_map.hxx:53-60
//   TRegionMap::init (adjacency lists a -> [(b, pts*)])      type/region_map.hxx:79-95
//   mutual-pair filter of TBoundaryTable::init               type/boundary_table.hxx:99-102
// Regions become dense indices in ascending label order (prefix sum over label presence);
// directed boundaries are compacted out of the hash table, sorted by (a,b) index pair
// (64-bit radix sort-by-key) into a CSR edge list with reverse-edge links; mutual pairs with
// a < b are the initial merge candidates, already in the reference's table order.
#include <cub/cub.cuh>
#include "handle.cuh"

namespace glia {

__global__ void mark_present_kernel(const uint32_t* __restrict__ regAdd, int nAdd, uint32_t n,
                                    uint32_t* __restrict__ flags) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) flags[i] = regAdd[(size_t)i * nAdd] > 0 ? 1u : 0u;
}

// one warp per label: copy the three record parts to the compact slot
__global__ void gather_regions_kernel(const uint32_t* __restrict__ flags,
                                      const uint32_t* __restrict__ rank, uint32_t n, RecLayout L,
                                      const uint32_t* __restrict__ gAdd,
                                      const uint32_t* __restrict__ gMax,
                                      const double* __restrict__ gF64, uint32_t* __restrict__ label,
                                      uint32_t* __restrict__ cAdd, uint32_t* __restrict__ cMax,
                                      double* __restrict__ cF64, const unsigned char* __restrict__ interest,
                                      int world, int me) {
  uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (w >= n || !flags[w]) return;
  uint32_t r = rank[w];
  if (lane == 0) label[r] = w;
  // sharded build: only the regions this rank featurises (its own and their neighbours) need their records here
  if (interest && !interest[(size_t)r * world + me]) return;
  for (int i = lane; i < L.nAdd; i += 32) cAdd[(size_t)r * L.nAdd + i] = gAdd[(size_t)w * L.nAdd + i];
  for (int i = lane; i < L.nMax; i += 32) cMax[(size_t)r * L.nMax + i] = gMax[(size_t)w * L.nMax + i];
  for (int i = lane; i < L.nF64; i += 32) cF64[(size_t)r * L.nF64 + i] = gF64[(size_t)w * L.nF64 + i];
}

// occupied hash slots -> (index-pair key, slot).  A CTA takes 1024 consecutive slots (4 coalesced rounds of 256), counts
// its occupied ones with ballots and a 32-entry scan, and appends them with ONE atomicAdd: a warp-aggregated append put
// one atomic per 32 slots on the same address (1 M at C4's 2^25 slots, ~0.5 ms of same-address serialisation in L2).
#define COMPACT_ROUNDS 4
__global__ void __launch_bounds__(256)
compact_edges_kernel(const unsigned long long* __restrict__ keys, uint64_t cap,
                     const uint32_t* __restrict__ rank,
                     const uint32_t* __restrict__ flags, uint32_t labelBound,
                     unsigned long long* __restrict__ outKey,
                     uint32_t* __restrict__ outSlot, unsigned long long* counter,
                     int* errFlag, int shift) {
  __shared__ uint32_t cnt[COMPACT_ROUNDS * 8];
  __shared__ unsigned long long blockBase;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint64_t first = blockIdx.x * (uint64_t)(256 * COMPACT_ROUNDS) + threadIdx.x;
  unsigned long long k[COMPACT_ROUNDS];
  unsigned m[COMPACT_ROUNDS];
#pragma unroll
  for (int j = 0; j < COMPACT_ROUNDS; ++j) {
    const uint64_t i = first + 256u * j;
    k[j] = (i < cap) ? keys[i] : EDGE_EMPTY;
    m[j] = __ballot_sync(0xffffffffu, k[j] != EDGE_EMPTY);
    if (lane == 0) cnt[j * 8 + warp] = __popc(m[j]);
  }
  __syncthreads();
  if (warp == 0) {  // exclusive scan of the 32 counts, one append for the CTA
    const uint32_t c = cnt[lane];
    uint32_t inc = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { uint32_t v = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += v; }
    cnt[lane] = inc - c;
    if (lane == 31) blockBase = inc ? atomicAdd(counter, (unsigned long long)inc) : 0ull;
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < COMPACT_ROUNDS; ++j) {
    if (k[j] == EDGE_EMPTY) continue;
    const uint32_t a = (uint32_t)(k[j] >> 32), b = (uint32_t)k[j];
    if (a >= labelBound || b >= labelBound || !flags[a] || !flags[b]) { atomicExch(errFlag, 3); continue; }
    const unsigned long long pos = blockBase + cnt[j * 8 + warp] + __popc(m[j] & ((1u << lane) - 1));
    outKey[pos] = ((unsigned long long)rank[a] << shift) | rank[b];  // 2 * shift key bits: fewer radix passes
    outSlot[pos] = (uint32_t)(first + 256u * j);
  }
}

__global__ void gather_edges_kernel(const unsigned long long* __restrict__ keySorted,
                                    const uint32_t* __restrict__ slotSorted, long long E, RecLayout L,
                                    const uint32_t* __restrict__ gAdd,
                                    const uint32_t* __restrict__ gMax,
                                    const double* __restrict__ gF64, int* __restrict__ ea,
                                    int* __restrict__ eb, uint32_t* __restrict__ sAdd,
                                    uint32_t* __restrict__ sMax, double* __restrict__ sF64, int shift) {
  long long w = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  int lane = threadIdx.x & 31;
  if (w >= E) return;
  size_t s = slotSorted[w];
  if (lane == 0) { unsigned long long k = keySorted[w]; ea[w] = (int)(k >> shift); eb[w] = (int)(k & ((1ull << shift) - 1)); }
  for (int i = lane; i < L.nAdd; i += 32) sAdd[(size_t)w * L.nAdd + i] = gAdd[s * L.nAdd + i];
  for (int i = lane; i < L.nMax; i += 32) sMax[(size_t)w * L.nMax + i] = gMax[s * L.nMax + i];
  for (int i = lane; i < L.nF64; i += 32) sF64[(size_t)w * L.nF64 + i] = gF64[s * L.nF64 + i];
}

// Vectorised variant for record layouts made of whole 16-byte units (default hmt layout: 20 add words,
// 2 max words, 2 doubles): 8 lanes move one record with uint4 / uint2 / double2 accesses, 4 records per
// warp, so four times as many bytes are in flight per warp as in the word-per-lane kernel above (which
// ran at 0.7 TB/s on the 1.6 GB of scattered records of C4).
__global__ void gather_edges_vec_kernel(const unsigned long long* __restrict__ keySorted,
                                        const uint32_t* __restrict__ slotSorted, long long E, RecLayout L,
                                        const uint32_t* __restrict__ gAdd, const uint32_t* __restrict__ gMax,
                                        const double* __restrict__ gF64, int* __restrict__ ea,
                                        int* __restrict__ eb, uint32_t* __restrict__ sAdd,
                                        uint32_t* __restrict__ sMax, double* __restrict__ sF64, int shift) {
  const long long w = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 3;
  const int sub = threadIdx.x & 7;
  if (w >= E) return;
  const size_t s = slotSorted[w];
  const int nA4 = L.nAdd >> 2, nM2 = L.nMax >> 1, nF2 = L.nF64 >> 1;
  const uint4* a4 = (const uint4*)(gAdd + s * L.nAdd);
  uint4* o4 = (uint4*)(sAdd + (size_t)w * L.nAdd);
  for (int i = sub; i < nA4; i += 8) o4[i] = a4[i];
  const uint2* m2 = (const uint2*)(gMax + s * L.nMax);
  uint2* om = (uint2*)(sMax + (size_t)w * L.nMax);
  for (int i = sub; i < nM2; i += 8) om[i] = m2[i];
  const double2* f2 = (const double2*)(gF64 + s * L.nF64);
  double2* of = (double2*)(sF64 + (size_t)w * L.nF64);
  for (int i = 7 - sub; i < nF2; i += 8) of[i] = f2[i];
  if (sub == 6) { unsigned long long k = keySorted[w]; ea[w] = (int)(k >> shift); eb[w] = (int)(k & ((1ull << shift) - 1)); }
}

// row[r] = first edge with a >= r (r in 0..R)
__global__ void csr_rows_kernel(const int* __restrict__ ea, long long E, long long R,
                                int* __restrict__ row) {
  long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (r > R) return;
  long long lo = 0, hi = E;
  while (lo < hi) { long long m = (lo + hi) >> 1; if (ea[m] < r) lo = m + 1; else hi = m; }
  row[r] = (int)lo;
}

__global__ void reverse_edges_kernel(const int* __restrict__ ea, const int* __restrict__ eb,
                                     const int* __restrict__ row, long long E,
                                     int* __restrict__ rev) {
  long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (e >= E) return;
  int a = ea[e], b = eb[e];
  int lo = row[b], hi = row[b + 1], r = -1;
  while (lo < hi) { int m = (lo + hi) >> 1; int v = eb[m]; if (v == a) { r = m; break; } if (v < a) lo = m + 1; else hi = m; }
  rev[e] = r;
}

struct IsCandidate {
  const int* ea; const int* eb; const int* rev;
  int lo, hi;  // owned region-index range (multi-GPU: a candidate belongs to the owner of its smaller end)
  __device__ bool operator()(int e) const { int a = ea[e]; return a < eb[e] && rev[e] >= 0 && a >= lo && a < hi; }
};

// Per-region aggregate over all of its directed boundaries (entries of a base region):
// one warp per region, word-parallel.
__global__ void boundary_aggregate_kernel(const int* __restrict__ row, long long R, RecLayout L,
                                          const uint32_t* __restrict__ sAdd,
                                          const uint32_t* __restrict__ sMax,
                                          const double* __restrict__ sF64,
                                          uint32_t* __restrict__ aAdd, uint32_t* __restrict__ aMax,
                                          double* __restrict__ aF64, uint32_t* __restrict__ aMax2,
                                          int* __restrict__ aOwn, long long lo, long long hi,
                                          const unsigned char* __restrict__ interest, int world, int me) {
  long long r = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  int lane = threadIdx.x & 31;
  if (r >= R) return;
  // sharded build: the boundaries of a region this rank neither owns nor neighbours are incomplete here and never read
  if (interest && !interest[(size_t)r * world + me]) return;
  int e0 = row[r], e1 = row[r + 1];
  (void)lo; (void)hi;  // every region whose boundaries are complete here (owned or neighbour of owned) is used
  // pointers stepped by the record width: the kernel is bound by instruction issue (65 %), not by its loads
  for (int i = lane; i < L.nAdd; i += 32) {
    uint32_t s = 0;
    const uint32_t* p = sAdd + (size_t)e0 * L.nAdd + i;
#pragma unroll 4
    for (int e = e0; e < e1; ++e, p += L.nAdd) s += *p;
    aAdd[(size_t)r * L.nAdd + i] = s;
  }
  for (int i = lane; i < L.nMax; i += 32) {  // best, its boundary, and the best of the others
    uint32_t b1 = 0, b2 = 0; int own = -1;
    const uint32_t* p = sMax + (size_t)e0 * L.nMax + i;
#pragma unroll 4
    for (int e = e0; e < e1; ++e, p += L.nMax) {
      const uint32_t v = *p;
      if (v > b1) { b2 = b1; b1 = v; own = e; } else if (v > b2) b2 = v;
    }
    aMax[(size_t)r * L.nMax + i] = b1;
    aMax2[(size_t)r * L.nMax + i] = b2;
    aOwn[(size_t)r * L.nMax + i] = own;
  }
  for (int i = lane; i < L.nF64; i += 32) {
    double s = 0.0;
    const double* p = sF64 + (size_t)e0 * L.nF64 + i;
#pragma unroll 4
    for (int e = e0; e < e1; ++e, p += L.nF64) s += *p;
    aF64[(size_t)r * L.nF64 + i] = s;
  }
}

}  // namespace glia

using namespace glia;

#define CK(expr) GLIA_CUDA_CHECK(h, expr)
#include <chrono>
#include <stdio.h>
#include <stdlib.h>
// GLIA_B200_TRACE=1: host-side wall clock of each finalisation phase (stream synchronised) on stderr
static inline unsigned nblk(long long n, int per) { return (unsigned)((n + per - 1) / per); }

int glia_finalize_regions(glia_handle* h);
int glia_finalize_edges(glia_handle* h, long long ownLo, long long ownHi);
int glia_gather_regions_some(glia_handle* h, const unsigned char* interest, int world, int me);

// Builds the finalised graph from the raw tables.  Returns a GLIA_* status.
int glia_finalize_graph(glia_handle* h) {
  int s = glia_finalize_regions(h);
  if (s != GLIA_OK) return s;
  return glia_finalize_edges(h, 0, h->R);
}

// regions: presence -> dense index in ascending label order, compact records.  Three steps so that a multi-GPU
// caller can reduce the 4-byte presence flags first, index the regions and exchange the boundary records while
// the all-reduce of the (30x larger) record tables is still in flight (glia_b200/parallel.py).
int glia_region_presence(glia_handle* h) {
  cudaStream_t st = h->stream;
  const VoxParams& P = h->vp;
  const uint32_t LB = P.labelBound;
  CK(h->flags.ensure(4 * (size_t)LB + 4));
  mark_present_kernel<<<nblk(LB, 256), 256, 0, st>>>(h->regAdd.as<uint32_t>(), P.reg.nAdd, LB,
                                                     h->flags.as<uint32_t>());
  h->launches++;
  return cudaGetLastError() == cudaSuccess ? GLIA_OK : GLIA_E_CUDA;
}

int glia_index_regions(glia_handle* h) {
  cudaStream_t st = h->stream;
  PhaseTrace tr(st);
  const VoxParams& P = h->vp;
  const uint32_t LB = P.labelBound;
  // scalars: [0] edge counter (u64), [1] cand count (int as u64 slot), [2] err flag
  CK(h->scalars.ensure(64));
  CK(cudaMemsetAsync(h->scalars.p, 0, 64, st));
  CK(h->rank.ensure(4 * (size_t)LB + 4));
  size_t tmpBytes = 0;
  CK(cub::DeviceScan::ExclusiveSum(nullptr, tmpBytes, h->flags.as<uint32_t>(), h->rank.as<uint32_t>(), (int)LB, st));
  CK(h->cubTmp.ensure(tmpBytes));
  CK(cub::DeviceScan::ExclusiveSum(h->cubTmp.p, tmpBytes, h->flags.as<uint32_t>(), h->rank.as<uint32_t>(), (int)LB, st));
  h->launches += 2;
  uint32_t lastRank = 0, lastFlag = 0;
  CK(cudaMemcpyAsync(&lastRank, h->rank.as<uint32_t>() + (LB - 1), 4, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(&lastFlag, h->flags.as<uint32_t>() + (LB - 1), 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  h->R = (long long)lastRank + lastFlag;
  tr.mark("regions: scan+readback");
  return GLIA_OK;
}

int glia_gather_regions(glia_handle* h) { return glia_gather_regions_some(h, nullptr, 1, 0); }

// A handle whose region records were compacted for the regions of interest only gets the rest on demand (dumps, the
// bc_feat stage on a sharded build: they read every region).
int glia_ensure_full_regions(glia_handle* h) {
  if (!h->regionsPartial) return GLIA_OK;
  return glia_gather_regions(h);
}

int glia_gather_regions_some(glia_handle* h, const unsigned char* interest, int world, int me) {
  cudaStream_t st = h->stream;
  PhaseTrace tr(st);
  const VoxParams& P = h->vp;
  const uint32_t LB = P.labelBound;
  const long long R = h->R;
  CK(h->regLabel.ensure(4 * (size_t)(R + 1)));
  CK(h->cAdd.ensure(4 * (size_t)(R + 1) * P.reg.nAdd));
  CK(h->cMax.ensure(4 * (size_t)(R + 1) * P.reg.nMax));
  CK(h->cF64.ensure(8 * (size_t)(R + 1) * P.reg.nF64));
  gather_regions_kernel<<<nblk((long long)LB * 32, 256), 256, 0, st>>>(
      h->flags.as<uint32_t>(), h->rank.as<uint32_t>(), LB, P.reg, h->regAdd.as<uint32_t>(),
      h->regMax.as<uint32_t>(), h->regF64.as<double>(), h->regLabel.as<uint32_t>(),
      h->cAdd.as<uint32_t>(), h->cMax.as<uint32_t>(), h->cF64.as<double>(), interest, world, me);
  h->launches++;
  h->regionsPartial = interest != nullptr;
  CK(cudaGetLastError());
  tr.mark("regions: gather");
  return GLIA_OK;
}

int glia_finalize_regions(glia_handle* h) {
  int s = glia_region_presence(h);
  if (s == GLIA_OK) s = glia_index_regions(h);
  if (s == GLIA_OK) s = glia_gather_regions(h);
  return s;
}

// directed boundaries of the local table -> CSR, reverse links, candidates whose smaller end lies
// in [ownLo, ownHi), per-region boundary aggregate (identity outside the owned range)
int glia_finalize_edges(glia_handle* h, long long ownLo, long long ownHi) {
  cudaStream_t st = h->stream;
  const VoxParams& P = h->vp;
  const uint32_t LB = P.labelBound;
  const long long R = h->R;
  h->ownLo = ownLo; h->ownHi = ownHi;
  size_t tmpBytes = 0;
  PhaseTrace tr(st);
  CK(cudaMemsetAsync(h->scalars.p, 0, 64, st));
  unsigned long long* dCounter = h->scalars.as<unsigned long long>();
  int* dNumSel = (int*)(dCounter + 1);
  int* dErr = (int*)(dCounter + 2);
  // ---- directed boundaries: compact, sort by (a,b) index pair
  // Upper bound on E is not known before compaction: count first with a null-output pass
  // would cost a second sweep, so size the outputs by the table capacity bound instead.
  size_t eCap = (size_t)P.edgeCap;
  int rbits = 1; while ((1ll << rbits) < R + 1) ++rbits;  // key = index pair in 2 * rbits bits (<= 64: R < 2^32)
  CK(h->eKey.ensure(8 * eCap));
  CK(h->eSlot.ensure(4 * eCap));
  compact_edges_kernel<<<nblk((long long)eCap, 256 * COMPACT_ROUNDS), 256, 0, st>>>(
      (const unsigned long long*)h->edgeKeys.p, P.edgeCap, h->rank.as<uint32_t>(),
      h->flags.as<uint32_t>(), LB, h->eKey.as<unsigned long long>(), h->eSlot.as<uint32_t>(),
      dCounter, dErr, rbits);
  h->launches++;
  unsigned long long E64 = 0; int errv = 0;
  CK(cudaMemcpyAsync(&E64, dCounter, 8, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(&errv, dErr, 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  if (errv) { h->err = "directed boundary refers to a label with no voxels (reduce the partial tables first)"; return GLIA_E_STATE; }
  if (E64 > 0x7fffffffull) { h->err = "more than 2^31 directed boundaries"; return GLIA_E_ARG; }
  h->E = (long long)E64;
  const long long E = h->E;
  tr.mark("edges: compact+readback");
  CK(h->eKeySorted.ensure(8 * (size_t)(E + 1)));
  CK(h->eSlotSorted.ensure(4 * (size_t)(E + 1)));
  const int endBit = 2 * rbits > 64 ? 64 : 2 * rbits;
  if (E > 0) {
    tmpBytes = 0;
    CK(cub::DeviceRadixSort::SortPairs(nullptr, tmpBytes, h->eKey.as<unsigned long long>(),
                                       h->eKeySorted.as<unsigned long long>(), h->eSlot.as<uint32_t>(),
                                       h->eSlotSorted.as<uint32_t>(), (int)E, 0, endBit, st));
    CK(h->cubTmp.ensure(tmpBytes));
    CK(cub::DeviceRadixSort::SortPairs(h->cubTmp.p, tmpBytes, h->eKey.as<unsigned long long>(),
                                       h->eKeySorted.as<unsigned long long>(), h->eSlot.as<uint32_t>(),
                                       h->eSlotSorted.as<uint32_t>(), (int)E, 0, endBit, st));
    h->launches += 8;
  }
  tr.mark("edges: radix sort");
  CK(h->ea.ensure(4 * (size_t)(E + 1)));
  CK(h->eb.ensure(4 * (size_t)(E + 1)));
  CK(h->rev.ensure(4 * (size_t)(E + 1)));
  CK(h->row.ensure(4 * (size_t)(R + 2)));
  CK(h->sAdd.ensure(4 * (size_t)(E + 1) * P.edge.nAdd));
  CK(h->sMax.ensure(4 * (size_t)(E + 1) * (P.edge.nMax + 1)));
  CK(h->sF64.ensure(8 * (size_t)(E + 1) * (P.edge.nF64 + 1)));
  const bool vecLayout = (P.edge.nAdd % 4 == 0) && (P.edge.nMax % 2 == 0) && (P.edge.nF64 % 2 == 0) && P.edge.nMax > 0 &&
                         P.edge.nF64 > 0;
  if (E > 0 && vecLayout) {
    gather_edges_vec_kernel<<<nblk(E * 8, 256), 256, 0, st>>>(
        h->eKeySorted.as<unsigned long long>(), h->eSlotSorted.as<uint32_t>(), E, P.edge,
        h->edgeAdd.as<uint32_t>(), h->edgeMax.as<uint32_t>(), h->edgeF64.as<double>(),
        h->ea.as<int>(), h->eb.as<int>(), h->sAdd.as<uint32_t>(), h->sMax.as<uint32_t>(),
        h->sF64.as<double>(), rbits);
    h->launches++;
  } else if (E > 0) {
    gather_edges_kernel<<<nblk(E * 32, 256), 256, 0, st>>>(
        h->eKeySorted.as<unsigned long long>(), h->eSlotSorted.as<uint32_t>(), E, P.edge,
        h->edgeAdd.as<uint32_t>(), h->edgeMax.as<uint32_t>(), h->edgeF64.as<double>(),
        h->ea.as<int>(), h->eb.as<int>(), h->sAdd.as<uint32_t>(), h->sMax.as<uint32_t>(),
        h->sF64.as<double>(), rbits);
    h->launches++;
  }
  tr.mark("edges: gather records");
  csr_rows_kernel<<<nblk(R + 1, 256), 256, 0, st>>>(h->ea.as<int>(), E, R, h->row.as<int>());
  h->launches++;
  CK(h->cand.ensure(4 * (size_t)(E + 1)));
  h->C = 0;
  if (E > 0) {
    reverse_edges_kernel<<<nblk(E, 256), 256, 0, st>>>(h->ea.as<int>(), h->eb.as<int>(),
                                                       h->row.as<int>(), E, h->rev.as<int>());
    h->launches++;
    IsCandidate pred{h->ea.as<int>(), h->eb.as<int>(), h->rev.as<int>(), (int)ownLo, (int)ownHi};
    cub::CountingInputIterator<int> it(0);
    tmpBytes = 0;
    CK(cub::DeviceSelect::If(nullptr, tmpBytes, it, h->cand.as<int>(), dNumSel, (int)E, pred, st));
    CK(h->cubTmp.ensure(tmpBytes));
    CK(cub::DeviceSelect::If(h->cubTmp.p, tmpBytes, it, h->cand.as<int>(), dNumSel, (int)E, pred, st));
    h->launches += 2;
    int c = 0;
    CK(cudaMemcpyAsync(&c, dNumSel, 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    h->C = c;
  }
  tr.mark("edges: csr+rev+select");
  // ---- per-region aggregate over its boundaries
  CK(h->bAggAdd.ensure(4 * (size_t)(R + 1) * P.edge.nAdd));
  CK(h->bAggMax.ensure(4 * (size_t)(R + 1) * (P.edge.nMax + 1)));
  CK(h->bAggF64.ensure(8 * (size_t)(R + 1) * (P.edge.nF64 + 1)));
  CK(h->bAggMax2.ensure(4 * (size_t)(R + 1) * (P.edge.nMax + 1)));
  CK(h->bAggOwn.ensure(4 * (size_t)(R + 1) * (P.edge.nMax + 1)));
  if (R > 0) {
    boundary_aggregate_kernel<<<nblk(R * 32, 256), 256, 0, st>>>(
        h->row.as<int>(), R, P.edge, h->sAdd.as<uint32_t>(), h->sMax.as<uint32_t>(),
        h->sF64.as<double>(), h->bAggAdd.as<uint32_t>(), h->bAggMax.as<uint32_t>(),
        h->bAggF64.as<double>(), h->bAggMax2.as<uint32_t>(), h->bAggOwn.as<int>(), ownLo, ownHi,
        h->shardWorld > 1 ? h->interest.as<unsigned char>() : nullptr, h->shardWorld, h->shardMe);
    h->launches++;
  }
  CK(cudaGetLastError());
  tr.mark("edges: boundary aggregate");
  return GLIA_OK;
}
