// EXPERIMENTAL (GLIA_B200_VOXEL=queue; parity-green, 5.24 ms vs 3.40 ms for v3 at 512^3).  v5b: two typed
// queues per warp, drains of exactly 32 runs, adjacent equal keys combined by segmented shuffles so that only
// segment heads touch the tables.  It halves the shared atomics of v3 but needs 126 registers (2 CTAs/SM) and
// 714 warp-instructions per 32 voxels; profiles/r1_v4_notes.txt has the numbers and what round 2 should keep.
// K1+K3+K4 fused, v5 "run queue" (default hmt configuration: one image = the boundary probability
// map as region image, boundary image and threshold source; <= 16 bins, <= 3 thresholds, no mask).
// Same results as voxel_pass_generic / voxel_pass_fast; reference semantics as cited there
// (type/neighbor.hxx:72-126, util/struct.hxx:95-144, type/feat.hxx:71-90, :485-502, :632-638,
// :706-738, util/image_stats.hxx:12-36).
//
// v3 (voxel_fast.cu) issues 596 warp-instructions per 32 voxels, most of them per-voxel hashing and
// shared atomics for the integer statistics (profiles/r1_v3_ncu_c4s.txt).  Here a thread keeps ALL
// statistics of its current run of equal label / equal directed pair in registers (count, border,
// 16 x 4-bit histogram, threshold counts, sum, sum^2, min, max, z-range); per voxel there is no
// hashing and no atomic.  A finished run (label change, or 15 voxels) is appended to the warp's queue
// with a ballot-computed position (no atomic, queue lengths are warp-uniform registers); when a
// queue holds >= 32 runs the whole warp drains 32 of them, one run per lane: segmented combine, then
// slot lookup in the CTA's shared-memory tables and a handful of shared atomics per segment head.  Tables are flushed to the global
// records once per tile, as in v3.
#include "voxel_fast.cuh"

namespace glia {
namespace {

#define RQ_CAP 128     // per warp: 64 region runs + 64 boundary runs (drained 32 at a time; a step appends <= 32 each)
#define RQ_EDGE 0x80000000u

// one finished run.  Region: a = label, b = RQ_REGION; boundary: (a -> b)
struct __align__(16) RqEntry {
  uint32_t a, b;
  uint32_t cnt;        // region: voxels | border << 16 ; boundary: voxels | thr0 << 8 | thr1 << 16 | thr2 << 24
  uint32_t hlo, hhi;   // 16 bins x 4 bits
  uint32_t kmin, kmax; // ~ord(min), ord(max); 0 == none
  uint32_t aux;        // region: zminOff | zmaxOff << 8 | hasNonInterior << 16 | lane << 24
  double s, s2;
};
#define RQ_REGION 0xFFFFFFFFu

__device__ __forceinline__ int rq_bin(float v, float lo, float hi, float inv, int bins, const float* __restrict__ bound) {
  if (v > lo && v < hi) {  // util/image_stats.hxx:17-36
    int i = (int)(v * inv);
    i = i < 0 ? 0 : (i > bins - 1 ? bins - 1 : i);
    while (i > 0 && v < bound[i - 1]) --i;
    while (i < bins && !(v < bound[i])) ++i;
    return i < bins ? i : -1;
  }
  return (v <= lo) ? 0 : bins - 1;
}

__device__ __forceinline__ void rq_smax(uint32_t* w, uint32_t v) {
  if (*(volatile uint32_t*)w < v) atomicMax(w, v);
}

__device__ __forceinline__ long long rq_edge_slot_global(const FastParams& P, unsigned long long key) {
  const uint64_t mask = P.edgeCap - 1;
  uint64_t h = mix64(key) & mask;
  for (int probe = 0; probe < 8192; ++probe) {
    unsigned long long cur = P.edgeKeys[h];
    if (cur == key) return (long long)h;
    if (cur == EDGE_EMPTY) {
      unsigned long long prev = atomicCAS(P.edgeKeys + h, EDGE_EMPTY, key);
      if (prev == EDGE_EMPTY || prev == key) return (long long)h;
    }
    h = (h + 1) & mask;
  }
  atomicExch(P.overflow, 1);
  return -1;
}

// shared-memory tables of a CTA (same layout as v3):
//   region slot: key u32 | cnt = count | border << 16 | hist[8] (2 x 16-bit bins per word) |
//                sum, sum2 (double) | mx[8] = ~bbminx, ~bbminy, ~bbminz, bbmax+1 x3, ~min, max
//   edge slot  : key u64 | cnt[2] = {count | thr0 << 16, thr1 | thr2 << 16} | hist[8] |
//                sum, sum2 (double) | mx[2] = ~min, max
struct RqSmem {
  unsigned long long* ekeys; double* esum; double* rsum;
  RqEntry* queue;
  uint32_t *ecnt, *ehist, *emx, *rkeys, *rcnt, *rhist, *rmx;
  float *sRB, *sEB;
};

__device__ __forceinline__ RqSmem rq_carve(unsigned char* raw, const FastParams& P, int nwarps) {
  RqSmem S;
  unsigned char* p = raw;
  S.queue = (RqEntry*)p; p += sizeof(RqEntry) * RQ_CAP * nwarps;
  S.ekeys = (unsigned long long*)p; p += 8 * (size_t)P.edgeSlots;
  S.esum = (double*)p; p += 16 * (size_t)P.edgeSlots;
  S.rsum = (double*)p; p += 16 * (size_t)P.regSlots;
  S.ecnt = (uint32_t*)p; p += 8 * (size_t)P.edgeSlots;
  S.ehist = (uint32_t*)p; p += 32 * (size_t)P.edgeSlots;
  S.emx = (uint32_t*)p; p += 8 * (size_t)P.edgeSlots;
  S.rkeys = (uint32_t*)p; p += 4 * (size_t)P.regSlots;
  S.rcnt = (uint32_t*)p; p += 4 * (size_t)P.regSlots;
  S.rhist = (uint32_t*)p; p += 32 * (size_t)P.regSlots;
  S.rmx = (uint32_t*)p; p += 32 * (size_t)P.regSlots;
  S.sRB = (float*)p; p += 64;
  S.sEB = (float*)p;
  return S;
}

}  // namespace

size_t voxel_rq_smem_bytes(const FastParams& P) {
  int nwarps = P.warpsX * P.warpsY;
  return sizeof(RqEntry) * RQ_CAP * nwarps + (size_t)P.edgeSlots * (8 + 16 + 8 + 32 + 8) +
         (size_t)P.regSlots * (16 + 4 + 4 + 32 + 32) + 128 + 64;
}

namespace {

// ---- drain: 32 queued runs of one kind, one per lane.  Runs finished in the same step by neighbouring
// lanes sit next to each other in the queue and mostly carry the same key, so adjacent equal keys are
// combined first (5-step segmented shuffle reduction, profiles/r1_v4_notes.txt) and only segment heads
// look up a slot and touch the tables: no region/boundary branch, no second round for leftovers, no
// contended CAS loops on the doubles.
__device__ __forceinline__ uint32_t rq_segment(bool head, int lane, int& end) {
  const uint32_t heads = __ballot_sync(0xffffffffu, head);
  const uint32_t above = lane == 31 ? 0u : (heads & (0xffffffffu << (lane + 1)));
  end = above ? (__ffs(above) - 2) : 31;
  return heads;
}
#define RQ_SEG_BEGIN for (int d_ = 1; d_ < 32; d_ <<= 1) { const bool ok_ = lane + d_ <= end;
#define RQ_SEG_ADD(v) { auto t_ = __shfl_down_sync(0xffffffffu, v, d_); if (ok_) v += t_; }
#define RQ_SEG_MAX(v) { uint32_t t_ = __shfl_down_sync(0xffffffffu, v, d_); if (ok_) v = max(v, t_); }
#define RQ_SEG_END }

__device__ __forceinline__ void rq_widen(uint32_t hlo, uint32_t hhi, uint32_t (&w)[8]) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {  // 2 x 16-bit bins per word: the layout of the shared tables
    w[j] = ((hlo >> (8 * j)) & 0xFu) | (((hlo >> (8 * j + 4)) & 0xFu) << 16);
    w[4 + j] = ((hhi >> (8 * j)) & 0xFu) | (((hhi >> (8 * j + 4)) & 0xFu) << 16);
  }
}

__device__ __forceinline__ void rq_drain_regions(const FastParams& P, const RqSmem& S, const RqEntry* q, int n,
                                                 int lane, int x0warp, int gy, int z0) {
  __syncwarp();
  const bool valid = lane < n;
  RqEntry e;
  if (valid) e = q[lane]; else { e.a = 0; e.b = RQ_REGION; e.cnt = 0; e.hlo = e.hhi = 0; e.kmin = e.kmax = 0; e.aux = 0; e.s = 0.0; e.s2 = 0.0; }
  const uint32_t pl = __shfl_up_sync(0xffffffffu, e.a, 1);
  int end;
  rq_segment(lane == 0 || !valid || lane == n || pl != e.a, lane, end);
  const bool head = valid && (lane == 0 || pl != e.a);
  uint32_t w[8];
  rq_widen(e.hlo, e.hhi, w);
  const bool fl = (e.aux & 0x10000u) != 0;
  const uint32_t gx = (uint32_t)(x0warp + (int)(e.aux >> 24));
  uint32_t cx0 = fl ? ~gx : 0u, cx1 = fl ? gx + 1u : 0u;
  uint32_t cz0 = fl ? ~(uint32_t)(z0 + (int)(e.aux & 0xFFu)) : 0u, cz1 = fl ? (uint32_t)(z0 + (int)((e.aux >> 8) & 0xFFu)) + 1u : 0u;
  uint32_t c = e.cnt, kmin = e.kmin, kmax = e.kmax;
  double s = e.s, s2 = e.s2;
  RQ_SEG_BEGIN
    RQ_SEG_ADD(c) RQ_SEG_ADD(s) RQ_SEG_ADD(s2)
    RQ_SEG_ADD(w[0]) RQ_SEG_ADD(w[1]) RQ_SEG_ADD(w[2]) RQ_SEG_ADD(w[3]) RQ_SEG_ADD(w[4]) RQ_SEG_ADD(w[5]) RQ_SEG_ADD(w[6]) RQ_SEG_ADD(w[7])
    RQ_SEG_MAX(kmin) RQ_SEG_MAX(kmax) RQ_SEG_MAX(cx0) RQ_SEG_MAX(cx1) RQ_SEG_MAX(cz0) RQ_SEG_MAX(cz1)
  RQ_SEG_END
  if (head) {
    const uint32_t rmask = (uint32_t)(P.regSlots - 1);
    int slot = -1;
    uint32_t hsh = ((e.a * 2654435761u) >> 9) & rmask;
    for (int probe = 0; probe < 8; ++probe) {
      uint32_t cur = S.rkeys[hsh];
      if (cur == 0xFFFFFFFFu) cur = atomicCAS(S.rkeys + hsh, 0xFFFFFFFFu, e.a);
      if (cur == e.a || cur == 0xFFFFFFFFu) { slot = (int)hsh; break; }
      hsh = (hsh + 1) & rmask;
    }
    if (slot >= 0) {
      atomicAdd(S.rcnt + slot, c);
#pragma unroll
      for (int j = 0; j < 8; ++j) if (w[j]) atomicAdd(S.rhist + 8 * slot + j, w[j]);
      atomicAdd(S.rsum + 2 * slot, s); atomicAdd(S.rsum + 2 * slot + 1, s2);
      uint32_t* m = S.rmx + 8 * slot;
      if (cx1) {
        rq_smax(m + 0, cx0); rq_smax(m + 1, ~(uint32_t)gy); rq_smax(m + 2, cz0);
        rq_smax(m + 3, cx1); rq_smax(m + 4, (uint32_t)gy + 1); rq_smax(m + 5, cz1);
      }
      if (kmin) rq_smax(m + 6, kmin);
      if (kmax) rq_smax(m + 7, kmax);
    } else if (e.a >= P.labelBound) {
      atomicExch(P.overflow, 2);
    } else {  // table full: the combined run goes straight to the global record
      uint32_t* ga = P.regAdd + (size_t)e.a * P.reg.nAdd;
      uint32_t* gm = P.regMax + (size_t)e.a * P.reg.nMax;
      double* gf = P.regF64 + (size_t)e.a * P.reg.nF64;
      atomicAdd(ga, c & 0xFFFFu);
      if (c >> 16) atomicAdd(ga + 1, c >> 16);
      for (int j = 0; j < 8; ++j) {
        if (w[j] & 0xFFFFu) atomicAdd(ga + 2 + 2 * j, w[j] & 0xFFFFu);
        if (w[j] >> 16) atomicAdd(ga + 3 + 2 * j, w[j] >> 16);
      }
      atomicAdd(gf, s); atomicAdd(gf + 1, s2);
      if (cx1) {
        atomicMax(gm + 0, cx0); atomicMax(gm + 1, ~(uint32_t)gy); atomicMax(gm + 2, cz0);
        atomicMax(gm + 3, cx1); atomicMax(gm + 4, (uint32_t)gy + 1); atomicMax(gm + 5, cz1);
      }
      if (kmin) atomicMax(gm + 6, kmin);
      if (kmax) atomicMax(gm + 7, kmax);
    }
  }
  __syncwarp();
}

__device__ __forceinline__ void rq_drain_edges(const FastParams& P, const RqSmem& S, const RqEntry* q, int n, int lane) {
  __syncwarp();
  const bool valid = lane < n;
  RqEntry e;
  if (valid) e = q[lane]; else { e.a = 0; e.b = 0; e.cnt = 0; e.hlo = e.hhi = 0; e.kmin = e.kmax = 0; e.aux = 0; e.s = 0.0; e.s2 = 0.0; }
  const uint32_t pa = __shfl_up_sync(0xffffffffu, e.a, 1), pb = __shfl_up_sync(0xffffffffu, e.b, 1);
  int end;
  rq_segment(lane == 0 || !valid || lane == n || pa != e.a || pb != e.b, lane, end);
  const bool head = valid && (lane == 0 || pa != e.a || pb != e.b);
  uint32_t w[8];
  rq_widen(e.hlo, e.hhi, w);
  uint32_t c01 = (e.cnt & 0xFFu) | ((e.cnt & 0xFF00u) << 8);         // voxels | thr0 << 16 (table layout)
  uint32_t c23 = ((e.cnt >> 16) & 0xFFu) | ((e.cnt >> 24) << 16);     // thr1 | thr2 << 16
  uint32_t kmin = e.kmin, kmax = e.kmax;
  double s = e.s, s2 = e.s2;
  RQ_SEG_BEGIN
    RQ_SEG_ADD(c01) RQ_SEG_ADD(c23) RQ_SEG_ADD(s) RQ_SEG_ADD(s2)
    RQ_SEG_ADD(w[0]) RQ_SEG_ADD(w[1]) RQ_SEG_ADD(w[2]) RQ_SEG_ADD(w[3]) RQ_SEG_ADD(w[4]) RQ_SEG_ADD(w[5]) RQ_SEG_ADD(w[6]) RQ_SEG_ADD(w[7])
    RQ_SEG_MAX(kmin) RQ_SEG_MAX(kmax)
  RQ_SEG_END
  if (head) {
    const uint32_t emask = (uint32_t)(P.edgeSlots - 1);
    const unsigned long long key = ((unsigned long long)e.a << 32) | e.b;
    int slot = -1;
    uint32_t hsh = ((e.a * 0x9E3779B1u) ^ (e.b * 0x85EBCA77u)) >> 11 & emask;
    for (int probe = 0; probe < 8; ++probe) {
      unsigned long long cur = S.ekeys[hsh];
      if (cur == EDGE_EMPTY) cur = atomicCAS(S.ekeys + hsh, EDGE_EMPTY, key);
      if (cur == key || cur == EDGE_EMPTY) { slot = (int)hsh; break; }
      hsh = (hsh + 1) & emask;
    }
    if (slot >= 0) {
      atomicAdd(S.ecnt + 2 * slot, c01);
      if (c23) atomicAdd(S.ecnt + 2 * slot + 1, c23);
#pragma unroll
      for (int j = 0; j < 8; ++j) if (w[j]) atomicAdd(S.ehist + 8 * slot + j, w[j]);
      atomicAdd(S.esum + 2 * slot, s); atomicAdd(S.esum + 2 * slot + 1, s2);
      if (kmin) rq_smax(S.emx + 2 * slot, kmin);
      if (kmax) rq_smax(S.emx + 2 * slot + 1, kmax);
    } else {
      const long long g = rq_edge_slot_global(P, key);
      if (g >= 0) {
        uint32_t* ga = P.edgeAdd + (size_t)g * P.edge.nAdd;
        uint32_t* gm = P.edgeMax + (size_t)g * P.edge.nMax;
        double* gf = P.edgeF64 + (size_t)g * P.edge.nF64;
        atomicAdd(ga, c01 & 0xFFFFu);
        if (P.nThr > 0 && (c01 >> 16)) atomicAdd(ga + 1, c01 >> 16);
        if (P.nThr > 1 && (c23 & 0xFFFFu)) atomicAdd(ga + 2, c23 & 0xFFFFu);
        if (P.nThr > 2 && (c23 >> 16)) atomicAdd(ga + 3, c23 >> 16);
        uint32_t* gh = ga + 1 + P.nThr;
        for (int j = 0; j < 8; ++j) {
          if (w[j] & 0xFFFFu) atomicAdd(gh + 2 * j, w[j] & 0xFFFFu);
          if (w[j] >> 16) atomicAdd(gh + 2 * j + 1, w[j] >> 16);
        }
        atomicAdd(gf, s); atomicAdd(gf + 1, s2);
        if (kmin) atomicMax(gm + 0, kmin);
        if (kmax) atomicMax(gm + 1, kmax);
      }
    }
  }
  __syncwarp();
}

// after draining the first 32 entries: move the remaining rem (< 32) to the front
__device__ __forceinline__ void rq_shift(RqEntry* q, int rem, int lane) {
  RqEntry t;
  if (lane < rem) t = q[32 + lane];
  __syncwarp();
  if (lane < rem) q[lane] = t;
  __syncwarp();
}

}  // namespace

__global__ void __launch_bounds__(256)
voxel_pass_rq(const __grid_constant__ FastParams P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nthreads = blockDim.x;
  const RqSmem S = rq_carve(smem_raw, P, nthreads >> 5);
  for (int i = tid; i < P.edgeSlots; i += nthreads) { S.ekeys[i] = EDGE_EMPTY; S.esum[2 * i] = 0.0; S.esum[2 * i + 1] = 0.0; S.ecnt[2 * i] = 0; S.ecnt[2 * i + 1] = 0; S.emx[2 * i] = 0; S.emx[2 * i + 1] = 0; }
  for (int i = tid; i < P.edgeSlots * 8; i += nthreads) S.ehist[i] = 0;
  for (int i = tid; i < P.regSlots; i += nthreads) { S.rkeys[i] = 0xFFFFFFFFu; S.rcnt[i] = 0; S.rsum[2 * i] = 0.0; S.rsum[2 * i + 1] = 0.0; }
  for (int i = tid; i < P.regSlots * 8; i += nthreads) { S.rhist[i] = 0; S.rmx[i] = 0; }
  if (tid < 16) { S.sRB[tid] = P.rBound[tid]; S.sEB[tid] = P.eBound[tid]; }
  __syncthreads();

  int t = blockIdx.x;
  const int tileX = t % P.tilesX; t /= P.tilesX;
  const int tileY = t % P.tilesY; t /= P.tilesY;
  const int wx = warp % P.warpsX, wy = warp / P.warpsX;
  const int x0warp = (tileX * P.warpsX + wx) * 32;
  const int gx = x0warp + lane;
  const int gy = tileY * P.warpsY + wy;
  const int z0 = P.zBegin + t * P.TZ;
  const int z1 = min(z0 + P.TZ, P.zEnd);
  const bool threeD = P.ndim == 3;
  const bool rowOk = gy < P.Y;
  const bool active = rowOk && gx < P.X;
  const size_t strideY = (size_t)P.X, strideZ = (size_t)P.X * P.Y;
  const bool vxm = gx > 0, vxp = gx < P.X - 1;
  const bool vym = threeD && gy > 0, vyp = threeD && gy < P.Y - 1;
  const int nvalidXY = (int)vxm + (int)vxp + (int)vym + (int)vyp;
  const int twoD = 2 * P.ndim;
  const bool edgeLane = active && ((lane == 0 && vxm) || (lane == 31 && vxp));
  const int edgeOff = (lane == 0) ? -1 : 1;
  const uint32_t ltmask = (1u << lane) - 1u;
  RqEntry* myq = S.queue + (size_t)warp * RQ_CAP;  // region runs
  RqEntry* myqe = myq + 64;                         // boundary runs
  int qn = 0, qe = 0;  // warp-uniform queue lengths

  uint32_t rLabel = 0, rCnt = 0, rHlo = 0, rHhi = 0, rKmin = 0, rKmax = 0; int rZmin = 0, rZmax = -1; double rS = 0.0, rS2 = 0.0;
  uint32_t eA = 0, eB = 0, eCnt = 0, eHlo = 0, eHhi = 0, eKmin = 0, eKmax = 0; double eS = 0.0, eS2 = 0.0;

#define RQ_PUSH_REGION(pos)                                                                                  \
  { RqEntry* d_ = myq + (pos);                                                                                \
    *(uint4*)d_ = make_uint4(rLabel, RQ_REGION, rCnt, rHlo);                                                  \
    *((uint4*)d_ + 1) = make_uint4(rHhi, rKmin, rKmax,                                                        \
        (rZmax >= 0 ? (uint32_t)(rZmin - z0) | ((uint32_t)(rZmax - z0) << 8) | 0x10000u : 0u) | ((uint32_t)lane << 24)); \
    *((double2*)d_ + 2) = make_double2(rS, rS2); }
#define RQ_PUSH_EDGE(pos)                                                                                    \
  { RqEntry* d_ = myqe + (pos);                                                                               \
    *(uint4*)d_ = make_uint4(eA, eB, eCnt, eHlo);                                                             \
    *((uint4*)d_ + 1) = make_uint4(eHhi, eKmin, eKmax, 0u);                                                   \
    *((double2*)d_ + 2) = make_double2(eS, eS2); }

  if (rowOk && z0 < z1) {  // warp-uniform
    const size_t col = (size_t)gy * strideY + (size_t)(active ? gx : 0);
    uint32_t lm = 0, l0 = 0, lp = 0, ym = 0, yp = 0, xe = 0; float pv = 0.f;
    if (active) {
      size_t p = (size_t)(z0 - P.zBuf0) * strideZ + col;
      l0 = __ldg(P.labels + p);
      if (z0 > 0) lm = __ldg(P.labels + p - strideZ);
      if (z0 + 1 < P.Z) lp = __ldg(P.labels + p + strideZ);
      if (vym) ym = __ldg(P.labels + p - strideY);
      if (vyp) yp = __ldg(P.labels + p + strideY);
      pv = __ldg(P.pb + p);
      if (edgeLane) xe = __ldg(P.labels + p + edgeOff);
    }
    for (int gz = z0; gz < z1; ++gz) {
      uint32_t nlp = 0, nym = 0, nyp = 0, nxe = 0; float npv = 0.f;
      if (active && gz + 1 < z1) {  // next step's loads first
        size_t p = (size_t)(gz + 1 - P.zBuf0) * strideZ + col;
        if (gz + 2 < P.Z) nlp = __ldg(P.labels + p + strideZ);
        if (vym) nym = __ldg(P.labels + p - strideY);
        if (vyp) nyp = __ldg(P.labels + p + strideY);
        npv = __ldg(P.pb + p);
        if (edgeLane) nxe = __ldg(P.labels + p + edgeOff);
      }
      uint32_t xm = __shfl_up_sync(0xffffffffu, l0, 1), xp = __shfl_down_sync(0xffffffffu, l0, 1);
      if (lane == 0) xm = xe;
      if (lane == 31) xp = xe;
      // ---- classify (type/neighbor.hxx:78-88 order x-, x+, y-, y+, z-, z+: assigned in reverse)
      const uint32_t me = l0;
      const bool vzm = gz > 0, vzp = gz < P.Z - 1;
      bool found = false; uint32_t nb = 0;
      if (vzp && lp != me) { found = true; nb = lp; }
      if (vzm && lm != me) { found = true; nb = lm; }
      if (vyp && yp != me) { found = true; nb = yp; }
      if (vym && ym != me) { found = true; nb = ym; }
      if (vxp && xp != me) { found = true; nb = xp; }
      if (vxm && xm != me) { found = true; nb = xm; }
      found = found && active;
      const bool isBorder = !found && (nvalidXY + (int)vzm + (int)vzp) < twoD;  // util/struct.hxx:103-106
      const bool nonInterior = found || isBorder;
      const int binR = rq_bin(pv, P.rLo, P.rHi, P.rInv, P.rBins, S.sRB);
      const int binE = P.sameSpec ? binR : rq_bin(pv, P.eLo, P.eHi, P.eInv, P.eBins, S.sEB);
      const double dv = (double)pv, dv2 = dv * dv;
      const bool notNan = pv == pv;
      const uint32_t kord = f2ord(pv);
      // ---- region run
      const bool endR = active && rCnt != 0 && (me != rLabel || (rCnt & 0xFFFFu) >= 15u);
      const uint32_t mR = __ballot_sync(0xffffffffu, endR);
      if (endR) { RQ_PUSH_REGION(qn + __popc(mR & ltmask)); rCnt = 0; }
      qn += __popc(mR);
      if (active) {
        if (rCnt == 0) { rLabel = me; rHlo = rHhi = 0; rKmin = rKmax = 0; rZmax = -1; rS = 0.0; rS2 = 0.0; }
        rCnt += isBorder ? 0x10001u : 1u;
        if (binR >= 0) { if (binR < 8) rHlo += 1u << (4 * binR); else rHhi += 1u << (4 * (binR - 8)); }
        rS += dv; rS2 += dv2;
        if (notNan) { rKmin = max(rKmin, ~kord); rKmax = max(rKmax, kord); }
        if (nonInterior) { if (rZmax < 0) rZmin = gz; rZmax = gz; }
      }
      // ---- directed boundary run (me -> nb); voxels that are not boundary voxels do not end it
      const bool endE = found && eCnt != 0 && (eA != me || eB != nb || (eCnt & 0xFFu) >= 15u);
      const uint32_t mE = __ballot_sync(0xffffffffu, endE);
      if (endE) { RQ_PUSH_EDGE(qe + __popc(mE & ltmask)); eCnt = 0; }
      qe += __popc(mE);
      if (found) {
        if (eCnt == 0) { eA = me; eB = nb; eHlo = eHhi = 0; eKmin = eKmax = 0; eS = 0.0; eS2 = 0.0; }
        eCnt += 1u | ((P.nThr > 0 && pv >= P.thr_up[0]) ? 0x100u : 0u) | ((P.nThr > 1 && pv >= P.thr_up[1]) ? 0x10000u : 0u) |
                ((P.nThr > 2 && pv >= P.thr_up[2]) ? 0x1000000u : 0u);  // type/feat.hxx:498
        if (binE >= 0) { if (binE < 8) eHlo += 1u << (4 * binE); else eHhi += 1u << (4 * (binE - 8)); }
        eS += dv; eS2 += dv2;
        if (notNan) { eKmin = max(eKmin, ~kord); eKmax = max(eKmax, kord); }
      }
      if (qn >= 32) { rq_drain_regions(P, S, myq, 32, lane, x0warp, gy, z0); qn -= 32; rq_shift(myq, qn, lane); }
      if (qe >= 32) { rq_drain_edges(P, S, myqe, 32, lane); qe -= 32; rq_shift(myqe, qe, lane); }
      lm = l0; l0 = lp; lp = nlp; ym = nym; yp = nyp; xe = nxe; pv = npv;
    }
    // close the open runs and drain
    const uint32_t mR = __ballot_sync(0xffffffffu, rCnt != 0);
    if (rCnt != 0) RQ_PUSH_REGION(qn + __popc(mR & ltmask));
    qn += __popc(mR);
    const uint32_t mE = __ballot_sync(0xffffffffu, eCnt != 0);
    if (eCnt != 0) RQ_PUSH_EDGE(qe + __popc(mE & ltmask));
    qe += __popc(mE);
    for (int o = 0; o < qn; o += 32) rq_drain_regions(P, S, myq + o, min(32, qn - o), lane, x0warp, gy, z0);
    for (int o = 0; o < qe; o += 32) rq_drain_edges(P, S, myqe + o, min(32, qe - o), lane);
  }
  __syncthreads();
  // ---- flush the CTA's tables into the global records
  for (int s = tid; s < P.regSlots; s += nthreads) {
    uint32_t l = S.rkeys[s];
    if (l == 0xFFFFFFFFu) continue;
    if (l >= P.labelBound) { atomicExch(P.overflow, 2); continue; }
    uint32_t* ga = P.regAdd + (size_t)l * P.reg.nAdd;
    uint32_t* gm = P.regMax + (size_t)l * P.reg.nMax;
    double* gf = P.regF64 + (size_t)l * P.reg.nF64;
    atomicAdd(gf, S.rsum[2 * s]); atomicAdd(gf + 1, S.rsum[2 * s + 1]);
    uint32_t cb = S.rcnt[s];
    atomicAdd(ga, cb & 0xFFFFu);
    if (cb >> 16) atomicAdd(ga + 1, cb >> 16);
    for (int b = 0; b < P.rBins; ++b) { uint32_t c = (S.rhist[8 * s + (b >> 1)] >> (16 * (b & 1))) & 0xFFFFu; if (c) atomicAdd(ga + 2 + b, c); }
    for (int i = 0; i < 8; ++i) { uint32_t v = S.rmx[8 * s + i]; if (v) atomicMax(gm + i, v); }
  }
  for (int s = tid; s < P.edgeSlots; s += nthreads) {
    unsigned long long key = S.ekeys[s];
    if (key == EDGE_EMPTY) continue;
    long long g = rq_edge_slot_global(P, key);
    if (g < 0) continue;
    uint32_t* ga = P.edgeAdd + (size_t)g * P.edge.nAdd;
    uint32_t* gm = P.edgeMax + (size_t)g * P.edge.nMax;
    double* gf = P.edgeF64 + (size_t)g * P.edge.nF64;
    atomicAdd(gf, S.esum[2 * s]); atomicAdd(gf + 1, S.esum[2 * s + 1]);
    const uint32_t c0 = S.ecnt[2 * s], c1 = S.ecnt[2 * s + 1];
    atomicAdd(ga, c0 & 0xFFFFu);
    if (P.nThr > 0 && (c0 >> 16)) atomicAdd(ga + 1, c0 >> 16);
    if (P.nThr > 1 && (c1 & 0xFFFFu)) atomicAdd(ga + 2, c1 & 0xFFFFu);
    if (P.nThr > 2 && (c1 >> 16)) atomicAdd(ga + 3, c1 >> 16);
    for (int b = 0; b < P.eBins; ++b) { uint32_t c = (S.ehist[8 * s + (b >> 1)] >> (16 * (b & 1))) & 0xFFFFu; if (c) atomicAdd(ga + 1 + P.nThr + b, c); }
    if (S.emx[2 * s]) atomicMax(gm + 0, S.emx[2 * s]);
    if (S.emx[2 * s + 1]) atomicMax(gm + 1, S.emx[2 * s + 1]);
  }
}

cudaError_t launch_voxel_pass_rq(const FastParams& P, cudaStream_t st) {
  size_t smem = voxel_rq_smem_bytes(P);
  cudaError_t e = cudaFuncSetAttribute(voxel_pass_rq, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  long long tiles = (long long)P.tilesX * P.tilesY * P.tilesZ;
  if (tiles <= 0) return cudaSuccess;
  voxel_pass_rq<<<(unsigned)tiles, 32 * P.warpsX * P.warpsY, smem, st>>>(P);
  return cudaGetLastError();
}

}  // namespace glia
