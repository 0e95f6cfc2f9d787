// EXPERIMENTAL (GLIA_B200_VOXEL=queue; parity-green, 8.97 ms vs 3.57 ms for v3 at 512^3: the drain
// costs ~670 warp-instructions per step at ~3 active lanes - region/boundary branches, contended CAS
// loops on doubles, a second round for n - 32 leftovers; profiles/r1_v4_notes.txt has the numbers
// and the fix: two typed queues, drain exactly 32, combine adjacent equal keys by segmented shuffles).
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
// with a ballot-computed position (no atomic, queue length is a warp-uniform register); when the
// queue holds >= 32 runs the whole warp drains it, one run per lane: slot lookup in the CTA's
// shared-memory tables, a handful of shared atomics per run.  Tables are flushed to the global
// records once per tile, as in v3.
#include "voxel_fast.cuh"

namespace glia {
namespace {

#define RQ_CAP 96      // runs per warp queue (drained at >= 32; one step appends at most 64)
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

// a run whose key found no shared slot goes straight to the global records
__device__ __noinline__ void rq_region_run_global(const FastParams& P, const RqEntry& e, int x0warp, int gy, int z0) {
  if (e.a >= P.labelBound) { atomicExch(P.overflow, 2); return; }
  uint32_t* ga = P.regAdd + (size_t)e.a * P.reg.nAdd;
  uint32_t* gm = P.regMax + (size_t)e.a * P.reg.nMax;
  double* gf = P.regF64 + (size_t)e.a * P.reg.nF64;
  atomicAdd(ga, e.cnt & 0xFFFFu);
  if (e.cnt >> 16) atomicAdd(ga + 1, e.cnt >> 16);
  for (int k = 0; k < 16; ++k) { uint32_t v = ((k < 8 ? e.hlo : e.hhi) >> (4 * (k & 7))) & 15u; if (v) atomicAdd(ga + 2 + k, v); }
  atomicAdd(gf, e.s); atomicAdd(gf + 1, e.s2);
  if (e.aux & 0x10000u) {
    const uint32_t gx = (uint32_t)(x0warp + (int)(e.aux >> 24));
    const uint32_t zmn = (uint32_t)(z0 + (int)(e.aux & 0xFFu)), zmx = (uint32_t)(z0 + (int)((e.aux >> 8) & 0xFFu));
    atomicMax(gm + 0, ~gx); atomicMax(gm + 1, ~(uint32_t)gy); atomicMax(gm + 2, ~zmn);
    atomicMax(gm + 3, gx + 1); atomicMax(gm + 4, (uint32_t)gy + 1); atomicMax(gm + 5, zmx + 1);
  }
  if (e.kmin) atomicMax(gm + 6, e.kmin);
  if (e.kmax) atomicMax(gm + 7, e.kmax);
}
__device__ __noinline__ void rq_edge_run_global(const FastParams& P, const RqEntry& e) {
  long long g = rq_edge_slot_global(P, ((unsigned long long)e.a << 32) | e.b);
  if (g < 0) return;
  uint32_t* ga = P.edgeAdd + (size_t)g * P.edge.nAdd;
  uint32_t* gm = P.edgeMax + (size_t)g * P.edge.nMax;
  double* gf = P.edgeF64 + (size_t)g * P.edge.nF64;
  atomicAdd(ga, e.cnt & 0xFFu);
  for (int t = 0; t < P.nThr; ++t) { uint32_t v = (e.cnt >> (8 * (t + 1))) & 0xFFu; if (v) atomicAdd(ga + 1 + t, v); }
  for (int k = 0; k < 16; ++k) { uint32_t v = ((k < 8 ? e.hlo : e.hhi) >> (4 * (k & 7))) & 15u; if (v) atomicAdd(ga + 1 + P.nThr + k, v); }
  atomicAdd(gf, e.s); atomicAdd(gf + 1, e.s2);
  if (e.kmin) atomicMax(gm + 0, e.kmin);
  if (e.kmax) atomicMax(gm + 1, e.kmax);
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

// the whole warp applies the queued runs, one per lane: slot lookup + a handful of shared atomics
__device__ __forceinline__ void rq_drain(const FastParams& P, const RqSmem& S, const RqEntry* q, int n, int lane,
                                         int x0warp, int gy, int z0) {
  __syncwarp();
  const uint32_t rmask = (uint32_t)(P.regSlots - 1), emask = (uint32_t)(P.edgeSlots - 1);
  for (int i = lane; i < n; i += 32) {
    const RqEntry e = q[i];
    unsigned long long h = (unsigned long long)e.hlo | ((unsigned long long)e.hhi << 32);
    if (e.b == RQ_REGION) {
      int slot = -1;
      uint32_t hsh = ((e.a * 2654435761u) >> 9) & rmask;
      for (int probe = 0; probe < 8; ++probe) {
        uint32_t cur = S.rkeys[hsh];
        if (cur == 0xFFFFFFFFu) cur = atomicCAS(S.rkeys + hsh, 0xFFFFFFFFu, e.a);
        if (cur == e.a || cur == 0xFFFFFFFFu) { slot = (int)hsh; break; }
        hsh = (hsh + 1) & rmask;
      }
      if (slot < 0) { rq_region_run_global(P, e, x0warp, gy, z0); continue; }
      atomicAdd(S.rcnt + slot, e.cnt);
      while (h) {
        const int b = (__ffsll((long long)h) - 1) >> 2;
        const uint32_t v = (uint32_t)(h >> (4 * b)) & 15u;
        atomicAdd(S.rhist + 8 * slot + (b >> 1), v << (16 * (b & 1)));
        h &= ~(15ull << (4 * b));
      }
      atomicAdd(S.rsum + 2 * slot, e.s); atomicAdd(S.rsum + 2 * slot + 1, e.s2);
      uint32_t* m = S.rmx + 8 * slot;
      if (e.aux & 0x10000u) {
        const uint32_t gx = (uint32_t)(x0warp + (int)(e.aux >> 24));
        const uint32_t zmn = (uint32_t)(z0 + (int)(e.aux & 0xFFu)), zmx = (uint32_t)(z0 + (int)((e.aux >> 8) & 0xFFu));
        rq_smax(m + 0, ~gx); rq_smax(m + 1, ~(uint32_t)gy); rq_smax(m + 2, ~zmn);
        rq_smax(m + 3, gx + 1); rq_smax(m + 4, (uint32_t)gy + 1); rq_smax(m + 5, zmx + 1);
      }
      if (e.kmin) rq_smax(m + 6, e.kmin);
      if (e.kmax) rq_smax(m + 7, e.kmax);
    } else {
      const unsigned long long key = ((unsigned long long)e.a << 32) | e.b;
      int slot = -1;
      uint32_t hsh = ((e.a * 0x9E3779B1u) ^ (e.b * 0x85EBCA77u)) >> 11 & emask;
      for (int probe = 0; probe < 8; ++probe) {
        unsigned long long cur = S.ekeys[hsh];
        if (cur == EDGE_EMPTY) cur = atomicCAS(S.ekeys + hsh, EDGE_EMPTY, key);
        if (cur == key || cur == EDGE_EMPTY) { slot = (int)hsh; break; }
        hsh = (hsh + 1) & emask;
      }
      if (slot < 0) { rq_edge_run_global(P, e); continue; }
      // shared layout: {count | thr0 << 16, thr1 | thr2 << 16}
      atomicAdd(S.ecnt + 2 * slot, (e.cnt & 0xFFu) | ((e.cnt & 0xFF00u) << 8));
      const uint32_t t12 = ((e.cnt >> 16) & 0xFFu) | ((e.cnt >> 24) << 16);
      if (t12) atomicAdd(S.ecnt + 2 * slot + 1, t12);
      while (h) {
        const int b = (__ffsll((long long)h) - 1) >> 2;
        const uint32_t v = (uint32_t)(h >> (4 * b)) & 15u;
        atomicAdd(S.ehist + 8 * slot + (b >> 1), v << (16 * (b & 1)));
        h &= ~(15ull << (4 * b));
      }
      atomicAdd(S.esum + 2 * slot, e.s); atomicAdd(S.esum + 2 * slot + 1, e.s2);
      if (e.kmin) rq_smax(S.emx + 2 * slot, e.kmin);
      if (e.kmax) rq_smax(S.emx + 2 * slot + 1, e.kmax);
    }
  }
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
  RqEntry* myq = S.queue + (size_t)warp * RQ_CAP;
  int qn = 0;  // warp-uniform queue length

  uint32_t rLabel = 0, rCnt = 0, rHlo = 0, rHhi = 0, rKmin = 0, rKmax = 0; int rZmin = 0, rZmax = -1; double rS = 0.0, rS2 = 0.0;
  uint32_t eA = 0, eB = 0, eCnt = 0, eHlo = 0, eHhi = 0, eKmin = 0, eKmax = 0; double eS = 0.0, eS2 = 0.0;

#define RQ_PUSH_REGION(pos)                                                                                  \
  { RqEntry* d_ = myq + (pos);                                                                                \
    *(uint4*)d_ = make_uint4(rLabel, RQ_REGION, rCnt, rHlo);                                                  \
    *((uint4*)d_ + 1) = make_uint4(rHhi, rKmin, rKmax,                                                        \
        (rZmax >= 0 ? (uint32_t)(rZmin - z0) | ((uint32_t)(rZmax - z0) << 8) | 0x10000u : 0u) | ((uint32_t)lane << 24)); \
    *((double2*)d_ + 2) = make_double2(rS, rS2); }
#define RQ_PUSH_EDGE(pos)                                                                                    \
  { RqEntry* d_ = myq + (pos);                                                                                \
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
      if (endE) { RQ_PUSH_EDGE(qn + __popc(mE & ltmask)); eCnt = 0; }
      qn += __popc(mE);
      if (found) {
        if (eCnt == 0) { eA = me; eB = nb; eHlo = eHhi = 0; eKmin = eKmax = 0; eS = 0.0; eS2 = 0.0; }
        eCnt += 1u | ((P.nThr > 0 && pv >= P.thr_up[0]) ? 0x100u : 0u) | ((P.nThr > 1 && pv >= P.thr_up[1]) ? 0x10000u : 0u) |
                ((P.nThr > 2 && pv >= P.thr_up[2]) ? 0x1000000u : 0u);  // type/feat.hxx:498
        if (binE >= 0) { if (binE < 8) eHlo += 1u << (4 * binE); else eHhi += 1u << (4 * (binE - 8)); }
        eS += dv; eS2 += dv2;
        if (notNan) { eKmin = max(eKmin, ~kord); eKmax = max(eKmax, kord); }
      }
      if (qn >= 32) { rq_drain(P, S, myq, qn, lane, x0warp, gy, z0); qn = 0; }
      lm = l0; l0 = lp; lp = nlp; ym = nym; yp = nyp; xe = nxe; pv = npv;
    }
    // close the open runs and drain
    const uint32_t mR = __ballot_sync(0xffffffffu, rCnt != 0);
    if (rCnt != 0) RQ_PUSH_REGION(qn + __popc(mR & ltmask));
    qn += __popc(mR);
    const uint32_t mE = __ballot_sync(0xffffffffu, eCnt != 0);
    if (eCnt != 0) RQ_PUSH_EDGE(qn + __popc(mE & ltmask));
    qn += __popc(mE);
    rq_drain(P, S, myq, qn, lane, x0warp, gy, z0);
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
