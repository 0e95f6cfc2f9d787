// EXPERIMENTAL (GLIA_B200_VOXEL=runs; parity-green, currently 2.3x slower than voxel_fast.cu - the
// warp flushes nearly every step because ONE lane with a second finished run triggers it; see
// profiles/r1_v4_notes.txt for the measurements and the per-lane queue that should fix it).
// K1+K3+K4 fused, v4 "runs" (default hmt configuration: one image = the boundary-probability map
// as region image, boundary image and threshold source; <= 16 bins, <= 3 thresholds, no mask).
// Same results as voxel_pass_generic / voxel_pass_fast; reference semantics:
//   getContourTraits / genContourMap        type/neighbor.hxx:72-126, util/struct.hxx:95-144
//   RegionShapeFeats, ImageRegionShapeFeats type/feat.hxx:71-90, :485-502
//   ImageLabelFeats, ImageRealFeats         type/feat.hxx:632-638, :706-738, util/image_stats.hxx:12-36
//
// Why (measured on B200, profiles/r1_v3_ncu_c4s.txt + profiles/microbench/r1_red_tput.txt): v3 is
// bound by atomic lane-operations, not by HBM: ~4 shared/global atomic lane-ops per voxel at
// 1.4-2 LSU cycles each.  v4 keeps EVERY statistic in registers and aggregates in two directions
// before anything atomic happens:
//   * along z: a thread owns one column and accumulates its current run of equal label (region)
//     and of equal directed pair (boundary) in registers: count, 16 x 4-bit histogram, threshold
//     counts, sum, sum^2, min, max, bbox evidence;
//   * along x: a finished run is parked in a second register set ("pending"); when any lane has
//     to park a second one, the WHOLE warp flushes its pending runs together: neighbouring lanes
//     whose pending runs carry the same label (pair) form a segment, integers and doubles are combined with
//     a 5-step segmented shuffle reduction (redux.sync with per-lane masks is emulated), and only
//     the segment head issues global RED operations (native f64 add / u32 max in L2).
// No shared memory, no per-CTA tables, no hash lookups for regions (records are indexed by label);
// boundary records are found by the segment head in the global open-addressing table.
#include "voxel_fast.cuh"

namespace glia {

namespace {

struct RegAcc {
  uint32_t label, cnt;   // cnt = voxels | border voxels << 16 ; 0 == empty
  uint32_t hlo, hhi;     // 16 bins x 4 bits (a run is closed after 15 voxels)
  uint32_t kmin, kmax;   // ~ord(min), ord(max); 0 == none
  uint32_t flags;        // bit0/1: x-1 / x+1 neighbour outside the region seen, bit2/3: y-1 / y+1
  uint32_t czmin, czmax; // ~z of a voxel whose z-1 neighbour is outside the region, z+1 likewise
  double s, s2;
};
struct EdgeAcc {
  uint32_t a, b, cnt;    // cnt = voxels | thr0 << 8 | thr1 << 16 | thr2 << 24 ; 0 == empty
  uint32_t hlo, hhi, kmin, kmax;
  double s, s2;
};

__device__ __forceinline__ int bin_of(float v, float lo, float hi, float inv, int bins, const float* bound) {
  if (v > lo && v < hi) {  // util/image_stats.hxx:17-36 (cumulative bounds, no lo offset)
    int i = (int)(v * inv);
    i = i < 0 ? 0 : (i > bins - 1 ? bins - 1 : i);
    while (i > 0 && v < bound[i - 1]) --i;
    while (i < bins && !(v < bound[i])) ++i;
    return i < bins ? i : -1;
  }
  return (v <= lo) ? 0 : bins - 1;
}

__device__ __forceinline__ long long edge_slot(const FastParams& P, unsigned long long key) {
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

// segment of equal keys among neighbouring lanes: mask of the lanes of my segment, its last lane
__device__ __forceinline__ uint32_t segment_mask(bool head, int lane, int& end) {
  const uint32_t heads = __ballot_sync(0xffffffffu, head);
  const int start = 31 - __clz(heads & (0xffffffffu >> (31 - lane)));  // lane 0 is always a head
  const uint32_t above = lane == 31 ? 0u : (heads & (0xffffffffu << (lane + 1)));
  end = above ? (__ffs(above) - 2) : 31;
  return (0xffffffffu >> (31 - end)) & (0xffffffffu << start);
}

// Segmented reductions towards the segment head by shuffles (redux.sync with per-lane masks is
// emulated by a loop over the distinct masks - 25x slower here, profiles/r1_v4_notes.txt).
// After the 5 steps lane i holds the combination of lanes [i, end]; only heads use it.
#define SEG_REDUCE_BEGIN for (int d_ = 1; d_ < 32; d_ <<= 1) { const bool ok_ = lane + d_ <= end;
#define SEG_ADD(v) { auto t_ = __shfl_down_sync(0xffffffffu, v, d_); if (ok_) v += t_; }
#define SEG_MAX(v) { uint32_t t_ = __shfl_down_sync(0xffffffffu, v, d_); if (ok_) v = max(v, t_); }
#define SEG_REDUCE_END }

// 16 x 4-bit bins -> 8 words of 2 x 16-bit bins (a segment sums at most 32 x 15 per bin)
__device__ __forceinline__ void widen_hist(uint32_t hlo, uint32_t hhi, uint32_t (&w)[8]) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    w[j] = ((hlo >> (8 * j)) & 0xFu) | (((hlo >> (8 * j + 4)) & 0xFu) << 16);
    w[4 + j] = ((hhi >> (8 * j)) & 0xFu) | (((hhi >> (8 * j + 4)) & 0xFu) << 16);
  }
}

// whole warp: combine the pending region runs of neighbouring lanes and add them to the records
__device__ __forceinline__ void flush_regions(const FastParams& P, RegAcc& r, int lane, int gx, int gy) {
  const bool valid = r.cnt != 0;
  const uint32_t pl = __shfl_up_sync(0xffffffffu, r.label, 1);
  const bool pv = __shfl_up_sync(0xffffffffu, (int)valid, 1) != 0;
  int end;
  const uint32_t seg = segment_mask(lane == 0 || !valid || !pv || pl != r.label, lane, end);
  const bool head = valid && (seg & ((1u << lane) - 1)) == 0;
  // bbox along x / y from ballots: first / last lane of the segment that saw an outside neighbour
  const uint32_t f0 = __ballot_sync(0xffffffffu, valid && (r.flags & 1u)) & seg;
  const uint32_t f1 = __ballot_sync(0xffffffffu, valid && (r.flags & 2u)) & seg;
  const uint32_t f2 = __ballot_sync(0xffffffffu, valid && (r.flags & 4u)) & seg;
  const uint32_t f3 = __ballot_sync(0xffffffffu, valid && (r.flags & 8u)) & seg;
  uint32_t w[8];
  widen_hist(r.hlo, r.hhi, w);
  uint32_t c = r.cnt, kmin = r.kmin, kmax = r.kmax, z0 = r.czmin, z1 = r.czmax;
  double s = r.s, s2 = r.s2;
  SEG_REDUCE_BEGIN
    SEG_ADD(c) SEG_ADD(s) SEG_ADD(s2)
    SEG_ADD(w[0]) SEG_ADD(w[1]) SEG_ADD(w[2]) SEG_ADD(w[3]) SEG_ADD(w[4]) SEG_ADD(w[5]) SEG_ADD(w[6]) SEG_ADD(w[7])
    SEG_MAX(kmin) SEG_MAX(kmax) SEG_MAX(z0) SEG_MAX(z1)
  SEG_REDUCE_END
  r.cnt = 0;
  if (!head) return;
  if (r.label >= P.labelBound) { atomicExch(P.overflow, 2); return; }
  uint32_t* ga = P.regAdd + (size_t)r.label * P.reg.nAdd;
  uint32_t* gm = P.regMax + (size_t)r.label * P.reg.nMax;
  double* gf = P.regF64 + (size_t)r.label * P.reg.nF64;
  atomicAdd(ga, c & 0xFFFFu);
  if (c >> 16) atomicAdd(ga + 1, c >> 16);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    if (w[j] & 0xFFFFu) atomicAdd(ga + 2 + 2 * j, w[j] & 0xFFFFu);
    if (w[j] >> 16) atomicAdd(ga + 3 + 2 * j, w[j] >> 16);
  }
  atomicAdd(gf, s); atomicAdd(gf + 1, s2);
  const int x0lane = gx - lane;
  if (f0) atomicMax(gm + 0, ~(uint32_t)(x0lane + __ffs(f0) - 1));
  if (f2) atomicMax(gm + 1, ~(uint32_t)gy);
  if (z0) atomicMax(gm + 2, z0);
  if (f1) atomicMax(gm + 3, (uint32_t)(x0lane + 31 - __clz(f1)) + 1u);
  if (f3) atomicMax(gm + 4, (uint32_t)gy + 1u);
  if (z1) atomicMax(gm + 5, z1);
  if (kmin) atomicMax(gm + 6, kmin);
  if (kmax) atomicMax(gm + 7, kmax);
}

__device__ __forceinline__ void flush_edges(const FastParams& P, EdgeAcc& e, int lane) {
  const bool valid = e.cnt != 0;
  const uint32_t pa = __shfl_up_sync(0xffffffffu, e.a, 1), pb = __shfl_up_sync(0xffffffffu, e.b, 1);
  const bool pv = __shfl_up_sync(0xffffffffu, (int)valid, 1) != 0;
  int end;
  const uint32_t seg = segment_mask(lane == 0 || !valid || !pv || pa != e.a || pb != e.b, lane, end);
  const bool head = valid && (seg & ((1u << lane) - 1)) == 0;
  uint32_t w[8];
  widen_hist(e.hlo, e.hhi, w);
  uint32_t c01 = (e.cnt & 0xFFu) | ((e.cnt & 0xFF00u) << 8);            // voxels | thr0 << 16
  uint32_t c23 = ((e.cnt >> 16) & 0xFFu) | ((e.cnt >> 24) << 16);        // thr1 | thr2 << 16
  uint32_t kmin = e.kmin, kmax = e.kmax;
  double s = e.s, s2 = e.s2;
  SEG_REDUCE_BEGIN
    SEG_ADD(c01) SEG_ADD(c23) SEG_ADD(s) SEG_ADD(s2)
    SEG_ADD(w[0]) SEG_ADD(w[1]) SEG_ADD(w[2]) SEG_ADD(w[3]) SEG_ADD(w[4]) SEG_ADD(w[5]) SEG_ADD(w[6]) SEG_ADD(w[7])
    SEG_MAX(kmin) SEG_MAX(kmax)
  SEG_REDUCE_END
  const uint32_t ea = e.a, eb = e.b;
  e.cnt = 0;
  if (!head) return;
  const long long g = edge_slot(P, ((unsigned long long)ea << 32) | eb);
  if (g < 0) return;
  uint32_t* ga = P.edgeAdd + (size_t)g * P.edge.nAdd;
  uint32_t* gm = P.edgeMax + (size_t)g * P.edge.nMax;
  double* gf = P.edgeF64 + (size_t)g * P.edge.nF64;
  atomicAdd(ga, c01 & 0xFFFFu);
  if (P.nThr > 0 && (c01 >> 16)) atomicAdd(ga + 1, c01 >> 16);  // type/feat.hxx:498
  if (P.nThr > 1 && (c23 & 0xFFFFu)) atomicAdd(ga + 2, c23 & 0xFFFFu);
  if (P.nThr > 2 && (c23 >> 16)) atomicAdd(ga + 3, c23 >> 16);
  uint32_t* gh = ga + 1 + P.nThr;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    if (w[j] & 0xFFFFu) atomicAdd(gh + 2 * j, w[j] & 0xFFFFu);
    if (w[j] >> 16) atomicAdd(gh + 2 * j + 1, w[j] >> 16);
  }
  atomicAdd(gf, s); atomicAdd(gf + 1, s2);
  if (kmin) atomicMax(gm + 0, kmin);
  if (kmax) atomicMax(gm + 1, kmax);
}

}  // namespace

__global__ void __launch_bounds__(256, 2)
voxel_pass_runs(const __grid_constant__ FastParams P) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int t = blockIdx.x;
  const int tileX = t % P.tilesX; t /= P.tilesX;
  const int tileY = t % P.tilesY; t /= P.tilesY;
  const int wx = warp % P.warpsX, wy = warp / P.warpsX;
  const int gx = (tileX * P.warpsX + wx) * 32 + lane;
  const int gy = tileY * P.warpsY + wy;
  const int z0 = P.zBegin + t * P.TZ;
  const int z1 = min(z0 + P.TZ, P.zEnd);
  if (gy >= P.Y || z0 >= z1 || gx - lane >= P.X) return;  // warp-uniform
  const bool threeD = P.ndim == 3;
  const bool active = gx < P.X;
  const size_t strideY = (size_t)P.X, strideZ = (size_t)P.X * P.Y;
  const bool vxm = gx > 0, vxp = gx < P.X - 1;
  const bool vym = threeD && gy > 0, vyp = threeD && gy < P.Y - 1;
  const int nvalidXY = (int)vxm + (int)vxp + (int)vym + (int)vyp;
  const int twoD = 2 * P.ndim;
  const bool edgeLane = active && ((lane == 0 && vxm) || (lane == 31 && vxp));
  const int edgeOff = (lane == 0) ? -1 : 1;

  RegAcc rc, rp; EdgeAcc ec, ep;
  rc.cnt = rp.cnt = 0; ec.cnt = ep.cnt = 0;
  rc.label = rp.label = 0; rc.hlo = rc.hhi = rp.hlo = rp.hhi = 0; rc.kmin = rc.kmax = rp.kmin = rp.kmax = 0;
  rc.flags = rp.flags = 0; rc.czmin = rc.czmax = rp.czmin = rp.czmax = 0; rc.s = rc.s2 = rp.s = rp.s2 = 0.0;
  ec.a = ec.b = ep.a = ep.b = 0; ec.hlo = ec.hhi = ep.hlo = ep.hhi = 0; ec.kmin = ec.kmax = ep.kmin = ep.kmax = 0;
  ec.s = ec.s2 = ep.s = ep.s2 = 0.0;

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
    const bool oxm = !vxm || xm != me, oxp = !vxp || xp != me;  // neighbour outside the region (or image)
    const bool oym = !vym || ym != me, oyp = !vyp || yp != me;
    const bool ozm = !vzm || lm != me, ozp = !vzp || lp != me;
    bool found = false; uint32_t nb = 0;
    if (vzp && lp != me) { found = true; nb = lp; }
    if (vzm && lm != me) { found = true; nb = lm; }
    if (vyp && yp != me) { found = true; nb = yp; }
    if (vym && ym != me) { found = true; nb = ym; }
    if (vxp && xp != me) { found = true; nb = xp; }
    if (vxm && xm != me) { found = true; nb = xm; }
    found = found && active;
    const bool isBorder = !found && (nvalidXY + (int)vzm + (int)vzp) < twoD;  // util/struct.hxx:103-106
    const int binR = bin_of(pv, P.rLo, P.rHi, P.rInv, P.rBins, P.rBound);
    const int binE = P.sameSpec ? binR : bin_of(pv, P.eLo, P.eHi, P.eInv, P.eBins, P.eBound);
    const double dv = (double)pv, dv2 = dv * dv;
    const bool notNan = pv == pv;
    const uint32_t kord = f2ord(pv);

    // ---- region run
    const bool changedR = active && rc.cnt != 0 && (me != rc.label || (rc.cnt & 0xFFFFu) >= 15u);
    if (__any_sync(0xffffffffu, changedR && rp.cnt != 0)) flush_regions(P, rp, lane, gx, gy);
    if (changedR) { rp = rc; rc.cnt = 0; }
    if (active) {
      if (rc.cnt == 0) { rc.label = me; rc.hlo = rc.hhi = 0; rc.kmin = rc.kmax = 0; rc.flags = 0; rc.czmin = rc.czmax = 0; rc.s = 0.0; rc.s2 = 0.0; }
      rc.cnt += isBorder ? 0x10001u : 1u;
      if (binR >= 0) { if (binR < 8) rc.hlo += 1u << (4 * binR); else rc.hhi += 1u << (4 * (binR - 8)); }
      rc.s += dv; rc.s2 += dv2;
      if (notNan) { rc.kmin = max(rc.kmin, ~kord); rc.kmax = max(rc.kmax, kord); }
      rc.flags |= (oxm ? 1u : 0u) | (oxp ? 2u : 0u) | (oym ? 4u : 0u) | (oyp ? 8u : 0u);
      if (ozm) rc.czmin = max(rc.czmin, ~(uint32_t)gz);
      if (ozp) rc.czmax = max(rc.czmax, (uint32_t)gz + 1u);
    }
    // ---- directed boundary run (me -> nb); voxels that are not boundary voxels do not end it
    const bool changedE = found && ec.cnt != 0 && (ec.a != me || ec.b != nb || (ec.cnt & 0xFFu) >= 15u);
    if (__any_sync(0xffffffffu, changedE && ep.cnt != 0)) flush_edges(P, ep, lane);
    if (changedE) { ep = ec; ec.cnt = 0; }
    if (found) {
      if (ec.cnt == 0) { ec.a = me; ec.b = nb; ec.hlo = ec.hhi = 0; ec.kmin = ec.kmax = 0; ec.s = 0.0; ec.s2 = 0.0; }
      ec.cnt += 1u | ((P.nThr > 0 && pv >= P.thr_up[0]) ? 0x100u : 0u) | ((P.nThr > 1 && pv >= P.thr_up[1]) ? 0x10000u : 0u) |
                ((P.nThr > 2 && pv >= P.thr_up[2]) ? 0x1000000u : 0u);
      if (binE >= 0) { if (binE < 8) ec.hlo += 1u << (4 * binE); else ec.hhi += 1u << (4 * (binE - 8)); }
      ec.s += dv; ec.s2 += dv2;
      if (notNan) { ec.kmin = max(ec.kmin, ~kord); ec.kmax = max(ec.kmax, kord); }
    }
    lm = l0; l0 = lp; lp = nlp; ym = nym; yp = nyp; xe = nxe; pv = npv;
  }
  // ---- close the tile: pending first, then the open runs
  flush_regions(P, rp, lane, gx, gy);
  flush_regions(P, rc, lane, gx, gy);
  flush_edges(P, ep, lane);
  flush_edges(P, ec, lane);
}

cudaError_t launch_voxel_pass_runs(const FastParams& P, cudaStream_t st) {
  long long tiles = (long long)P.tilesX * P.tilesY * P.tilesZ;
  if (tiles <= 0) return cudaSuccess;
  voxel_pass_runs<<<(unsigned)tiles, 32 * P.warpsX * P.warpsY, 0, st>>>(P);
  return cudaGetLastError();
}

}  // namespace glia
