// K1+K3+K4 fused, FAST path (the common hmt configuration: one image = the boundary
// probability map used as region image, boundary image and threshold source; <= 16 bins,
// <= 3 thresholds, no mask, values in [0, 1]).  Same results as voxel_pass_generic (voxel_pass.cu),
// same reference semantics (type/neighbor.hxx:72-126, util/struct.hxx:95-144, type/feat.hxx:71-90,
// :485-502, :632-638, :706-738, util/image_stats.hxx:12-36).
//
// Design (B200, round 2 - "one record per voxel, 32-bit shared atomics only"):
//  * A warp owns 32 consecutive x voxels of one (y) row and MARCHES along z (2D: along y): the
//    z-1 / z / z+1 labels of a thread's column live in registers, x+-1 come from warp shuffles,
//    y+-1 rows are coalesced 128-byte loads that hit L1, the next step's loads are issued before the
//    current step is processed.  The march is warp-convergent: every lane executes every step, lanes
//    outside the image only have their atomics predicated off.
//  * Every voxel of region a is exactly one of: boundary voxel of ONE directed pair (a -> b), border
//    voxel, interior voxel.  So a voxel updates exactly ONE record of the CTA's shared-memory table,
//    keyed (a, b) for boundary voxels and (a, a) otherwise, and the region totals are formed when the
//    tile is flushed: record(a) = (a, a) + sum_b (a, b).  That halves the work of boundary voxels
//    (about half of all voxels in 3D) and removes every divergent "if boundary" section from the march.
//  * All accumulators are 32-bit words updated with non-returning shared atomics (1-2.6 LSU cycles per
//    warp instruction on B200, same-address lanes merged in hardware; f64 / u64 shared atomics cost
//    20-150 and were 25 % of round 1's kernel, profiles/microbench/r1_prims.txt):
//      - one count per VALUE CLASS (voxel_fast.cuh): histogram bins of both channels and the
//        threshold counts are sums of class counts, so they cost one atomic per voxel together;
//      - sum v and sum v^2 as exact fixed-point integers (v * 2^48 truncated) split into three
//        16-bit chunks each, so a tile's chunk sums cannot overflow and need no carries.  The sums are
//        therefore independent of the order of the atomics: for values on a 2^-48 grid (every float
//        >= 2^-25, in particular all 8/16-bit derived probabilities) they are the exact real sums;
//      - min / max as order-preserving keys (test, then atomicMax), bounding box as x / y / z
//        occupancy bit masks of the tile (atomicOr), border count.
//  * Tile flush, one thread per OCCUPIED slot (slots are compacted into lists first so that warps stay
//    full): boundary slots are folded into their region slot inside shared memory, then every slot is
//    added to the global records with RED atomics (class counts are mapped to record words on the way,
//    fixed-point sums become doubles).
//  * Values outside [0, 1] (or NaN) raise overflow flag 4: the caller re-runs the volume through the
//    generic kernel.
#include <cstdlib>
#include "voxel_fast.cuh"

namespace glia {

#define NCW (GLIA_FAST_MAX_CLS / 2)  // class-count words (two 16-bit counters per word)
#define W_SUM NCW                    // 3 chunks of sum v, 3 chunks of sum v^2
#define W_MIN (NCW + 6)
#define W_MAX (NCW + 7)
#define W_BORDER (NCW + 8)           // region slots: border voxels; boundary slots (after the march): global slot index
#define W_FOLD (NCW + 9)             // boundary slots (after the march): region slot of the tile it was folded into, or ~0
#define W_YZ (NCW + 10)              // bits [0, WY): tile rows, bits [WY, WY + TZ): tile planes
#define W_X (NCW + 11)               // WX words: bit = lane
#define KEY_EMPTY (~0ull)

__device__ __forceinline__ long long edge_slot_global_fast(const FastParams& P, unsigned long long key) {
  uint64_t mask = P.edgeCap - 1;
  uint64_t h = mix64(key) & mask;
  for (int probe = 0; probe < 4096; ++probe) {
    unsigned long long cur = P.edgeKeys[h];
    if (cur == key) return (long long)h;
    if (cur == EDGE_EMPTY) {
      unsigned long long prev = atomicCAS(P.edgeKeys + h, EDGE_EMPTY, key);
      if (prev == EDGE_EMPTY || prev == key) return (long long)h;
    }
    h = (h + 1) & mask;
  }
  atomicOr(P.overflow, 1);
  return -1;
}

// slow path: a voxel whose key found no shared slot goes straight to the global records
__device__ __noinline__ void voxel_global(const FastParams& P, uint32_t me, uint32_t nb, bool found, float v,
                                          int cls, bool isBorder, int gx, int gy, int gz) {
  if (me >= P.labelBound) { atomicOr(P.overflow, 2); return; }
  const int rb = P.vc.rBin[cls];
  uint32_t* ga = P.regAdd + (size_t)me * P.reg.nAdd;
  uint32_t* gm = P.regMax + (size_t)me * P.reg.nMax;
  double* gf = P.regF64 + (size_t)me * P.reg.nF64;
  const uint32_t k = f2ord(v);
  atomicAdd(ga, 1u);
  if (isBorder) atomicAdd(ga + 1, 1u);
  if (rb >= 0) atomicAdd(ga + 2 + rb, 1u);
  atomicAdd(gf, (double)v); atomicAdd(gf + 1, (double)v * (double)v);
  if (found || isBorder) {
    atomicMax(gm + 0, ~(uint32_t)gx); atomicMax(gm + 1, ~(uint32_t)gy); atomicMax(gm + 2, ~(uint32_t)gz);
    atomicMax(gm + 3, (uint32_t)gx + 1); atomicMax(gm + 4, (uint32_t)gy + 1); atomicMax(gm + 5, (uint32_t)gz + 1);
  }
  atomicMax(gm + 6, ~k); atomicMax(gm + 7, k);
  if (!found) return;
  long long g = edge_slot_global_fast(P, ((unsigned long long)me << 32) | nb);
  if (g < 0) return;
  uint32_t* ea = P.edgeAdd + (size_t)g * P.edge.nAdd;
  uint32_t* em = P.edgeMax + (size_t)g * P.edge.nMax;
  double* ef = P.edgeF64 + (size_t)g * P.edge.nF64;
  const int eb = P.vc.eBin[cls];
  const uint32_t tm = P.vc.thr[cls];
  atomicAdd(ea, 1u);
  for (int t = 0; t < P.nThr; ++t) if (tm >> t & 1u) atomicAdd(ea + 1 + t, 1u);
  if (eb >= 0) atomicAdd(ea + 1 + P.nThr + eb, 1u);
  atomicAdd(ef, (double)v); atomicAdd(ef + 1, (double)v * (double)v);
  atomicMax(em + 0, ~k); atomicMax(em + 1, k);
}

__device__ __forceinline__ uint32_t key_hash(uint32_t a, uint32_t b) {
  return ((a * 0x9E3779B1u) ^ (b * 0x85EBCA77u)) >> 10;
}

// The CTA's table is probed in BUCKETS of two adjacent keys (one 16-byte load): bucket hash, hash + 1, ... (4 at most).
// A key goes into the first empty slot of the first bucket that has one; slots never become empty again, so two
// inserters of one key always meet in the same slot.
// find-or-insert (used by the flush; the march has its own warp-convergent copy of this loop); -1 = no room.
// *inserted is set when this call claimed the slot.
__device__ __forceinline__ int slot_of(unsigned long long* keys, uint32_t bmask, uint32_t a, uint32_t b, bool* inserted) {
  const unsigned long long key = ((unsigned long long)a << 32) | b;
  uint32_t h = key_hash(a, b) & bmask;
  int adv = 0;
  *inserted = false;
#pragma unroll 1
  for (int probe = 0; probe < 8; ++probe) {
    const ulonglong2 kk = *reinterpret_cast<const ulonglong2*>(keys + 2 * h);
    if (kk.x == key) return (int)(2 * h);
    if (kk.y == key) return (int)(2 * h + 1);
    if (kk.x == KEY_EMPTY || kk.y == KEY_EMPTY) {
      const int j = kk.x == KEY_EMPTY ? 0 : 1;
      const unsigned long long cur = atomicCAS(keys + 2 * h + j, KEY_EMPTY, key);
      if (cur == KEY_EMPTY) { *inserted = true; return (int)(2 * h + j); }
      if (cur == key) return (int)(2 * h + j);
      continue;  // another key took the slot: look at this bucket again
    }
    h = (h + 1) & bmask;
    if (++adv >= 4) return -1;  // four full buckets
  }
  return -1;
}

// exact fixed-point images of v and v^2 for a float v in [0, 1]: floor(v * 2^48), floor(v^2 * 2^48)
__device__ __forceinline__ void fixed48(float v, uint32_t& lo, uint32_t& hi, uint32_t& lo2, uint32_t& hi2) {
  const uint32_t bits = __float_as_uint(v) & 0x7FFFFFFFu;
  const uint32_t e = bits >> 23;
  const uint32_t m = (bits & 0x7FFFFFu) | (e ? 0x800000u : 0u);
  const int ee = (int)(e ? e : 1u);       // v = m * 2^(ee - 150)
  const int s = ee - 102;                 // v * 2^48 = m * 2^s, s <= 25 for v <= 1
  if (s >= 0) { lo = m << s; hi = __funnelshift_l(m, 0u, s); }
  else { lo = s > -24 ? m >> (-s) : 0u; hi = 0u; }
  const unsigned long long m2 = (unsigned long long)m * m;  // v^2 * 2^48 = m2 * 2^(2 ee - 252)
  const int s2 = 2 * ee - 252;            // <= 2
  const unsigned long long q2 = s2 >= 0 ? m2 << s2 : (s2 > -48 ? m2 >> (-s2) : 0ull);
  lo2 = (uint32_t)q2; hi2 = (uint32_t)(q2 >> 32);
}

struct TileGeom { int x0, y0, z0; };

template <int LD>
__device__ __forceinline__ double slot_sum(const uint32_t* __restrict__ T, int s, int w) {
  const unsigned long long q = (unsigned long long)T[w * LD + s] + ((unsigned long long)T[(w + 1) * LD + s] << 16) +
                               ((unsigned long long)T[(w + 2) * LD + s] << 32);
  return __ull2double_rn(q) * 0x1p-48;
}

// adds the statistics words of shared slot s to the global record of region a
template <int LD, int WX, int WY>
__device__ __forceinline__ void flush_region_words(const FastParams& P, const uint32_t* __restrict__ T, int s,
                                                   uint32_t a, const TileGeom& G, bool withBorder) {
  if (a >= P.labelBound) { atomicOr(P.overflow, 2); return; }
  uint32_t* ga = P.regAdd + (size_t)a * P.reg.nAdd;
  uint32_t* gm = P.regMax + (size_t)a * P.reg.nMax;
  double* gf = P.regF64 + (size_t)a * P.reg.nF64;
  uint32_t n = 0, run = 0; int cur = -1;
#pragma unroll
  for (int j = 0; j < GLIA_FAST_MAX_CLS; ++j) {  // j is a compile-time constant; the class -> bin map is warp-uniform
    if (j < P.vc.nCls) {
      const uint32_t c = (T[(j >> 1) * LD + s] >> (16 * (j & 1))) & 0xFFFFu;
      const int b = P.vc.rBin[j];
      if (b != cur) { if (run && cur >= 0) atomicAdd(ga + 2 + cur, run); run = 0; cur = b; }
      run += c; n += c;
    }
  }
  if (run && cur >= 0) atomicAdd(ga + 2 + cur, run);
  if (n) atomicAdd(ga, n);
  if (withBorder) { const uint32_t bord = T[W_BORDER * LD + s]; if (bord) atomicAdd(ga + 1, bord); }
  const double sv = slot_sum<LD>(T, s, W_SUM), sv2 = slot_sum<LD>(T, s, W_SUM + 3);
  if (sv != 0.0) atomicAdd(gf, sv);
  if (sv2 != 0.0) atomicAdd(gf + 1, sv2);
  const uint32_t kmin = T[W_MIN * LD + s], kmax = T[W_MAX * LD + s];
  if (kmin) atomicMax(gm + 6, kmin);
  if (kmax) atomicMax(gm + 7, kmax);
  const uint32_t yz = T[W_YZ * LD + s];
  if (yz) {  // bounding box of the non-interior voxels (alg/geometry.hxx:21-39 needs only the extremes)
    int xmn = -1, xmx = -1;
#pragma unroll
    for (int w = 0; w < WX; ++w) {
      const uint32_t m = T[(W_X + w) * LD + s];
      if (m) { if (xmn < 0) xmn = 32 * w + __ffs(m) - 1; xmx = 32 * w + 31 - __clz(m); }
    }
    const uint32_t ym = yz & ((1u << WY) - 1u), zm = yz >> WY;
    atomicMax(gm + 0, ~(uint32_t)(G.x0 + xmn)); atomicMax(gm + 3, (uint32_t)(G.x0 + xmx) + 1);
    atomicMax(gm + 1, ~(uint32_t)(G.y0 + __ffs(ym) - 1)); atomicMax(gm + 4, (uint32_t)(G.y0 + 31 - __clz(ym)) + 1);
    atomicMax(gm + 2, ~(uint32_t)(G.z0 + __ffs(zm) - 1)); atomicMax(gm + 5, (uint32_t)(G.z0 + 31 - __clz(zm)) + 1);
  }
}

template <int LD>
__device__ __forceinline__ void flush_edge_words(const FastParams& P, const uint32_t* __restrict__ T, int s, uint32_t g) {
  uint32_t* ga = P.edgeAdd + (size_t)g * P.edge.nAdd;
  uint32_t* gm = P.edgeMax + (size_t)g * P.edge.nMax;
  double* gf = P.edgeF64 + (size_t)g * P.edge.nF64;
  uint32_t n = 0, run = 0, t0 = 0, t1 = 0, t2 = 0; int cur = -1;
  const int hb = 1 + P.nThr;
#pragma unroll
  for (int j = 0; j < GLIA_FAST_MAX_CLS; ++j) {
    if (j < P.vc.nCls) {
      const uint32_t c = (T[(j >> 1) * LD + s] >> (16 * (j & 1))) & 0xFFFFu;
      const int b = P.vc.eBin[j];
      const uint32_t tm = P.vc.thr[j];
      if (b != cur) { if (run && cur >= 0) atomicAdd(ga + hb + cur, run); run = 0; cur = b; }
      run += c; n += c;
      if (tm & 1u) t0 += c;
      if (tm & 2u) t1 += c;
      if (tm & 4u) t2 += c;
    }
  }
  if (run && cur >= 0) atomicAdd(ga + hb + cur, run);
  if (n) atomicAdd(ga, n);
  if (t0) atomicAdd(ga + 1, t0);  // type/feat.hxx:498
  if (t1) atomicAdd(ga + 2, t1);
  if (t2) atomicAdd(ga + 3, t2);
  const double sv = slot_sum<LD>(T, s, W_SUM), sv2 = slot_sum<LD>(T, s, W_SUM + 3);
  if (sv != 0.0) atomicAdd(gf, sv);
  if (sv2 != 0.0) atomicAdd(gf + 1, sv2);
  const uint32_t kmin = T[W_MIN * LD + s], kmax = T[W_MAX * LD + s];
  if (kmin) atomicMax(gm + 0, kmin);
  if (kmax) atomicMax(gm + 1, kmax);
}

template <bool THREED, int NS>
struct FastShape {
  static constexpr int WX = THREED ? 1 : 8, WY = THREED ? 8 : 1;
  static constexpr int NW = W_X + WX, LD = NS + 1;
  static constexpr int TAB_WORDS = ((NW * LD + 3) / 4) * 4;
  // keys | table | lists (2 x NS u16) | list counters | LUT | class bounds
  static constexpr size_t SMEM = 8 * (size_t)NS + 4 * (size_t)TAB_WORDS + 4 * (size_t)NS + 16 + 256 + 4 * GLIA_FAST_MAX_CLS;
};

template <bool THREED, int NS>
__global__ void __launch_bounds__(256, THREED ? 4 : 3)
voxel_pass_fast(const __grid_constant__ FastParams P) {
  using S = FastShape<THREED, NS>;
  constexpr int WX = S::WX, WY = S::WY, LD = S::LD, NW = S::NW;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int nthreads = 256;
  unsigned long long* keys = reinterpret_cast<unsigned long long*>(smem_raw);
  uint32_t* T = reinterpret_cast<uint32_t*>(smem_raw + 8 * (size_t)NS);
  unsigned short* listE = reinterpret_cast<unsigned short*>(T + S::TAB_WORDS);
  unsigned short* listR = listE + NS;
  uint32_t* listN = reinterpret_cast<uint32_t*>(listR + NS);  // [0] boundary slots, [1] region slots
  unsigned char* sLut = reinterpret_cast<unsigned char*>(listN + 4);
  float* sBound = reinterpret_cast<float*>(sLut + 256);
  {
    uint4* k4 = reinterpret_cast<uint4*>(keys);
    for (int i = tid; i < NS / 2; i += nthreads) k4[i] = make_uint4(~0u, ~0u, ~0u, ~0u);
    uint4* t4 = reinterpret_cast<uint4*>(T);
    for (int i = tid; i < S::TAB_WORDS / 4; i += nthreads) t4[i] = make_uint4(0u, 0u, 0u, 0u);
    if (tid < 64) reinterpret_cast<uint32_t*>(sLut)[tid] = reinterpret_cast<const uint32_t*>(P.vc.lut)[tid];
    if (tid < GLIA_FAST_MAX_CLS) sBound[tid] = P.vc.bound[tid];
    if (tid < 4) listN[tid] = 0;
  }
  __syncthreads();

  const int tileX = blockIdx.x, tileY = blockIdx.y, t = blockIdx.z;
  const int wx = warp % WX, wy = warp / WX;
  const int x0warp = (tileX * WX + wx) * 32;
  const int gx = x0warp + lane;
  const int gy = tileY * WY + wy;
  const int z0 = P.zBegin + t * P.TZ;
  const int z1 = min(z0 + P.TZ, P.zEnd);
  const bool rowOk = gy < P.Y;
  const bool inImg = rowOk && gx < P.X;
  const bool vxm = gx > 0, vxp = gx < P.X - 1;
  const bool vym = THREED && gy > 0, vyp = THREED && gy < P.Y - 1;  // warp-uniform
  const int nvalidXY = (int)vxm + (int)vxp + (int)vym + (int)vyp;
  constexpr int twoD = THREED ? 6 : 4;
  constexpr uint32_t bmask = (uint32_t)(NS / 2 - 1);  // buckets of two keys
  const int nB = P.vc.nCls - 1;
  const float lutLo = P.vc.lutLo, lutScale = P.vc.lutScale;
  uint32_t* const Tx = T + (W_X + wx) * LD;
  const uint32_t laneBit = 1u << lane;
  const uint32_t yBit = 1u << wy;
  bool badValue = false;

  if (rowOk && z0 < z1) {  // warp-uniform
    // One pointer per load stream, advanced by a plane per step.  Lanes beyond the right image edge read the
    // last voxel of the row (their results are discarded), so no load depends on a per-lane predicate
    // except the two lanes that fetch the x neighbours outside the warp's 32 voxels.
    const long long sY = P.X, sZ = (long long)P.X * P.Y;
    const int xi = min(gx, P.X - 1);
    const uint32_t* pC = P.labels + ((long long)(z0 - P.zBuf0) * sZ + (long long)gy * sY + xi);
    const float* pP = P.pb + ((long long)(z0 - P.zBuf0) * sZ + (long long)gy * sY + xi);
    const bool edgeLane = inImg && ((lane == 0 && vxm) || (lane == 31 && vxp));
    const int edgeOff = (lane == 0) ? -1 : 1;
    uint32_t lm = 0, l0, lp = 0, ym = 0, yp = 0, xe = 0; float pv;
    l0 = __ldg(pC);
    if (z0 > 0) lm = __ldg(pC - sZ);
    if (z0 + 1 < P.Z) lp = __ldg(pC + sZ);
    if (vym) ym = __ldg(pC - sY);
    if (vyp) yp = __ldg(pC + sY);
    pv = __ldg(pP);
    if (edgeLane) xe = __ldg(pC + edgeOff);
    // streams of the NEXT step
    const uint32_t* qZ = pC + 2 * sZ;   // label two planes ahead
    const uint32_t* qYm = pC + sZ - sY;
    const uint32_t* qYp = pC + sZ + sY;
    const uint32_t* qX = pC + sZ + edgeOff;
    const float* qP = pP + sZ;
#pragma unroll 1
    for (int gz = z0; gz < z1; ++gz) {
      // ---- issue the next step's loads first
      uint32_t nlp = 0, nym = 0, nyp = 0, nxe = 0; float npv = 0.f;
      if (gz + 1 < z1) {  // warp-uniform
        if (gz + 2 < P.Z) nlp = __ldg(qZ);
        if (vym) nym = __ldg(qYm);
        if (vyp) nyp = __ldg(qYp);
        npv = __ldg(qP);
        if (edgeLane) nxe = __ldg(qX);
      }
      qZ += sZ; qYm += sZ; qYp += sZ; qX += sZ; qP += sZ;
      uint32_t xm = __shfl_up_sync(0xffffffffu, l0, 1), xp = __shfl_down_sync(0xffffffffu, l0, 1);
      if (lane == 0) xm = xe;
      if (lane == 31) xp = xe;
      // ---- classify (type/neighbor.hxx:78-88 order x-, x+, y-, y+, z-, z+: assigned in reverse)
      const uint32_t me = l0;
      const bool vzm = gz > 0, vzp = gz < P.Z - 1;
      uint32_t nb = me;
      if (vzp && lp != me) nb = lp;
      if (vzm && lm != me) nb = lm;
      if (vyp && yp != me) nb = yp;
      if (vym && ym != me) nb = ym;
      if (vxp && xp != me) nb = xp;
      if (vxm && xm != me) nb = xm;
      const bool found = nb != me;
      const bool isBorder = !found && (nvalidXY + (int)vzm + (int)vzp) < twoD;  // util/struct.hxx:103-106
      const bool inDomain = pv >= 0.f && pv <= 1.f;
      if (inImg && me == 0xFFFFFFFFu) atomicOr(P.overflow, 2);  // reserved: the empty-slot marker of the tables
      badValue = badValue || (inImg && !inDomain);
      bool need = inImg && inDomain && me != 0xFFFFFFFFu;
      const bool ok = need;
      // ---- value class: start from the table entry one cell below the value's cell, count upwards
      int q = (int)((pv - lutLo) * lutScale) - 1;
      q = max(0, min(q, 255));
      int cls = sLut[q];
      cls += (cls < nB && pv >= sBound[cls]) ? 1 : 0;  // at most two steps: build_value_classes checks the table
      cls += (cls < nB && pv >= sBound[cls]) ? 1 : 0;
      uint32_t lo, hi, lo2, hi2;
      fixed48(pv, lo, hi, lo2, hi2);
      const uint32_t kord = __float_as_uint(pv) | 0x80000000u;  // f2ord of a non-negative float
      // ---- slot of the key: warp-convergent linear probing
      const unsigned long long key = ((unsigned long long)me << 32) | nb;
      uint32_t h = key_hash(me, nb) & bmask;
      int s = -1, adv = 0;
#pragma unroll 1
      for (int probe = 0; probe < 8; ++probe) {
        if (need) {
          const ulonglong2 kk = *reinterpret_cast<const ulonglong2*>(keys + 2 * h);
          if (kk.x == key) s = (int)(2 * h);
          else if (kk.y == key) s = (int)(2 * h + 1);
          else if (kk.x == KEY_EMPTY || kk.y == KEY_EMPTY) {
            const int j = kk.x == KEY_EMPTY ? 0 : 1;
            const unsigned long long cur = atomicCAS(keys + 2 * h + j, KEY_EMPTY, key);
            if (cur == KEY_EMPTY || cur == key) s = (int)(2 * h + j);  // else another key took it: same bucket again
          } else { h = (h + 1) & bmask; if (++adv >= 4) need = false; }  // four full buckets: no room (global fallback)
          if (s >= 0) need = false;
        }
        if (!__any_sync(0xffffffffu, need)) break;
      }
      if (s >= 0) {
        uint32_t* Ts = T + s;
        atomicAdd(Ts + (cls >> 1) * LD, 1u << (16 * (cls & 1)));
        const uint32_t c0 = lo & 0xFFFFu, c1 = lo >> 16, d0 = lo2 & 0xFFFFu, d1 = lo2 >> 16;
        if (c0) atomicAdd(Ts + W_SUM * LD, c0);
        if (c1) atomicAdd(Ts + (W_SUM + 1) * LD, c1);
        if (hi) atomicAdd(Ts + (W_SUM + 2) * LD, hi);
        if (d0) atomicAdd(Ts + (W_SUM + 3) * LD, d0);
        if (d1) atomicAdd(Ts + (W_SUM + 4) * LD, d1);
        if (hi2) atomicAdd(Ts + (W_SUM + 5) * LD, hi2);
        if (Ts[W_MIN * LD] < ~kord) atomicMax(Ts + W_MIN * LD, ~kord);
        if (Ts[W_MAX * LD] < kord) atomicMax(Ts + W_MAX * LD, kord);
        if (found || isBorder) {
          atomicOr(Tx + s, laneBit);
          atomicOr(Ts + W_YZ * LD, yBit | (1u << (WY + (gz - z0))));
          if (isBorder) atomicAdd(Ts + W_BORDER * LD, 1u);
        }
      } else if (ok) {
        voxel_global(P, me, nb, found, pv, cls, isBorder, gx, gy, gz);
      }
      lm = l0; l0 = lp; lp = nlp; ym = nym; yp = nyp; xe = nxe; pv = npv;
    }
  }
  if (badValue) atomicOr(P.overflow, 4);
  __syncthreads();
  // ---- flush.  Phase 0: compact the occupied slots into a boundary list and a region list
  for (int s0 = warp * 32; s0 < NS; s0 += nthreads) {
    const int s = s0 + lane;
    const unsigned long long key = keys[s];
    const bool occ = key != KEY_EMPTY;
    const bool isE = occ && (uint32_t)(key >> 32) != (uint32_t)key;
    const uint32_t mE = __ballot_sync(0xffffffffu, isE), mR = __ballot_sync(0xffffffffu, occ && !isE);
    uint32_t bE = 0, bR = 0;
    if (lane == 0) { if (mE) bE = atomicAdd(listN, (uint32_t)__popc(mE)); if (mR) bR = atomicAdd(listN + 1, (uint32_t)__popc(mR)); }
    bE = __shfl_sync(0xffffffffu, bE, 0); bR = __shfl_sync(0xffffffffu, bR, 0);
    const uint32_t below = laneBit - 1u;
    if (isE) listE[bE + __popc(mE & below)] = (unsigned short)s;
    else if (occ) listR[bR + __popc(mR & below)] = (unsigned short)s;
  }
  __syncthreads();
  const int nE = (int)listN[0];
  const TileGeom G{tileX * WX * 32, tileY * WY, z0};
  // Phase A: every boundary slot (a, b) looks up (or creates) the tile's region slot (a, a) and its slot in
  // the global table (32 independent L2 probes in flight per warp)
  for (int i = tid; i < nE; i += nthreads) {
    const int s = listE[i];
    const unsigned long long key = keys[s];
    const uint32_t a = (uint32_t)(key >> 32);
    bool ins;
    const int r = slot_of(keys, bmask, a, a, &ins);
    if (ins) listR[atomicAdd(listN + 1, 1u)] = (unsigned short)r;
    T[W_FOLD * LD + s] = (uint32_t)r;
    const long long g = edge_slot_global_fast(P, key);
    T[W_BORDER * LD + s] = g < 0 ? 0xFFFFFFFFu : (uint32_t)g;
  }
  __syncthreads();
  // Phase B: fold the boundary slots into their region slots
  for (int i = tid; i < nE; i += nthreads) {
    const int s = listE[i];
    const int r = (int)T[W_FOLD * LD + s];
    if (r < 0) continue;  // no room in the table: phase C sends this part to the global region record directly
#pragma unroll
    for (int w = 0; w < W_MIN; ++w) { const uint32_t v = T[w * LD + s]; if (v) atomicAdd(T + w * LD + r, v); }
    { const uint32_t v = T[W_MIN * LD + s]; if (v) atomicMax(T + W_MIN * LD + r, v); }
    { const uint32_t v = T[W_MAX * LD + s]; if (v) atomicMax(T + W_MAX * LD + r, v); }
#pragma unroll
    for (int w = W_YZ; w < NW; ++w) { const uint32_t v = T[w * LD + s]; if (v) atomicOr(T + w * LD + r, v); }
  }
  __syncthreads();
  // Phase C: add the tile's records to the global tables
  const int nR = (int)listN[1];
  for (int i = tid; i < nE; i += nthreads) {
    const int s = listE[i];
    const uint32_t g = T[W_BORDER * LD + s];
    if (g != 0xFFFFFFFFu) flush_edge_words<LD>(P, T, s, g);
    if ((int)T[W_FOLD * LD + s] < 0) flush_region_words<LD, WX, WY>(P, T, s, (uint32_t)(keys[s] >> 32), G, false);
  }
  for (int i = tid; i < nR; i += nthreads) {
    const int s = listR[i];
    flush_region_words<LD, WX, WY>(P, T, s, (uint32_t)(keys[s] >> 32), G, true);
  }
}

template <bool THREED, int NS>
static cudaError_t launch_fast(const FastParams& P, cudaStream_t st) {
  const size_t smem = FastShape<THREED, NS>::SMEM;
  cudaError_t e = cudaFuncSetAttribute(voxel_pass_fast<THREED, NS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const long long tiles = (long long)P.tilesX * P.tilesY * P.tilesZ;
  if (tiles <= 0) return cudaSuccess;
  if (P.tilesY > 65535 || P.tilesZ > 65535) return cudaErrorInvalidConfiguration;
  voxel_pass_fast<THREED, NS><<<dim3((unsigned)P.tilesX, (unsigned)P.tilesY, (unsigned)P.tilesZ), 256, smem, st>>>(P);
  return cudaGetLastError();
}

// The CTA footprint is fixed by the dimension: 256 x 1 voxels per step in 2D, 32 x 8 in 3D (api_build.cu sets
// warpsX / warpsY accordingly); P.TZ planes per tile with warpsY + TZ <= 32.
cudaError_t launch_voxel_pass_fast(const FastParams& P, cudaStream_t st) {
  if (P.ndim == 3) return P.slots == 1024 ? launch_fast<true, 1024>(P, st) : launch_fast<true, 512>(P, st);
  return P.slots == 256 ? launch_fast<false, 256>(P, st) : launch_fast<false, 512>(P, st);
}

}  // namespace glia
