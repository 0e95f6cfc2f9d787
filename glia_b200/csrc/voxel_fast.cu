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
//    current step is processed.
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
//  * Tile flush: boundary slots are first folded into their region slot inside shared memory, then
//    every occupied slot is added to the global records with RED atomics (class counts are mapped to
//    record words on the way, fixed-point sums become doubles).
//  * Values outside [0, 1] (or NaN) raise overflow flag 4: the caller re-runs the volume through the
//    generic kernel.
#include <cstdlib>
#include "voxel_fast.cuh"

namespace glia {

#define NCW (GLIA_FAST_MAX_CLS / 2)  // class-count words (two 16-bit counters per word)
#define W_SUM NCW                    // 3 chunks of sum v, 3 chunks of sum v^2
#define W_MIN (NCW + 6)
#define W_MAX (NCW + 7)
#define W_BORDER (NCW + 8)
#define W_YZ (NCW + 9)               // bits [0, warpsY): tile rows, bits [warpsY, warpsY + TZ): tile planes
#define W_X (NCW + 10)               // warpsX words: bit = lane
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

// find-or-insert in the CTA's table; -1 when 8 probes found neither the key nor a free slot
__device__ __forceinline__ int slot_of(unsigned long long* keys, uint32_t mask, uint32_t a, uint32_t b) {
  const unsigned long long key = ((unsigned long long)a << 32) | b;
  uint32_t h = (((a * 0x9E3779B1u) ^ (b * 0x85EBCA77u)) >> 10) & mask;
#pragma unroll 1
  for (int probe = 0; probe < 8; ++probe) {
    unsigned long long cur = keys[h];
    if (cur == KEY_EMPTY) cur = atomicCAS(keys + h, KEY_EMPTY, key);
    if (cur == key || cur == KEY_EMPTY) return (int)h;
    h = (h + 1) & mask;
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

// adds the statistics words of shared slot s to the global record of region a
__device__ __forceinline__ void flush_region_words(const FastParams& P, const uint32_t* __restrict__ T, int LD, int s,
                                                   uint32_t a, const TileGeom& G) {
  if (a >= P.labelBound) { atomicOr(P.overflow, 2); return; }
  uint32_t* ga = P.regAdd + (size_t)a * P.reg.nAdd;
  uint32_t* gm = P.regMax + (size_t)a * P.reg.nMax;
  double* gf = P.regF64 + (size_t)a * P.reg.nF64;
  uint32_t n = 0, run = 0; int cur = -1;
  for (int j = 0; j < P.vc.nCls; ++j) {
    const uint32_t c = (T[(j >> 1) * LD + s] >> (16 * (j & 1))) & 0xFFFFu;
    const int b = P.vc.rBin[j];
    if (b != cur) { if (run && cur >= 0) atomicAdd(ga + 2 + cur, run); run = 0; cur = b; }
    run += c; n += c;
  }
  if (run && cur >= 0) atomicAdd(ga + 2 + cur, run);
  if (n) atomicAdd(ga, n);
  const uint32_t bord = T[W_BORDER * LD + s];
  if (bord) atomicAdd(ga + 1, bord);
  const unsigned long long q = (unsigned long long)T[W_SUM * LD + s] + ((unsigned long long)T[(W_SUM + 1) * LD + s] << 16) +
                               ((unsigned long long)T[(W_SUM + 2) * LD + s] << 32);
  const unsigned long long q2 = (unsigned long long)T[(W_SUM + 3) * LD + s] + ((unsigned long long)T[(W_SUM + 4) * LD + s] << 16) +
                                ((unsigned long long)T[(W_SUM + 5) * LD + s] << 32);
  if (q) atomicAdd(gf, __ull2double_rn(q) * 0x1p-48);
  if (q2) atomicAdd(gf + 1, __ull2double_rn(q2) * 0x1p-48);
  const uint32_t kmin = T[W_MIN * LD + s], kmax = T[W_MAX * LD + s];
  if (kmin) atomicMax(gm + 6, kmin);
  if (kmax) atomicMax(gm + 7, kmax);
  const uint32_t yz = T[W_YZ * LD + s];
  if (yz) {  // bounding box of the non-interior voxels (alg/geometry.hxx:21-39 needs only the extremes)
    int xmn = -1, xmx = -1;
    for (int w = 0; w < P.warpsX; ++w) {
      const uint32_t m = T[(W_X + w) * LD + s];
      if (m) { if (xmn < 0) xmn = 32 * w + __ffs(m) - 1; xmx = 32 * w + 31 - __clz(m); }
    }
    const uint32_t ym = yz & ((1u << P.warpsY) - 1u), zm = yz >> P.warpsY;
    atomicMax(gm + 0, ~(uint32_t)(G.x0 + xmn)); atomicMax(gm + 3, (uint32_t)(G.x0 + xmx) + 1);
    atomicMax(gm + 1, ~(uint32_t)(G.y0 + __ffs(ym) - 1)); atomicMax(gm + 4, (uint32_t)(G.y0 + 31 - __clz(ym)) + 1);
    atomicMax(gm + 2, ~(uint32_t)(G.z0 + __ffs(zm) - 1)); atomicMax(gm + 5, (uint32_t)(G.z0 + 31 - __clz(zm)) + 1);
  }
}

__device__ __forceinline__ void flush_edge_words(const FastParams& P, const uint32_t* __restrict__ T, int LD, int s,
                                                 unsigned long long key) {
  const long long g = edge_slot_global_fast(P, key);
  if (g < 0) return;
  uint32_t* ga = P.edgeAdd + (size_t)g * P.edge.nAdd;
  uint32_t* gm = P.edgeMax + (size_t)g * P.edge.nMax;
  double* gf = P.edgeF64 + (size_t)g * P.edge.nF64;
  uint32_t n = 0, run = 0, t0 = 0, t1 = 0, t2 = 0; int cur = -1;
  for (int j = 0; j < P.vc.nCls; ++j) {
    const uint32_t c = (T[(j >> 1) * LD + s] >> (16 * (j & 1))) & 0xFFFFu;
    const int b = P.vc.eBin[j];
    const uint32_t tm = P.vc.thr[j];
    if (b != cur) { if (run && cur >= 0) atomicAdd(ga + 1 + P.nThr + cur, run); run = 0; cur = b; }
    run += c; n += c;
    t0 += (tm & 1u) ? c : 0u; t1 += (tm & 2u) ? c : 0u; t2 += (tm & 4u) ? c : 0u;
  }
  if (run && cur >= 0) atomicAdd(ga + 1 + P.nThr + cur, run);
  if (n) atomicAdd(ga, n);
  if (t0) atomicAdd(ga + 1, t0);  // type/feat.hxx:498
  if (t1) atomicAdd(ga + 2, t1);
  if (t2) atomicAdd(ga + 3, t2);
  const unsigned long long q = (unsigned long long)T[W_SUM * LD + s] + ((unsigned long long)T[(W_SUM + 1) * LD + s] << 16) +
                               ((unsigned long long)T[(W_SUM + 2) * LD + s] << 32);
  const unsigned long long q2 = (unsigned long long)T[(W_SUM + 3) * LD + s] + ((unsigned long long)T[(W_SUM + 4) * LD + s] << 16) +
                                ((unsigned long long)T[(W_SUM + 5) * LD + s] << 32);
  if (q) atomicAdd(gf, __ull2double_rn(q) * 0x1p-48);
  if (q2) atomicAdd(gf + 1, __ull2double_rn(q2) * 0x1p-48);
  const uint32_t kmin = T[W_MIN * LD + s], kmax = T[W_MAX * LD + s];
  if (kmin) atomicMax(gm + 0, kmin);
  if (kmax) atomicMax(gm + 1, kmax);
}

size_t voxel_fast_smem_bytes(const FastParams& P) {
  const size_t NW = W_X + P.warpsX, LD = (size_t)P.slots + 1;
  const size_t tab = (4 * NW * LD + 15) & ~(size_t)15;
  return 8 * (size_t)P.slots + tab + 256 + 4 * GLIA_FAST_MAX_CLS + 16;
}

template <int MINB>
__global__ void __launch_bounds__(256, MINB)
voxel_pass_fast(const __grid_constant__ FastParams P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nthreads = blockDim.x;
  const int NS = P.slots, LD = NS + 1, NW = W_X + P.warpsX;
  unsigned long long* keys = reinterpret_cast<unsigned long long*>(smem_raw);
  uint32_t* T = reinterpret_cast<uint32_t*>(smem_raw + 8 * (size_t)NS);
  const int tabWords = ((NW * LD + 3) >> 2) << 2;
  unsigned char* sLut = reinterpret_cast<unsigned char*>(T + tabWords);
  float* sBound = reinterpret_cast<float*>(sLut + 256);
  {
    uint4* k4 = reinterpret_cast<uint4*>(keys);
    for (int i = tid; i < NS / 2; i += nthreads) k4[i] = make_uint4(~0u, ~0u, ~0u, ~0u);
    uint4* t4 = reinterpret_cast<uint4*>(T);
    for (int i = tid; i < tabWords / 4; i += nthreads) t4[i] = make_uint4(0u, 0u, 0u, 0u);
    if (tid < 64) reinterpret_cast<uint32_t*>(sLut)[tid] = reinterpret_cast<const uint32_t*>(P.vc.lut)[tid];
    if (tid < GLIA_FAST_MAX_CLS) sBound[tid] = P.vc.bound[tid];
  }
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
  const uint32_t smask = (uint32_t)(NS - 1);
  const int nB = P.vc.nCls - 1;
  const float lutLo = P.vc.lutLo, lutScale = P.vc.lutScale;
  uint32_t* const Tx = T + (W_X + wx) * LD;
  const uint32_t laneBit = 1u << lane;
  const uint32_t yBit = 1u << wy;
  bool badValue = false;

  if (rowOk && z0 < z1) {  // warp-uniform
    // 32-bit voxel offsets from the tile's first voxel of this row (launch code keeps (TZ + 2) planes < 2^31)
    const size_t rowBase = (size_t)(z0 - P.zBuf0) * strideZ + (size_t)gy * strideY + (size_t)x0warp;
    const uint32_t* __restrict__ L = P.labels + rowBase;
    const float* __restrict__ PB = P.pb + rowBase;
    const int sZ = (int)strideZ, sY = (int)strideY;
    int o = active ? lane : 0;
    uint32_t lm = 0, l0 = 0, lp = 0, ym = 0, yp = 0, xe = 0; float pv = 0.f;
    if (active) {
      l0 = __ldg(L + o);
      if (z0 > 0) lm = __ldg(L + (o - sZ));
      if (z0 + 1 < P.Z) lp = __ldg(L + (o + sZ));
      if (vym) ym = __ldg(L + (o - sY));
      if (vyp) yp = __ldg(L + (o + sY));
      pv = __ldg(PB + o);
      if (edgeLane) xe = __ldg(L + (o + edgeOff));
    }
    for (int gz = z0; gz < z1; ++gz) {
      // ---- issue the next step's loads first
      uint32_t nlp = 0, nym = 0, nyp = 0, nxe = 0; float npv = 0.f;
      o += sZ;
      if (active && gz + 1 < z1) {
        if (gz + 2 < P.Z) nlp = __ldg(L + (o + sZ));
        if (vym) nym = __ldg(L + (o - sY));
        if (vyp) nyp = __ldg(L + (o + sY));
        npv = __ldg(PB + o);
        if (edgeLane) nxe = __ldg(L + (o + edgeOff));
      }
      uint32_t xm = __shfl_up_sync(0xffffffffu, l0, 1), xp = __shfl_down_sync(0xffffffffu, l0, 1);
      if (lane == 0) xm = xe;
      if (lane == 31) xp = xe;
      if (active) {
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
        if (me == 0xFFFFFFFFu) atomicOr(P.overflow, 2);  // reserved: the empty-slot marker of the tables
        else if (!(pv >= 0.f && pv <= 1.f)) badValue = true;
        else {
          // ---- value class: start from the table entry one cell below the value's cell, count upwards
          int q = (int)((pv - lutLo) * lutScale) - 1;
          q = max(0, min(q, 255));
          int cls = sLut[q];
          while (cls < nB && pv >= sBound[cls]) ++cls;
          uint32_t lo, hi, lo2, hi2;
          fixed48(pv, lo, hi, lo2, hi2);
          const uint32_t kord = __float_as_uint(pv) | 0x80000000u;  // f2ord of a non-negative float
          const int s = slot_of(keys, smask, me, nb);
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
              atomicOr(Ts + W_YZ * LD, yBit | (1u << (P.warpsY + (gz - z0))));
              if (isBorder) atomicAdd(Ts + W_BORDER * LD, 1u);
            }
          } else {
            voxel_global(P, me, nb, found, pv, cls, isBorder, gx, gy, gz);
          }
        }
      }
      lm = l0; l0 = lp; lp = nlp; ym = nym; yp = nyp; xe = nxe; pv = npv;
    }
  }
  if (badValue) atomicOr(P.overflow, 4);
  __syncthreads();
  // ---- fold every boundary slot (a, b) into the region slot (a, a) of the tile
  const TileGeom G{tileX * P.warpsX * 32, tileY * P.warpsY, z0};
  for (int s = tid; s < NS; s += nthreads) {
    const unsigned long long key = keys[s];
    if (key == KEY_EMPTY) continue;
    const uint32_t a = (uint32_t)(key >> 32), b = (uint32_t)key;
    if (a == b) continue;
    const int r = slot_of(keys, smask, a, a);
    if (r < 0) { flush_region_words(P, T, LD, s, a, G); continue; }  // no room: this part goes to the global record directly
    for (int w = 0; w < W_MIN; ++w) { const uint32_t v = T[w * LD + s]; if (v) atomicAdd(T + w * LD + r, v); }
    { const uint32_t v = T[W_MIN * LD + s]; if (v) atomicMax(T + W_MIN * LD + r, v); }
    { const uint32_t v = T[W_MAX * LD + s]; if (v) atomicMax(T + W_MAX * LD + r, v); }
    for (int w = W_YZ; w < NW; ++w) { const uint32_t v = T[w * LD + s]; if (v) atomicOr(T + w * LD + r, v); }
  }
  __syncthreads();
  // ---- add the tile's records to the global tables
  for (int s = tid; s < NS; s += nthreads) {
    const unsigned long long key = keys[s];
    if (key == KEY_EMPTY) continue;
    const uint32_t a = (uint32_t)(key >> 32), b = (uint32_t)key;
    if (a == b) flush_region_words(P, T, LD, s, a, G);
    else flush_edge_words(P, T, LD, s, key);
  }
}

cudaError_t launch_voxel_pass_fast(const FastParams& P, cudaStream_t st) {
  const size_t smem = voxel_fast_smem_bytes(P);
  cudaError_t e = cudaFuncSetAttribute(voxel_pass_fast<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const long long tiles = (long long)P.tilesX * P.tilesY * P.tilesZ;
  if (tiles <= 0) return cudaSuccess;
  voxel_pass_fast<4><<<(unsigned)tiles, 32 * P.warpsX * P.warpsY, smem, st>>>(P);
  return cudaGetLastError();
}

}  // namespace glia
