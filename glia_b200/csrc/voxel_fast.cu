// K1+K3+K4 fused, FAST path (the common hmt configuration: one image = the boundary
// probability map used as region image, boundary image and threshold source; <= 16 bins,
// <= 3 thresholds, no mask).  Same results as voxel_pass_generic (voxel_pass.cu), same
// reference semantics (type/neighbor.hxx:72-126, util/struct.hxx:95-144, type/feat.hxx:71-90,
// :485-502, :632-638, :706-738, util/image_stats.hxx:12-36).
//
// Design (B200): no per-voxel atomics on doubles and no staging tile.
//  * A warp owns 32 consecutive x voxels of one (y) row and MARCHES along z (2D: along y):
//    the z-1 / z / z+1 labels of a thread's column live in registers, x+-1 come from warp
//    shuffles, y+-1 rows are coalesced 128-byte loads that hit L1 (the neighbouring warps of
//    the CTA load the same rows as their own), the next step's loads are issued before the
//    current step is processed.
//  * Integer statistics (counts, thresholds, histogram bins) go per voxel, convergently, into
//    shared-memory tables privatised per CTA with 32-bit shared atomics (~1 warp-instruction
//    per cycle on B200 and same-address lanes are merged, profiles/microbench/prims.cu).
//  * Double sums and min/max have no fast shared atomic.  A thread accumulates them in
//    registers over its run of equal slot and, when the slot changes, appends one 32-byte
//    entry to its warp's queue; the queue is drained by the WHOLE warp (one entry per lane, CAS
//    loop on doubles) whenever it holds >= 32 entries.  v2 flushed divergently inside the march
//    (8.7 of 32 lanes active on average, 1195 warp-instructions per 32 voxels:
//    profiles/r1_v2_ncu_voxel_fast.txt); the queue turns that into convergent work.
//  * The CTA's tables are flushed once per tile into the global records with RED atomics.
#include <cstdlib>
#include "voxel_fast.cuh"

namespace glia {

// shared-memory tables of a CTA (all 32-bit words unless noted)
//   region slot: key u32 | cnt = count | border << 16 | hist[8] (2 x 16-bit bins per word) |
//                sum, sum2 (double) | mx[8] = ~bbminx, ~bbminy, ~bbminz, bbmax+1 x3, ~min, max
//   edge slot  : key u64 | cnt[2] = {count | thr0 << 16, thr1 | thr2 << 16} | hist[8] |
//                sum, sum2 (double) | mx[2] = ~min, max
#define QCAP 96      // queue entries per warp (drained at >= 32; one step adds at most 64)
#define Q_EDGE 0x80000000u

struct __align__(16) QEntry {
  uint32_t slot;   // table slot | Q_EDGE
  uint32_t kmin, kmax;
  uint32_t aux;    // region: zminOff | zmaxOff << 8 | hasNonInterior << 16 | lane << 24
  double s, s2;
};

__device__ __forceinline__ int bin_fast(float v, float lo, float hi, float inv, int bins,
                                        const float* __restrict__ bound) {
  if (v > lo && v < hi) {
    int i = (int)(v * inv);
    i = i < 0 ? 0 : (i > bins - 1 ? bins - 1 : i);
    while (i > 0 && v < bound[i - 1]) --i;
    while (i < bins && !(v < bound[i])) ++i;
    return i < bins ? i : -1;
  }
  return (v <= lo) ? 0 : bins - 1;
}

__device__ __forceinline__ void smax(uint32_t* w, uint32_t v) {
  if (*(volatile uint32_t*)w < v) atomicMax(w, v);
}

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
  atomicExch(P.overflow, 1);
  return -1;
}

// slow path: a voxel whose key found no shared slot goes straight to the global records
__device__ __noinline__ void region_voxel_global(const FastParams& P, uint32_t label, float v, int bin,
                                                 bool isBorder, bool nonInterior, int gx, int gy, int gz) {
  if (label >= P.labelBound) { atomicExch(P.overflow, 2); return; }
  uint32_t* ga = P.regAdd + (size_t)label * P.reg.nAdd;
  uint32_t* gm = P.regMax + (size_t)label * P.reg.nMax;
  double* gf = P.regF64 + (size_t)label * P.reg.nF64;
  atomicAdd(ga, 1u);
  if (isBorder) atomicAdd(ga + 1, 1u);
  if (bin >= 0) atomicAdd(ga + 2 + bin, 1u);
  atomicAdd(gf, (double)v); atomicAdd(gf + 1, (double)v * (double)v);
  if (nonInterior) {
    atomicMax(gm + 0, ~(uint32_t)gx); atomicMax(gm + 1, ~(uint32_t)gy); atomicMax(gm + 2, ~(uint32_t)gz);
    atomicMax(gm + 3, (uint32_t)gx + 1); atomicMax(gm + 4, (uint32_t)gy + 1); atomicMax(gm + 5, (uint32_t)gz + 1);
  }
  if (v == v) { uint32_t k = f2ord(v); atomicMax(gm + 6, ~k); atomicMax(gm + 7, k); }
}
__device__ __noinline__ void edge_voxel_global(const FastParams& P, unsigned long long key, float v, int bin) {
  long long g = edge_slot_global_fast(P, key);
  if (g < 0) return;
  uint32_t* ga = P.edgeAdd + (size_t)g * P.edge.nAdd;
  uint32_t* gm = P.edgeMax + (size_t)g * P.edge.nMax;
  double* gf = P.edgeF64 + (size_t)g * P.edge.nF64;
  atomicAdd(ga, 1u);
  for (int k = 0; k < P.nThr; ++k) if (v >= P.thr_up[k]) atomicAdd(ga + 1 + k, 1u);
  if (bin >= 0) atomicAdd(ga + 1 + P.nThr + bin, 1u);
  atomicAdd(gf, (double)v); atomicAdd(gf + 1, (double)v * (double)v);
  if (v == v) { uint32_t k = f2ord(v); atomicMax(gm + 0, ~k); atomicMax(gm + 1, k); }
}

}  // namespace glia

namespace glia {

struct SmemView {
  unsigned long long* ekeys; double* esum; double* rsum;
  QEntry* queue;
  uint32_t *ecnt, *ehist, *emx, *rkeys, *rcnt, *rhist, *rmx, *qn;
  float *sRB, *sEB;
};

__device__ __forceinline__ SmemView carve(unsigned char* raw, const FastParams& P, int nwarps) {
  SmemView S;
  unsigned char* p = raw;
  S.queue = (QEntry*)p; p += sizeof(QEntry) * QCAP * nwarps;
  // keys (filled with ones) first, then everything that starts at zero: one vectorised fill each
  S.ekeys = (unsigned long long*)p; p += 8 * (size_t)P.edgeSlots;
  S.rkeys = (uint32_t*)p; p += 4 * (size_t)P.regSlots;
  S.esum = (double*)p; p += 16 * (size_t)P.edgeSlots;
  S.rsum = (double*)p; p += 16 * (size_t)P.regSlots;
  S.ecnt = (uint32_t*)p; p += 8 * (size_t)P.edgeSlots;
  S.ehist = (uint32_t*)p; p += 32 * (size_t)P.edgeSlots;
  S.emx = (uint32_t*)p; p += 8 * (size_t)P.edgeSlots;
  S.rcnt = (uint32_t*)p; p += 4 * (size_t)P.regSlots;
  S.rhist = (uint32_t*)p; p += 32 * (size_t)P.regSlots;
  S.rmx = (uint32_t*)p; p += 32 * (size_t)P.regSlots;
  S.qn = (uint32_t*)p; p += 4 * 32;
  S.sRB = (float*)p; p += 64;
  S.sEB = (float*)p;
  return S;
}

size_t voxel_fast_smem_bytes(const FastParams& P) {
  int nwarps = P.warpsX * P.warpsY;
  return sizeof(QEntry) * QCAP * nwarps + (size_t)P.edgeSlots * (8 + 16 + 8 + 32 + 8) +
         (size_t)P.regSlots * (16 + 4 + 4 + 32 + 32) + 128 + 128 + 64;
}

// the whole warp applies its queued run records (one entry per lane)
__device__ __forceinline__ void drain_queue(const SmemView& S, int warp, int lane, int x0warp, int gy, int z0) {
  __syncwarp();
  const uint32_t n = S.qn[warp];
  const QEntry* q = S.queue + (size_t)warp * QCAP;
  for (uint32_t i = lane; i < n; i += 32) {
    const QEntry e = q[i];
    const uint32_t slot = e.slot & ~Q_EDGE;
    if (e.slot & Q_EDGE) {
      atomicAdd(S.esum + 2 * slot, e.s); atomicAdd(S.esum + 2 * slot + 1, e.s2);
      if (e.kmin) smax(S.emx + 2 * slot, e.kmin);
      if (e.kmax) smax(S.emx + 2 * slot + 1, e.kmax);
    } else {
      atomicAdd(S.rsum + 2 * slot, e.s); atomicAdd(S.rsum + 2 * slot + 1, e.s2);
      uint32_t* m = S.rmx + 8 * slot;
      if (e.aux & 0x10000u) {
        const uint32_t gx = (uint32_t)(x0warp + (int)(e.aux >> 24));
        const uint32_t zmn = (uint32_t)(z0 + (int)(e.aux & 0xFFu)), zmx = (uint32_t)(z0 + (int)((e.aux >> 8) & 0xFFu));
        smax(m + 0, ~gx); smax(m + 1, ~(uint32_t)gy); smax(m + 2, ~zmn);
        smax(m + 3, gx + 1); smax(m + 4, (uint32_t)gy + 1); smax(m + 5, zmx + 1);
      }
      if (e.kmin) smax(m + 6, e.kmin);
      if (e.kmax) smax(m + 7, e.kmax);
    }
  }
  __syncwarp();
  if (lane == 0) S.qn[warp] = 0;
  __syncwarp();
}

template <int MINB>
__global__ void __launch_bounds__(256, MINB)
voxel_pass_fast(const __grid_constant__ FastParams P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nthreads = blockDim.x;
  const SmemView S = carve(smem_raw, P, nthreads >> 5);
  {  // slot counts are multiples of 4, so both blocks are whole 16-byte words (carve)
    uint4* ones = reinterpret_cast<uint4*>(S.ekeys);
    const int nOnes = (8 * P.edgeSlots + 4 * P.regSlots) >> 4;
    for (int i = tid; i < nOnes; i += nthreads) ones[i] = make_uint4(~0u, ~0u, ~0u, ~0u);
    uint4* zeros = reinterpret_cast<uint4*>(S.esum);
    const int nZeros = (64 * P.edgeSlots + 84 * P.regSlots) >> 4;
    for (int i = tid; i < nZeros; i += nthreads) zeros[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  if (tid < 32) S.qn[tid] = 0;
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
  const uint32_t rmask = (uint32_t)(P.regSlots - 1), emask = (uint32_t)(P.edgeSlots - 1);
  QEntry* myq = S.queue + (size_t)warp * QCAP;

  // run state: double sums / min / max of the current region slot and of the current edge slot
  int rCur = -1; double rS = 0.0, rS2 = 0.0; uint32_t rKmin = 0, rKmax = 0; int rZmin = 0, rZmax = -1;
  int eCur = -1; double eS = 0.0, eS2 = 0.0; uint32_t eKmin = 0, eKmax = 0;

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
        bool found = false; uint32_t nb = 0;
        if (vzp && lp != me) { found = true; nb = lp; }
        if (vzm && lm != me) { found = true; nb = lm; }
        if (vyp && yp != me) { found = true; nb = yp; }
        if (vym && ym != me) { found = true; nb = ym; }
        if (vxp && xp != me) { found = true; nb = xp; }
        if (vxm && xm != me) { found = true; nb = xm; }
        const bool isBorder = !found && (nvalidXY + (int)vzm + (int)vzp) < twoD;  // util/struct.hxx:103-106
        const bool nonInterior = found || isBorder;
        const int binR = bin_fast(pv, P.rLo, P.rHi, P.rInv, P.rBins, S.sRB);
        const double dv = (double)pv, dv2 = dv * dv;
        const bool notNan = pv == pv;
        const uint32_t kord = f2ord(pv);
        // ---- region: slot lookup, integer statistics per voxel
        int rslot = -1;
        {
          uint32_t hsh = ((me * 2654435761u) >> 9) & rmask;
          for (int probe = 0; probe < 8; ++probe) {
            uint32_t cur = S.rkeys[hsh];
            if (cur == 0xFFFFFFFFu) cur = atomicCAS(S.rkeys + hsh, 0xFFFFFFFFu, me);
            if (cur == me || cur == 0xFFFFFFFFu) { rslot = (int)hsh; break; }
            hsh = (hsh + 1) & rmask;
          }
        }
        if (rslot >= 0) {
          atomicAdd(S.rcnt + rslot, isBorder ? 0x10001u : 1u);
          if (binR >= 0) atomicAdd(S.rhist + 8 * rslot + (binR >> 1), 1u << (16 * (binR & 1)));
        } else {
          region_voxel_global(P, me, pv, binR, isBorder, nonInterior, gx, gy, gz);
        }
        if (rslot != rCur) {
          if (rCur >= 0) {
            uint32_t pos = atomicAdd(S.qn + warp, 1u);
            QEntry e; e.slot = (uint32_t)rCur; e.kmin = rKmin; e.kmax = rKmax;
            e.aux = (rZmax >= 0 ? (uint32_t)(rZmin - z0) | ((uint32_t)(rZmax - z0) << 8) | 0x10000u : 0u) | ((uint32_t)lane << 24);
            e.s = rS; e.s2 = rS2;
            myq[pos] = e;
          }
          rCur = rslot; rS = 0.0; rS2 = 0.0; rKmin = 0; rKmax = 0; rZmax = -1;
        }
        if (rslot >= 0) {
          rS += dv; rS2 += dv2;
          if (notNan) { rKmin = max(rKmin, ~kord); rKmax = max(rKmax, kord); }
          if (nonInterior) { if (rZmax < 0) rZmin = gz; rZmax = gz; }
        }
        // ---- directed boundary (me -> nb)
        int eslot = -1;
        if (found) {
          const unsigned long long key = ((unsigned long long)me << 32) | nb;
          uint32_t hsh = ((me * 0x9E3779B1u) ^ (nb * 0x85EBCA77u)) >> 11 & emask;
          for (int probe = 0; probe < 8; ++probe) {
            unsigned long long cur = S.ekeys[hsh];
            if (cur == EDGE_EMPTY) cur = atomicCAS(S.ekeys + hsh, EDGE_EMPTY, key);
            if (cur == key || cur == EDGE_EMPTY) { eslot = (int)hsh; break; }
            hsh = (hsh + 1) & emask;
          }
          const int binE = P.sameSpec ? binR : bin_fast(pv, P.eLo, P.eHi, P.eInv, P.eBins, S.sEB);
          if (eslot >= 0) {
            atomicAdd(S.ecnt + 2 * eslot, (P.nThr > 0 && pv >= P.thr_up[0]) ? 0x10001u : 1u);  // type/feat.hxx:498
            const uint32_t t12 = ((P.nThr > 1 && pv >= P.thr_up[1]) ? 1u : 0u) | ((P.nThr > 2 && pv >= P.thr_up[2]) ? 0x10000u : 0u);
            if (t12) atomicAdd(S.ecnt + 2 * eslot + 1, t12);
            if (binE >= 0) atomicAdd(S.ehist + 8 * eslot + (binE >> 1), 1u << (16 * (binE & 1)));
          } else {
            edge_voxel_global(P, key, pv, binE);
          }
        }
        if (found && eslot != eCur) {
          if (eCur >= 0) {
            uint32_t pos = atomicAdd(S.qn + warp, 1u);
            QEntry e; e.slot = (uint32_t)eCur | Q_EDGE; e.kmin = eKmin; e.kmax = eKmax; e.aux = 0; e.s = eS; e.s2 = eS2;
            myq[pos] = e;
          }
          eCur = eslot; eS = 0.0; eS2 = 0.0; eKmin = 0; eKmax = 0;
        }
        if (eslot >= 0) {
          eS += dv; eS2 += dv2;
          if (notNan) { eKmin = max(eKmin, ~kord); eKmax = max(eKmax, kord); }
        }
      }
      // one decision for the whole warp: lane 0's view, taken after every lane's pushes (the
      // shuffle is the barrier; a plain per-lane read races with fast lanes' next-step pushes)
      __syncwarp();
      if (__shfl_sync(0xffffffffu, S.qn[warp], 0) >= 32u) drain_queue(S, warp, lane, x0warp, gy, z0);
      lm = l0; l0 = lp; lp = nlp; ym = nym; yp = nyp; xe = nxe; pv = npv;
    }
    // close the open runs and drain
    if (active && rCur >= 0) {
      uint32_t pos = atomicAdd(S.qn + warp, 1u);
      QEntry e; e.slot = (uint32_t)rCur; e.kmin = rKmin; e.kmax = rKmax;
      e.aux = (rZmax >= 0 ? (uint32_t)(rZmin - z0) | ((uint32_t)(rZmax - z0) << 8) | 0x10000u : 0u) | ((uint32_t)lane << 24);
      e.s = rS; e.s2 = rS2;
      myq[pos] = e;
    }
    if (active && eCur >= 0) {
      uint32_t pos = atomicAdd(S.qn + warp, 1u);
      QEntry e; e.slot = (uint32_t)eCur | Q_EDGE; e.kmin = eKmin; e.kmax = eKmax; e.aux = 0; e.s = eS; e.s2 = eS2;
      myq[pos] = e;
    }
    drain_queue(S, warp, lane, x0warp, gy, z0);
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
    long long g = edge_slot_global_fast(P, key);
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

cudaError_t launch_voxel_pass_fast(const FastParams& P, cudaStream_t st) {
  size_t smem = voxel_fast_smem_bytes(P);
  static const int minb = getenv("GLIA_B200_MINB") ? atoi(getenv("GLIA_B200_MINB")) : 4;
  cudaError_t e = minb >= 4 ? cudaFuncSetAttribute(voxel_pass_fast<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)
                            : cudaFuncSetAttribute(voxel_pass_fast<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  long long tiles = (long long)P.tilesX * P.tilesY * P.tilesZ;
  if (tiles <= 0) return cudaSuccess;
  if (minb >= 4) voxel_pass_fast<4><<<(unsigned)tiles, 32 * P.warpsX * P.warpsY, smem, st>>>(P);
  else voxel_pass_fast<3><<<(unsigned)tiles, 32 * P.warpsX * P.warpsY, smem, st>>>(P);
  return cudaGetLastError();
}

}  // namespace glia
