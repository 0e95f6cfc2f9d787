// K1+K3+K4 fused: one pass over the label / probability volumes.
//
// Replaces (reference, /root/reference/code):
//   getContourTraits / getNeighborValues          type/neighbor.hxx:72-126
//   getContour / genContourMap                    util/struct.hxx:95-144
//   genCountMap / genPointMap                     util/struct.hxx:61-92
//   RegionShapeFeats::generate (area, bbox)       type/feat.hxx:71-90, alg/geometry.hxx:21-39
//   ImageLabelFeats / ImageRealFeats::generate    type/feat.hxx:632-638, :706-738
//   ImageRegionShapeFeats::generate (pb >= t)     type/feat.hxx:485-502
//   stats::hist bin rule                          util/image_stats.hxx:12-36
// The reference materialises a Point object per voxel and re-walks the lists per candidate;
// here every voxel is read once and folded into mergeable records (common.cuh).
//
// GENERIC kernel: any number of images/bins/thresholds, mask, 2D/3D, z-slab.  A CTA owns a
// TX x TY x TZ tile; the label tile (+1 halo) and a validity tile (in-image and unmasked) are
// staged in shared memory; statistics are privatised in shared-memory hash tables keyed by
// label and by directed label pair and flushed once per tile with global atomics.  Keys that
// do not fit the shared tables fall through to the global tables directly.
#include <string.h>
#include <cuda.h>  // CUtensorMap (type only: the encoder is fetched through cudaGetDriverEntryPoint, api_build.cu)
#include "common.cuh"

namespace glia {

#define GLIA_TILE_XOFF 4
// ---- TMA + mbarrier (sm_100a): the label (and mask) tile with its one-voxel halo is ONE bulk tensor copy issued by
// one thread; out-of-volume elements are zero-filled by the copy engine, completion is signalled on an mbarrier.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::
          "r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  } while (!ok);
}

__device__ __forceinline__ int bin_of(const Channel& c, float v) {
  if (v > c.lo_dn && v < c.hi_up) {
    for (int i = 0; i < c.bins; ++i)
      if (v < c.bound_up[i]) return i;
    return -1;  // in range but beyond the last cumulative bound: counted in n, in no bin
  }
  return (v <= c.lo_dn) ? 0 : c.bins - 1;
}

__device__ __forceinline__ void max_if(uint32_t* w, uint32_t v) {
  if (*(volatile uint32_t*)w < v) atomicMax(w, v);
}

__device__ __forceinline__ void acc_channel(const Channel& c, size_t p, uint32_t* add,
                                            uint32_t* mx, double* f64) {
  float v = __ldg(c.data + p);
  int b = bin_of(c, v);
  if (b >= 0) atomicAdd(add + c.addOff + b, 1u);
  atomicAdd(f64 + c.f64Off, (double)v);
  atomicAdd(f64 + c.f64Off + 1, (double)v * (double)v);
  if (v == v) {
    uint32_t k = f2ord(v);
    max_if(mx + c.maxOff, ~k);
    max_if(mx + c.maxOff + 1, k);
  }
}

__device__ __forceinline__ long long edge_slot_global(const VoxParams& P, unsigned long long key) {
  uint64_t mask = P.edgeCap - 1;
  uint64_t h = mix64(key) & mask;
  for (uint64_t probe = 0; probe < P.edgeCap; ++probe) {
    unsigned long long cur = P.edgeKeys[h];
    if (cur == key) return (long long)h;
    if (cur == EDGE_EMPTY) {
      unsigned long long prev = atomicCAS(P.edgeKeys + h, EDGE_EMPTY, key);
      if (prev == EDGE_EMPTY || prev == key) return (long long)h;
    }
    h = (h + 1) & mask;
    if (probe > 4096) break;
  }
  atomicOr(P.overflow, 1);
  return -1;
}

__device__ __forceinline__ void flush_rec(const RecLayout& L, const uint32_t* sadd,
                                          const uint32_t* smx, const double* sf, uint32_t* gadd,
                                          uint32_t* gmx, double* gf) {
  for (int i = 0; i < L.nAdd; ++i) { uint32_t v = sadd[i]; if (v) atomicAdd(gadd + i, v); }
  for (int i = 0; i < L.nMax; ++i) { uint32_t v = smx[i]; if (v) atomicMax(gmx + i, v); }
  for (int i = 0; i < L.nF64; ++i) { double v = sf[i]; if (v != 0.0) atomicAdd(gf + i, v); }
}

template <bool TMA>
__global__ void __launch_bounds__(256)
voxel_pass_generic(const __grid_constant__ VoxParams P, const __grid_constant__ VoxChannels C,
                   const __grid_constant__ CUtensorMap tmLab, const __grid_constant__ CUtensorMap tmMask) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int hy = (P.ndim == 2) ? 0 : 1;
  // The staged tile starts GLIA_TILE_XOFF = 4 voxels left of the CTA's first column and is TX + 8 wide: a TMA tensor copy
  // needs its innermost start coordinate on a 16-byte boundary (measured: profiles/microbench/tma_probe.cu - a start of
  // x0 - 1 raises "illegal instruction", x0 - 4 with zero-filled out-of-volume elements is fine).
  const int SX = P.SXp, SY = P.TY + 2 * hy, SZ = P.TZ + 2;
  const int ncell = SX * SY * SZ;
  // carve shared memory: the TMA destinations first (128-byte aligned), then 8-byte items
  unsigned char* sp = smem_raw;
  uint32_t* lab = (uint32_t*)sp; sp += ((4 * (size_t)ncell + 127) / 128) * 128;
  uint32_t* msk = (uint32_t*)sp; if (TMA && P.mask) sp += ((4 * (size_t)ncell + 127) / 128) * 128;
  __shared__ __align__(8) uint64_t mbar;
  unsigned long long* ekeys = (unsigned long long*)sp; sp += sizeof(unsigned long long) * P.edgeSlots;
  double* ef64 = (double*)sp; sp += sizeof(double) * (size_t)P.edgeSlots * P.edge.nF64;
  double* rf64 = (double*)sp; sp += sizeof(double) * (size_t)P.regSlots * P.reg.nF64;
  uint32_t* eadd = (uint32_t*)sp; sp += 4 * (size_t)P.edgeSlots * P.edge.nAdd;
  uint32_t* emax = (uint32_t*)sp; sp += 4 * (size_t)P.edgeSlots * P.edge.nMax;
  uint32_t* radd = (uint32_t*)sp; sp += 4 * (size_t)P.regSlots * P.reg.nAdd;
  uint32_t* rmax = (uint32_t*)sp; sp += 4 * (size_t)P.regSlots * P.reg.nMax;
  uint32_t* rkeys = (uint32_t*)sp; sp += 4 * (size_t)P.regSlots;
  unsigned char* valid = sp;

  const int nthreads = blockDim.x;
  const int tid = threadIdx.x;
  int t = blockIdx.x;
  const int tileX = t % P.tilesX; t /= P.tilesX;
  const int tileY = t % P.tilesY; t /= P.tilesY;
  const int tileZ = t;
  const int x0 = tileX * P.TX, y0 = tileY * P.TY, z0 = P.zBegin + tileZ * P.TZ;
  const size_t strideY = (size_t)P.X, strideZ = (size_t)P.X * P.Y;

  if (TMA) {
    if (tid == 0) mbar_init(&mbar, 1);
    __syncthreads();
    if (tid == 0) {  // one thread issues the bulk tensor copies; everybody else initialises the tables meanwhile
      const uint32_t bytes = 4u * (uint32_t)ncell;
      mbar_expect_tx(&mbar, P.mask ? 2u * bytes : bytes);
      tma_load_3d(lab, &tmLab, &mbar, x0 - GLIA_TILE_XOFF, y0 - hy, z0 - 1 - P.zBuf0);
      if (P.mask) tma_load_3d(msk, &tmMask, &mbar, x0 - GLIA_TILE_XOFF, y0 - hy, z0 - 1 - P.zBuf0);
    }
  }
  // ---- init shared tables
  for (int i = tid; i < P.edgeSlots; i += nthreads) ekeys[i] = EDGE_EMPTY;
  for (int i = tid; i < P.edgeSlots * P.edge.nF64; i += nthreads) ef64[i] = 0.0;
  for (int i = tid; i < P.regSlots * P.reg.nF64; i += nthreads) rf64[i] = 0.0;
  for (int i = tid; i < P.edgeSlots * P.edge.nAdd; i += nthreads) eadd[i] = 0;
  for (int i = tid; i < P.edgeSlots * P.edge.nMax; i += nthreads) emax[i] = 0;
  for (int i = tid; i < P.regSlots * P.reg.nAdd; i += nthreads) radd[i] = 0;
  for (int i = tid; i < P.regSlots * P.reg.nMax; i += nthreads) rmax[i] = 0;
  for (int i = tid; i < P.regSlots; i += nthreads) rkeys[i] = 0xFFFFFFFFu;
  if (TMA) {
    mbar_wait(&mbar, 0);
    // validity: inside the volume, inside the buffer (slab + halo), unmasked.  Labels outside were zero-filled.
    for (int i = tid; i < ncell; i += nthreads) {
      int cx = i % SX, r = i / SX;
      int cy = r % SY, cz = r / SY;
      int gx = x0 - GLIA_TILE_XOFF + cx, gy = y0 - hy + cy, gz = z0 - 1 + cz;
      int bz = gz - P.zBuf0;
      bool ok = gx >= 0 && gx < P.X && gy >= 0 && gy < P.Y && gz >= 0 && gz < P.Z && bz >= 0 && bz < P.zBufN;
      // 1 = in the volume and unmasked, 2 = in the volume but masked out (a source voxel only in contour-only mode)
      valid[i] = !ok ? 0 : ((P.mask && msk[i] == 0) ? 2 : 1);
    }
  } else {
    // ---- stage label + validity tile (halo 1)
    for (int i = tid; i < ncell; i += nthreads) {
      int cx = i % SX, r = i / SX;
      int cy = r % SY, cz = r / SY;
      int gx = x0 - GLIA_TILE_XOFF + cx, gy = y0 - hy + cy, gz = z0 - 1 + cz;
      int bz = gz - P.zBuf0;
      bool ok = gx >= 0 && gx < P.X && gy >= 0 && gy < P.Y && gz >= 0 && gz < P.Z && bz >= 0 &&
                bz < P.zBufN;
      uint32_t l = 0;
      unsigned char v = 0;
      if (ok) {
        size_t q = (size_t)bz * strideZ + (size_t)gy * strideY + gx;
        v = (P.mask && __ldg(P.mask + q) == 0) ? 2 : 1;
        if (v == 1 || P.contourOnly) l = __ldg(P.labels + q);
      }
      lab[i] = l;
      valid[i] = v;
    }
  }
  __syncthreads();

  // ---- march the thread's column through the tile
  const int tx = tid % P.TX, ty = tid / P.TX;
  const int gx = x0 + tx, gy = y0 + ty;
  if (gx < P.X && gy < P.Y && ty < P.TY) {
    uint32_t curLabel = 0xFFFFFFFFu;
    uint32_t* ra = nullptr; uint32_t* rm = nullptr; double* rf = nullptr;
    unsigned long long curKey = EDGE_EMPTY;
    uint32_t* ea = nullptr; uint32_t* em = nullptr; double* ef = nullptr;
    const int twoD = 2 * P.ndim;
    for (int zz = 0; zz < P.TZ; ++zz) {
      int gz = z0 + zz;
      if (gz >= P.zEnd) break;
      int c = ((zz + 1) * SY + (ty + hy)) * SX + tx + GLIA_TILE_XOFF;
      // contour-only region maps classify masked-out voxels too (genContourMap over the image, util/struct.hxx:129-146)
      if (!(valid[c] == 1 || (valid[c] == 2 && P.contourOnly))) continue;
      uint32_t me = lab[c];
      if (me == 0xFFFFFFFFu) { atomicOr(P.overflow, 2); continue; }  // reserved: the empty-slot marker of the tables
      // type/neighbor.hxx:78-88 order: d0-, d0+, d1-, d1+, d2-, d2+
      int nvalid = 0; bool found = false; uint32_t nb = 0;
#define GLIA_VISIT(off)                                                   \
  { int q_ = c + (off); if (valid[q_] == 1) { ++nvalid; uint32_t l_ = lab[q_]; \
      if (!found && l_ != me) { found = true; nb = l_; } } }
      GLIA_VISIT(-1) GLIA_VISIT(+1)
      if (hy) { GLIA_VISIT(-SX) GLIA_VISIT(+SX) }
      GLIA_VISIT(-SX * SY) GLIA_VISIT(+SX * SY)
#undef GLIA_VISIT
      const bool isBorder = !found && nvalid < twoD;  // util/struct.hxx:103-106
      size_t p = (size_t)(gz - P.zBuf0) * strideZ + (size_t)gy * strideY + gx;
      // ---- region record
      if (me != curLabel) {
        curLabel = me;
        ra = nullptr;
        uint32_t h = (me * 2654435761u) & (uint32_t)(P.regSlots - 1);
        for (int probe = 0; probe < 8; ++probe) {
          uint32_t cur = rkeys[h];
          if (cur != me && cur == 0xFFFFFFFFu) cur = atomicCAS(rkeys + h, 0xFFFFFFFFu, me);
          if (cur == me || cur == 0xFFFFFFFFu) {
            ra = radd + (size_t)h * P.reg.nAdd; rm = rmax + (size_t)h * P.reg.nMax;
            rf = rf64 + (size_t)h * P.reg.nF64;
            break;
          }
          h = (h + 1) & (uint32_t)(P.regSlots - 1);
        }
        if (!ra) {
          if (me >= P.labelBound) { atomicOr(P.overflow, 2); curLabel = 0xFFFFFFFFu; continue; }
          ra = P.regAdd + (size_t)me * P.reg.nAdd; rm = P.regMax + (size_t)me * P.reg.nMax;
          rf = P.regF64 + (size_t)me * P.reg.nF64;
        }
      }
      atomicAdd(ra + 0, 1u);
      if (isBorder) atomicAdd(ra + 1, 1u);
      if (found || isBorder) {  // the extreme voxels of a region are never interior
        max_if(rm + 0, ~(uint32_t)gx); max_if(rm + 1, ~(uint32_t)gy); max_if(rm + 2, ~(uint32_t)gz);
        max_if(rm + 3, (uint32_t)gx + 1); max_if(rm + 4, (uint32_t)gy + 1); max_if(rm + 5, (uint32_t)gz + 1);
      }
      for (int ch = 0; ch < P.nRegCh; ++ch) acc_channel(C.reg[ch], p, ra, rm, rf);
      // ---- directed boundary record (me -> nb)
      if (found) {
        unsigned long long key = ((unsigned long long)me << 32) | nb;
        if (key != curKey) {
          curKey = key;
          ea = nullptr;
          uint32_t h = (uint32_t)mix64(key) & (uint32_t)(P.edgeSlots - 1);
          for (int probe = 0; probe < 8; ++probe) {
            unsigned long long cur = ekeys[h];
            if (cur != key && cur == EDGE_EMPTY) cur = atomicCAS(ekeys + h, EDGE_EMPTY, key);
            if (cur == key || cur == EDGE_EMPTY) {
              ea = eadd + (size_t)h * P.edge.nAdd; em = emax + (size_t)h * P.edge.nMax;
              ef = ef64 + (size_t)h * P.edge.nF64;
              break;
            }
            h = (h + 1) & (uint32_t)(P.edgeSlots - 1);
          }
          if (!ea) {
            long long g = edge_slot_global(P, key);
            if (g < 0) { curKey = EDGE_EMPTY; continue; }
            ea = P.edgeAdd + (size_t)g * P.edge.nAdd; em = P.edgeMax + (size_t)g * P.edge.nMax;
            ef = P.edgeF64 + (size_t)g * P.edge.nF64;
          }
        }
        atomicAdd(ea + 0, 1u);
        float pbv = __ldg(P.pb + p);
        for (int k = 0; k < P.nThr; ++k)
          if (pbv >= P.thr_up[k]) atomicAdd(ea + 1 + k, 1u);  // type/feat.hxx:498
        for (int ch = 0; ch < P.nEdgeCh; ++ch) acc_channel(C.edge[ch], p, ea, em, ef);
      }
    }
  }
  __syncthreads();
  // ---- flush the shared tables
  for (int s = tid; s < P.regSlots; s += nthreads) {
    uint32_t l = rkeys[s];
    if (l == 0xFFFFFFFFu) continue;
    if (l >= P.labelBound) { atomicOr(P.overflow, 2); continue; }
    flush_rec(P.reg, radd + (size_t)s * P.reg.nAdd, rmax + (size_t)s * P.reg.nMax,
              rf64 + (size_t)s * P.reg.nF64, P.regAdd + (size_t)l * P.reg.nAdd,
              P.regMax + (size_t)l * P.reg.nMax, P.regF64 + (size_t)l * P.reg.nF64);
  }
  for (int s = tid; s < P.edgeSlots; s += nthreads) {
    unsigned long long key = ekeys[s];
    if (key == EDGE_EMPTY) continue;
    long long g = edge_slot_global(P, key);
    if (g < 0) continue;
    flush_rec(P.edge, eadd + (size_t)s * P.edge.nAdd, emax + (size_t)s * P.edge.nMax,
              ef64 + (size_t)s * P.edge.nF64, P.edgeAdd + (size_t)g * P.edge.nAdd,
              P.edgeMax + (size_t)g * P.edge.nMax, P.edgeF64 + (size_t)g * P.edge.nF64);
  }
}

// Largest label value (only used when the caller gives no label_bound).
__global__ void label_max_kernel(const uint32_t* __restrict__ labels, size_t n, uint32_t* out) {
  uint32_t m = 0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x)
    m = max(m, __ldg(labels + i));
  for (int o = 16; o; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0 && m) atomicMax(out, m);
}

size_t voxel_pass_smem_bytes(const VoxParams& P, bool tmaMask) {
  const int hy = (P.ndim == 2) ? 0 : 1;
  size_t ncell = (size_t)P.SXp * (P.TY + 2 * hy) * (P.TZ + 2);
  size_t tile = ((4 * ncell + 127) / 128) * 128;
  size_t b = tile * (tmaMask ? 2 : 1) + 8 * (size_t)P.edgeSlots + 8 * (size_t)P.edgeSlots * P.edge.nF64 +
             8 * (size_t)P.regSlots * P.reg.nF64 + 4 * (size_t)P.edgeSlots * (P.edge.nAdd + P.edge.nMax) +
             4 * (size_t)P.regSlots * (P.reg.nAdd + P.reg.nMax + 1) + ncell;
  return (b + 15) & ~(size_t)15;
}

// tm: label / mask tensor maps (api_build.cu) or null -> the tile is staged with plain loads
cudaError_t launch_voxel_pass_generic(const VoxParams& P, const VoxChannels& C, const CUtensorMap* tm, cudaStream_t st) {
  const bool tma = tm != nullptr;
  size_t smem = voxel_pass_smem_bytes(P, tma && P.mask);
  cudaError_t e = tma ? cudaFuncSetAttribute(voxel_pass_generic<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)
                      : cudaFuncSetAttribute(voxel_pass_generic<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  long long tiles = (long long)P.tilesX * P.tilesY * P.tilesZ;
  if (tiles <= 0) return cudaSuccess;
  if (tma) voxel_pass_generic<true><<<(unsigned)tiles, P.TX * P.TY, smem, st>>>(P, C, tm[0], tm[1]);
  else { CUtensorMap dummy; memset(&dummy, 0, sizeof(dummy)); voxel_pass_generic<false><<<(unsigned)tiles, P.TX * P.TY, smem, st>>>(P, C, dummy, dummy); }
  return cudaGetLastError();
}

cudaError_t launch_label_max(const uint32_t* labels, size_t n, uint32_t* out, cudaStream_t st) {
  if (!n) return cudaSuccess;
  int blocks = (int)((n + 1023) / 1024 < 148 * 8 ? (n + 1023) / 1024 : 148 * 8);
  label_max_kernel<<<blocks, 256, 0, st>>>(labels, n, out);
  return cudaGetLastError();
}

}  // namespace glia
