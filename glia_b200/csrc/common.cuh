// glia_b200 — internal definitions shared by the CUDA translation units.
//
// Data model in HBM.  Every statistic the reference derives by re-walking
// pixel lists is kept as a MERGEABLE record, one per base region and one per directed boundary
// (a -> b), split by reduction kind so that zero-initialised memory is the identity everywhere
// and a multi-GPU reduce is one collective per kind:
//   add  words (u32, +)   : counts and histogram bins
//   max  words (u32, max) : bbox and float min/max, encoded so that 0 == "no voxel yet"
//   f64  words (double, +): sum v, sum v^2
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include "../../include/glia_b200.h"

#define GLIA_MAX_CH (2 * GLIA_MAX_IMG)

namespace glia {

// One statistics channel = one (image, histogram spec) pair of a record.
struct Channel {
  const float* data;
  int bins;
  int addOff;  // first histogram bin inside the add words
  int maxOff;  // (~min, max) pair inside the max words
  int f64Off;  // (sum, sum2) pair inside the f64 words
  // util/image_stats.hxx:12-36 in float arithmetic: for a float v and a double bound b,
  // (double)v < b  <=>  v < ceil_f32(b), (double)v > b <=> v > floor_f32(b) (host-rounded).
  float lo_dn, hi_up;
  float bound_up[GLIA_MAX_BINS];
};

struct RecLayout {
  int nAdd, nMax, nF64;
};

// Kernel parameters of the voxel pass (passed by value, < 4 KB).
struct VoxParams {
  const uint32_t* labels;
  const uint32_t* mask;  // may be null
  const float* pb;
  int ndim;
  int contourOnly;  // glia_params.contour_only: masked-out voxels are still source voxels (the mask hides neighbours only)
  int X, Y, Z;      // full (global) extent; Z == 1 and Y == rows when ndim == 2 (see voxel_pass.cu)
  int zBegin, zEnd; // owned planes (global z)
  int zBuf0;        // global z of buffer plane 0
  int zBufN;        // planes in the buffer
  int TX, TY, TZ;   // tile extent (threads = TX*TY, each marches TZ planes)
  int SXp;          // row pitch of the staged label tile: TX + 8 (the tile starts 4 voxels left of the CTA's columns)
  int tilesX, tilesY, tilesZ;
  int nThr;
  float thr_up[GLIA_MAX_THR];  // (double)pb >= t  <=>  pb >= ceil_f32(t)
  int nRegCh, nEdgeCh;
  RecLayout reg, edge;
  // global tables
  uint32_t labelBound;
  uint32_t* regAdd; uint32_t* regMax; double* regF64;  // [labelBound][n*]
  uint64_t edgeCap;                                    // power of two
  unsigned long long* edgeKeys;                        // [edgeCap], EMPTY = ~0
  uint32_t* edgeAdd; uint32_t* edgeMax; double* edgeF64;
  int* overflow;   // bit flags: 1 = boundary table full, 2 = label >= labelBound (or 0xFFFFFFFF), 4 = value outside the fast kernel's domain
  // shared-memory privatisation
  int regSlots, edgeSlots;  // powers of two
};

struct VoxChannels {
  Channel reg[GLIA_MAX_CH];   // rimg then rlimg
  Channel edge[GLIA_MAX_IMG]; // bimg
};

static const unsigned long long EDGE_EMPTY = ~0ull;

// Order-preserving float <-> u32 key (NaN never enters: callers skip it).
__host__ __device__ inline uint32_t f2ord(float f) {
#ifdef __CUDA_ARCH__
  uint32_t b = __float_as_uint(f);
#else
  union { float f; uint32_t u; } c; c.f = f; uint32_t b = c.u;
#endif
  return b ^ ((b >> 31) ? 0xFFFFFFFFu : 0x80000000u);
}
__host__ __device__ inline float ord2f(uint32_t k) {
  uint32_t b = k ^ ((k >> 31) ? 0x80000000u : 0xFFFFFFFFu);
#ifdef __CUDA_ARCH__
  return __uint_as_float(b);
#else
  union { float f; uint32_t u; } c; c.u = b; return c.f;
#endif
}

__host__ __device__ inline uint64_t mix64(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return x;
}

#define GLIA_CUDA_CHECK(h, expr)                                                        \
  do {                                                                                  \
    cudaError_t e_ = (expr);                                                            \
    if (e_ != cudaSuccess) {                                                            \
      (h)->err = std::string(#expr) + ": " + cudaGetErrorString(e_);                    \
      return GLIA_E_CUDA;                                                               \
    }                                                                                   \
  } while (0)

}  // namespace glia
