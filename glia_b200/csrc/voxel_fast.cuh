// Parameters of the fast voxel pass (voxel_fast.cu).
#pragma once
#include "common.cuh"

namespace glia {

#define GLIA_FAST_MAX_CLS 24  // value classes of the joint (histogram x threshold) partition, 2 per table word

// Every per-voxel integer statistic of the default hmt configuration (histogram bin of the region
// channel, histogram bin of the boundary channel, pb >= t for each threshold) is a function of the
// position of the voxel's value among the sorted union of all comparison constants.  The host builds
// that partition once (api_build.cu, build_value_classes); the kernel counts voxels per class and the
// tile flush maps class counts back to the record words.
struct ValueClasses {
  int nCls;                         // classes 0..nCls-1; class = #{j : v >= bound[j]}
  float bound[GLIA_FAST_MAX_CLS];   // ascending, nCls-1 used
  float lutLo, lutScale;            // cell = (v - lutLo) * lutScale over 256 cells
  unsigned char lut[256];           // lower estimate of the class of a value in cell q+1 (see kernel)
  signed char rBin[GLIA_FAST_MAX_CLS];   // histogram bin of the region channel (-1: counted in no bin)
  signed char eBin[GLIA_FAST_MAX_CLS];   // histogram bin of the boundary channel
  unsigned char thr[GLIA_FAST_MAX_CLS];  // bit k: v >= threshold k
};

struct FastParams {
  const uint32_t* labels;
  const float* pb;
  int ndim;
  int X, Y, Z;
  int zBegin, zEnd, zBuf0;
  int warpsX, warpsY;  // CTA footprint: 32*warpsX by warpsY
  int TZ;              // planes (2D: rows) per tile; warpsY + TZ <= 32
  int tilesX, tilesY, tilesZ;
  int nThr;
  int rBins, eBins;    // histogram bins of the region / boundary channel
  ValueClasses vc;
  RecLayout reg, edge; // global layouts (region: add = [cnt, border, hist..], max = [bbox6, min~, max])
  uint32_t labelBound;
  uint32_t* regAdd; uint32_t* regMax; double* regF64;
  uint64_t edgeCap;
  unsigned long long* edgeKeys;
  uint32_t* edgeAdd; uint32_t* edgeMax; double* edgeF64;
  int* overflow;       // bit flags: 1 boundary table full, 2 label >= labelBound, 4 value outside [0, 1]
  int slots;           // shared-memory table slots (power of two)
};

}  // namespace glia
