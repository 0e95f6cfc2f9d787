// Parameters of the fast voxel pass (voxel_fast.cu).
#pragma once
#include "common.cuh"

namespace glia {

struct FastParams {
  const uint32_t* labels;
  const float* pb;
  int ndim;
  int X, Y, Z;
  int zBegin, zEnd, zBuf0;
  int warpsX, warpsY;  // CTA footprint: 32*warpsX by warpsY
  int TZ;
  int tilesX, tilesY, tilesZ;
  int nThr;
  float thr_up[4];
  int rBins, eBins;    // histogram bins of the region / boundary channel
  int sameSpec;        // both channels share one histogram spec
  float rLo, rHi, eLo, eHi, rInv, eInv;
  float rBound[16], eBound[16];
  RecLayout reg, edge; // global layouts (region: add = [cnt, border, hist..], max = [bbox6, min~, max])
  uint32_t labelBound;
  uint32_t* regAdd; uint32_t* regMax; double* regF64;
  uint64_t edgeCap;
  unsigned long long* edgeKeys;
  uint32_t* edgeAdd; uint32_t* edgeMax; double* edgeF64;
  int* overflow;
  int regSlots, edgeSlots;  // powers of two
};

}  // namespace glia
