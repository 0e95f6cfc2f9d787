// Seeded synthetic EM-like volume generator on the device: jittered-grid
// Voronoi superpixels + dyadic boundary-probability map.  Same integer hashing as
// glia_b200/synth.py::voronoi_volume, so labels are identical to the numpy generator (pb may
// differ by one 1/256 step at isolated voxels where exp() rounds differently).
// Bench/test utility only: it produces inputs, it is not part of the reference's path.
#include "common.cuh"

namespace glia {

__device__ __forceinline__ uint32_t hash32(uint32_t x) {
  x ^= x >> 16; x *= 0x85EBCA6Bu; x ^= x >> 13; x *= 0xC2B2AE35u; x ^= x >> 16;
  return x;
}

__global__ void synth_voronoi_kernel(int nd, int X, int Y, int Z, int cell, uint32_t seed,
                                     int ncx, int ncy, int ncz, int zOff, int nzOut,
                                     uint32_t* __restrict__ labels, float* __restrict__ pb) {
  // writes planes [zOff, zOff + nzOut) of the X x Y x Z volume (a z-slab, or everything)
  size_t n = (size_t)X * Y * nzOut;
  const size_t linOff = (size_t)X * Y * zOff;
  for (size_t out = blockIdx.x * (size_t)blockDim.x + threadIdx.x; out < n;
       out += (size_t)gridDim.x * blockDim.x) {
    const size_t lin = out + linOff;
    int x = (int)(lin % X); size_t r = lin / X;
    int y = (int)(r % Y), z = (int)(r / Y);
    int c[3] = {x / cell, y / cell, z / cell};
    long long best1 = 0x7fffffffffffffffLL, best2 = best1;
    long long lab = 0;
    const int z0 = nd == 3 ? -1 : 0, z1 = nd == 3 ? 1 : 0;
    for (int oz = z0; oz <= z1; ++oz)
      for (int oy = -1; oy <= 1; ++oy)
        for (int ox = -1; ox <= 1; ++ox) {
          int cx = c[0] + ox, cy = c[1] + oy, cz = c[2] + oz;
          if (cx < 0 || cx >= ncx || cy < 0 || cy >= ncy || cz < 0 || cz >= ncz) continue;
          unsigned long long ci = (unsigned long long)cx + (unsigned long long)ncx * (cy + (unsigned long long)ncy * cz);
          int cc[3] = {cx, cy, cz}, xyz[3] = {x, y, z};
          long long d2 = 0;
          for (int d = 0; d < nd; ++d) {
            uint32_t u = hash32((uint32_t)((unsigned long long)seed * 0x9E3779B9ull + ci * 4ull + (unsigned)d)) >> 8;
            long long pos = (long long)cc[d] * cell * 256 + (long long)(((unsigned long long)u * (unsigned long long)(cell * 256)) >> 24);
            long long dd = (long long)xyz[d] * 256 - pos;
            d2 += dd * dd;
          }
          bool closer = d2 < best1;
          best2 = closer ? best1 : (d2 < best2 ? d2 : best2);
          lab = closer ? (long long)ci : lab;
          best1 = closer ? d2 : best1;
        }
    double d1 = sqrt((double)best1) / 256.0;
    double d2v = best2 < 0x7fffffffffffffffLL ? sqrt((double)best2) / 256.0 : d1 + 1e9;
    uint32_t uh = hash32((uint32_t)((unsigned long long)seed * 0x85EBCA77ull + 0x1234567ull + (unsigned long long)lin));
    double u = (double)uh / 4294967296.0;
    double val = 0.8 * exp(-(d2v - d1) / 2.0) + 0.2 * u;
    val = val < 0.0 ? 0.0 : (val > 1.0 ? 1.0 : val);
    labels[out] = (uint32_t)(lab + 1);
    pb[out] = (float)(floor(255.0 * val) / 256.0);
  }
}

}  // namespace glia

// dims: x (fastest), y, z.  labels/pb: DEVICE buffers.  Returns the number of grid cells
// (= exclusive label bound - 1) in *n_cells.
extern "C" GLIA_API int glia_b200_synth_voronoi(int ndim, const int* dims, int cell, uint32_t seed,
                                                uint32_t* labels, float* pb, void* stream,
                                                uint64_t* n_cells) {
  if ((ndim != 2 && ndim != 3) || !dims || cell <= 0 || !labels || !pb) return GLIA_E_ARG;
  int X = dims[0], Y = dims[1], Z = ndim == 3 ? dims[2] : 1;
  int ncx = (X + cell - 1) / cell, ncy = (Y + cell - 1) / cell, ncz = ndim == 3 ? (Z + cell - 1) / cell : 1;
  if (n_cells) *n_cells = (uint64_t)ncx * ncy * ncz;
  glia::synth_voronoi_kernel<<<148 * 16, 256, 0, (cudaStream_t)stream>>>(ndim, X, Y, Z, cell, seed, ncx, ncy,
                                                                         ncz, 0, Z, labels, pb);
  return cudaGetLastError() == cudaSuccess ? GLIA_OK : GLIA_E_CUDA;
}

// Planes [z0, z0 + nz) of the same 3D volume (what one rank of the z-slab path holds, halo included).
extern "C" GLIA_API int glia_b200_synth_voronoi_slab(const int* dims, int cell, uint32_t seed, int z0, int nz,
                                                     uint32_t* labels, float* pb, void* stream,
                                                     uint64_t* n_cells) {
  if (!dims || cell <= 0 || !labels || !pb || z0 < 0 || nz <= 0 || z0 + nz > dims[2]) return GLIA_E_ARG;
  int X = dims[0], Y = dims[1], Z = dims[2];
  int ncx = (X + cell - 1) / cell, ncy = (Y + cell - 1) / cell, ncz = (Z + cell - 1) / cell;
  if (n_cells) *n_cells = (uint64_t)ncx * ncy * ncz;
  glia::synth_voronoi_kernel<<<148 * 16, 256, 0, (cudaStream_t)stream>>>(3, X, Y, Z, cell, seed, ncx, ncy, ncz,
                                                                         z0, nz, labels, pb);
  return cudaGetLastError() == cudaSuccess ? GLIA_OK : GLIA_E_CUDA;
}
