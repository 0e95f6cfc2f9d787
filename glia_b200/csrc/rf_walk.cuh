// RF walk over two-level blocks (device code of rf.cu).
#pragma once
#include "handle.cuh"

namespace glia {

#define RF_CPB 64  // candidates per CTA (their rows are staged in shared memory)

__device__ __forceinline__ double rf_lds(uint32_t a) {
  double v;
  asm("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
  return v;
}

// v5 (round 2): TWO tree levels per 16-byte load, integer compares.
//
// A block holds an internal node and its two children.  Word j (j = 0 node, 1 left child, 2 right child) is
//   K_j | f_j << 25      K_j = ord(floor_f32(threshold_j)) >> 7  (25 bits),  f_j = feature (7 bits)
// and word 3 is  base | leaf mask << 24 | class bits << 28  (base = index of the first of the four child blocks).
// ord() is the order-preserving map of float bit patterns to u32 (common.cuh); the CTA's rows are staged in shared
// memory as kx = ord(RN_f32(x)) >> 7, so a split is decided by ONE integer subtraction:
//   kx <  K      =>  RN(x) < floor_f32(t)  =>  x < t   (left)
//   kx >  K + 1  =>  RN(x) > ceil_f32(t)   =>  x > t   (right)
//   otherwise (x within 256 float ulps of t, ties included) the exact doubles decide: x from the row in global
//   memory, t from a side array - votes stay bit-exact (classForest: x <= t goes left, NaN goes right).
// The four outcomes o = 2 * right(node) + right(child) end in a leaf (bit o of the leaf mask, class = bit o of the
// class bits; a child that is itself a leaf sets both of its outcomes) or continue in block base + o.
// Why: the walk is bound by the SM's L1 data pipe - shared-memory wavefronts of the feature loads plus the
// write-back of the node loads (96 % of peak with double rows and one 16-byte node per level; profiles/r2_rf_notes.txt).
// This layout moves 8 bytes of node data and one conflict-free 4-byte feature per level, and needs no FP64.
// Two classes, at most 128 features and 2^20 blocks; other forests take rf_score_smem_kernel.
#define RFB_KBITS 25
#define RFB_KMASK 0x1FFFFFFu

__host__ __device__ inline uint32_t rf_xkey(double x) {  // kx of a feature value
  const float xf = (float)x;
  if (!(xf == xf)) return RFB_KMASK;  // NaN: right of every threshold
  return f2ord(xf) >> 7;
}

// The rare undecided compare, kept out of line: if-converted into the walk its ~25 predicated instructions would
// issue on every step (measured: 3.97 G instead of ~2 G warp instructions at c4s).
__device__ __noinline__ bool rf_exact_right(const double* __restrict__ xg, const double* __restrict__ thr, uint32_t f, uint32_t ti) {
  return !(__ldg(xg + f) <= __ldg(thr + ti));
}

// Walks one tree from block `k`; returns 1 when the leaf's class bit equals wantBit.  xs: shared-window address of
// the candidate's row of keys, xg: the same row as exact doubles in global memory, thr: exact thresholds [3 * block + j].
__device__ __forceinline__ uint32_t rf_walk_blocks(const uint4* __restrict__ bp, const double* __restrict__ thr, uint32_t k,
                                                   uint32_t xs, const double* __restrict__ xg, uint32_t wantBit) {
  for (;;) {
    const uint4 w = __ldg(bp + k);
    const uint32_t f0 = w.x >> RFB_KBITS;
    uint32_t kx;
    asm("ld.shared.u32 %0, [%1];" : "=r"(kx) : "r"(xs + 4u * f0));
    int d = (int)(kx - (w.x & RFB_KMASK));
    bool r0 = d > 1;
    if ((uint32_t)d <= 1u) r0 = rf_exact_right(xg, thr, f0, 3u * k);
    const uint32_t wc = r0 ? w.z : w.y;
    const uint32_t fc = wc >> RFB_KBITS;
    asm("ld.shared.u32 %0, [%1];" : "=r"(kx) : "r"(xs + 4u * fc));
    d = (int)(kx - (wc & RFB_KMASK));
    bool r1 = d > 1;
    if ((uint32_t)d <= 1u) r1 = rf_exact_right(xg, thr, fc, 3u * k + (r0 ? 2u : 1u));
    const uint32_t o = (r0 ? 2u : 0u) + (r1 ? 1u : 0u);
    if (w.w & (0x01000000u << o)) return ((w.w >> (28u + o)) & 1u) == wantBit ? 1u : 0u;  // leaf: one shift + one test
    k = (w.w & 0xFFFFFu) + o;
  }
}

}  // namespace glia
