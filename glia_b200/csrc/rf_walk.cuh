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

// v5 (round 2): TWO tree levels per load.  A 32-byte block holds an internal node and its two children:
//   {t0, t1, t2 : exact double thresholds of the node / left child / right child,
//    meta       : f0 | f1 << 8 | f2 << 16 (features of the three nodes),
//    base       : index of the first of the four child blocks | leaf mask << 24 | class bits << 28}
// The four outcomes o = 2 * [x[f0] > t0] + [x[f(child)] > t(child)] either end in a leaf (bit o of the leaf
// mask, its class = bit o of the class bits; a child that is itself a leaf sets both of its outcomes) or
// continue in block (base & 0xFFFFFF) + o.  One LDG.E.256 per lane fetches a block, so a walk issues half as
// many node loads and - with everything a step needs in one octet of registers - two thirds of the instructions of
// the one-node-per-level walk.  Two classes only (the boundary classifier's {-1, +1}), at most 2^24 blocks and 256
// features; other forests take rf_score_smem_kernel.
struct __align__(32) RfBlock {
  double t0, t1, t2;
  uint32_t meta, base;
};

// Walks one tree from block `k`; returns 1 when the leaf's class bit equals wantBit.  xs: shared-window address of
// the candidate's row.  alg/rf.hxx:35-36 -> classForest: x <= threshold goes left, anything else (NaN too) right.
__device__ __forceinline__ uint32_t rf_walk_blocks(const RfBlock* __restrict__ bp, uint32_t k, uint32_t xs, uint32_t wantBit) {
  for (;;) {
    uint32_t a0, a1, a2, a3, a4, a5, meta, base;
    asm("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r"(a0), "=r"(a1), "=r"(a2), "=r"(a3), "=r"(a4), "=r"(a5), "=r"(meta), "=r"(base)
        : "l"(bp + k));
    const double x0 = rf_lds(xs + 8u * (meta & 0xFFu));
    const bool r0 = !(x0 <= __hiloint2double((int)a1, (int)a0));
    const double tc = r0 ? __hiloint2double((int)a5, (int)a4) : __hiloint2double((int)a3, (int)a2);
    const uint32_t fc = __byte_perm(meta, 0u, r0 ? 0x4442u : 0x4441u);
    const double x1 = rf_lds(xs + 8u * fc);
    const uint32_t o = (r0 ? 2u : 0u) + (!(x1 <= tc) ? 1u : 0u);
    const uint32_t bits = base >> (24u + o);
    if (bits & 1u) return ((bits >> 4) & 1u) == wantBit ? 1u : 0u;
    k = (base & 0xFFFFFFu) + o;
  }
}

}  // namespace glia
