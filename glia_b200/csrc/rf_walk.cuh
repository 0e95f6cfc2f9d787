// RF walk over two-level blocks: device code shared by rf.cu (rows from HBM) and score_fused.cu (rows featurised
// straight into shared memory).
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
//    meta       : f0 | f1 << 8 | f2 << 16 (features) | leaf mask << 24 | class bits << 28,
//    base       : index of the first of the (up to four) blocks that follow}
// The four outcomes o = 2 * [x[f0] > t0] + [x[f(child)] > t(child)] either end in a leaf (bit o of the leaf
// mask, its class = bit o of the class bits; a child that is itself a leaf sets both of its outcomes) or
// continue in block base + (number of non-leaf outcomes below o).  One LDG.E.256 per lane fetches a block, so
// a walk issues half as many divergent loads as with one 16-byte node per level - the loads' L1 wavefronts
// (one per distinct line and instruction) are what bounds this kernel (profiles/r1_rf_v4_notes.txt).
// Two classes only (the boundary classifier's {-1, +1}); other forests take rf_score_smem_kernel.
struct __align__(32) RfBlock {
  double t0, t1, t2;
  uint32_t meta, base;
};

struct BlkRegs { double t0, t1, t2; uint32_t meta, base; };

__device__ __forceinline__ BlkRegs rf_block(const RfBlock* bp, uint32_t k) {
  BlkRegs r;
  uint32_t a0, a1, a2, a3, a4, a5;
  asm("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=r"(a0), "=r"(a1), "=r"(a2), "=r"(a3), "=r"(a4), "=r"(a5), "=r"(r.meta), "=r"(r.base)
      : "l"(bp + k));
  r.t0 = __hiloint2double((int)a1, (int)a0);
  r.t1 = __hiloint2double((int)a3, (int)a2);
  r.t2 = __hiloint2double((int)a5, (int)a4);
  return r;
}

// one block = two levels; done blocks (meta bit 31 after the step = reached a leaf) are skipped
#define RFB_DONE 0x80000000u
#define RFB_STEP(b, hitv)                                                                         \
  if (!(b.base & RFB_DONE)) {                                                                     \
    const double x0_ = rf_lds(xs + 8u * (b.meta & 0xFFu));                                        \
    const uint32_t c0_ = x0_ <= b.t0 ? 0u : 1u;                                                   \
    const double tc_ = c0_ ? b.t2 : b.t1;                                                         \
    const double x1_ = rf_lds(xs + 8u * ((b.meta >> (8u + 8u * c0_)) & 0xFFu));                   \
    const uint32_t o_ = 2u * c0_ + (x1_ <= tc_ ? 0u : 1u);                                        \
    const uint32_t leaf_ = (b.meta >> 24) & 0xFu;                                                 \
    if ((leaf_ >> o_) & 1u) { hitv += (((b.meta >> (28u + o_)) & 1u) == (uint32_t)labelIdx); b.base = RFB_DONE; } \
    else b = rf_block(bp, b.base + __popc(~leaf_ & ((1u << o_) - 1u) & 0xFu)); }


}  // namespace glia
