// K6 + K7 fused: the rows of a CTA's 64 initial candidates are featurised STRAIGHT INTO the shared-memory tile the
// forest walk reads (rf_walk.cuh), so the row matrix never exists in HBM.  Round 1 wrote C x dim doubles with one
// kernel and read them back with the next (6.2 GB each way at 1024^3); SURVEY §8(d) counts those bytes
// as algorithmic only when a feature file is requested - glia_b200_init_candidates / merge_order_bc -b still
// produce them through featurize.cu.
//
// Replaces TBoundaryTable::init's fb + fsal calls for every initial candidate (type/boundary_table.hxx:91-114,
// hmt/main_merge_order_bc.cxx:54-139): fBcFeat row -> fBcPred = votes(class -1) / ntree (alg/rf.hxx:35-36).
#include "featurize_init.cuh"
#include "rf_walk.cuh"

namespace glia {

template <bool ENS, int NS>
__global__ void __launch_bounds__(64 * NS, NS == 8 ? 2 : 4)
score_fused_kernel(const __grid_constant__ FeatCfg F, const __grid_constant__ GraphView G, int xi,
                   const double* __restrict__ blk, const unsigned char* __restrict__ orient,
                   const RfBlock* __restrict__ blocks, const int* __restrict__ treeRoot, const RfDist dist,
                   int labelIdx, int ldRow, double* __restrict__ prob) {
  extern __shared__ __align__(16) double srow[];
  __shared__ int svote[RF_CPB];
  __shared__ int sntree[RF_CPB];
  const long long c0 = (long long)blockIdx.x * RF_CPB;
  const int ncand = (int)min((long long)RF_CPB, G.C - c0);
  // ---- featurise: two threads per candidate (x0 + x1 block | x2 block + merged region)
  if (threadIdx.x < 2 * RF_CPB) {
    const int c = threadIdx.x >> 1, part = threadIdx.x & 1;
    if (c < ncand) candidate_row_from_blocks(F, G, xi, blk, orient, c0 + c, srow + (size_t)c * ldRow, 1 << part);
  }
  if (threadIdx.x < RF_CPB) svote[threadIdx.x] = 0;
  __syncthreads();
  // ---- walk (same mapping as rf_score_blk_kernel: a warp = 32 candidates on the same trees)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cand = (warp & 1) * 32 + lane, slice = warp >> 1;
  if (cand < ncand) {
    const double* x = srow + cand * ldRow;
    const int mdl = ENS ? dist.pick(x) : 0;
    const int tbeg = ENS ? dist.begin(mdl) : 0, ntree = ENS ? dist.end(mdl) : dist.te[0];
    if (slice == 0) sntree[cand] = ntree - tbeg;
    const uint32_t xs = (uint32_t)__cvta_generic_to_shared(x);
    const RfBlock* bp = blocks;
    int hits = 0, t = tbeg + slice;
    for (; t + NS < ntree; t += 2 * NS) {
      BlkRegs a = rf_block(bp, (uint32_t)__ldg(treeRoot + t)), b = rf_block(bp, (uint32_t)__ldg(treeRoot + t + NS));
      while (!(a.base & b.base & RFB_DONE)) { RFB_STEP(a, hits) RFB_STEP(b, hits) }
    }
    for (; t < ntree; t += NS) {
      BlkRegs a = rf_block(bp, (uint32_t)__ldg(treeRoot + t));
      while (!(a.base & RFB_DONE)) { RFB_STEP(a, hits) }
    }
    atomicAdd(&svote[cand], hits);
  }
  __syncthreads();
  if (threadIdx.x < ncand) prob[c0 + threadIdx.x] = (double)svote[threadIdx.x] / (double)sntree[threadIdx.x];
}

}  // namespace glia

using namespace glia;
GraphView glia_graph_view(const glia_handle* h);
int glia_region_blocks_device(glia_handle* h, int* xiOut);

// Featurise + score the initial candidates without materialising the rows.  Returns GLIA_E_STATE when this forest /
// configuration needs the unfused path (more than two classes, --simpf rows, rows too wide for shared memory).
int glia_score_fused_device(glia_handle* h, const glia_rf* rf, int labelIdx, double* dSal) {
  if (h->C == 0) return GLIA_OK;
  if (!rf->d_blocks || h->fc.simple) return GLIA_E_STATE;
  const int dim = h->fc.dim, ldRow = dim | 1;
  const size_t smem = sizeof(double) * (size_t)RF_CPB * ldRow;
  if (smem > 200 * 1024) return GLIA_E_STATE;
  int xi = 0;
  int s = glia_region_blocks_device(h, &xi);
  if (s != GLIA_OK) return s;
  GraphView G = glia_graph_view(h);
  const unsigned char* orient = (h->hasOrient && h->orientC == h->C) ? h->orient.as<unsigned char>() : nullptr;
  const unsigned nb = (unsigned)((h->C + RF_CPB - 1) / RF_CPB);
  const bool wide = smem > 28 * 1024;
  const RfBlock* bl = (const RfBlock*)rf->d_blocks;
#define FUSED_LAUNCH(ENS_, NS_)                                                                                  \
  { cudaFuncSetAttribute(score_fused_kernel<ENS_, NS_>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024); \
    score_fused_kernel<ENS_, NS_><<<nb, 64 * NS_, smem, h->stream>>>(h->fc, G, xi, h->regBlocks.as<double>(), orient, bl, \
                                                                     rf->d_blockRoot, rf->dist, labelIdx, ldRow, dSal); }
  if (rf->dist.n > 1) { if (wide) FUSED_LAUNCH(true, 8) else FUSED_LAUNCH(true, 4) }
  else { if (wide) FUSED_LAUNCH(false, 8) else FUSED_LAUNCH(false, 4) }
#undef FUSED_LAUNCH
  h->launches++;
  GLIA_CUDA_CHECK(h, cudaGetLastError());
  return GLIA_OK;
}
