// Allocation-free RF scoring: glia_b200_rf_score_noalloc.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is a variant of the vote-counting
// scorer (rf_score_kernel<true>, rf.cu; reference semantics: Model::predict -> classForest, ml/rf/rf.hxx:362-405): the
// per-row vote table [N][nclass] - the scratch of that kernel - lives in a file-scope __device__ array of fixed size
// instead of a cudaMalloc'ed buffer, the trees of a row are spread over the 32 lanes of a warp, each lane adds its
// leaves' classes to the row's counters, and a second kernel turns the counters into prob = votes[label] / ntree.
//
// Limits and rules (also in include/glia_b200.h):
//   * N * nclass <= GLIA_NOALLOC_MAX_VOTES (1 Mi counters = 4 MiB, held by every GPU that loads the library until exit);
//   * the entry point calls no cudaMalloc / cudaFree / allocating library routine;
//   * every call on a device shares the one array: it is cleared at the start of each call (stream-ordered memset), and
//     two calls that overlap on the GPU (different streams, no ordering between them) race on it.
#include <mutex>
#include <stdio.h>
#include "handle.cuh"

int glia_rf_label_index(const glia_rf* m, int label);

#define GLIA_NOALLOC_MAX_VOTES_ (1u << 20)
#define GLIA_MAX_DEVICES_ 64

// The scratch itself.  In host code this name is only a stand-in for the symbol: the device address is obtained once
// per device by cudaGetSymbolAddress (scratch_of below) and that pointer is what the kernels are given.
static __device__ __align__(256) int g_vote_scratch[GLIA_NOALLOC_MAX_VOTES_];

namespace glia {

struct __align__(16) RfNodeNA { double thr; int left; int feat; };  // == RfNode (rf.cu)

// one warp per row; lane l walks trees tb + l, tb + l + 32, ... and counts the class of every leaf it reaches
__global__ void __launch_bounds__(256)
rf_votes_scratch_kernel(const RfNodeNA* __restrict__ nodes, const int* __restrict__ treeOff, const RfDist dist, int nclass,
                        const double* __restrict__ X, size_t N, int D, int* __restrict__ votes) {
  const size_t row = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= N) return;
  const double* x = X + row * (size_t)D;
  const int mdl = dist.pick(x), tb = dist.begin(mdl), te = dist.end(mdl);  // alg/rf.hxx:97-98
  for (int t = tb + lane; t < te; t += 32) {
    int k = __ldg(treeOff + t);
    RfNodeNA nd = nodes[k];
    while (nd.feat >= 0) {  // x <= threshold goes left, anything else (NaN too) goes right
      const double v = __ldg(x + nd.feat);
      k = nd.left + (v <= nd.thr ? 0 : 1);
      nd = nodes[k];
    }
    atomicAdd(votes + row * (size_t)nclass + nd.left, 1);
  }
}

__global__ void rf_prob_from_votes_kernel(const int* __restrict__ votes, const RfDist dist, int nclass, int labelIdx,
                                          const double* __restrict__ X, size_t N, int D, double* __restrict__ prob) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= N) return;
  const int mdl = dist.pick(X + i * (size_t)D), ntree = dist.end(mdl) - dist.begin(mdl);
  prob[i] = (double)votes[i * (size_t)nclass + labelIdx] / (double)ntree;  // ml/rf/rf.hxx:366-369
}

}  // namespace glia

namespace {

// device address of g_vote_scratch on `device` (which must be current); looked up once per device
int scratch_of(int device, int** out, cudaError_t* errOut) {
  static std::mutex mu;
  static int* addr[GLIA_MAX_DEVICES_] = {nullptr};
  if (device < 0 || device >= GLIA_MAX_DEVICES_) return GLIA_E_ARG;
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[device]) {
    int* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_vote_scratch);
    if (e != cudaSuccess || !p) { if (errOut) *errOut = e; return GLIA_E_CUDA; }
    addr[device] = p;
  }
  *out = addr[device];
  return GLIA_OK;
}

void say(char* err, size_t errLen, const char* what, const char* detail = "") {
  if (err && errLen) snprintf(err, errLen, "%s%s", what, detail);
}

}  // namespace

extern "C" size_t glia_b200_rf_score_noalloc_max_votes(void) { return (size_t)GLIA_NOALLOC_MAX_VOTES_; }

extern "C" int glia_b200_rf_score_noalloc(const glia_rf* m, const double* d_rows, size_t N, int D, int label,
                                          double* d_prob, void* stream, char* err, size_t err_len) {
  if (err && err_len) err[0] = 0;
  if (!m || !d_rows || !d_prob || D < m->mdim) { say(err, err_len, "bad argument (null pointer, or rows narrower than the forest's features)"); return GLIA_E_ARG; }
  const int li = glia_rf_label_index(m, label);
  if (li < 0) { say(err, err_len, "the forest has no such class label"); return GLIA_E_ARG; }  // ml/rf/rf.hxx:370 perr
  if (N == 0) return GLIA_OK;
  const unsigned long long need = (unsigned long long)N * (unsigned long long)m->nclass;
  if (need > GLIA_NOALLOC_MAX_VOTES_) {
    if (err && err_len)
      snprintf(err, err_len, "glia_b200_rf_score_noalloc: N * nclass = %llu vote counters exceed the static scratch of %llu; "
                             "score in smaller batches or use glia_b200_rf_score", need, (unsigned long long)GLIA_NOALLOC_MAX_VOTES_);
    return GLIA_E_CAPACITY;
  }
  if (cudaSetDevice(m->device) != cudaSuccess) { say(err, err_len, "cudaSetDevice failed"); return GLIA_E_CUDA; }
  int* votes = nullptr;
  cudaError_t ce = cudaSuccess;
  int s = scratch_of(m->device, &votes, &ce);
  if (s != GLIA_OK) { say(err, err_len, "cudaGetSymbolAddress(vote scratch) failed: ", cudaGetErrorString(ce)); return s; }
  cudaStream_t st = (cudaStream_t)stream;
  // the array keeps the previous call's counters: the ones this call accumulates into must start at zero
  if ((ce = cudaMemsetAsync(votes, 0, sizeof(int) * (size_t)need, st)) != cudaSuccess) {
    say(err, err_len, "cudaMemsetAsync(vote scratch) failed: ", cudaGetErrorString(ce));
    return GLIA_E_CUDA;
  }
  const unsigned b1 = (unsigned)((N * 32 + 255) / 256), b2 = (unsigned)((N + 255) / 256);
  glia::rf_votes_scratch_kernel<<<b1, 256, 0, st>>>((const glia::RfNodeNA*)m->d_nodes, m->d_treeOff, m->dist, m->nclass, d_rows, N, D, votes);
  glia::rf_prob_from_votes_kernel<<<b2, 256, 0, st>>>(votes, m->dist, m->nclass, li, d_rows, N, D, d_prob);
  if ((ce = cudaGetLastError()) != cudaSuccess) { say(err, err_len, "kernel launch failed: ", cudaGetErrorString(ce)); return GLIA_E_CUDA; }
  return GLIA_OK;  // stream-ordered: d_prob is complete when the work queued on `stream` so far has finished
}
