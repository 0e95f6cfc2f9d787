// Multi-GPU (z-slab) support: the pieces of glia_b200_build split so that the partial tables of
// the voxel pass can be combined across ranks.
//
// The reference has no distributed path (one process, util/mp.hxx OpenMP only); what is kept is
// its result: every voxel is owned by exactly one slab, so the integer statistics add exactly.
//   1. glia_b200_build_partial      voxel pass over the owned planes (halo planes supply labels)
//   2. all-reduce of the dense per-label tables (sum / max / sum) by the caller
//   3. glia_b200_finalize_regions   label -> dense region index (identical on every rank)
//   4. glia_b200_edge_interest      which ranks need the boundaries of region x: the owner of x
//                                   and the owners of every neighbour of x   (all-reduce max)
//   5. glia_b200_export_edges       packed boundary records bucketed by destination rank
//   6. all-to-all by the caller, glia_b200_import_edges merges equal label pairs (the "reduce
//      keyed by label pair")
//   7. glia_b200_finalize_edges     CSR + candidates whose smaller end lies in the owned range
// A region index r is owned by rank min(world-1, r / ceil(R/world)).
#include <algorithm>
#include "handle.cuh"

int glia_finalize_regions(glia_handle* h);
int glia_region_presence(glia_handle* h);
int glia_index_regions(glia_handle* h);
int glia_gather_regions(glia_handle* h);
int glia_gather_regions_some(glia_handle* h, const unsigned char* interest, int world, int me);
int glia_finalize_edges(glia_handle* h, long long ownLo, long long ownHi);
int glia_build_raw(glia_handle* h, const uint32_t* labels, const uint32_t* mask, const float* pb);

namespace glia {

__device__ __forceinline__ int owner_of(uint32_t r, uint32_t chunk, int world) {
  int o = (int)(r / chunk);
  return o < world ? o : world - 1;
}

__global__ void edge_interest_kernel(const unsigned long long* __restrict__ keys, uint64_t cap,
                                     const uint32_t* __restrict__ rank, uint32_t chunk, int world,
                                     unsigned char* __restrict__ interest) {
  uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (i >= cap) return;
  unsigned long long k = keys[i];
  if (k == EDGE_EMPTY) return;
  uint32_t ra = rank[(uint32_t)(k >> 32)], rb = rank[(uint32_t)k];
  int oa = owner_of(ra, chunk, world), ob = owner_of(rb, chunk, world);
  interest[(size_t)ra * world + oa] = 1;
  interest[(size_t)ra * world + ob] = 1;
  interest[(size_t)rb * world + oa] = 1;
}

// pass 0: count records per destination; pass 1: write them (cursor = running position)
template <int PASS>
__global__ void export_edges_kernel(const unsigned long long* __restrict__ keys, uint64_t cap,
                                    const uint32_t* __restrict__ rank,
                                    const unsigned char* __restrict__ interest, int world,
                                    RecLayout L, const uint32_t* __restrict__ gAdd,
                                    const uint32_t* __restrict__ gMax, const double* __restrict__ gF64,
                                    unsigned long long* __restrict__ cursor, int recWords,
                                    uint32_t* __restrict__ out, int skipRank) {
  uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  unsigned long long k = (i < cap) ? keys[i] : EDGE_EMPTY;
  if (PASS == 0) {  // cursor[64] = occupied slots of the table (one atomic per block)
    int occ = __syncthreads_count(k != EDGE_EMPTY);
    if (threadIdx.x == 0 && occ) atomicAdd(cursor + 64, (unsigned long long)occ);
  }
  if (k == EDGE_EMPTY) return;
  uint32_t ra = rank[(uint32_t)(k >> 32)];
  const int padW = (L.nAdd + L.nMax) & 1;
  for (int d = 0; d < world; ++d) {
    if (d == skipRank || !interest[(size_t)ra * world + d]) continue;
    unsigned long long pos = atomicAdd(cursor + d, 1ull);
    if (PASS == 0) continue;
    uint32_t* o = out + (size_t)pos * recWords;
    o[0] = (uint32_t)k; o[1] = (uint32_t)(k >> 32);
    for (int w = 0; w < L.nAdd; ++w) o[2 + w] = gAdd[(size_t)i * L.nAdd + w];
    for (int w = 0; w < L.nMax; ++w) o[2 + L.nAdd + w] = gMax[(size_t)i * L.nMax + w];
    if (padW) o[2 + L.nAdd + L.nMax] = 0;
    double* f = (double*)(o + 2 + L.nAdd + L.nMax + padW);
    for (int w = 0; w < L.nF64; ++w) f[w] = gF64[(size_t)i * L.nF64 + w];
  }
}

__global__ void import_edges_kernel(const uint32_t* __restrict__ in, long long n, int recWords,
                                    RecLayout L, unsigned long long* __restrict__ keys, uint64_t cap,
                                    uint32_t* __restrict__ gAdd, uint32_t* __restrict__ gMax,
                                    double* __restrict__ gF64, int* __restrict__ overflow) {
  long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (r >= n) return;
  const uint32_t* p = in + (size_t)r * recWords;
  const unsigned long long key = ((unsigned long long)p[1] << 32) | p[0];
  const uint64_t mask = cap - 1;
  uint64_t hsh = mix64(key) & mask;
  long long slot = -1;
  for (uint64_t probe = 0; probe < cap; ++probe) {
    unsigned long long cur = keys[hsh];
    if (cur == EDGE_EMPTY) cur = atomicCAS(keys + hsh, EDGE_EMPTY, key);
    if (cur == key || cur == EDGE_EMPTY) { slot = (long long)hsh; break; }
    hsh = (hsh + 1) & mask;
  }
  if (slot < 0) { atomicExch(overflow, 1); return; }
  const int padW = (L.nAdd + L.nMax) & 1;
  for (int w = 0; w < L.nAdd; ++w) { uint32_t v = p[2 + w]; if (v) atomicAdd(gAdd + (size_t)slot * L.nAdd + w, v); }
  for (int w = 0; w < L.nMax; ++w) { uint32_t v = p[2 + L.nAdd + w]; if (v) atomicMax(gMax + (size_t)slot * L.nMax + w, v); }
  const double* f = (const double*)(p + 2 + L.nAdd + L.nMax + padW);
  for (int w = 0; w < L.nF64; ++w) atomicAdd(gF64 + (size_t)slot * L.nF64 + w, f[w]);
}

}  // namespace glia

using namespace glia;
#define CK(expr) GLIA_CUDA_CHECK(h, expr)
static inline unsigned nblk(unsigned long long n, int per) { return (unsigned)((n + per - 1) / per); }
static int rec_words(const RecLayout& L) { return 2 + L.nAdd + L.nMax + ((L.nAdd + L.nMax) & 1) + 2 * L.nF64; }

extern "C" int glia_b200_build_partial(glia_handle* h, const uint32_t* labels, const uint32_t* mask, const float* pb) {
  if (!h || !labels || !pb) return GLIA_E_ARG;
  if (h->prm.label_bound == 0) { h->err = "build_partial needs label_bound (identical on every rank)"; return GLIA_E_ARG; }
  int s = glia_build_raw(h, labels, mask, pb);
  if (s != GLIA_OK) return s;
  CK(cudaStreamSynchronize(h->stream));
  float ms = 0;
  cudaEventElapsedTime(&ms, h->ev[6], h->ev[1]); h->ms[0] = ms;
  cudaEventElapsedTime(&ms, h->ev[0], h->ev[6]); h->ms[4] = ms;
  h->partial = true;
  return GLIA_OK;
}

extern "C" int glia_b200_raw_region_tables(glia_handle* h, void** add, void** mx, void** f64, int64_t* nAdd,
                                           int64_t* nMax, int64_t* nF64) {
  if (!h) return GLIA_E_ARG;
  if (!h->partial) return GLIA_E_STATE;
  const VoxParams& P = h->vp;
  if (add) *add = h->regAdd.p;
  if (mx) *mx = h->regMax.p;
  if (f64) *f64 = h->regF64.p;
  if (nAdd) *nAdd = (int64_t)P.labelBound * P.reg.nAdd;
  if (nMax) *nMax = (int64_t)P.labelBound * P.reg.nMax;
  if (nF64) *nF64 = (int64_t)P.labelBound * P.reg.nF64;
  return GLIA_OK;
}

extern "C" int glia_b200_finalize_regions(glia_handle* h, int64_t* R) {
  if (!h) return GLIA_E_ARG;
  if (!h->partial) return GLIA_E_STATE;
  CK(cudaSetDevice(h->prm.device));
  int s = glia_finalize_regions(h);
  if (s != GLIA_OK) return s;
  if (R) *R = h->R;
  return GLIA_OK;
}

extern "C" int glia_b200_region_presence(glia_handle* h, void** flags, int64_t* n) {
  if (!h || !flags || !n) return GLIA_E_ARG;
  if (!h->partial) return GLIA_E_STATE;
  CK(cudaSetDevice(h->prm.device));
  int s = glia_region_presence(h);
  if (s != GLIA_OK) return s;
  *flags = h->flags.p; *n = (int64_t)h->vp.labelBound;
  return GLIA_OK;
}

extern "C" int glia_b200_index_regions(glia_handle* h, int64_t* R) {
  if (!h) return GLIA_E_ARG;
  if (!h->partial || !h->flags.p) return GLIA_E_STATE;
  CK(cudaSetDevice(h->prm.device));
  int s = glia_index_regions(h);
  if (s != GLIA_OK) return s;
  if (R) *R = h->R;
  return GLIA_OK;
}

extern "C" int glia_b200_gather_regions(glia_handle* h) {
  if (!h) return GLIA_E_ARG;
  if (!h->partial || !h->rank.p) return GLIA_E_STATE;
  CK(cudaSetDevice(h->prm.device));
  return glia_gather_regions(h);
}

extern "C" int glia_b200_edge_interest(glia_handle* h, int world, void** mask, int64_t* bytes) {
  if (!h || world < 1 || world > 64) return GLIA_E_ARG;
  if (!h->partial) return GLIA_E_STATE;
  CK(cudaSetDevice(h->prm.device));
  const VoxParams& P = h->vp;
  const size_t n = (size_t)std::max<long long>(h->R, 1) * world;
  CK(h->interest.ensure(n));
  CK(cudaMemsetAsync(h->interest.p, 0, n, h->stream));
  const uint32_t chunk = (uint32_t)std::max<long long>(1, (h->R + world - 1) / world);
  edge_interest_kernel<<<nblk(P.edgeCap, 256), 256, 0, h->stream>>>(
      (const unsigned long long*)h->edgeKeys.p, P.edgeCap, h->rank.as<uint32_t>(), chunk, world,
      h->interest.as<unsigned char>());
  h->launches += 2;
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(h->stream));
  if (mask) *mask = h->interest.p;
  if (bytes) *bytes = (int64_t)h->R * world;
  return GLIA_OK;
}

extern "C" int glia_b200_export_edges(glia_handle* h, int world, int self_rank, int64_t* counts, void** buf,
                                      int* rec_bytes) {
  if (!h || world < 1 || world > 64 || !counts) return GLIA_E_ARG;
  if (!h->partial || !h->interest.p) return GLIA_E_STATE;
  CK(cudaSetDevice(h->prm.device));
  const VoxParams& P = h->vp;
  cudaStream_t st = h->stream;
  const int rw = rec_words(P.edge);
  CK(h->exportCursor.ensure(8 * 65));
  unsigned long long* cursor = h->exportCursor.as<unsigned long long>();
  CK(cudaMemsetAsync(cursor, 0, 8 * 65, st));
  export_edges_kernel<0><<<nblk(P.edgeCap, 256), 256, 0, st>>>(
      (const unsigned long long*)h->edgeKeys.p, P.edgeCap, h->rank.as<uint32_t>(), h->interest.as<unsigned char>(),
      world, P.edge, nullptr, nullptr, nullptr, cursor, rw, nullptr, self_rank);
  unsigned long long cnt[65], off[64], total = 0;
  CK(cudaMemcpyAsync(cnt, cursor, 8 * 65, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  h->tableOccupied = (long long)cnt[64];
  for (int d = 0; d < world; ++d) { off[d] = total; total += cnt[d]; counts[d] = (int64_t)cnt[d]; }
  CK(h->exportBuf.ensure(4 * (size_t)rw * (size_t)std::max<unsigned long long>(total, 1)));
  CK(cudaMemcpyAsync(cursor, off, 8 * world, cudaMemcpyHostToDevice, st));
  export_edges_kernel<1><<<nblk(P.edgeCap, 256), 256, 0, st>>>(
      (const unsigned long long*)h->edgeKeys.p, P.edgeCap, h->rank.as<uint32_t>(), h->interest.as<unsigned char>(),
      world, P.edge, h->edgeAdd.as<uint32_t>(), h->edgeMax.as<uint32_t>(), h->edgeF64.as<double>(), cursor, rw,
      h->exportBuf.as<uint32_t>(), self_rank);
  h->launches += 3;
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(st));
  if (buf) *buf = h->exportBuf.p;
  if (rec_bytes) *rec_bytes = 4 * rw;
  return GLIA_OK;
}

extern "C" int glia_b200_table_load(const glia_handle* h, int64_t* occupied, int64_t* capacity) {
  if (!h) return GLIA_E_ARG;
  if (!h->partial) return GLIA_E_STATE;
  if (occupied) *occupied = h->tableOccupied;
  if (capacity) *capacity = (int64_t)h->vp.edgeCap;
  return GLIA_OK;
}

extern "C" int glia_b200_import_edges(glia_handle* h, const void* buf, int64_t n, int clear_first) {
  if (!h || n < 0 || (n > 0 && !buf)) return GLIA_E_ARG;
  if (!h->partial) return GLIA_E_STATE;
  CK(cudaSetDevice(h->prm.device));
  cudaStream_t st = h->stream;
  VoxParams& PW = h->vp;
  for (;;) {
    const size_t cap = (size_t)PW.edgeCap;
    if (clear_first) {
      CK(cudaMemsetAsync(h->edgeKeys.p, 0xFF, 8 * cap, st));
      CK(cudaMemsetAsync(h->edgeAdd.p, 0, 4 * cap * PW.edge.nAdd, st));
      CK(cudaMemsetAsync(h->edgeMax.p, 0, 4 * cap * PW.edge.nMax, st));
      CK(cudaMemsetAsync(h->edgeF64.p, 0, 8 * cap * PW.edge.nF64, st));
      CK(cudaMemsetAsync(PW.overflow, 0, 4, st));
      h->launches += 5;
    }
    if (n > 0) {
      import_edges_kernel<<<nblk((unsigned long long)n, 256), 256, 0, st>>>(
          (const uint32_t*)buf, n, rec_words(PW.edge), PW.edge, (unsigned long long*)h->edgeKeys.p, PW.edgeCap,
          h->edgeAdd.as<uint32_t>(), h->edgeMax.as<uint32_t>(), h->edgeF64.as<double>(), PW.overflow);
      h->launches++;
    }
    int ovf = 0;
    CK(cudaMemcpyAsync(&ovf, PW.overflow, 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (!ovf) break;
    // owned + neighbour boundaries do not fit (labels far from spatial order): the table is rebuilt
    // from the import buffer alone, so it can simply be enlarged and the import repeated
    if (!clear_first || h->prm.edge_capacity != 0) {
      h->err = "directed-boundary table full during import: raise edge_capacity"; return GLIA_E_CAPACITY;
    }
    const uint64_t ncap = PW.edgeCap * 4;
    CK(h->edgeKeys.ensure(8 * (size_t)ncap));
    CK(h->edgeAdd.ensure(4 * (size_t)ncap * PW.edge.nAdd));
    CK(h->edgeMax.ensure(4 * (size_t)ncap * std::max(1, PW.edge.nMax)));
    CK(h->edgeF64.ensure(8 * (size_t)ncap * std::max(1, PW.edge.nF64)));
    PW.edgeCap = ncap;
    h->capHint = ncap;
    PW.edgeKeys = (unsigned long long*)h->edgeKeys.p;
    PW.edgeAdd = h->edgeAdd.as<uint32_t>(); PW.edgeMax = h->edgeMax.as<uint32_t>(); PW.edgeF64 = h->edgeF64.as<double>();
  }
  return GLIA_OK;
}

extern "C" int glia_b200_finalize_edges(glia_handle* h, int64_t own_lo, int64_t own_hi) {
  if (!h || own_lo < 0 || own_hi < own_lo) return GLIA_E_ARG;
  if (!h->partial) return GLIA_E_STATE;
  CK(cudaSetDevice(h->prm.device));
  CK(cudaEventRecord(h->ev[1], h->stream));
  int s = glia_finalize_edges(h, own_lo, std::min<long long>(own_hi, h->R));
  if (s != GLIA_OK) return s;
  CK(cudaEventRecord(h->ev[2], h->stream));
  CK(cudaStreamSynchronize(h->stream));
  float ms = 0;
  cudaEventElapsedTime(&ms, h->ev[1], h->ev[2]); h->ms[1] = ms;
  h->built = true;
  h->err.clear();
  return GLIA_OK;
}

// ================================================================================================
// The exchange as ONE stream-ordered call over NCCL (C++; no Python between the steps).
//
// glia_b200/parallel.py drives the entry points above with torch.distributed collectives: every step there is a
// Python call and most of them end in a host synchronisation, which at 8 GPUs costs about as much as the voxel pass
// of a 128-plane slab (profiles/r2_bench_c4_n8.json).  glia_b200_build_sharded issues the same sequence from C++:
// NCCL calls go on the handle's stream, the all-reduce of the record tables runs on a second communicator and a side
// stream, and the host only waits where it needs a number to size a buffer (R, the send / receive counts, E, C).
// NCCL is bound at run time (dlopen "libnccl.so.2": the copy the process already holds - torch's in bench.py, the
// system's in a C++ host), so libglia_b200.so has no link-time dependency on it.
#include <dlfcn.h>
#include <nccl.h>
#include <string.h>
#include <vector>

namespace {
struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId*);
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
  ncclResult_t (*CommDestroy)(ncclComm_t);
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t);
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*GroupStart)();
  ncclResult_t (*GroupEnd)();
  const char* (*GetErrorString)(ncclResult_t);
  bool ok;
};
NcclApi* nccl_api() {
  static NcclApi a;
  static bool tried = false;
  if (!tried) {
    tried = true;
    memset(&a, 0, sizeof(a));
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (lib) {
#define GLIA_SYM(field, name) *(void**)(&a.field) = dlsym(lib, name)
      GLIA_SYM(GetUniqueId, "ncclGetUniqueId"); GLIA_SYM(CommInitRank, "ncclCommInitRank");
      GLIA_SYM(CommDestroy, "ncclCommDestroy"); GLIA_SYM(AllReduce, "ncclAllReduce");
      GLIA_SYM(AllGather, "ncclAllGather"); GLIA_SYM(Send, "ncclSend"); GLIA_SYM(Recv, "ncclRecv");
      GLIA_SYM(GroupStart, "ncclGroupStart"); GLIA_SYM(GroupEnd, "ncclGroupEnd");
      GLIA_SYM(GetErrorString, "ncclGetErrorString");
#undef GLIA_SYM
      a.ok = a.GetUniqueId && a.CommInitRank && a.CommDestroy && a.AllReduce && a.AllGather && a.Send && a.Recv &&
             a.GroupStart && a.GroupEnd && a.GetErrorString;
    }
  }
  return a.ok ? &a : nullptr;
}
}  // namespace

struct glia_comm {
  ncclComm_t main = nullptr, side = nullptr;  // boundary exchange | all-reduce of the region record tables
  int world = 1, rank = 0, device = 0;
  cudaStream_t sideStream = nullptr;
  cudaEvent_t evTables = nullptr, evSide = nullptr, evA = nullptr, evB = nullptr, evC = nullptr;
  glia::DevBuf recv, counts;
  double exchangeMs = -1.0;
};

#define NK(h, expr)                                                                                   \
  do {                                                                                                \
    ncclResult_t r_ = (expr);                                                                         \
    if (r_ != ncclSuccess) { (h)->err = std::string(#expr) + ": " + N->GetErrorString(r_); return GLIA_E_CUDA; } \
  } while (0)

extern "C" int glia_b200_comm_unique_id(void* id256) {
  NcclApi* N = nccl_api();
  if (!N || !id256) return GLIA_E_ARG;
  ncclUniqueId ids[2];
  if (N->GetUniqueId(&ids[0]) != ncclSuccess || N->GetUniqueId(&ids[1]) != ncclSuccess) return GLIA_E_CUDA;
  memcpy(id256, ids, sizeof(ids));
  return GLIA_OK;
}

extern "C" int glia_b200_comm_create(const void* id256, int world, int rank, int device, glia_comm** out) {
  NcclApi* N = nccl_api();
  if (!N || !id256 || !out || world < 1 || world > 64 || rank < 0 || rank >= world) return GLIA_E_ARG;
  if (cudaSetDevice(device) != cudaSuccess) return GLIA_E_CUDA;
  ncclUniqueId ids[2];
  memcpy(ids, id256, sizeof(ids));
  glia_comm* c = new glia_comm();
  c->world = world; c->rank = rank; c->device = device;
  if (N->CommInitRank(&c->main, world, ids[0], rank) != ncclSuccess || N->CommInitRank(&c->side, world, ids[1], rank) != ncclSuccess ||
      cudaStreamCreateWithFlags(&c->sideStream, cudaStreamNonBlocking) != cudaSuccess) {
    delete c; return GLIA_E_CUDA;
  }
  cudaEventCreateWithFlags(&c->evTables, cudaEventDisableTiming); cudaEventCreateWithFlags(&c->evSide, cudaEventDisableTiming);
  cudaEventCreate(&c->evA); cudaEventCreate(&c->evB); cudaEventCreate(&c->evC);
  *out = c;
  return GLIA_OK;
}

extern "C" void glia_b200_comm_destroy(glia_comm* c) {
  if (!c) return;
  NcclApi* N = nccl_api();
  cudaSetDevice(c->device);
  if (N) { if (c->main) N->CommDestroy(c->main); if (c->side) N->CommDestroy(c->side); }
  if (c->sideStream) cudaStreamDestroy(c->sideStream);
  cudaEvent_t evs[5] = {c->evTables, c->evSide, c->evA, c->evB, c->evC};
  for (auto e : evs) if (e) cudaEventDestroy(e);
  c->recv.release(); c->counts.release();
  delete c;
}

extern "C" double glia_b200_comm_last_exchange_ms(const glia_comm* c) { return c ? c->exchangeMs : -1.0; }

// labels / mask / pb hold planes [z_begin - halo_lo, z_end + halo_hi) of the volume, as for glia_b200_build_partial.
// On return the handle is finalised and answers init_candidates / score_candidates for the candidates this rank owns.
extern "C" int glia_b200_build_sharded(glia_handle* h, glia_comm* c, const uint32_t* labels, const uint32_t* mask,
                                       const float* pb) {
  NcclApi* N = nccl_api();
  if (!h || !c || !N) return GLIA_E_ARG;
  if (c->device != h->prm.device) { h->err = "communicator lives on another device"; return GLIA_E_ARG; }
  const int world = c->world, me = c->rank;
  int s = glia_b200_build_partial(h, labels, mask, pb);  // voxel pass; ends with the overflow check
  if (s != GLIA_OK) return s;
  cudaStream_t st = h->stream;
  const VoxParams& P = h->vp;
  const uint32_t LB = P.labelBound;
  CK(cudaEventRecord(c->evA, st));
  // presence flags from the LOCAL table first: the all-reduce below rewrites the table in place
  s = glia_region_presence(h);
  if (s != GLIA_OK) return s;
  // record tables: all-reduce on the side communicator / side stream, needed only by gather_regions at the very end
  CK(cudaEventRecord(c->evTables, st));
  CK(cudaStreamWaitEvent(c->sideStream, c->evTables, 0));
  NK(h, N->AllReduce(h->regAdd.p, h->regAdd.p, (size_t)LB * P.reg.nAdd, ncclUint32, ncclSum, c->side, c->sideStream));
  NK(h, N->AllReduce(h->regMax.p, h->regMax.p, (size_t)LB * P.reg.nMax, ncclUint32, ncclMax, c->side, c->sideStream));
  if (P.reg.nF64) NK(h, N->AllReduce(h->regF64.p, h->regF64.p, (size_t)LB * P.reg.nF64, ncclFloat64, ncclSum, c->side, c->sideStream));
  CK(cudaEventRecord(c->evSide, c->sideStream));
  NK(h, N->AllReduce(h->flags.p, h->flags.p, (size_t)LB, ncclUint32, ncclMax, c->main, st));
  s = glia_index_regions(h);  // scan + read-back of R
  if (s != GLIA_OK) return s;
  const long long R = h->R;
  // interest mask
  {
    const size_t n = (size_t)std::max<long long>(R, 1) * world;
    CK(h->interest.ensure(n));
    CK(cudaMemsetAsync(h->interest.p, 0, n, st));
    const uint32_t chunk = (uint32_t)std::max<long long>(1, (R + world - 1) / world);
    edge_interest_kernel<<<nblk(P.edgeCap, 256), 256, 0, st>>>((const unsigned long long*)h->edgeKeys.p, P.edgeCap,
                                                               h->rank.as<uint32_t>(), chunk, world, h->interest.as<unsigned char>());
    h->launches += 2;
    CK(cudaGetLastError());
    if (R > 0) NK(h, N->AllReduce(h->interest.p, h->interest.p, (size_t)R * world, ncclUint8, ncclMax, c->main, st));
  }
  // boundary records: count per destination, exchange the counts, pack, send / receive, merge
  const int rw = rec_words(P.edge);
  CK(h->exportCursor.ensure(8 * 65));
  CK(c->counts.ensure(8 * (size_t)(world + 2) * (world + 1) + 16));
  unsigned long long* cursor = h->exportCursor.as<unsigned long long>();
  unsigned long long* mine = c->counts.as<unsigned long long>();          // [world + 2]: counts, occupied, capacity
  unsigned long long* all = mine + (world + 2);                           // [world][world + 2]
  bool rebuild = false;
  std::vector<unsigned long long> info((size_t)world * (world + 2)), off(world), roff(world);
  unsigned long long total = 0, rtotal = 0;
  for (int attempt = 0; attempt < 2; ++attempt) {
    const int skip = rebuild ? -1 : me;
    CK(cudaMemsetAsync(cursor, 0, 8 * 65, st));
    export_edges_kernel<0><<<nblk(P.edgeCap, 256), 256, 0, st>>>(
        (const unsigned long long*)h->edgeKeys.p, P.edgeCap, h->rank.as<uint32_t>(), h->interest.as<unsigned char>(), world,
        P.edge, nullptr, nullptr, nullptr, cursor, rw, nullptr, skip);
    h->launches += 2;
    CK(cudaMemcpyAsync(mine, cursor, 8 * (size_t)world, cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(mine + world, cursor + 64, 8, cudaMemcpyDeviceToDevice, st));
    const unsigned long long capv = P.edgeCap;
    CK(cudaMemcpyAsync(mine + world + 1, &capv, 8, cudaMemcpyHostToDevice, st));
    NK(h, N->AllGather(mine, all, (size_t)world + 2, ncclUint64, c->main, st));
    CK(cudaMemcpyAsync(info.data(), all, 8 * info.size(), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    h->tableOccupied = (long long)info[(size_t)me * (world + 2) + world];
    bool tight = false;  // would some rank's table be more than 80 % full after the merge?
    for (int q = 0; q < world && !rebuild; ++q) {
      unsigned long long in = 0;
      for (int r = 0; r < world; ++r) in += info[(size_t)r * (world + 2) + q];
      tight = tight || (info[(size_t)q * (world + 2) + world] + in) * 10 > info[(size_t)q * (world + 2) + world + 1] * 8;
    }
    if (tight && !rebuild) { rebuild = true; continue; }
    total = rtotal = 0;
    for (int d = 0; d < world; ++d) { off[d] = total; total += info[(size_t)me * (world + 2) + d]; }
    for (int r = 0; r < world; ++r) { roff[r] = rtotal; rtotal += info[(size_t)r * (world + 2) + me]; }
    break;
  }
  CK(h->exportBuf.ensure(4 * (size_t)rw * (size_t)std::max<unsigned long long>(total, 1)));
  CK(c->recv.ensure(4 * (size_t)rw * (size_t)std::max<unsigned long long>(rtotal, 1)));
  CK(cudaMemcpyAsync(cursor, off.data(), 8 * (size_t)world, cudaMemcpyHostToDevice, st));
  export_edges_kernel<1><<<nblk(P.edgeCap, 256), 256, 0, st>>>(
      (const unsigned long long*)h->edgeKeys.p, P.edgeCap, h->rank.as<uint32_t>(), h->interest.as<unsigned char>(), world,
      P.edge, h->edgeAdd.as<uint32_t>(), h->edgeMax.as<uint32_t>(), h->edgeF64.as<double>(), cursor, rw,
      h->exportBuf.as<uint32_t>(), rebuild ? -1 : me);
  h->launches++;
  CK(cudaGetLastError());
  NK(h, N->GroupStart());
  for (int p = 0; p < world; ++p) {
    const unsigned long long ns = info[(size_t)me * (world + 2) + p], nr = info[(size_t)p * (world + 2) + me];
    if (ns) NK(h, N->Send(h->exportBuf.as<uint32_t>() + (size_t)off[p] * rw, (size_t)ns * rw, ncclUint32, p, c->main, st));
    if (nr) NK(h, N->Recv(c->recv.as<uint32_t>() + (size_t)roff[p] * rw, (size_t)nr * rw, ncclUint32, p, c->main, st));
  }
  NK(h, N->GroupEnd());
  CK(cudaEventRecord(c->evB, st));
  bool checkImport = false;
  if (rebuild) {  // rare path (labels far from spatial order): rebuild / grow the table from the received runs
    s = glia_b200_import_edges(h, rtotal ? c->recv.p : nullptr, (int64_t)rtotal, 1);
    if (s != GLIA_OK) return s;
  } else if (rtotal) {  // merge the received records into the table that already holds this rank's own; checked at the end
    VoxParams& PW = h->vp;
    CK(cudaMemsetAsync(PW.overflow, 0, 4, st));
    import_edges_kernel<<<nblk(rtotal, 256), 256, 0, st>>>(
        c->recv.as<uint32_t>(), (long long)rtotal, rw, PW.edge, (unsigned long long*)h->edgeKeys.p, PW.edgeCap,
        h->edgeAdd.as<uint32_t>(), h->edgeMax.as<uint32_t>(), h->edgeF64.as<double>(), PW.overflow);
    // glia_finalize_edges re-zeroes the scalar block that holds the flag: keep a copy for the check at the end
    CK(cudaMemcpyAsync(c->counts.as<unsigned long long>() + (size_t)(world + 2) * (world + 1), PW.overflow, 4, cudaMemcpyDeviceToDevice, st));
    h->launches += 2;
    checkImport = true;
  }
  const long long chunkR = std::max<long long>(1, (R + world - 1) / world);
  const long long lo = std::min<long long>(R, (long long)me * chunkR);
  const long long hi = me == world - 1 ? R : std::min<long long>(R, lo + chunkR);
  h->shardWorld = world; h->shardMe = me;
  s = glia_finalize_edges(h, lo, std::min<long long>(hi, R));
  h->shardWorld = 0;
  if (s != GLIA_OK) return s;
  CK(cudaStreamWaitEvent(st, c->evSide, 0));  // the reduced record tables
  s = glia_gather_regions_some(h, h->interest.as<unsigned char>(), world, me);
  if (s != GLIA_OK) return s;
  CK(cudaEventRecord(c->evC, st));
  int importOvf = 0;
  if (checkImport) CK(cudaMemcpyAsync(&importOvf, c->counts.as<unsigned long long>() + (size_t)(world + 2) * (world + 1), 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  if (importOvf) { h->err = "directed-boundary table full while merging received records: raise edge_capacity"; return GLIA_E_CAPACITY; }
  h->built = true;
  h->err.clear();
  float ms = 0;
  cudaEventElapsedTime(&ms, c->evA, c->evB); c->exchangeMs = ms;
  cudaEventElapsedTime(&ms, c->evB, c->evC); h->ms[1] = ms;  // graph: import + CSR + candidates + region gather
  return GLIA_OK;
}

// ---- K5 on a sharded build: the bc_feat stage for a given merge order, every rank featurising its share of the
// merges (hmt/main_bc_feat.cxx:76-95 is the reference's parfor over merges).  The directed boundaries of a sharded
// build are spread over the ranks, so the per-node tables merge.cu scatters them into are completed by all-reduces.
int glia_bc_feat_impl(glia_handle* h, const uint32_t* order, int64_t n64, const double* sal_in, double init_sal,
                      double sal_bias, double* rows, glia::TableReducer* red, int world, int rank, int64_t* firstOut,
                      int64_t* countOut);

namespace {
struct NcclReducer : glia::TableReducer {
  NcclApi* N; glia_comm* c; glia_handle* h;
  int all_reduce(void* buf, size_t count, Type t, Op op, cudaStream_t st) override {
    if (!count) return GLIA_OK;
    const ncclDataType_t dt = t == U32 ? ncclUint32 : (t == I32 ? ncclInt32 : ncclFloat64);
    const ncclRedOp_t ro = op == SUM ? ncclSum : (op == MAX ? ncclMax : ncclMin);
    NK(h, N->AllReduce(buf, buf, count, dt, ro, c->main, st));
    return GLIA_OK;
  }
};
}  // namespace

extern "C" int glia_b200_bc_feat_sharded(glia_handle* h, glia_comm* c, const uint32_t* order, int64_t n,
                                         const double* sal_in, double init_sal, double sal_bias, double* rows,
                                         int64_t* first, int64_t* count) {
  NcclApi* N = nccl_api();
  if (!h || !c || !N) return GLIA_E_ARG;
  if (c->device != h->prm.device) { h->err = "communicator lives on another device"; return GLIA_E_ARG; }
  NcclReducer red;
  red.N = N; red.c = c; red.h = h;
  return glia_bc_feat_impl(h, order, n, sal_in, init_sal, sal_bias, rows, &red, c->world, c->rank, first, count);
}
