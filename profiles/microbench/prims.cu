// Microbenchmarks of the primitives the voxel pass is built from (B200, sm_100a):
// cycles per warp-instruction for shared atomics (spread / same address, 32 and 64 bit),
// match.any, redux, ballot+popc, shuffles, global RED (spread / same line).
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o prims prims.cu ; run: ./prims
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 4096
template <int MODE>
__global__ void bench(unsigned long long* out, uint32_t* gbuf, int nwarps_active) {
  __shared__ uint32_t s32[8192];
  __shared__ unsigned long long s64[2048];
  int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < 8192; i += blockDim.x) s32[i] = 0;
  for (int i = tid; i < 2048; i += blockDim.x) s64[i] = 0;
  __syncthreads();
  uint32_t acc = 0, x = tid * 2654435761u;
  long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) {
    x = x * 1664525u + 1013904223u;
    if (MODE == 0) atomicAdd(&s32[(warp * 32 + lane) & 8191], 1u);                 // spread, conflict-free
    if (MODE == 1) atomicAdd(&s32[warp * 64], 1u);                                 // all lanes same address
    if (MODE == 2) atomicAdd(&s32[(x >> 8) & 8191], 1u);                           // random addresses
    if (MODE == 3) atomicAdd(&s64[(warp * 32 + lane) & 2047], 1ull);               // 64-bit spread
    if (MODE == 4) atomicAdd(&s64[(x >> 8) & 2047], 1ull);                         // 64-bit random
    if (MODE == 5) acc += __match_any_sync(0xffffffffu, (x >> 28));                // 16 groups
    if (MODE == 6) acc += __reduce_add_sync(0xffffffffu, x & 7);                   // redux
    if (MODE == 7) acc += __popc(__ballot_sync(0xffffffffu, x & 1));               // ballot+popc
    if (MODE == 8) acc += __shfl_xor_sync(0xffffffffu, x, 1);                      // shfl
    if (MODE == 9) atomicAdd(&gbuf[((size_t)blockIdx.x * blockDim.x + tid) * 33 % (1u << 24)], 1u);  // RED spread
    if (MODE == 10) atomicAdd(&gbuf[(blockIdx.x * 64 + (x >> 27)) % (1u << 24)], 1u);               // RED same 128B line
    if (MODE == 11) atomicMax(&s32[(x >> 8) & 8191], x);                           // smem max random
    if (MODE == 12) {                                                              // smem double add via CAS, spread
      double* p = (double*)&s64[(warp * 32 + lane) & 2047];
      atomicAdd(p, 1.0);
    }
    if (MODE == 13) { if (lane < 4) atomicAdd(&s32[(x >> 8) & 8191], 1u); }        // 4 active lanes random
    if (MODE == 14) { if (lane < 4) atomicAdd(&s64[(x >> 8) & 2047], 1ull); }
    if (MODE == 15) acc += __match_any_sync(0xffffffffu, (x >> 30));               // 4 groups
    if (MODE == 16) atomicAdd(&s32[((x >> 8) & 255) * 32 + (lane & 3)], 1u);       // random rows, 4 banks
    // ---- questions left open by round 1 (profiles/r1_v4_notes.txt): what an f64 sum costs per finished run
    if (MODE == 17) atomicAdd((double*)&s64[warp * 8], 1.0);                        // smem f64, all lanes same address
    if (MODE == 18) atomicAdd((double*)&s64[(warp * 8 + (lane >> 3)) & 2047], 1.0); // smem f64, 4 addresses x 8 lanes (a drain)
    if (MODE == 19) atomicAdd((double*)&s64[(x >> 8) & 2047], 1.0);                 // smem f64 random
    if (MODE == 20) { if (lane < 4) atomicAdd((double*)&s64[(x >> 8) & 2047], 1.0); }  // smem f64 random, 4 lanes
    if (MODE == 21) atomicAdd((double*)gbuf + (((size_t)blockIdx.x * 4096 + ((x >> 20) & 4095)) % (1u << 23)), 1.0);  // REDG.F64, 32 KB window per SM
    if (MODE == 22) { if (lane < 4) atomicAdd((double*)gbuf + (((size_t)blockIdx.x * 4096 + ((x >> 20) & 4095)) % (1u << 23)), 1.0); }
    if (MODE == 23) atomicAdd((double*)gbuf + ((x >> 9) & ((1u << 23) - 1)), 1.0);  // REDG.F64 random over 64 MB
  }
  long long t1 = clock64();
  if (tid == 0) out[blockIdx.x] = (unsigned long long)(t1 - t0);
  if (acc == 0x12345678u) gbuf[0] = acc;
}

template <int MODE>
void run(const char* name, int threads) {
  unsigned long long* out; uint32_t* g;
  cudaMalloc(&out, 8 * 148); cudaMalloc(&g, 4u << 24); cudaMemset(g, 0, 4u << 24);  // 64 MB
  bench<MODE><<<148, threads>>>(out, g, threads / 32);
  cudaDeviceSynchronize();
  bench<MODE><<<148, threads>>>(out, g, threads / 32);
  cudaDeviceSynchronize();
  unsigned long long h[148]; cudaMemcpy(h, out, 8 * 148, cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
  int warps = threads / 32;
  printf("%-44s threads=%4d  cycles/iter/SM=%8.2f  cycles per warp-instr=%7.2f\n", name, threads, avg / ITERS,
         avg / ITERS / warps);
  cudaFree(out); cudaFree(g);
}

int main() {
  for (int th : {128, 512, 1024}) {
    run<0>("ATOMS.ADD u32 spread (conflict-free)", th);
    run<1>("ATOMS.ADD u32 same address", th);
    run<2>("ATOMS.ADD u32 random", th);
    run<3>("ATOMS.ADD u64 spread", th);
    run<4>("ATOMS.ADD u64 random", th);
    run<5>("MATCH.ANY 16 groups", th);
    run<15>("MATCH.ANY 4 groups", th);
    run<6>("REDUX.ADD", th);
    run<7>("BALLOT+POPC", th);
    run<8>("SHFL", th);
    run<9>("RED.global u32 spread", th);
    run<10>("RED.global u32 same line", th);
    run<11>("ATOMS.MAX u32 random", th);
    run<12>("smem double atomicAdd spread", th);
    run<13>("ATOMS.ADD u32 random, 4 lanes active", th);
    run<14>("ATOMS.ADD u64 random, 4 lanes active", th);
    run<16>("ATOMS.ADD u32 random rows 4 banks", th);
    run<17>("smem f64 add, same address", th);
    run<18>("smem f64 add, 4 addresses x 8 lanes", th);
    run<19>("smem f64 add, random", th);
    run<20>("smem f64 add, random, 4 lanes active", th);
    run<21>("RED.global f64, 32 KB window per SM", th);
    run<22>("RED.global f64, 32 KB window, 4 lanes", th);
    run<23>("RED.global f64, random over 64 MB", th);
    printf("\n");
  }
  return 0;
}
