// Microbenchmark: TMEM -> register read throughput (tcgen05.ld) per SM for several shapes / warp counts.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I csrc -o /tmp/tmem_ld_bw tools/microbench/tmem_ld_bw.cu
#include "common.cuh"
void b200_set_error(const char*, ...) {}
using namespace b200;

__device__ __forceinline__ void ld_x64(uint32_t taddr, uint32_t (&r)[64]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,"
      "%32,%33,%34,%35,%36,%37,%38,%39,%40,%41,%42,%43,%44,%45,%46,%47,%48,%49,%50,%51,%52,%53,%54,%55,%56,%57,%58,%59,%60,%61,%62,%63}, [%64];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]),
        "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]),
        "=r"(r[30]), "=r"(r[31]), "=r"(r[32]), "=r"(r[33]), "=r"(r[34]), "=r"(r[35]), "=r"(r[36]), "=r"(r[37]), "=r"(r[38]), "=r"(r[39]),
        "=r"(r[40]), "=r"(r[41]), "=r"(r[42]), "=r"(r[43]), "=r"(r[44]), "=r"(r[45]), "=r"(r[46]), "=r"(r[47]), "=r"(r[48]), "=r"(r[49]),
        "=r"(r[50]), "=r"(r[51]), "=r"(r[52]), "=r"(r[53]), "=r"(r[54]), "=r"(r[55]), "=r"(r[56]), "=r"(r[57]), "=r"(r[58]), "=r"(r[59]),
        "=r"(r[60]), "=r"(r[61]), "=r"(r[62]), "=r"(r[63])
      : "r"(taddr) : "memory");
}

// mode 0: x32 + wait each; 1: 4 x x32 then wait; 2: x64 + wait; 3: 2 x x64 then wait; 4: x16 + wait
template <int MODE>
__global__ void bench(int iters, long long* out_cycles, uint32_t* sink) {
  __shared__ uint32_t tptr;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) { tmem_alloc<1>(smem_u32(&tptr), 512); tmem_relinquish<1>(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t base = tptr + (static_cast<uint32_t>((warp & 3) * 32) << 16);
  uint32_t acc = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {
#pragma unroll
      for (int c = 0; c < 4; ++c) { uint32_t r[32]; tmem_ld_32x32(base + c * 32, r); tmem_ld_wait(); acc += r[0] ^ r[31]; }
    } else if (MODE == 1) {
      uint32_t r0[32], r1[32], r2[32], r3[32];
      tmem_ld_32x32(base, r0); tmem_ld_32x32(base + 32, r1); tmem_ld_32x32(base + 64, r2); tmem_ld_32x32(base + 96, r3);
      tmem_ld_wait(); acc += r0[0] ^ r1[5] ^ r2[7] ^ r3[31];
    } else if (MODE == 2) {
#pragma unroll
      for (int c = 0; c < 2; ++c) { uint32_t r[64]; ld_x64(base + c * 64, r); tmem_ld_wait(); acc += r[0] ^ r[63]; }
    } else if (MODE == 3) {
      uint32_t r0[64], r1[64]; ld_x64(base, r0); ld_x64(base + 64, r1); tmem_ld_wait(); acc += r0[0] ^ r1[63];
    } else {
#pragma unroll
      for (int c = 0; c < 8; ++c) { uint32_t r[16]; tmem_ld_32x16(base + c * 16, r); tmem_ld_wait(); acc += r[0] ^ r[15]; }
    }
  }
  const long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) out_cycles[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  tc_fence_before(); __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc<1>(tptr, 512); }
}

int main() {
  long long* cyc; uint32_t* sink;
  cudaMalloc(&cyc, 1024 * 8); cudaMalloc(&sink, 148 * 1024 * 4);
  const int iters = 2000;
  const char* names[] = {"x32 + wait each", "4 x x32 then wait", "x64 + wait each", "2 x x64 then wait", "x16 + wait each"};
  for (int mode = 0; mode < 5; ++mode)
    for (int warps : {4, 8, 16}) {
      for (int rep = 0; rep < 2; ++rep) {
        switch (mode) {
          case 0: bench<0><<<148, warps * 32>>>(iters, cyc, sink); break;
          case 1: bench<1><<<148, warps * 32>>>(iters, cyc, sink); break;
          case 2: bench<2><<<148, warps * 32>>>(iters, cyc, sink); break;
          case 3: bench<3><<<148, warps * 32>>>(iters, cyc, sink); break;
          default: bench<4><<<148, warps * 32>>>(iters, cyc, sink); break;
        }
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
      }
      long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
      const double bytes = (double)iters * warps * 32 * 128 * 4;   // every warp reads 128 columns x 32 lanes x 4 B per iteration
      printf("%-20s warps=%2d : %8.1f cycles/iter  %6.1f B/cycle/SM\n", names[mode], warps, (double)h / iters, bytes / h);
    }
  return 0;
}
