// Microbenchmark: tcgen05.ld throughput while the tensor core is busy with tcgen05.mma (SS and TS forms).
#include "common.cuh"
void b200_set_error(const char*, ...) {}
using namespace b200;

// mode: 0 = no MMA, 1 = SS MMAs (A,B from smem), 2 = TS MMAs (A from TMEM)
__global__ void __launch_bounds__(320, 1) bench(int mode, int iters, long long* out, uint32_t* sink) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;
  __shared__ uint32_t tptr;
  __shared__ uint64_t bar;
  __shared__ volatile int stop;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); fence_mbar_init(); stop = 0; }
  for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem_raw + (sbase - smem_u32(smem_raw)))[i] = 0;
  if (warp == 0) { tmem_alloc<1>(smem_u32(&tptr), 512); tmem_relinquish<1>(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tb = tptr;
  long long mma_count = 0;
  if (warp == 1) {
    if (lane == 0 && mode > 0) {
      const uint32_t idesc = umma_idesc_bf16(128, 128, false, false);
      uint32_t phase = 0;
      while (!stop) {
        for (int k = 0; k < 8; ++k) {
          if (mode == 1)
            umma_f16_ss<1>(tb + 256, umma_smem_desc_sw128(sbase + k * 32, 16, 1024), umma_smem_desc_sw128(sbase + 32768 + k * 32, 16, 1024), idesc, 1);
          else
            umma_f16_ts<1>(tb + 256, tb + 384 + k * 8, umma_smem_desc_sw128(sbase + 32768 + k * 32, 16, 1024), idesc, 1);
        }
        umma_commit(smem_u32(&bar));
        mbar_wait(smem_u32(&bar), phase);
        phase ^= 1;
        mma_count += 8;
      }
    }
    __syncwarp();
  } else if (warp >= 2) {
    const uint32_t base = tb + (static_cast<uint32_t>((warp & 3) * 32) << 16);
    uint32_t acc = 0;
    asm volatile("bar.sync 1, 256;");
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      uint32_t r0[32], r1[32], r2[32], r3[32];
      tmem_ld_32x32(base, r0); tmem_ld_32x32(base + 32, r1); tmem_ld_32x32(base + 64, r2); tmem_ld_32x32(base + 96, r3);
      tmem_ld_wait(); acc += r0[0] ^ r1[5] ^ r2[7] ^ r3[31];
    }
    const long long t1 = clock64();
    asm volatile("bar.sync 1, 256;");
    if (threadIdx.x == 64) { out[blockIdx.x * 2] = t1 - t0; stop = 1; }
    sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  }
  tc_fence_before(); __syncthreads();
  if (warp == 1 && lane == 0) out[blockIdx.x * 2 + 1] = mma_count;
  if (warp == 0) { tc_fence_after(); tmem_dealloc<1>(tb, 512); }
}

int main() {
  long long* out; uint32_t* sink;
  cudaMalloc(&out, 148 * 16); cudaMalloc(&sink, 148 * 320 * 4);
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 70000);
  const int iters = 3000;
  const char* names[] = {"LDTM only", "LDTM + SS-MMA stream", "LDTM + TS-MMA stream"};
  for (int mode = 0; mode < 3; ++mode) {
    for (int rep = 0; rep < 2; ++rep) {
      bench<<<148, 320, 70000>>>(mode, iters, out, sink);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
    }
    long long h[2]; cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost);
    const double bytes = (double)iters * 8 * 32 * 128 * 4;
    printf("%-24s: %8.1f cycles per 8-warp x 16 KB read  -> %7.1f B/cycle/SM ; MMAs issued %lld (%.1f cycles per 128x128x16 MMA)\n", names[mode],
           (double)h[0] / iters, bytes / h[0], h[1], h[1] ? (double)h[0] / h[1] : 0.0);
  }
  return 0;
}
