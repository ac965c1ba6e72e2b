// Peer-memory data plane: symmetric buffers shared between the GPUs of one NVSwitch box and
// the collectives that run over them as ordinary CUDA kernels — every rank launches the same
// kernel on its own GPU, the kernels find each other through epoch-tagged flags in peer memory
// (st.release.sys / ld.acquire.sys), and the payload moves as 128-bit loads from mapped peer
// pointers over NVLink.  No NCCL on these paths.
//
// Protocol (one "channel" = one symmetric staging buffer + one pair of flag rows):
//   1. arrive : thread t of block 0 stores `seq` into peer t's arrive[channel][my_rank]
//   2. every block spins until its OWN arrive[channel][t] >= seq for all t  (inputs are ready)
//   3. payload: pull from peers (all-gather / reduce-scatter / broadcast / all-to-all) ...
//   4. depart : the last block to finish stores `seq` into every peer's depart[channel][my_rank]
//               and waits for all peers' depart flags, so when the kernel retires nobody is
//               still reading this rank's buffer and stream order protects the next writer.
// Flags only ever grow, so buffers are reused without a reset pass.  Spins carry a watchdog.
//
// Replaces the c10d/NCCL collectives the reference reaches through DDP, FSDP and DTensor
// (SURVEY.md §2.3, §2.4-D, §5.8): all_reduce (tests/torch_comm_bench.py:79, DDP buckets),
// broadcast (:49, DDP buffer sync), all_gather / reduce_scatter (FSDP, sequence parallel),
// barrier, all_to_all (Ulysses, docs/guide/08_sequence_parallel.md:41-80).
#include "vec.cuh"
#include <string.h>

namespace b200 {

constexpr int kMaxRanks = 8;
constexpr int kFlagStride = 32;  // uint32 slots per row (>= kMaxRanks, keeps rows 128-byte apart)

struct CommCtx {
  void* data[kMaxRanks];        // symmetric payload buffers (this channel), indexed by rank
  uint32_t* flags[kMaxRanks];   // signal pads: [channel][2][kFlagStride] uint32 per rank
  uint32_t* counter;            // local (not shared) per-channel block counter
  int rank, world, channel;
  uint32_t seq;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t* flag_row(const CommCtx& c, int owner, int which) {
  return c.flags[owner] + ((size_t)c.channel * 2 + which) * kFlagStride;
}

__device__ __forceinline__ void spin_until(const uint32_t* p, uint32_t seq, const char* what) {
  uint64_t t0 = 0;
  uint32_t spins = 0;
  while ((int32_t)(ld_acquire_sys(p) - seq) < 0) {
    if ((++spins & 0xfffu) == 0) {
      const uint64_t now = globaltimer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > 5 * B200_WATCHDOG_NS) {
        printf("b200dist watchdog: peer %s flag stuck (block %d thread %d want %u have %u)\n", what,
               blockIdx.x, threadIdx.x, seq, *p);
        __trap();
      }
    }
  }
}

// Steps 1+2.  Must be called by all threads of every block.
__device__ __forceinline__ void comm_arrive_and_wait(const CommCtx& c) {
  __syncthreads();
  if (blockIdx.x == 0 && threadIdx.x < c.world) {
    __threadfence_system();
    st_release_sys(flag_row(c, threadIdx.x, 0) + c.rank, c.seq);
  }
  if (threadIdx.x < c.world) spin_until(flag_row(c, c.rank, 0) + threadIdx.x, c.seq, "arrive");
  __syncthreads();
}

// Step 4.  Must be called by all threads of every block after their last peer access.
__device__ __forceinline__ void comm_depart(const CommCtx& c) {
  __shared__ int is_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const uint32_t done = atomicAdd(c.counter + c.channel, 1u) + 1u;
    is_last = (done == gridDim.x);
    if (is_last) c.counter[c.channel] = 0;
  }
  __syncthreads();
  if (is_last) {
    if (threadIdx.x < c.world) {
      __threadfence_system();
      st_release_sys(flag_row(c, threadIdx.x, 1) + c.rank, c.seq);
      spin_until(flag_row(c, c.rank, 1) + threadIdx.x, c.seq, "depart");
    }
    __syncthreads();
  }
}

// A mid-kernel barrier across ranks AND across all blocks of every rank (two-shot all-reduce).
// Uses flag row `which` (2 = extra row) with the same seq.  Grid must be co-resident.
__device__ __forceinline__ void comm_grid_barrier(const CommCtx& c, uint32_t* local_gate) {
  __shared__ int is_last2;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const uint32_t done = atomicAdd(c.counter + 64 + c.channel, 1u) + 1u;
    is_last2 = (done == gridDim.x);
    if (is_last2) c.counter[64 + c.channel] = 0;
  }
  __syncthreads();
  if (is_last2) {
    if (threadIdx.x < c.world) {
      __threadfence_system();
      uint32_t* mid = c.flags[threadIdx.x] + ((size_t)(32 + c.channel) * 2) * kFlagStride;  // dedicated rows
      st_release_sys(mid + c.rank, c.seq);
      const uint32_t* mine = c.flags[c.rank] + ((size_t)(32 + c.channel) * 2) * kFlagStride;
      spin_until(mine + threadIdx.x, c.seq, "mid");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      st_release_sys(local_gate, c.seq);
    }
  }
  if (threadIdx.x == 0) spin_until(local_gate, c.seq, "gate");
  __syncthreads();
}

// Weak (coherent) 128-bit load that does not allocate in L1: peer data is read exactly once.
__device__ __forceinline__ uint4 ld_peer_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}

template <typename T> struct Acc8 {
  float v[8];
  __device__ __forceinline__ void zero() {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = 0.f;
  }
  __device__ __forceinline__ void add(const T* p) {
    float t[8];
    Vec8<T>::load(p, t);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] += t[j];
  }
};

// Sum `world` peers' 8-element chunks at element offset `off` for kU consecutive grid-strided chunks:
// all kU * world 128-bit (bf16) / 2x128-bit (fp32) loads are issued before the first add, so one
// thread keeps kU * world NVLink requests in flight (latency ~2 us needs ~1 MB in flight per GPU).
template <typename T, int kU, int kW = kMaxRanks>
__device__ __forceinline__ void peer_sum(const CommCtx& c, long long off, const long long (&idx)[kU],
                                         const bool (&ok)[kU], float (&acc)[kU][8]) {
  constexpr int kV = sizeof(T) == 2 ? 1 : 2;
  uint4 raw[kU][kW][kV];
#pragma unroll
  for (int u = 0; u < kU; ++u)
#pragma unroll
    for (int t = 0; t < kW; ++t)
      if (t < c.world && ok[u]) {
        const T* p = reinterpret_cast<const T*>(c.data[(t + c.rank) % c.world]) + off + idx[u] * 8;
#pragma unroll
        for (int v = 0; v < kV; ++v) raw[u][t][v] = ld_peer_v4(reinterpret_cast<const uint4*>(p) + v);
      }
#pragma unroll
  for (int u = 0; u < kU; ++u) {
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[u][j] = 0.f;
#pragma unroll
    for (int t = 0; t < kW; ++t)
      if (t < c.world && ok[u]) {
        if constexpr (sizeof(T) == 2) {
          const uint4 r = raw[u][t][0];
          acc[u][0] += bf16lo(r.x); acc[u][1] += bf16hi(r.x); acc[u][2] += bf16lo(r.y); acc[u][3] += bf16hi(r.y);
          acc[u][4] += bf16lo(r.z); acc[u][5] += bf16hi(r.z); acc[u][6] += bf16lo(r.w); acc[u][7] += bf16hi(r.w);
        } else {
          const uint4 a = raw[u][t][0], b = raw[u][t][1];
          acc[u][0] += __uint_as_float(a.x); acc[u][1] += __uint_as_float(a.y); acc[u][2] += __uint_as_float(a.z);
          acc[u][3] += __uint_as_float(a.w); acc[u][4] += __uint_as_float(b.x); acc[u][5] += __uint_as_float(b.y);
          acc[u][6] += __uint_as_float(b.z); acc[u][7] += __uint_as_float(b.w);
        }
      }
  }
}

// ------------------------------ barrier -------------------------------------
__global__ void comm_barrier_kernel(CommCtx c) {
  comm_arrive_and_wait(c);
  comm_depart(c);
}

// ------------------------------ all-gather (pull) ----------------------------
// out[t * n .. (t+1) * n) = peer t's data[in_off .. in_off + n)   (elements of 16 bytes)
__global__ void __launch_bounds__(512, 1) comm_allgather_kernel(CommCtx c, uint4* __restrict__ out, long long in_off16,
                                      long long n16) {
  comm_arrive_and_wait(c);
  const long long total = n16 * c.world;
  const long long stride = (long long)gridDim.x * blockDim.x;
  constexpr int kAgU = 16;  // 256 bytes in flight per thread
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += stride * kAgU) {
    uint4 v[kAgU];
#pragma unroll
    for (int u = 0; u < kAgU; ++u) {
      const long long k = i + u * stride;
      if (k < total) {
        const int t = (int)(k / n16);
        const int src = (t + c.rank) % c.world;  // start with the local shard, spread peers
        v[u] = ld_peer_v4(reinterpret_cast<const uint4*>(c.data[src]) + in_off16 + (k - (long long)t * n16));
      }
    }
#pragma unroll
    for (int u = 0; u < kAgU; ++u) {
      const long long k = i + u * stride;
      if (k < total) {
        const int t = (int)(k / n16);
        const int src = (t + c.rank) % c.world;
        out[(long long)src * n16 + (k - (long long)t * n16)] = v[u];
      }
    }
  }
  comm_depart(c);
}

// ------------------------------ reduce-scatter (pull) ------------------------
// out[i] = scale * sum_t peer t's data[in_off + rank * n + i],  i in [0, n)  (n in elements, % 8 == 0)
template <typename TIn, typename TOut, int kW>
__global__ void __launch_bounds__(512, 1) comm_reducescatter_kernel(CommCtx c, TOut* __restrict__ out, long long in_off,
                                          long long n, float scale) {
  comm_arrive_and_wait(c);
  const long long n8 = n >> 3;
  constexpr int kU = (sizeof(TIn) == 2 ? 16 : 8) / kW;   // ~16 peer requests in flight per thread
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8; i += stride * kU) {
    long long idx[kU];
    bool ok[kU];
    float acc[kU][8];
#pragma unroll
    for (int u = 0; u < kU; ++u) { idx[u] = i + u * stride; ok[u] = idx[u] < n8; }
    peer_sum<TIn, kU, kW>(c, in_off + (long long)c.rank * n, idx, ok, acc);
#pragma unroll
    for (int u = 0; u < kU; ++u)
      if (ok[u]) {
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[u][j] *= scale;
        Vec8<TOut>::store(out + idx[u] * 8, acc[u]);
      }
  }
  comm_depart(c);
}

// reduce-scatter fused with the sharded AdamW step (FSDP gradient path):
// the summed gradient never touches HBM; master weight, moments and the bf16 shard are updated in place.
template <typename TIn, int kW>
__global__ void __launch_bounds__(512, 1) comm_rs_adamw_kernel(CommCtx c, long long in_off, long long n, float grad_scale,
                                     float* __restrict__ master, __nv_bfloat16* __restrict__ lowp,
                                     float* __restrict__ m, float* __restrict__ v, float lr, float beta1,
                                     float beta2, float eps, float wd, float bc1, float bc2_sqrt) {
  comm_arrive_and_wait(c);
  const long long n8 = n >> 3;
  const float step_size = lr / bc1, decay = 1.f - lr * wd;
  constexpr int kU = (sizeof(TIn) == 2 ? 16 : 8) / kW;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i0 = blockIdx.x * (long long)blockDim.x + threadIdx.x; i0 < n8; i0 += stride * kU) {
    long long idx[kU];
    bool ok[kU];
    float acc[kU][8];
#pragma unroll
    for (int u = 0; u < kU; ++u) { idx[u] = i0 + u * stride; ok[u] = idx[u] < n8; }
    peer_sum<TIn, kU, kW>(c, in_off + (long long)c.rank * n, idx, ok, acc);
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      if (!ok[u]) continue;
      const long long i = idx[u];
      float pv[8], mv[8], vv[8];
      Vec8<float>::load(master + i * 8, pv);
      Vec8<float>::load(m + i * 8, mv);
      Vec8<float>::load(v + i * 8, vv);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float g = acc[u][j] * grad_scale;
        mv[j] = beta1 * mv[j] + (1.f - beta1) * g;
        vv[j] = beta2 * vv[j] + (1.f - beta2) * g * g;
        pv[j] = pv[j] * decay - step_size * (mv[j] / (sqrtf(vv[j]) / bc2_sqrt + eps));
      }
      Vec8<float>::store(master + i * 8, pv);
      Vec8<float>::store(m + i * 8, mv);
      Vec8<float>::store(v + i * 8, vv);
      if (lowp) Vec8<__nv_bfloat16>::store(lowp + i * 8, pv);
    }
  }
  comm_depart(c);
}

// ------------------------------ all-reduce -----------------------------------
// one-shot: every rank sums all peers' buffers (small messages: one NVLink round trip)
template <typename TIn, typename TOut>
__global__ void __launch_bounds__(512, 1) comm_allreduce_oneshot_kernel(CommCtx c, TOut* __restrict__ out, long long in_off,
                                              long long n, float scale) {
  comm_arrive_and_wait(c);
  const long long n8 = n >> 3;
  constexpr int kU = sizeof(TIn) == 2 ? 2 : 1;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8; i += stride * kU) {
    long long idx[kU];
    bool ok[kU];
    float acc[kU][8];
#pragma unroll
    for (int u = 0; u < kU; ++u) { idx[u] = i + u * stride; ok[u] = idx[u] < n8; }
    peer_sum<TIn, kU>(c, in_off, idx, ok, acc);
#pragma unroll
    for (int u = 0; u < kU; ++u)
      if (ok[u]) {
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[u][j] *= scale;
        Vec8<TOut>::store(out + idx[u] * 8, acc[u]);
      }
  }
  comm_depart(c);
}

// two-shot: reduce own 1/world slice in place (scaled), grid+rank barrier, gather the reduced
// slices from their owners into `out`.  Per-GPU NVLink traffic ~ 2 * (W-1)/W * bytes.
// n must be a multiple of 8 * world.  The symmetric buffer is overwritten with the result slices.
template <typename T, typename TOut>
__global__ void __launch_bounds__(512, 1) comm_allreduce_twoshot_kernel(CommCtx c, TOut* __restrict__ out, long long in_off,
                                              long long n, float scale, uint32_t* local_gate) {
  comm_arrive_and_wait(c);
  const long long slice = n / c.world;
  const long long s8 = slice >> 3;
  T* mine = reinterpret_cast<T*>(c.data[c.rank]) + in_off + (long long)c.rank * slice;
  {
    constexpr int kU = sizeof(T) == 2 ? 2 : 1;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < s8; i += stride * kU) {
      long long idx[kU];
      bool ok[kU];
      float acc[kU][8];
#pragma unroll
      for (int u = 0; u < kU; ++u) { idx[u] = i + u * stride; ok[u] = idx[u] < s8; }
      peer_sum<T, kU>(c, in_off + (long long)c.rank * slice, idx, ok, acc);
#pragma unroll
      for (int u = 0; u < kU; ++u)
        if (ok[u]) {
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[u][j] *= scale;
          Vec8<T>::store(mine + idx[u] * 8, acc[u]);
        }
    }
  }
  comm_grid_barrier(c, local_gate);
  const long long total8 = s8 * c.world;
  for (long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x; k < total8;
       k += (long long)gridDim.x * blockDim.x) {
    const int t = (int)(k / s8);
    const int src = (t + c.rank) % c.world;
    const long long i = k - (long long)t * s8;
    float v8[8];
    // peers' reduced slices were written during this kernel: bypass the non-coherent path
    const T* p = reinterpret_cast<const T*>(c.data[src]) + in_off + (long long)src * slice + i * 8;
    if constexpr (sizeof(T) == 2) {
      uint4 u;
      asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w) : "l"(p));
      v8[0] = bf16lo(u.x); v8[1] = bf16hi(u.x); v8[2] = bf16lo(u.y); v8[3] = bf16hi(u.y);
      v8[4] = bf16lo(u.z); v8[5] = bf16hi(u.z); v8[6] = bf16lo(u.w); v8[7] = bf16hi(u.w);
    } else {
      uint4 u0, u1;
      asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(u0.x), "=r"(u0.y), "=r"(u0.z), "=r"(u0.w) : "l"(p));
      asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(u1.x), "=r"(u1.y), "=r"(u1.z), "=r"(u1.w) : "l"(p + 4));
      v8[0] = __uint_as_float(u0.x); v8[1] = __uint_as_float(u0.y); v8[2] = __uint_as_float(u0.z); v8[3] = __uint_as_float(u0.w);
      v8[4] = __uint_as_float(u1.x); v8[5] = __uint_as_float(u1.y); v8[6] = __uint_as_float(u1.z); v8[7] = __uint_as_float(u1.w);
    }
    Vec8<TOut>::store(out + (long long)src * slice + i * 8, v8);
  }
  comm_depart(c);
}

// ------------------------------ slab reduce (GEMM -> reduce-scatter, step 2) -------------------
// out[i] = sum_t local_stage[t * n + i] (+ residual[i]);  the slabs were written by the peers' GEMM epilogues.
__global__ void __launch_bounds__(512, 1) comm_slab_reduce_kernel(CommCtx c, __nv_bfloat16* __restrict__ out,
                                                                  const __nv_bfloat16* __restrict__ residual,
                                                                  long long n) {
  comm_arrive_and_wait(c);   // every rank's GEMM (and its peer stores) retired before it arrived here
  const long long n8 = n >> 3;
  const __nv_bfloat16* base = reinterpret_cast<const __nv_bfloat16*>(c.data[c.rank]);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8;
       i += (long long)gridDim.x * blockDim.x) {
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    if (residual) Vec8<__nv_bfloat16>::load(residual + i * 8, acc);
#pragma unroll
    for (int t = 0; t < kMaxRanks; ++t)
      if (t < c.world) {
        float v[8];
        Vec8<__nv_bfloat16>::load(base + (long long)t * n + i * 8, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += v[j];
      }
    Vec8<__nv_bfloat16>::store(out + i * 8, acc);
  }
  comm_depart(c);            // peers may refill our slabs only after we have read them
}

// ------------------------------ broadcast / all-to-all -----------------------
__global__ void __launch_bounds__(512, 1) comm_broadcast_kernel(CommCtx c, uint4* __restrict__ out, long long in_off16,
                                      long long n16, int root) {
  comm_arrive_and_wait(c);
  const uint4* src = reinterpret_cast<const uint4*>(c.data[root]) + in_off16;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n16;
       i += (long long)gridDim.x * blockDim.x)
    out[i] = ld_peer_v4(src + i);
  comm_depart(c);
}

// out[t * n ..] = peer t's data[in_off + rank * n ..]   (16-byte elements)
__global__ void __launch_bounds__(512, 1) comm_alltoall_kernel(CommCtx c, uint4* __restrict__ out, long long in_off16,
                                     long long n16) {
  comm_arrive_and_wait(c);
  const long long total = n16 * c.world;
  for (long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x; k < total;
       k += (long long)gridDim.x * blockDim.x) {
    const int t = (int)(k / n16);
    const int src = (t + c.rank) % c.world;
    const long long i = k - (long long)t * n16;
    out[(long long)src * n16 + i] =
        ld_peer_v4(reinterpret_cast<const uint4*>(c.data[src]) + in_off16 + (long long)c.rank * n16 + i);
  }
  comm_depart(c);
}

}  // namespace b200

using namespace b200;

// ------------------------------------------------------------------ peer memory (CUDA IPC)
B200_API int b200_peer_alloc(size_t bytes, void** out) {
  void* p = nullptr;
  B200_CUDA_CHECK(cudaMalloc(&p, bytes));
  B200_CUDA_CHECK(cudaMemset(p, 0, bytes));
  B200_CUDA_CHECK(cudaDeviceSynchronize());
  *out = p;
  return B200_OK;
}
B200_API int b200_peer_free(void* p) {
  B200_CUDA_CHECK(cudaFree(p));
  return B200_OK;
}
B200_API int b200_peer_export(void* p, unsigned char* handle64) {
  cudaIpcMemHandle_t h;
  B200_CUDA_CHECK(cudaIpcGetMemHandle(&h, p));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle64, &h, 64);
  return B200_OK;
}
B200_API int b200_peer_import(const unsigned char* handle64, void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  B200_CUDA_CHECK(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
  return B200_OK;
}
B200_API int b200_peer_close(void* p) {
  B200_CUDA_CHECK(cudaIpcCloseMemHandle(p));
  return B200_OK;
}

// Copy-engine transfer between (peer-mapped) device buffers: no SM is involved, so an all-gather built from
// barrier + peer copies leaves every SM to the GEMMs it overlaps with.
B200_API int b200_peer_memcpy(void* dst, const void* src, size_t bytes, cudaStream_t stream) {
  B200_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, stream));
  return B200_OK;
}

// ------------------------------------------------------------------ collectives
// ctx layout passed from Python (all pointers are device addresses valid on this GPU):
//   data_ptrs[world], flag_ptrs[world], counter, rank, world, channel, seq
static CommCtx make_ctx(const void* const* data_ptrs, const void* const* flag_ptrs, void* counter,
                        int rank, int world, int channel, unsigned seq) {
  CommCtx c = {};
  for (int i = 0; i < world && i < kMaxRanks; ++i) {
    c.data[i] = const_cast<void*>(data_ptrs[i]);
    c.flags[i] = reinterpret_cast<uint32_t*>(const_cast<void*>(flag_ptrs[i]));
  }
  c.counter = reinterpret_cast<uint32_t*>(counter);
  c.rank = rank; c.world = world; c.channel = channel; c.seq = seq;
  return c;
}

static int comm_grid(long long work_items, int threads, int max_ctas) {
  long long g = (work_items + threads - 1) / threads;
  int cap = max_ctas > 0 ? max_ctas : b200_sm_count();
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

#define COMM_ARGS const void* const* data_ptrs, const void* const* flag_ptrs, void* counter, int rank, \
                  int world, int channel, unsigned seq
#define COMM_CTX make_ctx(data_ptrs, flag_ptrs, counter, rank, world, channel, seq)
#define COMM_CHECK_WORLD()                                                             \
  if (world < 1 || world > kMaxRanks || channel < 0 || channel >= 32) {                \
    b200_set_error("comm: world %d / channel %d out of range", world, channel);        \
    return B200_ERR_LIMIT;                                                             \
  }

B200_API int b200_comm_barrier(COMM_ARGS, cudaStream_t stream) {
  COMM_CHECK_WORLD();
  comm_barrier_kernel<<<1, 32, 0, stream>>>(COMM_CTX);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// bytes and in_off_bytes must be multiples of 16
B200_API int b200_comm_allgather(COMM_ARGS, void* out, long long in_off_bytes, long long bytes,
                                 int max_ctas, cudaStream_t stream) {
  COMM_CHECK_WORLD();
  if ((bytes | in_off_bytes) & 15) { b200_set_error("allgather: sizes must be multiples of 16 bytes"); return B200_ERR_SHAPE; }
  const long long n16 = bytes >> 4;
  comm_allgather_kernel<<<comm_grid(n16 * world / 16, 512, max_ctas), 512, 0, stream>>>(
      COMM_CTX, (uint4*)out, in_off_bytes >> 4, n16);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// n = elements per output shard (multiple of 8); in_off in elements
B200_API int b200_comm_reducescatter(COMM_ARGS, void* out, long long in_off, long long n, float scale,
                                     int in_dtype, int out_dtype, int max_ctas, cudaStream_t stream) {
  COMM_CHECK_WORLD();
  if (n & 7) { b200_set_error("reducescatter: shard elements must be a multiple of 8"); return B200_ERR_SHAPE; }
  const int grid = comm_grid(n >> 3, 512, max_ctas);
#define RS(TI, TO)                                                                                            \
  do {                                                                                                        \
    if (world <= 2) comm_reducescatter_kernel<TI, TO, 2><<<grid, 512, 0, stream>>>(COMM_CTX, (TO*)out, in_off, n, scale);      \
    else if (world <= 4) comm_reducescatter_kernel<TI, TO, 4><<<grid, 512, 0, stream>>>(COMM_CTX, (TO*)out, in_off, n, scale); \
    else comm_reducescatter_kernel<TI, TO, 8><<<grid, 512, 0, stream>>>(COMM_CTX, (TO*)out, in_off, n, scale);                 \
  } while (0)
  if (in_dtype == 0 && out_dtype == 0) RS(__nv_bfloat16, __nv_bfloat16);
  else if (in_dtype == 0 && out_dtype == 1) RS(__nv_bfloat16, float);
  else if (in_dtype == 1 && out_dtype == 1) RS(float, float);
  else if (in_dtype == 1 && out_dtype == 0) RS(float, __nv_bfloat16);
  else { b200_set_error("reducescatter: bad dtype"); return B200_ERR_UNSUPPORTED; }
#undef RS
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_comm_rs_adamw(COMM_ARGS, long long in_off, long long n, float grad_scale, int in_dtype,
                                float* master, void* lowp, float* m, float* v, float lr, float beta1,
                                float beta2, float eps, float wd, float bc1, float bc2, int max_ctas,
                                cudaStream_t stream) {
  COMM_CHECK_WORLD();
  if (n & 7) { b200_set_error("rs_adamw: shard elements must be a multiple of 8"); return B200_ERR_SHAPE; }
  const int grid = comm_grid(n >> 3, 512, max_ctas);
#define RSA(TI, KW) comm_rs_adamw_kernel<TI, KW><<<grid, 512, 0, stream>>>(COMM_CTX, in_off, n, grad_scale, master, \
        (__nv_bfloat16*)lowp, m, v, lr, beta1, beta2, eps, wd, bc1, sqrtf(bc2))
  if (in_dtype == 0) { if (world <= 2) RSA(__nv_bfloat16, 2); else if (world <= 4) RSA(__nv_bfloat16, 4); else RSA(__nv_bfloat16, 8); }
  else { if (world <= 2) RSA(float, 2); else if (world <= 4) RSA(float, 4); else RSA(float, 8); }
#undef RSA
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// algo: 1 = one-shot, 2 = two-shot (n % (8*world) == 0; needs a co-resident grid, capped at the SM count)
B200_API int b200_comm_allreduce(COMM_ARGS, void* out, long long in_off, long long n, float scale,
                                 int in_dtype, int out_dtype, int algo, void* local_gate, int max_ctas,
                                 cudaStream_t stream) {
  COMM_CHECK_WORLD();
  if (n & 7) { b200_set_error("allreduce: elements must be a multiple of 8"); return B200_ERR_SHAPE; }
  if (algo == 2 && (n % (8LL * world))) algo = 1;
  if (algo == 1) {
    const int grid = comm_grid(n >> 3, 512, max_ctas);
#define AR1(TI, TO) comm_allreduce_oneshot_kernel<TI, TO><<<grid, 512, 0, stream>>>(COMM_CTX, (TO*)out, in_off, n, scale)
    if (in_dtype == 0 && out_dtype == 0) AR1(__nv_bfloat16, __nv_bfloat16);
    else if (in_dtype == 0 && out_dtype == 1) AR1(__nv_bfloat16, float);
    else if (in_dtype == 1 && out_dtype == 1) AR1(float, float);
    else if (in_dtype == 1 && out_dtype == 0) AR1(float, __nv_bfloat16);
    else { b200_set_error("allreduce: bad dtype"); return B200_ERR_UNSUPPORTED; }
#undef AR1
  } else {
    int cap = b200_sm_count();
    if (max_ctas > 0 && max_ctas < cap) cap = max_ctas;
    const int grid = comm_grid((n / world) >> 3, 512, cap);
    uint32_t* gate = reinterpret_cast<uint32_t*>(local_gate) + channel;
#define AR2(T, TO) comm_allreduce_twoshot_kernel<T, TO><<<grid, 512, 0, stream>>>(COMM_CTX, (TO*)out, in_off, n, scale, gate)
    if (in_dtype == 0 && out_dtype == 0) AR2(__nv_bfloat16, __nv_bfloat16);
    else if (in_dtype == 0 && out_dtype == 1) AR2(__nv_bfloat16, float);
    else if (in_dtype == 1 && out_dtype == 1) AR2(float, float);
    else if (in_dtype == 1 && out_dtype == 0) AR2(float, __nv_bfloat16);
    else { b200_set_error("allreduce: bad dtype"); return B200_ERR_UNSUPPORTED; }
#undef AR2
  }
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// n = elements of this rank's chunk (chunk_rows * N, multiple of 8); staging = data_ptrs[rank], [world, n] bf16
B200_API int b200_comm_slab_reduce(COMM_ARGS, void* out, const void* residual, long long n, int max_ctas,
                                   cudaStream_t stream) {
  COMM_CHECK_WORLD();
  if (n & 7) { b200_set_error("slab_reduce: elements must be a multiple of 8"); return B200_ERR_SHAPE; }
  comm_slab_reduce_kernel<<<comm_grid(n >> 3, 512, max_ctas), 512, 0, stream>>>(
      COMM_CTX, (__nv_bfloat16*)out, (const __nv_bfloat16*)residual, n);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_comm_broadcast(COMM_ARGS, void* out, long long in_off_bytes, long long bytes, int root,
                                 int max_ctas, cudaStream_t stream) {
  COMM_CHECK_WORLD();
  if ((bytes | in_off_bytes) & 15) { b200_set_error("broadcast: sizes must be multiples of 16 bytes"); return B200_ERR_SHAPE; }
  comm_broadcast_kernel<<<comm_grid(bytes >> 4, 512, max_ctas), 512, 0, stream>>>(COMM_CTX, (uint4*)out,
                                                                                  in_off_bytes >> 4, bytes >> 4, root);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_comm_alltoall(COMM_ARGS, void* out, long long in_off_bytes, long long bytes_per_peer,
                                int max_ctas, cudaStream_t stream) {
  COMM_CHECK_WORLD();
  if ((bytes_per_peer | in_off_bytes) & 15) { b200_set_error("alltoall: sizes must be multiples of 16 bytes"); return B200_ERR_SHAPE; }
  comm_alltoall_kernel<<<comm_grid((bytes_per_peer >> 4) * world, 512, max_ctas), 512, 0, stream>>>(
      COMM_CTX, (uint4*)out, in_off_bytes >> 4, bytes_per_peer >> 4);
  B200_LAUNCH_CHECK();
  return B200_OK;
}
