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
  while ((int32_t)(ld_acquire_sys(p) - seq) < 0) peer_spin_check(t0, spins, what, seq, p);
}

// Steps 1+2.  Must be called by all threads of every block.
__device__ __forceinline__ void comm_arrive_and_wait(const CommCtx& c) {
  if (c.channel < 0) return;  // local variant: the payload was staged by copy engines, ranks already synchronised
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
  if (c.channel < 0) return;
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

// 128-bit load of data a peer stored DURING this kernel (ordered after the flag acquire, never served from a stale line).
__device__ __forceinline__ uint4 ld_sys_v4(const void* p) {
  uint4 r;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
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

// fp32 vectors are 32 bytes per thread.  A lane-contiguous layout would make each 128-bit load instruction of a warp
// touch 32 half-used 32-byte sectors, i.e. every sector would cross NVLink twice.  So whenever a full warp's worth of
// vectors (32 x 8 floats = 1 KB) lies inside the range, lane l takes floats [4l, 4l+4) and [128 + 4l, 128 + 4l + 4) of that
// 1 KB run: both load instructions are fully coalesced 512-byte reads.  bf16 vectors (16 bytes) are contiguous anyway.
// Requires vector index i == lane (mod 32), true for every loop here (block size and strides are multiples of 32).
template <typename T>
__device__ __forceinline__ void vec_offsets(long long i, long long n8, long long& o0, long long& o1) {
  if constexpr (sizeof(T) == 2) { o0 = i * 8; o1 = o0 + 4; }
  else {
    const long long w = i & ~31LL;
    if (w + 32 <= n8) { o0 = w * 8 + (i & 31) * 4; o1 = o0 + 128; }
    else { o0 = i * 8; o1 = o0 + 4; }
  }
}
__device__ __forceinline__ void load4(const float* p, float* v) {
  const float4 t = *reinterpret_cast<const float4*>(p);
  v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
__device__ __forceinline__ void store4(float* p, const float* v) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
__device__ __forceinline__ void store4(__nv_bfloat16* p, const float* v) {
  const __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
  uint2 u;
  u.x = *reinterpret_cast<const uint32_t*>(&a); u.y = *reinterpret_cast<const uint32_t*>(&b);
  *reinterpret_cast<uint2*>(p) = u;
}
// store / load the 8 values of vector i in the layout that input type TIn dictates (see vec_offsets)
template <typename TIn, typename TOut>
__device__ __forceinline__ void store_vec(TOut* out, long long i, long long n8, const float (&v)[8]) {
  if constexpr (sizeof(TIn) == 2) Vec8<TOut>::store(out + i * 8, v);
  else {
    long long o0, o1;
    vec_offsets<TIn>(i, n8, o0, o1);
    store4(out + o0, v); store4(out + o1, v + 4);
  }
}
template <typename TIn>
__device__ __forceinline__ void load_vec_f32(const float* src, long long i, long long n8, float (&v)[8]) {
  if constexpr (sizeof(TIn) == 2) Vec8<float>::load(src + i * 8, v);
  else {
    long long o0, o1;
    vec_offsets<TIn>(i, n8, o0, o1);
    load4(src + o0, v); load4(src + o1, v + 4);
  }
}

// Sum `world` peers' 8-element vectors at element offset `off` for kU consecutive grid-strided vectors:
// all kU * world 128-bit (bf16) / 2x128-bit (fp32) loads are issued before the first add, so one
// thread keeps kU * world NVLink requests in flight (latency ~2 us needs ~1 MB in flight per GPU).
template <typename T, int kU, int kW = kMaxRanks>
__device__ __forceinline__ void peer_sum(const CommCtx& c, long long off, const long long (&idx)[kU],
                                         const bool (&ok)[kU], float (&acc)[kU][8], long long n8) {
  constexpr int kV = sizeof(T) == 2 ? 1 : 2;
  uint4 raw[kU][kW][kV];
#pragma unroll
  for (int u = 0; u < kU; ++u) {
    long long o0 = 0, o1 = 0;
    if (ok[u]) vec_offsets<T>(idx[u], n8, o0, o1);
#pragma unroll
    for (int t = 0; t < kW; ++t)
      if (t < c.world && ok[u]) {
        const T* p = reinterpret_cast<const T*>(c.data[(t + c.rank) % c.world]) + off;
        raw[u][t][0] = ld_peer_v4(p + o0);
        if constexpr (kV == 2) raw[u][t][1] = ld_peer_v4(p + o1);
      }
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
    peer_sum<TIn, kU, kW>(c, in_off + (long long)c.rank * n, idx, ok, acc, n8);
#pragma unroll
    for (int u = 0; u < kU; ++u)
      if (ok[u]) {
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[u][j] *= scale;
        store_vec<TIn, TOut>(out, idx[u], n8, acc[u]);
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
    peer_sum<TIn, kU, kW>(c, in_off + (long long)c.rank * n, idx, ok, acc, n8);
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      if (!ok[u]) continue;
      const long long i = idx[u];
      float pv[8], mv[8], vv[8];
      load_vec_f32<TIn>(master, i, n8, pv);
      load_vec_f32<TIn>(m, i, n8, mv);
      load_vec_f32<TIn>(v, i, n8, vv);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float g = acc[u][j] * grad_scale;
        mv[j] = beta1 * mv[j] + (1.f - beta1) * g;
        vv[j] = beta2 * vv[j] + (1.f - beta2) * g * g;
        pv[j] = pv[j] * decay - step_size * (mv[j] / (sqrtf(vv[j]) / bc2_sqrt + eps));
      }
      store_vec<TIn, float>(master, i, n8, pv);
      store_vec<TIn, float>(m, i, n8, mv);
      store_vec<TIn, float>(v, i, n8, vv);
      if (lowp) store_vec<TIn, __nv_bfloat16>(lowp, i, n8, pv);
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
    peer_sum<TIn, kU>(c, in_off, idx, ok, acc, n8);
#pragma unroll
    for (int u = 0; u < kU; ++u)
      if (ok[u]) {
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[u][j] *= scale;
        store_vec<TIn, TOut>(out, idx[u], n8, acc[u]);
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
      peer_sum<T, kU>(c, in_off + (long long)c.rank * slice, idx, ok, acc, s8);
#pragma unroll
      for (int u = 0; u < kU; ++u)
        if (ok[u]) {
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[u][j] *= scale;
          store_vec<T, T>(mine, idx[u], s8, acc[u]);
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
    const T* p = reinterpret_cast<const T*>(c.data[src]) + in_off + (long long)src * slice;
    if constexpr (sizeof(T) == 2) {
      uint4 u;
      asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w) : "l"(p + i * 8));
      v8[0] = bf16lo(u.x); v8[1] = bf16hi(u.x); v8[2] = bf16lo(u.y); v8[3] = bf16hi(u.y);
      v8[4] = bf16lo(u.z); v8[5] = bf16hi(u.z); v8[6] = bf16lo(u.w); v8[7] = bf16hi(u.w);
    } else {
      // s8 % 32 != 0 would break "i == lane (mod 32)" across slice boundaries: vec_offsets then stays on the safe path
      long long o0, o1;
      if ((s8 & 31) == 0) vec_offsets<T>(i, s8, o0, o1); else { o0 = i * 8; o1 = o0 + 4; }
      uint4 u0, u1;
      asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(u0.x), "=r"(u0.y), "=r"(u0.z), "=r"(u0.w) : "l"(p + o0));
      asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(u1.x), "=r"(u1.y), "=r"(u1.z), "=r"(u1.w) : "l"(p + o1));
      v8[0] = __uint_as_float(u0.x); v8[1] = __uint_as_float(u0.y); v8[2] = __uint_as_float(u0.z); v8[3] = __uint_as_float(u0.w);
      v8[4] = __uint_as_float(u1.x); v8[5] = __uint_as_float(u1.y); v8[6] = __uint_as_float(u1.z); v8[7] = __uint_as_float(u1.w);
      TOut* o = out + (long long)src * slice;
      store4(o + o0, v8); store4(o + o1, v8 + 4);
      continue;
    }
    Vec8<TOut>::store(out + (long long)src * slice + i * 8, v8);
  }
  comm_depart(c);
}

// End of the store phase of a push kernel: the last block of the grid raises this rank's flag at every peer (after all
// the grid's stores), then every block waits until all peers' flags have arrived.
__device__ __forceinline__ void push_publish_and_wait(const CommCtx& c, int* is_last_smem) {
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t done = atomicAdd(c.counter + 128 + c.channel, 1u) + 1u;
    *is_last_smem = (done == gridDim.x);
    if (*is_last_smem) c.counter[128 + c.channel] = 0;
  }
  __syncthreads();
  if (*is_last_smem && threadIdx.x < c.world) {
    __threadfence_system();
    st_release_sys(flag_row(c, threadIdx.x, 0) + c.rank, c.seq);
  }
  if (threadIdx.x < c.world) spin_until(flag_row(c, c.rank, 0) + threadIdx.x, c.seq, "push");
  __syncthreads();
}

// push one-shot (latency path for small messages): no staging copy, no depart handshake.
//   1. every rank stores its input straight from the user tensor into slot [parity][rank] of EVERY peer's slot buffer;
//   2. the last block to finish raises this rank's arrive flag at every peer (release: the stores are visible first);
//   3. each rank waits for all flags and sums the `world` slots out of its OWN memory in rank order, so the result is
//      bitwise identical on every rank.
// Two slot sets alternate (parity): a rank can run at most one collective ahead of the slowest peer (it needs that
// peer's flag to pass step 3), so the set it overwrites was fully consumed two collectives ago.
// RS = 1 turns it into a reduce-scatter: `in` holds world slices of n elements and peer t receives only slice t.
template <typename TIn, typename TOut, int RS = 0>
__global__ void __launch_bounds__(512, 1) comm_allreduce_push_kernel(CommCtx c, const TIn* __restrict__ in,
                                                                    TOut* __restrict__ out, long long n, float scale,
                                                                    long long slot_elems, int parity) {
  __shared__ int is_last_p;
  constexpr int kV = sizeof(TIn) == 2 ? 1 : 2;    // 16-byte pieces per 8-element vector
  const long long nv = (n >> 3) * kV;             // 16-byte pieces in the message
  const long long stride = (long long)gridDim.x * blockDim.x;
  const uint4* src = reinterpret_cast<const uint4*>(in);
  const long long my_slot16 = ((long long)(parity * c.world + c.rank) * slot_elems * sizeof(TIn)) >> 4;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nv; i += stride) {
    if constexpr (RS) {
      uint4 v[kMaxRanks];
#pragma unroll
      for (int t = 0; t < kMaxRanks; ++t)
        if (t < c.world) v[t] = src[(long long)((c.rank + t) % c.world) * nv + i];
#pragma unroll
      for (int t = 0; t < kMaxRanks; ++t)
        if (t < c.world) reinterpret_cast<uint4*>(c.data[(c.rank + t) % c.world])[my_slot16 + i] = v[t];
    } else {
      const uint4 v = src[i];
#pragma unroll
      for (int t = 0; t < kMaxRanks; ++t)
        if (t < c.world) reinterpret_cast<uint4*>(c.data[(c.rank + t) % c.world])[my_slot16 + i] = v;
    }
  }
  push_publish_and_wait(c, &is_last_p);
  const uint4* slots = reinterpret_cast<const uint4*>(c.data[c.rank]);
  const long long set16 = ((long long)parity * c.world * slot_elems * sizeof(TIn)) >> 4;
  const long long slot16 = (slot_elems * sizeof(TIn)) >> 4;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nv; i += stride) {
    float acc[sizeof(TIn) == 2 ? 8 : 4];
#pragma unroll
    for (int j = 0; j < (int)(sizeof(acc) / sizeof(float)); ++j) acc[j] = 0.f;
#pragma unroll
    for (int t = 0; t < kMaxRanks; ++t)
      if (t < c.world) {
        uint4 r;   // written by a peer during this kernel: coherent load
        asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(slots + set16 + t * slot16 + i));
        if constexpr (sizeof(TIn) == 2) {
          acc[0] += bf16lo(r.x); acc[1] += bf16hi(r.x); acc[2] += bf16lo(r.y); acc[3] += bf16hi(r.y);
          acc[4] += bf16lo(r.z); acc[5] += bf16hi(r.z); acc[6] += bf16lo(r.w); acc[7] += bf16hi(r.w);
        } else {
          acc[0] += __uint_as_float(r.x); acc[1] += __uint_as_float(r.y);
          acc[2] += __uint_as_float(r.z); acc[3] += __uint_as_float(r.w);
        }
      }
#pragma unroll
    for (int j = 0; j < (int)(sizeof(acc) / sizeof(float)); ++j) acc[j] *= scale;
    if constexpr (sizeof(TIn) == 2) Vec8<TOut>::store(out + i * 8, reinterpret_cast<const float(&)[8]>(acc));
    else store4(out + i * 4, acc);
  }
}

// The same protocol for pure data movement.  root < 0: all-gather (every rank stores its n16 pieces into slot [rank] of
// every peer, then copies the world slots out in rank order).  root >= 0: broadcast (only the root stores, into slot
// [root]; the others copy that slot out).  Every rank raises and awaits every flag in both modes: that is what bounds
// the skew between ranks to one collective and makes the two alternating slot sets sufficient.
__global__ void __launch_bounds__(512, 1) comm_gather_push_kernel(CommCtx c, const uint4* __restrict__ in,
                                                                 uint4* __restrict__ out, long long n16,
                                                                 long long slot16, int parity, int root) {
  __shared__ int is_last_p;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long set16 = (long long)parity * c.world * slot16;
  if (root < 0 || root == c.rank) {
    const long long my_slot16 = set16 + (long long)c.rank * slot16;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n16; i += stride) {
      const uint4 v = in[i];
#pragma unroll
      for (int t = 0; t < kMaxRanks; ++t)
        if (t < c.world && (root < 0 || t > 0)) reinterpret_cast<uint4*>(c.data[(c.rank + t) % c.world])[my_slot16 + i] = v;
    }
  }
  push_publish_and_wait(c, &is_last_p);
  const uint4* slots = reinterpret_cast<const uint4*>(c.data[c.rank]) + set16;
  if (root < 0) {
    for (int t = 0; t < c.world; ++t)
      for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n16; i += stride)
        out[(long long)t * n16 + i] = ld_sys_v4(slots + (long long)t * slot16 + i);
  } else if (root != c.rank) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n16; i += stride)
      out[i] = ld_sys_v4(slots + (long long)root * slot16 + i);
  }
}

// two-stage broadcast for large messages: rank r first pulls slice r from the root into its own symmetric buffer
// (root egress = one message), then after a rank+grid barrier every rank gathers the other slices from their new
// owners, so the root's links stop being the bottleneck (a flat pull costs the root (W-1) messages of egress).
// n16 must be a multiple of `world`; grid must be co-resident.
__global__ void __launch_bounds__(512, 1) comm_broadcast2_kernel(CommCtx c, uint4* __restrict__ out, long long in_off16,
                                                                long long n16, int root, uint32_t* local_gate) {
  comm_arrive_and_wait(c);
  const long long s16 = n16 / c.world;
  const long long stride = (long long)gridDim.x * blockDim.x;
  constexpr int kBU = 8;
  {
    const uint4* src = reinterpret_cast<const uint4*>(c.data[root]) + in_off16 + (long long)c.rank * s16;
    uint4* keep = reinterpret_cast<uint4*>(c.data[c.rank]) + in_off16 + (long long)c.rank * s16;
    uint4* dst = out + (long long)c.rank * s16;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < s16; i += stride * kBU) {
      uint4 v[kBU];
#pragma unroll
      for (int u = 0; u < kBU; ++u) if (i + u * stride < s16) v[u] = ld_peer_v4(src + i + u * stride);
#pragma unroll
      for (int u = 0; u < kBU; ++u)
        if (i + u * stride < s16) {
          if (c.rank != root) keep[i + u * stride] = v[u];
          dst[i + u * stride] = v[u];
        }
    }
  }
  comm_grid_barrier(c, local_gate);
  const long long rest = s16 * (c.world - 1);
  for (long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x; k < rest; k += stride * kBU) {
    uint4 v[kBU];
#pragma unroll
    for (int u = 0; u < kBU; ++u) {
      const long long kk = k + u * stride;
      if (kk < rest) {
        const int t = (int)(kk / s16) + 1;
        const int owner = (c.rank + t) % c.world;
        const uint4* p = reinterpret_cast<const uint4*>(c.data[owner]) + in_off16 + (long long)owner * s16 + (kk - (long long)(t - 1) * s16);
        asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(p));
      }
    }
#pragma unroll
    for (int u = 0; u < kBU; ++u) {
      const long long kk = k + u * stride;
      if (kk < rest) {
        const int t = (int)(kk / s16) + 1;
        const int owner = (c.rank + t) % c.world;
        out[(long long)owner * s16 + (kk - (long long)(t - 1) * s16)] = v[u];
      }
    }
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

// ------------------------------ neighbour (halo) exchange -------------------
// Each rank has placed its two edge strips in its symmetric buffer: slot 0 = what the PREVIOUS rank needs (my first rows),
// slot 1 = what the NEXT rank needs (my last rows), n16 16-byte elements each.  After the arrive handshake a rank reads
// only its two neighbours: from_prev <- rank-1's slot 1, from_next <- rank+1's slot 0 (zeros at the ends of the chain).
// Payload per rank is two strips whatever the world size (domain-parallel halo exchange, docs/guide/10_domain_parallel.md
// of the reference :69-103).
__global__ void __launch_bounds__(512, 1) comm_neighbor_kernel(CommCtx c, uint4* __restrict__ from_prev,
                                                              uint4* __restrict__ from_next, long long in_off16,
                                                              long long n16) {
  comm_arrive_and_wait(c);
  const uint4 zero = make_uint4(0, 0, 0, 0);
  const uint4* prev = c.rank > 0 ? reinterpret_cast<const uint4*>(c.data[c.rank - 1]) + in_off16 + n16 : nullptr;
  const uint4* next = c.rank + 1 < c.world ? reinterpret_cast<const uint4*>(c.data[c.rank + 1]) + in_off16 : nullptr;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n16;
       i += (long long)gridDim.x * blockDim.x) {
    from_prev[i] = prev ? ld_peer_v4(prev + i) : zero;
    from_next[i] = next ? ld_peer_v4(next + i) : zero;
  }
  comm_depart(c);
}

}  // namespace b200

using namespace b200;

B200_DEFINE_PEER_WD_SETTER(b200_peer_wd_set_comm)

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

// Same update, but the `world` gradient slabs are all LOCAL pointers (slab_ptrs[r] = n elements contributed by rank r):
// the caller moved the peers' slices here with copy-engine transfers (b200_peer_memcpy) between two flag barriers, so
// this kernel touches no peer memory, needs no rank synchronisation and runs at HBM speed instead of NVLink speed.
B200_API int b200_comm_rs_adamw_local(const void* const* slab_ptrs, int world, long long n, float grad_scale, int in_dtype,
                                      float* master, void* lowp, float* m, float* v, float lr, float beta1,
                                      float beta2, float eps, float wd, float bc1, float bc2, int max_ctas,
                                      cudaStream_t stream) {
  using namespace b200;
  if (world < 1 || world > kMaxRanks) { b200_set_error("rs_adamw_local: world %d out of range", world); return B200_ERR_LIMIT; }
  if (n & 7) { b200_set_error("rs_adamw_local: shard elements must be a multiple of 8"); return B200_ERR_SHAPE; }
  CommCtx c = {};
  for (int r = 0; r < world; ++r) c.data[r] = const_cast<void*>(slab_ptrs[r]);
  c.rank = 0; c.world = world; c.channel = -1;
  // 256-thread CTAs, twice as many: this kernel runs on FSDP's side stream, where a CTA must fit on an SM next to a CTA of a
  // main-stream kernel that is waiting for other GPUs (see nvls_sync_kernel in multicast.cu)
  const int grid = 2 * comm_grid(n >> 3, 512, max_ctas);
#define RSA(TI, KW) comm_rs_adamw_kernel<TI, KW><<<grid, 256, 0, stream>>>(c, 0, n, grad_scale, master, \
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

// ------------------------------ point-to-point signalling (pipeline stage boundaries) ------------------------------
// A stage boundary is a producer/consumer copy: the sender moves the tensor into a mailbox slot in the RECEIVER's
// symmetric memory (b200_peer_memcpy on a side stream, copy engines) and then raises the slot's flag; the receiver's
// compute stream waits on the flag and uses the slot in place.  Replaces dist.isend/irecv of
// scripts/04_pipeline_parallel_pp/*.py (torch.distributed.pipelining's batch_isend_irecv) without NCCL's per-communicator
// stream ordering, which serialises the two directions of a middle stage.
__global__ void p2p_signal_kernel(uint32_t* remote_flag, uint32_t value) {
  __threadfence_system();
  b200::st_release_sys(remote_flag, value);
}
__global__ void p2p_wait_kernel(const uint32_t* local_flag, uint32_t value) {
  b200::spin_until(local_flag, value, "p2p");
}
B200_API int b200_p2p_signal(void* remote_flag, unsigned value, cudaStream_t stream) {
  p2p_signal_kernel<<<1, 1, 0, stream>>>(reinterpret_cast<uint32_t*>(remote_flag), value);
  B200_LAUNCH_CHECK();
  return B200_OK;
}
B200_API int b200_p2p_wait(const void* local_flag, unsigned value, cudaStream_t stream) {
  p2p_wait_kernel<<<1, 1, 0, stream>>>(reinterpret_cast<const uint32_t*>(local_flag), value);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// Small-message all-reduce without staging copies (see comm_allreduce_push_kernel).  data_ptrs = the symmetric SLOT
// buffer [2][world][slot_elems] of in_dtype; `in` / `out` are ordinary device tensors (may alias); n % 8 == 0,
// n <= slot_elems, `in` and `out` 16-byte aligned.  parity must alternate between consecutive calls on a channel.
B200_API int b200_comm_allreduce_push(COMM_ARGS, const void* in, void* out, long long n, float scale, int in_dtype,
                                      int out_dtype, long long slot_elems, int parity, cudaStream_t stream) {
  COMM_CHECK_WORLD();
  if ((n & 7) || n > slot_elems || (slot_elems & 7) || ((uintptr_t)in & 15) || ((uintptr_t)out & 15)) {
    b200_set_error("allreduce_push: n=%lld slot=%lld must be multiples of 8 with n <= slot, 16-byte aligned tensors", n, slot_elems);
    return B200_ERR_SHAPE;
  }
  const long long nv = (n >> 3) * (in_dtype == 0 ? 1 : 2);
  const int grid = comm_grid(nv, 512, 16);
#define ARP(TI, TO) comm_allreduce_push_kernel<TI, TO><<<grid, 512, 0, stream>>>(COMM_CTX, (const TI*)in, (TO*)out, n, scale, \
                                                                               slot_elems, parity & 1)
  if (in_dtype == 0 && out_dtype == 0) ARP(__nv_bfloat16, __nv_bfloat16);
  else if (in_dtype == 0 && out_dtype == 1) ARP(__nv_bfloat16, float);
  else if (in_dtype == 1 && out_dtype == 1) ARP(float, float);
  else if (in_dtype == 1 && out_dtype == 0) ARP(float, __nv_bfloat16);
  else { b200_set_error("allreduce_push: bad dtype"); return B200_ERR_UNSUPPORTED; }
#undef ARP
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// Small-message reduce-scatter, same protocol: `in` = world slices of n elements, out = scale * sum of slice [rank].
B200_API int b200_comm_reducescatter_push(COMM_ARGS, const void* in, void* out, long long n, float scale, int in_dtype,
                                          int out_dtype, long long slot_elems, int parity, cudaStream_t stream) {
  COMM_CHECK_WORLD();
  if ((n & 7) || n > slot_elems || (slot_elems & 7) || ((uintptr_t)in & 15) || ((uintptr_t)out & 15)) {
    b200_set_error("reducescatter_push: n=%lld slot=%lld must be multiples of 8 with n <= slot, 16-byte aligned tensors", n, slot_elems);
    return B200_ERR_SHAPE;
  }
  const long long nv = (n >> 3) * (in_dtype == 0 ? 1 : 2);
  const int grid = comm_grid(nv, 512, 16);
#define RSP(TI, TO) comm_allreduce_push_kernel<TI, TO, 1><<<grid, 512, 0, stream>>>(COMM_CTX, (const TI*)in, (TO*)out, n, \
                                                                                  scale, slot_elems, parity & 1)
  if (in_dtype == 0 && out_dtype == 0) RSP(__nv_bfloat16, __nv_bfloat16);
  else if (in_dtype == 0 && out_dtype == 1) RSP(__nv_bfloat16, float);
  else if (in_dtype == 1 && out_dtype == 1) RSP(float, float);
  else if (in_dtype == 1 && out_dtype == 0) RSP(float, __nv_bfloat16);
  else { b200_set_error("reducescatter_push: bad dtype"); return B200_ERR_UNSUPPORTED; }
#undef RSP
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// Small-message all-gather (root < 0: out = world x bytes, rank order) or broadcast (root >= 0: `out` receives the
// root's `in`; on the root itself nothing is written).  data_ptrs = the symmetric slot buffer [2][world][slot_bytes].
B200_API int b200_comm_gather_push(COMM_ARGS, const void* in, void* out, long long bytes, long long slot_bytes, int root,
                                   int parity, cudaStream_t stream) {
  COMM_CHECK_WORLD();
  if (((bytes | slot_bytes) & 15) || bytes > slot_bytes || ((uintptr_t)in & 15) || ((uintptr_t)out & 15) || root >= world) {
    b200_set_error("gather_push: bytes=%lld slot=%lld must be multiples of 16 with bytes <= slot, 16-byte aligned tensors, root < world",
                   bytes, slot_bytes);
    return B200_ERR_SHAPE;
  }
  comm_gather_push_kernel<<<comm_grid(bytes >> 4, 512, 16), 512, 0, stream>>>(
      COMM_CTX, (const uint4*)in, (uint4*)out, bytes >> 4, slot_bytes >> 4, parity & 1, root);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// Large-message broadcast in two stages (see comm_broadcast2_kernel): bytes % (16 * world) == 0.
B200_API int b200_comm_broadcast2(COMM_ARGS, void* out, long long in_off_bytes, long long bytes, int root,
                                  void* local_gate, int max_ctas, cudaStream_t stream) {
  COMM_CHECK_WORLD();
  if (((bytes | in_off_bytes) & 15) || (bytes >> 4) % world) {
    b200_set_error("broadcast2: bytes must be a multiple of 16 * world");
    return B200_ERR_SHAPE;
  }
  int cap = b200_sm_count();
  if (max_ctas > 0 && max_ctas < cap) cap = max_ctas;
  uint32_t* gate = reinterpret_cast<uint32_t*>(local_gate) + channel;
  comm_broadcast2_kernel<<<comm_grid((bytes >> 4) / world, 512, cap), 512, 0, stream>>>(
      COMM_CTX, (uint4*)out, in_off_bytes >> 4, bytes >> 4, root, gate);
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

// Halo exchange along a chain of ranks (see comm_neighbor_kernel).  data_ptrs: symmetric buffer holding, at in_off_bytes,
// [slot 0: strip for rank-1][slot 1: strip for rank+1], `bytes` (multiple of 16) each.
B200_API int b200_comm_neighbor_exchange(COMM_ARGS, void* from_prev, void* from_next, long long in_off_bytes, long long bytes,
                                         int max_ctas, cudaStream_t stream) {
  COMM_CHECK_WORLD();
  if ((bytes & 15) || (in_off_bytes & 15)) { b200_set_error("neighbor_exchange: sizes must be multiples of 16 bytes"); return B200_ERR_SHAPE; }
  const long long n16 = bytes >> 4;
  const int grid = comm_grid(n16, 512, max_ctas);
  comm_neighbor_kernel<<<grid, 512, 0, stream>>>(COMM_CTX, (uint4*)from_prev, (uint4*)from_next, in_off_bytes >> 4, n16);
  B200_LAUNCH_CHECK();
  return B200_OK;
}
