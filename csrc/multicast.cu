// NVLS data plane: NVSwitch multicast objects (cuMulticast*) + multimem.ld_reduce / multimem.st kernels.
//
// A multicast object groups one physical allocation per GPU behind ONE virtual address.  A load with
// `multimem.ld_reduce` from that address is answered by the SWITCH with the element-wise sum of all GPUs' copies; a
// `multimem.st` is replicated by the switch into all copies.  An all-reduce therefore costs each GPU one read of its
// 1/world slice (reduced in the fabric) and one write of that slice (broadcast by the fabric): the SMs never see the
// peers' partial values.  Per GPU and direction the links carry (1 + 1/world) x bytes (every copy is read once by the
// switch, own slice included, and written once), against 2 (world-1)/world x bytes for the pull two-shot: NVLS wins
// from world = 4 and approaches 2x at world = 8 - the algorithm NCCL uses on this box.
// The pull kernels in comm_kernels.cu stay as the general path and the default below 4 ranks.
//
// Host side (all driver entry points are resolved with cudaGetDriverEntryPoint; the library does not link libcuda):
//   rank 0 : b200_mc_create  -> multicast handle + POSIX fd (the fd is passed to the peers over an AF_UNIX socket by
//            the Python layer, SCM_RIGHTS)
//   others : b200_mc_import(fd)
//   all    : b200_mc_add_device, [barrier], b200_mc_bind_map -> local unicast pointer + multicast pointer
// Replaces what NCCL does internally for the reference's dist.all_reduce call sites (tests/torch_comm_bench.py:79, DDP).
#include "vec.cuh"
#include <cuda.h>
#include <string.h>

namespace b200 {

template <typename F>
static F drv(const char* name) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
    return nullptr;
  return reinterpret_cast<F>(fn);
}

#define MC_DRV(var, name, ...)                                                                   \
  auto var = drv<CUresult (*)(__VA_ARGS__)>(name);                                               \
  if (!var) { b200_set_error("multicast: driver entry point %s unavailable", name); return B200_ERR_UNSUPPORTED; }
#define MC_CHECK(call, what)                                                                     \
  do {                                                                                           \
    CUresult _r = (call);                                                                        \
    if (_r != CUDA_SUCCESS) { b200_set_error("multicast: %s failed with CUresult %d", what, (int)_r); return B200_ERR_CUDA; } \
  } while (0)

static CUmulticastObjectProp mc_prop(int world, size_t bytes) {
  CUmulticastObjectProp p;
  memset(&p, 0, sizeof(p));
  p.numDevices = (unsigned)world;
  p.size = bytes;
  p.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  p.flags = 0;
  return p;
}

struct McState {   // one per b200_mc_bind_map, returned to the caller as an opaque handle for b200_mc_release
  CUmemGenericAllocationHandle mc, mem;
  CUdeviceptr local_va, mc_va;
  size_t bytes;
};

// ---------------------------------------------------------------- kernels
__device__ __forceinline__ void mm_ldred_f32(const void* p, float (&v)[4]) {
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]) : "l"(p) : "memory");
}
__device__ __forceinline__ void mm_st_f32(void* p, const float (&v)[4]) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]) : "memory");
}
__device__ __forceinline__ uint4 mm_ldred_bf16(const void* p) {   // 8 bf16, accumulated in fp32 inside the switch
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void mm_st_bf16(void* p, uint4 r) {
  asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "r"(r.x), "r"(r.y), "r"(r.z), "r"(r.w) : "memory");
}

__device__ __forceinline__ void mc_st_release(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t mc_ld_acquire(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// flags: the Communicator's flag pads (same layout as comm_kernels.cu: [channel][2][32] uint32 per rank)
struct McCtx {
  uint32_t* flags[8];
  uint32_t* counter;
  int rank, world, channel;
  uint32_t seq;
  int split;   // 1: the waits for other GPUs happen in nvls_sync_kernel launches around this kernel (see there)
};
__device__ __forceinline__ uint32_t* mc_flag_row(const McCtx& c, int owner, int which) {
  return c.flags[owner] + ((size_t)c.channel * 2 + which) * 32;
}
__device__ __forceinline__ void mc_spin(const uint32_t* p, uint32_t seq) {
  uint64_t t0 = 0;
  uint32_t spins = 0;
  while ((int32_t)(mc_ld_acquire(p) - seq) < 0) peer_spin_check(t0, spins, "nvls flag", seq, p);
}

// ---- shared prologue / epilogue of the NVLS kernels ----
// A kernel that holds SMs while it waits for another GPU can deadlock against a second such kernel on another stream: on
// rank A kernel X owns the SMs and waits for B's X, on rank B kernel Y owns them and waits for A's Y (seen at TP 4 x
// FSDP 2: the fused all-gather GEMM of the main stream against reduce-scatter + AdamW of FSDP's side stream).  Callers
// that run beside compute therefore ask for the SPLIT form: the waits move into one-warp nvls_sync_kernel launches (which
// always find room on an SM) before and after, and the bulk kernel only signals - it never waits, so it always finishes.
__device__ __forceinline__ void mc_arrive_and_wait(const McCtx& c) {
  if (c.split) return;
  if (blockIdx.x == 0 && threadIdx.x < c.world) {
    __threadfence_system();
    mc_st_release(mc_flag_row(c, threadIdx.x, 0) + c.rank, c.seq);
  }
  if (threadIdx.x < c.world) mc_spin(mc_flag_row(c, c.rank, 0) + threadIdx.x, c.seq);
  __syncthreads();
}
// Last CTA of the grid tells every peer "I am done" and (unless split) waits until every peer said so.
__device__ __forceinline__ void mc_depart(const McCtx& c, int* is_last_smem) {
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t done = atomicAdd(c.counter + 192 + c.channel, 1u) + 1u;
    *is_last_smem = (done == gridDim.x);
    if (*is_last_smem) c.counter[192 + c.channel] = 0;
  }
  __syncthreads();
  if (*is_last_smem) {
    if (threadIdx.x < c.world) {
      __threadfence_system();
      mc_st_release(mc_flag_row(c, threadIdx.x, 1) + c.rank, c.seq);
      if (!c.split) mc_spin(mc_flag_row(c, c.rank, 1) + threadIdx.x, c.seq);
    }
    __syncthreads();
  }
}
// Not waiting is not enough: a bulk kernel whose CTA needs a whole SM's registers (512 threads x 126) cannot START while a
// peer-waiting GEMM of the main stream holds one CTA on every SM, and the sync kernels queued behind it - which the other
// ranks wait for - are stuck with it.  Split-mode bulk kernels therefore run 256-thread CTAs (32 K registers, twice as many
// CTAs): they fit beside a 192-thread GEMM CTA (23 K registers, 226 KB of shared memory) on the same SM.
constexpr int kSplitThreads = 256;
// phase 0: raise my arrive flag at every peer and wait for theirs; phase 1: wait for every peer's depart flag.
__global__ void __launch_bounds__(32, 1) nvls_sync_kernel(McCtx c, int phase) {
  if ((int)threadIdx.x >= c.world) return;
  if (phase == 0) {
    __threadfence_system();
    mc_st_release(mc_flag_row(c, threadIdx.x, 0) + c.rank, c.seq);
    mc_spin(mc_flag_row(c, c.rank, 0) + threadIdx.x, c.seq);
  } else {
    mc_spin(mc_flag_row(c, c.rank, 1) + threadIdx.x, c.seq);
  }
}

// In-place all-reduce of the multicast buffer: rank r reduces and re-broadcasts elements [r * slice, (r+1) * slice).
//   arrive : every rank's input is final (flag row 0)
//   body   : ld_reduce own slice through the switch, scale, multimem.st it back to every copy
//   depart : every rank's stores have landed everywhere (flag row 1) -> the whole buffer is valid locally
template <typename T>
__global__ void __launch_bounds__(512, 1) nvls_allreduce_kernel(McCtx c, T* __restrict__ mc, long long n, float scale) {
  __shared__ int is_last;
  mc_arrive_and_wait(c);
  constexpr int kPer = 16 / sizeof(T);               // elements per 16-byte access
  const long long slice = n / c.world;               // n % (world * kPer) == 0 (host-checked)
  const long long nv = slice / kPer;
  T* base = mc + (long long)c.rank * slice;
  const long long stride = (long long)gridDim.x * blockDim.x;
  constexpr int kU = 8;   // 128 bytes of switch requests in flight per thread
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nv; i += stride * kU) {
    if constexpr (sizeof(T) == 4) {
      float v[kU][4];
#pragma unroll
      for (int u = 0; u < kU; ++u) if (i + u * stride < nv) mm_ldred_f32(base + (i + u * stride) * 4, v[u]);
#pragma unroll
      for (int u = 0; u < kU; ++u)
        if (i + u * stride < nv) {
#pragma unroll
          for (int j = 0; j < 4; ++j) v[u][j] *= scale;
          mm_st_f32(base + (i + u * stride) * 4, v[u]);
        }
    } else {
      uint4 r[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) if (i + u * stride < nv) r[u] = mm_ldred_bf16(base + (i + u * stride) * 8);
#pragma unroll
      for (int u = 0; u < kU; ++u)
        if (i + u * stride < nv) {
          if (scale != 1.f) {
            r[u].x = pack_bf16(bf16lo(r[u].x) * scale, bf16hi(r[u].x) * scale);
            r[u].y = pack_bf16(bf16lo(r[u].y) * scale, bf16hi(r[u].y) * scale);
            r[u].z = pack_bf16(bf16lo(r[u].z) * scale, bf16hi(r[u].z) * scale);
            r[u].w = pack_bf16(bf16lo(r[u].w) * scale, bf16hi(r[u].w) * scale);
          }
          mm_st_bf16(base + (i + u * stride) * 8, r[u]);
        }
    }
  }
  mc_depart(c, &is_last);
}

// Where vector i (8 elements) of an fp32 array lives: a warp's 32 vectors cover 256 consecutive floats and lane l takes
// floats [4l, 4l + 4) and [128 + 4l, 128 + 4l + 4), so each of its two 16-byte requests belongs to a fully used 512-byte
// warp request.  (A lane reading 32 contiguous bytes as two requests would send every 32-byte sector over NVLink twice.)
// The ragged tail (fewer than 32 vectors left) falls back to the contiguous layout.  16-bit arrays: o0 = 8 i.
template <typename T>
__device__ __forceinline__ void mm_offsets(long long i, long long n8, long long& o0, long long& o1) {
  if constexpr (sizeof(T) == 2) { o0 = i * 8; o1 = o0 + 4; }
  else {
    const long long w = i & ~31LL;
    if (w + 32 <= n8) { o0 = w * 8 + (i & 31) * 4; o1 = o0 + 128; }
    else { o0 = i * 8; o1 = o0 + 4; }
  }
}
// vector i of the sum over all ranks' replicas, as fp32 (bf16 inputs are accumulated in fp32 by the switch and arrive
// rounded to bf16 once; fp32 inputs take two 16-byte requests at the offsets above).
template <typename TIn>
__device__ __forceinline__ void mm_ldred8(const TIn* base, long long i, long long n8, float (&v)[8]) {
  if constexpr (sizeof(TIn) == 2) {
    const uint4 r = mm_ldred_bf16(base + i * 8);
    v[0] = bf16lo(r.x); v[1] = bf16hi(r.x); v[2] = bf16lo(r.y); v[3] = bf16hi(r.y);
    v[4] = bf16lo(r.z); v[5] = bf16hi(r.z); v[6] = bf16lo(r.w); v[7] = bf16hi(r.w);
  } else {
    long long o0, o1;
    mm_offsets<TIn>(i, n8, o0, o1);
    float a[4], b[4];
    mm_ldred_f32(base + o0, a);
    mm_ldred_f32(base + o1, b);
    v[0] = a[0]; v[1] = a[1]; v[2] = a[2]; v[3] = a[3]; v[4] = b[0]; v[5] = b[1]; v[6] = b[2]; v[7] = b[3];
  }
}
__device__ __forceinline__ void mm_ld4(const float* p, float* v) {
  const float4 t = *reinterpret_cast<const float4*>(p);
  v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
__device__ __forceinline__ void mm_st4(float* p, const float* v) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
__device__ __forceinline__ void mm_st4(__nv_bfloat16* p, const float* v) {
  const __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
  uint2 u;
  u.x = *reinterpret_cast<const uint32_t*>(&a); u.y = *reinterpret_cast<const uint32_t*>(&b);
  *reinterpret_cast<uint2*>(p) = u;
}
// the 8 values of vector i of an ordinary array that is processed alongside inputs of type TIn (same element positions)
template <typename TIn, typename T>
__device__ __forceinline__ void mm_store8(T* out, long long i, long long n8, const float (&v)[8]) {
  if constexpr (sizeof(TIn) == 2) Vec8<T>::store(out + i * 8, v);
  else {
    long long o0, o1;
    mm_offsets<TIn>(i, n8, o0, o1);
    mm_st4(out + o0, v); mm_st4(out + o1, v + 4);
  }
}
template <typename TIn>
__device__ __forceinline__ void mm_load8(const float* src, long long i, long long n8, float (&v)[8]) {
  if constexpr (sizeof(TIn) == 2) Vec8<float>::load(src + i * 8, v);
  else {
    long long o0, o1;
    mm_offsets<TIn>(i, n8, o0, o1);
    mm_ld4(src + o0, v); mm_ld4(src + o1, v + 4);
  }
}

// Reduce-scatter through the switch: every rank's input [world * n] sits at `mc` (multicast address of the region);
// rank r reads slice r with multimem.ld_reduce - ONE read per element answered by the NVSwitch with the sum over all
// replicas - scales and writes it to its ordinary output tensor.  The SMs never see the peers' partial values and each GPU
// receives n elements instead of (world - 1) * n.
// (Measured at 4 GPUs, 1 GiB fp32: 4 or 8 vectors in flight per thread and one or two CTAs per SM all give 1.57 ms - the
// rate is set by the switch / links, not by the request depth.)
template <typename TIn, typename TOut>
__global__ void __launch_bounds__(512, 1) nvls_reducescatter_kernel(McCtx c, const TIn* __restrict__ mc, TOut* __restrict__ out,
                                                                      long long n, float scale) {
  __shared__ int is_last;
  mc_arrive_and_wait(c);
  const long long n8 = n >> 3;
  const TIn* base = mc + (long long)c.rank * n;
  const long long stride = (long long)gridDim.x * blockDim.x;
  constexpr int kU = sizeof(TIn) == 2 ? 8 : 4;   // 128 bytes of switch requests in flight per thread
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8; i += stride * kU) {
    float v[kU][8];
#pragma unroll
    for (int u = 0; u < kU; ++u) if (i + u * stride < n8) mm_ldred8<TIn>(base, i + u * stride, n8, v[u]);
#pragma unroll
    for (int u = 0; u < kU; ++u)
      if (i + u * stride < n8) {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[u][j] *= scale;
        mm_store8<TIn>(out, i + u * stride, n8, v[u]);
      }
  }
  mc_depart(c, &is_last);   // the region may be refilled only after every rank has read its slice
}

// The FSDP gradient step as ONE kernel: reduce-scatter in the switch + 1/world scale + AdamW on this rank's shard
// (fp32 master, moments) + bf16 write-back of the updated shard.  The summed gradient never exists in memory.
template <typename TIn>
__global__ void __launch_bounds__(512, 1) nvls_rs_adamw_kernel(McCtx c, const TIn* __restrict__ mc, long long n, float grad_scale,
                                                              float* __restrict__ master, __nv_bfloat16* __restrict__ lowp,
                                                              float* __restrict__ m, float* __restrict__ v, float lr,
                                                              float beta1, float beta2, float eps, float wd, float bc1,
                                                              float bc2_sqrt) {
  __shared__ int is_last;
  mc_arrive_and_wait(c);
  const long long n8 = n >> 3;
  const TIn* base = mc + (long long)c.rank * n;
  const float step_size = lr / bc1, decay = 1.f - lr * wd;
  const long long stride = (long long)gridDim.x * blockDim.x;
  constexpr int kU = 4;
  for (long long i0 = blockIdx.x * (long long)blockDim.x + threadIdx.x; i0 < n8; i0 += stride * kU) {
    float g[kU][8];
#pragma unroll
    for (int u = 0; u < kU; ++u) if (i0 + u * stride < n8) mm_ldred8<TIn>(base, i0 + u * stride, n8, g[u]);
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long i = i0 + u * stride;
      if (i >= n8) continue;
      float pv[8], mv[8], vv[8];
      mm_load8<TIn>(master, i, n8, pv);
      mm_load8<TIn>(m, i, n8, mv);
      mm_load8<TIn>(v, i, n8, vv);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float gj = g[u][j] * grad_scale;
        mv[j] = beta1 * mv[j] + (1.f - beta1) * gj;
        vv[j] = beta2 * vv[j] + (1.f - beta2) * gj * gj;
        pv[j] = pv[j] * decay - step_size * (mv[j] / (sqrtf(vv[j]) / bc2_sqrt + eps));
      }
      mm_store8<TIn>(master, i, n8, pv);
      mm_store8<TIn>(m, i, n8, mv);
      mm_store8<TIn>(v, i, n8, vv);
      if (lowp) mm_store8<TIn>(lowp, i, n8, pv);
    }
  }
  mc_depart(c, &is_last);
}

// All-gather through the switch: rank r streams its shard (any local pointer) ONCE with multimem.st into slot r of the
// region; the NVSwitch replicates it into every rank's replica, so a GPU sends `bytes` instead of (world - 1) * bytes.
// When the kernel completes every replica holds all shards (the host entry point then copies the replica out if asked).
__global__ void __launch_bounds__(512, 1) nvls_allgather_kernel(McCtx c, const uint4* __restrict__ src, uint4* __restrict__ mc,
                                                               long long n16) {
  __shared__ int is_last;
  mc_arrive_and_wait(c);   // everybody has left the previous collective on this channel: the region is free
  const long long stride = (long long)gridDim.x * blockDim.x;
  uint4* dst = mc + (long long)c.rank * n16;
  constexpr int kU = 8;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n16; i += stride * kU) {
    uint4 v[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) if (i + u * stride < n16) v[u] = src[i + u * stride];
#pragma unroll
    for (int u = 0; u < kU; ++u)
      if (i + u * stride < n16)
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                     :: "l"(dst + i + u * stride), "f"(__uint_as_float(v[u].x)), "f"(__uint_as_float(v[u].y)),
                        "f"(__uint_as_float(v[u].z)), "f"(__uint_as_float(v[u].w)) : "memory");
  }
  // all ranks' stores have been issued and fenced: once the last CTA has seen every peer's depart flag - i.e. when this
  // kernel completes - every replica is complete.  (The copy-out is a separate stream-ordered copy: CTAs that spun here
  // for the peers' LAST CTA would hold their SM slots while those peers' unlaunched CTAs may be waiting for one.)
  mc_depart(c, &is_last);
}

// Broadcast through the switch: the root streams its payload with multimem.st (one egress copy, replicated by the
// NVSwitch into every rank's buffer), then releases a flag; the other ranks copy their local replica out.
//   row 0 arrive/wait (all ranks): everybody has left the previous collective on this channel, the buffer is free
//   row 1, slot `root`            : "payload landed"
// `src` (root only) may be any device pointer, including the root's own replica (registered in-place use);
// `out` == nullptr skips the copy-out (the caller reads the replica in place).
__global__ void __launch_bounds__(512, 1) nvls_broadcast_kernel(McCtx c, const uint4* __restrict__ src, uint4* __restrict__ mc,
                                                               const uint4* __restrict__ replica, uint4* __restrict__ out,
                                                               long long n16, int root) {
  __shared__ int is_last;
  if (blockIdx.x == 0 && threadIdx.x < c.world) {
    __threadfence_system();
    mc_st_release(mc_flag_row(c, threadIdx.x, 0) + c.rank, c.seq);
  }
  if (threadIdx.x < c.world) mc_spin(mc_flag_row(c, c.rank, 0) + threadIdx.x, c.seq);
  __syncthreads();
  const long long stride = (long long)gridDim.x * blockDim.x;
  if (c.rank == root) {
    constexpr int kU = 8;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n16; i += stride * kU) {
      uint4 v[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) if (i + u * stride < n16) v[u] = src[i + u * stride];
#pragma unroll
      for (int u = 0; u < kU; ++u)
        if (i + u * stride < n16)
          asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                       :: "l"(mc + i + u * stride), "f"(__uint_as_float(v[u].x)), "f"(__uint_as_float(v[u].y)),
                          "f"(__uint_as_float(v[u].z)), "f"(__uint_as_float(v[u].w)) : "memory");
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
      const uint32_t done = atomicAdd(c.counter + 192 + c.channel, 1u) + 1u;
      is_last = (done == gridDim.x);
      if (is_last) c.counter[192 + c.channel] = 0;
    }
    __syncthreads();
    if (is_last && threadIdx.x < c.world) {
      __threadfence_system();
      mc_st_release(mc_flag_row(c, threadIdx.x, 1) + root, c.seq);
    }
    return;
  }
  if (threadIdx.x == 0) mc_spin(mc_flag_row(c, c.rank, 1) + root, c.seq);
  __syncthreads();
  if (out == nullptr) return;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n16; i += stride) {
    uint4 v;   // written by the switch during this kernel: coherent load
    asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(replica + i));
    out[i] = v;
  }
}

}  // namespace b200

using namespace b200;

B200_DEFINE_PEER_WD_SETTER(b200_peer_wd_set_mc)

// *supported = 1 when the current device can join NVSwitch multicast teams
B200_API int b200_mc_supported(int* supported) {
  int dev = 0, v = 0;
  B200_CUDA_CHECK(cudaGetDevice(&dev));
  MC_DRV(getattr, "cuDeviceGetAttribute", int*, CUdevice_attribute, CUdevice);
  MC_CHECK(getattr(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, (CUdevice)dev), "cuDeviceGetAttribute");
  *supported = v;
  return B200_OK;
}

// bytes is rounded up to the recommended multicast granularity; *rounded returns the size every rank must use
B200_API int b200_mc_granularity(size_t bytes, int world, size_t* rounded) {
  MC_DRV(getgran, "cuMulticastGetGranularity", size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUmulticastObjectProp p = mc_prop(world, bytes);
  size_t g = 0;
  MC_CHECK(getgran(&g, &p, CU_MULTICAST_GRANULARITY_RECOMMENDED), "cuMulticastGetGranularity");
  if (g == 0) g = 2u << 20;
  *rounded = (bytes + g - 1) / g * g;
  return B200_OK;
}

// rank 0: create the multicast object (bytes must come from b200_mc_granularity) and export it as a POSIX fd
B200_API int b200_mc_create(size_t bytes, int world, unsigned long long* handle, int* fd) {
  MC_DRV(create, "cuMulticastCreate", CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  MC_DRV(exportfd, "cuMemExportToShareableHandle", void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUmulticastObjectProp p = mc_prop(world, bytes);
  CUmemGenericAllocationHandle h;
  MC_CHECK(create(&h, &p), "cuMulticastCreate");
  int f = -1;
  MC_CHECK(exportfd(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0), "cuMemExportToShareableHandle");
  *handle = (unsigned long long)h;
  *fd = f;
  return B200_OK;
}

B200_API int b200_mc_import(int fd, unsigned long long* handle) {
  MC_DRV(importfd, "cuMemImportFromShareableHandle", CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUmemGenericAllocationHandle h;
  MC_CHECK(importfd(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR), "cuMemImportFromShareableHandle");
  *handle = (unsigned long long)h;
  return B200_OK;
}

B200_API int b200_mc_add_device(unsigned long long handle) {
  int dev = 0;
  B200_CUDA_CHECK(cudaGetDevice(&dev));
  MC_DRV(adddev, "cuMulticastAddDevice", CUmemGenericAllocationHandle, CUdevice);
  MC_CHECK(adddev((CUmemGenericAllocationHandle)handle, (CUdevice)dev), "cuMulticastAddDevice");
  return B200_OK;
}

// After EVERY rank has added its device: allocate this GPU's physical backing, map it (local_ptr), bind it to the
// multicast object and map the multicast object (mc_ptr).  *state is an opaque handle for b200_mc_release.
B200_API int b200_mc_bind_map(unsigned long long handle, size_t bytes, void** local_ptr, void** mc_ptr, void** state) {
  int dev = 0;
  B200_CUDA_CHECK(cudaGetDevice(&dev));
  MC_DRV(memcreate, "cuMemCreate", CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  MC_DRV(reserve, "cuMemAddressReserve", CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  MC_DRV(map, "cuMemMap", CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  MC_DRV(setaccess, "cuMemSetAccess", CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  MC_DRV(bind, "cuMulticastBindMem", CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
  MC_DRV(allocgran, "cuMemGetAllocationGranularity", size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUmemAllocationProp ap;
  memset(&ap, 0, sizeof(ap));
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = dev;
  ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t ag = 0;
  MC_CHECK(allocgran(&ag, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED), "cuMemGetAllocationGranularity");
  if (ag && bytes % ag) { b200_set_error("multicast: size %zu is not a multiple of the allocation granularity %zu", bytes, ag); return B200_ERR_SHAPE; }
  McState* st = new McState();
  st->mc = (CUmemGenericAllocationHandle)handle;
  st->bytes = bytes;
  MC_CHECK(memcreate(&st->mem, bytes, &ap, 0), "cuMemCreate");
  CUmemAccessDesc ad;
  memset(&ad, 0, sizeof(ad));
  ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ad.location.id = dev;
  ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  MC_CHECK(reserve(&st->local_va, bytes, 0, 0, 0), "cuMemAddressReserve(local)");
  MC_CHECK(map(st->local_va, bytes, 0, st->mem, 0), "cuMemMap(local)");
  MC_CHECK(setaccess(st->local_va, bytes, &ad, 1), "cuMemSetAccess(local)");
  MC_CHECK(bind(st->mc, 0, st->mem, 0, bytes, 0), "cuMulticastBindMem");
  MC_CHECK(reserve(&st->mc_va, bytes, 0, 0, 0), "cuMemAddressReserve(multicast)");
  MC_CHECK(map(st->mc_va, bytes, 0, st->mc, 0), "cuMemMap(multicast)");
  MC_CHECK(setaccess(st->mc_va, bytes, &ad, 1), "cuMemSetAccess(multicast)");
  B200_CUDA_CHECK(cudaMemset((void*)st->local_va, 0, bytes));
  B200_CUDA_CHECK(cudaDeviceSynchronize());
  *local_ptr = (void*)st->local_va;
  *mc_ptr = (void*)st->mc_va;
  *state = st;
  return B200_OK;
}

B200_API int b200_mc_release(void* state) {
  McState* st = reinterpret_cast<McState*>(state);
  if (!st) return B200_OK;
  auto unmap = drv<CUresult (*)(CUdeviceptr, size_t)>("cuMemUnmap");
  auto afree = drv<CUresult (*)(CUdeviceptr, size_t)>("cuMemAddressFree");
  auto release = drv<CUresult (*)(CUmemGenericAllocationHandle)>("cuMemRelease");
  auto unbind = drv<CUresult (*)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t)>("cuMulticastUnbind");
  int dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceSynchronize();
  if (unmap && afree && release) {
    unmap(st->mc_va, st->bytes); afree(st->mc_va, st->bytes);
    if (unbind) unbind(st->mc, (CUdevice)dev, 0, st->bytes);
    unmap(st->local_va, st->bytes); afree(st->local_va, st->bytes);
    release(st->mem); release(st->mc);
  }
  delete st;
  return B200_OK;
}

// In-place all-reduce (sum, then * scale) of the first n elements of the multicast buffer.  flag_ptrs / counter are the
// Communicator's signal pads; n must be a multiple of world * (16 / element size).
B200_API int b200_comm_allreduce_nvls(const void* const* flag_ptrs, void* counter, int rank, int world, int channel,
                                      unsigned seq, void* mc_ptr, long long n, float scale, int dtype, int max_ctas,
                                      int split, cudaStream_t stream) {
  if (world < 2 || world > 8) { b200_set_error("allreduce_nvls: world %d out of range", world); return B200_ERR_LIMIT; }
  const int per = dtype == 0 ? 8 : 4;
  if (n % ((long long)world * per)) { b200_set_error("allreduce_nvls: n must be a multiple of world * %d", per); return B200_ERR_SHAPE; }
  McCtx c;
  memset(&c, 0, sizeof(c));
  for (int r = 0; r < world; ++r) c.flags[r] = reinterpret_cast<uint32_t*>(const_cast<void*>(flag_ptrs[r]));
  c.counter = reinterpret_cast<uint32_t*>(counter);
  c.rank = rank; c.world = world; c.channel = channel; c.seq = seq; c.split = split ? 1 : 0;
  if (dtype != 0 && dtype != 1) { b200_set_error("allreduce_nvls: bad dtype"); return B200_ERR_UNSUPPORTED; }
  const long long nv = n / world / per;
  long long g = (nv + 512 * 8 - 1) / (512 * 8);
  int cap = max_ctas > 0 ? max_ctas : b200_sm_count();
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  const int threads = split ? kSplitThreads : 512;
  if (split) { g = g * (512 / kSplitThreads); nvls_sync_kernel<<<1, 32, 0, stream>>>(c, 0); }
  if (dtype == 0) nvls_allreduce_kernel<__nv_bfloat16><<<(int)g, threads, 0, stream>>>(c, (__nv_bfloat16*)mc_ptr, n, scale);
  else nvls_allreduce_kernel<float><<<(int)g, threads, 0, stream>>>(c, (float*)mc_ptr, n, scale);
  if (split) nvls_sync_kernel<<<1, 32, 0, stream>>>(c, 1);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// Broadcast `bytes` (multiple of 16) from `root` through the multicast buffer.  src: the root's payload (ignored on
// other ranks); mc_ptr / replica_ptr: multicast address and this rank's local replica of the SAME region;
// out: destination on non-root ranks, or NULL to leave the data in the replica.
B200_API int b200_comm_broadcast_nvls(const void* const* flag_ptrs, void* counter, int rank, int world, int channel,
                                      unsigned seq, const void* src, void* mc_ptr, const void* replica_ptr, void* out,
                                      long long bytes, int root, int max_ctas, cudaStream_t stream) {
  if (world < 2 || world > 8) { b200_set_error("broadcast_nvls: world %d out of range", world); return B200_ERR_LIMIT; }
  if (bytes & 15) { b200_set_error("broadcast_nvls: bytes must be a multiple of 16"); return B200_ERR_SHAPE; }
  McCtx c;
  memset(&c, 0, sizeof(c));
  for (int r = 0; r < world; ++r) c.flags[r] = reinterpret_cast<uint32_t*>(const_cast<void*>(flag_ptrs[r]));
  c.counter = reinterpret_cast<uint32_t*>(counter);
  c.rank = rank; c.world = world; c.channel = channel; c.seq = seq;
  const long long n16 = bytes >> 4;
  long long g = (n16 + 512 * 8 - 1) / (512 * 8);
  int cap = max_ctas > 0 ? max_ctas : b200_sm_count();
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  nvls_broadcast_kernel<<<(int)g, 512, 0, stream>>>(c, (const uint4*)src, (uint4*)mc_ptr, (const uint4*)replica_ptr,
                                                     (uint4*)out, n16, root);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

static int mc_fill_ctx(McCtx& c, const void* const* flag_ptrs, void* counter, int rank, int world, int channel, unsigned seq,
                       const char* what) {
  if (world < 2 || world > 8) { b200_set_error("%s: world %d out of range", what, world); return B200_ERR_LIMIT; }
  memset(&c, 0, sizeof(c));
  for (int r = 0; r < world; ++r) c.flags[r] = reinterpret_cast<uint32_t*>(const_cast<void*>(flag_ptrs[r]));
  c.counter = reinterpret_cast<uint32_t*>(counter);
  c.rank = rank; c.world = world; c.channel = channel; c.seq = seq;
  return B200_OK;
}
static int mc_grid(long long items, int per_thread, int max_ctas) {
  long long g = (items + 512LL * per_thread - 1) / (512LL * per_thread);
  const int cap = max_ctas > 0 ? max_ctas : b200_sm_count();
  if (g > cap) g = cap;
  return g < 1 ? 1 : (int)g;
}

// Reduce-scatter: every rank's input [world * n] (in_dtype) lies at mc_ptr inside the multicast buffer; out[n] (out_dtype)
// = scale * sum over ranks of slice `rank`.  n % 8 == 0, mc_ptr 16-byte aligned.
// Replaces dist.reduce_scatter_tensor (the collective FSDP issues for gradients).
B200_API int b200_comm_reducescatter_nvls(const void* const* flag_ptrs, void* counter, int rank, int world, int channel,
                                          unsigned seq, const void* mc_ptr, void* out, long long n, float scale,
                                          int in_dtype, int out_dtype, int max_ctas, int split, cudaStream_t stream) {
  McCtx c;
  if (int rc = mc_fill_ctx(c, flag_ptrs, counter, rank, world, channel, seq, "reducescatter_nvls")) return rc;
  c.split = split ? 1 : 0;
  if ((n & 7) || ((uintptr_t)mc_ptr & 15) || ((uintptr_t)out & 15)) {
    b200_set_error("reducescatter_nvls: n %% 8 == 0 and 16-byte aligned pointers required"); return B200_ERR_SHAPE;
  }
  const int g = mc_grid(n >> 3, in_dtype == 0 ? 8 : 4, max_ctas);
  const int threads = split ? kSplitThreads : 512, gs = split ? g * (512 / kSplitThreads) : g;
#define RSN(TI, TO) nvls_reducescatter_kernel<TI, TO><<<gs, threads, 0, stream>>>(c, (const TI*)mc_ptr, (TO*)out, n, scale)
  if ((in_dtype | out_dtype) & ~1) { b200_set_error("reducescatter_nvls: bad dtype"); return B200_ERR_UNSUPPORTED; }
  if (split) nvls_sync_kernel<<<1, 32, 0, stream>>>(c, 0);
  if (in_dtype == 0 && out_dtype == 0) RSN(__nv_bfloat16, __nv_bfloat16);
  else if (in_dtype == 0 && out_dtype == 1) RSN(__nv_bfloat16, float);
  else if (in_dtype == 1 && out_dtype == 1) RSN(float, float);
  else RSN(float, __nv_bfloat16);
#undef RSN
  if (split) nvls_sync_kernel<<<1, 32, 0, stream>>>(c, 1);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// Fused FSDP gradient step through the switch (see nvls_rs_adamw_kernel): gradients [world * n] of in_dtype at mc_ptr.
B200_API int b200_comm_rs_adamw_nvls(const void* const* flag_ptrs, void* counter, int rank, int world, int channel,
                                     unsigned seq, const void* mc_ptr, long long n, float grad_scale, int in_dtype,
                                     float* master, void* lowp, float* m, float* v, float lr, float beta1, float beta2,
                                     float eps, float wd, float bc1, float bc2, int max_ctas, int split,
                                     cudaStream_t stream) {
  McCtx c;
  if (int rc = mc_fill_ctx(c, flag_ptrs, counter, rank, world, channel, seq, "rs_adamw_nvls")) return rc;
  if ((n & 7) || ((uintptr_t)mc_ptr & 15)) { b200_set_error("rs_adamw_nvls: n %% 8 == 0 and a 16-byte aligned buffer required"); return B200_ERR_SHAPE; }
  if (in_dtype != 0 && in_dtype != 1) { b200_set_error("rs_adamw_nvls: bad dtype"); return B200_ERR_UNSUPPORTED; }
  c.split = split ? 1 : 0;
  const int g = mc_grid(n >> 3, 4, max_ctas);
  const float bc2s = sqrtf(bc2);
  const int threads = split ? kSplitThreads : 512, gs = split ? g * (512 / kSplitThreads) : g;
  if (split) nvls_sync_kernel<<<1, 32, 0, stream>>>(c, 0);
  if (in_dtype == 0)
    nvls_rs_adamw_kernel<__nv_bfloat16><<<gs, threads, 0, stream>>>(c, (const __nv_bfloat16*)mc_ptr, n, grad_scale, master,
                                                                (__nv_bfloat16*)lowp, m, v, lr, beta1, beta2, eps, wd, bc1, bc2s);
  else
    nvls_rs_adamw_kernel<float><<<gs, threads, 0, stream>>>(c, (const float*)mc_ptr, n, grad_scale, master, (__nv_bfloat16*)lowp,
                                                        m, v, lr, beta1, beta2, eps, wd, bc1, bc2s);
  if (split) nvls_sync_kernel<<<1, 32, 0, stream>>>(c, 1);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// All-gather: src = this rank's shard (`bytes`, multiple of 16, any local pointer); mc_ptr / replica_ptr = multicast
// address and local replica of a [world * bytes] region; out = destination or NULL (use the replica in place).
// Replaces dist.all_gather_into_tensor (FSDP parameter gathers, SP activation gathers).
B200_API int b200_comm_allgather_nvls(const void* const* flag_ptrs, void* counter, int rank, int world, int channel,
                                      unsigned seq, const void* src, void* mc_ptr, const void* replica_ptr, void* out,
                                      long long bytes, int max_ctas, cudaStream_t stream) {
  McCtx c;
  if (int rc = mc_fill_ctx(c, flag_ptrs, counter, rank, world, channel, seq, "allgather_nvls")) return rc;
  if ((bytes & 15) || ((uintptr_t)src & 15) || ((uintptr_t)out & 15) || ((uintptr_t)mc_ptr & 15)) {
    b200_set_error("allgather_nvls: bytes %% 16 == 0 and 16-byte aligned pointers required"); return B200_ERR_SHAPE;
  }
  const long long n16 = bytes >> 4;
  const int g = mc_grid(n16, 8, max_ctas);
  nvls_allgather_kernel<<<g, 512, 0, stream>>>(c, (const uint4*)src, (uint4*)mc_ptr, n16);
  B200_LAUNCH_CHECK();
  if (out != nullptr)
    B200_CUDA_CHECK(cudaMemcpyAsync(out, replica_ptr, (size_t)bytes * world, cudaMemcpyDeviceToDevice, stream));
  return B200_OK;
}
