// Host-side runtime shared by all kernels: error reporting, device query and
// TMA tensor-map encoding through the driver entry point (so the library has
// no link-time dependency on libcuda and can be dlopen()ed on a CPU-only box).
// No counterpart in the reference (it has no native code).
#include "common.cuh"

#include <cudaTypedefs.h>
#include <stdarg.h>
#include <string.h>

static thread_local char g_err[1024] = "";

void b200_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

B200_API const char* b200_last_error() { return g_err; }

B200_API int b200_version() { return 101; }

// ---- cross-rank wait limit (see PeerWatchdog in common.cuh) ----
int b200_peer_wd_set_comm(const b200::PeerWatchdog*);
int b200_peer_wd_set_mc(const b200::PeerWatchdog*);
int b200_peer_wd_set_gemm(const b200::PeerWatchdog*);
static unsigned int* g_wd_host_flag = nullptr;   // mapped pinned word shared by every device of the process

// Applies to the CURRENT device (settings live in per-device __device__ variables): call once per device used.
// seconds <= 0 disables the limit; trap != 0 aborts the kernel (and with it the CUDA context) on expiry, trap == 0
// only reports - device printf plus the counter b200_peer_timeout_count() reads - and keeps waiting.
B200_API int b200_set_peer_timeout(double seconds, int trap) {
  if (g_wd_host_flag == nullptr) {
    void* h = nullptr;
    if (cudaHostAlloc(&h, sizeof(unsigned int), cudaHostAllocMapped | cudaHostAllocPortable) == cudaSuccess) {
      g_wd_host_flag = reinterpret_cast<unsigned int*>(h);
      *g_wd_host_flag = 0;
    } else {
      (void)cudaGetLastError();   // no pinned memory: the report is the device printf only
    }
  }
  b200::PeerWatchdog wd;
  wd.timeout_ns = seconds > 0 ? (unsigned long long)(seconds * 1e9) : 0ull;
  wd.trap = trap ? 1 : 0;
  wd.host_flag = nullptr;
  if (g_wd_host_flag != nullptr) {
    void* d = nullptr;
    if (cudaHostGetDevicePointer(&d, g_wd_host_flag, 0) == cudaSuccess) wd.host_flag = reinterpret_cast<unsigned int*>(d);
    else (void)cudaGetLastError();
  }
  if (b200_peer_wd_set_comm(&wd) || b200_peer_wd_set_mc(&wd) || b200_peer_wd_set_gemm(&wd)) {
    b200_set_error("set_peer_timeout: cudaMemcpyToSymbol failed: %s", cudaGetErrorString(cudaGetLastError()));
    return B200_ERR_CUDA;
  }
  return B200_OK;
}
// Number of expired cross-rank waits reported so far in this process (readable while a kernel is still spinning).
B200_API int b200_peer_timeout_count() {
  return g_wd_host_flag ? (int)*reinterpret_cast<volatile unsigned int*>(g_wd_host_flag) : 0;
}

B200_API int b200_device_info(int* sm_count, int* cc_major, int* cc_minor, size_t* smem_optin) {
  int dev = 0;
  B200_CUDA_CHECK(cudaGetDevice(&dev));
  cudaDeviceProp p;
  B200_CUDA_CHECK(cudaGetDeviceProperties(&p, dev));
  if (sm_count) *sm_count = p.multiProcessorCount;
  if (cc_major) *cc_major = p.major;
  if (cc_minor) *cc_minor = p.minor;
  if (smem_optin) *smem_optin = p.sharedMemPerBlockOptin;
  return B200_OK;
}

int b200_sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return b200::kNumSMs;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = b200::kNumSMs;
    cached[dev] = n;
  }
  return cached[dev];
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess || !p)
    return nullptr;
  fn = reinterpret_cast<EncodeTiledFn>(p);
  return fn;
}

int b200_make_tmap(CUtensorMap* out, const void* base, int elem_bytes, int rank,
                   const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box,
                   int swizzle_128b) {
  return b200_make_tmap_strided(out, base, elem_bytes, rank, dims, strides_bytes, box, nullptr, swizzle_128b);
}

// elem_strides (may be null = all 1): traversal stride per dimension.  With a stride s > 1 a box of extent b covers b
// tensor elements and delivers ceil(b / s) of them - every s-th one - packed in shared memory: a strided convolution reads
// exactly its taps through it.
int b200_make_tmap_strided(CUtensorMap* out, const void* base, int elem_bytes, int rank,
                           const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box,
                           const uint32_t* elem_strides, int swizzle_128b) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    b200_set_error("cuTensorMapEncodeTiled driver entry point unavailable");
    return B200_ERR_CUDA;
  }
  cuuint64_t gdim[5];
  cuuint64_t gstr[5];
  cuuint32_t gbox[5];
  cuuint32_t estr[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    gbox[i] = box[i];
    estr[i] = elem_strides ? elem_strides[i] : 1;
  }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_bytes[i];
  CUtensorMapDataType dt = elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                                           : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  CUresult r = fn(out, dt, static_cast<cuuint32_t>(rank), const_cast<void*>(base), gdim, gstr, gbox,
                  estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  swizzle_128b ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    b200_set_error(
        "cuTensorMapEncodeTiled failed (%d): base=%p rank=%d dims=[%llu,%llu,%llu] "
        "stride1=%llu box=[%u,%u,%u]",
        (int)r, base, rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
        (unsigned long long)(rank > 2 ? dims[2] : 0),
        (unsigned long long)(rank > 1 ? strides_bytes[0] : 0), box[0], rank > 1 ? box[1] : 0,
        rank > 2 ? box[2] : 0);
    return B200_ERR_SHAPE;
  }
  return B200_OK;
}
