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

B200_API int b200_version() { return 100; }

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
    estr[i] = 1;
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
