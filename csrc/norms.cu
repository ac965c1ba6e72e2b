// RMSNorm and LayerNorm, forward and backward, one CTA per row with the row held
// in registers (single pass over HBM).  Weight/bias gradients are reduced across
// rows in two steps: per-CTA partial sums into a caller-provided fp32 scratch
// [num_partials, D], then a column reduction.
//
// Reference semantics: RMSNorm  /root/reference/fsdp_tp/llama2_model.py:115-142
//   (fp32 upcast, x * rsqrt(mean(x^2) + eps), cast back, * weight);
// LayerNorm: nn.LayerNorm in /root/reference/scripts/03_tensor_parallel_tp/tensor_parallel_vit.py:141-143
//   and /root/reference/scripts/04_pipeline_parallel_pp/03_pipeline_training.py:56,59.
#include "vec.cuh"

namespace b200 {

constexpr int kMaxChunks = 4;  // 8-element chunks cached per thread

// ------------------------------- RMSNorm -----------------------------------
// Persistent CTAs (all resident: grid <= 6 per SM), one row at a time, the next row's loads in flight (unconverted, 16
// registers) while the current row is reduced, scaled and stored.
template <typename T>
__global__ void __launch_bounds__(256) rmsnorm_fwd_kernel(const T* __restrict__ x, const T* __restrict__ w, T* __restrict__ y,
                                                          float* __restrict__ rstd_out, int rows, int D, float eps) {
  __shared__ float red[33];
  const int nchunks = D >> 3;
  Raw8<T> nx[kMaxChunks];
  auto fetch = [&](int row) {
#pragma unroll
    for (int i = 0; i < kMaxChunks; ++i) {
      const int c = threadIdx.x + i * blockDim.x;
      if (c < nchunks) nx[i].load(x + (long long)row * D + c * 8);
    }
  };
  int row = blockIdx.x;
  if (row < rows) fetch(row);
  for (; row < rows; row += gridDim.x) {
    float v[kMaxChunks][8];
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < kMaxChunks; ++i) {
      const int c = threadIdx.x + i * blockDim.x;
      if (c < nchunks) {
        nx[i].unpack(v[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) ss += v[i][j] * v[i][j];
      }
    }
    if (row + (int)gridDim.x < rows) fetch(row + (int)gridDim.x);
    ss = block_sum(ss, red);
    const float rstd = rsqrtf(ss / (float)D + eps);
    if (threadIdx.x == 0 && rstd_out) rstd_out[row] = rstd;
    T* yr = y + (long long)row * D;
#pragma unroll
    for (int i = 0; i < kMaxChunks; ++i) {
      const int c = threadIdx.x + i * blockDim.x;
      if (c < nchunks) {
        float wv[8], o[8];
        Vec8<T>::load(w + c * 8, wv);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = to_f<T>(from_f<T>(v[i][j] * rstd)) * wv[j];
        Vec8<T>::store(yr + c * 8, o);
      }
    }
  }
}

// Persistent CTAs, one row at a time, software pipelined: the NEXT row's x / dy (/ residual gradient) are requested - and
// parked unconverted in registers - before the current row's block reduction, so the HBM round trip of row r + 1 overlaps the
// reduction, the arithmetic and the stores of row r (the unpipelined version was latency bound: 2.5 TB/s).  The weight is
// re-read per row (8-16 KB, L1 resident) instead of being held in registers.  CH = 8-element chunks per thread: rows of up to
// 2048 elements use 2, up to 4096 use 1 (512 threads for D = 4096: 32 warps per SM at two CTAs instead of 8 - with 4 chunks
// per thread the kernel was bound by instruction latency at two warps per scheduler, not by memory: 80.7 -> 64.4 us at CH = 2).
template <typename T, int CH>
__global__ void __launch_bounds__(CH == 1 ? 512 : 256) rmsnorm_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x,
                                   const T* __restrict__ w, const float* __restrict__ rstd_in,
                                   T* __restrict__ dx, float* __restrict__ dw_partial, int rows, int D,
                                   const T* addsrc) {   // addsrc: optional [rows, D] added to dx (may alias dx)
  __shared__ float red[33];
  const int nchunks = D >> 3;
  float dwacc[CH][8];
#pragma unroll
  for (int i = 0; i < CH; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) dwacc[i][j] = 0.f;
  Raw8<T> nx[CH], ng[CH], na[CH];
  auto fetch = [&](int row) {
#pragma unroll
    for (int i = 0; i < CH; ++i) {
      const int c = threadIdx.x + i * blockDim.x;
      if (c < nchunks) {
        const long long o = (long long)row * D + c * 8;
        nx[i].load(x + o);
        ng[i].load(dy + o);
        if (addsrc != nullptr) na[i].load(addsrc + o);   // rows are disjoint: no hazard with the stores of earlier rows
      }
    }
  };
  int row = blockIdx.x;
  if (row < rows) fetch(row);
  for (; row < rows; row += gridDim.x) {
    const float rstd = rstd_in[row];
    float xv[CH][8], gv[CH][8];
    Raw8<T> ca[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) {
      nx[i].unpack(xv[i]);
      ng[i].unpack(gv[i]);
      ca[i] = na[i];
    }
    if (row + (int)gridDim.x < rows) fetch(row + (int)gridDim.x);
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < CH; ++i) {
      const int c = threadIdx.x + i * blockDim.x;
      if (c < nchunks) {
        float wv[8];
        Vec8<T>::load(w + c * 8, wv);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          dwacc[i][j] += gv[i][j] * xv[i][j] * rstd;
          gv[i][j] *= wv[j];
          dot += gv[i][j] * xv[i][j];
        }
      }
    }
    dot = block_sum(dot, red);
    const float coef = dot * rstd * rstd * rstd / (float)D;
    T* dxr = dx + (long long)row * D;
#pragma unroll
    for (int i = 0; i < CH; ++i) {
      const int c = threadIdx.x + i * blockDim.x;
      if (c < nchunks) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = gv[i][j] * rstd - xv[i][j] * coef;
        if (addsrc != nullptr) {
          float prev[8];
          ca[i].unpack(prev);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += prev[j];
        }
        Vec8<T>::store(dxr + c * 8, o);
      }
    }
  }
  float* part = dw_partial + (long long)blockIdx.x * D;
#pragma unroll
  for (int i = 0; i < CH; ++i) {
    const int c = threadIdx.x + i * blockDim.x;
    if (c < nchunks) {
      *reinterpret_cast<float4*>(part + c * 8) = make_float4(dwacc[i][0], dwacc[i][1], dwacc[i][2], dwacc[i][3]);
      *reinterpret_cast<float4*>(part + c * 8 + 4) = make_float4(dwacc[i][4], dwacc[i][5], dwacc[i][6], dwacc[i][7]);
    }
  }
}

// out[c] (+)= sum_p partial[p][c].  One CTA owns 64 columns: 16 float4 column groups x 16 partial-row lanes, each lane
// strides over the partial rows with several independent loads in flight, then the 16 lanes meet in shared memory
// (a single thread walking all ~450 partial rows was a chain of dependent L2 round trips: 15-70 us per call).
constexpr int kColsumCols = 64;
template <typename T>
__global__ void __launch_bounds__(256) colsum_partials_kernel(const float* __restrict__ partial, T* __restrict__ out,
                                                              int nparts, int D, int accumulate) {
  __shared__ float4 red[16][17];
  const int cg = threadIdx.x & 15, pl = threadIdx.x >> 4;
  const int c = (blockIdx.x * 16 + cg) * 4;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  if (c < D) {
#pragma unroll 4
    for (int p = pl; p < nparts; p += 16) {
      const float4 v = *reinterpret_cast<const float4*>(partial + (long long)p * D + c);
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
  }
  red[pl][cg] = s;
  __syncthreads();
  if (pl == 0 && c < D) {
#pragma unroll
    for (int i = 1; i < 16; ++i) { const float4 v = red[i][cg]; s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w; }
    float r[4] = {s.x, s.y, s.z, s.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (accumulate) r[j] += to_f<T>(out[c + j]);
      out[c + j] = from_f<T>(r[j]);
    }
  }
}

// ------------------------------- LayerNorm ---------------------------------
template <typename T>
__global__ void layernorm_fwd_kernel(const T* __restrict__ x, const T* __restrict__ w,
                                     const T* __restrict__ b, T* __restrict__ y,
                                     float* __restrict__ mean_out, float* __restrict__ rstd_out,
                                     int rows, int D, float eps) {
  __shared__ float red[33];
  const int nchunks = D >> 3;
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const T* xr = x + (long long)row * D;
    float v[kMaxChunks][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < kMaxChunks; ++i) {
      const int c = threadIdx.x + i * blockDim.x;
      if (c < nchunks) {
        Vec8<T>::load(xr + c * 8, v[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) s += v[i][j];
      }
    }
    const float mean = block_sum(s, red) / (float)D;
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < kMaxChunks; ++i) {
      const int c = threadIdx.x + i * blockDim.x;
      if (c < nchunks) {
#pragma unroll
        for (int j = 0; j < 8; ++j) { const float d = v[i][j] - mean; ss += d * d; }
      }
    }
    const float rstd = rsqrtf(block_sum(ss, red) / (float)D + eps);
    if (threadIdx.x == 0) {
      if (mean_out) mean_out[row] = mean;
      if (rstd_out) rstd_out[row] = rstd;
    }
    T* yr = y + (long long)row * D;
#pragma unroll
    for (int i = 0; i < kMaxChunks; ++i) {
      const int c = threadIdx.x + i * blockDim.x;
      if (c < nchunks) {
        float wv[8], bv[8], o[8];
        Vec8<T>::load(w + c * 8, wv);
        if (b) Vec8<T>::load(b + c * 8, bv);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (v[i][j] - mean) * rstd * wv[j] + (b ? bv[j] : 0.f);
        Vec8<T>::store(yr + c * 8, o);
      }
    }
  }
}

template <typename T>
__global__ void layernorm_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x,
                                     const T* __restrict__ w, const float* __restrict__ mean_in,
                                     const float* __restrict__ rstd_in, T* __restrict__ dx,
                                     float* __restrict__ dw_partial, float* __restrict__ db_partial,
                                     int rows, int D) {
  __shared__ float red[33];
  const int nchunks = D >> 3;
  float dwacc[kMaxChunks][8], dbacc[kMaxChunks][8], wv[kMaxChunks][8];
#pragma unroll
  for (int i = 0; i < kMaxChunks; ++i) {
    const int c = threadIdx.x + i * blockDim.x;
#pragma unroll
    for (int j = 0; j < 8; ++j) { dwacc[i][j] = 0.f; dbacc[i][j] = 0.f; }
    if (c < nchunks) Vec8<T>::load(w + c * 8, wv[i]);
  }
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const T* xr = x + (long long)row * D;
    const T* gr = dy + (long long)row * D;
    const float mean = mean_in[row], rstd = rstd_in[row];
    float xh[kMaxChunks][8], gv[kMaxChunks][8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < kMaxChunks; ++i) {
      const int c = threadIdx.x + i * blockDim.x;
      if (c < nchunks) {
        Vec8<T>::load(xr + c * 8, xh[i]);
        Vec8<T>::load(gr + c * 8, gv[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[i][j] = (xh[i][j] - mean) * rstd;
          dwacc[i][j] += gv[i][j] * xh[i][j];
          dbacc[i][j] += gv[i][j];
          gv[i][j] *= wv[i][j];
          s1 += gv[i][j];
          s2 += gv[i][j] * xh[i][j];
        }
      }
    }
    s1 = block_sum(s1, red) / (float)D;
    s2 = block_sum(s2, red) / (float)D;
    T* dxr = dx + (long long)row * D;
#pragma unroll
    for (int i = 0; i < kMaxChunks; ++i) {
      const int c = threadIdx.x + i * blockDim.x;
      if (c < nchunks) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rstd * (gv[i][j] - s1 - xh[i][j] * s2);
        Vec8<T>::store(dxr + c * 8, o);
      }
    }
  }
  float* pw = dw_partial + (long long)blockIdx.x * D;
  float* pb = db_partial + (long long)blockIdx.x * D;
#pragma unroll
  for (int i = 0; i < kMaxChunks; ++i) {
    const int c = threadIdx.x + i * blockDim.x;
    if (c < nchunks) {
#pragma unroll
      for (int j = 0; j < 8; ++j) { pw[c * 8 + j] = dwacc[i][j]; pb[c * 8 + j] = dbacc[i][j]; }
    }
  }
}

static int norm_threads(int D, int* threads) {
  const int nchunks = D >> 3;
  int t = 32;
  while (t * kMaxChunks < nchunks && t < 256) t <<= 1;  // bwd kernels use ~200 regs: 256 threads max
  if (t * kMaxChunks < nchunks || (D & 7)) {
    b200_set_error("norm kernels need D %% 8 == 0 and D <= %d (got %d)", 256 * kMaxChunks * 8, D);
    return B200_ERR_SHAPE;
  }
  if (t < 64) t = 64;
  *threads = t;
  return B200_OK;
}

}  // namespace b200

using namespace b200;

// Number of per-CTA partial rows the backward kernels write into `scratch`
// ([b200_norm_bwd_partials(rows), D] fp32 for RMSNorm, twice that for LayerNorm).
B200_API int b200_norm_bwd_partials(int rows) {
  int n = b200_sm_count() * 3;   // = the CTAs resident at once (3 per SM at ~155 registers x 128-256 threads)
  return rows < n ? (rows < 1 ? 1 : rows) : n;
}

// The pipelined RMSNorm backward keeps two rows per thread in registers: whichever chunk count is chosen below, two CTAs
// per SM are resident (512 threads x 56 registers, 256 x 97, or 128-256 x 184), so a third per SM would only form a
// half-empty second wave.  The scratch is sized by b200_norm_bwd_partials (>= this).
static int rmsnorm_bwd_cfg(int D, int* threads, int* ch) {
  if (int rc = norm_threads(D, threads)) return rc;
  *ch = kMaxChunks;
  const int nchunks = D >> 3;
  if (nchunks > 256 && nchunks <= 512) {  // one chunk per thread on up to 512 threads (D = 4096: 32 warps per SM at two CTAs)
    *threads = (nchunks + 31) / 32 * 32;
    *ch = 1;
  } else if (nchunks > 64 && nchunks <= 256) {   // two chunks per thread
    int t = ((nchunks + 1) / 2 + 31) / 32 * 32;
    *threads = t < 64 ? 64 : t;
    *ch = 2;
  }
  return B200_OK;
}
static int rmsnorm_bwd_grid(int rows) {
  const int n = b200_norm_bwd_partials(rows), cap = 2 * b200_sm_count();
  return n < cap ? n : cap;
}

B200_API int b200_rmsnorm_fwd(const void* x, const void* w, void* y, float* rstd, int rows, int D,
                              float eps, int dtype, cudaStream_t stream) {
  if (rows <= 0) return B200_OK;
  int threads;
  if (int rc = norm_threads(D, &threads)) return rc;
  const int cap = 6 * b200_sm_count();   // all resident at once (<= 80 registers x <= 256 threads): one wave, no tail
  const int grid = rows < cap ? rows : cap;
  B200_DISPATCH_DTYPE(dtype, T, (rmsnorm_fwd_kernel<T><<<grid, threads, 0, stream>>>(
      (const T*)x, (const T*)w, (T*)y, rstd, rows, D, eps)));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// dx = d/dx, dw (+)= d/dw.  scratch: fp32 [b200_norm_bwd_partials(rows), D].
B200_API int b200_rmsnorm_bwd(const void* dy, const void* x, const void* w, const float* rstd,
                              void* dx, void* dw, float* scratch, int rows, int D, int add_to_dx,
                              int accumulate_dw, int dtype, cudaStream_t stream) {
  if (rows <= 0) return B200_OK;
  int threads, ch;
  if (int rc = rmsnorm_bwd_cfg(D, &threads, &ch)) return rc;
  const int grid = rmsnorm_bwd_grid(rows);
  B200_DISPATCH_DTYPE(dtype, T, {
    if (ch == 1)
      rmsnorm_bwd_kernel<T, 1><<<grid, threads, 0, stream>>>((const T*)dy, (const T*)x, (const T*)w, rstd,
                                                             (T*)dx, scratch, rows, D, add_to_dx ? (const T*)dx : (const T*)nullptr);
    else if (ch == 2)
      rmsnorm_bwd_kernel<T, 2><<<grid, threads, 0, stream>>>((const T*)dy, (const T*)x, (const T*)w, rstd,
                                                             (T*)dx, scratch, rows, D, add_to_dx ? (const T*)dx : (const T*)nullptr);
    else
      rmsnorm_bwd_kernel<T, kMaxChunks><<<grid, threads, 0, stream>>>((const T*)dy, (const T*)x, (const T*)w, rstd,
                                                        (T*)dx, scratch, rows, D, add_to_dx ? (const T*)dx : (const T*)nullptr);
    colsum_partials_kernel<T><<<ceil_div(D, kColsumCols), 256, 0, stream>>>(scratch, (T*)dw, grid, D,
                                                                    accumulate_dw);
  });
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// Same, with the residual branch of a pre-norm block folded in: dx = dres + d/dx (dres: [rows, D], may alias dx).
// Saves the separate elementwise add autograd would otherwise run for x -> {norm, residual}.
B200_API int b200_rmsnorm_bwd_res(const void* dy, const void* x, const void* w, const float* rstd, const void* dres,
                                  void* dx, void* dw, float* scratch, int rows, int D, int accumulate_dw, int dtype,
                                  cudaStream_t stream) {
  if (rows <= 0) return B200_OK;
  int threads, ch;
  if (int rc = rmsnorm_bwd_cfg(D, &threads, &ch)) return rc;
  const int grid = rmsnorm_bwd_grid(rows);
  B200_DISPATCH_DTYPE(dtype, T, {
    if (ch == 1)
      rmsnorm_bwd_kernel<T, 1><<<grid, threads, 0, stream>>>((const T*)dy, (const T*)x, (const T*)w, rstd,
                                                             (T*)dx, scratch, rows, D, (const T*)dres);
    else if (ch == 2)
      rmsnorm_bwd_kernel<T, 2><<<grid, threads, 0, stream>>>((const T*)dy, (const T*)x, (const T*)w, rstd,
                                                             (T*)dx, scratch, rows, D, (const T*)dres);
    else
      rmsnorm_bwd_kernel<T, kMaxChunks><<<grid, threads, 0, stream>>>((const T*)dy, (const T*)x, (const T*)w, rstd,
                                                        (T*)dx, scratch, rows, D, (const T*)dres);
    colsum_partials_kernel<T><<<ceil_div(D, kColsumCols), 256, 0, stream>>>(scratch, (T*)dw, grid, D, accumulate_dw);
  });
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_layernorm_fwd(const void* x, const void* w, const void* b, void* y, float* mean,
                                float* rstd, int rows, int D, float eps, int dtype,
                                cudaStream_t stream) {
  if (rows <= 0) return B200_OK;
  int threads;
  if (int rc = norm_threads(D, &threads)) return rc;
  const int grid = rows;   // one CTA per row: a capped persistent grid ran two unequal waves (7 CTAs/SM resident, 8/SM launched)
  B200_DISPATCH_DTYPE(dtype, T, (layernorm_fwd_kernel<T><<<grid, threads, 0, stream>>>(
      (const T*)x, (const T*)w, (const T*)b, (T*)y, mean, rstd, rows, D, eps)));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// scratch: fp32 [2 * b200_norm_bwd_partials(rows), D]
B200_API int b200_layernorm_bwd(const void* dy, const void* x, const void* w, const float* mean,
                                const float* rstd, void* dx, void* dw, void* db, float* scratch,
                                int rows, int D, int accumulate, int dtype, cudaStream_t stream) {
  if (rows <= 0) return B200_OK;
  int threads;
  if (int rc = norm_threads(D, &threads)) return rc;
  const int grid = b200_norm_bwd_partials(rows);
  float* pw = scratch;
  float* pb = scratch + (long long)grid * D;
  B200_DISPATCH_DTYPE(dtype, T, {
    layernorm_bwd_kernel<T><<<grid, threads, 0, stream>>>((const T*)dy, (const T*)x, (const T*)w,
                                                          mean, rstd, (T*)dx, pw, pb, rows, D);
    colsum_partials_kernel<T><<<ceil_div(D, kColsumCols), 256, 0, stream>>>(pw, (T*)dw, grid, D, accumulate);
    if (db)
      colsum_partials_kernel<T><<<ceil_div(D, kColsumCols), 256, 0, stream>>>(pb, (T*)db, grid, D, accumulate);
  });
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// ---------------------------------------------------------------------------------------------
// b200_rmsnorm_bwd_noalloc — the same backward as b200_rmsnorm_bwd, for hosts that have no memory
// to hand us and must not create any.
//
// Who needs this: the regular entry point asks the caller for an fp32 scratch of
// [b200_norm_bwd_partials(rows), D] floats.  A PyTorch caller takes that from the caching allocator,
// but some callers of this C ABI cannot do the equivalent: a C/C++ or Fortran host that links the
// library without any allocator of its own; a job whose whole HBM budget was carved up in advance
// (every byte of the 180 GB already belongs to a pool, so a late cudaMalloc may simply fail); and
// latency-sensitive loops where cudaMalloc / cudaFree are off limits because they serialise with
// all other work on the device.  For them the scratch lives inside the library image.
//
// How: a file-scope `static __device__` array below holds the per-CTA partial sums.  Its name is only a
// host-side handle, so before the first launch on each device (with that device current) we ask the
// runtime for its device address with cudaGetSymbolAddress, check the result and cache it; kernels get
// that pointer.  No cudaMalloc / cudaFree / torch call is made here, ever.
//
// Limits and rules for the caller:
//   * D <= 8192 and D % 8 == 0; larger rows are rejected with B200_ERR_LIMIT (text in b200_last_error()).
//     At most 148 partial rows are used, so the array is 148 * 8192 floats = 4.6 MiB, held by every GPU
//     that loads the library until the process exits.
//   * The first call on a device performs the address look-up described above; make that first call
//     before starting any stream capture on that device.
//   * All calls on a device share the one array and it keeps whatever the previous call left in it.
//     Nothing in it has to start at zero: each call overwrites every partial row it later reads, and
//     reads only those rows, so stale contents are never observed.
//   * Because the array is shared, two calls that may run at the same time on one GPU (different
//     streams, or different host threads) race on it.  Order them on one stream or with events.
namespace {
constexpr int kNoAllocMaxD = 8192;
constexpr int kNoAllocMaxPartials = 148;
constexpr int kNoAllocMaxDevices = 64;
}  // namespace

static __device__ __align__(256) float g_rmsnorm_bwd_scratch[(size_t)kNoAllocMaxPartials * kNoAllocMaxD];
static float* g_rmsnorm_bwd_scratch_addr[kNoAllocMaxDevices] = {nullptr};

B200_API int b200_rmsnorm_bwd_noalloc(const void* dy, const void* x, const void* w, const float* rstd,
                                      void* dx, void* dw, int rows, int D, int add_to_dx,
                                      int accumulate_dw, int dtype, cudaStream_t stream) {
  if (rows <= 0) return B200_OK;
  if (D > kNoAllocMaxD) {
    b200_set_error("rmsnorm_bwd_noalloc: D = %d exceeds the built-in scratch limit of %d columns "
                   "(use b200_rmsnorm_bwd with caller-provided scratch)", D, kNoAllocMaxD);
    return B200_ERR_LIMIT;
  }
  int threads, ch;
  if (int rc = rmsnorm_bwd_cfg(D, &threads, &ch)) return rc;
  int dev = -1;
  B200_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoAllocMaxDevices) {
    b200_set_error("rmsnorm_bwd_noalloc: device ordinal %d out of range", dev);
    return B200_ERR_LIMIT;
  }
  float* scratch = g_rmsnorm_bwd_scratch_addr[dev];
  if (scratch == nullptr) {  // once per device, with that device current
    void* p = nullptr;
    B200_CUDA_CHECK(cudaGetSymbolAddress(&p, g_rmsnorm_bwd_scratch));
    scratch = static_cast<float*>(p);
    g_rmsnorm_bwd_scratch_addr[dev] = scratch;
  }
  const int grid = rows < kNoAllocMaxPartials ? rows : kNoAllocMaxPartials;  // grid * D <= array size
  B200_DISPATCH_DTYPE(dtype, T, {
    if (ch == 1)
      rmsnorm_bwd_kernel<T, 1><<<grid, threads, 0, stream>>>((const T*)dy, (const T*)x, (const T*)w, rstd,
                                                             (T*)dx, scratch, rows, D, add_to_dx ? (const T*)dx : (const T*)nullptr);
    else if (ch == 2)
      rmsnorm_bwd_kernel<T, 2><<<grid, threads, 0, stream>>>((const T*)dy, (const T*)x, (const T*)w, rstd,
                                                             (T*)dx, scratch, rows, D, add_to_dx ? (const T*)dx : (const T*)nullptr);
    else
      rmsnorm_bwd_kernel<T, kMaxChunks><<<grid, threads, 0, stream>>>((const T*)dy, (const T*)x, (const T*)w, rstd,
                                                        (T*)dx, scratch, rows, D, add_to_dx ? (const T*)dx : (const T*)nullptr);
    colsum_partials_kernel<T><<<ceil_div(D, kColsumCols), 256, 0, stream>>>(scratch, (T*)dw, grid, D,
                                                                    accumulate_dw);
  });
  B200_LAUNCH_CHECK();
  return B200_OK;
}
