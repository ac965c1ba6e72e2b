// Convolution-network ops in NHWC (channels-last) layout: im2col / col2im around the tcgen05 GEMM,
// BatchNorm (batch statistics, fused scale/shift + residual + ReLU), max / average pooling,
// 2x2 transposed-conv pixel shuffle, bilinear resize and the latitude-weighted MSE.
//
// A convolution is  im2col(X)[N*Ho*Wo, KH*KW*C] x W[K, KH*KW*C]^T  on the tensor cores; 1x1 stride-1
// convolutions (two thirds of ResNet-50) skip im2col entirely because NHWC activations already
// ARE the GEMM operand.  dgrad = GEMM + col2im (gather form, no atomics); wgrad = GEMM with both
// operands MN-major.
//
// Replaces cuDNN conv / BN / pooling reached through torchvision ResNets (/root/reference/scripts/main.py:249),
// the UNet blocks (scripts/01_data_parallel_ddp/multinode_ddp_unet.py:171-214: Conv3x3-BN-ReLU,
// MaxPool2d(2), ConvTranspose2d(2,2), bilinear interpolate) and latitude_weighted_mse (:221-229).
#include "vec.cuh"

namespace b200 {

// ------------------------------------------------------------------ im2col / col2im
// col[(n, ho, wo), (kh, kw, c)] = x[n, ho*s - p + kh, wo*s - p + kw, c]  (zero outside); c padded to Cp
// A warp owns one output pixel at a time and sweeps that pixel's row of the patch matrix 32 vectors (512 bytes) per
// step, so the (n, ho, wo) decode is done once per row and the per-vector work is two 32-bit divisions, one 128-bit
// load and one fully coalesced 128-bit store (the first version decoded every vector with six 64-bit div/mods and was
// instruction-bound at ~1.4 TB/s).
template <typename T>
__global__ void __launch_bounds__(256) im2col_kernel(const T* __restrict__ x, T* __restrict__ col, int N, int H, int W,
                                                     int C, int Cp, int KH, int KW, int stride, int pad, int Ho, int Wo) {
  const unsigned c8n = Cp >> 3;
  const unsigned L = (unsigned)(KH * KW) * c8n;           // vectors per patch-matrix row
  const long long pixels = (long long)N * Ho * Wo;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const bool vec_ok = (C & 7) == 0;
  for (long long p = (long long)blockIdx.x * 8 + warp; p < pixels; p += (long long)gridDim.x * 8) {
    const unsigned pw = (unsigned)(p % ((long long)Ho * Wo));
    const int n = (int)(p / ((long long)Ho * Wo));
    const int ho = pw / Wo, wo = pw % Wo;
    const int h0 = ho * stride - pad, w0 = wo * stride - pad;
    const T* img = x + (long long)n * H * W * C;
    T* dst = col + p * (long long)L * 8;
    for (unsigned q = lane; q < L; q += 32) {
      const unsigned tap = q / c8n, c8 = q - tap * c8n;
      const int kh = tap / KW, kw = tap - kh * KW;
      const int h = h0 + kh, w = w0 + kw;
      float v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = 0.f;
      if (h >= 0 && h < H && w >= 0 && w < W) {
        const T* src = img + ((long long)h * W + w) * C + c8 * 8;
        if (vec_ok) Vec8<T>::load(src, v);
        else
          for (int j = 0; j < 8; ++j) if ((int)(c8 * 8 + j) < C) v[j] = to_f<T>(src[j]);
      }
      Vec8<T>::store(dst + (long long)q * 8, v);
    }
  }
}

// dx[n, h, w, c] = sum over (kh, kw) of dcol[(n, ho, wo), (kh, kw, c)] with ho*s - p + kh == h
template <typename T>
__global__ void col2im_kernel(const T* __restrict__ dcol, T* __restrict__ dx, int N, int H, int W, int C, int Cp,
                              int KH, int KW, int stride, int pad, int Ho, int Wo) {
  const int c8n = Cp >> 3;
  const long long total = (long long)N * H * W * c8n;
  const long long row_len = (long long)KH * KW * Cp;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    long long r = i;
    const int c8 = (int)(r % c8n); r /= c8n;
    const int w = (int)(r % W); r /= W;
    const int h = (int)(r % H);
    const int n = (int)(r / H);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    if (KH == 3 && KW == 3 && stride == 1) {
      // the common case (no divisions): one filter row = three independent 128-bit loads in flight at a time
#pragma unroll
      for (int kh = 0; kh < 3; ++kh) {
        const int ho = h + pad - kh;
        if (ho < 0 || ho >= Ho) continue;
        float v[3][8];
        bool ok[3];
#pragma unroll
        for (int kw = 0; kw < 3; ++kw) {
          const int wo = w + pad - kw;
          ok[kw] = wo >= 0 && wo < Wo;
          if (ok[kw])
            Vec8<T>::load(dcol + (((long long)n * Ho + ho) * Wo + wo) * row_len + (long long)(kh * 3 + kw) * Cp + c8 * 8, v[kw]);
        }
#pragma unroll
        for (int kw = 0; kw < 3; ++kw)
          if (ok[kw]) {
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] += v[kw][j];
          }
      }
    } else
    for (int kh = 0; kh < KH; ++kh) {
      const int hs = h + pad - kh;
      if (hs < 0 || hs % stride) continue;
      const int ho = hs / stride;
      if (ho >= Ho) continue;
      for (int kw = 0; kw < KW; ++kw) {
        const int ws = w + pad - kw;
        if (ws < 0 || ws % stride) continue;
        const int wo = ws / stride;
        if (wo >= Wo) continue;
        float v[8];
        Vec8<T>::load(dcol + (((long long)n * Ho + ho) * Wo + wo) * row_len + ((long long)kh * KW + kw) * Cp + c8 * 8, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += v[j];
      }
    }
    T* dst = dx + (((long long)n * H + h) * W + w) * C + c8 * 8;
    if ((C & 7) == 0) Vec8<T>::store(dst, acc);
    else
      for (int j = 0; j < 8; ++j) if (c8 * 8 + j < C) dst[j] = from_f<T>(acc[j]);
  }
}

// ------------------------------------------------------------------ per-channel reductions over rows
// sums[0][c] += sum_r f0(r, c);  sums[1][c] += sum_r f1(r, c)   (atomics: `sums` must be zeroed by the caller)
//   mode 0: stats       -> sum x, sum x^2
//   mode 1: bn backward -> sum dy_m, sum dy_m * (x - mean[c]) * rstd[c]   (dy_m = dy masked by relu(y))
// Each thread owns 8 consecutive channels (one 128-bit load per row); a block covers `nvx` channel vectors x 256/nvx rows.
template <typename T, int mode>
__global__ void __launch_bounds__(256) col_reduce_kernel(const T* __restrict__ a, const T* __restrict__ xin,
                                                         const T* __restrict__ y, const float* __restrict__ mean,
                                                         const float* __restrict__ rstd, float* __restrict__ sums,
                                                         long long rows, int C, int relu_mask, int nvx) {
  const int tx = threadIdx.x % nvx, ty = threadIdx.x / nvx, nty = 256 / nvx;
  const int c = (blockIdx.x * nvx + tx) * 8;
  const bool active = c < C;
  const long long rows_per = (rows + gridDim.y - 1) / gridDim.y;
  const long long r0 = blockIdx.y * rows_per;
  const long long r1 = r0 + rows_per < rows ? r0 + rows_per : rows;
  float s0[8], s1[8], mu[8], rs[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { s0[j] = 0.f; s1[j] = 0.f; mu[j] = 0.f; rs[j] = 0.f; }
  if (active && mode == 1) {
#pragma unroll
    for (int j = 0; j < 8; ++j) { mu[j] = mean[c + j]; rs[j] = rstd[c + j]; }
  }
  if (active) {
    // kB independent rows per iteration: all their 128-bit loads are issued before the first add, so a thread keeps
    // kB (stats) or 2-3 x kB (backward) requests in flight - the kernel is a pure HBM stream and needs ~6 MB in flight.
    constexpr int kB = mode == 0 ? 8 : 4;
    long long r = r0 + ty;
    for (; r + (long long)(kB - 1) * nty < r1; r += (long long)kB * nty) {
      float v[kB][8], xv[mode == 1 ? kB : 1][8], yv[mode == 1 ? kB : 1][8];
#pragma unroll
      for (int b = 0; b < kB; ++b) {
        const long long off = (r + (long long)b * nty) * C + c;
        Vec8<T>::load(a + off, v[b]);
        if constexpr (mode == 1) {
          Vec8<T>::load(xin + off, xv[b]);
          if (relu_mask) Vec8<T>::load(y + off, yv[b]);
        }
      }
#pragma unroll
      for (int b = 0; b < kB; ++b) {
        if constexpr (mode == 0) {
#pragma unroll
          for (int j = 0; j < 8; ++j) { s0[j] += v[b][j]; s1[j] += v[b][j] * v[b][j]; }
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float g = (relu_mask && !(yv[b][j] > 0.f)) ? 0.f : v[b][j];
            s0[j] += g;
            s1[j] += g * (xv[b][j] - mu[j]) * rs[j];
          }
        }
      }
    }
    for (; r < r1; r += nty) {
      float v[8];
      Vec8<T>::load(a + r * C + c, v);
      if (mode == 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j) { s0[j] += v[j]; s1[j] += v[j] * v[j]; }
      } else {
        float xv[8], yv[8];
        Vec8<T>::load(xin + r * C + c, xv);
        if (relu_mask) Vec8<T>::load(y + r * C + c, yv);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float g = (relu_mask && !(yv[j] > 0.f)) ? 0.f : v[j];
          s0[j] += g;
          s1[j] += g * (xv[j] - mu[j]) * rs[j];
        }
      }
    }
  }
  __shared__ float sh[2][256][9];
#pragma unroll
  for (int j = 0; j < 8; ++j) { sh[0][threadIdx.x][j] = s0[j]; sh[1][threadIdx.x][j] = s1[j]; }
  __syncthreads();
  if (ty == 0 && active) {
    for (int t = 1; t < nty; ++t) {
#pragma unroll
      for (int j = 0; j < 8; ++j) { s0[j] += sh[0][t * nvx + tx][j]; s1[j] += sh[1][t * nvx + tx][j]; }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) { atomicAdd(sums + c + j, s0[j]); atomicAdd(sums + C + c + j, s1[j]); }
  }
}

// mean/rstd from sums; update running stats (momentum, unbiased variance) like nn.BatchNorm2d
__global__ void bn_finalize_kernel(const float* __restrict__ sums, float* __restrict__ mean, float* __restrict__ rstd,
                                   float* __restrict__ running_mean, float* __restrict__ running_var, long long rows,
                                   int C, float eps, float momentum) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float m = sums[c] / (float)rows;
  const float var = fmaxf(sums[C + c] / (float)rows - m * m, 0.f);
  mean[c] = m;
  rstd[c] = rsqrtf(var + eps);
  if (running_mean) {
    const float unbiased = rows > 1 ? var * (float)rows / (float)(rows - 1) : var;
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * m;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
  }
}

// y = relu?( (x - mean) * rstd * gamma + beta + residual? )
template <typename T>
__global__ void bn_apply_kernel(const T* __restrict__ x, const float* __restrict__ mean, const float* __restrict__ rstd,
                                const float* __restrict__ gamma, const float* __restrict__ beta,
                                const T* __restrict__ residual, T* __restrict__ y, long long rows, int C, int relu) {
  // Each thread keeps ONE channel vector (8 channels) for its whole grid-stride walk, so the per-channel
  // scale / shift are computed once and the inner loop is load - fma - store.
  const int c8n = C >> 3;
  const long long nthreads = (long long)gridDim.x * blockDim.x;
  const long long tid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long lanes_per_vec = nthreads / c8n;     // host guarantees nthreads % c8n == 0
  if (lanes_per_vec == 0) return;
  const int cv = (int)(tid % c8n);
  const long long lane = tid / c8n;
  if (lane >= lanes_per_vec) return;
  float sc[8], sh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sc[j] = rstd[cv * 8 + j] * gamma[cv * 8 + j];
    sh[j] = beta[cv * 8 + j] - mean[cv * 8 + j] * sc[j];
  }
  for (long long r = lane; r < rows; r += lanes_per_vec) {
    const long long i = r * c8n + cv;
    float v[8], rr[8];
    Vec8<T>::load(x + i * 8, v);
    if (residual) Vec8<T>::load(residual + i * 8, rr);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float o = fmaf(v[j], sc[j], sh[j]);
      if (residual) o += rr[j];
      v[j] = relu ? fmaxf(o, 0.f) : o;
    }
    Vec8<T>::store(y + i * 8, v);
  }
}

// dx = gamma * rstd * (dy_m - sum_dy / M - xhat * sum_dy_xhat / M);  dres = dy_m (optional)
template <typename T>
__global__ void bn_bwd_apply_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ y,
                                    const float* __restrict__ mean, const float* __restrict__ rstd,
                                    const float* __restrict__ gamma, const float* __restrict__ sums,
                                    T* __restrict__ dx, T* __restrict__ dres, long long rows, int C, int relu_mask) {
  // dx = A * g + B * x + K with per-channel A = gamma*rstd, B = -A*rstd*s1/M, K = -A*s0/M - B*mean
  const int c8n = C >> 3;
  const float invM = 1.f / (float)rows;
  const long long nthreads = (long long)gridDim.x * blockDim.x;
  const long long tid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long lanes_per_vec = nthreads / c8n;
  if (lanes_per_vec == 0) return;
  const int cv = (int)(tid % c8n);
  const long long lane = tid / c8n;
  if (lane >= lanes_per_vec) return;
  float ca[8], cb[8], ck[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = cv * 8 + j;
    ca[j] = gamma[c] * rstd[c];
    cb[j] = -ca[j] * rstd[c] * sums[C + c] * invM;
    ck[j] = -ca[j] * sums[c] * invM - cb[j] * mean[c];
  }
  for (long long r = lane; r < rows; r += lanes_per_vec) {
    const long long i = r * c8n + cv;
    float g[8], xv[8], yv[8], o[8];
    Vec8<T>::load(dy + i * 8, g);
    Vec8<T>::load(x + i * 8, xv);
    if (relu_mask) Vec8<T>::load(y + i * 8, yv);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (relu_mask && !(yv[j] > 0.f)) g[j] = 0.f;
      o[j] = fmaf(ca[j], g[j], fmaf(cb[j], xv[j], ck[j]));
    }
    Vec8<T>::store(dx + i * 8, o);
    if (dres) Vec8<T>::store(dres + i * 8, g);
  }
}

// ------------------------------------------------------------------ pooling (NHWC)
template <typename T>
__global__ void maxpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, unsigned char* __restrict__ idx, int N,
                                   int H, int W, int C, int K, int stride, int pad, int Ho, int Wo) {
  const int c8n = C >> 3;   // host guarantees C % 8 == 0
  const long long total = (long long)N * Ho * Wo * c8n;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    long long r = i;
    const int c = (int)(r % c8n) * 8; r /= c8n;
    const int wo = (int)(r % Wo); r /= Wo;
    const int ho = (int)(r % Ho);
    const int n = (int)(r / Ho);
    float m[8];
    int arg[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { m[j] = -INFINITY; arg[j] = 0; }
    for (int kh = 0; kh < K; ++kh) {
      const int h = ho * stride - pad + kh;
      if (h < 0 || h >= H) continue;
      for (int kw = 0; kw < K; ++kw) {
        const int w = wo * stride - pad + kw;
        if (w < 0 || w >= W) continue;
        float v[8];
        Vec8<T>::load(x + (((long long)n * H + h) * W + w) * C + c, v);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (v[j] > m[j]) { m[j] = v[j]; arg[j] = kh * K + kw; }  // first maximum wins, like torch
      }
    }
    const long long o = (((long long)n * Ho + ho) * Wo + wo) * C + c;
    Vec8<T>::store(y + o, m);
    uint2 packed;
    packed.x = arg[0] | (arg[1] << 8) | (arg[2] << 16) | (arg[3] << 24);
    packed.y = arg[4] | (arg[5] << 8) | (arg[6] << 16) | (arg[7] << 24);
    *reinterpret_cast<uint2*>(idx + o) = packed;
  }
}

// gather form (no atomics): dx[p] = sum of dy over the windows whose recorded argmax is p
template <typename T>
__global__ void maxpool_bwd_kernel(const T* __restrict__ dy, const unsigned char* __restrict__ idx,
                                   T* __restrict__ dx, int N, int H, int W, int C, int K, int stride, int pad,
                                   int Ho, int Wo) {
  const int c8n = C >> 3;
  const long long total = (long long)N * H * W * c8n;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    long long r = i;
    const int c = (int)(r % c8n) * 8; r /= c8n;
    const int w = (int)(r % W); r /= W;
    const int h = (int)(r % H);
    const int n = (int)(r / H);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    for (int kh = 0; kh < K; ++kh) {
      const int hs = h + pad - kh;
      if (hs < 0 || hs % stride) continue;
      const int ho = hs / stride;
      if (ho >= Ho) continue;
      for (int kw = 0; kw < K; ++kw) {
        const int ws = w + pad - kw;
        if (ws < 0 || ws % stride) continue;
        const int wo = ws / stride;
        if (wo >= Wo) continue;
        const long long o = (((long long)n * Ho + ho) * Wo + wo) * C + c;
        const uint2 pk = *reinterpret_cast<const uint2*>(idx + o);
        float g[8];
        Vec8<T>::load(dy + o, g);
        const unsigned tag = kh * K + kw;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const unsigned a = ((j < 4 ? pk.x : pk.y) >> ((j & 3) * 8)) & 0xffu;
          if (a == tag) acc[j] += g[j];
        }
      }
    }
    Vec8<T>::store(dx + (((long long)n * H + h) * W + w) * C + c, acc);
  }
}

// y[n, c] = mean over hw of x[n, hw, c]   /   dx[n, hw, c] = dy[n, c] / HW
template <typename T>
__global__ void avgpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int N, int HW, int C) {
  const long long total = (long long)N * C;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const int n = (int)(i / C);
    float s = 0.f;
    for (int p = 0; p < HW; ++p) s += to_f<T>(x[((long long)n * HW + p) * C + c]);
    y[i] = from_f<T>(s / (float)HW);
  }
}
template <typename T>
__global__ void avgpool_bwd_kernel(const T* __restrict__ dy, T* __restrict__ dx, int N, int HW, int C) {
  const long long total = (long long)N * HW * C;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const int n = (int)(i / ((long long)HW * C));
    dx[i] = from_f<T>(to_f<T>(dy[(long long)n * C + c]) / (float)HW);
  }
}

// ------------------------------------------------------------------ 2x2 stride-2 transposed conv shuffle
// g[(n,h,w), (dh,dw,k)] <-> y[n, 2h+dh, 2w+dw, k]   (dir 0: g -> y ; dir 1: y -> g)
template <typename T>
__global__ void pixel_shuffle2_kernel(const T* __restrict__ src, T* __restrict__ dst, int N, int H, int W, int K, int dir) {
  const long long total = (long long)N * H * W * 4 * K;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    long long r = i;
    const int k = (int)(r % K); r /= K;
    const int dw = (int)(r % 2); r /= 2;
    const int dh = (int)(r % 2); r /= 2;
    const int w = (int)(r % W); r /= W;
    const int h = (int)(r % H);
    const int n = (int)(r / H);
    const long long yi = (((long long)n * 2 * H + 2 * h + dh) * 2 * W + 2 * w + dw) * K + k;
    if (dir == 0) dst[yi] = src[i]; else dst[i] = src[yi];
  }
}

// ------------------------------------------------------------------ bilinear resize (align_corners=False), NHWC
template <typename T>
__global__ void bilinear_kernel(const T* __restrict__ src, T* __restrict__ dst, int N, int Hi, int Wi, int Ho, int Wo,
                                int C, int backward) {
  // forward: dst[n,ho,wo,c] = sum of 4 taps of src; backward: scatter dy (src = dy [N,Ho,Wo,C]) into dst = dx (atomic fp32 not
  // needed: dst is T and we use a gather over output pixels via atomicAdd on float only when T == float)
  const long long total = (long long)N * Ho * Wo * C;
  const float sh = (float)Hi / (float)Ho, sw = (float)Wi / (float)Wo;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    long long r = i;
    const int c = (int)(r % C); r /= C;
    const int wo = (int)(r % Wo); r /= Wo;
    const int ho = (int)(r % Ho);
    const int n = (int)(r / Ho);
    float fy = fmaxf((ho + 0.5f) * sh - 0.5f, 0.f), fx = fmaxf((wo + 0.5f) * sw - 0.5f, 0.f);
    const int y0 = min((int)fy, Hi - 1), x0 = min((int)fx, Wi - 1);
    const int y1 = min(y0 + 1, Hi - 1), x1 = min(x0 + 1, Wi - 1);
    const float ly = fy - (float)y0, lx = fx - (float)x0;
    const float w00 = (1.f - ly) * (1.f - lx), w01 = (1.f - ly) * lx, w10 = ly * (1.f - lx), w11 = ly * lx;
    const long long base = (long long)n * Hi * Wi;
    if (!backward) {
      const float v = w00 * to_f<T>(src[((base + (long long)y0 * Wi + x0)) * C + c]) +
                      w01 * to_f<T>(src[((base + (long long)y0 * Wi + x1)) * C + c]) +
                      w10 * to_f<T>(src[((base + (long long)y1 * Wi + x0)) * C + c]) +
                      w11 * to_f<T>(src[((base + (long long)y1 * Wi + x1)) * C + c]);
      dst[i] = from_f<T>(v);
    } else {
      float* d = reinterpret_cast<float*>(dst);  // backward accumulates into an fp32 buffer
      const float g = to_f<T>(src[i]);
      atomicAdd(d + (base + (long long)y0 * Wi + x0) * C + c, w00 * g);
      atomicAdd(d + (base + (long long)y0 * Wi + x1) * C + c, w01 * g);
      atomicAdd(d + (base + (long long)y1 * Wi + x0) * C + c, w10 * g);
      atomicAdd(d + (base + (long long)y1 * Wi + x1) * C + c, w11 * g);
    }
  }
}

// ------------------------------------------------------------------ latitude-weighted MSE  (NCHW or NHWC: h index given)
// loss = mean( w[h] * (pred - target)^2 );  grad = 2 * w[h] * (pred - target) * gscale / numel
template <typename T>
__global__ void lat_mse_kernel(const T* __restrict__ pred, const T* __restrict__ target, const float* __restrict__ wlat,
                               float* __restrict__ loss_sum, T* __restrict__ grad, const float* __restrict__ gscale,
                               long long total, int H, long long h_div, int backward) {
  __shared__ float red[33];
  float s = 0.f;
  const float gs = backward ? (*gscale) * 2.f / (float)total : 0.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int h = (int)((i / h_div) % H);
    const float d = to_f<T>(pred[i]) - to_f<T>(target[i]);
    if (!backward) s += wlat[h] * d * d;
    else grad[i] = from_f<T>(wlat[h] * d * gs);
  }
  if (!backward) {
    s = block_sum(s, red);
    if (threadIdx.x == 0) atomicAdd(loss_sum, s / (float)total);
  }
}

}  // namespace b200

using namespace b200;

#define CONV_ARGS int N, int H, int W, int C, int Cp, int KH, int KW, int stride, int pad, int Ho, int Wo

B200_API int b200_im2col(const void* x, void* col, CONV_ARGS, int dtype, cudaStream_t stream) {
  const long long total = (long long)N * Ho * Wo * KH * KW * (Cp >> 3);
  if (total <= 0) return B200_OK;
  const long long pixels = (long long)N * Ho * Wo;
  B200_DISPATCH_DTYPE(dtype, T, (im2col_kernel<T><<<ew_grid(pixels * 32, 256, 16), 256, 0, stream>>>(
      (const T*)x, (T*)col, N, H, W, C, Cp, KH, KW, stride, pad, Ho, Wo)));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_col2im(const void* dcol, void* dx, CONV_ARGS, int dtype, cudaStream_t stream) {
  const long long total = (long long)N * H * W * (Cp >> 3);
  if (total <= 0) return B200_OK;
  B200_DISPATCH_DTYPE(dtype, T, (col2im_kernel<T><<<ew_grid(total, 256), 256, 0, stream>>>(
      (const T*)dcol, (T*)dx, N, H, W, C, Cp, KH, KW, stride, pad, Ho, Wo)));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// sums: fp32 [2, C], MUST be zeroed by the caller (kernel accumulates with atomics).
B200_API int b200_col_reduce(const void* a, const void* xin, const void* y, const float* mean, const float* rstd,
                             float* sums, long long rows, int C, int mode, int relu_mask, int dtype,
                             cudaStream_t stream) {
  if (rows <= 0 || C <= 0) return B200_OK;
  if (C & 7) { b200_set_error("col_reduce: C %% 8 != 0"); return B200_ERR_SHAPE; }
  const int nvc = C >> 3;
  int nvx = 1;
  while (nvx * 2 <= nvc && nvx < 32) nvx <<= 1;        // channel vectors per block row (power of two <= 32)
  const int gx = (nvc + nvx - 1) / nvx;
  const int rows_per_iter = 256 / nvx;
  long long gy = (rows + (long long)rows_per_iter * 8 - 1) / ((long long)rows_per_iter * 8);
  const long long cap = (long long)b200_sm_count() * 4 / gx + 1;   // few blocks: the final atomics serialise per address
  if (gy > cap) gy = cap;
  if (gy < 1) gy = 1;
  dim3 grd(gx, (unsigned)gy);
  if (mode == 0) {
    B200_DISPATCH_DTYPE(dtype, T, (col_reduce_kernel<T, 0><<<grd, 256, 0, stream>>>(
        (const T*)a, (const T*)xin, (const T*)y, mean, rstd, sums, rows, C, relu_mask, nvx)));
  } else {
    B200_DISPATCH_DTYPE(dtype, T, (col_reduce_kernel<T, 1><<<grd, 256, 0, stream>>>(
        (const T*)a, (const T*)xin, (const T*)y, mean, rstd, sums, rows, C, relu_mask, nvx)));
  }
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_bn_finalize(const float* sums, float* mean, float* rstd, float* running_mean, float* running_var,
                              long long rows, int C, float eps, float momentum, cudaStream_t stream) {
  bn_finalize_kernel<<<(C + 127) / 128, 128, 0, stream>>>(sums, mean, rstd, running_mean, running_var, rows, C, eps, momentum);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_bn_apply(const void* x, const float* mean, const float* rstd, const float* gamma, const float* beta,
                           const void* residual, void* y, long long rows, int C, int relu, int dtype,
                           cudaStream_t stream) {
  if (rows <= 0) return B200_OK;
  if (C & 7) { b200_set_error("bn_apply: C %% 8 != 0"); return B200_ERR_SHAPE; }
  const long long total = rows * (C >> 3);
  int grid_bn = ew_grid(total, 256);
  {  // every thread keeps one channel vector: make (grid * 256) a multiple of C/8
    const int c8n = C >> 3;
    int unit = c8n;
    while (unit % 2 == 0 && (unit / 2) * 256 % c8n == 0) unit /= 2;   // smallest block count whose threads tile c8n
    grid_bn = (grid_bn + unit - 1) / unit * unit;
  }
  B200_DISPATCH_DTYPE(dtype, T, (bn_apply_kernel<T><<<grid_bn, 256, 0, stream>>>(
      (const T*)x, mean, rstd, gamma, beta, (const T*)residual, (T*)y, rows, C, relu)));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_bn_bwd_apply(const void* dy, const void* x, const void* y, const float* mean, const float* rstd,
                               const float* gamma, const float* sums, void* dx, void* dres, long long rows, int C,
                               int relu_mask, int dtype, cudaStream_t stream) {
  if (rows <= 0) return B200_OK;
  if (C & 7) { b200_set_error("bn_bwd_apply: C %% 8 != 0"); return B200_ERR_SHAPE; }
  const long long total = rows * (C >> 3);
  int grid_bn = ew_grid(total, 256);
  {
    const int c8n = C >> 3;
    int unit = c8n;
    while (unit % 2 == 0 && (unit / 2) * 256 % c8n == 0) unit /= 2;
    grid_bn = (grid_bn + unit - 1) / unit * unit;
  }
  B200_DISPATCH_DTYPE(dtype, T, (bn_bwd_apply_kernel<T><<<grid_bn, 256, 0, stream>>>(
      (const T*)dy, (const T*)x, (const T*)y, mean, rstd, gamma, sums, (T*)dx, (T*)dres, rows, C, relu_mask)));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// forward: x -> out (+ idx, uint8 window offset of the max);  backward: a = dy, idx -> out = dx.  K <= 15.
B200_API int b200_maxpool(const void* a, const void* x, void* idx, void* out, int N, int H, int W, int C, int K,
                          int stride, int pad, int Ho, int Wo, int backward, int dtype, cudaStream_t stream) {
  if (C & 7) { b200_set_error("maxpool: C %% 8 != 0"); return B200_ERR_SHAPE; }
  const long long total = (backward ? (long long)N * H * W : (long long)N * Ho * Wo) * (C >> 3);
  if (total <= 0) return B200_OK;
  B200_DISPATCH_DTYPE(dtype, T, {
    if (!backward)
      maxpool_fwd_kernel<T><<<ew_grid(total, 256), 256, 0, stream>>>((const T*)x, (T*)out, (unsigned char*)idx, N, H, W, C, K,
                                                                   stride, pad, Ho, Wo);
    else
      maxpool_bwd_kernel<T><<<ew_grid(total, 256), 256, 0, stream>>>((const T*)a, (const unsigned char*)idx, (T*)out, N, H, W,
                                                                   C, K, stride, pad, Ho, Wo);
  });
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_avgpool(const void* in, void* out, int N, int HW, int C, int backward, int dtype, cudaStream_t stream) {
  const long long total = backward ? (long long)N * HW * C : (long long)N * C;
  if (total <= 0) return B200_OK;
  B200_DISPATCH_DTYPE(dtype, T, {
    if (!backward) avgpool_fwd_kernel<T><<<ew_grid(total, 128), 128, 0, stream>>>((const T*)in, (T*)out, N, HW, C);
    else avgpool_bwd_kernel<T><<<ew_grid(total, 256), 256, 0, stream>>>((const T*)in, (T*)out, N, HW, C);
  });
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_pixel_shuffle2(const void* src, void* dst, int N, int H, int W, int K, int dir, int dtype,
                                 cudaStream_t stream) {
  const long long total = (long long)N * H * W * 4 * K;
  if (total <= 0) return B200_OK;
  B200_DISPATCH_DTYPE(dtype, T, (pixel_shuffle2_kernel<T><<<ew_grid(total, 256), 256, 0, stream>>>(
      (const T*)src, (T*)dst, N, H, W, K, dir)));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// forward: src [N,Hi,Wi,C] -> dst [N,Ho,Wo,C].  backward: src = dy [N,Ho,Wo,C], dst = fp32 dx [N,Hi,Wi,C] (zeroed by caller)
B200_API int b200_bilinear(const void* src, void* dst, int N, int Hi, int Wi, int Ho, int Wo, int C, int backward,
                           int dtype, cudaStream_t stream) {
  const long long total = (long long)N * Ho * Wo * C;
  if (total <= 0) return B200_OK;
  B200_DISPATCH_DTYPE(dtype, T, (bilinear_kernel<T><<<ew_grid(total, 256), 256, 0, stream>>>(
      (const T*)src, (T*)dst, N, Hi, Wi, Ho, Wo, C, backward)));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// h index of flat element i is (i / h_div) % H  (NCHW contiguous: h_div = W; NHWC: h_div = W*C)
B200_API int b200_lat_mse(const void* pred, const void* target, const float* wlat, float* loss_sum, void* grad,
                          const float* gscale, long long total, int H, long long h_div, int backward, int dtype,
                          cudaStream_t stream) {
  if (total <= 0) return B200_OK;
  B200_DISPATCH_DTYPE(dtype, T, (lat_mse_kernel<T><<<ew_grid(total, 256, 4), 256, 0, stream>>>(
      (const T*)pred, (const T*)target, wlat, loss_sum, (T*)grad, gscale, total, H, h_div, backward)));
  B200_LAUNCH_CHECK();
  return B200_OK;
}
