// Memory-bound Llama/ViT ops: RoPE, SwiGLU, GELU/ReLU, residual add, cast/scale,
// embedding gather/scatter, softmax cross-entropy.  Each is a single pass over
// HBM with 128-bit accesses; the roofline for all of them is bytes / HBM bandwidth.
//
// Reference expressions being replaced:
//   RoPE      /root/reference/fsdp_tp/llama2_model.py:73-100  (interleaved pairs, fp32 math)
//   SwiGLU    /root/reference/fsdp_tp/llama2_model.py:266-267 (w2(silu(w1 x) * w3 x))
//   Embedding /root/reference/fsdp_tp/llama2_model.py:374,440
//   CE loss   /root/reference/scripts/main.py:308; scripts/04_pipeline_parallel_pp/03_pipeline_training.py:207
#include "vec.cuh"

namespace b200 {

// ------------------------------- RoPE --------------------------------------
// x: [rows = B*S*H, hd]; token position = pos_offset + (row / H) % S.
// freqs: [max_pos, hd/2] float2 (cos, sin) == torch.view_as_real(freqs_cis).
template <typename T>
__global__ void rope_kernel(const T* __restrict__ x, T* __restrict__ y, const float2* __restrict__ freqs,
                            long long n_chunks, int S, int H, int hd, int pos_offset, float sin_sign) {
  const int chunks_per_row = hd >> 3;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_chunks;
       i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / chunks_per_row;
    const int c = (int)(i - row * chunks_per_row);
    const int s = (int)((row / H) % S) + pos_offset;
    float v[8], o[8];
    Vec8<T>::load(x + i * 8, v);
    const float2* f = freqs + (long long)s * (hd >> 1) + c * 4;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 cs = f[j];
      const float sn = cs.y * sin_sign;
      o[2 * j] = v[2 * j] * cs.x - v[2 * j + 1] * sn;
      o[2 * j + 1] = v[2 * j] * sn + v[2 * j + 1] * cs.x;
    }
    Vec8<T>::store(y + i * 8, o);
  }
}

// Same rotation applied IN A FUSED QKV BUFFER: x is [tokens, ld] with the first n_heads * hd columns of every token
// row holding the heads to rotate (q heads followed by k heads; the v columns after them are left alone), so one
// launch covers q and k and nothing is copied out of the projection output.  y may alias x.
template <typename T>
__global__ void rope_strided_kernel(const T* __restrict__ x, T* __restrict__ y, const float2* __restrict__ freqs,
                                    long long tokens, int S, int n_heads, int hd, long long ldx, long long ldy,
                                    int pos_offset, float sin_sign) {
  const int cph = hd >> 3;               // 8-element chunks per head
  const int cpt = n_heads * cph;         // chunks per token
  const long long n = tokens * cpt;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    const long long tok = i / cpt;
    const int c = (int)(i - tok * cpt);
    const int s = (int)(tok % S) + pos_offset;
    float v[8], o[8];
    Vec8<T>::load(x + tok * ldx + c * 8, v);
    const float2* f = freqs + (long long)s * (hd >> 1) + (c % cph) * 4;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 cs = f[j];
      const float sn = cs.y * sin_sign;
      o[2 * j] = v[2 * j] * cs.x - v[2 * j + 1] * sn;
      o[2 * j + 1] = v[2 * j] * sn + v[2 * j + 1] * cs.x;
    }
    Vec8<T>::store(y + tok * ldy + c * 8, o);
  }
}

// ------------------------------- SwiGLU ------------------------------------
template <typename T>
__global__ void swiglu_fwd_kernel(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ y,
                                  long long rows, int F, long long lda, long long ldb, long long ldy) {
  // (row, chunk) advance incrementally: one 64-bit division per thread instead of one per 8 elements (the division and
  // its address arithmetic were a third of the instructions of these kernels, which sit close to the issue limit)
  const int cpr = F >> 3;
  const long long stride = (long long)gridDim.x * blockDim.x, i0 = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long dr = stride / cpr;
  const int dc = (int)(stride - dr * cpr);
  long long r = i0 / cpr;
  int cc = (int)(i0 - r * cpr);
  for (; r < rows; r += dr, cc += dc) {
    if (cc >= cpr) { cc -= cpr; ++r; if (r >= rows) break; }
    const int c = cc * 8;
    float av[8], bv[8], o[8];
    Vec8<T>::load(a + r * lda + c, av);
    Vec8<T>::load(b + r * ldb + c, bv);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = av[j] / (1.f + __expf(-av[j])) * bv[j];
    Vec8<T>::store(y + r * ldy + c, o);
  }
}

template <typename T>
__global__ void swiglu_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ a,
                                  const T* __restrict__ b, T* __restrict__ da, T* __restrict__ db,
                                  long long rows, int F, long long ldy, long long lda, long long ldb,
                                  long long ldda, long long lddb) {
  const int cpr = F >> 3;
  const long long stride = (long long)gridDim.x * blockDim.x, i0 = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long dr = stride / cpr;
  const int dc = (int)(stride - dr * cpr);
  long long r = i0 / cpr;
  int cc = (int)(i0 - r * cpr);
  for (; r < rows; r += dr, cc += dc) {
    if (cc >= cpr) { cc -= cpr; ++r; if (r >= rows) break; }
    const int c = cc * 8;
    float g[8], av[8], bv[8], oa[8], ob[8];
    Vec8<T>::load(dy + r * ldy + c, g);
    Vec8<T>::load(a + r * lda + c, av);
    Vec8<T>::load(b + r * ldb + c, bv);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float sg = 1.f / (1.f + __expf(-av[j]));
      const float silu = av[j] * sg;
      oa[j] = g[j] * bv[j] * (sg + silu * (1.f - sg));
      ob[j] = g[j] * silu;
    }
    Vec8<T>::store(da + r * ldda + c, oa);
    Vec8<T>::store(db + r * lddb + c, ob);
  }
}

// --------------------- unary / binary elementwise ---------------------------
// op: 0 gelu(erf) fwd, 1 gelu bwd (x, dy), 2 relu fwd, 3 relu bwd (x, dy), 4 add (x + y2),
//     5 scale (alpha * x), 6 axpy (x + alpha * y2), 7 silu fwd, 8 silu bwd
template <typename T>
__global__ void ew_kernel(const T* __restrict__ x, const T* __restrict__ y2, T* __restrict__ out,
                          long long n8, long long n, int op, float alpha) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8;
       i += (long long)gridDim.x * blockDim.x) {
    float a[8], b[8], o[8];
    Vec8<T>::load(x + i * 8, a);
    if (y2) Vec8<T>::load(y2 + i * 8, b);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float v = a[j];
      switch (op) {
        case 0: o[j] = 0.5f * v * (1.f + erff(v * 0.70710678118654752f)); break;
        case 1: {
          const float cdf = 0.5f * (1.f + erff(v * 0.70710678118654752f));
          const float pdf = 0.3989422804014327f * __expf(-0.5f * v * v);
          o[j] = b[j] * (cdf + v * pdf);
        } break;
        case 2: o[j] = fmaxf(v, 0.f); break;
        case 3: o[j] = v > 0.f ? b[j] : 0.f; break;
        case 4: o[j] = v + b[j]; break;
        case 5: o[j] = alpha * v; break;
        case 6: o[j] = v + alpha * b[j]; break;
        case 7: o[j] = v / (1.f + __expf(-v)); break;
        default: {
          const float sg = 1.f / (1.f + __expf(-v));
          o[j] = b[j] * sg * (1.f + v * (1.f - sg));
        } break;
      }
    }
    Vec8<T>::store(out + i * 8, o);
  }
  // scalar tail (n not a multiple of 8)
  if (blockIdx.x == 0) {
    for (long long i = n8 * 8 + threadIdx.x; i < n; i += blockDim.x) {
      const float v = to_f<T>(x[i]);
      const float w = y2 ? to_f<T>(y2[i]) : 0.f;
      float o;
      switch (op) {
        case 0: o = 0.5f * v * (1.f + erff(v * 0.70710678118654752f)); break;
        case 1: o = w * (0.5f * (1.f + erff(v * 0.70710678118654752f)) + v * 0.3989422804014327f * __expf(-0.5f * v * v)); break;
        case 2: o = fmaxf(v, 0.f); break;
        case 3: o = v > 0.f ? w : 0.f; break;
        case 4: o = v + w; break;
        case 5: o = alpha * v; break;
        case 6: o = v + alpha * w; break;
        case 7: o = v / (1.f + __expf(-v)); break;
        default: { const float sg = 1.f / (1.f + __expf(-v)); o = w * sg * (1.f + v * (1.f - sg)); } break;
      }
      out[i] = from_f<T>(o);
    }
  }
}

// dst = (TD)(alpha * src)
template <typename TS, typename TD>
__global__ void cast_scale_kernel(const TS* __restrict__ src, TD* __restrict__ dst, long long n,
                                  float alpha) {
  const long long n8 = n >> 3;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8;
       i += (long long)gridDim.x * blockDim.x) {
    float v[8];
    Vec8<TS>::load(src + i * 8, v);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] *= alpha;
    Vec8<TD>::store(dst + i * 8, v);
  }
  if (blockIdx.x == 0)
    for (long long i = n8 * 8 + threadIdx.x; i < n; i += blockDim.x)
      dst[i] = from_f<TD>(to_f<TS>(src[i]) * alpha);
}

// ------------------------------- Embedding ---------------------------------
// Rows whose index falls outside [vocab_start, vocab_start + vocab_local) produce zeros
// (vocab-parallel shards: exactly one TP rank owns each token).
template <typename T>
__global__ void embedding_fwd_kernel(const long long* __restrict__ idx, const T* __restrict__ w,
                                     T* __restrict__ out, long long ntok, int D, long long vocab_start,
                                     long long vocab_local) {
  const int cpr = D >> 3;
  const long long n = ntok * cpr;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    const long long t = i / cpr;
    const int c = (int)(i - t * cpr) * 8;
    const long long id = idx[t] - vocab_start;
    uint4 z = make_uint4(0, 0, 0, 0);
    if constexpr (sizeof(T) == 2) {
      if (id >= 0 && id < vocab_local) z = *reinterpret_cast<const uint4*>(w + id * D + c);
      *reinterpret_cast<uint4*>(out + t * D + c) = z;
    } else {
      uint4 z2 = z;
      if (id >= 0 && id < vocab_local) {
        z = *reinterpret_cast<const uint4*>(w + id * D + c);
        z2 = *reinterpret_cast<const uint4*>(w + id * D + c + 4);
      }
      *reinterpret_cast<uint4*>(out + t * D + c) = z;
      *reinterpret_cast<uint4*>(out + t * D + c + 4) = z2;
    }
  }
}

// dW[idx[t]] += dy[t]   (dW is fp32; atomics because tokens repeat)
template <typename T>
__global__ void embedding_bwd_kernel(const long long* __restrict__ idx, const T* __restrict__ dy,
                                     float* __restrict__ dw, long long ntok, int D,
                                     long long vocab_start, long long vocab_local) {
  const int cpr = D >> 3;
  const long long n = ntok * cpr;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    const long long t = i / cpr;
    const int c = (int)(i - t * cpr) * 8;
    const long long id = idx[t] - vocab_start;
    if (id < 0 || id >= vocab_local) continue;
    float g[8];
    Vec8<T>::load(dy + t * D + c, g);
    float* dst = dw + id * D + c;
#pragma unroll
    for (int j = 0; j < 8; ++j) atomicAdd(dst + j, g[j]);
  }
}

// ------------------------------- Cross entropy ------------------------------
// One CTA per row.  loss[row] = lse - logit[target]; rows with target == ignore_index give 0.
template <typename T>
__global__ void xent_fwd_kernel(const T* __restrict__ logits, const long long* __restrict__ target,
                                float* __restrict__ loss, float* __restrict__ lse_out, int rows, int V,
                                long long ld, long long ignore_index, long long vocab_start) {
  __shared__ float red[33];
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const T* lr = logits + (long long)row * ld;
    float mx = -INFINITY;
    for (int i = threadIdx.x; i < V; i += blockDim.x) mx = fmaxf(mx, to_f<T>(lr[i]));
    mx = block_max(mx, red);
    float s = 0.f;
    for (int i = threadIdx.x; i < V; i += blockDim.x) s += __expf(to_f<T>(lr[i]) - mx);
    s = block_sum(s, red);
    if (threadIdx.x == 0) {
      const float lse = mx + __logf(s);
      // vocabulary-parallel use: this rank holds columns [vocab_start, vocab_start + V); a target outside contributes 0
      const long long tg = target[row], t = tg - vocab_start;
      lse_out[row] = lse;
      loss[row] = (tg == ignore_index || t < 0 || t >= V) ? 0.f : lse - to_f<T>(lr[t]);
    }
  }
}

// dlogits = (softmax - onehot) * (*gscale)        (gscale lives on the device: no host sync)
template <typename T>
__global__ void xent_bwd_kernel(const T* __restrict__ logits, const long long* __restrict__ target,
                                const float* __restrict__ lse, const float* __restrict__ gscale,
                                T* __restrict__ dlogits, int rows, int V, long long ld, long long ldd,
                                long long ignore_index, long long vocab_start, long long v_total) {
  const float gs = *gscale;
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const T* lr = logits + (long long)row * ld;
    T* dr = dlogits + (long long)row * ldd;
    const long long tg = target[row], t = tg - vocab_start;   // softmax term for every kept row, one-hot only if local
    const bool ignored = (tg == ignore_index || tg < 0 || tg >= v_total);
    const float l = lse[row];
    for (int i = threadIdx.x; i < V; i += blockDim.x) {
      float g = 0.f;
      if (!ignored) g = (__expf(to_f<T>(lr[i]) - l) - (i == t ? 1.f : 0.f)) * gs;
      dr[i] = from_f<T>(g);
    }
  }
}

// out[0] = sum(x[0..n)) * scale — tiny single-CTA reduction (loss mean).
__global__ void sum_scale_kernel(const float* __restrict__ x, float* __restrict__ out, int n,
                                 float scale) {
  __shared__ float red[33];
  float s = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) s += x[i];
  s = block_sum(s, red);
  if (threadIdx.x == 0) out[0] = s * scale;
}

}  // namespace b200

using namespace b200;

B200_API int b200_rope(const void* x, void* y, const float* freqs, long long rows, int S, int H,
                       int hd, int pos_offset, int backward, int dtype, cudaStream_t stream) {
  if (rows <= 0) return B200_OK;
  if (hd % 8) { b200_set_error("rope: head_dim %% 8 != 0"); return B200_ERR_SHAPE; }
  const long long n = rows * (hd >> 3);
  B200_DISPATCH_DTYPE(dtype, T, (rope_kernel<T><<<ew_grid(n, 256), 256, 0, stream>>>(
      (const T*)x, (T*)y, (const float2*)freqs, n, S, H, hd, pos_offset, backward ? -1.f : 1.f)));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// Rotates the first n_heads heads of every [tokens, ld] row (see rope_strided_kernel); tokens = B * S, position = token % S.
B200_API int b200_rope_strided(const void* x, void* y, const float* freqs, long long tokens, int S, int n_heads, int hd,
                               long long ldx, long long ldy, int pos_offset, int backward, int dtype, cudaStream_t stream) {
  if (tokens <= 0 || n_heads <= 0) return B200_OK;
  if ((hd % 8) || (ldx % 8) || (ldy % 8)) { b200_set_error("rope_strided: head_dim and row strides must be multiples of 8"); return B200_ERR_SHAPE; }
  const long long n = tokens * n_heads * (hd >> 3);
  B200_DISPATCH_DTYPE(dtype, T, (rope_strided_kernel<T><<<ew_grid(n, 256), 256, 0, stream>>>(
      (const T*)x, (T*)y, (const float2*)freqs, tokens, S, n_heads, hd, ldx, ldy, pos_offset, backward ? -1.f : 1.f)));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_swiglu_fwd(const void* a, const void* b, void* y, long long rows, int F,
                             long long lda, long long ldb, long long ldy, int dtype,
                             cudaStream_t stream) {
  if (rows <= 0) return B200_OK;
  if (F % 8 || lda % 8 || ldb % 8 || ldy % 8) { b200_set_error("swiglu: sizes must be multiples of 8"); return B200_ERR_SHAPE; }
  const long long n = rows * (F >> 3);
  B200_DISPATCH_DTYPE(dtype, T, (swiglu_fwd_kernel<T><<<ew_grid(n, 256), 256, 0, stream>>>(
      (const T*)a, (const T*)b, (T*)y, rows, F, lda, ldb, ldy)));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_swiglu_bwd(const void* dy, const void* a, const void* b, void* da, void* db,
                             long long rows, int F, long long ldy, long long lda, long long ldb,
                             long long ldda, long long lddb, int dtype, cudaStream_t stream) {
  if (rows <= 0) return B200_OK;
  if (F % 8 || lda % 8 || ldb % 8 || ldy % 8 || ldda % 8 || lddb % 8) { b200_set_error("swiglu: sizes must be multiples of 8"); return B200_ERR_SHAPE; }
  const long long n = rows * (F >> 3);
  B200_DISPATCH_DTYPE(dtype, T, (swiglu_bwd_kernel<T><<<ew_grid(n, 256), 256, 0, stream>>>(
      (const T*)dy, (const T*)a, (const T*)b, (T*)da, (T*)db, rows, F, ldy, lda, ldb, ldda, lddb)));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_elementwise(const void* x, const void* y2, void* out, long long n, int op,
                              float alpha, int dtype, cudaStream_t stream) {
  if (n <= 0) return B200_OK;
  const long long n8 = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out) |
                         reinterpret_cast<uintptr_t>(y2)) & 15) ? 0 : (n >> 3);
  B200_DISPATCH_DTYPE(dtype, T, (ew_kernel<T><<<ew_grid(n8 > 0 ? n8 : 1, 256), 256, 0, stream>>>(
      (const T*)x, (const T*)y2, (T*)out, n8, n, op, alpha)));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// dst = cast(alpha * src); dtype codes: 0 bf16, 1 fp32.  Pointers must be 16-byte aligned.
B200_API int b200_cast_scale(const void* src, int src_dtype, void* dst, int dst_dtype, long long n,
                             float alpha, cudaStream_t stream) {
  if (n <= 0) return B200_OK;
  const int grid = ew_grid(n >> 3 ? n >> 3 : 1, 256);
  if (src_dtype == 0 && dst_dtype == 1)
    cast_scale_kernel<__nv_bfloat16, float><<<grid, 256, 0, stream>>>((const __nv_bfloat16*)src, (float*)dst, n, alpha);
  else if (src_dtype == 1 && dst_dtype == 0)
    cast_scale_kernel<float, __nv_bfloat16><<<grid, 256, 0, stream>>>((const float*)src, (__nv_bfloat16*)dst, n, alpha);
  else if (src_dtype == 0 && dst_dtype == 0)
    cast_scale_kernel<__nv_bfloat16, __nv_bfloat16><<<grid, 256, 0, stream>>>((const __nv_bfloat16*)src, (__nv_bfloat16*)dst, n, alpha);
  else if (src_dtype == 1 && dst_dtype == 1)
    cast_scale_kernel<float, float><<<grid, 256, 0, stream>>>((const float*)src, (float*)dst, n, alpha);
  else { b200_set_error("cast_scale: bad dtype codes"); return B200_ERR_UNSUPPORTED; }
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_embedding_fwd(const long long* idx, const void* w, void* out, long long ntok, int D,
                                long long vocab_start, long long vocab_local, int dtype,
                                cudaStream_t stream) {
  if (ntok <= 0) return B200_OK;
  if (D % 8) { b200_set_error("embedding: D %% 8 != 0"); return B200_ERR_SHAPE; }
  const long long n = ntok * (D >> 3);
  B200_DISPATCH_DTYPE(dtype, T, (embedding_fwd_kernel<T><<<ew_grid(n, 256), 256, 0, stream>>>(
      idx, (const T*)w, (T*)out, ntok, D, vocab_start, vocab_local)));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_embedding_bwd(const long long* idx, const void* dy, float* dw, long long ntok, int D,
                                long long vocab_start, long long vocab_local, int dtype,
                                cudaStream_t stream) {
  if (ntok <= 0) return B200_OK;
  if (D % 8) { b200_set_error("embedding: D %% 8 != 0"); return B200_ERR_SHAPE; }
  const long long n = ntok * (D >> 3);
  B200_DISPATCH_DTYPE(dtype, T, (embedding_bwd_kernel<T><<<ew_grid(n, 256), 256, 0, stream>>>(
      idx, (const T*)dy, dw, ntok, D, vocab_start, vocab_local)));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_xent_fwd(const void* logits, const long long* target, float* loss, float* lse,
                           int rows, int V, long long ld, long long ignore_index, int dtype,
                           cudaStream_t stream) {
  if (rows <= 0) return B200_OK;
  const int threads = V >= 4096 ? 1024 : 256;
  const int grid = rows < b200_sm_count() * 4 ? rows : b200_sm_count() * 4;
  B200_DISPATCH_DTYPE(dtype, T, (xent_fwd_kernel<T><<<grid, threads, 0, stream>>>(
      (const T*)logits, target, loss, lse, rows, V, ld, ignore_index, 0LL)));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// Vocabulary-parallel pieces (avoids the all-gather of [M, V] logits behind
// ColwiseParallel(output_layouts=Replicate()), 01_fsdp_tp_hybrid.py:119-122).  Forward: per-row log-sum-exp over the
// LOCAL columns [vocab_start, vocab_start + V) and `lse_local - logit[target]` where the target is local, else 0; the
// caller combines the per-rank pairs (log-sum-exp of the lse, sum of the target terms).  Backward: given the GLOBAL
// lse, dlogits_local = (exp(x - lse) - onehot_local) * gscale.
B200_API int b200_xent_fwd_vp(const void* logits, const long long* target, float* loss, float* lse, int rows, int V,
                              long long ld, long long ignore_index, long long vocab_start, int dtype,
                              cudaStream_t stream) {
  if (rows <= 0) return B200_OK;
  const int threads = V >= 4096 ? 1024 : 256;
  const int grid = rows < b200_sm_count() * 4 ? rows : b200_sm_count() * 4;
  B200_DISPATCH_DTYPE(dtype, T, (xent_fwd_kernel<T><<<grid, threads, 0, stream>>>(
      (const T*)logits, target, loss, lse, rows, V, ld, ignore_index, vocab_start)));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_xent_bwd_vp(const void* logits, const long long* target, const float* lse, const float* gscale,
                              void* dlogits, int rows, int V, long long ld, long long ldd, long long ignore_index,
                              long long vocab_start, long long v_total, int dtype, cudaStream_t stream) {
  if (rows <= 0) return B200_OK;
  const int threads = V >= 4096 ? 1024 : 256;
  const int grid = rows < b200_sm_count() * 4 ? rows : b200_sm_count() * 4;
  B200_DISPATCH_DTYPE(dtype, T, (xent_bwd_kernel<T><<<grid, threads, 0, stream>>>(
      (const T*)logits, target, lse, gscale, (T*)dlogits, rows, V, ld, ldd, ignore_index, vocab_start, v_total)));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_xent_bwd(const void* logits, const long long* target, const float* lse,
                           const float* gscale, void* dlogits, int rows, int V, long long ld,
                           long long ldd, long long ignore_index, int dtype, cudaStream_t stream) {
  if (rows <= 0) return B200_OK;
  const int threads = V >= 4096 ? 1024 : 256;
  const int grid = rows < b200_sm_count() * 4 ? rows : b200_sm_count() * 4;
  B200_DISPATCH_DTYPE(dtype, T, (xent_bwd_kernel<T><<<grid, threads, 0, stream>>>(
      (const T*)logits, target, lse, gscale, (T*)dlogits, rows, V, ld, ldd, ignore_index, 0LL, (long long)V)));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_sum_scale(const float* x, float* out, int n, float scale, cudaStream_t stream) {
  sum_scale_kernel<<<1, 1024, 0, stream>>>(x, out, n, scale);
  B200_LAUNCH_CHECK();
  return B200_OK;
}
