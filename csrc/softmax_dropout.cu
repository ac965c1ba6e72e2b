// Softmax over attention scores with dropout on the PROBABILITIES (what nn.MultiheadAttention(dropout=p) does in the
// reference's pipeline transformer, /root/reference/scripts/04_pipeline_parallel_pp/03_pipeline_training.py:57-58), for the
// explicit S = QK^T -> softmax -> PV path that small-head models take when dropout is on.
//
// The mask is never stored: element (row, col) of the [rows, cols] score matrix draws its bit from Philox4x32-10 with
// key = seed and counter = (linear element index / 4, offset), so forward and backward regenerate the same mask and a
// host-side Philox can reproduce it for the oracle.  One warp per row; a lane owns groups of 4 consecutive columns
// (one Philox call per group).
//   forward : P = softmax(scale * S [+ causal mask]) (bf16, kept for backward), Pd = P * mask / (1 - p) (bf16, feeds P V)
//   backward: dS = scale * P * (m dPd - sum_j P_j m_j dPd_j),  m = mask / (1 - p)
#include "common.cuh"

namespace b200 {

struct Philox4 { uint32_t x, y, z, w; };

__device__ __forceinline__ Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
  constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
    const uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += W0; k1 += W1;
  }
  return {c0, c1, c2, c3};
}

// keep-multipliers (0 or 1 / (1 - p)) of the 4 elements of group g (linear element index 4 g .. 4 g + 3)
__device__ __forceinline__ void dropout_mult4(long long group, uint64_t seed, uint64_t offset, uint32_t thresh, float inv_keep,
                                              float (&m)[4]) {
  const Philox4 r = philox4x32_10((uint32_t)group, (uint32_t)((uint64_t)group >> 32), (uint32_t)offset, (uint32_t)(offset >> 32),
                                  (uint32_t)seed, (uint32_t)(seed >> 32));
  m[0] = r.x >= thresh ? inv_keep : 0.f;   // P(drop) = thresh / 2^32
  m[1] = r.y >= thresh ? inv_keep : 0.f;
  m[2] = r.z >= thresh ? inv_keep : 0.f;
  m[3] = r.w >= thresh ? inv_keep : 0.f;
}

template <typename T> __device__ __forceinline__ void ld4(const T* p, float (&v)[4]);
template <> __device__ __forceinline__ void ld4<float>(const float* p, float (&v)[4]) {
  const float4 t = *reinterpret_cast<const float4*>(p);
  v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
template <> __device__ __forceinline__ void ld4<__nv_bfloat16>(const __nv_bfloat16* p, float (&v)[4]) {
  const uint2 u = *reinterpret_cast<const uint2*>(p);
  v[0] = bf16lo(u.x); v[1] = bf16hi(u.x); v[2] = bf16lo(u.y); v[3] = bf16hi(u.y);
}
__device__ __forceinline__ void st4_bf16(__nv_bfloat16* p, const float (&v)[4]) {
  const __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
  uint2 u;
  u.x = *reinterpret_cast<const uint32_t*>(&a); u.y = *reinterpret_cast<const uint32_t*>(&b);
  *reinterpret_cast<uint2*>(p) = u;
}

// rows = B * H * Sq (row r belongs to query position r % Sq), cols = Sk (multiple of 4).  Causal: key j is visible to
// query i when j <= i + (Sk - Sq) (the bottom-right aligned mask of the flash kernel with q_offset = Sk - Sq).
template <typename TIn>
__global__ void __launch_bounds__(256) softmax_dropout_fwd_kernel(const TIn* __restrict__ s, __nv_bfloat16* __restrict__ p,
                                                                  __nv_bfloat16* __restrict__ pd, long long rows, int cols,
                                                                  int sq, float scale, int causal, uint32_t thresh,
                                                                  float inv_keep, uint64_t seed, uint64_t offset) {
  const long long row = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const int limit = causal ? min(cols, (int)(row % sq) + (cols - sq) + 1) : cols;   // visible keys: [0, limit)
  const TIn* srow = s + row * cols;
  const float sl2 = scale * 1.4426950408889634f;
  float mx = -INFINITY;
  for (int c = lane * 4; c < limit; c += 128) {
    float v[4];
    ld4<TIn>(srow + c, v);
#pragma unroll
    for (int j = 0; j < 4; ++j) if (c + j < limit) mx = fmaxf(mx, v[j]);
  }
  mx = warp_max(mx);
  float sum = 0.f;
  for (int c = lane * 4; c < limit; c += 128) {
    float v[4];
    ld4<TIn>(srow + c, v);
#pragma unroll
    for (int j = 0; j < 4; ++j) if (c + j < limit) sum += exp2f((v[j] - mx) * sl2);
  }
  sum = warp_sum(sum);
  const float inv = sum > 0.f ? 1.f / sum : 0.f;
  for (int c = lane * 4; c < cols; c += 128) {
    float v[4], m[4], o[4], od[4];
    ld4<TIn>(srow + c, v);
    dropout_mult4((row * cols + c) >> 2, seed, offset, thresh, inv_keep, m);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      o[j] = (c + j < limit) ? exp2f((v[j] - mx) * sl2) * inv : 0.f;
      od[j] = o[j] * m[j];
    }
    st4_bf16(p + row * cols + c, o);
    st4_bf16(pd + row * cols + c, od);
  }
}

__global__ void __launch_bounds__(256) softmax_dropout_bwd_kernel(const __nv_bfloat16* __restrict__ p,
                                                                  const __nv_bfloat16* __restrict__ dpd,
                                                                  __nv_bfloat16* __restrict__ ds, long long rows, int cols,
                                                                  float scale, uint32_t thresh, float inv_keep, uint64_t seed,
                                                                  uint64_t offset) {
  const long long row = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const __nv_bfloat16* prow = p + row * cols;
  const __nv_bfloat16* grow = dpd + row * cols;
  float dot = 0.f;
  for (int c = lane * 4; c < cols; c += 128) {
    float pv[4], g[4], m[4];
    ld4<__nv_bfloat16>(prow + c, pv);
    ld4<__nv_bfloat16>(grow + c, g);
    dropout_mult4((row * cols + c) >> 2, seed, offset, thresh, inv_keep, m);
#pragma unroll
    for (int j = 0; j < 4; ++j) dot += pv[j] * m[j] * g[j];
  }
  dot = warp_sum(dot);
  for (int c = lane * 4; c < cols; c += 128) {
    float pv[4], g[4], m[4], o[4];
    ld4<__nv_bfloat16>(prow + c, pv);
    ld4<__nv_bfloat16>(grow + c, g);
    dropout_mult4((row * cols + c) >> 2, seed, offset, thresh, inv_keep, m);
#pragma unroll
    for (int j = 0; j < 4; ++j) o[j] = scale * pv[j] * (m[j] * g[j] - dot);
    st4_bf16(ds + row * cols + c, o);
  }
}

static uint32_t drop_threshold(float p) {
  const double t = (double)p * 4294967296.0;
  return t >= 4294967295.0 ? 0xFFFFFFFFu : (uint32_t)t;
}

}  // namespace b200

using namespace b200;

// scores [rows, cols] (dtype 0 = bf16, 1 = fp32; raw q.k, `scale` is applied here), rows = B * H * sq.
// probs / probs_dropped: bf16 [rows, cols].  0 <= p_drop < 1; cols % 4 == 0.
B200_API int b200_softmax_dropout_fwd(const void* scores, void* probs, void* probs_dropped, long long rows, int cols, int sq,
                                      float scale, int causal, float p_drop, unsigned long long seed,
                                      unsigned long long offset, int dtype, cudaStream_t stream) {
  if (rows <= 0) return B200_OK;
  if (cols % 4 || cols <= 0 || sq <= 0 || p_drop < 0.f || p_drop >= 1.f || ((uintptr_t)scores & 15) || ((uintptr_t)probs & 7) ||
      ((uintptr_t)probs_dropped & 7)) {
    b200_set_error("softmax_dropout_fwd: cols %% 4 == 0, 0 <= p < 1 and aligned tensors required (cols=%d p=%f)", cols, p_drop);
    return B200_ERR_SHAPE;
  }
  const unsigned grid = (unsigned)((rows * 32 + 255) / 256);
  const float inv_keep = 1.f / (1.f - p_drop);
  if (dtype == 0)
    softmax_dropout_fwd_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>((const __nv_bfloat16*)scores, (__nv_bfloat16*)probs,
        (__nv_bfloat16*)probs_dropped, rows, cols, sq, scale, causal, drop_threshold(p_drop), inv_keep, seed, offset);
  else if (dtype == 1)
    softmax_dropout_fwd_kernel<float><<<grid, 256, 0, stream>>>((const float*)scores, (__nv_bfloat16*)probs,
        (__nv_bfloat16*)probs_dropped, rows, cols, sq, scale, causal, drop_threshold(p_drop), inv_keep, seed, offset);
  else { b200_set_error("softmax_dropout_fwd: bad dtype %d", dtype); return B200_ERR_UNSUPPORTED; }
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// d_scores (bf16) = scale * P * (m dPd - rowsum(P m dPd)) with the mask of the forward call that used (seed, offset).
B200_API int b200_softmax_dropout_bwd(const void* probs, const void* d_probs_dropped, void* d_scores, long long rows, int cols,
                                      float scale, float p_drop, unsigned long long seed, unsigned long long offset,
                                      cudaStream_t stream) {
  if (rows <= 0) return B200_OK;
  if (cols % 4 || cols <= 0 || p_drop < 0.f || p_drop >= 1.f || ((uintptr_t)probs & 7) || ((uintptr_t)d_probs_dropped & 7) ||
      ((uintptr_t)d_scores & 7)) {
    b200_set_error("softmax_dropout_bwd: cols %% 4 == 0, 0 <= p < 1 and aligned tensors required (cols=%d p=%f)", cols, p_drop);
    return B200_ERR_SHAPE;
  }
  const unsigned grid = (unsigned)((rows * 32 + 255) / 256);
  softmax_dropout_bwd_kernel<<<grid, 256, 0, stream>>>((const __nv_bfloat16*)probs, (const __nv_bfloat16*)d_probs_dropped,
      (__nv_bfloat16*)d_scores, rows, cols, scale, drop_threshold(p_drop), 1.f / (1.f - p_drop), seed, offset);
  B200_LAUNCH_CHECK();
  return B200_OK;
}
