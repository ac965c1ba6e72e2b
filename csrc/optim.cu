// Fused optimizers over flat buffers: AdamW (decoupled weight decay) and SGD with
// momentum.  One pass: read grad (bf16 or fp32, with an optional scale such as
// 1/world for a summed gradient), update fp32 state, write the fp32 master weight
// and, optionally, the bf16 compute copy in the same kernel.
//
// Reference: torch.optim.AdamW(foreach=True)  /root/reference/scripts/06_hybrid_parallelism/01_fsdp_tp_hybrid.py:162
//            torch.optim.SGD(momentum=0.9, weight_decay=1e-5)  /root/reference/scripts/main.py:309-311
#include "vec.cuh"

namespace b200 {

template <typename TP, typename TG>
__global__ void adamw_kernel(TP* __restrict__ p, __nv_bfloat16* __restrict__ p_lowp,
                             const TG* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                             long long n, float lr, float beta1, float beta2, float eps, float wd,
                             float bc1, float bc2_sqrt, float grad_scale) {
  const long long n8 = n >> 3;
  const float step_size = lr / bc1;
  const float decay = 1.f - lr * wd;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8;
       i += (long long)gridDim.x * blockDim.x) {
    float pv[8], gv[8], mv[8], vv[8];
    Vec8<TP>::load(p + i * 8, pv);
    Vec8<TG>::load(g + i * 8, gv);
    Vec8<float>::load(m + i * 8, mv);
    Vec8<float>::load(v + i * 8, vv);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float gj = gv[j] * grad_scale;
      mv[j] = beta1 * mv[j] + (1.f - beta1) * gj;
      vv[j] = beta2 * vv[j] + (1.f - beta2) * gj * gj;
      const float denom = sqrtf(vv[j]) / bc2_sqrt + eps;
      pv[j] = pv[j] * decay - step_size * (mv[j] / denom);
    }
    Vec8<TP>::store(p + i * 8, pv);
    Vec8<float>::store(m + i * 8, mv);
    Vec8<float>::store(v + i * 8, vv);
    if (p_lowp) Vec8<__nv_bfloat16>::store(p_lowp + i * 8, pv);
  }
  if (blockIdx.x == 0) {
    for (long long i = n8 * 8 + threadIdx.x; i < n; i += blockDim.x) {
      const float gj = to_f<TG>(g[i]) * grad_scale;
      float mj = beta1 * m[i] + (1.f - beta1) * gj;
      float vj = beta2 * v[i] + (1.f - beta2) * gj * gj;
      const float denom = sqrtf(vj) / bc2_sqrt + eps;
      const float pj = to_f<TP>(p[i]) * decay - step_size * (mj / denom);
      p[i] = from_f<TP>(pj);
      m[i] = mj; v[i] = vj;
      if (p_lowp) p_lowp[i] = __float2bfloat16_rn(pj);
    }
  }
}

template <typename TP, typename TG>
__global__ void sgd_kernel(TP* __restrict__ p, __nv_bfloat16* __restrict__ p_lowp,
                           const TG* __restrict__ g, float* __restrict__ buf, long long n, float lr,
                           float momentum, float wd, float grad_scale, int first_step) {
  const long long n8 = n >> 3;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8;
       i += (long long)gridDim.x * blockDim.x) {
    float pv[8], gv[8], bv[8];
    Vec8<TP>::load(p + i * 8, pv);
    Vec8<TG>::load(g + i * 8, gv);
    if (buf && !first_step) Vec8<float>::load(buf + i * 8, bv);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float d = gv[j] * grad_scale + wd * pv[j];
      if (buf) { d = first_step ? d : momentum * bv[j] + d; bv[j] = d; }
      pv[j] -= lr * d;
    }
    Vec8<TP>::store(p + i * 8, pv);
    if (buf) Vec8<float>::store(buf + i * 8, bv);
    if (p_lowp) Vec8<__nv_bfloat16>::store(p_lowp + i * 8, pv);
  }
  if (blockIdx.x == 0) {
    for (long long i = n8 * 8 + threadIdx.x; i < n; i += blockDim.x) {
      float pj = to_f<TP>(p[i]);
      float d = to_f<TG>(g[i]) * grad_scale + wd * pj;
      if (buf) { d = first_step ? d : momentum * buf[i] + d; buf[i] = d; }
      pj -= lr * d;
      p[i] = from_f<TP>(pj);
      if (p_lowp) p_lowp[i] = __float2bfloat16_rn(pj);
    }
  }
}

// sum of squares of a flat buffer, atomically added into out[0] (grad-norm building block)
template <typename T>
__global__ void sumsq_kernel(const T* __restrict__ x, float* __restrict__ out, long long n) {
  __shared__ float red[33];
  float s = 0.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    const float v = to_f<T>(x[i]);
    s += v * v;
  }
  s = block_sum(s, red);
  if (threadIdx.x == 0) atomicAdd(out, s);
}

}  // namespace b200

using namespace b200;

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// p: master weights (p_dtype 0 bf16 / 1 fp32), p_lowp: optional bf16 copy written alongside,
// g: gradient (g_dtype), m/v: fp32 state.  bc1 = 1 - beta1^t, bc2 = 1 - beta2^t.
B200_API int b200_adamw(void* p, int p_dtype, void* p_lowp, const void* g, int g_dtype, float* m,
                        float* v, long long n, float lr, float beta1, float beta2, float eps,
                        float wd, float bc1, float bc2, float grad_scale, cudaStream_t stream) {
  if (n <= 0) return B200_OK;
  if (!aligned16(p) || !aligned16(g) || !aligned16(m) || !aligned16(v) || (p_lowp && !aligned16(p_lowp))) {
    b200_set_error("adamw: buffers must be 16-byte aligned");
    return B200_ERR_SHAPE;
  }
  const int grid = ew_grid((n >> 3) ? (n >> 3) : 1, 256);
  const float bc2s = sqrtf(bc2);
#define LAUNCH(TP, TG)                                                                      \
  adamw_kernel<TP, TG><<<grid, 256, 0, stream>>>((TP*)p, (__nv_bfloat16*)p_lowp, (const TG*)g, m, v, \
                                                 n, lr, beta1, beta2, eps, wd, bc1, bc2s, grad_scale)
  if (p_dtype == 1 && g_dtype == 1) LAUNCH(float, float);
  else if (p_dtype == 1 && g_dtype == 0) LAUNCH(float, __nv_bfloat16);
  else if (p_dtype == 0 && g_dtype == 0) LAUNCH(__nv_bfloat16, __nv_bfloat16);
  else if (p_dtype == 0 && g_dtype == 1) LAUNCH(__nv_bfloat16, float);
  else { b200_set_error("adamw: bad dtype codes"); return B200_ERR_UNSUPPORTED; }
#undef LAUNCH
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// AdamW whose step-dependent scalars live in DEVICE memory: scalars[0] = lr, [1] = 1 - beta1^t, [2] = sqrt(1 - beta2^t).
// A CUDA graph that contains this kernel stays valid from step to step: the host refreshes the three floats with a
// stream-ordered copy before each replay instead of baking them into the launch (engine.Trainer(cuda_graph=True)).
template <typename TP, typename TG>
__global__ void adamw_dev_kernel(TP* __restrict__ p, __nv_bfloat16* __restrict__ p_lowp, const TG* __restrict__ g,
                                 float* __restrict__ m, float* __restrict__ v, long long n, float beta1, float beta2,
                                 float eps, float wd, const float* __restrict__ scalars, float grad_scale) {
  const float lr = scalars[0], bc1 = scalars[1], bc2_sqrt = scalars[2];
  const long long n8 = n >> 3;
  const float step_size = lr / bc1;
  const float decay = 1.f - lr * wd;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8;
       i += (long long)gridDim.x * blockDim.x) {
    float pv[8], gv[8], mv[8], vv[8];
    Vec8<TP>::load(p + i * 8, pv);
    Vec8<TG>::load(g + i * 8, gv);
    Vec8<float>::load(m + i * 8, mv);
    Vec8<float>::load(v + i * 8, vv);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float gj = gv[j] * grad_scale;
      mv[j] = beta1 * mv[j] + (1.f - beta1) * gj;
      vv[j] = beta2 * vv[j] + (1.f - beta2) * gj * gj;
      const float denom = sqrtf(vv[j]) / bc2_sqrt + eps;
      pv[j] = pv[j] * decay - step_size * (mv[j] / denom);
    }
    Vec8<TP>::store(p + i * 8, pv);
    Vec8<float>::store(m + i * 8, mv);
    Vec8<float>::store(v + i * 8, vv);
    if (p_lowp) Vec8<__nv_bfloat16>::store(p_lowp + i * 8, pv);
  }
  if (blockIdx.x == 0) {
    for (long long i = n8 * 8 + threadIdx.x; i < n; i += blockDim.x) {
      const float gj = to_f<TG>(g[i]) * grad_scale;
      float mj = beta1 * m[i] + (1.f - beta1) * gj;
      float vj = beta2 * v[i] + (1.f - beta2) * gj * gj;
      const float denom = sqrtf(vj) / bc2_sqrt + eps;
      const float pj = to_f<TP>(p[i]) * decay - step_size * (mj / denom);
      p[i] = from_f<TP>(pj);
      m[i] = mj; v[i] = vj;
      if (p_lowp) p_lowp[i] = __float2bfloat16_rn(pj);
    }
  }
}

B200_API int b200_adamw_dev(void* p, int p_dtype, void* p_lowp, const void* g, int g_dtype, float* m, float* v,
                            long long n, float beta1, float beta2, float eps, float wd, const float* scalars,
                            float grad_scale, cudaStream_t stream) {
  if (n <= 0) return B200_OK;
  if (!aligned16(p) || !aligned16(g) || !aligned16(m) || !aligned16(v) || (p_lowp && !aligned16(p_lowp)) || !scalars) {
    b200_set_error("adamw_dev: buffers must be 16-byte aligned and scalars non-null");
    return B200_ERR_SHAPE;
  }
  const int grid = ew_grid((n >> 3) ? (n >> 3) : 1, 256);
#define LAUNCH(TP, TG)                                                                          \
  adamw_dev_kernel<TP, TG><<<grid, 256, 0, stream>>>((TP*)p, (__nv_bfloat16*)p_lowp, (const TG*)g, m, v, n, beta1, \
                                                     beta2, eps, wd, scalars, grad_scale)
  if (p_dtype == 1 && g_dtype == 1) LAUNCH(float, float);
  else if (p_dtype == 1 && g_dtype == 0) LAUNCH(float, __nv_bfloat16);
  else if (p_dtype == 0 && g_dtype == 0) LAUNCH(__nv_bfloat16, __nv_bfloat16);
  else if (p_dtype == 0 && g_dtype == 1) LAUNCH(__nv_bfloat16, float);
  else { b200_set_error("adamw_dev: bad dtype codes"); return B200_ERR_UNSUPPORTED; }
#undef LAUNCH
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// torch.optim.SGD semantics: d = g + wd*p; buf = d on the first step else momentum*buf + d; p -= lr*buf.
B200_API int b200_sgd(void* p, int p_dtype, void* p_lowp, const void* g, int g_dtype, float* buf,
                      long long n, float lr, float momentum, float wd, float grad_scale,
                      int first_step, cudaStream_t stream) {
  if (n <= 0) return B200_OK;
  if (!aligned16(p) || !aligned16(g) || (buf && !aligned16(buf)) || (p_lowp && !aligned16(p_lowp))) {
    b200_set_error("sgd: buffers must be 16-byte aligned");
    return B200_ERR_SHAPE;
  }
  const int grid = ew_grid((n >> 3) ? (n >> 3) : 1, 256);
#define LAUNCH(TP, TG)                                                                    \
  sgd_kernel<TP, TG><<<grid, 256, 0, stream>>>((TP*)p, (__nv_bfloat16*)p_lowp, (const TG*)g, buf, n, \
                                               lr, momentum, wd, grad_scale, first_step)
  if (p_dtype == 1 && g_dtype == 1) LAUNCH(float, float);
  else if (p_dtype == 1 && g_dtype == 0) LAUNCH(float, __nv_bfloat16);
  else if (p_dtype == 0 && g_dtype == 0) LAUNCH(__nv_bfloat16, __nv_bfloat16);
  else if (p_dtype == 0 && g_dtype == 1) LAUNCH(__nv_bfloat16, float);
  else { b200_set_error("sgd: bad dtype codes"); return B200_ERR_UNSUPPORTED; }
#undef LAUNCH
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_sumsq(const void* x, float* out, long long n, int dtype, cudaStream_t stream) {
  if (n <= 0) return B200_OK;
  B200_DISPATCH_DTYPE(dtype, T, (sumsq_kernel<T><<<ew_grid(n, 256, 4), 256, 0, stream>>>((const T*)x, out, n)));
  B200_LAUNCH_CHECK();
  return B200_OK;
}
