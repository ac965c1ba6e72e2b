// bf16 GEMM on 5th-gen tensor cores: TMA -> shared (128B swizzle) -> tcgen05.mma
// (accumulators in TMEM) -> tcgen05.ld epilogue.  Persistent, warp-specialised:
//   warp 0 : TMA producer (one elected lane) + TMEM allocator
//   warp 1 : MMA issuer   (one elected lane, leader CTA only in cta_group::2)
//   warp 2-5: epilogue    (TMEM -> registers -> fused epilogue -> global)
// Three pipelines: smem full/empty ring, double-buffered TMEM accumulators, and
// a static round-robin tile scheduler with grouped rasterisation for L2 reuse.
//
// This kernel replaces every cuBLAS call the reference makes through nn.Linear /
// torch.matmul:
//   forward  Y = X W^T        /root/reference/fsdp_tp/llama2_model.py:207,228,267,447
//   dgrad    dX = dY W        (autograd of the above)
//   wgrad    dW = dY^T X      (autograd of the above)
// All three operand-major combinations run on the same kernel through the UMMA
// descriptor major-ness bits, so no transposes are ever materialised.
#include "common.cuh"

namespace b200 {

struct GemmParams {
  int M, N, K;
  void* D;
  long long ldd;
  const __nv_bfloat16* bias;      // [N] or nullptr
  const __nv_bfloat16* residual;  // [M, N] (row stride ldr) or nullptr
  long long ldr;
  float alpha;
  int act;         // 0 none, 1 gelu(erf), 2 relu, 3 silu
  int accumulate;  // D += result (read-modify-write)
  int num_m_blocks, num_n_blocks, num_k_blocks, num_tiles;
};

constexpr int kBlockM = 128;   // rows per CTA (UMMA M = 128 * CTA_GROUP)
constexpr int kBlockK = 64;    // 64 bf16 = one 128-byte swizzle row
constexpr int kUmmaK = 16;
constexpr int kGroupM = 8;     // rasterisation group
constexpr int kGemmThreads = 192;
constexpr int kSmemBudget = 232448 - 2048;  // 227 KB minus alignment slack and barriers

template <int CG, int BN>
struct GemmCfg {
  static constexpr int kLoadN = BN / CG;  // B rows loaded by each CTA
  static constexpr int kABytes = kBlockM * kBlockK * 2;
  static constexpr int kBBytes = kLoadN * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStagesRaw = kSmemBudget / kStageBytes;
  static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
  static constexpr int kTmemCols = 2 * BN;  // double-buffered accumulators
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256;
};

__device__ __forceinline__ float apply_act(float v, int act) {
  if (act == 1) return 0.5f * v * (1.f + erff(v * 0.70710678118654752f));
  if (act == 2) return fmaxf(v, 0.f);
  if (act == 3) return v / (1.f + __expf(-v));
  return v;
}

__device__ __forceinline__ void tile_coords(int tile, int num_m, int num_n, int& m_blk, int& n_blk) {
  const int per_group = kGroupM * num_n;
  const int group = tile / per_group;
  const int first_m = group * kGroupM;
  const int gsz = min(num_m - first_m, kGroupM);
  const int in_group = tile - group * per_group;
  m_blk = first_m + in_group % gsz;
  n_blk = in_group / gsz;
}

template <int CG, int BN, bool A_MN, bool B_MN, typename OutT>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a,
                    const __grid_constant__ CUtensorMap tmap_b, const GemmParams p) {
  using Cfg = GemmCfg<CG, BN>;
  constexpr int kStages = Cfg::kStages;
  constexpr int kLoadN = Cfg::kLoadN;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar_base = smem_base + kStages * Cfg::kStageBytes;
  // barrier layout (8 bytes each): full[S] | empty[S] | tmem_full[2] | tmem_empty[2] | tmem_ptr
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kStages + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * kStages + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * kStages + 2 + s); };
  const uint32_t tmem_ptr_addr = bar_base + 8u * (2 * kStages + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta_rank = (CG == 2) ? cluster_ctarank() : 0u;
  const bool is_leader = cta_rank == 0;
  const int cluster_id = blockIdx.x / CG;
  const int num_clusters = gridDim.x / CG;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), 4 * CG);  // one arrival per epilogue warp per CTA
    }
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc<CG>(tmem_ptr_addr, Cfg::kTmemCols);
    tmem_relinquish<CG>();
  }
  tc_fence_before();
  if constexpr (CG == 2) cluster_sync(); else __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_ptr_addr));

  if (warp == 0) {
    if (lane == 0) {
      // ===================== TMA producer =====================
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t full0 = (CG == 2) ? mapa(full_bar(0), 0) : full_bar(0);
      for (int tile = cluster_id; tile < p.num_tiles; tile += num_clusters) {
        int m_blk, n_blk;
        tile_coords(tile, p.num_m_blocks, p.num_n_blocks, m_blk, n_blk);
        const int m0 = (m_blk * CG + (int)cta_rank) * kBlockM;
        const int n0 = n_blk * BN + (int)cta_rank * kLoadN;
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t sa = smem_base + stage * Cfg::kStageBytes;
          const uint32_t sb = sa + Cfg::kABytes;
          const uint32_t fb = full0 + 8u * stage;
          if (is_leader) mbar_expect_tx(full_bar(stage), Cfg::kStageBytes * CG);
          const int k0 = kb * kBlockK;
          if constexpr (!A_MN) {
            if constexpr (CG == 2) tma_load_2d_2sm(sa, &tmap_a, fb, k0, m0);
            else tma_load_2d(sa, &tmap_a, fb, k0, m0);
          } else {
#pragma unroll
            for (int j = 0; j < kBlockM / 64; ++j) {
              if constexpr (CG == 2) tma_load_2d_2sm(sa + j * 8192, &tmap_a, fb, m0 + j * 64, k0);
              else tma_load_2d(sa + j * 8192, &tmap_a, fb, m0 + j * 64, k0);
            }
          }
          if constexpr (!B_MN) {
            if constexpr (CG == 2) tma_load_2d_2sm(sb, &tmap_b, fb, k0, n0);
            else tma_load_2d(sb, &tmap_b, fb, k0, n0);
          } else {
#pragma unroll
            for (int j = 0; j < kLoadN / 64; ++j) {
              if constexpr (CG == 2) tma_load_2d_2sm(sb + j * 8192, &tmap_b, fb, n0 + j * 64, k0);
              else tma_load_2d(sb + j * 8192, &tmap_b, fb, n0 + j * 64, k0);
            }
          }
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && is_leader) {
      // ===================== MMA issuer =====================
      constexpr uint32_t idesc = umma_idesc_bf16(kBlockM * CG, BN, A_MN, B_MN);
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int tile = cluster_id; tile < p.num_tiles; tile += num_clusters) {
        mbar_wait(tempty_bar(as), aphase ^ 1u);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * BN;
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sa = smem_base + stage * Cfg::kStageBytes;
          const uint32_t sb = sa + Cfg::kABytes;
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k) {
            const uint64_t adesc = A_MN ? umma_smem_desc_sw128(sa + k * kUmmaK * 128, kBlockK * 128, 1024)
                                        : umma_smem_desc_sw128(sa + k * kUmmaK * 2, 16, 1024);
            const uint64_t bdesc = B_MN ? umma_smem_desc_sw128(sb + k * kUmmaK * 128, kBlockK * 128, 1024)
                                        : umma_smem_desc_sw128(sb + k * kUmmaK * 2, 16, 1024);
            umma_f16_ss<CG>(tmem_d, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          // frees this smem stage (in both CTAs for cta_group::2) once the MMAs retire
          if constexpr (CG == 2) umma_commit_2sm(empty_bar(stage), 0x3);
          else umma_commit(empty_bar(stage));
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
        if constexpr (CG == 2) umma_commit_2sm(tfull_bar(as), 0x3);
        else umma_commit(tfull_bar(as));
        if (++as == 2) { as = 0; aphase ^= 1u; }
      }
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const int quarter = warp & 3;  // TMEM lane quarter this warp may access
    int as = 0;
    uint32_t aphase = 0;
    const uint32_t tempty_remote = (CG == 2) ? mapa(tempty_bar(0), 0) : tempty_bar(0);
    for (int tile = cluster_id; tile < p.num_tiles; tile += num_clusters) {
      int m_blk, n_blk;
      tile_coords(tile, p.num_m_blocks, p.num_n_blocks, m_blk, n_blk);
      const int row = (m_blk * CG + (int)cta_rank) * kBlockM + quarter * 32 + lane;
      const int n_tile0 = n_blk * BN;
      mbar_wait(tfull_bar(as), aphase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + as * BN;
      const bool row_ok = row < p.M;
      OutT* drow = reinterpret_cast<OutT*>(p.D) + (long long)row * p.ldd;
      const __nv_bfloat16* rrow = p.residual ? p.residual + (long long)row * p.ldr : nullptr;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        const int n0 = n_tile0 + c * 32;
        if (n0 >= p.N) break;  // warp-uniform
        uint32_t r[32];
        tmem_ld_32x32(taddr + c * 32, r);
        tmem_ld_wait();
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * p.alpha;
        if (p.bias) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            if (n0 + g * 8 < p.N) {
              const uint4 b = ld_v4(p.bias + n0 + g * 8);
              v[g * 8 + 0] += bf16lo(b.x); v[g * 8 + 1] += bf16hi(b.x);
              v[g * 8 + 2] += bf16lo(b.y); v[g * 8 + 3] += bf16hi(b.y);
              v[g * 8 + 4] += bf16lo(b.z); v[g * 8 + 5] += bf16hi(b.z);
              v[g * 8 + 6] += bf16lo(b.w); v[g * 8 + 7] += bf16hi(b.w);
            }
          }
        }
        if (p.act != 0) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = apply_act(v[j], p.act);
        }
        if (row_ok) {
          if (rrow) {
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              if (n0 + g * 8 < p.N) {
                const uint4 b = ld_v4(rrow + n0 + g * 8);
                v[g * 8 + 0] += bf16lo(b.x); v[g * 8 + 1] += bf16hi(b.x);
                v[g * 8 + 2] += bf16lo(b.y); v[g * 8 + 3] += bf16hi(b.y);
                v[g * 8 + 4] += bf16lo(b.z); v[g * 8 + 5] += bf16hi(b.z);
                v[g * 8 + 6] += bf16lo(b.w); v[g * 8 + 7] += bf16hi(b.w);
              }
            }
          }
          if constexpr (sizeof(OutT) == 2) {
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              if (n0 + g * 8 < p.N) {
                OutT* dst = drow + n0 + g * 8;
                if (p.accumulate) {
                  const uint4 b = ld_v4(dst);
                  v[g * 8 + 0] += bf16lo(b.x); v[g * 8 + 1] += bf16hi(b.x);
                  v[g * 8 + 2] += bf16lo(b.y); v[g * 8 + 3] += bf16hi(b.y);
                  v[g * 8 + 4] += bf16lo(b.z); v[g * 8 + 5] += bf16hi(b.z);
                  v[g * 8 + 6] += bf16lo(b.w); v[g * 8 + 7] += bf16hi(b.w);
                }
                uint4 o;
                o.x = pack_bf16(v[g * 8 + 0], v[g * 8 + 1]);
                o.y = pack_bf16(v[g * 8 + 2], v[g * 8 + 3]);
                o.z = pack_bf16(v[g * 8 + 4], v[g * 8 + 5]);
                o.w = pack_bf16(v[g * 8 + 6], v[g * 8 + 7]);
                st_v4(dst, o);
              }
            }
          } else {
#pragma unroll
            for (int g = 0; g < 8; ++g) {
              if (n0 + g * 4 < p.N) {
                OutT* dst = drow + n0 + g * 4;
                float4 o = make_float4(v[g * 4 + 0], v[g * 4 + 1], v[g * 4 + 2], v[g * 4 + 3]);
                if (p.accumulate) {
                  const float4 b = *reinterpret_cast<const float4*>(dst);
                  o.x += b.x; o.y += b.y; o.z += b.z; o.w += b.w;
                }
                *reinterpret_cast<float4*>(dst) = o;
              }
            }
          }
        }
      }
      // hand the accumulator stage back to the MMA issuer
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (CG == 2) mbar_arrive_cluster(tempty_remote + 8u * as);
        else mbar_arrive(tempty_bar(as));
      }
      if (++as == 2) { as = 0; aphase ^= 1u; }
    }
  }

  // ---------------- teardown ----------------
  tc_fence_before();
  if constexpr (CG == 2) { __syncthreads(); cluster_sync(); } else __syncthreads();
  if (warp == 0) {
    __syncwarp();
    tc_fence_after();
    tmem_dealloc<CG>(tmem_base, Cfg::kTmemCols);
  }
}

// ---------------------------------------------------------------------------
// Small / unaligned shapes: plain SIMT tiled GEMM (fp32 accumulate).  Used only
// when TMA alignment rules cannot be met (toy models such as Linear(10,64)).
// ---------------------------------------------------------------------------
template <typename OutT>
__global__ void gemm_simt_kernel(const __nv_bfloat16* __restrict__ A, const __nv_bfloat16* __restrict__ B,
                                 OutT* __restrict__ D, int M, int N, int K, long long lda,
                                 long long ldb, long long ldd, int a_mn, int b_mn,
                                 const __nv_bfloat16* __restrict__ bias,
                                 const __nv_bfloat16* __restrict__ residual, long long ldr,
                                 float alpha, int act, int accumulate) {
  __shared__ float As[16][17];
  __shared__ float Bs[16][17];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int row = blockIdx.y * 16 + ty, col = blockIdx.x * 16 + tx;
  float acc = 0.f;
  for (int k0 = 0; k0 < K; k0 += 16) {
    const int ka = k0 + tx, kb = k0 + ty;
    As[ty][tx] = (row < M && ka < K)
                     ? __bfloat162float(a_mn ? A[(long long)ka * lda + row] : A[(long long)row * lda + ka])
                     : 0.f;
    const int bcol = blockIdx.x * 16 + tx;
    Bs[ty][tx] = (bcol < N && kb < K)
                     ? __bfloat162float(b_mn ? B[(long long)kb * ldb + bcol] : B[(long long)bcol * ldb + kb])
                     : 0.f;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k) acc += As[ty][k] * Bs[k][tx];
    __syncthreads();
  }
  if (row < M && col < N) {
    float v = acc * alpha;
    if (bias) v += __bfloat162float(bias[col]);
    v = apply_act(v, act);
    if (residual) v += __bfloat162float(residual[(long long)row * ldr + col]);
    OutT* dst = D + (long long)row * ldd + col;
    if (accumulate) v += static_cast<float>(*dst);
    *dst = static_cast<OutT>(v);
  }
}

template <int CG, int BN, bool A_MN, bool B_MN, typename OutT>
static int launch_gemm(const void* A, const void* B, const GemmParams& p0, long long lda,
                       long long ldb, int max_ctas, cudaStream_t stream) {
  using Cfg = GemmCfg<CG, BN>;
  GemmParams p = p0;
  p.num_m_blocks = ceil_div(p.M, kBlockM * CG);
  p.num_n_blocks = ceil_div(p.N, BN);
  p.num_k_blocks = ceil_div(p.K, kBlockK);
  p.num_tiles = p.num_m_blocks * p.num_n_blocks;

  CUtensorMap ta, tb;
  int rc;
  if (!A_MN) {
    const uint64_t dims[2] = {(uint64_t)p.K, (uint64_t)p.M};
    const uint64_t str[1] = {(uint64_t)lda * 2};
    const uint32_t box[2] = {(uint32_t)kBlockK, (uint32_t)kBlockM};
    rc = b200_make_tmap(&ta, A, 2, 2, dims, str, box, 1);
  } else {
    const uint64_t dims[2] = {(uint64_t)p.M, (uint64_t)p.K};
    const uint64_t str[1] = {(uint64_t)lda * 2};
    const uint32_t box[2] = {64u, (uint32_t)kBlockK};
    rc = b200_make_tmap(&ta, A, 2, 2, dims, str, box, 1);
  }
  if (rc) return rc;
  if (!B_MN) {
    const uint64_t dims[2] = {(uint64_t)p.K, (uint64_t)p.N};
    const uint64_t str[1] = {(uint64_t)ldb * 2};
    const uint32_t box[2] = {(uint32_t)kBlockK, (uint32_t)Cfg::kLoadN};
    rc = b200_make_tmap(&tb, B, 2, 2, dims, str, box, 1);
  } else {
    const uint64_t dims[2] = {(uint64_t)p.N, (uint64_t)p.K};
    const uint64_t str[1] = {(uint64_t)ldb * 2};
    const uint32_t box[2] = {64u, (uint32_t)kBlockK};
    rc = b200_make_tmap(&tb, B, 2, 2, dims, str, box, 1);
  }
  if (rc) return rc;

  auto kern = gemm_tcgen05_kernel<CG, BN, A_MN, B_MN, OutT>;
  static bool attr_set = false;  // one per template instantiation
  if (!attr_set) {
    B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Cfg::kSmemBytes));
    attr_set = true;
  }
  int sms = b200_sm_count();
  if (max_ctas > 0 && max_ctas < sms) sms = max_ctas;
  int clusters = sms / CG;
  if (clusters > p.num_tiles) clusters = p.num_tiles;
  if (clusters < 1) clusters = 1;

  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(clusters * CG);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  B200_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, ta, tb, p));
  return B200_OK;
}

template <int CG, int BN, typename OutT>
static int dispatch_major(int a_mn, int b_mn, const void* A, const void* B, const GemmParams& p,
                          long long lda, long long ldb, int max_ctas, cudaStream_t s) {
  if (!a_mn && !b_mn) return launch_gemm<CG, BN, false, false, OutT>(A, B, p, lda, ldb, max_ctas, s);
  if (!a_mn && b_mn) return launch_gemm<CG, BN, false, true, OutT>(A, B, p, lda, ldb, max_ctas, s);
  if (a_mn && b_mn) return launch_gemm<CG, BN, true, true, OutT>(A, B, p, lda, ldb, max_ctas, s);
  return launch_gemm<CG, BN, true, false, OutT>(A, B, p, lda, ldb, max_ctas, s);
}

template <typename OutT>
static int dispatch_cfg(int config, int a_mn, int b_mn, const void* A, const void* B,
                        const GemmParams& p, long long lda, long long ldb, int max_ctas,
                        cudaStream_t s) {
  switch (config) {
    case 1: return dispatch_major<1, 128, OutT>(a_mn, b_mn, A, B, p, lda, ldb, max_ctas, s);
    case 2: return dispatch_major<1, 256, OutT>(a_mn, b_mn, A, B, p, lda, ldb, max_ctas, s);
    default: return dispatch_major<2, 256, OutT>(a_mn, b_mn, A, B, p, lda, ldb, max_ctas, s);
  }
}

// Wave-quantisation-aware choice between the three tile configurations.
static int pick_config(int M, int N, int sms) {
  auto eff = [&](int tm, int tn, int units) {
    const long long tiles = (long long)ceil_div(M, tm) * ceil_div(N, tn);
    const long long waves = (tiles + units - 1) / units;
    const double useful = (double)M * N / ((double)tiles * tm * tn);
    return useful * (double)tiles / (double)(waves * units);
  };
  const double e3 = eff(256, 256, sms / 2) * 1.00;   // 2-CTA 256x256: best per-tile throughput
  const double e2 = eff(128, 256, sms) * 0.93;
  const double e1 = eff(128, 128, sms) * 0.85;
  if (e3 >= e2 && e3 >= e1) return 3;
  return e2 >= e1 ? 2 : 1;
}

}  // namespace b200

// D[M,N] = act(alpha * A.B^T + bias) + residual (+ D)         (fp32 accumulation)
//   A logical [M,K]: a_mn_major=0 -> stored row-major [M,K] (row stride lda)
//                    a_mn_major=1 -> stored row-major [K,M] (row stride lda)
//   B logical [N,K]: b_mn_major=0 -> stored row-major [N,K] (row stride ldb)
//                    b_mn_major=1 -> stored row-major [K,N] (row stride ldb)
//   config: 0 auto, 1 = 1-CTA 128x128, 2 = 1-CTA 128x256, 3 = 2-CTA 256x256, 9 = SIMT fallback
B200_API int b200_gemm_bf16(const void* A, const void* B, void* D, int M, int N, int K,
                            long long lda, long long ldb, long long ldd, int a_mn_major,
                            int b_mn_major, int out_fp32, const void* bias, const void* residual,
                            long long ldr, float alpha, int act, int accumulate, int config,
                            int max_ctas, cudaStream_t stream) {
  using namespace b200;
  if (M <= 0 || N <= 0) return B200_OK;
  if (K <= 0) {
    b200_set_error("b200_gemm_bf16: K must be positive");
    return B200_ERR_SHAPE;
  }
  GemmParams p = {};
  p.M = M; p.N = N; p.K = K;
  p.D = D; p.ldd = ldd;
  p.bias = reinterpret_cast<const __nv_bfloat16*>(bias);
  p.residual = reinterpret_cast<const __nv_bfloat16*>(residual);
  p.ldr = ldr;
  p.alpha = alpha; p.act = act; p.accumulate = accumulate;

  auto aligned16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  const bool tma_ok = aligned16(A) && aligned16(B) && aligned16(D) && (lda % 8 == 0) &&
                      (ldb % 8 == 0) && (ldd % 8 == 0) && (N % 8 == 0) && (K % 8 == 0) &&
                      (M % 8 == 0 || !a_mn_major) && (!bias || aligned16(bias)) &&
                      (!residual || (aligned16(residual) && ldr % 8 == 0));
  if (config == 9 || !tma_ok) {
    dim3 blk(16, 16), grd(ceil_div(N, 16), ceil_div(M, 16));
    if (out_fp32)
      gemm_simt_kernel<float><<<grd, blk, 0, stream>>>(
          (const __nv_bfloat16*)A, (const __nv_bfloat16*)B, (float*)D, M, N, K, lda, ldb, ldd,
          a_mn_major, b_mn_major, p.bias, p.residual, ldr, alpha, act, accumulate);
    else
      gemm_simt_kernel<__nv_bfloat16><<<grd, blk, 0, stream>>>(
          (const __nv_bfloat16*)A, (const __nv_bfloat16*)B, (__nv_bfloat16*)D, M, N, K, lda, ldb,
          ldd, a_mn_major, b_mn_major, p.bias, p.residual, ldr, alpha, act, accumulate);
    B200_LAUNCH_CHECK();
    return B200_OK;
  }
  if (config == 0) config = pick_config(M, N, (max_ctas > 0 && max_ctas < b200_sm_count()) ? max_ctas : b200_sm_count());
  if (out_fp32)
    return dispatch_cfg<float>(config, a_mn_major, b_mn_major, A, B, p, lda, ldb, max_ctas, stream);
  return dispatch_cfg<__nv_bfloat16>(config, a_mn_major, b_mn_major, A, B, p, lda, ldb, max_ctas,
                                     stream);
}
