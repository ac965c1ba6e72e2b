// Flash-attention backward on tcgen05, as two deterministic kernels (no atomics).  Both stream the non-resident
// dimension in 64-wide steps with S / dP DOUBLE-BUFFERED in TMEM, so the tensor cores compute step j+1 while the CUDA
// cores turn step j's scores into P / dS (the TMEM -> register read of S and dP is the binding resource):
//   dkv kernel: one CTA per (batch, kv-head, 128-key block); loops over query blocks (and over
//               the query heads of a GQA group) accumulating dK and dV in TMEM.
//               S^T = K Q^T, dP^T = V dO^T, dV += P^T dO, dK += dS^T Q
//   dq  kernel: one CTA per (batch, head, 128-query block); loops over key blocks.
//               S = Q K^T, dP = dO V^T, dQ += dS K
// P / dS are written as packed bf16 back into the TMEM columns of S / dP and consumed by
// TS-MMAs (A operand from tensor memory).  The same shared-memory tile serves as a K-major
// operand in one MMA and as an MN-major operand in another, so nothing is transposed.
// A small pre-pass computes delta = rowsum(dO * O).
//
// Replaces autograd of F.scaled_dot_product_attention (/root/reference/fsdp_tp/llama2_model.py:223).
#include "attn_common.cuh"

namespace b200 {

constexpr int kFbThreads = 320;  // loader + MMA warps, 8 compute warps (two per TMEM lane quarter, 64 columns each)
constexpr int kFbBlock = 128;

struct AttnBwdParams {
  AttnParams a;                       // Q/K/V/O geometry (a.O unused here)
  long long do_sb, do_ss, do_sh;      // dO strides
  long long dq_sb, dq_ss, dq_sh;
  long long dk_sb, dk_ss, dk_sh;
  long long dv_sb, dv_ss, dv_sh;
  const float* LSE;                   // [B, H, Sq]
  const float* Delta;                 // [B, H, Sq]
  void* dQ; void* dK; void* dV;       // bf16
};

constexpr int kFbSub = 64;      // width of the streamed dimension per pipeline step (keys in dq, queries in dkv)
constexpr int kFbStages = 3;    // smem ring depth for the streamed (sub-block) tiles

template <int HD>
struct FbCfg {
  static constexpr int kHalves = HD / 64;
  static constexpr int kTile = kFbBlock * HD * 2;       // resident 128-row tile
  static constexpr int kSubTile = kFbSub * HD * 2;      // streamed 64-row tile
  // 2 resident tiles + ring of (2 streamed tiles) + LSE/delta staging + barriers
  static constexpr int kSmemBytes = 2 * kTile + kFbStages * 2 * kSubTile + 2048 + 1024 + 256;
};

// delta[b,h,q] = sum_d dO[b,q,h,d] * O[b,q,h,d]    (one warp per row)
__global__ void attn_delta_kernel(const __nv_bfloat16* __restrict__ O, const __nv_bfloat16* __restrict__ dO,
                                  float* __restrict__ delta, int B, int H, int Sq, int hd,
                                  long long o_sb, long long o_ss, long long o_sh,
                                  long long do_sb, long long do_ss, long long do_sh) {
  const long long row = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const long long total = (long long)B * H * Sq;
  if (row >= total) return;
  const int q = (int)(row % Sq);
  const int h = (int)((row / Sq) % H);
  const int b = (int)(row / ((long long)Sq * H));
  const __nv_bfloat16* o = O + b * o_sb + q * o_ss + h * o_sh;
  const __nv_bfloat16* g = dO + b * do_sb + q * do_ss + h * do_sh;
  float s = 0.f;
  for (int c = lane * 8; c < hd; c += 256) {
    const uint4 a = ld_v4(o + c), d = ld_v4(g + c);
    s += bf16lo(a.x) * bf16lo(d.x) + bf16hi(a.x) * bf16hi(d.x) + bf16lo(a.y) * bf16lo(d.y) +
         bf16hi(a.y) * bf16hi(d.y) + bf16lo(a.z) * bf16lo(d.z) + bf16hi(a.z) * bf16hi(d.z) +
         bf16lo(a.w) * bf16lo(d.w) + bf16hi(a.w) * bf16hi(d.w);
  }
  s = warp_sum(s);
  if (lane == 0) delta[row] = s;
}

// Store one TMEM accumulator row chunk (32 fp32 columns) as bf16 with a scale.
__device__ __forceinline__ void store_chunk_bf16(__nv_bfloat16* dst, const uint32_t (&r)[32], float sc) {
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    uint4 o;
    o.x = pack_bf16(__uint_as_float(r[g * 8 + 0]) * sc, __uint_as_float(r[g * 8 + 1]) * sc);
    o.y = pack_bf16(__uint_as_float(r[g * 8 + 2]) * sc, __uint_as_float(r[g * 8 + 3]) * sc);
    o.z = pack_bf16(__uint_as_float(r[g * 8 + 4]) * sc, __uint_as_float(r[g * 8 + 5]) * sc);
    o.w = pack_bf16(__uint_as_float(r[g * 8 + 6]) * sc, __uint_as_float(r[g * 8 + 7]) * sc);
    st_v4(dst + g * 8, o);
  }
}

// ============================== dK / dV ======================================
// CTA = (batch, kv head, 128 keys).  TMEM: dV [0,128) | dK [128,256) | {S^T_b [64] , dP^T_b [64]} x 2 buffers.
// P^T / dS^T (bf16) are written back over the columns each thread has just read (first 16 of its 32), so the TS-MMA
// reads its A operand at tmem + {0, 8, 32, 40} for the four 16-query K steps.
template <int HD>
__global__ void __launch_bounds__(kFbThreads, 1)
flash_attn_bwd_dkv_kernel(const __grid_constant__ CUtensorMap tmap_q,
                          const __grid_constant__ CUtensorMap tmap_k,
                          const __grid_constant__ CUtensorMap tmap_v,
                          const __grid_constant__ CUtensorMap tmap_do, const AttnBwdParams p) {
  using Cfg = FbCfg<HD>;
  constexpr int kTile = Cfg::kTile, kSub = Cfg::kSubTile;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sK = smem_base, sV = smem_base + kTile;
  auto sQ = [&](int s) { return smem_base + 2 * kTile + (2 * s) * kSub; };
  auto sDO = [&](int s) { return smem_base + 2 * kTile + (2 * s + 1) * kSub; };
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  float* s_lse = reinterpret_cast<float*>(smem_gen + 2 * kTile + kFbStages * 2 * kSub);  // [2][64]
  float* s_delta = s_lse + 128;                                                            // [2][64]
  const uint32_t bar = smem_base + 2 * kTile + kFbStages * 2 * kSub + 2048;
  const uint32_t kv_full = bar;
  auto qdo_full = [&](int s) { return bar + 8u * (1 + s); };
  auto qdo_empty = [&](int s) { return bar + 8u * (1 + kFbStages + s); };
  auto sdp_full = [&](int b_) { return bar + 8u * (1 + 2 * kFbStages + b_); };
  auto pds_ready = [&](int b_) { return bar + 8u * (3 + 2 * kFbStages + b_); };
  const uint32_t acc_done = bar + 8u * (5 + 2 * kFbStages);
  const uint32_t tmem_ptr_addr = bar + 8u * (6 + 2 * kFbStages);

  const AttnParams& a = p.a;
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // warp index: provably uniform
  int kv_blk, hk, b;                                         // early key blocks see the most queries: already heavy first
  lpt_decode((int)gridDim.x, 8, kv_blk, hk, b);
  const int kv0 = kv_blk * kFbBlock;
  const int G = a.H / a.Hkv;
  const int nq = (a.Sq + kFbSub - 1) / kFbSub;              // 64-query steps
  int i_start = 0;
  if (a.causal) {
    const int d = kv0 + a.k_offset - a.q_offset;            // first query position that can see key kv0
    i_start = d <= 0 ? 0 : d / kFbSub;
  }
  const int per_head = nq > i_start ? nq - i_start : 0;
  const int n_iters = per_head * G;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmap_q); tma_prefetch_desc(&tmap_k);
    tma_prefetch_desc(&tmap_v); tma_prefetch_desc(&tmap_do);
    mbar_init(kv_full, 1);
    for (int s = 0; s < kFbStages; ++s) { mbar_init(qdo_full(s), 1); mbar_init(qdo_empty(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(sdp_full(s), 1); mbar_init(pds_ready(s), 8); }
    mbar_init(acc_done, 1);
    fence_mbar_init();
  }
  if (warp == 0) { tmem_alloc<1>(tmem_ptr_addr, 512); tmem_relinquish<1>(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_ptr_addr));
  const uint32_t tDV = tmem_base, tDK = tmem_base + 128;
  auto tS = [&](int b_) { return tmem_base + 256 + b_ * 128; };
  auto tDP = [&](int b_) { return tmem_base + 256 + b_ * 128 + 64; };

  if (warp == 0) {
    if (lane == 0 && n_iters > 0) {
      mbar_expect_tx(kv_full, 2 * kTile);
#pragma unroll
      for (int hh = 0; hh < Cfg::kHalves; ++hh) {
        tma_load_4d(sK + hh * 16384, &tmap_k, kv_full, hh * 64, kv0, hk, b, kEvictFirst);
        tma_load_4d(sV + hh * 16384, &tmap_v, kv_full, hh * 64, kv0, hk, b, kEvictFirst);
      }
      for (int it = 0; it < n_iters; ++it) {
        const int st = it % kFbStages;
        const int h = hk * G + it / per_head;
        const int q0 = (i_start + it % per_head) * kFbSub;
        mbar_wait(qdo_empty(st), ((it / kFbStages) & 1) ^ 1u);
        mbar_expect_tx(qdo_full(st), 2 * kSub);
#pragma unroll
        for (int hh = 0; hh < Cfg::kHalves; ++hh) {
          tma_load_4d(sQ(st) + hh * 8192, &tmap_q, qdo_full(st), hh * 64, q0, h, b, kEvictLast);
          tma_load_4d(sDO(st) + hh * 8192, &tmap_do, qdo_full(st), hh * 64, q0, h, b, kEvictLast);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (n_iters > 0) {
      // Whole warp in the schedule, one elected lane issues; shared-memory descriptors are a 64-bit add away from six
      // per-kernel bases (see elect_one_sync).  A 64-query step issues 24 MMAs worth ~1030 tensor cycles: with ~16
      // instructions of descriptor building per MMA on a single lane the ISSUE was the longer path.
      constexpr uint32_t idesc_s = umma_idesc_bf16(128, kFbSub, false, false);  // [128 keys] x [64 queries], K = hd
      constexpr uint32_t idesc_acc = umma_idesc_bf16(128, HD, false, true);     // TMEM A x MN-major B, K = 64 queries
      const uint64_t dK0 = umma_smem_desc_sw128(sK, 16, 1024), dV0 = umma_smem_desc_sw128(sV, 16, 1024);
      const uint64_t dQk = umma_smem_desc_sw128(sQ(0), 16, 1024), dDOk = umma_smem_desc_sw128(sDO(0), 16, 1024);          // K-major views
      const uint64_t dQm = umma_smem_desc_sw128(sQ(0), 8192, 1024), dDOm = umma_smem_desc_sw128(sDO(0), 8192, 1024);      // MN-major views
      constexpr uint32_t kRingStep = (uint32_t)(2 * kSub) >> 4;   // address-field distance between ring stages
      auto issue_sdp = [&](int it) {
        const int st = it % kFbStages, bb = it & 1;
        mbar_wait(qdo_full(st), (it / kFbStages) & 1);
        tc_fence_after();
        const uint64_t q = dQk + (uint64_t)(st * kRingStep), g = dDOk + (uint64_t)(st * kRingStep);
        const uint32_t ds_ = tS(bb), dp_ = tDP(bb);
        if (elect_one_sync()) {
#pragma unroll
          for (int kk = 0; kk < HD / 16; ++kk) {
            const uint32_t offa = ((kk >> 2) * 16384 + (kk & 3) * 32) >> 4;   // resident 128-row tiles: 16 KB per hd half
            const uint32_t offb = ((kk >> 2) * 8192 + (kk & 3) * 32) >> 4;    // streamed 64-row tiles: 8 KB per hd half
            umma_f16_ss<1>(ds_, dK0 + offa, q + offb, idesc_s, kk > 0 ? 1u : 0u);
          }
#pragma unroll
          for (int kk = 0; kk < HD / 16; ++kk) {
            const uint32_t offa = ((kk >> 2) * 16384 + (kk & 3) * 32) >> 4;
            const uint32_t offb = ((kk >> 2) * 8192 + (kk & 3) * 32) >> 4;
            umma_f16_ss<1>(dp_, dV0 + offa, g + offb, idesc_s, kk > 0 ? 1u : 0u);
          }
          umma_commit(sdp_full(bb));
        }
      };
      mbar_wait(kv_full, 0);
      issue_sdp(0);
      if (n_iters > 1) issue_sdp(1);
      for (int it = 0; it < n_iters; ++it) {
        const int st = it % kFbStages, bb = it & 1;
        mbar_wait(pds_ready(bb), (it >> 1) & 1);
        tc_fence_after();
        const uint64_t qm = dQm + (uint64_t)(st * kRingStep), gm = dDOm + (uint64_t)(st * kRingStep);
        const uint32_t ps = tS(bb), pd = tDP(bb);
        const uint32_t acc0 = it > 0 ? 1u : 0u;
        if (elect_one_sync()) {
#pragma unroll
          for (int kk = 0; kk < kFbSub / 16; ++kk) {
            const uint32_t acol = (kk >> 1) * 32 + (kk & 1) * 8;      // P^T / dS^T live where each half read its scores
            umma_f16_ts<1>(tDV, ps + acol, gm + (uint64_t)(kk * (2048 >> 4)), idesc_acc, kk > 0 ? 1u : acc0);
          }
#pragma unroll
          for (int kk = 0; kk < kFbSub / 16; ++kk) {
            const uint32_t acol = (kk >> 1) * 32 + (kk & 1) * 8;
            umma_f16_ts<1>(tDK, pd + acol, qm + (uint64_t)(kk * (2048 >> 4)), idesc_acc, kk > 0 ? 1u : acc0);
          }
          umma_commit(qdo_empty(st));
        }
        if (it + 2 < n_iters) issue_sdp(it + 2);   // reuses buffer bb: ordered after the two MMAs above
      }
      if (elect_one_sync()) umma_commit(acc_done);
    }
    __syncwarp();
  } else {
    const int quarter = warp & 3;
    const int half = (warp - 2) >> 2;            // which 32 of the 64 query columns this warp handles
    const int row = quarter * 32 + lane;         // key row inside the block
    const int ct = (warp - 2) * 32 + lane;       // 0..255 index among compute threads
    const uint32_t lane_off = static_cast<uint32_t>(quarter * 32) << 16;
    const float sl2 = a.scale * kLog2e;
    const int kv_pos = kv0 + row + a.k_offset - a.q_offset;  // visible to query q iff kv_pos <= q
    // LSE / delta of the 64 queries of a step travel global -> registers -> shared memory one step AHEAD: the loads of
    // step it + 1 are issued right after the barrier of step it and consumed at the top of the next iteration, so their
    // latency hides behind a whole step (ncu: the load-use stall plus the barrier behind it were ~18 % of the samples).
    auto fetch_stats = [&](int it_, float& l_out, float& d_out) {
      const int h_ = hk * G + it_ / per_head;
      const int q = (i_start + it_ % per_head) * kFbSub + ct;
      const long long off = ((long long)b * a.H + h_) * a.Sq + q;
      const float l_ = q < a.Sq ? p.LSE[off] : INFINITY;
      l_out = (l_ == -INFINITY) ? -INFINITY : -l_ * kLog2e;   // NEGATED; empty rows: P = exp2(-inf) = 0
      d_out = q < a.Sq ? -p.Delta[off] : 0.f;                 // NEGATED
    };
    float l_pre = 0.f, d_pre = 0.f;
    if (ct < kFbSub && n_iters > 0) fetch_stats(0, l_pre, d_pre);
    for (int it = 0; it < n_iters; ++it) {
      const int bb = it & 1;
      const int q0 = (i_start + it % per_head) * kFbSub;
      if (ct < kFbSub) {   // buffer bb was last read two steps ago
        s_lse[bb * 64 + ct] = l_pre;
        s_delta[bb * 64 + ct] = d_pre;
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (ct < kFbSub && it + 1 < n_iters) fetch_stats(it + 1, l_pre, d_pre);
      const bool need_mask = a.causal && (q0 + a.q_offset < kv0 + kFbBlock - 1 + a.k_offset);
      mbar_wait(sdp_full(bb), (it >> 1) & 1);
      tc_fence_after();
      const float* lse_s = s_lse + bb * 64 + half * 32;
      const float* del_s = s_delta + bb * 64 + half * 32;
      const uint32_t cs = tS(bb) + lane_off + half * 32, cd = tDP(bb) + lane_off + half * 32;
      uint32_t rs[32], rd[32];
      tmem_ld_32x32(cs, rs);
      tmem_ld_32x32(cd, rd);
      tmem_ld_wait();
      uint32_t pk[16], dk[16];
      const uint64_t sl2_2 = f2_pack(sl2, sl2);
      const uint64_t* lse2p = reinterpret_cast<const uint64_t*>(lse_s);   // (-lse[i], -lse[i+1]) pairs
      const uint64_t* del2p = reinterpret_cast<const uint64_t*>(del_s);   // (-delta[i], -delta[i+1]) pairs
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        float t0, t1, d0, d1;
        f2_unpack(ffma2(f2_pack_u(rs[i], rs[i + 1]), sl2_2, lse2p[i >> 1]), t0, t1);
        float p0 = ex2_approx(t0), p1 = ex2_approx(t1);
        if (need_mask) {
          if (kv_pos > q0 + half * 32 + i) p0 = 0.f;
          if (kv_pos > q0 + half * 32 + i + 1) p1 = 0.f;
        }
        f2_unpack(fmul2(f2_pack(p0, p1), fadd2(f2_pack_u(rd[i], rd[i + 1]), del2p[i >> 1])), d0, d1);
        pk[i >> 1] = pack_bf16(p0, p1);
        dk[i >> 1] = pack_bf16(d0, d1);
      }
      tmem_st_32x16(cs, pk);    // over the first 16 of the 32 columns this thread just read
      tmem_st_32x16(cd, dk);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(pds_ready(bb));
    }
    // ---- epilogue ----
    const int kv = kv0 + row;
    __nv_bfloat16* dkrow = reinterpret_cast<__nv_bfloat16*>(p.dK) + (long long)b * p.dk_sb + (long long)kv * p.dk_ss + (long long)hk * p.dk_sh;
    __nv_bfloat16* dvrow = reinterpret_cast<__nv_bfloat16*>(p.dV) + (long long)b * p.dv_sb + (long long)kv * p.dv_ss + (long long)hk * p.dv_sh;
    if (n_iters > 0) {
      mbar_wait(acc_done, 0);
      tc_fence_after();
#pragma unroll 1
      for (int c = half; c < HD / 32; c += 2) {
        uint32_t r[32];
        tmem_ld_32x32(tDV + lane_off + c * 32, r);
        tmem_ld_wait();
        if (kv < a.Sk) store_chunk_bf16(dvrow + c * 32, r, 1.f);
        tmem_ld_32x32(tDK + lane_off + c * 32, r);
        tmem_ld_wait();
        if (kv < a.Sk) store_chunk_bf16(dkrow + c * 32, r, a.scale);
      }
    } else if (kv < a.Sk && half == 0) {
      for (int c = 0; c < HD / 8; ++c) {
        st_v4(dkrow + c * 8, make_uint4(0, 0, 0, 0));
        st_v4(dvrow + c * 8, make_uint4(0, 0, 0, 0));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { __syncwarp(); tc_fence_after(); tmem_dealloc<1>(tmem_base, 512); }
}

// ============================== dQ ===========================================
// CTA = (batch, head, 128 queries); keys stream in 64-wide steps.  TMEM: dQ [0,128) | {S_b [64], dP_b [64]} x 2.
template <int HD>
__global__ void __launch_bounds__(kFbThreads, 1)
flash_attn_bwd_dq_kernel(const __grid_constant__ CUtensorMap tmap_q,
                         const __grid_constant__ CUtensorMap tmap_k,
                         const __grid_constant__ CUtensorMap tmap_v,
                         const __grid_constant__ CUtensorMap tmap_do, const AttnBwdParams p) {
  using Cfg = FbCfg<HD>;
  constexpr int kTile = Cfg::kTile, kSub = Cfg::kSubTile;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = smem_base, sDO = smem_base + kTile;
  auto sK = [&](int s) { return smem_base + 2 * kTile + (2 * s) * kSub; };
  auto sV = [&](int s) { return smem_base + 2 * kTile + (2 * s + 1) * kSub; };
  const uint32_t bar = smem_base + 2 * kTile + kFbStages * 2 * kSub + 2048;
  const uint32_t qdo_full = bar;
  auto kv_full = [&](int s) { return bar + 8u * (1 + s); };
  auto kv_empty = [&](int s) { return bar + 8u * (1 + kFbStages + s); };
  auto sdp_full = [&](int b_) { return bar + 8u * (1 + 2 * kFbStages + b_); };
  auto ds_ready = [&](int b_) { return bar + 8u * (3 + 2 * kFbStages + b_); };
  const uint32_t acc_done = bar + 8u * (5 + 2 * kFbStages);
  const uint32_t tmem_ptr_addr = bar + 8u * (6 + 2 * kFbStages);

  const AttnParams& a = p.a;
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // warp index: provably uniform
  int lpt_rank, h, b;
  lpt_decode((int)gridDim.x, 8, lpt_rank, h, b);
  const int q_blk = (int)gridDim.x - 1 - lpt_rank;
  const int hk = h / (a.H / a.Hkv);
  const int q0 = q_blk * kFbBlock;
  int kv_limit = a.Sk;
  if (a.causal) {
    const int lim = q0 + kFbBlock + a.q_offset - a.k_offset;
    kv_limit = lim < kv_limit ? lim : kv_limit;
  }
  const int n_blocks = kv_limit > 0 ? (kv_limit + kFbSub - 1) / kFbSub : 0;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmap_q); tma_prefetch_desc(&tmap_k);
    tma_prefetch_desc(&tmap_v); tma_prefetch_desc(&tmap_do);
    mbar_init(qdo_full, 1);
    for (int s = 0; s < kFbStages; ++s) { mbar_init(kv_full(s), 1); mbar_init(kv_empty(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(sdp_full(s), 1); mbar_init(ds_ready(s), 8); }
    mbar_init(acc_done, 1);
    fence_mbar_init();
  }
  if (warp == 0) { tmem_alloc<1>(tmem_ptr_addr, 512); tmem_relinquish<1>(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_ptr_addr));
  const uint32_t tDQ = tmem_base;
  auto tS = [&](int b_) { return tmem_base + 128 + b_ * 128; };
  auto tDP = [&](int b_) { return tmem_base + 128 + b_ * 128 + 64; };

  if (warp == 0) {
    if (lane == 0 && n_blocks > 0) {
      mbar_expect_tx(qdo_full, 2 * kTile);
#pragma unroll
      for (int hh = 0; hh < Cfg::kHalves; ++hh) {
        tma_load_4d(sQ + hh * 16384, &tmap_q, qdo_full, hh * 64, q0, h, b, kEvictFirst);
        tma_load_4d(sDO + hh * 16384, &tmap_do, qdo_full, hh * 64, q0, h, b, kEvictFirst);
      }
      for (int j = 0; j < n_blocks; ++j) {
        const int st = j % kFbStages;
        mbar_wait(kv_empty(st), ((j / kFbStages) & 1) ^ 1u);
        mbar_expect_tx(kv_full(st), 2 * kSub);
#pragma unroll
        for (int hh = 0; hh < Cfg::kHalves; ++hh) {
          tma_load_4d(sK(st) + hh * 8192, &tmap_k, kv_full(st), hh * 64, j * kFbSub, hk, b, kEvictLast);
          tma_load_4d(sV(st) + hh * 8192, &tmap_v, kv_full(st), hh * 64, j * kFbSub, hk, b, kEvictLast);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (n_blocks > 0) {
      // whole warp in the schedule, elected lane issues, descriptors by addition (see the dK / dV kernel)
      constexpr uint32_t idesc_s = umma_idesc_bf16(128, kFbSub, false, false);
      constexpr uint32_t idesc_acc = umma_idesc_bf16(128, HD, false, true);
      const uint64_t dQ0 = umma_smem_desc_sw128(sQ, 16, 1024), dDO0 = umma_smem_desc_sw128(sDO, 16, 1024);
      const uint64_t dKk = umma_smem_desc_sw128(sK(0), 16, 1024), dVk = umma_smem_desc_sw128(sV(0), 16, 1024);
      const uint64_t dKm = umma_smem_desc_sw128(sK(0), 8192, 1024);
      constexpr uint32_t kRingStep = (uint32_t)(2 * kSub) >> 4;
      auto issue_sdp = [&](int j) {
        const int st = j % kFbStages, bb = j & 1;
        mbar_wait(kv_full(st), (j / kFbStages) & 1);
        tc_fence_after();
        const uint64_t k_ = dKk + (uint64_t)(st * kRingStep), v_ = dVk + (uint64_t)(st * kRingStep);
        const uint32_t ds_ = tS(bb), dp_ = tDP(bb);
        if (elect_one_sync()) {
#pragma unroll
          for (int kk = 0; kk < HD / 16; ++kk) {
            const uint32_t offa = ((kk >> 2) * 16384 + (kk & 3) * 32) >> 4;
            const uint32_t offb = ((kk >> 2) * 8192 + (kk & 3) * 32) >> 4;
            umma_f16_ss<1>(ds_, dQ0 + offa, k_ + offb, idesc_s, kk > 0 ? 1u : 0u);
          }
#pragma unroll
          for (int kk = 0; kk < HD / 16; ++kk) {
            const uint32_t offa = ((kk >> 2) * 16384 + (kk & 3) * 32) >> 4;
            const uint32_t offb = ((kk >> 2) * 8192 + (kk & 3) * 32) >> 4;
            umma_f16_ss<1>(dp_, dDO0 + offa, v_ + offb, idesc_s, kk > 0 ? 1u : 0u);
          }
          umma_commit(sdp_full(bb));
        }
      };
      mbar_wait(qdo_full, 0);
      issue_sdp(0);
      if (n_blocks > 1) issue_sdp(1);
      for (int j = 0; j < n_blocks; ++j) {
        const int st = j % kFbStages, bb = j & 1;
        mbar_wait(ds_ready(bb), (j >> 1) & 1);
        tc_fence_after();
        const uint64_t km = dKm + (uint64_t)(st * kRingStep);
        const uint32_t pd = tDP(bb);
        const uint32_t acc0 = j > 0 ? 1u : 0u;
        if (elect_one_sync()) {
#pragma unroll
          for (int kk = 0; kk < kFbSub / 16; ++kk) {
            const uint32_t acol = (kk >> 1) * 32 + (kk & 1) * 8;
            umma_f16_ts<1>(tDQ, pd + acol, km + (uint64_t)(kk * (2048 >> 4)), idesc_acc, kk > 0 ? 1u : acc0);
          }
          umma_commit(kv_empty(st));
        }
        if (j + 2 < n_blocks) issue_sdp(j + 2);
      }
      if (elect_one_sync()) umma_commit(acc_done);
    }
    __syncwarp();
  } else {
    const int quarter = warp & 3;
    const int half = (warp - 2) >> 2;
    const int row = quarter * 32 + lane;
    const int q_idx = q0 + row;
    const uint32_t lane_off = static_cast<uint32_t>(quarter * 32) << 16;
    const float sl2 = a.scale * kLog2e;
    const long long roff = ((long long)b * a.H + h) * a.Sq + q_idx;
    const float lse_raw = q_idx < a.Sq ? p.LSE[roff] : INFINITY;
    const float lse2 = (lse_raw == -INFINITY) ? INFINITY : lse_raw * kLog2e;  // empty rows: P = 0
    const float delta = q_idx < a.Sq ? p.Delta[roff] : 0.f;
    const int q_pos = q_idx + a.q_offset - a.k_offset;
    for (int j = 0; j < n_blocks; ++j) {
      const int bb = j & 1;
      const int kv0 = j * kFbSub;
      const bool need_mask = (kv0 + kFbSub > a.Sk) ||
                             (a.causal && (kv0 + kFbSub - 1 > q0 + a.q_offset - a.k_offset));
      mbar_wait(sdp_full(bb), (j >> 1) & 1);
      tc_fence_after();
      const uint32_t cs = tS(bb) + lane_off + half * 32, cd = tDP(bb) + lane_off + half * 32;
      uint32_t rs[32], rd[32];
      tmem_ld_32x32(cs, rs);
      tmem_ld_32x32(cd, rd);
      tmem_ld_wait();
      uint32_t dk[16];
      const uint64_t sl2_2 = f2_pack(sl2, sl2), nlse_2 = f2_pack(-lse2, -lse2), ndel_2 = f2_pack(-delta, -delta);
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        float t0, t1, d0, d1;
        f2_unpack(ffma2(f2_pack_u(rs[i], rs[i + 1]), sl2_2, nlse_2), t0, t1);
        float p0 = ex2_approx(t0), p1 = ex2_approx(t1);
        if (need_mask) {
          const int col = kv0 + half * 32 + i;
          if (!(col < a.Sk && (!a.causal || col <= q_pos))) p0 = 0.f;
          if (!(col + 1 < a.Sk && (!a.causal || col + 1 <= q_pos))) p1 = 0.f;
        }
        f2_unpack(fmul2(f2_pack(p0, p1), fadd2(f2_pack_u(rd[i], rd[i + 1]), ndel_2)), d0, d1);
        dk[i >> 1] = pack_bf16(d0, d1);
      }
      tmem_st_32x16(cd, dk);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(ds_ready(bb));
    }
    __nv_bfloat16* dqrow = reinterpret_cast<__nv_bfloat16*>(p.dQ) + (long long)b * p.dq_sb + (long long)q_idx * p.dq_ss + (long long)h * p.dq_sh;
    if (n_blocks > 0) {
      mbar_wait(acc_done, 0);
      tc_fence_after();
#pragma unroll 1
      for (int c = half; c < HD / 32; c += 2) {
        uint32_t r[32];
        tmem_ld_32x32(tDQ + lane_off + c * 32, r);
        tmem_ld_wait();
        if (q_idx < a.Sq) store_chunk_bf16(dqrow + c * 32, r, a.scale);
      }
    } else if (q_idx < a.Sq && half == 0) {
      for (int c = 0; c < HD / 8; ++c) st_v4(dqrow + c * 8, make_uint4(0, 0, 0, 0));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { __syncwarp(); tc_fence_after(); tmem_dealloc<1>(tmem_base, 512); }
}

template <int HD>
static int launch_fa_bwd(const void* Q, const void* K, const void* V, const void* dO,
                         const AttnBwdParams& p, cudaStream_t stream) {
  using Cfg = FbCfg<HD>;
  const AttnParams& a = p.a;
  // resident tiles are 128 rows, streamed tiles 64 rows: each tensor needs both box shapes
  CUtensorMap tq, tk, tv, tdo, tq64, tk64, tv64, tdo64;
  if (int rc = make_attn_tmap(&tq, Q, a.B, a.Sq, a.H, HD, a.q_sb, a.q_ss, a.q_sh, kFbBlock)) return rc;
  if (int rc = make_attn_tmap(&tk, K, a.B, a.Sk, a.Hkv, HD, a.k_sb, a.k_ss, a.k_sh, kFbBlock)) return rc;
  if (int rc = make_attn_tmap(&tv, V, a.B, a.Sk, a.Hkv, HD, a.v_sb, a.v_ss, a.v_sh, kFbBlock)) return rc;
  if (int rc = make_attn_tmap(&tdo, dO, a.B, a.Sq, a.H, HD, p.do_sb, p.do_ss, p.do_sh, kFbBlock)) return rc;
  if (int rc = make_attn_tmap(&tq64, Q, a.B, a.Sq, a.H, HD, a.q_sb, a.q_ss, a.q_sh, kFbSub)) return rc;
  if (int rc = make_attn_tmap(&tk64, K, a.B, a.Sk, a.Hkv, HD, a.k_sb, a.k_ss, a.k_sh, kFbSub)) return rc;
  if (int rc = make_attn_tmap(&tv64, V, a.B, a.Sk, a.Hkv, HD, a.v_sb, a.v_ss, a.v_sh, kFbSub)) return rc;
  if (int rc = make_attn_tmap(&tdo64, dO, a.B, a.Sq, a.H, HD, p.do_sb, p.do_ss, p.do_sh, kFbSub)) return rc;
  auto k1 = flash_attn_bwd_dkv_kernel<HD>;
  auto k2 = flash_attn_bwd_dq_kernel<HD>;
  static bool attr_set = false;
  if (!attr_set) {
    B200_CUDA_CHECK(cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    B200_CUDA_CHECK(cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    attr_set = true;
  }
  dim3 g1(ceil_div(a.Sk, kFbBlock), a.Hkv, a.B);
  k1<<<g1, kFbThreads, Cfg::kSmemBytes, stream>>>(tq64, tk, tv, tdo64, p);
  B200_LAUNCH_CHECK();
  dim3 g2(ceil_div(a.Sq, kFbBlock), a.H, a.B);
  k2<<<g2, kFbThreads, Cfg::kSmemBytes, stream>>>(tq, tk64, tv64, tdo, p);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

}  // namespace b200

using namespace b200;

// delta: fp32 scratch [B, H, Sq] written by the pre-pass (caller allocates).
B200_API int b200_flash_attn_bwd(const void* Q, const void* K, const void* V, const void* O,
                                 const void* dO, const float* LSE, float* delta, void* dQ, void* dK,
                                 void* dV, int B, int H, int Hkv, int Sq, int Sk, int hd,
                                 const long long* strides /* 24: q,k,v,o,do,dq,dk,dv x (b,s,h) */,
                                 float scale, int causal, int q_offset, int k_offset,
                                 cudaStream_t stream) {
  if (B <= 0 || H <= 0 || Sq <= 0) return B200_OK;
  AttnBwdParams p = {};
  AttnParams& a = p.a;
  a.B = B; a.H = H; a.Hkv = Hkv; a.Sq = Sq; a.Sk = Sk; a.hd = hd;
  a.q_sb = strides[0]; a.q_ss = strides[1]; a.q_sh = strides[2];
  a.k_sb = strides[3]; a.k_ss = strides[4]; a.k_sh = strides[5];
  a.v_sb = strides[6]; a.v_ss = strides[7]; a.v_sh = strides[8];
  a.o_sb = strides[9]; a.o_ss = strides[10]; a.o_sh = strides[11];
  p.do_sb = strides[12]; p.do_ss = strides[13]; p.do_sh = strides[14];
  p.dq_sb = strides[15]; p.dq_ss = strides[16]; p.dq_sh = strides[17];
  p.dk_sb = strides[18]; p.dk_ss = strides[19]; p.dk_sh = strides[20];
  p.dv_sb = strides[21]; p.dv_ss = strides[22]; p.dv_sh = strides[23];
  a.scale = scale; a.causal = causal; a.q_offset = q_offset; a.k_offset = k_offset;
  p.LSE = LSE; p.Delta = delta; p.dQ = dQ; p.dK = dK; p.dV = dV;
  long long ored = 0;
  for (int i = 0; i < 24; ++i) ored |= strides[i];
  if (!((hd == 64 || hd == 128) && !(ored & 7) && Hkv > 0 && H % Hkv == 0 && Sk > 0)) {
    b200_set_error("flash_attn_bwd: unsupported shape/stride (hd=%d)", hd);
    return B200_ERR_SHAPE;
  }
  const long long rows = (long long)B * H * Sq;
  attn_delta_kernel<<<(unsigned)ceil_div64(rows * 32, 256), 256, 0, stream>>>(
      (const __nv_bfloat16*)O, (const __nv_bfloat16*)dO, delta, B, H, Sq, hd, a.o_sb, a.o_ss, a.o_sh,
      p.do_sb, p.do_ss, p.do_sh);
  B200_LAUNCH_CHECK();
  if (hd == 128) return launch_fa_bwd<128>(Q, K, V, dO, p, stream);
  return launch_fa_bwd<64>(Q, K, V, dO, p, stream);
}
