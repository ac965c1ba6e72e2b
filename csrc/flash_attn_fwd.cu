// Flash-attention forward on tcgen05: S = Q K^T and O += P V both run on the 5th-gen tensor cores with S/P and O
// resident in TMEM; K/V tiles stream through a TMA ring.  One CTA owns 256 query rows as two 128-row tiles that
// ping-pong on the tensor cores (while tile A's softmax runs on the CUDA cores, tile B's MMAs run, and vice versa):
//   warp 0    : TMA loader (both Q tiles once, K/V double-buffered, shared by the two tiles)
//   warp 1    : MMA issuer: PV_A(j), QK_A(j+1), PV_B(j), QK_B(j+1), ... (one thread, in-order tensor pipe)
//   warp 2-5  : softmax of tile A, warp 6-9: softmax of tile B.  One thread per query row (TMEM lane == row, no
//               shuffles); the whole 128-wide score row is read into registers with four back-to-back tcgen05.ld,
//               P goes back INTO the S columns as packed bf16 (the PV MMA reads its A operand from TMEM), and O is
//               rescaled lazily (only when the running max moves by more than 2^8).
// TMEM: S_A | S_B | O_A | O_B = 4 x 128 columns.  Returns O (bf16) and LSE (fp32); q_offset / k_offset let ring
// attention call it on chunks.
//
// Replaces F.scaled_dot_product_attention(q, k, v, is_causal=True)
//   (/root/reference/fsdp_tp/llama2_model.py:218-223) and the explicit softmax(QK^T)V of
//   /root/reference/scripts/03_tensor_parallel_tp/tensor_parallel_vit.py:120-122.
#include "attn_common.cuh"

namespace b200 {

constexpr int kFaThreads = 320;
constexpr int kFaBlock = 128;   // rows per tile == keys per KV block
constexpr int kFaTiles = 2;     // query tiles per CTA
#ifndef B200_FA_POLY_EVERY
#define B200_FA_POLY_EVERY 0     // 1 pair in this many takes the polynomial exp2 (0 = MUFU only; measured: 2 is SLOWER, 778 vs 836 TFLOP/s - the softmax warps are issue bound, not MUFU bound)
#endif
constexpr int kFaPolyEvery = B200_FA_POLY_EVERY;
constexpr int kFaPolyMod = kFaPolyEvery > 0 ? kFaPolyEvery : 1;

template <int HD>
struct FaCfg {
  static constexpr int kHalves = HD / 64;
  static constexpr int kTileBytes = kFaBlock * HD * 2;
  static constexpr int kSmemBytes = 6 * kTileBytes + 1024 + 256;  // 2 Q + 2 K + 2 V
};

template <int HD>
__global__ void __launch_bounds__(kFaThreads, 1)
flash_attn_fwd_kernel(const __grid_constant__ CUtensorMap tmap_q,
                      const __grid_constant__ CUtensorMap tmap_k,
                      const __grid_constant__ CUtensorMap tmap_v, const AttnParams p,
                      const __grid_constant__ RingParams ring) {
  using Cfg = FaCfg<HD>;
  constexpr int kTile = Cfg::kTileBytes;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int n_qt = (int)gridDim.x - (ring.on ? ring.n_comm : 0);     // query-tile CTAs along x
  if (ring.on && (int)blockIdx.x >= n_qt) {
    // ===================== communication CTA (ring attention) =====================
    if (blockIdx.y != 0 || blockIdx.z != 0) return;
    if (threadIdx.x != 0) return;
    const int my = (int)blockIdx.x - n_qt, n_ctas = ring.n_comm, W = ring.world;
    // six 32 KB tiles of dynamic shared memory serve as the copy ring; one mbarrier per slot behind them
    constexpr int kSlots = 6;
    const uint32_t cbar = smem_base + kSlots * kTile;
    for (int i = 0; i < kSlots; ++i) mbar_init(cbar + 8u * i, 1);
    fence_mbar_init();
    auto sig_row = [&](int owner, int which) { return ring.sig[owner] + ((size_t)ring.channel * 2 + which) * 32; };
    auto spin = [&](const uint32_t* f, uint32_t want, bool sys) {
      uint64_t t0 = 0; uint32_t spins = 0;
      for (;;) {
        uint32_t v;
        if (sys) asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
        else asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
        if ((int32_t)(v - want) >= 0) break;
        peer_spin_check(t0, spins, "ring attention flag", want, f);
      }
    };
    if (my == 0) {   // my chunk sits in my buffer (stream order put the copy before this kernel): tell every peer
      __threadfence_system();
      for (int t = 0; t < W; ++t) asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(sig_row(t, 0) + ring.rank), "r"(ring.seq) : "memory");
    }
    const int blocks_per_chunk = ring.chunk_rows / kFaBlock;
    const int blocks_total = blocks_per_chunk * W;
    uint32_t phases = 0;
    // regions in ascending key order (that is the order the attention CTAs consume them): block g of (tensor x, batch bi)
    const long long n_regions = (long long)blocks_total * 2 * ring.batch;
    for (long long rg = my; rg < n_regions; rg += n_ctas) {
      const int g = (int)(rg / (2 * ring.batch));              // global key block
      const int xb = (int)(rg % (2 * ring.batch));
      const int x = xb / ring.batch, bi = xb % ring.batch;      // 0 = K, 1 = V
      const int owner = g / blocks_per_chunk;
      uint32_t* flag = ring.flags + ((long long)x * ring.batch + bi) * blocks_total + g;
      if (owner == ring.rank) continue;                        // local chunk: always present, never flagged
      spin(sig_row(ring.rank, 0) + owner, ring.seq, true);     // the owner's chunk is final
      const long long off = (long long)x * ring.v_offset_bytes + (long long)bi * ring.batch_bytes + (long long)g * ring.block_bytes;
      const char* src = reinterpret_cast<const char*>(ring.src[owner]) + off;
      char* dst = reinterpret_cast<char*>(ring.dst) + off;
      for (long long b0 = 0; b0 < ring.block_bytes; b0 += (long long)kTile * kSlots) {
        int nb = 0;
        for (; nb < kSlots && b0 + (long long)nb * kTile < ring.block_bytes; ++nb) {
          const long long bo = b0 + (long long)nb * kTile;
          const uint32_t n = (uint32_t)((ring.block_bytes - bo) < kTile ? (ring.block_bytes - bo) : kTile);
          mbar_expect_tx(cbar + 8u * nb, n);
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                       :: "r"(smem_base + nb * kTile), "l"(src + bo), "r"(n), "r"(cbar + 8u * nb) : "memory");
        }
        for (int i = 0; i < nb; ++i) {
          const long long bo = b0 + (long long)i * kTile;
          const uint32_t n = (uint32_t)((ring.block_bytes - bo) < kTile ? (ring.block_bytes - bo) : kTile);
          mbar_wait(cbar + 8u * i, (phases >> i) & 1u);
          phases ^= (1u << i);
          asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                       :: "l"(dst + bo), "r"(smem_base + i * kTile), "r"(n) : "memory");
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      }
      asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
      asm volatile("fence.proxy.async;" ::: "memory");
      __threadfence();
      asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(flag), "r"(ring.epoch) : "memory");
    }
    // depart: a peer may overwrite its chunk (next call) only after everybody stopped reading it
    __threadfence();
    const uint32_t done = atomicAdd(ring.counter + ring.channel, 1u) + 1u;
    if (done == (uint32_t)n_ctas) {
      ring.counter[ring.channel] = 0;
      __threadfence_system();
      for (int t = 0; t < W; ++t) asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(sig_row(t, 1) + ring.rank), "r"(ring.seq) : "memory");
      for (int t = 0; t < W; ++t) spin(sig_row(ring.rank, 1) + t, ring.seq, true);
    }
    return;
  }
  auto sQ = [&](int t) { return smem_base + kTile * t; };
  auto sK = [&](int s) { return smem_base + kTile * (2 + s); };
  auto sV = [&](int s) { return smem_base + kTile * (4 + s); };
  const uint32_t bar = smem_base + 6 * kTile;
  const uint32_t q_full = bar;
  auto k_full = [&](int s) { return bar + 8u * (1 + s); };
  auto k_empty = [&](int s) { return bar + 8u * (3 + s); };
  auto v_full = [&](int s) { return bar + 8u * (5 + s); };
  auto v_empty = [&](int s) { return bar + 8u * (7 + s); };
  auto s_full = [&](int t) { return bar + 8u * (9 + t); };
  auto p_ready = [&](int t) { return bar + 8u * (11 + t); };
  auto pv_done = [&](int t) { return bar + 8u * (13 + t); };
  const uint32_t tmem_ptr_addr = bar + 8u * 15;

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // warp index: provably uniform
  int lpt_rank, h, b;
  lpt_decode(n_qt, 16, lpt_rank, h, b);
  const int q_blk = n_qt - 1 - lpt_rank;         // heavy (late) causal blocks first, over all heads
  const int hk = h / (p.H / p.Hkv);
  const int q_base = q_blk * kFaBlock * kFaTiles;

  // KV blocks each tile attends to
  int n_blk[kFaTiles];
#pragma unroll
  for (int t = 0; t < kFaTiles; ++t) {
    const int q0 = q_base + t * kFaBlock;
    int lim = p.Sk;
    if (p.causal) {
      const int c = q0 + kFaBlock + p.q_offset - p.k_offset;
      lim = c < lim ? c : lim;
    }
    n_blk[t] = (q0 < p.Sq && lim > 0) ? (lim + kFaBlock - 1) / kFaBlock : 0;
  }
  const int n_max = n_blk[0] > n_blk[1] ? n_blk[0] : n_blk[1];

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmap_q);
    tma_prefetch_desc(&tmap_k);
    tma_prefetch_desc(&tmap_v);
    mbar_init(q_full, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(k_full(s), 1); mbar_init(k_empty(s), 1);
      mbar_init(v_full(s), 1); mbar_init(v_empty(s), 1);
      mbar_init(s_full(s), 1); mbar_init(p_ready(s), 4); mbar_init(pv_done(s), 1);
    }
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc<1>(tmem_ptr_addr, 512);
    tmem_relinquish<1>();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_ptr_addr));
  auto tS = [&](int t) { return tmem_base + t * 128; };
  auto tO = [&](int t) { return tmem_base + 256 + t * 128; };

  if (warp == 0) {
    if (lane == 0 && n_max > 0) {
      // ===================== loader =====================
      mbar_expect_tx(q_full, kFaTiles * kTile);   // rows past Sq are zero-filled by TMA
#pragma unroll
      for (int t = 0; t < kFaTiles; ++t)
#pragma unroll
        for (int hh = 0; hh < Cfg::kHalves; ++hh)
          tma_load_4d(sQ(t) + hh * 16384, &tmap_q, q_full, hh * 64, q_base + t * kFaBlock, h, b, kEvictFirst);
      for (int j = 0; j < n_max; ++j) {
        const int st = j & 1;
        const uint32_t par = ((j >> 1) & 1) ^ 1u;
        const int kv0 = j * kFaBlock;
        mbar_wait(k_empty(st), par);
        if (ring.on && kv0 / ring.chunk_rows != ring.rank) {   // remote key block: wait until it has landed locally
          const int g = kv0 / kFaBlock, bt = ring.chunk_rows / kFaBlock * ring.world;
          const uint32_t* fk = ring.flags + (long long)b * bt + g;
          const uint32_t* fv = ring.flags + ((long long)ring.batch + b) * bt + g;
          uint64_t t0 = 0; uint32_t spins = 0;
          for (int w_ = 0; w_ < 2; ++w_) {
            const uint32_t* f = w_ ? fv : fk;
            for (;;) {
              uint32_t v_;
              asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v_) : "l"(f) : "memory");
              if ((int32_t)(v_ - ring.epoch) >= 0) break;
              peer_spin_check(t0, spins, "ring attention block", ring.epoch, f);
            }
          }
          asm volatile("fence.proxy.async;" ::: "memory");
        }
        mbar_expect_tx(k_full(st), kTile);
#pragma unroll
        for (int hh = 0; hh < Cfg::kHalves; ++hh)
          tma_load_4d(sK(st) + hh * 16384, &tmap_k, k_full(st), hh * 64, kv0, hk, b, kEvictLast);
        mbar_wait(v_empty(st), par);
        mbar_expect_tx(v_full(st), kTile);
#pragma unroll
        for (int hh = 0; hh < Cfg::kHalves; ++hh)
          tma_load_4d(sV(st) + hh * 16384, &tmap_v, v_full(st), hh * 64, kv0, hk, b, kEvictLast);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (n_max > 0) {
      // ===================== MMA issuer =====================
      // The whole warp walks this schedule and one elected lane issues (see elect_one_sync in common.cuh): descriptor
      // arithmetic stays in uniform registers and is one 64-bit add per operand off four per-kernel bases, so the 16
      // UTCHMMA of a QK^T + PV pair issue back to back.  (With a single-lane loop every UTCHMMA carried ~16 instructions
      // of descriptor building plus an ELECT / branch wrapper - about as long as the 65-cycle MMA it issues.)
      constexpr uint32_t idesc_qk = umma_idesc_bf16(128, 128, false, false);
      constexpr uint32_t idesc_pv = umma_idesc_bf16(128, HD, false, true);
      const uint64_t dq0 = umma_smem_desc_sw128(sQ(0), 16, 1024), dk0 = umma_smem_desc_sw128(sK(0), 16, 1024);
      const uint64_t dv0 = umma_smem_desc_sw128(sV(0), 16384, 1024);
      constexpr uint32_t kTileStep = (uint32_t)kTile >> 4;          // address-field distance between consecutive tiles
      auto issue_qk = [&](int t, int j) {
        const uint64_t qd = dq0 + (uint64_t)(t * kTileStep), kd = dk0 + (uint64_t)((j & 1) * kTileStep);
        const uint32_t d = tS(t);
        if (elect_one_sync()) {
#pragma unroll
          for (int kk = 0; kk < HD / 16; ++kk) {
            const uint32_t off = ((kk >> 2) * 16384 + (kk & 3) * 32) >> 4;
            umma_f16_ss<1>(d, qd + off, kd + off, idesc_qk, kk > 0 ? 1u : 0u);
          }
          umma_commit(s_full(t));
        }
      };
      auto issue_pv = [&](int t, int j) {
        const uint64_t vd = dv0 + (uint64_t)((j & 1) * kTileStep);
        const uint32_t d = tO(t), a0 = tS(t);
        const uint32_t acc0 = j > 0 ? 1u : 0u;
        if (elect_one_sync()) {
#pragma unroll
          for (int kk = 0; kk < kFaBlock / 16; ++kk)
            umma_f16_ts<1>(d, a0 + kk * 8, vd + (uint64_t)(kk * (2048 >> 4)), idesc_pv, kk > 0 ? 1u : acc0);
          umma_commit(pv_done(t));
        }
      };
      mbar_wait(q_full, 0);
      mbar_wait(k_full(0), 0);
      tc_fence_after();
#pragma unroll
      for (int t = 0; t < kFaTiles; ++t)
        if (n_blk[t] > 0) issue_qk(t, 0);
      if (elect_one_sync()) umma_commit(k_empty(0));
      for (int j = 0; j < n_max; ++j) {
        const int st = j & 1;
        const bool more = j + 1 < n_max;
        if (more) {
          mbar_wait(k_full((j + 1) & 1), ((j + 1) >> 1) & 1);
        }
        mbar_wait(v_full(st), (j >> 1) & 1);
#pragma unroll
        for (int t = 0; t < kFaTiles; ++t) {
          if (j < n_blk[t]) {
            mbar_wait(p_ready(t), j & 1);
            tc_fence_after();
            issue_pv(t, j);
            if (j + 1 < n_blk[t]) issue_qk(t, j + 1);   // overwrites S_t/P_t only after PV_t(j) (in-order pipe)
          }
        }
        if (elect_one_sync()) {
          umma_commit(v_empty(st));
          if (more) umma_commit(k_empty((j + 1) & 1));
        }
      }
    }
    __syncwarp();
  } else {
    // ===================== softmax / correction / epilogue =====================
    const int t = (warp - 2) >> 2;           // which query tile this warpgroup owns
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int q0 = q_base + t * kFaBlock;
    const int q_idx = q0 + row;
    const uint32_t lane_off = static_cast<uint32_t>(quarter * 32) << 16;
    const float sl2 = p.scale * kLog2e;
    float m_used = -INFINITY;  // max used for scaling (raw score units)
    float l_run = 0.f;
    const int q_pos = q_idx + p.q_offset - p.k_offset;  // key j visible iff j <= q_pos (causal)
    const int nb = n_blk[t];
    const uint32_t s_addr = tS(t) + lane_off;
    const uint32_t o_addr = tO(t) + lane_off;

    for (int j = 0; j < nb; ++j) {
      const int kv0 = j * kFaBlock;
      const bool need_mask = (kv0 + kFaBlock > p.Sk) ||
                             (p.causal && (kv0 + kFaBlock - 1 > q0 + p.q_offset - p.k_offset));
      mbar_wait(s_full(t), j & 1);
      tc_fence_after();
      // ---- whole score row into registers: four loads in flight, one wait ----
      uint32_t r0[32], r1[32], r2[32], r3[32];
      tmem_ld_32x32(s_addr, r0);
      tmem_ld_32x32(s_addr + 32, r1);
      tmem_ld_32x32(s_addr + 64, r2);
      tmem_ld_32x32(s_addr + 96, r3);
      tmem_ld_wait();
      if (need_mask) {
        const uint32_t ninf = __float_as_uint(-INFINITY);
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const int c = kv0 + i;
          if (!(c < p.Sk && (!p.causal || c <= q_pos))) r0[i] = ninf;
          if (!(c + 32 < p.Sk && (!p.causal || c + 32 <= q_pos))) r1[i] = ninf;
          if (!(c + 64 < p.Sk && (!p.causal || c + 64 <= q_pos))) r2[i] = ninf;
          if (!(c + 96 < p.Sk && (!p.causal || c + 96 <= q_pos))) r3[i] = ninf;
        }
      }
      float mx = -INFINITY, mx2 = -INFINITY;   // two chains of 3-input max (FMNMX3)
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        mx = fmax3(mx, __uint_as_float(r0[i]), __uint_as_float(r1[i]));
        mx2 = fmax3(mx2, __uint_as_float(r2[i]), __uint_as_float(r3[i]));
      }
      mx = fmaxf(mx, mx2);
      // ---- lazy rescale decision ----
      float alpha = 1.f;
      bool rescale = false;
      if (mx > m_used && (m_used == -INFINITY || (mx - m_used) * sl2 > 8.f)) {
        alpha = (m_used == -INFINITY) ? 0.f : exp2f((m_used - mx) * sl2);
        m_used = mx;
        rescale = true;
      }
      l_run *= alpha;
      if (j > 0 && __any_sync(0xffffffffu, rescale)) {
        mbar_wait(pv_done(t), (j - 1) & 1);
        tc_fence_after();
#pragma unroll 1
        for (int c = 0; c < HD / 32; ++c) {
          uint32_t r[32];
          tmem_ld_32x32(o_addr + c * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
          tmem_st_32x32(o_addr + c * 32, r);
        }
        tmem_st_wait();
      }
      // ---- P = exp2((S - m) * scale), row sum, packed bf16 back into the S columns ----
      const float mneg = (m_used == -INFINITY) ? 0.f : -m_used * sl2;
      const uint64_t sl2_2 = f2_pack(sl2, sl2), mneg_2 = f2_pack(mneg, mneg);
      uint64_t lsum2 = f2_pack(0.f, 0.f);
      auto do_chunk = [&](uint32_t (&r)[32], int c) {
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float t0, t1;
          f2_unpack(ffma2(f2_pack_u(r[i], r[i + 1]), sl2_2, mneg_2), t0, t1);   // one FFMA2 for two scores
          const float p0 = ex2_approx(t0), p1 = ex2_approx(t1);                  // exp2(-inf) == 0 for masked keys
          lsum2 = fadd2(lsum2, f2_pack(p0, p1));
          pk[i >> 1] = pack_bf16(p0, p1);
        }
        tmem_st_32x16(s_addr + c * 16, pk);
      };
      // unmasked blocks (the vast majority): every kFaPolyEvery-th pair of scores takes the polynomial exp2 on the FMA pipes,
      // the rest the MUFU unit.  Masked blocks keep MUFU throughout: exp2(-inf) must be exactly 0 for fully masked rows.
      auto do_chunk_mixed = [&](uint32_t (&r)[32], int c) {
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          const uint64_t t2 = ffma2(f2_pack_u(r[i], r[i + 1]), sl2_2, mneg_2);
          float p0, p1;
          if (kFaPolyEvery > 0 && ((i >> 1) % kFaPolyMod) == kFaPolyMod - 1) {
            ex2_poly2(t2, p0, p1);
          } else {
            float t0, t1;
            f2_unpack(t2, t0, t1);
            p0 = ex2_approx(t0); p1 = ex2_approx(t1);
          }
          lsum2 = fadd2(lsum2, f2_pack(p0, p1));
          pk[i >> 1] = pack_bf16(p0, p1);
        }
        tmem_st_32x16(s_addr + c * 16, pk);
      };
      if (need_mask || kFaPolyEvery <= 0) {
        do_chunk(r0, 0);
        do_chunk(r1, 1);
        do_chunk(r2, 2);
        do_chunk(r3, 3);
      } else {
        do_chunk_mixed(r0, 0);
        do_chunk_mixed(r1, 1);
        do_chunk_mixed(r2, 2);
        do_chunk_mixed(r3, 3);
      }
      float ls0, ls1;
      f2_unpack(lsum2, ls0, ls1);
      l_run += ls0 + ls1;
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_ready(t));
    }

    // ---- epilogue: O / l -> bf16 global, LSE ----
    __nv_bfloat16* orow = reinterpret_cast<__nv_bfloat16*>(p.O) + (long long)b * p.o_sb +
                          (long long)q_idx * p.o_ss + (long long)h * p.o_sh;
    if (nb > 0) {
      mbar_wait(pv_done(t), (nb - 1) & 1);
      tc_fence_after();
      const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
#pragma unroll 1
      for (int c = 0; c < HD / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(o_addr + c * 32, r);
        tmem_ld_wait();
        if (q_idx < p.Sq) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            uint4 o;
            o.x = pack_bf16(__uint_as_float(r[g * 8 + 0]) * inv, __uint_as_float(r[g * 8 + 1]) * inv);
            o.y = pack_bf16(__uint_as_float(r[g * 8 + 2]) * inv, __uint_as_float(r[g * 8 + 3]) * inv);
            o.z = pack_bf16(__uint_as_float(r[g * 8 + 4]) * inv, __uint_as_float(r[g * 8 + 5]) * inv);
            o.w = pack_bf16(__uint_as_float(r[g * 8 + 6]) * inv, __uint_as_float(r[g * 8 + 7]) * inv);
            st_v4(orow + c * 32 + g * 8, o);
          }
        }
      }
    } else if (q_idx < p.Sq) {
      for (int c = 0; c < HD / 8; ++c) st_v4(orow + c * 8, make_uint4(0, 0, 0, 0));
    }
    if (q_idx < p.Sq && p.LSE)
      p.LSE[((long long)b * p.H + h) * p.Sq + q_idx] =
          (l_run > 0.f) ? m_used * p.scale + __logf(l_run) : -INFINITY;
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    __syncwarp();
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 512);
  }
}

template <int HD>
static int launch_fa_fwd(const void* Q, const void* K, const void* V, const AttnParams& p,
                         cudaStream_t stream, const RingParams& ring = RingParams{}) {
  using Cfg = FaCfg<HD>;
  CUtensorMap tq, tk, tv;
  if (int rc = make_attn_tmap(&tq, Q, p.B, p.Sq, p.H, HD, p.q_sb, p.q_ss, p.q_sh, kFaBlock)) return rc;
  if (int rc = make_attn_tmap(&tk, K, p.B, p.Sk, p.Hkv, HD, p.k_sb, p.k_ss, p.k_sh, kFaBlock)) return rc;
  if (int rc = make_attn_tmap(&tv, V, p.B, p.Sk, p.Hkv, HD, p.v_sb, p.v_ss, p.v_sh, kFaBlock)) return rc;
  auto kern = flash_attn_fwd_kernel<HD>;
  static bool attr_set = false;
  if (!attr_set) {
    B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    attr_set = true;
  }
  dim3 grid(ceil_div(p.Sq, kFaBlock * kFaTiles) + (ring.on ? ring.n_comm : 0), p.H, p.B);
  kern<<<grid, kFaThreads, Cfg::kSmemBytes, stream>>>(tq, tk, tv, p, ring);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

}  // namespace b200

using namespace b200;

static bool al16(const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; }

// Strides are in elements.  Supported on the tensor-core path: hd in {64, 128}, bf16,
// all strides multiples of 8.  Returns B200_ERR_SHAPE otherwise.
B200_API int b200_flash_attn_fwd(const void* Q, const void* K, const void* V, void* O, float* LSE,
                                 int B, int H, int Hkv, int Sq, int Sk, int hd,
                                 long long q_sb, long long q_ss, long long q_sh,
                                 long long k_sb, long long k_ss, long long k_sh,
                                 long long v_sb, long long v_ss, long long v_sh,
                                 long long o_sb, long long o_ss, long long o_sh,
                                 float scale, int causal, int q_offset, int k_offset,
                                 cudaStream_t stream) {
  if (B <= 0 || H <= 0 || Sq <= 0) return B200_OK;
  AttnParams p = {};
  p.B = B; p.H = H; p.Hkv = Hkv; p.Sq = Sq; p.Sk = Sk; p.hd = hd;
  p.q_sb = q_sb; p.q_ss = q_ss; p.q_sh = q_sh;
  p.k_sb = k_sb; p.k_ss = k_ss; p.k_sh = k_sh;
  p.v_sb = v_sb; p.v_ss = v_ss; p.v_sh = v_sh;
  p.o_sb = o_sb; p.o_ss = o_ss; p.o_sh = o_sh;
  p.scale = scale; p.causal = causal; p.q_offset = q_offset; p.k_offset = k_offset;
  p.O = O; p.LSE = LSE;
  const bool ok = (hd == 64 || hd == 128) && al16(Q) && al16(K) && al16(V) && al16(O) &&
                  !((q_sb | q_ss | q_sh | k_sb | k_ss | k_sh | v_sb | v_ss | v_sh | o_sb | o_ss | o_sh) & 7) &&
                  Hkv > 0 && H % Hkv == 0 && Sk > 0;
  if (!ok) {
    b200_set_error("flash_attn_fwd: unsupported shape/stride (hd=%d H=%d Hkv=%d)", hd, H, Hkv);
    return B200_ERR_SHAPE;
  }
  if (hd == 128) return launch_fa_fwd<128>(Q, K, V, p, stream);
  return launch_fa_fwd<64>(Q, K, V, p, stream);
}

// Ring attention forward as one kernel (see RingParams in attn_common.cuh).  Q: this rank's queries [B, Sq, H, hd]
// (global positions q_offset ..); kv_ptrs[r]: rank r's gathered K/V buffer as mapped here - layout
// [2 (K, V)][B][S_total][Hkv][hd] bf16, S_total = world * chunk_rows - of which only chunk r is filled (by its owner, before
// the call); flags: local uint32 [2][B][S_total / 128], zero-initialised once; epoch / seq grow by one per call on the
// channel.  Communication CTAs stage the peers' chunks into kv_ptrs[rank] block by block while the attention CTAs run.
// O: [B, Sq, H, hd] (strided), LSE: [B, H, Sq].  hd in {64, 128}, chunk_rows % 128 == 0.
// No counterpart kernel in the reference (its guide describes the algorithm, docs/guide/08_sequence_parallel.md:82-143).
B200_API int b200_flash_attn_fwd_ring(const void* Q, const void* const* kv_ptrs, void* O, float* LSE, int B, int H, int Hkv,
                                      int Sq, int chunk_rows, int hd, long long q_sb, long long q_ss, long long q_sh,
                                      long long o_sb, long long o_ss, long long o_sh, float scale, int causal, void* flags,
                                      unsigned epoch, int n_comm, const void* const* sig_ptrs, void* counter, int channel,
                                      unsigned seq, int world, int rank, cudaStream_t stream) {
  if (B <= 0 || H <= 0 || Sq <= 0) return B200_OK;
  if (world < 1 || world > 8 || (chunk_rows % kFaBlock) || !(hd == 64 || hd == 128) || Hkv <= 0 || H % Hkv) {
    b200_set_error("flash_attn_fwd_ring: unsupported geometry (world=%d chunk_rows=%d hd=%d H=%d Hkv=%d)", world, chunk_rows, hd, H, Hkv);
    return B200_ERR_SHAPE;
  }
  const int S_total = world * chunk_rows;
  AttnParams p = {};
  p.B = B; p.H = H; p.Hkv = Hkv; p.Sq = Sq; p.Sk = S_total; p.hd = hd;
  p.q_sb = q_sb; p.q_ss = q_ss; p.q_sh = q_sh;
  p.k_sb = (long long)S_total * Hkv * hd; p.k_ss = (long long)Hkv * hd; p.k_sh = hd;
  p.v_sb = p.k_sb; p.v_ss = p.k_ss; p.v_sh = p.k_sh;
  p.o_sb = o_sb; p.o_ss = o_ss; p.o_sh = o_sh;
  p.scale = scale; p.causal = causal; p.q_offset = rank * chunk_rows; p.k_offset = 0;
  p.O = O; p.LSE = LSE;
  if (!al16(Q) || !al16(O) || ((q_sb | q_ss | q_sh | o_sb | o_ss | o_sh) & 7)) {
    b200_set_error("flash_attn_fwd_ring: Q / O must be 16-byte aligned with strides that are multiples of 8");
    return B200_ERR_SHAPE;
  }
  RingParams r = {};
  r.on = world > 1 ? 1 : 0;
  r.n_comm = n_comm > 0 ? n_comm : 8;
  r.world = world; r.rank = rank; r.chunk_rows = chunk_rows; r.batch = B;
  r.block_bytes = (long long)kFaBlock * Hkv * hd * 2;
  r.batch_bytes = (long long)S_total * Hkv * hd * 2;
  r.v_offset_bytes = (long long)B * r.batch_bytes;
  r.flags = reinterpret_cast<uint32_t*>(flags);
  r.epoch = epoch;
  for (int i = 0; i < world; ++i) {
    r.src[i] = kv_ptrs[i];
    r.sig[i] = reinterpret_cast<uint32_t*>(const_cast<void*>(sig_ptrs[i]));
  }
  r.dst = const_cast<void*>(kv_ptrs[rank]);
  r.counter = reinterpret_cast<uint32_t*>(counter);
  r.channel = channel; r.seq = seq;
  const char* kbase = reinterpret_cast<const char*>(kv_ptrs[rank]);
  const void* Kp = kbase;
  const void* Vp = kbase + r.v_offset_bytes;
  if (hd == 128) return launch_fa_fwd<128>(Q, Kp, Vp, p, stream, r);
  return launch_fa_fwd<64>(Q, Kp, Vp, p, stream, r);
}
