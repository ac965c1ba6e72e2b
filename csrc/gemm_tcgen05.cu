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
#include <stdlib.h>

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
  int splits;      // split-K factor: tile index = mn_tile * splits + split; fp32 output accumulated with red.add
  int kb_per_split;
};

// Extra state of the fused compute+collective variants (FUSED = true instantiations only).
//   all-gather -> GEMM : the A operand lives in a symmetric [M, K] buffer; rows of chunk r are produced by rank r.
//     The first `n_comm` clusters of the grid are COMMUNICATION CTAs: they pull the peers' chunks into the local
//     buffer with TMA bulk copies (peer global -> smem ring -> local global, megabytes in flight from one thread) and
//     publish a ready flag per sub-chunk; the GEMM CTAs' TMA producer waits on the flag of the rows it is about to
//     load, starting with the local chunk, so math overlaps the transfer tile by tile.
//   GEMM -> reduce-scatter : the epilogue stores every output row straight into the OWNER rank's staging slab
//     (peer st over NVLink); tiles are visited owner-rotated so all links are busy; a small reduce kernel sums the
//     slabs at the owner afterwards.
struct FusedParams {
  // ---- all-gather prologue ----
  uint32_t* a_flags;            // local, one uint32 per sub-chunk of `rows_per_flag` rows; ready when == epoch
  int rows_per_flag;
  uint32_t epoch;
  int n_comm;                   // clusters at the front of the grid that only move data
  const void* a_src[8];         // rank p's symmetric A buffer as mapped here (same layout as the local one)
  void* a_dst;                  // local symmetric A buffer [M, K], contiguous rows
  long long chunk_rows;         // M / world
  long long row_bytes;          // K * 2
  uint32_t* sig[8];             // peers' signal pads (arrive / depart rows, see comm_kernels.cu)
  uint32_t* counter;            // local counter for the depart handshake
  int channel; uint32_t seq;
  int world, rank;
  // ---- scatter epilogue ----
  void* d_ptrs[8];              // rank p's staging buffer [world, d_chunk_rows, N]
  long long d_chunk_rows;       // rows owned by each rank (0 = scatter disabled)
  int m_rot;                    // m-block rotation so that tiles start at a given chunk
  // ---- implicit-GEMM convolution (A operand = NHWC activation read through a 4-D tensor map) ----
  //   Stride-1 conv as a GEMM whose M rows are OUTPUT pixels and whose K index is (kh, kw, c): k-block (tap, 64-channel
  //   slice) of an m-block covering output pixels [n0.., h0.., w0..] is the activation box at (h0 + kh - pad,
  //   w0 + kw - pad): one shifted TMA box load per k-block, out-of-range coordinates zero-filled by the TMA unit
  //   (= the conv padding).  No patch matrix is ever written.  CG = 1 only.
  //   conv_on == 2 is the WEIGHT GRADIENT: dW[co, (kh, kw, c)] = sum over output pixels of dY[pixel, co] * X[pixel + tap, c].
  //   M = Cout, N = KH * KW * C, K = output pixels; k-block = a box of <= 64 output pixels; A (MN-major) = that box of dY
  //   (two 64-channel slabs), B (MN-major) = per 64-column slab the box of X shifted by the slab's tap.  Rows of a stage
  //   that no box ever writes are zeroed once at kernel start so they contribute nothing.
  int conv_on;
  int c_C;                      // input channels (weight gradient: column -> (tap, channel))
  int c_Ho, c_Wo, c_N;          // output extent
  int c_hbox, c_wbox, c_nbox;   // output-pixel box of one m-block (c_nbox * c_hbox * c_wbox <= 128 rows)
  int c_tiles_h, c_tiles_w;
  int c_KW, c_pad, c_cblocks;   // kernel width, padding, C / 64
  int c_stride;                 // convolution stride (1 or 2): input coordinate = output coordinate * stride + tap - pad
  int c_stem;                   // 1: small-channel stem (see launch_conv_stem): sliding-window activation boxes, k-block = one filter row
  int c_box_bytes;              // bytes one box load delivers (the rest of the 128-row A stage stays stale: rows unused)
};

constexpr int kBlockM = 128;   // rows per CTA (UMMA M = 128 * CTA_GROUP)
constexpr int kBlockK = 64;    // 64 bf16 = one 128-byte swizzle row
constexpr int kUmmaK = 16;
constexpr int kGroupM = 8;     // rasterisation group
constexpr int kGemmThreads = 192;
constexpr int kSmemBudget = 232448 - 2048;  // 227 KB minus alignment slack and barriers

// EPI = 1: the epilogue stages the bf16 output tile in shared memory (128-byte-swizzled 64-column slabs) and writes it with
// TMA tensor stores - used by the GEMM -> reduce-scatter variant, whose destination is the OWNER rank's slab over NVLink:
// full 128-byte lines leave the SM as bulk transfers instead of one 16-byte peer store per thread and row.
template <int CG, int BN, int EPI = 0>
struct GemmCfg {
  static constexpr int kLoadN = BN / CG;  // B rows loaded by each CTA
  static constexpr int kABytes = kBlockM * kBlockK * 2;
  static constexpr int kBBytes = kLoadN * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kEpiBytes = EPI ? kBlockM * BN * 2 : 0;   // [BN / 64 slabs][128 rows][128 bytes]
  static constexpr int kStagesRaw = (kSmemBudget - kEpiBytes) / kStageBytes;
  static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
  static constexpr int kTmemCols = 2 * BN;  // double-buffered accumulators
  static constexpr int kSmemBytes = kStages * kStageBytes + kEpiBytes + 1024 + 256;
};

__device__ __forceinline__ float apply_act(float v, int act) {
  if (act == 1) return 0.5f * v * (1.f + erff(v * 0.70710678118654752f));
  if (act == 2) return fmaxf(v, 0.f);
  if (act == 3) return v / (1.f + __expf(-v));
  return v;
}

__device__ __forceinline__ void st_release_gpu_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_gpu_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void bulk_g2s(uint32_t smem_dst, const void* gsrc, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"(smem_dst), "l"(gsrc), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* gdst, uint32_t smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
               :: "l"(gdst), "r"(smem_src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void spin_flag(const uint32_t* p, uint32_t want, bool sys) {
  uint64_t t0 = 0;
  uint32_t spins = 0;
  while ((int32_t)((sys ? ld_acquire_sys_u32(p) : ld_acquire_gpu_u32(p)) - want) < 0)
    peer_spin_check(t0, spins, "fused gemm flag", want, p);
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

struct ScatterMaps { CUtensorMap d[8]; };   // EPI = 1: tensor map over each rank's staging slab set [world * chunk_rows, N]

template <int CG, int BN, bool A_MN, bool B_MN, typename OutT, bool FUSED = false, int EPI = 0>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a,
                    const __grid_constant__ CUtensorMap tmap_b, const GemmParams p,
                    const __grid_constant__ FusedParams fp, const __grid_constant__ ScatterMaps smaps) {
  using Cfg = GemmCfg<CG, BN, EPI>;
  constexpr int kStages = Cfg::kStages;
  constexpr int kLoadN = Cfg::kLoadN;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t epi_base = smem_base + kStages * Cfg::kStageBytes;   // EPI = 1: output staging (1024-byte aligned)
  const uint32_t bar_base = epi_base + Cfg::kEpiBytes;
  // barrier layout (8 bytes each): full[S] | empty[S] | tmem_full[2] | tmem_empty[2] | tmem_ptr
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kStages + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * kStages + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * kStages + 2 + s); };
  const uint32_t tmem_ptr_addr = bar_base + 8u * (2 * kStages + 4);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // warp index: provably uniform
  const uint32_t cta_rank = (CG == 2) ? cluster_ctarank() : 0u;
  const bool is_leader = cta_rank == 0;
  const int n_comm = FUSED ? fp.n_comm : 0;
  const int cluster_id = (int)(blockIdx.x / CG) - n_comm;   // < 0: communication CTA
  const int num_clusters = (int)(gridDim.x / CG) - n_comm;

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
  if constexpr (FUSED && CG == 1 && A_MN && B_MN) {
    if (fp.conv_on == 2) {   // zero every operand stage: rows beyond a pixel box are never written by TMA
      uint8_t* base = smem_raw + (smem_base - smem_u32(smem_raw));
      uint4* z = reinterpret_cast<uint4*>(base);
      for (int i = threadIdx.x; i < kStages * Cfg::kStageBytes / 16; i += blockDim.x) z[i] = make_uint4(0, 0, 0, 0);
      fence_proxy_async_all();
    }
  }
  tc_fence_before();
  if constexpr (CG == 2) cluster_sync(); else __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_ptr_addr));

  if (FUSED && cluster_id < 0) {
    // ===================== communication CTAs (all-gather prologue) =====================
    // One thread streams peer chunks through the smem ring with TMA bulk copies; every sub-chunk of
    // `rows_per_flag` rows gets its ready flag once its bytes are in local memory.
    if (threadIdx.x == 0) {
      const int my = blockIdx.x;                      // 0 .. n_comm * CG - 1
      const int n_ctas = n_comm * CG;
      const int W = fp.world;
      // arrive: tell every peer that our chunk is complete (stream order put its producer before this kernel)
      if (my == 0) {
        __threadfence_system();
        for (int t = 0; t < W; ++t) st_release_sys_u32(fp.sig[t] + ((size_t)fp.channel * 2 + 0) * 32 + fp.rank, fp.seq);
      }
      const long long sub_per_chunk = fp.chunk_rows / fp.rows_per_flag;
      const long long sub_bytes = (long long)fp.rows_per_flag * fp.row_bytes;
      constexpr uint32_t kPiece = Cfg::kStageBytes;   // bytes per ring slot
      uint32_t phases = 0;                            // per-slot mbarrier parity
      for (int dp = 1; dp < W; ++dp) {
        const int peer = (fp.rank + dp) % W;
        spin_flag(fp.sig[fp.rank] + ((size_t)fp.channel * 2 + 0) * 32 + peer, fp.seq, true);  // peer's chunk is ready
        for (long long sc = my; sc < sub_per_chunk; sc += n_ctas) {
          const long long off = ((long long)peer * fp.chunk_rows + sc * fp.rows_per_flag) * fp.row_bytes;
          const char* src = reinterpret_cast<const char*>(fp.a_src[peer]) + off;
          char* dst = reinterpret_cast<char*>(fp.a_dst) + off;
          for (long long b0 = 0; b0 < sub_bytes; b0 += (long long)kPiece * kStages) {
            // batch: up to kStages peer loads in flight, then drain them to local memory in order
            int nb = 0;
            for (; nb < kStages && b0 + (long long)nb * kPiece < sub_bytes; ++nb) {
              const long long b = b0 + (long long)nb * kPiece;
              const uint32_t n = (uint32_t)((sub_bytes - b) < kPiece ? (sub_bytes - b) : kPiece);
              mbar_expect_tx(full_bar(nb), n);
              bulk_g2s(smem_base + nb * Cfg::kStageBytes, src + b, n, full_bar(nb));
            }
            for (int i = 0; i < nb; ++i) {
              const long long b = b0 + (long long)i * kPiece;
              const uint32_t n = (uint32_t)((sub_bytes - b) < kPiece ? (sub_bytes - b) : kPiece);
              mbar_wait(full_bar(i), (phases >> i) & 1u);
              phases ^= (1u << i);
              bulk_s2g(dst + b, smem_base + i * Cfg::kStageBytes, n);
              tma_store_commit();
            }
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // slots may be refilled
          }
          // all bytes of this sub-chunk are in local memory -> publish
          asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
          fence_proxy_async_all();
          __threadfence();
          st_release_gpu_u32(fp.a_flags + (long long)peer * sub_per_chunk + sc, fp.epoch);
        }
      }
      // depart: peers may overwrite their chunk only after everybody stopped reading it
      __threadfence();
      const uint32_t done = atomicAdd(fp.counter + fp.channel, 1u) + 1u;
      if (done == (uint32_t)n_ctas) {
        fp.counter[fp.channel] = 0;
        __threadfence_system();
        for (int t = 0; t < W; ++t) st_release_sys_u32(fp.sig[t] + ((size_t)fp.channel * 2 + 1) * 32 + fp.rank, fp.seq);
        for (int t = 0; t < W; ++t) spin_flag(fp.sig[fp.rank] + ((size_t)fp.channel * 2 + 1) * 32 + t, fp.seq, true);
      }
    }
    __syncwarp();
  } else if (warp == 0) {
    {
      // ===================== TMA producer =====================
      // (the whole warp walks the loop; only the elected lane issues - see elect_one_sync)
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t full0 = (CG == 2) ? mapa(full_bar(0), 0) : full_bar(0);
      for (int tile = cluster_id; tile < p.num_tiles; tile += num_clusters) {
        int m_blk, n_blk;
        tile_coords(tile / p.splits, p.num_m_blocks, p.num_n_blocks, m_blk, n_blk);
        const int kb0 = (tile % p.splits) * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, p.num_k_blocks);
        if constexpr (FUSED) m_blk = (m_blk + fp.m_rot) % p.num_m_blocks;
        const int m0 = (m_blk * CG + (int)cta_rank) * kBlockM;
        const int n0 = n_blk * BN + (int)cta_rank * kLoadN;
        if constexpr (FUSED) {
          if (fp.a_flags != nullptr && m0 < p.M) {
            // rows [m0, m0 + 128) of the gathered A: wait until their sub-chunk(s) have landed (local chunk: always)
            const long long r_last = (long long)(m0 + kBlockM - 1 < p.M ? m0 + kBlockM - 1 : p.M - 1);
            for (long long f = m0 / fp.rows_per_flag; f <= r_last / fp.rows_per_flag; ++f) {
              const long long owner = f * fp.rows_per_flag / fp.chunk_rows;
              if (owner != fp.rank) spin_flag(fp.a_flags + f, fp.epoch, false);
            }
            fence_proxy_async_all();
          }
        }
        // Implicit-GEMM convolution: everything that depends only on the tile is computed here, once, and the k loop
        // walks (channel block, kw, kh) resp. (w tile, h tile, image tile) with counters.  This loop runs on ONE thread
        // and feeds the whole CTA: with the divisions inside it, a k-block took ~1200 cycles to issue (ncu: the MMA
        // issuer spun on the full barriers while the producer never waited for a free stage).
        int cv_mode = 0, cv_w0 = 0, cv_h0 = 0, cv_n0 = 0, cv_a = 0, cv_b = 0, cv_c = 0;
        int cv_colc[kLoadN / 64 > 0 ? kLoadN / 64 : 1], cv_colw[kLoadN / 64 > 0 ? kLoadN / 64 : 1],
            cv_colh[kLoadN / 64 > 0 ? kLoadN / 64 : 1];
        uint32_t cv_tx = (uint32_t)Cfg::kStageBytes;
        if constexpr (FUSED && CG == 1) {
          cv_mode = fp.conv_on;
          if (cv_mode == 1) {
            const int tw = m_blk % fp.c_tiles_w, th = (m_blk / fp.c_tiles_w) % fp.c_tiles_h;
            cv_n0 = (m_blk / (fp.c_tiles_w * fp.c_tiles_h)) * fp.c_nbox;
            cv_w0 = fp.c_stem ? tw * fp.c_wbox : tw * fp.c_wbox * fp.c_stride - fp.c_pad;   // stem: w in OUTPUT pixels
            cv_h0 = th * fp.c_hbox * fp.c_stride - fp.c_pad;
            const int tap = kb0 / fp.c_cblocks;
            cv_a = kb0 - tap * fp.c_cblocks;          // channel block
            cv_c = tap / fp.c_KW;                     // kh
            cv_b = tap - cv_c * fp.c_KW;              // kw
            cv_tx = (uint32_t)(fp.c_box_bytes + Cfg::kBBytes);
          } else if (cv_mode == 2) {
            cv_a = kb0 % fp.c_tiles_w;                                  // w tile
            cv_b = (kb0 / fp.c_tiles_w) % fp.c_tiles_h;                 // h tile
            cv_c = kb0 / (fp.c_tiles_w * fp.c_tiles_h);                 // image tile
#pragma unroll
            for (int j = 0; j < kLoadN / 64; ++j) {
              const int col = n0 + j * 64;
              const int tap = col / fp.c_C;
              const int kh = tap / fp.c_KW;
              cv_colc[j] = col - tap * fp.c_C;
              cv_colw[j] = tap - kh * fp.c_KW - fp.c_pad;
              cv_colh[j] = kh - fp.c_pad;
            }
            cv_tx = (uint32_t)((kBlockM / 64 + kLoadN / 64) * fp.c_box_bytes);
          }
        }
        const int cv_wbox = FUSED ? fp.c_wbox : 0, cv_hbox = FUSED ? fp.c_hbox : 0, cv_nbox = FUSED ? fp.c_nbox : 0;
        const int cv_stride = FUSED ? fp.c_stride : 1, cv_cblocks = FUSED ? fp.c_cblocks : 1, cv_KW = FUSED ? fp.c_KW : 1;
        const int cv_tiles_w = FUSED ? fp.c_tiles_w : 1, cv_tiles_h = FUSED ? fp.c_tiles_h : 1;
        const bool cv_stem = FUSED ? fp.c_stem != 0 : false;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const bool el = elect_one_sync();   // the lane that issues; every lane computes the (uniform) operands
          const uint32_t sa = smem_base + stage * Cfg::kStageBytes;
          const uint32_t sb = sa + Cfg::kABytes;
          const uint32_t fb = full0 + 8u * stage;
          if constexpr (FUSED && CG == 1) {
            if (el) mbar_expect_tx(full_bar(stage), cv_tx);
          } else {
            if (is_leader && el) mbar_expect_tx(full_bar(stage), Cfg::kStageBytes * CG);
          }
          const int k0 = kb * kBlockK;
          bool a_done = false;
          if constexpr (FUSED && CG == 1 && A_MN && B_MN) {
            if (cv_mode == 2) {
              // k-block = box of output pixels (w fastest, then h, then image): counters cv_a / cv_b / cv_c
              const int pw = cv_a * cv_wbox, ph = cv_b * cv_hbox, pn = cv_c * cv_nbox;
              if (el) {
#pragma unroll
                for (int j = 0; j < kBlockM / 64; ++j) tma_load_4d(sa + j * 8192, &tmap_a, fb, m0 + j * 64, pw, ph, pn);
#pragma unroll
                for (int j = 0; j < kLoadN / 64; ++j) {
                  if (cv_stem) tma_load_4d(sb + j * 8192, &tmap_b, fb, 0, pw, ph * cv_stride + (n0 >> 6) + j, pn);  // slab = filter row
                  else tma_load_4d(sb + j * 8192, &tmap_b, fb, cv_colc[j], pw * cv_stride + cv_colw[j], ph * cv_stride + cv_colh[j], pn);
                }
              }
              if (++cv_a == cv_tiles_w) { cv_a = 0; if (++cv_b == cv_tiles_h) { cv_b = 0; ++cv_c; } }
              if (++stage == kStages) { stage = 0; phase ^= 1u; }
              continue;
            }
          }
          if constexpr (FUSED && CG == 1 && !A_MN) {
            if (cv_mode == 1) {
              if (el) {
                if (cv_stem) tma_load_4d(sa, &tmap_a, fb, 0, cv_w0, cv_h0 + cv_c, cv_n0);   // k-block = filter row kh
                else tma_load_4d(sa, &tmap_a, fb, cv_a * 64, cv_w0 + cv_b, cv_h0 + cv_c, cv_n0);   // (channel block, kw, kh)
              }
              if (++cv_a == cv_cblocks) { cv_a = 0; if (++cv_b == cv_KW) { cv_b = 0; ++cv_c; } }
              a_done = true;
            }
          }
          if (el) {
            if (a_done) {
            } else if constexpr (!A_MN) {
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
          }
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    if (is_leader) {
      // ===================== MMA issuer =====================
      // Whole warp in the loop, elected lane issues (see elect_one_sync).  The shared-memory descriptors differ between
      // stages and k-steps only in their 14-bit address field, so they are ONE add away from two per-kernel bases.
      constexpr uint32_t idesc = umma_idesc_bf16(kBlockM * CG, BN, A_MN, B_MN);
      const uint64_t adesc0 = A_MN ? umma_smem_desc_sw128(smem_base, kBlockK * 128, 1024) : umma_smem_desc_sw128(smem_base, 16, 1024);
      const uint64_t bdesc0 = B_MN ? umma_smem_desc_sw128(smem_base + Cfg::kABytes, kBlockK * 128, 1024)
                                   : umma_smem_desc_sw128(smem_base + Cfg::kABytes, 16, 1024);
      constexpr uint32_t kAStep = (A_MN ? kUmmaK * 128 : kUmmaK * 2) >> 4;   // address-field increment per k-step
      constexpr uint32_t kBStep = (B_MN ? kUmmaK * 128 : kUmmaK * 2) >> 4;
      constexpr uint32_t kStageStep = Cfg::kStageBytes >> 4;
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int tile = cluster_id; tile < p.num_tiles; tile += num_clusters) {
        mbar_wait(tempty_bar(as), aphase ^ 1u);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * BN;
        const int kb0 = (tile % p.splits) * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, p.num_k_blocks);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint64_t ad = adesc0 + (uint64_t)(stage * kStageStep);
          const uint64_t bd = bdesc0 + (uint64_t)(stage * kStageStep);
          const uint32_t acc0 = kb != kb0 ? 1u : 0u;
          if (elect_one_sync()) {
#pragma unroll
            for (int k = 0; k < kBlockK / kUmmaK; ++k)
              umma_f16_ss<CG>(tmem_d, ad + (uint64_t)(k * kAStep), bd + (uint64_t)(k * kBStep), idesc, k == 0 ? acc0 : 1u);
            // frees this smem stage (in both CTAs for cta_group::2) once the MMAs retire
            if constexpr (CG == 2) umma_commit_2sm(empty_bar(stage), 0x3);
            else umma_commit(empty_bar(stage));
          }
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
        if (elect_one_sync()) {
          if constexpr (CG == 2) umma_commit_2sm(tfull_bar(as), 0x3);
          else umma_commit(tfull_bar(as));
        }
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
      tile_coords(tile / p.splits, p.num_m_blocks, p.num_n_blocks, m_blk, n_blk);
      if constexpr (FUSED) m_blk = (m_blk + fp.m_rot) % p.num_m_blocks;
      long long row = (m_blk * CG + (int)cta_rank) * kBlockM + quarter * 32 + lane;
      bool row_ok = row < p.M;
      if constexpr (FUSED && CG == 1) {
        if (fp.conv_on == 1) {   // tile-local row r = (image, h, w) inside the output-pixel box -> NHWC pixel index
          const int r = quarter * 32 + lane;
          const int hw = fp.c_hbox * fp.c_wbox;
          const int in = r / hw, rem = r - in * hw, hh = rem / fp.c_wbox, ww = rem - hh * fp.c_wbox;
          const int tw = m_blk % fp.c_tiles_w, th = (m_blk / fp.c_tiles_w) % fp.c_tiles_h;
          const int tn = m_blk / (fp.c_tiles_w * fp.c_tiles_h);
          const int n = tn * fp.c_nbox + in, h = th * fp.c_hbox + hh, w = tw * fp.c_wbox + ww;
          row_ok = in < fp.c_nbox && n < fp.c_N && h < fp.c_Ho && w < fp.c_Wo;
          row = ((long long)n * fp.c_Ho + h) * fp.c_Wo + w;
        }
      }
      const int n_tile0 = n_blk * BN;
      mbar_wait(tfull_bar(as), aphase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + as * BN;
      OutT* drow = reinterpret_cast<OutT*>(p.D) + row * p.ldd;
      if constexpr (FUSED) {
        if (fp.d_chunk_rows > 0 && row_ok) {   // GEMM -> reduce-scatter: this row belongs to rank `owner`
          const long long owner = row / fp.d_chunk_rows;
          drow = reinterpret_cast<OutT*>(fp.d_ptrs[owner]) +
                 ((long long)fp.rank * fp.d_chunk_rows + (row - owner * fp.d_chunk_rows)) * p.ldd;
        }
      }
      const __nv_bfloat16* rrow = p.residual ? p.residual + row * p.ldr : nullptr;
      if constexpr (EPI == 1) {
        // ===== staged epilogue (GEMM -> reduce-scatter): TMEM -> registers -> 128-byte-swizzled shared-memory slabs of
        // 64 columns -> TMA tensor stores straight into the OWNER rank's slab (peer memory over NVLink) =====
        const int r_loc = quarter * 32 + lane;                          // row inside this CTA's 128-row tile
        const long long row0 = (long long)(m_blk * CG + (int)cta_rank) * kBlockM;
        // the previous tile's stores must have finished READING the staging buffer before it is rewritten
        if (warp == 2 && lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        asm volatile("bar.sync 1, 128;" ::: "memory");
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
          if (rrow && row_ok) {
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
          const uint32_t slab_row = epi_base + (uint32_t)(c >> 1) * 16384u + (uint32_t)r_loc * 128u;
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const uint32_t chunk = (uint32_t)((c & 1) * 4 + g) ^ (uint32_t)(r_loc & 7);   // SWIZZLE_128B: 16-byte chunk ^ (row % 8)
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};"
                         :: "r"(slab_row + (chunk << 4)), "r"(pack_bf16(v[g * 8 + 0], v[g * 8 + 1])),
                            "r"(pack_bf16(v[g * 8 + 2], v[g * 8 + 3])), "r"(pack_bf16(v[g * 8 + 4], v[g * 8 + 5])),
                            "r"(pack_bf16(v[g * 8 + 6], v[g * 8 + 7])) : "memory");
          }
        }
        // the accumulator now lives in shared memory: hand the TMEM stage back to the MMA issuer
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if constexpr (CG == 2) mbar_arrive_cluster(tempty_remote + 8u * as);
          else mbar_arrive(tempty_bar(as));
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy smem writes -> visible to the TMA unit
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (warp == 2 && lane == 0) {
          const long long owner = row0 / fp.d_chunk_rows;
          const int dest_row = (int)((long long)fp.rank * fp.d_chunk_rows + (row0 - owner * fp.d_chunk_rows));
#pragma unroll 1
          for (int j = 0; j < BN / 64; ++j) {
            const int col = n_tile0 + j * 64;
            if (col >= p.N) break;
            tma_store_2d(&smaps.d[owner], epi_base + (uint32_t)j * 16384u, col, dest_row);
          }
          tma_store_commit();
        }
        if (++as == 2) { as = 0; aphase ^= 1u; }
        continue;
      }
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
                if (p.splits > 1) {   // split-K partial: accumulate into the (pre-zeroed or accumulating) fp32 output
                  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};"
                               :: "l"(dst), "f"(o.x), "f"(o.y), "f"(o.z), "f"(o.w) : "memory");
                  continue;
                }
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
  if constexpr (EPI == 1) {
    if (warp == 2 && lane == 0) {   // every tensor store of this CTA has been written (not merely read out of smem)
      asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
      __threadfence_system();
    }
  }
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
  // split-K: skinny outputs with a long reduction (weight gradients of convolutions / small layers) would leave most SMs
  // idle; split the K loop over several CTAs and combine with fp32 atomics (plain fp32 output without other epilogue only)
  p.splits = 1;
  p.kb_per_split = p.num_k_blocks;
  {
    const int units = b200_sm_count() / CG;
    const bool plain = sizeof(OutT) == 4 && p.bias == nullptr && p.residual == nullptr && p.act == 0 && p.alpha == 1.f;
    if (plain && p.num_tiles * 2 <= units && p.num_k_blocks >= 16) {
      int s = units / p.num_tiles;
      const int max_by_k = p.num_k_blocks / 8;   // at least 8 k-blocks (512 reduction elements) per split
      if (s > max_by_k) s = max_by_k;
      if (s > 64) s = 64;
      if (s > 1) {
        p.kb_per_split = ceil_div(p.num_k_blocks, s);
        p.splits = ceil_div(p.num_k_blocks, p.kb_per_split);
        if (!p.accumulate)
          B200_CUDA_CHECK(cudaMemset2DAsync(p.D, p.ldd * 4, 0, (size_t)p.N * 4, p.M, stream));
        p.num_tiles *= p.splits;
      }
    }
  }

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
  FusedParams nofuse = {};
  static const ScatterMaps no_maps = {};
  B200_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, ta, tb, p, nofuse, no_maps));
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

// Fused variants always use the 2-CTA 256x256 configuration, bf16 output.
template <bool B_MN, int EPI = 0>
static int launch_gemm_fused(const void* A, const void* B, const GemmParams& p0, long long lda, long long ldb,
                             const FusedParams& fp, cudaStream_t stream) {
  constexpr int CG = 2, BN = 256;
  using Cfg = GemmCfg<CG, BN, EPI>;
  GemmParams p = p0;
  p.num_m_blocks = ceil_div(p.M, kBlockM * CG);
  p.num_n_blocks = ceil_div(p.N, BN);
  p.num_k_blocks = ceil_div(p.K, kBlockK);
  p.num_tiles = p.num_m_blocks * p.num_n_blocks;
  p.splits = 1;
  p.kb_per_split = p.num_k_blocks;
  CUtensorMap ta, tb;
  {
    const uint64_t dims[2] = {(uint64_t)p.K, (uint64_t)p.M};
    const uint64_t str[1] = {(uint64_t)lda * 2};
    const uint32_t box[2] = {(uint32_t)kBlockK, (uint32_t)kBlockM};
    if (int rc = b200_make_tmap(&ta, A, 2, 2, dims, str, box, 1)) return rc;
  }
  if (!B_MN) {
    const uint64_t dims[2] = {(uint64_t)p.K, (uint64_t)p.N};
    const uint64_t str[1] = {(uint64_t)ldb * 2};
    const uint32_t box[2] = {(uint32_t)kBlockK, (uint32_t)Cfg::kLoadN};
    if (int rc = b200_make_tmap(&tb, B, 2, 2, dims, str, box, 1)) return rc;
  } else {
    const uint64_t dims[2] = {(uint64_t)p.N, (uint64_t)p.K};
    const uint64_t str[1] = {(uint64_t)ldb * 2};
    const uint32_t box[2] = {64u, (uint32_t)kBlockK};
    if (int rc = b200_make_tmap(&tb, B, 2, 2, dims, str, box, 1)) return rc;
  }
  ScatterMaps maps = {};
  if constexpr (EPI == 1) {
    // one tensor map per destination rank: its staging buffer [world * chunk_rows, N] bf16, 64-column x 128-row boxes
    for (int o = 0; o < fp.world; ++o) {
      const uint64_t dims[2] = {(uint64_t)p.N, (uint64_t)((long long)fp.world * fp.d_chunk_rows)};
      const uint64_t str[1] = {(uint64_t)p.ldd * 2};
      const uint32_t box[2] = {64u, (uint32_t)kBlockM};
      if (int rc = b200_make_tmap(&maps.d[o], fp.d_ptrs[o], 2, 2, dims, str, box, 1)) return rc;
    }
  }
  auto kern = gemm_tcgen05_kernel<CG, BN, false, B_MN, __nv_bfloat16, true, EPI>;
  static bool attr_set = false;
  if (!attr_set) {
    B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    attr_set = true;
  }
  int clusters = b200_sm_count() / CG;  // comm clusters + GEMM clusters: all resident (one CTA per SM)
  if (clusters - fp.n_comm > p.num_tiles) clusters = p.num_tiles + fp.n_comm;
  if (clusters <= fp.n_comm) clusters = fp.n_comm + 1;
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
  B200_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, ta, tb, p, fp, maps));
  return B200_OK;
}

// Implicit-GEMM convolution launcher (stride 1): out[n, ho, wo, :] = sum_{kh, kw, c} x[n, ho + kh - pad, wo + kw - pad, c] *
// wmat[:, (kh, kw, c)]  (+ bias, + residual, activation) with x NHWC bf16 and wmat [Cout, KH * KW * C] K-major.
template <int BN>
static int launch_conv_igemm(const void* x, const void* wmat, void* out, int Nimg, int H, int W, int C, int Cout, int KH,
                             int KW, int pad, int stride, const void* bias, const void* residual, int act,
                             cudaStream_t stream) {
  constexpr int CG = 1;
  using Cfg = GemmCfg<CG, BN>;
  if (stride < 1 || stride > 2) { b200_set_error("conv_igemm: stride %d not supported (1 or 2)", stride); return B200_ERR_UNSUPPORTED; }
  const int Ho = (H + 2 * pad - KH) / stride + 1, Wo = (W + 2 * pad - KW) / stride + 1;
  if (Ho <= 0 || Wo <= 0 || (C % 64) || (Cout % 8) || Nimg <= 0) {
    b200_set_error("conv_igemm: needs C %% 64 == 0, Cout %% 8 == 0 and a non-empty output (C=%d Cout=%d Ho=%d Wo=%d)", C, Cout, Ho, Wo);
    return B200_ERR_SHAPE;
  }
  FusedParams fp = {};
  fp.conv_on = 1;
  fp.c_Ho = Ho; fp.c_Wo = Wo; fp.c_N = Nimg;
  fp.c_wbox = Wo < kBlockM ? Wo : kBlockM;
  fp.c_hbox = kBlockM / fp.c_wbox; if (fp.c_hbox > Ho) fp.c_hbox = Ho; if (fp.c_hbox < 1) fp.c_hbox = 1;
  fp.c_nbox = 1;
  if (fp.c_hbox == Ho && fp.c_wbox == Wo) {
    fp.c_nbox = kBlockM / (Ho * Wo); if (fp.c_nbox > Nimg) fp.c_nbox = Nimg; if (fp.c_nbox < 1) fp.c_nbox = 1;
  }
  fp.c_tiles_w = ceil_div(Wo, fp.c_wbox);
  fp.c_tiles_h = ceil_div(Ho, fp.c_hbox);
  fp.c_KW = KW; fp.c_pad = pad; fp.c_cblocks = C / 64; fp.c_stride = stride;
  fp.c_box_bytes = fp.c_nbox * fp.c_hbox * fp.c_wbox * 128;
  GemmParams p = {};
  p.M = Nimg * Ho * Wo; p.N = Cout; p.K = KH * KW * C;
  p.D = out; p.ldd = Cout;
  p.bias = reinterpret_cast<const __nv_bfloat16*>(bias);
  p.residual = reinterpret_cast<const __nv_bfloat16*>(residual); p.ldr = Cout;
  p.alpha = 1.f; p.act = act;
  p.num_m_blocks = ceil_div(Nimg, fp.c_nbox) * fp.c_tiles_h * fp.c_tiles_w;
  p.num_n_blocks = ceil_div(Cout, BN);
  p.num_k_blocks = p.K / kBlockK;
  p.num_tiles = p.num_m_blocks * p.num_n_blocks;
  p.splits = 1; p.kb_per_split = p.num_k_blocks;
  CUtensorMap ta, tb;
  {
    const uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)Nimg};
    const uint64_t str[3] = {(uint64_t)C * 2, (uint64_t)W * C * 2, (uint64_t)H * W * C * 2};
    // stride 2: the box spans 2 x as many input pixels and the tensor map steps over every second one
    const uint32_t box[4] = {64u, (uint32_t)(fp.c_wbox * stride), (uint32_t)(fp.c_hbox * stride), (uint32_t)fp.c_nbox};
    const uint32_t est[4] = {1u, (uint32_t)stride, (uint32_t)stride, 1u};
    if (int rc = b200_make_tmap_strided(&ta, x, 2, 4, dims, str, box, est, 1)) return rc;
  }
  {
    const uint64_t dims[2] = {(uint64_t)p.K, (uint64_t)p.N};
    const uint64_t str[1] = {(uint64_t)p.K * 2};
    const uint32_t box[2] = {(uint32_t)kBlockK, (uint32_t)Cfg::kLoadN};
    if (int rc = b200_make_tmap(&tb, wmat, 2, 2, dims, str, box, 1)) return rc;
  }
  auto kern = gemm_tcgen05_kernel<CG, BN, false, false, __nv_bfloat16, true>;
  static bool attr_set = false;
  if (!attr_set) {
    B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    attr_set = true;
  }
  int ctas = b200_sm_count();
  if (ctas > p.num_tiles) ctas = p.num_tiles;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(ctas);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  static const ScatterMaps no_maps = {};
  B200_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, ta, tb, p, fp, no_maps));
  return B200_OK;
}

// Implicit-GEMM weight gradient (stride 1): dW[co, (kh, kw, c)] (fp32, [Cout, KH*KW*C]) = sum over n, ho, wo of
// dY[n, ho, wo, co] * X[n, ho + kh - pad, wo + kw - pad, c].  Split-K over pixel boxes with fp32 red.add.
static int launch_conv_wgrad(const void* x, const void* dy, float* dw, int Nimg, int H, int W, int C, int Cout, int KH,
                             int KW, int pad, int stride, int accumulate, cudaStream_t stream) {
  constexpr int CG = 1, BN = 128;
  using Cfg = GemmCfg<CG, BN>;
  if (stride < 1 || stride > 2) { b200_set_error("conv_wgrad: stride %d not supported (1 or 2)", stride); return B200_ERR_UNSUPPORTED; }
  const int Ho = (H + 2 * pad - KH) / stride + 1, Wo = (W + 2 * pad - KW) / stride + 1;
  if (Ho <= 0 || Wo <= 0 || (C % 64) || (Cout % 8) || Nimg <= 0) {
    b200_set_error("conv_wgrad: needs C %% 64 == 0, Cout %% 8 == 0 and a non-empty output (C=%d Cout=%d)", C, Cout);
    return B200_ERR_SHAPE;
  }
  FusedParams fp = {};
  fp.conv_on = 2;
  fp.c_C = C;
  fp.c_Ho = Ho; fp.c_Wo = Wo; fp.c_N = Nimg;
  fp.c_wbox = Wo < kBlockK ? Wo : kBlockK;
  fp.c_hbox = kBlockK / fp.c_wbox; if (fp.c_hbox > Ho) fp.c_hbox = Ho; if (fp.c_hbox < 1) fp.c_hbox = 1;
  fp.c_nbox = 1;
  if (fp.c_hbox == Ho && fp.c_wbox == Wo) {
    fp.c_nbox = kBlockK / (Ho * Wo); if (fp.c_nbox > Nimg) fp.c_nbox = Nimg; if (fp.c_nbox < 1) fp.c_nbox = 1;
  }
  fp.c_tiles_w = ceil_div(Wo, fp.c_wbox);
  fp.c_tiles_h = ceil_div(Ho, fp.c_hbox);
  fp.c_KW = KW; fp.c_pad = pad; fp.c_cblocks = C / 64; fp.c_stride = stride;
  fp.c_box_bytes = fp.c_nbox * fp.c_hbox * fp.c_wbox * 128;
  GemmParams p = {};
  p.M = Cout; p.N = KH * KW * C;
  p.D = dw; p.ldd = p.N;
  p.alpha = 1.f; p.accumulate = accumulate;
  p.num_m_blocks = ceil_div(p.M, kBlockM);
  p.num_n_blocks = ceil_div(p.N, BN);
  p.num_k_blocks = ceil_div(Nimg, fp.c_nbox) * fp.c_tiles_h * fp.c_tiles_w;
  p.K = p.num_k_blocks * kBlockK;
  p.num_tiles = p.num_m_blocks * p.num_n_blocks;
  // always split-K (the reduction runs over every output pixel): the epilogue accumulates with red.global.add
  int s = b200_sm_count() / p.num_tiles; if (s < 2) s = 2;
  const int max_by_k = p.num_k_blocks / 4 > 0 ? p.num_k_blocks / 4 : 1;
  if (s > max_by_k) s = max_by_k;
  if (s < 2) s = 2;                      // splits > 1 selects the red.add epilogue (correct for any split count >= 2)
  if (s > p.num_k_blocks) s = p.num_k_blocks;
  if (s < 2) { b200_set_error("conv_wgrad: fewer than two pixel boxes"); return B200_ERR_UNSUPPORTED; }
  p.kb_per_split = ceil_div(p.num_k_blocks, s);
  p.splits = ceil_div(p.num_k_blocks, p.kb_per_split);
  if (p.splits < 2) { p.kb_per_split = ceil_div(p.num_k_blocks, 2); p.splits = ceil_div(p.num_k_blocks, p.kb_per_split); }
  if (!accumulate) B200_CUDA_CHECK(cudaMemsetAsync(dw, 0, (size_t)p.M * p.N * 4, stream));
  p.num_tiles *= p.splits;
  CUtensorMap ta, tb;
  {
    const uint64_t dims[4] = {(uint64_t)Cout, (uint64_t)Wo, (uint64_t)Ho, (uint64_t)Nimg};
    const uint64_t str[3] = {(uint64_t)Cout * 2, (uint64_t)Wo * Cout * 2, (uint64_t)Ho * Wo * Cout * 2};
    const uint32_t box[4] = {64u, (uint32_t)fp.c_wbox, (uint32_t)fp.c_hbox, (uint32_t)fp.c_nbox};
    if (int rc = b200_make_tmap(&ta, dy, 2, 4, dims, str, box, 1)) return rc;
  }
  {
    const uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)Nimg};
    const uint64_t str[3] = {(uint64_t)C * 2, (uint64_t)W * C * 2, (uint64_t)H * W * C * 2};
    const uint32_t box[4] = {64u, (uint32_t)(fp.c_wbox * stride), (uint32_t)(fp.c_hbox * stride), (uint32_t)fp.c_nbox};
    const uint32_t est[4] = {1u, (uint32_t)stride, (uint32_t)stride, 1u};
    if (int rc = b200_make_tmap_strided(&tb, x, 2, 4, dims, str, box, est, 1)) return rc;
  }
  auto kern = gemm_tcgen05_kernel<CG, BN, true, true, float, true>;
  static bool attr_set = false;
  if (!attr_set) {
    B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    attr_set = true;
  }
  int ctas = b200_sm_count();
  if (ctas > p.num_tiles) ctas = p.num_tiles;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(ctas);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  static const ScatterMaps no_maps = {};
  B200_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, ta, tb, p, fp, no_maps));
  return B200_OK;
}

// ---------------------------------------------------------------------------------------------------------
// Small-channel stem convolution (ResNet 7x7 stride-2 on RGB) as an implicit GEMM.
//   The caller hands over xp = the input zero-padded to 8 channels AND spatially ([N, Hp, Wp, 8] bf16, Hp >= 2 (Ho-1) + KH,
//   Wp >= 2 (Wo-1) + 8).  One output pixel and one filter row kh need the 8 consecutive input pixels
//   w = 2 wo .. 2 wo + 7 of row h = 2 ho + kh: 8 pixels x 8 channels = 128 contiguous bytes = exactly one swizzled smem row
//   of a k-block.  A tensor map {window: 64 elements, wo: Wo (stride two pixels = 32 bytes, i.e. windows overlap), h: Hp, n: N}
//   with box {64, wbox, 2 hbox (element stride 2), nbox} therefore delivers an A tile of 128 output pixels x 64 K-elements per TMA
//   instruction; the K index is (kh, kw in 0..7, c in 0..7) with zero weights for kw == KW..7 and c >= C.  No patch matrix
//   (the im2col path writes and re-reads 2.5 GB per step for ResNet-50 at batch 256).
static int stem_maps(CUtensorMap* tx, const void* xp, int Nimg, int Hp, int Wp, int Wo, int wbox, int hbox, int nbox) {
  // dimension 0 is the 64-element (8 pixels x 8 channels = 128 bytes) window itself, dimension 1 slides it by two pixels:
  // the dimensions overlap in memory, which the TMA unit handles (tools/microbench/tma5d_probe.cu); a 16-byte inner box
  // under SWIZZLE_128B does NOT work (it faults), so the window must be the innermost dimension as a whole.
  const uint64_t dims[4] = {64u, (uint64_t)Wo, (uint64_t)Hp, (uint64_t)Nimg};
  const uint64_t str[3] = {32u, (uint64_t)Wp * 16u, (uint64_t)Hp * Wp * 16u};
  const uint32_t box[4] = {64u, (uint32_t)wbox, (uint32_t)(hbox * 2), (uint32_t)nbox};
  const uint32_t est[4] = {1u, 1u, 2u, 1u};
  return b200_make_tmap_strided(tx, xp, 2, 4, dims, str, box, est, 1);
}

static int launch_conv_stem(const void* xp, const void* wmat, void* out, int Nimg, int Hp, int Wp, int Ho, int Wo, int Cout,
                            int KH, cudaStream_t stream) {
  constexpr int CG = 1, BN = 64;
  using Cfg = GemmCfg<CG, BN>;
  if (Cout > 64 || (Cout % 8) || Hp < 2 * (Ho - 1) + KH || Wp < 2 * (Wo - 1) + 8 || Wo > 256 || Nimg <= 0) {
    b200_set_error("conv_stem: needs Cout <= 64, Cout %% 8 == 0 and a padded input of at least %d x %d (Cout=%d Hp=%d Wp=%d)",
                   2 * (Ho - 1) + KH, 2 * (Wo - 1) + 8, Cout, Hp, Wp);
    return B200_ERR_SHAPE;
  }
  FusedParams fp = {};
  fp.conv_on = 1; fp.c_stem = 1;
  fp.c_Ho = Ho; fp.c_Wo = Wo; fp.c_N = Nimg;
  fp.c_wbox = Wo < kBlockM ? Wo : kBlockM;
  fp.c_hbox = kBlockM / fp.c_wbox; if (fp.c_hbox > Ho) fp.c_hbox = Ho; if (fp.c_hbox < 1) fp.c_hbox = 1;
  fp.c_nbox = 1;
  fp.c_tiles_w = ceil_div(Wo, fp.c_wbox);
  fp.c_tiles_h = ceil_div(Ho, fp.c_hbox);
  fp.c_KW = 1; fp.c_pad = 0; fp.c_cblocks = 1; fp.c_stride = 2;
  fp.c_box_bytes = fp.c_nbox * fp.c_hbox * fp.c_wbox * 128;
  GemmParams p = {};
  p.M = Nimg * Ho * Wo; p.N = Cout; p.K = KH * 64;
  p.D = out; p.ldd = Cout; p.alpha = 1.f;
  p.num_m_blocks = Nimg * fp.c_tiles_h * fp.c_tiles_w;
  p.num_n_blocks = 1;
  p.num_k_blocks = KH;
  p.num_tiles = p.num_m_blocks;
  p.splits = 1; p.kb_per_split = p.num_k_blocks;
  CUtensorMap ta, tb;
  if (int rc = stem_maps(&ta, xp, Nimg, Hp, Wp, Wo, fp.c_wbox, fp.c_hbox, fp.c_nbox)) return rc;
  {
    const uint64_t dims[2] = {(uint64_t)p.K, (uint64_t)p.N};
    const uint64_t str[1] = {(uint64_t)p.K * 2};
    const uint32_t box[2] = {(uint32_t)kBlockK, (uint32_t)Cfg::kLoadN};
    if (int rc = b200_make_tmap(&tb, wmat, 2, 2, dims, str, box, 1)) return rc;
  }
  auto kern = gemm_tcgen05_kernel<CG, BN, false, false, __nv_bfloat16, true>;
  static bool attr_set = false;
  if (!attr_set) {
    B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    attr_set = true;
  }
  int ctas = b200_sm_count();
  if (ctas > p.num_tiles) ctas = p.num_tiles;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(ctas);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  static const ScatterMaps no_maps = {};
  B200_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, ta, tb, p, fp, no_maps));
  return B200_OK;
}

// Weight gradient of the stem: dW[co, (kh, kw8, c8)] (fp32 [Cout, KH * 64]) = sum over output pixels of dY[pixel, co] *
// xp[n, 2 ho + kh, 2 wo + kw, c]; pixel-box split-K like launch_conv_wgrad, the B slab of filter row kh is one sliding-window box.
static int launch_conv_stem_wgrad(const void* xp, const void* dy, float* dw, int Nimg, int Hp, int Wp, int Ho, int Wo,
                                  int Cout, int KH, cudaStream_t stream) {
  constexpr int CG = 1, BN = 128;
  using Cfg = GemmCfg<CG, BN>;
  if ((Cout % 8) || Hp < 2 * (Ho - 1) + KH + 1 || Wp < 2 * (Wo - 1) + 8 || Nimg <= 0) {
    b200_set_error("conv_stem_wgrad: padded input too small or Cout %% 8 != 0 (Cout=%d Hp=%d Wp=%d)", Cout, Hp, Wp);
    return B200_ERR_SHAPE;
  }
  FusedParams fp = {};
  fp.conv_on = 2; fp.c_stem = 1;
  fp.c_C = 64;
  fp.c_Ho = Ho; fp.c_Wo = Wo; fp.c_N = Nimg;
  fp.c_wbox = Wo < kBlockK ? Wo : kBlockK;
  fp.c_hbox = kBlockK / fp.c_wbox; if (fp.c_hbox > Ho) fp.c_hbox = Ho; if (fp.c_hbox < 1) fp.c_hbox = 1;
  fp.c_nbox = 1;
  fp.c_tiles_w = ceil_div(Wo, fp.c_wbox);
  fp.c_tiles_h = ceil_div(Ho, fp.c_hbox);
  fp.c_KW = 1; fp.c_pad = 0; fp.c_cblocks = 1; fp.c_stride = 2;
  fp.c_box_bytes = fp.c_nbox * fp.c_hbox * fp.c_wbox * 128;
  GemmParams p = {};
  p.M = Cout; p.N = KH * 64;
  p.D = dw; p.ldd = p.N; p.alpha = 1.f;
  p.num_m_blocks = ceil_div(p.M, kBlockM);
  p.num_n_blocks = ceil_div(p.N, BN);
  p.num_k_blocks = Nimg * fp.c_tiles_h * fp.c_tiles_w;
  p.K = p.num_k_blocks * kBlockK;
  p.num_tiles = p.num_m_blocks * p.num_n_blocks;
  int sp = b200_sm_count() / p.num_tiles; if (sp < 2) sp = 2;
  const int max_by_k = p.num_k_blocks / 4 > 0 ? p.num_k_blocks / 4 : 1;
  if (sp > max_by_k) sp = max_by_k;
  if (sp < 2) sp = 2;
  if (sp > p.num_k_blocks) sp = p.num_k_blocks;
  if (sp < 2) { b200_set_error("conv_stem_wgrad: fewer than two pixel boxes"); return B200_ERR_UNSUPPORTED; }
  p.kb_per_split = ceil_div(p.num_k_blocks, sp);
  p.splits = ceil_div(p.num_k_blocks, p.kb_per_split);
  if (p.splits < 2) { p.kb_per_split = ceil_div(p.num_k_blocks, 2); p.splits = ceil_div(p.num_k_blocks, p.kb_per_split); }
  B200_CUDA_CHECK(cudaMemsetAsync(dw, 0, (size_t)p.M * p.N * 4, stream));
  p.num_tiles *= p.splits;
  CUtensorMap ta, tb;
  {
    const uint64_t dims[4] = {(uint64_t)Cout, (uint64_t)Wo, (uint64_t)Ho, (uint64_t)Nimg};
    const uint64_t str[3] = {(uint64_t)Cout * 2, (uint64_t)Wo * Cout * 2, (uint64_t)Ho * Wo * Cout * 2};
    const uint32_t box[4] = {64u, (uint32_t)fp.c_wbox, (uint32_t)fp.c_hbox, (uint32_t)fp.c_nbox};
    if (int rc = b200_make_tmap(&ta, dy, 2, 4, dims, str, box, 1)) return rc;
  }
  if (int rc = stem_maps(&tb, xp, Nimg, Hp, Wp, Wo, fp.c_wbox, fp.c_hbox, fp.c_nbox)) return rc;
  auto kern = gemm_tcgen05_kernel<CG, BN, true, true, float, true>;
  static bool attr_set = false;
  if (!attr_set) {
    B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    attr_set = true;
  }
  int ctas = b200_sm_count();
  if (ctas > p.num_tiles) ctas = p.num_tiles;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(ctas);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  static const ScatterMaps no_maps = {};
  B200_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, ta, tb, p, fp, no_maps));
  return B200_OK;
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

B200_DEFINE_PEER_WD_SETTER(b200_peer_wd_set_gemm)

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


// ---------------------------------------------------------------------------------------------------------
// Fused all-gather -> GEMM  (ColwiseParallel after a sequence-parallel region).
//   D[M, N] = act(X_full . B^T + bias) + residual, where X_full [M, K] is the concatenation over ranks of the
//   [M / world, K] chunks each rank wrote into ITS copy of a symmetric buffer (a_ptrs[r] = rank r's buffer as mapped
//   here; this rank's chunk must already sit at rows [rank * M/world, ...) of a_ptrs[rank]).  Communication CTAs inside
//   the GEMM grid pull the peers' chunks over NVLink while the tensor cores start on the local chunk.
//   flags: local uint32[M / rows_per_flag] (zero-initialised once); epoch/seq must increase on every call per channel.
//   Requirements: K % 8 == 0, (M / world) % 256 == 0, (M / world) % rows_per_flag == 0, bf16.
B200_API int b200_ag_gemm_bf16(const void* const* a_ptrs, const void* B, void* D, int M, int N, int K, long long ldb,
                               long long ldd, int b_mn_major, const void* bias, const void* residual, long long ldr,
                               int act, void* flags, int rows_per_flag, unsigned epoch, int n_comm_clusters,
                               const void* const* sig_ptrs, void* counter, int channel, unsigned seq, int world,
                               int rank, cudaStream_t stream) {
  using namespace b200;
  if (world < 1 || world > 8 || M % world || (M / world) % 256 || (M / world) % rows_per_flag || (K & 7) || (N & 7)) {
    b200_set_error("ag_gemm: unsupported geometry M=%d N=%d K=%d world=%d rows_per_flag=%d", M, N, K, world, rows_per_flag);
    return B200_ERR_SHAPE;
  }
  GemmParams p = {};
  p.M = M; p.N = N; p.K = K; p.D = D; p.ldd = ldd;
  p.bias = reinterpret_cast<const __nv_bfloat16*>(bias);
  p.residual = reinterpret_cast<const __nv_bfloat16*>(residual);
  p.ldr = ldr; p.alpha = 1.f; p.act = act; p.accumulate = 0;
  FusedParams fp = {};
  fp.a_flags = reinterpret_cast<uint32_t*>(flags);
  fp.rows_per_flag = rows_per_flag;
  fp.epoch = epoch;
  fp.n_comm = world > 1 ? (n_comm_clusters > 0 ? n_comm_clusters : 4) : 0;
  for (int i = 0; i < world; ++i) {
    fp.a_src[i] = a_ptrs[i];
    fp.sig[i] = reinterpret_cast<uint32_t*>(const_cast<void*>(sig_ptrs[i]));
  }
  fp.a_dst = const_cast<void*>(a_ptrs[rank]);
  fp.chunk_rows = M / world;
  fp.row_bytes = (long long)K * 2;
  fp.counter = reinterpret_cast<uint32_t*>(counter);
  fp.channel = channel; fp.seq = seq; fp.world = world; fp.rank = rank;
  fp.d_chunk_rows = 0;
  fp.m_rot = (int)((long long)rank * (M / world) / (kBlockM * 2));  // start with the local chunk
  if (world == 1) fp.a_flags = nullptr;
  if (b_mn_major) return launch_gemm_fused<true>(a_ptrs[rank], B, p, K, ldb, fp, stream);
  return launch_gemm_fused<false>(a_ptrs[rank], B, p, K, ldb, fp, stream);
}

// Fused GEMM -> reduce-scatter, step 1  (RowwiseParallel with sequence-sharded output; rows A7, A10, A12).
//   partial[M, N] = A . B^T is never materialised locally: each output row is stored into the staging buffer of the
//   rank that owns it, at slab `rank` (stage_ptrs[o] = rank o's staging [world, M / world, N] bf16, row stride N).
//   Follow with b200_comm_slab_reduce on every rank to obtain the reduced [M / world, N] chunk.
//   `residual` (optional, local [M, N] bf16, row stride ldr) is added to the partial before it is scattered.
B200_API int b200_gemm_rs_bf16(const void* A, const void* B, const void* const* stage_ptrs, int M, int N, int K,
                               long long lda, long long ldb, int b_mn_major, const void* residual, long long ldr,
                               int world, int rank, cudaStream_t stream) {
  using namespace b200;
  if (world < 1 || world > 8 || M % world || (K & 7) || (N & 7) || (lda & 7) || (ldb & 7)) {
    b200_set_error("gemm_rs: unsupported geometry M=%d N=%d K=%d world=%d", M, N, K, world);
    return B200_ERR_SHAPE;
  }
  GemmParams p = {};
  p.M = M; p.N = N; p.K = K; p.D = const_cast<void*>(stage_ptrs[rank]); p.ldd = N;
  p.alpha = 1.f;
  p.residual = reinterpret_cast<const __nv_bfloat16*>(residual);
  p.ldr = ldr;
  FusedParams fp = {};
  fp.n_comm = 0;
  fp.world = world; fp.rank = rank;
  for (int i = 0; i < world; ++i) fp.d_ptrs[i] = const_cast<void*>(stage_ptrs[i]);
  fp.d_chunk_rows = M / world;
  const int blocks_per_chunk = ceil_div(M / world, kBlockM * 2);
  fp.m_rot = ((rank + 1) % world) * blocks_per_chunk;  // first the neighbour's rows, own rows last
  // Scatter through shared memory + TMA tensor stores when every 256-row tile lies inside one owner's chunk (the normal
  // case: local sequence % 256 == 0); the per-thread peer-store epilogue remains for ragged chunks and as a reference
  // (B200DIST_GEMM_RS_DIRECT=1).
  static const bool direct = getenv("B200DIST_GEMM_RS_DIRECT") != nullptr && getenv("B200DIST_GEMM_RS_DIRECT")[0] == '1';
  if (!direct && world > 1 && (M / world) % (kBlockM * 2) == 0) {
    if (b_mn_major) return launch_gemm_fused<true, 1>(A, B, p, lda, ldb, fp, stream);
    return launch_gemm_fused<false, 1>(A, B, p, lda, ldb, fp, stream);
  }
  if (b_mn_major) return launch_gemm_fused<true>(A, B, p, lda, ldb, fp, stream);
  return launch_gemm_fused<false>(A, B, p, lda, ldb, fp, stream);
}

// Implicit-GEMM convolution, stride 1, NHWC bf16 (conv2d fprop, and dgrad when called with dY and
// the flipped / transposed filter matrix and pad' = K - 1 - pad).  x [N, H, W, C], wmat [Cout, KH * KW * C],
// out [N, Ho, Wo, Cout]; bias [Cout] and residual [N, Ho, Wo, Cout] optional; act as in b200_gemm_bf16.
// C % 64 == 0, Cout % 8 == 0.  Replaces the cuDNN convolution behind torchvision's ResNet (scripts/main.py:249) without
// materialising the im2col patch matrix.
B200_API int b200_conv2d_igemm_strided_bf16(const void* x, const void* wmat, void* out, int N, int H, int W, int C, int Cout,
                                            int KH, int KW, int pad, int stride, const void* bias, const void* residual,
                                            int act, cudaStream_t stream);
B200_API int b200_conv2d_igemm_bf16(const void* x, const void* wmat, void* out, int N, int H, int W, int C, int Cout,
                                    int KH, int KW, int pad, const void* bias, const void* residual, int act,
                                    cudaStream_t stream) {
  return b200_conv2d_igemm_strided_bf16(x, wmat, out, N, H, W, C, Cout, KH, KW, pad, 1, bias, residual, act, stream);
}
// Same with a convolution stride of 1 or 2 (ResNet stage transitions: 3x3 s2 and the 1x1 s2 shortcut): the activation
// tensor map traverses every `stride`-th pixel, so the strided taps still arrive as one TMA box per k-block.
B200_API int b200_conv2d_igemm_strided_bf16(const void* x, const void* wmat, void* out, int N, int H, int W, int C, int Cout,
                                            int KH, int KW, int pad, int stride, const void* bias, const void* residual,
                                            int act, cudaStream_t stream) {
  // 64-wide output tiles for the 64-channel layers (ResNet layer1): a 128-wide tile would spend half its MMAs and half
  // its weight loads on columns that do not exist
  if (Cout <= 64) return b200::launch_conv_igemm<64>(x, wmat, out, N, H, W, C, Cout, KH, KW, pad, stride, bias, residual, act, stream);
  return b200::launch_conv_igemm<128>(x, wmat, out, N, H, W, C, Cout, KH, KW, pad, stride, bias, residual, act, stream);
}

// Implicit-GEMM weight gradient of a stride-1 convolution (wgrad): x [N, H, W, C] and
// dy [N, Ho, Wo, Cout] NHWC bf16 -> dw [Cout, KH * KW * C] fp32 (zeroed first unless accumulate).  C % 64 == 0.
B200_API int b200_conv2d_igemm_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int C, int Cout,
                                     int KH, int KW, int pad, int accumulate, cudaStream_t stream) {
  return b200::launch_conv_wgrad(x, dy, dw, N, H, W, C, Cout, KH, KW, pad, 1, accumulate, stream);
}
B200_API int b200_conv2d_igemm_wgrad_strided(const void* x, const void* dy, float* dw, int N, int H, int W, int C, int Cout,
                                             int KH, int KW, int pad, int stride, int accumulate, cudaStream_t stream) {
  return b200::launch_conv_wgrad(x, dy, dw, N, H, W, C, Cout, KH, KW, pad, stride, accumulate, stream);
}

// Small-channel stride-2 stem convolution (torchvision ResNet conv1: 7x7, stride 2, pad 3, 3 input channels) without a patch
// matrix.  xp: input zero-padded to 8 channels and spatially by `pad` (+ enough columns for the 8-wide tap window),
// [N, Hp, Wp, 8] bf16; wmat: [Cout, KH * 64] bf16 in (kh, kw 0..7, c 0..7) order, zeros where kw >= KW or c >= C;
// out: [N, Ho, Wo, Cout] bf16.  Replaces the cuDNN stem convolution of /root/reference/scripts/main.py:249.
B200_API int b200_conv2d_stem_bf16(const void* xp, const void* wmat, void* out, int N, int Hp, int Wp, int Ho, int Wo,
                                   int Cout, int KH, cudaStream_t stream) {
  return b200::launch_conv_stem(xp, wmat, out, N, Hp, Wp, Ho, Wo, Cout, KH, stream);
}
// dw: fp32 [Cout, KH * 64] in the same column order (the caller drops the padded taps / channels).
B200_API int b200_conv2d_stem_wgrad(const void* xp, const void* dy, float* dw, int N, int Hp, int Wp, int Ho, int Wo, int Cout,
                                    int KH, cudaStream_t stream) {
  return b200::launch_conv_stem_wgrad(xp, dy, dw, N, Hp, Wp, Ho, Wo, Cout, KH, stream);
}
