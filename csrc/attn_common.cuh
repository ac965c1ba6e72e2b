// Shared declarations for the attention kernels (forward, backward, SIMT fallback).
#pragma once
#include "common.cuh"

namespace b200 {

// Q/K/V/O are addressed as [B, S, H, hd] with arbitrary (16-byte aligned) batch/seq/head
// strides and unit stride on hd, so the kernels read straight out of a fused QKV GEMM
// output and write the layout the output projection consumes: no transposes.
struct AttnParams {
  int B, H, Hkv, Sq, Sk, hd;
  long long q_sb, q_ss, q_sh;      // element strides: batch, seq, head
  long long k_sb, k_ss, k_sh;
  long long v_sb, v_ss, v_sh;
  long long o_sb, o_ss, o_sh;
  float scale;                     // softmax scale (1/sqrt(hd))
  int causal;
  int q_offset, k_offset;          // global positions of row 0 / key 0 (ring attention chunks)
  void* O;                         // bf16 [B, Sq, H, hd] (strided)
  float* LSE;                      // fp32 [B, H, Sq]
};

// Ring (context-parallel) attention as ONE forward kernel: K and V of the whole sequence are addressed in a LOCAL gathered
// buffer [B, S_total, Hkv, hd] whose remote chunks are still arriving while the kernel runs.  A few extra CTAs of the grid
// ("communication CTAs") pull the peers' chunks out of their symmetric buffers with plain (non-multicast) TMA bulk copies,
// 128-key block by block in ascending key order, and publish one ready flag per (tensor, batch, block); the loader warp of
// every attention CTA waits for the flag of the block it is about to load.  (O, m, l) never leave the chip: no per-chunk
// partial outputs, no log-sum-exp merge.  ring.on == 0 disables all of it.
struct RingParams {
  int on;
  int n_comm;                      // communication CTAs appended to grid.x (blockIdx.y == blockIdx.z == 0 only)
  int world, rank;
  int chunk_rows;                  // keys per rank (multiple of 128)
  long long block_bytes;           // 128 * Hkv * hd * 2: one key block of one tensor of one batch entry
  long long batch_bytes;           // S_total * Hkv * hd * 2
  long long v_offset_bytes;        // offset of V behind K inside the gathered buffer (B * batch_bytes)
  int batch;
  uint32_t* flags;                 // local, [2 (K, V)][B][S_total / 128], ready when >= epoch
  uint32_t epoch;
  const void* src[8];              // rank p's gathered buffer as mapped here (same layout as the local one)
  void* dst;                       // local gathered buffer
  uint32_t* sig[8];                // peers' signal pads (arrive / depart rows of the Communicator)
  uint32_t* counter;               // local counter for the depart handshake
  int channel; uint32_t seq;
};

// Work per CTA of a causal kernel falls along grid.x, so the ORDER in which CTAs are handed out decides the tail: the
// hardware dispatches x-fastest, which on a (tile, head, batch) grid runs each head's heavy tiles right before its light
// ones and leaves a heavy tile of the last head for the very end (measured: SMs active 77 % of the forward kernel at
// 512 CTAs over 148 SMs).  Read the linear dispatch index tile-major inside GROUPS of `group` heads instead: the
// heaviest tiles of all heads of a group first, their lightest last (longest-processing-time-first), one group after
// the other.  A group is large enough that its light tiles keep the SMs busy while the next group's heavy ones start
// (group * work per head >= heaviest tile * SMs), and small enough that the K / V (and Q / dO) of the heads in flight
// stay in L2 - fully tile-major order (one group) made the backward kernels slower, its working set exceeds L2.
// `rank` = position in heavy -> light order (the caller maps it to its tile index), nx = tiles along x (without any
// extra CTAs appended to grid.x).
__device__ __forceinline__ void lpt_decode(int nx, int group, int& rank, int& h, int& b) {
  const int ny = (int)gridDim.y, per = ny * (int)gridDim.z;
  const int lin = (int)blockIdx.x + nx * ((int)blockIdx.y + ny * (int)blockIdx.z);
  const int full = nx * group;
  const int g = lin / full, base = g * group;
  const int gsz = min(group, per - base);
  const int within = lin - g * full;
  rank = within / gsz;
  const int hb = base + (within - rank * gsz);
  b = hb / ny;
  h = hb - b * ny;
}

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

// 4-D tensor map over a [B, S, H, hd] bf16 tensor: dims {hd, S, H, B}, box {64, rows, 1, 1},
// 128-byte swizzle.  One box = `rows` sequence positions x 64 head-dim elements.
inline int make_attn_tmap(CUtensorMap* out, const void* base, int B, int S, int H, int hd,
                          long long sb, long long ss, long long sh, int box_rows) {
  const uint64_t dims[4] = {(uint64_t)hd, (uint64_t)S, (uint64_t)H, (uint64_t)B};
  const uint64_t strides[3] = {(uint64_t)ss * 2, (uint64_t)sh * 2, (uint64_t)sb * 2};
  const uint32_t box[4] = {64u, (uint32_t)box_rows, 1u, 1u};
  return b200_make_tmap(out, base, 2, 4, dims, strides, box, 1);
}

}  // namespace b200
