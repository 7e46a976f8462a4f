// attention_bwd.cu -- flash-style self-attention backward on tcgen05/TMEM.
// Gradient of a = softmax(q k^T / sqrt(ch)) v (QKVAttentionLegacy / QKVAttention, unet.py:340-405)
// w.r.t. the packed qkv tensor, recomputing P from the saved log-sum-exp (no L x L matrix in HBM;
// replaces the reference's checkpoint-recompute of the whole AttentionBlock, utils.py:179-207).
//
//   delta_i = sum_c dO_ic O_ic                       (attn_delta_kernel)
//   kernel A (one CTA per 128-key tile):  S^T = K Q^T, dP^T = V dO^T,  P^T = exp2(c S^T - lse),
//            dS^T = P^T (dP^T - delta);   dV += P^T dO;   dK += dS^T Q
//   kernel B (one CTA per 128-query tile): S = Q K^T, dP = dO V^T, dS = P (dP - delta);  dQ += dS K
// All four "second" GEMMs read their B operand (dO, Q, K) in the pixel-major layout through an
// MN-major tcgen05 descriptor, so no transposed copies of q/k/v/dO are ever written.
#include <stdlib.h>
#include "common.cuh"
#include "tmap.cuh"
#include "umma.cuh"

namespace b2d {
using namespace umma;

struct AttnBwdParams {
  int L, heads;
  int q_row0, q_rows;  // query rows visited (sequence-parallel shard; 0, L otherwise)
  int q_off, k_off, v_off, head_stride;
  float scale_log2;   // log2(e)/sqrt(ch)
  float scale;        // 1/sqrt(ch)
  const float* lse;   // (N, heads, L) log2 domain
  const float* delta; // (N, heads, L)
  __half* dqkv; long long lddqkv;  // (N, L, 3C)
};

struct AttnBwdMaps {
  CUtensorMap qkv128;  // box (64, 128, 1) on qkv
  CUtensorMap qkvT;    // box (64, BT, 1) on qkv   (streamed operand tile)
  CUtensorMap doR;     // box (64, 128, 1) on dO
  CUtensorMap doT;     // box (64, BT, 1) on dO
};

// warps: 0 = TMA producer, 1 = tcgen05 issuer, 2..9 = compute: two warps per TMEM lane quarter, each one half of the
// streamed tile's columns (round 2; one warp per quarter left the exp / pack phase ~2x longer than the tile's MMAs)
constexpr int AB_CWARPS = 8;
constexpr int AB_CTHREADS = AB_CWARPS * 32;
constexpr int AB_THREADS = 64 + AB_CTHREADS;

__device__ __forceinline__ void named_bar_sync(int id, int count) {
  asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(count) : "memory");
}

// ------------------------------------------------------------------ delta = rowsum(dO * O)
__global__ void __launch_bounds__(256)
attn_delta_kernel(const __half* __restrict__ o, long long ldo, const __half* __restrict__ d_o,
                  long long lddo, int N, int L, int heads, int ch, float* __restrict__ delta) {
  // ch / 8 lanes per (n, l, head) row, one 16-byte load of O and of dO each; two rows per thread in flight
  const int G = ch >> 3;                       // 8, 16 or 32 lanes per row
  const int rows_per_warp = 32 / G;
  const int lane = threadIdx.x & 31;
  const int sub = lane / G, li = lane - sub * G;
  const long long total = (long long)N * L * heads;
  const long long warp_id = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long r0 = warp_id * (2 * rows_per_warp) + sub;
  float acc[2] = {0.f, 0.f};
  uint4 ra[2], rb[2];
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    const long long r = r0 + u * rows_per_warp;
    if (r < total) {
      const int h = (int)(r % heads);
      const long long nl = r / heads;
      ra[u] = ldg_nc_u4(o + nl * ldo + h * ch + li * 8);
      rb[u] = ldg_nc_u4(d_o + nl * lddo + h * ch + li * 8);
    }
  }
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    const long long r = r0 + u * rows_per_warp;
    if (r < total) {
      float a[8], b[8];
      unpack8(ra[u], a);
      unpack8(rb[u], b);
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[u] = fmaf(a[i], b[i], acc[u]);
    }
  }
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    for (int off = G >> 1; off > 0; off >>= 1) acc[u] += __shfl_xor_sync(0xffffffffu, acc[u], off);
    const long long r = r0 + u * rows_per_warp;
    if (li == 0 && r < total) {
      const int h = (int)(r % heads);
      const long long nl = r / heads;
      const int n = (int)(nl / L), l = (int)(nl - (long long)n * L);
      delta[((long long)n * heads + h) * L + l] = acc[u];
    }
  }
}

// ------------------------------------------------------------------ kernel A: dK, dV
template <int D, int DH, int BQ>
struct DkvCfg {
  static constexpr int ST = (D == 256) ? 1 : (D == 128 ? 3 : 2);
  // S^T | dP^T of step j + 1 are computed into a second pair of TMEM column blocks while the compute warps work on
  // step j (round 2): needs two free stages and BQ <= 64 (columns [BQ, 128) and [128 + BQ, 256) are free then)
  static constexpr bool PIPE = ST >= 2 && BQ <= 64;
  static constexpr int KV_BYTES = 128 * D * 2;
  static constexpr int QT_BYTES = BQ * D * 2;
  static constexpr int STAGE_BYTES = 2 * QT_BYTES;   // Q tile + dO tile
  static constexpr int PT_BYTES = 128 * BQ * 2;
  static constexpr int SMEM_DATA = 2 * KV_BYTES + ST * STAGE_BYTES + 2 * PT_BYTES;
  static constexpr int SMEM_TOTAL = SMEM_DATA + 2 * 2 * BQ * 4 + (2 * ST + 5) * 8 + 16 + 1024;
  static constexpr uint32_t ST_COL = 0, DPT_COL = 128, DV_COL = 256, DK_COL = 256 + DH;
};

template <int D, int DH, int BQ>
__global__ void __launch_bounds__(AB_THREADS)
attn_bwd_dkv_kernel(const __grid_constant__ AttnBwdMaps maps, const AttnBwdParams p) {
  using Cfg = DkvCfg<D, DH, BQ>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* k_s = smem;
  uint8_t* v_s = k_s + Cfg::KV_BYTES;
  uint8_t* st_s = v_s + Cfg::KV_BYTES;                       // ring of {Q, dO}
  uint8_t* pt_s = st_s + Cfg::ST * Cfg::STAGE_BYTES;
  uint8_t* dst_s = pt_s + Cfg::PT_BYTES;
  float* lse_s = (float*)(dst_s + Cfg::PT_BYTES);            // [2][BQ]
  float* del_s = lse_s + 2 * BQ;                             // [2][BQ]
  uint64_t* bars = (uint64_t*)(del_s + 2 * BQ);
  uint64_t* kv_full = bars;
  uint64_t* full = bars + 1;
  uint64_t* empty = full + Cfg::ST;
  uint64_t* s_full = empty + Cfg::ST;      // [2]
  uint64_t* p_ready = s_full + 2;
  uint64_t* acc_done = p_ready + 1;
  uint32_t* tmem_slot = (uint32_t*)(acc_done + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int HALVES = D / DH;
  const int half = blockIdx.x % HALVES;
  const int kv0 = (blockIdx.x / HALVES) * 128;
  const int head = blockIdx.y, n = blockIdx.z;
  const int n_q = (p.q_rows + BQ - 1) / BQ;
  const int qc = p.q_off + head * p.head_stride, kc = p.k_off + head * p.head_stride, vc = p.v_off + head * p.head_stride;

  if (threadIdx.x == 0) {
    mbar_init(kv_full, 1);
    for (int s = 0; s < Cfg::ST; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(&s_full[0], 1); mbar_init(&s_full[1], 1); mbar_init(p_ready, AB_CTHREADS); mbar_init(acc_done, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      mbar_expect_tx(kv_full, 2 * Cfg::KV_BYTES);
#pragma unroll
      for (int b = 0; b < D / 64; ++b) {
        tma_load_3d(k_s + b * (128 * 128), &maps.qkv128, kv_full, kc + b * 64, kv0, n);
        tma_load_3d(v_s + b * (128 * 128), &maps.qkv128, kv_full, vc + b * 64, kv0, n);
      }
      for (int j = 0; j < n_q; ++j) {
        const int s = j % Cfg::ST;
        mbar_wait(&empty[s], (uint32_t)(((j / Cfg::ST) & 1) ^ 1));
        uint8_t* q_dst = st_s + s * Cfg::STAGE_BYTES;
        uint8_t* do_dst = q_dst + Cfg::QT_BYTES;
        mbar_expect_tx(&full[s], Cfg::STAGE_BYTES);
#pragma unroll
        for (int b = 0; b < D / 64; ++b) {
          tma_load_3d(q_dst + b * (BQ * 128), &maps.qkvT, &full[s], qc + b * 64, p.q_row0 + j * BQ, n);
          tma_load_3d(do_dst + b * (BQ * 128), &maps.doT, &full[s], head * D + b * 64, p.q_row0 + j * BQ, n);
        }
      }
    }
  } else if (warp == 1) {
    const uint32_t idesc_s = make_idesc(128, BQ, 0, 0, FMT_F16, FMT_F16);    // S^T  = K Q^T   (fp16 x fp16)
    const uint32_t idesc_p = make_idesc(128, BQ, 0, 0, FMT_F16, FMT_F16);   // dP^T = V dO^T  (fp16 x bf16)
    const uint32_t idesc_v = make_idesc(128, DH, 0, 1, FMT_F16, FMT_F16);   // dV  += P^T dO  (fp16 x bf16)
    const uint32_t idesc_k = make_idesc(128, DH, 0, 1, FMT_F16, FMT_F16);   // dK  += dS^T Q  (bf16 x fp16)
    const uint32_t k_addr = smem_u32(k_s), v_addr = smem_u32(v_s);
    const uint32_t pt_addr = smem_u32(pt_s), dst_addr = smem_u32(dst_s);
    mbar_wait(kv_full, 0);
    auto issue_sdp = [&](int j) {             // S^T | dP^T of step j into column block (PIPE ? j & 1 : 0)
      const int s = j % Cfg::ST;
      mbar_wait(&full[s], (uint32_t)((j / Cfg::ST) & 1));
      tc_fence_after();
      const uint32_t q_addr = smem_u32(st_s + s * Cfg::STAGE_BYTES);
      const uint32_t do_addr = q_addr + Cfg::QT_BYTES;
      const uint32_t buf = Cfg::PIPE ? (uint32_t)(j & 1) : 0u;
      if (elect_one()) {
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
          const uint32_t ao = (kk >> 2) * (128 * 128) + (kk & 3) * 32;
          const uint32_t bo = (kk >> 2) * (BQ * 128) + (kk & 3) * 32;
          mma_f16_ss(tmem_base + Cfg::ST_COL + buf * BQ, make_sdesc_sw128(k_addr + ao, 16, 1024),
                      make_sdesc_sw128(q_addr + bo, 16, 1024), idesc_s, (uint32_t)(kk != 0));
        }
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
          const uint32_t ao = (kk >> 2) * (128 * 128) + (kk & 3) * 32;
          const uint32_t bo = (kk >> 2) * (BQ * 128) + (kk & 3) * 32;
          mma_f16_ss(tmem_base + Cfg::DPT_COL + buf * BQ, make_sdesc_sw128(v_addr + ao, 16, 1024),
                      make_sdesc_sw128(do_addr + bo, 16, 1024), idesc_p, (uint32_t)(kk != 0));
        }
        mma_commit(&s_full[buf]);
      }
      __syncwarp();
    };
    if (Cfg::PIPE) issue_sdp(0);
    for (int j = 0; j < n_q; ++j) {
      const int s = j % Cfg::ST;
      // the column block of step j + 1 was last read by the compute warps in step j - 1, whose p_ready this warp has
      // waited for; the stage of step j + 1 is free since the dV / dK MMAs of step j + 1 - ST completed
      if (Cfg::PIPE) { if (j + 1 < n_q) issue_sdp(j + 1); } else issue_sdp(j);
      const uint32_t q_addr = smem_u32(st_s + s * Cfg::STAGE_BYTES);
      const uint32_t do_addr = q_addr + Cfg::QT_BYTES;
      mbar_wait(p_ready, (uint32_t)(j & 1));
      tc_fence_after();
      if (elect_one()) {
        const uint32_t goff = (uint32_t)(half * (DH / 64)) * (BQ * 128);
#pragma unroll
        for (int kk = 0; kk < BQ / 16; ++kk) {
          const uint32_t ao = (kk >> 2) * (128 * 128) + (kk & 3) * 32;
          mma_f16_ss(tmem_base + Cfg::DV_COL, make_sdesc_sw128(pt_addr + ao, 16, 1024),
                      make_sdesc_sw128(do_addr + goff + kk * 2048, BQ * 128, 1024), idesc_v, (uint32_t)((j | kk) != 0));
        }
#pragma unroll
        for (int kk = 0; kk < BQ / 16; ++kk) {
          const uint32_t ao = (kk >> 2) * (128 * 128) + (kk & 3) * 32;
          mma_f16_ss(tmem_base + Cfg::DK_COL, make_sdesc_sw128(dst_addr + ao, 16, 1024),
                      make_sdesc_sw128(q_addr + goff + kk * 2048, BQ * 128, 1024), idesc_k, (uint32_t)((j | kk) != 0));
        }
        mma_commit(&empty[s]);
        mma_commit(acc_done);
      }
      __syncwarp();
    }
  } else {
    const int qd = warp & 3;
    const int chalf = (warp - 2) >> 2;         // which half of the step's BQ query columns this warp takes
    const int row = qd * 32 + lane;            // key index within the tile
    const int ct = threadIdx.x - 64;           // 0..255 among compute threads
    const uint32_t lane_base = tmem_base + ((uint32_t)(qd * 32) << 16);
    const float c = p.scale_log2;
    const float* lse_g = p.lse + ((long long)n * p.heads + head) * p.L;
    const float* del_g = p.delta + ((long long)n * p.heads + head) * p.L;
    for (int j = 0; j < n_q; ++j) {
      float* ls = lse_s + (j & 1) * BQ;
      float* ds = del_s + (j & 1) * BQ;
      if (ct < BQ) {
        const int qi = p.q_row0 + j * BQ + ct;
        ls[ct] = qi < p.L ? __ldg(lse_g + qi) : INFINITY;
        ds[ct] = qi < p.L ? __ldg(del_g + qi) : 0.f;
      }
      named_bar_sync(1, AB_CTHREADS);
      const uint32_t buf = Cfg::PIPE ? (uint32_t)(j & 1) : 0u;
      mbar_wait(&s_full[buf], (uint32_t)(Cfg::PIPE ? ((j >> 1) & 1) : (j & 1)));
      tc_fence_after();
      if (j > 0) mbar_wait(acc_done, (uint32_t)((j - 1) & 1));   // previous dV/dK MMAs have read P^T / dS^T
#pragma unroll
      for (int cc = 0; cc < BQ / 2; cc += 32) {
        const int c0 = chalf * (BQ / 2) + cc;
        uint32_t sr[32], dr[32];
        tmem_ld32(lane_base + Cfg::ST_COL + buf * BQ + c0, sr);
        tmem_ld32(lane_base + Cfg::DPT_COL + buf * BQ + c0, dr);
        tmem_ld_wait();
        float pv[32], dv[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float pe = fast_exp2(fmaf(__uint_as_float(sr[i]), c, -ls[c0 + i]));
          pv[i] = pe;
          dv[i] = pe * (__uint_as_float(dr[i]) - ds[c0 + i]);
        }
        uint8_t* pblk = pt_s + (c0 >> 6) * (128 * 128);
        uint8_t* dblk = dst_s + (c0 >> 6) * (128 * 128);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 u, w;
          u.x = pack_f16x2_nosat(pv[8 * g + 0], pv[8 * g + 1]); u.y = pack_f16x2_nosat(pv[8 * g + 2], pv[8 * g + 3]);
          u.z = pack_f16x2_nosat(pv[8 * g + 4], pv[8 * g + 5]); u.w = pack_f16x2_nosat(pv[8 * g + 6], pv[8 * g + 7]);
          w.x = pack_f16x2(dv[8 * g + 0], dv[8 * g + 1]); w.y = pack_f16x2(dv[8 * g + 2], dv[8 * g + 3]);
          w.z = pack_f16x2(dv[8 * g + 4], dv[8 * g + 5]); w.w = pack_f16x2(dv[8 * g + 6], dv[8 * g + 7]);
          const uint32_t off = sw128_offset((uint32_t)row, (uint32_t)((c0 & 63) + 8 * g));
          sts128(smem_u32(pblk) + off, u);
          sts128(smem_u32(dblk) + off, w);
        }
      }
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(p_ready);
    }
    mbar_wait(acc_done, (uint32_t)((n_q - 1) & 1));
    tc_fence_after();
    const int ki = kv0 + row;
    const bool valid = ki < p.L;
    __half* base = p.dqkv + ((long long)n * p.L + ki) * p.lddqkv;
#pragma unroll
    for (int c0 = chalf * 16; c0 < DH; c0 += 32) {
      uint32_t a[16], b[16];
      tmem_ld16(lane_base + Cfg::DV_COL + c0, a);
      tmem_ld16(lane_base + Cfg::DK_COL + c0, b);
      tmem_ld_wait();
      if (valid) {
        uint4 o0, o1, k0, k1;
        o0.x = pack_f16x2(__uint_as_float(a[0]), __uint_as_float(a[1]));   o0.y = pack_f16x2(__uint_as_float(a[2]), __uint_as_float(a[3]));
        o0.z = pack_f16x2(__uint_as_float(a[4]), __uint_as_float(a[5]));   o0.w = pack_f16x2(__uint_as_float(a[6]), __uint_as_float(a[7]));
        o1.x = pack_f16x2(__uint_as_float(a[8]), __uint_as_float(a[9]));   o1.y = pack_f16x2(__uint_as_float(a[10]), __uint_as_float(a[11]));
        o1.z = pack_f16x2(__uint_as_float(a[12]), __uint_as_float(a[13])); o1.w = pack_f16x2(__uint_as_float(a[14]), __uint_as_float(a[15]));
        const float sc = p.scale;
        k0.x = pack_f16x2(__uint_as_float(b[0]) * sc, __uint_as_float(b[1]) * sc);   k0.y = pack_f16x2(__uint_as_float(b[2]) * sc, __uint_as_float(b[3]) * sc);
        k0.z = pack_f16x2(__uint_as_float(b[4]) * sc, __uint_as_float(b[5]) * sc);   k0.w = pack_f16x2(__uint_as_float(b[6]) * sc, __uint_as_float(b[7]) * sc);
        k1.x = pack_f16x2(__uint_as_float(b[8]) * sc, __uint_as_float(b[9]) * sc);   k1.y = pack_f16x2(__uint_as_float(b[10]) * sc, __uint_as_float(b[11]) * sc);
        k1.z = pack_f16x2(__uint_as_float(b[12]) * sc, __uint_as_float(b[13]) * sc); k1.w = pack_f16x2(__uint_as_float(b[14]) * sc, __uint_as_float(b[15]) * sc);
        __half* pv_ = base + vc + half * DH + c0;
        __half* pk_ = base + kc + half * DH + c0;
        *reinterpret_cast<uint4*>(pv_) = o0; *reinterpret_cast<uint4*>(pv_ + 8) = o1;
        *reinterpret_cast<uint4*>(pk_) = k0; *reinterpret_cast<uint4*>(pk_ + 8) = k1;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

// ------------------------------------------------------------------ kernel B: dQ
template <int D, int BKV>
struct DqCfg {
  static constexpr int ST = (D == 256) ? 1 : (D == 128 ? 3 : 2);
  static constexpr bool PIPE = ST >= 2 && BKV <= 64;   // S | dP of step j + 1 in a second pair of TMEM column blocks
  static constexpr int Q_BYTES = 128 * D * 2;
  static constexpr int T_BYTES = BKV * D * 2;
  static constexpr int STAGE_BYTES = 2 * T_BYTES;   // K tile + V tile
  static constexpr int DS_BYTES = 128 * BKV * 2;
  static constexpr int SMEM_DATA = 2 * Q_BYTES + ST * STAGE_BYTES + DS_BYTES;
  static constexpr int SMEM_TOTAL = SMEM_DATA + (2 * ST + 5) * 8 + 16 + 1024;
  static constexpr uint32_t S_COL = 0, DP_COL = 128, DQ_COL = 256;
};

template <int D, int BKV>
__global__ void __launch_bounds__(AB_THREADS)
attn_bwd_dq_kernel(const __grid_constant__ AttnBwdMaps maps, const AttnBwdParams p) {
  using Cfg = DqCfg<D, BKV>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* q_s = smem;
  uint8_t* do_s = q_s + Cfg::Q_BYTES;
  uint8_t* st_s = do_s + Cfg::Q_BYTES;                 // ring of {K, V}
  uint8_t* ds_s = st_s + Cfg::ST * Cfg::STAGE_BYTES;
  uint64_t* bars = (uint64_t*)(ds_s + Cfg::DS_BYTES);
  uint64_t* q_full = bars;
  uint64_t* full = bars + 1;
  uint64_t* empty = full + Cfg::ST;
  uint64_t* s_full = empty + Cfg::ST;      // [2]
  uint64_t* p_ready = s_full + 2;
  uint64_t* acc_done = p_ready + 1;
  uint32_t* tmem_slot = (uint32_t*)(acc_done + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = p.q_row0 + blockIdx.x * 128;
  const int head = blockIdx.y, n = blockIdx.z;
  const int n_kv = (p.L + BKV - 1) / BKV;
  const int qc = p.q_off + head * p.head_stride, kc = p.k_off + head * p.head_stride, vc = p.v_off + head * p.head_stride;

  if (threadIdx.x == 0) {
    mbar_init(q_full, 1);
    for (int s = 0; s < Cfg::ST; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(&s_full[0], 1); mbar_init(&s_full[1], 1); mbar_init(p_ready, AB_CTHREADS); mbar_init(acc_done, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      mbar_expect_tx(q_full, 2 * Cfg::Q_BYTES);
#pragma unroll
      for (int b = 0; b < D / 64; ++b) {
        tma_load_3d(q_s + b * (128 * 128), &maps.qkv128, q_full, qc + b * 64, q0, n);
        tma_load_3d(do_s + b * (128 * 128), &maps.doR, q_full, head * D + b * 64, q0, n);
      }
      for (int j = 0; j < n_kv; ++j) {
        const int s = j % Cfg::ST;
        mbar_wait(&empty[s], (uint32_t)(((j / Cfg::ST) & 1) ^ 1));
        uint8_t* k_dst = st_s + s * Cfg::STAGE_BYTES;
        uint8_t* v_dst = k_dst + Cfg::T_BYTES;
        mbar_expect_tx(&full[s], Cfg::STAGE_BYTES);
#pragma unroll
        for (int b = 0; b < D / 64; ++b) {
          tma_load_3d(k_dst + b * (BKV * 128), &maps.qkvT, &full[s], kc + b * 64, j * BKV, n);
          tma_load_3d(v_dst + b * (BKV * 128), &maps.qkvT, &full[s], vc + b * 64, j * BKV, n);
        }
      }
    }
  } else if (warp == 1) {
    const uint32_t idesc_s = make_idesc(128, BKV, 0, 0, FMT_F16, FMT_F16);   // S   = Q K^T   (fp16 x fp16)
    const uint32_t idesc_p = make_idesc(128, BKV, 0, 0, FMT_F16, FMT_F16);  // dP  = dO V^T  (bf16 x fp16)
    const uint32_t idesc_g = make_idesc(128, D, 0, 1, FMT_F16, FMT_F16);    // dQ += dS K    (bf16 x fp16)
    const uint32_t q_addr = smem_u32(q_s), do_addr = smem_u32(do_s), ds_addr = smem_u32(ds_s);
    mbar_wait(q_full, 0);
    auto issue_sdp = [&](int j) {             // S | dP of step j into column block (PIPE ? j & 1 : 0)
      const int s = j % Cfg::ST;
      mbar_wait(&full[s], (uint32_t)((j / Cfg::ST) & 1));
      tc_fence_after();
      const uint32_t k_addr = smem_u32(st_s + s * Cfg::STAGE_BYTES);
      const uint32_t v_addr = k_addr + Cfg::T_BYTES;
      const uint32_t buf = Cfg::PIPE ? (uint32_t)(j & 1) : 0u;
      if (elect_one()) {
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
          const uint32_t ao = (kk >> 2) * (128 * 128) + (kk & 3) * 32;
          const uint32_t bo = (kk >> 2) * (BKV * 128) + (kk & 3) * 32;
          mma_f16_ss(tmem_base + Cfg::S_COL + buf * BKV, make_sdesc_sw128(q_addr + ao, 16, 1024),
                      make_sdesc_sw128(k_addr + bo, 16, 1024), idesc_s, (uint32_t)(kk != 0));
        }
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
          const uint32_t ao = (kk >> 2) * (128 * 128) + (kk & 3) * 32;
          const uint32_t bo = (kk >> 2) * (BKV * 128) + (kk & 3) * 32;
          mma_f16_ss(tmem_base + Cfg::DP_COL + buf * BKV, make_sdesc_sw128(do_addr + ao, 16, 1024),
                      make_sdesc_sw128(v_addr + bo, 16, 1024), idesc_p, (uint32_t)(kk != 0));
        }
        mma_commit(&s_full[buf]);
      }
      __syncwarp();
    };
    if (Cfg::PIPE) issue_sdp(0);
    for (int j = 0; j < n_kv; ++j) {
      const int s = j % Cfg::ST;
      if (Cfg::PIPE) { if (j + 1 < n_kv) issue_sdp(j + 1); } else issue_sdp(j);
      const uint32_t k_addr = smem_u32(st_s + s * Cfg::STAGE_BYTES);
      mbar_wait(p_ready, (uint32_t)(j & 1));
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int kk = 0; kk < BKV / 16; ++kk) {
          const uint32_t ao = (kk >> 2) * (128 * 128) + (kk & 3) * 32;
          mma_f16_ss(tmem_base + Cfg::DQ_COL, make_sdesc_sw128(ds_addr + ao, 16, 1024),
                      make_sdesc_sw128(k_addr + kk * 2048, BKV * 128, 1024), idesc_g, (uint32_t)((j | kk) != 0));
        }
        mma_commit(&empty[s]);
        mma_commit(acc_done);
      }
      __syncwarp();
    }
  } else {
    const int qd = warp & 3;
    const int chalf = (warp - 2) >> 2;         // which half of the step's BKV key columns this warp takes
    const int row = qd * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(qd * 32) << 16);
    const float c = p.scale_log2;
    const int qi = q0 + row;
    const bool valid = qi < p.L;
    const long long sidx = ((long long)n * p.heads + head) * p.L + qi;
    const float lse_i = valid ? __ldg(p.lse + sidx) : INFINITY;
    const float del_i = valid ? __ldg(p.delta + sidx) : 0.f;
    for (int j = 0; j < n_kv; ++j) {
      const uint32_t buf = Cfg::PIPE ? (uint32_t)(j & 1) : 0u;
      mbar_wait(&s_full[buf], (uint32_t)(Cfg::PIPE ? ((j >> 1) & 1) : (j & 1)));
      tc_fence_after();
      if (j > 0) mbar_wait(acc_done, (uint32_t)((j - 1) & 1));
#pragma unroll
      for (int cc = 0; cc < BKV / 2; cc += 32) {
        const int c0 = chalf * (BKV / 2) + cc;
        uint32_t sr[32], dr[32];
        tmem_ld32(lane_base + Cfg::S_COL + buf * BKV + c0, sr);
        tmem_ld32(lane_base + Cfg::DP_COL + buf * BKV + c0, dr);
        tmem_ld_wait();
        float dv[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float pe = fast_exp2(fmaf(__uint_as_float(sr[i]), c, -lse_i));
          dv[i] = pe * (__uint_as_float(dr[i]) - del_i);
        }
        uint8_t* dblk = ds_s + (c0 >> 6) * (128 * 128);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 w;
          w.x = pack_f16x2(dv[8 * g + 0], dv[8 * g + 1]); w.y = pack_f16x2(dv[8 * g + 2], dv[8 * g + 3]);
          w.z = pack_f16x2(dv[8 * g + 4], dv[8 * g + 5]); w.w = pack_f16x2(dv[8 * g + 6], dv[8 * g + 7]);
          sts128(smem_u32(dblk) + sw128_offset((uint32_t)row, (uint32_t)((c0 & 63) + 8 * g)), w);
        }
      }
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(p_ready);
    }
    mbar_wait(acc_done, (uint32_t)((n_kv - 1) & 1));
    tc_fence_after();
    __half* dst = p.dqkv + ((long long)n * p.L + qi) * p.lddqkv + qc;
    const float sc = p.scale;
#pragma unroll
    for (int c0 = chalf * 16; c0 < D; c0 += 32) {
      uint32_t a[16];
      tmem_ld16(lane_base + Cfg::DQ_COL + c0, a);
      tmem_ld_wait();
      if (valid) {
        uint4 o0, o1;
        o0.x = pack_f16x2(__uint_as_float(a[0]) * sc, __uint_as_float(a[1]) * sc);   o0.y = pack_f16x2(__uint_as_float(a[2]) * sc, __uint_as_float(a[3]) * sc);
        o0.z = pack_f16x2(__uint_as_float(a[4]) * sc, __uint_as_float(a[5]) * sc);   o0.w = pack_f16x2(__uint_as_float(a[6]) * sc, __uint_as_float(a[7]) * sc);
        o1.x = pack_f16x2(__uint_as_float(a[8]) * sc, __uint_as_float(a[9]) * sc);   o1.y = pack_f16x2(__uint_as_float(a[10]) * sc, __uint_as_float(a[11]) * sc);
        o1.z = pack_f16x2(__uint_as_float(a[12]) * sc, __uint_as_float(a[13]) * sc); o1.w = pack_f16x2(__uint_as_float(a[14]) * sc, __uint_as_float(a[15]) * sc);
        *reinterpret_cast<uint4*>(dst + c0) = o0;
        *reinterpret_cast<uint4*>(dst + c0 + 8) = o1;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

template <int D, int DH, int BQ, int BKV>
static int launch_attn_bwd(const void* qkv, int64_t ldqkv, const void* d_o, int64_t lddo, int N, int L, int heads,
                           AttnBwdParams p, void* stream) {
  using CA = DkvCfg<D, DH, BQ>;
  using CB = DqCfg<D, BKV>;
  B2D_SET_SMEM_ONCE((attn_bwd_dkv_kernel<D, DH, BQ>), CA::SMEM_TOTAL);
  B2D_SET_SMEM_ONCE((attn_bwd_dq_kernel<D, BKV>), CB::SMEM_TOTAL);
  const int C = heads * D;
  const uint64_t dims[3] = {(uint64_t)(3 * C), (uint64_t)L, (uint64_t)N};
  const uint64_t str[2] = {(uint64_t)ldqkv * 2, (uint64_t)ldqkv * 2 * (uint64_t)L};
  const uint64_t ddims[3] = {(uint64_t)C, (uint64_t)L, (uint64_t)N};
  const uint64_t dstr[2] = {(uint64_t)lddo * 2, (uint64_t)lddo * 2 * (uint64_t)L};
  const uint32_t box128[3] = {64u, 128u, 1u};
  int rc;
  cudaStream_t st = (cudaStream_t)stream;
  {  // kernel A: streamed tiles are BQ queries
    AttnBwdMaps m;
    const uint32_t boxT[3] = {64u, (uint32_t)BQ, 1u};
    if ((rc = make_tmap_16b(&m.qkv128, qkv, 3, dims, str, box128))) return rc;
    if ((rc = make_tmap_16b(&m.qkvT, qkv, 3, dims, str, boxT))) return rc;
    if ((rc = make_tmap_16b(&m.doR, d_o, 3, ddims, dstr, box128))) return rc;
    if ((rc = make_tmap_16b(&m.doT, d_o, 3, ddims, dstr, boxT))) return rc;
    dim3 grid((unsigned)(((L + 127) / 128) * (D / DH)), (unsigned)heads, (unsigned)N);
    attn_bwd_dkv_kernel<D, DH, BQ><<<grid, AB_THREADS, CA::SMEM_TOTAL, st>>>(m, p);
    B2D_CHECK_LAUNCH();
  }
  {  // kernel B: streamed tiles are BKV keys
    AttnBwdMaps m;
    const uint32_t boxT[3] = {64u, (uint32_t)BKV, 1u};
    if ((rc = make_tmap_16b(&m.qkv128, qkv, 3, dims, str, box128))) return rc;
    if ((rc = make_tmap_16b(&m.qkvT, qkv, 3, dims, str, boxT))) return rc;
    if ((rc = make_tmap_16b(&m.doR, d_o, 3, ddims, dstr, box128))) return rc;
    m.doT = m.doR;
    dim3 grid((unsigned)((p.q_rows + 127) / 128), (unsigned)heads, (unsigned)N);
    attn_bwd_dq_kernel<D, BKV><<<grid, AB_THREADS, CB::SMEM_TOTAL, st>>>(m, p);
    B2D_CHECK_LAUNCH();
  }
  return B2D_OK;
}

int attention_bwd2_dispatch(const void* qkv, int64_t ldqkv, const void* d_o, int64_t lddo, int N, int L, int heads,
                            int new_order, const float* lse, const float* delta, void* dqkv, int64_t lddqkv,
                            void* stream);   // attention_bwd2.cu
int attention_bwd3_dispatch(const void* qkv, int64_t ldqkv, const void* d_o, int64_t lddo, int N, int L, int heads,
                            int new_order, const float* lse, const float* delta, float* dq_acc, void* dqkv,
                            int64_t lddqkv, int q_row0, int q_rows, void* stream);   // attention_bwd3.cu

}  // namespace b2d

using namespace b2d;

extern "C" int64_t b2d_attention_bwd_ws_floats(int N, int L, int heads, int ch) {
  const int64_t rows = (int64_t)N * heads * L;
  return (ch == 64 && L > 128) ? rows * 65 : rows;   // delta + the fp32 dQ accumulator of the fused kernel
}

extern "C" int b2d_attention_bwd(const void* qkv, int64_t ldqkv, const void* out, int64_t ldo, const void* d_out,
                                 int64_t lddo, const float* lse, int N, int L, int heads, int ch, int new_order,
                                 void* dqkv, int64_t lddqkv, float* delta_ws, void* stream) {
  B2D_REQUIRE(qkv && out && d_out && lse && dqkv && delta_ws, B2D_ERR_SHAPE, "attention_bwd: null pointer");
  B2D_REQUIRE(ch == 64 || ch == 128 || ch == 256, B2D_ERR_SHAPE, "attention_bwd: head dim %d not in {64,128,256}", ch);
  B2D_REQUIRE(N > 0 && L > 0 && heads > 0 && N <= 65535 && heads <= 65535, B2D_ERR_SHAPE, "attention_bwd: bad sizes");
  B2D_REQUIRE(lddqkv % 8 == 0 && aligned16(dqkv), B2D_ERR_ALIGN, "attention_bwd: dqkv rows must be 16-byte aligned");
  const int C = heads * ch;
  AttnBwdParams p;
  p.L = L; p.heads = heads;
  p.q_row0 = 0; p.q_rows = L;
  if (new_order) { p.q_off = 0; p.k_off = C; p.v_off = 2 * C; p.head_stride = ch; }
  else           { p.q_off = 0; p.k_off = ch; p.v_off = 2 * ch; p.head_stride = 3 * ch; }
  p.scale = 1.0f / sqrtf((float)ch);
  p.scale_log2 = 1.4426950408889634f * p.scale;
  p.lse = lse; p.delta = delta_ws;
  p.dqkv = (__half*)dqkv; p.lddqkv = lddqkv;
  {
    const long long rows_per_warp = 2 * (32 / (ch >> 3));
    const long long warps = ((long long)N * L * heads + rows_per_warp - 1) / rows_per_warp;
    const long long blocks = (warps * 32 + 255) / 256;
    attn_delta_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        (const __half*)out, ldo, (const __half*)d_out, lddo, N, L, heads, ch, delta_ws);
    B2D_CHECK_LAUNCH();
  }
  static const bool force_v1 = getenv("B2D_ATTN_BWD_V1") != nullptr;   // A/B timing of the single-tile kernels
  static const bool force_v2 = getenv("B2D_ATTN_BWD_V2") != nullptr;   // A/B timing of the two-kernel variant
  if (!force_v1 && !force_v2 && ch == 64 && L > 128 && L % 4 == 0)   // L % 4: 16-byte pitch of the lse / delta maps
    return attention_bwd3_dispatch(qkv, ldqkv, d_out, lddo, N, L, heads, new_order, lse, delta_ws,
                                   delta_ws + (long long)N * heads * L, dqkv, lddqkv, 0, L, stream);
  if (!force_v1 && ch == 64 && L > 128)
    return attention_bwd2_dispatch(qkv, ldqkv, d_out, lddo, N, L, heads, new_order, lse, delta_ws, dqkv, lddqkv, stream);
  if (ch == 64) return launch_attn_bwd<64, 64, 128, 128>(qkv, ldqkv, d_out, lddo, N, L, heads, p, stream);
  if (ch == 128) return launch_attn_bwd<128, 128, 64, 64>(qkv, ldqkv, d_out, lddo, N, L, heads, p, stream);
  return launch_attn_bwd<256, 128, 64, 64>(qkv, ldqkv, d_out, lddo, N, L, heads, p, stream);
}

// Sequence-parallel shard of the fused single-pass backward (head dim 64): only query rows
// [q_row0, q_row0 + q_rows) are visited.  dqkv is fully written: the Q columns hold dQ for those rows and ZERO elsewhere,
// the K / V columns hold this shard's PARTIAL dK / dV -- the sum of the shards' dqkv tensors is the full gradient.
extern "C" int b2d_attention_bwd_rows(const void* qkv, int64_t ldqkv, const void* out, int64_t ldo, const void* d_out,
                                      int64_t lddo, const float* lse, int N, int L, int heads, int ch, int new_order,
                                      int q_row0, int q_rows, void* dqkv, int64_t lddqkv, float* delta_ws, void* stream) {
  B2D_REQUIRE(qkv && out && d_out && lse && dqkv && delta_ws, B2D_ERR_SHAPE, "attention_bwd_rows: null pointer");
  B2D_REQUIRE((ch == 64 && L > 128 && L % 4 == 0) || ch == 128 || ch == 256, B2D_ERR_SHAPE,
              "attention_bwd_rows: head dim 64 (L > 128) / 128 / 256 only (ch=%d L=%d)", ch, L);
  B2D_REQUIRE(N > 0 && heads > 0 && N <= 65535 && heads <= 65535, B2D_ERR_SHAPE, "attention_bwd_rows: bad sizes");
  B2D_REQUIRE(q_row0 >= 0 && q_rows > 0 && q_row0 % 128 == 0 && q_rows % 128 == 0 && q_row0 + q_rows <= L, B2D_ERR_SHAPE,
              "attention_bwd_rows: rows [%d, +%d) must be 128-aligned inside L=%d", q_row0, q_rows, L);
  B2D_REQUIRE(lddqkv % 8 == 0 && aligned16(dqkv), B2D_ERR_ALIGN, "attention_bwd_rows: dqkv rows must be 16-byte aligned");
  {
    const long long rows_per_warp = 2 * (32 / (ch >> 3));
    const long long warps = ((long long)N * L * heads + rows_per_warp - 1) / rows_per_warp;
    const long long blocks = (warps * 32 + 255) / 256;
    attn_delta_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        (const __half*)out, ldo, (const __half*)d_out, lddo, N, L, heads, ch, delta_ws);
    B2D_CHECK_LAUNCH();
  }
  if (ch == 64)
    return attention_bwd3_dispatch(qkv, ldqkv, d_out, lddo, N, L, heads, new_order, lse, delta_ws,
                                   delta_ws + (long long)N * heads * L, dqkv, lddqkv, q_row0, q_rows, stream);
  // head dim 128 / 256: the two-kernel variant -- dK / dV kernel over the shard's query tiles, dQ kernel on the shard's
  // query tiles only; the Q columns of the other rows are zeroed first
  B2D_CHECK_CUDA(cudaMemsetAsync(dqkv, 0, (size_t)N * L * lddqkv * 2, (cudaStream_t)stream));
  const int C = heads * ch;
  AttnBwdParams p;
  p.L = L; p.heads = heads;
  p.q_row0 = q_row0; p.q_rows = q_rows;
  if (new_order) { p.q_off = 0; p.k_off = C; p.v_off = 2 * C; p.head_stride = ch; }
  else           { p.q_off = 0; p.k_off = ch; p.v_off = 2 * ch; p.head_stride = 3 * ch; }
  p.scale = 1.0f / sqrtf((float)ch);
  p.scale_log2 = 1.4426950408889634f * p.scale;
  p.lse = lse; p.delta = delta_ws;
  p.dqkv = (__half*)dqkv; p.lddqkv = lddqkv;
  if (ch == 128) return launch_attn_bwd<128, 128, 64, 64>(qkv, ldqkv, d_out, lddo, N, L, heads, p, stream);
  return launch_attn_bwd<256, 128, 64, 64>(qkv, ldqkv, d_out, lddo, N, L, heads, p, stream);
}
