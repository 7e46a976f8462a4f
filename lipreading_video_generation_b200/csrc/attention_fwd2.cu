// attention_fwd2.cu -- flash-style self-attention forward, two query tiles per CTA (QKVAttentionLegacy /
// QKVAttention, unet.py:340-405).  Used for head dim 64 / 128 when L > 128; attention_fwd.cu keeps the
// single-tile variant for head dim 256 and tiny L.
//
// One CTA = 256 queries (two 128-row tiles) of one (sample, head); K/V tiles are streamed ONCE for both.
// Warp roles (576 threads): warp 0 = TMA producer, warp 1 = tcgen05 issuer, warps 2..9 = softmax group 0
// (query tile 0), warps 10..17 = softmax group 1 (query tile 1).  TMEM: S0 | S1 | O0 | O1 | P0 | P1.  While one group
// runs exp2 on its S tile the tensor pipe computes the other tile's S = Q K^T and O += P V, so the MUFU and
// the tensor pipe overlap instead of alternating (the single-tile kernel spent ~7.9k clk per 128x128 tile
// against a ~1k clk MUFU floor: profiles/r1_attention_ncu.md).
//
// TWO threads per query row (round 2): a group is 8 warps, warp (qd, half) owns rows 32 qd .. 32 qd + 31 and the key
// columns [half * BKV/2, (half + 1) * BKV/2) of the tile.  Head dim 64 is bound by the 16 ex2/clk/SM of the MUFU
// unit (1024 clk per 128x128 tile vs 512 clk of MMA), and one warp per sub-partition issuing FFMA -> EX2 pairs reaches
// only 9.5-10 clk per warp-ex2 against the unit's 8 (scripts/micro/mufu_rate.cu); with one thread per row each
// sub-partition held two softmax warps whose load / max / pack phases left the unit idle ~1/3 of the time (clock64
// trace, profiles/r2_attention_fwd.md).  Four warps per sub-partition with half the columns each keep it fed, and the
// row fits 64 instead of 128 registers.  The two halves of a row exchange their tile maximum through shared memory
// (one 64-thread named barrier per tile); sums stay per thread until the epilogue.
//
// Per KV tile j and query tile t:   s_full[t] (S ready) -> group t loads S into registers -> s_free[t]
// (issuer may overwrite S_t with tile j+1) -> max / lazy O rescale / exp2 / P -> tensor memory -> p_full[t]
// -> issuer: O_t += P_t V_j (A from tensor memory) -> pv_done[t] (P_t columns and O_t reusable).
#include "common.cuh"
#include "tmap.cuh"
#include "umma.cuh"
#include <stdlib.h>

namespace b2d {
using namespace umma;

struct AttnFwd2Params {
  int L, heads;
  int q_row0, q_rows;   // query rows [q_row0, q_row0 + q_rows) are computed (sequence-parallel shard); keys: all L
  int q_off, k_off, v_off, head_stride;
  float scale_log2;
  __half* out; long long ldo;
  float* lse;
  // split-KV (small grids: batch-1 sampling): blockIdx.z = n * kv_splits + split, split s visits key tiles
  // [s * kv_tiles_per_split, ...) and writes its normalised fp32 output / lse to part_o / part_lse; attn_combine_kernel
  // merges the splits.  kv_splits == 1: the kernel writes out / lse itself.
  int kv_splits, kv_tiles_per_split;
  float* part_o;      // [split][n][L][heads * D]
  float* part_lse;    // [split][n][heads][L]
  long long* trace;     // developer builds (-DFWD2_TRACE): clock64 stamps of CTA (0,0,0), [17 roles][64 steps][8]
};

#ifdef FWD2_TRACE
#define FWD2_STAMP(role, j, k, dep)                                                                        \
  do {                                                                                                     \
    if (p.trace && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (j) < 64 && lane == 0) {      \
      long long _t;                                                                                        \
      asm volatile("mov.u64 %0, %%clock64;" : "=l"(_t) : "r"(dep) : "memory");                             \
      p.trace[((role) * 64 + (j)) * 8 + (k)] = _t;                                                         \
    }                                                                                                      \
  } while (0)
#else
#define FWD2_STAMP(role, j, k, dep) do { } while (0)
#endif

struct AttnFwd2Maps {
  CUtensorMap q;   // box (64, 128, 1)
  CUtensorMap kv;  // box (64, BKV, 1)
};

constexpr int ATT2_THREADS = 576;

template <int D, int BKV>
struct Fwd2Cfg {
  static constexpr int ST = 3;
  static constexpr int QT_BYTES = 128 * D * 2;          // one query tile
  static constexpr int K_BYTES = BKV * D * 2;
  static constexpr int STAGE_BYTES = 2 * K_BYTES;       // K + V
  static constexpr int PT_BYTES = 128 * BKV * 2;        // one P tile
  static constexpr int SMEM_DATA = 2 * QT_BYTES + ST * STAGE_BYTES;
  static constexpr int NBARS = 1 + 3 * ST + 8;
  static constexpr int XCH_BYTES = 2 * 2 * 2 * 128 * 4;   // [group][tile parity][half][row] fp32: row-max exchange
  static constexpr int SMEM_TOTAL = SMEM_DATA + NBARS * 8 + 16 + XCH_BYTES + 1024;
  static constexpr uint32_t S_COL = 0;                  // S_t at t * 128
  static constexpr uint32_t O_COL = 256;                // O_t at 256 + t * D
  // P_t as fp16 pairs (BKV / 2 columns), the tensor-memory A operand of O += P V: behind O for D = 64, in the
  // unused half of the S_t slot for D = 128
  static constexpr uint32_t P_COL = (D == 64) ? 384 : 64;
  static constexpr uint32_t P_STRIDE = (D == 64) ? 64 : 128;
  static_assert(BKV <= 128 && 256 + 2 * D <= 512, "TMEM budget");
  static_assert((D == 64 && BKV == 128) || (D == 128 && BKV == 64), "P placement");
};

template <int D, int BKV>
__global__ void __launch_bounds__(ATT2_THREADS, 1)
attn_fwd2_kernel(const __grid_constant__ AttnFwd2Maps maps, const AttnFwd2Params p) {
  using Cfg = Fwd2Cfg<D, BKV>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* q_s = smem;
  uint8_t* kv_s = q_s + 2 * Cfg::QT_BYTES;
  uint64_t* bars = (uint64_t*)(kv_s + Cfg::ST * Cfg::STAGE_BYTES);
  uint64_t* q_full = bars;
  uint64_t* k_full = bars + 1;
  uint64_t* v_full = k_full + Cfg::ST;
  uint64_t* kv_empty = v_full + Cfg::ST;
  uint64_t* s_full = kv_empty + Cfg::ST;   // [2]
  uint64_t* s_free = s_full + 2;           // [2]
  uint64_t* p_full = s_free + 2;           // [2]
  uint64_t* pv_done = p_full + 2;          // [2]
  uint32_t* tmem_slot = (uint32_t*)(pv_done + 2);
  float* xch = (float*)(tmem_slot + 4);      // 16-byte aligned behind the barriers

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int q0 = p.q_row0 + blockIdx.x * 256;
  const int head = blockIdx.y;
  const int n = (int)blockIdx.z / p.kv_splits;
  const int split = (int)blockIdx.z - n * p.kv_splits;
  const int j0 = split * p.kv_tiles_per_split;                      // first key tile of this CTA
  const int n_kv = min(p.kv_tiles_per_split, (p.L + BKV - 1) / BKV - j0);
  const int qc = p.q_off + head * p.head_stride;
  const int kc = p.k_off + head * p.head_stride;
  const int vc = p.v_off + head * p.head_stride;

  if (threadIdx.x == 0) {
    mbar_init(q_full, 1);
    for (int s = 0; s < Cfg::ST; ++s) { mbar_init(&k_full[s], 1); mbar_init(&v_full[s], 1); mbar_init(&kv_empty[s], 1); }
    for (int t = 0; t < 2; ++t) {
      mbar_init(&s_full[t], 1); mbar_init(&s_free[t], 256); mbar_init(&p_full[t], 256); mbar_init(&pv_done[t], 1);
    }
    fence_barrier_init();
    prefetch_tmap(&maps.q);
    prefetch_tmap(&maps.kv);
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ---------------------------------------------------------------- TMA producer
    if (elect_one()) {
      mbar_expect_tx(q_full, 2 * Cfg::QT_BYTES);
#pragma unroll
      for (int t = 0; t < 2; ++t)
#pragma unroll
        for (int b = 0; b < D / 64; ++b)
          tma_load_3d(q_s + t * Cfg::QT_BYTES + b * (128 * 128), &maps.q, q_full, qc + b * 64, q0 + t * 128, n);
      for (int j = 0; j < n_kv; ++j) {
        const int s = j % Cfg::ST;
        mbar_wait_sleep(&kv_empty[s], (uint32_t)(((j / Cfg::ST) & 1) ^ 1));
        uint8_t* k_dst = kv_s + s * Cfg::STAGE_BYTES;
        uint8_t* v_dst = k_dst + Cfg::K_BYTES;
        mbar_expect_tx(&k_full[s], Cfg::K_BYTES);
#pragma unroll
        for (int b = 0; b < D / 64; ++b) tma_load_3d(k_dst + b * (BKV * 128), &maps.kv, &k_full[s], kc + b * 64, (j0 + j) * BKV, n);
        mbar_expect_tx(&v_full[s], Cfg::K_BYTES);
#pragma unroll
        for (int b = 0; b < D / 64; ++b) tma_load_3d(v_dst + b * (BKV * 128), &maps.kv, &v_full[s], vc + b * 64, (j0 + j) * BKV, n);
      }
    }
  } else if (warp == 1) {
    // ---------------------------------------------------------------- tcgen05 issuer
    const uint32_t idesc_s = make_idesc(128, BKV, 0, 0, FMT_F16, FMT_F16);  // S = Q K^T, both K-major
    const uint32_t idesc_o = make_idesc(128, D, 0, 1, FMT_F16, FMT_F16);    // O += P V, V MN-major
    const uint32_t q_addr = smem_u32(q_s);
    auto issue_s = [&](int j, int t) {     // S_t(j) = Q_t K_j^T ; caller has waited for K_j and for S_t to be free
      if (elect_one()) {
        const uint32_t k_addr = smem_u32(kv_s + (j % Cfg::ST) * Cfg::STAGE_BYTES);
        const uint32_t d_tmem = tmem_base + Cfg::S_COL + (uint32_t)t * 128u;
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
          const uint32_t a = q_addr + t * Cfg::QT_BYTES + (kk >> 2) * (128 * 128) + (kk & 3) * 32;
          const uint32_t b = k_addr + (kk >> 2) * (BKV * 128) + (kk & 3) * 32;
          mma_f16_ss(d_tmem, make_sdesc_sw128(a, 16, 1024), make_sdesc_sw128(b, 16, 1024), idesc_s, (uint32_t)(kk != 0));
        }
        mma_commit(&s_full[t]);
      }
      __syncwarp();
    };
    auto issue_pv = [&](int j, int t, bool last_reader) {
      if (elect_one()) {
        const int s = j % Cfg::ST;
        const uint32_t v_addr = smem_u32(kv_s + s * Cfg::STAGE_BYTES + Cfg::K_BYTES);
        const uint32_t d_tmem = tmem_base + Cfg::O_COL + (uint32_t)(t * D);
        // A = P_t from tensor memory (an N = 64 instruction with both operands in shared memory wants 192 B/clk of
        // a 128 B/clk port: profiles/r1_attention_ncu.md), B = V_j MN-major from shared memory
        const uint32_t a_tmem = tmem_base + Cfg::P_COL + (uint32_t)t * Cfg::P_STRIDE;
#pragma unroll
        for (int kk = 0; kk < BKV / 16; ++kk) {
          const uint32_t b = v_addr + kk * 2048;
          mma_f16_ts(d_tmem, a_tmem + (uint32_t)kk * 8u, make_sdesc_sw128(b, BKV * 128, 1024), idesc_o,
                     (uint32_t)((j | kk) != 0));
        }
        mma_commit(&pv_done[t]);
        if (last_reader) mma_commit(&kv_empty[s]);
      }
      __syncwarp();
    };
    mbar_wait_sleep(q_full, 0);
    mbar_wait_sleep(&k_full[0], 0);
    tc_fence_after();
    issue_s(0, 0);
    issue_s(0, 1);
    // (An event loop that served the two groups independently, whichever barrier completed first, was tried: no gain at
    // head dim 64, -6 % at head dim 128 -- profiles/r2_attention_fwd.md.)
    for (int j = 0; j < n_kv; ++j) {
      const uint32_t par = (uint32_t)(j & 1);
      const bool more = j + 1 < n_kv;
      if (more) mbar_wait_sleep(&k_full[(j + 1) % Cfg::ST], (uint32_t)(((j + 1) / Cfg::ST) & 1));
      mbar_wait_sleep(&v_full[j % Cfg::ST], (uint32_t)((j / Cfg::ST) & 1));
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        FWD2_STAMP(16, j, 4 * t + 0, 0u);
        if (more) {
          mbar_wait_sleep(&s_free[t], par);
          tc_fence_after();
          FWD2_STAMP(16, j, 4 * t + 1, 0u);
          issue_s(j + 1, t);
        }
        FWD2_STAMP(16, j, 4 * t + 2, 0u);
        mbar_wait_sleep(&p_full[t], par);
        tc_fence_after();
        FWD2_STAMP(16, j, 4 * t + 3, 0u);
        issue_pv(j, t, t == 1);
      }
    }
  } else {
    // ---------------------------------------------------------------- softmax groups / epilogue
    constexpr int HC = BKV / 2;             // key columns of the tile owned by this thread
    const int t = (warp - 2) >> 3;          // query tile of this group
    const int g = (warp - 2) & 7;
    const int qd = warp & 3;                // TMEM lane quarter this warp may touch
    const int half = g >> 2;                // which half of the tile's key columns
    const int row = qd * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(qd * 32) << 16);
    const uint32_t s_addr = lane_base + Cfg::S_COL + (uint32_t)t * 128u + (uint32_t)(half * HC);
    const uint32_t o_addr = lane_base + Cfg::O_COL + (uint32_t)(t * D) + (uint32_t)(half * (D / 2));
    const uint32_t p_taddr = lane_base + Cfg::P_COL + (uint32_t)t * Cfg::P_STRIDE + (uint32_t)(half * (HC / 2));
    const int pair_bar = 1 + t * 4 + qd;    // named barrier of the two warps that share these 32 rows
    float* x_mine = xch + (t * 2 + half) * 128 + row;           // + (tile parity) * 512
    float* x_peer = xch + (t * 2 + (half ^ 1)) * 128 + row;
    const float c = p.scale_log2;
    float m_used = -INFINITY;
    float l_sum = 0.f;                      // this thread's half of the row sum
    for (int j = 0; j < n_kv; ++j) {
      FWD2_STAMP(t * 8 + g, j, 0, 0u);
      mbar_wait(&s_full[t], (uint32_t)(j & 1));
      tc_fence_after();
      FWD2_STAMP(t * 8 + g, j, 1, 0u);
      uint32_t sr[HC];                     // this thread's part of the row of S, kept in registers (raw fp32 bits)
#pragma unroll
      for (int c0 = 0; c0 < HC; c0 += 32) tmem_ld32(s_addr + c0, sr + c0);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&s_free[t]);             // S_t may be overwritten by the next tile's Q K^T
      FWD2_STAMP(t * 8 + g, j, 2, sr[0]);
      const int kv_valid = p.L - (j0 + j) * BKV - half * HC;  // keys of this part that exist (>= HC except on the last tile)
      if (kv_valid < HC) {
#pragma unroll
        for (int i = 0; i < HC; ++i) if (i >= kv_valid) sr[i] = 0xff800000u;   // -inf
      }
      float mxa[8];                        // 8 independent chains
#pragma unroll
      for (int i = 0; i < 8; ++i) mxa[i] = __uint_as_float(sr[i]);
#pragma unroll
      for (int i = 8; i < HC; i += 8) {
#pragma unroll
        for (int u = 0; u < 8; ++u) mxa[u] = fmaxf(mxa[u], __uint_as_float(sr[i + u]));
      }
      float mx = fmaxf(fmaxf(fmaxf(mxa[0], mxa[1]), fmaxf(mxa[2], mxa[3])),
                       fmaxf(fmaxf(mxa[4], mxa[5]), fmaxf(mxa[6], mxa[7])));
      // row maximum over both halves: exchange through shared memory (slot alternates with the tile parity, so one
      // barrier per tile suffices: a slot is rewritten two tiles later, after the peer has passed the next barrier)
      x_mine[(j & 1) * 512] = mx;
      asm volatile("bar.sync %0, 64;" :: "r"(pair_bar) : "memory");
      mx = fmaxf(mx, x_peer[(j & 1) * 512]);
      FWD2_STAMP(t * 8 + g, j, 3, __float_as_uint(mx));
      // previous O_t += P_t V must be complete before P_t is overwritten or O_t rescaled.  (Waiting only after the
      // exponentials -- the MUFU work does not need it -- was tried and LOST 5 %: this wait is what staggers the two
      // groups' exp2 phases; profiles/r2_attention_fwd.md.)
      if (j > 0) { mbar_wait(&pv_done[t], (uint32_t)((j - 1) & 1)); tc_fence_after(); }
      FWD2_STAMP(t * 8 + g, j, 4, 0u);
      const bool grow = (mx - m_used) * c > 8.0f;   // lazy rescale: move the offset only for > 2^8 growth
      if (j == 0) {
        m_used = mx;
      } else if (__any_sync(0xffffffffu, grow)) {   // both warps of the pair see the same 32 row maxima
        const float m_new = fmaxf(m_used, mx);
        const float f = fast_exp2((m_used - m_new) * c);
        m_used = m_new;
        l_sum *= f;
#pragma unroll
        for (int c0 = 0; c0 < D / 2; c0 += 16) {    // this half's columns of O_t
          uint32_t r[16];
          tmem_ld16(o_addr + c0, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * f);
          tmem_st16(o_addr + c0, r);
        }
        tmem_st_wait();
      }
      const float mc = m_used * c;
      FWD2_STAMP(t * 8 + g, j, 5, 0u);
#pragma unroll
      for (int i = 0; i < HC; ++i)         // exp2(-inf) = 0 for masked keys
        sr[i] = __float_as_uint(fast_exp2(fmaf(__uint_as_float(sr[i]), c, -mc)));
      FWD2_STAMP(t * 8 + g, j, 6, sr[0] ^ sr[HC - 1] ^ sr[HC / 2]);
      float lsa[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int i = 0; i < HC; i += 8) {
#pragma unroll
        for (int u = 0; u < 8; ++u) lsa[u] += __uint_as_float(sr[i + u]);
      }
      l_sum += ((lsa[0] + lsa[1]) + (lsa[2] + lsa[3])) + ((lsa[4] + lsa[5]) + (lsa[6] + lsa[7]));
#pragma unroll
      for (int c0 = 0; c0 < HC / 2; c0 += 16) {        // P_t -> tensor memory, two fp16 per column
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 16; ++i)
          pk[i] = pack_f16x2_nosat(__uint_as_float(sr[2 * (c0 + i)]), __uint_as_float(sr[2 * (c0 + i) + 1]));
        tmem_st16(p_taddr + (uint32_t)c0, pk);
      }
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(&p_full[t]);
      FWD2_STAMP(t * 8 + g, j, 7, 0u);
    }
    // epilogue: O / l -> fp16 global, lse.  Row sum = this thread's half + the peer's.
    x_mine[(n_kv & 1) * 512] = l_sum;
    asm volatile("bar.sync %0, 64;" :: "r"(pair_bar) : "memory");
    l_sum += x_peer[(n_kv & 1) * 512];
    mbar_wait(&pv_done[t], (uint32_t)((n_kv - 1) & 1));
    tc_fence_after();
    const int qi = q0 + t * 128 + row;
    const bool valid = qi < p.L;
    const float inv_l = 1.0f / l_sum;
    if (p.kv_splits > 1) {
      const int n_img = (int)gridDim.z / p.kv_splits;
      // this split's normalised output in fp32 + its log-sum-exp; merged by attn_combine_kernel
      float* op = p.part_o + (((long long)split * n_img + n) * p.L + qi) * (p.heads * D) + head * D + half * (D / 2);
#pragma unroll
      for (int c0 = 0; c0 < D / 2; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(o_addr + c0, r);
        tmem_ld_wait();
        if (valid) {
#pragma unroll
          for (int gg = 0; gg < 8; ++gg)
            *reinterpret_cast<float4*>(op + c0 + 4 * gg) =
                make_float4(__uint_as_float(r[4 * gg]) * inv_l, __uint_as_float(r[4 * gg + 1]) * inv_l,
                            __uint_as_float(r[4 * gg + 2]) * inv_l, __uint_as_float(r[4 * gg + 3]) * inv_l);
        }
      }
      if (half == 0 && valid)
        p.part_lse[(((long long)split * n_img + n) * p.heads + head) * p.L + qi] = m_used * c + log2f(l_sum);
    } else {
    __half* op = p.out + ((long long)n * p.L + qi) * p.ldo + head * D + half * (D / 2);
#pragma unroll
    for (int c0 = 0; c0 < D / 2; c0 += 32) {
      uint32_t r[32];
      tmem_ld32(o_addr + c0, r);
      tmem_ld_wait();
      if (valid) {
#pragma unroll
        for (int gg = 0; gg < 4; ++gg) {
          uint4 o;
          o.x = pack_f16x2(__uint_as_float(r[8 * gg + 0]) * inv_l, __uint_as_float(r[8 * gg + 1]) * inv_l);
          o.y = pack_f16x2(__uint_as_float(r[8 * gg + 2]) * inv_l, __uint_as_float(r[8 * gg + 3]) * inv_l);
          o.z = pack_f16x2(__uint_as_float(r[8 * gg + 4]) * inv_l, __uint_as_float(r[8 * gg + 5]) * inv_l);
          o.w = pack_f16x2(__uint_as_float(r[8 * gg + 6]) * inv_l, __uint_as_float(r[8 * gg + 7]) * inv_l);
          *reinterpret_cast<uint4*>(op + c0 + 8 * gg) = o;
        }
      }
    }
    if (half == 0 && valid && p.lse) p.lse[((long long)n * p.heads + head) * p.L + qi] = m_used * c + log2f(l_sum);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

// out[n, l, head, :] = sum_s w_s o_s,  w_s = 2^(lse_s - lse),  lse = log2 sum_s 2^lse_s  (one thread per 8 channels)
__global__ void __launch_bounds__(256)
attn_combine_kernel(const float* __restrict__ part_o, const float* __restrict__ part_lse, int splits, int N, int L, int heads,
                    int D, __half* __restrict__ out, long long ldo, float* __restrict__ lse) {
  const int G = D >> 3;
  const long long total = (long long)N * L * heads * G;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int g = (int)(i % G);
  const long long r = i / G;                 // (n, l, head)
  const int head = (int)(r % heads);
  const long long nl = r / heads;
  const int n = (int)(nl / L), l = (int)(nl - (long long)n * L);
  float ls[8];
  float m = -INFINITY;
  for (int s = 0; s < splits; ++s) {
    ls[s] = part_lse[(((long long)s * N + n) * heads + head) * L + l];
    m = fmaxf(m, ls[s]);
  }
  float tot = 0.f;
  for (int s = 0; s < splits; ++s) { ls[s] = exp2f(ls[s] - m); tot += ls[s]; }
  const float inv = 1.0f / tot;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int s = 0; s < splits; ++s) {
    const float* src = part_o + (((long long)s * N + n) * L + l) * (heads * D) + head * D + g * 8;
    const float4 a = *reinterpret_cast<const float4*>(src), b = *reinterpret_cast<const float4*>(src + 4);
    const float w = ls[s] * inv;
    acc[0] = fmaf(w, a.x, acc[0]); acc[1] = fmaf(w, a.y, acc[1]); acc[2] = fmaf(w, a.z, acc[2]); acc[3] = fmaf(w, a.w, acc[3]);
    acc[4] = fmaf(w, b.x, acc[4]); acc[5] = fmaf(w, b.y, acc[5]); acc[6] = fmaf(w, b.z, acc[6]); acc[7] = fmaf(w, b.w, acc[7]);
  }
  uint4 o;
  o.x = pack_f16x2(acc[0], acc[1]); o.y = pack_f16x2(acc[2], acc[3]);
  o.z = pack_f16x2(acc[4], acc[5]); o.w = pack_f16x2(acc[6], acc[7]);
  *reinterpret_cast<uint4*>(out + ((long long)n * L + l) * ldo + head * D + g * 8) = o;
  if (g == 0 && lse) lse[((long long)n * heads + head) * L + l] = m + log2f(tot);
}

template <int D, int BKV>
static int launch_attn_fwd2(const void* qkv, int64_t ldqkv, int N, int C3, const AttnFwd2Params& p, void* stream) {
  using Cfg = Fwd2Cfg<D, BKV>;
  B2D_SET_SMEM_ONCE((attn_fwd2_kernel<D, BKV>), Cfg::SMEM_TOTAL);
  AttnFwd2Maps maps;
  const uint64_t dims[3] = {(uint64_t)C3, (uint64_t)p.L, (uint64_t)N};
  const uint64_t str[2] = {(uint64_t)ldqkv * 2, (uint64_t)ldqkv * 2 * (uint64_t)p.L};
  const uint32_t boxq[3] = {64u, 128u, 1u};
  const uint32_t boxkv[3] = {64u, (uint32_t)BKV, 1u};
  int rc = make_tmap_16b(&maps.q, qkv, 3, dims, str, boxq);
  if (rc) return rc;
  rc = make_tmap_16b(&maps.kv, qkv, 3, dims, str, boxkv);
  if (rc) return rc;
  B2D_REQUIRE((long long)N * p.kv_splits <= 65535, B2D_ERR_SHAPE, "attention_fwd: N * splits = %lld exceeds 65535",
              (long long)N * p.kv_splits);
  dim3 grid((unsigned)((p.q_rows + 255) / 256), (unsigned)p.heads, (unsigned)(N * p.kv_splits));
  attn_fwd2_kernel<D, BKV><<<grid, ATT2_THREADS, Cfg::SMEM_TOTAL, (cudaStream_t)stream>>>(maps, p);
  B2D_CHECK_LAUNCH();
  if (p.kv_splits > 1) {
    const long long total = (long long)N * p.L * p.heads * (D / 8);
    attn_combine_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        p.part_o, p.part_lse, p.kv_splits, N, p.L, p.heads, D, p.out, p.ldo, p.lse);
    B2D_CHECK_LAUNCH();
  }
  return B2D_OK;
}

// called by b2d_attention_fwd (attention_fwd.cu) for ch in {64, 128} and L > 128
int attention_fwd2_dispatch(const void* qkv, int64_t ldqkv, int N, int L, int heads, int ch, int new_order, void* out,
                            int64_t ldo, float* lse, int q_row0, int q_rows, void* stream, int kv_splits, float* ws) {
  const int C = heads * ch;
  AttnFwd2Params p;
  p.L = L; p.heads = heads;
  p.q_row0 = q_row0; p.q_rows = q_rows;
  if (new_order) { p.q_off = 0; p.k_off = C; p.v_off = 2 * C; p.head_stride = ch; }          // unet.py:392
  else           { p.q_off = 0; p.k_off = ch; p.v_off = 2 * ch; p.head_stride = 3 * ch; }    // unet.py:359
  p.scale_log2 = 1.4426950408889634f / sqrtf((float)ch);
  p.out = (__half*)out; p.ldo = ldo; p.lse = lse;
  {
    const int bkv = ch == 64 ? 128 : 64;
    const int tiles = (L + bkv - 1) / bkv;
    if (kv_splits < 1 || !ws) kv_splits = 1;
    if (kv_splits > 8) kv_splits = 8;
    if (kv_splits > tiles) kv_splits = tiles;
    p.kv_tiles_per_split = (tiles + kv_splits - 1) / kv_splits;
    p.kv_splits = (tiles + p.kv_tiles_per_split - 1) / p.kv_tiles_per_split;     // no empty split
    p.part_o = ws;
    p.part_lse = ws ? ws + (size_t)p.kv_splits * N * L * heads * ch : nullptr;
  }
  p.trace = nullptr;
#ifdef FWD2_TRACE
  { const char* e = getenv("B2D_FWD2_TRACE_PTR"); if (e) p.trace = (long long*)strtoull(e, nullptr, 0); }
#endif
  if (ch == 64) return launch_attn_fwd2<64, 128>(qkv, ldqkv, N, 3 * C, p, stream);
  return launch_attn_fwd2<128, 64>(qkv, ldqkv, N, 3 * C, p, stream);
}

}  // namespace b2d
