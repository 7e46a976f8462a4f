// attention_bwd3.cu -- single-pass self-attention backward for head dim 64 (L > 128): one kernel produces dQ, dK
// and dV from one evaluation of S and dP per (key tile, query tile) pair -- 5 GEMMs per pair, the algorithmic
// minimum -- instead of the 7 of the dkv2 + dq2 pair in attention_bwd2.cu, which recomputes S and dP once per
// kernel.  (QKVAttentionLegacy unet.py:349-366 under autograd + the checkpoint recompute of utils.py:179-207.)
//
//   CTA = 128 keys (K_t, V_t resident), streams 64-query tiles j; compute group b = j & 1 owns the step:
//        S^T = K_t Q_j^T, dP^T = V_t dO_j^T                 (M = 128 keys, N = 64 queries, K = 64 ch)
//        P^T = exp2(c S^T - lse_q), dS^T = P^T (dP^T - delta_q)   -> fp16 smem (per group buffers)
//        dV += P^T dO_j, dK += dS^T Q_j                      (M = 128 keys, N = 64 ch, K = 64 queries)
//   after every odd step (one "pair" m = j >> 1 of 128 queries) the two dS^T buffers, read through an
//   MN-major descriptor, form the 128 x 128 operand of
//        dQ_m = dS_m K_t                                     (M = 128 queries, N = 64 ch, K = 128 keys)
//   whose fp32 tile is staged in shared memory and added to a global fp32 dQ accumulator by TMA bulk reductions
//   (every key-tile CTA of the image contributes; per-thread red.global of row-strided 16-byte pieces cost 32
//   L1 sectors per instruction and held the loop at ~6k clk per pair, profiles/r1_attention_ncu.md);
//   dq_finalize_kernel scales the accumulator into the fp16 dqkv columns afterwards.
// TMEM (512 columns): S^T|dP^T of group b at b*128 (64 + 64), overwritten in place by the fp16 P^T | dS^T operands;
// dV 256; dK 320; dQ 384; K_t 448 and V_t 480 as fp16 pairs (A operands of S^T / dP^T).
// Warps: 0 = TMA producer, 1 = tcgen05 issuer, 2..17 = compute: group (w-2)/4 % 2, column half (w-2)/8.
#include "common.cuh"
#include "tmap.cuh"
#include "umma.cuh"
#include <stdlib.h>

// Developer timeline: compile with -DB3_TRACE and set B2D_BWD3_TRACE_PTR=<device address of 1024 int64> to record clock64() stamps of CTA (0,0,0):
// [(group*32 + step)*8 + k] for the compute groups, [512 + j*6 + k] for the tcgen05 issuer (scripts/dev_bwd3_trace.py).
#ifdef B3_TRACE
#define B3_TR(cond, idx) do { if (cond) p.trace[(idx)] = clock64(); } while (0)
#else
#define B3_TR(cond, idx) do { } while (0)
#endif

namespace b2d {
using namespace umma;

struct AttnBwd3Params {
  int L, heads;
  int q_row0, q_rows;  // query rows [q_row0, q_row0 + q_rows) are visited (sequence-parallel shard): dQ of those rows is
                       // complete, dK / dV are this shard's partial sums over them
  int q_off, k_off, v_off, head_stride;
  float scale_log2, scale;
  const float* lse;    // (N, heads, L) log2 domain
  const float* delta;  // (N, heads, L)
  float* dq_acc;       // (N, L, heads, 64) fp32, zeroed before the launch
  __half* dqkv; long long lddqkv;
  long long* trace;    // developer timeline buffer (B3_TRACE builds), else null
};

struct AttnBwd3Maps {
  CUtensorMap qkv128;  // box (64, 128, 1) on qkv : resident K_t, V_t
  CUtensorMap qkv64;   // box (64, 64, 1)  on qkv : streamed Q_j
  CUtensorMap do64;    // box (64, 64, 1)  on dO  : streamed dO_j
  CUtensorMap dq;      // box (32, 128, 1) on the fp32 dQ accumulator (N, L, heads*64)
  CUtensorMap lse;     // box (64, 1) on lse   (L, N*heads) fp32, no swizzle
  CUtensorMap delta;   // box (64, 1) on delta (L, N*heads) fp32, no swizzle
};

constexpr int AB3_THREADS = 576;   // TMA warp + MMA warp + 16 compute warps
constexpr int AB3_ST = 4;
constexpr int B3_T128 = 128 * 128;   // [128 rows][64 ch] fp16
constexpr int B3_T64 = 64 * 128;     // [64 rows][64 ch] fp16

struct Bwd3Smem {
  static constexpr int K_OFF = 0;
  static constexpr int V_OFF = B3_T128;
  static constexpr int STAGE = 2 * B3_T64 + 1024;                 // {Q_j, dO_j, lse_j[64] | delta_j[64] (+ pad)}
  static constexpr int ST_OFF = 2 * B3_T128;                      // ring of stages
  static constexpr int DS_OFF = ST_OFF + AB3_ST * STAGE;          // dS^T of group b : [128 keys][64 queries], 16 KB apart
  static constexpr int DQ_OFF = DS_OFF + 2 * B3_T128;             // fp32 dQ staging [2 groups][128 q][32 ch], SW128
  static constexpr int BAR_OFF = DQ_OFF + 2 * B3_T128;
  static constexpr int NBARS = 1 + 2 * AB3_ST + 13;
  static constexpr int TOTAL = BAR_OFF + NBARS * 8 + 16 + 1024;
};

__device__ __forceinline__ void named_bar_sync3(int id, int count) {
  asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(count) : "memory");
}

__global__ void __launch_bounds__(AB3_THREADS, 1)
attn_bwd_fused_kernel(const __grid_constant__ AttnBwd3Maps maps, const AttnBwd3Params p) {
  using SM = Bwd3Smem;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = (uint64_t*)(smem + SM::BAR_OFF);
  uint64_t* kv_full = bars;
  uint64_t* full = bars + 1;
  uint64_t* empty = full + AB3_ST;
  uint64_t* sdp_full = empty + AB3_ST;   // [2] per group: S^T | dP^T of its step are in TMEM
  uint64_t* s_free = sdp_full + 2;       // [2] per group: S^T | dP^T loaded to registers
  uint64_t* p_ready = s_free + 2;        // [2] per group: P^T, dS^T written to smem
  uint64_t* pair_done = p_ready + 2;     // [2] per dQ buffer: every MMA up to and including dQ_m has completed
  uint64_t* dq_free = pair_done + 2;     // [2] per dQ buffer: both groups have read the dQ tile
  uint64_t* stg_free = dq_free + 2;      // [2] per group: its fp32 dQ staging buffer may be overwritten
  uint64_t* kv_tmem = stg_free + 2;      // K_t, V_t copied to tensor memory by the compute warps
  uint32_t* tmem_slot = (uint32_t*)(kv_tmem + 1);

  // S2R shares the MUFU issue path with ex2: an opaque copy keeps the compiler from re-reading %tid.x in the loop
  uint32_t tid;
  asm volatile("mov.u32 %0, %%tid.x;" : "=r"(tid));
  const int warp = tid >> 5, lane = tid & 31;
  const int kv0 = blockIdx.x * 128;
  const int head = blockIdx.y, n = blockIdx.z;
  const int n_pairs = (p.q_rows + 127) / 128;
  const int n_steps = 2 * n_pairs;       // an odd trailing 64-query tile is padded with an all-masked step
  const int qc = p.q_off + head * p.head_stride, kc = p.k_off + head * p.head_stride, vc = p.v_off + head * p.head_stride;

  if (tid == 0) {
    mbar_init(kv_full, 1);
    mbar_init(kv_tmem, 512);
    for (int s = 0; s < AB3_ST; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&sdp_full[b], 1); mbar_init(&s_free[b], 1); mbar_init(&p_ready[b], 256);
      mbar_init(&pair_done[b], 1); mbar_init(&dq_free[b], 512); mbar_init(&stg_free[b], 1);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (elect_one()) {
      mbar_expect_tx(kv_full, 2 * B3_T128);
      tma_load_3d(smem + SM::K_OFF, &maps.qkv128, kv_full, kc, kv0, n);
      tma_load_3d(smem + SM::V_OFF, &maps.qkv128, kv_full, vc, kv0, n);
      for (int j = 0; j < n_steps; ++j) {
        const int s = j % AB3_ST;
        mbar_wait(&empty[s], (uint32_t)(((j / AB3_ST) & 1) ^ 1));
        uint8_t* q_dst = smem + SM::ST_OFF + s * SM::STAGE;
        mbar_expect_tx(&full[s], 2 * B3_T64 + 512);
        tma_load_3d(q_dst, &maps.qkv64, &full[s], qc, p.q_row0 + j * 64, n);            // rows past L arrive as zeros
        tma_load_3d(q_dst + B3_T64, &maps.do64, &full[s], head * 64, p.q_row0 + j * 64, n);
        tma_load_2d(q_dst + 2 * B3_T64, &maps.lse, &full[s], p.q_row0 + j * 64, n * p.heads + head);
        tma_load_2d(q_dst + 2 * B3_T64 + 256, &maps.delta, &full[s], p.q_row0 + j * 64, n * p.heads + head);
      }
    }
  } else if (warp == 1) {
    const uint32_t idesc_s = make_idesc(128, 64, 0, 0, FMT_F16, FMT_F16);   // S^T, dP^T : both operands K-major
    const uint32_t idesc_g = make_idesc(128, 64, 0, 1, FMT_F16, FMT_F16);   // dV, dK    : B operand MN-major
    const uint32_t idesc_q = make_idesc(128, 64, 1, 1, FMT_F16, FMT_F16);   // dQ        : both operands MN-major
    const uint32_t k_addr = smem_u32(smem + SM::K_OFF);
    const uint32_t ds_addr = smem_u32(smem + SM::DS_OFF);
    const uint32_t st_addr = smem_u32(smem + SM::ST_OFF);
    auto issue_sdp = [&](int j) {
      const int b = j & 1;
      mbar_wait(&full[j % AB3_ST], (uint32_t)((j / AB3_ST) & 1));
      tc_fence_after();
      if (elect_one()) {
        const uint32_t q_addr = st_addr + (j % AB3_ST) * SM::STAGE, do_addr = q_addr + B3_T64;
        const uint32_t d_s = tmem_base + (uint32_t)b * 128u, d_dp = d_s + 64u;
        // A = K_t / V_t from tensor memory (8 columns per K = 16 channels), B = Q_j / dO_j K-major from smem
        // (the two accumulators alternate so that consecutive instructions are independent)
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          mma_f16_ts(d_s, tmem_base + 448u + (uint32_t)kk * 8u,
                     make_sdesc_sw128(q_addr + kk * 32, 16, 1024), idesc_s, (uint32_t)(kk != 0));
          mma_f16_ts(d_dp, tmem_base + 480u + (uint32_t)kk * 8u,
                     make_sdesc_sw128(do_addr + kk * 32, 16, 1024), idesc_s, (uint32_t)(kk != 0));
        }
        mma_commit(&sdp_full[b]);
      }
      __syncwarp();
    };
    mbar_wait(kv_tmem, 0);
    tc_fence_after();
    issue_sdp(0);
    issue_sdp(1);
    const bool tr = p.trace && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && lane == 0;
    for (int j = 0; j < n_steps; ++j) {
      const int b = j & 1, m = j >> 1;
      const uint32_t par = (uint32_t)(m & 1);
      B3_TR(tr && j < 64, 512 + j * 6 + 0);
      mbar_wait(&p_ready[b], par);
      B3_TR(tr && j < 64, 512 + j * 6 + 1);
      tc_fence_after();
      if (elect_one()) {
        const int s = j % AB3_ST;
        const uint32_t q_addr = st_addr + s * SM::STAGE, do_addr = q_addr + B3_T64;
        const uint32_t d_dv = tmem_base + 256u, d_dk = tmem_base + 320u;
        // A = P^T / dS^T from tensor memory: fp16 pairs packed over the group's own S^T / dP^T columns, 16 columns
        // per column half (kk = 0,1 -> half 0, kk = 2,3 -> half 1); B = dO_j / Q_j MN-major from shared memory
        const uint32_t a_p = tmem_base + (uint32_t)b * 128u, a_ds = a_p + 64u;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          mma_f16_ts(d_dv, a_p + (uint32_t)((kk >> 1) * 32 + (kk & 1) * 8),
                     make_sdesc_sw128(do_addr + kk * 2048, B3_T64, 1024), idesc_g, (uint32_t)((j | kk) != 0));
          mma_f16_ts(d_dk, a_ds + (uint32_t)((kk >> 1) * 32 + (kk & 1) * 8),
                     make_sdesc_sw128(q_addr + kk * 2048, B3_T64, 1024), idesc_g, (uint32_t)((j | kk) != 0));
        }
        mma_commit(&empty[s]);
      }
      __syncwarp();
      B3_TR(tr && j < 64, 512 + j * 6 + 2);
      if (b == 1) {
        const int db = m & 1;     // barrier slot only: the dQ tile itself is single-buffered
        if (m >= 1) { mbar_wait(&dq_free[(m - 1) & 1], (uint32_t)(((m - 1) >> 1) & 1)); tc_fence_after(); }
        if (elect_one()) {
          const uint32_t d_dq = tmem_base + 384u;
          // A = dS (M = 128 queries: two 64-query groups = the two dS^T buffers, K = keys = smem rows)
#pragma unroll
          for (int kk = 0; kk < 8; ++kk)
            mma_f16_ss(d_dq, make_sdesc_sw128(ds_addr + kk * 2048, B3_T128, 1024),
                       make_sdesc_sw128(k_addr + kk * 2048, B3_T128, 1024), idesc_q, (uint32_t)(kk != 0));
          mma_commit(&pair_done[db]);
        }
        __syncwarp();
      }
      // S^T | dP^T of the group's next step overwrite the columns P^T / dS^T were read from: issued behind the
      // dV / dK MMAs of this step (the tensor pipe executes in issue order); the group stages its dQ tile meanwhile
      B3_TR(tr && j < 64, 512 + j * 6 + 3);
      if (j + 2 < n_steps) issue_sdp(j + 2);
      B3_TR(tr && j < 64, 512 + j * 6 + 4);
    }
  } else {
    // 16 compute warps: group b (which steps) x column half h (which 32 of the step's 64 queries); the TMEM lane
    // quarter is fixed by warp % 4.  Four warps per scheduler keep the exp / pack chains of one step overlapped
    // with the TMEM loads, barrier waits and staging of the others (two warps per scheduler ran at ~5k clk/pair).
    const int cw = warp - 2;
    const int b = (cw >> 2) & 1, h = cw >> 3;
    const int qd = warp & 3;
    const int row = qd * 32 + lane;                 // TMEM lane: key (S^T, dV, dK) or query within the pair (dQ)
    const int gt = h * 128 + (cw & 3) * 32 + lane;  // 0..255 within the group
    const uint32_t lane_base = tmem_base + ((uint32_t)(qd * 32) << 16);
    const uint32_t s_addr = lane_base + (uint32_t)b * 128u + (uint32_t)h * 32u;
    const uint32_t ds_u32 = smem_u32(smem + SM::DS_OFF + b * B3_T128);
    const uint32_t st_u32 = smem_u32(smem + SM::ST_OFF);
    const float c = p.scale_log2;
    // this thread's 16 columns of the dQ tile of pair m (group b owns columns b*32.., half h 16 of them) -> smem
    // staging -> TMA reduce-add into the global fp32 accumulator.
    const uint32_t stg_u32 = smem_u32(smem + SM::DQ_OFF + b * B3_T128);
    auto stage_dq = [&](int m) {
      const int db = m & 1;
      uint32_t a[16];
      tmem_ld16(lane_base + 384u + (uint32_t)b * 32u + (uint32_t)h * 16u, a);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&dq_free[db]);
      // the previous bulk reduction (issued one whole step ago) has finished reading the staging buffer
      if (gt == 0) { tma_store_wait_read(); mbar_arrive(&stg_free[b]); }
      mbar_wait(&stg_free[b], (uint32_t)(m & 1));
#pragma unroll
      for (int g = 0; g < 4; ++g)
        sts128(stg_u32 + (uint32_t)row * 128u + (uint32_t)(((h * 4 + g) ^ (row & 7)) << 4),
               make_uint4(a[4 * g], a[4 * g + 1], a[4 * g + 2], a[4 * g + 3]));
    };
    // after a proxy fence by every thread of the group
    auto issue_dq = [&](int m) {
      named_bar_sync3(1 + b, 256);
      if (gt == 0) {
        tma_reduce_add_3d(&maps.dq, stg_u32, head * 64 + b * 32, p.q_row0 + m * 128, n);   // rows past L are clipped
        tma_store_commit();
      }
    };
    {
      // K_t and V_t also live in tensor memory (columns 448..479 / 480..511, fp16 pairs) as the A operands of S^T and
      // dP^T: an N = 64 MMA with both operands in shared memory wants 192 B/clk of a 128 B/clk port (measured ~72 clk
      // per instruction against 47 with A in tensor memory).  Warp (b, h) copies 32 channels of one of them.
      mbar_wait(kv_full, 0);
      const uint32_t src = smem_u32(smem + (b == 0 ? SM::K_OFF : SM::V_OFF));
      uint32_t kv[16];
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        const uint4 u = lds128(src + sw128_offset((uint32_t)row, (uint32_t)(8 * (h * 4 + g))));
        kv[4 * g] = u.x; kv[4 * g + 1] = u.y; kv[4 * g + 2] = u.z; kv[4 * g + 3] = u.w;
      }
      tmem_st16(lane_base + 448u + (uint32_t)b * 32u + (uint32_t)h * 16u, kv);
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(kv_tmem);
    }
    const bool tr = p.trace && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && lane == 0 && qd == 0 && h == 0;
    for (int m = 0; m < n_pairs; ++m) {
      B3_TR(tr && m < 32, (b * 32 + m) * 8 + 0);
      // lse / delta of the step's 64 queries arrived with its Q / dO stage (rows past L read as 0: with Q = dO = 0
      // that gives P = 1, dS = 0, and the zero dO / Q rows keep dV / dK untouched)
      const int j = 2 * m + b, stg = j % AB3_ST;
      const uint32_t ls_u32 = st_u32 + (uint32_t)stg * SM::STAGE + 2 * B3_T64 + (uint32_t)h * 128u, dl_u32 = ls_u32 + 256u;
      mbar_wait(&full[stg], (uint32_t)((j / AB3_ST) & 1));
      mbar_wait(&sdp_full[b], (uint32_t)(m & 1));
      B3_TR(tr && m < 32, (b * 32 + m) * 8 + 1);
      tc_fence_after();
      uint32_t sr[32], dr[32];
      tmem_ld32(s_addr, sr);
      tmem_ld32(s_addr + 64, dr);
      tmem_ld_wait();
      B3_TR(tr && m < 32, (b * 32 + m) * 8 + 2);
#pragma unroll
      for (int i4 = 0; i4 < 32; i4 += 4) {
        const float4 l4 = lds128f(ls_u32 + i4 * 4);   // warp-uniform address: broadcast LDS.128
        const float4 d4 = lds128f(dl_u32 + i4 * 4);
        const float lv[4] = {l4.x, l4.y, l4.z, l4.w}, dv[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int i = i4 + u;
          const float pe = fast_exp2(fmaf(__uint_as_float(sr[i]), c, -lv[u]));
          sr[i] = __float_as_uint(pe);
          dr[i] = __float_as_uint(pe * (__uint_as_float(dr[i]) - dv[u]));
        }
      }
      // dV/dK of this group's previous step and dQ of the previous pair have read P^T / dS^T
      B3_TR(tr && m < 32, (b * 32 + m) * 8 + 3);
      if (m > 0) { mbar_wait(&pair_done[(m - 1) & 1], (uint32_t)(((m - 1) >> 1) & 1)); tc_fence_after(); }
      B3_TR(tr && m < 32, (b * 32 + m) * 8 + 4);
      // P^T -> tensor memory over this thread's own S^T columns, dS^T -> tensor memory over its dP^T columns (A
      // operands of dV / dK) and -> shared memory (MN-major A operand of dQ)
      uint32_t pp[16], dd[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        pp[i] = pack_f16x2_nosat(__uint_as_float(sr[2 * i]), __uint_as_float(sr[2 * i + 1]));
        dd[i] = pack_f16x2(__uint_as_float(dr[2 * i]), __uint_as_float(dr[2 * i + 1]));
      }
      tmem_st16(s_addr, pp);
      tmem_st16(s_addr + 64, dd);
#pragma unroll
      for (int g = 0; g < 4; ++g)
        sts128(ds_u32 + sw128_offset((uint32_t)row, (uint32_t)(8 * (h * 4 + g))),
               make_uint4(dd[4 * g], dd[4 * g + 1], dd[4 * g + 2], dd[4 * g + 3]));
      tmem_st_wait();
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(&p_ready[b]);
      B3_TR(tr && m < 32, (b * 32 + m) * 8 + 5);
      // dQ tile of the previous pair -> staging -> bulk reduction, off the p_ready path (staging it under the
      // fence above saved a fence but delayed p_ready: 4.33 ms vs 4.0 ms at the bench shape)
      if (m > 0) { stage_dq(m - 1); fence_proxy_async_smem(); B3_TR(tr && m < 32, (b * 32 + m) * 8 + 6); issue_dq(m - 1); }
      B3_TR(tr && m < 32, (b * 32 + m) * 8 + 7);
    }
    mbar_wait(&pair_done[(n_pairs - 1) & 1], (uint32_t)(((n_pairs - 1) >> 1) & 1));
    tc_fence_after();
    stage_dq(n_pairs - 1);
    fence_proxy_async_smem();
    issue_dq(n_pairs - 1);
    // group 0 writes dV, group 1 writes dK * scale; half h the columns h*32..
    const int ki = kv0 + row;
    const bool valid = ki < p.L;
    __half* dst = p.dqkv + ((long long)n * p.L + ki) * p.lddqkv + (b == 0 ? vc : kc) + h * 32;
    const float sc = b == 0 ? 1.f : p.scale;
    {
      uint32_t a[32];
      tmem_ld32(lane_base + 256u + (uint32_t)b * 64u + (uint32_t)h * 32u, a);   // warp-collective: outside the row guard
      tmem_ld_wait();
      if (valid) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 o;
          o.x = pack_f16x2(__uint_as_float(a[8 * g + 0]) * sc, __uint_as_float(a[8 * g + 1]) * sc);
          o.y = pack_f16x2(__uint_as_float(a[8 * g + 2]) * sc, __uint_as_float(a[8 * g + 3]) * sc);
          o.z = pack_f16x2(__uint_as_float(a[8 * g + 4]) * sc, __uint_as_float(a[8 * g + 5]) * sc);
          o.w = pack_f16x2(__uint_as_float(a[8 * g + 6]) * sc, __uint_as_float(a[8 * g + 7]) * sc);
          *reinterpret_cast<uint4*>(dst + 8 * g) = o;
        }
      }
    }
    if (gt == 0) tma_store_wait_all();   // bulk reductions read this CTA's smem
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

// dqkv[q columns] = fp16(scale * dq_acc)
__global__ void __launch_bounds__(256)
dq_finalize_kernel(const float* __restrict__ acc, __half* __restrict__ dqkv, long long lddqkv, long long rows,
                   int heads, int q_off, int head_stride, float scale) {
  const long long total = rows * heads * 8;   // 8-column groups
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(i & 7);
    const long long rh = i >> 3;
    const int head = (int)(rh % heads);
    const long long r = rh / heads;
    const float4 a = __ldcs(reinterpret_cast<const float4*>(acc + i * 8));
    const float4 b = __ldcs(reinterpret_cast<const float4*>(acc + i * 8 + 4));
    uint4 o;
    o.x = pack_f16x2(a.x * scale, a.y * scale);
    o.y = pack_f16x2(a.z * scale, a.w * scale);
    o.z = pack_f16x2(b.x * scale, b.y * scale);
    o.w = pack_f16x2(b.z * scale, b.w * scale);
    *reinterpret_cast<uint4*>(dqkv + r * lddqkv + q_off + head * head_stride + g * 8) = o;
  }
}

// called by b2d_attention_bwd (attention_bwd.cu) after the delta kernel, for ch == 64 and L > 128.
// dq_acc: N*L*heads*64 floats of workspace.
int attention_bwd3_dispatch(const void* qkv, int64_t ldqkv, const void* d_o, int64_t lddo, int N, int L, int heads,
                            int new_order, const float* lse, const float* delta, float* dq_acc, void* dqkv,
                            int64_t lddqkv, int q_row0, int q_rows, void* stream) {
  B2D_SET_SMEM_ONCE((attn_bwd_fused_kernel), Bwd3Smem::TOTAL);
  const int ch = 64, C = heads * ch;
  AttnBwd3Params p;
  p.L = L; p.heads = heads;
  p.q_row0 = q_row0; p.q_rows = q_rows;
  if (new_order) { p.q_off = 0; p.k_off = C; p.v_off = 2 * C; p.head_stride = ch; }
  else           { p.q_off = 0; p.k_off = ch; p.v_off = 2 * ch; p.head_stride = 3 * ch; }
  p.scale = 1.0f / sqrtf((float)ch);
  p.scale_log2 = 1.4426950408889634f * p.scale;
  p.lse = lse; p.delta = delta; p.dq_acc = dq_acc;
  p.dqkv = (__half*)dqkv; p.lddqkv = lddqkv;
  AttnBwd3Maps m;
  const uint64_t dims[3] = {(uint64_t)(3 * C), (uint64_t)L, (uint64_t)N};
  const uint64_t str[2] = {(uint64_t)ldqkv * 2, (uint64_t)ldqkv * 2 * (uint64_t)L};
  const uint64_t ddims[3] = {(uint64_t)C, (uint64_t)L, (uint64_t)N};
  const uint64_t dstr[2] = {(uint64_t)lddo * 2, (uint64_t)lddo * 2 * (uint64_t)L};
  const uint32_t box128[3] = {64u, 128u, 1u}, box64[3] = {64u, 64u, 1u};
  int rc;
  if ((rc = make_tmap_16b(&m.qkv128, qkv, 3, dims, str, box128))) return rc;
  if ((rc = make_tmap_16b(&m.qkv64, qkv, 3, dims, str, box64))) return rc;
  if ((rc = make_tmap_16b(&m.do64, d_o, 3, ddims, dstr, box64))) return rc;
  const uint64_t qdims[3] = {(uint64_t)C, (uint64_t)L, (uint64_t)N};
  const uint64_t qstr[2] = {(uint64_t)C * 4, (uint64_t)C * 4 * (uint64_t)L};
  const uint32_t qbox[3] = {32u, 128u, 1u};
  if ((rc = make_tmap_f32(&m.dq, dq_acc, 3, qdims, qstr, qbox))) return rc;
  const uint64_t vdims[2] = {(uint64_t)L, (uint64_t)N * heads};
  const uint64_t vstr[1] = {(uint64_t)L * 4};
  const uint32_t vbox[2] = {64u, 1u};
  if ((rc = make_tmap_f32_linear(&m.lse, lse, 2, vdims, vstr, vbox))) return rc;
  if ((rc = make_tmap_f32_linear(&m.delta, delta, 2, vdims, vstr, vbox))) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  p.trace = nullptr;
#ifdef B3_TRACE   // developer builds: the CALLER owns the 1024 x int64 device buffer (scripts/dev_bwd3_trace.py); the
  {               // library allocates nothing and never synchronises, also when tracing
    const char* e = getenv("B2D_BWD3_TRACE_PTR");
    if (e) p.trace = (long long*)strtoull(e, nullptr, 0);
  }
#endif
  const long long rows = (long long)N * L;
  B2D_CHECK_CUDA(cudaMemsetAsync(dq_acc, 0, (size_t)rows * heads * 64 * sizeof(float), st));
  dim3 grid((unsigned)((L + 127) / 128), (unsigned)heads, (unsigned)N);
  attn_bwd_fused_kernel<<<grid, AB3_THREADS, Bwd3Smem::TOTAL, st>>>(m, p);
  B2D_CHECK_LAUNCH();
  const long long groups = rows * heads * 8;
  const unsigned fin_blocks = (unsigned)((groups + 255) / 256 < 148 * 16 ? (groups + 255) / 256 : 148 * 16);
  dq_finalize_kernel<<<fin_blocks, 256, 0, st>>>(dq_acc, (__half*)dqkv, lddqkv, rows, heads, p.q_off, p.head_stride, p.scale);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

}  // namespace b2d
