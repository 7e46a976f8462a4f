// attention_bwd2.cu -- flash-style self-attention backward for head dim 64, two 128-row tiles per CTA with two
// compute warpgroups ping-ponging against the tensor pipe (same scheme as attention_fwd2.cu).  Replaces the
// single-tile kernels of attention_bwd.cu for the full-resolution layers (ch = 64, L > 128), which carry
// >85 % of the attention-backward FLOPs (QKVAttentionLegacy unet.py:349-366 under autograd + the
// checkpoint recompute of utils.py:179-207).
//
//   kernel A (dkv2): CTA = 256 keys (tiles t = 0,1), streams 64-query tiles j:
//        S^T_t = K_t Q_j^T, dP^T_t = V_t dO_j^T            (M = 128 keys, N = 64 queries, K = 64 ch)
//        P^T = exp2(c S^T - lse_q), dS^T = P^T (dP^T - delta_q)   -> fp16 smem
//        dV_t += P^T_t dO_j, dK_t += dS^T_t Q_j             (M = 128 keys, N = 64 ch, K = 64 queries)
//   kernel B (dq2):  CTA = 256 queries (tiles t = 0,1), streams 64-key tiles j:
//        S_t = Q_t K_j^T, dP_t = dO_t V_j^T;  dS = P (dP - delta) -> fp16 smem;  dQ_t += dS_t K_j
// TMEM (512 columns): per tile S | dP (64 + 64) at t*128; accumulators from column 256.
// Warps: 0 = TMA producer, 1 = tcgen05 issuer, 2..5 = compute group 0, 6..9 = compute group 1.
#include "common.cuh"
#include "tmap.cuh"
#include "umma.cuh"

namespace b2d {
using namespace umma;

struct AttnBwd2Params {
  int L, heads;
  int q_off, k_off, v_off, head_stride;
  float scale_log2, scale;
  const float* lse;    // (N, heads, L) log2 domain
  const float* delta;  // (N, heads, L)
  __half* dqkv; long long lddqkv;
};

struct AttnBwd2Maps {
  CUtensorMap qkv128;  // box (64, 128, 1) on qkv : resident tiles
  CUtensorMap qkv64;   // box (64, 64, 1)  on qkv : streamed tiles
  CUtensorMap do128;   // box (64, 128, 1) on dO
  CUtensorMap do64;    // box (64, 64, 1)  on dO
};

constexpr int AB2_THREADS = 320;
constexpr int AB2_ST = 4;                 // streamed-tile ring depth
constexpr int T128 = 128 * 128;           // [128 rows][64 ch] fp16
constexpr int T64 = 64 * 128;             // [64 rows][64 ch] fp16

__device__ __forceinline__ void named_bar_sync2(int id, int count) {
  asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(count) : "memory");
}

// ------------------------------------------------------------------------------------------------ kernel A
struct Dkv2Smem {
  static constexpr int K_OFF = 0;                       // 2 x T128
  static constexpr int V_OFF = 2 * T128;                // 2 x T128
  static constexpr int ST_OFF = 4 * T128;               // ring of {Q_j, dO_j}: AB2_ST x 2 x T64
  static constexpr int PT_OFF = ST_OFF + AB2_ST * 2 * T64;   // P^T_t  : 2 x T128 ([128 keys][64 queries])
  static constexpr int DS_OFF = PT_OFF + 2 * T128;           // dS^T_t : 2 x T128
  static constexpr int VEC_OFF = DS_OFF + 2 * T128;          // lse/delta staging: [2 groups][3 bufs][128] fp32
  static constexpr int BAR_OFF = VEC_OFF + 2 * 3 * 128 * 4;
  static constexpr int NBARS = 1 + 2 * AB2_ST + 8;
  static constexpr int TOTAL = BAR_OFF + NBARS * 8 + 16 + 1024;
};

__global__ void __launch_bounds__(AB2_THREADS, 1)
attn_bwd_dkv2_kernel(const __grid_constant__ AttnBwd2Maps maps, const AttnBwd2Params p) {
  using SM = Dkv2Smem;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = (uint64_t*)(smem + SM::BAR_OFF);
  uint64_t* kv_full = bars;
  uint64_t* full = bars + 1;
  uint64_t* empty = full + AB2_ST;
  uint64_t* sdp_full = empty + AB2_ST;   // [2]
  uint64_t* s_free = sdp_full + 2;       // [2]
  uint64_t* p_ready = s_free + 2;        // [2]
  uint64_t* acc_done = p_ready + 2;      // [2]
  uint32_t* tmem_slot = (uint32_t*)(acc_done + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kv0 = blockIdx.x * 256;
  const int head = blockIdx.y, n = blockIdx.z;
  const int n_q = (p.L + 63) / 64;
  const int qc = p.q_off + head * p.head_stride, kc = p.k_off + head * p.head_stride, vc = p.v_off + head * p.head_stride;

  if (threadIdx.x == 0) {
    mbar_init(kv_full, 1);
    for (int s = 0; s < AB2_ST; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int t = 0; t < 2; ++t) {
      mbar_init(&sdp_full[t], 1); mbar_init(&s_free[t], 128); mbar_init(&p_ready[t], 128); mbar_init(&acc_done[t], 1);
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
      mbar_expect_tx(kv_full, 4 * T128);
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        tma_load_3d(smem + SM::K_OFF + t * T128, &maps.qkv128, kv_full, kc, kv0 + t * 128, n);
        tma_load_3d(smem + SM::V_OFF + t * T128, &maps.qkv128, kv_full, vc, kv0 + t * 128, n);
      }
      for (int j = 0; j < n_q; ++j) {
        const int s = j % AB2_ST;
        mbar_wait(&empty[s], (uint32_t)(((j / AB2_ST) & 1) ^ 1));
        uint8_t* q_dst = smem + SM::ST_OFF + s * 2 * T64;
        mbar_expect_tx(&full[s], 2 * T64);
        tma_load_3d(q_dst, &maps.qkv64, &full[s], qc, j * 64, n);
        tma_load_3d(q_dst + T64, &maps.do64, &full[s], head * 64, j * 64, n);
      }
    }
  } else if (warp == 1) {
    const uint32_t idesc_s = make_idesc(128, 64, 0, 0, FMT_F16, FMT_F16);   // S^T, dP^T : both operands K-major
    const uint32_t idesc_g = make_idesc(128, 64, 0, 1, FMT_F16, FMT_F16);   // dV, dK    : B operand MN-major
    const uint32_t k_addr = smem_u32(smem + SM::K_OFF), v_addr = smem_u32(smem + SM::V_OFF);
    const uint32_t pt_addr = smem_u32(smem + SM::PT_OFF), ds_addr = smem_u32(smem + SM::DS_OFF);
    const uint32_t st_addr = smem_u32(smem + SM::ST_OFF);
    auto issue_sdp = [&](int j, int t) {
      if (elect_one()) {
        const uint32_t q_addr = st_addr + (j % AB2_ST) * 2 * T64, do_addr = q_addr + T64;
        const uint32_t d_s = tmem_base + (uint32_t)t * 128u, d_dp = d_s + 64u;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          mma_f16_ss(d_s, make_sdesc_sw128(k_addr + t * T128 + kk * 32, 16, 1024),
                     make_sdesc_sw128(q_addr + kk * 32, 16, 1024), idesc_s, (uint32_t)(kk != 0));
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          mma_f16_ss(d_dp, make_sdesc_sw128(v_addr + t * T128 + kk * 32, 16, 1024),
                     make_sdesc_sw128(do_addr + kk * 32, 16, 1024), idesc_s, (uint32_t)(kk != 0));
        mma_commit(&sdp_full[t]);
      }
      __syncwarp();
    };
    auto issue_acc = [&](int j, int t) {
      if (elect_one()) {
        const int s = j % AB2_ST;
        const uint32_t q_addr = st_addr + s * 2 * T64, do_addr = q_addr + T64;
        const uint32_t d_dv = tmem_base + 256u + (uint32_t)t * 128u, d_dk = d_dv + 64u;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          mma_f16_ss(d_dv, make_sdesc_sw128(pt_addr + t * T128 + kk * 32, 16, 1024),
                     make_sdesc_sw128(do_addr + kk * 2048, T64, 1024), idesc_g, (uint32_t)((j | kk) != 0));
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          mma_f16_ss(d_dk, make_sdesc_sw128(ds_addr + t * T128 + kk * 32, 16, 1024),
                     make_sdesc_sw128(q_addr + kk * 2048, T64, 1024), idesc_g, (uint32_t)((j | kk) != 0));
        mma_commit(&acc_done[t]);
        if (t == 1) mma_commit(&empty[s]);
      }
      __syncwarp();
    };
    mbar_wait(kv_full, 0);
    mbar_wait(&full[0], 0);
    tc_fence_after();
    issue_sdp(0, 0);
    issue_sdp(0, 1);
    for (int j = 0; j < n_q; ++j) {
      const uint32_t par = (uint32_t)(j & 1);
      const bool more = j + 1 < n_q;
      if (more) mbar_wait(&full[(j + 1) % AB2_ST], (uint32_t)(((j + 1) / AB2_ST) & 1));
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        if (more) { mbar_wait(&s_free[t], par); tc_fence_after(); issue_sdp(j + 1, t); }
        mbar_wait(&p_ready[t], par);
        tc_fence_after();
        issue_acc(j, t);
      }
    }
  } else {
    const int t = (warp - 2) >> 2;
    const int qd = warp & 3;
    const int row = qd * 32 + lane;                 // key within the tile
    const int gt = threadIdx.x - 64 - t * 128;      // 0..127 within the group
    const uint32_t lane_base = tmem_base + ((uint32_t)(qd * 32) << 16);
    const uint32_t s_addr = lane_base + (uint32_t)t * 128u;
    const uint32_t acc_addr = lane_base + 256u + (uint32_t)t * 128u;
    const uint32_t pt_u32 = smem_u32(smem + SM::PT_OFF + t * T128);
    const uint32_t ds_u32 = smem_u32(smem + SM::DS_OFF + t * T128);
    const uint32_t vec_u32 = smem_u32(smem + SM::VEC_OFF) + (uint32_t)t * 384u * 4u;   // [3 bufs][64 lse | 64 delta]
    const float c = p.scale_log2;
    const float* lse_g = p.lse + ((long long)n * p.heads + head) * p.L;
    const float* del_g = p.delta + ((long long)n * p.heads + head) * p.L;
    // lse / delta of query tile j -> smem, software pipelined: the global load for tile j+2 is issued a full
    // step before its value is stored (a load -> store in the same step stalled every step on the HBM latency)
    auto ld_vec = [&](int j) -> float {
      const int qi = j * 64 + (gt & 63);
      if (gt < 64) return qi < p.L ? __ldg(lse_g + qi) : INFINITY;
      return qi < p.L ? __ldg(del_g + qi) : 0.f;
    };
    auto st_vec = [&](int j, float v) { sts32f(vec_u32 + (uint32_t)((j % 3) * 128 + gt) * 4u, v); };
    st_vec(0, ld_vec(0));
    float pend = n_q > 1 ? ld_vec(1) : 0.f;
    for (int j = 0; j < n_q; ++j) {
      // three buffers: a thread past this barrier writes buffer (j+2)%3 at the top of step j+1 while a slower
      // thread may still read buffer j%3; buffer (j+1)%3 was written before this barrier
      if (j + 1 < n_q) st_vec(j + 1, pend);
      if (j + 2 < n_q) pend = ld_vec(j + 2);
      named_bar_sync2(1 + t, 128);
      const uint32_t ls_u32 = vec_u32 + (uint32_t)(j % 3) * 512u, dl_u32 = ls_u32 + 256u;
      mbar_wait(&sdp_full[t], (uint32_t)(j & 1));
      tc_fence_after();
      uint32_t sr[64], dr[64];
      tmem_ld32(s_addr, sr); tmem_ld32(s_addr + 32, sr + 32);
      tmem_ld32(s_addr + 64, dr); tmem_ld32(s_addr + 96, dr + 32);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&s_free[t]);
#pragma unroll
      for (int i4 = 0; i4 < 64; i4 += 4) {
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
      if (j > 0) mbar_wait(&acc_done[t], (uint32_t)((j - 1) & 1));   // previous dV/dK MMAs have read P^T / dS^T
#pragma unroll
      for (int g = 0; g < 8; ++g) {
        uint4 u, w;
        u.x = pack_f16x2_nosat(__uint_as_float(sr[8 * g + 0]), __uint_as_float(sr[8 * g + 1]));
        u.y = pack_f16x2_nosat(__uint_as_float(sr[8 * g + 2]), __uint_as_float(sr[8 * g + 3]));
        u.z = pack_f16x2_nosat(__uint_as_float(sr[8 * g + 4]), __uint_as_float(sr[8 * g + 5]));
        u.w = pack_f16x2_nosat(__uint_as_float(sr[8 * g + 6]), __uint_as_float(sr[8 * g + 7]));
        w.x = pack_f16x2(__uint_as_float(dr[8 * g + 0]), __uint_as_float(dr[8 * g + 1]));
        w.y = pack_f16x2(__uint_as_float(dr[8 * g + 2]), __uint_as_float(dr[8 * g + 3]));
        w.z = pack_f16x2(__uint_as_float(dr[8 * g + 4]), __uint_as_float(dr[8 * g + 5]));
        w.w = pack_f16x2(__uint_as_float(dr[8 * g + 6]), __uint_as_float(dr[8 * g + 7]));
        const uint32_t off = sw128_offset((uint32_t)row, (uint32_t)(8 * g));
        sts128(pt_u32 + off, u);
        sts128(ds_u32 + off, w);
      }
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(&p_ready[t]);
    }
    mbar_wait(&acc_done[t], (uint32_t)((n_q - 1) & 1));
    tc_fence_after();
    const int ki = kv0 + t * 128 + row;
    const bool valid = ki < p.L;
    __half* base = p.dqkv + ((long long)n * p.L + ki) * p.lddqkv;
    const float sc = p.scale;
#pragma unroll
    for (int c0 = 0; c0 < 64; c0 += 32) {
      uint32_t a[32], b[32];
      tmem_ld32(acc_addr + c0, a);
      tmem_ld32(acc_addr + 64 + c0, b);
      tmem_ld_wait();
      if (valid) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 o, k;
          o.x = pack_f16x2(__uint_as_float(a[8 * g + 0]), __uint_as_float(a[8 * g + 1]));
          o.y = pack_f16x2(__uint_as_float(a[8 * g + 2]), __uint_as_float(a[8 * g + 3]));
          o.z = pack_f16x2(__uint_as_float(a[8 * g + 4]), __uint_as_float(a[8 * g + 5]));
          o.w = pack_f16x2(__uint_as_float(a[8 * g + 6]), __uint_as_float(a[8 * g + 7]));
          k.x = pack_f16x2(__uint_as_float(b[8 * g + 0]) * sc, __uint_as_float(b[8 * g + 1]) * sc);
          k.y = pack_f16x2(__uint_as_float(b[8 * g + 2]) * sc, __uint_as_float(b[8 * g + 3]) * sc);
          k.z = pack_f16x2(__uint_as_float(b[8 * g + 4]) * sc, __uint_as_float(b[8 * g + 5]) * sc);
          k.w = pack_f16x2(__uint_as_float(b[8 * g + 6]) * sc, __uint_as_float(b[8 * g + 7]) * sc);
          *reinterpret_cast<uint4*>(base + vc + c0 + 8 * g) = o;
          *reinterpret_cast<uint4*>(base + kc + c0 + 8 * g) = k;
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

// ------------------------------------------------------------------------------------------------ kernel B
struct Dq2Smem {
  static constexpr int Q_OFF = 0;                        // 2 x T128
  static constexpr int DO_OFF = 2 * T128;                // 2 x T128
  static constexpr int ST_OFF = 4 * T128;                // ring of {K_j, V_j}: AB2_ST x 2 x T64
  static constexpr int DS_OFF = ST_OFF + AB2_ST * 2 * T64;   // dS_t : 2 x T128 ([128 queries][64 keys])
  static constexpr int BAR_OFF = DS_OFF + 2 * T128;
  static constexpr int NBARS = 1 + 2 * AB2_ST + 8;
  static constexpr int TOTAL = BAR_OFF + NBARS * 8 + 16 + 1024;
};

__global__ void __launch_bounds__(AB2_THREADS, 1)
attn_bwd_dq2_kernel(const __grid_constant__ AttnBwd2Maps maps, const AttnBwd2Params p) {
  using SM = Dq2Smem;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = (uint64_t*)(smem + SM::BAR_OFF);
  uint64_t* q_full = bars;
  uint64_t* full = bars + 1;
  uint64_t* empty = full + AB2_ST;
  uint64_t* sdp_full = empty + AB2_ST;   // [2]
  uint64_t* s_free = sdp_full + 2;       // [2]
  uint64_t* ds_ready = s_free + 2;       // [2]
  uint64_t* acc_done = ds_ready + 2;     // [2]
  uint32_t* tmem_slot = (uint32_t*)(acc_done + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * 256;
  const int head = blockIdx.y, n = blockIdx.z;
  const int n_kv = (p.L + 63) / 64;
  const int qc = p.q_off + head * p.head_stride, kc = p.k_off + head * p.head_stride, vc = p.v_off + head * p.head_stride;

  if (threadIdx.x == 0) {
    mbar_init(q_full, 1);
    for (int s = 0; s < AB2_ST; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int t = 0; t < 2; ++t) {
      mbar_init(&sdp_full[t], 1); mbar_init(&s_free[t], 128); mbar_init(&ds_ready[t], 128); mbar_init(&acc_done[t], 1);
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
      mbar_expect_tx(q_full, 4 * T128);
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        tma_load_3d(smem + SM::Q_OFF + t * T128, &maps.qkv128, q_full, qc, q0 + t * 128, n);
        tma_load_3d(smem + SM::DO_OFF + t * T128, &maps.do128, q_full, head * 64, q0 + t * 128, n);
      }
      for (int j = 0; j < n_kv; ++j) {
        const int s = j % AB2_ST;
        mbar_wait(&empty[s], (uint32_t)(((j / AB2_ST) & 1) ^ 1));
        uint8_t* k_dst = smem + SM::ST_OFF + s * 2 * T64;
        mbar_expect_tx(&full[s], 2 * T64);
        tma_load_3d(k_dst, &maps.qkv64, &full[s], kc, j * 64, n);
        tma_load_3d(k_dst + T64, &maps.qkv64, &full[s], vc, j * 64, n);
      }
    }
  } else if (warp == 1) {
    const uint32_t idesc_s = make_idesc(128, 64, 0, 0, FMT_F16, FMT_F16);   // S, dP : both operands K-major
    const uint32_t idesc_g = make_idesc(128, 64, 0, 1, FMT_F16, FMT_F16);   // dQ    : K_j MN-major
    const uint32_t q_addr = smem_u32(smem + SM::Q_OFF), do_addr = smem_u32(smem + SM::DO_OFF);
    const uint32_t ds_addr = smem_u32(smem + SM::DS_OFF), st_addr = smem_u32(smem + SM::ST_OFF);
    auto issue_sdp = [&](int j, int t) {
      if (elect_one()) {
        const uint32_t k_addr = st_addr + (j % AB2_ST) * 2 * T64, v_addr = k_addr + T64;
        const uint32_t d_s = tmem_base + (uint32_t)t * 128u, d_dp = d_s + 64u;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          mma_f16_ss(d_s, make_sdesc_sw128(q_addr + t * T128 + kk * 32, 16, 1024),
                     make_sdesc_sw128(k_addr + kk * 32, 16, 1024), idesc_s, (uint32_t)(kk != 0));
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          mma_f16_ss(d_dp, make_sdesc_sw128(do_addr + t * T128 + kk * 32, 16, 1024),
                     make_sdesc_sw128(v_addr + kk * 32, 16, 1024), idesc_s, (uint32_t)(kk != 0));
        mma_commit(&sdp_full[t]);
      }
      __syncwarp();
    };
    auto issue_acc = [&](int j, int t) {
      if (elect_one()) {
        const int s = j % AB2_ST;
        const uint32_t k_addr = st_addr + s * 2 * T64;
        const uint32_t d_dq = tmem_base + 256u + (uint32_t)t * 64u;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          mma_f16_ss(d_dq, make_sdesc_sw128(ds_addr + t * T128 + kk * 32, 16, 1024),
                     make_sdesc_sw128(k_addr + kk * 2048, T64, 1024), idesc_g, (uint32_t)((j | kk) != 0));
        mma_commit(&acc_done[t]);
        if (t == 1) mma_commit(&empty[s]);
      }
      __syncwarp();
    };
    mbar_wait(q_full, 0);
    mbar_wait(&full[0], 0);
    tc_fence_after();
    issue_sdp(0, 0);
    issue_sdp(0, 1);
    for (int j = 0; j < n_kv; ++j) {
      const uint32_t par = (uint32_t)(j & 1);
      const bool more = j + 1 < n_kv;
      if (more) mbar_wait(&full[(j + 1) % AB2_ST], (uint32_t)(((j + 1) / AB2_ST) & 1));
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        if (more) { mbar_wait(&s_free[t], par); tc_fence_after(); issue_sdp(j + 1, t); }
        mbar_wait(&ds_ready[t], par);
        tc_fence_after();
        issue_acc(j, t);
      }
    }
  } else {
    const int t = (warp - 2) >> 2;
    const int qd = warp & 3;
    const int row = qd * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(qd * 32) << 16);
    const uint32_t s_addr = lane_base + (uint32_t)t * 128u;
    const uint32_t dq_addr = lane_base + 256u + (uint32_t)t * 64u;
    const uint32_t ds_u32 = smem_u32(smem + SM::DS_OFF + t * T128);
    const float c = p.scale_log2;
    const int qi = q0 + t * 128 + row;
    const bool valid = qi < p.L;
    const long long sidx = ((long long)n * p.heads + head) * p.L + qi;
    const float lse_i = valid ? __ldg(p.lse + sidx) : INFINITY;
    const float del_i = valid ? __ldg(p.delta + sidx) : 0.f;
    for (int j = 0; j < n_kv; ++j) {
      mbar_wait(&sdp_full[t], (uint32_t)(j & 1));
      tc_fence_after();
      uint32_t sr[64], dr[64];
      tmem_ld32(s_addr, sr); tmem_ld32(s_addr + 32, sr + 32);
      tmem_ld32(s_addr + 64, dr); tmem_ld32(s_addr + 96, dr + 32);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&s_free[t]);
#pragma unroll
      for (int i = 0; i < 64; ++i) {
        const float pe = fast_exp2(fmaf(__uint_as_float(sr[i]), c, -lse_i));
        dr[i] = __float_as_uint(pe * (__uint_as_float(dr[i]) - del_i));
      }
      if (j > 0) mbar_wait(&acc_done[t], (uint32_t)((j - 1) & 1));   // previous dQ MMA has read dS_t
#pragma unroll
      for (int g = 0; g < 8; ++g) {
        uint4 w;
        w.x = pack_f16x2(__uint_as_float(dr[8 * g + 0]), __uint_as_float(dr[8 * g + 1]));
        w.y = pack_f16x2(__uint_as_float(dr[8 * g + 2]), __uint_as_float(dr[8 * g + 3]));
        w.z = pack_f16x2(__uint_as_float(dr[8 * g + 4]), __uint_as_float(dr[8 * g + 5]));
        w.w = pack_f16x2(__uint_as_float(dr[8 * g + 6]), __uint_as_float(dr[8 * g + 7]));
        sts128(ds_u32 + sw128_offset((uint32_t)row, (uint32_t)(8 * g)), w);
      }
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(&ds_ready[t]);
    }
    mbar_wait(&acc_done[t], (uint32_t)((n_kv - 1) & 1));
    tc_fence_after();
    __half* dst = p.dqkv + ((long long)n * p.L + qi) * p.lddqkv + qc;
    const float sc = p.scale;
#pragma unroll
    for (int c0 = 0; c0 < 64; c0 += 32) {
      uint32_t a[32];
      tmem_ld32(dq_addr + c0, a);
      tmem_ld_wait();
      if (valid) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 o;
          o.x = pack_f16x2(__uint_as_float(a[8 * g + 0]) * sc, __uint_as_float(a[8 * g + 1]) * sc);
          o.y = pack_f16x2(__uint_as_float(a[8 * g + 2]) * sc, __uint_as_float(a[8 * g + 3]) * sc);
          o.z = pack_f16x2(__uint_as_float(a[8 * g + 4]) * sc, __uint_as_float(a[8 * g + 5]) * sc);
          o.w = pack_f16x2(__uint_as_float(a[8 * g + 6]) * sc, __uint_as_float(a[8 * g + 7]) * sc);
          *reinterpret_cast<uint4*>(dst + c0 + 8 * g) = o;
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

// called by b2d_attention_bwd (attention_bwd.cu) after the delta kernel, for ch == 64 and L > 128
int attention_bwd2_dispatch(const void* qkv, int64_t ldqkv, const void* d_o, int64_t lddo, int N, int L, int heads,
                            int new_order, const float* lse, const float* delta, void* dqkv, int64_t lddqkv,
                            void* stream) {
  B2D_SET_SMEM_ONCE((attn_bwd_dkv2_kernel), Dkv2Smem::TOTAL);
  B2D_SET_SMEM_ONCE((attn_bwd_dq2_kernel), Dq2Smem::TOTAL);
  const int ch = 64, C = heads * ch;
  AttnBwd2Params p;
  p.L = L; p.heads = heads;
  if (new_order) { p.q_off = 0; p.k_off = C; p.v_off = 2 * C; p.head_stride = ch; }
  else           { p.q_off = 0; p.k_off = ch; p.v_off = 2 * ch; p.head_stride = 3 * ch; }
  p.scale = 1.0f / sqrtf((float)ch);
  p.scale_log2 = 1.4426950408889634f * p.scale;
  p.lse = lse; p.delta = delta;
  p.dqkv = (__half*)dqkv; p.lddqkv = lddqkv;
  AttnBwd2Maps m;
  const uint64_t dims[3] = {(uint64_t)(3 * C), (uint64_t)L, (uint64_t)N};
  const uint64_t str[2] = {(uint64_t)ldqkv * 2, (uint64_t)ldqkv * 2 * (uint64_t)L};
  const uint64_t ddims[3] = {(uint64_t)C, (uint64_t)L, (uint64_t)N};
  const uint64_t dstr[2] = {(uint64_t)lddo * 2, (uint64_t)lddo * 2 * (uint64_t)L};
  const uint32_t box128[3] = {64u, 128u, 1u}, box64[3] = {64u, 64u, 1u};
  int rc;
  if ((rc = make_tmap_16b(&m.qkv128, qkv, 3, dims, str, box128))) return rc;
  if ((rc = make_tmap_16b(&m.qkv64, qkv, 3, dims, str, box64))) return rc;
  if ((rc = make_tmap_16b(&m.do128, d_o, 3, ddims, dstr, box128))) return rc;
  if ((rc = make_tmap_16b(&m.do64, d_o, 3, ddims, dstr, box64))) return rc;
  dim3 grid((unsigned)((L + 255) / 256), (unsigned)heads, (unsigned)N);
  cudaStream_t st = (cudaStream_t)stream;
  attn_bwd_dkv2_kernel<<<grid, AB2_THREADS, Dkv2Smem::TOTAL, st>>>(m, p);
  B2D_CHECK_LAUNCH();
  attn_bwd_dq2_kernel<<<grid, AB2_THREADS, Dq2Smem::TOTAL, st>>>(m, p);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

}  // namespace b2d
