// attention_fwd.cu -- flash-style self-attention forward on tcgen05/TMEM (QKVAttentionLegacy /
// QKVAttention, unet.py:340-405):  a = softmax_fp32(q k^T / sqrt(ch)) v  per (sample, head),
// never materialising the L x L matrix.
//
// One CTA = 128 queries of one (sample, head).  S = Q K^T and O += P V run on the tensor cores
// with S (double buffered) and O resident in TMEM; the softmax runs in fp32 registers with one
// query row per thread; P goes back through 128B-swizzled shared memory as the A operand of the
// second GEMM; V is consumed directly in its pixel-major layout through an MN-major descriptor.
//
// Warp roles (192 threads): warp 0 = TMA producer (Q once, K/V ring), warp 1 = MMA issuer,
// warps 2..5 = softmax / O-rescale / epilogue.
#include <stdlib.h>
#include "common.cuh"
#include "tmap.cuh"
#include "umma.cuh"

namespace b2d {
using namespace umma;

struct AttnFwdParams {
  int L, heads;
  int q_off, k_off, v_off, head_stride;  // channel offsets inside the 3C row
  float scale_log2;                      // log2(e) / sqrt(ch)
  __half* out; long long ldo;     // (N, L, ldo), head h at channel h*D
  float* lse;                            // (N, heads, L): log2-domain log-sum-exp of the scaled logits
};

struct AttnMaps {
  CUtensorMap q;   // box (64, 128, 1)
  CUtensorMap kv;  // box (64, BKV, 1)
};

constexpr int ATT_THREADS = 192;

template <int D, int BKV>
struct AttnCfg {
  static constexpr int KV_STAGES = (D == 256) ? 2 : (D == 128 ? 2 : 3);
  static constexpr int Q_BYTES = 128 * D * 2;
  static constexpr int K_BYTES = BKV * D * 2;
  static constexpr int V_BYTES = BKV * D * 2;
  static constexpr int P_BYTES = 128 * BKV * 2;
  static constexpr int STAGE_BYTES = K_BYTES + V_BYTES;
  static constexpr int SMEM_DATA = Q_BYTES + KV_STAGES * STAGE_BYTES + P_BYTES;
  static constexpr int NBARS = 1 + 3 * KV_STAGES + 2 + 1 + 1;  // q, k_full/v_full/kv_empty, s_full[2], p_full, pv_done
  static constexpr int SMEM_TOTAL = SMEM_DATA + NBARS * 8 + 16 + 1024;
  static constexpr uint32_t TMEM_COLS = 512;
  static constexpr uint32_t S_COL0 = 0, S_COL1 = BKV, O_COL = 256;
};

template <int D, int BKV>
__global__ void __launch_bounds__(ATT_THREADS)
attn_fwd_kernel(const __grid_constant__ AttnMaps maps, const AttnFwdParams p) {
  using Cfg = AttnCfg<D, BKV>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t* q_s = smem;
  uint8_t* kv_s = smem + Cfg::Q_BYTES;
  uint8_t* p_s = kv_s + Cfg::KV_STAGES * Cfg::STAGE_BYTES;
  uint64_t* bars = (uint64_t*)(p_s + Cfg::P_BYTES);
  uint64_t* q_full = bars;
  uint64_t* k_full = bars + 1;
  uint64_t* v_full = k_full + Cfg::KV_STAGES;
  uint64_t* kv_empty = v_full + Cfg::KV_STAGES;
  uint64_t* s_full = kv_empty + Cfg::KV_STAGES;  // [2]
  uint64_t* p_full = s_full + 2;
  uint64_t* pv_done = p_full + 1;
  uint32_t* tmem_slot = (uint32_t*)(pv_done + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * 128;
  const int head = blockIdx.y;
  const int n = blockIdx.z;
  const int n_kv = (p.L + BKV - 1) / BKV;
  const int qc = p.q_off + head * p.head_stride;
  const int kc = p.k_off + head * p.head_stride;
  const int vc = p.v_off + head * p.head_stride;

  if (threadIdx.x == 0) {
    mbar_init(q_full, 1);
    for (int s = 0; s < Cfg::KV_STAGES; ++s) { mbar_init(&k_full[s], 1); mbar_init(&v_full[s], 1); mbar_init(&kv_empty[s], 1); }
    mbar_init(&s_full[0], 1); mbar_init(&s_full[1], 1);
    mbar_init(p_full, 128);
    mbar_init(pv_done, 1);
    fence_barrier_init();
    prefetch_tmap(&maps.q);
    prefetch_tmap(&maps.kv);
  }
  if (warp == 1) tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ---------------------------------------------------------------- TMA producer
    if (elect_one()) {
      mbar_expect_tx(q_full, Cfg::Q_BYTES);
#pragma unroll
      for (int b = 0; b < D / 64; ++b) tma_load_3d(q_s + b * (128 * 128), &maps.q, q_full, qc + b * 64, q0, n);
      for (int j = 0; j < n_kv; ++j) {
        const int s = j % Cfg::KV_STAGES;
        const uint32_t par = (uint32_t)((j / Cfg::KV_STAGES) & 1);
        mbar_wait(&kv_empty[s], par ^ 1u);
        uint8_t* k_dst = kv_s + s * Cfg::STAGE_BYTES;
        uint8_t* v_dst = k_dst + Cfg::K_BYTES;
        mbar_expect_tx(&k_full[s], Cfg::K_BYTES);
#pragma unroll
        for (int b = 0; b < D / 64; ++b) tma_load_3d(k_dst + b * (BKV * 128), &maps.kv, &k_full[s], kc + b * 64, j * BKV, n);
        mbar_expect_tx(&v_full[s], Cfg::V_BYTES);
#pragma unroll
        for (int b = 0; b < D / 64; ++b) tma_load_3d(v_dst + b * (BKV * 128), &maps.kv, &v_full[s], vc + b * 64, j * BKV, n);
      }
    }
  } else if (warp == 1) {
    // ---------------------------------------------------------------- MMA issuer
    const uint32_t idesc_s = make_idesc(128, BKV, 0, 0, FMT_F16, FMT_F16);  // S = Q K^T, both K-major
    const uint32_t idesc_o = make_idesc(128, D, 0, 1, FMT_F16, FMT_F16);    // O += P V, V MN-major
    const uint32_t q_addr = smem_u32(q_s);
    const uint32_t p_addr = smem_u32(p_s);
    auto issue_s = [&](int j) {
      const int s = j % Cfg::KV_STAGES;
      mbar_wait(&k_full[s], (uint32_t)((j / Cfg::KV_STAGES) & 1));
      tc_fence_after();
      if (elect_one()) {
        const uint32_t k_addr = smem_u32(kv_s + s * Cfg::STAGE_BYTES);
        const uint32_t d_tmem = tmem_base + ((j & 1) ? Cfg::S_COL1 : Cfg::S_COL0);
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
          const uint32_t a = q_addr + (kk >> 2) * (128 * 128) + (kk & 3) * 32;
          const uint32_t b = k_addr + (kk >> 2) * (BKV * 128) + (kk & 3) * 32;
          mma_f16_ss(d_tmem, make_sdesc_sw128(a, 16, 1024), make_sdesc_sw128(b, 16, 1024), idesc_s, (uint32_t)(kk != 0));
        }
        mma_commit(&s_full[j & 1]);
      }
      __syncwarp();
    };
    mbar_wait(q_full, 0);
    issue_s(0);
    for (int j = 0; j < n_kv; ++j) {
      if (j + 1 < n_kv) issue_s(j + 1);
      const int s = j % Cfg::KV_STAGES;
      mbar_wait(&v_full[s], (uint32_t)((j / Cfg::KV_STAGES) & 1));
      mbar_wait(p_full, (uint32_t)(j & 1));
      tc_fence_after();
      if (elect_one()) {
        const uint32_t v_addr = smem_u32(kv_s + s * Cfg::STAGE_BYTES + Cfg::K_BYTES);
#pragma unroll
        for (int kk = 0; kk < BKV / 16; ++kk) {
          const uint32_t a = p_addr + (kk >> 2) * (128 * 128) + (kk & 3) * 32;
          const uint32_t b = v_addr + kk * 2048;
          mma_f16_ss(tmem_base + Cfg::O_COL, make_sdesc_sw128(a, 16, 1024), make_sdesc_sw128(b, BKV * 128, 1024),
                      idesc_o, (uint32_t)((j | kk) != 0));
        }
        mma_commit(&kv_empty[s]);
        mma_commit(pv_done);
      }
      __syncwarp();
    }
  } else {
    // ---------------------------------------------------------------- softmax / epilogue
    const int qd = warp & 3;
    const int row = qd * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(qd * 32) << 16);
    const float c = p.scale_log2;
    float m_used = -INFINITY;  // raw-logit running max used as the exponent offset
    float l_sum = 0.f;
    for (int j = 0; j < n_kv; ++j) {
      mbar_wait(&s_full[j & 1], (uint32_t)((j >> 1) & 1));
      tc_fence_after();
      const uint32_t s_addr = lane_base + ((j & 1) ? Cfg::S_COL1 : Cfg::S_COL0);
      const int kv_valid = min(BKV, p.L - j * BKV);  // keys of this tile that exist
      // pass 1: row max of the raw logits
      float mx = -INFINITY;
#pragma unroll
      for (int c0 = 0; c0 < BKV; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(s_addr + c0, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float v = (c0 + i < kv_valid) ? __uint_as_float(r[i]) : -INFINITY;
          mx = fmaxf(mx, v);
        }
      }
      // previous PV must be complete before P is overwritten or O rescaled
      if (j > 0) { mbar_wait(pv_done, (uint32_t)((j - 1) & 1)); tc_fence_after(); }
      // lazy rescale: only move the offset when some row's max grew by more than 2^8
      const bool grow = (mx - m_used) * c > 8.0f;
      if (j == 0) {
        m_used = mx;
      } else if (__any_sync(0xffffffffu, grow)) {
        const float m_new = fmaxf(m_used, mx);
        const float f = fast_exp2((m_used - m_new) * c);
        m_used = m_new;
        l_sum *= f;
#pragma unroll
        for (int c0 = 0; c0 < D; c0 += 16) {
          uint32_t r[16];
          tmem_ld16(lane_base + Cfg::O_COL + c0, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * f);
          tmem_st16(lane_base + Cfg::O_COL + c0, r);
        }
        tmem_st_wait();
      }
      // pass 2: p = exp2(c*s - c*m), row sum, P -> swizzled smem (fp16)
      const float mc = m_used * c;
#pragma unroll
      for (int c0 = 0; c0 < BKV; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(s_addr + c0, r);
        tmem_ld_wait();
        float pv[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float e = fast_exp2(fmaf(__uint_as_float(r[i]), c, -mc));
          pv[i] = (c0 + i < kv_valid) ? e : 0.f;
          l_sum += pv[i];
        }
        uint8_t* blk = p_s + (c0 >> 6) * (128 * 128);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 u;
          u.x = pack_f16x2_nosat(pv[8 * g + 0], pv[8 * g + 1]);
          u.y = pack_f16x2_nosat(pv[8 * g + 2], pv[8 * g + 3]);
          u.z = pack_f16x2_nosat(pv[8 * g + 4], pv[8 * g + 5]);
          u.w = pack_f16x2_nosat(pv[8 * g + 6], pv[8 * g + 7]);
          const uint32_t col = (uint32_t)((c0 & 63) + 8 * g);
          sts128(smem_u32(blk) + sw128_offset((uint32_t)row, col), u);
        }
      }
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(p_full);
    }
    // epilogue: O / l -> bf16 global, lse
    mbar_wait(pv_done, (uint32_t)((n_kv - 1) & 1));
    tc_fence_after();
    const int qi = q0 + row;
    const bool valid = qi < p.L;
    const float inv_l = 1.0f / l_sum;
    __half* op = p.out + ((long long)n * p.L + qi) * p.ldo + head * D;
#pragma unroll
    for (int c0 = 0; c0 < D; c0 += 16) {
      uint32_t r[16];
      tmem_ld16(lane_base + Cfg::O_COL + c0, r);
      tmem_ld_wait();
      if (valid) {
        uint4 o0, o1;
        o0.x = pack_f16x2(__uint_as_float(r[0]) * inv_l, __uint_as_float(r[1]) * inv_l);
        o0.y = pack_f16x2(__uint_as_float(r[2]) * inv_l, __uint_as_float(r[3]) * inv_l);
        o0.z = pack_f16x2(__uint_as_float(r[4]) * inv_l, __uint_as_float(r[5]) * inv_l);
        o0.w = pack_f16x2(__uint_as_float(r[6]) * inv_l, __uint_as_float(r[7]) * inv_l);
        o1.x = pack_f16x2(__uint_as_float(r[8]) * inv_l, __uint_as_float(r[9]) * inv_l);
        o1.y = pack_f16x2(__uint_as_float(r[10]) * inv_l, __uint_as_float(r[11]) * inv_l);
        o1.z = pack_f16x2(__uint_as_float(r[12]) * inv_l, __uint_as_float(r[13]) * inv_l);
        o1.w = pack_f16x2(__uint_as_float(r[14]) * inv_l, __uint_as_float(r[15]) * inv_l);
        *reinterpret_cast<uint4*>(op + c0) = o0;
        *reinterpret_cast<uint4*>(op + c0 + 8) = o1;
      }
    }
    if (valid && p.lse) p.lse[((long long)n * p.heads + head) * p.L + qi] = m_used * c + log2f(l_sum);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

template <int D, int BKV>
static int launch_attn_fwd(const AttnMaps& maps, const AttnFwdParams& p, int N, void* stream) {
  using Cfg = AttnCfg<D, BKV>;
  B2D_SET_SMEM_ONCE((attn_fwd_kernel<D, BKV>), Cfg::SMEM_TOTAL);
  dim3 grid((unsigned)((p.L + 127) / 128), (unsigned)p.heads, (unsigned)N);
  attn_fwd_kernel<D, BKV><<<grid, ATT_THREADS, Cfg::SMEM_TOTAL, (cudaStream_t)stream>>>(maps, p);
  B2D_CHECK_LAUNCH();
  return B2D_OK;
}

int attention_fwd2_dispatch(const void* qkv, int64_t ldqkv, int N, int L, int heads, int ch, int new_order, void* out,
                            int64_t ldo, float* lse, int q_row0, int q_rows, void* stream, int kv_splits = 1,
                            float* ws = nullptr);   // attention_fwd2.cu

}  // namespace b2d

using namespace b2d;

extern "C" int b2d_attention_fwd(const void* qkv, int64_t ldqkv, int N, int L, int heads, int ch, int new_order,
                                 void* out, int64_t ldo, float* lse, void* stream) {
  B2D_REQUIRE(qkv && out, B2D_ERR_SHAPE, "attention_fwd: null pointer");
  B2D_REQUIRE(N > 0 && L > 0 && heads > 0, B2D_ERR_SHAPE, "attention_fwd: bad sizes N=%d L=%d heads=%d", N, L, heads);
  B2D_REQUIRE(ch == 64 || ch == 128 || ch == 256, B2D_ERR_SHAPE, "attention_fwd: head dim %d not in {64,128,256}", ch);
  B2D_REQUIRE(N <= 65535 && heads <= 65535, B2D_ERR_SHAPE, "attention_fwd: grid too large");
  B2D_REQUIRE(ldo % 8 == 0 && aligned16(out), B2D_ERR_ALIGN, "attention_fwd: output must be 16-byte aligned rows");
  // two-query-tile kernel for the large layers; B2D_ATTN_FWD_V1=1 forces the single-tile variant (A/B timing)
  static const bool force_v1 = getenv("B2D_ATTN_FWD_V1") != nullptr;
  if (!force_v1 && (ch == 64 || ch == 128) && L > 128)
    return attention_fwd2_dispatch(qkv, ldqkv, N, L, heads, ch, new_order, out, ldo, lse, 0, L, stream);
  const int C = heads * ch;
  AttnFwdParams p;
  p.L = L; p.heads = heads;
  if (new_order) { p.q_off = 0; p.k_off = C; p.v_off = 2 * C; p.head_stride = ch; }          // unet.py:392
  else           { p.q_off = 0; p.k_off = ch; p.v_off = 2 * ch; p.head_stride = 3 * ch; }    // unet.py:359
  p.scale_log2 = 1.4426950408889634f / sqrtf((float)ch);
  p.out = (__half*)out; p.ldo = ldo; p.lse = lse;
  const int bkv = (ch == 256) ? 64 : 128;
  AttnMaps maps;
  const uint64_t dims[3] = {(uint64_t)(3 * C), (uint64_t)L, (uint64_t)N};
  const uint64_t str[2] = {(uint64_t)ldqkv * 2, (uint64_t)ldqkv * 2 * (uint64_t)L};
  const uint32_t boxq[3] = {64u, 128u, 1u};
  const uint32_t boxkv[3] = {64u, (uint32_t)bkv, 1u};
  int rc = make_tmap_16b(&maps.q, qkv, 3, dims, str, boxq);
  if (rc) return rc;
  rc = make_tmap_16b(&maps.kv, qkv, 3, dims, str, boxkv);
  if (rc) return rc;
  if (ch == 64) return launch_attn_fwd<64, 128>(maps, p, N, stream);
  if (ch == 128) return launch_attn_fwd<128, 128>(maps, p, N, stream);
  return launch_attn_fwd<256, 64>(maps, p, N, stream);
}

// Sequence-parallel shard: only query rows [q_row0, q_row0 + q_rows) are computed, against all L
// keys; out / lse rows outside the range are not touched.
extern "C" int b2d_attention_fwd_rows(const void* qkv, int64_t ldqkv, int N, int L, int heads, int ch, int new_order,
                                      int q_row0, int q_rows, void* out, int64_t ldo, float* lse, void* stream) {
  B2D_REQUIRE(qkv && out, B2D_ERR_SHAPE, "attention_fwd_rows: null pointer");
  B2D_REQUIRE(N > 0 && L > 128 && heads > 0 && N <= 65535 && heads <= 65535, B2D_ERR_SHAPE, "attention_fwd_rows: bad sizes");
  B2D_REQUIRE(ch == 64 || ch == 128, B2D_ERR_SHAPE, "attention_fwd_rows: head dim %d not in {64,128}", ch);
  B2D_REQUIRE(q_row0 >= 0 && q_rows > 0 && q_row0 % 256 == 0 && q_rows % 256 == 0 && q_row0 + q_rows <= L, B2D_ERR_SHAPE,
              "attention_fwd_rows: rows [%d, +%d) must be 256-aligned inside L=%d", q_row0, q_rows, L);
  B2D_REQUIRE(ldo % 8 == 0 && aligned16(out), B2D_ERR_ALIGN, "attention_fwd_rows: output must be 16-byte aligned rows");
  return attention_fwd2_dispatch(qkv, ldqkv, N, L, heads, ch, new_order, out, ldo, lse, q_row0, q_rows, stream);
}

// Small grids (batch-1 sampling: one 128 x 128 frame = 64 CTAs of 256 queries on 148 SMs): the key range is split over
// `splits` CTAs per query block, each writes its normalised fp32 output and log-sum-exp to `ws`, a second kernel merges
// them.  Same result as b2d_attention_fwd up to the 16-bit rounding of the output.  ws: b2d_attention_fwd_splitkv_ws_floats.
extern "C" int64_t b2d_attention_fwd_splitkv_ws_floats(int N, int L, int heads, int ch, int splits) {
  if (splits < 1) splits = 1;
  if (splits > 8) splits = 8;
  return (int64_t)splits * N * L * heads * (ch + 1);
}

extern "C" int b2d_attention_fwd_splitkv(const void* qkv, int64_t ldqkv, int N, int L, int heads, int ch, int new_order,
                                         int splits, float* ws, int64_t ws_floats, void* out, int64_t ldo, float* lse,
                                         void* stream) {
  B2D_REQUIRE(qkv && out && ws, B2D_ERR_SHAPE, "attention_fwd_splitkv: null pointer");
  B2D_REQUIRE(N > 0 && L > 128 && heads > 0 && N <= 65535 && heads <= 65535, B2D_ERR_SHAPE, "attention_fwd_splitkv: bad sizes");
  B2D_REQUIRE(ch == 64 || ch == 128, B2D_ERR_SHAPE, "attention_fwd_splitkv: head dim %d not in {64,128}", ch);
  B2D_REQUIRE(splits >= 1 && splits <= 8, B2D_ERR_SHAPE, "attention_fwd_splitkv: splits=%d not in [1, 8]", splits);
  B2D_REQUIRE(ws_floats >= b2d_attention_fwd_splitkv_ws_floats(N, L, heads, ch, splits) && aligned16(ws), B2D_ERR_WORKSPACE,
              "attention_fwd_splitkv: workspace of %lld floats too small or misaligned", (long long)ws_floats);
  B2D_REQUIRE(ldo % 8 == 0 && aligned16(out), B2D_ERR_ALIGN, "attention_fwd_splitkv: output must be 16-byte aligned rows");
  return attention_fwd2_dispatch(qkv, ldqkv, N, L, heads, ch, new_order, out, ldo, lse, 0, L, stream, splits, ws);
}
