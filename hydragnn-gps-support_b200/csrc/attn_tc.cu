// GPS global attention on the 5th-generation tensor cores (tcgen05.mma kind::tf32, accumulators
// in TMEM), fp32-accurate through 3xTF32 operand splitting.  Same contract as the SIMT kernels in
// attn.cu (reference: hydragnn/globalAtt/gps.py:126-133 -- to_dense_batch + nn.MultiheadAttention
// core + unpad), selected by hgk_attn_varlen_* when the head dim is 16 or 32.
//
// Common shape of the three kernels: a CTA of 256 threads owns a 128-row tile (TMEM lane = row) and
// loops over 64-column tiles of the other sequence.  Warp w works on TMEM lanes 32*(w%4)..+31 (the
// only lanes the hardware lets it touch) and on column half w/4 of the current [128 x 64] block,
// i.e. two threads share a row.  One elected thread issues the MMAs; an mbarrier (tcgen05.commit)
// tells everyone when the accumulators are ready.  Global loads of the next tile are issued into
// registers before waiting on the MMAs of the current one.
//
//   forward   S = Q K^T (SS)  ->  online softmax in registers (exp2 domain)  ->  P (hi, lo) back to
//             TMEM  ->  O_tile = P V (TS, A = P from TMEM)  ->  acc = acc * corr + O_tile
//   backward  P = 2^(s - lse2),  dP = dO V^T,  dS = P (dP - delta),  delta = rowsum(dO O)
//     dQ kernel  : lanes = queries;  S, dP (SS)  ->  dS (hi, lo) to TMEM  ->  dQ += dS K     (TS)
//     dK/dV kern.: lanes = keys;   S^T, dP^T (SS) -> P^T, dS^T to TMEM -> dV += P^T dO, dK += dS^T Q (TS)
//   Deterministic: every output element is produced by exactly one CTA, no float atomics.
#include "hgk_common.cuh"
#include "tc_common.cuh"

namespace hgk {

struct AttnTcArgs {
  const float* q; int64_t ldq;
  const float* k; int64_t ldk;
  const float* v; int64_t ldv;
  const int32_t* cu; int32_t S;
  int32_t H;
  float scale;
};

namespace tc {

constexpr int BM = 128;          // rows per CTA (= TMEM lanes)
constexpr int BN = 64;           // columns per iteration
constexpr int HALF = BN / 2;     // columns per thread

__device__ __forceinline__ bool locate_tile128(const int32_t* cu, int32_t S, int32_t tile,
                                               int32_t* seg_beg, int32_t* seg_end, int32_t* row0) {
  int32_t acc = 0;
  for (int32_t s = 0; s < S; ++s) {
    const int32_t b = cu[s], e = cu[s + 1];
    const int32_t nt = (e - b + BM - 1) / BM;
    if (tile < acc + nt) {
      *seg_beg = b; *seg_end = e; *row0 = b + (tile - acc) * BM;
      return true;
    }
    acc += nt;
  }
  return false;
}

// ================================================================================= forward ===
template <int D>
__global__ void __launch_bounds__(kThreads, 2)
k_attn_fwd_tc(const AttnTcArgs p, float* __restrict__ o, int64_t ldo, float* __restrict__ lse) {
  extern __shared__ __align__(128) float smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  __shared__ float xchg[2][BM];                        // row max / row sum exchange between halves

  int32_t seg_beg, seg_end, row0;
  if (!locate_tile128(p.cu, p.S, blockIdx.x, &seg_beg, &seg_end, &row0)) return;
  const int h = blockIdx.y;
  const int warp_u = warp_id_uniform();

  float* Khi = smem;             float* Klo = Khi + BN * D;
  float* Vhi = Klo + BN * D;     float* Vlo = Vhi + BN * D;      // V^T tile [D x BN]

  constexpr uint32_t kCols = 256;
  // P_hi overwrites S in place; Q (hi, lo) lives in TMEM for the whole kernel as the A operand of S
  constexpr uint32_t kColS = 0, kColPlo = 64, kColO = 128, kColQhi = 128 + D, kColQlo = 128 + 2 * D;
  static_assert(128 + 3 * D <= 256, "TMEM budget");
  const CtaSetup cs = cta_setup(&bar, &tmem_slot, kCols);
  const uint32_t tmem = cs.tmem;

  // scores come out in the log2 domain: Q is pre-multiplied by scale * log2(e)
  row_to_tmem<D>(tmem + cs.lane_base + kColQhi, tmem + cs.lane_base + kColQlo, p.q + h * D, p.ldq,
                 row0 + cs.row_local, row0 + cs.row_local < seg_end, cs.half, p.scale * kLog2e);
  const OperandDesc dK = operand_desc<D>(Khi, Klo);
  const OperandDesc dVt = operand_desc<BN>(Vhi, Vlo);
  Tile<BN, D> tk, tv;                                  // next key / value tile, prefetched in registers
  tk.load(p.k + h * D, p.ldk, seg_beg, seg_end);
  tv.load(p.v + h * D, p.ldv, seg_beg, seg_end);

  constexpr uint32_t kIdescS = idesc_tf32(BM, BN);
  constexpr uint32_t kIdescO = idesc_tf32(BM, D);
  constexpr int DH = D / 2;                            // output columns per thread
  float acc[DH];
#pragma unroll
  for (int d = 0; d < DH; ++d) acc[d] = 0.f;
  float m_run = -INFINITY, l_run = 0.f;                // l_run: this thread's half of the row sum
  uint32_t parity = 0;

  for (int32_t k0 = seg_beg; k0 < seg_end; k0 += BN) {
    tk.store_rows(Khi, Klo);
    tv.store_transposed(Vhi, Vlo);
    publish_operands();
    if (warp_u == 0 && elect_one_sync()) {
      fence_after_sync();
      mma3_ts<D>(tmem + kColS, tmem + kColQhi, tmem + kColQlo, dK, kIdescS, false);
      mma_commit(&bar);
    }
    if (k0 + BN < seg_end) {                           // global loads of the next tile fly under the MMAs
      tk.load(p.k + h * D, p.ldk, k0 + BN, seg_end);
      tv.load(p.v + h * D, p.ldv, k0 + BN, seg_end);
    }
    mbar_wait(&bar, parity);
    parity ^= 1;
    fence_after_sync();

    uint32_t s[HALF];
    tmem_ld32(tmem + cs.lane_base + kColS + cs.half * HALF, s);
    tmem_ld_wait();
    const int32_t nvalid = seg_end - k0 - cs.half * HALF;   // my columns [0, nvalid) exist
    if (nvalid < HALF) {                               // last tile of the segment only (warp uniform)
#pragma unroll
      for (int j = 0; j < HALF; ++j)
        if (j >= nvalid) s[j] = 0xff800000u;           // -inf
    }
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < HALF; ++j) mx = fmaxf(mx, __uint_as_float(s[j]));
    xchg[cs.half][cs.row_local] = mx;
    __syncthreads();
    mx = fmaxf(mx, xchg[cs.half ^ 1][cs.row_local]);
    const float m_new = fmaxf(m_run, mx);              // finite: every key tile has a valid key
    const float corr = fast_exp2(m_run - m_new);
    float rs = 0.f;
    uint32_t lo[HALF];
#pragma unroll
    for (int j = 0; j < HALF; ++j) {
      const float pj = fast_exp2(__uint_as_float(s[j]) - m_new);
      rs += pj;
      float hi_, lo_;
      split_tf32(pj, hi_, lo_);
      s[j] = __float_as_uint(hi_);
      lo[j] = __float_as_uint(lo_);
    }
    tmem_st32(tmem + cs.lane_base + kColS + cs.half * HALF, s);
    tmem_st32(tmem + cs.lane_base + kColPlo + cs.half * HALF, lo);
    tmem_st_wait();
    fence_before_sync();
    __syncthreads();
    if (warp_u == 0 && elect_one_sync()) {
      fence_after_sync();
      mma3_ts<BN>(tmem + kColO, tmem + kColS, tmem + kColPlo, dVt, kIdescO, false);
      mma_commit(&bar);
    }
    mbar_wait(&bar, parity);
    parity ^= 1;
    fence_after_sync();

    l_run = l_run * corr + rs;
    m_run = m_new;
    uint32_t ot[16];
#pragma unroll
    for (int c0 = 0; c0 < DH; c0 += 16) {
      tmem_ld16(tmem + cs.lane_base + kColO + cs.half * DH + c0, ot);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 16; ++j)
        if (c0 + j < DH) acc[c0 + j] = fmaf(acc[c0 + j], corr, __uint_as_float(ot[j]));
    }
  }

  // full row sum = sum of the two halves (both carry the same running max)
  __syncthreads();
  xchg[cs.half][cs.row_local] = l_run;
  __syncthreads();
  const float l_tot = l_run + xchg[cs.half ^ 1][cs.row_local];
  const int32_t row = row0 + cs.row_local;
  if (row < seg_end) {
    const float inv = 1.f / l_tot;
    float* orow = o + (int64_t)row * ldo + h * D + cs.half * DH;
#pragma unroll
    for (int d = 0; d < DH; d += 4)
      *reinterpret_cast<float4*>(orow + d) =
          make_float4(acc[d] * inv, acc[d + 1] * inv, acc[d + 2] * inv, acc[d + 3] * inv);
    if (cs.half == 0) lse[(int64_t)row * p.H + h] = (m_run + log2f(l_tot)) * kLn2;
  }
  cta_teardown(tmem, kCols);
}

// ================================================================================ backward ===
struct AttnTcBwdArgs {
  const float* dout; int64_t lddo;
  const float* lse; const float* delta;
};

// ---- dQ: CTA owns 128 queries, loops over 64-key tiles ---------------------------------------
template <int D>
__global__ void __launch_bounds__(kThreads, 2)
k_attn_bwd_q_tc(const AttnTcArgs p, const AttnTcBwdArgs b, float* __restrict__ dq) {
  extern __shared__ __align__(128) float smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  int32_t seg_beg, seg_end, row0;
  if (!locate_tile128(p.cu, p.S, blockIdx.x, &seg_beg, &seg_end, &row0)) return;
  const int h = blockIdx.y;
  const int warp_u = warp_id_uniform();

  float* Ghi = smem;             float* Glo = Ghi + BM * D;      // dO tile
  float* Khi = Glo + BM * D;     float* Klo = Khi + BN * D;
  float* Vhi = Klo + BN * D;     float* Vlo = Vhi + BN * D;
  float* KThi = Vlo + BN * D;    float* KTlo = KThi + BN * D;    // K^T tile [D x BN]

  constexpr uint32_t kCols = 256;
  // dS_hi -> S region, dS_lo -> dP region; Q (hi, lo) resident in TMEM as the A operand of S
  constexpr uint32_t kColS = 0, kColDP = 64, kColDQ = 128, kColQhi = 128 + D, kColQlo = 128 + 2 * D;
  const CtaSetup cs = cta_setup(&bar, &tmem_slot, kCols);
  const uint32_t tmem = cs.tmem;

  row_to_tmem<D>(tmem + cs.lane_base + kColQhi, tmem + cs.lane_base + kColQlo, p.q + h * D, p.ldq,
                 row0 + cs.row_local, row0 + cs.row_local < seg_end, cs.half, p.scale * kLog2e);
  {
    Tile<BM, D> t;
    t.load(b.dout + h * D, b.lddo, row0, seg_end);
    t.store_rows(Ghi, Glo);
  }
  const OperandDesc dG = operand_desc<D>(Ghi, Glo);
  const OperandDesc dK = operand_desc<D>(Khi, Klo), dV = operand_desc<D>(Vhi, Vlo);
  const OperandDesc dKt = operand_desc<BN>(KThi, KTlo);
  Tile<BN, D> tk, tv;
  tk.load(p.k + h * D, p.ldk, seg_beg, seg_end);
  tv.load(p.v + h * D, p.ldv, seg_beg, seg_end);
  const int32_t row = row0 + cs.row_local;
  const bool row_ok = row < seg_end;
  const float lse2 = row_ok ? b.lse[(int64_t)row * p.H + h] * kLog2e : 0.f;
  const float delta = row_ok ? b.delta[(int64_t)row * p.H + h] : 0.f;

  constexpr uint32_t kIdescS = idesc_tf32(BM, BN);
  constexpr uint32_t kIdescQ = idesc_tf32(BM, D);
  uint32_t parity = 0;
  bool first = true;

  for (int32_t k0 = seg_beg; k0 < seg_end; k0 += BN) {
    tk.store_rows(Khi, Klo);
    tk.store_transposed(KThi, KTlo);
    tv.store_rows(Vhi, Vlo);
    publish_operands();
    if (warp_u == 0 && elect_one_sync()) {
      fence_after_sync();
      mma3_ts<D>(tmem + kColS, tmem + kColQhi, tmem + kColQlo, dK, kIdescS, false);   // S  = Q K^T
      mma3_ss<D>(tmem + kColDP, dG, dV, kIdescS, false);                 // dP = dO V^T
      mma_commit(&bar);
    }
    if (k0 + BN < seg_end) {
      tk.load(p.k + h * D, p.ldk, k0 + BN, seg_end);
      tv.load(p.v + h * D, p.ldv, k0 + BN, seg_end);
    }
    mbar_wait(&bar, parity);
    parity ^= 1;
    fence_after_sync();

    const int32_t nvalid = seg_end - k0 - cs.half * HALF;
    uint32_t sv[HALF], dv[HALF];
    tmem_ld32(tmem + cs.lane_base + kColS + cs.half * HALF, sv);
    tmem_ld32(tmem + cs.lane_base + kColDP + cs.half * HALF, dv);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < HALF; ++j) {
      const float pj = fast_exp2(__uint_as_float(sv[j]) - lse2);
      const float ds = pj * (__uint_as_float(dv[j]) - delta);
      float hi_, lo_;
      split_tf32(ds, hi_, lo_);
      sv[j] = __float_as_uint(hi_);
      dv[j] = __float_as_uint(lo_);
    }
    if (nvalid < HALF) {                               // keys past the segment end contribute nothing
#pragma unroll
      for (int j = 0; j < HALF; ++j)
        if (j >= nvalid) { sv[j] = 0u; dv[j] = 0u; }
    }
    tmem_st32(tmem + cs.lane_base + kColS + cs.half * HALF, sv);
    tmem_st32(tmem + cs.lane_base + kColDP + cs.half * HALF, dv);
    tmem_st_wait();
    fence_before_sync();
    __syncthreads();
    if (warp_u == 0 && elect_one_sync()) {
      fence_after_sync();
      mma3_ts<BN>(tmem + kColDQ, tmem + kColS, tmem + kColDP, dKt, kIdescQ, !first);           // dQ += dS K
      mma_commit(&bar);
    }
    first = false;
    mbar_wait(&bar, parity);                           // K / K^T / V tiles and S / dP are free again
    parity ^= 1;
    fence_after_sync();
  }

  // tcgen05.ld is warp-collective (.sync.aligned): every thread loads, valid rows store
  constexpr int DH = D / 2;
  float* drow = dq + ((int64_t)row * p.H + h) * D + cs.half * DH;
  uint32_t t[16];
#pragma unroll
  for (int c0 = 0; c0 < DH; c0 += 16) {
    tmem_ld16(tmem + cs.lane_base + kColDQ + cs.half * DH + c0, t);
    tmem_ld_wait();
    if (row_ok) {
#pragma unroll
      for (int j = 0; j < 16; j += 4)
        if (c0 + j < DH)
          *reinterpret_cast<float4*>(drow + c0 + j) =
              make_float4(__uint_as_float(t[j]) * p.scale, __uint_as_float(t[j + 1]) * p.scale,
                          __uint_as_float(t[j + 2]) * p.scale, __uint_as_float(t[j + 3]) * p.scale);
    }
  }
  cta_teardown(tmem, kCols);
}

// ---- dK, dV: CTA owns 128 keys (TMEM lane = key), loops over 64-query tiles -------------------
// 512 threads: one CTA per SM (it needs 448 TMEM columns), so the warps have to come from inside
// the CTA -- four threads share a key row, 16 of the 64 query columns each.
constexpr int kKvThreads = 512;

template <int D>
__global__ void __launch_bounds__(kKvThreads, 1)
k_attn_bwd_kv_tc(const AttnTcArgs p, const AttnTcBwdArgs b, float* __restrict__ dk, float* __restrict__ dv) {
  constexpr int BQ = BN;                               // queries per iteration
  constexpr int QC = BQ / 4;                           // query columns per thread
  extern __shared__ __align__(128) float smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  __shared__ float lse2_s[2][BQ], delta_s[2][BQ];
  int32_t seg_beg, seg_end, key0;
  if (!locate_tile128(p.cu, p.S, blockIdx.x, &seg_beg, &seg_end, &key0)) return;
  const int h = blockIdx.y;
  const int warp_u = warp_id_uniform();

  // two shared-memory stages of the per-iteration operands: stage (it+1) is written while the
  // second MMA batch of iteration `it` is still running
  constexpr int kStage = 8 * BQ * D;                   // floats per stage
  auto Qhi = [&](int st) { return smem + st * kStage; };                 // rows = queries (B of S^T)
  auto Qlo = [&](int st) { return smem + st * kStage + 1 * BQ * D; };
  auto Ghi = [&](int st) { return smem + st * kStage + 2 * BQ * D; };    // dO rows        (B of dP^T)
  auto Glo = [&](int st) { return smem + st * kStage + 3 * BQ * D; };
  auto QThi = [&](int st) { return smem + st * kStage + 4 * BQ * D; };   // Q^T [D x BQ]   (B of dK)
  auto QTlo = [&](int st) { return smem + st * kStage + 5 * BQ * D; };
  auto GThi = [&](int st) { return smem + st * kStage + 6 * BQ * D; };   // dO^T [D x BQ]  (B of dV)
  auto GTlo = [&](int st) { return smem + st * kStage + 7 * BQ * D; };

  constexpr uint32_t kCols = 512;
  constexpr uint32_t kColST = 0, kColDPT = 64, kColPlo = 128, kColDSlo = 192, kColDV = 256, kColDK = 256 + D;
  // this CTA's K and V rows (hi, lo) stay in TMEM as the A operands of S^T and dP^T
  constexpr uint32_t kColKhi = 256 + 2 * D, kColKlo = 256 + 3 * D, kColVhi = 256 + 4 * D, kColVlo = 256 + 5 * D;
  static_assert(256 + 6 * D <= 512, "TMEM budget");
  const CtaSetup cs = cta_setup(&bar, &tmem_slot, kCols);     // cs.half = column quarter 0..3
  const uint32_t tmem = cs.tmem;
  const int part = cs.half;

  {
    const int64_t krow_ld = key0 + cs.row_local;
    const bool kok = krow_ld < seg_end;
    if (part < 2)
      row_to_tmem<D>(tmem + cs.lane_base + kColKhi, tmem + cs.lane_base + kColKlo, p.k + h * D, p.ldk, krow_ld, kok, part, 1.f);
    else
      row_to_tmem<D>(tmem + cs.lane_base + kColVhi, tmem + cs.lane_base + kColVlo, p.v + h * D, p.ldv, krow_ld, kok, part - 2, 1.f);
  }
  const OperandDesc dQ0 = operand_desc<D>(Qhi(0), Qlo(0)), dQ1 = operand_desc<D>(Qhi(1), Qlo(1));
  const OperandDesc dG0 = operand_desc<D>(Ghi(0), Glo(0)), dG1 = operand_desc<D>(Ghi(1), Glo(1));
  const OperandDesc dQt0 = operand_desc<BQ>(QThi(0), QTlo(0)), dQt1 = operand_desc<BQ>(QThi(1), QTlo(1));
  const OperandDesc dGt0 = operand_desc<BQ>(GThi(0), GTlo(0)), dGt1 = operand_desc<BQ>(GThi(1), GTlo(1));
  Tile<BQ, D, kKvThreads> tq, tg;                      // next query / dO tile, prefetched in registers
  tq.load(p.q + h * D, p.ldq, seg_beg, seg_end);
  tg.load(b.dout + h * D, b.lddo, seg_beg, seg_end);

  constexpr uint32_t kIdescS = idesc_tf32(BM, BQ);
  constexpr uint32_t kIdescD = idesc_tf32(BM, D);
  const float qmul = p.scale * kLog2e;
  uint32_t parity = 0;
  bool first = true;

  // registers (tile starting at query q_at) -> shared-memory stage st
  auto stage = [&](int st, int32_t q_at) {
    tq.scale(qmul);
    tq.store_rows(Qhi(st), Qlo(st));
    tq.store_transposed(QThi(st), QTlo(st));
    tg.store_rows(Ghi(st), Glo(st));
    tg.store_transposed(GThi(st), GTlo(st));
    if (threadIdx.x < BQ) {
      const int32_t qrow = q_at + (int32_t)threadIdx.x;
      const bool ok = qrow < seg_end;
      lse2_s[st][threadIdx.x] = ok ? b.lse[(int64_t)qrow * p.H + h] * kLog2e : 0.f;
      delta_s[st][threadIdx.x] = ok ? b.delta[(int64_t)qrow * p.H + h] : 0.f;
    }
  };
  stage(0, seg_beg);

  int st = 0;
  for (int32_t q0 = seg_beg; q0 < seg_end; q0 += BQ, st ^= 1) {
    const bool has_next = q0 + BQ < seg_end;
    publish_operands();                                // stage `st` is complete and visible
    if (warp_u == 0 && elect_one_sync()) {
      fence_after_sync();
      mma3_ts<D>(tmem + kColST, tmem + kColKhi, tmem + kColKlo, st ? dQ1 : dQ0, kIdescS, false);    // S^T  = K Q^T
      mma3_ts<D>(tmem + kColDPT, tmem + kColVhi, tmem + kColVlo, st ? dG1 : dG0, kIdescS, false);   // dP^T = V dO^T
      mma_commit(&bar);
    }
    if (has_next) {                                    // global loads of the next tile fly under the MMAs
      tq.load(p.q + h * D, p.ldq, q0 + BQ, seg_end);
      tg.load(b.dout + h * D, b.lddo, q0 + BQ, seg_end);
    }
    mbar_wait(&bar, parity);
    parity ^= 1;
    fence_after_sync();

    const int col0 = part * QC;
    const int32_t nvalid = seg_end - q0 - col0;        // my query columns [0, nvalid) exist
    uint32_t sv[QC], dp[QC], plo[QC], dslo[QC];
    tmem_ld16(tmem + cs.lane_base + kColST + col0, sv);
    tmem_ld16(tmem + cs.lane_base + kColDPT + col0, dp);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < QC; ++j) {
      float pj = fast_exp2(__uint_as_float(sv[j]) - lse2_s[st][col0 + j]);
      if (j >= nvalid) pj = 0.f;
      const float ds = pj * (__uint_as_float(dp[j]) - delta_s[st][col0 + j]);
      float hi_, lo_;
      split_tf32(pj, hi_, lo_);
      sv[j] = __float_as_uint(hi_); plo[j] = __float_as_uint(lo_);
      split_tf32(ds, hi_, lo_);
      dp[j] = __float_as_uint(hi_); dslo[j] = __float_as_uint(lo_);
    }
    tmem_st16(tmem + cs.lane_base + kColST + col0, sv);
    tmem_st16(tmem + cs.lane_base + kColPlo + col0, plo);
    tmem_st16(tmem + cs.lane_base + kColDPT + col0, dp);
    tmem_st16(tmem + cs.lane_base + kColDSlo + col0, dslo);
    tmem_st_wait();
    fence_before_sync();
    __syncthreads();
    if (warp_u == 0 && elect_one_sync()) {
      fence_after_sync();
      mma3_ts<BQ>(tmem + kColDV, tmem + kColST, tmem + kColPlo, st ? dGt1 : dGt0, kIdescD, !first);       // dV += P^T dO
      mma3_ts<BQ>(tmem + kColDK, tmem + kColDPT, tmem + kColDSlo, st ? dQt1 : dQt0, kIdescD, !first);     // dK += dS^T Q
      mma_commit(&bar);
    }
    first = false;
    if (has_next) stage(st ^ 1, q0 + BQ);              // overlaps the MMAs just issued
    mbar_wait(&bar, parity);
    parity ^= 1;
    fence_after_sync();
  }

  // parts 0, 1 write the two halves of dV's row, parts 2, 3 those of dK (D = 32: 16 columns each)
  static_assert(D == 32, "epilogue column split");
  constexpr int EC = D / 2;                            // columns per thread
  const int32_t krow = key0 + cs.row_local;
  const bool ok = krow < seg_end;
  const bool is_k = part >= 2;
  const int chalf = part & 1;
  uint32_t t[16];
  tmem_ld16(tmem + cs.lane_base + (is_k ? kColDK : kColDV) + chalf * EC, t);
  tmem_ld_wait();
  if (ok) {
    // Q was staged as q * scale * log2(e):  dK = scale * dS^T Q = (dS^T Q_staged) * ln 2
    const float mul = is_k ? kLn2 : 1.f;
    float* orow = (is_k ? dk : dv) + ((int64_t)krow * p.H + h) * D + chalf * EC;
#pragma unroll
    for (int j = 0; j < EC; j += 4)
      *reinterpret_cast<float4*>(orow + j) =
          make_float4(__uint_as_float(t[j]) * mul, __uint_as_float(t[j + 1]) * mul,
                      __uint_as_float(t[j + 2]) * mul, __uint_as_float(t[j + 3]) * mul);
  }
  cta_teardown(tmem, kCols);
}

// ---------------------------------------------------------------------------------- launch ---
template <typename K>
int opt_in_smem(K kernel, size_t bytes, const char* what) {
  cudaError_t err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (err != cudaSuccess)
    return fail(HGK_ERR_CUDA, "%s: smem opt-in (%zu B) failed: %s", what, bytes, cudaGetErrorString(err));
  return HGK_OK;
}

template <int D>
int launch_fwd(const AttnTcArgs& a, float* o, int64_t ldo, float* lse, int64_t N, cudaStream_t st) {
  // 2 CTAs per SM: 256 TMEM columns each; the shared-memory request (> 1/3 of the SM) enforces it
  size_t smem = sizeof(float) * (4 * BN * D);
  if (smem < 80 * 1024) smem = 80 * 1024;
  int rc = opt_in_smem(k_attn_fwd_tc<D>, smem, "attn_fwd_tc");
  if (rc) return rc;
  dim3 grid((unsigned)(ceil_div(N, BM) + a.S), (unsigned)a.H);
  k_attn_fwd_tc<D><<<grid, kThreads, smem, st>>>(a, o, ldo, lse);
  return check_launch("attn_varlen_fwd(tcgen05)");
}

template <int D>
int launch_bwd(const AttnTcArgs& a, const AttnTcBwdArgs& b, float* dq, float* dk, float* dv, int64_t N,
               cudaStream_t st) {
  dim3 grid((unsigned)(ceil_div(N, BM) + a.S), (unsigned)a.H);
  // 512 TMEM columns: one CTA per SM; ask for more than half of the shared memory to enforce it
  size_t smem_kv = sizeof(float) * (2 * 8 * BN * D);
  if (smem_kv < 120 * 1024) smem_kv = 120 * 1024;
  int rc = opt_in_smem(k_attn_bwd_kv_tc<D>, smem_kv, "attn_bwd_kv_tc");
  if (rc) return rc;
  k_attn_bwd_kv_tc<D><<<grid, kKvThreads, smem_kv, st>>>(a, b, dk, dv);
  size_t smem_q = sizeof(float) * (2 * BM * D + 6 * BN * D);
  if (smem_q < 80 * 1024) smem_q = 80 * 1024;
  rc = opt_in_smem(k_attn_bwd_q_tc<D>, smem_q, "attn_bwd_q_tc");
  if (rc) return rc;
  k_attn_bwd_q_tc<D><<<grid, kThreads, smem_q, st>>>(a, b, dq);
  return check_launch("attn_varlen_bwd(tcgen05)");
}

}  // namespace tc

static bool tc_aligned(const void* ptr, int64_t ld) {
  return (reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && ld % 4 == 0;
}

// Called by hgk_attn_varlen_fwd.  Returns 1 if this shape is not handled here (caller falls back
// to the SIMT kernel), 0 on success, a negative hgk_status on error.
int attn_fwd_tc(float* o, int64_t ldo, float* lse, const float* q, int64_t ldq, const float* k,
                int64_t ldk, const float* v, int64_t ldv, const int32_t* cu, int64_t S, int64_t N,
                int32_t H, int32_t D, float scale, cudaStream_t st) {
  if (!(D == 16 || D == 32)) return 1;
  if (!tc_aligned(q, ldq) || !tc_aligned(k, ldk) || !tc_aligned(v, ldv) || !tc_aligned(o, ldo)) return 1;
  AttnTcArgs a{q, ldq, k, ldk, v, ldv, cu, (int32_t)S, H, scale};
  if (D == 16) return tc::launch_fwd<16>(a, o, ldo, lse, N, st);
  return tc::launch_fwd<32>(a, o, ldo, lse, N, st);
}

// Called by hgk_attn_varlen_bwd after delta = rowsum(dO * O) has been enqueued.  Same return
// convention as attn_fwd_tc.  dq, dk, dv are contiguous [N, H, D].
int attn_bwd_tc(float* dq, float* dk, float* dv, const float* delta, const float* dout, int64_t lddo,
                const float* lse, const float* q, int64_t ldq, const float* k, int64_t ldk,
                const float* v, int64_t ldv, const int32_t* cu, int64_t S, int64_t N, int32_t H,
                int32_t D, float scale, cudaStream_t st) {
  if (D != 32) return 1;    // the 512-thread dK/dV kernel is laid out for 32-wide heads
  if (!tc_aligned(q, ldq) || !tc_aligned(k, ldk) || !tc_aligned(v, ldv) || !tc_aligned(dout, lddo) ||
      !tc_aligned(dq, 4) || !tc_aligned(dk, 4) || !tc_aligned(dv, 4))
    return 1;
  AttnTcArgs a{q, ldq, k, ldk, v, ldv, cu, (int32_t)S, H, scale};
  tc::AttnTcBwdArgs b{dout, lddo, lse, delta};
  return tc::launch_bwd<32>(a, b, dq, dk, dv, N, st);
}

}  // namespace hgk
