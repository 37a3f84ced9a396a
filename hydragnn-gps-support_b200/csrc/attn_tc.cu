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
#include <cstring>
#include <mutex>

#include "hgk_common.cuh"
#include "tc_common.cuh"
#include "tma_host.cuh"

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

// Long accumulation chains in TMEM lose precision: every tcgen05.mma adds into the fp32 accumulator
// with truncation, so the error of a sum grows linearly with the number of MMAs behind it (measured,
// profiles/precision_probe.py: 6e-5 of the tensor's scale after the 3072 MMAs of an 8192-key segment,
// against 1e-6 for plain fp32).  The pipelined kernels therefore restart their output accumulators
// every kPromote of a group's steps (48 MMAs) and add the finished partial sums to fp32 totals in
// shared memory with ordinary (round-to-nearest) adds.
constexpr int kPromote = 4;

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

// ---- forward, software pipelined --------------------------------------------------------------
// CTA = 128 queries (TMEM lane = query), 544 threads: kFwdGroups compute groups of four warps (one
// thread per query row, all 32 key columns of a step) and one MMA-issuing warp.  Group g owns the
// 32-key steps s = g (mod kFwdGroups), its own TMEM column set (S -> P_hi in place, P_lo, its
// output accumulator) and its own online-softmax state (reference max, row sum) over its steps;
// the states are merged once at the end (the split-K form of the online softmax).
// Hand-offs as in the backward kernels: group -> issuer through named barrier 2+g (bar.arrive),
// issuer -> group through mbarrier g (tcgen05.commit).  For step s the issuer queues
// [O_g += P_s V_s,  S_{s+G} = Q K_{s+G}^T]; while a group waits for its next scores the other
// groups keep the ALUs and the TMEM ports busy.  Shared-memory ring of 2G stages: the stage
// written during step s (for step s+G) was last read by step s-G, which the mbarrier wait at the
// top of step s has covered.
constexpr int kFwdGroups = 4;
constexpr int kFwdGroup = 128;
constexpr int kFwdThreads = kFwdGroups * kFwdGroup + 32;

template <int D>
__global__ void __launch_bounds__(kFwdThreads, 1)
k_attn_fwd_pipe(const AttnTcArgs p, float* __restrict__ o, int64_t ldo, float* __restrict__ lse) {
  constexpr int G = kFwdGroups;
  constexpr int BK = 32;                               // keys per pipeline step
  constexpr int NST = 2 * G;                           // shared-memory stages
  extern __shared__ __align__(1024) float smem_raw[];
  // the MN-major operand tiles need 1024-byte alignment; the launchers request 1024 bytes more than the layout
  float* const smem = smem_raw + (((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u) >> 2);
  __shared__ __align__(8) uint64_t bar[G + 1];         // [g]: scores of group g are ready; [G]: everything done
  __shared__ uint32_t tmem_slot;
  int32_t seg_beg, seg_end, row0;
  if (!locate_tile128(p.cu, p.S, blockIdx.x, &seg_beg, &seg_end, &row0)) return;
  const int h = blockIdx.y;
  const int warp_u = warp_id_uniform();
  const bool issuer = warp_u == G * kFwdGroup / 32;    // the last warp

  constexpr int kTile = BK * D;
  constexpr int kStage = 4 * kTile;                    // K rows hi/lo (B = K of the scores), V rows hi/lo (B = V^T)
  constexpr int oKhi = 0, oKlo = kTile, oVThi = 2 * kTile, oVTlo = 3 * kTile;
  constexpr uint32_t kCols = 512;
  constexpr uint32_t kColS = 0, kColPlo = G * BK, kColO = 2 * G * BK, kColQhi = 2 * G * BK + G * D,
                     kColQlo = kColQhi + D;
  static_assert(D == 32, "column sets are laid out for 32-wide heads");
  static_assert(kColQlo + D <= kCols, "TMEM budget");
  if (threadIdx.x == 0) {
#pragma unroll
    for (int i = 1; i <= G; ++i) mbar_init(&bar[i], 1);
  }
  const CtaSetup cs = cta_setup(&bar[0], &tmem_slot, kCols);  // cs.half = warp / 4 = group (G: issuer)
  const uint32_t tmem = cs.tmem;
  const int nsub = (seg_end - seg_beg + BK - 1) / BK;
  constexpr uint32_t kIdescS = idesc_tf32(BM, BK);
  constexpr uint32_t kIdescO = idesc_tf32(BM, D) | kIdescBMajorMN;

  if (issuer) {
    const OperandDesc dK = operand_desc_k32(smem + oKhi, smem + oKlo);
    const OperandDesc dVt = operand_desc_mn(smem + oVThi, smem + oVTlo);
    auto at_stage = [](const OperandDesc& d, int st) {
      const uint64_t adv = (uint64_t)st * (kStage * sizeof(float) / 16);
      return OperandDesc{d.hi + adv, d.lo + adv};
    };
    auto issue_scores = [&](int s) {
      mma3_ts_k32<D>(tmem + kColS + (uint32_t)(s % G) * BK, tmem + kColQhi, tmem + kColQlo, at_stage(dK, s % NST),
                     kIdescS, false);
    };
    named_bar_sync(1, kFwdThreads);                    // Q rows are in TMEM, the first G stages are written
    fence_after_sync();
    if (elect_one_sync()) {
      for (int s = 0; s < G && s < nsub; ++s) { issue_scores(s); mma_commit(&bar[s]); }
    }
    __syncwarp();
    for (int s = 0; s < nsub; ++s) {
      const int g = s % G;
      named_bar_sync(2 + g, kFwdGroup + 32);           // P_s is in TMEM, stage of step s+G is written
      fence_after_sync();
      if (elect_one_sync()) {
        mma3_ts_mn<BK>(tmem + kColO + (uint32_t)g * D, tmem + kColS + (uint32_t)g * BK, tmem + kColPlo + (uint32_t)g * BK,
                       at_stage(dVt, s % NST), kIdescO, ((s / G) % kPromote) != 0);   // O_g (+)= P_s V_s
        if (s + G < nsub) issue_scores(s + G);
        mma_commit(&bar[g]);
        if (s == nsub - 1) mma_commit(&bar[G]);
      }
      __syncwarp();
    }
  } else {
    const int g = cs.half;                             // my group; it owns the steps s = g (mod G)
    const int32_t row = row0 + cs.row_local;
    const bool row_ok = row < seg_end;
    // scores come out in the log2 domain: Q is pre-multiplied by scale * log2(e).  Groups 0 and 1
    // write the two feature halves of every row.
    if (g < 2)
      row_to_tmem<D>(tmem + cs.lane_base + kColQhi, tmem + cs.lane_base + kColQlo, p.q + h * D, p.ldq, row, row_ok, g,
                     p.scale * kLog2e);
    Tile<BK, D, kFwdGroup, 0> tk, tv;                  // the group's next K and V tiles, in registers
    const int tl = (int)threadIdx.x - g * kFwdGroup;   // thread index inside the group
    auto load = [&](int32_t k_at) {
      tk.load_as(tl, p.k + h * D, p.ldk, k_at, seg_end);
      tv.load_as(tl, p.v + h * D, p.ldv, k_at, seg_end);
    };
    auto stage = [&](int st) {
      float* base = smem + st * kStage;
      tk.store_mn_as(tl, base + oKhi, base + oKlo);
      tv.store_mn_as(tl, base + oVThi, base + oVTlo);
    };
    if (g < nsub) { load(seg_beg + g * BK); stage(g); }
    fence_smem_to_async();
    fence_before_sync();
    named_bar_arrive(1, kFwdThreads);
    if (g + G < nsub) load(seg_beg + (g + G) * BK);

    // The group's output accumulator stays in TMEM (the P V products accumulate into it); it is
    // rescaled only when the running maximum has grown by more than 2^kLazy since the reference
    // m_run was set (then, and only then, the 32 columns are read, scaled and written back).
    // Until that happens P = 2^(s - m_run) may exceed 1 by up to 2^kLazy -- harmless in fp32.
    constexpr float kLazy = 8.f;
    float m_run = -INFINITY, l_run = 0.f;
    const uint32_t t_s = tmem + cs.lane_base + kColS + (uint32_t)g * BK;
    const uint32_t t_plo = tmem + cs.lane_base + kColPlo + (uint32_t)g * BK;
    const uint32_t t_o = tmem + cs.lane_base + kColO + (uint32_t)g * D;
    // fp32 totals of the finished accumulator chains of this group: [D][BM] floats behind the stages
    float* tot = smem + NST * kStage + g * (D * BM) + cs.row_local;
    uint32_t par = 0;
    for (int s = g; s < nsub; s += G) {
      mbar_wait(&bar[g], par);                         // S_s is ready; P V of my previous step is complete
      par ^= 1;
      fence_after_sync();
      const int ls = s / G;                            // index of this step among the group's steps
      if (ls > 0 && ls % kPromote == 0) {              // the issuer restarts the accumulator with this step's P V
#pragma unroll
        for (int c0 = 0; c0 < D; c0 += 16) {
          uint32_t ot[16];
          tmem_ld16(t_o + c0, ot);
          tmem_ld_wait();
#pragma unroll
          for (int d = 0; d < 16; ++d) {
            float* slot = tot + (c0 + d) * BM;
            *slot = ls == kPromote ? __uint_as_float(ot[d]) : *slot + __uint_as_float(ot[d]);
          }
        }
      }
      uint32_t sv[BK];
      tmem_ld32(t_s, sv);
      tmem_ld_wait();
      const int32_t nvalid = seg_end - (seg_beg + s * BK);   // key columns [0, nvalid) exist
      if (nvalid < BK) {
#pragma unroll
        for (int j = 0; j < BK; ++j)
          if (j >= nvalid) sv[j] = 0xff800000u;        // -inf
      }
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int j = 0; j < BK; ++j) mx4[j & 3] = fmaxf(mx4[j & 3], __uint_as_float(sv[j]));
      const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));   // finite: every step has a valid key
      if (s == g) {
        m_run = mx;                                    // first step: the issuer overwrites the accumulator
      } else if (__any_sync(0xffffffffu, mx > m_run + kLazy)) {
        const float m_new = mx > m_run + kLazy ? mx : m_run;
        const float f = fast_exp2(m_run - m_new);      // 1 for the lanes that keep their reference
        m_run = m_new;
        l_run *= f;
#pragma unroll
        for (int c0 = 0; c0 < D; c0 += 16) {
          uint32_t ot[16];
          tmem_ld16(t_o + c0, ot);
          tmem_ld_wait();
#pragma unroll
          for (int d = 0; d < 16; ++d) ot[d] = __float_as_uint(__uint_as_float(ot[d]) * f);
          tmem_st16(t_o + c0, ot);
        }
        if (ls >= kPromote) {                          // the totals follow the reference maximum
#pragma unroll
          for (int d = 0; d < D; ++d) tot[d * BM] *= f;
        }
      }
      float rs4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int c0 = 0; c0 < BK; c0 += 16) {
        uint32_t hi[16], lo[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const float pj = fast_exp2(__uint_as_float(sv[c0 + j]) - m_run);
          rs4[j & 3] += pj;
          float hi_, lo_;
          split_tf32(pj, hi_, lo_);
          hi[j] = __float_as_uint(hi_);
          lo[j] = __float_as_uint(lo_);
        }
        tmem_st16(t_s + c0, hi);
        tmem_st16(t_plo + c0, lo);
      }
      l_run += (rs4[0] + rs4[1]) + (rs4[2] + rs4[3]);
      tmem_st_wait();

      if (s + G < nsub) stage((s + G) % NST);
      fence_smem_to_async();
      fence_before_sync();
      named_bar_arrive(2 + g, kFwdGroup + 32);
      if (s + 2 * G < nsub) load(seg_beg + (s + 2 * G) * BK);
    }
    mbar_wait(&bar[G], 0);                             // every MMA has completed
    fence_after_sync();
    float acc[D];
    if (g < nsub) {
      uint32_t ot[D];
      tmem_ld32(t_o, ot);
      tmem_ld_wait();
      const bool promoted = (nsub - g + G - 1) / G > kPromote;     // the group had more than kPromote steps
#pragma unroll
      for (int d = 0; d < D; ++d) acc[d] = __uint_as_float(ot[d]) + (promoted ? tot[d * BM] : 0.f);
    } else {
#pragma unroll
      for (int d = 0; d < D; ++d) acc[d] = 0.f;
    }
    // merge the groups' states: groups 1.. hand (m, l, acc) over through shared memory (the
    // operand stages are free: all MMAs are complete)
    constexpr int kX = (2 + D) * BM;                   // floats per group: m[128], l[128], acc[D][128]
    if (g > 0) {
      float* x = smem + (g - 1) * kX;
      x[cs.row_local] = m_run;
      x[BM + cs.row_local] = l_run;
#pragma unroll
      for (int d = 0; d < D; ++d) x[(2 + d) * BM + cs.row_local] = acc[d];
    }
    named_bar_sync(2 + G, G * kFwdGroup);
    if (g == 0 && row_ok) {
      float m = m_run;
#pragma unroll
      for (int i = 0; i < G - 1; ++i) m = fmaxf(m, smem[i * kX + cs.row_local]);
      const float f0 = fast_exp2(m_run - m);
      float l_tot = l_run * f0;
#pragma unroll
      for (int d = 0; d < D; ++d) acc[d] *= f0;
#pragma unroll
      for (int i = 0; i < G - 1; ++i) {
        const float* x = smem + i * kX;
        const float fi = fast_exp2(x[cs.row_local] - m);      // 0 for a group without steps (m = -inf)
        l_tot = fmaf(x[BM + cs.row_local], fi, l_tot);
#pragma unroll
        for (int d = 0; d < D; ++d) acc[d] = fmaf(x[(2 + d) * BM + cs.row_local], fi, acc[d]);
      }
      const float inv = 1.f / l_tot;
      float* orow = o + (int64_t)row * ldo + h * D;
#pragma unroll
      for (int d = 0; d < D; d += 4)
        *reinterpret_cast<float4*>(orow + d) =
            make_float4(acc[d] * inv, acc[d + 1] * inv, acc[d + 2] * inv, acc[d + 3] * inv);
      lse[(int64_t)row * p.H + h] = (m + log2f(l_tot)) * kLn2;
    }
  }
  cta_teardown(tmem, kCols);
}

// ================================================================================ backward ===
struct AttnTcBwdArgs {
  const float* dout; int64_t lddo;
  const float* lse; const float* delta;
  int64_t ldg;                       // row stride of dq / dk / dv (floats)
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
  float* drow = dq + (int64_t)row * b.ldg + h * D + cs.half * DH;
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
    float* orow = (is_k ? dk : dv) + (int64_t)krow * b.ldg + h * D + chalf * EC;
#pragma unroll
    for (int j = 0; j < EC; j += 4)
      *reinterpret_cast<float4*>(orow + j) =
          make_float4(__uint_as_float(t[j]) * mul, __uint_as_float(t[j + 1]) * mul,
                      __uint_as_float(t[j + 2]) * mul, __uint_as_float(t[j + 3]) * mul);
  }
  cta_teardown(tmem, kCols);
}

#ifdef HGK_KV_PROFILE
// Phase timing of the pipelined dK/dV kernel (debug builds only: profiles/kv_phase_profile.py).
__device__ unsigned long long g_kv_prof[8];
__device__ __forceinline__ unsigned long long prof_clock() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%clock64;" : "=l"(t) :: "memory");
  return t;
}
// accumulated in registers, flushed once: a global read-modify-write per sample would cost more
// than the phases being measured
#define KV_PROF_DECL unsigned long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0}; unsigned long long pt = prof_clock();
#define KV_PROF(slot) { const unsigned long long now = prof_clock(); prof_acc[slot] += now - pt; pt = now; }
#define KV_PROF_FLUSH if (threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) { for (int i = 0; i < 8; ++i) g_kv_prof[i] += prof_acc[i]; }
#else
#define KV_PROF_DECL
#define KV_PROF(slot)
#define KV_PROF_FLUSH
#endif

// ---- dK, dV: software pipelined, two compute groups, dedicated MMA-issuing warp ---------------
// Same mapping as k_attn_bwd_kv_tc (CTA = 128 keys, TMEM lane = key).  What differs:
//  * the query dimension is walked in steps of 32 with TWO sets of S^T / dP^T / P / dS columns;
//  * compute group g (8 warps, two threads per key row, 16 query columns each) owns the steps
//    s = g (mod 2) and column set g: while one group is waiting (TMEM load latency, the mbarrier,
//    the proxy fence) the other one computes, and the tensor pipe runs the other group's
//    dV / dK updates and S^T / dP^T products meanwhile;
//  * a 17th warp does nothing but issue MMAs.  Issuing the 48 tcgen05.mma of one step costs the
//    issuing thread ~1000 cycles (measured, profiles/kv_phase_profile.py); when that thread was
//    also a compute thread every other warp waited for it at the next barrier.
// Hand-offs: group g -> issuer through named barrier 2+g (bar.arrive, the group does not wait);
// issuer -> group g through mbarrier g (tcgen05.commit).  For step s the issuer queues
// [dV += P_s^T dO_s,  dK += dS_s^T Q_s,  S^T_{s+2} = K Q_{s+2}^T,  dP^T_{s+2} = V dO_{s+2}^T];
// the MMAs of one thread execute in issue order, so S^T_{s+2} overwrites column set g only after
// dV / dK of step s have read P_s / dS_s from it.  Four shared-memory stages: the stage written
// during step s (for step s+2, same group) was last read by the MMAs of step s-2, which the
// mbarrier wait at the top of step s has covered.
constexpr int kKvPipeThreads = kKvThreads + 32;
constexpr int kKvGroup = 256;                          // threads per compute group

// TMA = true: the Q / dO tiles are not loaded by the compute threads but land by cp.async.bulk.tensor straight in
// the operand layout, three of the group's steps ahead (8 stages instead of 4); the raw fp32 tile is the hi
// operand (kind::tf32 ignores the 13 low mantissa bits), the threads read it back only to form lo = x - hi.
// Q is then used unscaled: the resident K rows carry scale * log2(e) instead, and dK is scaled on the way out.
// Rows past the end of the segment are whatever follows in memory (finite); their columns are masked below.
template <int D, bool TMA>
__global__ void __launch_bounds__(kKvPipeThreads, 1)
k_attn_bwd_kv_pipe(const AttnTcArgs p, const AttnTcBwdArgs b, float* __restrict__ dk, float* __restrict__ dv,
                   const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_g) {
  constexpr int BQ = 32;                               // queries per pipeline step
  constexpr int QC = BQ / 2;                           // query columns per thread
  constexpr int NST = TMA ? 8 : 4;                     // shared-memory stages
  extern __shared__ __align__(1024) float smem_raw[];
  // the MN-major operand tiles need 1024-byte alignment; the launchers request 1024 bytes more than the layout
  float* const smem = smem_raw + (((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u) >> 2);
  __shared__ __align__(8) uint64_t bar[3];             // [g]: column set g is ready; [2]: everything done
  __shared__ __align__(8) uint64_t full[NST];          // TMA: the tiles of stage st have landed
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(16) float lse2_s[NST][BQ], delta_s[NST][BQ];
  int32_t seg_beg, seg_end, key0;
  if (!locate_tile128(p.cu, p.S, blockIdx.x, &seg_beg, &seg_end, &key0)) return;
  const int h = blockIdx.y;
  const int warp_u = warp_id_uniform();
  const bool issuer = warp_u == kKvThreads / 32;       // warp 16

  constexpr int kTile = BQ * D;                        // floats per operand tile
  constexpr int kStage = 4 * kTile;                    // floats per stage
  // one tile pair per matrix: read K-major as B = Q / dO (S^T, dP^T) and MN-major as B = Q^T / dO^T (dK, dV)
  constexpr int oQhi = 0, oQlo = kTile;                // Q rows
  constexpr int oGhi = 2 * kTile, oGlo = 3 * kTile;    // dO rows

  constexpr uint32_t kCols = 512;
  // one (dV, dK) accumulator pair per compute group: columns kColAcc + g * 2D (dV) and + D (dK); a group's
  // chain of updates is ordered by its own hand-offs, so the group can move it into the fp32 totals
  // every kPromote steps without stalling the other group
  constexpr uint32_t kColST = 0, kColDPT = 64, kColPlo = 128, kColDSlo = 192, kColAcc = 256;
  constexpr uint32_t kColKhi = 256 + 4 * D, kColKlo = 256 + 5 * D, kColVhi = 256 + 6 * D, kColVlo = 256 + 7 * D;
  static_assert(256 + 8 * D <= 512, "TMEM budget");
  if (threadIdx.x == 0) {                              // bar[0]: cta_setup, which also publishes the others
    mbar_init(&bar[1], 1); mbar_init(&bar[2], 1);
    if (TMA) {
#pragma unroll
      for (int i = 0; i < NST; ++i) mbar_init(&full[i], 1);
    }
  }
  const CtaSetup cs = cta_setup(&bar[0], &tmem_slot, kCols);  // cs.half = warp / 4
  const uint32_t tmem = cs.tmem;
  const int nsub = (seg_end - seg_beg + BQ - 1) / BQ;
  constexpr uint32_t kIdescS = idesc_tf32(BM, BQ);
  constexpr uint32_t kIdescD = idesc_tf32(BM, D) | kIdescBMajorMN;

  if (issuer) {
    // ------------------------------------------------------------------ MMA-issuing warp ---
    const OperandDesc dQ = operand_desc_k32(smem + oQhi, smem + oQlo), dG = operand_desc_k32(smem + oGhi, smem + oGlo);
    const OperandDesc dQt = operand_desc_mn(smem + oQhi, smem + oQlo), dGt = operand_desc_mn(smem + oGhi, smem + oGlo);
    auto at_stage = [](const OperandDesc& d, int st) {   // start-address field is in 16-byte units
      const uint64_t adv = (uint64_t)st * (kStage * sizeof(float) / 16);
      return OperandDesc{d.hi + adv, d.lo + adv};
    };
    auto issue_scores = [&](int s) {                   // S^T, dP^T of step s into column set s & 1
      const uint32_t set = (uint32_t)(s & 1) * BQ;
      mma3_ts_k32<D>(tmem + kColST + set, tmem + kColKhi, tmem + kColKlo, at_stage(dQ, s & (NST - 1)), kIdescS, false);
      mma3_ts_k32<D>(tmem + kColDPT + set, tmem + kColVhi, tmem + kColVlo, at_stage(dG, s & (NST - 1)), kIdescS, false);
    };
    auto issue_updates = [&](int s) {                  // dV += P^T dO, dK += dS^T Q of step s
      const uint32_t set = (uint32_t)(s & 1) * BQ;
      const bool accumulate = ((s >> 1) % kPromote) != 0;
      const uint32_t acc = tmem + kColAcc + (uint32_t)(s & 1) * 2 * D;
      mma3_ts_mn<BQ>(acc, tmem + kColST + set, tmem + kColPlo + set, at_stage(dGt, s & (NST - 1)), kIdescD, accumulate);
      mma3_ts_mn<BQ>(acc + D, tmem + kColDPT + set, tmem + kColDSlo + set, at_stage(dQt, s & (NST - 1)), kIdescD, accumulate);
    };
    named_bar_sync(1, kKvPipeThreads);                 // K / V rows are in TMEM, stages 0 and 1 are written
    fence_after_sync();
    if (elect_one_sync()) {
      issue_scores(0);
      mma_commit(&bar[0]);
      if (nsub > 1) { issue_scores(1); mma_commit(&bar[1]); }
    }
    __syncwarp();
    for (int s = 0; s < nsub; ++s) {
      const int g = s & 1;
      named_bar_sync(2 + g, kKvGroup + 32);            // P_s, dS_s are in TMEM, stage of step s+2 is written
      fence_after_sync();
      if (elect_one_sync()) {
        issue_updates(s);
        if (s + 2 < nsub) issue_scores(s + 2);
        mma_commit(&bar[g]);
        if (s == nsub - 1) mma_commit(&bar[2]);
      }
      __syncwarp();
    }
  } else {
    // ------------------------------------------------------------- compute groups 0 and 1 ---
    const int g = warp_u >> 3;                         // my group; it owns the steps s = g (mod 2)
    const int part = cs.half;                          // warp / 4: 0..3 (K / V rows to TMEM, epilogue)
    const int ch = part & 1;                           // my 16-column half of the step's 32 queries
    {
      const int64_t krow_ld = key0 + cs.row_local;
      const bool kok = krow_ld < seg_end;
      if (part < 2)
        row_to_tmem<D>(tmem + cs.lane_base + kColKhi, tmem + cs.lane_base + kColKlo, p.k + h * D, p.ldk, krow_ld, kok, part,
                       TMA ? p.scale * kLog2e : 1.f);
      else
        row_to_tmem<D>(tmem + cs.lane_base + kColVhi, tmem + cs.lane_base + kColVlo, p.v + h * D, p.ldv, krow_ld, kok, part - 2, 1.f);
    }
    // the group's next Q and dO tiles, prefetched in registers (one 16-byte chunk of each per thread)
    Tile<BQ, D, kKvGroup, 0> tq0, tg0;
    Tile<BQ, D, kKvGroup, kKvGroup> tq1, tg1;
    float lse_r = 0.f, delta_r = 0.f;                  // next step's row statistics (first warp of the group)
    const int tl = (int)threadIdx.x - g * kKvGroup;    // thread index inside the group
    const float qmul = p.scale * kLog2e;
    // TMA: one thread of the group requests the two tiles of step s (stage s % NST)
    auto request = [&](int s) {
      const int st = s & (NST - 1);
      float* base = smem + st * kStage;
      mbar_expect_tx(&full[st], 2u * kTile * (uint32_t)sizeof(float));
      tma_load_2d(base + oQhi, &tm_q, &full[st], h * D, seg_beg + s * BQ);
      tma_load_2d(base + oGhi, &tm_g, &full[st], h * D, seg_beg + s * BQ);
    };
    auto load = [&](int32_t q_at) {
      if (!TMA) {
        if (g == 0) { tq0.load(p.q + h * D, p.ldq, q_at, seg_end); tg0.load(b.dout + h * D, b.lddo, q_at, seg_end); }
        else        { tq1.load(p.q + h * D, p.ldq, q_at, seg_end); tg1.load(b.dout + h * D, b.lddo, q_at, seg_end); }
      }
      if (!TMA && tl < BQ) {
        const int32_t qrow = q_at + tl;
        lse_r = 0.f; delta_r = 0.f;
        if (qrow < seg_end) {
          asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(lse_r) : "l"(b.lse + (int64_t)qrow * p.H + h));
          asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(delta_r) : "l"(b.delta + (int64_t)qrow * p.H + h));
        }
      }
    };
    auto stage = [&](int s) {
      const int st = s & (NST - 1);
      float* base = smem + st * kStage;
      if (TMA) {
        mbar_wait(&full[st], (uint32_t)(s / NST) & 1u);     // requested one of my steps ago (prologue: just now)
        int r, c;
        Tile<BQ, D, kKvGroup, 0>::coord_of(tl, 0, r, c);    // one 16-byte chunk of each tile per thread
        static_assert(Tile<BQ, D, kKvGroup, 0>::PER == 1, "one chunk per thread");
        const int off = Tile<BQ, D, kKvGroup, 0>::mn_off(r, c);
        const float4 q4 = *reinterpret_cast<const float4*>(base + oQhi + off);
        const float4 g4 = *reinterpret_cast<const float4*>(base + oGhi + off);
        float4 hq, lq, hg, lg;
        split_tf32(q4.x, hq.x, lq.x); split_tf32(q4.y, hq.y, lq.y); split_tf32(q4.z, hq.z, lq.z); split_tf32(q4.w, hq.w, lq.w);
        split_tf32(g4.x, hg.x, lg.x); split_tf32(g4.y, hg.y, lg.y); split_tf32(g4.z, hg.z, lg.z); split_tf32(g4.w, hg.w, lg.w);
        *reinterpret_cast<float4*>(base + oQlo + off) = lq;
        *reinterpret_cast<float4*>(base + oGlo + off) = lg;
      } else if (g == 0) {
        tq0.scale(qmul);
        tq0.store_mn_as(tl, base + oQhi, base + oQlo);
        tg0.store_mn_as(tl, base + oGhi, base + oGlo);
      } else {
        tq1.scale(qmul);
        tq1.store_mn_as(tl, base + oQhi, base + oQlo);
        tg1.store_mn_as(tl, base + oGhi, base + oGlo);
      }
      if (tl < BQ) {
        if (TMA) {                                     // staged in the group's idle time: loaded here, no prefetch
          const int32_t qrow = seg_beg + s * BQ + tl;
          lse_r = 0.f; delta_r = 0.f;
          if (qrow < seg_end) {
            lse_r = __ldg(b.lse + (int64_t)qrow * p.H + h);
            delta_r = __ldg(b.delta + (int64_t)qrow * p.H + h);
          }
        }
        lse2_s[st][tl] = lse_r * kLog2e;
        delta_s[st][tl] = delta_r;
      }
    };

    if (TMA && tl == 0) {                              // my first three steps' tiles are on their way
#pragma unroll
      for (int j = 0; j < 3; ++j)
        if (g + 2 * j < nsub) request(g + 2 * j);
    }
    if (g < nsub) { load(seg_beg + g * BQ); stage(g); }
    fence_smem_to_async();
    fence_before_sync();
    named_bar_arrive(1, kKvPipeThreads);
    if (TMA) {
      // TMA: the lo tiles are formed TWO of my steps ahead, after the hand-off to the issuer -- in the time the
      // group would otherwise spend waiting for its MMAs -- so staging is off the step's critical path
      if (g + 2 < nsub) stage(g + 2);
    } else if (g + 2 < nsub) {
      load(seg_beg + (g + 2) * BQ);                    // after the fence: it would wait for these loads
    }

    // fp32 totals of the group's finished accumulator chains: [2 groups][dV | dK][D][BM] floats behind the
    // stages; this thread takes dV (ch == 0) or dK (ch == 1) of its group, all D columns of its row
    float* tot = smem + NST * kStage + (g * 2 + ch) * (D * BM) + cs.row_local;
    uint32_t par = 0;
    KV_PROF_DECL
    for (int s = g; s < nsub; s += 2) {
      KV_PROF(7)
      mbar_wait(&bar[g], par);
      par ^= 1;
      fence_after_sync();
      KV_PROF(0)
      // the MMAs of my previous step are complete: its stage, (s - 2) % NST == (s + 6) % NST, is free again
      if (TMA && tl == 0 && s + 6 < nsub) request(s + 6);
      const int ls = s >> 1;                           // index of this step among the group's steps
      if (ls > 0 && ls % kPromote == 0) {              // the issuer restarts the group's dV / dK with this step
#pragma unroll
        for (int c0 = 0; c0 < D; c0 += 16) {
          uint32_t va[16];
          tmem_ld16(tmem + cs.lane_base + kColAcc + (uint32_t)g * 2 * D + (uint32_t)ch * D + c0, va);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            float* slot = tot + (c0 + j) * BM;
            *slot = ls == kPromote ? __uint_as_float(va[j]) : *slot + __uint_as_float(va[j]);
          }
        }
      }

      const int st = s & (NST - 1);
      const int col0 = ch * QC;
      const uint32_t set = (uint32_t)g * BQ;
      const int32_t nvalid = seg_end - (seg_beg + s * BQ) - col0;   // my query columns [0, nvalid) exist
      uint32_t sv[QC], dp[QC], plo[QC], dslo[QC];
      tmem_ld16(tmem + cs.lane_base + kColST + set + col0, sv);
      tmem_ld16(tmem + cs.lane_base + kColDPT + set + col0, dp);
      tmem_ld_wait();
      KV_PROF(1)
      if (nvalid < QC) {                               // the segment ends inside my columns: queries past the end
#pragma unroll                                         // get score -inf, i.e. P = 0 and dS = 0
        for (int j = 0; j < QC; ++j)
          if (j >= nvalid) sv[j] = 0xff800000u;
      }
      // P and dS (hi, lo) of my 16 query columns; their lse / delta come from shared memory as 16-byte loads
#pragma unroll
      for (int j4 = 0; j4 < QC; j4 += 4) {
        const float4 l4 = *reinterpret_cast<const float4*>(&lse2_s[st][col0 + j4]);
        const float4 d4 = *reinterpret_cast<const float4*>(&delta_s[st][col0 + j4]);
        const float lq[4] = {l4.x, l4.y, l4.z, l4.w}, dq_[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          const int j = j4 + jj;
          const float pj = fast_exp2(__uint_as_float(sv[j]) - lq[jj]);
          const float ds = pj * (__uint_as_float(dp[j]) - dq_[jj]);
          float hi_, lo_;
          split_tf32(pj, hi_, lo_);
          sv[j] = __float_as_uint(hi_); plo[j] = __float_as_uint(lo_);
          split_tf32(ds, hi_, lo_);
          dp[j] = __float_as_uint(hi_); dslo[j] = __float_as_uint(lo_);
        }
      }
      KV_PROF(2)
      tmem_st16(tmem + cs.lane_base + kColST + set + col0, sv);
      tmem_st16(tmem + cs.lane_base + kColPlo + set + col0, plo);
      tmem_st16(tmem + cs.lane_base + kColDPT + set + col0, dp);
      tmem_st16(tmem + cs.lane_base + kColDSlo + set + col0, dslo);
      tmem_st_wait();
      KV_PROF(3)

      if (!TMA && s + 2 < nsub) stage(s + 2);          // registers -> stage of my next step
      KV_PROF(4)
      fence_smem_to_async();                           // (TMA: also publishes the lo tiles written after the last arrive)
      fence_before_sync();
      named_bar_arrive(2 + g, kKvGroup + 32);
      KV_PROF(5)
      if (TMA) {
        if (s + 4 < nsub) stage(s + 4);                // lo tiles of the step after next, while my MMAs run
      } else if (s + 4 < nsub) {
        load(seg_beg + (s + 4) * BQ);                  // global -> registers for the step after next
      }
      KV_PROF(6)
    }
    mbar_wait(&bar[2], 0);                             // committed after the last step's MMAs
    fence_after_sync();
    KV_PROF_FLUSH

    named_bar_sync(9, kKvThreads);                     // both groups' totals are in shared memory
    static_assert(D == 32, "epilogue column split");
    constexpr int EC = D / 2;                          // columns per thread
    const int32_t krow = key0 + cs.row_local;
    const bool ok = krow < seg_end;
    const bool is_k = part >= 2;
    const int chalf = part & 1;
    // dv / dk = sum over the groups of (last accumulator chain in TMEM + fp32 totals of the earlier ones)
    float accv[EC];
#pragma unroll
    for (int j = 0; j < EC; ++j) accv[j] = 0.f;
#pragma unroll
    for (int gg = 0; gg < 2; ++gg) {
      const int steps = gg < nsub ? (nsub - gg + 1) / 2 : 0;       // steps of group gg (CTA uniform)
      if (steps == 0) continue;                        // its accumulators were never written
      uint32_t t[16];
      tmem_ld16(tmem + cs.lane_base + kColAcc + (uint32_t)gg * 2 * D + (is_k ? D : 0) + chalf * EC, t);
      tmem_ld_wait();
      const float* tg = smem + NST * kStage + (gg * 2 + (is_k ? 1 : 0)) * (D * BM) + (chalf * EC) * BM + cs.row_local;
#pragma unroll
      for (int j = 0; j < EC; ++j) accv[j] += __uint_as_float(t[j]) + (steps > kPromote ? tg[j * BM] : 0.f);
    }
    if (ok) {
      // Q was staged as q * scale * log2(e) (TMA: unscaled, the scores' scale sits on the K rows)
      const float mul = is_k ? (TMA ? p.scale : kLn2) : 1.f;
      float* orow = (is_k ? dk : dv) + (int64_t)krow * b.ldg + h * D + chalf * EC;
#pragma unroll
      for (int j = 0; j < EC; j += 4)
        *reinterpret_cast<float4*>(orow + j) = make_float4(accv[j] * mul, accv[j + 1] * mul, accv[j + 2] * mul, accv[j + 3] * mul);
    }
  }
  cta_teardown(tmem, kCols);
}

// ---- dQ: the same pipeline, CTA = 128 queries (TMEM lane = query), 32-key steps ----------------
// Q and dO rows (hi, lo) are resident in TMEM as the A operands of S = Q K^T and dP = dO V^T;
// dS (hi, lo) overwrites the step's S / dP columns and is the A operand of dQ += dS K.
// TMA = true: K / V tiles by cp.async.bulk.tensor, as in k_attn_bwd_kv_pipe.
template <int D, bool TMA>
__global__ void __launch_bounds__(kKvPipeThreads, 1)
k_attn_bwd_q_pipe(const AttnTcArgs p, const AttnTcBwdArgs b, float* __restrict__ dq,
                  const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v) {
  constexpr int BK = 32;                               // keys per pipeline step
  constexpr int KC = BK / 2;                           // key columns per thread
  constexpr int NST = TMA ? 8 : 4;                     // shared-memory stages
  extern __shared__ __align__(1024) float smem_raw[];
  // the MN-major operand tiles need 1024-byte alignment; the launchers request 1024 bytes more than the layout
  float* const smem = smem_raw + (((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u) >> 2);
  __shared__ __align__(8) uint64_t bar[3];             // [g]: column set g is ready; [2]: everything done
  __shared__ __align__(8) uint64_t full[NST];          // TMA: the tiles of stage st have landed
  __shared__ uint32_t tmem_slot;
  int32_t seg_beg, seg_end, row0;
  if (!locate_tile128(p.cu, p.S, blockIdx.x, &seg_beg, &seg_end, &row0)) return;
  const int h = blockIdx.y;
  const int warp_u = warp_id_uniform();
  const bool issuer = warp_u == kKvThreads / 32;       // warp 16

  constexpr int kTile = BK * D;                        // floats per operand tile
  constexpr int kStage = 4 * kTile;                    // floats per stage
  constexpr int oKhi = 0, oKlo = kTile;                // K rows: K-major as B = K (S), MN-major as B = K^T (dQ)
  constexpr int oVhi = 2 * kTile, oVlo = 3 * kTile;    // V rows: B = V (dP)

  constexpr uint32_t kCols = 512;
  // one dQ accumulator per compute group (columns kColDQ + g * D): a group's chain of updates is then
  // ordered by its own hand-offs, and the group can move it into the fp32 totals every kPromote steps
  constexpr uint32_t kColS = 0, kColDP = 64, kColDQ = 128;
  constexpr uint32_t kColQhi = 128 + 2 * D, kColQlo = 128 + 3 * D, kColGhi = 128 + 4 * D, kColGlo = 128 + 5 * D;
  static_assert(128 + 6 * D <= 512, "TMEM budget");
  if (threadIdx.x == 0) {
    mbar_init(&bar[1], 1); mbar_init(&bar[2], 1);
    if (TMA) {
#pragma unroll
      for (int k = 0; k < NST; ++k) mbar_init(&full[k], 1);
    }
  }
  const CtaSetup cs = cta_setup(&bar[0], &tmem_slot, kCols);  // cs.half = warp / 4
  const uint32_t tmem = cs.tmem;
  const int nsub = (seg_end - seg_beg + BK - 1) / BK;
  constexpr uint32_t kIdescS = idesc_tf32(BM, BK);
  constexpr uint32_t kIdescQ = idesc_tf32(BM, D) | kIdescBMajorMN;

  if (issuer) {
    const OperandDesc dK = operand_desc_k32(smem + oKhi, smem + oKlo), dV = operand_desc_k32(smem + oVhi, smem + oVlo);
    const OperandDesc dKt = operand_desc_mn(smem + oKhi, smem + oKlo);
    auto at_stage = [](const OperandDesc& d, int st) {
      const uint64_t adv = (uint64_t)st * (kStage * sizeof(float) / 16);
      return OperandDesc{d.hi + adv, d.lo + adv};
    };
    auto issue_scores = [&](int s) {                   // S, dP of step s into column set s & 1
      const uint32_t set = (uint32_t)(s & 1) * BK;
      mma3_ts_k32<D>(tmem + kColS + set, tmem + kColQhi, tmem + kColQlo, at_stage(dK, s & (NST - 1)), kIdescS, false);
      mma3_ts_k32<D>(tmem + kColDP + set, tmem + kColGhi, tmem + kColGlo, at_stage(dV, s & (NST - 1)), kIdescS, false);
    };
    named_bar_sync(1, kKvPipeThreads);                 // Q / dO rows are in TMEM, stages 0 and 1 are written
    fence_after_sync();
    if (elect_one_sync()) {
      issue_scores(0);
      mma_commit(&bar[0]);
      if (nsub > 1) { issue_scores(1); mma_commit(&bar[1]); }
    }
    __syncwarp();
    for (int s = 0; s < nsub; ++s) {
      const int g = s & 1;
      named_bar_sync(2 + g, kKvGroup + 32);            // dS_s is in TMEM, stage of step s+2 is written
      fence_after_sync();
      if (elect_one_sync()) {
        const uint32_t set = (uint32_t)g * BK;
        mma3_ts_mn<BK>(tmem + kColDQ + (uint32_t)g * D, tmem + kColS + set, tmem + kColDP + set,
                       at_stage(dKt, s & (NST - 1)), kIdescQ, ((s >> 1) % kPromote) != 0);
        if (s + 2 < nsub) issue_scores(s + 2);
        mma_commit(&bar[g]);
        if (s == nsub - 1) mma_commit(&bar[2]);
      }
      __syncwarp();
    }
  } else {
    const int g = warp_u >> 3;                         // my group; it owns the steps s = g (mod 2)
    const int part = cs.half;                          // warp / 4: 0..3
    const int ch = part & 1;                           // my 16-column half of the step's 32 keys
    const int32_t row = row0 + cs.row_local;
    const bool row_ok = row < seg_end;
    if (part < 2)
      row_to_tmem<D>(tmem + cs.lane_base + kColQhi, tmem + cs.lane_base + kColQlo, p.q + h * D, p.ldq, row, row_ok, part,
                     p.scale * kLog2e);
    else
      row_to_tmem<D>(tmem + cs.lane_base + kColGhi, tmem + cs.lane_base + kColGlo, b.dout + h * D, b.lddo, row, row_ok,
                     part - 2, 1.f);
    const float lse2 = row_ok ? b.lse[(int64_t)row * p.H + h] * kLog2e : 0.f;
    const float delta = row_ok ? b.delta[(int64_t)row * p.H + h] : 0.f;

    Tile<BK, D, kKvGroup, 0> tk0, tv0;                 // the group's next K and V tiles, in registers
    Tile<BK, D, kKvGroup, kKvGroup> tk1, tv1;
    const int tl = (int)threadIdx.x - g * kKvGroup;    // thread index inside the group
    // TMA: one thread of the group requests the two tiles of step s (stage s % NST)
    auto request = [&](int s) {
      const int st = s & (NST - 1);
      float* base = smem + st * kStage;
      mbar_expect_tx(&full[st], 2u * kTile * (uint32_t)sizeof(float));
      tma_load_2d(base + oKhi, &tm_k, &full[st], h * D, seg_beg + s * BK);
      tma_load_2d(base + oVhi, &tm_v, &full[st], h * D, seg_beg + s * BK);
    };
    auto load = [&](int32_t k_at) {
      if (TMA) return;
      if (g == 0) { tk0.load(p.k + h * D, p.ldk, k_at, seg_end); tv0.load(p.v + h * D, p.ldv, k_at, seg_end); }
      else        { tk1.load(p.k + h * D, p.ldk, k_at, seg_end); tv1.load(p.v + h * D, p.ldv, k_at, seg_end); }
    };
    auto stage = [&](int s) {
      const int st = s & (NST - 1);
      float* base = smem + st * kStage;
      if (TMA) {
        mbar_wait(&full[st], (uint32_t)(s / NST) & 1u);     // requested one of my steps ago (prologue: just now)
        int r, c;
        Tile<BK, D, kKvGroup, 0>::coord_of(tl, 0, r, c);    // one 16-byte chunk of each tile per thread
        static_assert(Tile<BK, D, kKvGroup, 0>::PER == 1, "one chunk per thread");
        const int off = Tile<BK, D, kKvGroup, 0>::mn_off(r, c);
        const float4 k4 = *reinterpret_cast<const float4*>(base + oKhi + off);
        const float4 v4 = *reinterpret_cast<const float4*>(base + oVhi + off);
        float4 hk, lk, hv, lv;
        split_tf32(k4.x, hk.x, lk.x); split_tf32(k4.y, hk.y, lk.y); split_tf32(k4.z, hk.z, lk.z); split_tf32(k4.w, hk.w, lk.w);
        split_tf32(v4.x, hv.x, lv.x); split_tf32(v4.y, hv.y, lv.y); split_tf32(v4.z, hv.z, lv.z); split_tf32(v4.w, hv.w, lv.w);
        *reinterpret_cast<float4*>(base + oKlo + off) = lk;
        *reinterpret_cast<float4*>(base + oVlo + off) = lv;
      } else if (g == 0) {
        tk0.store_mn(base + oKhi, base + oKlo);
        tv0.store_mn(base + oVhi, base + oVlo);
      } else {
        tk1.store_mn(base + oKhi, base + oKlo);
        tv1.store_mn(base + oVhi, base + oVlo);
      }
    };
    if (TMA && tl == 0) {                              // my first three steps' tiles are on their way
#pragma unroll
      for (int j = 0; j < 3; ++j)
        if (g + 2 * j < nsub) request(g + 2 * j);
    }
    if (g < nsub) { load(seg_beg + g * BK); stage(g); }
    fence_smem_to_async();
    fence_before_sync();
    named_bar_arrive(1, kKvPipeThreads);
    if (TMA) {                                         // lo tiles two of my steps ahead, in the idle time after a hand-off
      if (g + 2 < nsub) stage(g + 2);
    } else if (g + 2 < nsub) {
      load(seg_beg + (g + 2) * BK);
    }

    // fp32 totals of the group's finished accumulator chains: [2][D][BM] floats behind the stages;
    // this thread's 16 of the group's 32 columns
    float* tot = smem + NST * kStage + g * (D * BM) + (ch * KC) * BM + cs.row_local;
    uint32_t par = 0;
    for (int s = g; s < nsub; s += 2) {
      mbar_wait(&bar[g], par);
      par ^= 1;
      fence_after_sync();
      // the MMAs of my previous step are complete: its stage, (s - 2) % NST == (s + 6) % NST, is free again
      if (TMA && tl == 0 && s + 6 < nsub) request(s + 6);
      const int ls = s >> 1;                           // index of this step among the group's steps
      if (ls > 0 && ls % kPromote == 0) {              // the issuer restarts dQ_g with this step's update
        uint32_t qa[KC];
        tmem_ld16(tmem + cs.lane_base + kColDQ + (uint32_t)g * D + ch * KC, qa);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < KC; ++j)
          tot[j * BM] = ls == kPromote ? __uint_as_float(qa[j]) : tot[j * BM] + __uint_as_float(qa[j]);
      }

      const int col0 = ch * KC;
      const uint32_t set = (uint32_t)g * BK;
      const int32_t nvalid = seg_end - (seg_beg + s * BK) - col0;   // my key columns [0, nvalid) exist
      uint32_t sv[KC], dv[KC];
      tmem_ld16(tmem + cs.lane_base + kColS + set + col0, sv);
      tmem_ld16(tmem + cs.lane_base + kColDP + set + col0, dv);
      tmem_ld_wait();
      if (nvalid < KC) {                               // the segment ends inside my columns: keys past the end
#pragma unroll                                         // get score -inf, i.e. P = 0 and dS = 0
        for (int j = 0; j < KC; ++j)
          if (j >= nvalid) sv[j] = 0xff800000u;
      }
#pragma unroll
      for (int j = 0; j < KC; ++j) {
        const float pj = fast_exp2(__uint_as_float(sv[j]) - lse2);
        const float ds = pj * (__uint_as_float(dv[j]) - delta);
        float hi_, lo_;
        split_tf32(ds, hi_, lo_);
        sv[j] = __float_as_uint(hi_);
        dv[j] = __float_as_uint(lo_);
      }
      tmem_st16(tmem + cs.lane_base + kColS + set + col0, sv);
      tmem_st16(tmem + cs.lane_base + kColDP + set + col0, dv);
      tmem_st_wait();

      if (!TMA && s + 2 < nsub) stage(s + 2);
      fence_smem_to_async();
      fence_before_sync();
      named_bar_arrive(2 + g, kKvGroup + 32);
      if (TMA) {
        if (s + 4 < nsub) stage(s + 4);                // lo tiles of the step after next, while my MMAs run
      } else if (s + 4 < nsub) {
        load(seg_beg + (s + 4) * BK);
      }
    }
    mbar_wait(&bar[2], 0);
    fence_after_sync();

    named_bar_sync(9, kKvThreads);                     // both groups' totals are in shared memory
    // four threads per row: D/4 columns each.  tcgen05.ld is warp-collective: everyone loads.
    // dq = sum over the groups of (last accumulator chain in TMEM + fp32 totals of the earlier ones)
    constexpr int EC = D / 4;
    static_assert(EC == 8, "epilogue column split");
    float acc[EC];
#pragma unroll
    for (int j = 0; j < EC; ++j) acc[j] = 0.f;
#pragma unroll
    for (int gg = 0; gg < 2; ++gg) {
      const int steps = gg < nsub ? (nsub - gg + 1) / 2 : 0;       // steps of group gg (CTA uniform)
      if (steps == 0) continue;                        // its accumulator was never written
      uint32_t t[EC];
      tmem_ld8(tmem + cs.lane_base + kColDQ + (uint32_t)gg * D + part * EC, t);
      tmem_ld_wait();
      const float* tg = smem + NST * kStage + gg * (D * BM) + (part * EC) * BM + cs.row_local;
#pragma unroll
      for (int j = 0; j < EC; ++j) acc[j] += __uint_as_float(t[j]) + (steps > kPromote ? tg[j * BM] : 0.f);
    }
    if (row_ok) {
      float* drow = dq + (int64_t)row * b.ldg + h * D + part * EC;
      *reinterpret_cast<float4*>(drow) = make_float4(acc[0] * p.scale, acc[1] * p.scale, acc[2] * p.scale, acc[3] * p.scale);
      *reinterpret_cast<float4*>(drow + 4) = make_float4(acc[4] * p.scale, acc[5] * p.scale, acc[6] * p.scale, acc[7] * p.scale);
    }
  }
  cta_teardown(tmem, kCols);
}

// ---------------------------------------------------------------------------------- launch ---
// cudaFuncSetAttribute is per function and costs host microseconds on a host-bound step: done once per
// kernel (a handful of kernels: linear scan of a small table, guarded for concurrent host threads)
template <typename K>
int opt_in_smem(K kernel, size_t bytes, const char* what) {
  static std::mutex mu;
  static const void* done[16];
  static int n_done = 0;
  const void* key = reinterpret_cast<const void*>(kernel);
  {
    std::lock_guard<std::mutex> lock(mu);
    for (int i = 0; i < n_done; ++i)
      if (done[i] == key) return HGK_OK;
  }
  cudaError_t err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (err != cudaSuccess)
    return fail(HGK_ERR_CUDA, "%s: smem opt-in (%zu B) failed: %s", what, bytes, cudaGetErrorString(err));
  std::lock_guard<std::mutex> lock(mu);
  if (n_done < 16) done[n_done++] = key;
  return HGK_OK;
}

template <int D>
int launch_fwd_pipe(const AttnTcArgs& a, float* o, int64_t ldo, float* lse, int64_t N, cudaStream_t st) {
  // one CTA per SM (512 TMEM columns): 8 operand stages of 16 KB + the groups' fp32 totals [G][D][128]
  size_t smem = sizeof(float) * (2 * kFwdGroups * 4 * 32 * D + kFwdGroups * D * BM) + 1024;   // + alignment slack
  if (smem < 120 * 1024) smem = 120 * 1024;
  int rc = opt_in_smem(k_attn_fwd_pipe<D>, smem, "attn_fwd_pipe");
  if (rc) return rc;
  dim3 grid((unsigned)(ceil_div(N, BM) + a.S), (unsigned)a.H);
  k_attn_fwd_pipe<D><<<grid, kFwdThreads, smem, st>>>(a, o, ldo, lse);
  return check_launch("attn_varlen_fwd(tcgen05, pipelined)");
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

// one side stream + fork / join events per device, created on first use
struct SideStream {
  cudaStream_t stream;
  cudaEvent_t fork, join;
};
static SideStream* side_stream() {
  static std::mutex mu;
  static SideStream table[64];
  static bool ready[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  std::lock_guard<std::mutex> lock(mu);
  if (!ready[dev]) {
    SideStream s{};
    if (cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
    if (cudaEventCreateWithFlags(&s.fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&s.join, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    table[dev] = s;
    ready[dev] = true;
  }
  return &table[dev];
}

template <int D>
int launch_bwd(const AttnTcArgs& a, const AttnTcBwdArgs& b, float* dq, float* dk, float* dv, int64_t N,
               cudaStream_t st) {
  dim3 grid((unsigned)(ceil_div(N, BM) + a.S), (unsigned)a.H);
  // 512 TMEM columns: one CTA per SM; ask for more than half of the shared memory to enforce it
  size_t smem_kv = sizeof(float) * (2 * 8 * BN * D + 4 * D * BM);      // 4 stages + the groups' fp32 totals
  if (smem_kv < 120 * 1024) smem_kv = 120 * 1024;
  // HGK_ATTN_BWD_SERIAL=1 selects the unpipelined backward kernels (A/B runs); read once
  static const bool serial = [] { const char* e = getenv("HGK_ATTN_BWD_SERIAL"); return e && e[0] == '1'; }();
  int rc;
  if (serial) {
    rc = opt_in_smem(k_attn_bwd_kv_tc<D>, smem_kv, "attn_bwd_kv_tc");
    if (rc) return rc;
    k_attn_bwd_kv_tc<D><<<grid, kKvThreads, smem_kv, st>>>(a, b, dk, dv);
    size_t smem_q = sizeof(float) * (2 * BM * D + 6 * BN * D);
    if (smem_q < 80 * 1024) smem_q = 80 * 1024;
    rc = opt_in_smem(k_attn_bwd_q_tc<D>, smem_q, "attn_bwd_q_tc");
    if (rc) return rc;
    k_attn_bwd_q_tc<D><<<grid, kThreads, smem_q, st>>>(a, b, dq);
    return check_launch("attn_varlen_bwd(tcgen05)");
  }
  // pipelined kernels: stages of 4 tiles [32 x D] (hi / lo of two matrices) + the groups' fp32 totals + alignment
  // slack; 4 stages filled from registers, or 8 filled by TMA (HGK_ATTN_BWD_TMA=0: the register path, A/B runs)
  static const bool want_tma = [] { const char* e = getenv("HGK_ATTN_BWD_TMA"); return !(e && e[0] == '0'); }();
  CUtensorMap tm_q, tm_k, tm_v, tm_g;
  const int64_t cols = (int64_t)a.H * D;
  const bool use_tma = want_tma && tma::rows_map_32(&tm_q, a.q, N, cols, a.ldq, 32) && tma::rows_map_32(&tm_k, a.k, N, cols, a.ldk, 32) &&
                       tma::rows_map_32(&tm_v, a.v, N, cols, a.ldv, 32) && tma::rows_map_32(&tm_g, b.dout, N, cols, b.lddo, 32);
  const int nst = use_tma ? 8 : 4;
  smem_kv = sizeof(float) * (nst * 4 * 32 * D + 4 * D * BM) + 1024;
  if (smem_kv < 120 * 1024) smem_kv = 120 * 1024;                      // one CTA per SM either way (512 TMEM columns)
  size_t smem_q = sizeof(float) * (nst * 4 * 32 * D + 2 * D * BM) + 1024;
  if (smem_q < 120 * 1024) smem_q = 120 * 1024;
  if (use_tma) {
    rc = opt_in_smem(k_attn_bwd_kv_pipe<D, true>, smem_kv, "attn_bwd_kv_pipe(tma)");
    if (rc) return rc;
    rc = opt_in_smem(k_attn_bwd_q_pipe<D, true>, smem_q, "attn_bwd_q_pipe(tma)");
  } else {
    rc = opt_in_smem(k_attn_bwd_kv_pipe<D, false>, smem_kv, "attn_bwd_kv_pipe");
    if (rc) return rc;
    rc = opt_in_smem(k_attn_bwd_q_pipe<D, false>, smem_q, "attn_bwd_q_pipe");
  }
  if (rc) return rc;
  // The dK/dV and dQ kernels are independent and each runs one CTA per SM: 256 + 256 CTAs on 148 SMs are
  // two partly empty second waves when the kernels run back to back.  With the dQ kernel on a second
  // stream its CTAs start on the SMs that the dK/dV kernel's last wave leaves idle.  Fork / join with
  // events around the pair; the side stream and the events are created once per device (the first call
  // on a device should therefore not be inside a stream capture).  HGK_ATTN_BWD_OVERLAP=0: one stream.
  static const bool overlap = [] { const char* e = getenv("HGK_ATTN_BWD_OVERLAP"); return !(e && e[0] == '0'); }();
  SideStream* ss = overlap ? side_stream() : nullptr;
  const bool forked = ss != nullptr && cudaEventRecord(ss->fork, st) == cudaSuccess &&
                      cudaStreamWaitEvent(ss->stream, ss->fork, 0) == cudaSuccess;
  if (use_tma) {
    k_attn_bwd_kv_pipe<D, true><<<grid, kKvPipeThreads, smem_kv, st>>>(a, b, dk, dv, tm_q, tm_g);
    k_attn_bwd_q_pipe<D, true><<<grid, kKvPipeThreads, smem_q, forked ? ss->stream : st>>>(a, b, dq, tm_k, tm_v);
  } else {
    memset(&tm_q, 0, sizeof(tm_q));                    // unused by the register-staged kernels
    k_attn_bwd_kv_pipe<D, false><<<grid, kKvPipeThreads, smem_kv, st>>>(a, b, dk, dv, tm_q, tm_q);
    k_attn_bwd_q_pipe<D, false><<<grid, kKvPipeThreads, smem_q, forked ? ss->stream : st>>>(a, b, dq, tm_q, tm_q);
  }
  if (forked) {
    if (cudaEventRecord(ss->join, ss->stream) != cudaSuccess || cudaStreamWaitEvent(st, ss->join, 0) != cudaSuccess)
      return fail(HGK_ERR_CUDA, "attn_varlen_bwd: joining the side stream failed: %s", cudaGetErrorString(cudaGetLastError()));
  }
  return check_launch("attn_varlen_bwd(tcgen05)");
}

}  // namespace tc

#ifdef HGK_KV_PROFILE
extern "C" int hgk_debug_kv_profile(unsigned long long* out8, int reset) {
  unsigned long long zero[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  cudaDeviceSynchronize();
  if (out8) cudaMemcpyFromSymbol(out8, tc::g_kv_prof, sizeof(zero));
  if (reset) cudaMemcpyToSymbol(tc::g_kv_prof, zero, sizeof(zero));
  return 0;
}
#endif

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
  // HGK_ATTN_FWD_SERIAL=1 selects the unpipelined forward kernel (A/B runs); read once
  static const bool serial = [] { const char* e = getenv("HGK_ATTN_FWD_SERIAL"); return e && e[0] == '1'; }();
  if (serial) return tc::launch_fwd<32>(a, o, ldo, lse, N, st);
  return tc::launch_fwd_pipe<32>(a, o, ldo, lse, N, st);
}

// Called by hgk_attn_varlen_bwd after delta = rowsum(dO * O) has been enqueued.  Same return
// convention as attn_fwd_tc.  dq, dk, dv are [N, H, D] with row stride ldg.
int attn_bwd_tc(float* dq, float* dk, float* dv, int64_t ldg, const float* delta, const float* dout, int64_t lddo,
                const float* lse, const float* q, int64_t ldq, const float* k, int64_t ldk,
                const float* v, int64_t ldv, const int32_t* cu, int64_t S, int64_t N, int32_t H,
                int32_t D, float scale, cudaStream_t st) {
  if (D != 32) return 1;    // the 512-thread dK/dV kernel is laid out for 32-wide heads
  if (!tc_aligned(q, ldq) || !tc_aligned(k, ldk) || !tc_aligned(v, ldv) || !tc_aligned(dout, lddo) ||
      !tc_aligned(dq, ldg) || !tc_aligned(dk, ldg) || !tc_aligned(dv, ldg))
    return 1;
  AttnTcArgs a{q, ldq, k, ldk, v, ldv, cu, (int32_t)S, H, scale};
  tc::AttnTcBwdArgs b{dout, lddo, lse, delta, ldg};
  return tc::launch_bwd<32>(a, b, dq, dk, dv, N, st);
}

}  // namespace hgk
