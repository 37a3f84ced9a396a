// GPS global attention on the 5th-generation tensor cores (tcgen05.mma kind::tf32, accumulators
// in TMEM), fp32-accurate through 3xTF32 operand splitting.  Same contract as the SIMT kernels in
// attn.cu (reference: hydragnn/globalAtt/gps.py:126-133 -- to_dense_batch + nn.MultiheadAttention
// core + unpad), selected by hgk_attn_varlen_* when the head dim is a multiple of 16.
//
// Forward, one CTA (128 threads) per (128-query tile, head):
//   S  = Q K^T      tcgen05.mma SS: A = Q tile [128 x D] (smem, resident), B = K tile [64 x D]
//   P  = online softmax of S: thread t owns TMEM lane t (= query row t), reads its 64 scores with
//        tcgen05.ld, keeps running max / sum in registers (exp2 domain), writes P back to TMEM
//        split in hi / lo parts (tcgen05.st)
//   Ot = P V        tcgen05.mma TS: A = P (TMEM, lane = query, column = key), B = V^T tile
//        [D x 64] (smem, K-major = keys contiguous); the [128 x D] product is read back and
//        folded into the fp32 running output  acc = acc * corr + Ot  in registers.
// Two CTAs share an SM (256 TMEM columns each) and overlap each other's load / MMA / softmax
// phases; inside a CTA the phases are serialised by one mbarrier and __syncthreads.
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

constexpr int BM = 128;          // queries per CTA (= TMEM lanes)
constexpr int BN = 64;           // keys per iteration
constexpr int kThreads = 128;
constexpr uint32_t kTmemCols = 256;
constexpr uint32_t kColS = 0;    // S, overwritten in place by P_hi     [128 x 64]
constexpr uint32_t kColPlo = 64; // P_lo                               [128 x 64]
constexpr uint32_t kColO = 128;  // per-iteration P V product           [128 x D]

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

// ---- tile staging: global -> registers (prefetchable) -> shared memory --------------------------
// A [ROWS x D] fp32 tile is ROWS * D/4 float4 chunks; thread t owns chunks t, t+128, ...
// chunk idx -> row r = idx / (D/4), 16-byte column c = idx % (D/4).
template <int ROWS, int D>
struct Tile {
  static constexpr int CH = D / 4;
  static constexpr int PER = ROWS * CH / kThreads;
  static_assert(ROWS * CH % kThreads == 0, "tile must divide evenly over the CTA");
  float4 v[PER];

  __device__ __forceinline__ void load(const float* src, int64_t ld, int32_t row0, int32_t row_end) {
#pragma unroll
    for (int it = 0; it < PER; ++it) {
      const int idx = threadIdx.x + it * kThreads;
      const int r = idx / CH, c = idx - r * CH;
      v[it] = make_float4(0.f, 0.f, 0.f, 0.f);          // rows past the segment end are zero
      if (row0 + r < row_end) v[it] = __ldg(reinterpret_cast<const float4*>(src + (int64_t)(row0 + r) * ld + 4 * c));
    }
  }
  // hi / lo K-major tiles with MN index = row, K index = feature
  __device__ __forceinline__ void store_rows(float* hi, float* lo, float mul) const {
    constexpr int SBO = CH * 32;
#pragma unroll
    for (int it = 0; it < PER; ++it) {
      const int idx = threadIdx.x + it * kThreads;
      const int r = idx / CH, c = idx - r * CH;
      float4 h, l;
      split_tf32(v[it].x * mul, h.x, l.x); split_tf32(v[it].y * mul, h.y, l.y);
      split_tf32(v[it].z * mul, h.z, l.z); split_tf32(v[it].w * mul, h.w, l.w);
      const int off = core_off(r, c, SBO);
      *reinterpret_cast<float4*>(hi + off) = h;
      *reinterpret_cast<float4*>(lo + off) = l;
    }
  }
  // hi / lo K-major tiles of the TRANSPOSE: MN index = feature (D rows), K index = row (ROWS columns).
  // The 4 lanes that hold the same 16-byte column of 4 consecutive rows exchange their values with
  // 4 shuffles (a 4x4 register transpose), after which every lane writes one 16-byte chunk.
  __device__ __forceinline__ void store_transposed(float* hi, float* lo, float mul) const {
    constexpr int SBO = (ROWS / 4) * 32;
    const int lane = threadIdx.x & 31;
    const int rr = (lane / CH) & 3;                    // my row inside its group of 4
    const int base = lane - rr * CH;                   // lane of row 0 of the group, same column
#pragma unroll
    for (int it = 0; it < PER; ++it) {
      const int idx = threadIdx.x + it * kThreads;
      const int r = idx / CH, c = idx - r * CH;
      const float x[4] = {v[it].x * mul, v[it].y * mul, v[it].z * mul, v[it].w * mul};
      float w[4];
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        const int give = (rr - s) & 3;                 // the partner reading me at step s wants this feature
        const float send = give == 0 ? x[0] : (give == 1 ? x[1] : (give == 2 ? x[2] : x[3]));
        const int from = (rr + s) & 3;
        const float got = __shfl_sync(0xffffffffu, send, base + from * CH);
        if (from == 0) w[0] = got; else if (from == 1) w[1] = got; else if (from == 2) w[2] = got; else w[3] = got;
      }
      // w[j] = value of feature 4c + rr at row (r & ~3) + j
      float4 h, l;
      split_tf32(w[0], h.x, l.x); split_tf32(w[1], h.y, l.y);
      split_tf32(w[2], h.z, l.z); split_tf32(w[3], h.w, l.w);
      const int off = core_off(4 * c + rr, r >> 2, SBO);
      *reinterpret_cast<float4*>(hi + off) = h;
      *reinterpret_cast<float4*>(lo + off) = l;
    }
  }
};

template <int D>
__global__ void __launch_bounds__(kThreads)
k_attn_fwd_tc(const AttnTcArgs p, float* __restrict__ o, int64_t ldo, float* __restrict__ lse) {
  extern __shared__ __align__(128) float smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;

  int32_t seg_beg, seg_end, row0;
  if (!locate_tile128(p.cu, p.S, blockIdx.x, &seg_beg, &seg_end, &row0)) return;
  const int h = blockIdx.y;
  const int tid = threadIdx.x, warp = tid >> 5;

  float* Qhi = smem;
  float* Qlo = Qhi + BM * D;
  float* Khi = Qlo + BM * D;
  float* Klo = Khi + BN * D;
  float* Vhi = Klo + BN * D;     // V^T tile [D x BN]
  float* Vlo = Vhi + BN * D;

  if (warp == 0) tmem_alloc(&tmem_slot, kTmemCols);
  if (tid == 0) mbar_init(&bar, 1);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = tmem_slot;
  const uint32_t lane_base = static_cast<uint32_t>(warp * 32) << 16;

  constexpr float kLog2e = 1.4426950408889634f;
  {
    Tile<BM, D> tq;
    tq.load(p.q + h * D, p.ldq, row0, seg_end);
    tq.store_rows(Qhi, Qlo, p.scale * kLog2e);
  }
  Tile<BN, D> tk, tv;                                  // next key / value tile, prefetched in registers
  tk.load(p.k + h * D, p.ldk, seg_beg, seg_end);
  tv.load(p.v + h * D, p.ldv, seg_beg, seg_end);

  constexpr uint32_t kLbo = 128;                       // bytes between K chunks
  constexpr uint32_t kSboQK = (D / 4) * 128;           // bytes between 8-row groups (Q, K tiles)
  constexpr uint32_t kSboV = (BN / 4) * 128;           // ... of the V^T tile
  constexpr uint32_t kIdescS = idesc_tf32(BM, BN);
  constexpr uint32_t kIdescO = idesc_tf32(BM, D);

  float acc[D];
#pragma unroll
  for (int d = 0; d < D; ++d) acc[d] = 0.f;
  float m_run = -INFINITY, l_run = 0.f;
  uint32_t parity = 0;

  for (int32_t k0 = seg_beg; k0 < seg_end; k0 += BN) {
    tk.store_rows(Khi, Klo, 1.f);
    tv.store_transposed(Vhi, Vlo, 1.f);
    fence_smem_to_async();
    fence_before_sync();
    __syncthreads();
    if (tid == 0) {
      fence_after_sync();
      const uint32_t qh = smem_u32(Qhi), ql = smem_u32(Qlo), kh = smem_u32(Khi), kl = smem_u32(Klo);
#pragma unroll
      for (int j = 0; j < D / 8; ++j) {
        const uint32_t adv = j * 2 * kLbo;             // one MMA consumes two 16-byte K chunks
        mma_tf32_ss(tmem + kColS, smem_desc(ql + adv, kLbo, kSboQK), smem_desc(kh + adv, kLbo, kSboQK), kIdescS, j > 0);
        mma_tf32_ss(tmem + kColS, smem_desc(qh + adv, kLbo, kSboQK), smem_desc(kl + adv, kLbo, kSboQK), kIdescS, 1);
        mma_tf32_ss(tmem + kColS, smem_desc(qh + adv, kLbo, kSboQK), smem_desc(kh + adv, kLbo, kSboQK), kIdescS, 1);
      }
      mma_commit(&bar);
    }
    if (k0 + BN < seg_end) {                           // global loads of the next tile fly under the MMAs
      tk.load(p.k + h * D, p.ldk, k0 + BN, seg_end);
      tv.load(p.v + h * D, p.ldv, k0 + BN, seg_end);
    }
    mbar_wait(&bar, parity);
    parity ^= 1;
    fence_after_sync();

    // ---- online softmax on this thread's row (scores are already in the log2 domain) ----
    uint32_t s0[32], s1[32];
    tmem_ld32(tmem + lane_base + kColS, s0);
    tmem_ld32(tmem + lane_base + kColS + 32, s1);
    tmem_ld_wait();
    const int32_t nvalid = seg_end - k0;               // keys [0, nvalid) of this tile exist
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      if (j >= nvalid) s0[j] = 0xff800000u;            // -inf
      if (j + 32 >= nvalid) s1[j] = 0xff800000u;
      mx = fmaxf(mx, fmaxf(__uint_as_float(s0[j]), __uint_as_float(s1[j])));
    }
    const float m_new = fmaxf(m_run, mx);
    const float corr = exp2f(m_run - m_new);
    float rs = 0.f;
    uint32_t lo0[32], lo1[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const float p0 = exp2f(__uint_as_float(s0[j]) - m_new);
      const float p1 = exp2f(__uint_as_float(s1[j]) - m_new);
      rs += p0 + p1;
      float h0, l0, h1, l1;
      split_tf32(p0, h0, l0);
      split_tf32(p1, h1, l1);
      s0[j] = __float_as_uint(h0); lo0[j] = __float_as_uint(l0);
      s1[j] = __float_as_uint(h1); lo1[j] = __float_as_uint(l1);
    }
    tmem_st32(tmem + lane_base + kColS, s0);
    tmem_st32(tmem + lane_base + kColS + 32, s1);
    tmem_st32(tmem + lane_base + kColPlo, lo0);
    tmem_st32(tmem + lane_base + kColPlo + 32, lo1);
    tmem_st_wait();
    fence_before_sync();
    __syncthreads();
    if (tid == 0) {
      fence_after_sync();
      const uint32_t vh = smem_u32(Vhi), vl = smem_u32(Vlo);
#pragma unroll
      for (int j = 0; j < BN / 8; ++j) {
        const uint32_t adv = j * 2 * kLbo;
        const uint32_t pc = j * 8;                     // 8 keys = 8 TMEM columns per MMA
        mma_tf32_ts(tmem + kColO, tmem + kColPlo + pc, smem_desc(vh + adv, kLbo, kSboV), kIdescO, j > 0);
        mma_tf32_ts(tmem + kColO, tmem + kColS + pc, smem_desc(vl + adv, kLbo, kSboV), kIdescO, 1);
        mma_tf32_ts(tmem + kColO, tmem + kColS + pc, smem_desc(vh + adv, kLbo, kSboV), kIdescO, 1);
      }
      mma_commit(&bar);
    }
    mbar_wait(&bar, parity);
    parity ^= 1;
    fence_after_sync();

    l_run = l_run * corr + rs;
    m_run = m_new;
#pragma unroll
    for (int c0 = 0; c0 < D; c0 += 16) {
      uint32_t ot[16];
      tmem_ld16(tmem + lane_base + kColO + c0, ot);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 16; ++j) acc[c0 + j] = fmaf(acc[c0 + j], corr, __uint_as_float(ot[j]));
    }
  }

  const int32_t row = row0 + tid;
  if (row < seg_end) {
    const float inv = 1.f / l_run;
    float* orow = o + (int64_t)row * ldo + h * D;
#pragma unroll
    for (int d = 0; d < D; d += 4)
      *reinterpret_cast<float4*>(orow + d) =
          make_float4(acc[d] * inv, acc[d + 1] * inv, acc[d + 2] * inv, acc[d + 3] * inv);
    constexpr float kLn2 = 0.6931471805599453f;
    lse[(int64_t)row * p.H + h] = (m_run + log2f(l_run)) * kLn2;
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, kTmemCols);
}

template <int D>
int launch_fwd(const AttnTcArgs& a, float* o, int64_t ldo, float* lse, int64_t N, cudaStream_t st) {
  // 2 CTAs per SM: 256 TMEM columns each; the shared-memory request (>= 1/3 of the SM) enforces it
  size_t smem = sizeof(float) * (2 * BM * D + 4 * BN * D);
  if (smem < 80 * 1024) smem = 80 * 1024;
  cudaError_t err = cudaFuncSetAttribute(k_attn_fwd_tc<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (err != cudaSuccess) return fail(HGK_ERR_CUDA, "attn_fwd_tc: smem opt-in failed: %s", cudaGetErrorString(err));
  dim3 grid((unsigned)(ceil_div(N, BM) + a.S), (unsigned)a.H);
  k_attn_fwd_tc<D><<<grid, kThreads, smem, st>>>(a, o, ldo, lse);
  return check_launch("attn_varlen_fwd(tcgen05)");
}


// =============================================================================== backward ===
// Two deterministic kernels (no float atomics), both recomputing the probabilities from lse:
//   k_attn_bwd_q_tc  : CTA owns 128 queries, loops over 64-key tiles          -> dQ
//   k_attn_bwd_kv_tc : CTA owns 128 keys (TMEM lane = key), loops over 64-query tiles -> dK, dV
// With s = (q.k) scale log2(e)  (Q is staged pre-multiplied), lse2 = lse log2(e):
//   P = 2^(s - lse2),  dP = dO V^T,  dS = P (dP - delta),  delta = rowsum(dO O)
//   dQ = scale dS K,   dK = scale dS^T Q,   dV = P^T dO
struct AttnTcBwdArgs {
  const float* dout; int64_t lddo;
  const float* lse; const float* delta;
};

template <int D>
__global__ void __launch_bounds__(kThreads)
k_attn_bwd_q_tc(const AttnTcArgs p, const AttnTcBwdArgs b, float* __restrict__ dq) {
  extern __shared__ __align__(128) float smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  int32_t seg_beg, seg_end, row0;
  if (!locate_tile128(p.cu, p.S, blockIdx.x, &seg_beg, &seg_end, &row0)) return;
  const int h = blockIdx.y;
  const int tid = threadIdx.x, warp = tid >> 5;

  float* Qhi = smem;             float* Qlo = Qhi + BM * D;
  float* Ghi = Qlo + BM * D;     float* Glo = Ghi + BM * D;      // dO tile
  float* Khi = Glo + BM * D;     float* Klo = Khi + BN * D;
  float* Vhi = Klo + BN * D;     float* Vlo = Vhi + BN * D;
  float* KThi = Vlo + BN * D;    float* KTlo = KThi + BN * D;    // K^T tile [D x BN]

  constexpr uint32_t kColS = 0, kColDP = 64, kColDQ = 128;       // dS_hi -> S region, dS_lo -> dP region
  if (warp == 0) tmem_alloc(&tmem_slot, 256);
  if (tid == 0) mbar_init(&bar, 1);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = tmem_slot;
  const uint32_t lane_base = static_cast<uint32_t>(warp * 32) << 16;

  constexpr float kLog2e = 1.4426950408889634f;
  {
    Tile<BM, D> t;
    t.load(p.q + h * D, p.ldq, row0, seg_end);
    t.store_rows(Qhi, Qlo, p.scale * kLog2e);
    t.load(b.dout + h * D, b.lddo, row0, seg_end);
    t.store_rows(Ghi, Glo, 1.f);
  }
  Tile<BN, D> tk, tv;
  tk.load(p.k + h * D, p.ldk, seg_beg, seg_end);
  tv.load(p.v + h * D, p.ldv, seg_beg, seg_end);
  const int32_t row = row0 + tid;
  const bool row_ok = row < seg_end;
  const float lse2 = row_ok ? b.lse[(int64_t)row * p.H + h] * kLog2e : 0.f;
  const float delta = row_ok ? b.delta[(int64_t)row * p.H + h] : 0.f;

  constexpr uint32_t kLbo = 128;
  constexpr uint32_t kSboD = (D / 4) * 128;            // tiles whose K dim is the feature dim
  constexpr uint32_t kSboT = (BN / 4) * 128;           // transposed tiles (K dim = 64 keys)
  constexpr uint32_t kIdescS = idesc_tf32(BM, BN);
  constexpr uint32_t kIdescQ = idesc_tf32(BM, D);
  uint32_t parity = 0;
  bool first = true;

  for (int32_t k0 = seg_beg; k0 < seg_end; k0 += BN) {
    tk.store_rows(Khi, Klo, 1.f);
    tk.store_transposed(KThi, KTlo, 1.f);
    tv.store_rows(Vhi, Vlo, 1.f);
    fence_smem_to_async();
    fence_before_sync();
    __syncthreads();
    if (tid == 0) {
      fence_after_sync();
      const uint32_t qh = smem_u32(Qhi), ql = smem_u32(Qlo), gh = smem_u32(Ghi), gl = smem_u32(Glo);
      const uint32_t kh = smem_u32(Khi), kl = smem_u32(Klo), vh = smem_u32(Vhi), vl = smem_u32(Vlo);
#pragma unroll
      for (int j = 0; j < D / 8; ++j) {
        const uint32_t adv = j * 2 * kLbo;
        mma_tf32_ss(tmem + kColS, smem_desc(ql + adv, kLbo, kSboD), smem_desc(kh + adv, kLbo, kSboD), kIdescS, j > 0);
        mma_tf32_ss(tmem + kColS, smem_desc(qh + adv, kLbo, kSboD), smem_desc(kl + adv, kLbo, kSboD), kIdescS, 1);
        mma_tf32_ss(tmem + kColS, smem_desc(qh + adv, kLbo, kSboD), smem_desc(kh + adv, kLbo, kSboD), kIdescS, 1);
      }
#pragma unroll
      for (int j = 0; j < D / 8; ++j) {
        const uint32_t adv = j * 2 * kLbo;
        mma_tf32_ss(tmem + kColDP, smem_desc(gl + adv, kLbo, kSboD), smem_desc(vh + adv, kLbo, kSboD), kIdescS, j > 0);
        mma_tf32_ss(tmem + kColDP, smem_desc(gh + adv, kLbo, kSboD), smem_desc(vl + adv, kLbo, kSboD), kIdescS, 1);
        mma_tf32_ss(tmem + kColDP, smem_desc(gh + adv, kLbo, kSboD), smem_desc(vh + adv, kLbo, kSboD), kIdescS, 1);
      }
      mma_commit(&bar);
    }
    if (k0 + BN < seg_end) {
      tk.load(p.k + h * D, p.ldk, k0 + BN, seg_end);
      tv.load(p.v + h * D, p.ldv, k0 + BN, seg_end);
    }
    mbar_wait(&bar, parity);
    parity ^= 1;
    fence_after_sync();

    const int32_t nvalid = seg_end - k0;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      uint32_t sv[32], dv[32];
      tmem_ld32(tmem + lane_base + kColS + 32 * half, sv);
      tmem_ld32(tmem + lane_base + kColDP + 32 * half, dv);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const float pj = exp2f(__uint_as_float(sv[j]) - lse2);
        float ds = pj * (__uint_as_float(dv[j]) - delta);
        if (j + 32 * half >= nvalid) ds = 0.f;
        float hi, lo;
        split_tf32(ds, hi, lo);
        sv[j] = __float_as_uint(hi);
        dv[j] = __float_as_uint(lo);
      }
      tmem_st32(tmem + lane_base + kColS + 32 * half, sv);
      tmem_st32(tmem + lane_base + kColDP + 32 * half, dv);
    }
    tmem_st_wait();
    fence_before_sync();
    __syncthreads();
    if (tid == 0) {
      fence_after_sync();
      const uint32_t th = smem_u32(KThi), tl = smem_u32(KTlo);
#pragma unroll
      for (int j = 0; j < BN / 8; ++j) {
        const uint32_t adv = j * 2 * kLbo;
        const uint32_t pc = j * 8;
        mma_tf32_ts(tmem + kColDQ, tmem + kColDP + pc, smem_desc(th + adv, kLbo, kSboT), kIdescQ, !(first && j == 0));
        mma_tf32_ts(tmem + kColDQ, tmem + kColS + pc, smem_desc(tl + adv, kLbo, kSboT), kIdescQ, 1);
        mma_tf32_ts(tmem + kColDQ, tmem + kColS + pc, smem_desc(th + adv, kLbo, kSboT), kIdescQ, 1);
      }
      mma_commit(&bar);
    }
    first = false;
    mbar_wait(&bar, parity);
    parity ^= 1;
    fence_after_sync();
  }

  // tcgen05.ld is warp-collective (.sync.aligned): every thread loads, valid rows store
  float* drow = dq + ((int64_t)row * p.H + h) * D;
#pragma unroll
  for (int c0 = 0; c0 < D; c0 += 16) {
    uint32_t t[16];
    tmem_ld16(tmem + lane_base + kColDQ + c0, t);
    tmem_ld_wait();
    if (row_ok) {
#pragma unroll
      for (int j = 0; j < 16; j += 4)
        *reinterpret_cast<float4*>(drow + c0 + j) =
            make_float4(__uint_as_float(t[j]) * p.scale, __uint_as_float(t[j + 1]) * p.scale,
                        __uint_as_float(t[j + 2]) * p.scale, __uint_as_float(t[j + 3]) * p.scale);
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

template <int D>
__global__ void __launch_bounds__(kThreads)
k_attn_bwd_kv_tc(const AttnTcArgs p, const AttnTcBwdArgs b, float* __restrict__ dk, float* __restrict__ dv) {
  constexpr int BQ = 64;                                // queries per iteration
  extern __shared__ __align__(128) float smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  __shared__ float lse2_s[BQ], delta_s[BQ];
  int32_t seg_beg, seg_end, key0;
  if (!locate_tile128(p.cu, p.S, blockIdx.x, &seg_beg, &seg_end, &key0)) return;
  const int h = blockIdx.y;
  const int tid = threadIdx.x, warp = tid >> 5;

  float* Khi = smem;             float* Klo = Khi + BM * D;
  float* Vhi = Klo + BM * D;     float* Vlo = Vhi + BM * D;
  float* Qhi = Vlo + BM * D;     float* Qlo = Qhi + BQ * D;      // rows = queries (B of S^T)
  float* Ghi = Qlo + BQ * D;     float* Glo = Ghi + BQ * D;      // dO rows        (B of dP^T)
  float* QThi = Glo + BQ * D;    float* QTlo = QThi + BQ * D;    // Q^T [D x BQ]   (B of dK)
  float* GThi = QTlo + BQ * D;   float* GTlo = GThi + BQ * D;    // dO^T [D x BQ]  (B of dV)

  constexpr uint32_t kColST = 0, kColDPT = 64, kColPlo = 128, kColDSlo = 192, kColDV = 256, kColDK = 256 + D;
  constexpr uint32_t kCols = 512;
  if (warp == 0) tmem_alloc(&tmem_slot, kCols);
  if (tid == 0) mbar_init(&bar, 1);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = tmem_slot;
  const uint32_t lane_base = static_cast<uint32_t>(warp * 32) << 16;

  constexpr float kLog2e = 1.4426950408889634f;
  {
    Tile<BM, D> t;
    t.load(p.k + h * D, p.ldk, key0, seg_end);
    t.store_rows(Khi, Klo, 1.f);
    t.load(p.v + h * D, p.ldv, key0, seg_end);
    t.store_rows(Vhi, Vlo, 1.f);
  }
  Tile<BQ, D> tq, tg;                                  // next query / dO tile, prefetched in registers
  tq.load(p.q + h * D, p.ldq, seg_beg, seg_end);
  tg.load(b.dout + h * D, b.lddo, seg_beg, seg_end);

  constexpr uint32_t kLbo = 128;
  constexpr uint32_t kSboD = (D / 4) * 128;
  constexpr uint32_t kSboT = (BQ / 4) * 128;
  constexpr uint32_t kIdescS = idesc_tf32(BM, BQ);
  constexpr uint32_t kIdescD = idesc_tf32(BM, D);
  uint32_t parity = 0;
  bool first = true;

  for (int32_t q0 = seg_beg; q0 < seg_end; q0 += BQ) {
    const float qmul = p.scale * kLog2e;
    tq.store_rows(Qhi, Qlo, qmul);
    tq.store_transposed(QThi, QTlo, qmul);
    tg.store_rows(Ghi, Glo, 1.f);
    tg.store_transposed(GThi, GTlo, 1.f);
    if (tid < BQ) {
      const int32_t qrow = q0 + tid;
      const bool ok = qrow < seg_end;
      lse2_s[tid] = ok ? b.lse[(int64_t)qrow * p.H + h] * kLog2e : 0.f;
      delta_s[tid] = ok ? b.delta[(int64_t)qrow * p.H + h] : 0.f;
    }
    fence_smem_to_async();
    fence_before_sync();
    __syncthreads();
    if (tid == 0) {
      fence_after_sync();
      const uint32_t kh = smem_u32(Khi), kl = smem_u32(Klo), vh = smem_u32(Vhi), vl = smem_u32(Vlo);
      const uint32_t qh = smem_u32(Qhi), ql = smem_u32(Qlo), gh = smem_u32(Ghi), gl = smem_u32(Glo);
#pragma unroll
      for (int j = 0; j < D / 8; ++j) {
        const uint32_t adv = j * 2 * kLbo;
        mma_tf32_ss(tmem + kColST, smem_desc(kl + adv, kLbo, kSboD), smem_desc(qh + adv, kLbo, kSboD), kIdescS, j > 0);
        mma_tf32_ss(tmem + kColST, smem_desc(kh + adv, kLbo, kSboD), smem_desc(ql + adv, kLbo, kSboD), kIdescS, 1);
        mma_tf32_ss(tmem + kColST, smem_desc(kh + adv, kLbo, kSboD), smem_desc(qh + adv, kLbo, kSboD), kIdescS, 1);
      }
#pragma unroll
      for (int j = 0; j < D / 8; ++j) {
        const uint32_t adv = j * 2 * kLbo;
        mma_tf32_ss(tmem + kColDPT, smem_desc(vl + adv, kLbo, kSboD), smem_desc(gh + adv, kLbo, kSboD), kIdescS, j > 0);
        mma_tf32_ss(tmem + kColDPT, smem_desc(vh + adv, kLbo, kSboD), smem_desc(gl + adv, kLbo, kSboD), kIdescS, 1);
        mma_tf32_ss(tmem + kColDPT, smem_desc(vh + adv, kLbo, kSboD), smem_desc(gh + adv, kLbo, kSboD), kIdescS, 1);
      }
      mma_commit(&bar);
    }
    if (q0 + BQ < seg_end) {
      tq.load(p.q + h * D, p.ldq, q0 + BQ, seg_end);
      tg.load(b.dout + h * D, b.lddo, q0 + BQ, seg_end);
    }
    mbar_wait(&bar, parity);
    parity ^= 1;
    fence_after_sync();

    const int32_t nvalid = seg_end - q0;                 // query columns [0, nvalid) exist
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      uint32_t sv[32], dp[32], plo[32], dslo[32];
      tmem_ld32(tmem + lane_base + kColST + 32 * half, sv);
      tmem_ld32(tmem + lane_base + kColDPT + 32 * half, dp);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const int col = j + 32 * half;
        float pj = exp2f(__uint_as_float(sv[j]) - lse2_s[col]);
        if (col >= nvalid) pj = 0.f;
        const float ds = pj * (__uint_as_float(dp[j]) - delta_s[col]);
        float hi, lo;
        split_tf32(pj, hi, lo);
        sv[j] = __float_as_uint(hi); plo[j] = __float_as_uint(lo);
        split_tf32(ds, hi, lo);
        dp[j] = __float_as_uint(hi); dslo[j] = __float_as_uint(lo);
      }
      tmem_st32(tmem + lane_base + kColST + 32 * half, sv);
      tmem_st32(tmem + lane_base + kColPlo + 32 * half, plo);
      tmem_st32(tmem + lane_base + kColDPT + 32 * half, dp);
      tmem_st32(tmem + lane_base + kColDSlo + 32 * half, dslo);
    }
    tmem_st_wait();
    fence_before_sync();
    __syncthreads();
    if (tid == 0) {
      fence_after_sync();
      const uint32_t qth = smem_u32(QThi), qtl = smem_u32(QTlo), gth = smem_u32(GThi), gtl = smem_u32(GTlo);
#pragma unroll
      for (int j = 0; j < BQ / 8; ++j) {
        const uint32_t adv = j * 2 * kLbo;
        const uint32_t pc = j * 8;
        const uint32_t acc0 = !(first && j == 0);
        mma_tf32_ts(tmem + kColDV, tmem + kColPlo + pc, smem_desc(gth + adv, kLbo, kSboT), kIdescD, acc0);
        mma_tf32_ts(tmem + kColDV, tmem + kColST + pc, smem_desc(gtl + adv, kLbo, kSboT), kIdescD, 1);
        mma_tf32_ts(tmem + kColDV, tmem + kColST + pc, smem_desc(gth + adv, kLbo, kSboT), kIdescD, 1);
        mma_tf32_ts(tmem + kColDK, tmem + kColDSlo + pc, smem_desc(qth + adv, kLbo, kSboT), kIdescD, acc0);
        mma_tf32_ts(tmem + kColDK, tmem + kColDPT + pc, smem_desc(qtl + adv, kLbo, kSboT), kIdescD, 1);
        mma_tf32_ts(tmem + kColDK, tmem + kColDPT + pc, smem_desc(qth + adv, kLbo, kSboT), kIdescD, 1);
      }
      mma_commit(&bar);
    }
    first = false;
    mbar_wait(&bar, parity);
    parity ^= 1;
    fence_after_sync();
  }

  const int32_t krow = key0 + tid;
  const bool ok = krow < seg_end;
  constexpr float kLn2 = 0.6931471805599453f;
#pragma unroll
  for (int c0 = 0; c0 < D; c0 += 16) {
    uint32_t tv[16], tk[16];
    tmem_ld16(tmem + lane_base + kColDV + c0, tv);
    tmem_ld16(tmem + lane_base + kColDK + c0, tk);
    tmem_ld_wait();
    if (ok) {
      float* vrow = dv + ((int64_t)krow * p.H + h) * D + c0;
      float* krow_p = dk + ((int64_t)krow * p.H + h) * D + c0;
#pragma unroll
      for (int j = 0; j < 16; j += 4) {
        *reinterpret_cast<float4*>(vrow + j) = make_float4(__uint_as_float(tv[j]), __uint_as_float(tv[j + 1]),
                                                           __uint_as_float(tv[j + 2]), __uint_as_float(tv[j + 3]));
        *reinterpret_cast<float4*>(krow_p + j) =
            make_float4(__uint_as_float(tk[j]) * kLn2, __uint_as_float(tk[j + 1]) * kLn2,
                        __uint_as_float(tk[j + 2]) * kLn2, __uint_as_float(tk[j + 3]) * kLn2);
      }
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, kCols);
}

template <typename K>
int opt_in_smem(K kernel, size_t bytes, const char* what) {
  cudaError_t err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (err != cudaSuccess) return fail(HGK_ERR_CUDA, "%s: smem opt-in (%zu B) failed: %s", what, bytes, cudaGetErrorString(err));
  return HGK_OK;
}

template <int D>
int launch_bwd(const AttnTcArgs& a, const AttnTcBwdArgs& b, float* dq, float* dk, float* dv, int64_t N,
               cudaStream_t st) {
  dim3 grid((unsigned)(ceil_div(N, BM) + a.S), (unsigned)a.H);
  const size_t smem_kv = sizeof(float) * (4 * BM * D + 8 * 64 * D);
  int rc = opt_in_smem(k_attn_bwd_kv_tc<D>, smem_kv, "attn_bwd_kv_tc");
  if (rc) return rc;
  k_attn_bwd_kv_tc<D><<<grid, kThreads, smem_kv, st>>>(a, b, dk, dv);
  const size_t smem_q = sizeof(float) * (4 * BM * D + 6 * BN * D);
  rc = opt_in_smem(k_attn_bwd_q_tc<D>, smem_q, "attn_bwd_q_tc");
  if (rc) return rc;
  k_attn_bwd_q_tc<D><<<grid, kThreads, smem_q, st>>>(a, b, dq);
  return check_launch("attn_varlen_bwd(tcgen05)");
}

}  // namespace tc

// Called by hgk_attn_varlen_fwd.  Returns 1 if this shape is not handled here (caller falls back
// to the SIMT kernel), 0 on success, a negative hgk_status on error.
int attn_fwd_tc(float* o, int64_t ldo, float* lse, const float* q, int64_t ldq, const float* k,
                int64_t ldk, const float* v, int64_t ldv, const int32_t* cu, int64_t S, int64_t N,
                int32_t H, int32_t D, float scale, cudaStream_t st) {
  auto ok16 = [](const void* ptr, int64_t ld) { return (reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && ld % 4 == 0; };
  if (!(D == 16 || D == 32 || D == 64)) return 1;
  if (!ok16(q, ldq) || !ok16(k, ldk) || !ok16(v, ldv) || !ok16(o, ldo)) return 1;
  AttnTcArgs a{q, ldq, k, ldk, v, ldv, cu, (int32_t)S, H, scale};
  switch (D) {
    case 16: return tc::launch_fwd<16>(a, o, ldo, lse, N, st);
    case 32: return tc::launch_fwd<32>(a, o, ldo, lse, N, st);
    default: return tc::launch_fwd<64>(a, o, ldo, lse, N, st);
  }
}

// Called by hgk_attn_varlen_bwd after delta = rowsum(dO * O) has been enqueued.  Same return
// convention as attn_fwd_tc.  dq, dk, dv are contiguous [N, H, D].
int attn_bwd_tc(float* dq, float* dk, float* dv, const float* delta, const float* dout, int64_t lddo,
                const float* lse, const float* q, int64_t ldq, const float* k, int64_t ldk,
                const float* v, int64_t ldv, const int32_t* cu, int64_t S, int64_t N, int32_t H,
                int32_t D, float scale, cudaStream_t st) {
  auto ok16 = [](const void* ptr, int64_t ld) { return (reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && ld % 4 == 0; };
  if (!(D == 16 || D == 32)) return 1;     // D = 64 would need 224+ KB of shared memory in the kv kernel
  if (!ok16(q, ldq) || !ok16(k, ldk) || !ok16(v, ldv) || !ok16(dout, lddo) || !ok16(dq, 4) || !ok16(dk, 4) || !ok16(dv, 4))
    return 1;
  AttnTcArgs a{q, ldq, k, ldk, v, ldv, cu, (int32_t)S, H, scale};
  tc::AttnTcBwdArgs b{dout, lddo, lse, delta};
  if (D == 16) return tc::launch_bwd<16>(a, b, dq, dk, dv, N, st);
  return tc::launch_bwd<32>(a, b, dq, dk, dv, N, st);
}

}  // namespace hgk
