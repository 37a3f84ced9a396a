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

// rows [row0, row0+ROWS) x D of a [*, ld] matrix -> hi / lo K-major core-matrix tiles
// (row = MN index, feature = K index), each value multiplied by `mul`; rows >= row_end are zero.
template <int ROWS, int D>
__device__ __forceinline__ void stage_rows(float* hi, float* lo, const float* src, int64_t ld,
                                           int32_t row0, int32_t row_end, float mul) {
  constexpr int CH = D / 4;                 // 16-byte chunks per row
  constexpr int SBO = CH * 32;              // floats per 8-row group
  for (int idx = threadIdx.x; idx < ROWS * CH; idx += kThreads) {
    const int r = idx / CH, c = idx - r * CH;
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row0 + r < row_end) x = __ldg(reinterpret_cast<const float4*>(src + (int64_t)(row0 + r) * ld + 4 * c));
    float4 h, l;
    split_tf32(x.x * mul, h.x, l.x); split_tf32(x.y * mul, h.y, l.y);
    split_tf32(x.z * mul, h.z, l.z); split_tf32(x.w * mul, h.w, l.w);
    const int off = core_off(r, c, SBO);
    *reinterpret_cast<float4*>(hi + off) = h;
    *reinterpret_cast<float4*>(lo + off) = l;
  }
}

// rows [row0, row0+COLS) x D of a [*, ld] matrix -> TRANSPOSED hi / lo tiles: MN index = feature
// (D rows), K index = source row (COLS columns).  Used for the B operand of P V (reduction over keys).
template <int COLS, int D>
__device__ __forceinline__ void stage_transposed(float* hi, float* lo, const float* src, int64_t ld,
                                                 int32_t row0, int32_t row_end) {
  constexpr int CH = D / 4;
  constexpr int SBO = (COLS / 4) * 32;      // floats per 8-feature group
  for (int idx = threadIdx.x; idx < COLS * CH; idx += kThreads) {
    const int r = idx / CH, c = idx - r * CH;     // r: source row (key), c: feature chunk
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row0 + r < row_end) x = __ldg(reinterpret_cast<const float4*>(src + (int64_t)(row0 + r) * ld + 4 * c));
    const float vals[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float h, l;
      split_tf32(vals[i], h, l);
      const int off = core_off(4 * c + i, r >> 2, SBO) + (r & 3);
      hi[off] = h;
      lo[off] = l;
    }
  }
}

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
  stage_rows<BM, D>(Qhi, Qlo, p.q + h * D, p.ldq, row0, seg_end, p.scale * kLog2e);

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
    stage_rows<BN, D>(Khi, Klo, p.k + h * D, p.ldk, k0, seg_end, 1.f);
    stage_transposed<BN, D>(Vhi, Vlo, p.v + h * D, p.ldv, k0, seg_end);
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

}  // namespace hgk
