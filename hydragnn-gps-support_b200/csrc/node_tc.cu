// Node-row kernels of the GPS block (reference: hydragnn/globalAtt/gps.py:72-83,116-150): the Linears
// around the attention kernel on the tensor cores, and BatchNorm with batch statistics.
//
//   hgk_linear_tc   Y = residual + scale * mask * act(X W^T + bias)         tcgen05, 3xTF32
//   hgk_bn_stats    per-column mean / 1/sqrt(var + eps) of a [N, C] matrix, running statistics update
//   hgk_bn_apply    out = act( BN_a(ta) [+ BN_b(tb)] )
//   hgk_bn_bwd      gradients of the above (column sums in a fixed order, then the elementwise part)
//   hgk_mask_axpy   out = residual + scale * mask * z
//
// The reference runs these as nn.Linear (cuBLAS fp32), PyG BatchNorm (= nn.BatchNorm1d) and F.dropout;
// the dropout masks stay torch's (bernoulli_ draws), handed in as uint8 so that forward and backward
// see the same mask.  All reductions over the N rows use a fixed slab / order: bit-reproducible.
#include "hgk_common.cuh"
#include "tc_common.cuh"
#include "tma_host.cuh"

namespace hgk {
namespace tc {

// ================================================================================ linear ======
// One CTA = one 128 x 128 output tile; K walked in 32-wide chunks: global -> registers (prefetched
// one chunk ahead, in flight during the MMAs) -> K-major hi / lo tiles in shared memory -> 3 MMAs
// per 8 k.  Two CTAs per SM (64 KB of shared memory, 128 TMEM columns each) hide each other's
// staging.  Same structure as k_wgrad_tn (filter_tc.cu), with the roles of rows and K swapped.
struct LinearArgs {
  const float* x; int64_t ldx;       // [M, K]
  const float* w; int64_t ldw;       // [N, K] (or, TRANS_B, [K, N])
  const float* bias;                 // [N] or NULL
  const uint8_t* mask;               // [M, N] (contiguous) or NULL: multiply by scale * mask
  const float* nz; int64_t ldnz;     // [M, N] or NULL: multiply by scale where nz != 0, by 0 elsewhere
  const float* residual; int64_t ldr;// [M, N] or NULL: added last
  float scale;
  int relu;
  float* y; int64_t ldy;             // [M, N]
  int64_t M; int N; int K;
  // K is cut into `chunks` pieces of `chunk_k` (a multiple of 32), each accumulated in its own 128
  // TMEM columns and added up in fp32 registers at the end: a tcgen05 accumulation chain loses
  // ~2^-24 of the sum per MMA (truncating adds), so chains are kept to 128 k (48 MMAs)
  int chunk_k; int chunks; uint32_t tmem_cols;
};

constexpr int kBK = 32;

// cp.async (LDGSTS): 16 bytes global -> shared without a register; src_bytes = 0 zero-fills
__device__ __forceinline__ void cp_async16(float* smem_dst, const float* gsrc, bool ok) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;"
               :: "r"(smem_u32(smem_dst)), "l"(gsrc), "r"(ok ? 16 : 0) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory"); }

// ---- epilogue shared by the two kernels: lane = row, this thread's 64 of the 128 columns ----
// `stage` / `tm_y` (TMA kernel): the tile leaves through shared memory ([4 column blocks][128 rows][32 floats],
// 128-byte swizzle: row-per-thread 16-byte stores without bank conflicts) and four bulk tensor stores, which clip
// at the matrix edge; NULL: row-per-thread global stores (one LSU wavefront per 16 bytes).
__device__ __forceinline__ void linear_epilogue(const LinearArgs& p, const CtaSetup& cs, uint32_t tmem, int32_t row0, int32_t col0,
                                                float* stage = nullptr, const CUtensorMap* tm_y = nullptr) {
  const int64_t row = (int64_t)row0 + cs.row_local;
  const bool row_ok = row < p.M;
#pragma unroll
  for (int cc = 0; cc < 64; cc += 32) {
    const int c_tile = cs.half * 64 + cc;               // column inside the tile
    float v[32];
    for (int chunk = 0; chunk < p.chunks; ++chunk) {    // CTA uniform: tcgen05.ld is warp collective
      uint32_t t[32];
      tmem_ld32(tmem + cs.lane_base + (uint32_t)chunk * 128 + c_tile, t);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = chunk == 0 ? __uint_as_float(t[j]) : v[j] + __uint_as_float(t[j]);
    }
    const int col = col0 + c_tile;
    if (row_ok && col < p.N) {                          // N is a multiple of 32: whole chunk in or out
      if (p.bias != nullptr) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + col + j));
          v[j] += b.x; v[j + 1] += b.y; v[j + 2] += b.z; v[j + 3] += b.w;
        }
      }
      if (p.relu) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
      }
      if (p.mask != nullptr) {
        const uint4* mp = reinterpret_cast<const uint4*>(p.mask + row * p.N + col);
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          const uint4 m = __ldg(mp + q);
          const uint32_t w4[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const uint32_t byte = (w4[j >> 2] >> (8 * (j & 3))) & 0xffu;
            v[16 * q + j] = byte ? v[16 * q + j] * p.scale : 0.f;
          }
        }
      }
      if (p.nz != nullptr) {
        const float* np_ = p.nz + row * p.ldnz + col;
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const float4 z = __ldg(reinterpret_cast<const float4*>(np_ + j));
          v[j] = z.x != 0.f ? v[j] * p.scale : 0.f;
          v[j + 1] = z.y != 0.f ? v[j + 1] * p.scale : 0.f;
          v[j + 2] = z.z != 0.f ? v[j + 2] * p.scale : 0.f;
          v[j + 3] = z.w != 0.f ? v[j + 3] * p.scale : 0.f;
        }
      }
      if (p.residual != nullptr) {
        const float* rp = p.residual + row * p.ldr + col;
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const float4 r = __ldg(reinterpret_cast<const float4*>(rp + j));
          v[j] += r.x; v[j + 1] += r.y; v[j + 2] += r.z; v[j + 3] += r.w;
        }
      }
      if (stage == nullptr) {
        float* yp = p.y + row * p.ldy + col;
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          *reinterpret_cast<float4*>(yp + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
      }
    }
    if (stage != nullptr) {                             // rows / columns outside the matrix are clipped by the store
      float* rowp = stage + (c_tile >> 5) * (128 * 32) + cs.row_local * 32;
#pragma unroll
      for (int j = 0; j < 8; ++j)
        *reinterpret_cast<float4*>(rowp + ((j ^ (cs.row_local & 7)) << 2)) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    }
  }
  if (stage != nullptr) {
    fence_smem_to_async();
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
      for (int blk = 0; blk < 4; ++blk)
        if (col0 + blk * 32 < p.N) tma_store_2d(tm_y, stage + blk * (128 * 32), col0 + blk * 32, row0);
      tma_store_commit();
      tma_store_wait();
    }
  }
}

constexpr int kLinDepth = 4;                             // K chunks in flight (raw copies in shared memory)

template <bool TRANS_B>
__global__ void __launch_bounds__(kThreads, 1)
k_linear_tc(const LinearArgs p) {
  extern __shared__ __align__(1024) float smem_raw[];
  // the MN-major B tile (TRANS_B) needs 1024-byte alignment; the launcher requests 1024 bytes more than the layout
  float* const smem = smem_raw + (((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u) >> 2);
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  // shared memory: operand tiles (A hi, A lo, B hi, B lo: 64 KB), then kLinDepth raw chunk slots of
  // 32 KB (A rows [128 x 32] and B rows [128 x 32] or, TRANS_B, [32 x 128])
  float* Ahi = smem;              float* Alo = Ahi + 128 * kBK;
  float* Bhi = Alo + 128 * kBK;   float* Blo = Bhi + 128 * kBK;
  float* raw = Blo + 128 * kBK;
  constexpr int kRawSlot = 2 * 128 * kBK;                // floats per raw chunk slot
  const uint32_t kCols = p.tmem_cols;
  const CtaSetup cs = cta_setup(&bar, &tmem_slot, kCols);
  const uint32_t tmem = cs.tmem;
  const int warp_u = warp_id_uniform();

  const int32_t row0 = (int32_t)blockIdx.x * 128;
  const int32_t col0 = (int32_t)blockIdx.y * 128;
  // TRANS_B: w is [K, N] row-major, i.e. its rows are the k index: an MN-major B tile, nothing to transpose
  const OperandDesc dA = operand_desc<kBK>(Ahi, Alo);
  const OperandDesc dB = TRANS_B ? operand_desc_mn<kBK>(Bhi, Blo) : operand_desc<kBK>(Bhi, Blo);
  constexpr uint32_t kIdesc = idesc_tf32(128, 128) | (TRANS_B ? kIdescBMajorMN : 0u);

  // The global-load latency used to be paid in every K iteration (ncu: long-scoreboard stalls dominate,
  // 16 us for [8192 x 128] x [128 x 128]).  Now the first kLinDepth chunks are requested at once with
  // cp.async into shared-memory slots; every thread copies exactly the 16-byte pieces it will itself
  // split into the hi / lo operand tiles, so its own cp.async.wait_group is the only synchronisation the
  // raw slots need, and a slot is refilled (chunk k + kLinDepth) as soon as its owner has read it.
  using TA = Tile<128, kBK>;
  using TB = Tile<128, kBK>;            // !TRANS_B: rows = output features
  using TBT = Tile<kBK, 128>;           // TRANS_B : rows = k
  const int nchunks = p.K / kBK;
  auto request = [&](int chunk) {       // one commit group per call (possibly empty: keeps the count uniform)
    if (chunk < nchunks) {
      const int k0 = chunk * kBK;
      float* slot = raw + (chunk % kLinDepth) * kRawSlot;
#pragma unroll
      for (int it = 0; it < TA::PER; ++it) {
        int r, c;
        TA::coord(it, r, c);
        const bool ok = row0 + r < (int32_t)p.M;
        cp_async16(slot + r * kBK + 4 * c, p.x + (int64_t)(ok ? row0 + r : 0) * p.ldx + k0 + 4 * c, ok);
      }
      float* bslot = slot + 128 * kBK;
      if (!TRANS_B) {
#pragma unroll
        for (int it = 0; it < TB::PER; ++it) {
          int r, c;
          TB::coord(it, r, c);
          const bool ok = col0 + r < p.N;
          cp_async16(bslot + r * kBK + 4 * c, p.w + (int64_t)(ok ? col0 + r : 0) * p.ldw + k0 + 4 * c, ok);
        }
      } else {
#pragma unroll
        for (int it = 0; it < TBT::PER; ++it) {
          int r, c;
          TBT::coord(it, r, c);            // r: k row of the chunk, c: 16-byte column of the 128 outputs
          cp_async16(bslot + r * 128 + 4 * c, p.w + (int64_t)(k0 + r) * p.ldw + col0 + 4 * c, true);
        }
      }
    }
    cp_async_commit();
  };

#pragma unroll
  for (int d = 0; d < kLinDepth; ++d) request(d);
  uint32_t parity = 0;
  for (int chunk = 0; chunk < nchunks; ++chunk) {
    cp_async_wait<kLinDepth - 1>();                     // this thread's pieces of `chunk` have landed
    const float* slot = raw + (chunk % kLinDepth) * kRawSlot;
    TA ta;
    TB tb;
    TBT tbt;
#pragma unroll
    for (int it = 0; it < TA::PER; ++it) {
      int r, c;
      TA::coord(it, r, c);
      ta.v[it] = *reinterpret_cast<const float4*>(slot + r * kBK + 4 * c);
    }
    if (!TRANS_B) {
#pragma unroll
      for (int it = 0; it < TB::PER; ++it) {
        int r, c;
        TB::coord(it, r, c);
        tb.v[it] = *reinterpret_cast<const float4*>(slot + 128 * kBK + r * kBK + 4 * c);
      }
    } else {
#pragma unroll
      for (int it = 0; it < TBT::PER; ++it) {
        int r, c;
        TBT::coord(it, r, c);
        tbt.v[it] = *reinterpret_cast<const float4*>(slot + 128 * kBK + r * 128 + 4 * c);
      }
    }
    if (chunk > 0) {                                     // the MMAs of the previous chunk have read the tiles
      mbar_wait(&bar, parity);
      parity ^= 1;
      fence_after_sync();
    }
    ta.store_rows(Ahi, Alo);
    if (!TRANS_B) tb.store_rows(Bhi, Blo);
    else tbt.store_mn(Bhi, Blo);
    // the slot is free again: its only reader has consumed the values (the stores above depend on them)
    request(chunk + kLinDepth);
    publish_operands();
    if (warp_u == 0 && elect_one_sync()) {
      fence_after_sync();
      const int k0 = chunk * kBK;
      const int acc = k0 / p.chunk_k;
      if (TRANS_B) mma3_ss_bmn<kBK>(tmem + (uint32_t)acc * 128, dA, dB, kIdesc, k0 > acc * p.chunk_k);
      else mma3_ss<kBK>(tmem + (uint32_t)acc * 128, dA, dB, kIdesc, k0 > acc * p.chunk_k);
      mma_commit(&bar);
    }
  }
  cp_async_wait<0>();
  mbar_wait(&bar, parity);                               // the last chunk's MMAs (K >= 32: at least one chunk)
  fence_after_sync();

  linear_epilogue(p, cs, tmem, row0, col0);
  cta_teardown(tmem, kCols);
}

// ---- the same with TMA loads -----------------------------------------------------------------------------
// x and w chunks ([128 x 32] boxes, or [32 x 128] of a transposed weight) land by cp.async.bulk.tensor in a ring of
// four hi slots, all requested up front; the raw tile is the hi operand (kind::tf32 truncates), the threads form the
// lo tiles in one of two buffers, so lo(c + 1) is computed while the MMAs of chunk c run.  Before, every chunk went
// cp.async -> registers -> split -> hi and lo stores -> MMA -> wait, one after the other.
template <bool TRANS_B>
__global__ void __launch_bounds__(kThreads, 1)
k_linear_tma(const LinearArgs p, const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_w,
             const __grid_constant__ CUtensorMap tm_y) {
  extern __shared__ __align__(1024) float smem_raw[];
  float* const smem = smem_raw + (((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u) >> 2);
  __shared__ __align__(8) uint64_t bar;                 // cta_setup
  __shared__ __align__(8) uint64_t full[4];             // hi slot s has landed
  __shared__ __align__(8) uint64_t done[2];             // the MMAs that read lo buffer b are complete
  __shared__ uint32_t tmem_slot;
  constexpr int kT = 128 * kBK;                         // floats per tile
  float* const Ahi = smem;                              // [4][kT]
  float* const Bhi = smem + 4 * kT;                     // [4][kT]
  float* const Alo = smem + 8 * kT;                     // [2][kT]
  float* const Blo = smem + 10 * kT;                    // [2][kT]
  if (threadIdx.x == 0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) mbar_init(&full[i], 1);
    mbar_init(&done[0], 1); mbar_init(&done[1], 1);
  }
  const CtaSetup cs = cta_setup(&bar, &tmem_slot, p.tmem_cols);
  const uint32_t tmem = cs.tmem;
  const int warp_u = warp_id_uniform();
  const int32_t row0 = (int32_t)blockIdx.x * 128;
  const int32_t col0 = (int32_t)blockIdx.y * 128;
  const int nchunks = p.K / kBK;
  constexpr uint32_t kIdesc = idesc_tf32(128, 128) | (TRANS_B ? kIdescBMajorMN : 0u);

  auto request = [&](int c) {                           // one thread: chunk c into slot c % 4
    const int sl = c & 3;
    mbar_expect_tx(&full[sl], 2u * kT * (uint32_t)sizeof(float));
    tma_load_2d(Ahi + sl * kT, &tm_x, &full[sl], c * kBK, row0);
    if (!TRANS_B) {
      tma_load_2d(Bhi + sl * kT, &tm_w, &full[sl], c * kBK, col0);
    } else {
#pragma unroll
      for (int blk = 0; blk < 4; ++blk) tma_load_2d(Bhi + sl * kT + blk * (kBK * 32), &tm_w, &full[sl], col0 + blk * 32, c * kBK);
    }
  };
  if (threadIdx.x == 0) {
    for (int c = 0; c < 4 && c < nchunks; ++c) request(c);
  }
  using TA = Tile<128, kBK>;                            // rows = output rows (or features), 32 k per chunk
  using TBT = Tile<kBK, 128>;                           // TRANS_B: rows = k
  uint32_t done_phase[2] = {0, 0};
  for (int c = 0; c < nchunks; ++c) {
    const int sl = c & 3, lb = c & 1;
    if (c >= 2) {                                       // lo buffer lb was read by the MMAs of chunk c - 2
      mbar_wait(&done[lb], done_phase[lb]);
      done_phase[lb] ^= 1;
      // ... and so was hi slot (c - 2) % 4: it takes chunk c + 2
      if (threadIdx.x == 0 && c + 2 >= 4 && c + 2 < nchunks) request(c + 2);
    }
    mbar_wait(&full[sl], (uint32_t)(c >> 2) & 1u);
    const float* ah = Ahi + sl * kT;
    const float* bh = Bhi + sl * kT;
    float* al = Alo + lb * kT;
    float* bl = Blo + lb * kT;
#pragma unroll
    for (int it = 0; it < TA::PER; ++it) {
      int r, cc;
      TA::coord(it, r, cc);
      const int off = TA::mn_off(r, cc);
      const float4 v = *reinterpret_cast<const float4*>(ah + off);
      float4 h, l;
      split_tf32(v.x, h.x, l.x); split_tf32(v.y, h.y, l.y); split_tf32(v.z, h.z, l.z); split_tf32(v.w, h.w, l.w);
      *reinterpret_cast<float4*>(al + off) = l;
    }
    constexpr int kPerB = TRANS_B ? TBT::PER : TA::PER;
#pragma unroll
    for (int it = 0; it < kPerB; ++it) {
      int r, cc;
      if (TRANS_B) TBT::coord(it, r, cc); else TA::coord(it, r, cc);
      const int off = TRANS_B ? TBT::mn_off(r, cc) : TA::mn_off(r, cc);
      const float4 v = *reinterpret_cast<const float4*>(bh + off);
      float4 h, l;
      split_tf32(v.x, h.x, l.x); split_tf32(v.y, h.y, l.y); split_tf32(v.z, h.z, l.z); split_tf32(v.w, h.w, l.w);
      *reinterpret_cast<float4*>(bl + off) = l;
    }
    publish_operands();
    if (warp_u == 0 && elect_one_sync()) {
      fence_after_sync();
      const int k0 = c * kBK;
      const int acc = k0 / p.chunk_k;
      const OperandDesc dA = operand_desc_k32(ah, al);
      const bool accumulate = k0 > acc * p.chunk_k;
      if (TRANS_B) {
        const OperandDesc dB = operand_desc_mn<kBK>(bh, bl);
        mma3_ss_k32_mn<kBK>(tmem + (uint32_t)acc * 128, dA, dB, kIdesc, accumulate);
      } else {
        const OperandDesc dB = operand_desc_k32(bh, bl);
        mma3_ss_k32<kBK>(tmem + (uint32_t)acc * 128, dA, dB, kIdesc, accumulate);
      }
      mma_commit(&done[lb]);
    }
  }
  // the last (up to two) chunks' MMAs
  if (nchunks >= 2) { const int lb = (nchunks - 2) & 1; mbar_wait(&done[lb], done_phase[lb]); done_phase[lb] ^= 1; }
  { const int lb = (nchunks - 1) & 1; mbar_wait(&done[lb], done_phase[lb]); }
  fence_after_sync();
  linear_epilogue(p, cs, tmem, row0, col0, Ahi, &tm_y);   // every MMA is complete: the hi slots are free
  cta_teardown(tmem, p.tmem_cols);
}

}  // namespace tc

// ============================================================================= BatchNorm ======
// Statistics of column c over the N rows: slabs of rows are reduced independently (mean of the slab,
// then the sum of squared deviations from that mean), the slabs are merged in slab order with the
// pairwise update of Chan et al.  Numerically this is the two-pass algorithm inside a slab and a
// stable merge across slabs; the order is fixed by (N, C) alone.
constexpr int kBnSlabRows = 256;

// Rows that count: the first min(N, *n_valid) (n_valid: device scalar or NULL).  Padded mini-batches
// (train/step_graph.py) keep their pad rows behind the real ones; statistics and gradients ignore them.
__device__ __forceinline__ int64_t valid_rows(int64_t N, const long long* n_valid) {
  if (n_valid == nullptr) return N;
  const int64_t v = (int64_t)__ldg(n_valid);
  return v < N ? (v > 0 ? v : 0) : N;
}

__global__ void __launch_bounds__(256)
k_bn_stats_partial(float* __restrict__ part_mean, float* __restrict__ part_m2, const float* __restrict__ t,
                   int64_t ldt, int64_t N_all, int C, const long long* __restrict__ n_valid) {
  __shared__ float red[8][32];
  __shared__ float mean_s[32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  const int64_t N = valid_rows(N_all, n_valid);
  const int64_t beg = (int64_t)blockIdx.y * kBnSlabRows;
  const int64_t end = beg + kBnSlabRows < N ? beg + kBnSlabRows : (N > beg ? N : beg);
  const float cnt = end > beg ? (float)(end - beg) : 1.f;
  float acc = 0.f;
  if (c < C)
    for (int64_t n = beg + warp; n < end; n += 8) acc += __ldg(t + n * ldt + c);
  red[warp][lane] = acc;
  __syncthreads();
  if (warp == 0) {
    float s = red[0][lane];
#pragma unroll
    for (int w = 1; w < 8; ++w) s += red[w][lane];
    mean_s[lane] = s / cnt;
  }
  __syncthreads();
  const float mean = mean_s[lane];
  acc = 0.f;
  if (c < C)
    for (int64_t n = beg + warp; n < end; n += 8) {
      const float d = __ldg(t + n * ldt + c) - mean;
      acc += d * d;
    }
  __syncthreads();
  red[warp][lane] = acc;
  __syncthreads();
  if (warp == 0 && c < C) {
    float s = red[0][lane];
#pragma unroll
    for (int w = 1; w < 8; ++w) s += red[w][lane];
    part_mean[(int64_t)blockIdx.y * C + c] = mean;
    part_m2[(int64_t)blockIdx.y * C + c] = s;
  }
}

__global__ void __launch_bounds__(128)
k_bn_stats_final(float* __restrict__ mean_out, float* __restrict__ rstd_out, float* __restrict__ running_mean,
                 float* __restrict__ running_var, long long* __restrict__ num_batches, float momentum, float eps,
                 const float* __restrict__ part_mean, const float* __restrict__ part_m2, int S, int64_t N_all, int C,
                 const long long* __restrict__ n_valid) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c == 0 && num_batches != nullptr) *num_batches += 1;
  if (c >= C) return;
  const int64_t N = valid_rows(N_all, n_valid);
  float n_a = 0.f, mean = 0.f, m2 = 0.f;
  for (int s0 = 0; s0 < S; s0 += 8) {                   // the slabs' statistics are fetched eight at a time
    float mb[8], qb[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const bool in = s0 + j < S;
      mb[j] = in ? __ldg(part_mean + (int64_t)(s0 + j) * C + c) : 0.f;
      qb[j] = in ? __ldg(part_m2 + (int64_t)(s0 + j) * C + c) : 0.f;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int s = s0 + j;
      if (s < S && (int64_t)s * kBnSlabRows < N) {
        const int64_t beg = (int64_t)s * kBnSlabRows;
        const float n_b = (float)((beg + kBnSlabRows < N ? beg + kBnSlabRows : N) - beg);
        const float n = n_a + n_b;
        const float delta = mb[j] - mean;
        mean += delta * (n_b / n);
        m2 += qb[j] + delta * delta * (n_a * n_b / n);
        n_a = n;
      }
    }
  }
  const float var = N > 0 ? m2 / (float)N : 0.f;        // biased: what normalises the batch
  mean_out[c] = mean;
  rstd_out[c] = rsqrtf(var + eps);
  if (running_mean != nullptr) {
    const float unbiased = N > 1 ? m2 / (float)(N - 1) : var;
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
  }
}

struct BnSide {
  const float* t; int64_t ldt;
  const float* mean; const float* rstd; const float* gamma; const float* beta;
};

// out = act( BN_a(ta) [+ BN_b(tb)] ); one float4 per thread
__global__ void __launch_bounds__(256)
k_bn_apply(float* __restrict__ out, int64_t ldo, const BnSide a, const BnSide b, int relu, int64_t N, int C4) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N * C4) return;
  const int64_t n = i / C4;
  const int c = (int)(i % C4) * 4;
  auto side = [&](const BnSide& s) {
    const float4 x = __ldg(reinterpret_cast<const float4*>(s.t + n * s.ldt + c));
    const float4 m = __ldg(reinterpret_cast<const float4*>(s.mean + c));
    const float4 r = __ldg(reinterpret_cast<const float4*>(s.rstd + c));
    float4 g = make_float4(1.f, 1.f, 1.f, 1.f), be = make_float4(0.f, 0.f, 0.f, 0.f);
    if (s.gamma != nullptr) g = __ldg(reinterpret_cast<const float4*>(s.gamma + c));
    if (s.beta != nullptr) be = __ldg(reinterpret_cast<const float4*>(s.beta + c));
    return make_float4((x.x - m.x) * r.x * g.x + be.x, (x.y - m.y) * r.y * g.y + be.y,
                       (x.z - m.z) * r.z * g.z + be.z, (x.w - m.w) * r.w * g.w + be.w);
  };
  float4 v = side(a);
  if (b.t != nullptr) {
    const float4 w = side(b);
    v.x += w.x; v.y += w.y; v.z += w.z; v.w += w.w;
  }
  if (relu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
  *reinterpret_cast<float4*>(out + n * ldo + c) = v;
}

// column sums of the backward: S0 = sum g, Sa = sum g * xhat_a, Sb = sum g * xhat_b (slab partials)
__global__ void __launch_bounds__(256)
k_bn_bwd_partial(float* __restrict__ partial, const float* __restrict__ g, int64_t ldg, const BnSide a,
                 const BnSide b, int64_t N_all, int C, int S, const long long* __restrict__ n_valid) {
  __shared__ float red[3][8][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  const int64_t N = valid_rows(N_all, n_valid);
  const int64_t beg = (int64_t)blockIdx.y * kBnSlabRows;
  const int64_t end = beg + kBnSlabRows < N ? beg + kBnSlabRows : N;
  float s0 = 0.f, sa = 0.f, sb = 0.f;
  if (c < C) {
    const float ma = __ldg(a.mean + c), ra = __ldg(a.rstd + c);
    const float mb = b.t ? __ldg(b.mean + c) : 0.f, rb = b.t ? __ldg(b.rstd + c) : 0.f;
    for (int64_t n = beg + warp; n < end; n += 8) {
      const float gv = __ldg(g + n * ldg + c);
      s0 += gv;
      sa += gv * ((__ldg(a.t + n * a.ldt + c) - ma) * ra);
      if (b.t) sb += gv * ((__ldg(b.t + n * b.ldt + c) - mb) * rb);
    }
  }
  red[0][warp][lane] = s0; red[1][warp][lane] = sa; red[2][warp][lane] = sb;
  __syncthreads();
  if (warp < 3 && c < C) {
    float s = red[warp][0][lane];
#pragma unroll
    for (int w = 1; w < 8; ++w) s += red[warp][w][lane];
    partial[((int64_t)warp * S + blockIdx.y) * C + c] = s;
  }
}

// sums[k][c] = sum over slabs (in order), k = 0 (g), 1 (g xhat_a), 2 (g xhat_b);
// dgamma_a = sums[1], dbeta_a = sums[0] (and dgamma_b = sums[2], dbeta_b = sums[0])
__global__ void __launch_bounds__(128)
k_bn_bwd_final(float* __restrict__ sums, float* __restrict__ dgamma_a, float* __restrict__ dbeta_a,
               float* __restrict__ dgamma_b, float* __restrict__ dbeta_b, const float* __restrict__ partial,
               int S, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float acc[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    float t = 0.f;
#pragma unroll 8
    for (int s = 0; s < S; ++s) t += __ldg(partial + ((int64_t)k * S + s) * C + c);
    acc[k] = t;
    sums[(int64_t)k * C + c] = t;
  }
  if (dbeta_a) dbeta_a[c] = acc[0];
  if (dgamma_a) dgamma_a[c] = acc[1];
  if (dbeta_b) dbeta_b[c] = acc[0];
  if (dgamma_b) dgamma_b[c] = acc[2];
}

struct BnBwdOut {
  float* dt;              // [N, C] gradient at the BatchNorm input
  float* dz; const uint8_t* mask; float scale;   // optional: dz = scale * mask * dt
};

// dt = gamma * rstd * (g - S0 / N - xhat * Sx / N)
__global__ void __launch_bounds__(256)
k_bn_bwd_apply(const float* __restrict__ g, int64_t ldg, const BnSide a, const BnSide b, const BnBwdOut oa,
               const BnBwdOut ob, const float* __restrict__ sums, int64_t N_all, int C,
               const long long* __restrict__ n_valid) {
  const int C4 = C / 4;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N_all * C4) return;
  const int64_t n = i / C4;
  const int c = (int)(i % C4) * 4;
  const int64_t N = valid_rows(N_all, n_valid);
  if (n >= N) {                                          // pad rows: no gradient (they never reach the loss)
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    *reinterpret_cast<float4*>(oa.dt + n * C + c) = z;
    if (oa.dz != nullptr) *reinterpret_cast<float4*>(oa.dz + n * C + c) = z;
    if (b.t != nullptr) {
      *reinterpret_cast<float4*>(ob.dt + n * C + c) = z;
      if (ob.dz != nullptr) *reinterpret_cast<float4*>(ob.dz + n * C + c) = z;
    }
    return;
  }
  const float inv_n = 1.f / (float)N;
  const float4 gv = __ldg(reinterpret_cast<const float4*>(g + n * ldg + c));
  const float4 s0 = __ldg(reinterpret_cast<const float4*>(sums + c));
  auto side = [&](const BnSide& s, const BnBwdOut& o, const float* sx_ptr) {
    const float4 x = __ldg(reinterpret_cast<const float4*>(s.t + n * s.ldt + c));
    const float4 m = __ldg(reinterpret_cast<const float4*>(s.mean + c));
    const float4 r = __ldg(reinterpret_cast<const float4*>(s.rstd + c));
    const float4 sx = __ldg(reinterpret_cast<const float4*>(sx_ptr + c));
    float4 ga = make_float4(1.f, 1.f, 1.f, 1.f);
    if (s.gamma != nullptr) ga = __ldg(reinterpret_cast<const float4*>(s.gamma + c));
    float4 d;
    d.x = ga.x * r.x * (gv.x - s0.x * inv_n - (x.x - m.x) * r.x * (sx.x * inv_n));
    d.y = ga.y * r.y * (gv.y - s0.y * inv_n - (x.y - m.y) * r.y * (sx.y * inv_n));
    d.z = ga.z * r.z * (gv.z - s0.z * inv_n - (x.z - m.z) * r.z * (sx.z * inv_n));
    d.w = ga.w * r.w * (gv.w - s0.w * inv_n - (x.w - m.w) * r.w * (sx.w * inv_n));
    *reinterpret_cast<float4*>(o.dt + n * C + c) = d;
    if (o.dz != nullptr) {
      const uint32_t mk = __ldg(reinterpret_cast<const uint32_t*>(o.mask + n * C + c));
      float4 z;
      z.x = (mk & 0xffu) ? d.x * o.scale : 0.f;
      z.y = (mk & 0xff00u) ? d.y * o.scale : 0.f;
      z.z = (mk & 0xff0000u) ? d.z * o.scale : 0.f;
      z.w = (mk & 0xff000000u) ? d.w * o.scale : 0.f;
      *reinterpret_cast<float4*>(o.dz + n * C + c) = z;
    }
  };
  side(a, oa, sums + C);
  if (b.t != nullptr) side(b, ob, sums + 2 * (int64_t)C);
}

// out = residual + scale * mask * z   (any of residual / mask may be NULL)
__global__ void __launch_bounds__(256)
k_mask_axpy(float* __restrict__ out, const float* __restrict__ z, int64_t ldz, const uint8_t* __restrict__ mask,
            float scale, const float* __restrict__ residual, int64_t ldr, int64_t N, int C4) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N * C4) return;
  const int64_t n = i / C4;
  const int c = (int)(i % C4) * 4;
  float4 v = __ldg(reinterpret_cast<const float4*>(z + n * ldz + c));
  if (mask != nullptr) {
    const uint32_t mk = __ldg(reinterpret_cast<const uint32_t*>(mask + n * (int64_t)C4 * 4 + c));
    v.x = (mk & 0xffu) ? v.x * scale : 0.f;
    v.y = (mk & 0xff00u) ? v.y * scale : 0.f;
    v.z = (mk & 0xff0000u) ? v.z * scale : 0.f;
    v.w = (mk & 0xff000000u) ? v.w * scale : 0.f;
  }
  if (residual != nullptr) {
    const float4 r = __ldg(reinterpret_cast<const float4*>(residual + n * ldr + c));
    v.x += r.x; v.y += r.y; v.z += r.z; v.w += r.w;
  }
  *reinterpret_cast<float4*>(out + n * (int64_t)C4 * 4 + c) = v;
}

}  // namespace hgk

using namespace hgk;

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

extern "C" int hgk_linear_tc(float* y, int64_t ldy, const float* x, int64_t ldx, const float* w, int64_t ldw,
                             int32_t trans_w, const float* bias, int32_t relu, const uint8_t* mask,
                             const float* nz, int64_t ldnz, float scale, const float* residual, int64_t ldr,
                             int64_t M, int32_t N, int32_t K, void* stream) {
  HGK_REQUIRE(M >= 0 && N >= 1 && K >= 1, "linear_tc: bad sizes M=%lld N=%d K=%d", (long long)M, N, K);
  if (M == 0) return HGK_OK;
  HGK_REQUIRE(y && x && w, "linear_tc: NULL operand");
  if (K % tc::kBK != 0 || N % 32 != 0 || (trans_w && N % 128 != 0))
    return fail(HGK_ERR_UNSUPPORTED, "linear_tc: needs K %% 32 == 0 and N %% 32 == 0 (N %% 128 == 0 with a transposed "
                "weight); got N=%d K=%d trans=%d", N, K, trans_w);
  HGK_REQUIRE(M < (1ll << 31) - 256, "linear_tc: M too large");
  HGK_REQUIRE(ldx % 4 == 0 && ldw % 4 == 0 && ldy % 4 == 0 && aligned16(x) && aligned16(w) && aligned16(y),
              "linear_tc: rows of x, w, y must be 16-byte aligned");
  HGK_REQUIRE(ldx >= K && ldy >= N && ldw >= (trans_w ? N : K), "linear_tc: row stride smaller than the row");
  HGK_REQUIRE(!bias || aligned16(bias), "linear_tc: bias must be 16-byte aligned");
  HGK_REQUIRE(!mask || aligned16(mask), "linear_tc: mask must be 16-byte aligned");
  HGK_REQUIRE(!nz || (ldnz % 4 == 0 && ldnz >= N && aligned16(nz)), "linear_tc: nz rows must be 16-byte aligned");
  HGK_REQUIRE(!residual || (ldr % 4 == 0 && ldr >= N && aligned16(residual)), "linear_tc: residual rows must be 16-byte aligned");
  // accumulation chains of at most 128 k while four accumulators suffice (K <= 512), longer beyond
  int chunks = (int)ceil_div(K, 128);
  if (chunks > 4) chunks = 4;
  const int chunk_k = (int)(ceil_div(ceil_div(K, chunks), tc::kBK) * tc::kBK);
  chunks = (int)ceil_div(K, chunk_k);
  const uint32_t tmem_cols = chunks == 1 ? 128u : (chunks == 2 ? 256u : 512u);
  tc::LinearArgs a{x, ldx, w, ldw, bias, mask, nz, ldnz, residual, ldr, scale, relu, y, ldy, M, N, K,
                   chunk_k, chunks, tmem_cols};
  // 64 KB of operand tiles + kLinDepth raw chunk slots of 32 KB: one CTA per SM
  const size_t smem_max = sizeof(float) * (4 * 128 * tc::kBK + tc::kLinDepth * 2 * 128 * tc::kBK) + 1024;   // + alignment slack
  const size_t smem = smem_max;
  static bool configured[2] = {false, false};                   // the attribute is per function, set once
  cudaError_t err = cudaSuccess;
  if (!configured[trans_w ? 1 : 0]) {
    err = trans_w ? cudaFuncSetAttribute(tc::k_linear_tc<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_max)
                  : cudaFuncSetAttribute(tc::k_linear_tc<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_max);
    if (err != cudaSuccess) return fail(HGK_ERR_CUDA, "linear_tc: smem opt-in: %s", cudaGetErrorString(err));
    configured[trans_w ? 1 : 0] = true;
  }
  dim3 grid((unsigned)ceil_div(M, 128), (unsigned)ceil_div(N, 128));
  // TMA-fed kernel (HGK_LINEAR_TMA=0: the cp.async one, for A/B runs; also when the driver will not encode the maps)
  static const bool want_tma = [] { const char* e = getenv("HGK_LINEAR_TMA"); return !(e && e[0] == '0'); }();
  CUtensorMap tm_x, tm_w, tm_y;
  if (want_tma && tma::rows_map_32(&tm_x, x, M, K, ldx, 128) &&
      (trans_w ? tma::rows_map_32(&tm_w, w, K, N, ldw, 32) : tma::rows_map_32(&tm_w, w, N, K, ldw, 128)) &&
      tma::rows_map_128b(&tm_y, y, M, N, ldy, 128)) {
    const size_t smem_tma = sizeof(float) * 12 * 128 * tc::kBK + 1024;   // 4 + 4 hi slots, 2 + 2 lo buffers, alignment slack
    static bool configured_tma[2] = {false, false};
    if (!configured_tma[trans_w ? 1 : 0]) {
      err = trans_w ? cudaFuncSetAttribute(tc::k_linear_tma<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_tma)
                    : cudaFuncSetAttribute(tc::k_linear_tma<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_tma);
      if (err != cudaSuccess) return fail(HGK_ERR_CUDA, "linear_tc: smem opt-in: %s", cudaGetErrorString(err));
      configured_tma[trans_w ? 1 : 0] = true;
    }
    if (trans_w) tc::k_linear_tma<true><<<grid, tc::kThreads, smem_tma, (cudaStream_t)stream>>>(a, tm_x, tm_w, tm_y);
    else tc::k_linear_tma<false><<<grid, tc::kThreads, smem_tma, (cudaStream_t)stream>>>(a, tm_x, tm_w, tm_y);
    return check_launch("linear_tc(tma)");
  }
  if (trans_w) tc::k_linear_tc<true><<<grid, tc::kThreads, smem, (cudaStream_t)stream>>>(a);
  else tc::k_linear_tc<false><<<grid, tc::kThreads, smem, (cudaStream_t)stream>>>(a);
  return check_launch("linear_tc");
}

extern "C" int64_t hgk_bn_slabs(int64_t N) { return N <= 0 ? 1 : ceil_div(N, kBnSlabRows); }

static int bn_check(const char* what, const float* t, int64_t ldt, int64_t N, int32_t C) {
  HGK_REQUIRE(N >= 1 && C >= 4 && C % 4 == 0, "%s: needs N >= 1 and C %% 4 == 0 (N=%lld C=%d)", what, (long long)N, C);
  HGK_REQUIRE(t && ldt >= C && ldt % 4 == 0 && aligned16(t), "%s: rows must be 16-byte aligned", what);
  return HGK_OK;
}

extern "C" int hgk_bn_stats(float* mean, float* rstd, float* running_mean, float* running_var,
                            int64_t* num_batches_tracked, float momentum, float eps, float* partial,
                            const float* t, int64_t ldt, int64_t N, int32_t C, const int64_t* n_valid,
                            void* stream) {
  int rc = bn_check("bn_stats", t, ldt, N, C);
  if (rc) return rc;
  HGK_REQUIRE(mean && rstd && partial, "bn_stats: NULL output");
  HGK_REQUIRE((running_mean == nullptr) == (running_var == nullptr), "bn_stats: running_mean and running_var go together");
  const int S = (int)hgk_bn_slabs(N);
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid((unsigned)ceil_div(C, 32), (unsigned)S);
  k_bn_stats_partial<<<grid, 256, 0, st>>>(partial, partial + (int64_t)S * C, t, ldt, N, C, (const long long*)n_valid);
  k_bn_stats_final<<<(unsigned)ceil_div(C, 32), 32, 0, st>>>(mean, rstd, running_mean, running_var,
                                                               (long long*)num_batches_tracked, momentum, eps, partial,
                                                               partial + (int64_t)S * C, S, N, C,
                                                               (const long long*)n_valid);
  return check_launch("bn_stats");
}

extern "C" int hgk_bn_apply(float* out, int64_t ldo, const float* ta, int64_t ldta, const float* mean_a,
                            const float* rstd_a, const float* gamma_a, const float* beta_a, const float* tb,
                            int64_t ldtb, const float* mean_b, const float* rstd_b, const float* gamma_b,
                            const float* beta_b, int32_t relu, int64_t N, int32_t C, void* stream) {
  int rc = bn_check("bn_apply", ta, ldta, N, C);
  if (rc) return rc;
  if (tb) { rc = bn_check("bn_apply", tb, ldtb, N, C); if (rc) return rc; }
  HGK_REQUIRE(out && ldo >= C && ldo % 4 == 0 && aligned16(out) && mean_a && rstd_a && (!tb || (mean_b && rstd_b)),
              "bn_apply: NULL / misaligned operand");
  const BnSide a{ta, ldta, mean_a, rstd_a, gamma_a, beta_a}, b{tb, ldtb, mean_b, rstd_b, gamma_b, beta_b};
  const int64_t total = N * (C / 4);
  k_bn_apply<<<(unsigned)ceil_div(total, 256), 256, 0, (cudaStream_t)stream>>>(out, ldo, a, b, relu, N, C / 4);
  return check_launch("bn_apply");
}

extern "C" int hgk_bn_bwd(float* dta, float* dza, const uint8_t* mask_a, float scale_a, float* dgamma_a,
                          float* dbeta_a, float* dtb, float* dzb, const uint8_t* mask_b, float scale_b,
                          float* dgamma_b, float* dbeta_b, float* workspace, const float* g, int64_t ldg,
                          const float* ta, int64_t ldta, const float* mean_a, const float* rstd_a,
                          const float* gamma_a, const float* tb, int64_t ldtb, const float* mean_b,
                          const float* rstd_b, const float* gamma_b, int64_t N, int32_t C, const int64_t* n_valid,
                          void* stream) {
  int rc = bn_check("bn_bwd", ta, ldta, N, C);
  if (rc) return rc;
  rc = bn_check("bn_bwd", g, ldg, N, C);
  if (rc) return rc;
  if (tb) { rc = bn_check("bn_bwd", tb, ldtb, N, C); if (rc) return rc; }
  HGK_REQUIRE(dta && workspace && mean_a && rstd_a && (!tb || (dtb && mean_b && rstd_b)), "bn_bwd: NULL operand");
  HGK_REQUIRE((dza == nullptr || mask_a) && (dzb == nullptr || mask_b), "bn_bwd: dz needs its mask");
  const int S = (int)hgk_bn_slabs(N);
  cudaStream_t st = (cudaStream_t)stream;
  const BnSide a{ta, ldta, mean_a, rstd_a, gamma_a, nullptr}, b{tb, ldtb, mean_b, rstd_b, gamma_b, nullptr};
  float* partial = workspace;                       // [3, S, C]
  float* sums = workspace + 3 * (int64_t)S * C;     // [3, C]
  dim3 grid((unsigned)ceil_div(C, 32), (unsigned)S);
  k_bn_bwd_partial<<<grid, 256, 0, st>>>(partial, g, ldg, a, b, N, C, S, (const long long*)n_valid);
  k_bn_bwd_final<<<(unsigned)ceil_div(C, 32), 32, 0, st>>>(sums, dgamma_a, dbeta_a, tb ? dgamma_b : nullptr,
                                                             tb ? dbeta_b : nullptr, partial, S, C);
  const BnBwdOut oa{dta, dza, mask_a, scale_a}, ob{dtb, dzb, mask_b, scale_b};
  const int64_t total = N * (C / 4);
  k_bn_bwd_apply<<<(unsigned)ceil_div(total, 256), 256, 0, st>>>(g, ldg, a, b, oa, ob, sums, N, C,
                                                                 (const long long*)n_valid);
  return check_launch("bn_bwd");
}

extern "C" int64_t hgk_bn_bwd_workspace_floats(int64_t N, int32_t C) { return 3 * (hgk_bn_slabs(N) + 1) * (int64_t)C; }

extern "C" int hgk_mask_axpy(float* out, const float* z, int64_t ldz, const uint8_t* mask, float scale,
                             const float* residual, int64_t ldr, int64_t N, int32_t C, void* stream) {
  HGK_REQUIRE(N >= 0 && C >= 4 && C % 4 == 0, "mask_axpy: needs C %% 4 == 0");
  if (N == 0) return HGK_OK;
  HGK_REQUIRE(out && z && ldz >= C && ldz % 4 == 0 && aligned16(out) && aligned16(z), "mask_axpy: misaligned operand");
  HGK_REQUIRE(!residual || (ldr >= C && ldr % 4 == 0 && aligned16(residual)), "mask_axpy: misaligned residual");
  HGK_REQUIRE(!mask || (reinterpret_cast<uintptr_t>(mask) & 3) == 0, "mask_axpy: misaligned mask");
  const int64_t total = N * (C / 4);
  k_mask_axpy<<<(unsigned)ceil_div(total, 256), 256, 0, (cudaStream_t)stream>>>(out, z, ldz, mask, scale, residual,
                                                                                ldr, N, C / 4);
  return check_launch("mask_axpy");
}
