// SchNet filter-generating network on the tensor cores, one pass over the edges.
//
// Reference (hydragnn/models/SCFStack.py:259-272 with PyG GaussianSmearing / ShiftedSoftplus):
//     rbf_e = exp(coeff (d_e - mu_k)^2)                         [E, G]
//     C_e   = 0.5 (cos(d_e pi / r_c) + 1)
//     W_e   = ( Linear1( ssp( Linear0( [rbf_e | edge_attr_e] ) ) ) ) C_e    [E, F]
// evaluated there as five full-width [E, .] torch kernels plus two cuBLAS GEMMs per layer.  Here a
// persistent CTA walks 128-edge tiles and keeps everything on chip between the two GEMMs:
//     X  = [gauss(d) | raw]      built in registers, written to TMEM as the A operand (hi, lo)
//     Z0 = X W0'^T               tcgen05.mma TS, W0' (hi, lo) resident in shared memory
//     H  = ssp(Z0 + b0)          registers; stored to HBM for the backward and back to TMEM (hi, lo)
//     Z1 = H W1^T                tcgen05.mma TS, W1 (hi, lo) resident in shared memory
//     W  = (Z1 + b1) C           registers -> HBM
// `raw` is the un-embedded edge attribute of nn.FactoredEdgeAttr (e.g. rel_pe, r = 8 columns) whose
// embedding matrix is folded into W0' on the host; W0' columns are [G gaussians | pad to 4 | r raw |
// pad to 64].  fp32 accuracy through 3xTF32 operand splitting as in attn_tc.cu.
#include <cstring>
#include "hgk_common.cuh"
#include "tc_common.cuh"
#include "tma_host.cuh"

namespace hgk {
namespace tc {

constexpr int kFilt = 128;        // filters (N of both GEMMs); also the tile height
constexpr int kK0 = 64;           // padded input width of the first Linear

struct FilterArgs {
  const float* d;                 // [E] edge lengths
  const float* raw; int64_t ldr;  // [E, R] raw edge attribute (may be NULL when R == 0)
  const float* offset;            // [G] gaussian centres
  float coeff;                    // -0.5 / (mu_1 - mu_0)^2
  float pi_over_cutoff;
  const float* w0;                // [128, 64]  folded + padded first-layer weight
  const float* b0;                // [128]
  const float* w1;                // [128, 128]
  const float* b1;                // [128]
  float* h;                       // [E, 128] out: ssp(Z0 + b0)   (kept for the backward)
  float* w;                       // [E, 128] out: filter
  int64_t E;
  int tma_out;                    // TMA variant: bit 0: h through the bulk store, bit 1: w (else row-per-thread stores)
};

__device__ __forceinline__ float fast_log2(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// softplus(z) - ln 2 = max(z, 0) + ln(1 + e^{-|z|}) - ln 2, with two MUFU ops (ex2, lg2); absolute
// error ~1e-7 (torch: z > 20 ? z : log1p(exp(z)), same value to fp32 rounding)
__device__ __forceinline__ float shifted_softplus(float z) {
  const float u = fast_exp2(-fabsf(z) * kLog2e);
  return fmaxf(z, 0.f) + (fast_log2(1.f + u) - 1.f) * kLn2;
}

constexpr int kFiltThreads = 512;   // 16 warps: 4 threads per row, 32 of the 128 columns each

#ifdef HGK_KV_PROFILE
// Phase timing of the forward filter kernel (debug builds only: profiles/filter_phase_profile.py)
__device__ unsigned long long g_filt_prof[8];
__device__ __forceinline__ unsigned long long fprof_clock() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%clock64;" : "=l"(t) :: "memory");
  return t;
}
#define F_PROF_DECL unsigned long long fprof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0}; unsigned long long fpt = fprof_clock();
#define F_PROF(slot) { const unsigned long long now = fprof_clock(); fprof_acc[slot] += now - fpt; fpt = now; }
#define F_PROF_FLUSH if (threadIdx.x == 0 && blockIdx.x == 0) { for (int i = 0; i < 8; ++i) g_filt_prof[i] += fprof_acc[i]; }
#else
#define F_PROF_DECL
#define F_PROF(slot)
#define F_PROF_FLUSH
#endif

// TMA = true: h and w leave through shared memory and cp.async.bulk.tensor stores.  A thread owns 32 columns of one
// row; written straight to global memory that is one LSU wavefront per 32 bytes (the two epilogues' stores cost
// ~8k of the 17k cycles of a tile, profiles/filter_phase_profile.py).  Instead each column quarter (4 warps) has an
// 8 KB buffer [128 rows x 16 columns] (64-byte swizzle: row-per-thread 16-byte stores without bank conflicts),
// filled and handed to one bulk store twice per epilogue; the global writes drain while the next MMAs run.
template <int G, int R, bool TMA>
__global__ void __launch_bounds__(kFiltThreads, 1)
k_schnet_filter_fwd(const FilterArgs p, const __grid_constant__ CUtensorMap tm_h, const __grid_constant__ CUtensorMap tm_w) {
  constexpr int GP = (G + 3) / 4 * 4;
  static_assert(GP + R <= kK0 && R % 4 == 0 && G > 48 && GP <= 52, "input layout: gaussians end in the last quarter");
  extern __shared__ __align__(512) float smem_raw[];
  float* const smem = smem_raw + (((512u - (smem_u32(smem_raw) & 511u)) & 511u) >> 2);
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  __shared__ float off_s[kK0], b0_s[kFilt], b1_s[kFilt];

  float* sbuf = smem;                 // TMA: [4 quarters][128 rows][16 floats] staging of the outputs
  float* W0hi = smem + (TMA ? 4 * kFilt * 16 : 0);   float* W0lo = W0hi + kFilt * kK0;
  float* W1hi = W0lo + kFilt * kK0;   float* W1lo = W1hi + kFilt * kFilt;

  constexpr uint32_t kCols = 512;
  constexpr uint32_t kColXhi = 0, kColXlo = 64, kColZ = 128, kColHhi = 256, kColHlo = 384;
  const CtaSetup cs = cta_setup(&bar, &tmem_slot, kCols);     // cs.half = column quarter 0..3
  const uint32_t tmem = cs.tmem;
  const int warp_u = warp_id_uniform();
  const int quarter = cs.half;

  // resident operands: weights (hi, lo) in shared memory, biases and gaussian centres
  if (threadIdx.x < kThreads) {
    Tile<kFilt, kK0> t0;
    t0.load(p.w0, kK0, 0, kFilt);
    t0.store_rows(W0hi, W0lo);
    Tile<kFilt, kFilt> t1;
    t1.load(p.w1, kFilt, 0, kFilt);
    t1.store_rows(W1hi, W1lo);
  }
  if (threadIdx.x < kK0) off_s[threadIdx.x] = threadIdx.x < G ? p.offset[threadIdx.x] : 0.f;
  if (threadIdx.x < kFilt) { b0_s[threadIdx.x] = p.b0[threadIdx.x]; b1_s[threadIdx.x] = p.b1[threadIdx.x]; }
  const OperandDesc dW0 = operand_desc<kK0>(W0hi, W0lo);
  const OperandDesc dW1 = operand_desc<kFilt>(W1hi, W1lo);
  constexpr uint32_t kIdesc = idesc_tf32(kFilt, kFilt);
  publish_operands();

  uint32_t parity = 0;
  const float c2 = p.coeff * kLog2e;
  const int64_t num_tiles = (p.E + kFilt - 1) / kFilt;
  // TMA: this thread's 32 columns v[0..31] of row (tile, row_local) -> two bulk stores of the quarter's buffer
  float* const qbuf = sbuf + quarter * (kFilt * 16);
  const bool q_issuer = cs.row_local == 0;             // first thread of the quarter's four warps
  auto store_quarter = [&](const CUtensorMap* map, const uint32_t (&v)[32], int64_t tile) {
#pragma unroll
    for (int ph = 0; ph < 2; ++ph) {
      if (q_issuer) tma_store_wait_read();             // the previous store has read the buffer
      named_bar_sync(4 + quarter, 128);
      float* rowp = qbuf + cs.row_local * 16;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        *reinterpret_cast<uint4*>(rowp + ((j ^ ((cs.row_local >> 1) & 3)) << 2)) =
            make_uint4(v[ph * 16 + 4 * j], v[ph * 16 + 4 * j + 1], v[ph * 16 + 4 * j + 2], v[ph * 16 + 4 * j + 3]);
      fence_smem_to_async();
      named_bar_sync(4 + quarter, 128);
      if (q_issuer) {
        tma_store_2d(map, qbuf, quarter * 32 + ph * 16, (int32_t)(tile * kFilt));
        tma_store_commit();
      }
    }
  };
  F_PROF_DECL
  // X = [gauss | raw] of one edge row: this thread's 16 of the 64 columns -> TMEM (hi, lo)
  auto build_x = [&](int64_t e, bool ok, float dist) {
    uint32_t xh[16], xl[16];
    float val[16];
    if (quarter < 3) {                                  // columns 0..47: gaussians only
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const float t = dist - off_s[quarter * 16 + j];
        val[j] = fast_exp2(c2 * (t * t));
      }
    } else {                                            // columns 48..63: last gaussians | pad | raw | pad
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int k = 48 + j;                           // compile-time after unrolling
        if (k < G) { const float t = dist - off_s[k]; val[j] = fast_exp2(c2 * (t * t)); }
        else val[j] = 0.f;
      }
      if (R > 0 && ok) {
#pragma unroll
        for (int c = 0; c < R; c += 4) {
          const float4 r4 = __ldg(reinterpret_cast<const float4*>(p.raw + e * p.ldr + c));
          val[GP - 48 + c] = r4.x; val[GP - 48 + c + 1] = r4.y; val[GP - 48 + c + 2] = r4.z; val[GP - 48 + c + 3] = r4.w;
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      float hi_, lo_;
      split_tf32(ok ? val[j] : 0.f, hi_, lo_);
      xh[j] = __float_as_uint(hi_);
      xl[j] = __float_as_uint(lo_);
    }
    tmem_st16(tmem + cs.lane_base + kColXhi + quarter * 16, xh);
    tmem_st16(tmem + cs.lane_base + kColXlo + quarter * 16, xl);
    tmem_st_wait();
  };
  auto issue_mma0 = [&]() {                             // Z0 = X W0'^T (all threads call; one issues)
    if (warp_u == 0 && elect_one_sync()) {
      fence_after_sync();
      mma3_ts<kK0>(tmem + kColZ, tmem + kColXhi, tmem + kColXlo, dW0, kIdesc, false);
      mma_commit(&bar);
    }
  };
  // Software pipeline over the CTA's tiles: X of the next tile is built while the second GEMM of the current one
  // runs (the X columns are free once Z0 is complete), and the next tile's first GEMM is issued as soon as every
  // thread has taken its Z1 columns out of TMEM, so the w epilogue runs under it.
  int64_t e = (int64_t)blockIdx.x * kFilt + cs.row_local;
  bool ok = e < p.E;
  float dist = ok ? __ldg(p.d + e) : 0.f;
  if ((int64_t)blockIdx.x < num_tiles) {
    build_x(e, ok, dist);
    fence_before_sync();
    __syncthreads();
    issue_mma0();
  }
  const int c0 = quarter * 32;
  for (int64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
    F_PROF(7)
    mbar_wait(&bar, parity);                            // Z0 of this tile
    parity ^= 1;
    fence_after_sync();
    F_PROF(3)

    // ---- H = ssp(Z0 + b0): my 32 columns.  They go to TMEM first (the raw value is the hi operand: kind::tf32
    // ignores the 13 low mantissa bits of an A operand in TMEM too, profiles/_scratch/tf32_round_probe.cu), the
    // second GEMM is started, and only then is h written out -- under its MMAs ----
    uint32_t z[32];
    {
      uint32_t lo[32];
      tmem_ld32(tmem + cs.lane_base + kColZ + c0, z);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const float hv = shifted_softplus(__uint_as_float(z[j]) + b0_s[c0 + j]);
        float hi_, lo_;
        split_tf32(hv, hi_, lo_);
        z[j] = __float_as_uint(hv);
        lo[j] = __float_as_uint(lo_);
      }
      tmem_st32(tmem + cs.lane_base + kColHhi + c0, z);
      tmem_st32(tmem + cs.lane_base + kColHlo + c0, lo);
      tmem_st_wait();
    }
    F_PROF(4)
    fence_before_sync();
    __syncthreads();
    if (warp_u == 0 && elect_one_sync()) {
      fence_after_sync();
      mma3_ts<kFilt>(tmem + kColZ, tmem + kColHhi, tmem + kColHlo, dW1, kIdesc, false);  // Z1 = H W1^T
      mma_commit(&bar);
    }
    if (TMA && (p.tma_out & 1)) {
      store_quarter(&tm_h, z, tile);
    } else if (ok) {
      float* hrow = p.h + e * kFilt + c0;
#pragma unroll
      for (int j = 0; j < 32; j += 8)
        st_global_v8(hrow + j,
                     make_float4(__uint_as_float(z[j]), __uint_as_float(z[j + 1]), __uint_as_float(z[j + 2]), __uint_as_float(z[j + 3])),
                     make_float4(__uint_as_float(z[j + 4]), __uint_as_float(z[j + 5]), __uint_as_float(z[j + 6]), __uint_as_float(z[j + 7])));
    }
    const float env = 0.5f * (__cosf(dist * p.pi_over_cutoff) + 1.0f);   // argument in [0, pi]
    const bool ok_w = ok;
    const int64_t e_w = e;
    // the next tile's X, under the second GEMM
    const int64_t next = tile + gridDim.x;
    if (next < num_tiles) {
      e = next * kFilt + cs.row_local;
      ok = e < p.E;
      dist = ok ? __ldg(p.d + e) : 0.f;
      build_x(e, ok, dist);
    }
    F_PROF(5)
    mbar_wait(&bar, parity);                            // Z1 of this tile
    parity ^= 1;
    fence_after_sync();
    F_PROF(6)

    // ---- W = (Z1 + b1) * C ----
    tmem_ld32(tmem + cs.lane_base + kColZ + c0, z);
    tmem_ld_wait();
    fence_before_sync();
    __syncthreads();                                    // everybody holds its Z1 columns, the next X is in TMEM
    if (next < num_tiles) issue_mma0();                 // overwrites the Z columns: runs under the w epilogue
    if (TMA && (p.tma_out & 2)) {
#pragma unroll
      for (int j = 0; j < 32; ++j) z[j] = __float_as_uint((__uint_as_float(z[j]) + b1_s[c0 + j]) * env);
      store_quarter(&tm_w, z, tile);
    } else if (ok_w) {
      float* wrow = p.w + e_w * kFilt + c0;
#pragma unroll
      for (int j = 0; j < 32; j += 8) {
        float o[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = (__uint_as_float(z[j + i]) + b1_s[c0 + j + i]) * env;
        st_global_v8(wrow + j, make_float4(o[0], o[1], o[2], o[3]), make_float4(o[4], o[5], o[6], o[7]));
      }
    }
  }
  F_PROF(7)
  F_PROF_FLUSH
  if (TMA && q_issuer) tma_store_wait();               // my bulk stores are complete before the CTA retires
  cta_teardown(tmem, kCols);
}

}  // namespace tc
}  // namespace hgk
namespace hgk {
namespace tc {

// ================================================================================== backward ===
// dZ1 = gW * C(d);  dH = dZ1 W1;  dZ0 = dH * ssp'(Z0),  ssp'(z) = sigmoid(z) = 1 - exp(-(H + ln 2)).
// One persistent CTA per SM walks 128-edge tiles: dZ1 (hi, lo) goes to TMEM as the A operand,
// W1^T (hi, lo) is resident in shared memory, dZ0 is written to HBM for the weight-gradient kernel.
struct FilterBwdArgs {
  const float* gw;     // [E, 128] gradient of the filter, or NULL: gw[e] = g[gi[e]] * x[xi[e]] formed here
  const float* g; int64_t ldg; const int32_t* gi;     // (the two node tables stay in L2; nothing [E, 128]
  const float* x; int64_t ldx; const int32_t* xi;     //  is read for the filter gradient)
  const float* h;      // [E, 128] saved hidden activations
  const float* d;      // [E]
  float pi_over_cutoff;
  const float* w1t;    // [128, 128] = W1^T (row = input feature, column = output filter)
  float* dz0;          // [E, 128] out
  int64_t E;
};

// TMA = true (a materialised gw): the [128 x 128] gradient tile of the NEXT tile is requested with four
// cp.async.bulk.tensor boxes as soon as every thread has consumed the current one, and lands (128-byte swizzle:
// a thread reads its own row without bank conflicts) while the MMAs and the epilogue run.  Before, each tile began
// with a full DRAM round trip, fetched as row-per-lane 32-byte pieces (one LSU wavefront per 32 bytes).
// (dz0 through shared memory and bulk stores, as in the forward kernel, was tried: 0.155 -> 0.174 ms, the two
// store phases per tile wait for each other; its row-per-thread stores already overlap the next tile's start.)
template <bool TMA>
__global__ void __launch_bounds__(kFiltThreads, 1)
k_schnet_filter_bwd(const FilterBwdArgs p, const __grid_constant__ CUtensorMap tm_gw) {
  extern __shared__ __align__(1024) float smem_raw[];
  float* const smem = smem_raw + (((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u) >> 2);
  __shared__ __align__(8) uint64_t bar;
  __shared__ __align__(8) uint64_t full;               // TMA: the gw tile has landed
  __shared__ uint32_t tmem_slot;
  float* Whi = smem;
  float* Wlo = Whi + kFilt * kFilt;
  float* gbuf = Wlo + kFilt * kFilt;                   // TMA: [4 column blocks][128 rows][32 floats], swizzled
  if (TMA && threadIdx.x == 0) mbar_init(&full, 1);
  constexpr uint32_t kCols = 512;
  constexpr uint32_t kColAhi = 0, kColAlo = 128, kColDH = 256;
  const CtaSetup cs = cta_setup(&bar, &tmem_slot, kCols);     // cs.half = column quarter 0..3
  const uint32_t tmem = cs.tmem;
  const int warp_u = warp_id_uniform();
  const int c0 = cs.half * 32;
  if (threadIdx.x < kThreads) {
    Tile<kFilt, kFilt> t;
    t.load(p.w1t, kFilt, 0, kFilt);
    t.store_rows(Whi, Wlo);
  }
  const OperandDesc dW = operand_desc<kFilt>(Whi, Wlo);
  constexpr uint32_t kIdesc = idesc_tf32(kFilt, kFilt);
  publish_operands();

  uint32_t parity = 0;
  const int64_t num_tiles = (p.E + kFilt - 1) / kFilt;
  auto request = [&](int64_t tile) {                   // one thread
    mbar_expect_tx(&full, (uint32_t)sizeof(float) * kFilt * kFilt);
#pragma unroll
    for (int q = 0; q < 4; ++q) tma_load_2d(gbuf + q * (kFilt * 32), &tm_gw, &full, q * 32, (int32_t)(tile * kFilt));
  };
  if (TMA && threadIdx.x == 0 && (int64_t)blockIdx.x < num_tiles) request(blockIdx.x);
  uint32_t full_parity = 0;
  for (int64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
    const int64_t e = tile * kFilt + cs.row_local;
    const bool ok = e < p.E;
    const float dist = ok ? __ldg(p.d + e) : 0.f;
    const float env = 0.5f * (__cosf(dist * p.pi_over_cutoff) + 1.0f);
    if (TMA) {
      mbar_wait(&full, full_parity);
      full_parity ^= 1;
      // my row's 32 columns: block cs.half, 16-byte chunk j stored at chunk j ^ (row & 7)
      const float* rowp = gbuf + cs.half * (kFilt * 32) + cs.row_local * 32;
      uint32_t hi[32], lo[32];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 g4 = *reinterpret_cast<const float4*>(rowp + ((j ^ (cs.row_local & 7)) << 2));
        const float x[4] = {g4.x * env, g4.y * env, g4.z * env, g4.w * env};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float h_, l_;
          split_tf32(ok ? x[i] : 0.f, h_, l_);
          hi[4 * j + i] = __float_as_uint(h_);
          lo[4 * j + i] = __float_as_uint(l_);
        }
      }
      tmem_st32(tmem + cs.lane_base + kColAhi + c0, hi);
      tmem_st32(tmem + cs.lane_base + kColAlo + c0, lo);
      tmem_st_wait();
    } else {
      const bool gathered = p.gw == nullptr;
      const float* grow = nullptr;
      const float* xrow = nullptr;
      if (ok) {
        if (gathered) {
          grow = p.g + (int64_t)__ldg(p.gi + e) * p.ldg + c0;
          xrow = p.x + (int64_t)__ldg(p.xi + e) * p.ldx + c0;
        } else {
          grow = p.gw + e * kFilt + c0;
        }
      }
      uint32_t hi[32], lo[32];
#pragma unroll
      for (int j = 0; j < 32; j += 8) {
        float4 g0 = make_float4(0.f, 0.f, 0.f, 0.f), g1 = g0;
        if (ok) {
          ld_global_nc_v8(grow + j, g0, g1);
          if (gathered) {
            float4 x0, x1;
            ld_global_nc_v8(xrow + j, x0, x1);
            g0.x *= x0.x; g0.y *= x0.y; g0.z *= x0.z; g0.w *= x0.w;
            g1.x *= x1.x; g1.y *= x1.y; g1.z *= x1.z; g1.w *= x1.w;
          }
        }
        const float x[8] = {g0.x * env, g0.y * env, g0.z * env, g0.w * env, g1.x * env, g1.y * env, g1.z * env, g1.w * env};
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float h_, l_;
          split_tf32(x[i], h_, l_);
          hi[j + i] = __float_as_uint(h_);
          lo[j + i] = __float_as_uint(l_);
        }
      }
      tmem_st32(tmem + cs.lane_base + kColAhi + c0, hi);
      tmem_st32(tmem + cs.lane_base + kColAlo + c0, lo);
      tmem_st_wait();
    }
    fence_before_sync();
    __syncthreads();
    if (warp_u == 0 && elect_one_sync()) {
      fence_after_sync();
      mma3_ts<kFilt>(tmem + kColDH, tmem + kColAhi, tmem + kColAlo, dW, kIdesc, false);   // dH = dZ1 W1
      mma_commit(&bar);
      // every thread has read its row of the gw tile (the barrier above): the buffer takes the next tile's
      if (TMA && tile + gridDim.x < num_tiles) request(tile + gridDim.x);
    }
    // the saved activations of this thread's columns arrive while the MMAs run
    float4 hv[8];
    if (ok) {
      const float* hrow = p.h + e * kFilt + c0;
#pragma unroll
      for (int j = 0; j < 8; j += 2) ld_global_nc_v8(hrow + 4 * j, hv[j], hv[j + 1]);
    }
    mbar_wait(&bar, parity);
    parity ^= 1;
    fence_after_sync();

    uint32_t dh[32];
    tmem_ld32(tmem + cs.lane_base + kColDH + c0, dh);
    tmem_ld_wait();
    if (ok) {
      float* orow = p.dz0 + e * kFilt + c0;
      float4 o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        o[j].x = __uint_as_float(dh[4 * j]) * (1.f - fast_exp2(-(hv[j].x + kLn2) * kLog2e));
        o[j].y = __uint_as_float(dh[4 * j + 1]) * (1.f - fast_exp2(-(hv[j].y + kLn2) * kLog2e));
        o[j].z = __uint_as_float(dh[4 * j + 2]) * (1.f - fast_exp2(-(hv[j].z + kLn2) * kLog2e));
        o[j].w = __uint_as_float(dh[4 * j + 3]) * (1.f - fast_exp2(-(hv[j].w + kLn2) * kLog2e));
      }
#pragma unroll
      for (int j = 0; j < 8; j += 2) st_global_v8(orow + 4 * j, o[j], o[j + 1]);
    }
    // the next tile's barrier after its TMEM stores orders these dH reads before the next MMA
  }
  cta_teardown(tmem, kCols);
}

// ---- weight gradients: dW[n, k] = sum_e A[e, n] B[e, k],  db[n] = sum_e A[e, n] ----------------------
// A = rows of [E, 128] (optionally scaled per row by the cosine envelope of d_e); B = rows of
// [E, KB], or, with GAUSS, the filter network's input [gauss(d_e) | raw_e] rebuilt on the fly.
// Each CTA reduces its share of 64-edge tiles into a TMEM accumulator and writes ONE partial; the
// partials are summed in a fixed order afterwards (deterministic, no atomics).
struct WgradArgs {
  const float* a; int64_t lda;   // [E, 128], row stride lda; with a_idx: A[e] = a[a_idx[e]] * a2[a2_idx[e]]
  const float* scale_d;      // [E] or NULL: multiply row e of A by 0.5 (cos(d_e pi / r_c) + 1)
  float pi_over_cutoff;
  const float* b; int64_t ldb;                      // [E, KB] (GAUSS == false)
  const float* d; const float* raw; int64_t ldr;    // GAUSS == true
  const float* offset; float coeff;
  float* dw_partial;         // [gridDim.y, gridDim.x, 128, KB]
  float* db_partial;         // [gridDim.y, gridDim.x, 128]
  int64_t E;
  // gridDim.y > 1 (hgk_wgrad_rows): block row y works on the 128-column slice y % num_a_slices of A
  // and the KB-column slice y / num_a_slices of B
  int num_a_slices;
  // gathered A operand (CFConv filter gradient formed on the fly): both NULL = plain rows of a
  const int32_t* a_idx; const float* a2; int64_t lda2; const int32_t* a2_idx;
};

template <int KB, bool GAUSS, int G, int R>
__global__ void __launch_bounds__(kThreads, 2)
k_wgrad_tn(const WgradArgs p) {
  constexpr int BE = 32;                               // edges per iteration (the MMA K extent)
  constexpr int GP = (G + 3) / 4 * 4;
  extern __shared__ __align__(1024) float smem_raw[];
  // the MN-major operand tiles need 1024-byte alignment; the launchers request 1024 bytes more than the layout
  float* const smem = smem_raw + (((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u) >> 2);
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  __shared__ float off_s[64];
  __shared__ __align__(16) float cs_s[2][kFilt];
  // Two CTAs share an SM (64 KB of shared memory and <= 128 TMEM columns each): while one waits
  // for its global loads or its MMAs, the other stages its tile.  Both operands are MN-major tiles (rows =
  // edges = the K index, as they lie in global memory): dW = A^T B needs no transposed staging.
  float* AThi = smem;                float* ATlo = AThi + kFilt * BE;
  float* BThi = ATlo + kFilt * BE;   float* BTlo = BThi + KB * BE;
  // Columns [0, KB): the running accumulator of the tensor cores; [KB, 2 KB): fp32 totals.  A tcgen05
  // accumulation chain loses ~2^-24 of the sum per MMA (truncating adds; profiles/precision_probe.py),
  // so every kWgradPromote iterations (48 MMAs) the accumulator is added to the totals by the CUDA cores
  // (round-to-nearest) and restarted.
  constexpr int kWgradPromote = 4;
  constexpr uint32_t kCols = 2 * KB;
  constexpr uint32_t kColTot = KB;
  const CtaSetup cs = cta_setup(&bar, &tmem_slot, kCols);
  const uint32_t tmem = cs.tmem;
  const int warp_u = warp_id_uniform();
  // this thread's KB / 2 columns of the accumulator -> totals (first time: the totals are undefined)
  auto promote = [&](bool first_time) {
#pragma unroll
    for (int c0 = 0; c0 < KB / 2; c0 += 16) {
      const uint32_t col = cs.lane_base + cs.half * (KB / 2) + c0;
      uint32_t a16[16], t16[16];
      tmem_ld16(tmem + col, a16);
      if (!first_time) tmem_ld16(tmem + kColTot + col, t16);
      tmem_ld_wait();
      if (!first_time) {
#pragma unroll
        for (int j = 0; j < 16; ++j) a16[j] = __float_as_uint(__uint_as_float(a16[j]) + __uint_as_float(t16[j]));
      }
      tmem_st16(tmem + kColTot + col, a16);
    }
    tmem_st_wait();
  };
  if (GAUSS && threadIdx.x < 64) off_s[threadIdx.x] = threadIdx.x < G ? p.offset[threadIdx.x] : 0.f;
  __syncthreads();

  const OperandDesc dA = operand_desc_mn<BE>(AThi, ATlo), dB = operand_desc_mn<BE>(BThi, BTlo);
  constexpr uint32_t kIdesc = idesc_tf32(kFilt, KB) | kIdescAMajorMN | kIdescBMajorMN;
  const int a_slice = gridDim.y > 1 ? (int)blockIdx.y % p.num_a_slices : 0;
  const int b_slice = gridDim.y > 1 ? (int)blockIdx.y / p.num_a_slices : 0;
  const float* a_base = p.a + a_slice * kFilt;
  const float* b_base = GAUSS ? nullptr : p.b + b_slice * KB;
  const int64_t part = (int64_t)blockIdx.y * gridDim.x + blockIdx.x;     // this CTA's partial slot
  Tile<BE, kFilt> ta;                                  // next tile, prefetched in registers
  Tile<BE, KB> tb;
  float dsc[Tile<BE, kFilt>::PER];                     // its row-scale arguments (d_e)
  float4 colsum = make_float4(0.f, 0.f, 0.f, 0.f);     // this thread's 4 columns of A, summed over its rows
  const float c2 = p.coeff * kLog2e;

  auto fetch = [&](int64_t tile) {
    const int32_t row0 = (int32_t)(tile * BE);
    const int32_t row_end = (int32_t)p.E;
    if (p.a_idx == nullptr) {
      ta.load(a_base, p.lda, row0, row_end);
    } else {
#pragma unroll
      for (int it = 0; it < Tile<BE, kFilt>::PER; ++it) {
        int r, c;
        Tile<BE, kFilt>::coord(it, r, c);
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row0 + r < row_end) {
          const float4 ga = __ldg(reinterpret_cast<const float4*>(a_base + (int64_t)__ldg(p.a_idx + row0 + r) * p.lda + 4 * c));
          const float4 xa = __ldg(reinterpret_cast<const float4*>(p.a2 + (int64_t)__ldg(p.a2_idx + row0 + r) * p.lda2 + 4 * c));
          v = make_float4(ga.x * xa.x, ga.y * xa.y, ga.z * xa.z, ga.w * xa.w);
        }
        ta.v[it] = v;
      }
    }
    // the row scale is only FETCHED here and applied in stage(): nothing in this function depends
    // on the loads it has just issued
    if (p.scale_d != nullptr) {
#pragma unroll
      for (int it = 0; it < Tile<BE, kFilt>::PER; ++it) {
        int r, c;
        Tile<BE, kFilt>::coord(it, r, c);
        dsc[it] = 0.f;
        if (row0 + r < row_end) asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(dsc[it]) : "l"(p.scale_d + row0 + r));
      }
    }
    if (!GAUSS) {
      tb.load(b_base, p.ldb, row0, row_end);
    } else {
#pragma unroll
      for (int it = 0; it < Tile<BE, KB>::PER; ++it) {
        int r, c;
        Tile<BE, KB>::coord(it, r, c);
        float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
        const int64_t e = (int64_t)row0 + r;
        if (e < p.E) {
          const int k0 = 4 * c;
          if (k0 < G) {
            const float dd = __ldg(p.d + e);
            float g[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float t = dd - off_s[k0 + i];
              g[i] = (k0 + i < G) ? fast_exp2(c2 * (t * t)) : 0.f;
            }
            x = make_float4(g[0], g[1], g[2], g[3]);
          } else if (R > 0 && k0 >= GP && k0 < GP + R) {
            x = __ldg(reinterpret_cast<const float4*>(p.raw + e * p.ldr + (k0 - GP)));
          }
        }
        tb.v[it] = x;
      }
    }
  };
  // registers -> shared memory (hi / lo) and the bias-gradient partial sums of A
  auto stage = [&]() {
    if (p.scale_d != nullptr) {
#pragma unroll
      for (int it = 0; it < Tile<BE, kFilt>::PER; ++it) {
        const float env = 0.5f * (__cosf(dsc[it] * p.pi_over_cutoff) + 1.0f);
        ta.v[it].x *= env; ta.v[it].y *= env; ta.v[it].z *= env; ta.v[it].w *= env;
      }
    }
#pragma unroll
    for (int it = 0; it < Tile<BE, kFilt>::PER; ++it) {
      colsum.x += ta.v[it].x; colsum.y += ta.v[it].y; colsum.z += ta.v[it].z; colsum.w += ta.v[it].w;
    }
    ta.store_mn(AThi, ATlo);
    tb.store_mn(BThi, BTlo);
  };

  const int64_t num_tiles = (p.E + BE - 1) / BE;
  const int64_t stride = gridDim.x;
  uint32_t parity = 0;
  int in_chain = 0;            // iterations accumulated since the accumulator was last restarted
  int promotions = 0;
  if ((int64_t)blockIdx.x < num_tiles) fetch(blockIdx.x);
  for (int64_t tile = blockIdx.x; tile < num_tiles; tile += stride) {
    stage();
    publish_operands();        // also orders the promotion's TMEM reads before the restarting MMA
    if (warp_u == 0 && elect_one_sync()) {
      fence_after_sync();
      mma3_ss_mn<BE>(tmem, dA, dB, kIdesc, in_chain > 0);
      mma_commit(&bar);
    }
    if (tile + stride < num_tiles) fetch(tile + stride);       // in flight during the MMAs
    mbar_wait(&bar, parity);
    parity ^= 1;
    fence_after_sync();
    if (++in_chain == kWgradPromote) {
      promote(promotions == 0);
      ++promotions;
      in_chain = 0;
    }
  }
  if (in_chain > 0) {          // the unfinished chain
    promote(promotions == 0);
    ++promotions;
  }

  // ---- partial dW: lane n, this thread's KB/2 columns ----
  float* out = p.dw_partial + (part * kFilt + cs.row_local) * KB + cs.half * (KB / 2);
  const bool any = (int64_t)blockIdx.x < num_tiles;
#pragma unroll
  for (int c0 = 0; c0 < KB / 2; c0 += 32) {
    uint32_t t[32];
    tmem_ld32(tmem + kColTot + cs.lane_base + cs.half * (KB / 2) + c0, t);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 32; j += 4)
      *reinterpret_cast<float4*>(out + c0 + j) =
          any ? make_float4(__uint_as_float(t[j]), __uint_as_float(t[j + 1]), __uint_as_float(t[j + 2]), __uint_as_float(t[j + 3]))
              : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  // ---- partial db: the 4 lanes that share a 16-byte column add up their rows; threads t and
  // t + 128 hold the same column ----
#pragma unroll
  for (int off = 1; off < 4; off <<= 1) {
    colsum.x += __shfl_xor_sync(0xffffffffu, colsum.x, off);
    colsum.y += __shfl_xor_sync(0xffffffffu, colsum.y, off);
    colsum.z += __shfl_xor_sync(0xffffffffu, colsum.z, off);
    colsum.w += __shfl_xor_sync(0xffffffffu, colsum.w, off);
  }
  if ((threadIdx.x & 3) == 0) {
    const int c = (threadIdx.x >> 2) % (kFilt / 4);
    *reinterpret_cast<float4*>(&cs_s[threadIdx.x >> 7][4 * c]) = colsum;
  }
  __syncthreads();
  if (threadIdx.x < kFilt) p.db_partial[part * kFilt + threadIdx.x] = cs_s[0][threadIdx.x] + cs_s[1][threadIdx.x];
  cta_teardown(tmem, kCols);
}

// ---- the same with TMA loads ------------------------------------------------------------------------
// The streamed operands arrive by cp.async.bulk.tensor straight in the MN-major operand layout (tensor-map
// swizzle 128B_ATOM_32B = the matrix descriptor's SWIZZLE_128B_BASE32B), one iteration ahead, into one of two
// buffers.  kind::tf32 ignores the 13 low mantissa bits of its operands (profiles/_scratch/tf32_round_probe.cu),
// so the raw fp32 tile IS the `hi` operand; the threads only read it back to form lo = x - hi (and the
// bias-gradient sums), half of the staging work of k_wgrad_tn and no global-load latency in the loop.  A row
// scale (cosine envelope) rewrites hi as well; the gaussian B operand is still built by the threads.
template <int KB, bool GAUSS, int G, int R>
__global__ void __launch_bounds__(kThreads, 2)
k_wgrad_tma(const WgradArgs p, const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b) {
  constexpr int BE = 32;                               // edges per iteration (the MMA K extent)
  constexpr int GP = (G + 3) / 4 * 4;
  extern __shared__ __align__(1024) float smem_raw[];
  float* const smem = smem_raw + (((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u) >> 2);
  __shared__ __align__(8) uint64_t bar;                // the iteration's MMAs are complete
  __shared__ __align__(8) uint64_t full[2];            // the TMA loads of buffer b have landed
  __shared__ uint32_t tmem_slot;
  __shared__ float off_s[64];
  __shared__ __align__(16) float cs_s[2][kFilt];
  constexpr int kA = kFilt * BE, kB = KB * BE;         // floats per tile
  float* const Ahi0 = smem;                            // [2][kA]
  float* const Bhi0 = smem + 2 * kA;                   // [2][kB]
  float* const Alo = Bhi0 + 2 * kB;
  float* const Blo = Alo + kA;
  constexpr int kWgradPromote = 4;
  constexpr uint32_t kCols = 2 * KB;
  constexpr uint32_t kColTot = KB;
  if (threadIdx.x == 0) { mbar_init(&full[0], 1); mbar_init(&full[1], 1); }
  const CtaSetup cs = cta_setup(&bar, &tmem_slot, kCols);
  const uint32_t tmem = cs.tmem;
  const int warp_u = warp_id_uniform();
  auto promote = [&](bool first_time) {
#pragma unroll
    for (int c0 = 0; c0 < KB / 2; c0 += 16) {
      const uint32_t col = cs.lane_base + cs.half * (KB / 2) + c0;
      uint32_t a16[16], t16[16];
      tmem_ld16(tmem + col, a16);
      if (!first_time) tmem_ld16(tmem + kColTot + col, t16);
      tmem_ld_wait();
      if (!first_time) {
#pragma unroll
        for (int j = 0; j < 16; ++j) a16[j] = __float_as_uint(__uint_as_float(a16[j]) + __uint_as_float(t16[j]));
      }
      tmem_st16(tmem + kColTot + col, a16);
    }
    tmem_st_wait();
  };
  if (GAUSS && threadIdx.x < 64) off_s[threadIdx.x] = threadIdx.x < G ? p.offset[threadIdx.x] : 0.f;
  __syncthreads();

  const OperandDesc dA0 = operand_desc_mn<BE>(Ahi0, Alo), dA1 = operand_desc_mn<BE>(Ahi0 + kA, Alo);
  const OperandDesc dB0 = operand_desc_mn<BE>(Bhi0, Blo), dB1 = operand_desc_mn<BE>(Bhi0 + kB, Blo);
  constexpr uint32_t kIdesc = idesc_tf32(kFilt, KB) | kIdescAMajorMN | kIdescBMajorMN;
  const int a_slice = gridDim.y > 1 ? (int)blockIdx.y % p.num_a_slices : 0;
  const int b_slice = gridDim.y > 1 ? (int)blockIdx.y / p.num_a_slices : 0;
  const int64_t part = (int64_t)blockIdx.y * gridDim.x + blockIdx.x;     // this CTA's partial slot
  using TA = Tile<BE, kFilt>;
  using TB = Tile<BE, KB>;
  TB tb;                                               // GAUSS: the next B tile, built in registers
  float dsc[TA::PER];                                  // the next tile's row-scale arguments (d_e)
  float4 colsum = make_float4(0.f, 0.f, 0.f, 0.f);     // this thread's 4 columns of A, summed over its rows
  const float c2 = p.coeff * kLog2e;

  const int64_t num_tiles = (p.E + BE - 1) / BE;
  const int64_t stride = gridDim.x;
  // one thread: the boxes of [32 rows x 32 columns] of tile `tile` into buffer b
  auto issue_tma = [&](int64_t tile, int b) {
    const int32_t row0 = (int32_t)(tile * BE);
    mbar_expect_tx(&full[b], (uint32_t)sizeof(float) * (kA + (GAUSS ? 0 : kB)));
#pragma unroll
    for (int blk = 0; blk < kFilt / 32; ++blk)
      tma_load_2d(Ahi0 + b * kA + blk * (BE * 32), &tm_a, &full[b], a_slice * kFilt + blk * 32, row0);
    if (!GAUSS) {
#pragma unroll
      for (int blk = 0; blk < KB / 32; ++blk)
        tma_load_2d(Bhi0 + b * kB + blk * (BE * 32), &tm_b, &full[b], b_slice * KB + blk * 32, row0);
    }
  };
  // what still goes through registers, one tile ahead: the row scale and the gaussian operand
  auto fetch = [&](int64_t tile) {
    const int32_t row0 = (int32_t)(tile * BE);
    const int32_t row_end = (int32_t)p.E;
    if (p.scale_d != nullptr) {
#pragma unroll
      for (int it = 0; it < TA::PER; ++it) {
        int r, c;
        TA::coord(it, r, c);
        dsc[it] = 0.f;
        if (row0 + r < row_end) asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(dsc[it]) : "l"(p.scale_d + row0 + r));
      }
    }
    if (GAUSS) {
#pragma unroll
      for (int it = 0; it < TB::PER; ++it) {
        int r, c;
        TB::coord(it, r, c);
        float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
        const int64_t e = (int64_t)row0 + r;
        if (e < p.E) {
          const int k0 = 4 * c;
          if (k0 < G) {
            const float dd = __ldg(p.d + e);
            float g[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float t = dd - off_s[k0 + i];
              g[i] = (k0 + i < G) ? fast_exp2(c2 * (t * t)) : 0.f;
            }
            x = make_float4(g[0], g[1], g[2], g[3]);
          } else if (R > 0 && k0 >= GP && k0 < GP + R) {
            x = __ldg(reinterpret_cast<const float4*>(p.raw + e * p.ldr + (k0 - GP)));
          }
        }
        tb.v[it] = x;
      }
    }
  };
  // lo tiles (and hi where the threads own it) of the landed tile in buffer b
  auto stage = [&](int b) {
    float* Ahi = Ahi0 + b * kA;
    const bool scaled = p.scale_d != nullptr;
#pragma unroll
    for (int it = 0; it < TA::PER; ++it) {
      int r, c;
      TA::coord(it, r, c);
      const int off = TA::mn_off(r, c);
      float4 v = *reinterpret_cast<const float4*>(Ahi + off);
      if (scaled) {
        const float env = 0.5f * (__cosf(dsc[it] * p.pi_over_cutoff) + 1.0f);
        v.x *= env; v.y *= env; v.z *= env; v.w *= env;
      }
      colsum.x += v.x; colsum.y += v.y; colsum.z += v.z; colsum.w += v.w;
      float4 h, l;
      split_tf32(v.x, h.x, l.x); split_tf32(v.y, h.y, l.y);
      split_tf32(v.z, h.z, l.z); split_tf32(v.w, h.w, l.w);
      if (scaled) *reinterpret_cast<float4*>(Ahi + off) = h;
      *reinterpret_cast<float4*>(Alo + off) = l;
    }
    if (GAUSS) {
      tb.store_mn(Bhi0 + b * kB, Blo);
    } else {
      const float* Bhi = Bhi0 + b * kB;
#pragma unroll
      for (int it = 0; it < TB::PER; ++it) {
        int r, c;
        TB::coord(it, r, c);
        const int off = TB::mn_off(r, c);
        const float4 v = *reinterpret_cast<const float4*>(Bhi + off);
        float4 h, l;
        split_tf32(v.x, h.x, l.x); split_tf32(v.y, h.y, l.y);
        split_tf32(v.z, h.z, l.z); split_tf32(v.w, h.w, l.w);
        *reinterpret_cast<float4*>(Blo + off) = l;
      }
    }
  };

  uint32_t parity = 0;
  int in_chain = 0;            // iterations accumulated since the accumulator was last restarted
  int promotions = 0;
  if ((int64_t)blockIdx.x < num_tiles) {
    if (threadIdx.x == 0) issue_tma(blockIdx.x, 0);
    fetch(blockIdx.x);
  }
  int it = 0;
  for (int64_t tile = blockIdx.x; tile < num_tiles; tile += stride, ++it) {
    const int b = it & 1;
    // the other buffer was last read by the MMAs of the previous iteration, which have completed
    if (threadIdx.x == 0 && tile + stride < num_tiles) issue_tma(tile + stride, b ^ 1);
    mbar_wait(&full[b], (uint32_t)(it >> 1) & 1u);
    stage(b);
    publish_operands();        // also orders the promotion's TMEM reads before the restarting MMA
    if (warp_u == 0 && elect_one_sync()) {
      fence_after_sync();
      mma3_ss_mn<BE>(tmem, b ? dA1 : dA0, b ? dB1 : dB0, kIdesc, in_chain > 0);
      mma_commit(&bar);
    }
    if (tile + stride < num_tiles) fetch(tile + stride);
    mbar_wait(&bar, parity);
    parity ^= 1;
    fence_after_sync();
    if (++in_chain == kWgradPromote) {
      promote(promotions == 0);
      ++promotions;
      in_chain = 0;
    }
  }
  if (in_chain > 0) {          // the unfinished chain
    promote(promotions == 0);
    ++promotions;
  }

  // ---- partial dW: lane n, this thread's KB/2 columns ----
  float* out = p.dw_partial + (part * kFilt + cs.row_local) * KB + cs.half * (KB / 2);
  const bool any = (int64_t)blockIdx.x < num_tiles;
#pragma unroll
  for (int c0 = 0; c0 < KB / 2; c0 += 32) {
    uint32_t t[32];
    tmem_ld32(tmem + kColTot + cs.lane_base + cs.half * (KB / 2) + c0, t);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 32; j += 4)
      *reinterpret_cast<float4*>(out + c0 + j) =
          any ? make_float4(__uint_as_float(t[j]), __uint_as_float(t[j + 1]), __uint_as_float(t[j + 2]), __uint_as_float(t[j + 3]))
              : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  // ---- partial db: as in k_wgrad_tn ----
#pragma unroll
  for (int off = 1; off < 4; off <<= 1) {
    colsum.x += __shfl_xor_sync(0xffffffffu, colsum.x, off);
    colsum.y += __shfl_xor_sync(0xffffffffu, colsum.y, off);
    colsum.z += __shfl_xor_sync(0xffffffffu, colsum.z, off);
    colsum.w += __shfl_xor_sync(0xffffffffu, colsum.w, off);
  }
  if ((threadIdx.x & 3) == 0) {
    const int c = (threadIdx.x >> 2) % (kFilt / 4);
    *reinterpret_cast<float4*>(&cs_s[threadIdx.x >> 7][4 * c]) = colsum;
  }
  __syncthreads();
  if (threadIdx.x < kFilt) p.db_partial[part * kFilt + threadIdx.x] = cs_s[0][threadIdx.x] + cs_s[1][threadIdx.x];
  cta_teardown(tmem, kCols);
}

// Sum of the per-CTA partials of k_wgrad_tn in CTA order (fixed, so the result is reproducible):
// element i of [dw | db] = sum_p partial[p][i].  Consecutive threads read consecutive addresses.
__global__ void __launch_bounds__(256)
k_wgrad_reduce(float* __restrict__ dw, float* __restrict__ db, const float* __restrict__ dw_partial,
               const float* __restrict__ db_partial, int P, int n_w, int n_b) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_w + n_b) return;
  const bool is_w = i < n_w;
  const float* src = is_w ? dw_partial + i : db_partial + (i - n_w);
  const int64_t stride = is_w ? n_w : n_b;
  float acc = 0.f;
  int p = 0;
  for (; p + 8 <= P; p += 8) {
    float v[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) v[q] = __ldg(src + (int64_t)(p + q) * stride);
#pragma unroll
    for (int q = 0; q < 8; ++q) acc += v[q];
  }
  for (; p < P; ++p) acc += __ldg(src + (int64_t)p * stride);
  if (is_w) dw[i] = acc; else db[i - n_w] = acc;
}

// hgk_wgrad_rows: dW[(ia, n), (ib, k)] = sum_p partial[ib * nA + ia][p][n][k] (CTA order), db[(ia, n)] from
// the block rows with ib == 0.
__global__ void __launch_bounds__(256)
k_wgrad_rows_reduce(float* __restrict__ dw, int64_t lddw, float* __restrict__ db, const float* __restrict__ dw_partial,
                    const float* __restrict__ db_partial, int P, int nA, int nB) {
  const int64_t n_w = (int64_t)nA * nB * kFilt * kFilt;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_w + (db != nullptr ? nA * kFilt : 0)) return;
  if (i < n_w) {
    const int slice = (int)(i / (kFilt * kFilt)), rem = (int)(i % (kFilt * kFilt));
    const int ia = slice % nA, ib = slice / nA, n = rem / kFilt, k = rem % kFilt;
    const float* src = dw_partial + (int64_t)slice * P * kFilt * kFilt + rem;
    float acc = 0.f;
    for (int q = 0; q < P; ++q) acc += __ldg(src + (int64_t)q * kFilt * kFilt);
    dw[(int64_t)(ia * kFilt + n) * lddw + ib * kFilt + k] = acc;
  } else {
    const int j = (int)(i - n_w), ia = j / kFilt, n = j % kFilt;
    const float* src = db_partial + (int64_t)ia * P * kFilt + n;      // slices with ib == 0 come first
    float acc = 0.f;
    for (int q = 0; q < P; ++q) acc += __ldg(src + (int64_t)q * kFilt);
    db[j] = acc;
  }
}

}  // namespace tc
}  // namespace hgk

using namespace hgk;

#ifdef HGK_KV_PROFILE
extern "C" int hgk_debug_filter_profile(unsigned long long* out8, int reset) {
  unsigned long long zero[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  cudaDeviceSynchronize();
  if (out8) cudaMemcpyFromSymbol(out8, tc::g_filt_prof, sizeof(zero));
  if (reset) cudaMemcpyToSymbol(tc::g_filt_prof, zero, sizeof(zero));
  return 0;
}
#endif

extern "C" int hgk_schnet_filter_fwd(float* w, float* h, const float* d, const float* raw, int64_t ldr,
                                     int32_t R, const float* offset, int32_t G, float coeff, float cutoff,
                                     const float* w0p, const float* b0, const float* w1, const float* b1,
                                     int64_t E, int32_t F, void* stream) {
  HGK_REQUIRE(E >= 0, "schnet_filter_fwd: negative E");
  if (E == 0) return HGK_OK;
  HGK_REQUIRE(w && h && d && offset && w0p && b0 && w1 && b1, "schnet_filter_fwd: NULL operand");
  HGK_REQUIRE(R == 0 || raw, "schnet_filter_fwd: raw NULL with R > 0");
  if (F != tc::kFilt || G != 50 || !(R == 0 || R == 8))
    return fail(HGK_ERR_UNSUPPORTED, "schnet_filter_fwd: built for F=128, G=50, R in {0, 8} (got F=%d G=%d R=%d)", F, G, R);
  HGK_REQUIRE(R == 0 || (ldr % 4 == 0 && (reinterpret_cast<uintptr_t>(raw) & 15) == 0), "schnet_filter_fwd: raw must be 16-byte aligned rows");
  tc::FilterArgs a{d, raw, ldr, offset, coeff, 3.14159265358979323846f / cutoff, w0p, b0, w1, b1, h, w, E, 3};
  // torch evaluates d * pi / cutoff; keep that operation order exactly in the kernel argument
  a.pi_over_cutoff = 3.14159265358979323846f / cutoff;
  // resident weights (hi, lo) + alignment slack; the TMA variant adds the 32 KB output staging
  // HGK_FILTER_FWD_TMA = 0 (row-per-thread stores) | 1 (h by bulk store) | 2 (w) | 3 (both, default)
  static const int tma_out = [] { const char* e = getenv("HGK_FILTER_FWD_TMA"); return e && e[0] >= '0' && e[0] <= '3' ? e[0] - '0' : 3; }();
  a.tma_out = tma_out;
  CUtensorMap tm_h, tm_w;
  const bool use_tma = tma_out != 0 && tma::rows_map_store16(&tm_h, h, E, tc::kFilt, tc::kFilt, 128) &&
                       tma::rows_map_store16(&tm_w, w, E, tc::kFilt, tc::kFilt, 128);
  if (!use_tma) memset(&tm_h, 0, sizeof(tm_h)), memset(&tm_w, 0, sizeof(tm_w));
  const size_t smem = sizeof(float) * (2 * tc::kFilt * tc::kK0 + 2 * tc::kFilt * tc::kFilt + (use_tma ? 4 * tc::kFilt * 16 : 0)) + 512;
  const int64_t tiles = ceil_div(E, tc::kFilt);
  const unsigned grid = (unsigned)imin64(tiles, kNumSM);
  cudaStream_t st = (cudaStream_t)stream;
  static bool configured[4] = {false, false, false, false};
  auto launch = [&](auto kernel, int slot) -> int {
    if (!configured[slot]) {
      cudaError_t err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (err != cudaSuccess) return fail(HGK_ERR_CUDA, "schnet_filter_fwd: smem opt-in (%zu B): %s", smem, cudaGetErrorString(err));
      configured[slot] = true;
    }
    kernel<<<grid, tc::kFiltThreads, smem, st>>>(a, tm_h, tm_w);
    return check_launch("schnet_filter_fwd");
  };
  if (use_tma) return R == 8 ? launch(tc::k_schnet_filter_fwd<50, 8, true>, 0) : launch(tc::k_schnet_filter_fwd<50, 0, true>, 1);
  return R == 8 ? launch(tc::k_schnet_filter_fwd<50, 8, false>, 2) : launch(tc::k_schnet_filter_fwd<50, 0, false>, 3);
}

static int filter_bwd_launch(const tc::FilterBwdArgs& a, void* stream) {
  // resident W1^T (hi, lo) + alignment slack; the TMA variant adds the [128 x 128] gradient tile buffer
  static const bool want_tma = [] { const char* e = getenv("HGK_FILTER_BWD_TMA"); return !(e && e[0] == '0'); }();
  CUtensorMap tm_gw;
  const bool use_tma = want_tma && a.gw != nullptr && tma::rows_map_128b(&tm_gw, a.gw, a.E, tc::kFilt, tc::kFilt, 128);
  const size_t smem = sizeof(float) * (2 + (use_tma ? 1 : 0)) * tc::kFilt * tc::kFilt + 1024;
  static bool configured[2] = {false, false};
  if (!configured[use_tma]) {
    cudaError_t err = use_tma
        ? cudaFuncSetAttribute(tc::k_schnet_filter_bwd<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)
        : cudaFuncSetAttribute(tc::k_schnet_filter_bwd<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return fail(HGK_ERR_CUDA, "schnet_filter_bwd: smem opt-in: %s", cudaGetErrorString(err));
    configured[use_tma] = true;
  }
  const unsigned grid = (unsigned)imin64(ceil_div(a.E, tc::kFilt), kNumSM);
  if (use_tma) {
    tc::k_schnet_filter_bwd<true><<<grid, tc::kFiltThreads, smem, (cudaStream_t)stream>>>(a, tm_gw);
  } else {
    memset(&tm_gw, 0, sizeof(tm_gw));
    tc::k_schnet_filter_bwd<false><<<grid, tc::kFiltThreads, smem, (cudaStream_t)stream>>>(a, tm_gw);
  }
  return check_launch("schnet_filter_bwd");
}

extern "C" int hgk_schnet_filter_bwd(float* dz0, const float* gw, const float* h, const float* d,
                                     float cutoff, const float* w1t, int64_t E, int32_t F, void* stream) {
  HGK_REQUIRE(E >= 0, "schnet_filter_bwd: negative E");
  if (E == 0) return HGK_OK;
  HGK_REQUIRE(dz0 && gw && h && d && w1t, "schnet_filter_bwd: NULL operand");
  if (F != tc::kFilt) return fail(HGK_ERR_UNSUPPORTED, "schnet_filter_bwd: built for F=128 (got %d)", F);
  tc::FilterBwdArgs a{gw, nullptr, 0, nullptr, nullptr, 0, nullptr, h, d, 3.14159265358979323846f / cutoff, w1t, dz0, E};
  return filter_bwd_launch(a, stream);
}

extern "C" int hgk_cfconv_filter_bwd(float* dz0, const float* g, int64_t ldg, const int32_t* gi, const float* x,
                                     int64_t ldx, const int32_t* xi, const float* h, const float* d, float cutoff,
                                     const float* w1t, int64_t E, int32_t F, void* stream) {
  HGK_REQUIRE(E >= 0, "cfconv_filter_bwd: negative E");
  if (E == 0) return HGK_OK;
  HGK_REQUIRE(dz0 && g && gi && x && xi && h && d && w1t, "cfconv_filter_bwd: NULL operand");
  if (F != tc::kFilt) return fail(HGK_ERR_UNSUPPORTED, "cfconv_filter_bwd: built for F=128 (got %d)", F);
  HGK_REQUIRE(ldg % 8 == 0 && ldx % 8 == 0 && (reinterpret_cast<uintptr_t>(g) & 31) == 0 &&
              (reinterpret_cast<uintptr_t>(x) & 31) == 0, "cfconv_filter_bwd: node rows must be 32-byte aligned");
  tc::FilterBwdArgs a{nullptr, g, ldg, gi, x, ldx, xi, h, d, 3.14159265358979323846f / cutoff, w1t, dz0, E};
  return filter_bwd_launch(a, stream);
}

extern "C" int64_t hgk_wgrad_tn_partials(int64_t E) {
  if (E <= 0) return 0;
  // two CTAs per SM; at least four 32-row tiles per CTA so that small inputs (node rows: 8192) do not write more
  // partial sums than they read operands
  return imin64(ceil_div(E, 128), 2 * (int64_t)kNumSM);
}

static int wgrad_tn_impl(float* dw_partial, float* db_partial, const float* a, int64_t lda, const int32_t* a_idx,
                         const float* a2, int64_t lda2, const int32_t* a2_idx,
                         const float* scale_d, float cutoff, const float* b, int64_t ldb, int32_t KB, const float* d,
                         const float* raw, int64_t ldr, int32_t R, const float* offset, int32_t G,
                         float coeff, int64_t E, int32_t F, void* stream);

extern "C" int hgk_wgrad_tn(float* dw_partial, float* db_partial, const float* a, int64_t lda,
                            const float* scale_d, float cutoff, const float* b, int64_t ldb, int32_t KB, const float* d,
                            const float* raw, int64_t ldr, int32_t R, const float* offset, int32_t G,
                            float coeff, int64_t E, int32_t F, void* stream) {
  return wgrad_tn_impl(dw_partial, db_partial, a, lda, nullptr, nullptr, 0, nullptr, scale_d, cutoff, b, ldb, KB, d, raw,
                       ldr, R, offset, G, coeff, E, F, stream);
}

/* A[e] = g[gi[e]] * x[xi[e]] * envelope(scale_d[e]) formed on the fly, B = h [E, 128] */
extern "C" int hgk_cfconv_wgrad(float* dw_partial, float* db_partial, const float* g, int64_t ldg, const int32_t* gi,
                                const float* x, int64_t ldx, const int32_t* xi, const float* scale_d, float cutoff,
                                const float* b, int64_t ldb, int64_t E, int32_t F, void* stream) {
  HGK_REQUIRE(g && gi && x && xi && b, "cfconv_wgrad: NULL operand");
  HGK_REQUIRE(ldx % 4 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0, "cfconv_wgrad: x rows must be 16-byte aligned");
  return wgrad_tn_impl(dw_partial, db_partial, g, ldg, gi, x, ldx, xi, scale_d, cutoff, b, ldb, 128, nullptr, nullptr, 0,
                       0, nullptr, 0, 0.f, E, F, stream);
}

static bool wgrad_use_tma() {
  static const bool v = [] { const char* e = getenv("HGK_WGRAD_TMA"); return !(e && e[0] == '0'); }();
  return v;
}

static int wgrad_tn_impl(float* dw_partial, float* db_partial, const float* a, int64_t lda, const int32_t* a_idx,
                         const float* a2, int64_t lda2, const int32_t* a2_idx,
                         const float* scale_d, float cutoff, const float* b, int64_t ldb, int32_t KB, const float* d,
                         const float* raw, int64_t ldr, int32_t R, const float* offset, int32_t G,
                         float coeff, int64_t E, int32_t F, void* stream) {
  HGK_REQUIRE(E >= 0, "wgrad_tn: negative E");
  if (E == 0) return HGK_OK;
  HGK_REQUIRE(dw_partial && db_partial && a, "wgrad_tn: NULL operand");
  if (F != tc::kFilt) return fail(HGK_ERR_UNSUPPORTED, "wgrad_tn: built for 128-wide A (got %d)", F);
  HGK_REQUIRE(lda >= F && lda % 4 == 0 && (reinterpret_cast<uintptr_t>(a) & 15) == 0, "wgrad_tn: A rows must be 16-byte aligned");
  const bool gauss = (b == nullptr);
  if (gauss) {
    HGK_REQUIRE(d && offset && (R == 0 || raw), "wgrad_tn: gaussian input needs d / offset / raw");
    if (KB != 64 || G != 50 || !(R == 0 || R == 8))
      return fail(HGK_ERR_UNSUPPORTED, "wgrad_tn: gaussian input built for KB=64, G=50, R in {0,8}");
  } else {
    if (KB != 128) return fail(HGK_ERR_UNSUPPORTED, "wgrad_tn: built for KB=128 (got %d)", KB);
    HGK_REQUIRE(ldb % 4 == 0 && (reinterpret_cast<uintptr_t>(b) & 15) == 0, "wgrad_tn: B rows must be 16-byte aligned");
  }
  tc::WgradArgs args{a, lda, scale_d, 3.14159265358979323846f / cutoff, b, ldb, d, raw, ldr, offset, coeff,
                     dw_partial, db_partial, E, 1, a_idx, a2, lda2, a2_idx};
  const unsigned grid = (unsigned)hgk_wgrad_tn_partials(E);
  cudaStream_t st = (cudaStream_t)stream;
  static bool configured[3] = {false, false, false};      // the smem opt-in is per function: once is enough
  auto launch = [&](auto kernel, int kb, int slot) -> int {
    size_t smem = sizeof(float) * 2 * 32 * (tc::kFilt + kb) + 1024;     // + alignment slack of the MN-major tiles
    if (smem < 80 * 1024) smem = 80 * 1024;    // caps the occupancy at 2 CTAs per SM (register budget)
    if (!configured[slot]) {
      cudaError_t err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (err != cudaSuccess) return fail(HGK_ERR_CUDA, "wgrad_tn: smem opt-in: %s", cudaGetErrorString(err));
      configured[slot] = true;
    }
    kernel<<<grid, tc::kThreads, smem, st>>>(args);
    return check_launch("wgrad_tn");
  };
  // plain streamed rows: the TMA kernel (HGK_WGRAD_TMA=0: the register-staged one, for A/B runs); a gathered A
  // operand, or a driver that will not encode the tensor maps, takes the register-staged kernel
  if (a_idx == nullptr && wgrad_use_tma()) {
    CUtensorMap tm_a, tm_b;
    if (tma::rows_map_32(&tm_a, a, E, F, lda, 32) && (gauss ? true : tma::rows_map_32(&tm_b, b, E, KB, ldb, 32))) {
      if (gauss) tm_b = tm_a;
      static bool configured_tma[3] = {false, false, false};
      auto launch_tma = [&](auto kernel, int kb, int slot) -> int {
        size_t smem = sizeof(float) * 3 * 32 * (tc::kFilt + kb) + 1024;   // two hi buffers + lo, + alignment slack
        if (smem < 80 * 1024) smem = 80 * 1024;
        if (!configured_tma[slot]) {
          cudaError_t err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
          if (err != cudaSuccess) return fail(HGK_ERR_CUDA, "wgrad_tn: smem opt-in: %s", cudaGetErrorString(err));
          configured_tma[slot] = true;
        }
        kernel<<<grid, tc::kThreads, smem, st>>>(args, tm_a, tm_b);
        return check_launch("wgrad_tn(tma)");
      };
      if (!gauss) return launch_tma(tc::k_wgrad_tma<128, false, 50, 0>, 128, 0);
      if (R == 8) return launch_tma(tc::k_wgrad_tma<64, true, 50, 8>, 64, 1);
      return launch_tma(tc::k_wgrad_tma<64, true, 50, 0>, 64, 2);
    }
  }
  if (!gauss) return launch(tc::k_wgrad_tn<128, false, 50, 0>, 128, 0);
  if (R == 8) return launch(tc::k_wgrad_tn<64, true, 50, 8>, 64, 1);
  return launch(tc::k_wgrad_tn<64, true, 50, 0>, 64, 2);
}

extern "C" int hgk_wgrad_reduce(float* dw, float* db, const float* dw_partial, const float* db_partial,
                                int64_t P, int32_t KB, int32_t F, void* stream) {
  HGK_REQUIRE(P >= 0 && KB >= 1 && F >= 1, "wgrad_reduce: bad sizes");
  HGK_REQUIRE(dw && db && (P == 0 || (dw_partial && db_partial)), "wgrad_reduce: NULL operand");
  const int n_w = F * KB, n_b = F;
  const unsigned blocks = (unsigned)ceil_div((int64_t)n_w + n_b, 256);
  tc::k_wgrad_reduce<<<blocks, 256, 0, (cudaStream_t)stream>>>(dw, db, dw_partial, db_partial, (int)P, n_w, n_b);
  return check_launch("wgrad_reduce");
}

// ---- weight / bias gradients of a node-row Linear of any width (multiples of 128) in two launches ----
extern "C" int64_t hgk_wgrad_rows_workspace_floats(int64_t M, int32_t NA, int32_t NB) {
  if (M <= 0 || NA <= 0 || NB <= 0) return 0;
  const int64_t slices = (int64_t)(NA / tc::kFilt) * (NB / tc::kFilt);
  return hgk_wgrad_tn_partials(M) * slices * (tc::kFilt * tc::kFilt + tc::kFilt);
}

extern "C" int hgk_wgrad_rows(float* dw, int64_t lddw, float* db, float* workspace, const float* a, int64_t lda,
                              int32_t NA, const float* b, int64_t ldb, int32_t NB, int64_t M, void* stream) {
  HGK_REQUIRE(M >= 1 && NA >= 1 && NB >= 1, "wgrad_rows: bad sizes");
  if (NA % tc::kFilt != 0 || NB % tc::kFilt != 0)
    return fail(HGK_ERR_UNSUPPORTED, "wgrad_rows: widths must be multiples of 128 (got %d, %d)", NA, NB);
  HGK_REQUIRE(dw && workspace && a && b && lddw >= NB, "wgrad_rows: NULL operand");
  HGK_REQUIRE(lda >= NA && lda % 4 == 0 && ldb >= NB && ldb % 4 == 0 && (reinterpret_cast<uintptr_t>(a) & 15) == 0 &&
              (reinterpret_cast<uintptr_t>(b) & 15) == 0, "wgrad_rows: rows must be 16-byte aligned");
  const int nA = NA / tc::kFilt, nB = NB / tc::kFilt;
  const int64_t P = hgk_wgrad_tn_partials(M);
  float* dw_partial = workspace;
  float* db_partial = workspace + P * nA * nB * tc::kFilt * tc::kFilt;
  tc::WgradArgs args{a, lda, nullptr, 0.f, b, ldb, nullptr, nullptr, 0, nullptr, 0.f, dw_partial, db_partial, M, nA,
                     nullptr, nullptr, 0, nullptr};
  size_t smem = sizeof(float) * 2 * 32 * (tc::kFilt + 128) + 1024;      // + alignment slack of the MN-major tiles
  if (smem < 80 * 1024) smem = 80 * 1024;
  static bool configured = false;
  if (!configured) {
    cudaError_t err = cudaFuncSetAttribute(tc::k_wgrad_tn<128, false, 50, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return fail(HGK_ERR_CUDA, "wgrad_rows: smem opt-in: %s", cudaGetErrorString(err));
    configured = true;
  }
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid((unsigned)P, (unsigned)(nA * nB));
  CUtensorMap tm_a, tm_b;
  if (wgrad_use_tma() && tma::rows_map_32(&tm_a, a, M, NA, lda, 32) && tma::rows_map_32(&tm_b, b, M, NB, ldb, 32)) {
    size_t smem_tma = sizeof(float) * 3 * 32 * (tc::kFilt + 128) + 1024;
    static bool configured_tma = false;
    if (!configured_tma) {
      cudaError_t err = cudaFuncSetAttribute(tc::k_wgrad_tma<128, false, 50, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_tma);
      if (err != cudaSuccess) return fail(HGK_ERR_CUDA, "wgrad_rows: smem opt-in: %s", cudaGetErrorString(err));
      configured_tma = true;
    }
    tc::k_wgrad_tma<128, false, 50, 0><<<grid, tc::kThreads, smem_tma, st>>>(args, tm_a, tm_b);
  } else {
    tc::k_wgrad_tn<128, false, 50, 0><<<grid, tc::kThreads, smem, st>>>(args);
  }
  const int64_t total = (int64_t)nA * nB * tc::kFilt * tc::kFilt + (db ? nA * tc::kFilt : 0);
  tc::k_wgrad_rows_reduce<<<(unsigned)ceil_div(total, 256), 256, 0, st>>>(dw, lddw, db, dw_partial, db_partial, (int)P, nA, nB);
  return check_launch("wgrad_rows");
}
