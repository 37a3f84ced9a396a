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
#include "hgk_common.cuh"
#include "tc_common.cuh"

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

template <int G, int R>
__global__ void __launch_bounds__(kThreads, 1)
k_schnet_filter_fwd(const FilterArgs p) {
  constexpr int GP = (G + 3) / 4 * 4;
  static_assert(GP + R <= kK0 && R % 4 == 0, "input layout must fit 64 columns");
  extern __shared__ __align__(128) float smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  __shared__ float off_s[kK0], b0_s[kFilt], b1_s[kFilt];

  float* W0hi = smem;                 float* W0lo = W0hi + kFilt * kK0;
  float* W1hi = W0lo + kFilt * kK0;   float* W1lo = W1hi + kFilt * kFilt;

  constexpr uint32_t kCols = 512;
  constexpr uint32_t kColXhi = 0, kColXlo = 64, kColZ = 128, kColHhi = 256, kColHlo = 384;
  const CtaSetup cs = cta_setup(&bar, &tmem_slot, kCols);
  const uint32_t tmem = cs.tmem;
  const int warp_u = warp_id_uniform();

  // resident operands: weights (hi, lo) in shared memory, biases and gaussian centres
  {
    Tile<kFilt, kK0> t0;
    t0.load(p.w0, kK0, 0, kFilt);
    t0.store_rows(W0hi, W0lo);
  }
  {
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
  const int64_t num_tiles = (p.E + kFilt - 1) / kFilt;
  for (int64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
    const int64_t e = tile * kFilt + cs.row_local;
    const bool ok = e < p.E;
    const float dist = ok ? __ldg(p.d + e) : 0.f;
    const float c2 = p.coeff * kLog2e;

    // ---- X = [gauss | raw]: this thread's 32 of the 64 columns -> TMEM (hi, lo) ----
    {
      uint32_t xh[32], xl[32];
      float rawv[R > 0 ? R : 1];
      if (R > 0 && cs.half == 1) {
#pragma unroll
        for (int c = 0; c < R; c += 4) {
          float4 r4 = make_float4(0.f, 0.f, 0.f, 0.f);
          if (ok) r4 = __ldg(reinterpret_cast<const float4*>(p.raw + e * p.ldr + c));
          rawv[c] = r4.x; rawv[c + 1] = r4.y; rawv[c + 2] = r4.z; rawv[c + 3] = r4.w;
        }
      }
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        float val = 0.f;
        if (cs.half == 0) {
          if (j < G) { const float t = dist - off_s[j]; val = fast_exp2(c2 * (t * t)); }
        } else {
          const int k = 32 + j;                         // compile-time after unrolling
          if (k < G) { const float t = dist - off_s[k]; val = fast_exp2(c2 * (t * t)); }
          else if (R > 0 && k >= GP && k < GP + R) val = rawv[(k - GP) < R ? (k - GP) : 0];
        }
        if (!ok) val = 0.f;
        float hi_, lo_;
        split_tf32(val, hi_, lo_);
        xh[j] = __float_as_uint(hi_);
        xl[j] = __float_as_uint(lo_);
      }
      tmem_st32(tmem + cs.lane_base + kColXhi + cs.half * 32, xh);
      tmem_st32(tmem + cs.lane_base + kColXlo + cs.half * 32, xl);
      tmem_st_wait();
    }
    fence_before_sync();
    __syncthreads();
    if (warp_u == 0 && elect_one_sync()) {
      fence_after_sync();
      mma3_ts<kK0>(tmem + kColZ, tmem + kColXhi, tmem + kColXlo, dW0, kIdesc, false);   // Z0 = X W0'^T
      mma_commit(&bar);
    }
    mbar_wait(&bar, parity);
    parity ^= 1;
    fence_after_sync();

    // ---- H = ssp(Z0 + b0): my 64 columns, two passes of 32 ----
    float* hrow = p.h + e * kFilt + cs.half * 64;
#pragma unroll
    for (int part = 0; part < 2; ++part) {
      uint32_t z[32], lo[32];
      const int c0 = cs.half * 64 + part * 32;
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
      if (ok) {
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          *reinterpret_cast<float4*>(hrow + part * 32 + j) =
              make_float4(__uint_as_float(z[j]), __uint_as_float(z[j + 1]), __uint_as_float(z[j + 2]), __uint_as_float(z[j + 3]));
      }
#pragma unroll
      for (int j = 0; j < 32; ++j) z[j] &= 0xffffe000u;            // hi part
      tmem_st32(tmem + cs.lane_base + kColHhi + c0, z);
      tmem_st32(tmem + cs.lane_base + kColHlo + c0, lo);
    }
    tmem_st_wait();
    fence_before_sync();
    __syncthreads();
    if (warp_u == 0 && elect_one_sync()) {
      fence_after_sync();
      mma3_ts<kFilt>(tmem + kColZ, tmem + kColHhi, tmem + kColHlo, dW1, kIdesc, false);  // Z1 = H W1^T
      mma_commit(&bar);
    }
    const float env = 0.5f * (__cosf(dist * p.pi_over_cutoff) + 1.0f);   // argument in [0, pi]
    mbar_wait(&bar, parity);
    parity ^= 1;
    fence_after_sync();

    // ---- W = (Z1 + b1) * C ----
    float* wrow = p.w + e * kFilt + cs.half * 64;
#pragma unroll
    for (int part = 0; part < 2; ++part) {
      uint32_t z[32];
      const int c0 = cs.half * 64 + part * 32;
      tmem_ld32(tmem + cs.lane_base + kColZ + c0, z);
      tmem_ld_wait();
      if (ok) {
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          *reinterpret_cast<float4*>(wrow + part * 32 + j) =
              make_float4((__uint_as_float(z[j]) + b1_s[c0 + j]) * env,
                          (__uint_as_float(z[j + 1]) + b1_s[c0 + j + 1]) * env,
                          (__uint_as_float(z[j + 2]) + b1_s[c0 + j + 2]) * env,
                          (__uint_as_float(z[j + 3]) + b1_s[c0 + j + 3]) * env);
      }
    }
    // the next tile's first barrier (after its X store) orders these Z reads before the next MMA
  }
  cta_teardown(tmem, kCols);
}

}  // namespace tc
}  // namespace hgk

using namespace hgk;

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
  tc::FilterArgs a{d, raw, ldr, offset, coeff, 3.14159265358979323846f / cutoff, w0p, b0, w1, b1, h, w, E};
  // torch evaluates d * pi / cutoff; keep that operation order exactly in the kernel argument
  a.pi_over_cutoff = 3.14159265358979323846f / cutoff;
  const size_t smem = sizeof(float) * (2 * tc::kFilt * tc::kK0 + 2 * tc::kFilt * tc::kFilt);
  const int64_t tiles = ceil_div(E, tc::kFilt);
  const unsigned grid = (unsigned)imin64(tiles, kNumSM);
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t err;
  if (R == 8) {
    err = cudaFuncSetAttribute(tc::k_schnet_filter_fwd<50, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return fail(HGK_ERR_CUDA, "schnet_filter_fwd: smem opt-in: %s", cudaGetErrorString(err));
    tc::k_schnet_filter_fwd<50, 8><<<grid, tc::kThreads, smem, st>>>(a);
  } else {
    err = cudaFuncSetAttribute(tc::k_schnet_filter_fwd<50, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return fail(HGK_ERR_CUDA, "schnet_filter_fwd: smem opt-in: %s", cudaGetErrorString(err));
    tc::k_schnet_filter_fwd<50, 0><<<grid, tc::kThreads, smem, st>>>(a);
  }
  return check_launch("schnet_filter_fwd");
}
