// PNA multi-aggregation: mean / min / max / std of the per-edge message
// h[e,:] = (P[dst(e),:] + Q[src[e],:] (+ R[e,:])) (* M[e,:]) over each destination's edges.
// The optional per-edge multiplier M is PNAPlus's Hadamard product with the projected radial basis
// (hydragnn/models/PNAPlusStack.py:275-283).
//
// Replaces PyG PNAConv.propagate + DegreeScalerAggregation as configured by
// the reference (hydragnn/models/PNAStack.py:30-53: aggregators mean,min,max,
// std; pre_layers=1 so the message is one Linear on cat[x_i, x_j(, e)] which
// splits into node-level products P, Q (and an edge-level R)).  h is formed in
// registers: the [E,F] message tensor and the four scatter passes of the
// reference are never materialised.
//
// Semantics kept from PyG 2.5.2 / torch.scatter_reduce:
//   mean = sum/deg;  std = sqrt(clamp(E[h^2]-E[h]^2, 1e-5)), set to 0 where
//   <= sqrt(1e-5);  empty segment -> 0 for all four;
//   backward of min/max splits the gradient evenly among tied entries.
#include "hgk_common.cuh"

namespace hgk {

constexpr float kVarEps = 1e-5f;

template <int TPR>
__global__ void __launch_bounds__(256)
k_pna_fwd(float* __restrict__ out, int32_t* __restrict__ cntmin, int32_t* __restrict__ cntmax,
          const float* __restrict__ P, const float* __restrict__ Q, const float* __restrict__ R,
          const float* __restrict__ M,
          const int32_t* __restrict__ rowptr, const int32_t* __restrict__ perm,
          const int32_t* __restrict__ src, int64_t N, int F) {
  const int sub = (threadIdx.x & 31) % TPR;
  const int64_t n = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / TPR;
  if (n >= N) return;
  const int32_t beg = rowptr[n], end = rowptr[n + 1];
  const int32_t deg = end - beg;
  const float sqrt_eps = sqrtf(kVarEps);
  for (int f = sub; f < F; f += TPR) {
    const float pn = P ? P[n * F + f] : 0.f;
    float sum = 0.f, sumsq = 0.f, mn = INFINITY, mx = -INFINITY;
    int32_t cmn = 0, cmx = 0;
    for (int32_t k = beg; k < end; ++k) {
      const int32_t e = perm ? perm[k] : k;
      float h = Q ? add_rn(pn, Q[(int64_t)src[e] * F + f]) : pn;
      if (R) h = add_rn(h, R[(int64_t)e * F + f]);
      if (M) h = mul_rn(h, M[(int64_t)e * F + f]);
      sum = add_rn(sum, h);
      sumsq = add_rn(sumsq, mul_rn(h, h));
      if (h < mn) { mn = h; cmn = 1; } else if (h == mn) { ++cmn; }
      if (h > mx) { mx = h; cmx = 1; } else if (h == mx) { ++cmx; }
    }
    float mean = 0.f, sd = 0.f;
    if (deg > 0) {
      const float cnt = (float)deg;
      mean = sum / cnt;
      const float mean2 = sumsq / cnt;
      float var = add_rn(mean2, -mul_rn(mean, mean));
      var = fmaxf(var, kVarEps);
      sd = sqrtf(var);
      if (sd <= sqrt_eps) sd = 0.f;
    } else {
      mn = 0.f; mx = 0.f;
    }
    float* o = out + n * 4 * F;
    o[f] = mean; o[F + f] = mn; o[2 * F + f] = mx; o[3 * F + f] = sd;
    cntmin[n * F + f] = cmn;
    cntmax[n * F + f] = cmx;
  }
}

// Backward, destination-centric like the forward: the eleven per-node rows (gout x4, out x4, cntmin,
// cntmax, P) are read once per destination and kept in registers while its edges stream by, instead
// of once per edge (which made the kernel L2-bandwidth bound: 11 F floats per edge).
template <int TPR>
__global__ void __launch_bounds__(256)
k_pna_bwd(float* __restrict__ gh, float* __restrict__ gm, const float* __restrict__ gout,
          const float* __restrict__ out, const int32_t* __restrict__ cntmin, const int32_t* __restrict__ cntmax,
          const float* __restrict__ P, const float* __restrict__ Q, const float* __restrict__ R,
          const float* __restrict__ M, const int32_t* __restrict__ rowptr, const int32_t* __restrict__ perm,
          const int32_t* __restrict__ src, int64_t N, int F) {
  const int sub = (threadIdx.x & 31) % TPR;
  const int64_t n = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / TPR;
  if (n >= N) return;
  const int32_t beg = rowptr[n], end = rowptr[n + 1];
  if (end == beg) return;
  const float cnt = (float)(end - beg);
  const float* o = out + n * 4 * F;
  const float* g = gout + n * 4 * F;
  for (int f = sub; f < F; f += TPR) {
    const float pn = P ? P[n * F + f] : 0.f;
    const float mean = o[f], omin = o[F + f], omax = o[2 * F + f], sd = o[3 * F + f];
    const float g_mean = g[f] / cnt;
    const float g_min = g[F + f] / (float)cntmin[n * F + f];
    const float g_max = g[2 * F + f] / (float)cntmax[n * F + f];
    const float g_sd = g[3 * F + f];
    const float den = cnt * sd;
    for (int32_t k = beg; k < end; ++k) {
      const int32_t e = perm ? perm[k] : k;
      float h = Q ? add_rn(pn, Q[(int64_t)src[e] * F + f]) : pn;
      if (R) h = add_rn(h, R[(int64_t)e * F + f]);
      const float pre = h;
      float m = 1.f;
      if (M) { m = M[(int64_t)e * F + f]; h = mul_rn(h, m); }
      float acc = g_mean;
      if (h == omin) acc += g_min;
      if (h == omax) acc += g_max;
      if (sd > 0.f) acc += g_sd * (h - mean) / den;
      if (M) {                                   // h = pre * m: hand back both factors' gradients
        gm[(int64_t)e * F + f] = acc * pre;
        acc *= m;
      }
      gh[(int64_t)e * F + f] = acc;
    }
  }
}

inline int pna_tpr(int F) {
  int t = 1;
  while (t < F && t < 32) t <<= 1;
  return t;
}

}  // namespace hgk

using namespace hgk;

#define HGK_PNA_DISPATCH(TPRV, ...)                          \
  switch (TPRV) {                                            \
    case 1: { constexpr int TPR = 1; __VA_ARGS__; } break;   \
    case 2: { constexpr int TPR = 2; __VA_ARGS__; } break;   \
    case 4: { constexpr int TPR = 4; __VA_ARGS__; } break;   \
    case 8: { constexpr int TPR = 8; __VA_ARGS__; } break;   \
    case 16: { constexpr int TPR = 16; __VA_ARGS__; } break; \
    default: { constexpr int TPR = 32; __VA_ARGS__; } break; \
  }

extern "C" int hgk_pna_aggregate_fwd(float* out, int32_t* cntmin, int32_t* cntmax,
                                     const float* P, const float* Q, const float* R, const float* M,
                                     const int32_t* rowptr, const int32_t* perm,
                                     const int32_t* src, int64_t N, int64_t E, int32_t F,
                                     void* stream) {
  HGK_REQUIRE(N >= 0 && E >= 0 && F >= 1, "pna_aggregate_fwd: bad sizes");
  if (N == 0) return HGK_OK;
  HGK_REQUIRE(out && cntmin && cntmax && rowptr, "pna_aggregate_fwd: NULL operand");
  HGK_REQUIRE(P || Q || R, "pna_aggregate_fwd: P, Q and R all NULL");
  HGK_REQUIRE(E == 0 || !Q || src, "pna_aggregate_fwd: src NULL");
  const int tpr = pna_tpr(F);
  const unsigned blocks = (unsigned)ceil_div(N * tpr, 256);
  cudaStream_t st = (cudaStream_t)stream;
  HGK_PNA_DISPATCH(tpr, (k_pna_fwd<TPR><<<blocks, 256, 0, st>>>(out, cntmin, cntmax, P, Q, R, M, rowptr, perm, src, N, F)));
  return check_launch("pna_aggregate_fwd");
}

extern "C" int hgk_pna_aggregate_bwd(float* gh, float* gm, const float* gout, const float* out,
                                     const int32_t* cntmin, const int32_t* cntmax,
                                     const float* P, const float* Q, const float* R, const float* M,
                                     const int32_t* rowptr, const int32_t* perm,
                                     const int32_t* src, int64_t N, int64_t E, int32_t F,
                                     void* stream) {
  HGK_REQUIRE(N >= 0 && E >= 0 && F >= 1, "pna_aggregate_bwd: bad sizes");
  if (E == 0 || N == 0) return HGK_OK;
  HGK_REQUIRE(gh && gout && out && cntmin && cntmax && rowptr && (src || !Q) && (P || Q || R),
              "pna_aggregate_bwd: NULL operand");
  HGK_REQUIRE(!M || gm, "pna_aggregate_bwd: M given without gm");
  const int tpr = pna_tpr(F);
  const unsigned blocks = (unsigned)ceil_div(N * tpr, 256);
  cudaStream_t st = (cudaStream_t)stream;
  HGK_PNA_DISPATCH(tpr, (k_pna_bwd<TPR><<<blocks, 256, 0, st>>>(gh, gm, gout, out, cntmin, cntmax, P, Q, R, M, rowptr, perm, src, N, F)));
  return check_launch("pna_aggregate_bwd");
}
