// The contraction family (see include/hgk.h): node / edge_w / edge_u partial
// derivatives of  Phi = sum_e sum_c sum_f G[i_e,c,f] X[j_e,c,f] W[e,f] U[e,c].
//
// These kernels replace, for the reference's convolutions,
//   index_select + mul + scatter(sum)     SCFStack.py:285,292-293
//   scatter_add_ / unsorted_segment_*     EGCLStack.py:253,263-271,292-298
//   index_select + mul + index_add_       PAINNStack.py:238-257
//   global_mean_pool / scatter_add        Base.py:478,598
// and all of their first and second derivatives (Base.py:614-620).
//
// All three are HBM/L2-bandwidth kernels (< 1 FLOP per byte): the design rules
// are coalesced 128-bit row accesses, several independent row loads in flight
// per lane, streaming (L1::no_allocate) access to the read-once [E,F] operands
// so the gathered node rows keep L1/L2, and no atomics.
//
// Thread mapping: a row (one node or one edge, F floats per channel) is owned
// by a group of TPR lanes (TPR = 32 for F = 128 with 128-bit accesses), lane
// `sub` holding columns [VW*sub, VW*sub+VW) of every F-chunk.  Node rows are
// reduced sequentially in CSR order by their group, which makes the sum order
// identical to a sequential CPU scatter_add_ over a stable CSR.
#include "hgk_common.cuh"

namespace hgk {

template <int VW> struct Vec;
template <> struct Vec<4> {
  float4 v;
  __device__ __forceinline__ static Vec zero() { return {make_float4(0.f, 0.f, 0.f, 0.f)}; }
  __device__ __forceinline__ static Vec one() { return {make_float4(1.f, 1.f, 1.f, 1.f)}; }
  __device__ __forceinline__ static Vec load(const float* p) { return {__ldg(reinterpret_cast<const float4*>(p))}; }
  __device__ __forceinline__ static Vec load_stream(const float* p) { return {ld_stream4(p)}; }
  __device__ __forceinline__ void store(float* p) const { *reinterpret_cast<float4*>(p) = v; }
  __device__ __forceinline__ void store_stream(float* p) const { st_stream4(p, v); }
  __device__ __forceinline__ Vec mul(const Vec& o) const {
    return {make_float4(mul_rn(v.x, o.v.x), mul_rn(v.y, o.v.y), mul_rn(v.z, o.v.z), mul_rn(v.w, o.v.w))};
  }
  __device__ __forceinline__ Vec mul(float s) const {
    return {make_float4(mul_rn(v.x, s), mul_rn(v.y, s), mul_rn(v.z, s), mul_rn(v.w, s))};
  }
  __device__ __forceinline__ Vec add(const Vec& o) const {
    return {make_float4(add_rn(v.x, o.v.x), add_rn(v.y, o.v.y), add_rn(v.z, o.v.z), add_rn(v.w, o.v.w))};
  }
  // sequential sum of the lanes' own columns (x, then y, z, w)
  __device__ __forceinline__ float hsum() const { return add_rn(add_rn(add_rn(v.x, v.y), v.z), v.w); }
};
template <> struct Vec<1> {
  float v;
  __device__ __forceinline__ static Vec zero() { return {0.f}; }
  __device__ __forceinline__ static Vec one() { return {1.f}; }
  __device__ __forceinline__ static Vec load(const float* p) { return {__ldg(p)}; }
  __device__ __forceinline__ static Vec load_stream(const float* p) { return {ld_stream1(p)}; }
  __device__ __forceinline__ void store(float* p) const { *p = v; }
  __device__ __forceinline__ void store_stream(float* p) const { *p = v; }
  __device__ __forceinline__ Vec mul(const Vec& o) const { return {mul_rn(v, o.v)}; }
  __device__ __forceinline__ Vec mul(float s) const { return {mul_rn(v, s)}; }
  __device__ __forceinline__ Vec add(const Vec& o) const { return {add_rn(v, o.v)}; }
  __device__ __forceinline__ float hsum() const { return v; }
};

template <int TPR>
__device__ __forceinline__ unsigned group_mask(int lane) {
  if constexpr (TPR == 32) {
    return 0xffffffffu;
  } else {
    return ((1u << TPR) - 1u) << ((lane / TPR) * TPR);
  }
}

// ------------------------------------------------------------------ node --
struct NodeArgs {
  float* out; int64_t ldo; int32_t CO;
  const float* A; int64_t lda; int32_t CA;
  const float* W; int64_t ldw;
  const float* U;
  const int32_t* rowptr; const int32_t* perm; const int32_t* gidx;
  int64_t N; int32_t F;
};

template <int VW, int C, int TPR, int UNROLL>
__global__ void __launch_bounds__(256)
k_contract_node(const NodeArgs p) {
  using V = Vec<VW>;
  const int lane = threadIdx.x & 31;
  const int sub = lane % TPR;
  const unsigned mask = group_mask<TPR>(lane);
  const int64_t n = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / TPR;
  if (n >= p.N) return;  // whole groups leave together
  const int32_t beg = __ldg(p.rowptr + n), end = __ldg(p.rowptr + n + 1);
  const int FV = p.F / VW;  // vector columns per channel
  const bool hasA = p.A != nullptr, hasW = p.W != nullptr, hasU = p.U != nullptr;
  const bool a_per_c = (p.CA == C);

  for (int chunk = 0; chunk < FV; chunk += TPR) {
    const int col = (chunk + sub) * VW;
    const bool active = (chunk + sub) < FV;
    V acc[C];
#pragma unroll
    for (int c = 0; c < C; ++c) acc[c] = V::zero();

    for (int32_t base = beg; base < end; base += TPR) {
      int32_t my_e = -1, my_j = 0;
      if (base + sub < end) {
        my_e = p.perm ? __ldg(p.perm + base + sub) : base + sub;
        my_j = p.gidx ? __ldg(p.gidx + my_e) : 0;
      }
      const int cnt = min((int32_t)TPR, end - base);
      for (int t = 0; t < cnt; t += UNROLL) {
        V a[UNROLL][C];
        V w[UNROLL];
        float u[UNROLL][C];
#pragma unroll
        for (int q = 0; q < UNROLL; ++q) {
          const int tt = t + q;
          const int32_t e = __shfl_sync(mask, my_e, tt, TPR);
          const int32_t j = __shfl_sync(mask, my_j, tt, TPR);
          const bool ok = active && (tt < cnt);
          w[q] = V::one();
#pragma unroll
          for (int c = 0; c < C; ++c) { a[q][c] = V::one(); u[q][c] = 1.f; }
          if (ok) {
            if (hasW) w[q] = V::load_stream(p.W + (int64_t)e * p.ldw + col);
            if (hasA) {
              const float* arow = p.A + (int64_t)j * p.lda + col;
              if (a_per_c) {
#pragma unroll
                for (int c = 0; c < C; ++c) a[q][c] = V::load(arow + c * p.F);
              } else {
                a[q][0] = V::load(arow);
#pragma unroll
                for (int c = 1; c < C; ++c) a[q][c] = a[q][0];
              }
            }
            if (hasU) {
#pragma unroll
              for (int c = 0; c < C; ++c) u[q][c] = __ldg(p.U + (int64_t)e * C + c);
            }
          }
        }
#pragma unroll
        for (int q = 0; q < UNROLL; ++q) {
          if (t + q < cnt) {
#pragma unroll
            for (int c = 0; c < C; ++c) {
              V term = hasA ? (hasW ? a[q][c].mul(w[q]) : a[q][c]) : w[q];
              if (hasU) term = term.mul(u[q][c]);
              if (p.CO == C) acc[c] = acc[c].add(term);
              else acc[0] = acc[0].add(term);
            }
          }
        }
      }
    }
    if (active) {
      float* orow = p.out + n * p.ldo + col;
      if (p.CO == C) {
#pragma unroll
        for (int c = 0; c < C; ++c) acc[c].store(orow + c * p.F);
      } else {
        acc[0].store(orow);
      }
    }
  }
}

// ---------------------------------------------------- node, single channel (fast path) --
// The shapes that carry the traffic (CFConv aggregate and its derivative w.r.t. x, segment sums,
// pooling): C = CO = 1, no U.  Operand presence and the permutation are template parameters so the
// inner loop is straight-line: 2 shuffles (1 when edges are already grouped), 2 row loads, 4 mul +
// 4 add per lane and edge.  Same summation order as the generic kernel (sequential in CSR order).
template <int VW, int TPR, bool HAS_A, bool HAS_W, bool HAS_PERM, int UNROLL>
__global__ void __launch_bounds__(256)
k_contract_node_c1(const NodeArgs p) {
  using V = Vec<VW>;
  const int lane = threadIdx.x & 31;
  const int sub = lane % TPR;
  const unsigned mask = group_mask<TPR>(lane);
  const int64_t n = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / TPR;
  if (n >= p.N) return;  // whole groups leave together
  const int32_t beg = __ldg(p.rowptr + n), end = __ldg(p.rowptr + n + 1);
  const int FV = p.F / VW;
  for (int chunk = 0; chunk < FV; chunk += TPR) {
    const int col = (chunk + sub) * VW;
    const bool active = (chunk + sub) < FV;
    const float* wcol = p.W + col;
    const float* acol = p.A + col;
    V acc = V::zero();
    for (int32_t base = beg; base < end; base += TPR) {
      int32_t my_e = base + sub, my_j = 0;
      if (base + sub < end) {
        if (HAS_PERM) my_e = __ldg(p.perm + base + sub);
        if (HAS_A) my_j = __ldg(p.gidx + my_e);
      }
      const int cnt = min((int32_t)TPR, end - base);
      for (int t = 0; t < cnt; t += UNROLL) {
        V a[UNROLL], w[UNROLL];
#pragma unroll
        for (int q = 0; q < UNROLL; ++q) {
          const int tt = t + q;
          const int32_t e = HAS_PERM ? __shfl_sync(mask, my_e, tt, TPR) : base + tt;
          const int32_t j = HAS_A ? __shfl_sync(mask, my_j, tt, TPR) : 0;
          a[q] = V::zero();
          w[q] = V::zero();
          if (active && tt < cnt) {
            if (HAS_W) w[q] = V::load_stream(wcol + (int64_t)e * p.ldw);
            if (HAS_A) a[q] = V::load(acol + (int64_t)j * p.lda);
          }
        }
#pragma unroll
        for (int q = 0; q < UNROLL; ++q) {
          // a past-the-end slot contributes (+0) * (+0) = +0, which leaves the sum unchanged
          const V term = (HAS_A && HAS_W) ? a[q].mul(w[q]) : (HAS_A ? a[q] : w[q]);
          acc = acc.add(term);
        }
      }
    }
    if (active) acc.store(p.out + n * p.ldo + col);
  }
}

// ---------------------------------------------------------------- edge_w --
struct EdgeArgs {
  float* out; int64_t ldout;           // EW [E,F] (ldout) or EU [E,C]
  const float* G; int64_t ldg; int32_t CG; const int32_t* gi;
  const float* X; int64_t ldx; int32_t CX; const int32_t* xi;
  const float* W; int64_t ldw;         // edge_u only
  const float* U;                      // edge_w only
  int64_t E; int32_t F;
};

template <int VW, int C, int TPR>
__global__ void __launch_bounds__(256)
k_contract_edge_w(const EdgeArgs p) {
  using V = Vec<VW>;
  const int lane = threadIdx.x & 31;
  const int sub = lane % TPR;
  const int64_t group = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / TPR;
  const int64_t ngroups = ((int64_t)gridDim.x * blockDim.x) / TPR;
  const int FV = p.F / VW;
  const bool hasG = p.G != nullptr, hasX = p.X != nullptr, hasU = p.U != nullptr;
  const bool g_per_c = (p.CG == C), x_per_c = (p.CX == C);
  // number of distinct (g,x) channel products that actually differ
  const bool any_per_c = (hasG && g_per_c) || (hasX && x_per_c) || hasU;

  for (int64_t e = group; e < p.E; e += ngroups) {
    const int32_t ig = hasG ? __ldg(p.gi + e) : 0;
    const int32_t ix = hasX ? __ldg(p.xi + e) : 0;
    float u[C];
#pragma unroll
    for (int c = 0; c < C; ++c) u[c] = hasU ? __ldg(p.U + e * C + c) : 1.f;
    for (int chunk = sub; chunk < FV; chunk += TPR) {
      const int col = chunk * VW;
      V g[C], x[C];
#pragma unroll
      for (int c = 0; c < C; ++c) { g[c] = V::one(); x[c] = V::one(); }
      if (hasG) {
        const float* grow = p.G + (int64_t)ig * p.ldg + col;
        g[0] = V::load(grow);
#pragma unroll
        for (int c = 1; c < C; ++c) g[c] = g_per_c ? V::load(grow + c * p.F) : g[0];
      }
      if (hasX) {
        const float* xrow = p.X + (int64_t)ix * p.ldx + col;
        x[0] = V::load(xrow);
#pragma unroll
        for (int c = 1; c < C; ++c) x[c] = x_per_c ? V::load(xrow + c * p.F) : x[0];
      }
      V acc;
      {
        V term = hasG ? (hasX ? g[0].mul(x[0]) : g[0]) : x[0];
        if (hasU) term = term.mul(u[0]);
        acc = term;
      }
      if (C > 1) {
        // With no per-channel operand the C terms are equal: still add them
        // (Phi sums over c) so the value matches the definition.
#pragma unroll
        for (int c = 1; c < C; ++c) {
          V term = hasG ? (hasX ? g[c].mul(x[c]) : g[c]) : x[c];
          if (hasU) term = term.mul(u[c]);
          acc = acc.add(term);
        }
        (void)any_per_c;
      }
      acc.store_stream(p.out + e * p.ldout + col);
    }
  }
}

// ---------------------------------------------------------------- edge_u --
template <int VW, int C, int TPR>
__global__ void __launch_bounds__(256)
k_contract_edge_u(const EdgeArgs p) {
  using V = Vec<VW>;
  const int lane = threadIdx.x & 31;
  const int sub = lane % TPR;
  const unsigned mask = group_mask<TPR>(lane);
  const int64_t group = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / TPR;
  const int64_t ngroups = ((int64_t)gridDim.x * blockDim.x) / TPR;
  const int FV = p.F / VW;
  const bool hasG = p.G != nullptr, hasX = p.X != nullptr, hasW = p.W != nullptr;
  const bool g_per_c = (p.CG == C), x_per_c = (p.CX == C);
  // every group runs the same number of iterations (shuffles stay converged)
  const int64_t iters = (p.E + ngroups - 1) / ngroups;
  for (int64_t it = 0; it < iters; ++it) {
    const int64_t e = group + it * ngroups;
    const bool live = e < p.E;
    float part[C];
#pragma unroll
    for (int c = 0; c < C; ++c) part[c] = 0.f;
    if (live) {
      const int32_t ig = hasG ? __ldg(p.gi + e) : 0;
      const int32_t ix = hasX ? __ldg(p.xi + e) : 0;
      for (int chunk = sub; chunk < FV; chunk += TPR) {
        const int col = chunk * VW;
        V w = hasW ? V::load_stream(p.W + e * p.ldw + col) : V::one();
        V g0 = V::one(), x0 = V::one();
        const float* grow = hasG ? p.G + (int64_t)ig * p.ldg + col : nullptr;
        const float* xrow = hasX ? p.X + (int64_t)ix * p.ldx + col : nullptr;
        if (hasG) g0 = V::load(grow);
        if (hasX) x0 = V::load(xrow);
#pragma unroll
        for (int c = 0; c < C; ++c) {
          V g = (hasG && g_per_c && c > 0) ? V::load(grow + c * p.F) : g0;
          V x = (hasX && x_per_c && c > 0) ? V::load(xrow + c * p.F) : x0;
          V term = hasG ? (hasX ? g.mul(x) : g) : x;
          if (hasW) term = term.mul(w);
          part[c] = add_rn(part[c], term.hsum());
        }
      }
    }
#pragma unroll
    for (int c = 0; c < C; ++c) {
#pragma unroll
      for (int off = TPR / 2; off > 0; off >>= 1)
        part[c] = add_rn(part[c], __shfl_xor_sync(mask, part[c], off, TPR));
    }
    if (live && sub == 0) {
#pragma unroll
      for (int c = 0; c < C; ++c) p.out[e * C + c] = part[c];
    }
  }
}

// ------------------------------------------------------- edge_gather_add --
template <int VW, int TPR>
__global__ void __launch_bounds__(256)
k_edge_gather_add(float* __restrict__ out, int64_t ldo, const float* __restrict__ P,
                  int64_t ldp, const int32_t* __restrict__ pi,
                  const float* __restrict__ Q, int64_t ldq,
                  const int32_t* __restrict__ qi, int64_t E, int F) {
  using V = Vec<VW>;
  const int sub = (threadIdx.x & 31) % TPR;
  const int64_t group = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / TPR;
  const int64_t ngroups = ((int64_t)gridDim.x * blockDim.x) / TPR;
  const int FV = F / VW;
  for (int64_t e = group; e < E; e += ngroups) {
    const int32_t ip = __ldg(pi + e);
    const int32_t iq = Q ? __ldg(qi + e) : 0;
    for (int chunk = sub; chunk < FV; chunk += TPR) {
      const int col = chunk * VW;
      V r = V::load(P + (int64_t)ip * ldp + col);
      if (Q) r = r.add(V::load(Q + (int64_t)iq * ldq + col));
      r.store_stream(out + e * ldo + col);
    }
  }
}

// -------------------------------------------------------------- dispatch --
inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
inline int pick_tpr(int fv) {
  int t = 1;
  while (t < fv && t < 32) t <<= 1;
  return t;
}

#define HGK_DISPATCH_TPR(TPRV, ...)                 \
  switch (TPRV) {                                   \
    case 1: { constexpr int TPR = 1; __VA_ARGS__; } break;   \
    case 2: { constexpr int TPR = 2; __VA_ARGS__; } break;   \
    case 4: { constexpr int TPR = 4; __VA_ARGS__; } break;   \
    case 8: { constexpr int TPR = 8; __VA_ARGS__; } break;   \
    case 16: { constexpr int TPR = 16; __VA_ARGS__; } break; \
    default: { constexpr int TPR = 32; __VA_ARGS__; } break; \
  }

template <int VW, bool HAS_A, bool HAS_W>
void launch_node_c1(const NodeArgs& a, cudaStream_t st) {
  const int tpr = pick_tpr(a.F / VW);
  const unsigned blocks = (unsigned)ceil_div(a.N * tpr, 256);
  if (a.perm) {
    HGK_DISPATCH_TPR(tpr, (k_contract_node_c1<VW, TPR, HAS_A, HAS_W, true, (TPR < 8 ? TPR : 8)>
                           <<<blocks, 256, 0, st>>>(a)));
  } else {
    HGK_DISPATCH_TPR(tpr, (k_contract_node_c1<VW, TPR, HAS_A, HAS_W, false, (TPR < 8 ? TPR : 8)>
                           <<<blocks, 256, 0, st>>>(a)));
  }
}

template <int VW, int C>
void launch_node(const NodeArgs& a, cudaStream_t st) {
  if (C == 1 && a.CO == 1 && a.U == nullptr) {
    if (a.A && a.W) return launch_node_c1<VW, true, true>(a, st);
    if (a.W) return launch_node_c1<VW, false, true>(a, st);
    if (a.A) return launch_node_c1<VW, true, false>(a, st);
  }
  const int tpr = pick_tpr(a.F / VW);
  const int64_t threads = a.N * tpr;
  const unsigned blocks = (unsigned)ceil_div(threads, 256);
  constexpr int UNROLL = (C == 1) ? 8 : 4;
  HGK_DISPATCH_TPR(tpr, (k_contract_node<VW, C, TPR, (TPR < UNROLL ? TPR : UNROLL)>
                         <<<blocks, 256, 0, st>>>(a)));
}
template <int VW, int C>
void launch_edge_w(const EdgeArgs& a, cudaStream_t st) {
  const int tpr = pick_tpr(a.F / VW);
  const int64_t want = ceil_div(a.E * tpr, 256);
  const unsigned blocks = (unsigned)imin64(want, (int64_t)kNumSM * 8 * 4);
  HGK_DISPATCH_TPR(tpr, (k_contract_edge_w<VW, C, TPR><<<blocks, 256, 0, st>>>(a)));
}
template <int VW, int C>
void launch_edge_u(const EdgeArgs& a, cudaStream_t st) {
  const int tpr = pick_tpr(a.F / VW);
  const int64_t want = ceil_div(a.E * tpr, 256);
  const unsigned blocks = (unsigned)imin64(want, (int64_t)kNumSM * 8 * 4);
  HGK_DISPATCH_TPR(tpr, (k_contract_edge_u<VW, C, TPR><<<blocks, 256, 0, st>>>(a)));
}

}  // namespace hgk

using namespace hgk;

static bool channels_ok(int c, int C) { return c == 1 || c == C; }

extern "C" int hgk_contract_node(float* out, int64_t ldo, int32_t CO, const float* A,
                                 int64_t lda, int32_t CA, const float* W, int64_t ldw,
                                 const float* U, const int32_t* rowptr,
                                 const int32_t* perm, const int32_t* gidx, int64_t N,
                                 int64_t E, int32_t F, int32_t C, void* stream) {
  HGK_REQUIRE(N >= 0 && E >= 0 && F >= 1, "contract_node: bad sizes N=%lld E=%lld F=%d",
              (long long)N, (long long)E, F);
  if (N == 0) return HGK_OK;
  HGK_REQUIRE(out && rowptr, "contract_node: out/rowptr is NULL");
  if (E == 0) {  // every segment is empty: zeros (gidx / W / U may be NULL-sized)
    HGK_REQUIRE(CO >= 1 && ldo >= (int64_t)CO * F, "contract_node: bad ldo/CO");
    cudaError_t err = cudaMemset2DAsync(out, sizeof(float) * ldo, 0, sizeof(float) * CO * F, N,
                                        (cudaStream_t)stream);
    if (err != cudaSuccess) return fail(HGK_ERR_CUDA, "contract_node: memset: %s", cudaGetErrorString(err));
    return HGK_OK;
  }
  HGK_REQUIRE(C == 1 || C == 3, "contract_node: C must be 1 or 3 (got %d)", C);
  HGK_REQUIRE(channels_ok(CO, C), "contract_node: CO=%d incompatible with C=%d", CO, C);
  HGK_REQUIRE(!A || channels_ok(CA, C), "contract_node: CA=%d incompatible with C=%d", CA, C);
  HGK_REQUIRE(!A || gidx, "contract_node: A given without gidx");
  HGK_REQUIRE(A || W || U, "contract_node: at least one operand required");
  HGK_REQUIRE(U || C == 1 || (A && CA == C), "contract_node: C=%d needs U or a %d-channel A", C, C);
  NodeArgs a{out, ldo, CO, A, lda, A ? CA : 1, W, ldw, U, rowptr, perm, gidx, N, F};
  const bool v4 = (F % 4 == 0) && (ldo % 4 == 0) && aligned16(out) &&
                  (!A || ((lda % 4 == 0) && aligned16(A))) &&
                  (!W || ((ldw % 4 == 0) && aligned16(W)));
  cudaStream_t st = (cudaStream_t)stream;
  if (v4) { if (C == 1) launch_node<4, 1>(a, st); else launch_node<4, 3>(a, st); }
  else    { if (C == 1) launch_node<1, 1>(a, st); else launch_node<1, 3>(a, st); }
  return check_launch("contract_node");
}

extern "C" int hgk_contract_edge_w(float* EW, int64_t ldew, const float* G, int64_t ldg,
                                   int32_t CG, const int32_t* gi, const float* X,
                                   int64_t ldx, int32_t CX, const int32_t* xi,
                                   const float* U, int64_t E, int32_t F, int32_t C,
                                   void* stream) {
  HGK_REQUIRE(E >= 0 && F >= 1, "contract_edge_w: bad sizes E=%lld F=%d", (long long)E, F);
  if (E == 0) return HGK_OK;
  HGK_REQUIRE(EW, "contract_edge_w: EW is NULL");
  HGK_REQUIRE(C == 1 || C == 3, "contract_edge_w: C must be 1 or 3 (got %d)", C);
  HGK_REQUIRE(G || X, "contract_edge_w: G and X both NULL");
  HGK_REQUIRE(!G || (gi && channels_ok(CG, C)), "contract_edge_w: bad G operand");
  HGK_REQUIRE(!X || (xi && channels_ok(CX, C)), "contract_edge_w: bad X operand");
  EdgeArgs a{EW, ldew, G, ldg, G ? CG : 1, gi, X, ldx, X ? CX : 1, xi, nullptr, 0, U, E, F};
  const bool v4 = (F % 4 == 0) && (ldew % 4 == 0) && aligned16(EW) &&
                  (!G || ((ldg % 4 == 0) && aligned16(G))) &&
                  (!X || ((ldx % 4 == 0) && aligned16(X)));
  cudaStream_t st = (cudaStream_t)stream;
  if (v4) { if (C == 1) launch_edge_w<4, 1>(a, st); else launch_edge_w<4, 3>(a, st); }
  else    { if (C == 1) launch_edge_w<1, 1>(a, st); else launch_edge_w<1, 3>(a, st); }
  return check_launch("contract_edge_w");
}

extern "C" int hgk_contract_edge_u(float* EU, const float* G, int64_t ldg, int32_t CG,
                                   const int32_t* gi, const float* X, int64_t ldx,
                                   int32_t CX, const int32_t* xi, const float* W,
                                   int64_t ldw, int64_t E, int32_t F, int32_t C,
                                   void* stream) {
  HGK_REQUIRE(E >= 0 && F >= 1, "contract_edge_u: bad sizes E=%lld F=%d", (long long)E, F);
  if (E == 0) return HGK_OK;
  HGK_REQUIRE(EU, "contract_edge_u: EU is NULL");
  HGK_REQUIRE(C == 1 || C == 3, "contract_edge_u: C must be 1 or 3 (got %d)", C);
  HGK_REQUIRE(G || X, "contract_edge_u: G and X both NULL");
  HGK_REQUIRE(!G || (gi && channels_ok(CG, C)), "contract_edge_u: bad G operand");
  HGK_REQUIRE(!X || (xi && channels_ok(CX, C)), "contract_edge_u: bad X operand");
  EdgeArgs a{EU, C, G, ldg, G ? CG : 1, gi, X, ldx, X ? CX : 1, xi, W, ldw, nullptr, E, F};
  const bool v4 = (F % 4 == 0) && (!G || ((ldg % 4 == 0) && aligned16(G))) &&
                  (!X || ((ldx % 4 == 0) && aligned16(X))) &&
                  (!W || ((ldw % 4 == 0) && aligned16(W)));
  cudaStream_t st = (cudaStream_t)stream;
  if (v4) { if (C == 1) launch_edge_u<4, 1>(a, st); else launch_edge_u<4, 3>(a, st); }
  else    { if (C == 1) launch_edge_u<1, 1>(a, st); else launch_edge_u<1, 3>(a, st); }
  return check_launch("contract_edge_u");
}

extern "C" int hgk_edge_gather_add(float* out, int64_t ldo, const float* P, int64_t ldp,
                                   const int32_t* pi, const float* Q, int64_t ldq,
                                   const int32_t* qi, int64_t E, int32_t F, void* stream) {
  HGK_REQUIRE(E >= 0 && F >= 1, "edge_gather_add: bad sizes E=%lld F=%d", (long long)E, F);
  if (E == 0) return HGK_OK;
  HGK_REQUIRE(out && P && pi, "edge_gather_add: out/P/pi is NULL");
  HGK_REQUIRE(!Q || qi, "edge_gather_add: Q given without qi");
  const bool v4 = (F % 4 == 0) && (ldo % 4 == 0) && (ldp % 4 == 0) && aligned16(out) &&
                  aligned16(P) && (!Q || ((ldq % 4 == 0) && aligned16(Q)));
  cudaStream_t st = (cudaStream_t)stream;
  const int tpr = pick_tpr(v4 ? F / 4 : F);
  const unsigned blocks =
      (unsigned)imin64(ceil_div(E * tpr, 256), (int64_t)kNumSM * 8 * 4);
  if (v4) {
    HGK_DISPATCH_TPR(tpr, (k_edge_gather_add<4, TPR><<<blocks, 256, 0, st>>>(out, ldo, P, ldp, pi, Q, ldq, qi, E, F)));
  } else {
    HGK_DISPATCH_TPR(tpr, (k_edge_gather_add<1, TPR><<<blocks, 256, 0, st>>>(out, ldo, P, ldp, pi, Q, ldq, qi, E, F)));
  }
  return check_launch("edge_gather_add");
}

// --------------------------------------------------------------------------- column sums ---
// out[c] = sum_n x[n, c] (the bias gradient of a node-row Linear).  torch's column reduction runs
// 128-block grids for 20 us on a [8192, 128] matrix; here S row slabs x C/32 column groups of CTAs
// write partials (warp = rows, lane = column: 128-byte coalesced reads; the eight warps of a CTA are
// combined in warp order), and a second tiny kernel adds the S partials in slab order: fixed
// summation order, reproducible.
namespace hgk {
__global__ void __launch_bounds__(256)
k_colsum_partial(float* __restrict__ partial, const float* __restrict__ x, int64_t ldx, int64_t N, int C,
                 int64_t rows_per_slab) {
  __shared__ float red[8][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  const int64_t beg = (int64_t)blockIdx.y * rows_per_slab;
  const int64_t end = beg + rows_per_slab < N ? beg + rows_per_slab : N;
  float acc = 0.f;
  if (c < C)
    for (int64_t n = beg + warp; n < end; n += 8) acc += __ldg(x + n * ldx + c);
  red[warp][lane] = acc;
  __syncthreads();
  if (warp == 0 && c < C) {
    float t = red[0][lane];
#pragma unroll
    for (int w = 1; w < 8; ++w) t += red[w][lane];
    partial[(int64_t)blockIdx.y * C + c] = t;
  }
}
__global__ void __launch_bounds__(256)
k_colsum_final(float* __restrict__ out, const float* __restrict__ partial, int S, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float t = 0.f;
  for (int s = 0; s < S; ++s) t += partial[(int64_t)s * C + c];
  out[c] = t;
}
}  // namespace hgk

extern "C" int64_t hgk_colsum_slabs(int64_t N) {
  if (N <= 0) return 1;
  return imin64(ceil_div(N, 128), 64);
}

extern "C" int hgk_colsum(float* out, float* partial, const float* x, int64_t ldx, int64_t N, int32_t C,
                          void* stream) {
  HGK_REQUIRE(N >= 0 && C >= 1 && ldx >= C, "colsum: bad sizes N=%lld C=%d", (long long)N, C);
  HGK_REQUIRE(out && partial && (N == 0 || x), "colsum: NULL operand");
  const int S = (int)hgk_colsum_slabs(N);
  const int64_t rows = N > 0 ? ceil_div(N, S) : 1;
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid((unsigned)ceil_div(C, 32), (unsigned)S);
  k_colsum_partial<<<grid, 256, 0, st>>>(partial, x, ldx, N, C, rows);
  k_colsum_final<<<(unsigned)ceil_div(C, 256), 256, 0, st>>>(out, partial, S, C);
  return check_launch("colsum");
}
