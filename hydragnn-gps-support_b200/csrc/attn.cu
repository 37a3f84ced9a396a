// GPS global attention over variable-length node segments (fp32, SIMT path).
//
// Replaces hydragnn/globalAtt/gps.py:126-133 of the reference:
//   to_dense_batch -> nn.MultiheadAttention core (softmax(QK^T/sqrt(d)) V) -> unpad
// with a streaming (online-softmax) kernel that never pads and never stores
// the [L,L] score matrix.  Segments are given by cu_seqlens; S = 1 reproduces
// the reference's whole-batch attention, S = num_graphs is the per-graph mode.
//
// Tiling: one CTA (256 threads as a 16x16 grid) owns a 64-row tile; thread
// (ty,tx) owns rows ty+16i and columns tx+16j (i,j < 4) of every 64x64 score
// tile, and rows ty+16i / feature columns tx+16jj of the [64,D] accumulators.
// Shared-memory rows are padded to D+4 floats so that the 128-bit operand
// reads of the interleaved rows are bank-conflict free.
// The backward is split in two deterministic kernels (no float atomics):
//   k_attn_bwd_kv : CTA owns a key/value tile, loops over query tiles -> dK,dV
//   k_attn_bwd_q  : CTA owns a query tile, loops over key tiles      -> dQ
#include "hgk_common.cuh"
#include <stdlib.h>

namespace hgk {

constexpr int kTile = 64;
constexpr int kAttnThreads = 256;
constexpr int kPS = kTile + 4;  // stride of the 64x64 probability tile

struct AttnArgs {
  const float* q; int64_t ldq;
  const float* k; int64_t ldk;
  const float* v; int64_t ldv;
  const int32_t* cu; int32_t S; int64_t N;
  int32_t H, D, Dp;
  float scale;
};

// Map a linear tile id to (segment start, segment end, first row of the tile).
__device__ __forceinline__ bool locate_tile(const int32_t* cu, int32_t S, int32_t tile,
                                            int32_t* seg_beg, int32_t* seg_end, int32_t* row0) {
  int32_t acc = 0;
  for (int32_t s = 0; s < S; ++s) {
    const int32_t b = cu[s], e = cu[s + 1];
    const int32_t nt = (e - b + kTile - 1) / kTile;
    if (tile < acc + nt) {
      *seg_beg = b; *seg_end = e; *row0 = b + (tile - acc) * kTile;
      return true;
    }
    acc += nt;
  }
  return false;
}

// smem tile [64][Dp+4] <- rows row0.. of a [N, H*D]-strided matrix, head offset folded in `src`.
__device__ __forceinline__ void load_tile(float* dst, const float* src, int64_t ld, int32_t row0,
                                          int32_t row_end, int32_t D, int32_t Dp) {
  const int lds = Dp + 4;
  for (int idx = threadIdx.x; idx < kTile * Dp; idx += kAttnThreads) {
    const int r = idx / Dp, d = idx - r * Dp;
    float val = 0.f;
    if (row0 + r < row_end && d < D) val = __ldg(src + (int64_t)(row0 + r) * ld + d);
    dst[r * lds + d] = val;
  }
}

// acc[i][j] += sum_d A[ty+16i][d] * B[tx+16j][d]
__device__ __forceinline__ void mma_nt(float acc[4][4], const float* A, const float* B, int Dp,
                                       int ty, int tx) {
  const int lds = Dp + 4;
  for (int d = 0; d < Dp; d += 4) {
    float4 a[4], b[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) a[i] = *reinterpret_cast<const float4*>(A + (ty + 16 * i) * lds + d);
#pragma unroll
    for (int j = 0; j < 4; ++j) b[j] = *reinterpret_cast<const float4*>(B + (tx + 16 * j) * lds + d);
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        acc[i][j] = fmaf(a[i].x, b[j].x, acc[i][j]);
        acc[i][j] = fmaf(a[i].y, b[j].y, acc[i][j]);
        acc[i][j] = fmaf(a[i].z, b[j].z, acc[i][j]);
        acc[i][j] = fmaf(a[i].w, b[j].w, acc[i][j]);
      }
  }
}

// acc[i][jj] += sum_c P[ty+16i][c] * V[c][tx+16jj]        (P: [64][kPS], V: [64][Dp+4])
template <int JJ>
__device__ __forceinline__ void mma_pv(float acc[4][JJ], const float* P, const float* V, int Dp,
                                       int ty, int tx) {
  const int lds = Dp + 4;
  for (int c = 0; c < kTile; c += 4) {
    float4 p[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) p[i] = *reinterpret_cast<const float4*>(P + (ty + 16 * i) * kPS + c);
#pragma unroll
    for (int cc = 0; cc < 4; ++cc) {
      float vv[JJ];
#pragma unroll
      for (int jj = 0; jj < JJ; ++jj) {
        const int col = tx + 16 * jj;
        vv[jj] = (col < Dp) ? V[(c + cc) * lds + col] : 0.f;
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float pv = cc == 0 ? p[i].x : (cc == 1 ? p[i].y : (cc == 2 ? p[i].z : p[i].w));
#pragma unroll
        for (int jj = 0; jj < JJ; ++jj) acc[i][jj] = fmaf(pv, vv[jj], acc[i][jj]);
      }
    }
  }
}

// acc[i][jj] += sum_r P[r][ty+16i] * X[r][tx+16jj]        (transposed use of P)
template <int JJ>
__device__ __forceinline__ void mma_ptx(float acc[4][JJ], const float* P, const float* X, int Dp,
                                        int ty, int tx) {
  const int lds = Dp + 4;
  for (int r = 0; r < kTile; ++r) {
    float pv[4], xv[JJ];
#pragma unroll
    for (int i = 0; i < 4; ++i) pv[i] = P[r * kPS + ty + 16 * i];
#pragma unroll
    for (int jj = 0; jj < JJ; ++jj) {
      const int col = tx + 16 * jj;
      xv[jj] = (col < Dp) ? X[r * lds + col] : 0.f;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int jj = 0; jj < JJ; ++jj) acc[i][jj] = fmaf(pv[i], xv[jj], acc[i][jj]);
  }
}

__device__ __forceinline__ float row_reduce_max(float v) {
#pragma unroll
  for (int off = 8; off > 0; off >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, off));
  return v;
}
__device__ __forceinline__ float row_reduce_sum(float v) {
#pragma unroll
  for (int off = 8; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  return v;
}

// ------------------------------------------------------------- forward --
template <int JJ>
__global__ void __launch_bounds__(kAttnThreads)
k_attn_fwd(const AttnArgs p, float* __restrict__ o, int64_t ldo, float* __restrict__ lse) {
  extern __shared__ __align__(16) float smem[];
  int32_t seg_beg, seg_end, row0;
  if (!locate_tile(p.cu, p.S, blockIdx.x, &seg_beg, &seg_end, &row0)) return;
  const int h = blockIdx.y;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int lds = p.Dp + 4;
  float* Qs = smem;
  float* Ks = Qs + kTile * lds;
  float* Vs = Ks + kTile * lds;
  float* Ps = Vs + kTile * lds;

  load_tile(Qs, p.q + h * p.D, p.ldq, row0, seg_end, p.D, p.Dp);
  float m[4], l[4], acc[4][JJ];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    m[i] = -INFINITY; l[i] = 0.f;
#pragma unroll
    for (int jj = 0; jj < JJ; ++jj) acc[i][jj] = 0.f;
  }
  for (int32_t k0 = seg_beg; k0 < seg_end; k0 += kTile) {
    __syncthreads();  // previous tile fully consumed (also orders the Q load)
    load_tile(Ks, p.k + h * p.D, p.ldk, k0, seg_end, p.D, p.Dp);
    load_tile(Vs, p.v + h * p.D, p.ldv, k0, seg_end, p.D, p.Dp);
    __syncthreads();
    float s[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) s[i][j] = 0.f;
    mma_nt(s, Qs, Ks, p.Dp, ty, tx);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float mx = -INFINITY;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const bool valid = (k0 + tx + 16 * j) < seg_end;
        s[i][j] = valid ? s[i][j] * p.scale : -INFINITY;
        mx = fmaxf(mx, s[i][j]);
      }
      mx = row_reduce_max(mx);
      const float m_new = fmaxf(m[i], mx);  // finite: every key tile has >= 1 valid key
      const float corr = expf(m[i] - m_new);
      float rs = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float pij = expf(s[i][j] - m_new);
        Ps[(ty + 16 * i) * kPS + tx + 16 * j] = pij;
        rs += pij;
      }
      rs = row_reduce_sum(rs);
      l[i] = l[i] * corr + rs;
      m[i] = m_new;
#pragma unroll
      for (int jj = 0; jj < JJ; ++jj) acc[i][jj] *= corr;
    }
    __syncthreads();
    mma_pv<JJ>(acc, Ps, Vs, p.Dp, ty, tx);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int32_t row = row0 + ty + 16 * i;
    if (row < seg_end) {
      const float inv = 1.f / l[i];
#pragma unroll
      for (int jj = 0; jj < JJ; ++jj) {
        const int col = tx + 16 * jj;
        if (col < p.D) o[(int64_t)row * ldo + h * p.D + col] = acc[i][jj] * inv;
      }
      if (tx == 0) lse[(int64_t)row * p.H + h] = m[i] + logf(l[i]);
    }
  }
}

// delta[n,h] = sum_d dO[n,h,d] * O[n,h,d]
__global__ void __launch_bounds__(256)
k_attn_delta(const float* __restrict__ dout, int64_t lddo, const float* __restrict__ o,
             int64_t ldo, float* __restrict__ delta, int64_t N, int H, int D) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= N * H) return;
  const int64_t n = idx / H;
  const int h = (int)(idx - n * H);
  float acc = 0.f;
  for (int d = 0; d < D; ++d)
    acc = fmaf(dout[n * lddo + h * D + d], o[n * ldo + h * D + d], acc);
  delta[idx] = acc;
}

struct AttnBwdArgs {
  const float* dout; int64_t lddo;
  const float* lse; const float* delta;
  int64_t ldg;                       // row stride of dq / dk / dv (floats)
};

// ---------------------------------------------------- backward: dK, dV --
template <int JJ>
__global__ void __launch_bounds__(kAttnThreads)
k_attn_bwd_kv(const AttnArgs p, const AttnBwdArgs b, float* __restrict__ dk, float* __restrict__ dv) {
  extern __shared__ __align__(16) float smem[];
  int32_t seg_beg, seg_end, k0;
  if (!locate_tile(p.cu, p.S, blockIdx.x, &seg_beg, &seg_end, &k0)) return;
  const int h = blockIdx.y;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int lds = p.Dp + 4;
  float* Ks = smem;
  float* Vs = Ks + kTile * lds;
  float* Qs = Vs + kTile * lds;
  float* Gs = Qs + kTile * lds;   // dO tile
  float* Ps = Gs + kTile * lds;   // P  [q][k]
  float* Ds = Ps + kTile * kPS;   // dS [q][k]
  load_tile(Ks, p.k + h * p.D, p.ldk, k0, seg_end, p.D, p.Dp);
  load_tile(Vs, p.v + h * p.D, p.ldv, k0, seg_end, p.D, p.Dp);
  float adk[4][JJ], adv[4][JJ];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int jj = 0; jj < JJ; ++jj) { adk[i][jj] = 0.f; adv[i][jj] = 0.f; }

  for (int32_t q0 = seg_beg; q0 < seg_end; q0 += kTile) {
    __syncthreads();
    load_tile(Qs, p.q + h * p.D, p.ldq, q0, seg_end, p.D, p.Dp);
    load_tile(Gs, b.dout + h * p.D, b.lddo, q0, seg_end, p.D, p.Dp);
    __syncthreads();
    float s[4][4], dp[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) { s[i][j] = 0.f; dp[i][j] = 0.f; }
    mma_nt(s, Qs, Ks, p.Dp, ty, tx);    // rows = queries, cols = keys
    mma_nt(dp, Gs, Vs, p.Dp, ty, tx);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int32_t qrow = q0 + ty + 16 * i;
      const bool qok = qrow < seg_end;
      const float lse_q = qok ? b.lse[(int64_t)qrow * p.H + h] : 0.f;
      const float del_q = qok ? b.delta[(int64_t)qrow * p.H + h] : 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const bool ok = qok && (k0 + tx + 16 * j) < seg_end;
        const float pij = ok ? expf(s[i][j] * p.scale - lse_q) : 0.f;
        Ps[(ty + 16 * i) * kPS + tx + 16 * j] = pij;
        Ds[(ty + 16 * i) * kPS + tx + 16 * j] = pij * (dp[i][j] - del_q);
      }
    }
    __syncthreads();
    mma_ptx<JJ>(adv, Ps, Gs, p.Dp, ty, tx);  // dV[k][d] += P[q][k] dO[q][d]
    mma_ptx<JJ>(adk, Ds, Qs, p.Dp, ty, tx);  // dK[k][d] += dS[q][k] Q[q][d]
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int32_t row = k0 + ty + 16 * i;
    if (row < seg_end) {
#pragma unroll
      for (int jj = 0; jj < JJ; ++jj) {
        const int col = tx + 16 * jj;
        if (col < p.D) {
          const int64_t off = (int64_t)row * b.ldg + h * p.D + col;
          dk[off] = adk[i][jj] * p.scale;
          dv[off] = adv[i][jj];
        }
      }
    }
  }
}

// -------------------------------------------------------- backward: dQ --
template <int JJ>
__global__ void __launch_bounds__(kAttnThreads)
k_attn_bwd_q(const AttnArgs p, const AttnBwdArgs b, float* __restrict__ dq) {
  extern __shared__ __align__(16) float smem[];
  int32_t seg_beg, seg_end, q0;
  if (!locate_tile(p.cu, p.S, blockIdx.x, &seg_beg, &seg_end, &q0)) return;
  const int h = blockIdx.y;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int lds = p.Dp + 4;
  float* Qs = smem;
  float* Gs = Qs + kTile * lds;
  float* Ks = Gs + kTile * lds;
  float* Vs = Ks + kTile * lds;
  float* Ds = Vs + kTile * lds;  // dS [q][k]
  load_tile(Qs, p.q + h * p.D, p.ldq, q0, seg_end, p.D, p.Dp);
  load_tile(Gs, b.dout + h * p.D, b.lddo, q0, seg_end, p.D, p.Dp);
  float lse_q[4], del_q[4], adq[4][JJ];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int32_t qrow = q0 + ty + 16 * i;
    const bool qok = qrow < seg_end;
    lse_q[i] = qok ? b.lse[(int64_t)qrow * p.H + h] : 0.f;
    del_q[i] = qok ? b.delta[(int64_t)qrow * p.H + h] : 0.f;
#pragma unroll
    for (int jj = 0; jj < JJ; ++jj) adq[i][jj] = 0.f;
  }
  for (int32_t k0 = seg_beg; k0 < seg_end; k0 += kTile) {
    __syncthreads();
    load_tile(Ks, p.k + h * p.D, p.ldk, k0, seg_end, p.D, p.Dp);
    load_tile(Vs, p.v + h * p.D, p.ldv, k0, seg_end, p.D, p.Dp);
    __syncthreads();
    float s[4][4], dp[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) { s[i][j] = 0.f; dp[i][j] = 0.f; }
    mma_nt(s, Qs, Ks, p.Dp, ty, tx);
    mma_nt(dp, Gs, Vs, p.Dp, ty, tx);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const bool qok = (q0 + ty + 16 * i) < seg_end;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const bool ok = qok && (k0 + tx + 16 * j) < seg_end;
        const float pij = ok ? expf(s[i][j] * p.scale - lse_q[i]) : 0.f;
        Ds[(ty + 16 * i) * kPS + tx + 16 * j] = pij * (dp[i][j] - del_q[i]);
      }
    }
    __syncthreads();
    mma_pv<JJ>(adq, Ds, Ks, p.Dp, ty, tx);  // dQ[q][d] += dS[q][k] K[k][d]
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int32_t row = q0 + ty + 16 * i;
    if (row < seg_end) {
#pragma unroll
      for (int jj = 0; jj < JJ; ++jj) {
        const int col = tx + 16 * jj;
        if (col < p.D) dq[(int64_t)row * b.ldg + h * p.D + col] = adq[i][jj] * p.scale;
      }
    }
  }
}

inline int pick_jj(int Dp) { return Dp <= 16 ? 1 : (Dp <= 32 ? 2 : (Dp <= 64 ? 4 : 8)); }

#define HGK_DISPATCH_JJ(JJV, ...)                          \
  switch (JJV) {                                           \
    case 1: { constexpr int JJ = 1; __VA_ARGS__; } break;  \
    case 2: { constexpr int JJ = 2; __VA_ARGS__; } break;  \
    case 4: { constexpr int JJ = 4; __VA_ARGS__; } break;  \
    default: { constexpr int JJ = 8; __VA_ARGS__; } break; \
  }

template <typename K>
int set_smem(K kernel, size_t bytes) {
  if (bytes > 48 * 1024) {
    cudaError_t err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (err != cudaSuccess) return fail(HGK_ERR_CUDA, "attn: smem opt-in failed: %s", cudaGetErrorString(err));
  }
  return HGK_OK;
}

}  // namespace hgk

namespace hgk {
int attn_fwd_tc(float* o, int64_t ldo, float* lse, const float* q, int64_t ldq, const float* k,
                int64_t ldk, const float* v, int64_t ldv, const int32_t* cu, int64_t S, int64_t N,
                int32_t H, int32_t D, float scale, cudaStream_t st);
int attn_bwd_tc(float* dq, float* dk, float* dv, int64_t ldg, const float* delta, const float* dout, int64_t lddo,
                const float* lse, const float* q, int64_t ldq, const float* k, int64_t ldk,
                const float* v, int64_t ldv, const int32_t* cu, int64_t S, int64_t N, int32_t H,
                int32_t D, float scale, cudaStream_t st);
// HGK_ATTN_SIMT=1 forces the SIMT kernels (A/B runs); read once.
static bool attn_force_simt() {
  static const bool v = [] { const char* e = getenv("HGK_ATTN_SIMT"); return e && e[0] == '1'; }();
  return v;
}
}  // namespace hgk

using namespace hgk;

static int attn_check(const char* what, const void* q, const void* k, const void* v,
                      const int32_t* cu, int64_t S, int64_t N, int32_t H, int32_t D) {
  HGK_REQUIRE(N >= 0 && S >= 0 && H >= 1 && D >= 1, "%s: bad sizes N=%lld S=%lld H=%d D=%d", what,
              (long long)N, (long long)S, H, D);
  HGK_REQUIRE(D <= 128, "%s: head dim %d > 128 unsupported", what, D);
  HGK_REQUIRE(H <= 65535, "%s: too many heads", what);
  HGK_REQUIRE(N == 0 || (q && k && v && cu && S > 0), "%s: NULL operand", what);
  return HGK_OK;
}

extern "C" int hgk_attn_varlen_fwd(float* o, int64_t ldo, float* lse, const float* q,
                                   int64_t ldq, const float* k, int64_t ldk, const float* v,
                                   int64_t ldv, const int32_t* cu_seqlens, int64_t S, int64_t N,
                                   int32_t H, int32_t D, float scale, void* stream) {
  int rc = attn_check("attn_varlen_fwd", q, k, v, cu_seqlens, S, N, H, D);
  if (rc) return rc;
  if (N == 0) return HGK_OK;
  HGK_REQUIRE(o && lse, "attn_varlen_fwd: o/lse NULL");
  if (!attn_force_simt()) {
    rc = attn_fwd_tc(o, ldo, lse, q, ldq, k, ldk, v, ldv, cu_seqlens, S, N, H, D, scale, (cudaStream_t)stream);
    if (rc <= 0) return rc;   // handled (or failed) on the tensor-core path; 1 = shape not covered
  }
  const int Dp = (D + 3) / 4 * 4;
  AttnArgs a{q, ldq, k, ldk, v, ldv, cu_seqlens, (int32_t)S, N, H, D, Dp, scale};
  const size_t smem = sizeof(float) * (3 * kTile * (Dp + 4) + kTile * kPS);
  dim3 grid((unsigned)(ceil_div(N, kTile) + S), (unsigned)H);
  cudaStream_t st = (cudaStream_t)stream;
  HGK_DISPATCH_JJ(pick_jj(Dp), {
    rc = set_smem(k_attn_fwd<JJ>, smem);
    if (rc) return rc;
    k_attn_fwd<JJ><<<grid, kAttnThreads, smem, st>>>(a, o, ldo, lse);
  });
  return check_launch("attn_varlen_fwd");
}

extern "C" int hgk_attn_varlen_bwd_ld(float* dq, float* dk, float* dv, int64_t ldg, float* delta,
                                      const float* dout, int64_t lddo, const float* o, int64_t ldo,
                                   const float* lse, const float* q, int64_t ldq, const float* k,
                                   int64_t ldk, const float* v, int64_t ldv,
                                   const int32_t* cu_seqlens, int64_t S, int64_t N, int32_t H,
                                   int32_t D, float scale, void* stream) {
  int rc = attn_check("attn_varlen_bwd", q, k, v, cu_seqlens, S, N, H, D);
  if (rc) return rc;
  if (N == 0) return HGK_OK;
  HGK_REQUIRE(dq && dk && dv && delta && dout && o && lse, "attn_varlen_bwd: NULL operand");
  HGK_REQUIRE(ldg >= (int64_t)H * D, "attn_varlen_bwd: gradient row stride %lld < H*D", (long long)ldg);
  const int Dp = (D + 3) / 4 * 4;
  AttnArgs a{q, ldq, k, ldk, v, ldv, cu_seqlens, (int32_t)S, N, H, D, Dp, scale};
  AttnBwdArgs b{dout, lddo, lse, delta, ldg};
  cudaStream_t st = (cudaStream_t)stream;
  k_attn_delta<<<(unsigned)ceil_div(N * H, 256), 256, 0, st>>>(dout, lddo, o, ldo, delta, N, H, D);
  if (!attn_force_simt()) {
    rc = attn_bwd_tc(dq, dk, dv, ldg, delta, dout, lddo, lse, q, ldq, k, ldk, v, ldv, cu_seqlens, S, N, H, D, scale, st);
    if (rc <= 0) return rc;
  }
  dim3 grid((unsigned)(ceil_div(N, kTile) + S), (unsigned)H);
  const size_t smem_kv = sizeof(float) * (4 * kTile * (Dp + 4) + 2 * kTile * kPS);
  const size_t smem_q = sizeof(float) * (4 * kTile * (Dp + 4) + kTile * kPS);
  HGK_DISPATCH_JJ(pick_jj(Dp), {
    rc = set_smem(k_attn_bwd_kv<JJ>, smem_kv);
    if (rc) return rc;
    k_attn_bwd_kv<JJ><<<grid, kAttnThreads, smem_kv, st>>>(a, b, dk, dv);
    rc = set_smem(k_attn_bwd_q<JJ>, smem_q);
    if (rc) return rc;
    k_attn_bwd_q<JJ><<<grid, kAttnThreads, smem_q, st>>>(a, b, dq);
  });
  return check_launch("attn_varlen_bwd");
}

extern "C" int hgk_attn_varlen_bwd(float* dq, float* dk, float* dv, float* delta,
                                   const float* dout, int64_t lddo, const float* o, int64_t ldo,
                                   const float* lse, const float* q, int64_t ldq, const float* k,
                                   int64_t ldk, const float* v, int64_t ldv,
                                   const int32_t* cu_seqlens, int64_t S, int64_t N, int32_t H,
                                   int32_t D, float scale, void* stream) {
  return hgk_attn_varlen_bwd_ld(dq, dk, dv, (int64_t)H * D, delta, dout, lddo, o, ldo, lse, q, ldq, k, ldk, v, ldv,
                                cu_seqlens, S, N, H, D, scale, stream);
}
