// Edge geometry with explicit first and second derivatives (reference: get_edge_vectors_and_lengths,
// hydragnn/utils/model/operations.py:21-36):
//     v_e = pos[dst(e)] - pos[src(e)] + shift_e,   L_e = |v_e|,   u_e = v_e / (L_e + eps)
// The reference builds this from two index_selects, a subtraction, an addition, a vector norm and a
// division, and lets autograd derive the force loss's second derivatives (Base.py:614-620) through that
// chain.  Here: one kernel forward; one kernel for the vector-Jacobian product
//     t_e = g_v + g_L v / L + g_u / (L + eps) - v (g_u . v) / (L (L + eps)^2)        (d loss / d v_e)
// whose scatter to the two endpoints is a pair of segmented sums of the contraction family; and one
// kernel for the derivative of THAT map (what create_graph=True differentiates), written out by hand.
// Per-edge work on 3-vectors: latency / HBM bound, a few tens of bytes per edge.
#include "hgk_common.cuh"

namespace hgk {

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 ld3(const float* p) { return V3{p[0], p[1], p[2]}; }
__device__ __forceinline__ void st3(float* p, V3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }
__device__ __forceinline__ float dot3(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

__global__ void __launch_bounds__(256)
k_edge_geom_fwd(float* __restrict__ vec, float* __restrict__ len, float* __restrict__ unit,
                const float* __restrict__ pos, const int32_t* __restrict__ src, const int32_t* __restrict__ dst,
                const float* __restrict__ shifts, float eps, int64_t E) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  const V3 a = ld3(pos + 3 * (int64_t)src[e]), b = ld3(pos + 3 * (int64_t)dst[e]);
  // same operation order as the reference: (receiver - sender) + shift, separate roundings
  V3 v{__fsub_rn(b.x, a.x), __fsub_rn(b.y, a.y), __fsub_rn(b.z, a.z)};
  if (shifts != nullptr) {
    const V3 s = ld3(shifts + 3 * e);
    v = V3{__fadd_rn(v.x, s.x), __fadd_rn(v.y, s.y), __fadd_rn(v.z, s.z)};
  }
  const float L = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(v.x, v.x), __fmul_rn(v.y, v.y)), __fmul_rn(v.z, v.z)));
  st3(vec + 3 * e, v);
  len[e] = L;
  if (unit != nullptr) {
    const float d = __fadd_rn(L, eps);
    st3(unit + 3 * e, V3{v.x / d, v.y / d, v.z / d});
  }
}

// t = g_v + g_L v / L + g_u / (L + eps) - v (g_u . v) / (L (L + eps)^2); any of g_v, g_L, g_u may be NULL.
// At L = 0 the terms with 1 / L are dropped (the subgradient torch's norm backward uses).
__global__ void __launch_bounds__(256)
k_edge_geom_bwd(float* __restrict__ t, const float* __restrict__ g_vec, const float* __restrict__ g_len,
                const float* __restrict__ g_unit, const float* __restrict__ vec, const float* __restrict__ len,
                float eps, int64_t E) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  const V3 v = ld3(vec + 3 * e);
  const float L = len[e];
  const float invL = L > 0.f ? 1.f / L : 0.f;
  V3 out{0.f, 0.f, 0.f};
  if (g_vec != nullptr) out = ld3(g_vec + 3 * e);
  if (g_len != nullptr) {
    const float c = g_len[e] * invL;
    out.x += c * v.x; out.y += c * v.y; out.z += c * v.z;
  }
  if (g_unit != nullptr) {
    const V3 gu = ld3(g_unit + 3 * e);
    const float d = 1.f / (L + eps);
    const float c = dot3(gu, v) * invL * d * d;
    out.x += gu.x * d - c * v.x; out.y += gu.y * d - c * v.y; out.z += gu.z * d - c * v.z;
  }
  st3(t + 3 * e, out);
}

// Derivative of the map above: for h = d loss / d t,
//   d g_v = h (the caller reuses h),  d g_L = h . v / L,  d g_u = h / (L + eps) - v (h . v) / (L (L + eps)^2),
//   d v   = g_L (h / L - v (h . v) / L^3)
//           - [ (h . g_u) v + (g_u . v) h + (h . v) g_u ] c  -  (h . v)(g_u . v) c' v / L,
//   c = 1 / (L (L + eps)^2),  c' = dc/dL = -1 / (L^2 (L + eps)^2) - 2 / (L (L + eps)^3).
__global__ void __launch_bounds__(256)
k_edge_geom_bwd2(float* __restrict__ d_glen, float* __restrict__ d_gunit, float* __restrict__ d_vec,
                 const float* __restrict__ h, const float* __restrict__ g_len, const float* __restrict__ g_unit,
                 const float* __restrict__ vec, const float* __restrict__ len, float eps, int64_t E) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  const V3 v = ld3(vec + 3 * e), hh = ld3(h + 3 * e);
  const float L = len[e];
  const float invL = L > 0.f ? 1.f / L : 0.f;
  const float d = 1.f / (L + eps);
  const float c = invL * d * d;
  const float hv = dot3(hh, v);
  if (d_glen != nullptr) d_glen[e] = hv * invL;
  if (d_gunit != nullptr)
    st3(d_gunit + 3 * e, V3{hh.x * d - c * hv * v.x, hh.y * d - c * hv * v.y, hh.z * d - c * hv * v.z});
  V3 dv{0.f, 0.f, 0.f};
  if (g_len != nullptr) {
    const float gl = g_len[e];
    const float a = gl * invL, b = gl * hv * invL * invL * invL;
    dv.x += a * hh.x - b * v.x; dv.y += a * hh.y - b * v.y; dv.z += a * hh.z - b * v.z;
  }
  if (g_unit != nullptr) {
    const V3 gu = ld3(g_unit + 3 * e);
    const float hg = dot3(hh, gu), gv = dot3(gu, v);
    const float cp = -invL * invL * d * d - 2.f * invL * d * d * d;
    const float k = hg * c + hv * gv * cp * invL;
    dv.x -= k * v.x + gv * c * hh.x + hv * c * gu.x;
    dv.y -= k * v.y + gv * c * hh.y + hv * c * gu.y;
    dv.z -= k * v.z + gv * c * hh.z + hv * c * gu.z;
  }
  st3(d_vec + 3 * e, dv);
}

}  // namespace hgk

using namespace hgk;

extern "C" int hgk_edge_geom_fwd(float* vec, float* len, float* unit, const float* pos, const int32_t* src,
                                 const int32_t* dst, const float* shifts, float eps, int64_t E, void* stream) {
  HGK_REQUIRE(E >= 0, "edge_geom_fwd: negative E");
  if (E == 0) return HGK_OK;
  HGK_REQUIRE(vec && len && pos && src && dst, "edge_geom_fwd: NULL operand");
  k_edge_geom_fwd<<<(unsigned)ceil_div(E, 256), 256, 0, (cudaStream_t)stream>>>(vec, len, unit, pos, src, dst, shifts, eps, E);
  return check_launch("edge_geom_fwd");
}

extern "C" int hgk_edge_geom_bwd(float* t, const float* g_vec, const float* g_len, const float* g_unit,
                                 const float* vec, const float* len, float eps, int64_t E, void* stream) {
  HGK_REQUIRE(E >= 0, "edge_geom_bwd: negative E");
  if (E == 0) return HGK_OK;
  HGK_REQUIRE(t && vec && len, "edge_geom_bwd: NULL operand");
  k_edge_geom_bwd<<<(unsigned)ceil_div(E, 256), 256, 0, (cudaStream_t)stream>>>(t, g_vec, g_len, g_unit, vec, len, eps, E);
  return check_launch("edge_geom_bwd");
}

extern "C" int hgk_edge_geom_bwd2(float* d_glen, float* d_gunit, float* d_vec, const float* h, const float* g_len,
                                  const float* g_unit, const float* vec, const float* len, float eps, int64_t E,
                                  void* stream) {
  HGK_REQUIRE(E >= 0, "edge_geom_bwd2: negative E");
  if (E == 0) return HGK_OK;
  HGK_REQUIRE(d_vec && h && vec && len, "edge_geom_bwd2: NULL operand");
  k_edge_geom_bwd2<<<(unsigned)ceil_div(E, 256), 256, 0, (cudaStream_t)stream>>>(d_glen, d_gunit, d_vec, h, g_len, g_unit, vec,
                                                                                 len, eps, E);
  return check_launch("edge_geom_bwd2");
}
