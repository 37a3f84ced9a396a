// Batched radius graph on cell-list bins, with periodic images.
//
// Replaces (reference, CPU, one graph at a time):
//   ase.neighborlist.neighbor_list("ijdS") + _remove_true_self_loops +
//   _limit_neighbors        preprocess/graph_samples_checks_and_updates.py:146-282
//   PyG RadiusGraph / RadiusInteractionGraph -> torch_cluster.radius_graph
//                           same file :109-114, models/SCFStack.py:59-61
//
// Layout of one call (all graphs of a mini-batch at once):
//   frames  : per graph, fp64 cell, inverse cell, bin grid (k_rg_frames)
//   assign  : per atom, global bin id + integer wrap offset (k_rg_assign)
//   bins    : CSR of atoms by bin, ascending atom id inside a bin (csr.cu)
//   count   : one warp per destination atom scans the (2m+1)^3 neighbour bins
//   fill    : same scan, lane-ordered compaction of the matches, then the warp
//             orders its own segment by (d, src, Sx, Sy, Sz) and keeps the
//             first `deg` -- the reference's "nearest max_neighbours per dst".
// Integer/fp64 work, HBM/L2-latency bound; nothing here is GEMM-shaped.
#include "hgk_common.cuh"

namespace hgk {

int csr_build_impl(const int32_t* key, int64_t E, int64_t N, int32_t* rowptr,
                   int32_t* perm, void* workspace, cudaStream_t st);

struct Frame {
  double C[9];     // cell rows a0,a1,a2
  double Ci[9];    // inverse: frac = pos . Ci
  double fmin[3];  // lower corner of the bin grid in fractional coordinates
  double fext[3];  // extent of the bin grid in fractional coordinates
  int32_t nb[3];   // bins per axis
  int32_t m[3];    // search half-width in bins
  int32_t pbc[3];
  int32_t bin_offset;  // first global bin id of this graph
  int32_t nbins;
  int32_t pad;
};

struct Cand {
  double d;
  int32_t src, sx, sy, sz;
};

__device__ __forceinline__ bool cand_less(const Cand& a, const Cand& b) {
  if (a.d != b.d) return a.d < b.d;
  if (a.src != b.src) return a.src < b.src;
  if (a.sx != b.sx) return a.sx < b.sx;
  if (a.sy != b.sy) return a.sy < b.sy;
  return a.sz < b.sz;
}

__device__ __forceinline__ void inverse3(const double* C, double* Ci, double* h) {
  // cofactors; rows of C are lattice vectors a0,a1,a2
  const double c00 = C[4] * C[8] - C[5] * C[7];
  const double c01 = C[5] * C[6] - C[3] * C[8];
  const double c02 = C[3] * C[7] - C[4] * C[6];
  const double det = C[0] * c00 + C[1] * c01 + C[2] * c02;
  const double id = 1.0 / det;
  Ci[0] = c00 * id; Ci[1] = (C[2] * C[7] - C[1] * C[8]) * id; Ci[2] = (C[1] * C[5] - C[2] * C[4]) * id;
  Ci[3] = c01 * id; Ci[4] = (C[0] * C[8] - C[2] * C[6]) * id; Ci[5] = (C[2] * C[3] - C[0] * C[5]) * id;
  Ci[6] = c02 * id; Ci[7] = (C[1] * C[6] - C[0] * C[7]) * id; Ci[8] = (C[0] * C[4] - C[1] * C[3]) * id;
  // perpendicular height along axis k = 1 / |column k of Ci|
  for (int k = 0; k < 3; ++k) {
    const double n2 = Ci[k] * Ci[k] + Ci[3 + k] * Ci[3 + k] + Ci[6 + k] * Ci[6 + k];
    h[k] = 1.0 / sqrt(n2);
  }
}

__device__ __forceinline__ void to_frac(const Frame& fr, double x, double y, double z, double* f) {
  f[0] = x * fr.Ci[0] + y * fr.Ci[3] + z * fr.Ci[6];
  f[1] = x * fr.Ci[1] + y * fr.Ci[4] + z * fr.Ci[7];
  f[2] = x * fr.Ci[2] + y * fr.Ci[5] + z * fr.Ci[8];
}

// One CTA per graph.
__global__ void __launch_bounds__(128)
k_rg_frames(const float* __restrict__ pos, const int32_t* __restrict__ gptr,
            const float* __restrict__ cell, const uint8_t* __restrict__ pbc,
            double r, Frame* __restrict__ frames) {
  const int g = blockIdx.x;
  __shared__ Frame fr;
  __shared__ double h[3];
  __shared__ double smin[3][4], smax[3][4];
  if (threadIdx.x == 0) {
    for (int k = 0; k < 9; ++k)
      fr.C[k] = cell ? (double)cell[g * 9 + k] : ((k % 4 == 0) ? 1.0 : 0.0);
    for (int k = 0; k < 3; ++k) fr.pbc[k] = (cell && pbc) ? (pbc[g * 3 + k] ? 1 : 0) : 0;
    inverse3(fr.C, fr.Ci, h);
  }
  __syncthreads();
  const int32_t beg = gptr[g], end = gptr[g + 1];
  double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
  for (int32_t a = beg + threadIdx.x; a < end; a += blockDim.x) {
    double f[3];
    to_frac(fr, (double)pos[3 * a], (double)pos[3 * a + 1], (double)pos[3 * a + 2], f);
    for (int k = 0; k < 3; ++k) { lo[k] = fmin(lo[k], f[k]); hi[k] = fmax(hi[k], f[k]); }
  }
  for (int k = 0; k < 3; ++k) {
    for (int off = 16; off > 0; off >>= 1) {
      lo[k] = fmin(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], off));
      hi[k] = fmax(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], off));
    }
    if ((threadIdx.x & 31) == 0) { smin[k][threadIdx.x >> 5] = lo[k]; smax[k][threadIdx.x >> 5] = hi[k]; }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const int32_t n = end - beg;
    const double rbin = r * (1.0 + 1e-6) + 1e-12;  // guard band against bin-edge rounding
    double nbd[3];
    for (int k = 0; k < 3; ++k) {
      double l = fmin(fmin(smin[k][0], smin[k][1]), fmin(smin[k][2], smin[k][3]));
      double u = fmax(fmax(smax[k][0], smax[k][1]), fmax(smax[k][2], smax[k][3]));
      if (n == 0) { l = 0.0; u = 0.0; }
      if (fr.pbc[k]) {
        fr.fmin[k] = 0.0; fr.fext[k] = 1.0;
        nbd[k] = floor(h[k] / rbin);
        fr.m[k] = (nbd[k] >= 1.0) ? 1 : (int32_t)ceil(rbin / h[k]);
      } else {
        fr.fmin[k] = l; fr.fext[k] = u - l;
        nbd[k] = floor((u - l) * h[k] / rbin);
        fr.m[k] = 1;
      }
      if (nbd[k] < 1.0) nbd[k] = 1.0;
      if (nbd[k] > 1024.0) nbd[k] = 1024.0;
    }
    // keep the bin table O(atoms): at most max(1, 2n) bins per graph
    const double cap = (double)max(1, 2 * n);
    for (int it = 0; it < 64 && nbd[0] * nbd[1] * nbd[2] > cap; ++it) {
      int kmax = 0;
      if (nbd[1] > nbd[kmax]) kmax = 1;
      if (nbd[2] > nbd[kmax]) kmax = 2;
      nbd[kmax] = fmax(1.0, floor(nbd[kmax] * 0.5));
    }
    for (int k = 0; k < 3; ++k) fr.nb[k] = (int32_t)nbd[k];
    fr.nbins = fr.nb[0] * fr.nb[1] * fr.nb[2];
    fr.bin_offset = 0;
    fr.pad = 0;
    frames[g] = fr;
  }
}

// Exclusive scan of nbins over graphs (B is small: one thread).
__global__ void k_rg_bin_offsets(Frame* frames, int32_t B) {
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    int32_t acc = 0;
    for (int32_t g = 0; g < B; ++g) { frames[g].bin_offset = acc; acc += frames[g].nbins; }
  }
}

__global__ void __launch_bounds__(256)
k_rg_assign(const float* __restrict__ pos, const int32_t* __restrict__ gptr, int32_t B,
            int64_t N, const Frame* __restrict__ frames, int32_t* __restrict__ atom_graph,
            int32_t* __restrict__ atom_bin, int32_t* __restrict__ atom_w) {
  const int64_t a = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= N) return;
  // graph of atom a: last g with gptr[g] <= a
  int32_t lo = 0, hi = B;  // invariant gptr[lo] <= a < gptr[hi]
  while (hi - lo > 1) {
    const int32_t mid = (lo + hi) >> 1;
    if (gptr[mid] <= a) lo = mid; else hi = mid;
  }
  const int32_t g = lo;
  const Frame& fr = frames[g];
  double f[3];
  to_frac(fr, (double)pos[3 * a], (double)pos[3 * a + 1], (double)pos[3 * a + 2], f);
  int32_t b[3], w[3];
  for (int k = 0; k < 3; ++k) {
    if (fr.pbc[k]) {
      const double fl = floor(f[k]);
      w[k] = (int32_t)fl;
      double x = (f[k] - fl) * fr.nb[k];
      int32_t bi = (int32_t)x;
      b[k] = bi < 0 ? 0 : (bi >= fr.nb[k] ? fr.nb[k] - 1 : bi);
    } else {
      w[k] = 0;
      int32_t bi = 0;
      if (fr.fext[k] > 0.0) bi = (int32_t)(((f[k] - fr.fmin[k]) / fr.fext[k]) * fr.nb[k]);
      b[k] = bi < 0 ? 0 : (bi >= fr.nb[k] ? fr.nb[k] - 1 : bi);
    }
  }
  atom_graph[a] = g;
  atom_bin[a] = fr.bin_offset + (b[0] * fr.nb[1] + b[1]) * fr.nb[2] + b[2];
  atom_w[3 * a] = w[0]; atom_w[3 * a + 1] = w[1]; atom_w[3 * a + 2] = w[2];
}

struct ScanArgs {
  const float* pos;
  const Frame* frames;
  const int32_t* atom_graph;
  const int32_t* atom_bin;
  const int32_t* atom_w;
  const int32_t* bin_rowptr;
  const int32_t* bin_atoms;
  int64_t N;
  double r;
  float r2f;
  int32_t max_nbr, loop, mode;
};

// Visit every candidate (i, S) of destination atom j with all 32 lanes; the
// callback receives (lane_has_match, candidate) in a warp-converged context.
template <typename F>
__device__ __forceinline__ void scan_candidates(const ScanArgs& p, int64_t j, int lane, F&& on_match) {
  const Frame& fr = p.frames[p.atom_graph[j]];
  const float pjx = p.pos[3 * j], pjy = p.pos[3 * j + 1], pjz = p.pos[3 * j + 2];
  const int32_t wjx = p.atom_w[3 * j], wjy = p.atom_w[3 * j + 1], wjz = p.atom_w[3 * j + 2];
  int32_t lb = p.atom_bin[j] - fr.bin_offset;
  const int32_t bz = lb % fr.nb[2]; lb /= fr.nb[2];
  const int32_t by = lb % fr.nb[1];
  const int32_t bx = lb / fr.nb[1];
  for (int32_t ox = -fr.m[0]; ox <= fr.m[0]; ++ox) {
    int32_t tx = bx + ox, qx = 0;
    if (fr.pbc[0]) { qx = (tx >= 0) ? tx / fr.nb[0] : -((-tx + fr.nb[0] - 1) / fr.nb[0]); tx -= qx * fr.nb[0]; }
    else if (tx < 0 || tx >= fr.nb[0]) continue;
    for (int32_t oy = -fr.m[1]; oy <= fr.m[1]; ++oy) {
      int32_t ty = by + oy, qy = 0;
      if (fr.pbc[1]) { qy = (ty >= 0) ? ty / fr.nb[1] : -((-ty + fr.nb[1] - 1) / fr.nb[1]); ty -= qy * fr.nb[1]; }
      else if (ty < 0 || ty >= fr.nb[1]) continue;
      for (int32_t oz = -fr.m[2]; oz <= fr.m[2]; ++oz) {
        int32_t tz = bz + oz, qz = 0;
        if (fr.pbc[2]) { qz = (tz >= 0) ? tz / fr.nb[2] : -((-tz + fr.nb[2] - 1) / fr.nb[2]); tz -= qz * fr.nb[2]; }
        else if (tz < 0 || tz >= fr.nb[2]) continue;
        const int32_t bin = fr.bin_offset + (tx * fr.nb[1] + ty) * fr.nb[2] + tz;
        const int32_t beg = p.bin_rowptr[bin], end = p.bin_rowptr[bin + 1];
        for (int32_t base = beg; base < end; base += 32) {
          const int32_t idx = base + lane;
          bool match = false;
          Cand c;
          c.d = 0.0; c.src = 0; c.sx = c.sy = c.sz = 0;
          if (idx < end) {
            const int32_t i = p.bin_atoms[idx];
            const float pix = p.pos[3 * i], piy = p.pos[3 * i + 1], piz = p.pos[3 * i + 2];
            c.src = i;
            if (p.mode == HGK_RG_CLUSTER) {
              const float dx = __fsub_rn(pix, pjx), dy = __fsub_rn(piy, pjy), dz = __fsub_rn(piz, pjz);
              const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
              c.d = (double)d2;
              match = (d2 < p.r2f) && (p.loop || i != (int32_t)j);
            } else {
              c.sx = -qx - wjx + p.atom_w[3 * i];
              c.sy = -qy - wjy + p.atom_w[3 * i + 1];
              c.sz = -qz - wjz + p.atom_w[3 * i + 2];
              const double sx = (double)c.sx, sy = (double)c.sy, sz = (double)c.sz;
              // shift = S . cell, summed row by row (numpy dot order), no FMA
              const double tx_ = __dadd_rn(__dadd_rn(__dmul_rn(sx, fr.C[0]), __dmul_rn(sy, fr.C[3])), __dmul_rn(sz, fr.C[6]));
              const double ty_ = __dadd_rn(__dadd_rn(__dmul_rn(sx, fr.C[1]), __dmul_rn(sy, fr.C[4])), __dmul_rn(sz, fr.C[7]));
              const double tz_ = __dadd_rn(__dadd_rn(__dmul_rn(sx, fr.C[2]), __dmul_rn(sy, fr.C[5])), __dmul_rn(sz, fr.C[8]));
              const double Dx = __dadd_rn(__dsub_rn((double)pjx, (double)pix), tx_);
              const double Dy = __dadd_rn(__dsub_rn((double)pjy, (double)piy), ty_);
              const double Dz = __dadd_rn(__dsub_rn((double)pjz, (double)piz), tz_);
              const double d2 = __dadd_rn(__dadd_rn(__dmul_rn(Dx, Dx), __dmul_rn(Dy, Dy)), __dmul_rn(Dz, Dz));
              c.d = __dsqrt_rn(d2);
              const bool self = (i == (int32_t)j) && c.sx == 0 && c.sy == 0 && c.sz == 0;
              match = (c.d < p.r) && (p.loop || !self);
            }
          }
          on_match(match, c);
        }
      }
    }
  }
}

__global__ void __launch_bounds__(256)
k_rg_count(const ScanArgs p, int32_t* __restrict__ deg_raw, int32_t* __restrict__ deg) {
  const int lane = threadIdx.x & 31;
  const int64_t j = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (j >= p.N) return;
  int32_t count = 0;
  scan_candidates(p, j, lane, [&](bool match, const Cand&) {
    count += __popc(__ballot_sync(0xffffffffu, match));
  });
  if (lane == 0) {
    deg_raw[j] = count;
    deg[j] = (p.max_nbr > 0 && count > p.max_nbr) ? p.max_nbr : count;
  }
}

__global__ void __launch_bounds__(256)
k_rg_fill(const ScanArgs p, const int32_t* __restrict__ rowptr_raw,
          const int32_t* __restrict__ rowptr, Cand* __restrict__ cand, int64_t cand_cap,
          int32_t* __restrict__ src, int32_t* __restrict__ dst,
          int32_t* __restrict__ shift_int, double* __restrict__ dist) {
  const int lane = threadIdx.x & 31;
  const int64_t j = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (j >= p.N) return;
  const int32_t rbeg = rowptr_raw[j];
  int32_t rlen = rowptr_raw[j + 1] - rbeg;
  // never write past the caller's candidate buffer (cand_bytes / 24 records): a row that does not
  // fit is truncated instead -- its edge count then falls short of rowptr, which the caller sees
  if ((int64_t)rbeg + rlen > cand_cap) rlen = (int32_t)max((int64_t)0, cand_cap - rbeg);
  int32_t written = 0;
  scan_candidates(p, j, lane, [&](bool match, const Cand& c) {
    const unsigned b = __ballot_sync(0xffffffffu, match);
    if (match) {
      const int32_t slot = written + __popc(b & ((1u << lane) - 1u));
      if (slot < rlen) cand[rbeg + slot] = c;
    }
    written += __popc(b);
  });
  __syncwarp();
  // order the warp's own segment; keep the first `keep`
  const int32_t obeg = rowptr[j];
  const int32_t keep = rowptr[j + 1] - obeg;
  const int32_t len = min(written, rlen);
  for (int32_t t = lane; t < len; t += 32) {
    const Cand c = cand[rbeg + t];
    int32_t rank = 0;
    for (int32_t s = 0; s < len; ++s) rank += cand_less(cand[rbeg + s], c) ? 1 : 0;
    if (rank < keep) {
      const int64_t o = obeg + rank;
      src[o] = c.src;
      dst[o] = (int32_t)j;
      shift_int[3 * o] = c.sx; shift_int[3 * o + 1] = c.sy; shift_int[3 * o + 2] = c.sz;
      dist[o] = c.d;
    }
  }
}

struct RgWorkspace {
  Frame* frames;
  int32_t* atom_graph;
  int32_t* atom_bin;
  int32_t* atom_w;
  int32_t* bin_rowptr;
  int32_t* bin_atoms;
  void* csr_ws;
  int64_t nbins_bound;
};

inline int64_t rg_nbins_bound(int64_t N, int64_t B) { return 2 * N + B; }

inline int64_t rg_carve(int64_t N, int64_t B, char* base, RgWorkspace* w) {
  int64_t off = 0;
  auto take = [&](int64_t bytes) { int64_t o = off; off += align_up(bytes, 256); return base ? base + o : nullptr; };
  const int64_t nb = rg_nbins_bound(N, B);
  char* p;
  p = take(sizeof(Frame) * B); if (w) w->frames = (Frame*)p;
  p = take(4 * N); if (w) w->atom_graph = (int32_t*)p;
  p = take(4 * N); if (w) w->atom_bin = (int32_t*)p;
  p = take(12 * N); if (w) w->atom_w = (int32_t*)p;
  p = take(4 * (nb + 1)); if (w) w->bin_rowptr = (int32_t*)p;
  p = take(4 * N); if (w) w->bin_atoms = (int32_t*)p;
  p = take(hgk_csr_build_workspace_bytes(N, nb)); if (w) { w->csr_ws = p; w->nbins_bound = nb; }
  return off + 256;
}

}  // namespace hgk

using namespace hgk;

extern "C" int64_t hgk_radius_graph_workspace_bytes(int64_t N, int64_t B) {
  if (N < 0 || B < 0) return 0;
  return rg_carve(N, B, nullptr, nullptr);
}

static int rg_check(const float* pos, const int32_t* gptr, const float* cell,
                    const uint8_t* pbc, int64_t N, int64_t B, double r, int32_t mode,
                    void* ws, int64_t ws_bytes) {
  HGK_REQUIRE(N >= 0 && B >= 0, "radius_graph: negative size");
  HGK_REQUIRE(N < (1ll << 30), "radius_graph: N too large");
  HGK_REQUIRE(N == 0 || (pos && gptr && B > 0), "radius_graph: pos/gptr NULL or B=0 with N>0");
  HGK_REQUIRE(r > 0.0, "radius_graph: r must be positive");
  HGK_REQUIRE(mode == HGK_RG_ASE || mode == HGK_RG_CLUSTER, "radius_graph: unknown mode %d", mode);
  HGK_REQUIRE(mode == HGK_RG_ASE || cell == nullptr, "radius_graph: CLUSTER mode has no cell");
  HGK_REQUIRE((cell == nullptr) == (pbc == nullptr), "radius_graph: cell and pbc go together");
  if (!ws || ws_bytes < hgk_radius_graph_workspace_bytes(N, B))
    return fail(HGK_ERR_WORKSPACE, "radius_graph: workspace %lld < %lld bytes",
                (long long)ws_bytes, (long long)hgk_radius_graph_workspace_bytes(N, B));
  return HGK_OK;
}

static ScanArgs make_scan(const float* pos, const RgWorkspace& w, int64_t N, double r,
                          int32_t max_nbr, int32_t loop, int32_t mode) {
  ScanArgs s;
  s.pos = pos; s.frames = w.frames; s.atom_graph = w.atom_graph; s.atom_bin = w.atom_bin;
  s.atom_w = w.atom_w; s.bin_rowptr = w.bin_rowptr; s.bin_atoms = w.bin_atoms;
  s.N = N; s.r = r; s.r2f = (float)(r * r); s.max_nbr = max_nbr; s.loop = loop; s.mode = mode;
  return s;
}

extern "C" int hgk_radius_graph_count(const float* pos, const int32_t* gptr,
                                      const float* cell, const uint8_t* pbc, int64_t N,
                                      int64_t B, double r, int32_t max_nbr, int32_t loop,
                                      int32_t mode, int32_t* deg_raw, int32_t* deg,
                                      void* workspace, int64_t workspace_bytes,
                                      void* stream) {
  int rc = rg_check(pos, gptr, cell, pbc, N, B, r, mode, workspace, workspace_bytes);
  if (rc) return rc;
  if (N == 0) return HGK_OK;
  HGK_REQUIRE(deg_raw && deg, "radius_graph_count: deg outputs NULL");
  cudaStream_t st = (cudaStream_t)stream;
  RgWorkspace w;
  rg_carve(N, B, (char*)workspace, &w);
  k_rg_frames<<<(unsigned)B, 128, 0, st>>>(pos, gptr, cell, pbc, r, w.frames);
  k_rg_bin_offsets<<<1, 32, 0, st>>>(w.frames, (int32_t)B);
  k_rg_assign<<<(unsigned)ceil_div(N, 256), 256, 0, st>>>(pos, gptr, (int32_t)B, N, w.frames,
                                                          w.atom_graph, w.atom_bin, w.atom_w);
  rc = csr_build_impl(w.atom_bin, N, w.nbins_bound, w.bin_rowptr, w.bin_atoms, w.csr_ws, st);
  if (rc) return rc;
  const ScanArgs s = make_scan(pos, w, N, r, max_nbr, loop, mode);
  k_rg_count<<<(unsigned)ceil_div(N * 32, 256), 256, 0, st>>>(s, deg_raw, deg);
  return check_launch("radius_graph_count");
}

extern "C" int hgk_radius_graph_fill(const float* pos, const int32_t* gptr,
                                     const float* cell, const uint8_t* pbc, int64_t N,
                                     int64_t B, double r, int32_t max_nbr, int32_t loop,
                                     int32_t mode, const int32_t* rowptr_raw,
                                     const int32_t* rowptr, int32_t* src, int32_t* dst,
                                     int32_t* shift_int, double* dist, void* cand,
                                     int64_t cand_bytes, void* workspace,
                                     int64_t workspace_bytes, void* stream) {
  int rc = rg_check(pos, gptr, cell, pbc, N, B, r, mode, workspace, workspace_bytes);
  if (rc) return rc;
  if (N == 0) return HGK_OK;
  HGK_REQUIRE(rowptr_raw && rowptr, "radius_graph_fill: rowptr NULL");
  HGK_REQUIRE(cand_bytes >= 0 && (cand || cand_bytes == 0), "radius_graph_fill: cand NULL");
  static_assert(sizeof(Cand) == 24, "Cand must be 24 bytes (see hgk.h)");
  cudaStream_t st = (cudaStream_t)stream;
  RgWorkspace w;
  rg_carve(N, B, (char*)workspace, &w);
  const ScanArgs s = make_scan(pos, w, N, r, max_nbr, loop, mode);
  k_rg_fill<<<(unsigned)ceil_div(N * 32, 256), 256, 0, st>>>(s, rowptr_raw, rowptr, (Cand*)cand,
                                                            cand_bytes / (int64_t)sizeof(Cand),
                                                            src, dst, shift_int, dist);
  return check_launch("radius_graph_fill");
}
