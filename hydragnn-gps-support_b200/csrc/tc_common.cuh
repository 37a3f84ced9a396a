// tcgen05 / TMEM / mbarrier primitives for sm_100a (inline PTX, no CUTLASS dependency).
//
// Conventions used by the attention kernels (attn_tc.cu):
//   * MMA kind::tf32, cta_group::1, M = 128 (TMEM lane == tile row), fp32 accumulators in TMEM.
//   * Shared-memory operands are K-major, SWIZZLE_NONE "core matrix" layout: a core matrix is
//     8 rows x 16 bytes (4 tf32), stored as 128 contiguous bytes;
//         offset(row, k) = (row / 8) * SBO + (k / 4) * LBO + (row % 8) * 16 + (k % 4) * 4
//     LBO = byte distance between the two 16-byte K chunks one MMA (K = 8) consumes,
//     SBO = byte distance between consecutive 8-row groups.
//   * fp32 accuracy on tf32 tensor cores by operand splitting ("3xTF32"):
//         x = hi + lo,  hi = x with the 13 low mantissa bits cleared,  lo = x - hi  (exact),
//         a*b ~= hi_a*hi_b + hi_a*lo_b + lo_a*hi_b         (relative error ~2^-21)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace hgk {
namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ---- operand splitting -------------------------------------------------------
__device__ __forceinline__ float tf32_hi(float x) {
  return __uint_as_float(__float_as_uint(x) & 0xffffe000u);
}
__device__ __forceinline__ void split_tf32(float x, float& hi, float& lo) {
  hi = tf32_hi(x);
  lo = x - hi;  // exact: at most 13 significant bits
}

// ---- descriptors ---------------------------------------------------------------
// 64-bit shared-memory matrix descriptor (SWIZZLE_NONE, K-major, Blackwell version field = 1).
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr >> 4) & 0x3fffu);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3fffu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3fffu) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  return d;
}
// MN-major operands.  For 32-bit elements the hardware reads an MN-major operand only in the
// SWIZZLE_128B_BASE32B layout (layout type 1; measured with profiles/_scratch/mn_major_probe*.cu -- with
// any other layout type and the MN-major bit set the MMA reads zeros): rows of 128 bytes = 32 consecutive
// MN elements of ONE k, consecutive k rows 128 bytes apart, and the four 32-byte chunks of a row stored at
// chunk index c ^ (k & 3):
//     offset(mn, k) = (mn / 32) * LBO + k * 128 + ((((mn % 32) / 8) ^ (k % 4)) * 32) + (mn % 8) * 4   [bytes]
// SBO = 512 (distance between the 4-row atoms), one MMA (K = 8) covers 1024 bytes of rows.  This is the
// natural row-major tile of a [rows x 32] fp32 matrix, so "B = X^T" needs no transposed staging: the rows
// of X are the k index.  Tiles must start on a 1024-byte boundary (the swizzle is a function of the address).
__device__ __forceinline__ uint64_t smem_desc_mn(uint32_t saddr, uint32_t lbo_bytes) {
  return smem_desc(saddr, lbo_bytes, 512) | (static_cast<uint64_t>(1) << 61);
}
constexpr uint32_t kIdescBMajorMN = 1u << 16;   // OR into idesc_tf32: B is MN-major
constexpr uint32_t kIdescAMajorMN = 1u << 15;   // OR into idesc_tf32: A (shared memory) is MN-major
// 32-bit instruction descriptor: tf32 x tf32 -> f32, dense, A and B K-major.
__host__ __device__ constexpr uint32_t idesc_tf32(int M, int N) {
  return (1u << 4)            // D format  = F32
       | (2u << 7)            // A format  = TF32
       | (2u << 10)           // B format  = TF32
       | (static_cast<uint32_t>(N >> 3) << 17)
       | (static_cast<uint32_t>(M >> 4) << 24);
}

// ---- TMEM allocation (one full warp executes these) -----------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
               :: "r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(taddr), "r"(ncols) : "memory");
}

// ---- fences -----------------------------------------------------------------------
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// make generic-proxy shared-memory writes visible to the async proxy (the tensor core reads smem through it)
__device__ __forceinline__ void fence_smem_to_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- mbarrier -----------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
// Bounded wait: a barrier that never completes (a descriptor or phase bug) traps the kernel after
// ~2 s of wall time instead of hanging the GPU.
__device__ __forceinline__ uint64_t global_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const uint64_t t0 = global_ns();
  while (!mbar_try_wait(bar, parity)) {
    if (global_ns() - t0 > 2000000000ull) __trap();
  }
}

// ---- TMA: bulk tensor load global -> shared memory, completion counted in bytes on an mbarrier ----------
// (tensor maps: tma_host.cuh; they cover this GPU's memory only, and no load here is multicast)
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_2d(float* dst_smem, const void* tensor_map, uint64_t* bar, int32_t col, int32_t row) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      :: "r"(smem_u32(dst_smem)), "l"(reinterpret_cast<uint64_t>(tensor_map)), "r"(smem_u32(bar)), "r"(col), "r"(row)
      : "memory");
}

// bulk tensor store shared memory -> global (rows past the tensor's end are clipped), tracked in bulk groups
__device__ __forceinline__ void tma_store_2d(const void* tensor_map, const float* src_smem, int32_t col, int32_t row) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
               :: "l"(reinterpret_cast<uint64_t>(tensor_map)), "r"(smem_u32(src_smem)), "r"(col), "r"(row) : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// this thread's bulk stores have finished READING shared memory (the source may be overwritten)
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// this thread's bulk stores are complete
__device__ __forceinline__ void tma_store_wait() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// ---- named barriers: producers arrive and move on, the consumer warp waits ------------------------
__device__ __forceinline__ void named_bar_arrive(uint32_t id, uint32_t count) {
  asm volatile("bar.arrive %0, %1;" :: "r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ void named_bar_sync(uint32_t id, uint32_t count) {
  asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(count) : "memory");
}

// ---- single-lane election inside a converged warp ------------------------------------------------
// tcgen05.mma / commit take warp-uniform operands; issued under `if (elect_one_sync())` in a
// warp-uniform branch the compiler emits them straight-line (under `if (threadIdx.x == 0)` it
// wraps every MMA in an elect-and-branch loop over the possibly active lanes).
__device__ __forceinline__ bool elect_one_sync() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t.reg .pred P1;\n\t"
      "elect.sync _|P1, %1;\n\t"
      "@P1 mov.s32 %0, 1;\n\t}"
      : "+r"(pred) : "r"(0xffffffffu));
  return pred != 0;
}
__device__ __forceinline__ int warp_id_uniform() { return __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0); }

// ---- MMA issue (ONE thread) ------------------------------------------------------------
// D[tmem] (+)= A[smem] * B[smem]^T      (both K-major)
__device__ __forceinline__ void mma_tf32_ss(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc,
                                            uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      :: "r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem]^T      (A: lane = row, column = k)
__device__ __forceinline__ void mma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc,
                                            uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      :: "r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrive on `bar` once every MMA issued so far by this thread has completed
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
               :: "r"(smem_u32(bar)) : "memory");
}

// ---- TMEM <-> registers: 32 lanes x 32 consecutive columns per warp ---------------------
// thread i of warp w owns TMEM lane 32*(w%4)+i; r[j] <-> column (col0 + j)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&r)[8]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
      "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
      :: "r"(taddr),
         "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
         "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
         "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]),
         "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
      :: "r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
         "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&r)[8]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
      :: "r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// K-major core-matrix offset in floats (see header comment); lbo = 128 B, i.e. chunks contiguous.
__device__ __forceinline__ int core_off(int row, int kchunk, int sbo_floats) {
  return (row >> 3) * sbo_floats + kchunk * 32 + (row & 7) * 4;
}

// ================================================================================================
// Building blocks shared by the tensor-core kernels (attn_tc.cu, filter_tc.cu): 256-thread CTAs,
// TMEM lane = tile row, two threads per row.
constexpr int kThreads = 256;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr uint32_t kLbo = 128;   // bytes between the two 16-byte K chunks of one MMA

// ---- tile staging: global -> registers (prefetchable) -> shared memory --------------------------
// A [ROWS x D] fp32 tile is ROWS * D/4 float4 chunks, spread over the NT threads TOFF .. TOFF+NT-1
// of the CTA (TOFF a multiple of 32).  Chunk idx (thread t owns t, t+NT, ...) maps to
//   row r = idx % 4 + 4 * (idx / (4 * CH)),   16-byte column c = (idx / 4) % CH,      CH = D / 4
// so the 32 lanes of a warp cover 4 consecutive rows x 8 columns: a global read is four full
// 128-byte lines (the LSU data pipe, not HBM, was the limiter with 64-byte runs: ncu showed it 90 %
// busy in the weight-gradient kernel), and in the transposed store the 8 lanes of a store phase
// write 8 different features of one k chunk = 128 contiguous bytes (no bank conflicts); the row
// store pays a 2-way conflict instead (two columns of the same four rows share their banks).
template <int ROWS, int D, int NT = kThreads, int TOFF = 0>
struct Tile {
  static constexpr int CH = D / 4;
  static constexpr int PER = ROWS * CH / NT;
  static_assert(ROWS * CH % NT == 0, "tile must divide evenly over the CTA");
  static_assert(CH % 4 == 0, "rows are whole 64-byte runs");
  float4 v[PER];

  __device__ __forceinline__ static void coord_of(int tid, int it, int& r, int& c) {
    const int idx = tid + it * NT;
    r = (idx & 3) + 4 * (idx / (4 * CH));
    c = (idx >> 2) % CH;
  }
  __device__ __forceinline__ static void coord(int it, int& r, int& c) { coord_of((int)threadIdx.x - TOFF, it, r, c); }
  // the *_as forms take the thread's index inside its group of NT threads explicitly (groups whose
  // first thread is not a compile-time constant); the plain forms use threadIdx.x - TOFF
  __device__ __forceinline__ void load(const float* src, int64_t ld, int32_t row0, int32_t row_end) {
    load_as((int)threadIdx.x - TOFF, src, ld, row0, row_end);
  }
  __device__ __forceinline__ void store_rows(float* hi, float* lo) const { store_rows_as((int)threadIdx.x - TOFF, hi, lo); }
  __device__ __forceinline__ void store_transposed(float* hi, float* lo) const {
    store_transposed_as((int)threadIdx.x - TOFF, hi, lo);
  }
  __device__ __forceinline__ void load_as(int tid, const float* src, int64_t ld, int32_t row0, int32_t row_end) {
#pragma unroll
    for (int it = 0; it < PER; ++it) {
      int r, c;
      coord_of(tid, it, r, c);
      v[it] = make_float4(0.f, 0.f, 0.f, 0.f);          // rows past the segment end are zero
      // volatile asm: the load is issued HERE (a plain __ldg may be sunk to its first use, which
      // would put the whole L2 latency back on the critical path of the next iteration)
      if (row0 + r < row_end) {
        const float* ptr = src + (int64_t)(row0 + r) * ld + 4 * c;
        asm volatile("ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(v[it].x), "=f"(v[it].y), "=f"(v[it].z), "=f"(v[it].w) : "l"(ptr));
      }
    }
  }
  __device__ __forceinline__ void scale(float mul) {
#pragma unroll
    for (int it = 0; it < PER; ++it) {
      v[it].x *= mul; v[it].y *= mul; v[it].z *= mul; v[it].w *= mul;
    }
  }
  // hi / lo K-major tiles with MN index = row, K index = feature
  __device__ __forceinline__ void store_rows_as(int tid, float* hi, float* lo) const {
    constexpr int SBO = CH * 32;
#pragma unroll
    for (int it = 0; it < PER; ++it) {
      int r, c;
      coord_of(tid, it, r, c);
      float4 h, l;
      split_tf32(v[it].x, h.x, l.x); split_tf32(v[it].y, h.y, l.y);
      split_tf32(v[it].z, h.z, l.z); split_tf32(v[it].w, h.w, l.w);
      const int off = core_off(r, c, SBO);
      *reinterpret_cast<float4*>(hi + off) = h;
      *reinterpret_cast<float4*>(lo + off) = l;
    }
  }
  // float offset of 16-byte chunk c of row r in the MN-major (SWIZZLE_128B_BASE32B) tile of the matrix: one
  // [ROWS x 32] block per 32 columns (LBO = ROWS * 128 bytes), 32-byte chunks of a row XOR-swizzled with r & 3
  __device__ __forceinline__ static int mn_off(int r, int c) {
    return (c >> 3) * (ROWS * 32) + r * 32 + (((((c & 7) >> 1) ^ (r & 3)) << 3) | ((c & 1) << 2));
  }
  // hi / lo tiles for BOTH uses of a [ROWS x 32] matrix X: K-major (operand X: MN index = row, K = feature)
  // and MN-major (operand X^T: MN index = feature, K = row).  One split, four conflict-free 16-byte stores
  // (the lanes of a warp cover 4 rows x 128 bytes).
  __device__ __forceinline__ void store_rows_and_mn_as(int tid, float* hi, float* lo, float* hi_mn, float* lo_mn) const {
    static_assert(D == 32, "the MN-major tile is one 32-element atom wide");
    constexpr int SBO = CH * 32;
#pragma unroll
    for (int it = 0; it < PER; ++it) {
      int r, c;
      coord_of(tid, it, r, c);
      float4 h, l;
      split_tf32(v[it].x, h.x, l.x); split_tf32(v[it].y, h.y, l.y);
      split_tf32(v[it].z, h.z, l.z); split_tf32(v[it].w, h.w, l.w);
      const int off = core_off(r, c, SBO);
      *reinterpret_cast<float4*>(hi + off) = h;
      *reinterpret_cast<float4*>(lo + off) = l;
      const int moff = mn_off(r, c);
      *reinterpret_cast<float4*>(hi_mn + moff) = h;
      *reinterpret_cast<float4*>(lo_mn + moff) = l;
    }
  }
  // the MN-major tile alone
  __device__ __forceinline__ void store_mn(float* hi_mn, float* lo_mn) const { store_mn_as((int)threadIdx.x - TOFF, hi_mn, lo_mn); }
  __device__ __forceinline__ void store_mn_as(int tid, float* hi_mn, float* lo_mn) const {
    static_assert(D % 32 == 0, "MN-major tiles are made of 32-column blocks");
#pragma unroll
    for (int it = 0; it < PER; ++it) {
      int r, c;
      coord_of(tid, it, r, c);
      float4 h, l;
      split_tf32(v[it].x, h.x, l.x); split_tf32(v[it].y, h.y, l.y);
      split_tf32(v[it].z, h.z, l.z); split_tf32(v[it].w, h.w, l.w);
      const int moff = mn_off(r, c);
      *reinterpret_cast<float4*>(hi_mn + moff) = h;
      *reinterpret_cast<float4*>(lo_mn + moff) = l;
    }
  }
  // hi / lo K-major tiles of the TRANSPOSE: MN index = feature (D rows), K index = row (ROWS columns).
  // Lanes l, l^1, l^2, l^3 hold the same 16-byte column of 4 consecutive rows; a 4x4 register
  // transpose (4 shuffles) lets every lane write one 16-byte chunk: feature 4c + (l & 3), 4 rows.
  __device__ __forceinline__ void store_transposed_as(int tid, float* hi, float* lo) const {
    constexpr int SBO = (ROWS / 4) * 32;
    const int lane = threadIdx.x & 31;
    const int rr = lane & 3;                           // my row inside its group of 4
#pragma unroll
    for (int it = 0; it < PER; ++it) {
      int r, c;
      coord_of(tid, it, r, c);
      // 4x4 transpose across lanes l, l^1, l^2, l^3 in two butterfly stages; only selects (no
      // lane-dependent register indexing, which the compiler would turn into divergent branches)
      float w[4] = {v[it].x, v[it].y, v[it].z, v[it].w};
      {
        const bool odd = (lane & 1) != 0;
        const float r0 = __shfl_xor_sync(0xffffffffu, odd ? w[0] : w[1], 1);
        const float r1 = __shfl_xor_sync(0xffffffffu, odd ? w[2] : w[3], 1);
        w[0] = odd ? r0 : w[0]; w[1] = odd ? w[1] : r0;
        w[2] = odd ? r1 : w[2]; w[3] = odd ? w[3] : r1;
      }
      {
        const bool up = (lane & 2) != 0;
        const float r0 = __shfl_xor_sync(0xffffffffu, up ? w[0] : w[2], 2);
        const float r1 = __shfl_xor_sync(0xffffffffu, up ? w[1] : w[3], 2);
        w[0] = up ? r0 : w[0]; w[2] = up ? w[2] : r0;
        w[1] = up ? r1 : w[1]; w[3] = up ? w[3] : r1;
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

// 2^x for x <= 0 (and -inf): one MUFU instruction
__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Operand descriptors are loop invariant: build them once, advance the 14-bit start-address field
// (units of 16 bytes) by a constant per K step.
struct OperandDesc {
  uint64_t hi, lo;
};
template <int KD>
__device__ __forceinline__ OperandDesc operand_desc(const float* hi, const float* lo) {
  constexpr uint32_t sbo = (KD / 4) * 128;
  return {smem_desc(smem_u32(hi), kLbo, sbo), smem_desc(smem_u32(lo), kLbo, sbo)};
}

// K-major descriptors of the same tile pair (operand X itself: MN index = row, K index = column < 32)
__device__ __forceinline__ OperandDesc operand_desc_k32(const float* hi, const float* lo) {
  return {smem_desc_mn(smem_u32(hi), 1024), smem_desc_mn(smem_u32(lo), 1024)};
}
// descriptors of an MN-major tile pair (see smem_desc_mn): [ROWS rows (k)] x [32-column blocks (mn)]
template <int ROWS = 32>
__device__ __forceinline__ OperandDesc operand_desc_mn(const float* hi, const float* lo) {
  return {smem_desc_mn(smem_u32(hi), ROWS * 128), smem_desc_mn(smem_u32(lo), ROWS * 128)};
}

// 3xTF32 product  D (+)= A B^T  with A, B in shared memory ([rows x KD] K-major tiles, hi / lo).
template <int KD>
__device__ __forceinline__ void mma3_ss(uint32_t d_tmem, const OperandDesc& a, const OperandDesc& b,
                                        uint32_t idesc, bool accumulate) {
#pragma unroll
  for (int j = 0; j < KD / 8; ++j) {
    const uint64_t adv = static_cast<uint64_t>(j * 2 * kLbo / 16);   // two 16-byte K chunks per MMA
    mma_tf32_ss(d_tmem, a.lo + adv, b.hi + adv, idesc, accumulate || j > 0);
    mma_tf32_ss(d_tmem, a.hi + adv, b.lo + adv, idesc, 1);
    mma_tf32_ss(d_tmem, a.hi + adv, b.hi + adv, idesc, 1);
  }
}
// 3xTF32 product  D (+)= A B^T  with A (hi, lo) in TMEM (lane = row, column = k, KD columns) and
// B = [N x KD] K-major tile in shared memory.
template <int KD>
__device__ __forceinline__ void mma3_ts(uint32_t d_tmem, uint32_t a_hi_tmem, uint32_t a_lo_tmem,
                                        const OperandDesc& b, uint32_t idesc, bool accumulate) {
#pragma unroll
  for (int j = 0; j < KD / 8; ++j) {
    const uint64_t adv = static_cast<uint64_t>(j * 2 * kLbo / 16);
    const uint32_t col = j * 8;                        // 8 k values = 8 TMEM columns per MMA
    mma_tf32_ts(d_tmem, a_lo_tmem + col, b.hi + adv, idesc, accumulate || j > 0);
    mma_tf32_ts(d_tmem, a_hi_tmem + col, b.lo + adv, idesc, 1);
    mma_tf32_ts(d_tmem, a_hi_tmem + col, b.hi + adv, idesc, 1);
  }
}

// D (+)= A^T B with A = [KD rows (k) x M] and B = [KD rows (k) x N] row-major MN-major tiles in shared memory
// (a "TN" product: nothing is transposed on the way in); `idesc` must carry both MN-major bits.
template <int KD>
__device__ __forceinline__ void mma3_ss_mn(uint32_t d_tmem, const OperandDesc& a, const OperandDesc& b,
                                           uint32_t idesc, bool accumulate) {
#pragma unroll
  for (int j = 0; j < KD / 8; ++j) {
    const uint64_t adv = static_cast<uint64_t>(j * (8 * 128 / 16));   // 8 rows of 128 bytes per MMA
    mma_tf32_ss(d_tmem, a.lo + adv, b.hi + adv, idesc, accumulate || j > 0);
    mma_tf32_ss(d_tmem, a.hi + adv, b.lo + adv, idesc, 1);
    mma_tf32_ss(d_tmem, a.hi + adv, b.hi + adv, idesc, 1);
  }
}
// D (+)= A B^T with A = [M x KD] and B = [N x KD] row tiles in the SWIZZLE_128B_BASE32B layout, both read K-major
// (operand_desc_k32): 32 bytes of every row per MMA.
template <int KD>
__device__ __forceinline__ void mma3_ss_k32(uint32_t d_tmem, const OperandDesc& a, const OperandDesc& b,
                                            uint32_t idesc, bool accumulate) {
  static_assert(KD <= 32, "one 128-byte row holds 32 k values");
#pragma unroll
  for (int j = 0; j < KD / 8; ++j) {
    const uint64_t adv = static_cast<uint64_t>(j * 2);
    mma_tf32_ss(d_tmem, a.lo + adv, b.hi + adv, idesc, accumulate || j > 0);
    mma_tf32_ss(d_tmem, a.hi + adv, b.lo + adv, idesc, 1);
    mma_tf32_ss(d_tmem, a.hi + adv, b.hi + adv, idesc, 1);
  }
}
// D (+)= A B with A = [M x KD] row tile read K-major and B = [KD rows (k) x N] MN-major blocks; `idesc` must carry
// kIdescBMajorMN.
template <int KD>
__device__ __forceinline__ void mma3_ss_k32_mn(uint32_t d_tmem, const OperandDesc& a, const OperandDesc& b,
                                               uint32_t idesc, bool accumulate) {
  static_assert(KD <= 32, "one 128-byte row holds 32 k values");
#pragma unroll
  for (int j = 0; j < KD / 8; ++j) {
    const uint64_t adv_a = static_cast<uint64_t>(j * 2);
    const uint64_t adv_b = static_cast<uint64_t>(j * (8 * 128 / 16));
    mma_tf32_ss(d_tmem, a.lo + adv_a, b.hi + adv_b, idesc, accumulate || j > 0);
    mma_tf32_ss(d_tmem, a.hi + adv_a, b.lo + adv_b, idesc, 1);
    mma_tf32_ss(d_tmem, a.hi + adv_a, b.hi + adv_b, idesc, 1);
  }
}
// D (+)= A B with A = [M x KD] K-major and B = [KD rows (k) x N] MN-major; `idesc` must carry kIdescBMajorMN.
template <int KD>
__device__ __forceinline__ void mma3_ss_bmn(uint32_t d_tmem, const OperandDesc& a, const OperandDesc& b,
                                            uint32_t idesc, bool accumulate) {
#pragma unroll
  for (int j = 0; j < KD / 8; ++j) {
    const uint64_t adv_a = static_cast<uint64_t>(j * 2 * kLbo / 16);
    const uint64_t adv_b = static_cast<uint64_t>(j * (8 * 128 / 16));
    mma_tf32_ss(d_tmem, a.lo + adv_a, b.hi + adv_b, idesc, accumulate || j > 0);
    mma_tf32_ss(d_tmem, a.hi + adv_a, b.lo + adv_b, idesc, 1);
    mma_tf32_ss(d_tmem, a.hi + adv_a, b.hi + adv_b, idesc, 1);
  }
}
// The same with B = an MN-major tile ([KD rows = k] x [32 = n]); `idesc` must carry kIdescBMajorMN.
template <int KD>
__device__ __forceinline__ void mma3_ts_mn(uint32_t d_tmem, uint32_t a_hi_tmem, uint32_t a_lo_tmem,
                                           const OperandDesc& b, uint32_t idesc, bool accumulate) {
#pragma unroll
  for (int j = 0; j < KD / 8; ++j) {
    const uint64_t adv = static_cast<uint64_t>(j * (8 * 128 / 16));   // 8 rows of 128 bytes per MMA
    const uint32_t col = j * 8;
    mma_tf32_ts(d_tmem, a_lo_tmem + col, b.hi + adv, idesc, accumulate || j > 0);
    mma_tf32_ts(d_tmem, a_hi_tmem + col, b.lo + adv, idesc, 1);
    mma_tf32_ts(d_tmem, a_hi_tmem + col, b.hi + adv, idesc, 1);
  }
}

// D (+)= A B^T with A (hi, lo) in TMEM and B = a [N x KD <= 32] row tile in the SWIZZLE_128B_BASE32B layout used
// K-major (operand_desc_k32); plain K-major `idesc`.
template <int KD>
__device__ __forceinline__ void mma3_ts_k32(uint32_t d_tmem, uint32_t a_hi_tmem, uint32_t a_lo_tmem,
                                            const OperandDesc& b, uint32_t idesc, bool accumulate) {
  static_assert(KD <= 32, "one 128-byte row holds 32 k values");
#pragma unroll
  for (int j = 0; j < KD / 8; ++j) {
    const uint64_t adv = static_cast<uint64_t>(j * 2);                // 32 bytes of every row per MMA
    const uint32_t col = j * 8;
    mma_tf32_ts(d_tmem, a_lo_tmem + col, b.hi + adv, idesc, accumulate || j > 0);
    mma_tf32_ts(d_tmem, a_hi_tmem + col, b.lo + adv, idesc, 1);
    mma_tf32_ts(d_tmem, a_hi_tmem + col, b.hi + adv, idesc, 1);
  }
}

// This thread's half (D/2 features) of row `row` of a [*, ld] matrix, times `mul`, split and written
// to TMEM as the A operand of later MMAs (lane = row, column = feature).  Warp-collective.
template <int D>
__device__ __forceinline__ void row_to_tmem(uint32_t t_hi, uint32_t t_lo, const float* src, int64_t ld,
                                            int64_t row, bool ok, int half, float mul) {
  constexpr int DH = D / 2;
  const float* ptr = src + row * ld + half * DH;
#pragma unroll
  for (int c0 = 0; c0 < DH; c0 += 8) {
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
    if (ok) {
      a = __ldg(reinterpret_cast<const float4*>(ptr + c0));
      b = __ldg(reinterpret_cast<const float4*>(ptr + c0 + 4));
    }
    const float x[8] = {a.x * mul, a.y * mul, a.z * mul, a.w * mul, b.x * mul, b.y * mul, b.z * mul, b.w * mul};
    uint32_t hi[8], lo[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float h_, l_;
      split_tf32(x[j], h_, l_);
      hi[j] = __float_as_uint(h_);
      lo[j] = __float_as_uint(l_);
    }
    tmem_st8(t_hi + half * DH + c0, hi);
    tmem_st8(t_lo + half * DH + c0, lo);
  }
  tmem_st_wait();
}

struct CtaSetup {
  uint32_t tmem;        // base of this CTA's TMEM allocation
  uint32_t lane_base;   // this warp's lane quarter, in the address's lane field
  int row_local;        // TMEM lane (= tile row) of this thread
  int half;             // which 32-column half of a 64-column block this thread owns
};

__device__ __forceinline__ CtaSetup cta_setup(uint64_t* bar, uint32_t* tmem_slot, uint32_t ncols) {
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc(tmem_slot, ncols);
  if (threadIdx.x == 0) mbar_init(bar, 1);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  CtaSetup s;
  s.tmem = *tmem_slot;
  s.lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
  s.row_local = (warp & 3) * 32 + (threadIdx.x & 31);
  s.half = warp >> 2;
  return s;
}
__device__ __forceinline__ void cta_teardown(uint32_t tmem, uint32_t ncols) {
  fence_before_sync();
  __syncthreads();
  if ((threadIdx.x >> 5) == 0) tmem_dealloc(tmem, ncols);
}
// everyone's shared-memory / TMEM writes done -> the elected thread may issue MMAs
__device__ __forceinline__ void publish_operands() {
  fence_smem_to_async();
  fence_before_sync();
  __syncthreads();
}

}  // namespace tc
}  // namespace hgk
