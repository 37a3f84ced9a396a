// Probe: how does tcgen05.mma kind::tf32 convert an fp32 operand: truncation of the 13 low mantissa bits, or
// rounding?  A is one-hot (exact), B holds fp32 values with low bits set; D = the value the hardware used.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "tc_common.cuh"
using namespace hgk::tc;
constexpr int KTOT = 32, N = 32;

__global__ void __launch_bounds__(128, 1) k_probe(float* out, const float* bvals, int ts) {
  __shared__ __align__(1024) float sA[128 * KTOT];
  __shared__ __align__(1024) float sB[KTOT * N];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const CtaSetup cs = cta_setup(&bar, &tmem_slot, 64);
  const uint32_t tmem = cs.tmem;
  const int row = cs.row_local;
  for (int k = 0; k < KTOT; ++k) sA[core_off(row, k >> 2, (KTOT / 4) * 32) + (k & 3)] = (k == row % KTOT) ? 1.f : 0.f;
  // B(n, k) = bvals[k] for every n  (K-major tile: rows = n); TS mode: B(n, k) = [k == 0]
  for (int i = threadIdx.x; i < KTOT * N; i += 128) {
    const int n = i / KTOT, k = i % KTOT;
    sB[core_off(n, k >> 2, (KTOT / 4) * 32) + (k & 3)] = ts ? (k == 0 ? 1.f : 0.f) : bvals[k];
  }
  if (ts) {                                            // A from TMEM: lane m holds bvals[m % 32] at k = 0
    uint32_t v[32];
    for (int k = 0; k < 32; ++k) v[k] = k == 0 ? __float_as_uint(bvals[row % KTOT]) : 0u;
    tmem_st32(tmem + cs.lane_base + 32, v);
    tmem_st_wait();
  }
  publish_operands();
  if (threadIdx.x == 0) {
    fence_after_sync();
    const uint32_t idesc = idesc_tf32(128, N);
    const uint64_t a0 = smem_desc(smem_u32(sA), kLbo, (KTOT / 4) * 128);
    const uint64_t b0 = smem_desc(smem_u32(sB), kLbo, (KTOT / 4) * 128);
    for (int j = 0; j < KTOT / 8; ++j) {
      if (ts) mma_tf32_ts(tmem, tmem + 32 + j * 8, b0 + (uint64_t)(j * 2 * kLbo / 16), idesc, j > 0);
      else mma_tf32_ss(tmem, a0 + (uint64_t)(j * 2 * kLbo / 16), b0 + (uint64_t)(j * 2 * kLbo / 16), idesc, j > 0);
    }
    mma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  fence_after_sync();
  uint32_t r[32];
  tmem_ld32(tmem + cs.lane_base, r);
  tmem_ld_wait();
  if (row < KTOT) out[row] = __uint_as_float(r[0]);       // D[m = k][n = 0] = converted bvals[k]
  cta_teardown(tmem, 64);
}

int main(int argc, char** argv) {
  std::vector<float> b(KTOT), h(KTOT);
  for (int k = 0; k < KTOT; ++k) {
    uint32_t bits = 0x3f800000u | ((uint32_t)k << 8) | 0x1fffu * (k & 1) | ((k & 2) ? 0x1000u : 0u);  // assorted low bits
    if (k >= 16) bits |= 0x80000000u;
    memcpy(&b[k], &bits, 4);
  }
  float *d, *db;
  cudaMalloc(&d, KTOT * 4); cudaMalloc(&db, KTOT * 4);
  cudaMemcpy(db, b.data(), KTOT * 4, cudaMemcpyHostToDevice);
  const int ts = argc > 1 ? atoi(argv[1]) : 0;
  k_probe<<<1, 128>>>(d, db, ts);
  cudaError_t e = cudaDeviceSynchronize();
  printf("operand under test: %s\n", ts ? "A from TMEM (TS)" : "B from shared memory (SS)");
  if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); return 1; }
  cudaMemcpy(h.data(), d, KTOT * 4, cudaMemcpyDeviceToHost);
  int trunc_ok = 0, rn_ok = 0;
  for (int k = 0; k < KTOT; ++k) {
    uint32_t in, out;
    memcpy(&in, &b[k], 4); memcpy(&out, &h[k], 4);
    const uint32_t t = in & 0xffffe000u;
    const uint32_t rn = (in + 0x1000u) & 0xffffe000u;
    trunc_ok += out == t; rn_ok += out == rn;
    printf("in %08x -> used %08x   (truncate %08x, round-half-up %08x)\n", in, out, t, rn);
  }
  printf("matches: truncation %d / %d, rounding %d / %d\n", trunc_ok, KTOT, rn_ok, KTOT);
  return 0;
}
