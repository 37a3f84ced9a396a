// Probe 2: MN-major operands wider than one 32-element atom (SWIZZLE_128B_BASE32B), for B (N = 128) and for A
// from shared memory (M = 128).  One operand is one-hot so that D reveals which shared-memory float the other
// operand's element (mn, k) was read from.
//   mode 0: B MN-major, N = 128, K = 32; A one-hot K-major: D[m][n] = offset of B(n, k = m % 32)
//   mode 1: A MN-major, M = 128, K = 32; B one-hot K-major (N = 32): D[m][n] = offset of A(m, k = n)
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "tc_common.cuh"
using namespace hgk::tc;

constexpr int KTOT = 32;

__global__ void __launch_bounds__(128, 1) k_probe(float* out, int mode, uint32_t lbo, uint32_t sbo, uint32_t adv) {
  __shared__ __align__(1024) float sX[4096];       // the MN-major operand: holds its own offsets (16 KB)
  __shared__ __align__(1024) float sH[128 * KTOT]; // the one-hot K-major operand
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const CtaSetup cs = cta_setup(&bar, &tmem_slot, 128);
  const uint32_t tmem = cs.tmem;
  const int row = cs.row_local;
  // one-hot [128 rows x 32 k]: row r has a one at k = r % 32
  for (int k = 0; k < KTOT; ++k) sH[core_off(row, k >> 2, (KTOT / 4) * 32) + (k & 3)] = (k == row % KTOT) ? 1.f : 0.f;
  for (int i = threadIdx.x; i < 4096; i += 128) sX[i] = (float)(i % 2048) + (i >= 2048 ? 0.5f : 0.f);  // tf32-exact
  publish_operands();
  const int N = mode == 0 ? 128 : 32;
  if (threadIdx.x == 0) {
    fence_after_sync();
    const uint32_t idesc = idesc_tf32(128, N) | (mode == 0 ? (1u << 16) : (1u << 15));
    for (int j = 0; j < KTOT / 8; ++j) {
      const uint64_t xd = smem_desc(smem_u32(sX) + j * adv, lbo, sbo) | ((uint64_t)1 << 61);
      const uint64_t hd = smem_desc(smem_u32(sH), kLbo, (KTOT / 4) * 128) + (uint64_t)(j * 2 * kLbo / 16);
      if (mode == 0) mma_tf32_ss(tmem, hd, xd, idesc, j > 0);
      else mma_tf32_ss(tmem, xd, hd, idesc, j > 0);
    }
    mma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  fence_after_sync();
  for (int c0 = 0; c0 < N; c0 += 32) {
    uint32_t r[32];
    tmem_ld32(tmem + cs.lane_base + c0, r);
    tmem_ld_wait();
    for (int n = 0; n < 32; ++n) out[row * 128 + c0 + n] = __uint_as_float(r[n]);
  }
  cta_teardown(tmem, 128);
}

int main(int argc, char** argv) {
  const int mode = atoi(argv[1]);
  const uint32_t lbo = atoi(argv[2]), sbo = atoi(argv[3]), adv = atoi(argv[4]);
  float* d;
  std::vector<float> h(128 * 128);
  cudaMalloc(&d, 128 * 128 * 4);
  cudaMemset(d, 0, 128 * 128 * 4);
  k_probe<<<1, 128>>>(d, mode, lbo, sbo, adv);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("== mode %d lbo %u sbo %u adv %u: CUDA error %s\n", mode, lbo, sbo, adv, cudaGetErrorString(e)); return 0; }
  cudaMemcpy(h.data(), d, 128 * 128 * 4, cudaMemcpyDeviceToHost);
  // expectation: element (mn, k) at float offset (mn / 32) * lbo/4 + k * 32 + (((mn % 32) / 8) ^ (k % 4)) * 8 + mn % 8
  int bad = 0;
  for (int mn = 0; mn < 128; ++mn)
    for (int k = 0; k < 32; ++k) {
      const int off = (mn / 32) * (int)(lbo / 4) + (k / 8) * (int)(adv / 4) + (k % 8 / 4) * (int)(sbo / 4) + (k % 4) * 32 +
                      ((((mn % 32) / 8) ^ (k % 4)) * 8) + mn % 8;
      const float want = (float)(off % 2048) + (off >= 2048 ? 0.5f : 0.f);
      const float got = mode == 0 ? h[k * 128 + mn] : h[mn * 128 + k];
      bad += got != want;
    }
  printf("== mode %d (%s MN-major) lbo %u sbo %u adv %u: %d mismatches vs the expected mapping\n", mode,
         mode == 0 ? "B, N=128" : "A, M=128", lbo, sbo, adv, bad);
  const int ks[] = {0, 1, 4, 8};
  const int ms[] = {0, 1, 8, 31, 32, 33, 64, 96, 127};
  for (int k : ks) {
    printf("  k=%2d:", k);
    for (int mn : ms) printf(" mn%-3d=%-7g", mn, mode == 0 ? h[k * 128 + mn] : h[mn * 128 + k]);
    printf("\n");
  }
  return 0;
}
