// Probe: which shared-memory element does tcgen05.mma kind::tf32 read for B element (n, k) when the
// instruction descriptor says B is MN-major and the matrix descriptor is SWIZZLE_NONE with given LBO / SBO?
// A is one-hot (A[m][k] = [k == m % KTOT]), B's shared memory holds its own float offsets, so
// D[m][n] = offset of the element the hardware took as B(n, k = m % KTOT).
//   nvcc -gencode arch=compute_100a,code=sm_100a -I../../hydragnn-gps-support_b200/csrc -o mn_major_probe mn_major_probe.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "tc_common.cuh"
using namespace hgk::tc;

constexpr int KTOT = 32, N = 32;

__global__ void __launch_bounds__(128, 1) k_probe(float* out, uint32_t b_lbo, uint32_t b_sbo, uint32_t b_adv,
                                                  uint32_t b_major, uint32_t a_from_tmem, uint32_t layout, uint32_t extra_idesc) {
  __shared__ __align__(1024) float sA[128 * KTOT];
  __shared__ __align__(1024) float sB[2048];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const CtaSetup cs = cta_setup(&bar, &tmem_slot, 64);
  const uint32_t tmem = cs.tmem;
  const int row = cs.row_local;
  for (int k = 0; k < KTOT; ++k) sA[core_off(row, k >> 2, (KTOT / 4) * 32) + (k & 3)] = (k == row % KTOT) ? 1.f : 0.f;
  for (int i = threadIdx.x; i < 2048; i += 128) sB[i] = (float)i;
  if (a_from_tmem) {
    uint32_t v[32];
    for (int k = 0; k < 32; ++k) v[k] = __float_as_uint((k == row % KTOT) ? 1.f : 0.f);
    tmem_st32(tmem + cs.lane_base + 32, v);
    tmem_st_wait();
  }
  {
    uint32_t v[32];
    for (int k = 0; k < 32; ++k) v[k] = __float_as_uint(-7.f);
    tmem_st32(tmem + cs.lane_base, v);
    tmem_st_wait();
  }
  publish_operands();
  if (threadIdx.x == 0) {
    fence_after_sync();
    const uint32_t idesc = idesc_tf32(128, N) | (b_major << 16) | extra_idesc;
    const uint64_t a0 = smem_desc(smem_u32(sA), kLbo, (KTOT / 4) * 128);
    for (int j = 0; j < KTOT / 8; ++j) {
      const uint64_t bd = smem_desc(smem_u32(sB) + j * b_adv, b_lbo, b_sbo) | ((uint64_t)layout << 61);
      if (a_from_tmem) mma_tf32_ts(tmem, tmem + 32 + j * 8, bd, idesc, j > 0);
      else mma_tf32_ss(tmem, a0 + (uint64_t)(j * 2 * kLbo / 16), bd, idesc, j > 0);
    }
    mma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  fence_after_sync();
  uint32_t r[32];
  tmem_ld32(tmem + cs.lane_base, r);
  tmem_ld_wait();
  for (int n = 0; n < N; ++n) out[row * N + n] = __uint_as_float(r[n]);
  cta_teardown(tmem, 64);
}

int main(int argc, char** argv) {
  float* d;
  std::vector<float> h(128 * N);
  struct Cfg { uint32_t lbo, sbo, adv, major, ts, layout, extra; const char* name; };
  const Cfg cfgs[] = {
      {1024, 1024, 32, 0, 0, 1, 0, "K-major BASE32B lbo=1024 sbo=1024 adv=32"},
      {4096, 1024, 32, 0, 0, 1, 0, "K-major BASE32B lbo=4096 sbo=1024 adv=32"},
      {1024, 512, 32, 0, 0, 1, 0, "K-major BASE32B lbo=1024 sbo=512 adv=32"},
      {128, 1024, 32, 0, 0, 1, 0, "K-major BASE32B lbo=128 sbo=1024 adv=32"},
      {1024, 1024, 32, 0, 1, 1, 0, "K-major BASE32B lbo=1024 sbo=1024 adv=32 (ts)"},
      {1024, 1024, 64, 0, 0, 1, 0, "K-major BASE32B lbo=1024 sbo=1024 adv=64"},
  };
  const int which = argc > 1 ? atoi(argv[1]) : 0;
  if (which >= (int)(sizeof(cfgs) / sizeof(cfgs[0]))) return 0;
  const Cfg& c = cfgs[which];
  const int ks[] = {0, 1, 2, 3, 4, 5, 7, 8, 9, 16, 31};
  const int ns[] = {0, 1, 2, 3, 4, 5, 7, 8, 12, 16, 24, 31};
  cudaMalloc(&d, 128 * N * 4);
  cudaMemset(d, 0, 128 * N * 4);
  k_probe<<<1, 128>>>(d, c.lbo, c.sbo, c.adv, c.major, c.ts, c.layout, c.extra);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("== %s: CUDA error %s\n", c.name, cudaGetErrorString(e)); return 0; }
  cudaMemcpy(h.data(), d, 128 * N * 4, cudaMemcpyDeviceToHost);
  printf("== %s\n      n:", c.name);
  for (int n : ns) printf("%5d", n);
  printf("\n");
  for (int k : ks) {
    printf("  k=%2d: ", k);
    for (int n : ns) printf("%5g", h[k * N + n]);
    printf("\n");
  }
  return 0;
}
