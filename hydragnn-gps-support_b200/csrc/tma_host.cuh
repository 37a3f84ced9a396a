// Host side of the TMA loads (cp.async.bulk.tensor): tensor maps over row matrices in THIS GPU's memory.
// cuTensorMapEncodeTiled is a driver entry point; it is fetched through the runtime
// (cudaGetDriverEntryPoint), so libhgk.so does not link against libcuda.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace hgk {
namespace tma {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult status;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &status) != cudaSuccess ||
        status != cudaDriverEntryPointSuccess)
      ptr = nullptr;
    (void)cudaGetLastError();
    return reinterpret_cast<EncodeTiledFn>(ptr);
  }();
  return fn;
}

// fp32 row matrix [rows, cols] with row stride ld (floats), boxes of [box_rows x 32 columns] = rows of 128 bytes in
// shared memory; rows past the end read as zeros.  false: the driver refused (the caller falls back).
inline bool rows_map(CUtensorMap* map, const float* base, int64_t rows, int64_t cols, int64_t ld, uint32_t box_rows,
                     CUtensorMapSwizzle swizzle) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (fn == nullptr || rows <= 0 || cols < 32) return false;
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)ld * sizeof(float)};
  const cuuint32_t box[2] = {32, box_rows};
  const cuuint32_t estr[2] = {1, 1};
  return fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
// ... as the SWIZZLE_128B_BASE32B operand block of tc_common.cuh (32-byte chunks of a row XOR-ed with row & 3)
inline bool rows_map_32(CUtensorMap* map, const float* base, int64_t rows, int64_t cols, int64_t ld, uint32_t box_rows) {
  return rows_map(map, base, rows, cols, ld, box_rows, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
}
// ... with the classic 128-byte swizzle (16-byte chunks XOR-ed with row & 7): conflict-free for threads that each
// read one row (8 consecutive rows hit 8 different chunks)
inline bool rows_map_128b(CUtensorMap* map, const float* base, int64_t rows, int64_t cols, int64_t ld, uint32_t box_rows) {
  return rows_map(map, base, rows, cols, ld, box_rows, CU_TENSOR_MAP_SWIZZLE_128B);
}

// output row matrix written in boxes of [box_rows x 16 columns] = rows of 64 bytes in shared memory, 64-byte swizzle
// (16-byte chunks XOR-ed with (row / 2) & 3): conflict-free for threads that each write one row
inline bool rows_map_store16(CUtensorMap* map, float* base, int64_t rows, int64_t cols, int64_t ld, uint32_t box_rows) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (fn == nullptr || rows <= 0 || cols < 16) return false;
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)ld * sizeof(float)};
  const cuuint32_t box[2] = {16, box_rows};
  const cuuint32_t estr[2] = {1, 1};
  return fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace tma
}  // namespace hgk
