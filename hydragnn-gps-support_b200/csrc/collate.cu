// Device-side collate: a mini-batch is gathered out of a packed arena that holds every sample of a
// split (reference: torch_geometric Batch.from_data_list on the host for every batch,
// hydragnn/preprocess/load_data.py:225-326, serialized_dataset_loader.py:110-212 -- SURVEY 8(f) rank 1).
//
// A field of the dataset is a row matrix [R, W] (4-byte words) with per-sample row ranges
// arena_off[s] .. arena_off[s+1]; the batch takes the samples idx[0..B) in order:
//     out rows out_ptr[b] .. out_ptr[b+1]  =  arena rows src_off[b] .. src_off[b] + (out_ptr[b+1] - out_ptr[b])
// One launch per field; integer / byte work only, bit-exact.
#include "hgk_common.cuh"

namespace hgk {

__device__ __forceinline__ int find_segment(const int64_t* __restrict__ ptr, int B, int64_t row) {
  int lo = 0, hi = B;                       // ptr[lo] <= row < ptr[hi]
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(ptr + mid) <= row) lo = mid; else hi = mid;
  }
  return lo;
}

// one thread per 4-byte word of the output
__global__ void __launch_bounds__(256)
k_collate_rows(uint32_t* __restrict__ out, const uint32_t* __restrict__ arena, const int64_t* __restrict__ out_ptr,
               const int64_t* __restrict__ src_off, int B, int W, int64_t total_words, int64_t* __restrict__ seg_id) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total_words) return;
  const int64_t row = i / W;
  const int w = (int)(i - row * W);
  const int b = find_segment(out_ptr, B, row);
  const int64_t src_row = __ldg(src_off + b) + (row - __ldg(out_ptr + b));
  out[i] = __ldg(arena + src_row * W + w);
  if (seg_id != nullptr && w == 0) seg_id[row] = b;
}

// edge lists: the arena stores (src, dst) pairs with node ids local to their sample; the batch needs
// edge_index [2, E_out] with the batch's node offsets added
__global__ void __launch_bounds__(256)
k_collate_edges(int64_t* __restrict__ out, const int64_t* __restrict__ pairs, const int64_t* __restrict__ out_ptr,
                const int64_t* __restrict__ src_off, const int64_t* __restrict__ node_ptr, int B, int64_t E_out) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E_out) return;
  const int b = find_segment(out_ptr, B, e);
  const int64_t src_e = __ldg(src_off + b) + (e - __ldg(out_ptr + b));
  const int64_t base = __ldg(node_ptr + b);
  out[e] = __ldg(pairs + 2 * src_e) + base;
  out[E_out + e] = __ldg(pairs + 2 * src_e + 1) + base;
}

}  // namespace hgk

using namespace hgk;

extern "C" int hgk_collate_rows(void* out, const void* arena, const int64_t* out_ptr, const int64_t* src_off,
                                int64_t B, int32_t row_bytes, int64_t* seg_id, int64_t total_rows, void* stream) {
  HGK_REQUIRE(B >= 0 && total_rows >= 0 && row_bytes > 0 && row_bytes % 4 == 0, "collate_rows: bad sizes");
  if (total_rows == 0 || B == 0) return HGK_OK;
  HGK_REQUIRE(out && arena && out_ptr && src_off, "collate_rows: NULL operand");
  HGK_REQUIRE(B < (1ll << 31), "collate_rows: too many segments");
  const int W = row_bytes / 4;
  const int64_t words = total_rows * W;
  k_collate_rows<<<(unsigned)ceil_div(words, 256), 256, 0, (cudaStream_t)stream>>>(
      (uint32_t*)out, (const uint32_t*)arena, out_ptr, src_off, (int)B, W, words, seg_id);
  return check_launch("collate_rows");
}

extern "C" int hgk_collate_edges(int64_t* edge_index, const int64_t* pairs, const int64_t* out_ptr,
                                 const int64_t* src_off, const int64_t* node_ptr, int64_t B, int64_t E_out,
                                 void* stream) {
  HGK_REQUIRE(B >= 0 && E_out >= 0, "collate_edges: bad sizes");
  if (E_out == 0 || B == 0) return HGK_OK;
  HGK_REQUIRE(edge_index && pairs && out_ptr && src_off && node_ptr, "collate_edges: NULL operand");
  k_collate_edges<<<(unsigned)ceil_div(E_out, 256), 256, 0, (cudaStream_t)stream>>>(edge_index, pairs, out_ptr, src_off,
                                                                                  node_ptr, (int)B, E_out);
  return check_launch("collate_edges");
}
