// 2-D+ TMA (cp.async.bulk.tensor, SASS UTMALDG) helpers: host-side tensor-map encoding through the driver entry point
// (no link against libcuda: cudaGetDriverEntryPoint) and the device-side tile load.
#pragma once
#include <cuda.h>

#include "m3r_common.cuh"
#include "tma.cuh"

namespace m3r {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static inline EncodeTiledFn tensormap_encoder() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess &&
        qr == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

// Image (b, h, w, c) of `elem_bytes`-byte elements, channels innermost, as a rank-4 tensor {c, w, h, b}; box = {c, box_w,
// box_h, 1}: one box is a box_h x box_w window of whole pixels, dense in shared memory (row pitch box_w * c * elem_bytes),
// out-of-image pixels zero-filled by the copy engine.  Returns false if the driver refuses (then callers keep their
// gather path).
static inline bool encode_image_tiles(CUtensorMap* map, const void* base, int b, int h, int w, int c, int elem_bytes,
                                      int box_w, int box_h) {
  EncodeTiledFn enc = tensormap_encoder();
  if (enc == nullptr) return false;
  const CUtensorMapDataType dt = elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  const cuuint64_t dims[4] = {(cuuint64_t)c, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)b};
  const cuuint64_t pix = (cuuint64_t)c * elem_bytes;
  const cuuint64_t strides[3] = {pix, pix * w, pix * w * h};  // bytes, dims 1..3 (multiples of 16 required)
  const cuuint32_t box[4] = {(cuuint32_t)c, (cuuint32_t)box_w, (cuuint32_t)box_h, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  return enc(map, dt, 4, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// Plane (b, h, row) of fp32 with the whole image row as the innermost dimension: rank-4 tensor {row, h, b, 1}, box =
// {box_elems, box_h, 1, 1}.  One request of the copy engine is then box_elems * 4 contiguous bytes (at most 256 elements)
// instead of one pixel: a box over {c, w, h, b} with c = 4 floats costs a request per pixel, and the engine's request rate,
// not its bandwidth, then sets the pace.  row * 4 and box_elems * 4 must be multiples of 16.
static inline bool encode_row_tiles_f32(CUtensorMap* map, const void* base, int b, int h, int row, int box_elems,
                                        int box_h) {
  EncodeTiledFn enc = tensormap_encoder();
  if (enc == nullptr || (row & 3) != 0 || (box_elems & 3) != 0 || box_elems > 256) return false;
  const cuuint64_t dims[4] = {(cuuint64_t)row, (cuuint64_t)h, (cuuint64_t)b, 1};
  const cuuint64_t rb = (cuuint64_t)row * 4;
  const cuuint64_t strides[3] = {rb, rb * h, rb * h * b};
  const cuuint32_t box[4] = {(cuuint32_t)box_elems, (cuuint32_t)box_h, 1, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  return enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<void*>(base), dims, strides, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// one box -> shared memory (128-byte aligned), completion on the mbarrier; coordinates may be negative / past the edge
__device__ __forceinline__ void tma_load_4d(void* dst, const CUtensorMap* map, int c0, int c1, int c2, int c3,
                                            unsigned long long* bar) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

}  // namespace m3r
