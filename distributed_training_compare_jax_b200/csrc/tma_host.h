// Host-side construction of TMA tensor maps.  The driver entry point is fetched through the runtime
// (cudaGetDriverEntryPoint) so the library does not link against libcuda directly.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "common.cuh"

namespace dtc {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn get_encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = (EncodeTiledFn)p;
  }
  return fn;
}

// 2-D bf16 tensor [rows, cols] with `ld` elements between rows (cols contiguous).  The box is
// box_cols x box_rows; swizzle must match box_cols*2 bytes (128/64/32) — out-of-bounds reads are zero.
inline int make_tmap_2d_bf16(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld,
                             uint32_t box_rows, uint32_t box_cols, CUtensorMapSwizzle swz) {
  EncodeTiledFn fn = get_encode_tiled();
  if (!fn) {
    dtc_set_error("cuTensorMapEncodeTiled entry point unavailable");
    return -2;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    dtc_set_error("cuTensorMapEncodeTiled failed (%d): ptr=%p rows=%llu cols=%llu ld=%llu box=%ux%u", (int)r, ptr,
                  (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld, box_rows, box_cols);
    return -3;
  }
  return 0;
}

// 2-D fp32 tensor (row-major, `ld` elements between rows; ld * 4 bytes must be a multiple of 16): the target of TMA
// reduce-add epilogues.  box_cols * 4 bytes <= 128 with the 128-byte swizzle.
inline int make_tmap_2d_f32(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                            uint32_t box_cols, CUtensorMapSwizzle swz) {
  EncodeTiledFn fn = get_encode_tiled();
  if (!fn) {
    dtc_set_error("cuTensorMapEncodeTiled entry point unavailable");
    return -2;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 4};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    dtc_set_error("cuTensorMapEncodeTiled (fp32) failed (%d): ptr=%p rows=%llu cols=%llu ld=%llu box=%ux%u", (int)r, ptr,
                  (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld, box_rows, box_cols);
    return -3;
  }
  return 0;
}

// 3-D bf16 tensor [d2, d1, d0] (d0 contiguous) with element strides ld1 (between d1 rows) and ld2.
inline int make_tmap_3d_bf16(CUtensorMap* map, const void* ptr, uint64_t d2, uint64_t d1, uint64_t d0, uint64_t ld2,
                             uint64_t ld1, uint32_t box2, uint32_t box1, uint32_t box0, CUtensorMapSwizzle swz) {
  EncodeTiledFn fn = get_encode_tiled();
  if (!fn) {
    dtc_set_error("cuTensorMapEncodeTiled entry point unavailable");
    return -2;
  }
  cuuint64_t dims[3] = {d0, d1, d2};
  cuuint64_t strides[2] = {ld1 * 2, ld2 * 2};
  cuuint32_t box[3] = {box0, box1, box2};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    dtc_set_error("cuTensorMapEncodeTiled(3d) failed (%d)", (int)r);
    return -3;
  }
  return 0;
}

}  // namespace dtc
