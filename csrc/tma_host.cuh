// Host-side construction of TMA tensor maps.  cuTensorMapEncodeTiled lives in the
// driver (libcuda); it is fetched through the runtime's entry-point lookup so the
// library links against cudart only and still loads on a machine without a GPU.
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace tma {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// Generic N-d map.  dims/box innermost first; strides_bytes[i] is the stride of dim i+1.
inline bool make(CUtensorMap* out, CUtensorMapDataType dt, int rank, const void* base, const uint64_t* dims,
                 const uint64_t* strides_bytes, const uint32_t* box, CUtensorMapSwizzle swz) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  cuuint64_t gd[5], gs[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) gs[i] = strides_bytes[i];
  CUresult r = fn(out, dt, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// Row-major matrix [rows, cols] (cols contiguous), box = [box_rows, box_cols], 128B swizzle:
// box_cols * elem_size must be exactly 128 bytes.
inline bool make_2d_sw128(CUtensorMap* out, CUtensorMapDataType dt, int elem_bytes, const void* base, uint64_t rows,
                          uint64_t cols, uint64_t row_pitch_bytes, uint32_t box_rows, uint32_t box_cols,
                          CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
  uint64_t dims[2] = {cols, rows};
  uint64_t strides[1] = {row_pitch_bytes};
  uint32_t box[2] = {box_cols, box_rows};
  (void)elem_bytes;
  return make(out, dt, 2, base, dims, strides, box, swz);
}

}  // namespace tma
