// Probe of the tcgen05 shared-memory descriptor: can an operand start at an arbitrary ROW of a
// swizzled tile (start address + row * pitch, with the descriptor's base-offset field set to
// the row phase)?  The implicit-GEMM convolutions rely on it: one halo patch per image lands in
// shared memory once and each of the 25 filter taps is the same patch read from a shifted
// start row, instead of 25 separate loads.  tests/test_gpu_kernels.py pins the behaviour.
// No counterpart in the reference (its convolutions are cuDNN calls, mpipy.py:156,159).
#include "common.cuh"
#include "sm100.cuh"
#include "tma_host.cuh"

namespace {

using namespace sm100;

__device__ __forceinline__ uint64_t desc_with_base(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t layout,
                                                   uint32_t base_off) {
  return make_smem_desc(addr, lbo, sbo, layout) | ((uint64_t)(base_off & 7) << 49);
}

// mode 0: A K-major, rows of `row_bytes` (64 -> SW64, K=32; 128 -> SW128, K=64); A tile = rows [shift, shift+128)
// mode 1: A MN-major SW128, stored [K rows][128 M]; K = 32 rows [shift, shift+32)
// B is always K-major [64 N rows][K], same swizzle width as A in mode 0, SW128 with K=32.. see below.
__global__ void __launch_bounds__(128, 1)
shift_probe_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                   float* __restrict__ out, int mode, int row_bytes, int rows_total, int shift, int base_off) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                       // up to 64 KB
  uint8_t* sB = smem + 65536;               // 8 KB
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 65536 + 8192);
  uint64_t* done = bar + 1;
  uint32_t* holder = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    mbar_init(done, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc<64>(holder);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *holder;
  if (threadIdx.x == 0) {
    if (mode == 0) {
      const uint32_t bytes = (uint32_t)rows_total * row_bytes + 64u * row_bytes;
      mbar_arrive_expect_tx(bar, bytes);
      tma_load_2d(sA, &tma_a, bar, 0, 0);
      tma_load_2d(sB, &tma_b, bar, 0, 0);
    } else {
      const uint32_t bytes = 2u * rows_total * 128u + 64u * 64u;
      mbar_arrive_expect_tx(bar, bytes);
      tma_load_2d(sA, &tma_a, bar, 0, 0);
      tma_load_2d(sA + rows_total * 128, &tma_a, bar, 64, 0);
      tma_load_2d(sB, &tma_b, bar, 0, 0);
    }
    mbar_wait(bar, 0);
    tc_fence_after();
    if (mode == 0) {
      const uint32_t layout = row_bytes == 128 ? kSwizzle128B : kSwizzle64B;
      const uint32_t sbo = 8u * row_bytes;
      const uint32_t idesc = make_idesc(kFmtBF16, 128, 64, false, false);
      const int ksteps = row_bytes / 32;
      for (int k = 0; k < ksteps; ++k) {
        const uint64_t da = desc_with_base(smem_u32(sA) + shift * row_bytes + k * 32, 16, sbo, layout, base_off);
        const uint64_t db = make_smem_desc(smem_u32(sB) + k * 32, 16, sbo, layout);
        mma_f16(tmem, da, db, idesc, k != 0);
      }
    } else {
      const uint32_t idesc = make_idesc(kFmtBF16, 128, 64, true, false);
      for (int k = 0; k < 2; ++k) {
        const uint64_t da = desc_with_base(smem_u32(sA) + (shift + 16 * k) * 128, rows_total * 128, 1024, kSwizzle128B,
                                           base_off);
        const uint64_t db = make_smem_desc(smem_u32(sB) + k * 32, 16, 512, kSwizzle64B);
        mma_f16(tmem, da, db, idesc, k != 0);
      }
    }
    mma_commit(done);
  }
  __syncwarp();
  mbar_wait(done, 0);
  tc_fence_after();
  uint32_t r[32];
  const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
  for (int c = 0; c < 2; ++c) {
    tmem_ld_32x32(taddr + c * 32, r);
    tmem_ld_wait();
    for (int j = 0; j < 32; ++j) out[(warp * 32 + lane) * 64 + c * 32 + j] = __uint_as_float(r[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc<64>(tmem); }
}

// Hand-written (no TMA) 32-byte-row tiles, the layout the first convolution layer builds in
// shared memory itself:
//   mode 2: A K-major SW32 [rows][16 bf16], tile = rows [shift, shift+128); B K-major SW32 [32][16]
//   mode 3: A MN-major SW32: K = 16 rows [shift, shift+16), M = 8 atoms of 16, atom a = the same
//           tile moved down by a*atom_rows rows (LBO = atom_rows*32 bytes); B MN-major SW64 [16][32]
__global__ void __launch_bounds__(128, 1)
shift_probe32_kernel(const __nv_bfloat16* __restrict__ A, const __nv_bfloat16* __restrict__ B, float* __restrict__ out,
                     int mode, int rows_total, int shift, int atom_rows) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                       // rows_total * 32 B
  uint8_t* sB = smem + 16384;
  uint64_t* done = reinterpret_cast<uint64_t*>(smem + 16384 + 2048);
  uint32_t* holder = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < rows_total * 2; i += blockDim.x) {
    const int r = i >> 1, c = i & 1;
    *reinterpret_cast<uint4*>(sA + r * 32 + ((c ^ ((r >> 2) & 1)) << 4)) =
        *reinterpret_cast<const uint4*>(A + r * 16 + c * 8);
  }
  if (mode == 2) {
    for (int i = threadIdx.x; i < 32 * 2; i += blockDim.x) {
      const int r = i >> 1, c = i & 1;
      *reinterpret_cast<uint4*>(sB + r * 32 + ((c ^ ((r >> 2) & 1)) << 4)) =
          *reinterpret_cast<const uint4*>(B + r * 16 + c * 8);
    }
  } else {
    for (int i = threadIdx.x; i < 16 * 4; i += blockDim.x) {
      const int r = i >> 2, c = i & 3;
      *reinterpret_cast<uint4*>(sB + r * 64 + ((c ^ ((r >> 1) & 3)) << 4)) =
          *reinterpret_cast<const uint4*>(B + r * 32 + c * 8);
    }
  }
  fence_proxy_async_smem();
  if (threadIdx.x == 0) { mbar_init(done, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc<32>(holder);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *holder;
  if (warp == 1 && elect_one()) {
    if (mode == 2) {
      const uint32_t idesc = make_idesc(kFmtBF16, 128, 32, false, false);
      mma_f16_first(tmem, make_smem_desc(smem_u32(sA) + shift * 32, 16, 256, kSwizzle32B),
                    make_smem_desc(smem_u32(sB), 16, 256, kSwizzle32B), idesc);
    } else {
      const uint32_t idesc = make_idesc(kFmtBF16, 128, 32, true, true);
      mma_f16_first(tmem, make_smem_desc(smem_u32(sA) + shift * 32, atom_rows * 32, 256, kSwizzle32B),
                    make_smem_desc(smem_u32(sB), 64, 512, kSwizzle64B), idesc);
    }
    mma_commit(done);
  }
  __syncwarp();
  mbar_wait(done, 0);
  tc_fence_after();
  uint32_t r[32];
  tmem_ld_32x32(tmem + ((uint32_t)(warp * 32) << 16), r);
  tmem_ld_wait();
  for (int j = 0; j < 32; ++j) out[(warp * 32 + lane) * 32 + j] = __uint_as_float(r[j]);
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc<32>(tmem); }
}

}  // namespace

extern "C" int mtb_debug_shift_mma32(const void* A, const void* B, float* out, int mode, int rows_total, int shift,
                                     int atom_rows, cudaStream_t stream) {
  MTB_REQUIRE(mode == 2 || mode == 3, "shift probe32: mode 2 or 3");
  MTB_REQUIRE(rows_total <= 512, "shift probe32: at most 512 rows");
  shift_probe32_kernel<<<1, 128, 1024 + 16384 + 2048 + 64, stream>>>(
      reinterpret_cast<const __nv_bfloat16*>(A), reinterpret_cast<const __nv_bfloat16*>(B), out, mode, rows_total, shift,
      atom_rows);
  MTB_LAUNCH_CHECK();
  return 0;
}

extern "C" int mtb_debug_shift_mma(const void* A, const void* B, float* out, int mode, int row_bytes, int rows_total,
                                   int shift, int base_off, cudaStream_t stream) {
  MTB_REQUIRE(mode == 0 || mode == 1, "shift probe: mode 0 or 1");
  MTB_REQUIRE(rows_total <= 256 && rows_total >= 32, "shift probe: 32..256 rows");
  CUtensorMap ma, mb;
  bool ok;
  if (mode == 0) {
    MTB_REQUIRE(row_bytes == 64 || row_bytes == 128, "shift probe: 64 or 128 byte rows");
    MTB_REQUIRE(shift + 128 <= rows_total, "shift probe: shift out of range");
    const uint64_t kel = row_bytes / 2;
    uint64_t dims[2] = {kel, (uint64_t)rows_total};
    uint64_t str[1] = {(uint64_t)row_bytes};
    uint32_t box[2] = {(uint32_t)kel, (uint32_t)rows_total};
    const CUtensorMapSwizzle sw = row_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
    ok = tma::make(&ma, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, A, dims, str, box, sw);
    uint64_t dimsb[2] = {kel, 64};
    uint32_t boxb[2] = {(uint32_t)kel, 64};
    ok = ok && tma::make(&mb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, B, dimsb, str, boxb, sw);
  } else {
    MTB_REQUIRE(shift + 32 <= rows_total, "shift probe: shift out of range");
    uint64_t dims[2] = {128, (uint64_t)rows_total};
    uint64_t str[1] = {256};
    uint32_t box[2] = {64, (uint32_t)rows_total};
    ok = tma::make(&ma, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, A, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
    uint64_t dimsb[2] = {32, 64};
    uint64_t strb[1] = {64};
    uint32_t boxb[2] = {32, 64};
    ok = ok && tma::make(&mb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, B, dimsb, strb, boxb, CU_TENSOR_MAP_SWIZZLE_64B);
  }
  MTB_REQUIRE(ok, "shift probe: tensor map encode failed");
  static bool attr = false;
  const int smem = 1024 + 65536 + 8192 + 64;
  if (!attr) {
    MTB_CUDA_CHECK(cudaFuncSetAttribute(shift_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr = true;
  }
  shift_probe_kernel<<<1, 128, smem, stream>>>(ma, mb, out, mode, row_bytes, rows_total, shift, base_off);
  MTB_LAUNCH_CHECK();
  return 0;
}
