// Shared device/host helpers for libmtb200.so (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../include/mtb200.h"

#define MTB_CUDA_CHECK(expr)                                  \
  do {                                                        \
    cudaError_t _e = (expr);                                  \
    if (_e != cudaSuccess) return mtb_set_error(_e, #expr, __FILE__, __LINE__); \
  } while (0)

#define MTB_LAUNCH_CHECK() MTB_CUDA_CHECK(cudaGetLastError())

#define MTB_REQUIRE(cond, msg)                                \
  do {                                                        \
    if (!(cond)) return mtb_set_error_msg(MTB_E_ARG, msg, __FILE__, __LINE__); \
  } while (0)

int mtb_set_error(cudaError_t e, const char* expr, const char* file, int line);
int mtb_set_error_msg(int code, const char* msg, const char* file, int line);

namespace mtb {

constexpr int kNumSMs = 148;

// ---- programmatic dependent launch -------------------------------------------------------------
// Every kernel of the training step is launched with the programmatic-stream-serialization
// attribute and calls pdl_sync() once, by all threads, after the part of its prologue that
// touches no global data (barrier init, TMEM allocation, shared-memory clears, tensor-map
// prefetch).  The next kernel's CTAs can then be placed, and run that prologue, while this
// kernel's tail is still executing; pdl_sync() returns only when the previous kernel has
// completed and its writes are visible, so everything after it sees the usual stream order.
__device__ __forceinline__ void pdl_sync() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

bool pdl_enabled();      // MTB_PDL=0 turns the launch attribute off (runtime.cu)

template <typename... KArgs, typename... Args>
inline cudaError_t launch(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                          Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

// Same with a thread-block cluster shape (1x1x1 = no cluster attribute).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_cluster(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                  dim3 cluster, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[2];
  int n = 0;
  if (cluster.x * cluster.y * cluster.z > 1) {
    at[n].id = cudaLaunchAttributeClusterDimension;
    at[n].val.clusterDim.x = cluster.x;
    at[n].val.clusterDim.y = cluster.y;
    at[n].val.clusterDim.z = cluster.z;
    ++n;
  }
  if (pdl_enabled()) {
    at[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = at;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

__host__ __device__ inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

// ---- Philox4x32-10: identical to utils/philox.py ------------------------------
struct Philox {
  static constexpr uint32_t kM0 = 0xD2511F53u, kM1 = 0xCD9E8D57u;
  static constexpr uint32_t kW0 = 0x9E3779B9u, kW1 = 0xBB67AE85u;
  __device__ static inline uint4 gen(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                     uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
      uint32_t hi0 = __umulhi(kM0, c0), lo0 = kM0 * c0;
      uint32_t hi1 = __umulhi(kM1, c2), lo1 = kM1 * c2;
      uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
      c0 = n0; c1 = n1; c2 = n2; c3 = n3;
      k0 += kW0; k1 += kW1;
    }
    return make_uint4(c0, c1, c2, c3);
  }
};

enum : uint32_t { kPurposeDropout = 1, kPurposeInit = 2, kPurposeData = 3 };

__device__ inline float u32_to_unit(uint32_t w) { return (float)(w >> 8) * (1.0f / 16777216.0f); }

// Four consecutive keep decisions for elements 4*blk .. 4*blk+3 of dropout stream `step`.
__device__ inline uint4 dropout_words(uint32_t blk, uint64_t seed, uint64_t step) {
  return Philox::gen(blk, (uint32_t)step, (uint32_t)(step >> 32), kPurposeDropout, (uint32_t)seed,
                     (uint32_t)(seed >> 32));
}

// element access for kernels instantiated for both activation types (bf16 and fp32)
__device__ __forceinline__ float ldf(const __nv_bfloat16* p) { return __bfloat162float(*p); }
__device__ __forceinline__ float ldf(const float* p) { return *p; }
__device__ __forceinline__ void stf(__nv_bfloat16* p, float v) { *p = __float2bfloat16(v); }
__device__ __forceinline__ void stf(float* p, float v) { *p = v; }

__device__ inline float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ inline float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ inline uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

// MaxPoolGrad + ReluGrad on 8 channels with byte/halfword SIMD: g, a = 4 bf16x2 words each
// (pooled gradient, pooled activation >= 0), k = 8 argmax bytes.  Returns the gradient words of
// window position `pos`: the gradient where the argmax is `pos` and the activation is non-zero,
// else 0.  A handful of integer ops instead of unpack / compare / select per channel.
__device__ __forceinline__ void relu_mask_bf16x8(const uint4& g, const uint4& a, uint32_t (&gm)[4]) {
  constexpr uint32_t kAbs = 0x7fff7fffu;          // a -0.0 out of a ReLU counts as zero
  gm[0] = g.x & __vcmpne2(a.x & kAbs, 0u); gm[1] = g.y & __vcmpne2(a.y & kAbs, 0u);
  gm[2] = g.z & __vcmpne2(a.z & kAbs, 0u); gm[3] = g.w & __vcmpne2(a.w & kAbs, 0u);
}
__device__ __forceinline__ uint4 route_to_pos_bf16x8(const uint32_t (&gm)[4], const uint2& k, uint32_t pos) {
  const uint32_t p4 = pos * 0x01010101u;
  const uint32_t m0 = __vcmpeq4(k.x, p4), m1 = __vcmpeq4(k.y, p4);        // 0xff per channel whose argmax is pos
  uint4 v;
  v.x = gm[0] & __byte_perm(m0, 0u, 0x1100);                               // bytes (b0,b0,b1,b1): one halfword per channel
  v.y = gm[1] & __byte_perm(m0, 0u, 0x3322);
  v.z = gm[2] & __byte_perm(m1, 0u, 0x1100);
  v.w = gm[3] & __byte_perm(m1, 0u, 0x3322);
  return v;
}

__device__ inline float2 unpack_bf16x2(uint32_t w) {
  __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&w);
  return __bfloat1622float2(v);
}

}  // namespace mtb
