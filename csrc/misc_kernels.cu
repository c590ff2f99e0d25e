// Small elementwise / reduction kernels: parameter init, synthetic data, casts, L2 value.
#include "common.cuh"

namespace {

using namespace mtb;

// K10: tf.truncated_normal(stddev, seed) (mpipy.py:38-52).  Same algorithm as
// utils/philox.py::truncated_normal so host and device initialise identically.
__global__ void trunc_normal_kernel(float* __restrict__ dst, long long n, float stddev, unsigned long long seed,
                                    uint32_t stream_id) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n) return;
  float z = 0.f;
  for (uint32_t t = 0; t < 16; ++t) {
    uint4 w = Philox::gen((uint32_t)i, stream_id, t, kPurposeInit, (uint32_t)seed, (uint32_t)(seed >> 32));
    float u1 = ((float)(w.x >> 8) + 1.0f) * (1.0f / 16777216.0f);
    float u2 = (float)(w.y >> 8) * (1.0f / 16777216.0f);
    float r = sqrtf(-2.0f * logf(u1));
    z = r * cosf(6.283185307179586f * u2);
    if (fabsf(z) <= 2.0f) break;
    z = 0.f;
  }
  dst[i] = z * stddev;
}

__global__ void fill_kernel(float* __restrict__ dst, long long n, float v) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i < n) dst[i] = v;
}

// K12: (x - 127.5) / 255  (convolutional.extract_data, used at mpipy.py:215)
__global__ void normalize_u8_kernel(const uint8_t* __restrict__ src, float* __restrict__ dst, long long n) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i < n) dst[i] = ((float)src[i] - 127.5f) / 255.0f;
}

// K12: deterministic fake MNIST, same streams as utils/data.py::synthetic_mnist.
__global__ void synthetic_kernel(float* __restrict__ images, int* __restrict__ labels, long long n_img,
                                 long long row_offset, unsigned long long seed, int learnable) {
  const int npx = 28 * 28;
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n_img * npx) return;
  long long img = i / npx;
  int px = (int)(i % npx);
  unsigned long long gl = (unsigned long long)(row_offset + img);
  uint4 wl = Philox::gen((uint32_t)(gl >> 2), 0u, 0u, kPurposeData, (uint32_t)seed, (uint32_t)(seed >> 32));
  uint32_t lw = (gl & 3) == 0 ? wl.x : (gl & 3) == 1 ? wl.y : (gl & 3) == 2 ? wl.z : wl.w;
  int lab = (int)(lw % 10u);
  unsigned long long ge = gl * (unsigned long long)npx + (unsigned)px;
  uint4 wn = Philox::gen((uint32_t)(ge >> 2), 1u, 0u, kPurposeData, (uint32_t)seed, (uint32_t)(seed >> 32));
  uint32_t nw = (ge & 3) == 0 ? wn.x : (ge & 3) == 1 ? wn.y : (ge & 3) == 2 ? wn.z : wn.w;
  float v = (float)(nw >> 24);
  if (learnable) {
    float cy = 6.0f + 16.0f * (float)((lab * 7) % 10) / 10.0f;
    float cx = 6.0f + 16.0f * (float)((lab * 3) % 10) / 10.0f;
    float y = (float)(px / 28), x = (float)(px % 28);
    float proto = 255.0f * expf(-((y - cy) * (y - cy) + (x - cx) * (x - cx)) / 18.0f);
    v = 0.35f * v + 0.65f * proto;
  }
  v = fminf(fmaxf(v, 0.f), 255.f);
  float u8 = floorf(v);                      // numpy .astype(uint8) truncates
  images[i] = (u8 - 127.5f) / 255.0f;
  if (px == 0) labels[img] = lab;
}

__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long n) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i < n) dst[i] = __float2bfloat16(src[i]);
}

__global__ void cast_bf16_f32_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst, long long n) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i < n) dst[i] = __bfloat162float(src[i]);
}

// K11: sum(x^2)/2 accumulated into *out (tf.nn.l2_loss, mpipy.py:57-58).  Logging only: the
// gradient of the L2 term lives inside the fused update kernel.
__global__ void l2_half_sumsq_kernel(const float* __restrict__ x, long long n, float* __restrict__ out) {
  float acc = 0.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float v = x[i];
    acc += v * v;
  }
  acc = warp_sum(acc);
  __shared__ float s[32];
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < (blockDim.x >> 5) ? s[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) atomicAdd(out, 0.5f * v);
  }
}

// K12: the reference's sequential batch window (mpipy.py:80-82), offset = (step*B) % (rows-B),
// evaluated on the device from the device-resident step counter so a captured CUDA graph
// needs no per-step host input.
__global__ void batch_window_kernel(const float4* __restrict__ data, const int* __restrict__ labels,
                                    const long long* __restrict__ step_ptr, long long rows, int B,
                                    float4* __restrict__ x_out, int* __restrict__ y_out) {
  pdl_sync();
  const long long step = *step_ptr;
  const long long off = (step * (long long)B) % (rows - B);
  const long long n4 = (long long)B * 196;            // 784 floats = 196 float4 per image
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x)
    x_out[i] = data[off * 196 + i];
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < B; i += (long long)gridDim.x * blockDim.x)
    y_out[i] = labels[off + i];
}

// One launch that clears every per-step accumulator (gradient ranges whose kernels accumulate,
// the loss and error counters): replaces a handful of separate memsets in the captured step.
struct ZeroList { void* ptr[8]; long long n_words[8]; int count; };
__global__ void zero_multi_kernel(const ZeroList z) {
  pdl_sync();
  for (int k = 0; k < z.count; ++k) {
    uint32_t* p = reinterpret_cast<uint32_t*>(z.ptr[k]);
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < z.n_words[k];
         i += (long long)gridDim.x * blockDim.x)
      p[i] = 0u;
  }
}

inline unsigned blocks_for(long long n, int threads) { return (unsigned)((n + threads - 1) / threads); }

}  // namespace

// Occupancy probe (scripts/, not the product): blocks that sit on the SMs for a given time without touching memory,
// lean (a handful of registers) or fat (about 56 live registers per thread), to separate placement / register
// co-residency effects from memory-system interference when a kernel runs under the persistent convolution kernels.
template <bool FAT>
__global__ void __launch_bounds__(256) occupy_kernel(long long cycles, float* sink) {
  const long long t0 = clock64();
  float acc[FAT ? 48 : 1];
#pragma unroll
  for (int i = 0; i < (FAT ? 48 : 1); ++i) acc[i] = (float)(threadIdx.x + i);
  while (clock64() - t0 < cycles) {
#pragma unroll
    for (int i = 0; i < (FAT ? 48 : 1); ++i) acc[i] = fmaf(acc[i], 1.0001f, 0.5f);
  }
  float t = 0.f;
#pragma unroll
  for (int i = 0; i < (FAT ? 48 : 1); ++i) t += acc[i];
  if (t == 12345.678f) *sink = t;
}

extern "C" {

int mtb_debug_occupy(int grid, int block, int fat, long long cycles, float* sink, cudaStream_t stream) {
  if (fat) occupy_kernel<true><<<grid, block, 0, stream>>>(cycles, sink);
  else occupy_kernel<false><<<grid, block, 0, stream>>>(cycles, sink);
  MTB_LAUNCH_CHECK();
  return 0;
}

int mtb_init_trunc_normal(float* dst, long long n, float stddev, unsigned long long seed, unsigned stream_id,
                          cudaStream_t stream) {
  if (n <= 0) return 0;
  trunc_normal_kernel<<<blocks_for(n, 256), 256, 0, stream>>>(dst, n, stddev, seed, stream_id);
  MTB_LAUNCH_CHECK();
  return 0;
}

int mtb_zero_multi(void* const* ptrs, const long long* n_words, int count, cudaStream_t stream) {
  MTB_REQUIRE(count >= 0 && count <= 8, "zero_multi: at most 8 ranges");
  if (count == 0) return 0;
  ZeroList z;
  long long total = 0;
  for (int k = 0; k < 8; ++k) { z.ptr[k] = k < count ? ptrs[k] : nullptr; z.n_words[k] = k < count ? n_words[k] : 0; total += z.n_words[k]; }
  z.count = count;
  unsigned blocks = blocks_for(total, 256 * 4);
  if (blocks > 148) blocks = 148;
  if (blocks < 1) blocks = 1;
  MTB_CUDA_CHECK(mtb::launch(zero_multi_kernel, dim3(blocks), dim3(256), 0, stream, z));
  return 0;
}

int mtb_fill_f32(float* dst, long long n, float value, cudaStream_t stream) {
  if (n <= 0) return 0;
  fill_kernel<<<blocks_for(n, 256), 256, 0, stream>>>(dst, n, value);
  MTB_LAUNCH_CHECK();
  return 0;
}

int mtb_normalize_u8_dev(const uint8_t* src, float* dst, long long n, cudaStream_t stream) {
  if (n <= 0) return 0;
  normalize_u8_kernel<<<blocks_for(n, 256), 256, 0, stream>>>(src, dst, n);
  MTB_LAUNCH_CHECK();
  return 0;
}

int mtb_synthetic_batch(float* images, int* labels, long long n_img, long long row_offset,
                        unsigned long long seed, int learnable, cudaStream_t stream) {
  if (n_img <= 0) return 0;
  synthetic_kernel<<<blocks_for(n_img * 784, 256), 256, 0, stream>>>(images, labels, n_img, row_offset, seed, learnable);
  MTB_LAUNCH_CHECK();
  return 0;
}

int mtb_batch_window(const float* data, const int* labels, const long long* step_ptr, long long rows, int B,
                     float* x_out, int* y_out, cudaStream_t stream) {
  MTB_REQUIRE(rows > B && B > 0, "batch_window: shard must be larger than the batch");
  unsigned blocks = blocks_for((long long)B * 196, 256);
  if (blocks > 1184) blocks = 1184;
  MTB_CUDA_CHECK(mtb::launch(batch_window_kernel, dim3(blocks), dim3(256), 0, stream,
                             reinterpret_cast<const float4*>(data), labels, step_ptr, rows, B,
                             reinterpret_cast<float4*>(x_out), y_out));
  return 0;
}

int mtb_cast_f32_to_bf16(const float* src, void* dst, long long n, cudaStream_t stream) {
  if (n <= 0) return 0;
  cast_f32_bf16_kernel<<<blocks_for(n, 256), 256, 0, stream>>>(src, reinterpret_cast<__nv_bfloat16*>(dst), n);
  MTB_LAUNCH_CHECK();
  return 0;
}

int mtb_cast_bf16_to_f32(const void* src, float* dst, long long n, cudaStream_t stream) {
  if (n <= 0) return 0;
  cast_bf16_f32_kernel<<<blocks_for(n, 256), 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(src), dst, n);
  MTB_LAUNCH_CHECK();
  return 0;
}

int mtb_l2_half_sumsq(const float* x, long long n, float* out_accum, cudaStream_t stream) {
  if (n <= 0) return 0;
  unsigned blocks = blocks_for(n, 256 * 8);
  if (blocks > 592) blocks = 592;
  l2_half_sumsq_kernel<<<blocks, 256, 0, stream>>>(x, n, out_accum);
  MTB_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
