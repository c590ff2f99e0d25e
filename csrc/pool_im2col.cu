// Max-pool forward/backward (mpipy.py:158,161 and the MaxPoolGrad/ReluGrad ops autodiff adds,
// SURVEY O3/O6/O18/O19) and the im2col / col2im pair of the generic convolution path.
// All tensors NHWC, bf16 or fp32 (the `_f32` entry points), 8 channels per thread.
#include "common.cuh"

namespace {

using namespace mtb;

struct alignas(16) BF8 { __nv_bfloat162 v[4]; };

__device__ __forceinline__ void unpack8(const BF8& a, float (&f)[8]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) { float2 t = __bfloat1622float2(a.v[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
}
struct alignas(32) F8 { float4 a, b; };
__device__ __forceinline__ void unpack8(const F8& v, float (&f)[8]) {
  f[0] = v.a.x; f[1] = v.a.y; f[2] = v.a.z; f[3] = v.a.w; f[4] = v.b.x; f[5] = v.b.y; f[6] = v.b.z; f[7] = v.b.w;
}
template <class V> __device__ __forceinline__ V pack8(const float (&f)[8]);
template <> __device__ __forceinline__ BF8 pack8<BF8>(const float (&f)[8]) {
  BF8 r;
#pragma unroll
  for (int i = 0; i < 4; ++i) r.v[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return r;
}
template <> __device__ __forceinline__ F8 pack8<F8>(const float (&f)[8]) {
  F8 r;
  r.a = make_float4(f[0], f[1], f[2], f[3]);
  r.b = make_float4(f[4], f[5], f[6], f[7]);
  return r;
}

// in [B,H,W,C] -> out [B,H/2,W/2,C], idx u8 = 2*dy+dx (first max wins)
template <class V>
__global__ void maxpool_fwd_kernel(const V* __restrict__ in, V* __restrict__ out, uint2* __restrict__ idx,
                                   long long total, int H, int W, int C8) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int c8 = (int)(i % C8);
  long long t = i / C8;
  const int pw = (int)(t % (W / 2)); t /= (W / 2);
  const int ph = (int)(t % (H / 2));
  const long long b = t / (H / 2);
  float m[8];
  uint8_t k[8];
#pragma unroll
  for (int pos = 0; pos < 4; ++pos) {
    const int h = 2 * ph + (pos >> 1), w = 2 * pw + (pos & 1);
    float f[8];
    unpack8(in[((b * H + h) * W + w) * C8 + c8], f);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (pos == 0 || f[j] > m[j]) { m[j] = f[j]; k[j] = (uint8_t)pos; }
    }
  }
  out[i] = pack8<V>(m);
  uint2 kk;
  kk.x = k[0] | (k[1] << 8) | (k[2] << 16) | (k[3] << 24);
  kk.y = k[4] | (k[5] << 8) | (k[6] << 16) | (k[7] << 24);
  idx[i] = kk;
}

// dA[b,2ph+dy,2pw+dx,c] = dP * [P>0] at the argmax position, 0 elsewhere  (pool grad + ReLU grad)
template <class V>
__global__ void unpool_relu_bwd_kernel(const V* __restrict__ dP, const V* __restrict__ P,
                                       const uint2* __restrict__ idx, V* __restrict__ dA, long long total, int H,
                                       int W, int C8) {
  pdl_sync();
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int c8 = (int)(i % C8);
  long long t = i / C8;
  const int pw = (int)(t % (W / 2)); t /= (W / 2);
  const int ph = (int)(t % (H / 2));
  const long long b = t / (H / 2);
  float g[8], p[8];
  unpack8(dP[i], g);
  unpack8(P[i], p);
  const uint2 kk = idx[i];
  uint8_t k[8];
#pragma unroll
  for (int j = 0; j < 4; ++j) { k[j] = (kk.x >> (8 * j)) & 0xff; k[4 + j] = (kk.y >> (8 * j)) & 0xff; }
#pragma unroll
  for (int j = 0; j < 8; ++j) if (!(p[j] > 0.f)) g[j] = 0.f;
#pragma unroll
  for (int pos = 0; pos < 4; ++pos) {
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = (k[j] == pos) ? g[j] : 0.f;
    const int h = 2 * ph + (pos >> 1), w = 2 * pw + (pos & 1);
    dA[((b * H + h) * W + w) * C8 + c8] = pack8<V>(o);
  }
}

// col[(b,h,w), (kh,kw,c)] = x[b,h+kh-2,w+kw-2,c]  (zero outside)
template <class V>
__global__ void im2col5_kernel(const V* __restrict__ x, V* __restrict__ col, long long total, int H, int W,
                               int C8) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int c8 = (int)(i % C8);
  long long t = i / C8;
  const int tap = (int)(t % 25); t /= 25;
  const int w = (int)(t % W); t /= W;
  const int h = (int)(t % H);
  const long long b = t / H;
  const int hh = h + tap / 5 - 2, ww = w + tap % 5 - 2;
  V v;
  if (hh >= 0 && hh < H && ww >= 0 && ww < W) {
    v = x[((b * H + hh) * W + ww) * C8 + c8];
  } else {
    const float zero[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    v = pack8<V>(zero);
  }
  col[i] = v;
}

// dx[b,h,w,c] = sum_taps dcol[(b,h-kh+2,w-kw+2), (kh,kw,c)]   (gather form, no atomics)
template <class V>
__global__ void col2im5_kernel(const V* __restrict__ dcol, V* __restrict__ dx, long long total, int H, int W,
                               int C8) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int c8 = (int)(i % C8);
  long long t = i / C8;
  const int w = (int)(t % W); t /= W;
  const int h = (int)(t % H);
  const long long b = t / H;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  for (int tap = 0; tap < 25; ++tap) {
    const int hh = h - (tap / 5 - 2), ww = w - (tap % 5 - 2);
    if (hh < 0 || hh >= H || ww < 0 || ww >= W) continue;
    float f[8];
    unpack8(dcol[(((b * H + hh) * W + ww) * 25 + tap) * C8 + c8], f);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] += f[j];
  }
  dx[i] = pack8<V>(acc);
}

inline unsigned nblk(long long n) { return (unsigned)((n + 255) / 256); }

template <class V>
int run_maxpool(const void* in, void* out, uint8_t* idx, int B, int H, int W, int C, cudaStream_t stream) {
  MTB_REQUIRE(C % 8 == 0 && H % 2 == 0 && W % 2 == 0, "maxpool: C%8, even H/W required");
  long long total = (long long)B * (H / 2) * (W / 2) * (C / 8);
  maxpool_fwd_kernel<V><<<nblk(total), 256, 0, stream>>>(reinterpret_cast<const V*>(in), reinterpret_cast<V*>(out),
                                                         reinterpret_cast<uint2*>(idx), total, H, W, C / 8);
  MTB_LAUNCH_CHECK();
  return 0;
}

template <class V>
int run_unpool(const void* dP, const void* P, const uint8_t* idx, void* dA, int B, int H, int W, int C,
               cudaStream_t stream) {
  MTB_REQUIRE(C % 8 == 0 && H % 2 == 0 && W % 2 == 0, "unpool: C%8, even H/W required");
  long long total = (long long)B * (H / 2) * (W / 2) * (C / 8);
  MTB_CUDA_CHECK(mtb::launch(unpool_relu_bwd_kernel<V>, dim3(nblk(total)), dim3(256), 0, stream,
                             reinterpret_cast<const V*>(dP), reinterpret_cast<const V*>(P),
                             reinterpret_cast<const uint2*>(idx), reinterpret_cast<V*>(dA), total, H, W, C / 8));
  return 0;
}

template <class V>
int run_im2col(const void* x, void* col, int B, int H, int W, int C, cudaStream_t stream) {
  MTB_REQUIRE(C % 8 == 0, "im2col: C%8 required");
  long long total = (long long)B * H * W * 25 * (C / 8);
  im2col5_kernel<V><<<nblk(total), 256, 0, stream>>>(reinterpret_cast<const V*>(x), reinterpret_cast<V*>(col), total, H,
                                                     W, C / 8);
  MTB_LAUNCH_CHECK();
  return 0;
}

template <class V>
int run_col2im(const void* dcol, void* dx, int B, int H, int W, int C, cudaStream_t stream) {
  MTB_REQUIRE(C % 8 == 0, "col2im: C%8 required");
  long long total = (long long)B * H * W * (C / 8);
  col2im5_kernel<V><<<nblk(total), 256, 0, stream>>>(reinterpret_cast<const V*>(dcol), reinterpret_cast<V*>(dx), total,
                                                     H, W, C / 8);
  MTB_LAUNCH_CHECK();
  return 0;
}

}  // namespace

extern "C" {

int mtb_maxpool2x2_fwd(const void* in, void* out, uint8_t* idx, int B, int H, int W, int C, cudaStream_t stream) {
  return run_maxpool<BF8>(in, out, idx, B, H, W, C, stream);
}
int mtb_maxpool2x2_fwd_f32(const float* in, float* out, uint8_t* idx, int B, int H, int W, int C, cudaStream_t stream) {
  return run_maxpool<F8>(in, out, idx, B, H, W, C, stream);
}
int mtb_unpool_relu_bwd(const void* dP, const void* P, const uint8_t* idx, void* dA, int B, int H, int W, int C,
                        cudaStream_t stream) {
  return run_unpool<BF8>(dP, P, idx, dA, B, H, W, C, stream);
}
int mtb_unpool_relu_bwd_f32(const float* dP, const float* P, const uint8_t* idx, float* dA, int B, int H, int W, int C,
                            cudaStream_t stream) {
  return run_unpool<F8>(dP, P, idx, dA, B, H, W, C, stream);
}
int mtb_im2col_5x5(const void* x, void* col, int B, int H, int W, int C, cudaStream_t stream) {
  return run_im2col<BF8>(x, col, B, H, W, C, stream);
}
int mtb_im2col_5x5_f32(const float* x, float* col, int B, int H, int W, int C, cudaStream_t stream) {
  return run_im2col<F8>(x, col, B, H, W, C, stream);
}
int mtb_col2im_5x5(const void* dcol, void* dx, int B, int H, int W, int C, cudaStream_t stream) {
  return run_col2im<BF8>(dcol, dx, B, H, W, C, stream);
}
int mtb_col2im_5x5_f32(const float* dcol, float* dx, int B, int H, int W, int C, cudaStream_t stream) {
  return run_col2im<F8>(dcol, dx, B, H, W, C, stream);
}

}  // extern "C"
