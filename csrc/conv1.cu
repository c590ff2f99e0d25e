// First convolution layer, Cin = 1: conv5x5 SAME + bias + ReLU + 2x2 max-pool in one kernel
// (mpipy.py:156-158), and its weight/bias gradient fused with the pool/ReLU backward.
//
// With one input channel the implicit GEMM has K = 25: the layer is bound by moving the
// [B,28,28,C1] activation, which this kernel never writes — only the pooled bf16 map and a
// 2-bit argmax leave the SM.  Lanes map to output channels, so the 6x6 input patch of a
// pooled pixel is a shared-memory broadcast and each lane keeps its 25 filter taps in
// registers.  The backward kernel scatters nothing: dW[tap,c] += x[patch] * g routed by the
// stored argmax, accumulated in registers over a persistent block's images.
#include "common.cuh"

namespace {

using namespace mtb;

constexpr int IMG = 28, PAD = 2, PW = 32;      // padded row pitch (28 + 2*2), 32x32 floats
constexpr int POOLED = 14;

__device__ __forceinline__ void load_padded_image(float* s_img, const float* __restrict__ x) {
  for (int i = threadIdx.x; i < PW * PW; i += blockDim.x) {
    int r = i / PW - PAD, c = i % PW - PAD;
    s_img[i] = (r >= 0 && r < IMG && c >= 0 && c < IMG) ? __ldg(x + r * IMG + c) : 0.f;
  }
}

// out [B,14,14,C1] bf16 or fp32, idx [B,14,14,C1] u8 (2*dy+dx, first max wins), x [B,28,28] fp32
template <class T>
__global__ void __launch_bounds__(256)
conv1_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w /*[25,C1]*/, const float* __restrict__ bias,
                 T* __restrict__ out, uint8_t* __restrict__ idx, int B, int C1, int parts) {
  __shared__ float s_img[PW * PW];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int groups = C1 / 32;
  // small batches: `parts` blocks share an image (each takes a slice of the pooled pixels) so the grid covers the chip
  const int part = blockIdx.x % parts, per = (POOLED * POOLED + parts - 1) / parts;
  const int item0 = part * per, item1 = min(POOLED * POOLED, item0 + per);
  for (int b = blockIdx.x / parts; b < B; b += gridDim.x / parts) {
    __syncthreads();
    load_padded_image(s_img, x + (size_t)b * IMG * IMG);
    __syncthreads();
    for (int g = 0; g < groups; ++g) {
      const int c = g * 32 + lane;
      float wt[25];
#pragma unroll
      for (int t = 0; t < 25; ++t) wt[t] = __ldg(w + t * C1 + c);
      const float bs = __ldg(bias + c);
      for (int item = item0 + warp; item < item1; item += nwarps) {
        const int pi = item / POOLED, pj = item % POOLED;
        const float* p = s_img + (2 * pi) * PW + 2 * pj;     // top-left of the 6x6 patch (padded coords)
        float a00 = bs, a01 = bs, a10 = bs, a11 = bs;
#pragma unroll
        for (int r = 0; r < 6; ++r) {
          float v[6];
#pragma unroll
          for (int q = 0; q < 6; ++q) v[q] = p[r * PW + q];
#pragma unroll
          for (int kw = 0; kw < 5; ++kw) {
            if (r < 5) { a00 = fmaf(v[kw], wt[r * 5 + kw], a00); a01 = fmaf(v[kw + 1], wt[r * 5 + kw], a01); }
            if (r >= 1) { a10 = fmaf(v[kw], wt[(r - 1) * 5 + kw], a10); a11 = fmaf(v[kw + 1], wt[(r - 1) * 5 + kw], a11); }
          }
        }
        a00 = fmaxf(a00, 0.f); a01 = fmaxf(a01, 0.f); a10 = fmaxf(a10, 0.f); a11 = fmaxf(a11, 0.f);
        float m = a00; int k = 0;
        if (a01 > m) { m = a01; k = 1; }
        if (a10 > m) { m = a10; k = 2; }
        if (a11 > m) { m = a11; k = 3; }
        const size_t o = (((size_t)b * POOLED + pi) * POOLED + pj) * C1 + c;
        stf(out + o, m);
        idx[o] = (uint8_t)k;
      }
    }
  }
}

// dW[25,C1] += sum x[patch at argmax] * g,  db[C1] += sum g, with g = dP * [P > 0]
// dP, P: [B,14,14,C1] bf16 or fp32; idx u8.  Persistent blocks; one atomic per (tap, c) per block.
template <class T>
__global__ void __launch_bounds__(256)
conv1_wgrad_kernel(const float* __restrict__ x, const T* __restrict__ dP, const T* __restrict__ P, const uint8_t* __restrict__ idx, float* __restrict__ dW,
                   float* __restrict__ db, int B, int C1, int parts) {
  __shared__ float s_img[PW * PW];
  __shared__ float s_red[8][26][33];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int groups = C1 / 32;
  const int part = blockIdx.x % parts, per = (POOLED * POOLED + parts - 1) / parts;
  const int item0 = part * per, item1 = min(POOLED * POOLED, item0 + per);
  for (int g = 0; g < groups; ++g) {
    const int c = g * 32 + lane;
    float acc[25];
#pragma unroll
    for (int t = 0; t < 25; ++t) acc[t] = 0.f;
    float accb = 0.f;
    for (int b = blockIdx.x / parts; b < B; b += gridDim.x / parts) {
      __syncthreads();
      load_padded_image(s_img, x + (size_t)b * IMG * IMG);
      __syncthreads();
      for (int item = item0 + warp; item < item1; item += nwarps) {
        const int pi = item / POOLED, pj = item % POOLED;
        const size_t o = (((size_t)b * POOLED + pi) * POOLED + pj) * C1 + c;
        float gv = ldf(dP + o);
        if (!(ldf(P + o) > 0.f)) gv = 0.f;
        const int k = idx[o];
        accb += gv;
        const float* p = s_img + (2 * pi) * PW + 2 * pj;
        const float g00 = k == 0 ? gv : 0.f, g01 = k == 1 ? gv : 0.f, g10 = k == 2 ? gv : 0.f, g11 = k == 3 ? gv : 0.f;
#pragma unroll
        for (int r = 0; r < 6; ++r) {
          float v[6];
#pragma unroll
          for (int q = 0; q < 6; ++q) v[q] = p[r * PW + q];
#pragma unroll
          for (int kw = 0; kw < 5; ++kw) {
            if (r < 5) acc[r * 5 + kw] = fmaf(v[kw], g00, fmaf(v[kw + 1], g01, acc[r * 5 + kw]));
            if (r >= 1) acc[(r - 1) * 5 + kw] = fmaf(v[kw], g10, fmaf(v[kw + 1], g11, acc[(r - 1) * 5 + kw]));
          }
        }
      }
    }
    __syncthreads();
#pragma unroll
    for (int t = 0; t < 25; ++t) s_red[warp][t][lane] = acc[t];
    s_red[warp][25][lane] = accb;
    __syncthreads();
    for (int i = threadIdx.x; i < 26 * 32; i += blockDim.x) {
      const int t = i / 32, l = i % 32;
      float s = 0.f;
      for (int wv = 0; wv < nwarps; ++wv) s += s_red[wv][t][l];
      if (t < 25) atomicAdd(dW + t * C1 + g * 32 + l, s);
      else atomicAdd(db + g * 32 + l, s);
    }
  }
}

// grid = parts * (images in flight): up to `cap` blocks; when the batch alone cannot fill it, 2 or 4 blocks per image
inline void conv1_grid(int B, int cap, int* parts, int* grid) {
  int p = 1;
  while (p < 4 && B * p * 2 <= cap) p *= 2;
  int imgs = B * p <= cap ? B : cap / p;
  *parts = p;
  *grid = imgs * p;
}

}  // namespace

extern "C" {

int mtb_conv1_fwd(const float* x, const float* w, const float* bias, void* out_bf16, uint8_t* idx, int B, int C1,
                  cudaStream_t stream) {
  MTB_REQUIRE(C1 % 32 == 0 && B > 0, "conv1_fwd: C1 must be a multiple of 32");
  int parts, grid;
  conv1_grid(B, 148 * 4, &parts, &grid);
  conv1_fwd_kernel<<<grid, 256, 0, stream>>>(x, w, bias, reinterpret_cast<__nv_bfloat16*>(out_bf16), idx, B, C1, parts);
  MTB_LAUNCH_CHECK();
  return 0;
}

int mtb_conv1_fwd_f32(const float* x, const float* w, const float* bias, float* out, uint8_t* idx, int B, int C1,
                      cudaStream_t stream) {
  MTB_REQUIRE(C1 % 32 == 0 && B > 0, "conv1_fwd: C1 must be a multiple of 32");
  int parts, grid;
  conv1_grid(B, 148 * 4, &parts, &grid);
  conv1_fwd_kernel<<<grid, 256, 0, stream>>>(x, w, bias, out, idx, B, C1, parts);
  MTB_LAUNCH_CHECK();
  return 0;
}

// dW and db are ACCUMULATED into (caller zeroes them once per step).
int mtb_conv1_wgrad(const float* x, const void* dP_bf16, const void* P_bf16, const uint8_t* idx, float* dW, float* db,
                    int B, int C1, cudaStream_t stream) {
  MTB_REQUIRE(C1 % 32 == 0 && B > 0, "conv1_wgrad: C1 must be a multiple of 32");
  int parts, grid;
  conv1_grid(B, 148 * 2, &parts, &grid);
  conv1_wgrad_kernel<<<grid, 256, 0, stream>>>(x, reinterpret_cast<const __nv_bfloat16*>(dP_bf16),
                                               reinterpret_cast<const __nv_bfloat16*>(P_bf16), idx, dW, db, B, C1,
                                               parts);
  MTB_LAUNCH_CHECK();
  return 0;
}

int mtb_conv1_wgrad_f32(const float* x, const float* dP, const float* P, const uint8_t* idx, float* dW, float* db,
                        int B, int C1, cudaStream_t stream) {
  MTB_REQUIRE(C1 % 32 == 0 && B > 0, "conv1_wgrad: C1 must be a multiple of 32");
  int parts, grid;
  conv1_grid(B, 148 * 2, &parts, &grid);
  conv1_wgrad_kernel<<<grid, 256, 0, stream>>>(x, dP, P, idx, dW, db, B, C1, parts);
  MTB_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
