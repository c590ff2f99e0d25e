// Last layer + loss: FC2, softmax, cross-entropy, their gradients (mpipy.py:167, :55-56, :67-68).
//
// N = 10 classes is far too narrow for a tensor-core tile, and the whole layer is a few
// MFLOP, so it runs on CUDA cores — but as ONE forward kernel (logits, softmax probabilities,
// mean CE, dlogits = (softmax - onehot)/B, error count) and two small backward kernels,
// instead of the ~10 TF ops the reference instantiates.
#include "common.cuh"

namespace {

using namespace mtb;
constexpr int kMaxClasses = 16;

// one warp per row
__global__ void __launch_bounds__(256)
fc2_softmax_ce_kernel(const __nv_bfloat16* __restrict__ h, const float* __restrict__ W /*[F,NC]*/,
                      const float* __restrict__ bias, const int* __restrict__ labels, float* __restrict__ logits,
                      float* __restrict__ probs, float* __restrict__ dlogits, float* __restrict__ loss_accum,
                      int* __restrict__ err_accum, int B, int F, int NC) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= B) return;
  float acc[kMaxClasses];
#pragma unroll
  for (int c = 0; c < kMaxClasses; ++c) acc[c] = 0.f;
  const __nv_bfloat16* hr = h + (size_t)warp * F;
  for (int f = lane; f < F; f += 32) {
    const float hv = __bfloat162float(hr[f]);
    const float* wr = W + (size_t)f * NC;
#pragma unroll
    for (int c = 0; c < kMaxClasses; ++c)
      if (c < NC) acc[c] = fmaf(hv, __ldg(wr + c), acc[c]);
  }
  float mx = -INFINITY;
#pragma unroll
  for (int c = 0; c < kMaxClasses; ++c) {
    if (c < NC) {
      acc[c] = warp_sum(acc[c]) + __ldg(bias + c);
      mx = fmaxf(mx, acc[c]);
    }
  }
  float s = 0.f;
  float e[kMaxClasses];
#pragma unroll
  for (int c = 0; c < kMaxClasses; ++c) {
    e[c] = c < NC ? expf(acc[c] - mx) : 0.f;
    s += e[c];
  }
  const float inv = 1.0f / s;
  const int y = labels ? labels[warp] : -1;
  int arg = 0;
  float best = -INFINITY, ly = 0.f;
#pragma unroll
  for (int c = 0; c < kMaxClasses; ++c) {
    if (c < NC) {
      if (acc[c] > best) { best = acc[c]; arg = c; }
      if (c == y) ly = acc[c];
    }
  }
  if (lane == 0) {
    const float invB = 1.0f / (float)B;
#pragma unroll
    for (int c = 0; c < kMaxClasses; ++c) {
      if (c < NC) {
        const float p = e[c] * inv;
        if (logits) logits[(size_t)warp * NC + c] = acc[c];
        if (probs) probs[(size_t)warp * NC + c] = p;
        if (dlogits) dlogits[(size_t)warp * NC + c] = (p - (c == y ? 1.f : 0.f)) * invB;
      }
    }
    if (y >= 0) {
      if (loss_accum) atomicAdd(loss_accum, -(ly - mx - logf(s)) * invB);
      if (err_accum && arg != y) atomicAdd(err_accum, 1);
    }
  }
}

// dh[b,f] = (sum_c dlogits[b,c] * W[f,c]) * (h[b,f] > 0 ? scale : 0)   -> bf16
__global__ void __launch_bounds__(256)
fc2_dgrad_kernel(const float* __restrict__ dlogits, const float* __restrict__ W, const __nv_bfloat16* __restrict__ h,
                 __nv_bfloat16* __restrict__ dh, int B, int F, int NC, float scale) {
  const int b = blockIdx.y;
  __shared__ float dl[kMaxClasses];
  if (threadIdx.x < NC) dl[threadIdx.x] = dlogits[(size_t)b * NC + threadIdx.x];
  __syncthreads();
  for (int f = blockIdx.x * blockDim.x + threadIdx.x; f < F; f += gridDim.x * blockDim.x) {
    float acc = 0.f;
    const float* wr = W + (size_t)f * NC;
    for (int c = 0; c < NC; ++c) acc = fmaf(dl[c], __ldg(wr + c), acc);
    const float hv = __bfloat162float(h[(size_t)b * F + f]);
    dh[(size_t)b * F + f] = __float2bfloat16(hv > 0.f ? acc * scale : 0.f);
  }
}

// dW[f,c] += sum_b h[b,f]*dlogits[b,c];  db[c] += sum_b dlogits[b,c].  lanes <-> f.
__global__ void __launch_bounds__(128)
fc2_wgrad_kernel(const __nv_bfloat16* __restrict__ h, const float* __restrict__ dlogits, float* __restrict__ dW,
                 float* __restrict__ db, int B, int F, int NC, int rows_per_block) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  const int b0 = blockIdx.y * rows_per_block, b1 = min(B, b0 + rows_per_block);
  float acc[kMaxClasses];
#pragma unroll
  for (int c = 0; c < kMaxClasses; ++c) acc[c] = 0.f;
  float accb = 0.f;
  const int cb = threadIdx.x;            // thread cb (< NC) of block x==0 also sums the bias column
  for (int b = b0; b < b1; ++b) {
    const float* dl = dlogits + (size_t)b * NC;
    if (f < F) {
      const float hv = __bfloat162float(h[(size_t)b * F + f]);
#pragma unroll
      for (int c = 0; c < kMaxClasses; ++c)
        if (c < NC) acc[c] = fmaf(hv, __ldg(dl + c), acc[c]);
    }
    if (blockIdx.x == 0 && cb < NC) accb += __ldg(dl + cb);
  }
  if (f < F) {
#pragma unroll
    for (int c = 0; c < kMaxClasses; ++c)
      if (c < NC) atomicAdd(dW + (size_t)f * NC + c, acc[c]);
  }
  if (blockIdx.x == 0 && cb < NC) atomicAdd(db + cb, accb);
}

// out[c] += sum_r x[r,c]   (bias gradients, SURVEY O18/O20 "BiasAddGrad")
__global__ void __launch_bounds__(256)
colsum_bf16_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ out, long long R, int C,
                   int rows_per_block) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const long long r0 = (long long)blockIdx.y * rows_per_block;
  const long long r1 = r0 + rows_per_block < R ? r0 + rows_per_block : R;
  float acc = 0.f;
  for (long long r = r0; r < r1; ++r) acc += __bfloat162float(x[r * C + c]);
  atomicAdd(out + c, acc);
}

}  // namespace

extern "C" {

int mtb_fc2_softmax_ce(const void* h_bf16, const float* W, const float* bias, const int* labels, float* logits,
                       float* probs, float* dlogits, float* loss_accum, int* err_accum, int B, int F, int NC,
                       cudaStream_t stream) {
  MTB_REQUIRE(NC > 0 && NC <= kMaxClasses, "fc2_softmax_ce: at most 16 classes");
  const int warps_per_block = 8;
  fc2_softmax_ce_kernel<<<(B + warps_per_block - 1) / warps_per_block, 256, 0, stream>>>(
      reinterpret_cast<const __nv_bfloat16*>(h_bf16), W, bias, labels, logits, probs, dlogits, loss_accum, err_accum,
      B, F, NC);
  MTB_LAUNCH_CHECK();
  return 0;
}

int mtb_fc2_dgrad(const float* dlogits, const float* W, const void* h_bf16, void* dh_bf16, int B, int F, int NC,
                  float scale, cudaStream_t stream) {
  MTB_REQUIRE(NC > 0 && NC <= kMaxClasses, "fc2_dgrad: at most 16 classes");
  dim3 grid((F + 255) / 256, B);
  fc2_dgrad_kernel<<<grid, 256, 0, stream>>>(dlogits, W, reinterpret_cast<const __nv_bfloat16*>(h_bf16),
                                             reinterpret_cast<__nv_bfloat16*>(dh_bf16), B, F, NC, scale);
  MTB_LAUNCH_CHECK();
  return 0;
}

// dW, db are accumulated into.
int mtb_fc2_wgrad(const void* h_bf16, const float* dlogits, float* dW, float* db, int B, int F, int NC,
                  cudaStream_t stream) {
  MTB_REQUIRE(NC > 0 && NC <= kMaxClasses, "fc2_wgrad: at most 16 classes");
  int rows_per_block = 32;
  dim3 grid((F + 127) / 128, (B + rows_per_block - 1) / rows_per_block);
  fc2_wgrad_kernel<<<grid, 128, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(h_bf16), dlogits, dW, db, B, F,
                                             NC, rows_per_block);
  MTB_LAUNCH_CHECK();
  return 0;
}

int mtb_colsum_bf16(const void* x_bf16, float* out, long long R, int C, cudaStream_t stream) {
  int rows_per_block = 64;
  dim3 grid((C + 255) / 256, (unsigned)((R + rows_per_block - 1) / rows_per_block));
  colsum_bf16_kernel<<<grid, 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(x_bf16), out, R, C,
                                               rows_per_block);
  MTB_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
