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

#ifdef MTB_TRACE
__device__ long long g_trace2[64];
#define TRACE2(i) do { if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) g_trace2[i] = clock64(); } while (0)
#else
#define TRACE2(i) do { } while (0)
#endif

// one warp per row; T = activation type (bf16 or fp32)
template <class T>
__global__ void __launch_bounds__(256)
fc2_softmax_ce_kernel(const T* __restrict__ h, const float* __restrict__ W /*[F,NC]*/,
                      const float* __restrict__ bias, const int* __restrict__ labels, float* __restrict__ logits,
                      float* __restrict__ probs, float* __restrict__ dlogits, float* __restrict__ loss_accum,
                      int* __restrict__ err_accum, int B, int F, int NC) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= B) return;
  float acc[kMaxClasses];
#pragma unroll
  for (int c = 0; c < kMaxClasses; ++c) acc[c] = 0.f;
  const T* hr = h + (size_t)warp * F;
  for (int f = lane; f < F; f += 32) {
    const float hv = ldf(hr + f);
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
template <class T>
__global__ void __launch_bounds__(256)
fc2_dgrad_kernel(const float* __restrict__ dlogits, const float* __restrict__ W, const T* __restrict__ h,
                 T* __restrict__ dh, int B, int F, int NC, float scale) {
  const int b = blockIdx.y;
  __shared__ float dl[kMaxClasses];
  if (threadIdx.x < NC) dl[threadIdx.x] = dlogits[(size_t)b * NC + threadIdx.x];
  __syncthreads();
  for (int f = blockIdx.x * blockDim.x + threadIdx.x; f < F; f += gridDim.x * blockDim.x) {
    float acc = 0.f;
    const float* wr = W + (size_t)f * NC;
    for (int c = 0; c < NC; ++c) acc = fmaf(dl[c], __ldg(wr + c), acc);
    const float hv = ldf(h + (size_t)b * F + f);
    stf(dh + (size_t)b * F + f, hv > 0.f ? acc * scale : 0.f);
  }
}

// dW[f,c] += sum_b h[b,f]*dlogits[b,c];  db[c] += sum_b dlogits[b,c].  lanes <-> f.
template <class T>
__global__ void __launch_bounds__(128)
fc2_wgrad_kernel(const T* __restrict__ h, const float* __restrict__ dlogits, float* __restrict__ dW,
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
      const float hv = ldf(h + (size_t)b * F + f);
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
template <class T>
__global__ void __launch_bounds__(256)
colsum_bf16_kernel(const T* __restrict__ x, float* __restrict__ out, long long R, int C,
                   int rows_per_block) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const long long r0 = (long long)blockIdx.y * rows_per_block;
  const long long r1 = r0 + rows_per_block < R ? r0 + rows_per_block : R;
  float acc = 0.f;
  for (long long r = r0; r < r1; ++r) acc += ldf(x + r * C + c);
  atomicAdd(out + c, acc);
}

// ---------------------------------------------------------------------------------------------
// Faster variants used by the engine.
//
// Forward: 32 rows per block, lane <-> row, the 8 warps split the features; W sits in shared
// memory and every read of it is a warp-wide broadcast, so the loop is pure FMA.  Loss and error
// count are reduced inside the block first: one atomic per block instead of one per row.
constexpr int kHeadWarps = 16;

// 8 consecutive activations as floats; 2 consecutive activations in / out (bf16 or fp32 storage)
__device__ __forceinline__ void load8f(const __nv_bfloat16* p, float (&f)[8]) {
  const uint4 pk = __ldg(reinterpret_cast<const uint4*>(p));
  const uint32_t w4[4] = {pk.x, pk.y, pk.z, pk.w};
#pragma unroll
  for (int e = 0; e < 4; ++e) { const float2 t = unpack_bf16x2(w4[e]); f[2 * e] = t.x; f[2 * e + 1] = t.y; }
}
__device__ __forceinline__ void load8f(const float* p, float (&f)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
__device__ __forceinline__ float2 load2f(const __nv_bfloat16* p) {
  return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p));
}
__device__ __forceinline__ float2 load2f(const float* p) { return *reinterpret_cast<const float2*>(p); }
// returns what was stored (after rounding to the storage type)
__device__ __forceinline__ float2 store2f(__nv_bfloat16* p, float a, float b) {
  const __nv_bfloat162 q = __floats2bfloat162_rn(a, b);
  *reinterpret_cast<__nv_bfloat162*>(p) = q;
  return __bfloat1622float2(q);
}
__device__ __forceinline__ float2 store2f(float* p, float a, float b) {
  *reinterpret_cast<float2*>(p) = make_float2(a, b);
  return make_float2(a, b);
}

template <int NC, class T>
__global__ void __launch_bounds__(kHeadWarps * 32)
fc2_softmax_ce_rows_kernel(const T* __restrict__ h, const float* __restrict__ W /*[F,NC]*/,
                           const float* __restrict__ bias, const int* __restrict__ labels, float* __restrict__ logits,
                           float* __restrict__ probs, float* __restrict__ dlogits, float* __restrict__ loss_accum,
                           int* __restrict__ err_accum, int B, int F, const long long* win_step, long long win_rows) {
  extern __shared__ float sm[];
  pdl_sync();
  if (labels != nullptr && win_step != nullptr) labels += ((*win_step) * (long long)B) % (win_rows - B);
  float* sW = sm;                                      // [F][NC]
  float* red = sm + (size_t)F * NC;                    // [kHeadWarps][NC][32 rows]
  {
    const float4* src = reinterpret_cast<const float4*>(W);
    float4* dst = reinterpret_cast<float4*>(sW);
    for (int i = threadIdx.x; i < F * NC / 4; i += blockDim.x) dst[i] = __ldg(src + i);
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 32 + lane;
  const int fper = F / kHeadWarps, fbeg = warp * fper;
  float acc[NC];
#pragma unroll
  for (int c = 0; c < NC; ++c) acc[c] = 0.f;
  if (row < B) {
    const T* hr = h + (size_t)row * F + fbeg;
#pragma unroll 8
    for (int i = 0; i < fper / 8; ++i) {
      float hv[8];
      load8f(hr + i * 8, hv);
      const float* wr = sW + (size_t)(fbeg + i * 8) * NC;
#pragma unroll
      for (int e = 0; e < 8; ++e) {
#pragma unroll
        for (int c = 0; c < NC; ++c) acc[c] = fmaf(hv[e], wr[e * NC + c], acc[c]);
      }
    }
  }
#pragma unroll
  for (int c = 0; c < NC; ++c) red[(warp * NC + c) * 32 + lane] = acc[c];
  __syncthreads();
  if (warp != 0) return;
  float lg[NC];
  float mx = -INFINITY;
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    float s = __ldg(bias + c);
#pragma unroll
    for (int w = 0; w < kHeadWarps; ++w) s += red[(w * NC + c) * 32 + lane];
    lg[c] = s;
    mx = fmaxf(mx, s);
  }
  float s = 0.f, e[NC];
#pragma unroll
  for (int c = 0; c < NC; ++c) { e[c] = expf(lg[c] - mx); s += e[c]; }
  const float inv = 1.0f / s, invB = 1.0f / (float)B;
  const bool valid = row < B;
  const int y = (labels && valid) ? labels[row] : -1;
  int arg = 0;
  float best = -INFINITY, ly = 0.f;
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    if (lg[c] > best) { best = lg[c]; arg = c; }
    if (c == y) ly = lg[c];
  }
  if (valid) {
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      const float p = e[c] * inv;
      if (logits) logits[(size_t)row * NC + c] = lg[c];
      if (probs) probs[(size_t)row * NC + c] = p;
      if (dlogits) dlogits[(size_t)row * NC + c] = (p - (c == y ? 1.f : 0.f)) * invB;
    }
  }
  float l = (valid && y >= 0) ? -(ly - mx - logf(s)) * invB : 0.f;
  l = warp_sum(l);
  const unsigned wrong = __ballot_sync(0xffffffffu, valid && y >= 0 && arg != y);
  if (lane == 0) {
    if (loss_accum && labels) atomicAdd(loss_accum, l);
    if (err_accum && wrong) atomicAdd(err_accum, __popc(wrong));
  }
}

// Backward of FC2 in one pass over h: weight gradient, bias gradient, input gradient with the
// dropout/ReLU mask, and the column sums of that input gradient (= FC1's bias gradient).
//   grid (F/64, splits), 256 threads: lane <-> 2 features, warps and blockIdx.y split the rows.
template <int NC, class T>
__global__ void __launch_bounds__(256)
fc2_bwd_fused_kernel(const T* __restrict__ h, const float* __restrict__ dlogits,
                     const float* __restrict__ W /*[F,NC]*/, T* __restrict__ dh, float* __restrict__ dW,
                     float* __restrict__ db2, float* __restrict__ db1, int B, int F, float scale, int rows_per_block) {
  extern __shared__ float sm[];                        // dl [rows_per_block][NC], then the reduction buffer
  TRACE2(0);
  pdl_sync();
  TRACE2(1);
  float* sdl = sm;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int f0 = blockIdx.x * 64 + lane * 2;
  const int b0 = blockIdx.y * rows_per_block;
  const int nrows = min(rows_per_block, B - b0);
  for (int i = threadIdx.x; i < nrows * NC; i += blockDim.x) sdl[i] = dlogits[(size_t)b0 * NC + i];
  float w0[NC], w1[NC], a0[NC], a1[NC];
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    w0[c] = f0 < F ? __ldg(W + (size_t)f0 * NC + c) : 0.f;
    w1[c] = f0 + 1 < F ? __ldg(W + (size_t)(f0 + 1) * NC + c) : 0.f;
    a0[c] = 0.f; a1[c] = 0.f;
  }
  float cs0 = 0.f, cs1 = 0.f;
  __syncthreads();
  TRACE2(2);
  if (f0 < F) {
    // 4 rows per pass, their loads issued together (the loop is latency-bound otherwise)
    for (int r0 = warp; r0 < nrows; r0 += 32) {
      float2 hq[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int r = r0 + 8 * u;
        if (r < nrows) hq[u] = load2f(h + (size_t)(b0 + r) * F + f0);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int r = r0 + 8 * u;
        if (r >= nrows) break;
        const float* dl = sdl + r * NC;
        const size_t o = (size_t)(b0 + r) * F + f0;
        const float2 hv = hq[u];
        float t0 = 0.f, t1 = 0.f;
#pragma unroll
        for (int c = 0; c < NC; ++c) {
          const float d = dl[c];
          t0 = fmaf(d, w0[c], t0);
          t1 = fmaf(d, w1[c], t1);
          a0[c] = fmaf(hv.x, d, a0[c]);
          a1[c] = fmaf(hv.y, d, a1[c]);
        }
        t0 = hv.x > 0.f ? t0 * scale : 0.f;
        t1 = hv.y > 0.f ? t1 * scale : 0.f;
        const float2 qf = store2f(dh + o, t0, t1);       // bias gradient of FC1 sums what FC1's wgrad will see
        cs0 += qf.x;
        cs1 += qf.y;
      }
    }
  }
  __syncthreads();                                      // everyone is done with sdl
  TRACE2(3);
  float* red = sm;                                      // [8 warps][2*NC+2][32 lanes]
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    red[(warp * (2 * NC + 2) + c) * 32 + lane] = a0[c];
    red[(warp * (2 * NC + 2) + NC + c) * 32 + lane] = a1[c];
  }
  red[(warp * (2 * NC + 2) + 2 * NC) * 32 + lane] = cs0;
  red[(warp * (2 * NC + 2) + 2 * NC + 1) * 32 + lane] = cs1;
  __syncthreads();
  for (int i = threadIdx.x; i < (2 * NC + 2) * 32; i += blockDim.x) {
    const int k = i >> 5, l = i & 31;
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red[(w * (2 * NC + 2) + k) * 32 + l];
    const int f = blockIdx.x * 64 + l * 2;
    if (f >= F) continue;
    if (k < NC) atomicAdd(dW + (size_t)f * NC + k, s);
    else if (k < 2 * NC) { if (f + 1 < F) atomicAdd(dW + (size_t)(f + 1) * NC + (k - NC), s); }
    else if (k == 2 * NC) atomicAdd(db1 + f, s);
    else if (f + 1 < F) atomicAdd(db1 + f + 1, s);
  }
  TRACE2(4);
  if (blockIdx.x == 0 && threadIdx.x < NC) {            // FC2 bias gradient: column sums of dlogits
    float s = 0.f;
    for (int r = 0; r < nrows; ++r) s += dlogits[(size_t)(b0 + r) * NC + threadIdx.x];
    atomicAdd(db2 + threadIdx.x, s);
  }
  TRACE2(5);
}

// Same head, 8 rows per block instead of 32: four times as many blocks (128 at batch 1024, 13 at batch 100 - the
// 32-row form ran on 32 and 4 SMs) and an eighth of the shared-memory reads per block.  Lane = (row r = lane & 7,
// quarter q = lane >> 3); warp w owns features [w F/16, (w+1) F/16) and each lane a contiguous quarter of them, so a lane
// loads whole 16-byte vectors of its row.  W rows are stored in shared memory with 8 extra words per 16 feature chunk
// group so that the four quarters of a warp read four different banks; products are read as float2 (LDS.64).
template <int NC, class T>
__global__ void __launch_bounds__(kHeadWarps * 32)
fc2_softmax_ce_rows8_kernel(const T* __restrict__ h, const float* __restrict__ W /*[F,NC]*/,
                            const float* __restrict__ bias, const int* __restrict__ labels, float* __restrict__ logits,
                            float* __restrict__ probs, float* __restrict__ dlogits, float* __restrict__ loss_accum,
                            int* __restrict__ err_accum, int B, int F, const long long* win_step, long long win_rows) {
  static_assert(NC % 2 == 0, "float2 reads of a weight row");
  extern __shared__ float sm[];
  pdl_sync();
  if (labels != nullptr && win_step != nullptr) labels += ((*win_step) * (long long)B) % (win_rows - B);
  const int chunk = F / (kHeadWarps * 4);              // features per lane
  // feature f lives at sW[f * NC + (f / chunk) * 8]: consecutive quarters are 8 words apart modulo 32 banks
  float* sW = sm;
  float* red = sm + (size_t)F * NC + (F / chunk) * 8;  // [kHeadWarps][8 rows][NC]
  for (int i = threadIdx.x; i < F * NC / 2; i += blockDim.x) {
    const int f = (2 * i) / NC;
    *reinterpret_cast<float2*>(sW + 2 * i + (f / chunk) * 8) = __ldg(reinterpret_cast<const float2*>(W) + i);
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = lane & 7, q = lane >> 3;
  const int row = blockIdx.x * 8 + r;
  const int f0 = (warp * 4 + q) * chunk;
  float acc[NC];
#pragma unroll
  for (int c = 0; c < NC; ++c) acc[c] = 0.f;
  if (row < B) {
    const T* hr = h + (size_t)row * F + f0;
    const float* wbase = sW + (size_t)f0 * NC + (f0 / chunk) * 8;
    for (int i = 0; i < chunk / 8; ++i) {
      float hv[8];
      load8f(hr + i * 8, hv);
      const float2* wr = reinterpret_cast<const float2*>(wbase + (size_t)i * 8 * NC);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
#pragma unroll
        for (int c = 0; c < NC / 2; ++c) {
          const float2 w2 = wr[e * (NC / 2) + c];
          acc[2 * c] = fmaf(hv[e], w2.x, acc[2 * c]);
          acc[2 * c + 1] = fmaf(hv[e], w2.y, acc[2 * c + 1]);
        }
      }
    }
  }
#pragma unroll
  for (int c = 0; c < NC; ++c) {                       // the four quarters of a row, then the 16 warps through smem
    acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], 8);
    acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], 16);
  }
  if (q == 0) {
#pragma unroll
    for (int c = 0; c < NC; ++c) red[(warp * 8 + r) * NC + c] = acc[c];
  }
  __syncthreads();
  if (warp != 0) return;
  const bool valid = lane < 8 && row < B;
  float lg[NC];
  float mx = -INFINITY;
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    float s = __ldg(bias + c);
    if (lane < 8) {
#pragma unroll
      for (int w = 0; w < kHeadWarps; ++w) s += red[(w * 8 + r) * NC + c];
    }
    lg[c] = s;
    mx = fmaxf(mx, s);
  }
  float s = 0.f, e[NC];
#pragma unroll
  for (int c = 0; c < NC; ++c) { e[c] = expf(lg[c] - mx); s += e[c]; }
  const float inv = 1.0f / s, invB = 1.0f / (float)B;
  const int y = (labels && valid) ? labels[row] : -1;
  int arg = 0;
  float best = -INFINITY, ly = 0.f;
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    if (lg[c] > best) { best = lg[c]; arg = c; }
    if (c == y) ly = lg[c];
  }
  if (valid) {
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      const float p = e[c] * inv;
      if (logits) logits[(size_t)row * NC + c] = lg[c];
      if (probs) probs[(size_t)row * NC + c] = p;
      if (dlogits) dlogits[(size_t)row * NC + c] = (p - (c == y ? 1.f : 0.f)) * invB;
    }
  }
  float l = (valid && y >= 0) ? -(ly - mx - logf(s)) * invB : 0.f;
  l = warp_sum(l);
  const unsigned wrong = __ballot_sync(0xffffffffu, valid && y >= 0 && arg != y);
  if (lane == 0) {
    if (loss_accum && labels) atomicAdd(loss_accum, l);
    if (err_accum && wrong) atomicAdd(err_accum, __popc(wrong));
  }
}

template <class T>
int run_head_fast(const T* h, const float* W, const float* bias, const int* labels, float* logits, float* probs,
                  float* dlogits, float* loss_accum, int* err_accum, int B, int F, int NC, const long long* win_step,
                  long long win_rows, cudaStream_t stream) {
  MTB_REQUIRE(NC == 10 && F % (8 * kHeadWarps) == 0, "fc2_softmax_ce_fast: 10 classes, F a multiple of 128");
  MTB_REQUIRE(win_step == nullptr || win_rows > B, "fc2_softmax_ce_fast: shard must be larger than the batch");
  if (F % (kHeadWarps * 4 * 8) == 0) {                   // every lane owns whole 8-feature vectors: the 8-row form
    const int smem8 = (F * NC + kHeadWarps * 4 * 8 + kHeadWarps * 8 * NC) * (int)sizeof(float);
    if (smem8 <= 200 * 1024) {
      static int attr8[64] = {};
      int dev = 0;
      MTB_CUDA_CHECK(cudaGetDevice(&dev));
      if (dev < 0 || dev >= 64 || smem8 > attr8[dev]) {
        MTB_CUDA_CHECK(cudaFuncSetAttribute(fc2_softmax_ce_rows8_kernel<10, T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            smem8));
        if (dev >= 0 && dev < 64) attr8[dev] = smem8;
      }
      MTB_CUDA_CHECK(mtb::launch(fc2_softmax_ce_rows8_kernel<10, T>, dim3((B + 7) / 8), dim3(kHeadWarps * 32), (size_t)smem8,
                                 stream, h, W, bias, labels, logits, probs, dlogits, loss_accum, err_accum, B, F, win_step,
                                 win_rows));
      return 0;
    }
  }
  const int smem = (F * NC + kHeadWarps * NC * 32) * (int)sizeof(float);
  MTB_REQUIRE(smem <= 200 * 1024, "fc2_softmax_ce_fast: weight matrix does not fit in shared memory");
  static int attr_bytes = 0;
  if (smem > attr_bytes) {
    MTB_CUDA_CHECK(cudaFuncSetAttribute(fc2_softmax_ce_rows_kernel<10, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_bytes = smem;
  }
  MTB_CUDA_CHECK(mtb::launch(fc2_softmax_ce_rows_kernel<10, T>, dim3((B + 31) / 32), dim3(kHeadWarps * 32), (size_t)smem,
                             stream, h, W, bias, labels, logits, probs, dlogits, loss_accum, err_accum, B, F, win_step,
                             win_rows));
  return 0;
}

template <class T>
int run_bwd_fused(const T* h, const float* dlogits, const float* W, T* dh, float* dW, float* db2, float* db1, int B,
                  int F, int NC, float scale, cudaStream_t stream) {
  MTB_REQUIRE(NC == 10 && F % 2 == 0, "fc2_bwd_fused: 10 classes, even F");
  int splits = 1;
  const int fx = (F + 63) / 64;
  // two blocks (16 warps) per SM: the row loop is a dependent FMA chain and needs the extra warps to hide latency
  while (fx * splits < 256 && B / (splits * 2) >= 32) splits *= 2;
  const int rows_per_block = (B + splits - 1) / splits;
  const int smem = (int)sizeof(float) * (rows_per_block * NC > 8 * 22 * 32 ? rows_per_block * NC : 8 * 22 * 32);
  MTB_REQUIRE(smem <= 48 * 1024, "fc2_bwd_fused: batch slice too large");
  dim3 grid(fx, splits);
  MTB_CUDA_CHECK(mtb::launch(fc2_bwd_fused_kernel<10, T>, grid, dim3(256), (size_t)smem, stream, h, dlogits, W, dh, dW,
                             db2, db1, B, F, scale, rows_per_block));
  return 0;
}

}  // namespace

extern "C" {

#ifdef MTB_TRACE
int mtb_debug_trace2_read(long long* host_out) {
  MTB_CUDA_CHECK(cudaMemcpyFromSymbol(host_out, g_trace2, sizeof(long long) * 64));
  return 0;
}
#endif

int mtb_fc2_softmax_ce_fast(const void* h_bf16, const float* W, const float* bias, const int* labels, float* logits,
                            float* probs, float* dlogits, float* loss_accum, int* err_accum, int B, int F, int NC,
                            cudaStream_t stream) {
  return mtb_fc2_softmax_ce_fast_win(h_bf16, W, bias, labels, logits, probs, dlogits, loss_accum, err_accum, B, F, NC,
                                     nullptr, 0, stream);
}

// labels may be a resident shard of win_rows labels: the batch is the window of step *win_step (see conv1_sm100.cu)
int mtb_fc2_softmax_ce_fast_win(const void* h_bf16, const float* W, const float* bias, const int* labels,
                                float* logits, float* probs, float* dlogits, float* loss_accum, int* err_accum, int B,
                                int F, int NC, const long long* win_step, long long win_rows, cudaStream_t stream) {
  return run_head_fast(reinterpret_cast<const __nv_bfloat16*>(h_bf16), W, bias, labels, logits, probs, dlogits,
                       loss_accum, err_accum, B, F, NC, win_step, win_rows, stream);
}
int mtb_fc2_softmax_ce_fast_win_f32(const float* h, const float* W, const float* bias, const int* labels,
                                    float* logits, float* probs, float* dlogits, float* loss_accum, int* err_accum,
                                    int B, int F, int NC, const long long* win_step, long long win_rows,
                                    cudaStream_t stream) {
  return run_head_fast(h, W, bias, labels, logits, probs, dlogits, loss_accum, err_accum, B, F, NC, win_step, win_rows,
                       stream);
}

// dW, db2 and db1 are accumulated into; dh is overwritten.  NC must be 10.
int mtb_fc2_bwd_fused(const void* h_bf16, const float* dlogits, const float* W, void* dh_bf16, float* dW, float* db2,
                      float* db1, int B, int F, int NC, float scale, cudaStream_t stream) {
  return run_bwd_fused(reinterpret_cast<const __nv_bfloat16*>(h_bf16), dlogits, W,
                       reinterpret_cast<__nv_bfloat16*>(dh_bf16), dW, db2, db1, B, F, NC, scale, stream);
}
int mtb_fc2_bwd_fused_f32(const float* h, const float* dlogits, const float* W, float* dh, float* dW, float* db2,
                          float* db1, int B, int F, int NC, float scale, cudaStream_t stream) {
  return run_bwd_fused(h, dlogits, W, dh, dW, db2, db1, B, F, NC, scale, stream);
}

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

// ---- fp32-activation twins of the four generic kernels (TF32 / fp32 training path) ----
int mtb_fc2_softmax_ce_f32(const float* h, const float* W, const float* bias, const int* labels, float* logits,
                           float* probs, float* dlogits, float* loss_accum, int* err_accum, int B, int F, int NC,
                           cudaStream_t stream) {
  MTB_REQUIRE(NC > 0 && NC <= kMaxClasses, "fc2_softmax_ce: at most 16 classes");
  const int warps_per_block = 8;
  fc2_softmax_ce_kernel<<<(B + warps_per_block - 1) / warps_per_block, 256, 0, stream>>>(
      h, W, bias, labels, logits, probs, dlogits, loss_accum, err_accum, B, F, NC);
  MTB_LAUNCH_CHECK();
  return 0;
}

int mtb_fc2_dgrad_f32(const float* dlogits, const float* W, const float* h, float* dh, int B, int F, int NC,
                      float scale, cudaStream_t stream) {
  MTB_REQUIRE(NC > 0 && NC <= kMaxClasses, "fc2_dgrad: at most 16 classes");
  dim3 grid((F + 255) / 256, B);
  fc2_dgrad_kernel<<<grid, 256, 0, stream>>>(dlogits, W, h, dh, B, F, NC, scale);
  MTB_LAUNCH_CHECK();
  return 0;
}

int mtb_fc2_wgrad_f32(const float* h, const float* dlogits, float* dW, float* db, int B, int F, int NC,
                      cudaStream_t stream) {
  MTB_REQUIRE(NC > 0 && NC <= kMaxClasses, "fc2_wgrad: at most 16 classes");
  int rows_per_block = 32;
  dim3 grid((F + 127) / 128, (B + rows_per_block - 1) / rows_per_block);
  fc2_wgrad_kernel<<<grid, 128, 0, stream>>>(h, dlogits, dW, db, B, F, NC, rows_per_block);
  MTB_LAUNCH_CHECK();
  return 0;
}

int mtb_colsum_f32(const float* x, float* out, long long R, int C, cudaStream_t stream) {
  int rows_per_block = 64;
  dim3 grid((C + 255) / 256, (unsigned)((R + rows_per_block - 1) / rows_per_block));
  colsum_bf16_kernel<<<grid, 256, 0, stream>>>(x, out, R, C, rows_per_block);
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
