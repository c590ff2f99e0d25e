// First convolution layer (5x5, SAME, 1 -> 32 channels on 28x28 images; mpipy.py:156-158 and its
// Conv2DBackpropFilter / ReluGrad / MaxPoolGrad ops) on tcgen05.
//
// One input channel means the implicit GEMM has K = 25, far too thin for TMA tiles (a pixel is
// 2 bytes).  Instead each image is expanded 5x, not 25x, by CUDA cores into a shared-memory
// operand "X5": one 32-byte row per pixel holding the pixel's five vertical neighbours
//     X5[row(h,w)][kh] = x_pad[h+kh][w],  kh = 0..4;   X5[..][5] = 1;   X5[..][6..15] = 0
// so the vertical taps are the K dimension (K = 16 = one bf16 MMA) and the horizontal tap kw is
// a start-address shift of the same operand, as in conv2_sm100.cu.  Rows are ordered
//     row = 128*(h/4) + 4*w + (h%4)          (w = 0..31, columns 28..31 are padding)
// so (a) tap kw is a shift of 4*kw rows and (b) the four pixels of a 2x2 pooling window sit in
// lanes L, L^1, L^4, L^5 of one warp: bias + ReLU + max-pool + argmax happen in registers with
// 24 shuffles per 32 channels, and only the pooled bf16 map and the 2-bit argmax leave the SM.
//
//   fwd    D[pixel, c] = sum_kw X5[pixel + 4kw][0..15] . Wk[kw][c][0..15]      35 MMAs (128x32x16) per image
//   wgrad  D[(kw,kh), c] = sum_pixel X5[pixel + 4kw][kh] . dA[pixel][c]        56 MMAs per image; both operands
//          MN-major; the 8 M-atoms of the A operand are X5 moved down 4 rows each (kw = 0..7, 5 used).
//          dA = pooled gradient routed to the argmax and gated by ReLU, rebuilt in shared memory
//          from the pooled maps (never materialised in global memory).  The ones column of X5 makes
//          lane (kw=0, kh=5) the bias gradient.
#include "common.cuh"
#include "sm100.cuh"

namespace {

using namespace sm100;
using bf16 = __nv_bfloat16;

// Optional role-level timeline of CTA 0 (build with -DMTB_TRACE): 8 slots per event row.
#ifdef MTB_TRACE
__device__ long long g_trace[64 * 16];
#define TRACE(row, col) do { if (blockIdx.x == 0) g_trace[(row) * 16 + (col)] = clock64(); } while (0)
#else
#define TRACE(row, col) do { } while (0)
#endif

constexpr int IMG = 28, C1 = 32;
constexpr int X5_ROWS = 928;                 // 896 + 32: kw shifts (<= 28 rows) of the last tile stay in bounds
constexpr int X5_BYTES = X5_ROWS * 32;       // 29696
constexpr int XS_PITCH = 33;
constexpr int XS_FLOATS = 32 * XS_PITCH;     // padded 32x32 image, pitch 33

// The reference's sequential batch window (mpipy.py:80-82) evaluated where the data is read:
// rows [off, off+B) of the resident shard with off = (step*B) % (rows-B), step on the device.
__device__ __forceinline__ long long window_offset(const long long* win_step, long long win_rows, int B) {
  return win_step != nullptr ? ((*win_step) * (long long)B) % (win_rows - B) : 0;
}

__device__ __forceinline__ void named_bar(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

// padded image -> shared floats; xs[i][j] = x[i-2][j-2] (zero outside).  NT threads, all loads issued first.
template <int NT>
__device__ __forceinline__ void load_xs(float* xs, const float* __restrict__ x, int t) {
  constexpr int PER = 1024 / NT;
  float v[PER];
#pragma unroll
  for (int k = 0; k < PER; ++k) {
    const int i = t + k * NT;
    const int r = (i >> 5) - 2, c = (i & 31) - 2;
    v[k] = (r >= 0 && r < IMG && c >= 0 && c < IMG) ? __ldg(x + r * IMG + c) : 0.f;
  }
#pragma unroll
  for (int k = 0; k < PER; ++k) {
    const int i = t + k * NT;
    xs[(i >> 5) * XS_PITCH + (i & 31)] = v[k];
  }
}

// X5 rows [0, 896) from the padded image.  Row R = 128t + 4w + hl  <->  pixel (h = 4t + hl, w).
__device__ __forceinline__ void build_x5(uint8_t* x5, const float* xs, int t, int nthreads) {
  for (int R = t; R < 896; R += nthreads) {
    const int tile = R >> 7, w = (R >> 2) & 31, hl = R & 3;
    const int h = tile * 4 + hl;
    const float* p = xs + h * XS_PITCH + w;
    uint4 v;
    v.x = mtb::pack_bf16x2(p[0], p[XS_PITCH]);
    v.y = mtb::pack_bf16x2(p[2 * XS_PITCH], p[3 * XS_PITCH]);
    v.z = mtb::pack_bf16x2(p[4 * XS_PITCH], 1.0f);
    v.w = 0u;
    const int sw = w & 1;                                   // SW32: chunk ^= bit 7 of the byte address = (R>>2)&1
    *reinterpret_cast<uint4*>(x5 + R * 32 + (sw << 4)) = v;
    *reinterpret_cast<uint4*>(x5 + R * 32 + ((sw ^ 1) << 4)) = make_uint4(0, 0, 0, 0);
  }
}

// ======================================================================================= forward
// Forward-pass row order of X5 ("planes"): a 2x2 pooling window must end up in ONE thread, so the
// four window positions (dy,dx) become four separate accumulator tiles whose row r is the POOLED
// pixel (ph = r/16, pw = r%16).  X5 is stored as four planes (dy, p): plane row 16*ph + q holds
// pixel row h = 2ph+dy, padded column 2q+p.  Output position (dy,dx), tap kw reads padded column
// 2pw+dx+kw, i.e. plane (dy, (dx+kw)&1) shifted by (dx+kw)>>1 rows: still a pure start-address
// shift.  The epilogue thread of pooled pixel r loads its four accumulators and pools in
// registers: no shuffles, no shared staging.
namespace f1 {
constexpr int kProducers = 128, kEpilogue = 256;            // kProducers per stage; one producer group per stage
constexpr int kThreads = 32 + 2 * kProducers + kEpilogue;    // warp 0 MMA, 1-8 producers, 9-16 epilogue
constexpr int kPlaneRows = 14 * 16;                          // 224
constexpr int kX5Rows = 4 * kPlaneRows + 40;                 // tile 1 of the last plane reads up to row 3*224+255+2
constexpr int kX5Bytes = ((kX5Rows * 32 + 1023) / 1024) * 1024;   // 30720
constexpr int kOffX5 = 0;                                    // 2 stages
constexpr int kOffW = 2 * kX5Bytes;                          // 5 x [32 c][16 kh] SW32 = 5 KB
constexpr int kOffXs = kOffW + 5 * 1024;                     // 2 x padded image
constexpr int kOffBar = kOffXs + 2 * XS_FLOATS * 4;
constexpr int kSmem = 1024 + kOffBar + 128;
}  // namespace f1

// X5 planes from the padded image.  Row R = 224*(2dy+p) + 16ph + q  <->  pixel row 2ph+dy, padded column 2q+p.
__device__ __forceinline__ void build_x5_planes(uint8_t* x5, const float* xs, int t, int nthreads) {
  for (int R = t; R < 4 * f1::kPlaneRows; R += nthreads) {
    const int pl = R / f1::kPlaneRows, rr = R % f1::kPlaneRows;
    const int dy = pl >> 1, p = pl & 1, ph = rr >> 4, q = rr & 15;
    const float* src = xs + (2 * ph + dy) * XS_PITCH + 2 * q + p;
    uint4 v;
    v.x = mtb::pack_bf16x2(src[0], src[XS_PITCH]);
    v.y = mtb::pack_bf16x2(src[2 * XS_PITCH], src[3 * XS_PITCH]);
    v.z = mtb::pack_bf16x2(src[4 * XS_PITCH], 1.0f);
    v.w = 0u;
    const int sw = (R >> 2) & 1;                             // SW32: chunk ^= bit 7 of the byte address
    *reinterpret_cast<uint4*>(x5 + R * 32 + (sw << 4)) = v;
    *reinterpret_cast<uint4*>(x5 + R * 32 + ((sw ^ 1) << 4)) = make_uint4(0, 0, 0, 0);
  }
}

struct ZeroList { void* ptr[8]; long long n_words[8]; int count; };

__global__ void __launch_bounds__(f1::kThreads, 1)
conv1_fwd_tc_kernel(const float* __restrict__ x, const float* __restrict__ w /*[25][32]*/, const float* __restrict__ bias,
                    bf16* __restrict__ out, uint8_t* __restrict__ idx, int B, const long long* win_step,
                    long long win_rows, const ZeroList zl) {
  using namespace f1;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sX5 = smem + kOffX5;
  uint8_t* sW = smem + kOffW;
  float* sXs = reinterpret_cast<float*>(smem + kOffXs);
  uint64_t* x_full = reinterpret_cast<uint64_t*>(smem + kOffBar);
  uint64_t* x_empty = x_full + 2;
  uint64_t* a_full = x_empty + 2;
  uint64_t* a_empty = a_full + 2;
  uint32_t* holder = reinterpret_cast<uint32_t*>(a_empty + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // zero both X5 stages (tail rows stay zero for the kernel's lifetime)
  for (int i = threadIdx.x; i < 2 * kX5Bytes / 16; i += kThreads) reinterpret_cast<uint4*>(sX5)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(&x_full[s], kProducers);
      mbar_init(&x_empty[s], 1);
      mbar_init(&a_full[s], 1);
      mbar_init(&a_empty[s], kEpilogue / 32);
    }
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc<512>(holder);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *holder;
  mtb::pdl_sync();
  // first kernel of the step: clear the accumulators later kernels add into (they all run after this one)
  for (int k = 0; k < zl.count; ++k) {
    uint32_t* p = reinterpret_cast<uint32_t*>(zl.ptr[k]);
    for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < zl.n_words[k]; i += (long long)gridDim.x * kThreads)
      p[i] = 0u;
  }
  x += window_offset(win_step, win_rows, B) * (IMG * IMG);
  // per-kw weight tiles [32 c][16 kh] (K-major, SW32) from the fp32 master weights
  for (int i = threadIdx.x; i < 5 * C1; i += kThreads) {
    const int kw = i / C1, c = i % C1;
    uint4 v;
    v.x = mtb::pack_bf16x2(__ldg(w + (0 * 5 + kw) * C1 + c), __ldg(w + (1 * 5 + kw) * C1 + c));
    v.y = mtb::pack_bf16x2(__ldg(w + (2 * 5 + kw) * C1 + c), __ldg(w + (3 * 5 + kw) * C1 + c));
    v.z = mtb::pack_bf16x2(__ldg(w + (4 * 5 + kw) * C1 + c), 0.f);
    v.w = 0u;
    const int sw = (c >> 2) & 1;
    *reinterpret_cast<uint4*>(sW + kw * 1024 + c * 32 + (sw << 4)) = v;
    *reinterpret_cast<uint4*>(sW + kw * 1024 + c * 32 + ((sw ^ 1) << 4)) = make_uint4(0, 0, 0, 0);
  }
  fence_proxy_async_smem();
  __syncthreads();

  if (warp == 0) {
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc(kFmtBF16, 128, C1, false, false);
      int it = 0;
      for (int img = blockIdx.x; img < B; img += gridDim.x, ++it) {
        const int s = it & 1, ph = (it >> 1) & 1;
        TRACE(it, 0);
        mbar_wait(&x_full[s], ph);
        TRACE(it, 1);
        mbar_wait(&a_empty[s], ph ^ 1);
        TRACE(it, 2);
        tc_fence_after();
        const uint64_t a0 = make_smem_desc(smem_u32(sX5 + s * kX5Bytes), 16, 256, kSwizzle32B);
        const uint64_t b0 = make_smem_desc(smem_u32(sW), 16, 256, kSwizzle32B);
        // accumulator (dy,dx,tile) lives in columns 32*(4dy + 2dx + tile) of this image's TMEM half
#pragma unroll
        for (int pos = 0; pos < 4; ++pos) {
          const int dy = pos >> 1, dx = pos & 1;
#pragma unroll
          for (int tile = 0; tile < 2; ++tile) {
            const uint32_t d = tmem + s * 256 + (pos * 2 + tile) * 32;
#pragma unroll
            for (int kw = 0; kw < 5; ++kw) {
              const int row = (2 * dy + ((dx + kw) & 1)) * kPlaneRows + tile * 128 + ((dx + kw) >> 1);
              const uint64_t da = a0 + (uint64_t)((row * 32) >> 4);
              const uint64_t db = b0 + (uint64_t)((kw * 1024) >> 4);
              if (kw == 0) mma_f16_first(d, da, db, idesc);
              else mma_f16_acc(d, da, db, idesc);
            }
          }
        }
        mma_commit(&x_empty[s]);
        mma_commit(&a_full[s]);
        TRACE(it, 3);
      }
    }
    __syncwarp();
  } else if (warp <= 8) {
    // producer group g owns stage g and every second image of this CTA, so two images are in flight
    const int g = (warp - 1) >> 2;
    const int t = threadIdx.x - 32 - g * kProducers;
    float* xs = sXs + g * XS_FLOATS;
    int it = g;
    for (int img = blockIdx.x + g * gridDim.x; img < B; img += 2 * gridDim.x, it += 2) {
      const int ph = (it >> 1) & 1;
      if (t == 0) TRACE(it, 4);
      load_xs<kProducers>(xs, x + (size_t)img * IMG * IMG, t);   // xs is this group's own: free since the last bar
      named_bar(2 + g, kProducers);                          // padded image complete
      mbar_wait(&x_empty[g], ph ^ 1);                        // MMAs that read this X5 stage have retired
      if (t == 0) TRACE(it, 5);
      build_x5_planes(sX5 + g * kX5Bytes, xs, t, kProducers);
      fence_proxy_async_smem();
      mbar_arrive(&x_full[g]);
      if (t == 0) TRACE(it, 7);
      named_bar(2 + g, kProducers);                          // everyone is done reading xs
    }
  } else {
    const int q = warp & 3;
    const int tile = (warp - 9) >> 2;
    const int r = tile * 128 + q * 32 + lane;                // pooled pixel of this thread
    const int php = r >> 4, pwp = r & 15;
    const bool valid = php < 14 && pwp < 14;
    int it = 0;
    for (int img = blockIdx.x; img < B; img += gridDim.x, ++it) {
      const int s = it & 1, ph = (it >> 1) & 1;
      if (warp == 9 && lane == 0) TRACE(it, 8);
      mbar_wait(&a_full[s], ph);
      if (warp == 9 && lane == 0) TRACE(it, 9);
      tc_fence_after();
      const uint32_t tbase = tmem + ((uint32_t)(q * 32) << 16) + s * 256 + tile * 32;
      const size_t o = ((size_t)img * 196 + php * 14 + pwp) * C1;
#pragma unroll
      for (int half = 0; half < 2; ++half) {                 // 16 channels at a time keeps registers in check
        uint32_t v[4][16];
#pragma unroll
        for (int pos = 0; pos < 4; ++pos) tmem_ld_32x16(tbase + pos * 64 + half * 16, v[pos]);
        tmem_ld_wait();
        uint32_t pk[8], ik[4];
#pragma unroll
        for (int j = 0; j < 16; j += 2) {
          float val[2];
          uint32_t arg[2];
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            // value with its two low mantissa bits replaced by (3 - position): a float max picks the
            // largest value and, among equal non-negative values, the first position
            const float k0 = __uint_as_float((v[0][j + e] & ~3u) | 3u), k1 = __uint_as_float((v[1][j + e] & ~3u) | 2u);
            const float k2 = __uint_as_float((v[2][j + e] & ~3u) | 1u), k3 = __uint_as_float(v[3][j + e] & ~3u);
            const uint32_t m = __float_as_uint(fmaxf(fmaxf(k0, k1), fmaxf(k2, k3)));
            val[e] = fmaxf(__uint_as_float(m & ~3u) + __ldg(bias + half * 16 + j + e), 0.f);   // relu(max + b) = max relu(. + b)
            arg[e] = val[e] > 0.f ? 3u - (m & 3u) : 0u;      // an all-zero window reports position 0, like a first-max scan
          }
          pk[j >> 1] = mtb::pack_bf16x2(val[0], val[1]);
          if ((j & 2) == 0) ik[j >> 2] = arg[0] | (arg[1] << 8);
          else ik[j >> 2] |= (arg[0] << 16) | (arg[1] << 24);
        }
        if (valid) {
          *reinterpret_cast<uint4*>(out + o + half * 16) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
          *reinterpret_cast<uint4*>(out + o + half * 16 + 8) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
          *reinterpret_cast<uint4*>(idx + o + half * 16) = make_uint4(ik[0], ik[1], ik[2], ik[3]);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&a_empty[s]);
      if (warp == 9 && lane == 0) TRACE(it, 10);
    }
  }
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(tmem); }
}

// ======================================================================================= wgrad
namespace g1 {
constexpr int kProducers = 256;                              // per stage; one producer group per stage
constexpr int kThreads = 32 + 2 * kProducers + 128;          // warp 0 MMA, 1-16 producers, 17-20 final epilogue
constexpr int DA_BYTES = 896 * 64;                           // [pixel row][32 c] bf16, SW64
constexpr int kOffX5 = 0;
constexpr int kOffDA = 2 * X5_BYTES;                         // 59392 (1 KiB aligned)
constexpr int kOffXs = kOffDA + 2 * DA_BYTES;
constexpr int kOffBar = kOffXs + 2 * XS_FLOATS * 4;
constexpr int kSmem = 1024 + kOffBar + 128;
}  // namespace g1

__global__ void __launch_bounds__(g1::kThreads, 1)
conv1_wgrad_tc_kernel(const float* __restrict__ x, const bf16* __restrict__ dP, const bf16* __restrict__ P,
                      const uint8_t* __restrict__ idx, float* __restrict__ dW, float* __restrict__ db, int B,
                      const long long* win_step, long long win_rows) {
  using namespace g1;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sX5 = smem + kOffX5;
  uint8_t* sDA = smem + kOffDA;
  float* sXs = reinterpret_cast<float*>(smem + kOffXs);
  uint64_t* x_full = reinterpret_cast<uint64_t*>(smem + kOffBar);
  uint64_t* x_empty = x_full + 2;
  uint64_t* a_full = x_empty + 2;
  uint32_t* holder = reinterpret_cast<uint32_t*>(a_full + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) TRACE(30, 0);
  // X5 tails and the padding columns (w >= 28) of dA must read as zero for the whole kernel
  for (int i = threadIdx.x; i < (2 * X5_BYTES + 2 * DA_BYTES) / 16; i += kThreads)
    reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) { mbar_init(&x_full[s], kProducers); mbar_init(&x_empty[s], 1); }
    mbar_init(a_full, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc<32>(holder);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *holder;
  if (threadIdx.x == 0) TRACE(30, 1);
  mtb::pdl_sync();
  if (threadIdx.x == 0) TRACE(30, 2);
  x += window_offset(win_step, win_rows, B) * (IMG * IMG);

  if (warp == 0) {
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc(kFmtBF16, 128, C1, true, true);
      int it = 0;
      for (int img = blockIdx.x; img < B; img += gridDim.x, ++it) {
        const int s = it & 1, ph = (it >> 1) & 1;
        TRACE(16 + it, 0);
        mbar_wait(&x_full[s], ph);
        TRACE(16 + it, 1);
        tc_fence_after();
        // A: 8 atoms of 16 (kh) one kw = 4 rows = 128 bytes apart; B: one 32-channel atom
        const uint64_t a0 = make_smem_desc(smem_u32(sX5 + s * X5_BYTES), 128, 256, kSwizzle32B);
        const uint64_t b0 = make_smem_desc(smem_u32(sDA + s * DA_BYTES), 64, 512, kSwizzle64B);
        if (it == 0) mma_f16_first(tmem, a0, b0, idesc);
        else mma_f16_acc(tmem, a0, b0, idesc);
#pragma unroll 8
        for (int ks = 1; ks < 56; ++ks)
          mma_f16_acc(tmem, a0 + (uint64_t)((ks * 16 * 32) >> 4), b0 + (uint64_t)((ks * 16 * 64) >> 4), idesc);
        mma_commit(&x_empty[s]);
        TRACE(16 + it, 2);
      }
      mma_commit(a_full);
    }
    __syncwarp();
  } else if (warp <= 16) {
    const int g = (warp - 1) >> 3;
    const int t = threadIdx.x - 32 - g * kProducers;
    float* xs = sXs + g * XS_FLOATS;
    uint8_t* da = sDA + g * DA_BYTES;
    int it = g;
    for (int img = blockIdx.x + g * gridDim.x; img < B; img += 2 * gridDim.x, it += 2) {
      const int ph = (it >> 1) & 1;
      if (t == 0) TRACE(16 + it, 4);
      // all global loads of this image first: pooled gradient, pooled activation, argmax (16+16+8 bytes
      // per 8 channels of a pooled pixel), then the padded input image
      constexpr int PER = (196 * 4 + kProducers - 1) / kProducers;      // 4 (the last one only for t < 16)
      uint4 gq[PER], aq[PER];
      uint2 kq[PER];
#pragma unroll
      for (int k = 0; k < PER; ++k) {
        const int item = t + k * kProducers;
        if (item < 196 * 4) {
          const size_t o = ((size_t)img * 196 + (item >> 2)) * C1 + (item & 3) * 8;
          gq[k] = __ldg(reinterpret_cast<const uint4*>(dP + o));
          aq[k] = __ldg(reinterpret_cast<const uint4*>(P + o));
          kq[k] = __ldg(reinterpret_cast<const uint2*>(idx + o));
        }
      }
      load_xs<kProducers>(xs, x + (size_t)img * IMG * IMG, t);   // xs is this group's own: free since the last bar
      mbar_wait(&x_empty[g], ph ^ 1);
      if (t == 0) TRACE(16 + it, 5);
      if (t == 0) TRACE(16 + it, 6);
      // dA rows of one pooling window: gradient goes to the argmax position if the pooled
      // activation is positive (ReLU), the other three positions get zero
#pragma unroll
      for (int k = 0; k < PER; ++k) {
        const int item = t + k * kProducers;
        if (item < 196 * 4) {
          const int pp = item >> 2, chunk = item & 3;
          const int ph2 = pp / 14, pw = pp % 14;
          uint32_t gm[4];                                    // gradient with the ReLU mask applied, bf16x2
          mtb::relu_mask_bf16x8(gq[k], aq[k], gm);
#pragma unroll
          for (int pos = 0; pos < 4; ++pos) {
            const uint4 v = mtb::route_to_pos_bf16x8(gm, kq[k], (uint32_t)pos);
            const int h = 2 * ph2 + (pos >> 1), wc = 2 * pw + (pos & 1);
            const int R = (h >> 2) * 128 + wc * 4 + (h & 3);
            *reinterpret_cast<uint4*>(da + R * 64 + ((chunk ^ ((R >> 1) & 3)) << 4)) = v;
          }
        }
      }
      if (t == 0) TRACE(16 + it, 7);
      named_bar(2 + g, kProducers);                          // padded image complete
      build_x5(sX5 + g * X5_BYTES, xs, t, kProducers);
      fence_proxy_async_smem();
      mbar_arrive(&x_full[g]);
      if (t == 0) TRACE(16 + it, 8);
      named_bar(2 + g, kProducers);                          // everyone is done reading xs
    }
  } else {
    const int q = warp & 3;
    const int L = q * 32 + lane;
    const int kw = L >> 4, kh = L & 15;
    if (blockIdx.x < B) {
      mbar_wait(a_full, 0);
      if (L == 0) TRACE(30, 3);
      tc_fence_after();
      uint32_t r[32];
      tmem_ld_32x32(tmem + ((uint32_t)(q * 32) << 16), r);
      tmem_ld_wait();
      float* dst = (kw < 5 && kh < 5) ? dW + (kh * 5 + kw) * C1 : ((kw == 0 && kh == 5) ? db : nullptr);
      if (dst != nullptr) {
#pragma unroll
        for (int c = 0; c < 32; c += 4)
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + c), "f"(__uint_as_float(r[c])),
                       "f"(__uint_as_float(r[c + 1])), "f"(__uint_as_float(r[c + 2])), "f"(__uint_as_float(r[c + 3]))
                       : "memory");
      }
      tc_fence_before();
      if (L == 0) TRACE(30, 4);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) TRACE(30, 5);
  if (warp == 0) { tc_fence_after(); tmem_dealloc<32>(tmem); }
}

// The opt-in to more than 48 KB of dynamic shared memory is per device: `done` is indexed by device ordinal.
template <typename K>
int set_smem(K kern, int bytes, bool* done) {
  int dev = 0;
  MTB_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !done[dev]) {
    MTB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    if (dev >= 0 && dev < 64) done[dev] = true;
  }
  return 0;
}

}  // namespace

extern "C" {

#ifdef MTB_TRACE
int mtb_debug_trace_read(long long* host_out) {
  MTB_CUDA_CHECK(cudaMemcpyFromSymbol(host_out, g_trace, sizeof(long long) * 64 * 16));
  return 0;
}
#endif

// Tensor-core versions of mtb_conv1_fwd / mtb_conv1_wgrad (same arguments and output formats), C1 = 32.
int mtb_conv1_fwd_tc(const float* x, const float* w, const float* bias, void* out_bf16, uint8_t* idx, int B,
                     cudaStream_t stream) {
  return mtb_conv1_fwd_tc_win(x, w, bias, out_bf16, idx, B, nullptr, 0, stream);
}

// As above, but x is a resident shard of win_rows images and the batch is the reference's
// sequential window for the step stored at *win_step (win_step == NULL: x is the batch itself).
int mtb_conv1_fwd_tc_win(const float* x, const float* w, const float* bias, void* out_bf16, uint8_t* idx, int B,
                         const long long* win_step, long long win_rows, cudaStream_t stream) {
  return mtb_conv1_fwd_tc_win_zero(x, w, bias, out_bf16, idx, B, win_step, win_rows, nullptr, nullptr, 0, stream);
}

int mtb_conv1_fwd_tc_win_zero(const float* x, const float* w, const float* bias, void* out_bf16, uint8_t* idx, int B,
                              const long long* win_step, long long win_rows, void* const* zero_ptrs,
                              const long long* zero_words, int zero_count, cudaStream_t stream) {
  MTB_REQUIRE(x && w && bias && out_bf16 && idx && B > 0, "conv1_fwd_tc: bad argument");
  MTB_REQUIRE(zero_count >= 0 && zero_count <= 8, "conv1_fwd_tc: at most 8 ranges to clear");
  ZeroList zl;
  for (int k = 0; k < 8; ++k) { zl.ptr[k] = k < zero_count ? zero_ptrs[k] : nullptr; zl.n_words[k] = k < zero_count ? zero_words[k] : 0; }
  zl.count = zero_count;
  MTB_REQUIRE(win_step == nullptr || win_rows > B, "conv1_fwd_tc: shard must be larger than the batch");
  static bool attr[64] = {};
  if (int e = set_smem(conv1_fwd_tc_kernel, f1::kSmem, attr)) return e;
  const int grid = B < mtb::kNumSMs ? B : mtb::kNumSMs;
  MTB_CUDA_CHECK(mtb::launch(conv1_fwd_tc_kernel, dim3(grid), dim3(f1::kThreads), (size_t)f1::kSmem, stream, x, w, bias,
                             reinterpret_cast<bf16*>(out_bf16), idx, B, win_step, win_rows, zl));
  return 0;
}

// dW and db are ACCUMULATED into (caller zeroes them once per step).
int mtb_conv1_wgrad_tc(const float* x, const void* dP_bf16, const void* P_bf16, const uint8_t* idx, float* dW, float* db,
                       int B, cudaStream_t stream) {
  return mtb_conv1_wgrad_tc_win(x, dP_bf16, P_bf16, idx, dW, db, B, nullptr, 0, stream);
}

int mtb_conv1_wgrad_tc_win(const float* x, const void* dP_bf16, const void* P_bf16, const uint8_t* idx, float* dW,
                           float* db, int B, const long long* win_step, long long win_rows, cudaStream_t stream) {
  MTB_REQUIRE(x && dP_bf16 && P_bf16 && idx && dW && db && B > 0, "conv1_wgrad_tc: bad argument");
  MTB_REQUIRE(win_step == nullptr || win_rows > B, "conv1_wgrad_tc: shard must be larger than the batch");
  static bool attr[64] = {};
  if (int e = set_smem(conv1_wgrad_tc_kernel, g1::kSmem, attr)) return e;
  const int grid = B < mtb::kNumSMs ? B : mtb::kNumSMs;
  MTB_CUDA_CHECK(mtb::launch(conv1_wgrad_tc_kernel, dim3(grid), dim3(g1::kThreads), (size_t)g1::kSmem, stream, x,
                             reinterpret_cast<const bf16*>(dP_bf16), reinterpret_cast<const bf16*>(P_bf16), idx, dW, db,
                             B, win_step, win_rows));
  return 0;
}

}  // extern "C"
