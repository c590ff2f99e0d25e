// Second convolution layer (5x5, SAME, 32 -> 64 channels on 14x14 maps; mpipy.py:159-161 and the
// Conv2DBackpropInput / Conv2DBackpropFilter / MaxPoolGrad ops autodiff adds)
// as implicit GEMM on tcgen05 with NO im2col anywhere, not even in shared memory.
//
// "Shift-GEMM": TMA drops one zero-padded 18x18 halo patch of an image into shared memory
// (out-of-bounds box coordinates supply the SAME padding), pixels as rows, channels contiguous,
// hardware-swizzled.  If output pixel (h,w) is GEMM row m = 18h + w, then filter tap (kh,kw)
// reads patch row m + 18kh + kw: every tap is the SAME patch viewed from a start address moved
// by whole rows.  tcgen05 shared-memory descriptors swizzle on absolute address bits (pinned by
// csrc/debug_shift.cu and its test), so the 25 taps are 25 descriptor start addresses and the
// patch crosses L2 -> SM once instead of 25 times.  4 of every 18 GEMM rows are halo columns
// whose results are dropped: 1.3x spare MMA work, which is not the bottleneck.
//
//   fprop   D[pixel, co]   = sum_tap patch_x [pixel + shift(tap)] . W[tap]      A K-major SW64, B MN-major SW128
//           epilogue: + bias, ReLU, 2x2 max-pool + argmax through a shared staging tile;
//           the un-pooled activation never exists in global memory.
//   dgrad   D[pixel, ci]   = sum_tap patch_dy[pixel + shift(tap)] . W[24-tap]^T A K-major SW128, B K-major SW128
//   wgrad   D[co, (kw,ci)] = sum_pixel dy[pixel] . patch_x[pixel + 18kh + kw]   both operands MN-major, K = pixels.
//           One MMA covers all five kw taps: the five B "atoms" are the patch one row apart
//           (leading-dimension byte offset = one 64-byte row).  Two kh taps share one MMA: the
//           A operand's second 64-lane atom is dy itself, the first is dy moved back by 36 rows,
//           i.e. tap kh+2.  Three N=160 MMAs per 16 pixels give all 25 taps in 480 TMEM columns.
//           Per-CTA partial sums go to a workspace and a second kernel adds them in CTA order
//           (deterministic); the bias gradient is summed from the same shared tile.
//
// All three are persistent, warp-specialised kernels (TMA warp, MMA warp, 4 epilogue warps) with
// mbarrier rings, one CTA per SM, weights resident in shared memory for the CTA's lifetime.
#include <map>
#include <mutex>
#include <tuple>

#include "common.cuh"
#include "sm100.cuh"
#include "tma_host.cuh"

namespace {

using namespace sm100;
using bf16 = __nv_bfloat16;

constexpr int HW = 14;                 // feature map is 14 x 14
constexpr int PWID = 18;               // padded width = 14 + 2*2
constexpr int CIN = 32, COUT = 64, TAPS = 25;
constexpr int kThreads = 192;
constexpr int M_ROWS = HW * PWID;      // 252 GEMM rows per image (incl. 4 halo columns per row)
constexpr int PATCH_ROWS = PWID * PWID;  // 324
constexpr int W_BYTES = TAPS * CIN * COUT * 2;      // 102400: every tap's [32 ci][64 co] tile

__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 128;" ::: "memory"); }
__device__ __forceinline__ void epi_bar256() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// ======================================================================================= fprop
namespace fp {
constexpr int kStages = 3;
constexpr int kPatchBytes = 336 * 64;                      // 324 rows land, slack keeps tile 1's taps in bounds
constexpr int kOffW = 0;
constexpr int kOffPatch = W_BYTES;
constexpr int kOffStage = kOffPatch + kStages * kPatchBytes;   // bf16 [256][64] relu'd tile for pooling
constexpr int kOffBar = kOffStage + 256 * 128;
constexpr int kSmem = 1024 + kOffBar + 128 + 256;
constexpr int kThreadsFp = 64 + 256;                       // TMA warp, MMA warp, 8 epilogue warps
}  // namespace fp

__global__ void __launch_bounds__(fp::kThreadsFp, 1)
conv2_fprop_pool_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                        const float* __restrict__ bias, bf16* __restrict__ out, uint8_t* __restrict__ idx, int B) {
  using namespace fp;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sW = smem + kOffW;
  uint8_t* sPatch = smem + kOffPatch;
  uint8_t* sStage = smem + kOffStage;
  uint64_t* w_full = reinterpret_cast<uint64_t*>(smem + kOffBar);
  uint64_t* p_full = w_full + 1;
  uint64_t* p_empty = p_full + kStages;
  uint64_t* a_full = p_empty + kStages;
  uint64_t* a_empty = a_full + 2;
  uint32_t* holder = reinterpret_cast<uint32_t*>(a_empty + 2);
  float* sBias = reinterpret_cast<float*>(smem + kOffBar + 128);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&map_x);
    tma_prefetch_desc(&map_w);
    mbar_init(w_full, 1);
    for (int s = 0; s < kStages; ++s) { mbar_init(&p_full[s], 1); mbar_init(&p_empty[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&a_full[a], 1); mbar_init(&a_empty[a], 8); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<256>(holder);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *holder;
  mtb::pdl_sync();
  if (warp >= 2 && threadIdx.x - 64 < COUT) sBias[threadIdx.x - 64] = __ldg(bias + threadIdx.x - 64);
  if (warp >= 2) epi_bar256();                       // bias visible to all epilogue warps

  if (warp == 0) {
    if (lane == 0) {
      mbar_arrive_expect_tx(w_full, W_BYTES);
      for (int t = 0; t < TAPS; ++t) tma_load_2d(sW + t * 4096, &map_w, w_full, 0, t * CIN);
      int it = 0;
      for (int img = blockIdx.x; img < B; img += gridDim.x, ++it) {
        const int s = it % kStages, ph = (it / kStages) & 1;
        mbar_wait(&p_empty[s], ph ^ 1);
        mbar_arrive_expect_tx(&p_full[s], PATCH_ROWS * 64);
        tma_load_4d(sPatch + s * kPatchBytes, &map_x, &p_full[s], 0, -2, -2, img);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc(kFmtBF16, 128, COUT, false, true);
      mbar_wait(w_full, 0);
      int it = 0;
      for (int img = blockIdx.x; img < B; img += gridDim.x, ++it) {
        const int s = it % kStages, ph = (it / kStages) & 1;
        const int a = it & 1, aph = (it >> 1) & 1;
        mbar_wait(&p_full[s], ph);
        mbar_wait(&a_empty[a], aph ^ 1);
        tc_fence_after();
        // descriptors differ only in their 14-bit start-address field: build one per operand,
        // then every MMA is a 64-bit add with a compile-time offset (the single issuing thread
        // is the critical path of this kernel)
        const uint64_t a0 = make_smem_desc(smem_u32(sPatch + s * kPatchBytes), 16, 512, kSwizzle64B);
        const uint64_t b0 = make_smem_desc(smem_u32(sW), 4096, 1024, kSwizzle128B);
#pragma unroll
        for (int tile = 0; tile < 2; ++tile) {
          const uint32_t d = tmem + a * 128 + tile * 64;
#pragma unroll
          for (int tap = 0; tap < TAPS; ++tap) {
            const int shift = (tap / 5) * PWID + (tap % 5);
#pragma unroll
            for (int k = 0; k < 2; ++k) {
              const uint64_t da = a0 + (uint64_t)(((tile * 128 + shift) * 64 + k * 32) >> 4);
              const uint64_t db = b0 + (uint64_t)((tap * 4096 + k * 2048) >> 4);
              if (tap == 0 && k == 0) mma_f16_first(d, da, db, idesc);
              else mma_f16_acc(d, da, db, idesc);
            }
          }
        }
        mma_commit(&p_empty[s]);
        mma_commit(&a_full[a]);
      }
    }
    __syncwarp();
  } else {
    // 8 warps: warp%4 picks the TMEM lane quadrant, (warp-2)/4 picks which of the two M tiles
    const int q = warp & 3;
    const int tile = (warp - 2) >> 2;
    const int t128 = q * 32 + lane;                 // 0..127: TMEM lane = row within a tile
    const int t256 = tile * 128 + t128;
    int it = 0;
    for (int img = blockIdx.x; img < B; img += gridDim.x, ++it) {
      const int a = it & 1, aph = (it >> 1) & 1;
      mbar_wait(&a_full[a], aph);
      tc_fence_after();
      {
        const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16) + a * 128 + tile * 64;
        const int row = t256;
        uint8_t* dst = sStage + row * 128;
        uint32_t r0[32], r1[32];
        tmem_ld_32x32(taddr, r0);
        tmem_ld_32x32(taddr + 32, r1);
        tmem_ld_wait();
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          const uint32_t* r = c == 0 ? r0 : r1;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 b0 = *reinterpret_cast<const float4*>(sBias + c * 32 + 8 * j);
            const float4 b1 = *reinterpret_cast<const float4*>(sBias + c * 32 + 8 * j + 4);
            uint4 pk;
            pk.x = mtb::pack_bf16x2(fmaxf(__uint_as_float(r[8 * j + 0]) + b0.x, 0.f),
                                    fmaxf(__uint_as_float(r[8 * j + 1]) + b0.y, 0.f));
            pk.y = mtb::pack_bf16x2(fmaxf(__uint_as_float(r[8 * j + 2]) + b0.z, 0.f),
                                    fmaxf(__uint_as_float(r[8 * j + 3]) + b0.w, 0.f));
            pk.z = mtb::pack_bf16x2(fmaxf(__uint_as_float(r[8 * j + 4]) + b1.x, 0.f),
                                    fmaxf(__uint_as_float(r[8 * j + 5]) + b1.y, 0.f));
            pk.w = mtb::pack_bf16x2(fmaxf(__uint_as_float(r[8 * j + 6]) + b1.z, 0.f),
                                    fmaxf(__uint_as_float(r[8 * j + 7]) + b1.w, 0.f));
            const int chunk = c * 4 + j;
            *reinterpret_cast<uint4*>(dst + ((chunk ^ (row & 7)) << 4)) = pk;
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&a_empty[a]);
      epi_bar256();                                  // staging tile complete
      // 2x2 max-pool + argmax (first max wins, order (0,0),(0,1),(1,0),(1,1))
      for (int item = t256; item < 49 * 8; item += 256) {
        const int pp = item >> 3, chunk = item & 7;
        const int ph = pp / 7, pw = pp % 7;
        const int m00 = (2 * ph) * PWID + 2 * pw;
        float best[8];
        uint32_t arg[8];
#pragma unroll
        for (int pos = 0; pos < 4; ++pos) {
          const int row = m00 + (pos >> 1) * PWID + (pos & 1);
          const uint4 pk = *reinterpret_cast<const uint4*>(sStage + row * 128 + ((chunk ^ (row & 7)) << 4));
          const uint32_t w4[4] = {pk.x, pk.y, pk.z, pk.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 f = mtb::unpack_bf16x2(w4[e]);
            if (pos == 0 || f.x > best[2 * e]) { best[2 * e] = f.x; arg[2 * e] = pos; }
            if (pos == 0 || f.y > best[2 * e + 1]) { best[2 * e + 1] = f.y; arg[2 * e + 1] = pos; }
          }
        }
        uint4 o;
        o.x = mtb::pack_bf16x2(best[0], best[1]); o.y = mtb::pack_bf16x2(best[2], best[3]);
        o.z = mtb::pack_bf16x2(best[4], best[5]); o.w = mtb::pack_bf16x2(best[6], best[7]);
        const size_t off = ((size_t)img * 49 + pp) * COUT + chunk * 8;
        *reinterpret_cast<uint4*>(out + off) = o;
        uint2 k2;
        k2.x = arg[0] | (arg[1] << 8) | (arg[2] << 16) | (arg[3] << 24);
        k2.y = arg[4] | (arg[5] << 8) | (arg[6] << 16) | (arg[7] << 24);
        *reinterpret_cast<uint2*>(idx + off) = k2;
      }
      epi_bar256();                                  // staging tile free for the next image
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc<256>(tmem); }
}

// The un-pooled gradient of one image, rebuilt in shared memory from the pooled maps (MaxPoolGrad
// + ReluGrad, mpipy.py:160-161 backward): dflat / P / idx are [B,7,7,64]; the pooled gradient goes
// to its argmax position if the pooled activation is positive, the other three positions of the
// window get zero.  Pixel (h,w) lands in row base_row + 18h + w of a 128-byte-row SW128 tile whose
// row 0 is 1 KiB aligned.  Replaces a separate kernel and a 25 MB round trip through HBM per step.
struct PooledGrad { const bf16* dflat; const bf16* P; const uint8_t* idx; };

template <int NT>
struct DyItems {                                          // one image's share of a builder thread, in registers
  static constexpr int PER = (49 * 8 + NT - 1) / NT;
  uint4 g[PER], a[PER];
  uint2 k[PER];
  __device__ __forceinline__ void load(const PooledGrad& pg, int img, int t) {
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      const int item = t + j * NT;
      if (item < 49 * 8) {
        const size_t o = ((size_t)img * 49 + (item >> 3)) * COUT + (item & 7) * 8;
        g[j] = __ldg(reinterpret_cast<const uint4*>(pg.dflat + o));
        a[j] = __ldg(reinterpret_cast<const uint4*>(pg.P + o));
        k[j] = __ldg(reinterpret_cast<const uint2*>(pg.idx + o));
      }
    }
  }
  __device__ __forceinline__ void store(uint8_t* tile, int base_row, int t) const {
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      const int item = t + j * NT;
      if (item < 49 * 8) {
        const int pp = item >> 3, chunk = item & 7;
        const int ph = pp / 7, pw = pp % 7;
        uint32_t gm[4];
        mtb::relu_mask_bf16x8(g[j], a[j], gm);
#pragma unroll
        for (int pos = 0; pos < 4; ++pos) {
          const uint4 v = mtb::route_to_pos_bf16x8(gm, k[j], (uint32_t)pos);
          const int row = base_row + (2 * ph + (pos >> 1)) * PWID + 2 * pw + (pos & 1);
          *reinterpret_cast<uint4*>(tile + row * 128 + ((chunk ^ (row & 7)) << 4)) = v;
        }
      }
    }
  }
};

constexpr int kBuilders = 256;                          // extra warps that build the dy tile in the fused variants

// ======================================================================================= dgrad
namespace dg {
constexpr int kStages = 2;
constexpr int kPatchBytes = 336 * 128;
constexpr int kOffW = 0;
constexpr int kOffPatch = W_BYTES;
constexpr int kOffBnd = kOffPatch + kStages * kPatchBytes;   // fp32 [5 sources][4 kw][4 lanes][32 ci] boundary rows
constexpr int kBndFloats = 4 * 4 * 32;                       // per source (4 warps of the tile + the carry from tile 1)
constexpr int kOffBar = kOffBnd + 5 * kBndFloats * 4;
constexpr int kSmem = 1024 + kOffBar + 128;
}  // namespace dg

// dgrad with the horizontal taps folded into the MMA's N dimension:
//   E[r, (kw,ci)] = sum_kh sum_co patch[r + 18 kh][co] . W'[kh][kw][ci][co]         20 MMAs (128 x 160 x 16) per tile
//   dx[m, ci]     = sum_kw E[m + kw, (kw,ci)]
// instead of 100 MMAs with N = 32: shared-memory operand traffic per FLOP drops ~2.4x, which is what
// bounds these kernels.  The row shift of the final sum is a lane shift: shuffles inside a warp, a
// small shared-memory hand-over for the last four lanes of each warp (tile 1 is issued first so
// that its leading rows are there when tile 0's last warp needs them).
constexpr int kThreadsD = 64 + 256;                      // dgrad: TMA warp, MMA warp, 8 epilogue warps
constexpr int kBuildersD = 128;

template <bool kFused>
__global__ void __launch_bounds__(kThreadsD + (kFused ? kBuildersD : 0), 1)
conv2_dgrad_kernel(const __grid_constant__ CUtensorMap map_dy, const __grid_constant__ CUtensorMap map_w,
                   bf16* __restrict__ dx, int B, const PooledGrad pg) {
  using namespace dg;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sW = smem + kOffW;
  uint8_t* sPatch = smem + kOffPatch;
  float* sBnd = reinterpret_cast<float*>(smem + kOffBnd);
  uint64_t* w_full = reinterpret_cast<uint64_t*>(smem + kOffBar);
  uint64_t* p_full = w_full + 1;
  uint64_t* p_empty = p_full + kStages;
  uint64_t* a_full = p_empty + kStages;         // one per accumulator buffer (a buffer holds ONE tile, 160 columns)
  uint64_t* a_empty = a_full + 2;
  uint32_t* holder = reinterpret_cast<uint32_t*>(a_empty + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&map_dy);
    tma_prefetch_desc(&map_w);
    mbar_init(w_full, 1);
    for (int s = 0; s < kStages; ++s) { mbar_init(&p_full[s], kFused ? kBuildersD : 1); mbar_init(&p_empty[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&a_full[a], 1); mbar_init(&a_empty[a], 8); }
    fence_barrier_init();
  }
  if (kFused) {       // halo rows and columns of both patch stages stay zero for the kernel's lifetime
    for (int i = threadIdx.x; i < kStages * kPatchBytes / 16; i += blockDim.x)
      reinterpret_cast<uint4*>(sPatch)[i] = make_uint4(0, 0, 0, 0);
    fence_proxy_async_smem();
  }
  if (warp == 1) tmem_alloc<512>(holder);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *holder;
  mtb::pdl_sync();

  if (warp == 0) {
    if (lane == 0) {
      // tap t goes to slot 24 - t: slot s then holds the FLIPPED tap s, and the five kw of one kh are
      // 160 consecutive K-major rows (one N = 160 operand)
      mbar_arrive_expect_tx(w_full, W_BYTES);
      for (int t = 0; t < TAPS; ++t) tma_load_2d(sW + (TAPS - 1 - t) * 4096, &map_w, w_full, 0, t * CIN);
      int it = 0;
      for (int img = blockIdx.x; !kFused && img < B; img += gridDim.x, ++it) {
        const int s = it % kStages, ph = (it / kStages) & 1;
        mbar_wait(&p_empty[s], ph ^ 1);
        mbar_arrive_expect_tx(&p_full[s], PATCH_ROWS * 128);
        tma_load_4d(sPatch + s * kPatchBytes, &map_dy, &p_full[s], 0, -2, -2, img);
      }
    }
    __syncwarp();
  } else if (kFused && warp >= kThreadsD / 32) {
    // software pipeline: the next image's global loads are in flight while this one is written
    const int t = threadIdx.x - kThreadsD;
    DyItems<kBuildersD> cur, nxt;
    if ((int)blockIdx.x < B) nxt.load(pg, blockIdx.x, t);
    int it = 0;
    for (int img = blockIdx.x; img < B; img += gridDim.x, ++it) {
      const int s = it % kStages, ph = (it / kStages) & 1;
      cur = nxt;
      if (img + (int)gridDim.x < B) nxt.load(pg, img + gridDim.x, t);
      mbar_wait(&p_empty[s], ph ^ 1);
      cur.store(sPatch + s * kPatchBytes, 2 * PWID + 2, t);
      fence_proxy_async_smem();
      mbar_arrive(&p_full[s]);
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc(kFmtBF16, 128, 5 * CIN, false, false);
      mbar_wait(w_full, 0);
      int it = 0, n = 0;                               // n counts tiles: buffer n & 1
      for (int img = blockIdx.x; img < B; img += gridDim.x, ++it) {
        const int s = it % kStages, ph = (it / kStages) & 1;
        mbar_wait(&p_full[s], ph);
        const uint64_t a0 = make_smem_desc(smem_u32(sPatch + s * kPatchBytes), 16, 1024, kSwizzle128B);
        const uint64_t b0 = make_smem_desc(smem_u32(sW), 16, 1024, kSwizzle128B);
#pragma unroll 1
        for (int tt = 0; tt < 2; ++tt, ++n) {          // tile 1 first, then tile 0
          const int tile = 1 - tt;
          const int a = n & 1, aph = (n >> 1) & 1;
          mbar_wait(&a_empty[a], aph ^ 1);
          tc_fence_after();
          const uint32_t d = tmem + a * 256;
#pragma unroll
          for (int kh = 0; kh < 5; ++kh) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint64_t da = a0 + (uint64_t)(((tile * 128 + kh * PWID) * 128 + k * 32) >> 4);
              const uint64_t db = b0 + (uint64_t)((kh * 5 * 4096 + k * 32) >> 4);
              if (kh == 0 && k == 0) mma_f16_first(d, da, db, idesc);
              else mma_f16_acc(d, da, db, idesc);
            }
          }
          if (tt == 1) mma_commit(&p_empty[s]);
          mma_commit(&a_full[a]);
        }
      }
    }
    __syncwarp();
  } else {
    // 8 epilogue warps: warp % 4 = TMEM lane quadrant, (warp - 2) / 4 = which 16 of the 32 channels
    const int q = warp & 3;
    const int hc = (warp - 2) >> 2;
    const int t128 = q * 32 + lane;
    float* my_bnd = sBnd + q * kBndFloats;             // what the warp below me needs: my first four rows
    const float* next_bnd = sBnd + (q + 1) * kBndFloats;   // q == 3: slot 4 = carry from tile 1
    int n = 0;
    for (int img = blockIdx.x; img < B; img += gridDim.x) {
#pragma unroll 1
      for (int tt = 0; tt < 2; ++tt, ++n) {
        const int tile = 1 - tt;
        const int a = n & 1, aph = (n >> 1) & 1;
        mbar_wait(&a_full[a], aph);
        tc_fence_after();
        const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16) + a * 256 + hc * 16;
        float acc[16];
        {
          uint32_t r[16];
          tmem_ld_32x16(taddr, r);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) acc[j] = __uint_as_float(r[j]);
        }
        uint32_t r[4][16];
#pragma unroll
        for (int kw = 1; kw < 5; ++kw) tmem_ld_32x16(taddr + kw * 32, r[kw - 1]);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&a_empty[a]);
#pragma unroll
        for (int kw = 1; kw < 5; ++kw) {
          // rows 0..3 of this warp are the rows 32..35 of the warp above it in the tile (for tile 1, warp 0:
          // of tile 0's last warp, through the carry slot)
          if (lane < 4) {
            float* dst = ((tile == 1 && q == 0) ? sBnd + 4 * kBndFloats : my_bnd) + ((kw - 1) * 4 + lane) * 32 + hc * 16;
#pragma unroll
            for (int j = 0; j < 16; j += 4)
              *reinterpret_cast<float4*>(dst + j) =
                  make_float4(__uint_as_float(r[kw - 1][j]), __uint_as_float(r[kw - 1][j + 1]),
                              __uint_as_float(r[kw - 1][j + 2]), __uint_as_float(r[kw - 1][j + 3]));
          }
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float v = __shfl_down_sync(0xffffffffu, __uint_as_float(r[kw - 1][j]), kw);
            if (lane + kw < 32) acc[j] += v;
          }
        }
        epi_bar256();                                  // boundary rows of all warps (and the carry) are in place
        if (lane >= 28 && !(tile == 1 && q == 3)) {
          // q == 3 of tile 0 continues into tile 1 (carry slot); tile 1's last rows are beyond the image
          const float* src = (q == 3) ? sBnd + 4 * kBndFloats : next_bnd;
#pragma unroll
          for (int kw = 1; kw < 5; ++kw) {
            const int l2 = lane + kw - 32;
            if (l2 >= 0) {
              const float* p = src + ((kw - 1) * 4 + l2) * 32 + hc * 16;
#pragma unroll
              for (int j = 0; j < 16; j += 4) {
                const float4 v = *reinterpret_cast<const float4*>(p + j);
                acc[j] += v.x; acc[j + 1] += v.y; acc[j + 2] += v.z; acc[j + 3] += v.w;
              }
            }
          }
        }
        const int m = tile * 128 + t128;
        const int h = m / PWID, w = m % PWID;
        if (h < HW && w < HW) {
          uint4* dst = reinterpret_cast<uint4*>(dx + (((size_t)img * HW + h) * HW + w) * CIN + hc * 16);
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            uint4 pk;
            pk.x = mtb::pack_bf16x2(acc[8 * j + 0], acc[8 * j + 1]);
            pk.y = mtb::pack_bf16x2(acc[8 * j + 2], acc[8 * j + 3]);
            pk.z = mtb::pack_bf16x2(acc[8 * j + 4], acc[8 * j + 5]);
            pk.w = mtb::pack_bf16x2(acc[8 * j + 6], acc[8 * j + 7]);
            dst[j] = pk;
          }
        }
        epi_bar256();                                  // boundary buffers free for the next tile
      }
    }
  }
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc<512>(tmem); }
}

// ======================================================================================= wgrad
namespace wg {
constexpr int kStages = 3;
constexpr int kLead = 40;                                  // zero rows in front of dy (36 needed; 40 keeps 1 KiB alignment)
constexpr int kDyRows = kLead + M_ROWS + 36;               // 328
constexpr int kDyBytes = kDyRows * 128;                    // 41984
constexpr int kXRows = 368;                                // 324 land; zero slack up to row 287 + 76
constexpr int kXBytes = kXRows * 64;                       // 23552
constexpr int kStageBytes = kDyBytes + kXBytes;            // 65536
constexpr int kOffBar = kStages * kStageBytes;
constexpr int kOffRed = kOffBar + 128;                     // fp32 [16][64] bias partials
constexpr int kSmem = 1024 + kOffRed + 16 * 64 * 4;
constexpr int kKSteps = 18;                                // 288 pixel rows / 16
constexpr int kCols = 160;                                 // one N=160 MMA per CTA (see the group split below)
constexpr int kPartialFloats = kCols * 128;
constexpr int kMaxPerGroup = 49;                           // 3 groups x 49 CTAs = 147 SMs
}  // namespace wg

template <bool kFused>
__global__ void __launch_bounds__(kThreads + (kFused ? kBuilders : 0), 1)
conv2_wgrad_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_dy,
                   float* __restrict__ partial, float* __restrict__ db, int B, const PooledGrad pg) {
  using namespace wg;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* p_full = reinterpret_cast<uint64_t*>(smem + kOffBar);
  uint64_t* p_empty = p_full + kStages;
  uint64_t* a_full = p_empty + kStages;
  uint32_t* holder = reinterpret_cast<uint32_t*>(a_full + 1);
  float* sRed = reinterpret_cast<float*>(smem + kOffRed);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // guard rows must read as zero (0 * garbage could be NaN): clear everything once, TMA only ever
  // rewrites the data rows
  for (int i = threadIdx.x; i < kStages * kStageBytes / 16; i += blockDim.x)
    reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&map_x);
    tma_prefetch_desc(&map_dy);
    for (int s = 0; s < kStages; ++s) { mbar_init(&p_full[s], kFused ? 1 + kBuilders : 1); mbar_init(&p_empty[s], 5); }
    mbar_init(a_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<256>(holder);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *holder;
  mtb::pdl_sync();
  // CTA group j = blockIdx % 3 owns filter rows kh_j (TMEM lanes 64..127) and kh_j + 2 (lanes 0..63),
  // kh_j = 0, 1, 4, over a third of the CTAs each: three times fewer per-CTA partial sums to
  // write and reduce than letting every CTA hold all 25 taps.
  const int grp = blockIdx.x % 3, g_idx = blockIdx.x / 3, g_cnt = gridDim.x / 3;
  const int kh_j = grp == 2 ? 4 : grp;
  const int n_img = g_idx < B ? (B - 1 - g_idx) / g_cnt + 1 : 0;

  if (warp == 0) {
    if (lane == 0) {
      int it = 0;
      for (int img = g_idx; img < B; img += g_cnt, ++it) {
        const int s = it % kStages, ph = (it / kStages) & 1;
        mbar_wait(&p_empty[s], ph ^ 1);
        uint8_t* st = smem + s * kStageBytes;
        if (kFused) {
          mbar_arrive_expect_tx(&p_full[s], PATCH_ROWS * 64);
        } else {
          mbar_arrive_expect_tx(&p_full[s], M_ROWS * 128 + PATCH_ROWS * 64);
          tma_load_4d(st + kLead * 128, &map_dy, &p_full[s], 0, 0, 0, img);
        }
        tma_load_4d(st + kDyBytes, &map_x, &p_full[s], 0, -2, -2, img);
      }
    }
    __syncwarp();
  } else if (kFused && warp >= kThreads / 32) {
    const int t = threadIdx.x - kThreads;
    DyItems<kBuilders> cur, nxt;
    if (g_idx < B) nxt.load(pg, g_idx, t);
    int it = 0;
    for (int img = g_idx; img < B; img += g_cnt, ++it) {
      const int s = it % kStages, ph = (it / kStages) & 1;
      cur = nxt;
      if (img + g_cnt < B) nxt.load(pg, img + g_cnt, t);
      mbar_wait(&p_empty[s], ph ^ 1);
      cur.store(smem + s * kStageBytes, kLead, t);
      fence_proxy_async_smem();
      mbar_arrive(&p_full[s]);
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc(kFmtBF16, 128, 160, true, true);
      int it = 0;
      for (int img = g_idx; img < B; img += g_cnt, ++it) {
        const int s = it % kStages, ph = (it / kStages) & 1;
        mbar_wait(&p_full[s], ph);
        tc_fence_after();
        // A starts at row p-36 of dy; B at row p + 18 kh of the x patch
        const uint64_t a0 = make_smem_desc(smem_u32(smem + s * kStageBytes) + (kLead - 36) * 128, 36 * 128, 1024,
                                           kSwizzle128B);
        const uint64_t b0 = make_smem_desc(smem_u32(smem + s * kStageBytes + kDyBytes) + kh_j * PWID * 64, 64, 512,
                                           kSwizzle64B);
        if (it == 0) mma_f16_first(tmem, a0, b0, idesc);
        else mma_f16_acc(tmem, a0, b0, idesc);
#pragma unroll
        for (int ks = 1; ks < kKSteps; ++ks)
          mma_f16_acc(tmem, a0 + (uint64_t)((ks * 2048) >> 4), b0 + (uint64_t)((ks * 16 * 64) >> 4), idesc);
        mma_commit(&p_empty[s]);
      }
      mma_commit(a_full);
    }
    __syncwarp();
  } else {
    const int q = warp & 3;
    const int t128 = q * 32 + lane;
    // ---- bias gradient: column sums of dy straight from the shared tile ----
    const int chunk = t128 & 7, sub = t128 >> 3;
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
    int it = 0;
    for (int img = g_idx; img < B; img += g_cnt, ++it) {
      const int s = it % kStages, ph = (it / kStages) & 1;
      mbar_wait(&p_full[s], ph);
      const uint8_t* dy = smem + s * kStageBytes + kLead * 128;
      for (int r = sub; grp == 0 && r < M_ROWS; r += 16) {
        const uint4 pk = *reinterpret_cast<const uint4*>(dy + r * 128 + ((chunk ^ (r & 7)) << 4));
        const uint32_t w4[4] = {pk.x, pk.y, pk.z, pk.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 f = mtb::unpack_bf16x2(w4[e]);
          acc[2 * e] += f.x;
          acc[2 * e + 1] += f.y;
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_empty[s]);
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) sRed[sub * 64 + chunk * 8 + e] = acc[e];
    epi_bar();
    if (t128 < 64 && n_img > 0 && grp == 0) {
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < 16; ++i) s += sRed[i * 64 + t128];
      atomicAdd(db + t128, s);
    }
    // ---- weight-gradient partial: TMEM -> workspace [col][lane] ----
    if (n_img > 0) {
      mbar_wait(a_full, 0);
      tc_fence_after();
      float* dst = partial + (size_t)blockIdx.x * kPartialFloats + t128;
#pragma unroll 1
      for (int c = 0; c < kCols / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(tmem + ((uint32_t)(q * 32) << 16) + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) dst[(size_t)(c * 32 + j) * 128] = __uint_as_float(r[j]);
      }
      tc_fence_before();
    }
  }
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc<256>(tmem); }
}

// dW[(kh*5+kw)*32+ci][co] = sum over the CTAs c of group j of partial[c][kw*32 + ci][lane], in CTA order
//   kh 0,1,4 -> group 0,1,2, lane 64+co   |   kh 2,3 -> group 0,1, lane co
__global__ void __launch_bounds__(256)
conv2_wgrad_reduce_kernel(const float* __restrict__ partial, float* __restrict__ dW, int n_cta) {
  mtb::pdl_sync();
  const int o = blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= TAPS * CIN * COUT) return;
  const int co = o % COUT, ci = (o / COUT) % CIN, tap = o / (COUT * CIN);
  const int kh = tap / 5, kw = tap % 5;
  int j, lane;
  if (kh == 2 || kh == 3) { j = kh - 2; lane = co; }
  else { j = kh == 4 ? 2 : kh; lane = 64 + co; }
  const float* src = partial + (size_t)j * wg::kPartialFloats + (size_t)(kw * 32 + ci) * 128 + lane;
  const size_t stride = (size_t)3 * wg::kPartialFloats;
  const int n = n_cta / 3;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  int c = 0;
  for (; c + 4 <= n; c += 4) {
    s0 += __ldcs(src + (size_t)(c + 0) * stride);
    s1 += __ldcs(src + (size_t)(c + 1) * stride);
    s2 += __ldcs(src + (size_t)(c + 2) * stride);
    s3 += __ldcs(src + (size_t)(c + 3) * stride);
  }
  for (; c < n; ++c) s0 += __ldcs(src + (size_t)c * stride);
  dW[o] = (s0 + s1) + (s2 + s3);
}

// ---------------------------------------------------------------------------- tensor maps
struct Key {
  const void* p; int B; int kind;
  bool operator<(const Key& o) const { return std::tie(p, B, kind) < std::tie(o.p, o.B, o.kind); }
};
std::map<Key, CUtensorMap> g_maps;
std::mutex g_mu;

// kind 0: x [B,14,14,32] box (32,18,18,1) SW64 | 1: dy [B,14,14,64] box (64,18,18,1) SW128
// kind 2: dy box (64,18,14,1) SW128            | 3: W [800,64] box (64,32) SW128
bool get_map(CUtensorMap* out, const void* p, int B, int kind) {
  std::lock_guard<std::mutex> lk(g_mu);
  Key k{p, B, kind};
  auto it = g_maps.find(k);
  if (it != g_maps.end()) { *out = it->second; return true; }
  CUtensorMap m;
  bool ok;
  if (kind == 3) {
    uint64_t dims[2] = {COUT, (uint64_t)TAPS * CIN};
    uint64_t str[1] = {COUT * 2};
    uint32_t box[2] = {COUT, CIN};
    ok = tma::make(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, p, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
  } else {
    const uint64_t C = kind == 0 ? CIN : COUT;
    uint64_t dims[4] = {C, HW, HW, (uint64_t)B};
    uint64_t str[3] = {C * 2, C * 2 * HW, C * 2 * HW * HW};
    uint32_t box[4] = {(uint32_t)C, PWID, kind == 2 ? (uint32_t)HW : (uint32_t)PWID, 1};
    ok = tma::make(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, p, dims, str, box,
                   kind == 0 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B);
  }
  if (!ok) return false;
  if (g_maps.size() > 1024) g_maps.clear();
  g_maps[k] = m;
  *out = m;
  return true;
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

inline int grid_for(int B) { return B < mtb::kNumSMs ? B : mtb::kNumSMs; }

}  // namespace

extern "C" {

int mtb_conv2_supported(int H, int cin, int cout) { return H == HW && cin == CIN && cout == COUT; }

int mtb_conv2_fprop_pool(const void* x_bf16, const void* w_bf16, const float* bias, void* out_bf16, uint8_t* idx, int B,
                         cudaStream_t stream) {
  MTB_REQUIRE(x_bf16 && w_bf16 && bias && out_bf16 && idx && B > 0, "conv2_fprop: bad argument");
  CUtensorMap mx, mw;
  MTB_REQUIRE(get_map(&mx, x_bf16, B, 0) && get_map(&mw, w_bf16, 0, 3), "conv2_fprop: tensor map encode failed");
  static bool attr[64] = {};
  if (int e = set_smem(conv2_fprop_pool_kernel, fp::kSmem, attr)) return e;
  MTB_CUDA_CHECK(mtb::launch(conv2_fprop_pool_kernel, dim3(grid_for(B)), dim3(fp::kThreadsFp), (size_t)fp::kSmem, stream,
                             mx, mw, bias, reinterpret_cast<bf16*>(out_bf16), idx, B));
  return 0;
}

int mtb_conv2_dgrad(const void* dy_bf16, const void* w_bf16, void* dx_bf16, int B, cudaStream_t stream) {
  MTB_REQUIRE(dy_bf16 && w_bf16 && dx_bf16 && B > 0, "conv2_dgrad: bad argument");
  CUtensorMap md, mw;
  MTB_REQUIRE(get_map(&md, dy_bf16, B, 1) && get_map(&mw, w_bf16, 0, 3), "conv2_dgrad: tensor map encode failed");
  static bool attr[64] = {};
  if (int e = set_smem(conv2_dgrad_kernel<false>, dg::kSmem, attr)) return e;
  MTB_CUDA_CHECK(mtb::launch(conv2_dgrad_kernel<false>, dim3(grid_for(B)), dim3(kThreadsD), (size_t)dg::kSmem, stream, md,
                             mw, reinterpret_cast<bf16*>(dx_bf16), B, PooledGrad{nullptr, nullptr, nullptr}));
  return 0;
}

// Same, with the MaxPoolGrad + ReluGrad in front of it fused in: takes the POOLED gradient dflat and the pooled
// activation / argmax of the forward pass ([B,7,7,64] each) instead of the un-pooled gradient map.
int mtb_conv2_dgrad_unpool(const void* dflat_bf16, const void* pooled_bf16, const uint8_t* idx, const void* w_bf16,
                           void* dx_bf16, int B, cudaStream_t stream) {
  MTB_REQUIRE(dflat_bf16 && pooled_bf16 && idx && w_bf16 && dx_bf16 && B > 0, "conv2_dgrad_unpool: bad argument");
  CUtensorMap mw;
  MTB_REQUIRE(get_map(&mw, w_bf16, 0, 3), "conv2_dgrad_unpool: tensor map encode failed");
  static bool attr[64] = {};
  if (int e = set_smem(conv2_dgrad_kernel<true>, dg::kSmem, attr)) return e;
  const PooledGrad pg{reinterpret_cast<const bf16*>(dflat_bf16), reinterpret_cast<const bf16*>(pooled_bf16), idx};
  MTB_CUDA_CHECK(mtb::launch(conv2_dgrad_kernel<true>, dim3(grid_for(B)), dim3(kThreadsD + kBuildersD), (size_t)dg::kSmem,
                             stream, mw, mw, reinterpret_cast<bf16*>(dx_bf16), B, pg));
  return 0;
}

static int wgrad_grid(int B) { return 3 * (B < wg::kMaxPerGroup ? B : wg::kMaxPerGroup); }
size_t mtb_conv2_wgrad_workspace_bytes(int B) { return (size_t)wgrad_grid(B) * wg::kPartialFloats * sizeof(float); }

// dW is overwritten; db is accumulated into (the caller zeroes it once per step).
int mtb_conv2_wgrad(const void* x_bf16, const void* dy_bf16, float* dW, float* db, void* workspace, int B,
                    cudaStream_t stream) {
  MTB_REQUIRE(x_bf16 && dy_bf16 && dW && db && workspace && B > 0, "conv2_wgrad: bad argument");
  CUtensorMap mx, md;
  MTB_REQUIRE(get_map(&mx, x_bf16, B, 0) && get_map(&md, dy_bf16, B, 2), "conv2_wgrad: tensor map encode failed");
  static bool attr[64] = {};
  if (int e = set_smem(conv2_wgrad_kernel<false>, wg::kSmem, attr)) return e;
  const int grid = wgrad_grid(B);
  MTB_CUDA_CHECK(mtb::launch(conv2_wgrad_kernel<false>, dim3(grid), dim3(kThreads), (size_t)wg::kSmem, stream, mx, md,
                             reinterpret_cast<float*>(workspace), db, B, PooledGrad{nullptr, nullptr, nullptr}));
  MTB_CUDA_CHECK(mtb::launch(conv2_wgrad_reduce_kernel, dim3((TAPS * CIN * COUT + 255) / 256), dim3(256), 0, stream,
                             reinterpret_cast<const float*>(workspace), dW, grid));
  return 0;
}

// Same, fed with the pooled gradient and the forward pass's pooled activation / argmax (see mtb_conv2_dgrad_unpool).
int mtb_conv2_wgrad_unpool(const void* x_bf16, const void* dflat_bf16, const void* pooled_bf16, const uint8_t* idx,
                           float* dW, float* db, void* workspace, int B, cudaStream_t stream) {
  MTB_REQUIRE(x_bf16 && dflat_bf16 && pooled_bf16 && idx && dW && db && workspace && B > 0,
              "conv2_wgrad_unpool: bad argument");
  CUtensorMap mx;
  MTB_REQUIRE(get_map(&mx, x_bf16, B, 0), "conv2_wgrad_unpool: tensor map encode failed");
  static bool attr[64] = {};
  if (int e = set_smem(conv2_wgrad_kernel<true>, wg::kSmem, attr)) return e;
  const int grid = wgrad_grid(B);
  const PooledGrad pg{reinterpret_cast<const bf16*>(dflat_bf16), reinterpret_cast<const bf16*>(pooled_bf16), idx};
  MTB_CUDA_CHECK(mtb::launch(conv2_wgrad_kernel<true>, dim3(grid), dim3(kThreads + kBuilders), (size_t)wg::kSmem, stream,
                             mx, mx, reinterpret_cast<float*>(workspace), db, B, pg));
  MTB_CUDA_CHECK(mtb::launch(conv2_wgrad_reduce_kernel, dim3((TAPS * CIN * COUT + 255) / 256), dim3(256), 0, stream,
                             reinterpret_cast<const float*>(workspace), dW, grid));
  return 0;
}

}  // extern "C"
