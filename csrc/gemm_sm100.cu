// tcgen05 GEMM engine: D[M,N] = op(A) * op(B) with bf16 operands and fp32 accumulation in
// tensor memory.  One kernel, warp-specialised:
//   warp 0     TMA producer   global -> 128B-swizzled shared tiles, multi-stage mbarrier ring
//   warp 1     MMA issuer     one elected thread issues tcgen05.mma (M=128, N=BN, K=16) x4 per stage
//   warps 2-5  epilogue       tcgen05.ld TMEM -> registers -> bias / ReLU / dropout / cast -> global
// Either operand may be K-major ("row = M/N index, K contiguous") or MN-major ("row = K index,
// M/N contiguous"); TMA lands both in the canonical SWIZZLE_128B layouts so no transposes are
// ever materialised.  That covers every FC product of the network with ONE bf16 weight copy:
//   fprop  Y = X * W          A K-major [B,in],  B MN-major W[in,out]      (mpipy.py:165,167)
//   dgrad  dX = dY * W^T      A K-major [B,out], B K-major  W[in,out]      (TF MatMul grad)
//   wgrad  dW = X^T * dY      A MN-major X[B,in], B MN-major dY[B,out]
// and, through im2col, the generic convolution path (mpipy.py:156,159).
// Split-K: each split writes its fp32 partial tile to a workspace; the last CTA to arrive at a
// per-tile counter sums the partials in split order (deterministic) and runs the epilogue.
#include <cstdlib>
#include <map>
#include <mutex>
#include <tuple>

#include "common.cuh"
#include "sm100.cuh"
#include "tma_host.cuh"

namespace {

using namespace sm100;

constexpr int BM = 128;
constexpr int BK = 64;                 // bf16 elements per stage along K = one 128-byte swizzle row
constexpr int kThreads = 192;

struct GemmParams {
  int M, N, K;
  int kb_total, kb_per_split, splits;
  void* out;
  long long ldc;
  int out_bf16;
  const float* bias;
  int relu;
  float keep_prob;
  unsigned long long seed;
  const long long* step_ptr;
  float* workspace;
  unsigned int* counters;
#ifdef MTB_GEMM_RS_TU
  // Reduce-scatter epilogue (gemm_rs_sm100.cu compiles this file a second time with MTB_GEMM_RS_TU): instead of storing
  // the tile, add it into the gradient buffer of the rank that owns each 16-byte vector (slice r = vectors
  // [r*rs_per/4, (r+1)*rs_per/4) of the flat buffer), over NVLink.
  long long rs_base, rs_per;           // flat element offset of out[0][0]; elements per owner slice (multiple of 4)
  float* rs_out[16];                   // every rank's flat gradient buffer (peer-mapped)
  int rs_staged;                       // 1: transpose each 32x32 chunk through smem so a warp adds whole 128-byte row segments
#endif
  // Implicit-GEMM 5x5 SAME convolution (CONV template parameter): the im2col matrix is never built; the producer
  // addresses the NHWC activation tensor directly through a 4-D tensor map (out-of-bounds box coordinates give the zero
  // padding).  Image rows are walked 16 columns wide (14 real + 2 zero-filled), so that a box is a rectangle of pixels.
  //   CONV 1 (fprop / dgrad): GEMM row = output pixel.  M tile = 8 image rows x 16 columns of one image (2 tiles per
  //           image), K block = one filter tap x one 128-byte channel block of the A tensor.  conv_sign -1 walks the taps
  //           mirrored (dx[p] = sum dy[p - off] W[tap]^T).  Rows are stored at their NHWC position; pad rows are dropped.
  //   CONV 2 (wgrad): GEMM K = pixels.  K block = conv_rpb image rows x 16 columns of one image; A column (tap, ci) is the
  //           activation tensor moved by the tap, B is the output-gradient tensor, both MN-major.
  int conv_c;                          // channels of the tensor behind A
  int conv_cpb;                        // CONV 1: K blocks per tap (= conv_c / elements per 128-byte row)
  int conv_sign;                       // CONV 1: +1 fprop, -1 dgrad
  int conv_rpb, conv_kpi;              // CONV 2: image rows per K block, K blocks per image
  int conv_brow;                       // CONV 1 with a K-major B (dgrad): rows of B per tap (= channels of the output)
  int tiles_m;                         // grid.x when it is not ceil(M / 128) (CONV 1: two tiles per image)
  int ext_reduce;                      // split-K: only write the partial tiles; splitk_reduce_kernel sums them afterwards
  int stages;                          // smem ring depth actually used (<= Cfg::kStages): short-K problems take
                                       // fewer so that several CTAs fit on one SM and their epilogues overlap
};

template <int BN>
struct Cfg {
  static constexpr int kStages = BN == 256 ? 4 : (BN == 128 ? 4 : 6);      // 256: 4 x 48 KB
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = BN * BK * 2;
  static constexpr int kSmemBytes = 1024 + kStages * (kABytes + kBBytes) + 256;
};

__device__ __forceinline__ void epilogue_store(const GemmParams& p, const float (&v)[32], int grow, int gcol0,
                                               unsigned long long step) {
  if (grow >= p.M) return;
  float o[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    float x = v[j];
    int c = gcol0 + j;
    if (p.bias != nullptr && c < p.N) x += __ldg(p.bias + c);
    if (p.relu) x = fmaxf(x, 0.f);
    o[j] = x;
  }
  if (p.keep_prob < 1.0f) {
    const float inv = 1.0f / p.keep_prob;
#pragma unroll
    for (int j4 = 0; j4 < 8; ++j4) {
      unsigned long long e = (unsigned long long)grow * (unsigned long long)p.N + (unsigned)(gcol0 + 4 * j4);
      uint4 w = mtb::dropout_words((uint32_t)(e >> 2), p.seed, step);
      o[4 * j4 + 0] = mtb::u32_to_unit(w.x) < p.keep_prob ? o[4 * j4 + 0] * inv : 0.f;
      o[4 * j4 + 1] = mtb::u32_to_unit(w.y) < p.keep_prob ? o[4 * j4 + 1] * inv : 0.f;
      o[4 * j4 + 2] = mtb::u32_to_unit(w.z) < p.keep_prob ? o[4 * j4 + 2] * inv : 0.f;
      o[4 * j4 + 3] = mtb::u32_to_unit(w.w) < p.keep_prob ? o[4 * j4 + 3] * inv : 0.f;
    }
  }
  const bool full = (gcol0 + 32 <= p.N);
  if (p.out_bf16) {
    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.out) + (long long)grow * p.ldc + gcol0;
    if (full && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint4 q;
        q.x = mtb::pack_bf16x2(o[8 * j + 0], o[8 * j + 1]);
        q.y = mtb::pack_bf16x2(o[8 * j + 2], o[8 * j + 3]);
        q.z = mtb::pack_bf16x2(o[8 * j + 4], o[8 * j + 5]);
        q.w = mtb::pack_bf16x2(o[8 * j + 6], o[8 * j + 7]);
        reinterpret_cast<uint4*>(dst)[j] = q;
      }
    } else {
      for (int j = 0; j < 32 && gcol0 + j < p.N; ++j) dst[j] = __float2bfloat16(o[j]);
    }
#ifdef MTB_GEMM_RS_TU
  } else if (true) {
    // GEMM -> reduce-scatter in one kernel: every rank adds its tile into the owner's slice while the other CTAs
    // are still multiplying.  red.v4.f32 at system scope: the sum is read only after the consumer's flag barrier.
    const long long flat0 = p.rs_base + (long long)grow * p.ldc + gcol0;
    if (full) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const long long flat = flat0 + 4 * j;
        float* dst = p.rs_out[(flat >> 2) / (p.rs_per >> 2)] + flat;
        asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(o[4 * j]),
                     "f"(o[4 * j + 1]), "f"(o[4 * j + 2]), "f"(o[4 * j + 3])
                     : "memory");
      }
    } else {
      for (int j = 0; j < 32 && gcol0 + j < p.N; ++j) {
        const long long flat = flat0 + j;
        float* dst = p.rs_out[(flat >> 2) / (p.rs_per >> 2)] + flat;
        asm volatile("red.relaxed.sys.global.add.f32 [%0], %1;" ::"l"(dst), "f"(o[j]) : "memory");
      }
    }
#endif
  } else {
    float* dst = reinterpret_cast<float*>(p.out) + (long long)grow * p.ldc + gcol0;
    if (full && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        reinterpret_cast<float4*>(dst)[j] = make_float4(o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
    } else {
      for (int j = 0; j < 32 && gcol0 + j < p.N; ++j) dst[j] = o[j];
    }
  }
}

#ifdef MTB_GEMM_RS_TU
// The TMEM read gives each lane one ROW of a 32x32 chunk, so adding it as is sends 32 separate 16-byte atomics per
// instruction across NVLink.  Transposed through a padded shared-memory tile, 8 lanes cover 128 contiguous bytes of
// one row and an instruction touches 4 row segments: the remote additions coalesce.  `stg` is this warp's 32x33 tile
// in the (by now idle) operand ring.
__device__ __forceinline__ void rs_store_staged(const GemmParams& p, const float (&v)[32], float* stg, int grow0, int lane,
                                                int gcol0) {
#pragma unroll
  for (int j = 0; j < 32; ++j) stg[lane * 33 + j] = v[j];
  __syncwarp();
  const int sub = lane >> 3, c4 = (lane & 7) * 4;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int r = 4 * k + sub;
    const int grow = grow0 + r;
    if (grow < p.M) {
      const float* src = stg + r * 33 + c4;
      const long long flat = p.rs_base + (long long)grow * p.ldc + gcol0 + c4;
      float* dst = p.rs_out[(flat >> 2) / (p.rs_per >> 2)] + flat;
      asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(src[0]), "f"(src[1]),
                   "f"(src[2]), "f"(src[3])
                   : "memory");
    }
  }
  __syncwarp();
}
#endif

// TF32 = true: fp32 operands in shared memory, kind::tf32 MMAs (K = 8 per instruction).  A 128-byte swizzle row
// then holds 32 elements instead of 64, so a stage covers 32 K and an MN-major atom is 32 elements wide; the
// byte geometry (16 KB of A per stage, 32 bytes of K per MMA in K-major form) is the same as for bf16.
// TWO: built for two CTAs per SM (168 registers, ring depth chosen at launch): one CTA's epilogue runs under the other's
// main loop, and a grid of up to 296 tiles is a single wave.
template <int BN, bool A_MN, bool B_MN, bool TF32 = false, int CONV = 0, bool TWO = false>
__global__ void __launch_bounds__(kThreads, (TF32 || TWO) ? 2 : 1)      // fp32 path: two CTAs per SM when the smem ring is short
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                 const GemmParams p) {
  using C = Cfg<BN>;
  const int S = (TF32 || TWO) ? p.stages : C::kStages;        // compile-time for the one-CTA-per-SM bf16 instantiations
  constexpr int kBKe = TF32 ? 32 : 64;             // K elements per stage (one 128-byte row)
  constexpr int kBlkE = TF32 ? 32 : 64;            // MN elements per MN-major atom column
  constexpr int kBlkBytes = kBKe * 128;            // one MN-major block: kBKe K-rows of 128 bytes
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + S * C::kABytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(sB + S * C::kBBytes);
  uint64_t* empty_bar = full_bar + S;
  uint64_t* tmem_full = empty_bar + S;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tmem_full + 1);
  volatile int* last_flag = reinterpret_cast<volatile int*>(tmem_holder + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile_m = blockIdx.x, tile_n = blockIdx.y, split = blockIdx.z;
  const int kb0 = split * p.kb_per_split;
  const int nkb = min(p.kb_per_split, p.kb_total - kb0);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    for (int s = 0; s < S; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<BN>(tmem_holder);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  mtb::pdl_sync();

  if (warp == 0) {
    if (elect_one()) {
      int s = 0, ph = 0;
      for (int i = 0; i < nkb; ++i, s = (s + 1 == S ? 0 : s + 1), ph ^= (s == 0)) {
        mbar_wait(&empty_bar[s], ph ^ 1);
        mbar_arrive_expect_tx(&full_bar[s], C::kABytes + C::kBBytes);
        const int kc = (kb0 + i) * kBKe;
        uint8_t* a = sA + s * C::kABytes;
        uint8_t* b = sB + s * C::kBBytes;
        if (CONV == 1) {
          // A tile: 8 image rows x 16 columns of image tile_m / 2, moved by the tap; one 128-byte channel block
          const int kb = kb0 + i, tap = kb / p.conv_cpb, cb = kb - tap * p.conv_cpb;
          const int dh = p.conv_sign * (tap / 5 - 2), dw = p.conv_sign * (tap % 5 - 2);
          tma_load_4d(a, &tma_a, &full_bar[s], cb * kBKe, dw, (tile_m & 1) * 8 + dh, tile_m >> 1);
          if (!B_MN) {                 // dgrad: B[n = ci][k = co] of this tap
            tma_load_2d(b, &tma_b, &full_bar[s], cb * kBKe, tap * p.conv_brow + tile_n * BN);
          } else {                     // fprop: rows (tap, ci) of the filter are the GEMM's K index as they are
#pragma unroll
            for (int h = 0; h < BN / kBlkE; ++h)
              tma_load_2d(b + h * kBlkBytes, &tma_b, &full_bar[s], tile_n * BN + h * kBlkE, kc);
          }
          continue;
        }
        if (CONV == 2) {
          const int kb = kb0 + i, img = kb / p.conv_kpi, h0 = (kb - img * p.conv_kpi) * p.conv_rpb;
#pragma unroll
          for (int h = 0; h < BM / kBlkE; ++h) {
            const int col = tile_m * BM + h * kBlkE, tap = col / p.conv_c, c0 = col - tap * p.conv_c;
            // columns past the 25th tap: a box that lies outside the tensor lands as zeros (and counts its bytes)
            if (tap < 25) tma_load_4d(a + h * kBlkBytes, &tma_a, &full_bar[s], c0, tap % 5 - 2, h0 + tap / 5 - 2, img);
            else tma_load_4d(a + h * kBlkBytes, &tma_a, &full_bar[s], 0, 0, 64, img);
          }
#pragma unroll
          for (int h = 0; h < BN / kBlkE; ++h)
            tma_load_4d(b + h * kBlkBytes, &tma_b, &full_bar[s], tile_n * BN + h * kBlkE, 0, h0, img);
          continue;
        }
        if (!A_MN) {
          tma_load_2d(a, &tma_a, &full_bar[s], kc, tile_m * BM);
        } else {
#pragma unroll
          for (int h = 0; h < BM / kBlkE; ++h)
            tma_load_2d(a + h * kBlkBytes, &tma_a, &full_bar[s], tile_m * BM + h * kBlkE, kc);
        }
        if (!B_MN) {
          tma_load_2d(b, &tma_b, &full_bar[s], kc, tile_n * BN);
        } else {
#pragma unroll
          for (int h = 0; h < BN / kBlkE; ++h)
            tma_load_2d(b + h * kBlkBytes, &tma_b, &full_bar[s], tile_n * BN + h * kBlkE, kc);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc(TF32 ? kFmtTF32 : kFmtBF16, BM, BN, A_MN, B_MN);
      // one descriptor per operand; stages and K steps only move the 14-bit start-address field.
      // MN-major: LBO = stride between atom columns, SBO = stride between 8-row K groups; one MMA covers
      // 16 (bf16) or 8 (tf32) K rows = two or one such group.
      // tf32 MN-major operands only exist in the 32-byte-chunk flavour of the 128-byte swizzle (atoms of 4 K rows,
      // SBO = 512): TMA writes it with SWIZZLE_128B_ATOM_32B.
      constexpr uint32_t kMnSwz = TF32 ? kSwizzle128B_Base32B : kSwizzle128B;
      constexpr uint32_t kMnSbo = TF32 ? 512 : 1024;
      const uint64_t da0 = A_MN ? make_smem_desc(smem_u32(sA), kBlkBytes, kMnSbo, kMnSwz)
                                : make_smem_desc(smem_u32(sA), 16, 1024, kSwizzle128B);
      const uint64_t db0 = B_MN ? make_smem_desc(smem_u32(sB), kBlkBytes, kMnSbo, kMnSwz)
                                : make_smem_desc(smem_u32(sB), 16, 1024, kSwizzle128B);
      constexpr uint32_t kMnStep = TF32 ? 1024 : 2048;
      constexpr uint32_t a_kstep = (A_MN ? kMnStep : 32) >> 4, b_kstep = (B_MN ? kMnStep : 32) >> 4;
      int s = 0, ph = 0;
      for (int i = 0; i < nkb; ++i) {
        mbar_wait(&full_bar[s], ph);
        tc_fence_after();
        const uint64_t da = da0 + (uint64_t)((uint32_t)(s * C::kABytes) >> 4);
        const uint64_t db = db0 + (uint64_t)((uint32_t)(s * C::kBBytes) >> 4);
        if (TF32) {
#pragma unroll
          for (int k = 0; k < 4; ++k) mma_tf32(tmem_base, da + k * a_kstep, db + k * b_kstep, idesc, i > 0 || k > 0);
        } else {
          if (i == 0) mma_f16_first(tmem_base, da, db, idesc);
          else mma_f16_acc(tmem_base, da, db, idesc);
#pragma unroll
          for (int k = 1; k < BK / 16; ++k) mma_f16_acc(tmem_base, da + k * a_kstep, db + k * b_kstep, idesc);
        }
        mma_commit(&empty_bar[s]);       // smem slot reusable once these MMAs retire
        if (++s == S) { s = 0; ph ^= 1; }
      }
      mma_commit(tmem_full);             // accumulator complete
    }
    __syncwarp();
  } else {
    const int q = warp & 3;              // TMEM lane quadrant this warp may read
    const int row = q * 32 + lane;
    int grow = tile_m * BM + row;
    if (CONV == 1) {                     // tile row -> NHWC pixel of image tile_m / 2; pad rows / columns are dropped
      const int hh = (tile_m & 1) * 8 + (row >> 4), ww = row & 15;
      grow = (hh < 14 && ww < 14) ? (tile_m >> 1) * 196 + hh * 14 + ww : p.M;
    }
    const unsigned long long step = p.step_ptr ? (unsigned long long)*p.step_ptr : 0ull;
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
    if (p.splits == 1) {
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(taddr + c * 32, r);
        tmem_ld_wait();
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
#ifdef MTB_GEMM_RS_TU
        if (p.rs_staged && tile_n * BN + c * 32 + 32 <= p.N) {
          rs_store_staged(p, v, reinterpret_cast<float*>(sA) + q * (32 * 33), tile_m * BM + q * 32, lane,
                          tile_n * BN + c * 32);
          continue;
        }
#endif
        epilogue_store(p, v, grow, tile_n * BN + c * 32, step);
      }
    } else {
      const int tile_id = tile_m * gridDim.y + tile_n;
      float* ws = p.workspace + ((size_t)(tile_id * p.splits + split) * BM + row) * BN;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(taddr + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 8; ++j)
          __stcg(reinterpret_cast<float4*>(ws + c * 32) + j,
                 make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                             __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3])));
      }
      __threadfence();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (warp == 2 && lane == 0) {
        if (p.ext_reduce) {
          *last_flag = 0;
        } else {
          unsigned old = atomicAdd(&p.counters[tile_id], 1u);
          *last_flag = (old == (unsigned)(p.splits - 1)) ? 1 : 0;
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (*last_flag) {
        __threadfence();
        const float* base = p.workspace + ((size_t)tile_id * p.splits * BM + row) * BN;
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0.f;
          for (int s = 0; s < p.splits; ++s) {
            const float4* src = reinterpret_cast<const float4*>(base + (size_t)s * BM * BN + c * 32);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              float4 t = __ldcg(src + j);
              v[4 * j] += t.x; v[4 * j + 1] += t.y; v[4 * j + 2] += t.z; v[4 * j + 3] += t.w;
            }
          }
          epilogue_store(p, v, grow, tile_n * BN + c * 32, step);
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (warp == 2 && lane == 0) p.counters[tile_id] = 0u;     // self-cleaning for the next launch
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<BN>(tmem_base);
  }
}

// ---- tensor-map cache ------------------------------------------------------------------------
struct MapKey {
  const void* base; uint64_t rows, cols, pitch; uint32_t box_rows, box_cols; int elem_bytes; bool atom32;
  bool operator<(const MapKey& o) const {
    return std::tie(base, rows, cols, pitch, box_rows, box_cols, elem_bytes, atom32) <
           std::tie(o.base, o.rows, o.cols, o.pitch, o.box_rows, o.box_cols, o.elem_bytes, o.atom32);
  }
};
std::map<MapKey, CUtensorMap> g_maps;
std::mutex g_maps_mu;

bool get_map(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t pitch_bytes,
             uint32_t box_rows, uint32_t box_cols, int elem_bytes = 2, bool atom32 = false) {
  std::lock_guard<std::mutex> lk(g_maps_mu);
  MapKey k{base, rows, cols, pitch_bytes, box_rows, box_cols, elem_bytes, atom32};
  auto it = g_maps.find(k);
  if (it != g_maps.end()) { *out = it->second; return true; }
  CUtensorMap m;
  const CUtensorMapDataType dt = elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  const CUtensorMapSwizzle swz = atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B;
  if (!tma::make_2d_sw128(&m, dt, elem_bytes, base, rows, cols, pitch_bytes, box_rows, box_cols, swz)) return false;
  if (g_maps.size() > 4096) g_maps.clear();
  g_maps[k] = m;
  *out = m;
  return true;
}

template <int BN, bool A_MN, bool B_MN, bool TF32 = false, int CONV = 0, bool TWO = false>
int launch(const CUtensorMap& ma, const CUtensorMap& mb, const GemmParams& p, cudaStream_t st) {
  auto kern = gemm_bf16_kernel<BN, A_MN, B_MN, TF32, CONV, TWO>;
  // the opt-in to more than 48 KB of dynamic shared memory is a per-device property of the function
  static bool attr_set[64] = {};
  int dev = 0;
  MTB_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    MTB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::kSmemBytes));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  dim3 grid((unsigned)(CONV == 1 ? p.tiles_m : mtb::ceil_div(p.M, BM)), (unsigned)mtb::ceil_div(p.N, BN),
            (unsigned)p.splits);
  GemmParams q = p;
  q.stages = Cfg<BN>::kStages;
  if (TF32 && p.kb_per_split < q.stages) q.stages = p.kb_per_split < 2 ? 2 : p.kb_per_split;
  // implicit convolution at small batch: a 3-deep ring lets two CTAs share an SM, so all tiles are resident at once
  if (TF32 && CONV == 1 && BN <= 128 && q.stages > 3) q.stages = 3;
  if (TWO) q.stages = 3;                                 // 3 x 32 KB (BN 128): two CTAs fit in the 227 KB of an SM
  const size_t smem = 1024 + (size_t)q.stages * (Cfg<BN>::kABytes + Cfg<BN>::kBBytes) + 256;
  MTB_CUDA_CHECK(mtb::launch(kern, grid, dim3(kThreads), smem, st, ma, mb, q));
  return 0;
}

}  // namespace

#ifdef MTB_GEMM_RS_TU

extern "C" {

// D = op(A) op(B) ADDED into the gradient buffers of the ranks that own each slice (GEMM -> reduce-scatter in one
// kernel over NVLink peer memory).  Element (i, j) of the product lands at flat index rs->base + i*ldc + j of rank
// ((flat / 4) / (rs->per / 4))'s buffer rs->out[rank].  Owners' slices must be zero (or hold earlier partial sums) and
// nobody may read them until every rank's call has completed and a cross-GPU barrier has passed: that is what
// mtb_fused_allreduce_update does for ranges declared in mtb_fused_args.prered_*.  Instantiated for the
// weight-gradient form only (X^T * dY: both operands MN-major, bn 128 or 256).
int mtb_gemm_bf16_rs(const void* A, const void* B, int M, int N, int K, int a_mn_major, int b_mn_major, long long lda,
                     long long ldb, long long ldc, int bn, const mtb_gemm_rs* rs, cudaStream_t stream) {
  MTB_REQUIRE(A && B && rs && M > 0 && N > 0 && K > 0, "gemm_rs: null pointer or empty shape");
  MTB_REQUIRE((lda * 2) % 16 == 0 && (ldb * 2) % 16 == 0, "gemm_rs: operand row pitch must be a multiple of 16 bytes");
  MTB_REQUIRE(((uintptr_t)A & 15) == 0 && ((uintptr_t)B & 15) == 0, "gemm_rs: operands must be 16-byte aligned");
  MTB_REQUIRE(a_mn_major && b_mn_major && (bn == 128 || bn == 256), "gemm_rs: MN-major x MN-major, bn 128 or 256");
  MTB_REQUIRE(rs->world > 1 && rs->world <= 16 && rs->rank >= 0 && rs->rank < rs->world, "gemm_rs: bad world / rank");
  MTB_REQUIRE(rs->base % 4 == 0 && rs->per > 0 && rs->per % 4 == 0 && ldc % 4 == 0,
              "gemm_rs: base, slice size and ldc must be multiples of 4 elements");
  MTB_REQUIRE((rs->base + (long long)(M - 1) * ldc + N - 1) / rs->per < rs->world, "gemm_rs: output exceeds the slices");
  GemmParams p = {};
  p.M = M; p.N = N; p.K = K;
  p.kb_total = (int)mtb::ceil_div(K, BK);
  p.kb_per_split = p.kb_total;
  p.splits = 1;
  p.out = rs->out[rs->rank]; p.ldc = ldc; p.out_bf16 = 0; p.bias = nullptr; p.relu = 0;
  p.keep_prob = 1.0f; p.seed = 0; p.step_ptr = nullptr;
  p.workspace = nullptr; p.counters = nullptr;
  p.rs_base = rs->base; p.rs_per = rs->per;
  p.rs_staged = (rs->flags & 1) ? 0 : 1;
  for (int r = 0; r < 16; ++r) p.rs_out[r] = nullptr;
  for (int r = 0; r < rs->world; ++r) {
    MTB_REQUIRE(rs->out[r] != nullptr, "gemm_rs: null peer buffer");
    p.rs_out[r] = reinterpret_cast<float*>(rs->out[r]);
  }
  CUtensorMap ma, mb;
  MTB_REQUIRE(get_map(&ma, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda * 2, 64, 64), "gemm_rs: tensor map for A failed");
  MTB_REQUIRE(get_map(&mb, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb * 2, 64, 64), "gemm_rs: tensor map for B failed");
  return bn == 128 ? launch<128, true, true>(ma, mb, p, stream) : launch<256, true, true>(ma, mb, p, stream);
}

}  // extern "C"

#else  // the regular library entry points

extern "C" {

size_t mtb_gemm_workspace_bytes(int M, int N, int bn, int splits) {
  if (splits <= 1) return 0;
  if (bn != 64 && bn != 128 && bn != 256) bn = 128;
  size_t tiles = (size_t)mtb::ceil_div(M, BM) * (size_t)mtb::ceil_div(N, bn);
  return tiles * (size_t)splits * BM * bn * sizeof(float);
}

size_t mtb_gemm_counter_bytes(int M, int N, int bn) {
  if (bn != 64 && bn != 128 && bn != 256) bn = 128;
  return (size_t)mtb::ceil_div(M, BM) * (size_t)mtb::ceil_div(N, bn) * sizeof(unsigned int);
}

int mtb_gemm_bf16(const void* A, const void* B, void* out, int M, int N, int K, int a_mn_major, int b_mn_major,
                  long long lda, long long ldb, long long ldc, int out_bf16, const float* bias, int relu,
                  float keep_prob, unsigned long long seed, const long long* step_ptr, int bn, int splits,
                  void* workspace, void* counters, cudaStream_t stream) {
  MTB_REQUIRE(A && B && out && M > 0 && N > 0 && K > 0, "gemm: null pointer or empty shape");
  MTB_REQUIRE((lda * 2) % 16 == 0 && (ldb * 2) % 16 == 0, "gemm: operand row pitch must be a multiple of 16 bytes");
  MTB_REQUIRE(((uintptr_t)A & 15) == 0 && ((uintptr_t)B & 15) == 0, "gemm: operands must be 16-byte aligned");
  const bool two = bn == 129 && a_mn_major == b_mn_major && splits <= 1;      // BN 128, two CTAs per SM
  if (bn == 129) bn = 128;
  if (bn != 64 && bn != 128 && bn != 256) bn = (N % 128 == 0 && mtb::ceil_div(M, BM) * (N / 128) >= 96) ? 128 : 64;
  if (splits < 1) splits = 1;
  GemmParams p = {};
  p.M = M; p.N = N; p.K = K;
  p.kb_total = (int)mtb::ceil_div(K, BK);
  if (splits > p.kb_total) splits = p.kb_total;
  p.kb_per_split = (int)mtb::ceil_div(p.kb_total, splits);
  splits = (int)mtb::ceil_div(p.kb_total, p.kb_per_split);     // no empty split
  p.splits = splits;
  MTB_REQUIRE(splits == 1 || (workspace && counters), "gemm: split-K needs workspace and counters");
  p.out = out; p.ldc = ldc; p.out_bf16 = out_bf16; p.bias = bias; p.relu = relu;
  p.keep_prob = keep_prob; p.seed = seed; p.step_ptr = step_ptr;
  p.workspace = reinterpret_cast<float*>(workspace);
  p.counters = reinterpret_cast<unsigned int*>(counters);
  CUtensorMap ma, mb;
  bool ok;
  // K-major operand stored [MN rows, K cols]; MN-major operand stored [K rows, MN cols].
  ok = a_mn_major ? get_map(&ma, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda * 2, 64, 64)
                  : get_map(&ma, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda * 2, BM, 64);
  MTB_REQUIRE(ok, "gemm: cuTensorMapEncodeTiled failed for A");
  ok = b_mn_major ? get_map(&mb, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb * 2, 64, 64)
                  : get_map(&mb, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb * 2, (uint32_t)bn, 64);
  MTB_REQUIRE(ok, "gemm: cuTensorMapEncodeTiled failed for B");
  // bn = 129: BN 128 tiles built for two CTAs per SM (dgrad / wgrad forms only)
  if (two && !a_mn_major) return launch<128, false, false, false, 0, true>(ma, mb, p, stream);
  if (two && a_mn_major) return launch<128, true, true, false, 0, true>(ma, mb, p, stream);
#define MTB_GEMM_CASE(BN_, AM_, BM_)                                            \
  if (bn == BN_ && (a_mn_major != 0) == AM_ && (b_mn_major != 0) == BM_) return launch<BN_, AM_, BM_>(ma, mb, p, stream);
  MTB_GEMM_CASE(128, false, false) MTB_GEMM_CASE(128, false, true) MTB_GEMM_CASE(128, true, false)
  MTB_GEMM_CASE(128, true, true) MTB_GEMM_CASE(64, false, false) MTB_GEMM_CASE(64, false, true)
  MTB_GEMM_CASE(64, true, false) MTB_GEMM_CASE(64, true, true)
  MTB_GEMM_CASE(256, false, false) MTB_GEMM_CASE(256, false, true) MTB_GEMM_CASE(256, true, false)
  MTB_GEMM_CASE(256, true, true)
#undef MTB_GEMM_CASE
  return mtb_set_error_msg(MTB_E_ARG, "gemm: no kernel variant", __FILE__, __LINE__);
}

// fp32 operands, TF32 tensor-core math, fp32 accumulation and output (the "fp32" configs of BASELINE.json; the
// reference's own matmuls are fp32 cuBLAS sgemm, mpipy.py:165,167).  Same operand-major options, split-K and
// epilogue as mtb_gemm_bf16; lda/ldb/ldc in elements.
int mtb_gemm_tf32(const float* A, const float* B, float* out, int M, int N, int K, int a_mn_major, int b_mn_major,
                  long long lda, long long ldb, long long ldc, const float* bias, int relu, float keep_prob,
                  unsigned long long seed, const long long* step_ptr, int bn, int splits, void* workspace,
                  void* counters, cudaStream_t stream) {
  MTB_REQUIRE(A && B && out && M > 0 && N > 0 && K > 0, "gemm_tf32: null pointer or empty shape");
  MTB_REQUIRE((lda * 4) % 16 == 0 && (ldb * 4) % 16 == 0, "gemm_tf32: operand row pitch must be a multiple of 16 bytes");
  MTB_REQUIRE(((uintptr_t)A & 15) == 0 && ((uintptr_t)B & 15) == 0, "gemm_tf32: operands must be 16-byte aligned");
  if (bn != 64 && bn != 128 && bn != 256) bn = (N % 128 == 0 && mtb::ceil_div(M, BM) * (N / 128) >= 96) ? 128 : 64;
  if (splits < 1) splits = 1;
  constexpr int kBKe = 32;
  GemmParams p = {};
  p.M = M; p.N = N; p.K = K;
  p.kb_total = (int)mtb::ceil_div(K, kBKe);
  if (splits > p.kb_total) splits = p.kb_total;
  p.kb_per_split = (int)mtb::ceil_div(p.kb_total, splits);
  splits = (int)mtb::ceil_div(p.kb_total, p.kb_per_split);
  p.splits = splits;
  MTB_REQUIRE(splits == 1 || (workspace && counters), "gemm_tf32: split-K needs workspace and counters");
  p.out = out; p.ldc = ldc; p.out_bf16 = 0; p.bias = bias; p.relu = relu;
  p.keep_prob = keep_prob; p.seed = seed; p.step_ptr = step_ptr;
  p.workspace = reinterpret_cast<float*>(workspace);
  p.counters = reinterpret_cast<unsigned int*>(counters);
  CUtensorMap ma, mb;
  bool ok;
  ok = a_mn_major ? get_map(&ma, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda * 4, kBKe, 32, 4, true)
                  : get_map(&ma, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda * 4, BM, 32, 4);
  MTB_REQUIRE(ok, "gemm_tf32: cuTensorMapEncodeTiled failed for A");
  ok = b_mn_major ? get_map(&mb, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb * 4, kBKe, 32, 4, true)
                  : get_map(&mb, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb * 4, (uint32_t)bn, 32, 4);
  MTB_REQUIRE(ok, "gemm_tf32: cuTensorMapEncodeTiled failed for B");
#define MTB_GEMM_CASE(BN_, AM_, BM_)                                            \
  if (bn == BN_ && (a_mn_major != 0) == AM_ && (b_mn_major != 0) == BM_)        \
    return launch<BN_, AM_, BM_, true>(ma, mb, p, stream);
  MTB_GEMM_CASE(128, false, false) MTB_GEMM_CASE(128, false, true) MTB_GEMM_CASE(128, true, false)
  MTB_GEMM_CASE(128, true, true) MTB_GEMM_CASE(64, false, false) MTB_GEMM_CASE(64, false, true)
  MTB_GEMM_CASE(64, true, false) MTB_GEMM_CASE(64, true, true)
  MTB_GEMM_CASE(256, false, false) MTB_GEMM_CASE(256, false, true) MTB_GEMM_CASE(256, true, false)
  MTB_GEMM_CASE(256, true, true)
#undef MTB_GEMM_CASE
  return mtb_set_error_msg(MTB_E_ARG, "gemm_tf32: no kernel variant", __FILE__, __LINE__);
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// 5x5 SAME convolution on 14x14 maps as an implicit GEMM through the engine above (CONV producer): any channel counts
// that are multiples of one 128-byte row (32 fp32 / 64 bf16 elements), bf16 or fp32 (TF32 math) tensors.  This is the
// path of every configuration the specialised 32 -> 64 bf16 shift-GEMM kernels (conv2_sm100.cu) do not cover: the fp32
// configuration and the widened model.  No im2col / col2im matrix exists anywhere.
namespace {

// Split-K partial tiles [tile][split][128][bn] summed in split order (deterministic) into out[M][ldc]: one thread per
// float4 of the result, so a reduction over 21 splits is a 21-deep chain in 12,800 threads instead of in 128.
__global__ void splitk_reduce_kernel(const float* __restrict__ ws, float* __restrict__ out, int M, int N, int bn, int splits,
                                     int tiles_n, long long ldc) {
  mtb::pdl_sync();
  const int n4 = N >> 2;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)M * n4) return;
  const int row = (int)(idx / n4), col = (int)(idx - (long long)row * n4) * 4;
  const int tile_id = (row / BM) * tiles_n + col / bn;
  const float* src = ws + ((size_t)tile_id * splits * BM + (row % BM)) * bn + (col % bn);
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int s = 0; s < splits; ++s) {
    const float4 t = __ldcg(reinterpret_cast<const float4*>(src + (size_t)s * BM * bn));
    acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
  }
  *reinterpret_cast<float4*>(out + (long long)row * ldc + col) = acc;
}

struct Map4Key {
  const void* base; uint64_t c, b; uint32_t box_c, box_rows; int elem_bytes; bool atom32;
  bool operator<(const Map4Key& o) const {
    return std::tie(base, c, b, box_c, box_rows, elem_bytes, atom32) <
           std::tie(o.base, o.c, o.b, o.box_c, o.box_rows, o.elem_bytes, o.atom32);
  }
};
std::map<Map4Key, CUtensorMap> g_maps4;

// NHWC [B,14,14,C] tensor, box = box_c channels x 16 columns x box_rows rows x 1 image
bool get_map4(CUtensorMap* out, const void* base, uint64_t C, uint64_t B, uint32_t box_c, uint32_t box_rows, int elem_bytes,
              bool atom32) {
  std::lock_guard<std::mutex> lk(g_maps_mu);
  Map4Key k{base, C, B, box_c, box_rows, elem_bytes, atom32};
  auto it = g_maps4.find(k);
  if (it != g_maps4.end()) { *out = it->second; return true; }
  const uint64_t dims[4] = {C, 14, 14, B};
  const uint64_t strides[3] = {C * elem_bytes, 14 * C * elem_bytes, 196 * C * elem_bytes};
  const uint32_t box[4] = {box_c, 16, box_rows, 1};
  const CUtensorMapDataType dt = elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  const CUtensorMapSwizzle swz = atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B;
  CUtensorMap m;
  if (!tma::make(&m, dt, 4, base, dims, strides, box, swz)) return false;
  if (g_maps4.size() > 1024) g_maps4.clear();
  g_maps4[k] = m;
  *out = m;
  return true;
}

}  // namespace

extern "C" {

// mode 0 fprop : out[B,14,14,Cout] = conv5x5(x[B,14,14,Cin], w[5,5,Cin,Cout]) (+ bias, ReLU)      (mpipy.py:159-160)
// mode 1 dgrad : out[B,14,14,Cin]  = conv5x5 transposed of dy[B,14,14,Cout] with w                 (Conv2DBackpropInput)
// mode 2 wgrad : out[5,5,Cin,Cout] (fp32) = sum over pixels of x moved by the tap (x) dy            (Conv2DBackpropFilter)
//                `x` is the layer input, `w` is dy; split-K over the batch with `workspace` / `counters` as mtb_gemm_bf16.
// tf32 = 1: every tensor fp32, TF32 tensor-core math; else bf16 tensors (wgrad output always fp32).
int mtb_conv5x5_gemm(int mode, const void* x, const void* w, void* out, int B, int Cin, int Cout, int tf32,
                     const float* bias, int relu, int splits, void* workspace, void* counters, cudaStream_t stream) {
  MTB_REQUIRE(x && w && out && B > 0, "conv5x5_gemm: null pointer or empty batch");
  MTB_REQUIRE(mode >= 0 && mode <= 2, "conv5x5_gemm: mode must be 0 (fprop), 1 (dgrad) or 2 (wgrad)");
  const int es = tf32 ? 4 : 2, row_e = 128 / es;                  // elements per 128-byte row
  MTB_REQUIRE(Cin % row_e == 0 && Cout % row_e == 0, "conv5x5_gemm: channel counts must be multiples of one 128-byte row");
  GemmParams p = {};
  p.out_bf16 = tf32 ? 0 : 1; p.bias = nullptr; p.relu = 0; p.keep_prob = 1.0f; p.splits = 1;
  p.out = out;
  CUtensorMap ma, mb;
  if (mode == 0 || mode == 1) {
    const int Ca = mode == 0 ? Cin : Cout;                        // channels of the tensor behind A
    const int Nn = mode == 0 ? Cout : Cin;
    const int bn = Nn % 256 == 0 ? 256 : (Nn % 128 == 0 ? 128 : 64);
    MTB_REQUIRE(Nn <= 64 || Nn % bn == 0, "conv5x5_gemm: output channels must be <= 64 or a multiple of 128");
    p.M = B * 196; p.N = Nn; p.K = 25 * Ca; p.ldc = Nn;
    p.conv_c = Ca; p.conv_cpb = Ca / row_e; p.conv_sign = mode == 0 ? 1 : -1; p.conv_brow = Cin;
    p.kb_total = 25 * p.conv_cpb; p.kb_per_split = p.kb_total; p.tiles_m = 2 * B;
    if (mode == 0) { p.bias = bias; p.relu = relu; }
    MTB_REQUIRE(get_map4(&ma, x, (uint64_t)Ca, (uint64_t)B, (uint32_t)row_e, 8, es, false), "conv5x5_gemm: tensor map for A failed");
    bool ok;
    // the filter as a matrix: [25*Cin rows (tap, ci)][Cout]; fprop reads it MN-major (K = rows), dgrad K-major (K = columns)
    if (mode == 0) ok = tf32 ? get_map(&mb, w, 25ull * Cin, (uint64_t)Cout, (uint64_t)Cout * 4, 32, 32, 4, true)
                             : get_map(&mb, w, 25ull * Cin, (uint64_t)Cout, (uint64_t)Cout * 2, 64, 64);
    else ok = tf32 ? get_map(&mb, w, 25ull * Cin, (uint64_t)Cout, (uint64_t)Cout * 4, (uint32_t)bn, 32, 4)
                   : get_map(&mb, w, 25ull * Cin, (uint64_t)Cout, (uint64_t)Cout * 2, (uint32_t)bn, 64);
    MTB_REQUIRE(ok, "conv5x5_gemm: tensor map for the filter failed");
#define MTB_CONV_CASE(BN_, BMN_, TF_) if (bn == BN_ && (mode == 0) == BMN_ && (tf32 != 0) == TF_) return launch<BN_, false, BMN_, TF_, 1>(ma, mb, p, stream);
    MTB_CONV_CASE(64, true, true) MTB_CONV_CASE(64, false, true) MTB_CONV_CASE(128, true, true) MTB_CONV_CASE(128, false, true)
    MTB_CONV_CASE(256, true, true) MTB_CONV_CASE(256, false, true)
    MTB_CONV_CASE(64, true, false) MTB_CONV_CASE(64, false, false) MTB_CONV_CASE(128, true, false) MTB_CONV_CASE(128, false, false)
    MTB_CONV_CASE(256, true, false) MTB_CONV_CASE(256, false, false)
#undef MTB_CONV_CASE
    return mtb_set_error_msg(MTB_E_ARG, "conv5x5_gemm: no kernel variant", __FILE__, __LINE__);
  }
  // wgrad: M = (tap, ci), N = co, K = pixels walked as (image, rows-of-16)
  const int bn = Cout % 256 == 0 ? 256 : (Cout % 128 == 0 ? 128 : 64);
  MTB_REQUIRE(Cout <= 64 || Cout % bn == 0, "conv5x5_gemm: output channels must be <= 64 or a multiple of 128");
  const int kbe = tf32 ? 32 : 64;                                  // K rows (pixels) per block
  p.M = 25 * Cin; p.N = Cout; p.ldc = Cout; p.out_bf16 = 0;
  p.conv_c = Cin; p.conv_rpb = kbe / 16; p.conv_kpi = (14 + p.conv_rpb - 1) / p.conv_rpb;
  p.kb_total = B * p.conv_kpi; p.K = p.kb_total * kbe;
  if (splits < 1) splits = 1;
  if (splits > p.kb_total) splits = p.kb_total;
  p.kb_per_split = (int)mtb::ceil_div(p.kb_total, splits);
  p.splits = (int)mtb::ceil_div(p.kb_total, p.kb_per_split);
  MTB_REQUIRE(p.splits == 1 || (workspace && counters), "conv5x5_gemm: split-K needs workspace and counters");
  p.workspace = reinterpret_cast<float*>(workspace);
  p.counters = reinterpret_cast<unsigned int*>(counters);
  MTB_REQUIRE(get_map4(&ma, x, (uint64_t)Cin, (uint64_t)B, (uint32_t)row_e, (uint32_t)p.conv_rpb, es, tf32 != 0),
              "conv5x5_gemm: tensor map for the activations failed");
  MTB_REQUIRE(get_map4(&mb, w, (uint64_t)Cout, (uint64_t)B, (uint32_t)row_e, (uint32_t)p.conv_rpb, es, tf32 != 0),
              "conv5x5_gemm: tensor map for the output gradient failed");
  p.ext_reduce = p.splits > 1 ? 1 : 0;                  // partial tiles summed by a wide second launch, not by the last CTA
  int rc = MTB_E_ARG;
#define MTB_CONV_CASE(BN_, TF_) if (bn == BN_ && (tf32 != 0) == TF_) rc = launch<BN_, true, true, TF_, 2>(ma, mb, p, stream);
  MTB_CONV_CASE(64, true) MTB_CONV_CASE(128, true) MTB_CONV_CASE(256, true)
  MTB_CONV_CASE(64, false) MTB_CONV_CASE(128, false) MTB_CONV_CASE(256, false)
#undef MTB_CONV_CASE
  if (rc != 0 || !p.ext_reduce) return rc;
  const long long n_vec = (long long)p.M * (p.N / 4);
  MTB_CUDA_CHECK(mtb::launch(splitk_reduce_kernel, dim3((unsigned)((n_vec + 255) / 256)), dim3(256), 0, stream,
                             (const float*)p.workspace, reinterpret_cast<float*>(out), p.M, p.N, bn, p.splits,
                             (int)mtb::ceil_div(p.N, bn), p.ldc));
  return 0;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// Split-K GEMM for callers that cannot hand us scratch memory.
//
// The split-K path above needs a workspace for the per-split partial tiles and a counter per
// output tile, both supplied by the caller.  Some users of this C ABI have no way to supply
// them: a plug-in loaded into a host process whose own allocator has already reserved the
// whole GPU (a raw cudaMalloc from the plug-in fails or steals from the host's budget), a
// control loop that must not call the allocator once it is running because cudaMalloc
// synchronises the device, or a small C program that links this library alone and has no
// tensor library to allocate with.  For them the scratch lives in the library image itself: a
// fixed array in device global memory, one copy on every GPU that loads libmtb200.so, held
// until the process exits.
namespace {

constexpr size_t kNoAllocWorkspaceBytes = 4u << 20;      // partial tiles: tiles * splits * 128 * bn * 4 bytes
constexpr size_t kNoAllocCounterBytes = 4u << 10;        // one 32-bit counter per output tile
constexpr int kNoAllocMaxDevices = 64;

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocWorkspaceBytes + kNoAllocCounterBytes];

// In host code `g_noalloc_scratch` is only a handle for the runtime, not a device address, so
// the address is looked up once per device (with that device current) and remembered.
void* g_noalloc_ptr[kNoAllocMaxDevices] = {};
std::mutex g_noalloc_mu;

int noalloc_scratch(void** out) {
  int dev = -1;
  MTB_CUDA_CHECK(cudaGetDevice(&dev));
  MTB_REQUIRE(dev >= 0 && dev < kNoAllocMaxDevices, "gemm_noalloc: device ordinal out of range");
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (g_noalloc_ptr[dev] == nullptr) {
    void* p = nullptr;
    MTB_CUDA_CHECK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}

}  // namespace

extern "C" {

size_t mtb_gemm_noalloc_max_workspace_bytes(void) { return kNoAllocWorkspaceBytes; }

// Same product and epilogue (bias, ReLU, bf16 or fp32 output) as mtb_gemm_bf16 with split-K,
// but with no workspace / counter arguments and no allocation of any kind: no cudaMalloc, no
// cudaFree, no tensor-library call.  Limits: the partial tiles (see
// mtb_gemm_workspace_bytes) must fit in 4 MiB and the tile counters in 4 KiB; larger
// problems are refused with MTB_E_LIMIT.
//
// The first call on each device looks up the scratch array's device address.  Make that first
// call during start-up, before any stream capture is begun on that device.
//
// Every call on a device uses the same array and finds whatever the previous call left in it.
// The tile counters have to start at zero, so each call clears them on `stream` before the
// kernel; the partial tiles are fully overwritten before they are read.  Two calls that can
// execute at the same time on one GPU (different streams, or different host threads) would
// share the array and corrupt each other: callers must serialise them.
int mtb_gemm_bf16_splitk_noalloc(const void* A, const void* B, void* out, int M, int N, int K, int a_mn_major,
                                 int b_mn_major, long long lda, long long ldb, long long ldc, int out_bf16,
                                 const float* bias, int relu, int bn, int splits, cudaStream_t stream) {
  MTB_REQUIRE(A && B && out && M > 0 && N > 0 && K > 0, "gemm_noalloc: null pointer or empty shape");
  MTB_REQUIRE((lda * 2) % 16 == 0 && (ldb * 2) % 16 == 0, "gemm_noalloc: operand row pitch must be a multiple of 16 bytes");
  MTB_REQUIRE(((uintptr_t)A & 15) == 0 && ((uintptr_t)B & 15) == 0, "gemm_noalloc: operands must be 16-byte aligned");
  if (bn == 0) bn = 64;                                  // 0 = default tile width
  MTB_REQUIRE(bn == 64 || bn == 128, "gemm_noalloc: bn must be 0 (default), 64 or 128");
  if (splits < 1) splits = 1;
  GemmParams p = {};
  p.M = M; p.N = N; p.K = K;
  p.kb_total = (int)mtb::ceil_div(K, BK);
  if (splits > p.kb_total) splits = p.kb_total;
  p.kb_per_split = (int)mtb::ceil_div(p.kb_total, splits);
  splits = (int)mtb::ceil_div(p.kb_total, p.kb_per_split);
  p.splits = splits;
  const size_t tiles = (size_t)mtb::ceil_div(M, BM) * (size_t)mtb::ceil_div(N, bn);
  const size_t ws_bytes = splits > 1 ? tiles * (size_t)splits * BM * bn * sizeof(float) : 0;
  const size_t cnt_bytes = tiles * sizeof(unsigned int);
  if (ws_bytes > kNoAllocWorkspaceBytes || cnt_bytes > kNoAllocCounterBytes)
    return mtb_set_error_msg(MTB_E_LIMIT,
                             "gemm_noalloc: split-K scratch for this shape exceeds the built-in 4 MiB + 4 KiB array; "
                             "use mtb_gemm_bf16 with caller-provided workspace", __FILE__, __LINE__);
  void* scratch = nullptr;
  if (int e = noalloc_scratch(&scratch)) return e;
  unsigned char* base = reinterpret_cast<unsigned char*>(scratch);
  p.workspace = reinterpret_cast<float*>(base);
  p.counters = reinterpret_cast<unsigned int*>(base + kNoAllocWorkspaceBytes);
  MTB_CUDA_CHECK(cudaMemsetAsync(p.counters, 0, cnt_bytes, stream));
  p.out = out; p.ldc = ldc; p.out_bf16 = out_bf16; p.bias = bias; p.relu = relu;
  p.keep_prob = 1.0f; p.seed = 0; p.step_ptr = nullptr;
  // tensor maps are encoded on the stack here: the cached path of mtb_gemm_bf16 would grow a host map
  CUtensorMap ma, mb;
  bool ok;
  ok = a_mn_major ? tma::make_2d_sw128(&ma, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda * 2, 64, 64)
                  : tma::make_2d_sw128(&ma, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda * 2, BM, 64);
  MTB_REQUIRE(ok, "gemm_noalloc: cuTensorMapEncodeTiled failed for A");
  ok = b_mn_major ? tma::make_2d_sw128(&mb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb * 2, 64, 64)
                  : tma::make_2d_sw128(&mb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb * 2, (uint32_t)bn, 64);
  MTB_REQUIRE(ok, "gemm_noalloc: cuTensorMapEncodeTiled failed for B");
#define MTB_GEMM_CASE(BN_, AM_, BM_)                                            \
  if (bn == BN_ && (a_mn_major != 0) == AM_ && (b_mn_major != 0) == BM_) return launch<BN_, AM_, BM_>(ma, mb, p, stream);
  MTB_GEMM_CASE(128, false, false) MTB_GEMM_CASE(128, false, true) MTB_GEMM_CASE(128, true, false)
  MTB_GEMM_CASE(128, true, true) MTB_GEMM_CASE(64, false, false) MTB_GEMM_CASE(64, false, true)
  MTB_GEMM_CASE(64, true, false) MTB_GEMM_CASE(64, true, true)
#undef MTB_GEMM_CASE
  return mtb_set_error_msg(MTB_E_ARG, "gemm_noalloc: no kernel variant", __FILE__, __LINE__);
}

}  // extern "C"

#endif  // MTB_GEMM_RS_TU
