// K1 — the data-parallel hot path: cross-GPU gradient allreduce fused with the 1/N average,
// the L2 weight-decay term, the optimizer update (SGD / momentum / Adam), the staircase LR
// schedule and the fp32 -> bf16 shadow-weight cast, in ONE kernel that reads and writes peer
// GPUs' buffers directly over NVLink (CUDA-IPC mapped pointers).  No NCCL call and no separate
// elementwise kernel runs on this path.
//
// What it replaces in the reference: Gather to rank 0 over host MPI (mpipy.py:121-127), the
// NumPy mean on one CPU thread (mpipy.py:130-136), the H2D tf.assign on root only
// (mpipy.py:138-152), the L2 gradient (mpipy.py:57-58), exponential_decay (mpipy.py:60-64) and
// ApplyMomentum (mpipy.py:65-66).
//
// Algorithms (chosen per bucket size by the caller):
//   one-shot  every rank loads the bucket from all N peers, reduces in rank order in fp32
//             registers and updates its own replica.  (N-1)*S bytes inbound, one signal round
//             before and one after.  Optimizer slots replicated.
//   two-shot  rank r owns slice r: reduce-scatter by peer loads, update the slice, all-gather
//             by peer stores of the new fp32 master + bf16 shadow into every replica.
//             2*(N-1)/N*S bytes per GPU.  Optimizer slots sharded (ZeRO-1 style).
//   nvls      as two-shot, but multimem.ld_reduce pulls the switch-reduced slice in one load
//             and multimem.st broadcasts the result (needs a multicast mapping).
// Every rank sums in the same order, so replicas stay bit-identical.
//
// Flag protocol: one 32-bit slot per (block, peer) in each rank's signal pad, written with
// st.release.sys after a system fence and polled with ld.acquire.sys.  Values are a per-block
// sequence number that only grows (2 per launch), so nothing is ever reset and CUDA-graph
// replays are safe.  Every spin is bounded; on expiry the status word is set and the kernel
// finishes instead of hanging the GPU (SURVEY section 5 "failure detection").
#include "common.cuh"

namespace {

using namespace mtb;

constexpr int kMaxRanks = 16;
constexpr int kMaxBlocks = 512;
constexpr int kMaxRanges = 8;
constexpr long long kSpinLimit = 30000000000ll;     // ~15 s at 2 GHz

struct FusedParams {
  int world, rank, algo, optimizer;                 // optimizer: -1 none (plain allreduce), 0 sgd, 1 momentum, 2 adam
  const float* grads[kMaxRanks];                    // per-rank flat fp32 gradient buffers (peer mapped)
  float* params[kMaxRanks];                         // per-rank flat fp32 master parameters / allreduce output
  __nv_bfloat16* shadow[kMaxRanks];                 // per-rank bf16 shadow weights (may be null)
  const float* mc_grads;                            // multicast views (nvls only)
  float* mc_params;
  __nv_bfloat16* mc_shadow;
  float* slot_m;
  float* slot_v;
  uint32_t* flags[kMaxRanks];                       // signal pads [kMaxBlocks][kMaxRanks]
  uint32_t* seq;                                    // local per-block sequence numbers [kMaxBlocks]
  int* status;
  long long offset, count;                          // bucket range in elements, multiples of 4
  float scale;                                      // 1/N (or 1 for a plain sum)
  float lr0, lr_decay, lr_override;
  int batch;
  long long decay_rows;
  const long long* step_ptr;
  float momentum, beta1, beta2, eps, wd;
  int n_ranges;
  long long range_begin[kMaxRanges], range_end[kMaxRanges];   // where the L2 term applies
  int n_sharded;                                    // ranges whose fp32 master stays with the slice owner
  long long shard_begin[4], shard_end[4];
  long long* advance_step;                          // bumped by the last block to finish (may be null)
  unsigned int* done_counter;
  // Ranges whose gradient slices arrive ALREADY summed in the owner's buffer: the producing GEMM's epilogue
  // added every rank's tile straight into the owner over NVLink (mtb_gemm_bf16_rs).  The owner reads its slice
  // locally, skips the peer reduction, and clears it for the next step.
  int n_prered;
  long long prered_begin[2], prered_end[2];
  const float* loss_src;                            // step result mailed to the host by the last block (may be null)
  unsigned long long* loss_mail;                    // pinned host ring of {f32 loss, u32 step} words
  int loss_ring;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_f4(float* p, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void multimem_st_bf16x4(__nv_bfloat16* p, uint2 v) {
  asm volatile("multimem.st.relaxed.sys.global.v2.bf16x2 [%0], {%1,%2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}

// All ranks' block `b` meet here.  With `fence`, every thread first makes its own stores (to
// peer memory) visible system-wide; without, the barrier only orders control flow, which is
// all the opening round needs: the gradients it guards were written by earlier kernels, and a
// kernel boundary already made them visible to peer loads.
__device__ void cross_gpu_barrier(const FusedParams& p, int b, uint32_t val, bool fence) {
  if (fence) __threadfence_system();
  __syncthreads();
  if ((int)threadIdx.x < p.world) {
    const int peer = threadIdx.x;
    st_release_sys(p.flags[peer] + b * kMaxRanks + p.rank, val);
    const uint32_t* mine = p.flags[p.rank] + b * kMaxRanks + peer;
    const long long t0 = clock64();
    while ((int32_t)(ld_acquire_sys(mine) - val) < 0) {
      if (clock64() - t0 > kSpinLimit) {
        atomicExch(p.status, MTB_E_TIMEOUT);
        break;
      }
    }
  }
  __syncthreads();
}

struct Hyper {
  float lr, bc_lr;      // bc_lr: Adam's bias-corrected step size
};

__device__ __forceinline__ Hyper make_hyper(const FusedParams& p) {
  Hyper h;
  const long long step = p.step_ptr ? *p.step_ptr : 0;
  float lr = p.lr0;
  if (p.lr_override >= 0.f) {
    lr = p.lr_override;
  } else if (p.decay_rows > 0 && p.lr_decay != 1.0f) {
    // tf.train.exponential_decay(lr0, step*batch, decay_rows, decay, staircase=True), fp32 like TF
    const float e = floorf((float)step * (float)p.batch / (float)p.decay_rows);
    lr = p.lr0 * powf(p.lr_decay, e);
  }
  h.lr = lr;
  const float t = (float)(step + 1);
  h.bc_lr = lr * sqrtf(1.0f - powf(p.beta2, t)) / (1.0f - powf(p.beta1, t));
  return h;
}

__device__ __forceinline__ bool decays(const FusedParams& p, long long i) {
  bool d = false;
#pragma unroll
  for (int r = 0; r < kMaxRanges; ++r)
    if (r < p.n_ranges && i >= p.range_begin[r] && i < p.range_end[r]) d = true;
  return d;
}

__device__ __forceinline__ bool master_sharded(const FusedParams& p, long long i) {
  bool s = false;
#pragma unroll
  for (int r = 0; r < 4; ++r)
    if (r < p.n_sharded && i >= p.shard_begin[r] && i < p.shard_end[r]) s = true;
  return s;
}

__device__ __forceinline__ bool prereduced(const FusedParams& p, long long i) {
  bool s = false;
#pragma unroll
  for (int r = 0; r < 2; ++r)
    if (r < p.n_prered && i >= p.prered_begin[r] && i < p.prered_end[r]) s = true;
  return s;
}

// One float4 of the optimizer.  g is the already averaged gradient.
__device__ __forceinline__ float4 apply_update(const FusedParams& p, const Hyper& h, long long i, float4 g) {
  float4 w = *reinterpret_cast<const float4*>(p.params[p.rank] + i);
  if (p.optimizer < 0) return g;
  if (p.wd != 0.f && decays(p, i)) {            // L2 term: grad += wd * w (mpipy.py:57-58)
    g.x = fmaf(p.wd, w.x, g.x); g.y = fmaf(p.wd, w.y, g.y); g.z = fmaf(p.wd, w.z, g.z); g.w = fmaf(p.wd, w.w, g.w);
  }
  if (p.optimizer == 0) {
    w.x -= h.lr * g.x; w.y -= h.lr * g.y; w.z -= h.lr * g.z; w.w -= h.lr * g.w;
  } else if (p.optimizer == 1) {                // accum = m*accum + g; w -= lr*accum (mpipy.py:65)
    float4 m = *reinterpret_cast<float4*>(p.slot_m + i);
    m.x = fmaf(p.momentum, m.x, g.x); m.y = fmaf(p.momentum, m.y, g.y);
    m.z = fmaf(p.momentum, m.z, g.z); m.w = fmaf(p.momentum, m.w, g.w);
    *reinterpret_cast<float4*>(p.slot_m + i) = m;
    w.x -= h.lr * m.x; w.y -= h.lr * m.y; w.z -= h.lr * m.z; w.w -= h.lr * m.w;
  } else {
    float4 m = *reinterpret_cast<float4*>(p.slot_m + i);
    float4 v = *reinterpret_cast<float4*>(p.slot_v + i);
    const float b1 = p.beta1, b2 = p.beta2, c1 = 1.f - b1, c2 = 1.f - b2;
    m.x = b1 * m.x + c1 * g.x; m.y = b1 * m.y + c1 * g.y; m.z = b1 * m.z + c1 * g.z; m.w = b1 * m.w + c1 * g.w;
    v.x = b2 * v.x + c2 * g.x * g.x; v.y = b2 * v.y + c2 * g.y * g.y;
    v.z = b2 * v.z + c2 * g.z * g.z; v.w = b2 * v.w + c2 * g.w * g.w;
    *reinterpret_cast<float4*>(p.slot_m + i) = m;
    *reinterpret_cast<float4*>(p.slot_v + i) = v;
    w.x -= h.bc_lr * m.x / (sqrtf(v.x) + p.eps); w.y -= h.bc_lr * m.y / (sqrtf(v.y) + p.eps);
    w.z -= h.bc_lr * m.z / (sqrtf(v.z) + p.eps); w.w -= h.bc_lr * m.w / (sqrtf(v.w) + p.eps);
  }
  return w;
}

__device__ __forceinline__ uint2 to_bf16x4(float4 w) {
  return make_uint2(pack_bf16x2(w.x, w.y), pack_bf16x2(w.z, w.w));
}

enum Kind { kLocal = 0, kOneShot = 1, kTwoShot = 2, kNvls = 3 };

// Vectors [lo, hi) of the bucket, kUnroll per thread per pass: all reductions (peer loads or
// switch-reduced multimem loads) of a pass are issued before any of them is used, which is what
// keeps enough bytes in flight to cover NVLink latency.
template <int KIND>
__device__ __forceinline__ void run_range(const FusedParams& p, const Hyper& h, long long lo, long long hi) {
  constexpr int kUnroll = 4;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long v0 = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; v0 < hi; v0 += stride * kUnroll) {
    float4 g[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const long long v = v0 + u * stride;
      g[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (v < hi) {
        const long long i = p.offset + (v << 2);
        if (KIND == kLocal) {
          g[u] = *reinterpret_cast<const float4*>(p.grads[0] + i);
        } else if ((KIND == kNvls || KIND == kTwoShot) && p.n_prered > 0 && prereduced(p, i)) {
          g[u] = ld_peer_f4(p.grads[p.rank] + i);            // summed in place by the producers' epilogues
          *reinterpret_cast<float4*>(const_cast<float*>(p.grads[p.rank]) + i) = make_float4(0.f, 0.f, 0.f, 0.f);
        } else if (KIND == kNvls) {
          g[u] = multimem_ld_reduce_f4(p.mc_grads + i);
        } else {
          for (int r = 0; r < p.world; ++r) {       // rank order: every replica sums identically
            const float4 t = ld_peer_f4(p.grads[r] + i);
            g[u].x += t.x; g[u].y += t.y; g[u].z += t.z; g[u].w += t.w;
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const long long v = v0 + u * stride;
      if (v >= hi) continue;
      const long long i = p.offset + (v << 2);
      float4 gs = g[u];
      gs.x *= p.scale; gs.y *= p.scale; gs.z *= p.scale; gs.w *= p.scale;
      const float4 w = apply_update(p, h, i, gs);
      const uint2 wb = to_bf16x4(w);
      if (KIND == kLocal || KIND == kOneShot) {
        *reinterpret_cast<float4*>(p.params[p.rank] + i) = w;
        if (p.shadow[p.rank]) *reinterpret_cast<uint2*>(p.shadow[p.rank] + i) = wb;
      } else if (KIND == kNvls) {
        // sharded master: my fp32 copy is the only current one, peers get the bf16 shadow
        if (master_sharded(p, i)) *reinterpret_cast<float4*>(p.params[p.rank] + i) = w;
        else multimem_st_f4(p.mc_params + i, w);
        if (p.mc_shadow) multimem_st_bf16x4(p.mc_shadow + i, wb);
      } else {
        const bool sharded = master_sharded(p, i);
        for (int r = 0; r < p.world; ++r) {
          if (!sharded || r == p.rank) *reinterpret_cast<float4*>(p.params[r] + i) = w;
          if (p.shadow[r]) *reinterpret_cast<uint2*>(p.shadow[r] + i) = wb;
        }
      }
    }
  }
}

__global__ void __launch_bounds__(256)
fused_allreduce_update_kernel(const FusedParams p) {
  const int b = blockIdx.x;
  __shared__ uint32_t s_seq;
  pdl_sync();
  if (threadIdx.x == 0) s_seq = p.seq[b];
  __syncthreads();
  const uint32_t seq = s_seq;
  const Hyper h = make_hyper(p);
  const bool multi = p.world > 1;
  // every peer's gradients are complete (with pre-reduced ranges some of them were written into MY memory by
  // peers' kernels: the fence makes this rank's own remote additions globally visible before it signals)
  if (multi) cross_gpu_barrier(p, b, seq + 1, p.n_prered > 0);

  const long long vec_total = p.count >> 2;
  if (!multi) {
    run_range<kLocal>(p, h, 0, vec_total);
  } else if (p.algo == 0) {
    run_range<kOneShot>(p, h, 0, vec_total);
  } else {
    // two-shot / nvls: this rank owns one contiguous slice
    const long long per = (vec_total + p.world - 1) / p.world;
    const long long lo = per * p.rank < vec_total ? per * p.rank : vec_total;
    const long long hi = lo + per < vec_total ? lo + per : vec_total;
    if (p.algo == 2) run_range<kNvls>(p, h, lo, hi);
    else run_range<kTwoShot>(p, h, lo, hi);
  }
  // peers are done reading my grads (loads have retired: their values were consumed) and, in the
  // two-shot / nvls forms, done writing my params (needs the fence)
  if (multi) cross_gpu_barrier(p, b, seq + 2, p.algo != 0);
  if (threadIdx.x == 0) {
    p.seq[b] = seq + 2;
    if (p.advance_step != nullptr) {
      // every block read the step counter at its start; the last one to finish moves it on
      __threadfence();
      if (atomicAdd(p.done_counter, 1u) == gridDim.x - 1) {
        *p.done_counter = 0u;
        const long long next = *p.advance_step + 1;
        *p.advance_step = next;
        if (p.loss_mail != nullptr) {
          // one 8-byte store = one PCIe write: the host sees the loss and its step number together
          const unsigned long long word = (static_cast<unsigned long long>(static_cast<uint32_t>(next)) << 32) |
                                          __float_as_uint(*p.loss_src);
          asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p.loss_mail + ((next - 1) & (p.loss_ring - 1))),
                       "l"(word) : "memory");
        }
      }
    }
  }
}

__global__ void barrier_kernel(const FusedParams p) {
  __shared__ uint32_t s_seq;
  if (threadIdx.x == 0) s_seq = p.seq[0];
  __syncthreads();
  const uint32_t seq = s_seq;
  cross_gpu_barrier(p, 0, seq + 2, true);
  if (threadIdx.x == 0) p.seq[0] = seq + 2;
}

__global__ void advance_step_kernel(long long* step) { *step += 1; }

}  // namespace

extern "C" {

static int fill(const mtb_fused_args* a, FusedParams* p) {
  MTB_REQUIRE(a->world >= 1 && a->world <= kMaxRanks && a->rank >= 0 && a->rank < a->world, "fused: bad world/rank");
  MTB_REQUIRE(a->offset % 4 == 0 && a->count % 4 == 0 && a->count >= 0, "fused: offset/count must be multiples of 4");
  MTB_REQUIRE(a->n_ranges >= 0 && a->n_ranges <= kMaxRanges, "fused: too many decay ranges");
  MTB_REQUIRE(a->algo != 2 || (a->mc_grads && a->mc_params), "fused: nvls needs multicast pointers");
  p->world = a->world; p->rank = a->rank; p->algo = a->algo; p->optimizer = a->optimizer;
  for (int r = 0; r < kMaxRanks; ++r) {
    p->grads[r] = reinterpret_cast<const float*>(a->grads[r]);
    p->params[r] = reinterpret_cast<float*>(a->params[r]);
    p->shadow[r] = reinterpret_cast<__nv_bfloat16*>(a->shadow[r]);
    p->flags[r] = reinterpret_cast<uint32_t*>(a->flags[r]);
  }
  p->mc_grads = reinterpret_cast<const float*>(a->mc_grads);
  p->mc_params = reinterpret_cast<float*>(a->mc_params);
  p->mc_shadow = reinterpret_cast<__nv_bfloat16*>(a->mc_shadow);
  p->slot_m = reinterpret_cast<float*>(a->slot_m);
  p->slot_v = reinterpret_cast<float*>(a->slot_v);
  p->seq = reinterpret_cast<uint32_t*>(a->seq);
  p->status = reinterpret_cast<int*>(a->status);
  p->offset = a->offset; p->count = a->count; p->scale = a->scale;
  p->lr0 = a->lr0; p->lr_decay = a->lr_decay; p->lr_override = a->lr_override;
  p->batch = a->batch; p->decay_rows = a->decay_rows;
  p->step_ptr = reinterpret_cast<const long long*>(a->step_ptr);
  p->momentum = a->momentum; p->beta1 = a->beta1; p->beta2 = a->beta2; p->eps = a->eps; p->wd = a->wd;
  p->n_ranges = a->n_ranges;
  for (int r = 0; r < kMaxRanges; ++r) { p->range_begin[r] = a->range_begin[r]; p->range_end[r] = a->range_end[r]; }
  MTB_REQUIRE(a->n_sharded >= 0 && a->n_sharded <= 4, "fused: at most 4 sharded-master ranges");
  p->n_sharded = a->n_sharded;
  for (int r = 0; r < 4; ++r) { p->shard_begin[r] = a->shard_begin[r]; p->shard_end[r] = a->shard_end[r]; }
  MTB_REQUIRE(p->n_sharded == 0 || (a->shadow[a->rank] != nullptr), "fused: sharded masters need the bf16 shadow");
  p->advance_step = reinterpret_cast<long long*>(a->advance_step);
  p->done_counter = reinterpret_cast<unsigned int*>(a->done_counter);
  MTB_REQUIRE(p->advance_step == nullptr || p->done_counter != nullptr, "fused: advance_step needs done_counter");
  MTB_REQUIRE(a->n_prered >= 0 && a->n_prered <= 2, "fused: at most 2 pre-reduced ranges");
  MTB_REQUIRE(a->n_prered == 0 || (a->world > 1 && a->algo != 0),
              "fused: pre-reduced ranges need an owner-sliced algorithm (two-shot or nvls) and peers");
  p->n_prered = a->n_prered;
  for (int r = 0; r < 2; ++r) { p->prered_begin[r] = a->prered_begin[r]; p->prered_end[r] = a->prered_end[r]; }
  p->loss_src = reinterpret_cast<const float*>(a->loss_src);
  p->loss_mail = reinterpret_cast<unsigned long long*>(a->loss_mail);
  p->loss_ring = a->loss_ring;
  MTB_REQUIRE(p->loss_mail == nullptr || (p->advance_step != nullptr && p->loss_src != nullptr && p->loss_ring > 0 &&
                                          (p->loss_ring & (p->loss_ring - 1)) == 0),
              "fused: loss_mail needs advance_step, loss_src and a power-of-two loss_ring");
  return 0;
}

int mtb_fused_allreduce_update(const mtb_fused_args* a, cudaStream_t stream) {
  FusedParams p;
  int e = fill(a, &p);
  if (e) return e;
  MTB_REQUIRE(p.optimizer < 1 || p.slot_m, "fused: momentum/adam need slot_m");
  MTB_REQUIRE(p.optimizer < 2 || p.slot_v, "fused: adam needs slot_v");
  MTB_REQUIRE(p.world == 1 || (p.seq && p.status && p.flags[p.rank]), "fused: multi-GPU needs flags/seq/status");
  MTB_REQUIRE(p.seq, "fused: seq buffer required");
  if (p.count == 0) return 0;
  // 256-thread blocks without shared memory fit next to the persistent tensor-core kernels, so
  // a bucket's update can run on a second stream underneath the rest of backward
  int block = a->block > 0 ? (a->block > 256 ? 256 : a->block) : 256;
  long long vecs = p.count >> 2;
  if (p.world > 1 && p.algo != 0) vecs = (vecs + p.world - 1) / p.world;
  int grid = a->grid > 0 ? a->grid : 3 * 148;
  // one pass = 4 vectors (64 bytes) per thread; another block (and its flag round) pays off after that
  long long need = (vecs + 4 * block - 1) / (4 * block);
  if (need < 1) need = 1;
  if (grid > need) grid = (int)need;
  if (grid > kMaxBlocks) grid = kMaxBlocks;
  MTB_CUDA_CHECK(mtb::launch(fused_allreduce_update_kernel, dim3(grid), dim3(block), 0, stream, p));
  return 0;
}

int mtb_p2p_barrier(const mtb_fused_args* a, cudaStream_t stream) {
  FusedParams p;
  int e = fill(a, &p);
  if (e) return e;
  if (p.world == 1) return 0;
  barrier_kernel<<<1, 32, 0, stream>>>(p);
  MTB_LAUNCH_CHECK();
  return 0;
}

int mtb_advance_step(long long* step_ptr, cudaStream_t stream) {
  advance_step_kernel<<<1, 1, 0, stream>>>(step_ptr);
  MTB_LAUNCH_CHECK();
  return 0;
}

int mtb_fused_max_blocks(void) { return kMaxBlocks; }
int mtb_fused_flag_bytes(void) { return kMaxBlocks * kMaxRanks * (int)sizeof(uint32_t); }

}  // extern "C"
