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
//   push      one-shot for the small bucket that closes a step: every rank STORES its gradients
//             into a receive row on each peer as flag-in-data words (value + step tag in one
//             16-byte store) and the receiver polls the words it needs.  No flag round at all:
//             one one-way NVLink latency.  Rows are double buffered by step parity (a row is
//             rewritten two steps later, which a rank only reaches after every peer has met it
//             in the hidden bucket's barrier of the step in between).
// Every rank sums in the same order, so replicas stay bit-identical.
//
// The same kernel averages PARAMETERS in place (the reference's own strategy, mpipy.py:95-153):
// grads := params, optimizer = none, two-shot / nvls (an owner is the only reader of its slice
// on every peer, so in-place is race free), dst_mask = which ranks receive the mean
// (bit 0 only = the reference's root-only assign, mpipy.py:129).
//
// It also returns the step's loss the way the reference defines it (mpipy.py:55-58): each launch
// sums 0.5*w^2 over the weight-decay ranges it walks (block partials in block order, so the value is
// reproducible), slice owners exchange their partial through tagged 8-byte slots, and the launch that
// advances the step mails CE + wd*L2 and the batch's error count to the host.
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
constexpr int kChan1Base = 448;                     // flag / sequence blocks: channel 0 = [0, 448), channel 1 = [448, 511),
constexpr int kChan1Blocks = 63;                    // block 511 = the stand-alone barrier (concurrent launches never share)
constexpr int kMaxL2Blocks = 4096;                  // per-block L2 partials of one launch
constexpr int kL2Rows = 4;                          // launches of one step that can contribute an L2 partial
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
  int chan_base;                                    // first flag / sequence block of this launch's channel
  float* push_recv[kMaxRanks];                      // push: every rank's receive area [2 parity][world][2 * push_stride]
  long long push_stride;
  const __nv_bfloat16* grads_lp[kMaxRanks];         // bf16 wire format: gradients of [lp_begin, lp_end) live here
  const __nv_bfloat16* mc_grads_lp;
  long long lp_begin, lp_end;
  unsigned dst_mask;                                // two-shot: ranks that receive the result (0 = all)
  float* l2_blocks;                                 // per-block partials of 0.5*sum(w^2) (null: no L2 value wanted)
  unsigned long long* l2_slots[kMaxRanks];          // every rank's [kL2Rows][kMaxRanks] tagged {step+1 : f32} words
  int l2_row;
  int l2_rows_n[kL2Rows];                           // mailing launch: entries to add per row (0 unused, 1 local, world sliced)
  int no_flags;                                     // the caller brackets this launch with barrier launches instead
  unsigned long long* trace;                        // optional: globaltimer at the first block's start [0] and at the close [1]
  const int* err_src;                               // batch error count, mailed next to the loss
  unsigned long long* aux_mail;                     // pinned host ring of {u32 errors, u32 step} words
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
__device__ __forceinline__ float4 ld_peer_bf16x4(const __nv_bfloat16* p) {
  uint32_t a, b;
  asm volatile("ld.relaxed.sys.global.v2.b32 {%0,%1}, [%2];" : "=r"(a), "=r"(b) : "l"(p) : "memory");
  const float2 lo = unpack_bf16x2(a), hi = unpack_bf16x2(b);
  return make_float4(lo.x, lo.y, hi.x, hi.y);
}
// switch-reduced bf16 gradients: the switch accumulates in fp32 and rounds the sum once
__device__ __forceinline__ float4 multimem_ld_reduce_bf16x4(const __nv_bfloat16* p) {
  uint32_t a, b;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v2.bf16x2 {%0,%1}, [%2];"
               : "=r"(a), "=r"(b)
               : "l"(p)
               : "memory");
  const float2 lo = unpack_bf16x2(a), hi = unpack_bf16x2(b);
  return make_float4(lo.x, lo.y, hi.x, hi.y);
}
__device__ __forceinline__ void st_peer_f4(float* p, float4 v) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
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
// kernel boundary already made them visible to peer loads.  `b` is the absolute flag block
// (channel base + block index).
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

__device__ __forceinline__ Hyper make_hyper(const FusedParams& p, long long step) {
  Hyper h;
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

__device__ __forceinline__ bool low_precision(const FusedParams& p, long long i) {
  return i >= p.lp_begin && i < p.lp_end;
}

// One float4 of the optimizer.  g is the already averaged gradient.  l2 collects 0.5*w^2 of the weights BEFORE the
// update over the weight-decay ranges (the L2 term of the loss, mpipy.py:57-58).
// (Loading w / m / v for all vectors of a pass up front, next to the gradients, was measured: 128 registers, half
// the resident threads, the 13 MB bucket 41 -> 105 us.  The loads stay here.)
// PRE (the small, latency-bound buckets: one pass per thread, register budget irrelevant): w / m / v were loaded by the
// caller together with the gradient, so the pass costs one memory round trip instead of two.
template <bool PRE>
__device__ __forceinline__ float4 apply_update(const FusedParams& p, const Hyper& h, long long i, float4 g, float& l2,
                                               float4 w, float4 m, float4 v) {
  if (p.optimizer < 0) return g;
  if (!PRE) w = *reinterpret_cast<const float4*>(p.params[p.rank] + i);
  if (p.n_ranges > 0 && decays(p, i)) {         // L2 term: grad += wd * w (mpipy.py:57-58)
    l2 += 0.5f * (w.x * w.x + w.y * w.y + w.z * w.z + w.w * w.w);
    g.x = fmaf(p.wd, w.x, g.x); g.y = fmaf(p.wd, w.y, g.y); g.z = fmaf(p.wd, w.z, g.z); g.w = fmaf(p.wd, w.w, g.w);
  }
  if (p.optimizer == 0) {
    w.x -= h.lr * g.x; w.y -= h.lr * g.y; w.z -= h.lr * g.z; w.w -= h.lr * g.w;
  } else if (p.optimizer == 1) {                // accum = m*accum + g; w -= lr*accum (mpipy.py:65)
    if (!PRE) m = *reinterpret_cast<float4*>(p.slot_m + i);
    m.x = fmaf(p.momentum, m.x, g.x); m.y = fmaf(p.momentum, m.y, g.y);
    m.z = fmaf(p.momentum, m.z, g.z); m.w = fmaf(p.momentum, m.w, g.w);
    *reinterpret_cast<float4*>(p.slot_m + i) = m;
    w.x -= h.lr * m.x; w.y -= h.lr * m.y; w.z -= h.lr * m.z; w.w -= h.lr * m.w;
  } else {
    if (!PRE) {
      m = *reinterpret_cast<float4*>(p.slot_m + i);
      v = *reinterpret_cast<float4*>(p.slot_v + i);
    }
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

// Flag-in-data words of the push bucket (the "LL" idea): a 16-byte store carries two floats and the step tag twice,
// {d0, tag, d1, tag}; the receiver polls the data itself until both tags match, so there is no fence, no separate
// flag and no second trip: one one-way NVLink latency from the sender's store to the receiver's use.
__device__ __forceinline__ void st_ll(uint4* dst, float a, float b, uint32_t tag) {
  asm volatile("st.relaxed.sys.global.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(dst), "r"(__float_as_uint(a)), "r"(tag),
               "r"(__float_as_uint(b)), "r"(tag) : "memory");
}
__device__ __forceinline__ uint4 ld_ll(const uint4* src) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(src) : "memory");
  return v;
}

// Vectors [lo, hi) of the bucket, kUnroll per thread per pass: all reductions (peer loads or switch-reduced multimem
// loads) of a pass are issued before any of them is used, which is what keeps enough bytes in flight to cover NVLink
// latency.
template <int KIND, bool PRE>
__device__ __forceinline__ void run_range(const FusedParams& p, const Hyper& h, long long lo, long long hi, float& l2) {
  constexpr int kUnroll = 4;
  constexpr int kPre = PRE ? kUnroll : 1;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const bool has_lp = p.lp_end > p.lp_begin;
  for (long long v0 = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; v0 < hi; v0 += stride * kUnroll) {
    float4 g[kUnroll], pw[kPre], pm[kPre], pv[kPre];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const long long v = v0 + u * stride;
      g[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (PRE) pw[u % kPre] = pm[u % kPre] = pv[u % kPre] = g[u];
      if (v < hi) {
        const long long i = p.offset + (v << 2);
        if (PRE && p.optimizer >= 0) {
          pw[u % kPre] = *reinterpret_cast<const float4*>(p.params[p.rank] + i);
          if (p.optimizer >= 1) pm[u % kPre] = *reinterpret_cast<const float4*>(p.slot_m + i);
          if (p.optimizer == 2) pv[u % kPre] = *reinterpret_cast<const float4*>(p.slot_v + i);
        }
        if (KIND == kLocal) {
          g[u] = (has_lp && low_precision(p, i)) ? ld_peer_bf16x4(p.grads_lp[0] + i)
                                                 : *reinterpret_cast<const float4*>(p.grads[0] + i);
        } else if ((KIND == kNvls || KIND == kTwoShot) && p.n_prered > 0 && prereduced(p, i)) {
          g[u] = ld_peer_f4(p.grads[p.rank] + i);            // summed in place by the producers' epilogues
          *reinterpret_cast<float4*>(const_cast<float*>(p.grads[p.rank]) + i) = make_float4(0.f, 0.f, 0.f, 0.f);
        } else if (has_lp && low_precision(p, i)) {          // bf16 on the wire, fp32 from here on
          if (KIND == kNvls) {
            g[u] = multimem_ld_reduce_bf16x4(p.mc_grads_lp + i);
          } else {
            for (int r = 0; r < p.world; ++r) {
              const float4 t = ld_peer_bf16x4(p.grads_lp[r] + i);
              g[u].x += t.x; g[u].y += t.y; g[u].z += t.z; g[u].w += t.w;
            }
          }
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
      const float4 wn = apply_update<PRE>(p, h, i, gs, l2, pw[u % kPre], pm[u % kPre], pv[u % kPre]);
      const uint2 wb = to_bf16x4(wn);
      if (KIND == kLocal || KIND == kOneShot) {
        *reinterpret_cast<float4*>(p.params[p.rank] + i) = wn;
        if (p.shadow[p.rank]) *reinterpret_cast<uint2*>(p.shadow[p.rank] + i) = wb;
      } else if (KIND == kNvls) {
        // sharded master: my fp32 copy is the only current one, peers get the bf16 shadow
        if (master_sharded(p, i)) *reinterpret_cast<float4*>(p.params[p.rank] + i) = wn;
        else multimem_st_f4(p.mc_params + i, wn);
        if (p.mc_shadow) multimem_st_bf16x4(p.mc_shadow + i, wb);
      } else {
        const bool sharded = master_sharded(p, i);
        for (int r = 0; r < p.world; ++r) {
          if (p.dst_mask != 0u && !((p.dst_mask >> r) & 1u)) continue;
          if (!sharded || r == p.rank) *reinterpret_cast<float4*>(p.params[r] + i) = wn;
          if (p.shadow[r]) *reinterpret_cast<uint2*>(p.shadow[r] + i) = wb;
        }
      }
    }
  }
}

// The push bucket, ONE vector per thread per iteration: store my gradient (as LL words) into my row on every peer,
// request my own parameter state, then poll the peers' words for this vector - all of them at once, eight peers per
// round trip, not one after the other (7 peers x 4 vectors polled in sequence cost 35 us on 8 GPUs) - and sum in rank
// order with my own gradient in its place.  Rows: [parity][source rank][push_stride / 2] uint4.
template <bool PRE>
__device__ __forceinline__ void push_update_range(const FusedParams& p, const Hyper& h, long long vec_total, int parity,
                                                  uint32_t tag, float& l2) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long rowq = p.push_stride >> 1;
  const uint4* recv = reinterpret_cast<const uint4*>(p.push_recv[p.rank]) + (long long)parity * p.world * rowq;
  const long long myrow = ((long long)parity * p.world + p.rank) * rowq;
  for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < vec_total; v += stride) {
    const long long i = p.offset + (v << 2);
    const float4 mine = *reinterpret_cast<const float4*>(p.grads[p.rank] + i);
    for (int r = 0; r < p.world; ++r)
      if (r != p.rank) {
        uint4* dst = reinterpret_cast<uint4*>(p.push_recv[r]) + myrow + (v << 1);
        st_ll(dst, mine.x, mine.y, tag);
        st_ll(dst + 1, mine.z, mine.w, tag);
      }
    float4 w = make_float4(0.f, 0.f, 0.f, 0.f), m = w, sv = w;
    if (p.optimizer >= 0) {
      w = *reinterpret_cast<const float4*>(p.params[p.rank] + i);
      if (p.optimizer >= 1) m = *reinterpret_cast<const float4*>(p.slot_m + i);
      if (p.optimizer == 2) sv = *reinterpret_cast<const float4*>(p.slot_v + i);
    }
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    const long long t0 = clock64();
    for (int r0 = 0; r0 < p.world; r0 += 8) {
      uint4 a[8], b[8];
      bool ok;
      do {
        ok = true;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int r = r0 + k;
          if (r < p.world && r != p.rank) {
            const uint4* src = recv + (long long)r * rowq + (v << 1);
            a[k] = ld_ll(src);
            b[k] = ld_ll(src + 1);
          }
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int r = r0 + k;
          if (r < p.world && r != p.rank)
            ok = ok && a[k].y == tag && a[k].w == tag && b[k].y == tag && b[k].w == tag;
        }
        if (!ok && clock64() - t0 > kSpinLimit) {
          atomicExch(p.status, MTB_E_TIMEOUT);
          break;
        }
      } while (!ok);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int r = r0 + k;
        if (r < p.world) {
          float4 t = mine;
          if (r != p.rank)
            t = make_float4(__uint_as_float(a[k].x), __uint_as_float(a[k].z), __uint_as_float(b[k].x),
                            __uint_as_float(b[k].z));
          acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
        }
      }
    }
    acc.x *= p.scale; acc.y *= p.scale; acc.z *= p.scale; acc.w *= p.scale;
    const float4 wn = apply_update<true>(p, h, i, acc, l2, w, m, sv);
    *reinterpret_cast<float4*>(p.params[p.rank] + i) = wn;
    if (p.shadow[p.rank]) *reinterpret_cast<uint2*>(p.shadow[p.rank] + i) = to_bf16x4(wn);
  }
}

__device__ __forceinline__ unsigned long long tagged(uint32_t tag, float v) {
  return (static_cast<unsigned long long>(tag) << 32) | __float_as_uint(v);
}

template <bool PRE>
__device__ __forceinline__ void fused_body(const FusedParams& p) {
  const int b = blockIdx.x;
  const int fb = p.chan_base + b;                    // this block's flag / sequence slot
  __shared__ uint32_t s_seq;
  __shared__ float s_red[8];
  __shared__ int s_last;
  pdl_sync();
  if (threadIdx.x == 0) s_seq = (p.world > 1 && !p.no_flags) ? p.seq[fb] : 0u;
  __syncthreads();
  const uint32_t seq = s_seq;
  if (p.trace != nullptr && b == 0 && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    p.trace[0] = t;
  }
  const long long step = p.step_ptr ? *p.step_ptr : 0;
  // what the closing warp mails was final before this launch began: fetch it now, off the closing chain's critical path
  float loss_in = 0.f;
  int err_in = 0;
  if (threadIdx.x == 0 && p.loss_mail != nullptr) {
    loss_in = __ldcg(p.loss_src);
    if (p.err_src != nullptr) err_in = __ldcg(p.err_src);
  }
  const Hyper h = make_hyper(p, step);
  const bool multi = p.world > 1;
  const bool sliced = multi && (p.algo == 1 || p.algo == 2);
  const long long vec_total = p.count >> 2;
  float l2 = 0.f;
  uint32_t rounds = 0;

  if (!multi) {
    run_range<kLocal, PRE>(p, h, 0, vec_total, l2);
  } else if (p.algo == 3) {
    // no flag round at all: the pushed words carry the step's tag and the receiver polls them
    const int parity = (int)(step & 1);
    const uint32_t tag = static_cast<uint32_t>(step + 1);
    if constexpr (PRE) push_update_range<PRE>(p, h, vec_total, parity, tag, l2);      // (only built into the small-bucket kernel)
  } else {
    // every peer's gradients are complete (with pre-reduced ranges some of them were written into MY memory by
    // peers' kernels: the fence makes this rank's own remote additions globally visible before it signals)
    if (!p.no_flags) cross_gpu_barrier(p, fb, seq + 1, p.n_prered > 0);
    if (p.algo == 0) {
      run_range<kOneShot, PRE>(p, h, 0, vec_total, l2);
    } else {
      // two-shot / nvls: this rank owns one contiguous slice
      const long long per = (vec_total + p.world - 1) / p.world;
      const long long lo = per * p.rank < vec_total ? per * p.rank : vec_total;
      const long long hi = lo + per < vec_total ? lo + per : vec_total;
      if (p.algo == 2) run_range<kNvls, PRE>(p, h, lo, hi, l2);
      else run_range<kTwoShot, PRE>(p, h, lo, hi, l2);
    }
    // peers are done reading my grads (loads have retired: their values were consumed) and, in the
    // two-shot / nvls forms, done writing my params (needs the fence)
    if (!p.no_flags) {
      cross_gpu_barrier(p, fb, seq + 2, p.algo != 0);
      rounds = 2;
    }
  }

  const bool want_l2 = p.l2_blocks != nullptr;
  if (want_l2) {                                     // block partial, summed in warp order
    l2 = warp_sum(l2);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = l2;
    __syncthreads();
    if (threadIdx.x == 0) {
      float t = 0.f;
      for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_red[w];
      p.l2_blocks[b] = t;
    }
  }
  if (threadIdx.x == 0) {
    if (multi && rounds) p.seq[fb] = seq + rounds;
    s_last = 0;
    if (p.done_counter != nullptr) {
      // every block read the step counter at its start; the last one to finish closes the launch
      if (gridDim.x == 1) {
        s_last = 1;
      } else {
        __threadfence();
        if (atomicAdd(p.done_counter, 1u) == gridDim.x - 1) {
          *p.done_counter = 0u;
          s_last = 1;
        }
      }
    }
  }
  __syncthreads();
  if (!s_last || threadIdx.x >= 32) return;
  // The last block's first warp closes the launch.  Everything below is spread over its 32 lanes: a single thread
  // walking hundreds of block partials (or 24 peer slots) pays one L2 round trip per element, tens of microseconds.
  const int lane = threadIdx.x;
  __threadfence();
  if (p.trace != nullptr && lane == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    p.trace[1] = t;
  }
  const uint32_t tag = static_cast<uint32_t>(step + 1);
  float total = 0.f;
  if (want_l2) {
    float part = 0.f;
    for (int k = lane; k < (int)gridDim.x; k += 32) part += __ldcg(p.l2_blocks + k);   // fixed order: reproducible
    total = warp_sum(part);
    const unsigned long long word = tagged(tag, total);
    if (sliced) {                                    // I only walked my slice: every rank needs every partial
      if (lane < p.world)
        asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p.l2_slots[lane] + p.l2_row * kMaxRanks + p.rank),
                     "l"(word) : "memory");
    } else if (lane == 0) {
      asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p.l2_slots[p.rank] + p.l2_row * kMaxRanks), "l"(word)
                   : "memory");
    }
  }
  if (p.advance_step == nullptr) return;
  const long long next = step + 1;
  if (lane == 0) *p.advance_step = next;
  if (p.loss_mail == nullptr) return;
  float l2sum = 0.f;
  if (want_l2) {
    // partials of this step, mine and the slice owners'; each word carries its step, so a lane simply waits for its own
    const long long t0 = clock64();
#pragma unroll
    for (int j = 0; j < (kL2Rows * kMaxRanks) / 32; ++j) {
      const int e = lane + 32 * j, row = e / kMaxRanks, ent = e % kMaxRanks;
      if (row == p.l2_row && !sliced) {
        if (ent == 0) l2sum += total;                // my own complete partial: no need to read it back
      } else if (ent < p.l2_rows_n[row]) {
        const unsigned long long* src = p.l2_slots[p.rank] + e;
        unsigned long long w;
        while (true) {
          asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(w) : "l"(src) : "memory");
          if (static_cast<uint32_t>(w >> 32) == tag) break;
          if (clock64() - t0 > kSpinLimit) { atomicExch(p.status, MTB_E_TIMEOUT); break; }
        }
        l2sum += __uint_as_float(static_cast<uint32_t>(w));
      }
    }
    l2sum = warp_sum(l2sum);
  }
  if (lane == 0) {
    const float loss = fmaf(p.wd, l2sum, loss_in);
    // one 8-byte store = one PCIe write: the host sees the value and its step number together
    const long long slot = (next - 1) & (p.loss_ring - 1);
    if (p.aux_mail != nullptr && p.err_src != nullptr)
      asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p.aux_mail + slot),
                   "l"((static_cast<unsigned long long>(tag) << 32) | static_cast<uint32_t>(err_in)) : "memory");
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p.loss_mail + slot), "l"(tagged(tag, loss)) : "memory");
  }
}

// Two builds of the same body.  The bulk kernel is capped at 64 registers so that four 256-thread blocks (or many
// narrow ones next to a convolution CTA) share an SM: a 13 MB bucket is bandwidth-delay bound and wants threads.  The
// small-bucket kernel runs one pass per thread and is pure latency: it loads the parameter state with the gradients.
__global__ void __launch_bounds__(256, 4) fused_allreduce_update_kernel(const FusedParams p) { fused_body<false>(p); }
__global__ void __launch_bounds__(256, 2) fused_small_bucket_kernel(const FusedParams p) { fused_body<true>(p); }

// The stand-alone barrier has a flag block of its own (the last one), so it never shares a sequence number with a
// fused launch that may be in flight on another stream.
constexpr int kBarrierBlock = kMaxBlocks - 1;

__global__ void barrier_kernel(const FusedParams p) {
  __shared__ uint32_t s_seq;
  if (threadIdx.x == 0) s_seq = p.seq[kBarrierBlock];
  __syncthreads();
  const uint32_t seq = s_seq;
  cross_gpu_barrier(p, kBarrierBlock, seq + 2, true);
  if (threadIdx.x == 0) p.seq[kBarrierBlock] = seq + 2;
}

__global__ void advance_step_kernel(long long* step) { *step += 1; }

}  // namespace

extern "C" {

static int fill(const mtb_fused_args* a, FusedParams* p) {
  MTB_REQUIRE(a->world >= 1 && a->world <= kMaxRanks && a->rank >= 0 && a->rank < a->world, "fused: bad world/rank");
  MTB_REQUIRE(a->offset % 4 == 0 && a->count % 4 == 0 && a->count >= 0, "fused: offset/count must be multiples of 4");
  MTB_REQUIRE(a->n_ranges >= 0 && a->n_ranges <= kMaxRanges, "fused: too many decay ranges");
  MTB_REQUIRE(a->algo >= 0 && a->algo <= 3, "fused: algo must be 0 (one-shot), 1 (two-shot), 2 (nvls) or 3 (push)");
  MTB_REQUIRE(a->algo != 2 || (a->mc_grads && a->mc_params), "fused: nvls needs multicast pointers");
  MTB_REQUIRE(a->chan == 0 || a->chan == 1, "fused: chan must be 0 or 1");
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
  p->chan_base = a->chan ? kChan1Base : 0;
  p->no_flags = a->no_flags;
  p->trace = reinterpret_cast<unsigned long long*>(a->trace);
  MTB_REQUIRE(!a->no_flags || a->algo != 3, "fused: push carries its own flag round");
  p->push_stride = a->push_stride;
  for (int r = 0; r < kMaxRanks; ++r) {
    p->push_recv[r] = reinterpret_cast<float*>(a->push_recv[r]);
    p->grads_lp[r] = reinterpret_cast<const __nv_bfloat16*>(a->grads_lp[r]);
    p->l2_slots[r] = reinterpret_cast<unsigned long long*>(a->l2_slots[r]);
  }
  if (a->algo == 3 && a->world > 1) {
    MTB_REQUIRE(a->push_stride >= a->count && a->push_stride % 4 == 0, "fused: push_stride must cover the bucket");
    for (int r = 0; r < a->world; ++r) MTB_REQUIRE(a->push_recv[r] != nullptr, "fused: push needs every rank's receive area");
    MTB_REQUIRE(a->n_prered == 0 && a->lp_end <= a->lp_begin, "fused: push takes plain fp32 gradients");
  }
  p->mc_grads_lp = reinterpret_cast<const __nv_bfloat16*>(a->mc_grads_lp);
  p->lp_begin = a->lp_begin; p->lp_end = a->lp_end;
  if (a->lp_end > a->lp_begin) {
    MTB_REQUIRE(a->lp_begin % 4 == 0 && a->lp_end % 4 == 0, "fused: bf16-wire range must be a multiple of 4");
    for (int r = 0; r < a->world; ++r) MTB_REQUIRE(a->grads_lp[r] != nullptr, "fused: bf16-wire range without bf16 buffers");
    MTB_REQUIRE(a->algo != 2 || a->world == 1 || a->mc_grads_lp != nullptr, "fused: nvls with a bf16-wire range needs its multicast view");
  }
  p->dst_mask = a->dst_mask;
  MTB_REQUIRE(a->dst_mask == 0u || a->algo == 1, "fused: dst_mask is a two-shot option");
  p->l2_blocks = reinterpret_cast<float*>(a->l2_blocks);
  p->l2_row = a->l2_row;
  for (int r = 0; r < kL2Rows; ++r) p->l2_rows_n[r] = a->l2_rows_n[r];
  if (p->l2_blocks != nullptr) {
    MTB_REQUIRE(a->l2_row >= 0 && a->l2_row < kL2Rows, "fused: l2_row out of range");
    MTB_REQUIRE(p->done_counter != nullptr && p->step_ptr != nullptr, "fused: the L2 value needs done_counter and step_ptr");
    for (int r = 0; r < a->world; ++r) MTB_REQUIRE(a->l2_slots[r] != nullptr, "fused: the L2 value needs every rank's slots");
    for (int r = 0; r < kL2Rows; ++r)
      MTB_REQUIRE(a->l2_rows_n[r] >= 0 && a->l2_rows_n[r] <= a->world, "fused: l2_rows_n out of range");
  }
  p->err_src = reinterpret_cast<const int*>(a->err_src);
  p->aux_mail = reinterpret_cast<unsigned long long*>(a->aux_mail);
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
  if (p.count == 0) return 0;
  // 256-thread blocks without shared memory fit next to the persistent tensor-core kernels, so
  // a bucket's update can run on a second stream underneath the rest of backward
  int block = a->block > 0 ? (a->block > 256 ? 256 : a->block) : 256;
  long long vecs = p.count >> 2;
  if (p.world > 1 && (p.algo == 1 || p.algo == 2)) vecs = (vecs + p.world - 1) / p.world;     // owner slices
  if (block < 32) block = 32;
  block = (block + 31) / 32 * 32;
  // one pass = 4 vectors (64 bytes) per thread; another block (and its flag round) pays off after that
  const int per_thread = (p.world > 1 && p.algo == 3) ? 1 : 4;      // push polls all peers per vector: one vector each
  long long need = (vecs + per_thread * block - 1) / (per_thread * block);
  if (need < 1) need = 1;
  // grid < 0: one pass per block ("short-lived blocks": a launch that runs UNDER other kernels must not park blocks
  // on an SM for its whole duration; leftover register space is handed back every few microseconds)
  int grid = a->grid > 0 ? a->grid : (a->grid < 0 ? kMaxL2Blocks : 3 * 148);
  if (grid > need) grid = (int)need;
  const bool flags = p.world > 1 && !p.no_flags;
  if (flags && grid > (a->chan ? kChan1Blocks : kChan1Base)) grid = a->chan ? kChan1Blocks : kChan1Base;
  if (grid > kMaxL2Blocks) grid = kMaxL2Blocks;
  const bool small = (p.world > 1 && p.algo == 3) ||
                     (p.count * 4 <= (1ll << 20) && grid * block * 4ll >= vecs);       // at most one pass per thread
  MTB_CUDA_CHECK(mtb::launch(small ? fused_small_bucket_kernel : fused_allreduce_update_kernel, dim3(grid), dim3(block), 0,
                             stream, p));
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
int mtb_fused_args_size(void) { return (int)sizeof(mtb_fused_args); }
int mtb_fused_max_l2_blocks(void) { return kMaxL2Blocks; }
int mtb_fused_flag_bytes(void) { return kMaxBlocks * kMaxRanks * (int)sizeof(uint32_t); }

}  // extern "C"
