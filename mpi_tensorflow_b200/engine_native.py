"""GPU engine: the whole train step on hand-written sm_100a kernels.

Replaces the TensorFlow session (``tf.Session`` + ``sess.run``, ``mpipy.py:72-73,85``),
its autodiff (``minimize``, ``mpipy.py:65``) and its allocator (``mpipy.py:69-71``):

* a static set of buffers sized once from the config (all shapes are static,
  ``mpipy.py:33-37``) — parameters / gradients / bf16 shadow weights live in
  symmetric, peer-mapped memory when the world has more than one rank;
* forward and backward wired kernel by kernel (no autograd on the hot path);
* the fused allreduce + optimizer kernel K1 as the only synchronisation;
* the step captured in a CUDA graph: one launch per step, no per-step H2D, the
  step counter / LR schedule / dropout stream all live on the device.

Precision: ``dtype="bf16"`` - bf16 activations and bf16 tensor-core operands, fp32 accumulation,
fp32 master weights, optimizer state and gradients; ``dtype="fp32"`` - fp32 everywhere, the
GEMM-shaped products as TF32 tensor-core MMAs (the generic kernel set, ``*_f32`` entry points).
"""
from __future__ import annotations

import ctypes as C
import os
import time
from typing import Dict, List, Optional

import numpy as np
import torch

from .config import TrainConfig
from .models.spec import ModelSpec, flat_size
from .ops import native as N
from .parallel.comm import Communicator

_OPT = {"sgd": 0, "momentum": 1, "adam": 2}


class NativeEngine:
    name = "native-sm100a"

    def __init__(self, spec: ModelSpec, cfg: TrainConfig, comm: Communicator, device: torch.device,
                 conv_impl: str = "auto"):
        if device.type != "cuda":
            raise RuntimeError("NativeEngine needs a CUDA device")
        N.lib()                                   # fail loudly if the extension is missing
        if cfg.comm_dtype not in ("fp32", "bf16"):
            raise ValueError(f"comm_dtype must be 'fp32' or 'bf16', got {cfg.comm_dtype!r}")
        if cfg.dtype not in ("bf16", "fp32"):
            raise ValueError(f"dtype must be 'bf16' or 'fp32', got {cfg.dtype!r}")
        if cfg.comm_dtype == "bf16" and cfg.dtype != "bf16":
            raise ValueError("comm_dtype='bf16' (bf16 gradients on the wire) needs dtype='bf16'")
        # fp32: every activation, weight operand and gradient is fp32; the GEMM-shaped products run as TF32
        # tensor-core MMAs with fp32 accumulation, conv1 (K = 25) and the loss head on fp32 CUDA cores.
        self.fp32 = cfg.dtype == "fp32"
        self.spec, self.cfg, self.comm, self.device = spec, cfg, comm, device
        torch.cuda.set_device(device)
        self.B = cfg.batch
        self.n = flat_size(spec)
        self.p2p = comm if getattr(comm, "backend", "") == "p2p" else None
        self.world = comm.size if self.p2p is not None else 1
        self.conv_impl = conv_impl
        # the fused step is valid only when K1 itself reaches every rank of the world
        self.can_fuse = comm.size == 1 or self.p2p is not None
        if spec.image_size != 28 or spec.c1 % 32 or spec.c2 % 8 or spec.fc % 8:
            raise ValueError("NativeEngine supports 28x28 inputs, c1 % 32 == 0, c2 % 8 == 0, fc % 8 == 0")
        self._alloc_state()
        self._alloc_activations(max(self.B, cfg.eval_batch))
        # GEMM -> reduce-scatter for the FC1 weight gradient (cfg.fuse_wgrad_rs): needs peers, an owner-sliced K1
        # over the whole buffer (so the GEMM and K1 agree on who owns what) and the bf16 GEMM engine
        self._rs_mode = (bool(cfg.fuse_wgrad_rs) and self.p2p is not None and self.world > 1 and not self.fp32
                         and cfg.comm_algo not in ("oneshot", "push") and not self._lp)
        # bucketed overlap (default): the fc bucket's fused allreduce+update runs on the comm stream under the
        # convolution backward; the reduce-scatter GEMM needs the single whole-buffer launch instead
        # (a lone GPU has no flag round or NVLink latency to hide, but the 72 MB of the update itself still hide under
        # the convolutions: 213.6 vs 219.3 us per step at batch 1024, 111.3 vs 113.4 at batch 100; MTB_OVERLAP_N1=0
        # gives the single launch back)
        self._overlap = (bool(cfg.overlap) and not self._rs_mode
                         and (self.world > 1 or os.environ.get("MTB_OVERLAP_N1", "1") != "0"))
        self._rs_active = False                  # True only inside the fused step body
        self._rs_dirty = True                    # owners' fc1 slices must be zero when a fused step starts
        if self._rs_mode:
            fc1w = spec.param("fc1_weight")
            per = -(-(self.n // 4) // self.world) * 4
            self._rs = N.GemmRS()
            self._rs.world, self._rs.rank, self._rs.base, self._rs.per = self.world, self.p2p.rank, fc1w.offset, per
            self._rs.flags = 1 if os.environ.get("MTB_RS_STAGED", "1") == "0" else 0
            for r in range(self.world):
                self._rs.out[r] = self._grad_ptrs[r]
        self.init_params(cfg.seed)
        self.opt_step = 0
        self._graph: Optional[torch.cuda.CUDAGraph] = None
        self._graph_key = None
        self._graphs: Dict[str, torch.cuda.CUDAGraph] = {}
        self._stream_state = None
        self.data_x = self.data_y = None
        self.rows = 0
        self.decay_rows = cfg.lr_decay_rows
        # The step's own kernels run (and are captured) at high priority and the hidden bucket's launches at normal
        # priority: when SM resources free up the next convolution CTA is placed first and the update's blocks take
        # what is left over.
        self._side = torch.cuda.Stream(device=device, priority=-1)
        self._comm_stream = torch.cuda.Stream(device=device, priority=int(os.environ.get("MTB_COMM_PRIO", "0")))
        self._ev_fc, self._ev_c2, self._ev_done = torch.cuda.Event(), torch.cuda.Event(), torch.cuda.Event()
        self._eval_cache: Dict = {}
        self._marks = None                       # list of (name, event) while a step is being phase-timed
        self._nvtx = os.environ.get("MTB_NVTX") == "1"
        self.last_aux: Dict = {}

    # ------------------------------------------------------------------ buffers
    def _sym(self, nbytes: int):
        """Symmetric (peer-mapped) allocation under p2p, plain device memory otherwise."""
        if self.p2p is not None:
            sb = self.p2p.alloc_symmetric(nbytes)
            self._sym_keep.append(sb)
            return sb.local, sb.ptrs, sb.mc_ptr
        t = torch.zeros(nbytes, dtype=torch.uint8, device=self.device)
        return t, [t.data_ptr()], 0

    def _alloc_state(self):
        n, dev = self.n, self.device
        self._sym_keep = []
        raw, self._param_ptrs, self._param_mc = self._sym(n * 4)
        self.flat_params = raw[:n * 4].view(torch.float32)
        raw, self._grad_ptrs, self._grad_mc = self._sym(n * 4)
        self.flat_grads = raw[:n * 4].view(torch.float32)
        raw, self._shadow_ptrs, self._shadow_mc = self._sym(n * 2)
        self.shadow = raw[:n * 2].view(torch.bfloat16)
        self.has_nvls = bool(self._param_mc and self._grad_mc and self._shadow_mc)
        multi = self.p2p is not None and self.world > 1
        # bf16 gradient wire format (cfg.comm_dtype="bf16"): the FC1 weight gradient (98 % of the bytes) is stored
        # as bf16 by its GEMM's epilogue into a second symmetric bucket and crosses NVLink at 2 bytes per element
        self._lp = self.cfg.comm_dtype == "bf16"
        self.flat_grads_lp = None
        if self._lp:
            raw, self._grad_lp_ptrs, self._grad_lp_mc = self._sym(n * 2)
            self.flat_grads_lp = raw[:n * 2].view(torch.bfloat16)
            self.has_nvls = self.has_nvls and bool(self._grad_lp_mc)
        # optimizer slots are peer-mapped only when the sync strategy averages them (weight_avg + include_momentum)
        sym_slots = multi and self.cfg.sync == "weight_avg" and self.cfg.sync_include_momentum
        self.slots: List[torch.Tensor] = []
        self._slot_ptrs, self._slot_mc = [], []
        n_slots = {"momentum": 1, "adam": 2}.get(self.cfg.optimizer, 0)
        for _ in range(n_slots):
            if sym_slots:
                raw, ptrs, mc = self._sym(n * 4)
                self.slots.append(raw[:n * 4].view(torch.float32))
                self._slot_ptrs.append(ptrs)
                self._slot_mc.append(mc)
            else:
                self.slots.append(torch.zeros(n, device=dev))
        # receive rows of the "push" tail bucket: [2 parity][world][stride] (value, tag) pairs on every rank
        self._tail_hi = self.spec.param("fc1_weight").offset
        self._push_stride = (self._tail_hi + 3) // 4 * 4
        self._push = self.p2p.alloc_symmetric(2 * self.world * self._push_stride * 8) if multi else None
        self._sliced: Dict = {}                  # (lo, hi) of every owner-sliced peer launch -> algo
        self._l2_blocks = torch.zeros(2, N.lib().mtb_fused_max_l2_blocks(), device=dev)
        self._trace = torch.zeros(3, 2, dtype=torch.int64, device=dev) if os.environ.get("MTB_K1_TRACE") == "1" else None
        self._l2_slots = torch.zeros(4 * 16, dtype=torch.int64, device=dev)          # world-1 stand-in for the p2p slots
        self.dev_step = torch.zeros(1, dtype=torch.int64, device=dev)
        self.loss_acc = torch.zeros(1, device=dev)
        self.err_acc = torch.zeros(1, dtype=torch.int32, device=dev)
        self._zero_args = None
        self._scratch = {}
        self._masters_dirty = self._slots_dirty = False
        # fp32 path: consumers read the fp32 master itself, so every rank needs all of it
        self._shard_master = (bool(self.cfg.shard_master) and self.p2p is not None and self.world > 1
                              and not self.fp32)
        self._done = torch.zeros(2, dtype=torch.int32, device=dev)                  # one per channel
        # loss mailbox: K1's last block stores {step, loss} words straight into this pinned host ring, and the
        # batch's error count into a second one
        self._mail = torch.zeros(self._MAIL_RING, dtype=torch.int64).pin_memory()
        self._mail_np = self._mail.numpy()
        self._aux = torch.zeros(self._MAIL_RING, dtype=torch.int64).pin_memory()
        self._aux_np = self._aux.numpy()
        self._seq = torch.zeros(N.lib().mtb_fused_max_blocks(), dtype=torch.int32, device=dev)
        self._status = torch.zeros(1, dtype=torch.int32, device=dev)
        self.P = {p.key: self.flat_params[p.offset:p.offset + p.numel].view(p.shape) for p in self.spec.params}
        self.G = {p.key: self.flat_grads[p.offset:p.offset + p.numel].view(p.shape) for p in self.spec.params}
        self.Glp = ({p.key: self.flat_grads_lp[p.offset:p.offset + p.numel].view(p.shape) for p in self.spec.params}
                    if self._lp else None)
        self.S = {p.key: self.shadow[p.offset:p.offset + p.numel].view(p.shape) for p in self.spec.params}
        # ranges that must start each step at zero because their wgrad kernels accumulate
        fc1w = self.spec.param("fc1_weight")
        c2w = self.spec.param("conv2_weight")
        self._zero_ranges = [(0, c2w.offset), (c2w.offset + c2w.numel, fc1w.offset),
                             (fc1w.offset + fc1w.numel, n)]

    def _alloc_activations(self, Bmax: int):
        s, dev = self.spec, self.device
        f32 = torch.float32
        bf = f32 if self.fp32 else torch.bfloat16                       # activation dtype
        z = lambda *shape, dtype=bf: torch.zeros(*shape, dtype=dtype, device=dev)  # noqa: E731
        self.Bmax = Bmax
        self.x_in = z(Bmax, 784, dtype=f32)
        self.y_in = z(Bmax, dtype=torch.int32)
        self.a1p, self.idx1 = z(Bmax, 14, 14, s.c1), z(Bmax, 14, 14, s.c1, dtype=torch.uint8)
        self.a2p, self.idx2 = z(Bmax, 7, 7, s.c2), z(Bmax, 7, 7, s.c2, dtype=torch.uint8)
        self.h = z(Bmax, s.fc)
        self.logits, self.probs = z(Bmax, s.num_classes, dtype=f32), z(Bmax, s.num_classes, dtype=f32)
        self.dlogits = z(Bmax, s.num_classes, dtype=f32)
        self.dh, self.dflat = z(Bmax, s.fc), z(Bmax, s.flat_dim)
        self.da2 = z(Bmax, 14, 14, s.c2)
        self.dp1 = z(Bmax, 14, 14, s.c1)
        # conv1 on tensor cores when it has 32 output channels, else the CUDA-core kernel
        self.conv1_tc = s.c1 == 32 and self.conv_impl != "im2col" and not self.fp32
        self._fast_head = s.num_classes == 10 and s.fc % 128 == 0 and s.fc * 40 <= 170 * 1024
        # the generic layer kernels exist for both activation types; GEMMs are bf16 or TF32
        self._sfx = "_f32" if self.fp32 else ""
        self._gemm = N.gemm_tf32 if self.fp32 else N.gemm_bf16
        # batch window fused into the kernels that read the batch (needs both fast paths)
        self._can_window = self.conv1_tc and self._fast_head
        self._windowed = False
        self._stream_buf = None
        self._stream_u8 = False
        # conv2: implicit GEMM (no im2col) when the shape has a shift-GEMM kernel, else im2col + GEMM
        want = self.conv_impl
        have = bool(N.lib().mtb_conv2_supported(14, s.c1, s.c2)) and not self.fp32
        self.conv2_implicit = have and want != "im2col"
        # every other shape whose channel counts fill whole 128-byte rows: implicit GEMM through the GEMM engine's
        # convolution producer (4-D TMA over the NHWC tensor), still no im2col.  What is left for im2col + GEMM is
        # conv_impl="im2col" (kept as a second implementation for the tests) and odd channel counts.
        row_e = 32 if self.fp32 else 64
        self.conv2_igemm = (not self.conv2_implicit and want != "im2col" and s.c1 % row_e == 0 and s.c2 % row_e == 0
                            and (s.c2 <= 64 or s.c2 % 128 == 0) and (s.c1 <= 64 or s.c1 % 128 == 0))
        if want == "implicit" and not (have or self.conv2_igemm):
            raise ValueError(f"no implicit-GEMM conv2 kernel for {s.c1}->{s.c2} channels")
        if self.conv2_implicit:
            self.c2_ws = torch.zeros(N.lib().mtb_conv2_wgrad_workspace_bytes(Bmax), dtype=torch.uint8, device=dev)
        else:
            self.a2 = z(Bmax * 196, s.c2)
            bn = 256 if s.c2 % 256 == 0 else (128 if s.c2 % 128 == 0 else 64)
            if self.conv2_igemm:
                # split-K over the batch so that the weight gradient covers the chip
                tiles = -(-25 * s.c1 // 128) * -(-s.c2 // bn)
                kb_total = self.B * (7 if self.fp32 else 4)
                self.wg_splits = max(1, min(kb_total, 148 // tiles))
            else:
                self.col = z(Bmax * 196, 25 * s.c1)
                self.dcol = z(Bmax * 196, 25 * s.c1)
                bn = 64
                self.wg_splits = 16        # 21 splits (one full wave at 32->64 channels) measured slower: 43.7 vs 38.2 us
            # split-K scratch for the conv2 weight gradient (K = B*196 is the long axis)
            ws, cn = N.gemm_scratch_sizes(25 * s.c1, s.c2, bn, self.wg_splits)
            self.wg_ws = torch.zeros(max(ws, 16), dtype=torch.uint8, device=dev)
            self.wg_cnt = torch.zeros(max(cn, 16), dtype=torch.uint8, device=dev)

    # ------------------------------------------------------------------ parameters
    def init_params(self, seed: int):
        """Seeded init, identical on every rank (``mpipy.py:38-53``), on the device (K10)."""
        for i, p in enumerate(self.spec.params):
            dst = self.flat_params[p.offset:p.offset + p.numel]
            if p.init == "trunc_normal":
                N.call("mtb_init_trunc_normal", dst.data_ptr(), p.numel, p.init_value, seed, i)
            else:
                N.call("mtb_fill_f32", dst.data_ptr(), p.numel, p.init_value if p.init == "const" else 0.0)
        self.params_changed()

    def params_changed(self):
        """Refresh the bf16 shadow after the fp32 master was written from outside K1."""
        if self.fp32:
            return                                # no shadow: the kernels read the master
        N.call("mtb_cast_f32_to_bf16", self.flat_params.data_ptr(), self.shadow.data_ptr(), self.n)

    def synchronize(self):
        torch.cuda.current_stream().synchronize()

    def _reset_push(self):
        """Clear the push bucket's receive rows: their words are tagged with the step number, so whenever the step
        counter moves backwards (checkpoint restore, the rolled-back warm-up step before a graph capture) a stale
        row could pass for the new step's.  Callers put a barrier between this and the next step."""
        if self._push is not None:
            self._push.local.zero_()
            torch.cuda.synchronize()

    # -- state that the two-shot / nvls update keeps only with the owner of each slice ------------
    @staticmethod
    def owned_slice(lo: int, hi: int, rank: int, world: int):
        """Rank ``rank``'s slice of an owner-sliced K1 launch over elements [lo, hi): the kernel's own arithmetic
        (``allreduce_fused.cu``: ``per = ceil(vectors / world)`` vectors of 4 elements per rank)."""
        vecs = (hi - lo) // 4
        per = (vecs + world - 1) // world
        olo = lo + min(per * rank, vecs) * 4
        return olo, min(olo + per * 4, hi)

    @staticmethod
    def contribution_pieces(rank: int, world: int, sliced, lo: int, hi: int):
        """Which parts of [lo, hi) rank ``rank`` contributes when a buffer is completed across ranks: inside each
        owner-sliced bucket of ``sliced`` (a collection of (blo, bhi)) its own slice; everything outside those buckets
        (updated in full by every rank) comes from rank 0.  Over all ranks the pieces tile [lo, hi) exactly once."""
        out = []
        buckets = sorted(sliced)
        for blo, bhi in buckets:
            olo, ohi = NativeEngine.owned_slice(blo, bhi, rank, world)
            a, b = max(lo, olo), min(hi, ohi)
            if b > a:
                out.append((a, b))
        if rank == 0:
            cur = lo
            for blo, bhi in buckets + [(hi, hi)]:
                a, b = max(cur, lo), min(blo, hi)
                if b > a:
                    out.append((a, b))
                cur = max(cur, bhi)
        return sorted(out)

    def _owned_ranges(self):
        """Element ranges of the flat buffers whose optimizer state (and sharded fp32 master) only THIS rank
        maintains: its slice of every owner-sliced K1 launch made so far.  With bucketed launches ownership is per
        bucket, not a slice of the whole buffer."""
        return [self.owned_slice(lo, hi, self.p2p.rank, self.world) for (lo, hi) in sorted(self._sliced)]

    def _gather_owned(self, buf: torch.Tensor, ranges):
        """Make ``buf`` complete on every rank: every element has exactly one contributing rank
        (:meth:`contribution_pieces`) and the parts are summed.  Collective; used at checkpoint / inspection time."""
        for lo, hi in ranges:
            seg = torch.zeros(hi - lo, device=self.device)
            for a, b in self.contribution_pieces(self.p2p.rank, self.world, list(self._sliced), lo, hi):
                seg[a - lo:b - lo] = buf[a:b]
            self.p2p.allreduce(seg, average=False)
            buf[lo:hi] = seg

    def sync_state(self):
        """COLLECTIVE.  After sharded-master / sharded-slot updates, complete the fp32 master weights
        and the optimizer slots on every rank (no-op otherwise).  Call on all ranks before
        ``get_params`` / ``get_slots`` are used on any of them."""
        if self.p2p is None or self.world == 1:
            return
        self.synchronize()
        flags = self.p2p.boot.allgather_object((self._masters_dirty, self._slots_dirty))
        if any(f[0] for f in flags):
            big = [self.spec.param("conv2_weight"), self.spec.param("fc1_weight")]
            self._gather_owned(self.flat_params, [(p.offset, p.offset + p.numel) for p in big])
            self._masters_dirty = False
        if any(f[1] for f in flags):
            for sl in self.slots:
                self._gather_owned(sl, [(0, self.n)])
            self._slots_dirty = False
        self.synchronize()

    def get_params(self):
        self.synchronize()
        return {k: v.detach().cpu().numpy().copy() for k, v in self.P.items()}

    def get_slots(self):
        self.synchronize()
        names = ["m", "v"]
        return {p.key: {names[i]: s[p.offset:p.offset + p.numel].view(p.shape).cpu().numpy().copy()
                        for i, s in enumerate(self.slots)} for p in self.spec.params}

    def set_state(self, params, slots, step: int):
        names = ["m", "v"]
        for p in self.spec.params:
            self.P[p.key].copy_(torch.from_numpy(np.asarray(params[p.key], dtype=np.float32).reshape(p.shape)))
            for i, s in enumerate(self.slots):
                if slots and p.key in slots and names[i] in slots[p.key]:
                    s[p.offset:p.offset + p.numel].copy_(
                        torch.from_numpy(np.asarray(slots[p.key][names[i]], dtype=np.float32).reshape(-1)))
        self.dev_step.fill_(step)
        self.opt_step = step
        self.params_changed()
        self.synchronize()
        self._mail.zero_()                        # a rewound counter must not match losses mailed before
        self._aux.zero_()
        self._reset_push()

    # ------------------------------------------------------------------ data
    def attach_dataset(self, shards):
        """Keep this rank's train shard resident in HBM (fp32 [rows,784] + int32 labels): the
        reference re-feeds every batch from host memory (``mpipy.py:83-85``)."""
        x = torch.from_numpy(np.ascontiguousarray(shards.train_x.reshape(shards.train_x.shape[0], -1)))
        self.data_x = x.to(self.device, dtype=torch.float32)
        self.data_y = torch.from_numpy(np.ascontiguousarray(shards.train_y)).to(self.device, dtype=torch.int32)
        self.rows = self.data_x.shape[0]
        self._graphs.pop("window", None)         # that graph holds the old shard's address
        if not self.decay_rows:
            self.decay_rows = self.rows

    def attach_device_dataset(self, x: torch.Tensor, y: torch.Tensor):
        self.data_x, self.data_y, self.rows = x, y, x.shape[0]
        self._graphs.pop("window", None)
        if not self.decay_rows:
            self.decay_rows = self.rows

    def resident_batch(self, sampler, step: int):
        return None, None                       # fused_step pulls the window on the device

    def stage_batch(self, images, labels):
        """Copy a caller-supplied batch into the static input buffers (H2D if on the host)."""
        B = self.B
        if isinstance(images, np.ndarray):
            images = torch.from_numpy(images)
        if isinstance(labels, np.ndarray):
            labels = torch.from_numpy(labels)
        self.x_in[:B].copy_(images.reshape(B, -1), non_blocking=True)
        self.y_in[:B].copy_(labels.reshape(B), non_blocking=True)

    # ------------------------------------------------------------------ forward
    def _win(self):
        """(images, labels, step pointer, rows) of the fused device-side batch window, or the
        staged batch with a null step pointer."""
        if self._windowed:
            return self.data_x.data_ptr(), self.data_y.data_ptr(), self.dev_step.data_ptr(), self.rows
        if self._stream_buf is not None:         # host-streamed training: read the slot of the input ring directly
            st = self._stream_state
            return st["x"][self._stream_buf].data_ptr(), st["y"][self._stream_buf].data_ptr(), None, 0
        return self.x_in.data_ptr(), self.y_in.data_ptr(), None, 0

    def _forward(self, B: int, train: bool, with_loss: bool, zero: bool = False):
        s, c = self.spec, self.cfg
        P = self.P
        S = self.P if self.fp32 else self.S       # weight operands of the GEMM-shaped products
        sfx, gemm = self._sfx, self._gemm
        keep = s.keep_prob if (train and c.dropout) else 1.0
        wx, wy, wstep, wrows = self._win()
        # conv1 + bias + ReLU + pool (mpipy.py:156-158)
        if self.conv1_tc and zero:
            # the step's first kernel also clears the accumulators the backward kernels add into
            ptrs, words, nz = self._zero_list()
            N.call("mtb_conv1_fwd_tc_win_zero", wx, P["conv1_weight"].data_ptr(), P["conv1_bias"].data_ptr(),
                   self.a1p.data_ptr(), self.idx1.data_ptr(), B, wstep, wrows, ptrs, words, nz)
        elif self.conv1_tc:
            N.call("mtb_conv1_fwd_tc_win", wx, P["conv1_weight"].data_ptr(),
                   P["conv1_bias"].data_ptr(), self.a1p.data_ptr(), self.idx1.data_ptr(), B, wstep, wrows)
        else:
            N.call("mtb_conv1_fwd" + sfx, wx, P["conv1_weight"].data_ptr(), P["conv1_bias"].data_ptr(),
                   self.a1p.data_ptr(), self.idx1.data_ptr(), B, s.c1)
        # conv2 + bias + ReLU + pool (mpipy.py:159-161)
        M2, K2 = B * 196, 25 * s.c1
        if self.conv2_implicit:
            N.call("mtb_conv2_fprop_pool", self.a1p.data_ptr(), S["conv2_weight"].data_ptr(),
                   P["conv2_bias"].data_ptr(), self.a2p.data_ptr(), self.idx2.data_ptr(), B)
        elif self.conv2_igemm:
            N.call("mtb_conv5x5_gemm", 0, self.a1p.data_ptr(), S["conv2_weight"].data_ptr(), self.a2.data_ptr(), B, s.c1,
                   s.c2, int(self.fp32), P["conv2_bias"].data_ptr(), 1, 1, None, None)
            N.call("mtb_maxpool2x2_fwd" + sfx, self.a2.data_ptr(), self.a2p.data_ptr(), self.idx2.data_ptr(), B, 14, 14,
                   s.c2)
        else:
            N.call("mtb_im2col_5x5" + sfx, self.a1p.data_ptr(), self.col.data_ptr(), B, 14, 14, s.c1)
            gemm(self.col, S["conv2_weight"].view(K2, s.c2), self.a2, M2, s.c2, K2, False, True,
                 bias=P["conv2_bias"], relu=True)
            N.call("mtb_maxpool2x2_fwd" + sfx, self.a2.data_ptr(), self.a2p.data_ptr(), self.idx2.data_ptr(), B, 14, 14,
                   s.c2)
        # FC1 + bias + ReLU + dropout in the GEMM epilogue (mpipy.py:165-166)
        flat = self.a2p.view(self.Bmax, s.flat_dim)
        bn, splits = 64 if B * s.fc < 96 * 128 * 128 else 128, 1
        tiles = -(-B // 128) * -(-s.fc // bn)
        if tiles < 64 and (self.fp32 or os.environ.get("MTB_FC_SPLITK") == "1"):
            # small batch: split K so the GEMM covers the chip.  With bf16 weights it measured slower at
            # batch 100 (the last-arriver reduction costs more than the idle SMs), so opt-in only; with
            # fp32 weights 16 CTAs stream 12.8 MB at 0.46 TB/s (28 us) and the split pays.
            splits = max(1, min(8, 128 // tiles))
        ws, cnt = self._splitk_scratch("fc1_fprop", B, s.fc, bn, splits)
        gemm(flat, S["fc1_weight"], self.h, B, s.fc, s.flat_dim, False, True, bias=P["fc1_bias"], relu=True,
             keep_prob=keep, seed=c.seed, step_ptr=self.dev_step, bn=bn, splits=splits, workspace=ws, counters=cnt)
        # FC2 + softmax + CE + dlogits (mpipy.py:167, :55-56, :67)
        head = (self.h.data_ptr(), P["fc2_weight"].data_ptr(), P["fc2_bias"].data_ptr(),
                wy if with_loss else None, self.logits.data_ptr(), self.probs.data_ptr(),
                self.dlogits.data_ptr() if with_loss else None, self.loss_acc.data_ptr() if with_loss else None,
                self.err_acc.data_ptr() if with_loss else None, B, s.fc, s.num_classes)
        if self._fast_head:
            N.call("mtb_fc2_softmax_ce_fast_win" + sfx, *head, wstep, wrows)
        else:
            N.call("mtb_fc2_softmax_ce" + sfx, *head)

    # ------------------------------------------------------------------ backward
    def _backward(self, B: int):
        self._backward_fc(B)
        self._backward_conv(B)

    def _backward_fc(self, B: int):
        """FC2 and FC1 backward.  When it returns, every gradient of the fc bucket is final and no
        later kernel of the step reads an FC weight, so the bucket's fused update may start."""
        s, c = self.spec, self.cfg
        P, G = self.P, self.G
        S = self.P if self.fp32 else self.S
        sfx, gemm = self._sfx, self._gemm
        scale = 1.0 / s.keep_prob if c.dropout else 1.0
        M2, K2 = B * 196, 25 * s.c1
        flat = self.a2p.view(self.Bmax, s.flat_dim)
        if s.num_classes == 10:
            # dW2, db2, dh (with the dropout / ReLU mask) and db of FC1 in one pass over h
            N.call("mtb_fc2_bwd_fused" + sfx, self.h.data_ptr(), self.dlogits.data_ptr(), P["fc2_weight"].data_ptr(),
                   self.dh.data_ptr(), G["fc2_weight"].data_ptr(), G["fc2_bias"].data_ptr(),
                   G["fc1_bias"].data_ptr(), B, s.fc, s.num_classes, scale)
        else:
            N.call("mtb_fc2_wgrad" + sfx, self.h.data_ptr(), self.dlogits.data_ptr(), G["fc2_weight"].data_ptr(),
                   G["fc2_bias"].data_ptr(), B, s.fc, s.num_classes)
            N.call("mtb_fc2_dgrad" + sfx, self.dlogits.data_ptr(), P["fc2_weight"].data_ptr(), self.h.data_ptr(),
                   self.dh.data_ptr(), B, s.fc, s.num_classes, scale)
            N.call("mtb_colsum_f32" if self.fp32 else "mtb_colsum_bf16", self.dh.data_ptr(), G["fc1_bias"].data_ptr(), B, s.fc)
        # dflat = dh W1^T : W1 [in,out] row-major is already the K-major B operand.  Runs BEFORE the
        # weight gradient so that the last reader of the FC1 weights is done when the bucket closes.
        # bf16 at large batch: 128-wide tiles in the build that puts two CTAs on an SM (bn 129): 200 tiles in one wave, one
        # CTA's epilogue under the other's main loop: 16.0 us vs 18.7 (256-wide) / 21.3 (128-wide, one per SM); cuBLAS 14.5
        bn = (129 if not self.fp32 else 256) if B >= 512 else 128
        tiles = -(-B // 128) * -(-s.flat_dim // min(bn, 256))
        # (split-K measured slower here in both precisions: fp32 batch 100, 4 splits, 25.2 vs 16.1 us)
        splits = max(1, min(4, 128 // tiles)) if (tiles < 64 and os.environ.get("MTB_FC_SPLITK") == "1") else 1
        ws, cnt = self._splitk_scratch("fc1_dgrad", B, s.flat_dim, bn, splits)
        gemm(self.dh, S["fc1_weight"], self.dflat, B, s.flat_dim, s.fc, False, False, bn=bn, splits=splits,
             workspace=ws, counters=cnt)
        # dW1 = flat^T dh : both operands MN-major, fp32 straight into the gradient bucket (K = batch: wide tiles
        # only pay off when K is long enough to amortise their epilogue)
        wbn = (129 if not self.fp32 else 256) if (s.fc % 256 == 0 and B >= 512) else 128       # 17.6 vs 19.7 us
        if self._rs_active:
            # ... added tile by tile into the slice owners' buckets over NVLink: the reduce-scatter half of the
            # gradient allreduce happens inside this GEMM, K1 only updates and broadcasts (prered range)
            N.gemm_bf16_rs(flat, self.dh, s.flat_dim, s.fc, B, True, True, s.fc, 256 if wbn == 129 else wbn, self._rs)
        elif self._lp:
            # bf16 wire format: the epilogue rounds once and K1 reads 2 bytes per element from every peer
            gemm(flat, self.dh, self.Glp["fc1_weight"], s.flat_dim, s.fc, B, True, True, bn=wbn)
        else:
            gemm(flat, self.dh, G["fc1_weight"], s.flat_dim, s.fc, B, True, True, bn=wbn)

    def _backward_conv(self, B: int, after_conv2_wgrad=None):
        s = self.spec
        P, G = self.P, self.G
        S = self.P if self.fp32 else self.S
        sfx, gemm = self._sfx, self._gemm
        M2, K2 = B * 196, 25 * s.c1
        if self.conv2_implicit:
            # wgrad (+ bias grad) and dgrad as shift-GEMMs; pool2 + ReLU backward (O19, O18) is rebuilt
            # inside both kernels from the pooled maps, so the un-pooled gradient never exists in HBM
            N.call("mtb_conv2_wgrad_unpool", self.a1p.data_ptr(), self.dflat.data_ptr(), self.a2p.data_ptr(),
                   self.idx2.data_ptr(), G["conv2_weight"].data_ptr(), G["conv2_bias"].data_ptr(),
                   self.c2_ws.data_ptr(), B)
            if after_conv2_wgrad is not None:
                after_conv2_wgrad()
            N.call("mtb_conv2_dgrad_unpool", self.dflat.data_ptr(), self.a2p.data_ptr(), self.idx2.data_ptr(),
                   S["conv2_weight"].data_ptr(), self.dp1.data_ptr(), B)
        else:
            # pool2 + ReLU backward (O19, O18), conv2 bias gradient
            N.call("mtb_unpool_relu_bwd" + sfx, self.dflat.data_ptr(), self.a2p.data_ptr(), self.idx2.data_ptr(),
                   self.da2.data_ptr(), B, 14, 14, s.c2)
            da2 = self.da2.view(self.Bmax * 196, s.c2)
            N.call("mtb_colsum_f32" if self.fp32 else "mtb_colsum_bf16", da2.data_ptr(), G["conv2_bias"].data_ptr(), M2, s.c2)
            if self.conv2_igemm:
                # wgrad and dgrad as implicit GEMMs straight from the NHWC tensors (split-K over the batch for wgrad)
                N.call("mtb_conv5x5_gemm", 2, self.a1p.data_ptr(), da2.data_ptr(), G["conv2_weight"].data_ptr(), B, s.c1,
                       s.c2, int(self.fp32), None, 0, self.wg_splits, self.wg_ws.data_ptr(), self.wg_cnt.data_ptr())
                if after_conv2_wgrad is not None:
                    after_conv2_wgrad()
                N.call("mtb_conv5x5_gemm", 1, da2.data_ptr(), S["conv2_weight"].data_ptr(), self.dp1.data_ptr(), B, s.c1,
                       s.c2, int(self.fp32), None, 0, 1, None, None)
            else:
                # conv2 wgrad: dW = col^T da2, K = B*196 -> split-K with last-arriver reduction
                gemm(self.col, da2, G["conv2_weight"].view(K2, s.c2), K2, s.c2, M2, True, True, bn=64,
                     splits=self.wg_splits, workspace=self.wg_ws, counters=self.wg_cnt)
                if after_conv2_wgrad is not None:
                    after_conv2_wgrad()
                # conv2 dgrad: dcol = da2 W2^T, then col2im
                gemm(da2, S["conv2_weight"].view(K2, s.c2), self.dcol, M2, K2, s.c2, False, False, bn=128)
                N.call("mtb_col2im_5x5" + sfx, self.dcol.data_ptr(), self.dp1.data_ptr(), B, 14, 14, s.c1)
        # pool1 + ReLU backward fused into the conv1 weight gradient
        if self.conv1_tc:
            wx, _, wstep, wrows = self._win()
            N.call("mtb_conv1_wgrad_tc_win", wx, self.dp1.data_ptr(), self.a1p.data_ptr(),
                   self.idx1.data_ptr(), G["conv1_weight"].data_ptr(), G["conv1_bias"].data_ptr(), B, wstep, wrows)
        else:
            N.call("mtb_conv1_wgrad" + sfx, self._win()[0], self.dp1.data_ptr(), self.a1p.data_ptr(),
                   self.idx1.data_ptr(), G["conv1_weight"].data_ptr(), G["conv1_bias"].data_ptr(), B, s.c1)

    def _splitk_scratch(self, key: str, M: int, Nn: int, bn: int, splits: int):
        """Workspace and tile counters of a split-K GEMM, allocated once per call site and shape."""
        if splits <= 1:
            return None, None
        k = (key, M, Nn, bn, splits)
        if k not in self._scratch:
            ws, cn = N.gemm_scratch_sizes(M, Nn, bn, splits)
            self._scratch[k] = (torch.zeros(max(ws, 16), dtype=torch.uint8, device=self.device),
                                torch.zeros(max(cn, 16), dtype=torch.uint8, device=self.device))
        return self._scratch[k]

    def _zero_list(self):
        if self._zero_args is None:
            self._zero_step_accumulators(build_only=True)
        return self._zero_args

    def _zero_step_accumulators(self, build_only: bool = False):
        if self._zero_args is None:
            items = [(self.flat_grads[lo:hi].data_ptr(), hi - lo) for lo, hi in self._zero_ranges if hi > lo]
            items += [(self.loss_acc.data_ptr(), 1), (self.err_acc.data_ptr(), 1)]
            ptrs = (C.c_void_p * 8)(*[p for p, _ in items])
            words = (C.c_longlong * 8)(*[n for _, n in items])
            self._zero_args = (ptrs, words, len(items))
        if build_only:
            return
        ptrs, words, n = self._zero_args
        N.call("mtb_zero_multi", ptrs, words, n)

    # ------------------------------------------------------------------ update (K1)
    def _fused_args(self, peers: bool, lr_override: float = -1.0) -> N.FusedArgs:
        c, s = self.cfg, self.spec
        if peers and self.p2p is not None:
            a = self.p2p.fused_args()
            for r in range(self.world):
                a.grads[r], a.params[r], a.shadow[r] = self._grad_ptrs[r], self._param_ptrs[r], self._shadow_ptrs[r]
                if self._lp:
                    a.grads_lp[r] = self._grad_lp_ptrs[r]
                if self._push is not None:
                    a.push_recv[r] = self._push.ptrs[r]
            a.push_stride = self._push_stride
            if self.has_nvls:
                a.mc_grads, a.mc_params, a.mc_shadow = self._grad_mc, self._param_mc, self._shadow_mc
                if self._lp:
                    a.mc_grads_lp = self._grad_lp_mc
            if self.fp32:                          # no bf16 shadow on the fp32 path
                for r in range(self.world):
                    a.shadow[r] = None
                a.mc_shadow = None
            a.scale = 1.0 / self.world
        else:
            a = N.FusedArgs()
            a.world, a.rank = 1, 0
            a.grads[0], a.params[0], a.shadow[0] = (self.flat_grads.data_ptr(), self.flat_params.data_ptr(),
                                                    None if self.fp32 else self.shadow.data_ptr())
            if self._lp:
                a.grads_lp[0] = self.flat_grads_lp.data_ptr()
            a.l2_slots[0] = self._l2_slots.data_ptr()
            a.seq, a.status = self._seq.data_ptr(), self._status.data_ptr()
            a.scale = 1.0
        if self._lp:
            fc1w = s.param("fc1_weight")
            a.lp_begin, a.lp_end = fc1w.offset, fc1w.offset + fc1w.numel
        a.optimizer = _OPT[c.optimizer]
        a.slot_m = self.slots[0].data_ptr() if self.slots else None
        a.slot_v = self.slots[1].data_ptr() if len(self.slots) > 1 else None
        a.lr0, a.lr_decay, a.lr_override = c.lr, c.lr_decay, lr_override
        a.batch, a.decay_rows = c.batch, int(self.decay_rows or 0)
        a.step_ptr = self.dev_step.data_ptr()
        a.momentum, a.beta1, a.beta2, a.eps, a.wd = c.momentum, c.beta1, c.beta2, c.adam_eps, c.weight_decay
        dec = [p for p in s.params if p.decay]
        a.n_ranges = len(dec) if c.weight_decay else 0
        for i, p in enumerate(dec):
            a.range_begin[i], a.range_end[i] = p.offset, p.offset + p.numel
        if peers and self._shard_master:
            big = [s.param("conv2_weight"), s.param("fc1_weight")]
            a.n_sharded = len(big)
            for i, p in enumerate(big):
                a.shard_begin[i], a.shard_end[i] = p.offset, p.offset + p.numel
        return a

    def _pick_algo(self, count: int, tail: bool = False) -> int:
        """0 one-shot, 1 two-shot, 2 nvls, 3 push.  ``tail``: the bucket that closes the step on the main stream."""
        algo = self.cfg.comm_algo
        if self.world == 1 or algo == "oneshot":
            return 0
        push_ok = tail and self._push is not None and count <= self._push_stride and not self._rs_active
        if algo == "push":
            if push_ok:
                return 3
            algo = "auto"
        if algo == "twoshot":
            return 1
        if algo == "nvls":
            if not self.has_nvls:
                raise RuntimeError("comm_algo='nvls' but no multicast mapping is available on this box")
            return 2
        # The closing bucket is exposed (nothing runs under it): stores + ONE flag round beat load-based one-shot's two
        if push_ok and count * 4 <= (4 << 20):
            return 3
        # measured crossovers (bench/allreduce_sweep.py, profiles/allreduce_sweep_*.jsonl): below
        # 256 KB one flag round less wins; the in-switch reduction wins from 4 GPUs up, with 2 GPUs
        # it has nothing to save and the plain two-shot is faster
        if count * 4 <= (256 << 10):
            return 0
        # NVLS loses to the plain two-shot again at 1 GB (3.40 vs 2.95 ms, profiles/late_n8/sweep_1gb.jsonl)
        return 2 if (self.has_nvls and self.world >= 4 and count * 4 <= (512 << 20)) else 1

    def _update(self, peers: bool, lr_override: float = -1.0, lo: int = 0, hi: Optional[int] = None,
                advance: bool = True, grid: int = 0, block: int = 0, chan: int = 0, l2_row: int = 0,
                l2_rows=None, tail: bool = False, no_flags: bool = False):
        """K1 over elements [lo, hi) of the flat buffers.  ``advance``: this launch closes the step (bumps the
        counter, mails loss + error count).  ``l2_row``: where its 0.5*sum(w^2) partial goes; ``l2_rows``: for the
        closing launch, {row: sliced?} of every launch of the step (its own included)."""
        hi = self.n if hi is None else hi
        peers = peers and self.p2p is not None and self.world > 1
        a = self._fused_args(peers, lr_override)
        a.offset, a.count = lo, hi - lo
        a.algo = self._pick_algo(hi - lo, tail) if peers else 0
        sliced = peers and a.algo in (1, 2)
        if not sliced:
            a.n_sharded = 0                                            # every rank computes everything
        else:
            self._sliced[(lo, hi)] = a.algo
            self._masters_dirty = self._masters_dirty or a.n_sharded > 0
            self._slots_dirty = True                                   # slots of non-owned slices are not maintained
        a.grid, a.block, a.chan = grid, block, chan
        a.no_flags = int(no_flags and peers and a.algo != 3)
        if self._trace is not None:
            a.trace = self._trace[l2_row if (advance or chan == 0) else 2].data_ptr()
        if self._rs_active and sliced:
            fc1w = self.spec.param("fc1_weight")
            a.n_prered = 1
            a.prered_begin[0], a.prered_end[0] = fc1w.offset, fc1w.offset + fc1w.numel
        if a.algo == 3:
            a.lp_begin = a.lp_end = 0                                  # the tail holds no bf16-wire tensor
        want_l2 = bool(self.cfg.weight_decay) and lr_override < 0
        if want_l2:
            a.l2_blocks, a.l2_row = self._l2_blocks[chan].data_ptr(), l2_row
            a.done_counter = self._done[chan:].data_ptr()
        if advance:                                                     # iter_ += 1 in the same launch
            a.advance_step, a.done_counter = self.dev_step.data_ptr(), self._done[chan:].data_ptr()
            a.loss_src, a.loss_mail, a.loss_ring = self.loss_acc.data_ptr(), self._mail.data_ptr(), self._MAIL_RING
            a.err_src, a.aux_mail = self.err_acc.data_ptr(), self._aux.data_ptr()
            if want_l2:
                rows = l2_rows if l2_rows is not None else {l2_row: sliced}
                for row, is_sliced in rows.items():
                    a.l2_rows_n[row] = self.world if (is_sliced and peers) else 1
        N.call("mtb_fused_allreduce_update", C.byref(a))
        return sliced

    # ------------------------------------------------------------------ engine interface
    def compute_grads(self, x, y, step: int) -> float:
        """Eager forward+backward into ``flat_grads`` (generic path used with non-fused sync modes)."""
        if x is not None:
            self.stage_batch(x, y)
        else:
            self._window()
        self._windowed = False
        self._stream_buf = None
        self._zero_step_accumulators()
        self._forward(self.B, True, True)
        self._backward(self.B)
        self._rs_dirty = True
        return self._read_loss()

    def apply(self, lr: float) -> None:
        self._update(peers=False, lr_override=float(lr))
        self.opt_step += 1

    def _window(self):
        if self.data_x is None:
            raise RuntimeError("no resident dataset: call attach_dataset() or pass a batch")
        N.call("mtb_batch_window", self.data_x.data_ptr(), self.data_y.data_ptr(), self.dev_step.data_ptr(),
               self.rows, self.B, self.x_in.data_ptr(), self.y_in.data_ptr())

    def _read_loss(self) -> float:
        loss = float(self.loss_acc.item())
        if self.cfg.weight_decay:
            l2 = torch.zeros(1, device=self.device)
            for p in self.spec.params:
                if p.decay:
                    N.call("mtb_l2_half_sumsq", self.P[p.key].data_ptr(), p.numel, l2.data_ptr())
            loss += self.cfg.weight_decay * float(l2.item())
        return loss

    _MAIL_RING = 16

    def _await_loss(self, step_no: int, timeout: Optional[float] = None) -> float:
        """Loss of optimizer step ``step_no`` (1-based value of the device counter after it), read
        from the pinned mailbox K1 writes: no copy, no stream synchronize, just a poll."""
        slot, want = (step_no - 1) % self._MAIL_RING, step_no & 0xFFFFFFFF
        timeout = float(self.cfg.step_timeout_s) if timeout is None else timeout
        mail, t0, spins = self._mail_np, None, 0
        while True:
            word = int(mail[slot])
            if (word >> 32) & 0xFFFFFFFF == want:
                return float(np.uint32(word & 0xFFFFFFFF).view(np.float32))
            spins += 1
            if spins & 1023 == 0:
                now = time.monotonic()
                t0 = now if t0 is None else t0
                if now - t0 > timeout:
                    raise RuntimeError(f"step {step_no}: no loss arrived from the device within {timeout:.0f}s")

    def _mark(self, name: str, stream=None):
        """Phase boundary: an NVTX marker (MTB_NVTX=1, eager steps) and, while :meth:`timed_step` runs, a CUDA event
        on the stream the phase was launched on."""
        if self._nvtx and not torch.cuda.is_current_stream_capturing():
            torch.cuda.nvtx.mark(name)
        if self._marks is not None:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(stream if stream is not None else torch.cuda.current_stream())
            self._marks.append((name, ev))

    def timed_step(self, images=None, labels=None, peers: bool = True) -> Dict[str, float]:
        """One REAL optimizer step run eagerly with CUDA events between its phases (fwd, bwd,
        allreduce, update timers).  Returns microseconds: ``fwd``, ``bwd_fc``, ``bwd_conv``, ``k1_hidden`` (the fc
        bucket's fused allreduce+update, measured on the comm stream where it overlaps ``bwd_conv``), ``k1_tail`` (the
        exposed closing bucket), ``step``, plus ``allreduce_bytes`` and ``bus_gbs`` = 2(N-1)/N * bytes / (k1_hidden +
        k1_tail).  Collective when ``peers``: every rank must call it for the same step."""
        if images is not None:
            self.stage_batch(images, labels)
            mode = "staged"
        elif self._can_window and self.data_x is not None:
            mode = "window"
        else:
            self._window()
            mode = "staged"
        self._windowed = mode == "window"
        self._stream_buf = None
        peers = peers and self.p2p is not None and self.world > 1
        if self._rs_mode and self._rs_dirty and peers:
            self.G["fc1_weight"].zero_()
            torch.cuda.synchronize()
            self.p2p.barrier()
            self._rs_dirty = False
        self._marks = []
        try:
            self._mark("start")
            self._step_body(peers)
            self._mark("end")
            torch.cuda.synchronize()
            ev = dict(self._marks)
        finally:
            self._marks = None
        self.opt_step += 1
        us = lambda a, b: ev[a].elapsed_time(ev[b]) * 1e3 if (a in ev and b in ev) else 0.0   # noqa: E731
        out = {"fwd": us("start", "fwd"), "step": us("start", "end")}
        if "bwd_fc" in ev:
            out.update(bwd_fc=us("fwd", "bwd_fc"), bwd_conv=us("bwd_fc", "bwd_conv"),
                       k1_hidden=us("k1_hidden_begin", "k1_hidden_end"), k1_tail=us("bwd_conv", "end"))
        else:
            out.update(bwd=us("fwd", "bwd"), k1_tail=us("bwd", "end"), k1_hidden=0.0)
        nbytes = self.n * (2 if self._lp else 4)
        t = out["k1_hidden"] + out["k1_tail"]
        out["allreduce_bytes"] = nbytes
        out["bus_gbs"] = (2 * (self.world - 1) / self.world * nbytes / (t * 1e-6) / 1e9) if (peers and t > 0) else 0.0
        return out

    def _hidden_launch(self, peers: bool, lo: int, hi: int, row: int) -> bool:
        """One bucket's fused allreduce+update on the comm stream, shaped to run UNDER the persistent convolution
        kernels: many short-lived narrow blocks (one warp, one 4-vector pass each) that fit into whatever registers
        the resident convolution CTA leaves free and give them back every few microseconds, instead of blocks parked on
        the SMs for the whole transfer.  That needs more blocks than there are flag slots, so the two flag rounds
        are their own one-warp launches before and after the data launch."""
        multi = peers and self.p2p is not None and self.world > 1
        grid = int(os.environ.get("MTB_K1_GRID_A", "-1"))
        block = int(os.environ.get("MTB_K1_BLOCK_A", "32"))
        probe = os.environ.get("MTB_K1_PROBE")              # "fat:60" / "lean:60": occupancy probe instead of K1 (scripts)
        with torch.cuda.stream(self._comm_stream):
            if probe:
                kind, us = probe.split(":")
                N.call("mtb_debug_occupy", 148 if grid < 0 else grid, block, int(kind == "fat"),
                       int(float(us) * 1900), self._status.data_ptr())
                return None                                # no bucket was updated: the closing launch expects no L2 row
            if row == 0:
                self._mark("k1_hidden_begin")
            own_flags = os.environ.get("MTB_K1_INNER_FLAGS") == "1" or not multi
            if not own_flags:
                self.p2p.device_barrier()                  # every rank's gradients of this bucket are complete
            sliced = self._update(peers=peers, lo=lo, hi=hi, advance=False, grid=grid, block=block, chan=0, l2_row=row,
                                  no_flags=not own_flags)
            if not own_flags:
                self.p2p.device_barrier()                  # peers are done reading my gradients and writing my weights
            self._mark("k1_hidden_end")
        return sliced

    def train_errors(self, step_no: int, timeout: Optional[float] = None) -> int:
        """Misclassified rows of the batch of optimizer step ``step_no`` (1-based), from the second mailbox ring:
        the train error the reference computes and throws away (``mpipy.py:89``)."""
        slot, want = (step_no - 1) % self._MAIL_RING, step_no & 0xFFFFFFFF
        timeout = float(self.cfg.step_timeout_s) if timeout is None else timeout
        t0 = time.monotonic()
        while True:
            word = int(self._aux_np[slot])
            if (word >> 32) & 0xFFFFFFFF == want:
                return int(word & 0xFFFFFFFF)
            if time.monotonic() - t0 > timeout:
                raise RuntimeError(f"step {step_no}: no error count arrived from the device within {timeout:.0f}s")

    def _step_body(self, peers: bool = True):
        if self._stream_buf is not None and self._stream_u8:
            # host-streamed raw pixels: (u8 - 127.5) / 255 as extract_data does on the host (mpipy.py:215), on the device
            st = self._stream_state
            N.call("mtb_normalize_u8_dev", st["u8"][self._stream_buf].data_ptr(), st["x"][self._stream_buf].data_ptr(),
                   self.B * 784)
        if self.conv1_tc:
            self._forward(self.B, True, True, zero=True)
        else:
            self._zero_step_accumulators()
            self._forward(self.B, True, True)
        self._mark("fwd")
        if not self._overlap:
            self._rs_active = self._rs_mode and peers
            try:
                self._backward(self.B)
                self._mark("bwd")
                self._update(peers=peers, tail=True)
            finally:
                self._rs_active = False
            return
        # Buckets in gradient-ready order (SURVEY 5.5 "Overlap").  The fc bucket (98% of the bytes) closes after the
        # FC1 weight gradient; its allreduce+update (both flag rounds included) hides under the convolution backward.
        # A conv2 bucket follows the same way when it is big (the widened model).  What is left for the end of the
        # step is a few KB: one launch, one flag round (push), which also advances the step and mails the loss.
        main = torch.cuda.current_stream()
        comm = self._comm_stream
        split = self.spec.param("fc1_weight").offset
        c2 = self.spec.param("conv2_weight").offset
        rows = {}
        self._backward_fc(self.B)
        self._mark("bwd_fc")
        self._ev_fc.record(main)
        comm.wait_event(self._ev_fc)
        sliced = self._hidden_launch(peers, split, self.n, 0)
        if sliced is not None:
            rows[0] = sliced
        tail_hi = split
        hook = None
        big_conv2 = (split - c2) * 4 > int(os.environ.get("MTB_CONV2_BUCKET_BYTES", str(256 << 10)))
        if big_conv2:
            def hook():
                self._ev_c2.record(main)
                comm.wait_event(self._ev_c2)
                sliced = self._hidden_launch(peers, c2, split, 1)
                if sliced is not None:
                    rows[1] = sliced
            tail_hi = c2
        self._backward_conv(self.B, hook)
        self._mark("bwd_conv")
        self._ev_done.record(comm)
        main.wait_event(self._ev_done)
        rows[2] = False                                       # the tail: every rank updates all of it
        self._update(peers=peers, lo=0, hi=tail_hi, advance=True, chan=1, l2_row=2, l2_rows=rows, tail=True)

    def fused_step(self, images, labels, lr, step: int, read_loss: bool = True, peers: bool = True):
        """One full step: (window | staged batch) -> graph replay.  ``lr`` is ignored: the
        schedule runs on the device from the device step counter.  ``peers=False``: a LOCAL step
        (same graph, K1 without the cross-GPU reduction) - what runs between two parameter averagings of the
        reference's strategy (mpipy.py:83-91) and under ``sync="none"``."""
        if images is not None:
            self.stage_batch(images, labels)
            mode = "staged"
        elif self._can_window:
            if self.data_x is None:
                raise RuntimeError("no resident dataset: call attach_dataset() or pass a batch")
            mode = "window"                     # the kernels gather the window themselves
        else:
            self._window()
            mode = "staged"
        self._run_step(mode, peers)
        self.opt_step += 1
        return self._await_loss(self.opt_step) if read_loss else 0.0

    def _run_step(self, mode: str, peers: bool = True):
        peers = peers and self.p2p is not None and self.world > 1
        if self._rs_mode and self._rs_dirty and peers:
            # an eager backward (compute_grads) stored a whole local gradient: put the owners' slices back to zero
            self.G["fc1_weight"].zero_()
            torch.cuda.synchronize()
            self.p2p.barrier()
            self._rs_dirty = False
        self._windowed = mode == "window"
        self._stream_buf = int(mode[-1]) if "stream" in mode else None
        self._stream_u8 = mode.startswith("ustream")
        if self.cfg.cuda_graph:
            key = mode if peers or self.world == 1 else mode + "/local"
            if key not in self._graphs:
                self._capture(mode, peers, key)
            self._graphs[key].replay()
        else:
            self._step_body(peers)

    # ------------------------------------------------------------------ host-streamed training
    # Input ring of the host-streamed loop: a batch may be copied up to three steps ahead of the step that reads it.
    # With two buffers (one step ahead) 8 ranks ran 8 % below their device-resident rate: one late host->device copy
    # on any rank stalls all of them at the next allreduce; a deeper ring absorbs that jitter.
    _STREAM_BUFS = 4

    def stream_steps(self, hx: torch.Tensor, hy: torch.Tensor, steps: int, ring: int = 4,
                     peers: bool = True) -> List[float]:
        """``steps`` optimizer steps over a dataset that lives in PINNED HOST memory
        (``hx`` [rows, 784] fp32, ``hy`` [rows] int32), walked with the reference's sequential
        window (``mpipy.py:80-82``).  This is the reference's feed loop (``mpipy.py:83-85``: one
        host->device feed per step) done the B200 way: step i+1's batch is copied on a copy
        stream into the next slot of a small ring of input buffers while step i computes (one CUDA
        graph per slot, so the kernels read it in place), and each step's loss is mailed by the update
        kernel into a pinned host ring (:meth:`_await_loss`) that the host polls a few steps
        later, so neither transfer ever stalls the GPU.  Returns the loss of every step."""
        B, dev = self.B, self.device
        if not (hx.is_pinned() and hy.is_pinned()):
            raise ValueError("stream_steps needs pinned host tensors (tensor.pin_memory())")
        u8 = hx.dtype == torch.uint8                 # raw pixels as the IDX files hold them: a quarter of the bytes per copy
        if hy.dtype != torch.int32 or hx.dtype not in (torch.float32, torch.uint8):
            raise ValueError("stream_steps needs float32 (normalised) or uint8 (raw) images and int32 labels")
        hx = hx.view(hx.shape[0], -1)
        rows = hx.shape[0]
        if rows < B:
            raise ValueError("dataset smaller than one batch")
        if self._stream_state is None:
            self._stream_state = dict(
                x=[torch.empty(B, 784, device=dev) for _ in range(self._STREAM_BUFS)],
                y=[torch.empty(B, dtype=torch.int32, device=dev) for _ in range(self._STREAM_BUFS)],
                ready=[torch.cuda.Event() for _ in range(self._STREAM_BUFS)],
                consumed=[torch.cuda.Event() for _ in range(self._STREAM_BUFS)],
                copy=torch.cuda.Stream(device=dev),
            )
        if u8 and "u8" not in self._stream_state:
            self._stream_state["u8"] = [torch.empty(B, 784, dtype=torch.uint8, device=dev)
                                        for _ in range(self._STREAM_BUFS)]
        st = self._stream_state
        ring = max(1, min(ring, self._MAIL_RING - 1))
        main = torch.cuda.current_stream()
        copy = st["copy"]
        for ev in st["consumed"]:
            ev.record(main)
        direct = self._can_window                  # the kernels that read the batch take any pointer
        losses: List[float] = []
        first = self.opt_step + 1
        for i in range(steps):
            k = i % self._STREAM_BUFS
            # the reference's window, on the GLOBAL step so that a resumed or continued run walks on (mpipy.py:80-82)
            o = ((self.opt_step * B) % (rows - B)) if rows > B else 0
            copy.wait_event(st["consumed"][k])                 # the step that used this half has copied it out
            with torch.cuda.stream(copy):
                (st["u8"] if u8 else st["x"])[k].copy_(hx[o:o + B], non_blocking=True)
                st["y"][k].copy_(hy[o:o + B], non_blocking=True)
                st["ready"][k].record(copy)
            main.wait_event(st["ready"][k])
            if direct:                                         # one graph per slot: no device-to-device copy
                self._run_step(f"ustream{k}" if u8 else f"stream{k}", peers)
            else:
                if u8:
                    N.call("mtb_normalize_u8_dev", st["u8"][k].data_ptr(), st["x"][k].data_ptr(), B * 784)
                self.x_in[:B].copy_(st["x"][k], non_blocking=True)
                self.y_in[:B].copy_(st["y"][k], non_blocking=True)
                self._run_step("staged", peers)
            st["consumed"][k].record(main)
            self.opt_step += 1
            if i >= ring:                                      # the device mailed it `ring` steps ago
                losses.append(self._await_loss(first + len(losses)))
        while len(losses) < steps:
            losses.append(self._await_loss(first + len(losses)))
        return losses

    def _capture(self, mode: str = "staged", peers: bool = True, key: Optional[str] = None):
        """Warm up eagerly (sets kernel attributes, builds tensor maps), roll the state back, capture.
        One graph per input mode: ``staged`` reads the batch buffers, ``window`` gathers the
        resident shard's window inside the kernels; ``/local`` variants skip the cross-GPU reduction."""
        key = key or mode
        self._windowed = mode == "window"
        self._stream_buf = int(mode[-1]) if "stream" in mode else None
        self._stream_u8 = mode.startswith("ustream")
        snap = [t.clone() for t in (self.flat_params, self.shadow, self.dev_step, *self.slots)]
        if self.p2p is not None:
            self.p2p.barrier()
        self._step_body(peers)
        torch.cuda.synchronize()
        if self.p2p is not None:
            self.p2p.barrier()
        for t, s in zip((self.flat_params, self.shadow, self.dev_step, *self.slots), snap):
            t.copy_(s)
        torch.cuda.synchronize()
        self._mail[self.opt_step % self._MAIL_RING] = 0        # the rolled-back warm-up step mailed a loss too
        self._aux[self.opt_step % self._MAIL_RING] = 0
        self._reset_push()
        if self.p2p is not None:
            self.p2p.barrier()
        g = torch.cuda.CUDAGraph()
        launches = N.LAUNCHES
        with torch.cuda.graph(g, stream=self._side):
            self._step_body(peers)
        self.launches_per_step = N.LAUNCHES - launches
        self._graphs[key] = g
        self._graph = g
        torch.cuda.synchronize()

    # ------------------------------------------------------------------ parameter averaging (mpipy.py:95-153)
    def average_params(self, ranges, root_only: bool = False, include_momentum: bool = False) -> None:
        """COLLECTIVE.  The reference's ``bcast_parameters`` as ONE kind of kernel over peer memory: the element
        ranges ``[(offset, numel), ...]`` of the fp32 parameters (optionally of the optimizer slots too) are replaced
        by their mean over ranks, in place, on every rank - or on rank 0 only with ``root_only`` (the literal
        ``mpipy.py:129`` behaviour).  No host staging, no extra copy, no synchronize: K1 in plain-allreduce mode with
        grads := params.  Owner-sliced algorithms only: rank r is the sole reader of slice r on every peer and its
        sole writer afterwards, which is what makes in-place safe; the bf16 shadow is rewritten in the same pass."""
        if self.p2p is None or self.world == 1:
            return
        bufs = [(self._param_ptrs, self._param_mc, True)]
        if include_momentum:
            if not self._slot_ptrs and self.slots:
                raise RuntimeError("optimizer slots are not peer-mapped: construct the engine with "
                                   "sync='weight_avg' and sync_include_momentum=True")
            bufs += [(ptrs, mc, False) for ptrs, mc in zip(self._slot_ptrs, self._slot_mc)]
        merged = []
        for off, numel in sorted(ranges):
            lo, hi = off // 4 * 4, min(self.n, (off + numel + 3) // 4 * 4)     # padding between tensors is zero everywhere
            if merged and lo <= merged[-1][1]:
                merged[-1][1] = max(merged[-1][1], hi)
            else:
                merged.append([lo, hi])
        for ptrs, mc, is_param in bufs:
            for lo, hi in merged:
                a = self.p2p.fused_args()
                a.optimizer = -1
                nvls = bool(mc) and self.world >= 4 and not root_only and self.cfg.comm_algo in ("auto", "nvls")
                a.algo = 2 if nvls else 1
                for r in range(self.world):
                    a.grads[r] = a.params[r] = ptrs[r]
                    if is_param and not self.fp32:
                        a.shadow[r] = self._shadow_ptrs[r]
                if nvls:
                    a.mc_grads = a.mc_params = mc
                    if is_param and not self.fp32 and self._shadow_mc:
                        a.mc_shadow = self._shadow_mc
                    elif is_param and not self.fp32:
                        a.algo = 1
                a.offset, a.count, a.scale = lo, hi - lo, 1.0 / self.world
                a.dst_mask = 1 if root_only else 0
                a.chan = 0
                N.call("mtb_fused_allreduce_update", C.byref(a))

    # ------------------------------------------------------------------ inference
    def predict(self, x, dropout_step: Optional[int] = None) -> torch.Tensor:
        if isinstance(x, np.ndarray):
            x = torch.from_numpy(x)
        B = x.shape[0]
        if B > self.Bmax:
            raise ValueError(f"batch {B} exceeds the engine's static capacity {self.Bmax}")
        self.x_in[:B].copy_(x.reshape(B, -1), non_blocking=True)
        self._windowed = False
        self._stream_buf = None
        self._forward(B, dropout_step is not None, False)
        return self.probs[:B].clone()

    def _eval_body(self, dev: torch.Tensor, out: torch.Tensor, B: int, dropout: bool):
        """The reference's ``eval_in_batches`` loop (``mpipy.py:174-182``) as a sequence of kernel launches that read
        the resident split in place and write the probabilities straight into ``out``."""
        size = dev.shape[0]
        self._windowed, self._stream_buf = False, None
        keep_x, keep_probs = self.x_in, self.probs
        try:
            for begin in range(0, size, B):
                if begin + B <= size:
                    self.x_in, self.probs = dev[begin:begin + B], out[begin:begin + B]
                    self._forward(B, dropout, False)
                else:                                   # ragged tail: the last B rows again, keep the needed suffix
                    self.x_in, self.probs = dev[size - B:], keep_probs
                    self._forward(B, dropout, False)
                    out[begin:].copy_(keep_probs[begin - size + B:B])
        finally:
            self.x_in, self.probs = keep_x, keep_probs

    def predict_all(self, data, B: int, dropout_step: Optional[int] = None) -> np.ndarray:
        """``eval_in_batches`` (``mpipy.py:169-183``).  The split is uploaded ONCE and stays resident (the reference
        re-feeds it every step: ~157 host->device feeds per evaluation at N=1, ``mpipy.py:86,177``); the whole loop
        over it is captured in one CUDA graph per (split, batch, dropout) and replayed, so an evaluation is one
        launch plus one device->host copy of the probabilities.  Dropout at eval (the reference's behaviour,
        ``mpipy.py:166`` via ``:68``) uses the device step counter's stream, as training does."""
        if isinstance(data, np.ndarray):
            data = torch.from_numpy(data)
        size = data.shape[0]
        if B > self.Bmax:
            raise ValueError(f"batch {B} exceeds the engine's static capacity {self.Bmax}")
        key = (data.data_ptr(), tuple(data.shape), B, dropout_step is not None)
        ent = self._eval_cache.get(key)
        # The resident copy is keyed by the host array's address; a strided sample of its contents guards against the
        # caller having rewritten the array in place since the upload (then it is uploaded again into the same device
        # buffer, which the captured graph keeps reading).
        flat = data.reshape(-1)
        sample = flat[::max(1, flat.numel() // 4096)][:4096].to(torch.float64)
        mark = (float(sample.sum()), float(sample.abs().sum()), float(flat[0]), float(flat[-1]))
        if ent is not None and ent["mark"] != mark:
            ent["dev"].copy_(data.reshape(size, -1).to(torch.float32))
            ent["mark"] = mark
        if ent is None:
            if len(self._eval_cache) >= 8:
                self._eval_cache.clear()
            dev = data.reshape(size, -1).to(self.device, dtype=torch.float32).contiguous()
            out = torch.empty(size, self.spec.num_classes, device=self.device)
            ent = {"dev": dev, "out": out, "graph": None, "host": data, "mark": mark}      # `host` pins the key's address
            self._eval_cache[key] = ent
            if self.cfg.cuda_graph:
                self._eval_body(dev, out, B, dropout_step is not None)      # warm-up: attributes, tensor maps
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=self._side):
                    self._eval_body(dev, out, B, dropout_step is not None)
                ent["graph"] = g
        if ent["graph"] is not None:
            ent["graph"].replay()
        else:
            self._eval_body(ent["dev"], ent["out"], B, dropout_step is not None)
        return ent["out"].cpu().numpy()

    def close(self):
        self._graph = None
        self._graphs = {}
        self._eval_cache = {}
