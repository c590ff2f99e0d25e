"""GPU peer-memory communicator: collectives as kernels over NVLink / NVSwitch.

The reference stages every collective through host NumPy buffers and an MPI
library (``mpipy.py:121-127`` D2H ``.eval()`` + ``Gather``; ``:146-152`` H2D assign;
``:236-241`` host ``Scatter``).  Here each rank owns *symmetric* device
allocations that every peer maps through CUDA IPC; a collective is then a kernel
(or a peer D2D copy) on those mappings, bracketed by the flag barrier of
``csrc/allreduce_fused.cu``.  ``torch.distributed`` (gloo) is used only as the
bootstrap channel that exchanges the 64-byte IPC handles, and for CPU tensors.

The training hot path does not go through the generic methods below: the engine
hands its gradient / parameter buffers (allocated with :meth:`alloc_symmetric`)
straight to ``mtb_fused_allreduce_update``.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Tuple

import torch
import torch.distributed

from ..ops import native as N
from .comm import CommTimeout, Communicator, TorchComm


class SymmetricBuffer:
    """One allocation per rank, same size everywhere, all of them mapped locally."""

    def __init__(self, local_ptr: int, peer_ptrs: List[int], nbytes: int, rank: int, device: torch.device,
                 mc_ptr: int = 0, local: Optional[torch.Tensor] = None, keep=None):
        self.ptrs, self.nbytes, self.rank, self.device = peer_ptrs, nbytes, rank, device
        self.local_ptr = local_ptr
        self.mc_ptr = mc_ptr                      # NVLS multicast address of the same bytes on every rank (0: none)
        self.local = local if local is not None else N.tensor_from_ptr(local_ptr, nbytes, device)
        self._keep = keep                         # whatever owns the mapping

    def view(self, rank: int) -> torch.Tensor:
        """uint8 view of ``rank``'s copy (peer memory mapped into this process)."""
        return self.local if rank == self.rank else N.tensor_from_ptr(self.ptrs[rank], self.nbytes, self.device)

    def typed(self, dtype: torch.dtype, rank: Optional[int] = None) -> torch.Tensor:
        return self.view(self.rank if rank is None else rank).view(dtype)


class P2PComm(Communicator):
    backend = "p2p"

    def __init__(self, info, device: torch.device, timeout_s: float = 120.0, staging_bytes: int = 64 << 20,
                 boot: Optional[Communicator] = None):
        self.device = device
        self.rank, self.size = info.rank, info.world
        torch.cuda.set_device(device)
        self.boot = boot if boot is not None else (TorchComm("gloo", timeout_s) if self.size > 1 else None)
        self._L = N.lib()
        # Allocator: torch's symmetric memory (CUDA VMM handles; also binds an NVLS multicast
        # object when the fabric has one) if it works here, else plain CUDA-IPC arenas.
        self.allocator = os.environ.get("MTB_P2P_ALLOC", "auto")       # auto | symm | ipc
        self._symm = None
        if self.size > 1 and self.allocator in ("auto", "symm") and isinstance(self.boot, TorchComm):
            try:
                import torch.distributed._symmetric_memory as symm
                probe = symm.empty(256, dtype=torch.uint8, device=device)
                symm.rendezvous(probe, group=torch.distributed.group.WORLD)
                self._symm = symm
            except Exception as e:  # noqa: BLE001
                if self.allocator == "symm":
                    raise
                self._symm = None
            # every rank must take the same path
            oks = self.boot.allgather_object(self._symm is not None)
            if not all(oks):
                self._symm = None
        self.has_multicast = False
        self._allocs: List[Tuple[int, List[int]]] = []
        flag_bytes = self._L.mtb_fused_flag_bytes()
        # control block: [flags | seq (2 KiB) | status (256 B) | tagged L2 partial slots (4 rows x 16 ranks x 8 B)]
        self._seq_bytes = 4 * self._L.mtb_fused_max_blocks()
        self._l2_off = flag_bytes + self._seq_bytes + 256
        self.ctrl = self.alloc_symmetric(self._l2_off + 4 * 16 * 8)
        self._flag_bytes = flag_bytes
        self.staging = self.alloc_symmetric(staging_bytes)
        self.staging_bytes = staging_bytes
        self.barrier()

    # ------------------------------------------------------------------ memory
    def alloc_symmetric(self, nbytes: int) -> SymmetricBuffer:
        nbytes = (nbytes + 255) // 256 * 256
        if self._symm is not None:
            t = self._symm.empty(nbytes, dtype=torch.uint8, device=self.device)
            t.zero_()
            h = self._symm.rendezvous(t, group=torch.distributed.group.WORLD)
            ptrs = [int(p) for p in h.buffer_ptrs]
            mc = int(getattr(h, "multicast_ptr", 0) or 0)
            self.has_multicast = self.has_multicast or mc != 0
            torch.cuda.current_stream().synchronize()
            self._allocs.append((0, []))
            return SymmetricBuffer(ptrs[self.rank], ptrs, nbytes, self.rank, self.device, mc_ptr=mc, local=t,
                                   keep=(t, h))
        p = C.c_void_p()
        N.check(self._L.mtb_arena_alloc(nbytes, C.byref(p)), "arena_alloc")
        ptrs = [0] * self.size
        ptrs[self.rank] = p.value
        if self.size > 1:
            h = (C.c_ubyte * 64)()
            N.check(self._L.mtb_ipc_export(p, h), "ipc_export")
            handles = self.boot.allgather_object(bytes(h))
            for r, hb in enumerate(handles):
                if r == self.rank:
                    continue
                q = C.c_void_p()
                buf = (C.c_ubyte * 64).from_buffer_copy(hb)
                N.check(self._L.mtb_ipc_import(buf, C.byref(q)), f"ipc_import(rank {r})")
                ptrs[r] = q.value
        self._allocs.append((p.value, ptrs))
        return SymmetricBuffer(p.value, ptrs, nbytes, self.rank, self.device)

    # ------------------------------------------------------------------ kernels
    def fused_args(self) -> N.FusedArgs:
        """FusedArgs pre-filled with world/rank and the control-block pointers."""
        a = N.FusedArgs()
        a.world, a.rank = self.size, self.rank
        for r in range(self.size):
            a.flags[r] = self.ctrl.ptrs[r]
            a.l2_slots[r] = self.ctrl.ptrs[r] + self._l2_off
        a.seq = self.ctrl.local_ptr + self._flag_bytes
        a.status = self.ctrl.local_ptr + self._flag_bytes + self._seq_bytes
        a.lr_override = -1.0
        a.scale = 1.0
        return a

    def check_status(self) -> None:
        o = self._flag_bytes + self._seq_bytes
        st = self.ctrl.local[o:o + 4].view(torch.int32)
        code = int(st.item())
        if code != 0:
            st.zero_()
            raise CommTimeout(f"rank {self.rank}: a device-side wait on a peer timed out (code {code})")

    def barrier(self) -> None:
        """Host-level barrier.  Ranks first meet on the bootstrap channel (which may take as
        long as the slowest rank needs), then on the device flags, so the bounded device-side
        spin only ever covers launch skew."""
        if self.size > 1:
            torch.cuda.current_stream().synchronize()
            self.boot.barrier()
            a = self.fused_args()
            N.call("mtb_p2p_barrier", C.byref(a))
        torch.cuda.current_stream().synchronize()
        if self.size > 1:
            self.check_status()

    def device_barrier(self) -> None:
        """Stream-ordered barrier (no host sync) — capturable in a CUDA graph."""
        if self.size > 1:
            a = self.fused_args()
            N.call("mtb_p2p_barrier", C.byref(a))

    # ------------------------------------------------------------------ collectives
    def _is_dev(self, t: Optional[torch.Tensor]) -> bool:
        return t is not None and t.is_cuda

    def _auto_algo(self, nbytes: int) -> int:
        """One-shot below 256 KB (one flag round less), else the plain two-shot in place on the staging buffer (measured
        crossovers, bench/allreduce_sweep.py; at 4 GPUs it also beats the in-switch reduction, 63.7 vs 73.3 us at 16 MB,
        and at 8 they are within 2 %).  ``algo=2`` asks for the in-switch reduction explicitly."""
        return 0 if nbytes <= (256 << 10) else 1

    def allreduce(self, t: torch.Tensor, average: bool = False, algo: Optional[int] = None) -> None:
        if not t.is_cuda:
            return self.boot.allreduce(t, average) if self.size > 1 else None
        if self.size == 1:
            return
        if t.dtype != torch.float32 or not t.is_contiguous():
            tmp = t.to(torch.float32).contiguous()
            self.allreduce(tmp, average, algo)
            t.copy_(tmp.to(t.dtype))
            return
        flat = t.view(-1)
        cap = self.staging_bytes // 4
        stage = self.staging.typed(torch.float32)
        for off in range(0, flat.numel(), cap):
            n = min(cap, flat.numel() - off)
            n4 = (n + 3) // 4 * 4
            stage[:n].copy_(flat[off:off + n])
            if n4 > n:
                stage[n:n4].zero_()
            use = self._auto_algo(n4 * 4) if algo is None else algo
            a = self.fused_args()
            a.algo, a.optimizer = use, -1
            for r in range(self.size):
                a.grads[r] = self.staging.ptrs[r]
            a.offset, a.count = 0, n4
            a.scale = (1.0 / self.size) if average else 1.0
            if use == 0:
                # every rank reduces the whole message: the result goes straight to the caller's tensor
                out = flat[off:off + n] if n4 == n else torch.empty(n4, dtype=torch.float32, device=t.device)
                a.params[self.rank] = out.data_ptr()
                N.call("mtb_fused_allreduce_update", C.byref(a))
                if n4 != n:
                    flat[off:off + n].copy_(out[:n])
            else:
                # owner-sliced, in place on the staging buffer: rank r is the only reader of slice r on every peer and
                # the only writer of it afterwards, and the kernel's closing flag round covers the peer stores
                for r in range(self.size):
                    a.params[r] = self.staging.ptrs[r]
                if use == 2:
                    a.mc_grads = a.mc_params = self.staging.mc_ptr
                N.call("mtb_fused_allreduce_update", C.byref(a))
                flat[off:off + n].copy_(stage[:n])
        torch.cuda.current_stream().synchronize()
        self.check_status()

    def _staged(self, nbytes: int):
        if nbytes > self.staging_bytes:
            raise ValueError(f"message of {nbytes} B exceeds the {self.staging_bytes} B staging buffer; "
                             f"construct P2PComm with a larger staging_bytes")

    def bcast(self, t: torch.Tensor, root: int = 0) -> None:
        if not t.is_cuda:
            return self.boot.bcast(t, root) if self.size > 1 else None
        if self.size == 1:
            return
        nb = t.numel() * t.element_size()
        self._staged(nb)
        raw = t.contiguous().view(-1).view(torch.uint8)
        if self.rank == root:
            self.staging.local[:nb].copy_(raw)
        self.device_barrier()
        if self.rank != root:
            t.view(-1).view(torch.uint8).copy_(self.staging.view(root)[:nb])
        self.barrier()

    def scatter(self, send: Optional[torch.Tensor], recv: torch.Tensor, root: int = 0) -> None:
        if not recv.is_cuda:
            return self.boot.scatter(send, recv, root) if self.size > 1 else recv.copy_(send.reshape(recv.shape))
        nb = recv.numel() * recv.element_size()
        if self.size == 1:
            recv.view(-1).copy_(send.reshape(-1)[:recv.numel()])
            return
        self._staged(nb * self.size)
        if self.rank == root:
            src = send.to(recv.device, non_blocking=True).contiguous().view(-1).view(torch.uint8)
            self.staging.local[:nb * self.size].copy_(src)
        self.device_barrier()
        recv.view(-1).view(torch.uint8).copy_(self.staging.view(root)[self.rank * nb:(self.rank + 1) * nb])
        self.barrier()

    def gather(self, send: torch.Tensor, recv: Optional[torch.Tensor], root: int = 0) -> None:
        if not send.is_cuda:
            return self.boot.gather(send, recv, root) if self.size > 1 else recv.reshape(-1).copy_(send.reshape(-1))
        nb = send.numel() * send.element_size()
        if self.size == 1:
            recv.view(-1)[:send.numel()].copy_(send.reshape(-1))
            return
        self._staged(nb)
        self.staging.local[:nb].copy_(send.contiguous().view(-1).view(torch.uint8))
        self.device_barrier()
        if self.rank == root:
            out = recv.view(-1).view(torch.uint8)
            for r in range(self.size):
                out[r * nb:(r + 1) * nb].copy_(self.staging.view(r)[:nb])
        self.barrier()

    def allgather(self, send: torch.Tensor, recv: torch.Tensor) -> None:
        if not send.is_cuda:
            return self.boot.allgather(send, recv) if self.size > 1 else recv.reshape(-1).copy_(send.reshape(-1))
        nb = send.numel() * send.element_size()
        if self.size == 1:
            recv.view(-1).copy_(send.reshape(-1))
            return
        self._staged(nb)
        self.staging.local[:nb].copy_(send.contiguous().view(-1).view(torch.uint8))
        self.device_barrier()
        out = recv.view(-1).view(torch.uint8)
        for r in range(self.size):
            out[r * nb:(r + 1) * nb].copy_(self.staging.view(r)[:nb])
        self.barrier()

    def abort(self) -> None:
        if self.boot is not None:
            self.boot.abort()

    def close(self) -> None:
        torch.cuda.synchronize()
        for local, ptrs in self._allocs:
            for r, q in enumerate(ptrs):
                if r != self.rank and q:
                    self._L.mtb_ipc_close(C.c_void_p(q))
        for local, _ in self._allocs:
            if local:
                self._L.mtb_arena_free(C.c_void_p(local))
        self._allocs.clear()
        if self.boot is not None:
            self.boot.close()
