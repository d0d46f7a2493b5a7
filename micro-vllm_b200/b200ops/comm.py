"""PeerComm — NVLink peer-memory communicator behind the fused all-reduce + residual + RMSNorm
kernel (``b200_allreduce_residual_rmsnorm``, include/b200_abi.h).

Replaces, on the decode path, the NCCL ``all_reduce`` of core/parallel_config.py:65-69 and the
elementwise kernels that follow it at core/layer/model_graph.py:210-216 and :219/:143.

One process per GPU.  Every rank owns ONE symmetric slab, laid out as

    [ flags u32[1024][8] + u32[1024] + NVLS counters | epoch u32[1032] (local only) | pad | slot 0 |
      slot 1 | y | x staging (LL) | y staging (LL) | gather buffer (optional) ]

and maps every peer's slab.  Two plumbing back ends provide the slab (``B200_COMM_BACKEND`` =
``auto`` | ``symm`` | ``ipc``):

  * ``symm`` — ``torch.distributed._symmetric_memory`` (cuMemCreate + fabric / fd handles exchanged
    through the process group's store, plumbing only): peer pointers AND, when the node has an
    NVSwitch with multicast support, ONE multicast address range bound to all ranks' slabs.  The
    multicast range is what the NVLS form of the kernel (``multimem.ld_reduce`` / ``multimem.st`` /
    ``multimem.red``: the reduction happens inside the switch) runs on.
  * ``ipc`` — ``cudaMalloc`` + ``cudaIpcGetMemHandle`` exchanged through torch.distributed; peer
    pointers only (one-shot / row-sharded / LL forms).

The two data slots are the GEMM output buffers the kernel reads on every rank.  In the one-shot form
consecutive calls must alternate slots (the barrier of call k+1 proves that all peers finished reading
the slot of call k), which :meth:`allreduce_residual_rmsnorm` enforces on the host at call / capture
time; the other forms end with a barrier of their own.  ``y`` is the peer-visible output buffer of the
row-sharded and NVLS forms, in which the owner of a row pushes the normalised row to every rank.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional

import torch

from . import _lib
from ._lib import ArArgs, GatherArgs, check, lib

# ready flags [1024][8] + done flags [1024] + the two NVLS counters (128 bytes apart)
_FLAG_BYTES = _lib.AR_MAX_CTAS * (_lib.AR_MAX_WORLD + 1) * 4 + 256
# per-CTA epochs, error word, LL call + arrival counters, NVLS targets + arrival counter
_EPOCH_BYTES = ((_lib.AR_MAX_CTAS + 16) * 4 + 255) // 256 * 256        # words up to MAX_CTAS + 10 are in use
_HEADER = _FLAG_BYTES + _EPOCH_BYTES
_ERR_OFF = _FLAG_BYTES + _lib.AR_MAX_CTAS * 4

MODES = ("oneshot", "sharded", "ll", "nvls")


class _DevMem:
    """Wraps a raw device allocation so torch can view it (``__cuda_array_interface__``)."""

    def __init__(self, ptr: int, nbytes: int):
        self.ptr, self.nbytes = ptr, nbytes
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 2, "strides": None}


def _wdt(w: torch.Tensor) -> int:
    if w.dtype == torch.bfloat16:
        return _lib.DT_BF16
    if w.dtype == torch.float32:
        return _lib.DT_F32
    raise AssertionError(f"norm weight dtype {w.dtype} unsupported (bf16 / fp32)")


def allreduce_residual_rmsnorm_local(x, residual, weight, y, res_out, eps: float, ss_rounded: bool):
    """world == 1 form of the fused op (no communicator, no barrier): ``res_out = bf16(x + residual)``,
    ``y = rmsnorm(res_out) * weight`` with the sum of squares taken from the rounded (``ss_rounded``)
    or the unrounded sum.  Launches on the current stream; capturable."""
    for t in (x, residual, weight, y, res_out):
        if t is not None and not t.is_cuda:
            raise AssertionError("b200ops: all tensors must be CUDA tensors (no CPU fallback exists)")
    rows, n = x.shape
    a = ArArgs()
    a.x_peers[0] = x.data_ptr()
    _fill_common(a, x, residual, weight, y, res_out, eps, ss_rounded)
    a.rank, a.world = 0, 1
    check(lib.b200_allreduce_residual_rmsnorm(C.byref(a), torch.cuda.current_stream().cuda_stream),
          "b200_allreduce_residual_rmsnorm")
    return y, res_out


def _fill_common(a: ArArgs, x, residual, weight, y, res_out, eps, ss_rounded):
    rows, n = x.shape
    assert x.dtype == torch.bfloat16 and res_out.dtype == torch.bfloat16 and x.stride(1) == 1
    assert res_out.shape == x.shape and res_out.stride(1) == 1
    a.rows, a.n = rows, n
    a.stride_x, a.stride_res = x.stride(0), res_out.stride(0)
    a.res_out = res_out.data_ptr()
    if residual is not None:
        assert residual.dtype == torch.bfloat16 and residual.shape == x.shape and residual.stride(1) == 1
        a.residual, a.stride_r = residual.data_ptr(), residual.stride(0)
    if weight is not None:
        assert y is not None and y.dtype == torch.bfloat16 and y.shape == x.shape and y.stride(1) == 1
        assert weight.is_contiguous() and weight.numel() == n
        a.weight, a.y, a.stride_y, a.w_dtype = weight.data_ptr(), y.data_ptr(), y.stride(0), _wdt(weight)
    else:
        assert y is None
    a.eps, a.ss_rounded = float(eps), 1 if ss_rounded else 0


class PeerComm:
    """Two symmetric bf16 slots of ``slot_bytes`` each on every rank of ``group`` + flag arrays."""

    def __init__(self, slot_bytes: int, group=None, device: Optional[torch.device] = None, timeout_ms: int = 2000,
                 backend: Optional[str] = None, gather_bytes: int = 0):
        import torch.distributed as dist
        assert dist.is_initialized(), "PeerComm needs torch.distributed (one process per GPU)"
        self.group = group if group is not None else dist.group.WORLD
        self.rank = dist.get_rank(self.group)
        self.world = dist.get_world_size(self.group)
        assert self.world in (2, 4, 8), f"PeerComm: world size {self.world} (2, 4 or 8 GPUs of one node)"
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.slot_bytes = (int(slot_bytes) + 255) // 256 * 256
        self.timeout_ms = int(timeout_ms)
        # LL staging: packets double the bytes; x staging holds world * ceil(rows / world) rows (<= rows + world)
        self._xstage_off = _HEADER + 3 * self.slot_bytes
        self._xstage_bytes = 2 * self.slot_bytes + self.world * 16384 * 4
        self._ystage_off = self._xstage_off + self._xstage_bytes
        # optional gather buffer (vocab-parallel lm_head: [rows, vocab] logits assembled on every rank)
        self._gather_off = self._ystage_off + 2 * self.slot_bytes
        self.gather_bytes = (int(gather_bytes) + 255) // 256 * 256
        total = self._gather_off + self.gather_bytes
        self._total = total
        self._imported: List[int] = []
        self._symm = None
        self._mc_base = 0
        want = (backend or os.environ.get("B200_COMM_BACKEND", "auto")).lower()
        assert want in ("auto", "symm", "ipc"), f"B200_COMM_BACKEND={want} (auto | symm | ipc)"
        self.backend = None
        self.backend_note = ""
        if want in ("auto", "symm"):
            ok, why = self._try_symm(total)
            flags = [None] * self.world
            dist.all_gather_object(flags, bool(ok), group=self.group)
            if all(flags):
                self.backend = "symm"
            else:
                self._drop_symm()
                self.backend_note = why or "a peer could not set up symmetric memory"
                if want == "symm":
                    raise RuntimeError(f"PeerComm: torch symmetric memory unavailable: {self.backend_note}")
        if self.backend is None:
            self._init_ipc(total)
            self.backend = "ipc"
        # every rank must agree on whether the multicast range exists
        mc = [None] * self.world
        dist.all_gather_object(mc, bool(self._mc_base), group=self.group)
        if not all(mc):
            self._mc_base = 0
        self.has_multicast = bool(self._mc_base)
        assert self._bytes.data_ptr() == self._base and self._bytes.numel() >= total
        self._last_slot, self._last_mode = -1, None
        self._err_host = torch.zeros(1, dtype=torch.int32).pin_memory()
        self._err_event: Optional[torch.cuda.Event] = None
        self._err_calls = 0
        dist.barrier(group=self.group)          # nobody launches before every rank has mapped every slab

    # ------------------------------------------------------------------ plumbing back ends
    def _try_symm(self, total: int):
        try:
            import torch.distributed._symmetric_memory as symm_mem
            with torch.cuda.device(self.device):
                t = symm_mem.empty(total, dtype=torch.uint8, device=self.device)
                hdl = symm_mem.rendezvous(t, self.group)
                t.zero_()
                torch.cuda.synchronize(self.device)
            ptrs = [int(p) for p in hdl.buffer_ptrs]
            off = int(t.data_ptr()) - ptrs[self.rank]
            if len(ptrs) != self.world or off < 0:
                return False, "unexpected symmetric memory layout"
            self._symm = (t, hdl)
            self._bytes = t
            self._base = int(t.data_ptr())
            self._peer_base = [p + off for p in ptrs]
            mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
            self._mc_base = mc + off if mc else 0
            return True, ""
        except Exception as e:  # noqa: BLE001 - any failure selects the IPC back end on every rank
            return False, f"{type(e).__name__}: {e}"[:200]

    def _drop_symm(self):
        self._symm = None
        self._mc_base = 0

    def _init_ipc(self, total: int):
        import torch.distributed as dist
        with torch.cuda.device(self.device):
            p = C.c_void_p()
            check(lib.b200_comm_alloc(total, C.byref(p)), "b200_comm_alloc")
            self._base = p.value
            self._ipc_base = p.value
            handle = C.create_string_buffer(_lib.IPC_HANDLE_BYTES)
            check(lib.b200_ipc_export(self._base, handle), "b200_ipc_export")
            handles: List[Optional[bytes]] = [None] * self.world
            dist.all_gather_object(handles, bytes(handle.raw), group=self.group)
            self._peer_base = []
            for r, h in enumerate(handles):
                if r == self.rank:
                    self._peer_base.append(self._base)
                    continue
                q = C.c_void_p()
                check(lib.b200_ipc_import(C.create_string_buffer(h, _lib.IPC_HANDLE_BYTES), C.byref(q)),
                      f"b200_ipc_import(rank {r})")
                self._peer_base.append(q.value)
                self._imported.append(q.value)
        self._mem = _DevMem(self._base, total)
        self._bytes = torch.as_tensor(self._mem, device=self.device)

    # ------------------------------------------------------------------ buffers
    def slot(self, idx: int, rows: int, n: int) -> torch.Tensor:
        """bf16 [rows, n] view of local slot ``idx`` (0 / 1): the GEMM whose output is to be reduced
        writes here (``torch.matmul(..., out=comm.slot(i, rows, n))``)."""
        assert idx in (0, 1) and rows * n * 2 <= self.slot_bytes, "PeerComm slot too small"
        off = _HEADER + idx * self.slot_bytes
        return self._bytes[off: off + rows * n * 2].view(torch.bfloat16).view(rows, n)

    def ybuf(self, rows: int, n: int) -> torch.Tensor:
        """bf16 [rows, n] view of the local peer-visible y buffer (row-sharded / NVLS forms: every owner
        pushes its normalised rows into every rank's copy)."""
        assert rows * n * 2 <= self.slot_bytes, "PeerComm y buffer too small"
        off = _HEADER + 2 * self.slot_bytes
        return self._bytes[off: off + rows * n * 2].view(torch.bfloat16).view(rows, n)

    def gather_buf(self, rows: int, cols: int) -> torch.Tensor:
        """bf16 [rows, cols] view of the local gather buffer (:meth:`allgather_cols` fills it on every rank)."""
        assert rows * cols * 2 <= self.gather_bytes, "PeerComm gather buffer too small"
        return self._bytes[self._gather_off: self._gather_off + rows * cols * 2].view(torch.bfloat16).view(rows, cols)

    def allgather_cols(self, src: torch.Tensor, dst: torch.Tensor, pdl: bool = False) -> torch.Tensor:
        """``dst[:, r * c:(r + 1) * c] = src of rank r`` on EVERY rank (``b200_allgather_cols``): ``src`` bf16
        ``[rows, c]`` local, ``dst`` a ``[rows, world * c]`` prefix view of :meth:`gather_buf`.  Current stream;
        capturable; complete on this rank when the kernel ends.  Between two calls every rank must pass another
        barrier-carrying collective after its consumers of ``dst`` (the decode step's reductions do)."""
        assert src.is_cuda and src.dtype == torch.bfloat16 and src.dim() == 2 and src.stride(1) == 1
        rows, c = src.shape
        assert dst.shape == (rows, self.world * c) and dst.stride(1) == 1
        off = dst.data_ptr() - self._base
        assert self._gather_off <= off and off + (rows - 1) * dst.stride(0) * 2 + self.world * c * 2 <= self._gather_off + self.gather_bytes, \
            "PeerComm.allgather_cols: dst must live in gather_buf()"
        a = GatherArgs()
        a.src = src.data_ptr()
        for r in range(self.world):
            a.dst_peers[r] = self._peer_base[r] + off
            a.flag_peers[r] = self._peer_base[r]
        if self.has_multicast:
            a.dst_mc = self._mc_base + off
            a.flags_mc = self._mc_base
        a.epoch = self._base + _FLAG_BYTES
        a.rows, a.cols_local = rows, c
        a.stride_src, a.stride_dst = src.stride(0) if rows > 1 else max(src.stride(0), c), dst.stride(0) if rows > 1 else max(dst.stride(0), self.world * c)
        a.rank, a.world, a.timeout_ms, a.flags = self.rank, self.world, self.timeout_ms, 1 if pdl else 0
        check(lib.b200_allgather_cols(C.byref(a), torch.cuda.current_stream().cuda_stream), "b200_allgather_cols")
        return dst

    def fits(self, rows: int, n: int) -> bool:
        return rows * n * 2 <= self.slot_bytes and n % 8 == 0 and n <= 16384

    def error_flag(self) -> int:
        """1 if any flag wait timed out since creation (synchronises)."""
        return int(self._bytes[_ERR_OFF: _ERR_OFF + 4].view(torch.int32).item())

    def poll_error(self, every: int = 16) -> None:
        """Non-blocking check of the kernel's timeout word, for the product path: every ``every``
        calls an async 4-byte copy of the word is queued behind the work issued so far; the value an
        EARLIER copy brought back is examined without waiting for the device.  Raises when a flag
        wait timed out, i.e. a peer rank did not reach the reduction within ``timeout_ms`` and the
        residual stream of this rank is no longer trustworthy."""
        if self._err_event is not None and self._err_event.query():
            if int(self._err_host[0]) != 0:
                raise RuntimeError(
                    f"PeerComm (rank {self.rank}): a peer did not reach the fused all-reduce within {self.timeout_ms} ms; "
                    "results since then are invalid")
            self._err_event = None
        self._err_calls += 1
        if self._err_event is None and self._err_calls % max(1, every) == 0:
            self._err_host.copy_(self._bytes[_ERR_OFF: _ERR_OFF + 4].view(torch.int32), non_blocking=True)
            self._err_event = torch.cuda.Event()
            self._err_event.record()

    # ------------------------------------------------------------------ the fused op
    def allreduce_residual_rmsnorm(self, slot_idx: int, rows: int, n: int, residual: Optional[torch.Tensor],
                                   weight: Optional[torch.Tensor], y: Optional[torch.Tensor],
                                   res_out: torch.Tensor, eps: float = 1e-6, ss_rounded: bool = False,
                                   sharded: bool = False, trace: Optional[torch.Tensor] = None, ll: bool = False,
                                   nvls: bool = False, pdl: bool = False, nvls1: bool = False, packet_tail: bool = False):
        """res_out = bf16(bf16(sum over ranks of slot[slot_idx][:rows, :n]) + residual);
        y = rmsnorm(res_out) * weight (skipped when weight is None).  Current stream; capturable.
        ``sharded`` / ``nvls``: the row-sharded forms (B200_AR_MODE_ROW_SHARDED, B200_AR_MODE_NVLS) —
        ``y`` must be a prefix view of :meth:`ybuf`, and ``residual`` / ``res_out`` are only read /
        written for the rows this rank owns.  ``ll``: the flag-in-data push form (same residual contract).
        ``nvls1``: the one-hop NVLS form (B200_AR_MODE_NVLS_ONEHOP: the one-shot contract — replicated residual, any
        local ``y``, alternate slots).  ``pdl``: launch with programmatic stream serialization (B200_AR_FLAG_PDL).  ``packet_tail`` (with ``nvls``): y
        travels as flag-in-data packets (B200_AR_FLAG_PACKET_TAIL) and may be any local tensor."""
        assert int(sharded) + int(ll) + int(nvls) + int(nvls1) <= 1, "PeerComm: choose one of sharded / ll / nvls / nvls1"
        # "oneshot" here = no trailing barrier (peer-load one-shot and the one-hop NVLS form)
        mode = "sharded" if sharded else "ll" if ll else "nvls" if nvls else "oneshot"
        # The one-shot form has no trailing barrier: after a one-shot call on slot s, peers may still be READING
        # slot s, so the next call (of any form) must use the other slot.  The other forms end with a barrier.
        if self._last_mode == "oneshot" and slot_idx == self._last_slot:
            raise AssertionError("PeerComm: the call after a one-shot fused all-reduce must use the other slot (0 / 1)")
        self._last_slot, self._last_mode = slot_idx, mode
        x = self.slot(slot_idx, rows, n)
        a = ArArgs()
        off = _HEADER + slot_idx * self.slot_bytes
        for r in range(self.world):
            a.x_peers[r] = self._peer_base[r] + off
            a.flag_peers[r] = self._peer_base[r]
        a.epoch = self._base + _FLAG_BYTES
        _fill_common(a, x, residual, weight, y, res_out, eps, ss_rounded)
        a.rank, a.world, a.timeout_ms = self.rank, self.world, self.timeout_ms
        yoff = _HEADER + 2 * self.slot_bytes
        if sharded:
            assert y is not None and y.data_ptr() == self._base + yoff and y.stride(0) == n, \
                "PeerComm: the row-sharded form writes y into ybuf() on every rank"
            assert rows <= _lib.AR_MAX_CTAS
            a.mode = _lib.AR_MODE_ROW_SHARDED
            for r in range(self.world):
                a.y_peers[r] = self._peer_base[r] + yoff
        if nvls:
            assert self.has_multicast, "PeerComm: the NVLS form needs a multicast mapping (symm back end on an NVSwitch node)"
            assert y is not None and (packet_tail or (y.data_ptr() == self._base + yoff and y.stride(0) == n)), \
                "PeerComm: the NVLS form writes y into ybuf() on every rank"
            a.mode = _lib.AR_MODE_NVLS
            a.x_mc = self._mc_base + off
            a.y_mc = self._mc_base + yoff
            a.flags_mc = self._mc_base
            if packet_tail:                 # y travels as flag-in-data packets through the y staging; y itself is local
                assert rows * n * 4 <= 2 * self.slot_bytes
                a.mode |= _lib.AR_FLAG_PACKET_TAIL
                for r in range(self.world):
                    a.ystage_peers[r] = self._peer_base[r] + self._ystage_off
        if nvls1:
            assert self.has_multicast, "PeerComm: the NVLS forms need a multicast mapping (symm back end on an NVSwitch node)"
            a.mode = _lib.AR_MODE_NVLS_ONEHOP
            a.x_mc = self._mc_base + off
            a.flags_mc = self._mc_base
        if ll:
            # LL push form: flag-in-data packets through the staging areas; y is any local tensor
            assert y is not None
            R = -(-rows // self.world)
            assert self.world * R * n * 4 <= self._xstage_bytes and rows * n * 4 <= 2 * self.slot_bytes
            a.mode = _lib.AR_MODE_LL
            for r in range(self.world):
                a.xstage_peers[r] = self._peer_base[r] + self._xstage_off
                a.ystage_peers[r] = self._peer_base[r] + self._ystage_off
        if pdl:
            a.mode |= _lib.AR_FLAG_PDL
        if trace is not None:
            assert trace.dtype == torch.int64 and trace.is_cuda and trace.numel() >= 4 * min(rows, _lib.AR_MAX_CTAS)
            a.trace = trace.data_ptr()
        check(lib.b200_allreduce_residual_rmsnorm(C.byref(a), torch.cuda.current_stream().cuda_stream),
              "b200_allreduce_residual_rmsnorm")
        return y, res_out

    def close(self):
        if getattr(self, "_base", None) is None:
            return
        torch.cuda.synchronize(self.device)
        for p in self._imported:
            lib.b200_ipc_close(p)
        self._imported = []
        self._bytes = None
        if self.backend == "ipc":
            lib.b200_comm_free(self._ipc_base)
        self._symm = None
        self._base = None
