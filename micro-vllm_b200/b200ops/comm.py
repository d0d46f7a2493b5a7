"""PeerComm — NVLink peer-memory communicator behind the fused all-reduce + residual + RMSNorm
kernel (``b200_allreduce_residual_rmsnorm``, include/b200_abi.h).

Replaces, on the decode path, the NCCL ``all_reduce`` of core/parallel_config.py:65-69 and the
elementwise kernels that follow it at core/layer/model_graph.py:210-216 and :219/:143.

One process per GPU.  Every rank cudaMalloc's ONE slab (IPC-exportable, outside torch's caching
allocator), laid out as

    [ flags u32[1024][8] + u32[1024] | epoch u32[1028] (local only) | pad | slot 0 | slot 1 | y |
      x staging (LL) | y staging (LL) ]

exports it with cudaIpcGetMemHandle, exchanges the 64-byte handles through torch.distributed
(plumbing only) and maps every peer's slab.  The two data slots are the GEMM output buffers the
kernel reads on every rank; consecutive calls must alternate slots (the barrier of call k+1 proves
that all peers finished reading the slot of call k), which :meth:`allreduce_residual_rmsnorm`
enforces on the host at call / capture time.  ``y`` is the peer-visible output buffer of the
row-sharded form, in which the owner of a row pushes the normalised row to every rank.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch

from . import _lib
from ._lib import ArArgs, check, lib

_FLAG_BYTES = _lib.AR_MAX_CTAS * (_lib.AR_MAX_WORLD + 1) * 4        # ready flags [1024][8] + done flags [1024]
_EPOCH_BYTES = ((_lib.AR_MAX_CTAS + 4) * 4 + 255) // 256 * 256     # per-CTA epochs, error word, LL call + arrival counters
_HEADER = _FLAG_BYTES + _EPOCH_BYTES


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

    def __init__(self, slot_bytes: int, group=None, device: Optional[torch.device] = None, timeout_ms: int = 2000):
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
        total = self._ystage_off + 2 * self.slot_bytes
        with torch.cuda.device(self.device):
            p = C.c_void_p()
            check(lib.b200_comm_alloc(total, C.byref(p)), "b200_comm_alloc")
            self._base = p.value
            handle = C.create_string_buffer(_lib.IPC_HANDLE_BYTES)
            check(lib.b200_ipc_export(self._base, handle), "b200_ipc_export")
            handles: List[Optional[bytes]] = [None] * self.world
            dist.all_gather_object(handles, bytes(handle.raw), group=self.group)
            self._peer_base: List[int] = []
            self._imported: List[int] = []
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
        assert self._bytes.data_ptr() == self._base and self._bytes.numel() == total
        self._last_slot = -1
        dist.barrier(group=self.group)          # nobody launches before every rank has mapped every slab

    # ------------------------------------------------------------------ buffers
    def slot(self, idx: int, rows: int, n: int) -> torch.Tensor:
        """bf16 [rows, n] view of local slot ``idx`` (0 / 1): the GEMM whose output is to be reduced
        writes here (``torch.matmul(..., out=comm.slot(i, rows, n))``)."""
        assert idx in (0, 1) and rows * n * 2 <= self.slot_bytes, "PeerComm slot too small"
        off = _HEADER + idx * self.slot_bytes
        return self._bytes[off: off + rows * n * 2].view(torch.bfloat16).view(rows, n)

    def ybuf(self, rows: int, n: int) -> torch.Tensor:
        """bf16 [rows, n] view of the local peer-visible y buffer (row-sharded form: every owner
        pushes its normalised rows into every rank's copy)."""
        assert rows * n * 2 <= self.slot_bytes, "PeerComm y buffer too small"
        off = _HEADER + 2 * self.slot_bytes
        return self._bytes[off: off + rows * n * 2].view(torch.bfloat16).view(rows, n)

    def fits(self, rows: int, n: int) -> bool:
        return rows * n * 2 <= self.slot_bytes and n % 8 == 0 and n <= 16384

    def error_flag(self) -> int:
        """1 if any flag wait timed out since creation (synchronises)."""
        off = _FLAG_BYTES + _lib.AR_MAX_CTAS * 4
        return int(self._bytes[off: off + 4].view(torch.int32).item())

    # ------------------------------------------------------------------ the fused op
    def allreduce_residual_rmsnorm(self, slot_idx: int, rows: int, n: int, residual: Optional[torch.Tensor],
                                   weight: Optional[torch.Tensor], y: Optional[torch.Tensor],
                                   res_out: torch.Tensor, eps: float = 1e-6, ss_rounded: bool = False,
                                   sharded: bool = False, trace: Optional[torch.Tensor] = None, ll: bool = False):
        """res_out = bf16(bf16(sum over ranks of slot[slot_idx][:rows, :n]) + residual);
        y = rmsnorm(res_out) * weight (skipped when weight is None).  Current stream; capturable.
        ``sharded``: the row-sharded form (B200_AR_MODE_ROW_SHARDED) — ``y`` must be a prefix view of
        :meth:`ybuf`, rows <= 1024, and ``residual`` / ``res_out`` are only read / written for the rows
        this rank owns."""
        if slot_idx == self._last_slot and not (sharded or ll):
            raise AssertionError("PeerComm: consecutive fused all-reduces must alternate slots 0 / 1")
        self._last_slot = slot_idx
        x = self.slot(slot_idx, rows, n)
        a = ArArgs()
        off = _HEADER + slot_idx * self.slot_bytes
        for r in range(self.world):
            a.x_peers[r] = self._peer_base[r] + off
            a.flag_peers[r] = self._peer_base[r]
        a.epoch = self._base + _FLAG_BYTES
        _fill_common(a, x, residual, weight, y, res_out, eps, ss_rounded)
        a.rank, a.world, a.timeout_ms = self.rank, self.world, self.timeout_ms
        if sharded:
            yoff = _HEADER + 2 * self.slot_bytes
            assert y is not None and y.data_ptr() == self._base + yoff and y.stride(0) == n, \
                "PeerComm: the row-sharded form writes y into ybuf() on every rank"
            assert rows <= _lib.AR_MAX_CTAS
            a.mode = _lib.AR_MODE_ROW_SHARDED
            for r in range(self.world):
                a.y_peers[r] = self._peer_base[r] + yoff
        if ll:
            # LL push form: flag-in-data packets through the staging areas; y is any local tensor
            assert not sharded and y is not None
            R = -(-rows // self.world)
            assert self.world * R * n * 4 <= self._xstage_bytes and rows * n * 4 <= 2 * self.slot_bytes
            a.mode = _lib.AR_MODE_LL
            for r in range(self.world):
                a.xstage_peers[r] = self._peer_base[r] + self._xstage_off
                a.ystage_peers[r] = self._peer_base[r] + self._ystage_off
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
        lib.b200_comm_free(self._base)
        self._base = None
