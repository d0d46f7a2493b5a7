"""Paged KV-cache allocator with device-resident int32 block tables.

Drop-in for the reference's ``core/cache_manager.py`` (``KVCacheManager`` :262-691,
``store_kvcache`` :143-194, ``store_kvcache_batch`` :196-255): same constructor,
same methods and return conventions, same observable allocator state
(``_free`` FIFO deque, ``_blocks``, ``_pos``) — including the reference's quirks,
which the engine relies on:

  * ``append`` hands out offsets 0..254 of a block and then moves to a NEW block
    (offset 255 is never returned, cache_manager.py:425);
  * ``alloc`` with ``n_tokens % block_size == 0`` records the last block as EMPTY
    (``_pos == 0``, cache_manager.py:384-387);
  * ``free`` returns blocks to the right end of the FIFO in list order.

What is different is where the data lives: block tables and sequence lengths are
staged in pinned host mirrors and reach the device with ONE async copy each per
step (the reference issues one slice-assign kernel per batch row,
cache_manager.py:530-540); slot maps are built vectorised; the scatter kernel is
the sm_100a ``b200_store_kvcache`` instead of Triton.  There is no CPU scatter
path: ``put``/``store_kvcache`` require CUDA tensors.
"""
from __future__ import annotations

import collections
from typing import List, Optional, Tuple

import numpy as np
import torch

__all__ = ["KVCacheManager", "store_kvcache", "store_kvcache_batch"]


def store_kvcache(key: torch.Tensor, value: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                  slot_mapping: torch.Tensor, block_size: int) -> None:
    """Scatter ``key/value[t]`` to ``cache[slot // block_size, slot % block_size]``; slot -1 skips
    the token (reference cache_manager.py:143-194).  CUDA only."""
    num_tokens, num_heads, head_size = key.shape
    assert key.dim() == 3 and value.dim() == 3
    assert value.shape == (num_tokens, num_heads, head_size)
    assert slot_mapping.numel() == num_tokens
    assert key.is_cuda, "store_kvcache: CUDA tensors required (no CPU path)"
    from b200ops import store_kvcache as _store
    slots = slot_mapping.reshape(-1)
    if slots.dtype != torch.int32:
        slots = slots.to(torch.int32)
    _store(key, value, k_cache, v_cache, slots.contiguous(), block_size)


def store_kvcache_batch(key: torch.Tensor, value: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                        slot_mapping_batch: torch.Tensor, block_size: int) -> None:
    """Decode-step variant: ``key/value [B, 1, H, D]``, ``slot_mapping_batch [B, 1]``
    (reference cache_manager.py:196-255)."""
    batch_size, num_tokens, num_heads, head_size = key.shape
    assert key.dim() == 4 and value.dim() == 4
    assert value.shape == (batch_size, num_tokens, num_heads, head_size)
    assert slot_mapping_batch.shape == (batch_size, num_tokens)
    assert num_tokens == 1, "Only support num_tokens=1 for decoding stage"
    store_kvcache(key.view(-1, num_heads, head_size), value.view(-1, num_heads, head_size),
                  k_cache, v_cache, slot_mapping_batch.reshape(-1), block_size)


class KVCacheManager:
    def __init__(self, n_blocks: int, block_size: int, n_layers: int, n_heads: int, head_size: int,
                 dtype=torch.float16, device="cuda", max_tokens: int = 1024, max_batch_size: int = 32):
        self.n_blocks, self.block_size, self.n_layers = n_blocks, block_size, n_layers
        self.n_heads, self.head_size = n_heads, head_size
        self.dtype, self.device = dtype, device

        # one zero-filled slab, exposed per layer with the reference's shapes
        # [n_blocks, block_size, n_heads, head_size] (cache_manager.py:311-323)
        self._storage = torch.zeros((n_layers, 2, n_blocks, block_size, n_heads, head_size),
                                    dtype=dtype, device=device)
        self.k_caches = [self._storage[i, 0] for i in range(n_layers)]
        self.v_caches = [self._storage[i, 1] for i in range(n_layers)]

        self._free = collections.deque(range(n_blocks))      # FIFO of free block ids
        self._blocks = {}                                     # seq_id -> [block ids]
        self._pos = {}                                        # block id -> used slots

        self.cache_seqlens = torch.tensor([1], dtype=torch.int32, device=device)
        self.max_tokens = max_tokens
        self.max_seq_blocks = (max_tokens + block_size - 1) // block_size
        self.max_batch_size = max_batch_size
        self._block_table_buffer = torch.full((max_batch_size, self.max_seq_blocks), -1,
                                              dtype=torch.int32, device=device)
        self._cache_seqlens_buffer = torch.zeros(max_batch_size, dtype=torch.int32, device=device)

        # Pinned host mirrors: filled with numpy, shipped with one async copy each per step.
        # Two sets alternate and each carries the event of the copy that last read it, so the
        # host can run one step ahead of the device without overwriting bytes still in flight.
        self._on_gpu = torch.cuda.is_available() and torch.device(device).type == "cuda"
        self._stages = []
        for _ in range(2):
            bt = torch.full((max_batch_size, self.max_seq_blocks), -1, dtype=torch.int32, pin_memory=self._on_gpu)
            sl = torch.zeros(max_batch_size, dtype=torch.int32, pin_memory=self._on_gpu)
            self._stages.append({"bt": bt, "sl": sl, "bt_np": bt.numpy(), "sl_np": sl.numpy(), "event": None})
        self._stage_idx = 0
        self.len_max_over_mean = 1.0
        # Device-resident metadata between decode steps: a small device-side update list
        # [count, (row, col, block) * n] that b200_table_advance applies, and its pinned double-buffered mirror
        self.max_table_updates = 4 * max_batch_size
        self._updates_dev = torch.zeros(1 + 3 * self.max_table_updates, dtype=torch.int32, device=device) \
            if self._on_gpu else None
        self._upd_stages = []
        for _ in range(2):
            up = torch.zeros(1 + 3 * self.max_table_updates, dtype=torch.int32, pin_memory=self._on_gpu)
            self._upd_stages.append({"t": up, "np": up.numpy(), "event": None})
        self._upd_idx = 0
        self._updates_pending = False       # an uploaded update list no device launch has consumed yet

    # ------------------------------------------------------------------ allocator
    def alloc(self, seq_id: int, n_tokens: int) -> Tuple[bool, Optional[torch.Tensor]]:
        bs = self.block_size
        need = -(-n_tokens // bs)
        if len(self._free) < need:
            return False, None
        blocks = [self._free.popleft() for _ in range(need)]
        for b in blocks:
            self._pos[b] = bs
        if blocks:
            self._pos[blocks[-1]] = n_tokens % bs            # 0 when n_tokens is a multiple of bs (reference quirk)
        self._blocks[seq_id] = blocks
        idx = np.arange(n_tokens, dtype=np.int64)
        if n_tokens:
            slots = np.asarray(blocks, dtype=np.int64)[idx // bs] * bs + idx % bs
        else:
            slots = idx
        slot_mapping = torch.from_numpy(slots.astype(np.int32))
        if torch.device(self.device).type != "cpu":
            slot_mapping = slot_mapping.to(self.device)
        return True, slot_mapping

    def append(self, seq_id: int) -> int:
        blocks = self._blocks.get(seq_id)
        if blocks is None:
            return -1
        last = blocks[-1]
        used = self._pos[last]
        if used < self.block_size - 1:                        # offsets 0..254 only (reference quirk)
            self._pos[last] = used + 1
            return last * self.block_size + used
        if self._free:
            new = self._free.popleft()
            blocks.append(new)
            self._pos[new] = 1
            return new * self.block_size
        print(f"cache_manager.append failed for seq {seq_id}: blocks={blocks} pos={used} "
              f"block_size={self.block_size} free={len(self._free)}")
        return -1

    def free(self, seq_id: int) -> None:
        blocks = self._blocks.pop(seq_id, None)
        if blocks is None:
            return
        for b in blocks:
            self._free.append(b)
            self._pos.pop(b, None)

    def reset(self) -> None:
        self._free = collections.deque(range(self.n_blocks))
        self._blocks.clear()
        self._pos.clear()

    # ------------------------------------------------------------------ cache access
    def put(self, seq_id: int, key: torch.Tensor, value: torch.Tensor, layer: int, slot_map: torch.Tensor) -> None:
        store_kvcache(key, value, self.k_caches[layer], self.v_caches[layer], slot_map, self.block_size)

    def get(self, layer: int, block_id: int = None):
        k, v = self.k_caches[layer], self.v_caches[layer]
        return (k[block_id], v[block_id]) if block_id is not None else (k, v)

    def get_blocks(self, seq_id: int) -> List[int]:
        return self._blocks.get(seq_id, [])                  # live reference, as in the reference (:506)

    def get_slots(self, seq_id: int, token_positions: list) -> list:
        blocks = self._blocks.get(seq_id)
        if blocks is None:
            return [-1] * len(token_positions)
        bs = self.block_size
        out = []
        for pos in token_positions:
            bi, off = divmod(pos, bs)
            if bi >= len(blocks) or off >= self._pos.get(blocks[bi], 0):
                out.append(-1)
            else:
                out.append(blocks[bi] * bs + off)
        return out

    # ------------------------------------------------------------------ per-step device metadata
    def cache_batch_data(self, seq_ids: list, context_lens: list):
        """Refresh rows ``[:B]`` of the static device buffers (block ids, -1 padded) and
        ``cache_seqlens[:B]``; returns views of the same device memory every time
        (reference cache_manager.py:509-546)."""
        B = len(seq_ids)
        if B > 0:
            if self._updates_pending:
                # the full table below supersedes an update list nobody applied; left in place it would later be
                # written over rows of a DIFFERENT batch layout
                self._updates_dev[:1].zero_()
                self._updates_pending = False
            stage = self._stages[self._stage_idx]
            self._stage_idx ^= 1
            if stage["event"] is not None:
                stage["event"].synchronize()
            bt = stage["bt_np"]
            bt[:B] = -1
            for i, sid in enumerate(seq_ids):
                blocks = self._blocks.get(sid)
                if blocks:
                    bt[i, :len(blocks)] = blocks               # raises if a sequence outgrew max_tokens, like the reference
            stage["sl_np"][:B] = context_lens
            # host-side hint for the decode runner: longest / mean context of this batch (ragged batches take
            # the balanced attention schedule)
            tot = float(sum(context_lens))
            self.len_max_over_mean = (max(context_lens) * B / tot) if tot > 0 else 1.0
            self._block_table_buffer[:B].copy_(stage["bt"][:B], non_blocking=True)
            self._cache_seqlens_buffer[:B].copy_(stage["sl"][:B], non_blocking=True)
            if self._on_gpu:
                if stage["event"] is None:
                    stage["event"] = torch.cuda.Event()
                stage["event"].record()
        return self._block_table_buffer[:B], self._cache_seqlens_buffer[:B]

    def append_on_device(self, seq_ids: list, steps: int = 1) -> int:
        """Run ``append`` ``steps`` times for every sequence of the batch on the host (the allocator state stays
        exactly the reference's, cache_manager.py:401-438) and ship ONLY what changed to the device: each page the
        calls handed out is one ``(row, column, block id)`` entry of the device-side update list that
        ``b200_table_advance`` applies (one small async H2D, none at all when no page boundary was crossed).
        The sequence lengths are NOT uploaded: they advance on the device (``advance_lengths`` or the in-graph
        launch of a multi-step decode graph).  Rows must already be laid out by ``cache_batch_data`` in this order.
        Returns the number of new pages; -1 if the pool ran dry (like ``append``)."""
        assert self._on_gpu, "append_on_device: CUDA cache manager only"
        # Dry run first: nothing below may fail after the host allocator has moved, or pages handed to earlier rows
        # would never reach the device table.  A page holds block_size - 1 appended tokens (the reference's offset
        # quirk, see ``append``); a fresh page starts at position 1.
        per_page = self.block_size - 1
        need = 0
        for sid in seq_ids:
            blocks = self._blocks.get(sid)
            if blocks is None:
                return -1
            room = per_page - self._pos[blocks[-1]]
            new_pages = 0 if steps <= room else -(-(steps - room) // per_page)
            if len(blocks) + new_pages > self.max_seq_blocks:
                raise ValueError("sequence outgrew the block table")
            need += new_pages
        if need > len(self._free):
            print(f"cache_manager.append_on_device failed: {need} pages needed for {steps} steps, free={len(self._free)}")
            return -1
        if need > self.max_table_updates:
            raise ValueError(f"{need} new pages exceed the device update list ({self.max_table_updates} entries); use fewer steps")
        if need and self._updates_pending:
            raise RuntimeError("append_on_device: the previous update list has not been applied on the device yet "
                               "(advance_lengths or a multi-step graph replay consumes it); a second upload would "
                               "overwrite pages the device block table never received")
        # step-major like the engine (engine.py:248-262: every step appends to each sequence in batch order), so the
        # free list hands the same page to the same sequence as K single steps would
        before = [len(self._blocks[sid]) for sid in seq_ids]
        for _ in range(steps):
            for sid in seq_ids:
                got = self.append(sid)
                assert got >= 0, "append_on_device: the dry run admitted a request the allocator refused"
        ups = []
        for row, sid in enumerate(seq_ids):
            blocks = self._blocks[sid]
            for col in range(before[row], len(blocks)):
                ups.append((row, col, blocks[col]))
        assert len(ups) == need
        if ups:
            self._updates_pending = True
            st = self._upd_stages[self._upd_idx]
            self._upd_idx ^= 1
            if st["event"] is not None:
                st["event"].synchronize()
            st["np"][0] = len(ups)
            st["np"][1:1 + 3 * len(ups)] = np.asarray(ups, dtype=np.int32).reshape(-1)
            self._updates_dev[:1 + 3 * len(ups)].copy_(st["t"][:1 + 3 * len(ups)], non_blocking=True)
            if st["event"] is None:
                st["event"] = torch.cuda.Event()
            st["event"].record()
        return len(ups)

    def advance_lengths(self, batch_size: int, delta: int = 1) -> None:
        """``cache_seqlens[:B] += delta`` and apply pending table updates, on the device (one tiny launch)."""
        import b200ops
        b200ops.table_advance(self._cache_seqlens_buffer, batch_size, delta, self._block_table_buffer, self._updates_dev)
        self._updates_pending = False

    def mark_updates_applied(self) -> None:
        """Tell the host side that a launch outside this class (the first node of a multi-step decode graph) has been
        enqueued behind the upload and consumes the update list."""
        self._updates_pending = False

    def get_buffer_data(self, seq_ids: list):
        B = len(seq_ids)
        return self._block_table_buffer[:B], self._cache_seqlens_buffer[:B]

    # ------------------------------------------------------------------ monitoring
    @property
    def stats(self) -> dict:
        total, free = self.n_blocks, len(self._free)
        longest = run = 0
        prev = None
        for b in sorted(self._free):
            run = run + 1 if prev is not None and b == prev + 1 else 1
            longest = max(longest, run)
            prev = b
        frag = 1.0 - longest / max(free, 1) if free > 0 else 0.0
        return {"total": total, "free": free, "used": total - free, "frag": round(frag, 3)}
