"""Pure-Python restatement of the reference allocator (test infrastructure only).

Follows core/cache_manager.py: state :327-349, alloc :351-399, append :401-438,
get_slots :563-614, free :616-636, reset :638-652, stats :654-691, cache_batch_data :509-546.
No tensors for the cache pages — only the integer bookkeeping, returned as Python lists so a
test can compare it bit-for-bit with the product class and with the golden traces recorded
from the reference itself.
"""
import collections


class AllocatorOracle:
    def __init__(self, n_blocks, block_size, max_tokens=1024, max_batch_size=32):
        self.n_blocks, self.block_size = n_blocks, block_size
        self.free_q = collections.deque(range(n_blocks))                       # :327
        self.blocks = {}                                                        # :330
        self.pos = {}                                                           # :333
        self.max_seq_blocks = (max_tokens + block_size - 1) // block_size       # :341
        self.block_table = [[-1] * self.max_seq_blocks for _ in range(max_batch_size)]   # :343-346
        self.seqlens = [0] * max_batch_size                                     # :347-349

    def alloc(self, seq_id, n_tokens):
        need = (n_tokens + self.block_size - 1) // self.block_size              # :370
        if len(self.free_q) < need:                                             # :373
            return False, None
        blocks = [self.free_q.popleft() for _ in range(need)]                   # :377
        for i, b in enumerate(blocks):                                          # :384-387
            self.pos[b] = n_tokens % self.block_size if i == len(blocks) - 1 else self.block_size
        self.blocks[seq_id] = blocks                                            # :390
        slots = [blocks[i // self.block_size] * self.block_size + i % self.block_size
                 for i in range(n_tokens)]                                      # :394-397
        return True, slots

    def append(self, seq_id):
        if seq_id not in self.blocks:                                           # :417
            return -1
        blocks = self.blocks[seq_id]
        last = blocks[-1]
        cur = self.pos[last]
        if cur < self.block_size - 1:                                           # :425  (offset 255 is never handed out)
            self.pos[last] += 1
            return last * self.block_size + cur
        elif self.free_q:                                                       # :430
            new = self.free_q.popleft()
            blocks.append(new)
            self.pos[new] = 1                                                   # :434
            return new * self.block_size
        return -1                                                               # :438

    def free(self, seq_id):
        if seq_id in self.blocks:                                               # :629
            for b in self.blocks[seq_id]:
                self.free_q.append(b)                                           # :632
                self.pos.pop(b, None)
            del self.blocks[seq_id]

    def reset(self):
        self.free_q = collections.deque(range(self.n_blocks))
        self.blocks.clear()
        self.pos.clear()

    def get_slots(self, seq_id, token_positions):
        if seq_id not in self.blocks:
            return [-1] * len(token_positions)
        blocks, out = self.blocks[seq_id], []
        for p in token_positions:
            bi = p // self.block_size
            if bi >= len(blocks):
                out.append(-1)
                continue
            off = p % self.block_size
            out.append(-1 if off >= self.pos.get(blocks[bi], 0) else blocks[bi] * self.block_size + off)
        return out

    def cache_batch_data(self, seq_ids, context_lens):
        B = len(seq_ids)
        for i in range(B):                                                      # :517
            self.block_table[i] = [-1] * self.max_seq_blocks
        for i, sid in enumerate(seq_ids):                                       # :532-536
            blocks = self.blocks.get(sid, [])
            if len(blocks) > self.max_seq_blocks:
                raise ValueError("sequence outgrew the block table")
            self.block_table[i][:len(blocks)] = blocks
        self.seqlens[:B] = list(context_lens)                                   # :539-540
        return [row[:] for row in self.block_table[:B]], self.seqlens[:B]

    def stats(self):
        total, free = self.n_blocks, len(self.free_q)
        best = cur = 0
        s = sorted(self.free_q)
        for i, b in enumerate(s):                                               # :672-681
            if i == 0 or b == s[i - 1] + 1:
                cur += 1
            else:
                best = max(best, cur)
                cur = 1
        best = max(best, cur)
        frag = 1.0 - best / max(free, 1) if free > 0 else 0.0
        return {"total": total, "free": free, "used": total - free, "frag": round(frag, 3)}

    def snapshot(self):
        return {"free": list(self.free_q), "blocks": {int(k): list(v) for k, v in self.blocks.items()},
                "pos": {int(k): int(v) for k, v in self.pos.items()}}
