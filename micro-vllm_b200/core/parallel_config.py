"""Tensor-parallel runtime: one process per GPU, torch.distributed over NCCL (NVLink 5 /
NVSwitch) on CUDA and gloo on CPU.

Drop-in for the reference's ``core/parallel_config.py`` (:14-79): ``setup``, ``get_group``,
``get_rank``, ``get_world_size``, ``rank0``, ``all_reduce``, ``barrier`` with the same
behaviour — single-process runs never touch torch.distributed, ``all_reduce`` is an
in-place SUM that returns its argument, ``rank0`` trusts ``$RANK`` first.
Attention needs no collective (it shards by KV head, engine.py:108-114); the only data-plane
collectives on the path are the two all-reduces after ``o_proj`` and ``down_proj``
(model_graph.py:210, :219).
"""
import os

import torch
import torch.distributed as dist

_state = {"world_group": None, "parallel_group": None, "is_main_rank": False, "peer_comm": None}


def setup():
    rank = int(os.environ.get("RANK", 0))
    world_size = int(os.environ.get("WORLD_SIZE", 1))
    if world_size > 1 and not dist.is_initialized():
        use_cuda = torch.cuda.is_available()
        if use_cuda:
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)) % max(1, torch.cuda.device_count()))
        dist.init_process_group(backend="nccl" if use_cuda else "gloo", rank=rank, world_size=world_size)
    if dist.is_initialized():
        if _state["world_group"] is None:
            _state["world_group"] = dist.group.WORLD
        if _state["parallel_group"] is None:
            _state["parallel_group"] = dist.new_group(ranks=list(range(get_world_size())))
    _state["is_main_rank"] = rank0()


def get_group():
    return _state["parallel_group"] if dist.is_initialized() else None


def get_rank():
    return dist.get_rank() if dist.is_initialized() else 0


def get_world_size():
    return dist.get_world_size() if dist.is_initialized() else 1


def rank0():
    env = os.environ.get("RANK")
    if env is not None:
        return int(env) == 0
    return get_rank() == 0


def all_reduce(input_):
    if get_world_size() == 1:
        return input_
    dist.all_reduce(input_, group=get_group() or dist.group.WORLD)
    return input_


def barrier():
    if get_world_size() > 1 and dist.is_initialized():
        dist.barrier(group=get_group() or dist.group.WORLD)


def peer_comm(slot_bytes: int, gather_bytes: int = 0):
    """The node-local NVLink peer-memory communicator behind the fused all-reduce + residual +
    RMSNorm kernel (``b200ops.comm.PeerComm``), created on first use and grown on demand (``gather_bytes``:
    room for the logits all-gather of a vocab-parallel lm_head); ``None``
    for single-process runs, on CPU (gloo), for world sizes other than 2 / 4 / 8, or when
    ``B200_FUSED_ALLREDUCE=0`` selects the plain NCCL ``all_reduce`` + elementwise kernels
    (the reference's sequence, model_graph.py:210-219)."""
    if get_world_size() not in (2, 4, 8) or not torch.cuda.is_available():
        return None
    if os.environ.get("B200_FUSED_ALLREDUCE", "1") == "0":
        return None
    comm = _state["peer_comm"]
    if comm is None or comm.slot_bytes < slot_bytes or comm.gather_bytes < gather_bytes:
        from b200ops.comm import PeerComm
        # an outgrown communicator stays mapped: captured graphs may hold its addresses
        comm = PeerComm(max(slot_bytes, comm.slot_bytes if comm is not None else 0),
                        group=get_group() or dist.group.WORLD,
                        gather_bytes=max(gather_bytes, comm.gather_bytes if comm is not None else 0))
        _state["peer_comm"] = comm
    return comm
