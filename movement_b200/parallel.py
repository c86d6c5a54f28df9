"""One process per GPU: torch.distributed for rendezvous, the library's own NCCL communicator
for the data path (halo exchange + scalar allreduce inside libmovement_b200.so).

replaces: ``firedrake.COMM_WORLD`` and the implicit MPI traffic of PyOP2/PETSc.
"""
from __future__ import annotations

import ctypes as C
import glob
import os

_state = {"active": False, "rank": 0, "size": 1}


def is_active():
    return _state["active"]


def rank():
    return _state["rank"]


def size():
    return _state["size"]


def find_nccl():
    """Path of the NCCL library torch itself uses (so both share one NCCL build)."""
    try:
        import nvidia.nccl  # type: ignore

        for base in list(getattr(nvidia.nccl, "__path__", [])):
            hits = sorted(glob.glob(os.path.join(base, "lib", "libnccl.so*")))
            if hits:
                return hits[0]
    except Exception:
        pass
    return "libnccl.so.2"


def init(backend=None):
    """Initialise from the torchrun environment (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_*).

    Safe to call on one process (no-op).  With ``backend='gloo'`` (CPU tests) only the host-side
    partition logic is distributed; the NCCL communicator is created when CUDA is available.
    """
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world == 1:
        return False
    r = int(os.environ.get("RANK", "0"))
    lr = int(os.environ.get("LOCAL_RANK", str(r)))
    cuda = torch.cuda.is_available()
    if cuda:
        torch.cuda.set_device(lr)
    if not dist.is_initialized():
        dist.init_process_group(backend or ("nccl" if cuda else "gloo"), rank=r, world_size=world)
    _state.update(active=True, rank=r, size=world)
    if cuda:
        from . import _lib as L

        lib = L.load()
        L.check(lib.mv_set_device(lr))
        path = find_nccl().encode()
        uid = torch.zeros(128, dtype=torch.uint8)
        if r == 0:
            buf = (C.c_ubyte * 128)()
            L.check(lib.mv_comm_unique_id(path, buf))
            uid = torch.tensor(list(buf), dtype=torch.uint8)
        dev = torch.device("cuda", lr) if dist.get_backend() == "nccl" else torch.device("cpu")
        uid = uid.to(dev)
        dist.broadcast(uid, src=0)
        raw = bytes(uid.cpu().tolist())
        L.check(lib.mv_comm_init(path, raw, r, world))
    return True


def finalize():
    import torch.distributed as dist

    if _state["active"]:
        try:
            from . import _lib as L

            if L._lib is not None:
                L._lib.mv_comm_finalize()
        finally:
            _state.update(active=False, rank=0, size=1)
            if dist.is_initialized():
                dist.destroy_process_group()


def allreduce_host(value, op="sum"):
    """Small host-side reduction used by host logic (tag sets, counts)."""
    if not _state["active"]:
        return value
    import torch
    import torch.distributed as dist

    t = torch.tensor([float(value)], dtype=torch.float64)
    if dist.get_backend() == "nccl":
        t = t.cuda()
    dist.all_reduce(t, op={"sum": dist.ReduceOp.SUM, "max": dist.ReduceOp.MAX, "min": dist.ReduceOp.MIN}[op])
    return float(t.item())
