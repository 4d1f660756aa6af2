"""Process-group helpers with the interface of the reference's
``basic_utils/dist_util.py`` (same names, same argument meaning): the drop-in
``TrainLoop`` and the reference's ``run/train.py`` can use either module.

Behaviour kept from the reference: everything degrades to a single process
when ``torch.distributed`` is unavailable or ``LOCAL_RANK`` is unset
(``dist_util.py:66,82-85,92-106``); the backend is NCCL when CUDA is available,
else gloo (``:69``); ``dev()`` is ``cuda:{LOCAL_RANK}`` (``:109-115``).
What changes: ``sync_params`` on a list that is backed by one flat buffer is a
single coalesced broadcast instead of one collective per tensor (``:141-152``).
"""
import io
import os
import socket

import torch
import torch.distributed as dist

_state = {"setup_done": None}


def is_available() -> bool:
    """Whether torch was built with the c10d runtime."""
    ok = bool(dist.is_available())
    if not ok:
        os.environ.setdefault("LOCAL_RANK", "0")
    return ok


def is_initialized() -> bool:
    return is_available() and dist.is_initialized()


def setup_dist(backend=None, silent=False) -> bool:
    """Create the default process group from the ``env://`` variables torchrun
    sets.  Idempotent.  Returns True when running distributed."""
    if is_initialized():
        return True
    if _state["setup_done"] is not None:
        return _state["setup_done"]
    result = False
    # the reference's condition (basic_utils/dist_util.py:66): LOCAL_RANK alone decides; an environment
    # that lacks the other env:// variables fails inside init_process_group and falls back below
    if is_available() and os.environ.get("LOCAL_RANK") is not None:
        try:
            if backend is None:
                backend = "nccl" if torch.cuda.is_available() else "gloo"
            if torch.cuda.is_available():
                torch.cuda.set_device(dev())
            dist.init_process_group(backend=backend, init_method="env://")
            if os.environ["LOCAL_RANK"] == "0" and not silent:
                print("<INFO> torch.distributed setup success, using distributed setting..")
            result = True
        except Exception as exc:  # same policy as the reference: fall back to one process
            if not silent:
                print(f"<INFO> {type(exc).__qualname__}: {exc}")
    if not result:
        os.environ.setdefault("LOCAL_RANK", "0")
        if os.environ["LOCAL_RANK"] == "0" and not silent:
            print("<INFO> torch.distributed is not available, skipping distributed setting..")
    _state["setup_done"] = result
    return result


def get_rank(group=None) -> int:
    return dist.get_rank(group=group) if is_initialized() else 0


def get_world_size(group=None) -> int:
    return dist.get_world_size(group=group) if is_initialized() else 1


def barrier(*args, **kwargs):
    if is_initialized():
        return dist.barrier(*args, **kwargs)


def dev() -> torch.device:
    """Device of this rank."""
    if torch.cuda.is_available():
        return torch.device(f"cuda:{os.environ.get('LOCAL_RANK', '0')}")
    return torch.device("cpu")


def load_state_dict(path, **kwargs):
    """torch.load of a local checkpoint file (the reference goes through
    blobfile for remote paths; only local paths are supported here)."""
    with open(path, "rb") as f:
        data = f.read()
    return torch.load(io.BytesIO(data), **kwargs)


def broadcast(tensor, src=0, group=None, async_op=False):
    if not is_initialized():
        return
    with torch.no_grad():
        return dist.broadcast(tensor, src, group=group, async_op=async_op)


def _common_flat_base(params):
    """If `params` are consecutive views of one contiguous 1-D tensor, return
    the covering slice of it, else None."""
    params = list(params)
    if not params:
        return None
    base = params[0]._base if params[0]._base is not None else None
    if base is None or base.dim() != 1 or not base.is_contiguous():
        return None
    lo, hi = None, None
    for p in params:
        if p._base is not base or not p.is_contiguous():
            return None
        start = p.storage_offset() - base.storage_offset()
        lo = start if lo is None else min(lo, start)
        hi = start + p.numel() if hi is None else max(hi, start + p.numel())
    return base[lo:hi]


def sync_params(params, src=0, group=None, async_op=False):
    """Make every rank's tensors equal to rank `src`'s."""
    if not is_initialized():
        return
    params = list(params)
    flat = _common_flat_base(params)
    if flat is not None:
        return broadcast(flat, src, group=group, async_op=async_op)
    for p in params:
        broadcast(p, src, group=group, async_op=async_op)


def find_free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        s.bind(("", 0))
        return s.getsockname()[1]
