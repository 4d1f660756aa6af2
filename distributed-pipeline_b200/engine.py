"""Python face of one libdpk handle: plans tiles/buckets for a FlatLayout and
enqueues the kernels on torch's current CUDA stream.

All methods are asynchronous with respect to the device unless their name says
otherwise.  The reference lines each call replaces are listed in
``include/dpk.h``.
"""
import ctypes as C
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib
from .flat import FlatLayout, dtype_code


def _stream_ptr(stream: Optional[torch.cuda.Stream], device) -> int:
    s = stream if stream is not None else torch.cuda.current_stream(device)
    return s.cuda_stream


class _DeviceArray:
    """Minimal __cuda_array_interface__ carrier so torch can alias library memory."""

    def __init__(self, ptr: int, shape: Tuple[int, ...], typestr: str, owner):
        self.__cuda_array_interface__ = {
            "shape": shape, "typestr": typestr, "data": (ptr, False), "version": 3, "strides": None,
        }
        self._owner = owner


class Engine:
    def __init__(self, layout: FlatLayout, dtype: torch.dtype, device: torch.device,
                 ema_rates: Sequence[float] = (), betas=(0.9, 0.999), eps: float = 1e-8,
                 bucket_cap_bytes: int = 128 << 20, tile_elems: int = 0, edge_bucket_bytes: int = 0,
                 frozen: Optional[Sequence[bool]] = None):
        """``frozen[i]``: tensor i has no gradient (``requires_grad=False``): no AdamW step, not part
        of the norm, EMA copies still advanced (what torch.optim.AdamW and update_ema do)."""
        device = torch.device(device)
        if device.type != "cuda":
            raise RuntimeError(
                "the B200 update path has no CPU implementation: parameters must live on a CUDA device")
        if device.index is None:
            device = torch.device("cuda", torch.cuda.current_device())
        if len(ema_rates) > _lib.MAX_EMA:
            raise ValueError(f"at most {_lib.MAX_EMA} EMA rates are supported, got {len(ema_rates)}")
        self.lib = _lib.lib()
        self.layout = layout
        self.dtype = dtype
        self.device = device
        self.ema_rates = [float(r) for r in ema_rates]
        self.betas = (float(betas[0]), float(betas[1]))
        self.eps = float(eps)
        self._h = C.c_void_p()
        n_seg = len(layout.offsets)
        offs = (C.c_int64 * max(n_seg, 1))(*layout.offsets)
        nums = (C.c_int64 * max(n_seg, 1))(*layout.numels)
        rates = (C.c_double * max(len(self.ema_rates), 1))(*self.ema_rates)
        self.frozen = [bool(f) for f in frozen] if frozen is not None else [False] * n_seg
        if len(self.frozen) != n_seg:
            raise ValueError("frozen: one flag per tensor")
        frozen_arr = (C.c_int32 * max(n_seg, 1))(*[int(f) for f in self.frozen]) if any(self.frozen) else None
        cfg = _lib.Config(
            abi_version=_lib.ABI_VERSION, dtype=dtype_code(dtype),
            device=device.index,
            n_segments=n_seg, n_ema=len(self.ema_rates), tile_elems=tile_elems, n_elems=layout.n_elems,
            seg_offsets=offs, seg_numel=nums, ema_rates=rates,
            beta1=self.betas[0], beta2=self.betas[1], eps=self.eps, bucket_cap_bytes=bucket_cap_bytes,
            edge_bucket_bytes=edge_bucket_bytes, seg_frozen=frozen_arr)
        _lib.check(self.lib.dpk_create(C.byref(self._h), C.byref(cfg)))
        self._bound: List[torch.Tensor] = []
        self.world_size = 1
        self.nvls = False
        self.sharded = False
        self._results = None
        self._seg_norms = None

    # -- lifetime -------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self.lib.dpk_destroy(self._h)
            self._h = C.c_void_p()
            self._bound = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def bind(self, params, grads, exp_avg, exp_avg_sq, emas: Sequence[torch.Tensor] = ()):
        tensors = [params, grads, exp_avg, exp_avg_sq, *emas]
        for t in tensors:
            if t.device != self.device or t.dtype != self.dtype or t.numel() != self.layout.n_elems \
                    or not t.is_contiguous():
                raise ValueError("bind: every buffer must be a contiguous flat tensor of the engine's "
                                 "device, dtype and length")
        if len(emas) != len(self.ema_rates):
            raise ValueError("bind: one EMA buffer per rate")
        arr = (C.c_void_p * max(len(emas), 1))(*[e.data_ptr() for e in emas])
        _lib.check(self.lib.dpk_bind(self._h, params.data_ptr(), grads.data_ptr(), exp_avg.data_ptr(),
                                     exp_avg_sq.data_ptr(), arr))
        self._bound = tensors  # keep the storages alive as long as the handle borrows them

    # -- communicator ---------------------------------------------------------
    def init_comm(self, group=None):
        """Create the NCCL communicator, shipping the id through the existing
        torch.distributed process group (set up by ``dist_util.setup_dist``)."""
        import torch.distributed as dist
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        if world == 1:
            return
        box = [None]
        if rank == 0:
            buf = C.create_string_buffer(128)
            _lib.check(self.lib.dpk_comm_unique_id(buf))
            box[0] = bytes(buf.raw)
        dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0,
                                   group=group)
        idbuf = C.create_string_buffer(box[0], 128)
        _lib.check(self.lib.dpk_comm_init(self._h, idbuf, world, rank))
        self.world_size = world

    def enable_nvls(self, grads: torch.Tensor, group=None, n_slots: int = 65536):
        """Route the bucket all-reduce through the library's NVLS kernel.  `grads`
        is the bound gradient buffer and must come from symmetric memory
        (FlatState(symmetric_grads=True)).  Collective: every rank calls it."""
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        group = group if group is not None else dist.group.WORLD
        if grads.data_ptr() != self._bound[1].data_ptr():
            raise ValueError("enable_nvls: not the bound gradient buffer")
        hdl = symm_mem.rendezvous(grads, group)
        sig = symm_mem.empty(n_slots, dtype=torch.int32, device=self.device)
        sig.zero_()
        hsig = symm_mem.rendezvous(sig, group)
        if not hdl.multicast_ptr or not hsig.multicast_ptr:
            raise RuntimeError("enable_nvls: no NVSwitch multicast support on this system")
        mc_grads = hdl.multicast_ptr + (grads.data_ptr() - hdl.buffer_ptrs[hdl.rank])
        mc_sig = hsig.multicast_ptr + (sig.data_ptr() - hsig.buffer_ptrs[hsig.rank])
        torch.cuda.synchronize(self.device)
        dist.barrier(group)          # every rank's counters are zero before anyone signals
        off = grads.data_ptr() - hdl.buffer_ptrs[hdl.rank]
        peers = (C.c_void_p * hdl.world_size)(*[p + off for p in hdl.buffer_ptrs])
        _lib.check(self.lib.dpk_bind_nvls(self._h, mc_grads, peers, mc_sig, sig.data_ptr(), n_slots))
        self._nvls = (hdl, hsig, sig)  # keep the mappings alive
        self.nvls = True

    def enable_shard(self, params: torch.Tensor, group=None):
        """Sharded update (dpk_shard_update): map the (symmetric) parameter buffer on the NVSwitch.
        Collective; after enable_nvls."""
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        group = group if group is not None else dist.group.WORLD
        if params.data_ptr() != self._bound[0].data_ptr():
            raise ValueError("enable_shard: not the bound parameter buffer")
        hdl = symm_mem.rendezvous(params, group)
        if not hdl.multicast_ptr:
            raise RuntimeError("enable_shard: no NVSwitch multicast support on this system")
        mc = hdl.multicast_ptr + (params.data_ptr() - hdl.buffer_ptrs[hdl.rank])
        torch.cuda.synchronize(self.device)
        dist.barrier(group)
        _lib.check(self.lib.dpk_bind_shard(self._h, mc))
        self._shard = hdl
        self.sharded = True

    def shard_update(self, lr: float, weight_decay: float, adam_step: int, inv_scale: float = 1.0, stream=None):
        _lib.check(self.lib.dpk_shard_update(self._h, lr, weight_decay, adam_step, inv_scale,
                                             _stream_ptr(stream, self.device)))

    def shard_gather(self, stream=None):
        """Make exp_avg, exp_avg_sq and the EMA copies complete on every rank again (no-op if they are)."""
        _lib.check(self.lib.dpk_shard_gather(self._h, _stream_ptr(stream, self.device)))

    def enable_nvls_agreed(self, flat, group=None, n_slots: int = 65536):
        """Switch the gradient all-reduce to the library's symmetric-memory kernels only if EVERY
        rank can: the decision is taken collectively (MIN over ranks before and after the set-up),
        because a rank that quietly stayed on NCCL while its peers spin in the symmetric-memory
        kernels would hang the job.  Returns (enabled, reason-if-not).  Collective."""
        import torch.distributed as dist
        group = group if group is not None else dist.group.WORLD

        def all_ok(flag: bool) -> bool:
            t = torch.tensor([1 if flag else 0], device=self.device, dtype=torch.int32)
            dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
            return bool(t.item())

        why = None
        if not flat.symmetric:
            why = f"symmetric allocation failed on this rank: {flat.symmetric_error}"
        if not all_ok(flat.symmetric):
            return False, why or "symmetric allocation failed on another rank"
        # every rank holds symmetric memory: the rendezvous below is entered by all of them
        try:
            self.enable_nvls(flat.grads, group=group, n_slots=n_slots)
        except Exception as exc:  # no multicast support, mapping failure, ...
            why = str(exc)
        if not all_ok(self.nvls):
            self.nvls = False
            return False, why or "symmetric-memory set-up failed on another rank"
        return True, None

    @property
    def buckets(self) -> List[Tuple[int, int]]:
        out = []
        for b in range(self.lib.dpk_num_buckets(self._h)):
            s, c = C.c_int64(), C.c_int64()
            _lib.check(self.lib.dpk_bucket_range(self._h, b, C.byref(s), C.byref(c)))
            out.append((s.value, c.value))
        return out

    def bucket_ready(self, b: int, stream=None):
        _lib.check(self.lib.dpk_bucket_ready(self._h, b, _stream_ptr(stream, self.device)))

    def all_buckets_ready(self, stream=None):
        _lib.check(self.lib.dpk_all_buckets_ready(self._h, _stream_ptr(stream, self.device)))

    def wait_bucket(self, b: int, stream=None):
        _lib.check(self.lib.dpk_wait_bucket(self._h, b, _stream_ptr(stream, self.device)))

    def wait_allreduce(self, stream=None):
        _lib.check(self.lib.dpk_wait_allreduce(self._h, _stream_ptr(stream, self.device)))

    def broadcast_state(self, root: int = 0, include_ema: bool = True, stream=None):
        _lib.check(self.lib.dpk_broadcast_state(self._h, root, int(include_ema), _stream_ptr(stream, self.device)))

    # -- the update -----------------------------------------------------------
    def zero_grads(self, stream=None):
        _lib.check(self.lib.dpk_zero_grads(self._h, _stream_ptr(stream, self.device)))

    def grad_norm(self, inv_scale: float = 1.0, max_norm: float = 0.0, stream=None):
        _lib.check(self.lib.dpk_grad_norm(self._h, inv_scale, max_norm, _stream_ptr(stream, self.device)))

    def fused_update(self, lr: float, weight_decay: float, adam_step: int, inv_scale: float = 1.0,
                     use_clip_coef: bool = False, skip_if_nonfinite: bool = False,
                     grad_mode: int = _lib.GRAD_KEEP, stream=None):
        _lib.check(self.lib.dpk_fused_update(self._h, lr, weight_decay, adam_step, inv_scale, int(use_clip_coef),
                                             int(skip_if_nonfinite), grad_mode, _stream_ptr(stream, self.device)))

    # -- the same two passes, one all-reduce bucket at a time ---------------------
    def grad_norm_bucket(self, bucket: int, inv_scale: float = 1.0, stream=None):
        _lib.check(self.lib.dpk_grad_norm_bucket(self._h, bucket, inv_scale, _stream_ptr(stream, self.device)))

    def grad_norm_finish(self, max_norm: float = 0.0, stream=None):
        _lib.check(self.lib.dpk_grad_norm_finish(self._h, max_norm, _stream_ptr(stream, self.device)))

    def fused_update_bucket(self, bucket: int, lr: float, weight_decay: float, adam_step: int,
                            inv_scale: float = 1.0, use_clip_coef: bool = False, skip_if_nonfinite: bool = False,
                            grad_mode: int = _lib.GRAD_KEEP, stream=None):
        _lib.check(self.lib.dpk_fused_update_bucket(self._h, bucket, lr, weight_decay, adam_step, inv_scale,
                                                    int(use_clip_coef), int(skip_if_nonfinite), grad_mode,
                                                    _stream_ptr(stream, self.device)))

    def fused_update_finish(self, skip_if_nonfinite: bool = False, stream=None):
        _lib.check(self.lib.dpk_fused_update_finish(self._h, int(skip_if_nonfinite),
                                                    _stream_ptr(stream, self.device)))

    def reduce_update_pair(self, lr: float, weight_decay: float, adam_step: int, inv_scale: float = 1.0, stream=None):
        """Two ranks: all-reduce and update as one kernel (see dpk_reduce_update_pair)."""
        _lib.check(self.lib.dpk_reduce_update_pair(self._h, lr, weight_decay, adam_step, inv_scale,
                                                   _stream_ptr(stream, self.device)))

    def reduce_update(self, lr: float, weight_decay: float, adam_step: int, inv_scale: float = 1.0, stream=None):
        """Any world size >= 2: all-reduce and update as one kernel (see dpk_reduce_update)."""
        _lib.check(self.lib.dpk_reduce_update(self._h, lr, weight_decay, adam_step, inv_scale,
                                              _stream_ptr(stream, self.device)))

    def reduce_norm(self, inv_scale: float = 1.0, max_norm: float = 0.0, stream=None):
        """Any world size >= 2: all-reduce and the norm pass together (see dpk_reduce_norm)."""
        _lib.check(self.lib.dpk_reduce_norm(self._h, inv_scale, max_norm, _stream_ptr(stream, self.device)))

    @property
    def results(self) -> torch.Tensor:
        """Device tensor (float64[R_COUNT]) aliasing the library's result block."""
        if self._results is None:
            p = C.c_void_p()
            _lib.check(self.lib.dpk_result_ptr(self._h, C.byref(p)))
            with torch.cuda.device(self.device):
                self._results = torch.as_tensor(_DeviceArray(p.value, (_lib.R_COUNT,), "<f8", self),
                                                device=self.device)
        return self._results

    @property
    def segment_norms(self) -> torch.Tensor:
        if self._seg_norms is None:
            p = C.c_void_p()
            _lib.check(self.lib.dpk_segment_norms_ptr(self._h, C.byref(p)))
            n = max(len(self.layout.offsets), 1)
            with torch.cuda.device(self.device):
                self._seg_norms = torch.as_tensor(_DeviceArray(p.value, (n,), "<f8", self), device=self.device)
        return self._seg_norms[:len(self.layout.offsets)]

    def set_option(self, option: int, value: int):
        _lib.check(self.lib.dpk_set_option(self._h, option, value))

    # -- instrumentation -------------------------------------------------------
    def timing(self, on: bool):
        _lib.check(self.lib.dpk_timing_enable(self._h, int(on)))

    def read_timings_sync(self, which: int = 0) -> List[float]:
        """Per-launch device durations (ms) of the fused-update (0) or sum-of-squares (1) kernel."""
        buf = (C.c_float * 256)()
        n = C.c_int32()
        _lib.check(self.lib.dpk_timing_read_sync(self._h, which, buf, 256, C.byref(n)))
        return list(buf[:n.value])

    @property
    def launch_count(self) -> int:
        out = C.c_int64()
        _lib.check(self.lib.dpk_launch_count(self._h, C.byref(out)))
        return out.value

    def read_results_sync(self, stream=None) -> List[float]:
        out = (C.c_double * _lib.R_COUNT)()
        _lib.check(self.lib.dpk_read_result_sync(self._h, out, _stream_ptr(stream, self.device)))
        return list(out)


def ema_update(target: torch.Tensor, source: torch.Tensor, rate: float, stream=None):
    """``trg.mul_(rate).add_(src, alpha=1-rate)`` (reference utils/trainer.py:370) as one kernel."""
    if target.device.type != "cuda" or source.device != target.device:
        raise RuntimeError("ema_update: tensors must be on the same CUDA device (no CPU implementation)")
    if target.dtype != source.dtype or target.numel() != source.numel():
        raise ValueError("ema_update: dtype / size mismatch")
    if not (target.is_contiguous() and source.is_contiguous()):
        raise ValueError("ema_update: tensors must be contiguous")
    with torch.cuda.device(target.device):
        _lib.check(_lib.lib().dpk_ema_update(target.data_ptr(), source.data_ptr(), target.numel(),
                                             dtype_code(target.dtype), float(rate),
                                             _stream_ptr(stream, target.device)))


def plan_buckets(layout: FlatLayout, dtype: torch.dtype, bucket_cap_bytes: int = 128 << 20,
                 edge_bucket_bytes: int = 0) -> List[Tuple[int, int]]:
    """The library's all-reduce bucket plan for a layout, as (start, count) flat
    element ranges, tail first.  Host-only: works without a GPU."""
    lib = _lib.lib()
    n_seg = len(layout.offsets)
    offs = (C.c_int64 * max(n_seg, 1))(*layout.offsets)
    nums = (C.c_int64 * max(n_seg, 1))(*layout.numels)
    cfg = _lib.Config(abi_version=_lib.ABI_VERSION, dtype=dtype_code(dtype), device=0, n_segments=n_seg, n_ema=0,
                      tile_elems=0, n_elems=layout.n_elems, seg_offsets=offs, seg_numel=nums, ema_rates=None,
                      beta1=0.9, beta2=0.999, eps=1e-8, bucket_cap_bytes=bucket_cap_bytes,
                      edge_bucket_bytes=edge_bucket_bytes)
    h = C.c_void_p()
    _lib.check(lib.dpk_create(C.byref(h), C.byref(cfg)))
    try:
        out = []
        for b in range(lib.dpk_num_buckets(h)):
            s, c = C.c_int64(), C.c_int64()
            _lib.check(lib.dpk_bucket_range(h, b, C.byref(s), C.byref(c)))
            out.append((s.value, c.value))
        return out
    finally:
        lib.dpk_destroy(h)
