"""Flat, aligned state buffers behind a model's parameters.

The reference keeps 209 separate tensors per role (params, grads, two AdamW
moments, K EMA lists; ``utils/trainer.py:91-113``) and walks them in Python.
Here each role is ONE contiguous device buffer; every ``nn.Parameter`` (and its
``.grad``, moments and EMA copies) is a view into it, in
``model.parameters()`` order, so the reference's list-of-tensors attributes
(``model_params``, ``master_params``, ``ema_params``) and its checkpoint code
keep working unchanged while the kernels stream over whole buffers.
"""
from dataclasses import dataclass
from typing import List, Sequence

import torch

ALIGN_BYTES = 128  # each tensor starts on a cache line (the C ABI needs only 16)


@dataclass(frozen=True)
class FlatLayout:
    """Where each parameter tensor lives inside the flat buffers."""
    offsets: tuple      # element offset of every tensor
    numels: tuple       # elements of every tensor
    n_elems: int        # flat length, padding included
    elem_size: int

    @staticmethod
    def from_numels(numels: Sequence[int], elem_size: int, align_bytes: int = ALIGN_BYTES) -> "FlatLayout":
        if align_bytes % 16 or align_bytes % elem_size:
            raise ValueError("alignment must be a multiple of 16 bytes and of the element size")
        step = align_bytes // elem_size
        offsets, pos = [], 0
        for n in numels:
            if n < 0:
                raise ValueError("negative numel")
            offsets.append(pos)
            pos = (pos + int(n) + step - 1) // step * step
        return FlatLayout(tuple(offsets), tuple(int(n) for n in numels), pos, elem_size)

    @property
    def n_params(self) -> int:
        return sum(self.numels)

    def views(self, flat: torch.Tensor, shapes: Sequence[torch.Size]) -> List[torch.Tensor]:
        if flat.numel() != self.n_elems:
            raise ValueError("flat buffer has the wrong length")
        return [flat[o:o + n].view(s) for o, n, s in zip(self.offsets, self.numels, shapes)]

    def segment_of(self, start: int, count: int) -> List[int]:
        """Indices of the tensors that begin inside flat range [start, start+count).  A trailing
        empty tensor sits at offset == n_elems: it belongs to the range that ends there."""
        end = start + count
        return [i for i, o in enumerate(self.offsets)
                if start <= o < end or (o == end == self.n_elems and self.numels[i] == 0)]


_DTYPES = {torch.float32: 0, torch.bfloat16: 1}


def dtype_code(dtype: torch.dtype) -> int:
    try:
        return _DTYPES[dtype]
    except KeyError:
        raise TypeError(f"state dtype {dtype} is not supported (float32 or bfloat16)") from None


class FlatState:
    """Owns the flat buffers and re-points a model's parameters into them.

    After construction ``p.data`` and ``p.grad`` of every parameter are views of
    ``self.params`` / ``self.grads``; gradients therefore accumulate in place
    across microbatches exactly as with the reference's ``zero_()``-not-``None``
    policy (``utils/trainer.py:203-207``).
    """

    def __init__(self, parameters: Sequence[torch.nn.Parameter], n_ema: int, symmetric_grads: bool = False,
                 symmetric_params: bool = False):
        """``symmetric_grads``: allocate the gradient buffer from torch's symmetric
        memory allocator so that it can later be mapped on every rank and bound to
        an NVSwitch multicast address (Engine.enable_nvls); collective on the
        default process group at rendezvous time, plain device memory until then."""
        self.param_list = list(parameters)
        if not self.param_list:
            raise ValueError("no parameters")
        first = self.param_list[0]
        self.device, self.dtype = first.device, first.dtype
        dtype_code(self.dtype)
        for p in self.param_list:
            if p.device != self.device or p.dtype != self.dtype:
                raise TypeError("all parameters must share one device and dtype")
        self.shapes = [p.shape for p in self.param_list]
        self.frozen = [not p.requires_grad for p in self.param_list]
        self.layout = FlatLayout.from_numels([p.numel() for p in self.param_list], first.element_size())
        n = self.layout.n_elems

        def buf():
            return torch.zeros(n, dtype=self.dtype, device=self.device)

        self.params = buf()
        self.symmetric_params = False
        if symmetric_params:
            # the sharded update multicasts new parameters into every replica: the buffer must be mappable
            import torch.distributed._symmetric_memory as symm_mem
            self.params = symm_mem.empty(n, dtype=self.dtype, device=self.device)
            self.params.zero_()
            self.symmetric_params = True
        self.symmetric = False
        self.symmetric_error = None
        if symmetric_grads:
            try:
                import torch.distributed._symmetric_memory as symm_mem
                self.grads = symm_mem.empty(n, dtype=self.dtype, device=self.device)
                self.grads.zero_()
                self.symmetric = True
            except Exception as exc:  # allocator unavailable: ordinary device memory, NCCL carries the all-reduce
                self.symmetric_error = exc
                self.grads = buf()
        else:
            self.grads = buf()
        self.exp_avg, self.exp_avg_sq = buf(), buf()
        self.emas = [buf() for _ in range(n_ema)]

        self.param_views = self.layout.views(self.params, self.shapes)
        self.grad_views = self.layout.views(self.grads, self.shapes)
        self.exp_avg_views = self.layout.views(self.exp_avg, self.shapes)
        self.exp_avg_sq_views = self.layout.views(self.exp_avg_sq, self.shapes)
        self.ema_views = [self.layout.views(e, self.shapes) for e in self.emas]

        with torch.no_grad():
            for p, pv, gv in zip(self.param_list, self.param_views, self.grad_views):
                pv.copy_(p.data)
                if p.grad is not None:
                    gv.copy_(p.grad)
                p.data = pv
                # a frozen parameter keeps grad=None, as in torch (the optimizer and the norm skip it)
                p.grad = gv if p.requires_grad else None
            # EMA copies start as copies of the parameters (utils/trainer.py:111-113)
            for e in self.emas:
                e.copy_(self.params)

    def recheck_views(self):
        """Raise if something (e.g. ``zero_grad(set_to_none=True)``) detached a view."""
        for p, pv, gv in zip(self.param_list, self.param_views, self.grad_views):
            if p.data_ptr() != pv.data_ptr():
                raise RuntimeError("a parameter no longer aliases the flat buffer")
            if not p.requires_grad:
                continue
            if p.grad is None or p.grad.data_ptr() != gv.data_ptr():
                raise RuntimeError(
                    "a gradient no longer aliases the flat buffer (was it set to None?); "
                    "use TrainLoop._zero_grad / FusedAdamW.zero_grad")
