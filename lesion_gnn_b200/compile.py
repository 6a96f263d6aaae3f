"""torch.compile support: the reference wraps every model in `torch.compile(model, dynamic=True)`
(models/gat.py:84, gin.py:56, drgnet.py:95; `compile=True` is its default, configs/config.py:64).

The lgk layers call a C ABI through ctypes, which Dynamo cannot trace.  Each layer (and each pooling / activation
function) is therefore registered as an OPAQUE custom operator:

    lgk::opaque_fwd(key, need, n, tensors, ints) -> tensors       torch.library.custom_op + register_fake + register_autograd
    lgk::opaque_bwd(key, need, ticket, grads, like) -> tensors

`key` names the layer instance (a constant of the traced graph), `tensors` are its tensor inputs followed by its
parameters (so that autograd sees them), the fake implementation derives output shapes from the layer's own shape rule
(symbolic sizes pass through: `dynamic=True`), and the real implementation runs the layer's ordinary eager forward on the
CUDA kernels, keeping its autograd graph under a ticket that the backward operator redeems.  Dynamo / AOTAutograd thus
see ONE node per layer with exact shape propagation and gradients -- `backend="eager"`, `"aot_eager"` or `"inductor"` --
and every arithmetic instruction still executes in liblgk.
"""
from __future__ import annotations

import threading
from collections import OrderedDict
from typing import Callable, List, Optional, Sequence

import torch
from torch import Tensor

_LOCK = threading.Lock()
_ENTRIES: dict = {}        # key -> (eager(*tensor_inputs) -> Tensor | tuple, fake(*tensor_inputs) -> list of shapes/dtypes, params() -> list)
_TICKETS: "OrderedDict[int, tuple]" = OrderedDict()
_NEXT = [1]
_MAX_TICKETS = 4096


def is_compiling() -> bool:
    fn = getattr(torch.compiler, "is_compiling", None)
    return bool(fn()) if fn is not None else False


def register(eager: Callable, fake: Callable, params: Callable[[], Sequence[Tensor]] = lambda: (), key: Optional[int] = None) -> int:
    """-> key of a (new) opaque entry.  eager(*tensors) runs the layer; fake(*tensors) returns the outputs' (shape, dtype,
    device) as tensors created with new_empty(); params() lists the parameters the layer reads."""
    with _LOCK:
        if key is None:
            key = _NEXT[0]
            _NEXT[0] += 1
        _ENTRIES[key] = (eager, fake, params)
    return key


def _as_list(out) -> List[Tensor]:
    return [out] if isinstance(out, Tensor) else list(out)


@torch.library.custom_op("lgk::opaque_fwd", mutates_args=())
def opaque_fwd(key: int, need: int, n_inputs: int, tensors: List[Tensor], ints: List[int]) -> List[Tensor]:
    if key not in _ENTRIES:
        raise RuntimeError("lgk::opaque_fwd: unknown layer key -- call lesion_gnn_b200.compile.prepare(model) after copying / unpickling a model")
    eager, _, params = _ENTRIES[key]
    with torch.enable_grad():
        ins = []
        for i in range(n_inputs):
            t = tensors[i].detach()
            if (need >> i) & 1 and t.is_floating_point():
                t.requires_grad_(True)
            ins.append(t)
        outs = _as_list(eager(*ins, *ints))
    with _LOCK:
        ticket = _NEXT[0]
        _NEXT[0] += 1
        _TICKETS[ticket] = (ins, outs, list(params()))
        while len(_TICKETS) > _MAX_TICKETS:  # forwards whose backward never ran (inference under no_grad is cheaper: see call())
            _TICKETS.popitem(last=False)
    return [o.detach() for o in outs] + [torch.tensor([ticket], dtype=torch.int64)]


@opaque_fwd.register_fake
def _(key, need, n_inputs, tensors, ints):
    fake = _ENTRIES[key][1]
    outs = _as_list(fake(*tensors[:n_inputs], *ints))
    return outs + [torch.empty(1, dtype=torch.int64)]


@torch.library.custom_op("lgk::opaque_bwd", mutates_args=())
def opaque_bwd(key: int, need: int, n_inputs: int, ticket: Tensor, grads: List[Tensor], like: List[Tensor]) -> List[Tensor]:
    with _LOCK:
        ins, outs, params = _TICKETS.pop(int(ticket[0]))
    wrt = [t for i, t in enumerate(ins) if (need >> i) & 1 and t.requires_grad]
    wrt += [p for j, p in enumerate(params) if (need >> (n_inputs + j)) & 1]
    diff = [(o, g) for o, g in zip(outs, grads) if o.requires_grad]
    res = torch.autograd.grad([o for o, _ in diff], wrt, [g for _, g in diff], allow_unused=True) if wrt and diff else ()
    it = iter(res)
    out = []
    for i, t in enumerate(like):
        if (need >> i) & 1 and t.is_floating_point():
            g = next(it)
            out.append(g if g is not None else torch.zeros_like(t))
    return out


@opaque_bwd.register_fake
def _(key, need, n_inputs, ticket, grads, like):
    return [torch.empty_like(t) for i, t in enumerate(like) if (need >> i) & 1 and t.is_floating_point()]


def _setup(ctx, inputs, output):
    key, need, n_inputs, tensors, ints = inputs
    ctx.key, ctx.need, ctx.n_inputs, ctx.n_tensors, ctx.ints = key, need, n_inputs, len(tensors), list(ints)
    ctx.save_for_backward(output[-1], *tensors)


def _backward(ctx, grads):
    ticket, *tensors = ctx.saved_tensors
    gs = []
    for g, _ in zip(grads[:-1], range(len(grads) - 1)):
        gs.append(g)
    # outputs that took no part in the loss arrive as None: the backward operator wants tensors
    outs_like = None
    if any(g is None for g in gs):
        fake = _ENTRIES[ctx.key][1]
        outs_like = _as_list(fake(*tensors[:ctx.n_inputs], *ctx.ints))
        gs = [g if g is not None else torch.zeros_like(o) for g, o in zip(gs, outs_like)]
    res = torch.ops.lgk.opaque_bwd(ctx.key, ctx.need, ctx.n_inputs, ticket, gs, list(tensors))
    it = iter(res)
    full = [next(it) if (ctx.need >> i) & 1 and t.is_floating_point() else None for i, t in enumerate(tensors)]
    return None, None, None, full, ([] if not ctx.ints else None)  # pytree: an empty int list is a node, a filled one a leaf


opaque_fwd.register_autograd(_backward, setup_context=_setup)


def call(key: int, inputs: Sequence[Tensor], params: Sequence[Tensor] = (), ints: Sequence[int] = ()):
    """Run entry `key` as one opaque operator on `inputs` (tensors only; `params`: the parameters the entry reads, passed
    along so that autograd routes their gradients).  Returns the entry's tensor outputs (a single tensor when there is one)."""
    tensors = list(inputs) + list(params)
    need = 0
    if torch.is_grad_enabled():
        for i, t in enumerate(tensors):
            if t.requires_grad:
                need |= 1 << i
    res = torch.ops.lgk.opaque_fwd(key, need, len(inputs), tensors, list(ints))
    outs = res[:-1]
    return outs[0] if len(outs) == 1 else tuple(outs)


def _all_dense(args) -> bool:
    return all(isinstance(a, Tensor) and a.layout == torch.strided for a in args)


class OpaqueModule:
    """Mixin for the lgk layers: `_dispatch` routes to the opaque operator while Dynamo is tracing and every argument is a
    dense tensor, and to `_forward_eager` otherwise.  Subclasses implement `_forward_eager(*tensors)` and `_fake(*tensors)`
    and call `_register_opaque()` at the end of `__init__` (registration is a Python side effect that must not happen in a
    traced frame).  The key is the object's id: a copy (copy.deepcopy, unpickling) has none until `prepare(model)`."""

    def _register_opaque(self) -> int:
        import weakref
        ref = weakref.ref(self)
        return register(lambda *t: ref()._forward_eager(*t), lambda *t: ref()._fake(*t), lambda: list(ref().parameters()), key=id(self))

    def _dispatch(self, *args, ints: Sequence[int] = ()):
        """args: tensors, with trailing None allowed (optional tensor arguments); ints: integer arguments (sizes), which
        reach `_forward_eager` / `_fake` as extra positional arguments"""
        while args and args[-1] is None:
            args = args[:-1]
        if is_compiling() and _all_dense(args):
            return call(id(self), args, list(self.parameters()), ints)
        return self._forward_eager(*args, *ints)


def opaque_function(eager: Callable, fake: Callable) -> Callable:
    """a free function (global_mean_pool, elu, ...) as an opaque operator while compiling; tensors-only positional arguments"""
    key = register(eager, fake)

    def wrapper(*tensors, ints: Sequence[int] = ()):
        while tensors and tensors[-1] is None:
            tensors = tensors[:-1]
        if is_compiling() and _all_dense(tensors):
            return call(key, tensors, (), ints)
        return eager(*tensors, *ints)

    wrapper.__wrapped__ = eager
    return wrapper


def prepare(model: torch.nn.Module) -> torch.nn.Module:
    """(Re)create the opaque entries of every lgk layer of `model` ahead of tracing -- needed after copy.deepcopy / unpickling,
    harmless otherwise.  `torch.compile(prepare(model), dynamic=True)`."""
    for m in model.modules():
        if isinstance(m, OpaqueModule):
            m._register_opaque()
    return model
