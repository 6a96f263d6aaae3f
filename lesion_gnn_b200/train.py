"""The training step of reference models/base.py:196-201 (+ optimiser, base.py:162-170) on the lgk layers, with
the three things that make it fast on one B200 and scale over eight:

  FlatAdam      parameters, gradients and Adam moments live in four flat fp32 buffers; module parameters are
                views, backward accumulates straight into the flat gradient bucket, and one lgk_adam_step
                kernel updates everything (Adam / AdamW, base.py:162-170)
  GraphedStep   H2D copy of the batch + CSR/CSC build + forward + loss + backward (+ optimiser) captured once
                per batch shape into a CUDA graph and replayed: the small-batch configs are launch-bound
  data parallel one NCCL all-reduce over the single flat gradient bucket per step
"""
from __future__ import annotations

import os
from typing import Callable, Optional

import torch
import torch.nn as nn
import torch.nn.functional as TF
from torch import Tensor

from . import ops
from .data import HostBatch
from .graph import CSRGraph
from .nn.conv import dropout_clock
from .nn.functional import direct_param_grads, side_stream
from .nn.pool import Segments

_ALIGN = 32  # floats: every parameter starts on a 128-byte boundary inside the flat buffers


class FlatParams:
    """Parameters and gradients of a module in two flat fp32 buffers (any device): module parameters become
    views, backward accumulates straight into the gradient bucket, and the data-parallel all-reduce is ONE
    collective over `grad`."""

    def __init__(self, module: nn.Module, grad_alloc: Optional[Callable[[int, torch.device], Tensor]] = None):
        """grad_alloc(numel, device) -> the fp32 gradient bucket (default: a plain zeroed buffer; PeerBucket places it in
        memory that the other ranks can read)"""
        params = [p for p in module.parameters() if p.requires_grad]
        if not params:
            raise ValueError("no trainable parameters")
        dev = params[0].device
        offs, total = [], 0
        for p in params:
            if p.dtype != torch.float32:
                raise TypeError("master parameters must be float32")
            offs.append(total)
            total += (p.numel() + _ALIGN - 1) // _ALIGN * _ALIGN
        self.param = torch.zeros(total, dtype=torch.float32, device=dev)
        self.grad = torch.zeros_like(self.param) if grad_alloc is None else grad_alloc(total, dev)
        if self.grad.shape != self.param.shape or self.grad.dtype != torch.float32:
            raise ValueError("grad_alloc must return a float32 tensor of the bucket's length")
        self.params, self.offsets = params, offs
        for p, o in zip(params, offs):
            self.param[o:o + p.numel()].view_as(p).copy_(p.data)
            p.data = self.param[o:o + p.numel()].view_as(p)
            p.grad = self.grad[o:o + p.numel()].view_as(p)
        self.numel = total

    def zero_grad(self) -> None:
        if getattr(self, "_grad_is_zero", False):
            self._grad_is_zero = False  # cleared by the last optimiser step (FlatAdam.step); the backward dirties it
        else:
            self.grad.zero_()
        for p, o in zip(self.params, self.offsets):  # keep .grad bound to the bucket
            if p.grad is None or p.grad.data_ptr() != self.grad.data_ptr() + 4 * o:
                p.grad = self.grad[o:o + p.numel()].view_as(p)

    def all_reduce(self, group=None) -> None:
        """sum the gradient bucket over the data-parallel group (NCCL on GPUs, gloo in the CPU tests)"""
        import torch.distributed as dist
        dist.all_reduce(self.grad, op=dist.ReduceOp.SUM, group=group)


def dp_loss_weight(local_graphs: int, global_graphs: int, world_size: int) -> float:
    """Weight of a rank's mean loss such that (sum over ranks of weighted gradients) / world_size equals the
    gradient of the mean loss over the WHOLE batch, even when ranks hold different numbers of graphs (shards are
    balanced by node count, data.shard_counts)."""
    return float(local_graphs) * world_size / float(global_graphs)


class PeerBucket:
    """The gradient bucket of one rank inside a SYMMETRIC allocation that every peer GPU of the node maps over NVLink
    (torch.distributed._symmetric_memory), laid out as lgk_dp_allreduce_adam wants it (include/lgk.h): flags, the gradients,
    then the slots the peers push their gradients into.  `group=None` gives a single-rank bucket in ordinary device memory (tests, world size 1)."""

    def __init__(self, group=None):
        self.group = group
        self.buffer = self.grad = self.peer_base = self.handle = None
        self.rank, self.world = 0, 1

    def __call__(self, numel: int, device: torch.device) -> Tensor:
        from . import _lib
        L = _lib.lib()
        off = int(L.query("lgk_dp_grad_offset_bytes"))
        if self.group is not None:
            import torch.distributed as dist
            self.world = dist.get_world_size(self.group)
        nbytes = int(L.query("lgk_dp_buffer_bytes", numel, self.world))
        if self.group is None:
            self.buffer = torch.zeros(nbytes, dtype=torch.uint8, device=device)
            self.peer_base = torch.tensor([self.buffer.data_ptr()], dtype=torch.int64, device=device)
            self.peer_base_ptr = self.peer_base.data_ptr()
        else:
            import torch.distributed as dist
            import torch.distributed._symmetric_memory as symm_mem
            self.buffer = symm_mem.empty(nbytes, dtype=torch.uint8, device=device)
            self.handle = symm_mem.rendezvous(self.buffer, self.group)
            self.rank, self.world = self.handle.rank, self.handle.world_size
            self.buffer.zero_()
            torch.cuda.synchronize(device)
            dist.barrier(group=self.group)  # every pad is zero before the first signal can arrive
            self.peer_base_ptr = int(self.handle.buffer_ptrs_dev)
        self.grad = self.buffer[off:off + 4 * numel].view(torch.float32)
        self.epoch_dev = torch.zeros(2, dtype=torch.int64, device=device)  # never rewound (unlike the optimiser's step count)
        return self.grad


class FlatAdam(FlatParams):
    """torch.optim.Adam / AdamW (reference models/base.py:162-170) as one lgk_adam_step launch over the flat
    buffers.  CUDA only.  With `peer_bucket` (a PeerBucket) the gradients live in peer-visible memory and `step_dp()`
    is the data-parallel step: all-reduce over NVLink fused with the update, no NCCL call (csrc/dp.cu)."""

    def __init__(self, module: nn.Module, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0, adamw: bool = False, peer_bucket: Optional["PeerBucket"] = None):
        super().__init__(module, grad_alloc=peer_bucket)
        self.peer_bucket = peer_bucket
        if self.param.device.type != "cuda":
            raise RuntimeError("FlatAdam needs the module on a CUDA device: lesion_gnn_b200 has no CPU path")
        self.exp_avg = torch.zeros_like(self.param)
        self.exp_avg_sq = torch.zeros_like(self.param)
        self.step_dev = torch.zeros(2, dtype=torch.int64, device=self.param.device)  # {steps completed, kernel scratch}
        self.lr, self.betas, self.eps, self.weight_decay, self.adamw = lr, betas, eps, weight_decay, adamw

    def state_tensors(self):
        return [self.param, self.exp_avg, self.exp_avg_sq, self.step_dev]

    def step(self, grad_scale: float = 1.0) -> None:
        """one fused kernel; the step counter lives on the device (and is advanced by the kernel) so the call can be
        graph-captured"""
        if not _ADAM_ADVANCES_STEP:
            self.step_dev[:1].add_(1)
        ops.adam_step(self.param, self.grad, self.exp_avg, self.exp_avg_sq, lr=self.lr, betas=self.betas, eps=self.eps,
                      weight_decay=self.weight_decay, step=0, step_dev=self.step_dev, grad_scale=grad_scale,
                      adamw=self.adamw, zero_grad=_ADAM_ZEROES_GRAD, advance_step=_ADAM_ADVANCES_STEP)
        self._grad_is_zero = _ADAM_ZEROES_GRAD  # the kernel cleared the bucket: the next zero_grad() launches nothing

    def step_dp(self) -> None:
        """data-parallel step: mean of all ranks' gradient buckets (read through NVLink peer mappings, summed in rank order)
        and the Adam update in ONE kernel (pushed over NVLink into per-rank slots, summed in rank order); capturable, no host sync.  Every
        rank of the bucket's group must call it once per step."""
        pb = self.peer_bucket
        if pb is None:
            raise RuntimeError("step_dp() needs FlatAdam(..., peer_bucket=PeerBucket(group))")
        ops.dp_allreduce_adam(pb.peer_base_ptr, pb.rank, pb.world, self.param, self.exp_avg, self.exp_avg_sq, lr=self.lr,
                              betas=self.betas, eps=self.eps, weight_decay=self.weight_decay, step_dev=self.step_dev,
                              epoch_dev=pb.epoch_dev, adamw=self.adamw, advance_step=True)
        self._grad_is_zero = True


class FlatSGD(FlatParams):
    """torch.optim.SGD (reference models/base.py:167-168: lr and weight_decay) as one lgk_sgd_step launch over the flat
    buffers; momentum / nesterov (dampening 0) are available although the reference leaves them at 0.  CUDA only."""

    def __init__(self, module: nn.Module, lr: float = 1e-3, momentum: float = 0.0, weight_decay: float = 0.0,
                 nesterov: bool = False):
        super().__init__(module)
        if self.param.device.type != "cuda":
            raise RuntimeError("FlatSGD needs the module on a CUDA device: lesion_gnn_b200 has no CPU path")
        self.lr, self.momentum, self.weight_decay, self.nesterov = lr, momentum, weight_decay, nesterov
        self.momentum_buf = torch.zeros_like(self.param) if momentum != 0.0 else None
        self.step_dev = torch.zeros(2, dtype=torch.int64, device=self.param.device)  # step count: the dropout clock

    def state_tensors(self):
        return [t for t in (self.param, self.momentum_buf, self.step_dev) if t is not None]

    def step(self, grad_scale: float = 1.0) -> None:
        self.step_dev[:1].add_(1)
        ops.sgd_step(self.param, self.grad, self.momentum_buf, lr=self.lr, momentum=self.momentum,
                     weight_decay=self.weight_decay, nesterov=self.nesterov, grad_scale=grad_scale,
                     zero_grad=_ADAM_ZEROES_GRAD)
        self._grad_is_zero = _ADAM_ZEROES_GRAD


# 0 = data-parallel steps use one NCCL all-reduce of the bucket + lgk_adam_step instead of the fused peer-memory kernel
DP_FUSED = os.environ.get("LGK_DP_FUSED", "1") != "0"


def make_optimizer(module: nn.Module, algo: str = "adam", lr: float = 1e-3, weight_decay: float = 0.0, process_group=None,
                   world_size: int = 1):
    """reference models/base.py:162-170 `configure_optimizers`: OptimizerAlgo ADAM / ADAMW / SGD.  With world_size > 1 the
    Adam variants keep their gradients in a PeerBucket (fused NVLink all-reduce + update) unless LGK_DP_FUSED=0."""
    algo = algo.lower()
    pb = PeerBucket(process_group if process_group is not None else _world_group()) if world_size > 1 and DP_FUSED and algo != "sgd" else None
    if algo == "adam":
        return FlatAdam(module, lr=lr, weight_decay=weight_decay, peer_bucket=pb)
    if algo == "adamw":
        return FlatAdam(module, lr=lr, weight_decay=weight_decay, adamw=True, peer_bucket=pb)
    if algo == "sgd":
        return FlatSGD(module, lr=lr, weight_decay=weight_decay)
    raise ValueError(f"unknown optimizer {algo!r}")


def _world_group():
    import torch.distributed as dist
    return dist.group.WORLD


class _CrossEntropyLoss(torch.autograd.Function):
    """(weighted) cross-entropy and its gradient from one launch (ops.cross_entropy_loss)"""

    @staticmethod
    def forward(ctx, logits, y, weight):
        loss, dl = ops.cross_entropy_loss(logits, y, weight, want_grad=ctx.needs_input_grad[0])
        ctx.save_for_backward(dl)
        return loss

    @staticmethod
    def backward(ctx, g):
        (dl,) = ctx.saved_tensors
        return dl * g, None, None


class _ClampedRegressionLoss(torch.autograd.Function):
    """loss and its gradient from one launch (ops.clamped_regression_loss) instead of a dozen elementwise torch kernels"""

    @staticmethod
    def forward(ctx, logits, y, hi, kind):
        loss, dl = ops.clamped_regression_loss(logits.reshape(-1), y, 0.0, hi, kind, want_grad=ctx.needs_input_grad[0])
        ctx.save_for_backward(dl)
        ctx.shape = logits.shape
        return loss

    @staticmethod
    def backward(ctx, g):
        (dl,) = ctx.saved_tensors
        return (dl * g).view(ctx.shape), None, None, None


def _fused_regression(logits: Tensor, y: Tensor, loss_type: str) -> bool:
    return (loss_type in ("MSE", "SmoothL1") and logits.is_cuda and logits.dtype == torch.float32 and y.dtype == torch.int64
            and logits.dim() == 2 and logits.size(1) == 1)


# profiling knob: 0 = batches whose edges are grouped by destination still take the general CSR/CSC build
_SORTED_BUILD = os.environ.get("LGK_CSR_SORTED", "1") != "0"
# 1 = the Adam kernel advances the step counter itself (its last CTA, after every CTA has read it) instead of a 1-element add
# launch in front of it: headline step 0.2882 -> 0.2861 ms (slower, and off, before the main chain got its priority stream)
_ADAM_ADVANCES_STEP = os.environ.get("LGK_ADAM_ADVANCES_STEP", "1") != "0"
# profiling knob: 0 = zero_grad() is its own fill launch at the head of every step
_ADAM_ZEROES_GRAD = os.environ.get("LGK_ADAM_ZEROES_GRAD", "1") != "0"
# profiling knob: 0 = the loss goes through its autograd node (ones_like seed + multiply by the upstream gradient)
_DIRECT_LOSS_GRAD = os.environ.get("LGK_DIRECT_LOSS_GRAD", "1") != "0"


def _fused_ce(logits: Tensor, y: Tensor) -> bool:
    return logits.is_cuda and logits.dtype == torch.float32 and y.dtype == torch.int64 and logits.dim() == 2


def loss_fn(logits: Tensor, y: Tensor, loss_type: str = "MSE", num_classes: int = 5,
            class_weights: Optional[Tensor] = None) -> Tensor:
    """reference models/base.py:88-96,196-199 and the regression clamp of models/gat.py:94-95; class_weights is the
    `weight` of nn.CrossEntropyLoss (base.py:93-94)"""
    if _fused_regression(logits, y, loss_type):
        return _ClampedRegressionLoss.apply(logits.contiguous(), y.contiguous(), float(num_classes - 1),
                                            0 if loss_type == "MSE" else 1)
    if loss_type == "MSE":
        return TF.mse_loss(torch.clamp(logits.squeeze(1), min=0, max=num_classes - 1), y.float())
    if loss_type == "SmoothL1":
        return TF.smooth_l1_loss(torch.clamp(logits.squeeze(1), min=0, max=num_classes - 1), y.float())
    if loss_type == "CE":
        if _fused_ce(logits, y):
            w = None if class_weights is None else class_weights.to(device=logits.device, dtype=torch.float32).contiguous()
            return _CrossEntropyLoss.apply(logits.contiguous(), y.contiguous(), w)
        return TF.cross_entropy(logits, y, weight=class_weights)
    raise ValueError(loss_type)


class DeviceBatch:
    """static device buffers for one batch shape (what a captured CUDA graph reads)"""

    def __init__(self, hb: HostBatch, device):
        self.num_graphs, self.num_nodes = hb.num_graphs, hb.num_nodes
        self.loss_weight = 1.0  # dp_loss_weight(...) under data parallelism
        self.x = torch.empty_like(hb.x, device=device)
        self.batch = torch.empty_like(hb.batch, device=device)
        self.y = torch.empty_like(hb.y, device=device)
        self.edge_index = torch.empty_like(hb.edge_index, device=device)
        self.edge_weight = torch.empty_like(hb.edge_weight, device=device) if hb.edge_weight is not None else None
        # edges grouped by destination (collated kNN samples are): the CSR comes from boundary detection, not a bucket
        # sort.  A property of the loader's output, decided once on the host; every batch copied into this buffer
        # must share it (copy_from checks).
        self.dst_sorted = bool(getattr(hb, "dst_sorted", False)) and _SORTED_BUILD

    def copy_from(self, hb: HostBatch) -> None:
        if self.dst_sorted and not getattr(hb, "dst_sorted", False):
            raise ValueError("this DeviceBatch was sized for destination-sorted edges; the batch copied in is not marked so")
        self.x.copy_(hb.x, non_blocking=True)
        self.batch.copy_(hb.batch, non_blocking=True)
        self.y.copy_(hb.y, non_blocking=True)
        self.edge_index.copy_(hb.edge_index, non_blocking=True)
        if self.edge_weight is not None:
            self.edge_weight.copy_(hb.edge_weight, non_blocking=True)


def forward_backward(model: nn.Module, db: DeviceBatch, loss_type: str, num_classes: int, build_stream=None,
                     class_weights: Optional[Tensor] = None) -> Tensor:
    """CSR/CSC build + segment pointer + forward + loss + backward for the batch resident in `db`.  With
    `build_stream` the CSR/CSC build runs beside the model's first transform (in_proj needs no graph)."""
    graph = CSRGraph.from_edge_index(db.edge_index, db.num_nodes, db.edge_weight, stream=build_stream,
                                     dst_sorted=getattr(db, "dst_sorted", False))
    seg = Segments.from_batch(db.batch, db.num_graphs)
    logits = model(db.x, graph, seg)
    graph.join()  # a model that never touched the graph must still join the build stream
    if logits.dtype != torch.float32:  # bf16 activations: the [B, out] logits are widened, the loss is always fp32
        logits = logits.float()
    y = db.y
    valid = getattr(db, "valid_graphs", None)
    if valid is not None and valid < logits.size(0):  # PaddedBatch: the trailing graph holds the padding nodes
        logits, y = logits[:valid], y[:valid]
    if _DIRECT_LOSS_GRAD and _fused_regression(logits, y, loss_type) and logits.requires_grad:
        # loss and (weighted) d loss / d logits from one launch; the backward starts at the logits: no autograd node
        # for the loss, no ones_like seed, no multiply by the upstream gradient
        loss, dl = ops.clamped_regression_loss(logits.detach().reshape(-1), y, 0.0, float(num_classes - 1),
                                               0 if loss_type == "MSE" else 1, grad_scale=db.loss_weight)
        logits.backward(dl.view_as(logits))
        return loss
    if _DIRECT_LOSS_GRAD and loss_type == "CE" and _fused_ce(logits, y) and logits.requires_grad:
        loss, dl = ops.cross_entropy_loss(logits.detach().contiguous(), y, class_weights, grad_scale=db.loss_weight)
        logits.backward(dl)
        return loss
    loss = loss_fn(logits, y, loss_type, num_classes, class_weights)
    (loss * db.loss_weight if db.loss_weight != 1.0 else loss).backward()
    return loss.detach()


class TrainStep:
    """One optimisation step over a resident DeviceBatch; optionally CUDA-graph captured.

    world_size > 1: the optimiser's gradients live in a PeerBucket and the all-reduce is fused with Adam INSIDE the graph
    (FlatAdam.step_dp, csrc/dp.cu); without a PeerBucket (LGK_DP_FUSED=0, SGD) fwd+bwd (graph) -> one NCCL all-reduce of the
    flat gradient bucket -> fused Adam with grad_scale = 1/world_size.
    """

    def __init__(self, model: nn.Module, opt: FlatAdam, loss_type: str = "MSE", num_classes: int = 5,
                 use_cuda_graph: bool = True, process_group=None, world_size: int = 1,
                 class_weights: Optional[Tensor] = None):
        self.model, self.opt = model, opt
        self.loss_type, self.num_classes = loss_type, num_classes
        dev = opt.param.device
        self.class_weights = None if class_weights is None else class_weights.to(device=dev, dtype=torch.float32).contiguous()
        self.use_cuda_graph = use_cuda_graph
        self.pg, self.world_size = process_group, world_size
        self._fused_dp = world_size > 1 and getattr(opt, "peer_bucket", None) is not None and opt.peer_bucket.world == world_size
        self._graphs = {}
        # parameter-gradient kernels are dealt over this many side streams (tuning knob: LGK_SIDE_STREAMS, 0 = none)
        nside = int(os.environ.get("LGK_SIDE_STREAMS", "3"))
        self._side = [torch.cuda.Stream() for _ in range(nside)] if torch.cuda.is_available() and nside > 0 else None
        self._build = torch.cuda.Stream() if torch.cuda.is_available() else None
        # the main chain is captured on a high-priority stream: its thread blocks are placed before those of the
        # parameter-gradient kernels on the (default-priority) side streams whenever both are pending -- headline step
        # 0.3199 -> 0.3131 ms (LGK_MAIN_PRIORITY=0 captures on torch's default capture stream as before)
        self._capture = (torch.cuda.Stream(priority=-1)
                         if torch.cuda.is_available() and os.environ.get("LGK_MAIN_PRIORITY", "1") != "0" else None)

    def _body(self, db: DeviceBatch) -> Tensor:
        self.opt.zero_grad()
        # the backward kernels add weight / bias / attention gradients straight into the flat bucket (no per-parameter
        # autograd `add` launches); the bucket was just zeroed, so this equals plain accumulation
        # ... and run on a side stream, overlapping the main backward chain; joined before the optimiser reads them
        # ... attention dropout is keyed on the optimiser's device-side step counter, so graph replays draw fresh masks
        with direct_param_grads(), side_stream(self._side), dropout_clock(getattr(self.opt, "step_dev", None)):
            loss = forward_backward(self.model, db, self.loss_type, self.num_classes, self._build, self.class_weights)
        if self.world_size == 1:
            self.opt.step()
        elif self._fused_dp:
            self.opt.step_dp()  # inside the captured graph: no NCCL call, no eager tail
        return loss

    def _finish_dp(self) -> None:
        if self.world_size > 1 and not self._fused_dp:
            self.opt.all_reduce(self.pg)
            self.opt.step(grad_scale=1.0 / self.world_size)

    def __call__(self, db: DeviceBatch) -> Tensor:
        if not self.use_cuda_graph:
            loss = self._body(db)
            self._finish_dp()
            return loss
        entry = self._graphs.get(id(db))
        if entry is None:
            # warm up on a side stream (autograd and the allocator must have seen the shapes), then capture
            # -- the warm-up runs must not train: optimiser state, parameters and module buffers (BatchNorm
            # running statistics) are restored afterwards
            opt = self.opt
            state = opt.state_tensors()
            saved = [t.clone() for t in state]
            bufs = [(b, b.clone()) for b in self.model.buffers()]
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                for _ in range(2):
                    self._body(db)
                    self._finish_dp()
            torch.cuda.current_stream().wait_stream(s)
            for dst, src in zip(state, saved):
                dst.copy_(src)
            for b, src in bufs:
                b.copy_(src)
            g = torch.cuda.CUDAGraph()
            with (torch.cuda.graph(g, stream=self._capture) if self._capture is not None else torch.cuda.graph(g)):
                loss = self._body(db)
            entry = (g, loss, db)
            self._graphs[id(db)] = entry
        entry[0].replay()
        self._finish_dp()
        return entry[1]


class PaddedBatch(DeviceBatch):
    """Static device buffers for a BUCKET of batch shapes: room for `node_cap` nodes, `edge_cap` edges and `num_graphs` real
    graphs plus ONE trailing padding graph.  `load(hb)` copies a host batch of any smaller shape into the front and turns
    the rest into padding that cannot influence the real graphs: the spare nodes have zero features and belong to the
    padding graph, the spare edges are self loops on the spare nodes, and the padding graph's logit is left
    out of the loss (`valid_graphs`), so it receives a zero gradient and contributes exact zeros to every parameter
    gradient.  The spare edges are spread evenly over the spare nodes (a single hub of thousands of self loops would be one
    serial row in every per-node kernel: measured 70 ms per step instead of 0.4 ms).  Every kernel of the captured step sees the bucket's sizes: one CUDA graph serves all shapes of the bucket
    (SURVEY.md 7 hard part 10; reference training.py:64-72 steps over DataLoader batches of varying (N, E)).
    Not for models with BatchNorm in training mode: batch statistics would see the padding rows."""

    def __init__(self, node_cap: int, edge_cap: int, num_graphs: int, feat: int, device, x_dtype=torch.float32,
                 with_edge_weight: bool = False, dst_sorted: bool = True):
        self.num_nodes, self.num_graphs, self.valid_graphs = int(node_cap), int(num_graphs) + 1, int(num_graphs)
        self.num_edges = int(edge_cap)
        self.loss_weight = 1.0
        self.dst_sorted = bool(dst_sorted)
        self.x = torch.zeros(node_cap, feat, dtype=x_dtype, device=device)
        self.batch = torch.full((node_cap,), num_graphs, dtype=torch.int64, device=device)
        self.y = torch.zeros(num_graphs + 1, dtype=torch.int64, device=device)
        self.edge_index = torch.empty((2, edge_cap), dtype=torch.int64, device=device)
        self.edge_weight = torch.zeros(edge_cap, dtype=torch.float32, device=device) if with_edge_weight else None
        self._iota = torch.arange(edge_cap, dtype=torch.int64, device=device)
        self._dirty_nodes = 0   # rows of x / batch that the previous load wrote
        self._dirty_edges = 0
        self._pad_edges(0, 0)

    def _pad_edges(self, n: int, e: int) -> None:
        """spare edge slots e.. -> self loops (j, j) on the spare nodes n.., spread evenly and ascending (the edge list stays
        grouped by destination)"""
        spare_e, spare_n = self.num_edges - e, self.num_nodes - n
        if spare_e > 0:
            tail = (self._iota[:spare_e] * spare_n).div_(spare_e, rounding_mode="floor").add_(n)
            self.edge_index[0, e:].copy_(tail)
            self.edge_index[1, e:].copy_(tail)

    def fits(self, hb: HostBatch) -> bool:
        return (hb.num_nodes < self.num_nodes and hb.edge_index.size(1) <= self.num_edges and hb.num_graphs == self.valid_graphs
                and hb.x.size(1) == self.x.size(1))

    def load(self, hb: HostBatch) -> None:
        """host -> device copy of `hb` into the front of the buffers (stream-ordered, non-blocking from pinned memory); only
        the part of the padding that the previous, larger batch had overwritten is re-initialised"""
        if not self.fits(hb):
            raise ValueError("the batch does not fit this bucket (needs num_nodes < node_cap, E <= edge_cap, the same number "
                             "of graphs and feature width)")
        if self.dst_sorted and not hb.dst_sorted:
            raise ValueError("this bucket was declared destination-sorted; the batch is not")
        n, e = hb.num_nodes, hb.edge_index.size(1)
        self.x[:n].copy_(hb.x, non_blocking=True)
        self.batch[:n].copy_(hb.batch, non_blocking=True)
        self.y[:self.valid_graphs].copy_(hb.y, non_blocking=True)
        self.edge_index[:, :e].copy_(hb.edge_index, non_blocking=True)
        if self.edge_weight is not None:
            self.edge_weight[:e].copy_(hb.edge_weight, non_blocking=True)
        if self._dirty_nodes > n:
            self.x[n:self._dirty_nodes].zero_()
            self.batch[n:self._dirty_nodes].fill_(self.valid_graphs)
        self._pad_edges(n, e)  # depends on (n, e): recomputed per load, a handful of launches over the spare slots
        if self._dirty_edges > e and self.edge_weight is not None:
            self.edge_weight[e:self._dirty_edges].zero_()
        self._dirty_nodes, self._dirty_edges = n, e

    def copy_from(self, hb: HostBatch) -> None:
        self.load(hb)


class BucketedTrainStep:
    """TrainStep over batches of VARYING shape with a handful of captured graphs: a batch goes into the PaddedBatch of its
    bucket -- node capacity = num_nodes rounded up to `node_step` (plus the one node the padding needs), edge capacity =
    node capacity x `max_degree` -- and the bucket's CUDA graph, captured on first use, is replayed.  `captures` counts
    the graphs captured so far.  `buffers=2` + stage() / run(): double buffering -- the next batch is copied while the
    current one trains.
    Data parallel (world_size > 1): capturing a bucket's graph runs warm-up steps, which include the cross-rank gradient
    exchange -- every rank must therefore meet new buckets at the same steps (e.g. run one batch of every bucket through
    all ranks before training, or give all ranks the bucket of the largest shard)."""

    def __init__(self, model: nn.Module, opt, num_graphs: int, max_degree: int, loss_type: str = "MSE", num_classes: int = 5,
                 node_step: int = 4096, with_edge_weight: bool = False, x_dtype=torch.float32, dst_sorted: bool = True,
                 buffers: int = 1, **train_step_kwargs):
        for m in model.modules():
            if isinstance(m, nn.modules.batchnorm._BatchNorm) and model.training:
                raise NotImplementedError("padded batches would enter BatchNorm's batch statistics: use TrainStep per shape")
        self.step = TrainStep(model, opt, loss_type, num_classes, use_cuda_graph=True, **train_step_kwargs)
        self.num_graphs, self.max_degree, self.node_step = int(num_graphs), int(max_degree), int(node_step)
        self.with_edge_weight, self.x_dtype, self.dst_sorted = with_edge_weight, x_dtype, dst_sorted
        self.device = opt.param.device
        self.buffers = max(1, int(buffers))  # device buffers (and captured graphs) per bucket; 2 = copy / compute overlap
        self.buckets, self._turn, self._free, self._copy_stream = {}, {}, {}, None

    @property
    def captures(self) -> int:
        return len(self.step._graphs)

    def bucket_of(self, hb: HostBatch) -> int:
        cap = (hb.num_nodes + 1 + self.node_step - 1) // self.node_step * self.node_step
        while hb.edge_index.size(1) > cap * self.max_degree:  # a batch denser than declared: the next bucket up
            cap += self.node_step
        return cap

    def _buffer(self, hb: HostBatch) -> "PaddedBatch":
        if hb.num_graphs != self.num_graphs:
            raise ValueError(f"bucketed steps take batches of {self.num_graphs} graphs (the DataLoader's batch_size); "
                             f"got {hb.num_graphs}: run the last, smaller batch of an epoch through its own BucketedTrainStep")
        cap = self.bucket_of(hb)
        bufs = self.buckets.get(cap)
        if bufs is None:
            bufs = [PaddedBatch(cap, cap * self.max_degree, self.num_graphs, hb.x.size(1), self.device, self.x_dtype,
                                self.with_edge_weight, self.dst_sorted) for _ in range(self.buffers)]
            self.buckets[cap] = bufs
            self._turn[cap] = 0
            if torch.cuda.is_available():
                # the buffers were just initialised by kernels on the CURRENT stream: a first load on the copy stream must
                # not overtake them (a host that never synchronises would otherwise see the zero-fill land after the copy)
                born = torch.cuda.Event()
                born.record(torch.cuda.current_stream())
                for b in bufs:
                    self._free[id(b)] = born
        pb = bufs[self._turn[cap] % len(bufs)]
        self._turn[cap] += 1
        return pb

    def __call__(self, hb: HostBatch) -> Tensor:
        pb = self._buffer(hb)
        pb.load(hb)
        return self.step(pb)

    # ---- prefetching form: `h = stage(next_batch)` while the previous step runs, then `loss = run(h)` ------------------
    def stage(self, hb: HostBatch):
        """Start the host -> device copy of `hb` into a free buffer of its bucket on the copy stream and return a handle for
        run().  With `buffers=2` the copy of batch i + 1 overlaps the step of batch i (each buffer has its own captured
        graph: up to 2 graphs per bucket).  `hb` must stay alive (and pinned, for a truly asynchronous copy) until run()."""
        pb = self._buffer(hb)
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
        free = self._free.get(id(pb))
        with torch.cuda.stream(self._copy_stream):
            if free is not None:
                self._copy_stream.wait_event(free)  # the last step that read this buffer has finished
            pb.load(hb)
            ready = torch.cuda.Event()
            ready.record(self._copy_stream)
        return pb, ready

    def run(self, handle) -> Tensor:
        pb, ready = handle
        cur = torch.cuda.current_stream()
        cur.wait_event(ready)
        loss = self.step(pb)
        free = torch.cuda.Event()
        free.record(cur)
        self._free[id(pb)] = free
        return loss
