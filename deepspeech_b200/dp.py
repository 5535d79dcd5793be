"""Data-parallel plumbing: one process per GPU, batch sharded by utterance, gradients averaged with
bucketed asynchronous all-reduce launched from inside backward so that it overlaps with the backward of
earlier layers.  This is the NCCL form of the reference's (dead) in-graph tower averaging
`average_gradients` (src/deepSpeech_train.py:193-228; call site commented out at :453).

torch.distributed is the transport (NCCL over NVLink/NVSwitch on GPUs, gloo on CPU for tests)."""
from __future__ import annotations

from typing import Iterable, List

import os

import torch
import torch.distributed as dist

# timing experiments only (gradients stay local, results are wrong): what the step costs without any communication
_DEBUG_NO_COMM = bool(os.environ.get("DS2_DP_DEBUG_NO_COMM"))


class _Done(object):
    def wait(self):
        return True


def _all_reduce(t, group):
    if _DEBUG_NO_COMM:
        return _Done()
    return dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group, async_op=True)


def shard_batch(n_items: int, rank: int, world: int) -> slice:
    """Equal contiguous shards of the global batch (utterances are independent).  The global
    batch must divide evenly so that the mean of per-rank mean losses equals the global mean
    (src/deepSpeech_NCHW.py:226)."""
    if n_items % world:
        raise ValueError("global batch %d is not divisible by world size %d" % (n_items, world))
    per = n_items // world
    return slice(rank * per, (rank + 1) * per)


class GradAllReducer(object):
    """Registers post-accumulate hooks on `params`; each gradient is all-reduced (sum) asynchronously as
    soon as autograd has produced it - backward visits layers last-to-first, so layer l's reduction runs
    while layers < l are still back-propagating.  Small tensors are packed into buckets of
    `bucket_bytes` to amortise launch latency.  `finish()` waits for all reductions and applies 1/world - unless
    `prescaled`: the caller then back-propagates loss / world (`loss_scale`), the sums ARE the means and no scaling
    kernel runs at all."""

    def __init__(self, params: Iterable[torch.nn.Parameter], group=None, bucket_bytes: int = 32 << 20, prescaled: bool = False):
        self.params: List[torch.nn.Parameter] = [p for p in params if p.requires_grad]
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.bucket_bytes = bucket_bytes
        self._pending = []          # (work handle, flat buffer or None, [params])
        self._bucket: List[torch.nn.Parameter] = []
        self._bucket_size = 0
        self._seen = 0
        self._hooks = [p.register_post_accumulate_grad_hook(self._on_grad) for p in self.params]
        self.launched = 0
        self.copied = 0             # gradients copied out of a reduced buffer (autograd had cloned them)
        # The recurrent stack is one autograd node: its gradients reach the hooks only when the whole stack has been
        # back-propagated.  It offers them layer by layer instead (ops.STACK_GRAD_HOOK); they are all-reduced right
        # there; when autograd later installs its own copy as .grad, finish() overwrites that copy with the reduced
        # tensor.  This relies on .grad being a fresh tensor of this backward pass, i.e. on gradients having been reset
        # with zero_grad(set_to_none=True): accumulating into an older .grad would add the not-yet-reduced values.
        self._early = {}            # id(parameter) -> (work handle, reduced tensor)
        self._taken = set()         # ids of parameters whose .grad this reducer installed itself (autograd still runs their hooks)
        self._keep = []             # flat buffers whose views are in flight
        self.prescaled = prescaled
        if self.world > 1:
            # replicas must start from the same weights whatever seeded them: rank 0's values go to everybody
            with torch.no_grad():
                for q in self.params:
                    dist.broadcast(q.data, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
            from . import ops
            ops.STACK_GRAD_HOOK = self._on_stack_layer

    def _on_stack_layer(self, layer, tensors, params, flat=None):
        """All gradients of one recurrent layer, offered from inside the stack's backward.  When they are views of one
        flat buffer (ops._RecStackFn allocates them so) the layer goes out as ONE all-reduce."""
        live = [(t, p) for t, p in zip(tensors, params) if t is not None]
        for t, p in live:
            if p.grad is not None:
                raise RuntimeError("GradAllReducer: reset gradients with zero_grad(set_to_none=True) before backward")
        if flat is not None and sum(t.numel() for t, _ in live) == flat.numel():
            # The views of the flat buffer become the parameters' .grad right here and the stack's backward returns no
            # gradient for them (return value True): autograd neither clones nor accumulates, the in-place all-reduce
            # lands in .grad, finish() only waits.
            w = _all_reduce(flat, self.group)
            for t, p in live:
                p.grad = t
                self._taken.add(id(p))
            self._pending.append((w, None, [p for _, p in live]))
            self._keep.append(flat)
            self.launched += 1
            return True
        for t, p in live:
            w = _all_reduce(t, self.group)
            self._early[id(p)] = (w, t, 0, t.numel())
            self.launched += 1
        return False

    @property
    def loss_scale(self) -> float:
        """Factor to apply to the loss before backward() when the reducer was built with prescaled=True."""
        return 1.0 / self.world if self.prescaled else 1.0

    def _flush(self):
        if not self._bucket:
            return
        ps, self._bucket, self._bucket_size = self._bucket, [], 0
        if len(ps) == 1:
            w = _all_reduce(ps[0].grad, self.group)
            self._pending.append((w, None, ps))
        else:
            flat = torch.cat([p.grad.reshape(-1) for p in ps])
            w = _all_reduce(flat, self.group)
            self._pending.append((w, flat, ps))
        self.launched += 1

    def _on_grad(self, p):
        if self.world == 1:
            return
        self._seen += 1
        if id(p) in self._taken:                # installed and reduced from inside the stack's backward
            self._taken.discard(id(p))
            return
        e = self._early.pop(id(p), None)
        if e is not None:                       # already on its way (offered by the recurrent stack's backward)
            self._pending.append((e[0], e[1][e[2]:e[2] + e[3]], [p]))
            return
        nbytes = p.grad.numel() * p.grad.element_size()
        if nbytes >= self.bucket_bytes:
            self._flush()
            self._bucket = [p]
            self._flush()
        else:
            self._bucket.append(p)
            self._bucket_size += nbytes
            if self._bucket_size >= self.bucket_bytes:
                self._flush()

    def finish(self):
        """Call after backward(): flush the tail bucket, wait, unpack, scale by 1/world."""
        if self.world == 1:
            return
        self._flush()
        if self._early:
            raise RuntimeError("GradAllReducer: %d gradients were all-reduced early for parameters this reducer does not own"
                               % len(self._early))
        inv = 1.0 / self.world
        for w, flat, ps in self._pending:
            w.wait()
            if flat is not None:
                off = 0
                for p in ps:
                    n = p.grad.numel()
                    if p.grad.data_ptr() != flat.data_ptr() + off * flat.element_size():
                        p.grad.copy_(flat[off:off + n].view_as(p.grad))
                        self.copied += 1
                    off += n
            if not self.prescaled:
                for p in ps:
                    _scale_(p.grad, inv)
        self._pending = []
        self._keep = []
        self._taken = set()
        self._seen = 0

    def remove(self):
        for h in self._hooks:
            h.remove()
        from . import ops
        if ops.STACK_GRAD_HOOK == self._on_stack_layer:
            ops.STACK_GRAD_HOOK = None


def _scale_(t: torch.Tensor, s: float):
    if t.is_cuda and t.dtype == torch.float32 and t.is_contiguous():
        from . import ops
        ops.check(ops.lib.ds2_scale(ops._p(t), float(s), ops._p(t), t.numel(), ops._stream()), "ds2_scale")
    else:
        t.mul_(s)
