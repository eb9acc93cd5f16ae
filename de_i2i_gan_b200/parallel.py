"""Data parallelism for the training step: one process per GPU, batch-sharded, gradient all-reduce over NCCL (NVLink 5 /
NVSwitch) - by default ONE collective over the flat gradient vector after the backward, optionally (DEI2I_DDP_OVERLAP=1)
bucketed and overlapped with the backward (see OVERLAP below for the measurement behind the default).

The reference has no distributed code for DefectGAN (single device, utils/util.py:11-18): this layer is
added *under* the trainer.  Replicas keep local BatchNorm statistics (DDP semantics = a reference run at the per-GPU
batch); spectral-norm u/v stay identical across ranks without communication because they are a function of the
(identical) weights only.

Mechanics: FlatAdam keeps each network's gradients in one contiguous FP32 vector.  In the overlapped mode that vector is cut into
contiguous buckets from the END (backward produces gradients roughly in reverse parameter order); a
post-accumulate-grad hook per parameter counts its bucket down and, when a bucket is complete, enqueues
``all_reduce(AVG)`` of that slice on a side stream that waits for the backward stream at that point - so the collective
of early-finished buckets runs under the rest of backward.  ``finish()`` (called before ``optimizer.step()``) flushes
the buckets that never completed (parameters without gradient in this pass) and joins the streams.
"""
import os
from typing import List, Optional

import torch
import torch.distributed as dist


def init_distributed(backend: Optional[str] = None):
    """Reads RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* (torchrun); returns (rank, local_rank, world)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if torch.cuda.is_available():
            torch.cuda.set_device(local)
        backend = backend or ("nccl" if torch.cuda.is_available() else "gloo")
        kw = {}
        if backend == "nccl":
            kw["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend=backend, **kw)
    return rank, local, world


def broadcast_module(module: torch.nn.Module, src: int = 0, group=None):
    """Parameters AND buffers (spectral-norm u/v, BN running stats) from rank ``src`` - one-off at start-up."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    with torch.no_grad():
        for t in list(module.parameters()) + list(module.buffers()):
            dist.broadcast(t.data, src, group=group)


# DEI2I_DDP_OVERLAP=0 (the default): reduce the whole flat gradient vector in ONE all-reduce after the backward instead of
# bucket by bucket under it (=1).  The convolution kernels are persistent (one CTA per SM, ~200 KB of shared memory): NCCL CTAs running beside them
# take SMs away and push every 148-CTA kernel into a second wave while a bucket is in flight, which costs more than the
# < 1 ms of exposed NVLink time of a 50-180 MB all-reduce (8 GPUs: 2953.7 vs 2935.7 img/s, 2 GPUs: 747.1 vs 741.8).
OVERLAP = os.environ.get("DEI2I_DDP_OVERLAP", "0") != "0"
# Diagnostic only (tools/gpu_scale.sh): replicas run without exchanging gradients, which separates the cost of the collective from the
# max-over-ranks spread of N power-capped GPUs in a scaling measurement.  Never set it for training.
SKIP_ALLREDUCE = os.environ.get("DEI2I_DDP_SKIP_ALLREDUCE", "0") == "1"


class BucketReducer:
    """Overlapped all-reduce of one flat gradient vector (see module docstring)."""

    def __init__(self, flat_grad: torch.Tensor, slices, bucket_bytes: int = 32 << 20, group=None):
        """slices: [(param, offset, numel)] in parameter order (FlatAdam.slices())."""
        self.flat = flat_grad
        self.group = group
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.buckets: List[dict] = []
        self._hooks = []
        self._stream = torch.cuda.Stream() if flat_grad.is_cuda else None
        self.enabled = True
        per = max(1, bucket_bytes // 4)
        # cut from the end
        cur_end, cur_members = flat_grad.numel(), []
        for p, off, n in reversed(list(slices)):
            cur_members.append(p)
            if cur_end - off >= per:
                self.buckets.append(dict(lo=off, hi=cur_end, members=cur_members, pending=0, done=False))
                cur_end, cur_members = off, []
        if cur_members:
            self.buckets.append(dict(lo=0, hi=cur_end, members=cur_members, pending=0, done=False))
        self._bucket_of = {}
        for bi, b in enumerate(self.buckets):
            for p in b["members"]:
                self._bucket_of[id(p)] = bi
                if p.requires_grad:
                    self._hooks.append(p.register_post_accumulate_grad_hook(self._on_grad))
        self.reset()

    def reset(self):
        for b in self.buckets:
            b["pending"] = sum(1 for p in b["members"] if p.requires_grad)
            b["done"] = False

    def _launch(self, b):
        if b["done"]:
            return
        b["done"] = True
        if self.world == 1 or SKIP_ALLREDUCE:
            return
        view = self.flat[b["lo"]:b["hi"]]
        if self._stream is not None:
            self._stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self._stream):
                dist.all_reduce(view, op=dist.ReduceOp.AVG, group=self.group)
        else:  # gloo (CPU tests): no AVG
            dist.all_reduce(view, op=dist.ReduceOp.SUM, group=self.group)
            view.div_(self.world)

    def _on_grad(self, p):
        if not self.enabled or not OVERLAP:
            return
        b = self.buckets[self._bucket_of[id(p)]]
        b["pending"] -= 1
        if b["pending"] == 0:
            self._launch(b)

    def finish(self):
        """All buckets reduced and visible to the current stream; re-arms the counters for the next backward."""
        if not OVERLAP and self.world > 1 and not SKIP_ALLREDUCE and not all(b["done"] for b in self.buckets):
            # one all-reduce of the whole flat vector after the backward (see OVERLAP)
            for b in self.buckets:
                b["done"] = True
            if self._stream is not None:  # same stream choreography as the bucketed path (CUDA-graph capturable)
                self._stream.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(self._stream):
                    dist.all_reduce(self.flat, op=dist.ReduceOp.AVG, group=self.group)
            else:
                dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group)
                self.flat.div_(self.world)
        for b in self.buckets:
            self._launch(b)
        if self._stream is not None and self.world > 1 and not SKIP_ALLREDUCE:
            torch.cuda.current_stream().wait_stream(self._stream)
        self.reset()

    def remove(self):
        for h in self._hooks:
            h.remove()
        self._hooks = []
