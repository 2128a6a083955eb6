"""Data-parallel training: one process per GPU, full model replica, batch sharded.

The reference has no multi-GPU code (run/train.py:235-243 builds a single-device Lightning Trainer); the exchange step the
sharded path needs is a gradient average.  Gradients leave the backward pass in four flat buckets in the order the backward
produces them -- postnet -> decoder (+ prenet) -> conditioning / att_encoder -> encoder -- and each bucket's NCCL all-reduce
is enqueued on a side stream the moment the kernels that write it have been launched (``autograd._after_bucket``), so the
112 MB exchange overlaps with the backward of the modules upstream (the 71 MB decoder bucket overlaps the encoder backward;
only the 12 MB encoder bucket is exposed).  BatchNorm statistics stay per-GPU (standard DDP semantics).

``GradSync.finish()`` makes the compute stream wait for the exchange and installs the averaged gradients as ``p.grad``.
One backward pass per step is assumed (``optimizer.zero_grad(set_to_none=True)``); a parameter group whose bucket arrives
several times in that pass (decoder chunks of a batch above 64 utterances) is summed over its buckets.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


class GradSync:
    def __init__(self, model, process_group=None):
        self.model, self.pg = model, process_group
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.pending = []  # (flat, params, views, work)
        self.stream = None
        model._grad_bucket_hook = self.on_bucket

    def on_bucket(self, flat, params, views):
        """Called by the autograd glue right after a module's backward was launched on the current stream."""
        work = None
        if self.world > 1:
            if flat.is_cuda:
                if self.stream is None:
                    self.stream = torch.cuda.Stream(device=flat.device)
                self.stream.wait_stream(torch.cuda.current_stream(flat.device))  # the bucket's producer kernels
                flat.record_stream(self.stream)
                with torch.cuda.stream(self.stream):
                    dist.all_reduce(flat, op=dist.ReduceOp.AVG, group=self.pg)
            else:  # gloo (CPU tests of the host logic): no AVG, no streams
                work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.pg, async_op=True)
        self.pending.append((flat, params, views, work))

    def finish(self):
        """Wait for every bucket and install the averaged gradients."""
        n = 0
        seen = set()
        for flat, params, views, work in self.pending:
            if work is not None:
                work.wait()
                flat.div_(self.world)
            elif self.stream is not None and flat.is_cuda:
                torch.cuda.current_stream(flat.device).wait_stream(self.stream)
            for p, v in zip(params, views):
                # a parameter group can arrive more than once per backward (a rank with more than 64 utterances runs the decoder
                # in chunks, each with its own bucket): the contributions add up, as autograd's own accumulation would
                p.grad = p.grad + v if id(p) in seen else v
                seen.add(id(p))
            n += flat.numel()
        self.pending.clear()
        return n

    def detach(self):
        self.model._grad_bucket_hook = None
