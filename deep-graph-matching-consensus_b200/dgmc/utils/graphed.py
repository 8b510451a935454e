"""CUDA-graph capture of a whole DGMC training step (no counterpart in the reference, which
launches ~100 small kernels per consensus step from Python and is launch-bound on a B200).

At the DBP15K shape one step is ~1,000 kernel launches of 5-100 us each; issued eagerly from Python
the GPU waits on the host.  `GraphedTrainStep` captures zero_grad + forward + loss + backward +
optimizer step once and replays it with one launch.  What makes that possible: the sparse
path has static shapes and no host synchronisation (`__include_gt__` and `loss` avoid nonzero),
every kernel of libdgmc_b200 is stream-ordered on the current stream and allocates nothing, and
the random draws use torch's graph-safe Philox generator, so each replay draws fresh indicators
and negatives.

    step = GraphedTrainStep(model, optimizer, x_s, e_s, None, None, x_t, e_t, None, None, y)
    loss = step(x_s, e_s, None, None, x_t, e_t, None, None, y)   # tensors may live on the host

Inputs are copied into static device buffers on every call (host tensors should be pinned);
shapes must match the example.  `structure_is_static=True` additionally treats edge_index /
edge_attr as fixed (their CSR is built once, outside the graph) — the analogue of the reference
training on one fixed pair for 200 epochs (examples/dbp15k.py:63-76).
"""
import torch

from .. import ops


class GraphedTrainStep(object):
    def __init__(self, model, optimizer, *example_inputs, loss_fn=None, warmup=3,
                 structure_is_static=False):
        for group in optimizer.param_groups:
            if 'capturable' in group and not group['capturable']:
                raise ValueError('GraphedTrainStep needs an optimizer built with capturable=True')
        self.model, self.optimizer = model, optimizer
        self.loss_fn = loss_fn or (lambda S_0, S_L, y: model.loss(S_L, y))
        dev = next(model.parameters()).device
        self.static = [None if t is None else t.to(dev).clone() for t in example_inputs]
        for dst, src in zip(self.static, example_inputs):
            # (number of graphs, largest graph) of a batch vector, attached by dgmc.utils.collate_pairs: the
            # captured step keeps that padded layout, so every batch replayed through it must fit it
            if src is not None and hasattr(src, '__dgmc_layout__'):
                dst.__dgmc_layout__ = src.__dgmc_layout__
        self.structure_is_static = structure_is_static
        self.graph = torch.cuda.CUDAGraph()

        import contextlib
        dynamic = contextlib.nullcontext if structure_is_static else ops.dynamic_structure
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side), dynamic():
            for _ in range(warmup):
                self._step()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        if not structure_is_static:
            ops.clear_caches()  # so that the CSR builds are captured and re-run on new inputs
        # with NCCL collectives inside the step, the communicator's watchdog thread keeps polling
        # CUDA events while this thread captures: only this thread's calls may be checked
        import torch.distributed as dist
        mode = 'thread_local' if dist.is_available() and dist.is_initialized() else 'global'
        with torch.cuda.graph(self.graph, capture_error_mode=mode), dynamic():
            self.loss = self._step()
        if not structure_is_static:
            ops.clear_caches()  # entries made during capture point into the graph's private pool

    def _step(self):
        self.optimizer.zero_grad(set_to_none=True)
        S_0, S_L = self.model(*self.static)
        loss = self.loss_fn(S_0, S_L, self.static[-1])
        loss.backward()
        if hasattr(self.model, 'sync_gradients'):  # RowShardedDGMC: sum partial gradients
            self.model.sync_gradients()
        self.optimizer.step()
        return loss.detach()

    def __call__(self, *inputs):
        """Copy the inputs into the captured buffers and replay.  Returns the (static) loss tensor;
        reading it with .item() is the only host synchronisation.  Inputs that were announced with
        `prefetch()` are taken from the staging buffers (a device-to-device copy)."""
        staged = self._staged_for(inputs)
        if staged is not None:
            torch.cuda.current_stream().wait_event(self._staged_event)
            inputs = staged
        for dst, src in zip(self.static, inputs):
            if dst is not None and src is not dst:
                dst.copy_(src, non_blocking=True)
        if staged is not None:
            self._consumed_event.record()   # staging buffers are free again from here on
        self.graph.replay()
        return self.loss

    # ---- optional input pipelining ----------------------------------------------------------------
    def prefetch(self, *inputs):
        """Start copying the NEXT step's (pinned host) inputs to the device on a side stream while
        the current step is still running; the following __call__ with the same tensors then only
        pays a device-to-device copy.  Call it right after __call__ and before reading the loss."""
        if not hasattr(self, '_copy_stream'):
            self._copy_stream = torch.cuda.Stream()
            self._staging = [None if t is None else torch.empty_like(t) for t in self.static]
            self._staged_event = torch.cuda.Event()
            self._consumed_event = torch.cuda.Event()
            self._consumed_event.record()
        # the staging buffers may still feed the previous __call__'s device-to-device copies; do NOT
        # wait for the replay that follows them — overlapping it is the point
        self._copy_stream.wait_event(self._consumed_event)
        with torch.cuda.stream(self._copy_stream):
            for dst, src in zip(self._staging, inputs):
                if dst is not None:
                    dst.copy_(src, non_blocking=True)
            self._staged_event.record()
        self._staged_ids = tuple(id(t) for t in inputs)

    def _staged_for(self, inputs):
        if getattr(self, '_staged_ids', None) == tuple(id(t) for t in inputs):
            self._staged_ids = None
            return self._staging
        return None
