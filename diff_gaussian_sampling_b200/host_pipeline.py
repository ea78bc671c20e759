"""Host-resident inputs: double-buffered staging around the sampling step.

A caller whose Gaussians and query points live in (pinned) host memory pays, per step, the PCIe copy of the inputs
(cfg2: 170 MB, about 3 ms) before the 6 ms of device work and the copy of the gradients after it.  `HostPipeline`
overlaps the three: while step i computes on the caller's stream, the inputs of step i+1 are copied on a copy stream
into the other of two device buffer sets, and the results of step i-1 drain to the host on a third stream.  Every
step's bytes still cross the bus once in each direction; only the waiting is removed.  This is plumbing around the
path (torch streams and events), not part of the drop-in boundary: the reference has no counterpart because it only
accepts device tensors (sample_points.cu:37-44, CHECK_INPUT).
"""
import torch


class HostPipeline:
    def __init__(self, device, depth=2):
        if depth < 2:
            raise ValueError("depth must be >= 2 (one buffer set in use, one being filled)")
        self.device = torch.device(device)
        self.depth = depth
        self.h2d_stream = torch.cuda.Stream(self.device)
        self.d2h_stream = torch.cuda.Stream(self.device)
        self._sets = None          # `depth` dicts of device tensors
        self._filled = None        # event per set: H2D of the set complete
        self._released = None      # event per set: the step that read the set is complete
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def _ensure_sets(self, batch):
        ok = self._sets is not None and all(
            k in self._sets[0] and self._sets[0][k].shape == v.shape and self._sets[0][k].dtype == v.dtype
            for k, v in batch.items()) and len(self._sets[0]) == len(batch)
        if ok:
            return
        self._sets = [{k: torch.empty(v.shape, dtype=v.dtype, device=self.device) for k, v in batch.items()}
                      for _ in range(self.depth)]
        self._filled = [None] * self.depth
        self._released = [None] * self.depth

    def _prefetch(self, batch, slot):
        """Enqueue the H2D copy of `batch` into buffer set `slot` on the copy stream."""
        for k, v in batch.items():
            if not v.is_pinned():
                raise ValueError(f"host tensor '{k}' is not in pinned memory: the copy would not be asynchronous")
        with torch.cuda.stream(self.h2d_stream):
            if self._released[slot] is not None:
                self.h2d_stream.wait_event(self._released[slot])
            for k, v in batch.items():
                self._sets[slot][k].copy_(v, non_blocking=True)
                self.h2d_bytes += v.numel() * v.element_size()
            e = torch.cuda.Event()
            e.record(self.h2d_stream)
            self._filled[slot] = e

    def run(self, host_batches, step_fn, results_host):
        """host_batches: iterable of dicts name -> pinned CPU tensor (same shapes and dtypes throughout).
        step_fn(device_batch) -> tuple of device tensors, called on the current stream.
        results_host(i) -> tuple of pinned CPU tensors that receive the results of step i (may return the same
        tensors every time when only the last value is wanted, like a loss read).
        Returns the number of steps run.  On return all copies have been ENQUEUED and the current stream waits for
        them: synchronise the current stream (or the device) before reading the host results."""
        cur = torch.cuda.current_stream(self.device)
        it = iter(host_batches)
        try:
            nxt = next(it)
        except StopIteration:
            return 0
        self._ensure_sets(nxt)
        self._prefetch(nxt, 0)
        i = 0
        drained = None
        while nxt is not None:
            slot = i % self.depth
            try:
                following = next(it)
            except StopIteration:
                following = None
            if following is not None:
                self._prefetch(following, (i + 1) % self.depth)     # before step i is issued: the step syncs the host
            cur.wait_event(self._filled[slot])
            outs = step_fn(self._sets[slot])
            done = torch.cuda.Event()
            done.record(cur)
            self._released[slot] = done
            dst = results_host(i)
            if len(dst) != len(outs):
                raise ValueError("results_host(i) must return one host tensor per result of step_fn")
            with torch.cuda.stream(self.d2h_stream):
                self.d2h_stream.wait_event(done)
                for h, o in zip(dst, outs):
                    o.record_stream(self.d2h_stream)
                    h.copy_(o.reshape(h.shape), non_blocking=True)
                    self.d2h_bytes += h.numel() * h.element_size()
                drained = torch.cuda.Event()
                drained.record(self.d2h_stream)
            nxt = following
            i += 1
        if drained is not None:
            cur.wait_event(drained)
        return i
