"""Input side of the training / evaluation loops: pinned-memory prefetch.

The reference feeds its loops from a `torch.utils.data.DataLoader` over `.mat` files (dataset_preparation.py:242-273 reads
`<root>/<k>m/*.mat`, key 'data', one [100, 250] array per file -> [1, 100, 250] fp32 plus the two labels) and uploads
every batch synchronously inside the loop (`bacthX.cuda()`, utils.py:348-357): pageable memory, on the compute stream,
100 KB per patch.  `DevicePrefetcher` wraps any such iterable and runs it `depth` batches ahead:

  worker thread   next(loader) -> copy into a pinned staging slot (ring of `depth` slots per batch element)
                  -> asynchronous upload on a copy stream -> event
  consumer        makes its current stream wait for the batch's event and gets device tensors; `.cuda()` on them is a
                  no-op, so the reference's loop body runs unchanged (mtldas_b200.run_reference installs the wrapper
                  around DataLoader iteration; MTLDAS_PREFETCH=0 turns it off).

Batches are tuples / lists whose elements are tensors, numpy arrays or anything else (passed through untouched).  The
order of the batches, their values and dtypes are exactly the loader's.  Without a CUDA device the same machinery runs
with plain host staging (this is how the CPU tests exercise ordering, ragged batches and error propagation).
"""
import queue
import threading

import numpy as np
import torch

_END = object()


class DevicePrefetcher:
    def __init__(self, loader, device=None, depth=2, length=None):
        if depth < 1:
            raise ValueError("depth must be >= 1")
        self.loader, self.depth, self._length = loader, int(depth), length
        self.cuda = torch.cuda.is_available() and (device is None or torch.device(device).type == "cuda")
        if self.cuda:
            self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
            if self.device.index is None:
                self.device = torch.device("cuda", torch.cuda.current_device())
        else:
            self.device = torch.device("cpu")
        self.batches_staged = 0

    def __len__(self):
        return self._length if self._length is not None else len(self.loader)

    # ---- worker ------------------------------------------------------------------------------------------
    def _stage(self, slots, k, i, t):
        """Copy tensor t into slot k of element i (grown when a larger batch shows up) and start its upload."""
        key = (i, k)
        n = t.numel()
        buf = slots.get(key)
        if buf is None or buf.dtype != t.dtype or buf.numel() < n:
            buf = torch.empty(max(n, 1), dtype=t.dtype)
            if self.cuda:
                buf = buf.pin_memory()
            slots[key] = buf
        view = buf[:n].view(t.shape)
        view.copy_(t)
        if not self.cuda:
            return view.clone()
        with torch.cuda.stream(self._copy_stream):
            return view.to(self.device, non_blocking=True)

    def _worker(self, q, stop):
        try:
            if self.cuda:
                torch.cuda.set_device(self.device)
            slots, events = {}, [None] * self.depth
            for b, batch in enumerate(self.loader):
                if stop.is_set():
                    return
                k = b % self.depth
                if events[k] is not None:
                    events[k].synchronize()          # the upload that last read this slot's pinned memory is done
                single = not isinstance(batch, (tuple, list))
                items = [batch] if single else list(batch)
                out = []
                for i, x in enumerate(items):
                    if isinstance(x, np.ndarray):
                        x = torch.from_numpy(x)
                    out.append(self._stage(slots, k, i, x.contiguous()) if isinstance(x, torch.Tensor) and
                               x.device.type == "cpu" else x)
                ev = None
                if self.cuda:
                    ev = torch.cuda.Event()
                    ev.record(self._copy_stream)
                    events[k] = ev
                self.batches_staged += 1
                item = (out[0] if single else (tuple(out) if isinstance(batch, tuple) else out), ev)
                while not stop.is_set():
                    try:
                        q.put(item, timeout=0.1)
                        break
                    except queue.Full:
                        continue
            q.put((_END, None))
        except BaseException as e:  # noqa: BLE001  (re-raised in the consumer)
            q.put((e, None))

    # ---- consumer ----------------------------------------------------------------------------------------
    def __iter__(self):
        if self.cuda:
            with torch.cuda.device(self.device):
                self._copy_stream = torch.cuda.Stream(self.device)
        q, stop = queue.Queue(maxsize=self.depth), threading.Event()
        worker = threading.Thread(target=self._worker, args=(q, stop), daemon=True, name="mtldas-prefetch")
        worker.start()
        try:
            while True:
                item, ev = q.get()
                if item is _END:
                    return
                if isinstance(item, BaseException):
                    raise item
                if ev is not None:
                    cur = torch.cuda.current_stream(self.device)
                    cur.wait_event(ev)
                    for t in (item if isinstance(item, (tuple, list)) else [item]):
                        if isinstance(t, torch.Tensor) and t.is_cuda:
                            t.record_stream(cur)     # allocated on the copy stream, consumed on this one
                yield item
        finally:
            stop.set()
            while worker.is_alive():                 # unblock a worker waiting on a full queue, then let it finish
                try:
                    q.get_nowait()
                except queue.Empty:
                    worker.join(timeout=0.05)


def install_dataloader_prefetch(depth=2):
    """Make every torch DataLoader iteration run through DevicePrefetcher (used by mtldas_b200.run_reference so that the
    reference's unmodified train.py / test.py loops receive device tensors uploaded ahead of time).  Returns a function
    that restores the original behaviour."""
    from torch.utils.data import DataLoader
    orig = DataLoader.__iter__

    def __iter__(self):
        it = orig(self)
        if not torch.cuda.is_available():
            return it
        return iter(DevicePrefetcher(it, depth=depth, length=len(self)))

    DataLoader.__iter__ = __iter__

    def restore():
        DataLoader.__iter__ = orig

    return restore
