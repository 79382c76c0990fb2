"""Two host-side shortcuts for the launch paths (a FEATURE-state view issues ~110 launches; the Python around them is
what bounds that state, tools/host_vs_device.py): the raw handle of torch's current stream without building a
torch.cuda.Stream object, and a device guard that costs nothing when the device is already current."""
from __future__ import annotations

import torch


def raw_stream(dev) -> int:
    """cudaStream_t of torch's current stream on `dev`."""
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    return torch._C._cuda_getCurrentRawStream(idx)


class on_device:
    """`with torch.cuda.device(dev):` when dev is not the current device, nothing otherwise."""
    __slots__ = ("idx", "ctx")

    def __init__(self, dev):
        self.idx = -1 if dev.index is None else dev.index
        self.ctx = None

    def __enter__(self):
        if self.idx >= 0 and torch._C._cuda_getDevice() != self.idx:
            self.ctx = torch.cuda.device(self.idx)
            self.ctx.__enter__()
        return self

    def __exit__(self, *exc):
        if self.ctx is not None:
            ctx, self.ctx = self.ctx, None
            return ctx.__exit__(*exc)
        return False
