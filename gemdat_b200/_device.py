"""Device-buffer plumbing: torch tensors own HBM, streams come from torch, nothing else.

There is no CPU fallback: every helper raises when CUDA is not available.
"""

from __future__ import annotations

import numpy as np
import torch


def require_cuda() -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError(
            'gemdat_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback'
        )
    return torch.device('cuda', torch.cuda.current_device())


def stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def ptr(t) -> int | None:
    """Raw device (or host) address of a tensor / numpy array; None stays NULL."""
    if t is None:
        return None
    if isinstance(t, torch.Tensor):
        return t.data_ptr()
    if isinstance(t, np.ndarray):
        return t.ctypes.data
    raise TypeError(type(t))


def empty(shape, dtype) -> torch.Tensor:
    return torch.empty(shape, dtype=dtype, device=require_cuda())


def zeros(shape, dtype) -> torch.Tensor:
    return torch.zeros(shape, dtype=dtype, device=require_cuda())


def is_device(x) -> bool:
    return isinstance(x, torch.Tensor) and x.is_cuda


def to_device(x, dtype: torch.dtype | None = None) -> torch.Tensor:
    """Contiguous device tensor of ``x`` (numpy array, host tensor or device tensor)."""
    dev = require_cuda()
    if isinstance(x, torch.Tensor):
        t = x
    else:
        arr = np.ascontiguousarray(x)
        t = torch.from_numpy(arr)
    if dtype is not None and t.dtype != dtype:
        t = t.to(dtype)
    if not t.is_cuda:
        t = t.to(dev, non_blocking=True)
    return t.contiguous()


_PINNED_MIN_BYTES = 1 << 20


def to_host(t: torch.Tensor) -> np.ndarray:
    """Device tensor -> numpy.  Large results land in pinned memory (torch's caching host
    allocator recycles the blocks, so only the first call pays for cudaHostAlloc) and move by
    DMA at PCIe speed; the returned array keeps its pinned block alive."""
    t = t.detach()
    if not t.is_cuda:
        return t.numpy()
    if t.numel() * t.element_size() < _PINNED_MIN_BYTES:
        return t.cpu().numpy()
    out = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
    out.copy_(t, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    return out.numpy()


def host_f64(x) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(x, dtype=np.float64))
