"""Helpers shared by the GPU parity tests."""
import numpy as np
import torch

_NP2T = {np.dtype(np.uint8): torch.uint8, np.dtype(np.uint16): torch.uint16, np.dtype(np.uint32): torch.uint32,
         np.dtype(np.uint64): torch.uint64, np.dtype(np.int32): torch.int32, np.dtype(np.int64): torch.int64}
_SIGNED_VIEW = {1: np.int8, 2: np.int16, 4: np.int32, 8: np.int64}


def to_dev(x, offset_elems=0):
    """numpy -> CUDA tensor with the same bits.  offset_elems > 0 places the data
    at a deliberately misaligned address (a view into a larger allocation)."""
    x = np.ascontiguousarray(x)
    signed = x.view(_SIGNED_VIEW[x.dtype.itemsize])
    t = torch.from_numpy(signed.copy())
    tdt = _NP2T.get(x.dtype, t.dtype)
    if offset_elems:
        buf = torch.empty(t.numel() + offset_elems + 64, dtype=t.dtype, device="cuda")
        view = buf[offset_elems:offset_elems + t.numel()]
        view.copy_(t)
        return view.view(tdt)
    return t.cuda().view(tdt)


def to_np(t, dtype=None):
    es = t.element_size()
    signed_t = {1: torch.int8, 2: torch.int16, 4: torch.int32, 8: torch.int64}[es]
    a = t.view(signed_t).cpu().numpy()
    if dtype is None:
        dtype = {1: np.uint8, 2: np.uint16, 4: np.uint32, 8: np.uint64}[es]
    return a.view(dtype)
