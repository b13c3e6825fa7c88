"""Oracle for forward/backward fill and integer_remap: src/gemdat/utils.py:59-125.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).
"""

from __future__ import annotations

import numpy as np


def ffill(arr: np.ndarray, fill_val: int = -1, axis=-1) -> np.ndarray:
    """Last value != ``fill_val`` along ``axis``; leading fill values stay.

    Quirk kept from utils.py:76-77: with ``axis=0`` the recursive call drops
    ``fill_val`` and always uses -1.
    """
    if axis == 0:
        return ffill(arr.T).T
    if arr.ndim > 2:
        raise ValueError
    n = arr.shape[1]
    last = np.where(arr != fill_val, np.arange(n), 0)
    np.maximum.accumulate(last, axis=1, out=last)
    return np.take_along_axis(arr, last, axis=1)


def bfill(arr: np.ndarray, fill_val: int = -1, axis=-1) -> np.ndarray:
    if axis == 0:
        return bfill(arr.T).T
    if arr.ndim > 2:
        raise ValueError
    return ffill(arr[:, ::-1], fill_val=fill_val)[:, ::-1]


def integer_remap(a: np.ndarray, key: np.ndarray, palette: np.ndarray | None = None) -> np.ndarray:
    """Map ``palette[i] -> key[i]`` (palette sorted; default: unique values of ``a``)."""
    if palette is None:
        palette = np.unique(a)
    index = np.digitize(a, palette, right=True)
    return key[index].reshape(a.shape)
