"""Cyclic-MDS code construction on the device (semantics of util.py:64-103)."""
from __future__ import annotations

import itertools

import numpy as np
import torch

from ._lib import check, lib, stream_ptr


def cyclic_code(H: torch.Tensor | None, n: int, s: int, device="cuda") -> torch.Tensor:
    B = torch.empty((n, n), dtype=torch.float64, device=device)
    hp = 0
    if s > 0:
        H = H.to(device=device, dtype=torch.float64).contiguous()
        assert H.shape == (s, n)
        hp = H.data_ptr()
    check(lib().eh_code_cyclic(hp, n, s, B.data_ptr(), stream_ptr()), "eh_code_cyclic")
    return B


def decode_table(B: torch.Tensor, s: int) -> torch.Tensor:
    """``[C(n,s), n]`` decode rows in ``itertools.combinations`` order of the missing workers."""
    n = B.shape[0]
    miss = np.array(list(itertools.combinations(range(n), s)), dtype=np.int32).reshape(-1, max(s, 1))
    npat = max(1, len(miss)) if s > 0 else 1
    A = torch.empty((npat, n), dtype=torch.float64, device=B.device)
    mt = torch.from_numpy(miss).to(B.device) if s > 0 else None
    check(lib().eh_code_decode_table(B.contiguous().data_ptr(), n, s, 0 if mt is None else mt.data_ptr(), npat,
                                     A.data_ptr(), stream_ptr()), "eh_code_decode_table")
    return A
