"""One-hot feature construction on the device (K15): ELL index assembly and pair features."""
from __future__ import annotations

import numpy as np
import torch

from ._lib import check, lib, stream_ptr


def assemble_ell(codes: np.ndarray, offsets: np.ndarray, device="cuda") -> np.ndarray:
    """``idx[r, c] = offsets[c] + codes[r, c]`` as int32 (host in, host out; device compute)."""
    c = torch.from_numpy(np.ascontiguousarray(codes, dtype=np.int64)).to(device)
    o = torch.from_numpy(np.ascontiguousarray(offsets, dtype=np.int64)).to(device)
    idx = torch.empty(c.shape, dtype=torch.int32, device=device)
    check(lib().eh_onehot_assemble(c.data_ptr(), o.data_ptr(), c.shape[0], c.shape[1], idx.data_ptr(), stream_ptr()),
          "eh_onehot_assemble")
    return idx.cpu().numpy()


def pair_codes(codes: torch.Tensor, pairs) -> torch.Tensor:
    """Mixed-radix id of every listed column pair (degree-2 interaction features)."""
    codes = codes.contiguous().long()
    radix = (codes.max(dim=0).values + 1).contiguous()
    pt = torch.tensor(list(pairs), dtype=torch.int32, device=codes.device).contiguous()
    out = torch.empty((codes.shape[0], pt.shape[0]), dtype=torch.int64, device=codes.device)
    check(lib().eh_pair_codes(codes.data_ptr(), codes.shape[0], codes.shape[1], pt.data_ptr(), pt.shape[0],
                              radix.data_ptr(), out.data_ptr(), stream_ptr()), "eh_pair_codes")
    return out
