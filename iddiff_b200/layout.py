"""Host-side view of the slab layout (include/iddiff_b200.h) -- used to ingest/inspect operands
and by the tests; the hot path itself never reorders data on the host."""
from __future__ import annotations

import torch


def slab_index(R: int, C: int, device=None) -> torch.Tensor:
    """int64 [R, C]: bf16-element index (byte offset / 2) of element (r, c) in the slab buffer."""
    r = torch.arange(R, device=device, dtype=torch.int64)[:, None]
    c = torch.arange(C, device=device, dtype=torch.int64)[None, :]
    byte = ((c // 64) * R + r) * 128 + ((((c % 64) // 8) ^ (r % 8)) * 16) + (c % 8) * 2
    return byte // 2


def split_bf16(x: torch.Tensor):
    """x ~= hi + lo, both bf16 (round-to-nearest-even), as the kernels compute it."""
    hi = x.to(torch.bfloat16)
    lo = (x - hi.to(torch.float32)).to(torch.bfloat16)
    return hi, lo


def to_slab(mat_bf16: torch.Tensor) -> torch.Tensor:
    """[R, C] bf16 (R % 8 == 0, C % 64 == 0) -> uint8 slab buffer of R*C*2 bytes."""
    R, C = mat_bf16.shape
    assert R % 8 == 0 and C % 64 == 0
    out = torch.empty(R * C, dtype=torch.bfloat16, device=mat_bf16.device)
    out[slab_index(R, C, mat_bf16.device).reshape(-1)] = mat_bf16.reshape(-1)
    return out.view(torch.uint8)


def from_slab(buf_u8: torch.Tensor, R: int, C: int) -> torch.Tensor:
    """Inverse of to_slab: uint8 buffer -> [R, C] bf16."""
    flat = buf_u8.view(torch.bfloat16)
    return flat[slab_index(R, C, buf_u8.device).reshape(-1)].reshape(R, C)
