from __future__ import annotations

import torch


def ptr(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


def stream_ptr(device: torch.device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def as_points(t: torch.Tensor, name: str = "points") -> torch.Tensor:
    if t.dim() != 2:
        raise ValueError(f"{name} must be a [n, dim] tensor, got shape {tuple(t.shape)}")
    if t.dtype != torch.float32:
        raise TypeError(f"{name} must be float32 (the reference stores float coordinates, Node.hpp:13), got {t.dtype}")
    return t.contiguous()
