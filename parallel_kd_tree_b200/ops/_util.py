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


def on_device(t: torch.Tensor, device: torch.device, dtype: torch.dtype, name: str) -> torch.Tensor:
    """Contiguous ``dtype`` copy of ``t`` on ``device`` (the kernels take raw pointers: a tensor on another device
    would make them dereference a foreign address)."""
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a tensor")
    return t.to(device=device, dtype=dtype).contiguous()


def check_out(t: torch.Tensor | None, shape: tuple, dtype: torch.dtype, device: torch.device, name: str) -> None:
    """Output buffers are written through raw pointers: shape, dtype, device and contiguity must be exact."""
    if t is None:
        return
    if tuple(t.shape) != tuple(shape) or t.dtype != dtype or t.device != device or not t.is_contiguous():
        raise ValueError(f"{name} must be a contiguous {dtype} tensor of shape {tuple(shape)} on {device}, got "
                         f"{t.dtype} {tuple(t.shape)} on {t.device}")
