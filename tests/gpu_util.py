"""Helpers shared by the -m gpu parity tests."""
import numpy as np
import torch


def to_dev(a: np.ndarray) -> torch.Tensor:
    if a.dtype == np.float16:
        return torch.from_numpy(a.view(np.int16).copy()).view(torch.float16).cuda()
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def bits(t: torch.Tensor) -> np.ndarray:
    """fp16 tensor -> uint16 bit patterns on the host."""
    return t.detach().cpu().contiguous().view(torch.int16).numpy().view(np.uint16)


def gemm_err(y: torch.Tensor, Yf: np.ndarray):
    """max-abs and max-rel (relative to max(|ref|, 1)) error vs the fp32 oracle."""
    d = np.abs(y.detach().float().cpu().numpy() - Yf)
    return float(d.max(initial=0.0)), float((d / np.maximum(np.abs(Yf), 1.0)).max(initial=0.0))
