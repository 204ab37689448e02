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


def gemm_check(y: torch.Tensor, Yf: np.ndarray, tol: float = 1e-2, what=""):
    """The north-star bar (|y - ref| <= tol * max(1, |ref|)), plus the observed absolute error
    where an absolute bound is meaningful for fp16 outputs (|ref| < 16: output spacing <= 2^-7),
    asserted and returned so a drift shows up long before the relative bar trips."""
    yv = y.detach().float().cpu().numpy()
    d = np.abs(yv - Yf)
    rel = float((d / np.maximum(np.abs(Yf), 1.0)).max(initial=0.0))
    small = np.abs(Yf) < 16.0
    abs_small = float(d[small].max(initial=0.0))
    assert rel <= tol, (what, "max rel", rel, "max abs", float(d.max(initial=0.0)))
    assert abs_small <= 2e-2, (what, "max abs at |ref| < 16", abs_small)
    return rel, abs_small


def sampled_case(oracle, M, K, N, G, seed, n_words=32, edge_words=0):
    """A full-size GEMM problem generated on the GPU (SURVEY.md 8d recipe) plus the fp32 oracle
    result for a random sample of 8 * n_words output channels: only the sampled packed
    columns travel to the host, so any BASELINE shape is checked in well under a second."""
    g = torch.Generator(device="cuda").manual_seed(seed)
    qw = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device="cuda", generator=g)
    qz = torch.randint(-2**31, 2**31 - 1, (K // G, N // 8), dtype=torch.int32, device="cuda", generator=g)
    s = (torch.rand((K // G, N), device="cuda", generator=g) * 0.019 + 0.001).half()
    x = torch.randn((M, K), device="cuda", generator=g).half()
    words = torch.randperm(N // 8, device="cuda", generator=g)[:n_words]
    if edge_words:      # always include the first and the last packed columns (tile-edge cases)
        edge = torch.cat([torch.arange(edge_words, device="cuda"),
                          torch.arange(N // 8 - edge_words, N // 8, device="cuda")])
        words = torch.unique(torch.cat([words, edge]))
    words = words.sort().values
    cols = (words[:, None] * 8 + torch.arange(8, device="cuda")[None, :]).reshape(-1)
    W = oracle.dequantize(qw[:, words].contiguous().cpu().numpy(),
                          s[:, cols].contiguous().cpu().view(torch.int16).numpy().view(np.float16),
                          qz[:, words].contiguous().cpu().numpy())
    xf = x.cpu().view(torch.int16).numpy().view(np.float16)
    Yf, _ = oracle.gemm_f32(xf, W)
    return x, qw, s, qz, cols, Yf
