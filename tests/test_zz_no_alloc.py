"""b200_w4a16_gemm_noalloc (no counterpart in the reference):
the decode-batch GEMM on scratch built into the library, called through the C ABI
and compared with b200_w4a16_gemm on a caller-provided workspace -- same kernel and plan, so the
results are bit-identical, which is inside the GEMM tests' 1e-2 bar.  Collected last (file name),
must pass alone in a fresh process and after every other test."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

G = 128
TOL = 1e-2          # max-abs/rel bar of tests/test_gpu_gemm.py


def _with_workspace(L, dx, r, ds, dqz, M, K, N):
    y = torch.empty((M, N), dtype=torch.float16, device="cuda")
    ws = torch.empty(L.b200_w4a16_workspace_bytes(M, K, N, 1), dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    assert L.b200_w4a16_workspace_init(ws.data_ptr(), ws.numel(), st) == 0
    assert L.b200_w4a16_gemm(dx.data_ptr(), r.data_ptr(), ds.data_ptr(), dqz.data_ptr(), y.data_ptr(),
                             M, K, N, G, 1, 0, ws.data_ptr(), ws.numel(), st) == 0
    torch.cuda.synchronize()
    return y


@pytest.mark.gpu
@pytest.mark.parametrize("M,K,N", [(16, 4096, 4096), (1, 1024, 768), (7, 11008, 4096), (16, 512, 12288)])
def test_noalloc_gemm_matches_the_workspace_entry_point_twice_in_a_row(pkg, oracle, cuda_lib, M, K, N):
    from gpu_util import to_dev
    L = cuda_lib
    st = torch.cuda.current_stream().cuda_stream
    cases = []
    for call in range(2):                                   # other inputs each time
        x, qw, s, qz = oracle.make_inputs(M, K, N, G, seed=1000 + 10 * M + call)
        dx, ds, dqz = to_dev(x), to_dev(s), to_dev(qz)
        r = pkg.ops.quick_repack(to_dev(qw))
        want = _with_workspace(L, dx, r, ds, dqz, M, K, N)
        y = torch.full((M, N), float("nan"), dtype=torch.float16, device="cuda")
        cases.append((dx, r, ds, dqz, y, want))
    torch.cuda.synchronize()
    for (dx, r, ds, dqz, y, _) in cases:                    # two calls in a row on one stream, no sync between
        rc = L.b200_w4a16_gemm_noalloc(dx.data_ptr(), r.data_ptr(), ds.data_ptr(), dqz.data_ptr(),
                                       y.data_ptr(), M, K, N, G, st)
        assert rc == 0, L.b200_w4a16_strerror(rc)
    torch.cuda.synchronize()
    outs = []
    for (_, _, _, _, y, want) in cases:
        a, b = y.float(), want.float()
        assert not torch.isnan(a).any()
        err = (a - b).abs() / b.abs().clamp(min=1.0)
        assert float(err.max()) <= TOL
        assert torch.equal(y, want)                         # same kernel, same plan: the same bits
        outs.append(y)
    # and against the fp32 oracle on a few channels of the last call
    W = oracle.dequantize(qw[:, :8], s[:, :64], qz[:, :8])
    Yf, _ = oracle.gemm_f32(x, W)
    d = np.abs(outs[-1][:, :64].float().cpu().numpy() - Yf) / np.maximum(np.abs(Yf), 1.0)
    assert float(d.max()) <= TOL


@pytest.mark.gpu
def test_noalloc_gemm_limits_and_arguments(pkg, oracle, cuda_lib):
    from gpu_util import to_dev
    L = cuda_lib
    st = torch.cuda.current_stream().cuda_stream
    K, N = 512, 256
    x, qw, s, qz = oracle.make_inputs(17, K, N, G, seed=2)
    dx, ds, dqz = to_dev(x), to_dev(s), to_dev(qz)
    r = pkg.ops.quick_repack(to_dev(qw))
    y = torch.zeros((17, N), dtype=torch.float16, device="cuda")
    args = (dx.data_ptr(), r.data_ptr(), ds.data_ptr(), dqz.data_ptr(), y.data_ptr())
    rc = L.b200_w4a16_gemm_noalloc(*args, 17, K, N, G, st)              # beyond the built-in scratch
    assert rc == pkg._lib.B200_ERR_STATIC_LIMIT
    assert b"16 tokens" in L.b200_w4a16_strerror(rc)
    assert L.b200_w4a16_gemm_noalloc(*args, 4096, K, N, G, st) == pkg._lib.B200_ERR_STATIC_LIMIT
    assert L.b200_w4a16_gemm_noalloc(*args, 4, K, N, 64, st) == pkg._lib.B200_ERR_BAD_SHAPE   # G % 128
    assert L.b200_w4a16_gemm_noalloc(*args, 4, K, N + 8, G, st) == pkg._lib.B200_ERR_BAD_SHAPE
    assert L.b200_w4a16_gemm_noalloc(None, *args[1:], 4, K, N, G, st) == pkg._lib.B200_ERR_BAD_ARG
    assert L.b200_w4a16_gemm_noalloc(*args, 0, K, N, G, st) == 0
    torch.cuda.synchronize()
    assert not y.any()                                                  # nothing ran
    # the workspace entry point is unaffected by calls on the built-in scratch, and vice versa
    a = _with_workspace(L, dx[:16].contiguous(), r, ds, dqz, 16, K, N)
    assert L.b200_w4a16_gemm_noalloc(*args, 16, K, N, G, st) == 0
    torch.cuda.synchronize()
    assert torch.equal(y[:16], a)
    assert torch.equal(_with_workspace(L, dx[:16].contiguous(), r, ds, dqz, 16, K, N), a)
