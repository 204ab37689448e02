"""The reference-side binding of INTEGRATION.md section A as code that runs: integration/b200_shim.cpp
(the INTEGRATION.md snippet verbatim + module registration) built by integration/build_shim.sh exposes
the reference's op signatures -- awq_gemm(in_feats, kernel, scaling_factors, zeros, split_k_iters),
awq_dequantize(kernel, scaling_factors, zeros, split_k_iters, thx, thy) (csrc/ops.h:69-82) -- over
libb200_w4a16.so.  CPU: the file and the document agree.  GPU: the ops called through the shim give the
bits of the package's own operators and stay within the GEMM bar of the fp32 oracle.  The GPU test
skips when the shim has not been built (it is built in-tree and travels with the snapshot)."""
import glob
import importlib.util
import os
import re
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

G = 128


def test_shim_source_is_the_integration_snippet():
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    snippet = re.search(r"```cpp\n(// csrc/quantization/awq/b200_shim\.cpp.*?)```", doc, re.S).group(1)
    src = open(os.path.join(ROOT, "integration", "b200_shim.cpp")).read()
    body = src.split("// ---- INTEGRATION.md section A: begin\n")[1].split("// ---- INTEGRATION.md section A: end")[0]
    assert body == snippet


def _load_shim():
    found = glob.glob(os.path.join(ROOT, "integration", "_build", "b200_shim*.so"))
    if not found:
        pytest.skip("integration/_build/b200_shim*.so not built (bash integration/build_shim.sh)")
    try:
        spec = importlib.util.spec_from_file_location("b200_shim", found[0])
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    except Exception as e:      # another torch / python than the one it was built against
        pytest.skip(f"b200_shim does not load here: {e!r}")
    return mod


@pytest.mark.gpu
@pytest.mark.parametrize("M,K,N", [(16, 4096, 4096), (5, 1024, 768), (200, 2048, 1024)])
def test_reference_op_signatures_through_the_shim(pkg, oracle, cuda_lib, M, K, N):
    from gpu_util import to_dev
    shim = _load_shim()
    x, qw, s, qz = oracle.make_inputs(M, K, N, G, seed=M + K)
    dx, dqw, ds, dqz = to_dev(x), to_dev(qw), to_dev(s), to_dev(qz)
    y = shim.awq_gemm(dx, dqw, ds, dqz, 8)                       # the reference's call (awq.py:160-162)
    want = pkg.ops.awq_gemm(dx, dqw, ds, dqz, 8)
    assert y.shape == (M, N) and y.dtype == torch.float16 and torch.equal(y, want)
    W = shim.awq_dequantize(dqw, ds, dqz, 0, 0, 0)               # awq.py:155-156
    assert torch.equal(W, pkg.ops.awq_dequantize(dqw, ds, dqz, 0, 0, 0))
    Wo = oracle.dequantize(qw, s, qz)
    assert np.array_equal(W.cpu().view(torch.int16).numpy().view(np.uint16), Wo.view(np.uint16))   # bit-exact
    Yf, _ = oracle.gemm_f32(x, Wo)
    err = np.abs(y.float().cpu().numpy() - Yf) / np.maximum(np.abs(Yf), 1.0)
    assert float(err.max()) <= 1e-2
    r, rp = shim.quick_repack(dqw), pkg.ops.quick_repack(dqw)   # the same words (the shim returns them flat)
    assert torch.equal(r.reshape(-1), rp.reshape(-1))
    assert torch.equal(shim.quick_gemm(dx, r, ds, dqz, 8), pkg.ops.quick_gemm(dx, rp, ds, dqz, 8))
    with pytest.raises(ValueError):                              # std::invalid_argument, as the reference
        shim.awq_gemm(dx, dqw[:, :-1].contiguous(), ds[:, :-8].contiguous(), dqz[:, :-1].contiguous(), 8)
