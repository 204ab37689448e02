"""-m gpu tests of the pre-interleaved checkpoint format (quantization/quick.py):
``convert_awq_state_dict`` output loaded with ``"version": "quick"`` must give, bit for bit, the
weights and results of the AWQ checkpoint interleaved at start-up ("version": "gemm")."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu


def test_converted_checkpoint_loads_without_repack(pkg, oracle):
    from gpu_util import to_dev
    quick = importlib_quick(pkg)
    K, N = 1024, 768
    _, qw, s, qz = oracle.make_inputs(1, K, N, 128, seed=4)
    sd = {"mlp.down_proj.qweight": to_dev(qw), "mlp.down_proj.scales": to_dev(s),
          "mlp.down_proj.qzeros": to_dev(qz), "norm.weight": torch.ones(8, device="cuda")}
    conv = quick.convert_awq_state_dict(sd)
    assert tuple(conv["mlp.down_proj.qweight"].shape) == (N // 128, K // 128, 2048)
    assert conv["mlp.down_proj.scales"] is sd["mlp.down_proj.scales"] and conv["norm.weight"] is sd["norm.weight"]
    # the converter's tiles are the oracle's layout, word for word
    want = oracle.np_b200_repack(qw).reshape(N // 128, K // 128, 2048)
    assert np.array_equal(conv["mlp.down_proj.qweight"].cpu().numpy(), want)
    layers = {}
    for version, tensors in (("quick", conv), ("gemm", sd)):
        method = pkg.QuickConfig(4, 128, True, version).get_linear_method()
        lin = pkg.ReplicatedLinear(K, N, bias=False, linear_method=method, params_dtype=torch.float16).cuda()
        for name in ("qweight", "scales", "qzeros"):
            getattr(lin, name).data.copy_(tensors["mlp.down_proj." + name])
        layers[version] = lin
    assert layers["quick"].linear_weights["qweight_layout"] == "b200"
    for M in (1, 16, 64):
        x = torch.randn(M, K, device="cuda").half()
        a, _ = layers["quick"](x)
        b, _ = layers["gemm"](x)
        assert torch.equal(a, b)
    assert torch.equal(layers["quick"].qweight.data.reshape(-1), layers["gemm"].qweight.data.reshape(-1))
    # a tensor off the 128 grid stays in AWQ form
    odd = quick.convert_awq_state_dict({"x.qweight": torch.zeros(128, 24, dtype=torch.int32, device="cuda")})
    assert tuple(odd["x.qweight"].shape) == (128, 24)


def test_pre_interleaved_gate_up_can_be_paired(pkg, oracle):
    """The fused SiLU-and-multiply path on a pre-interleaved checkpoint: pairing un-interleaves,
    permutes and interleaves again -- same results as from the AWQ checkpoint."""
    from gpu_util import to_dev
    quick = importlib_quick(pkg)
    K, inter = 512, 1024
    _, gq, gs, gz = oracle.make_inputs(1, K, inter, 128, seed=5)
    _, uq, us, uz = oracle.make_inputs(1, K, inter, 128, seed=6)
    outs = []
    for version in ("quick", "gemm"):
        method = pkg.QuickConfig(4, 128, True, version).get_linear_method()
        gu = pkg.MergedColumnParallelLinear(K, [inter, inter], bias=False, linear_method=method,
                                            params_dtype=torch.float16).cuda()
        for sid, (q, s, z) in enumerate(((gq, gs, gz), (uq, us, uz))):
            qt = to_dev(q)
            if version == "quick":
                qt = quick.convert_awq_state_dict({"qweight": qt})["qweight"]
            gu.qweight.weight_loader(gu.qweight, qt, sid)
            gu.scales.weight_loader(gu.scales, to_dev(s), sid)
            gu.qzeros.weight_loader(gu.qzeros, to_dev(z), sid)
        x = torch.randn(8, K, device="cuda", generator=torch.Generator(device="cuda").manual_seed(1)).half()
        outs.append(gu.forward_silu_and_mul(x))
    assert outs[0].shape == (8, inter) and torch.equal(outs[0], outs[1])


def importlib_quick(pkg):
    import importlib
    return importlib.import_module(pkg.__name__ + ".quantization.quick")
