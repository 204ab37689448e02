"""Quantized Mixtral expert path (reference: vllm/model_executor/models/mixtral_quant.py:57-160).

CPU part: routing, masking, accumulation and the expert split of ``MixtralMoE`` against a
plain restatement of the reference's dense loop, with a test-only linear method that evaluates
the packed weights on the CPU oracle (same device as test_parallel.py).  GPU part (-m gpu):
the gated GEMM through the C ABI and the whole block on the CUDA path, idle experts skipped
or not, eager and under CUDA-graph replay."""
import importlib
import os
import sys
import types

import numpy as np
import pytest
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

G = 128


def _cfg(experts=8, top_k=2, hidden=256, ffn=512):
    return types.SimpleNamespace(num_local_experts=experts, num_experts_per_tok=top_k,
                                 hidden_size=hidden, intermediate_size=ffn)


def _fill(moe, oracle, seed, device="cpu"):
    """Random AWQ checkpoints for every local expert; returns the fp32 dequantized weights."""
    dense = {}
    for e, mlp in enumerate(moe.experts):
        if mlp is None:
            continue
        for j, (name, lin) in enumerate((("w1", mlp.w1), ("w2", mlp.w2), ("w3", mlp.w3))):
            K, N = lin.input_size, lin.output_size
            _, qw, s, qz = oracle.make_inputs(1, K, N, G, seed=seed + 10 * e + j)
            lin.qweight.data.copy_(torch.from_numpy(qw))
            lin.scales.data.copy_(torch.from_numpy(s.view(np.int16)).view(torch.float16))
            lin.qzeros.data.copy_(torch.from_numpy(qz))
            dense[(e, name)] = torch.from_numpy(oracle.dequantize(qw, s, qz).astype(np.float32))
    g = torch.Generator().manual_seed(seed)
    moe.gate.weight.data.copy_(torch.randn(moe.gate.weight.shape, generator=g) * 0.5)
    return dense


def _reference_block(x, gate_w, dense, experts, top_k, round_half, with_mag=False):
    """The reference's dense loop (mixtral_quant.py:133-157) on dequantized weights."""
    r = (lambda t: t.half().float()) if round_half else (lambda t: t)
    logits = r(x @ gate_w.float().t())
    w = F.softmax(logits, dim=1, dtype=torch.float)
    w, sel = torch.topk(w, top_k, dim=-1)
    w = w / w.sum(dim=-1, keepdim=True)
    out = torch.zeros(x.shape[0], dense[(experts[0], "w2")].shape[1])
    mag = torch.zeros_like(out)
    for e in experts:
        ew = (w * (sel == e)).sum(dim=-1, keepdim=True)
        h1 = r(F.silu(r(x @ dense[(e, "w1")])))
        h3 = r(x @ dense[(e, "w3")])
        y = r(r(h1 * h3) @ dense[(e, "w2")])
        out = r(out + r(y * ew))
        mag += (y * ew).abs()
    if with_mag:       # sum of the experts' absolute contributions: the scale rounding errors live on
        return out, sel, mag
    return out, sel


def _close(a, b, tol=4e-3):
    """Agreement up to the fp32 summation order inside the GEMMs (a few fp16 ulps after them)."""
    a, b = a.float(), b.float()
    return bool(((a - b).abs() <= tol * b.abs().clamp(min=1.0)).all())


def test_moe_cpu_matches_reference_loop(pkg, oracle):
    class OracleMethod(pkg.AWQLinearMethod):
        def apply_weights(self, weights, x, bias=None):
            W = oracle.dequantize(weights["qweight"].data.numpy(),
                                  weights["scales"].data.numpy().view(np.float16),
                                  weights["qzeros"].data.numpy())
            return x.float() @ torch.from_numpy(W.astype(np.float32))

    torch.set_default_dtype(torch.float16)
    try:
        moe = pkg.MixtralMoE(_cfg(), OracleMethod(pkg.AWQConfig(4, G, True)))
    finally:
        torch.set_default_dtype(torch.float32)
    assert moe.expert_indicies == list(range(8))           # world of one: every expert is local
    assert [n for n, _ in moe.named_parameters()][:4] == [
        "experts.0.w1.qweight", "experts.0.w1.qzeros", "experts.0.w1.scales", "experts.0.w2.qweight"]
    dense = _fill(moe, oracle, seed=3)
    g = torch.Generator().manual_seed(0)
    x = torch.randn(1, 5, 256, generator=g)
    moe.gate.weight.data = moe.gate.weight.data.float()
    got = moe(x).reshape(5, 256)
    want, sel = _reference_block(x.reshape(5, 256), moe.gate.weight.data, dense, list(range(8)), 2, False)
    assert torch.allclose(got.float(), want, rtol=1e-4, atol=1e-4)
    _, _, counts = moe.route(x.reshape(5, 256))
    assert counts.dtype == torch.int32 and int(counts.sum()) == 10
    assert counts.tolist() == [int((sel == e).sum()) for e in range(8)]


def test_path_policy_by_batch_size(pkg):
    """Which body a batch takes (MixtralMoE.select_path): decode batches grouped, then the gated
    dense loop, device-routed from 65 tokens, host-routed from 8192 -- never under graph capture."""
    torch.set_default_dtype(torch.float16)
    try:
        moe = pkg.MixtralMoE(_cfg(), pkg.AWQLinearMethod(pkg.AWQConfig(4, G, True)))
    finally:
        torch.set_default_dtype(torch.float32)
    yes, no = (lambda: True), (lambda: False)
    group = lambda t: (lambda: t <= moe.grouped_max_tokens)
    pick = lambda t, capturing=False, fused=True: moe.select_path(t, True, capturing, group(t), fused)
    assert [pick(t) for t in (1, 16, 17, 64, 65, 4096, 8191, 8192, 100000)] == [
        "grouped", "grouped", "dense", "dense", "routed_on_device", "routed_on_device", "routed_on_device",
        "routed", "routed"]
    assert pick(8192, capturing=True) == "routed_on_device" and pick(100000, capturing=True) == "routed_on_device"
    assert pick(300, fused=False) == "dense"                  # a linear method without the fused epilogue
    assert moe.select_path(8, True, False, no, True) == "dense"     # e.g. group size 64: no grouped launch
    assert moe.select_path(8, False, False, yes, True) == "dense"   # host tensors: the reference's loop
    moe.skip_idle_experts = False
    assert [pick(t) for t in (1, 300, 9000)] == ["dense"] * 3  # the reference's behaviour on request
    for name in ("grouped", "dense", "routed_on_device", "routed"):
        assert callable(getattr(moe, "_forward_" + name))


def test_expert_gate_context_validates(pkg):
    with pytest.raises(ValueError):
        with pkg.ops.expert_gate(torch.zeros(1, dtype=torch.int32)):   # not a CUDA tensor
            pass
    with pkg.ops.expert_gate(None):
        pass


# ------------------------------------------------------------------------------- GPU parity

@pytest.mark.gpu
@pytest.mark.parametrize("M", [1, 7, 16])
def test_gated_gemm_skips_and_computes(pkg, oracle, cuda_lib, M):
    from gpu_util import to_dev, bits
    K, N = 1024, 768
    x, qw, s, qz = oracle.make_inputs(M, K, N, G, seed=21 + M)
    dx, ds, dqz = to_dev(x), to_dev(s), to_dev(qz)
    r = pkg.ops.quick_repack(to_dev(qw))
    plain = pkg.ops.quick_gemm(dx, r, ds, dqz, 8)
    counts = torch.tensor([0, 3], dtype=torch.int32, device="cuda")
    with pkg.ops.expert_gate(counts[1:2]):
        on = pkg.ops.quick_gemm(dx, r, ds, dqz, 8)
    with pkg.ops.expert_gate(counts[0:1]):
        off = pkg.ops.quick_gemm(dx, r, ds, dqz, 8)
    after = pkg.ops.quick_gemm(dx, r, ds, dqz, 8)          # the pipeline state survives a skipped launch
    torch.cuda.synchronize()
    assert np.array_equal(bits(on), bits(plain))
    assert np.array_equal(bits(after), bits(plain))
    assert not off.any()


@pytest.mark.gpu
def test_gated_gemm_ignores_gate_above_decode_batch(pkg, oracle, cuda_lib):
    from gpu_util import to_dev, bits
    M, K, N = 40, 1024, 512
    x, qw, s, qz = oracle.make_inputs(M, K, N, G, seed=5)
    dx, ds, dqz = to_dev(x), to_dev(s), to_dev(qz)
    r = pkg.ops.quick_repack(to_dev(qw))
    plain = pkg.ops.quick_gemm(dx, r, ds, dqz, 8)
    zero = torch.zeros(1, dtype=torch.int32, device="cuda")
    with pkg.ops.expert_gate(zero):
        got = pkg.ops.quick_gemm(dx, r, ds, dqz, 8)
    assert np.array_equal(bits(got), bits(plain))


@pytest.mark.gpu
@pytest.mark.parametrize("tokens", [1, 5, 16])
def test_moe_block_parity_and_idle_expert_skipping(pkg, oracle, cuda_lib, tokens):
    torch.set_default_dtype(torch.float16)
    try:
        method = pkg.QuickConfig(4, G, True).get_linear_method()
        with torch.device("cuda"):
            moe = pkg.MixtralMoE(_cfg(), method)
            dense_moe = pkg.MixtralMoE(_cfg(), method, skip_idle_experts=False)
    finally:
        torch.set_default_dtype(torch.float32)
    dense = _fill(moe, oracle, seed=7)
    _fill(dense_moe, oracle, seed=7)
    g = torch.Generator().manual_seed(tokens)
    x = torch.randn(1, tokens, 256, generator=g).half()
    want, sel = _reference_block(x.reshape(tokens, 256).float(), moe.gate.weight.data.float().cpu(),
                                 dense, list(range(8)), 2, True)
    xd = x.cuda()
    got = moe(xd).reshape(tokens, 256)
    got_dense = dense_moe(xd).reshape(tokens, 256)
    torch.cuda.synchronize()
    # The grouped launch cuts the (live expert, tile, K block) space differently from eight
    # single launches: other fp32 summation order inside the GEMMs, same roundings after them.
    assert _close(got, got_dense)
    moe.grouped_max_tokens = 0                             # one gated launch per expert instead
    assert torch.equal(moe(xd).reshape(tokens, 256), got_dense)   # idle experts skipped or not: same bits
    moe.grouped_max_tokens = 16
    assert torch.equal(moe(xd).reshape(tokens, 256), got)          # deterministic
    err = (got.float().cpu() - want).abs()
    tol = 1e-2 * want.abs().clamp(min=1.0)                 # bar of the GEMM tests, three GEMMs deep: x2
    assert bool((err <= 2 * tol).all()), float((err / want.abs().clamp(min=1.0)).max())
    if tokens == 1:
        assert len(set(sel.reshape(-1).tolist())) == 2     # six of eight experts were idle

    # CUDA-graph capture: the gates are read on the device, replay with other tokens re-routes
    st = torch.cuda.Stream()
    static_x = xd.clone()
    with torch.cuda.stream(st):
        moe(static_x)
        st.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=st):
            static_y = moe(static_x)
        x2 = torch.randn(1, tokens, 256, generator=g).half().cuda()
        static_x.copy_(x2)
        graph.replay()
        st.synchronize()
    assert torch.equal(static_y, moe(x2))
    assert _close(static_y, dense_moe(x2))


@pytest.mark.gpu
def test_moe_gather_scatter_kernels(pkg, oracle, cuda_lib):
    """b200_w4a16_moe_gather / _scatter_add against NumPy: gather is a byte copy (exact), the
    scatter follows mul_ (fp32 product, one rounding) then add_ (fp16 sum, one rounding)
    of mixtral_quant.py:152-157 -- exact as well."""
    from gpu_util import bits, to_dev
    rng = np.random.default_rng(0)
    T, H = 300, 1024
    x = rng.standard_normal((T, H)).astype(np.float16)
    idx = np.sort(rng.choice(T, 77, replace=False)).astype(np.int32)
    g = pkg.ops.moe_gather(to_dev(x), torch.from_numpy(idx).cuda())
    assert np.array_equal(bits(g), x[idx].view(np.uint16))
    acc = rng.standard_normal((T, H)).astype(np.float16)
    y = rng.standard_normal((77, H)).astype(np.float16)
    w = rng.random(77).astype(np.float32)
    dacc = to_dev(acc)
    pkg.ops.moe_scatter_add(dacc, to_dev(y), torch.from_numpy(idx).cuda(), torch.from_numpy(w).cuda())
    want = acc.copy()
    scaled = (y.astype(np.float32) * w[:, None]).astype(np.float16)
    want[idx] = (acc[idx].astype(np.float32) + scaled.astype(np.float32)).astype(np.float16)
    assert np.array_equal(bits(dacc), want.view(np.uint16))
    # empty selection, bad arguments
    e = torch.zeros(0, dtype=torch.int32, device="cuda")
    assert pkg.ops.moe_gather(to_dev(x), e).shape == (0, H)
    with pytest.raises(ValueError):
        pkg.ops.moe_gather(to_dev(x), torch.zeros(3, dtype=torch.int64, device="cuda"))
    s = torch.cuda.current_stream().cuda_stream
    assert cuda_lib.b200_w4a16_moe_gather(dacc.data_ptr(), e.data_ptr(), dacc.data_ptr(), 1, 12, s) == pkg._lib.B200_ERR_BAD_SHAPE
    assert cuda_lib.b200_w4a16_moe_gather(None, e.data_ptr(), dacc.data_ptr(), 1, 16, s) == pkg._lib.B200_ERR_BAD_ARG


@pytest.mark.gpu
@pytest.mark.parametrize("tokens", [17, 64, 333])
def test_moe_block_routed_matches_dense_and_reference(pkg, oracle, cuda_lib, tokens):
    """Prefill batches take the gather -> per-expert M_e -> scatter-add path; it must agree
    with the dense all-tokens loop (the reference's structure) and with the CPU block."""
    torch.set_default_dtype(torch.float16)
    try:
        method = pkg.QuickConfig(4, G, True).get_linear_method()
        with torch.device("cuda"):
            moe = pkg.MixtralMoE(_cfg(), method)
            dense_moe = pkg.MixtralMoE(_cfg(), method, skip_idle_experts=False)
    finally:
        torch.set_default_dtype(torch.float32)
    moe.routed_min_tokens = 17              # the product threshold is a tuning value (512)
    dense = _fill(moe, oracle, seed=9)
    _fill(dense_moe, oracle, seed=9)
    g = torch.Generator().manual_seed(tokens)
    x = torch.randn(1, tokens, 256, generator=g).half()
    want, _ = _reference_block(x.reshape(tokens, 256).float(), moe.gate.weight.data.float().cpu(),
                               dense, list(range(8)), 2, True)
    calls = []
    orig = pkg.ops.moe_gather
    pkg.ops.moe_gather = lambda *a: (calls.append(a[1].numel()), orig(*a))[1]
    try:
        got = moe(x.cuda()).reshape(tokens, 256)
    finally:
        pkg.ops.moe_gather = orig
    assert sum(calls) == 2 * tokens and len(calls) <= 8        # top-2: every token in two batches
    got_dense = dense_moe(x.cuda()).reshape(tokens, 256)
    # the expert GEMMs run at another batch size (other tile plan, other fp32 summation order):
    # agreement to the GEMM bar, not bits
    ref = got_dense.float()
    assert bool(((got.float() - ref).abs() <= 2e-2 * ref.abs().clamp(min=1.0)).all())
    err = (got.float().cpu() - want).abs()
    assert bool((err <= 2e-2 * want.abs().clamp(min=1.0)).all())
    assert torch.equal(got, moe(x.cuda()).reshape(tokens, 256))   # deterministic


@pytest.mark.gpu
@pytest.mark.parametrize("tokens,routed", [(3, False), (16, False), (64, False), (40, True)])
def test_moe_block_at_mixtral_shapes(pkg, oracle, cuda_lib, tokens, routed):
    """BASELINE config 5 shapes (hidden 4096, ffn 14336; two local experts keep the host
    reference affordable): merged w1/w3 GEMM with the activation in its epilogue (<= 16 tokens)
    or GEMM + paired activation kernel (above), routing-weight accumulate kernel, gated experts;
    dense loop and routed path against the CPU restatement of the reference's loop."""
    cfg = _cfg(experts=2, top_k=2, hidden=4096, ffn=14336)
    torch.set_default_dtype(torch.float16)
    try:
        method = pkg.QuickConfig(4, G, True).get_linear_method()
        with torch.device("cuda"):
            moe = pkg.MixtralMoE(cfg, method)
    finally:
        torch.set_default_dtype(torch.float32)
    if routed:
        moe.routed_min_tokens = tokens
    dense = _fill(moe, oracle, seed=13)
    g = torch.Generator().manual_seed(tokens)
    x = (torch.randn(1, tokens, 4096, generator=g) * 0.25).half()
    want, _, mag = _reference_block(x.reshape(tokens, 4096).float(), moe.gate.weight.data.float().cpu(),
                                    dense, [0, 1], 2, True, with_mag=True)
    got = moe(x.cuda()).reshape(tokens, 4096)
    torch.cuda.synchronize()
    assert moe.experts[0].w1.qweight.numel() == 0          # merged into one gate/up weight, originals released
    assert moe.experts[0]._w13["scales"].shape == (4096 // G, 2 * 14336)
    # K = 14336 makes single expert outputs reach +-45 (fp16 spacing 3e-2) while their weighted
    # sum can cancel: twice the bar of the GEMM tests (three GEMMs deep), relative to the experts' absolute contributions
    err = (got.float().cpu() - want).abs()
    assert bool((err <= 2e-2 * mag.clamp(min=1.0)).all()), float((err / mag.clamp(min=1.0)).max())
    assert torch.equal(got, moe(x.cuda()).reshape(tokens, 4096))   # deterministic


@pytest.mark.gpu
@pytest.mark.parametrize("M", [1, 8, 16])
@pytest.mark.parametrize("act", [False, True])
def test_grouped_gemm_matches_per_expert_launches(pkg, oracle, cuda_lib, M, act):
    """b200_w4a16_moe_grouped_gemm: five experts in one launch against one launch per expert;
    experts without tokens are neither streamed nor written; one live expert = the single
    launch bit for bit (same walk)."""
    from gpu_util import to_dev, bits
    K, N, E = 1024, 768, 5
    ws = []
    for e in range(E):
        x, qw, s, qz = oracle.make_inputs(M, K, N, G, seed=100 + 7 * e + M)
        ws.append((to_dev(x), pkg.ops.quick_repack(to_dev(qw)), to_dev(s), to_dev(qz)))
    single = (lambda x, w: pkg.ops.quick_gemm_silu_mul(x, w[1], w[2], w[3])) if act else \
             (lambda x, w: pkg.ops.quick_gemm(x, w[1], w[2], w[3], 8))
    xs = torch.stack([w[0] for w in ws]).contiguous()      # [E, M, K]
    NO = N // 2 if act else N
    for shared_x in (True, False):
        xin = ws[0][0] if shared_x else xs
        for counts in ([2, 0, 1, 0, 3], [0, 0, 0, 4, 0], [1, 1, 1, 1, 1], [0, 0, 0, 0, 0]):
            dc = torch.tensor(counts, dtype=torch.int32, device="cuda")
            out = torch.full((E, M, NO), 7.0, dtype=torch.float16, device="cuda")
            got = pkg.ops.moe_grouped_gemm(xin, [w[1] for w in ws], [w[2] for w in ws], [w[3] for w in ws],
                                           dc, act=act, out=out)
            again = pkg.ops.moe_grouped_gemm(xin, [w[1] for w in ws], [w[2] for w in ws], [w[3] for w in ws],
                                             dc, act=act)
            torch.cuda.synchronize()
            assert got.data_ptr() == out.data_ptr()
            for e in range(E):
                if counts[e] == 0:
                    assert bool((got[e] == 7.0).all())     # untouched
                    continue
                want = single(ws[0][0] if shared_x else ws[e][0], ws[e])
                assert _close(got[e], want), (counts, e)
                assert torch.equal(got[e], again[e])
                if sum(c > 0 for c in counts) == 1:
                    assert np.array_equal(bits(got[e]), bits(want))
    # the plain launches still work after grouped ones (workspace handed back clean)
    assert np.array_equal(bits(single(ws[1][0], ws[1])), bits(single(ws[1][0], ws[1])))


@pytest.mark.gpu
def test_grouped_gemm_and_combine_arguments(pkg, oracle, cuda_lib):
    from gpu_util import to_dev
    K, N = 256, 256
    x, qw, s, qz = oracle.make_inputs(4, K, N, G, seed=3)
    w = (pkg.ops.quick_repack(to_dev(qw)), to_dev(s), to_dev(qz))
    c1 = torch.ones(1, dtype=torch.int32, device="cuda")
    with pytest.raises(ValueError):                        # nine experts
        pkg.ops.moe_grouped_gemm(to_dev(x), [w[0]] * 9, [w[1]] * 9, [w[2]] * 9,
                                 torch.ones(9, dtype=torch.int32, device="cuda"))
    x17 = torch.zeros((17, K), dtype=torch.float16, device="cuda")
    with pytest.raises(ValueError):                        # above the decode kernel's batch sizes
        pkg.ops.moe_grouped_gemm(x17, [w[0]], [w[1]], [w[2]], c1)
    with pytest.raises(ValueError):                        # counts must be int32 [E]
        pkg.ops.moe_grouped_gemm(to_dev(x), [w[0]], [w[1]], [w[2]], torch.ones(2, dtype=torch.int32, device="cuda"))
    with pytest.raises(ValueError):
        pkg.ops.moe_grouped_gemm(to_dev(x), [w[0]], [w[1]], [w[2]], torch.ones(1, dtype=torch.int64, device="cuda"))
    with pytest.raises(ValueError):                        # x per expert: [E, M, K]
        pkg.ops.moe_grouped_gemm(torch.zeros((2, 4, K), dtype=torch.float16, device="cuda"), [w[0]], [w[1]], [w[2]], c1)
    empty = pkg.ops.moe_grouped_gemm(torch.zeros((0, K), dtype=torch.float16, device="cuda"), [w[0]], [w[1]], [w[2]], c1)
    assert empty.shape == (1, 0, N)
    with pytest.raises(ValueError):
        pkg.ops.moe_combine(torch.zeros((2, 3, 64), dtype=torch.float16, device="cuda"),
                            torch.zeros((3, 1), device="cuda"), torch.ones(2, dtype=torch.int32, device="cuda"))


@pytest.mark.gpu
def test_moe_combine_kernel(pkg, cuda_lib):
    """b200_w4a16_moe_combine against NumPy: per live expert, in index order, mul_ (fp32 product,
    one rounding) then add_ (fp16 sum, one rounding) -- mixtral_quant.py:152-157, exact."""
    from gpu_util import bits, to_dev
    rng = np.random.default_rng(1)
    E, T, H, total = 4, 13, 512, 8
    y = rng.standard_normal((E, T, H)).astype(np.float16)
    y[2] = np.float16(np.nan)                              # expert 2 has no tokens: its slice is never read
    w = rng.random((T, total)).astype(np.float32)
    counts = np.array([3, 1, 0, 5], dtype=np.int32)
    first = 3
    dw = torch.from_numpy(w).cuda()
    got = pkg.ops.moe_combine(to_dev(y), dw[:, first:first + E], torch.from_numpy(counts).cuda())
    want = np.zeros((T, H), dtype=np.float16)
    for e in range(E):
        if counts[e] > 0:
            scaled = (y[e].astype(np.float32) * w[:, first + e, None]).astype(np.float16)
            want = (want.astype(np.float32) + scaled.astype(np.float32)).astype(np.float16)
    assert np.array_equal(bits(got), want.view(np.uint16))
    none = pkg.ops.moe_combine(to_dev(y), dw[:, first:first + E], torch.zeros(E, dtype=torch.int32, device="cuda"))
    assert not none.any()


@pytest.mark.gpu
@pytest.mark.parametrize("T,E,H,k", [(1, 8, 4096, 2), (16, 8, 4096, 2), (333, 8, 256, 2), (40, 20, 512, 3)])
def test_router_kernel_matches_the_reference_arithmetic(pkg, cuda_lib, T, E, H, k):
    """b200_w4a16_moe_route against mixtral_quant.py:133-142 in float64 on the host (fp16 logits,
    softmax, top-k, renormalise); outputs consistent with each other exactly."""
    g = torch.Generator().manual_seed(T + E)
    x = torch.randn(T, H, generator=g).half()
    gw = (torch.randn(E, H, generator=g) * (2.0 / H ** 0.5)).half()
    w, idx, counts, dense = pkg.ops.moe_route(x.cuda(), gw.cuda(), k)
    torch.cuda.synchronize()
    w, idx, counts, dense = w.cpu(), idx.cpu(), counts.cpu(), dense.cpu()
    logits = (x.double() @ gw.double().t()).half().double()
    p = torch.softmax(logits, dim=1)
    rw, ridx = torch.topk(p, k, dim=-1)
    nxt = torch.topk(p, k + 1, dim=-1)[0][:, k] if E > k else torch.zeros(T, dtype=torch.float64)
    # an fp16 logit can round the other way in another summation order (2^-11 relative):
    # compare where no two adjacent ranks are that close
    gaps = torch.cat([rw, nxt[:, None]], dim=1)
    clear = ((gaps[:, :-1] - gaps[:, 1:]) > 4e-3 * gaps[:, :-1]).all(dim=1)
    assert int(clear.sum()) >= T // 2 or T == 1
    assert torch.equal(idx[clear], ridx[clear])
    rw = rw / rw.sum(dim=-1, keepdim=True)
    assert torch.allclose(w[clear].double(), rw[clear], rtol=5e-3, atol=1e-6)
    # self-consistency, exact
    assert idx.dtype == torch.int64 and w.dtype == torch.float32 and counts.dtype == torch.int32
    assert torch.equal(counts.long(), torch.bincount(idx.reshape(-1), minlength=E))
    assert all(len(set(r)) == k for r in idx.tolist())               # an expert at most once per token
    assert bool((w[:, :-1] >= w[:, 1:]).all())                        # descending, as torch.topk
    assert torch.equal(torch.zeros(T, E).scatter_(1, idx, w), dense)
    assert torch.allclose(w.sum(dim=1), torch.ones(T), atol=1e-6)


@pytest.mark.gpu
def test_router_kernel_arguments(pkg, cuda_lib):
    x = torch.zeros((0, 64), dtype=torch.float16, device="cuda")
    gw = torch.zeros((4, 64), dtype=torch.float16, device="cuda")
    w, idx, counts, dense = pkg.ops.moe_route(x, gw, 2)
    assert w.shape == (0, 2) and not counts.any()
    with pytest.raises(ValueError):
        pkg.ops.moe_route(torch.zeros((2, 64), dtype=torch.float16, device="cuda"), gw, 5)      # top_k > experts
    with pytest.raises(ValueError):
        pkg.ops.moe_route(torch.zeros((2, 60), dtype=torch.float16, device="cuda"),
                          torch.zeros((4, 60), dtype=torch.float16, device="cuda"), 2)            # hidden % 8
    with pytest.raises(ValueError):
        pkg.ops.moe_route(torch.zeros((2, 64), dtype=torch.float16, device="cuda"),
                          torch.zeros((33, 64), dtype=torch.float16, device="cuda"), 2)           # > 32 experts


@pytest.mark.gpu
@pytest.mark.parametrize("M_max,K,N", [(32, 512, 384), (64, 1024, 768), (200, 512, 1024), (300, 2048, 256)])
def test_gemm_with_device_side_row_count(pkg, oracle, cuda_lib, M_max, K, N):
    """b200_w4a16_gemm_rows: the live rows equal b200_w4a16_gemm on exactly those rows up to the
    fp32 summation order (other stream-K cut); rows beyond the count are never written, whatever
    the input holds there (NaN here)."""
    from gpu_util import to_dev
    x, qw, s, qz = oracle.make_inputs(M_max, K, N, G, seed=M_max)
    r, ds, dqz = pkg.ops.quick_repack(to_dev(qw)), to_dev(s), to_dev(qz)
    dx = to_dev(x)
    for rows in sorted({0, 1, 17, 32, 33, min(M_max, 129), M_max - 1, M_max, M_max + 50}):
        live = min(rows, M_max)
        xin = dx.clone()
        xin[live:] = float("nan")
        cnt = torch.tensor([rows], dtype=torch.int32, device="cuda")
        # run it twice: the second call overwrites a result buffer pre-filled with a sentinel
        got = pkg.ops.quick_gemm_rows(xin, r, ds, dqz, cnt)
        torch.cuda.synchronize()
        if live:
            want = pkg.ops.quick_gemm(dx[:live].contiguous(), r, ds, dqz, 8)
            assert _close(got[:live], want), rows
            assert not torch.isnan(got[:live]).any()
        again = pkg.ops.quick_gemm_rows(xin, r, ds, dqz, cnt)
        assert torch.equal(got[:live], again[:live])
    # the workspace is handed back clean: a plain launch afterwards is unaffected
    a = pkg.ops.quick_gemm(dx, r, ds, dqz, 8)
    assert torch.equal(a, pkg.ops.quick_gemm(dx, r, ds, dqz, 8))
    with pytest.raises(ValueError):
        pkg.ops.quick_gemm_rows(dx, r, ds, dqz, torch.tensor([1], dtype=torch.int64, device="cuda"))


@pytest.mark.gpu
def test_gemm_rows_never_writes_beyond_the_count(pkg, oracle, cuda_lib):
    from gpu_util import to_dev
    M_max, K, N = 160, 512, 256
    x, qw, s, qz = oracle.make_inputs(M_max, K, N, G, seed=4)
    r, ds, dqz, dx = pkg.ops.quick_repack(to_dev(qw)), to_dev(s), to_dev(qz), to_dev(x)
    L = cuda_lib
    y = torch.full((M_max, N), 3.0, dtype=torch.float16, device="cuda")
    cnt = torch.tensor([70], dtype=torch.int32, device="cuda")
    ws = torch.empty(L.b200_w4a16_workspace_bytes(M_max, K, N, 1), dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    assert L.b200_w4a16_workspace_init(ws.data_ptr(), ws.numel(), st) == 0
    assert L.b200_w4a16_gemm_rows(dx.data_ptr(), r.data_ptr(), ds.data_ptr(), dqz.data_ptr(), y.data_ptr(),
                                  M_max, cnt.data_ptr(), K, N, G, ws.data_ptr(), ws.numel(), st) == 0
    torch.cuda.synchronize()
    assert bool((y[70:] == 3.0).all()) and not bool((y[:70] == 3.0).all())
    assert L.b200_w4a16_gemm_rows(dx.data_ptr(), r.data_ptr(), ds.data_ptr(), dqz.data_ptr(), y.data_ptr(),
                                  M_max, None, K, N, G, ws.data_ptr(), ws.numel(), st) == pkg._lib.B200_ERR_BAD_ARG


@pytest.mark.gpu
def test_moe_compact_and_row_counted_helpers(pkg, cuda_lib):
    from gpu_util import bits, to_dev
    rng = np.random.default_rng(5)
    T, E, k, H = 333, 8, 2, 256
    sel = np.stack([rng.choice(E, k, replace=False) for _ in range(T)]).astype(np.int64)
    wts = rng.random((T, k)).astype(np.float32)
    first, n_local = 2, 4
    tok, w = pkg.ops.moe_compact(torch.from_numpy(sel).cuda(), torch.from_numpy(wts).cuda(), first, n_local)
    torch.cuda.synchronize()
    x = rng.standard_normal((T, H)).astype(np.float16)
    for j in range(n_local):
        hits = np.argwhere(sel == first + j)
        n = len(hits)
        assert np.array_equal(tok[j, :n].cpu().numpy(), hits[:, 0].astype(np.int32))
        assert np.array_equal(w[j, :n].cpu().numpy(), wts[hits[:, 0], hits[:, 1]])
        cnt = torch.tensor([n], dtype=torch.int32, device="cuda")
        g = torch.full((T, H), 9.0, dtype=torch.float16, device="cuda")
        got = pkg.ops.moe_gather(to_dev(x), tok[j], cnt)
        assert np.array_equal(bits(got[:n]), x[hits[:, 0]].view(np.uint16))
        # scatter-add of the live rows only (the rest of tok / w is uninitialised)
        acc = torch.zeros((T, H), dtype=torch.float16, device="cuda")
        pkg.ops.moe_scatter_add(acc, got, tok[j], w[j], cnt)
        want = np.zeros((T, H), dtype=np.float16)
        want[hits[:, 0]] = (x[hits[:, 0]].astype(np.float32) * wts[hits[:, 0], hits[:, 1], None]).astype(np.float16)
        assert np.array_equal(bits(acc), want.view(np.uint16))
        # activation on the live rows only
        src = to_dev(rng.standard_normal((T, 2 * H)).astype(np.float16))
        dst = torch.full((T, H), 9.0, dtype=torch.float16, device="cuda")
        full = torch.empty((T, H), dtype=torch.float16, device="cuda")
        pkg.ops.silu_and_mul_paired(dst, src, cnt)
        pkg.ops.silu_and_mul_paired(full, src)
        assert torch.equal(dst[:n], full[:n]) and bool((dst[n:] == 9.0).all())
        del g


@pytest.mark.gpu
@pytest.mark.parametrize("tokens", [17, 64, 333])
def test_moe_block_routed_on_device_matches_dense_and_reference(pkg, oracle, cuda_lib, tokens):
    """The routed computation with every row count read on the device: against the dense loop,
    the CPU block, and under CUDA-graph replay with tokens that route differently."""
    torch.set_default_dtype(torch.float16)
    try:
        method = pkg.QuickConfig(4, G, True).get_linear_method()
        with torch.device("cuda"):
            moe = pkg.MixtralMoE(_cfg(), method)
            dense_moe = pkg.MixtralMoE(_cfg(), method, skip_idle_experts=False)
    finally:
        torch.set_default_dtype(torch.float32)
    moe.routed_on_device_min_tokens = 17
    dense = _fill(moe, oracle, seed=11)
    _fill(dense_moe, oracle, seed=11)
    g = torch.Generator().manual_seed(tokens)
    x = torch.randn(1, tokens, 256, generator=g).half()
    want, _ = _reference_block(x.reshape(tokens, 256).float(), moe.gate.weight.data.float().cpu(),
                               dense, list(range(8)), 2, True)
    calls = []
    orig = pkg.ops.quick_gemm_rows
    pkg.ops.quick_gemm_rows = lambda *a: (calls.append(1), orig(*a))[1]
    try:
        got = moe(x.cuda()).reshape(tokens, 256)
    finally:
        pkg.ops.quick_gemm_rows = orig
    assert len(calls) == 16                                  # two GEMMs per local expert
    ref = dense_moe(x.cuda()).reshape(tokens, 256).float()
    assert bool(((got.float() - ref).abs() <= 2e-2 * ref.abs().clamp(min=1.0)).all())
    err = (got.float().cpu() - want).abs()
    assert bool((err <= 2e-2 * want.abs().clamp(min=1.0)).all())
    assert torch.equal(got, moe(x.cuda()).reshape(tokens, 256))   # deterministic

    st = torch.cuda.Stream()
    static_x = x.cuda().clone()
    with torch.cuda.stream(st):
        moe(static_x)
        st.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=st):
            static_y = moe(static_x)
        x2 = torch.randn(1, tokens, 256, generator=g).half().cuda()
        static_x.copy_(x2)
        graph.replay()
        st.synchronize()
    assert torch.equal(static_y, moe(x2))
