"""-m gpu parity tests of the W4A16 GEMM through the C ABI / operator surface.

Bar (BASELINE.json north_star): max-abs / max-rel error <= 1e-2 against the fp32
dequant-then-matmul oracle, applied as |y - ref| <= 1e-2 * max(1, |ref|): absolute where
|ref| <= 1, relative above (the spacing of fp16 outputs is 1.6e-2 from |y| = 16 up, so a
pure absolute bound is unreachable for ANY fp16-output kernel, the reference's included).
Inputs are the seeded synthetic tensors of SURVEY.md 8(d)."""
import numpy as np
import pytest
import torch

from gpu_util import gemm_check, gemm_err, sampled_case, to_dev

pytestmark = pytest.mark.gpu
TOL = 1e-2


def _run(pkg, oracle, M, K, N, G, seed, awq_layout=False):
    x, qw, s, qz = oracle.make_inputs(M, K, N, G, seed=seed)
    Yf, _ = oracle.gemm_f32(x, oracle.dequantize(qw, s, qz))
    dx, dqw, ds, dqz = to_dev(x), to_dev(qw), to_dev(s), to_dev(qz)
    if awq_layout:
        y = pkg.ops.awq_gemm(dx, dqw, ds, dqz, 8)
    else:
        y = pkg.ops.quick_gemm(dx, pkg.ops.quick_repack(dqw), ds, dqz, 8)
    torch.cuda.synchronize()
    assert y.shape == (M, N) and y.dtype == torch.float16
    return gemm_err(y, Yf), y


@pytest.mark.parametrize("M,K,N,G", [
    (16, 128, 128, 128),     # one tile, one CTA
    (16, 256, 128, 128),     # two K blocks, stream-K split across two CTAs
    (16, 1024, 256, 128),
    (1, 512, 128, 128),
    (7, 512, 192, 64),       # ragged channel tile (N % 128 == 64), G = 64
    (3, 256, 64, 32),        # G = 32
    (16, 512, 256, 256),     # group spans two K blocks
])
def test_gemm_small(pkg, oracle, cuda_lib, M, K, N, G):
    (ea, er), _ = _run(pkg, oracle, M, K, N, G, seed=M + K + N)
    assert er <= TOL, (ea, er)


@pytest.mark.parametrize("M", [1, 2, 4, 8, 16, 17, 32, 33, 64, 65, 128, 200, 256])
def test_gemm_config1_shape_all_token_counts(pkg, oracle, cuda_lib, M):
    # BASELINE configs[0]: K = N = 4096, G = 128 (M = 16 is the quoted point)
    (ea, er), _ = _run(pkg, oracle, M, 4096, 4096, 128, seed=M)
    assert er <= TOL, (ea, er)


@pytest.mark.parametrize("K,N", [(4096, 12288), (4096, 22016), (11008, 4096)])
@pytest.mark.parametrize("M", [1, 16, 64])
def test_gemm_llama7b_shapes(pkg, oracle, cuda_lib, M, K, N):
    (ea, er), _ = _run(pkg, oracle, M, K, N, 128, seed=1)
    assert er <= TOL, (ea, er)


def test_awq_layout_entry_point_matches_quick(pkg, oracle, cuda_lib):
    (ea, er), y0 = _run(pkg, oracle, 16, 1024, 512, 128, seed=5, awq_layout=True)
    assert er <= TOL, (ea, er)
    _, y1 = _run(pkg, oracle, 16, 1024, 512, 128, seed=5, awq_layout=False)
    assert torch.equal(y0, y1)


def test_gemm_is_deterministic_and_workspace_self_cleans(pkg, oracle, cuda_lib):
    x, qw, s, qz = oracle.make_inputs(16, 4096, 4096, 128, seed=11)
    dx, ds, dqz = to_dev(x), to_dev(s), to_dev(qz)
    r = pkg.ops.quick_repack(to_dev(qw))
    ys = [pkg.ops.quick_gemm(dx, r, ds, dqz, 8) for _ in range(5)]
    torch.cuda.synchronize()
    for y in ys[1:]:
        assert torch.equal(ys[0], y)
    # every launch hands the split-K scratch back in its idle state (all words 0xFFFFFFFF)
    for ws in pkg.ops._workspaces.values():      # (beyond the slots: scratch of the AWQ-layout entry point)
        assert bool((ws[:148 * 2 * 2 * 128 * 128 * 4] == 0xFF).all())


@pytest.mark.parametrize("M", [16, 48, 200])
def test_two_graphs_on_the_default_capture_stream_replay_in_any_order(pkg, oracle, cuda_lib, M):
    """The reference captures its graphs with torch.cuda.graph(graph, pool=...) and no stream
    (worker/model_runner.py:781): all captures share the scratch created eagerly at load time,
    so a graph other than the first captured one may be the first to replay."""
    x, qw, s, qz = oracle.make_inputs(M, 4096, 4096, 128, seed=M)
    dx, ds, dqz = to_dev(x), to_dev(s), to_dev(qz)
    r = pkg.ops.quick_repack(to_dev(qw))
    pkg.ops.ensure_workspace(dx.device)
    eager = pkg.ops.quick_gemm(dx, r, ds, dqz, 8)
    torch.cuda.synchronize()
    graphs, outs = [], []
    for _ in range(2):
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            outs.append(pkg.ops.quick_gemm(dx, r, ds, dqz, 8))
        graphs.append(gr)
    graphs[1].replay()
    torch.cuda.synchronize()
    assert torch.equal(outs[1], eager)
    graphs[0].replay(); graphs[1].replay(); graphs[0].replay()
    torch.cuda.synchronize()
    assert torch.equal(outs[0], eager) and torch.equal(outs[1], eager)


def test_gemm_exactness_on_integer_inputs(pkg, oracle, cuda_lib):
    # small integers: every product and partial sum is exact in fp32 and in fp16 output,
    # so the GEMM must equal the oracle bit for bit whatever the accumulation order
    rng = np.random.default_rng(0)
    M, K, N, G = 16, 1024, 256, 128
    x = rng.integers(-2, 3, size=(M, K)).astype(np.float16)
    q = rng.integers(0, 16, size=(K, N), dtype=np.uint8)
    z = rng.integers(0, 16, size=(K // G, N), dtype=np.uint8)
    s = np.ones((K // G, N), dtype=np.float16)
    qw, qz = oracle.pack(q), oracle.pack(z)
    Yf, Yh = oracle.gemm_f32(x, oracle.dequantize(qw, s, qz))
    y = pkg.ops.quick_gemm(to_dev(x), pkg.ops.quick_repack(to_dev(qw)), to_dev(s), to_dev(qz), 8)
    assert np.array_equal(y.cpu().numpy().astype(np.float32), Yh.astype(np.float32))


def test_gemm_linearity_at_full_size(pkg, cuda_lib):
    # size-independent property at a BASELINE full size (13B gate_up, M = 512):
    # y(x1 + x2) == y(x1) + y(x2) within fp16 rounding, and y(0) == 0
    M, K, N, G = 512, 5120, 27648, 128
    g = torch.Generator(device="cuda").manual_seed(0)
    qw = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device="cuda", generator=g)
    qz = torch.randint(-2**31, 2**31 - 1, (K // G, N // 8), dtype=torch.int32, device="cuda", generator=g)
    s = (torch.rand((K // G, N), device="cuda", generator=g) * 0.019 + 0.001).half()
    x1 = torch.randn((M, K), device="cuda", generator=g).half()
    x2 = torch.randn((M, K), device="cuda", generator=g).half()
    r = pkg.ops.quick_repack(qw)
    y1 = pkg.ops.quick_gemm(x1, r, s, qz, 8).float()
    y2 = pkg.ops.quick_gemm(x2, r, s, qz, 8).float()
    y12 = pkg.ops.quick_gemm((x1.float() + x2.float()).half(), r, s, qz, 8).float()
    assert torch.count_nonzero(pkg.ops.quick_gemm(torch.zeros_like(x1), r, s, qz, 8)) == 0
    scale = y12.abs().max().item()
    assert (y12 - (y1 + y2)).abs().max().item() <= 2e-2 * max(scale, 1.0)
    # spot-check 64 columns against dequantize + fp32 matmul
    W = pkg.ops.quick_dequantize(r, s, qz)
    ref = x1.float() @ W[:, :64].float()
    assert (y1[:, :64] - ref).abs().max().item() <= TOL * max(1.0, ref.abs().max().item())


def test_empty_batch(pkg, oracle, cuda_lib):
    _, qw, s, qz = oracle.make_inputs(1, 128, 128, 128, seed=0)
    x = torch.empty((0, 128), dtype=torch.float16, device="cuda")
    y = pkg.ops.quick_gemm(x, pkg.ops.quick_repack(to_dev(qw)), to_dev(s), to_dev(qz), 8)
    assert y.shape == (0, 128)


def test_bad_shapes_raise_value_error(pkg, cuda_lib):
    x = torch.zeros((4, 96), dtype=torch.float16, device="cuda")
    qw = torch.zeros((96, 16), dtype=torch.int32, device="cuda")
    s = torch.zeros((3, 128), dtype=torch.float16, device="cuda")
    qz = torch.zeros((3, 16), dtype=torch.int32, device="cuda")
    with pytest.raises(ValueError):
        pkg.ops.awq_gemm(x, qw, s, qz, 8)  # K % 128 != 0


def test_linear_method_apply_weights(pkg, oracle, cuda_lib):
    M, K, N, G = 5, 512, 384, 128
    x, qw, s, qz = oracle.make_inputs(M * 2, K, N, G, seed=21)
    Yf, _ = oracle.gemm_f32(x, oracle.dequantize(qw, s, qz))
    for name in ("awq", "quick"):
        method = pkg.get_quantization_config(name)(4, G, True).get_linear_method()
        w = method.create_weights(K, N, K, N, torch.float16)
        w = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in w.items()}
        w["qweight"].data.copy_(to_dev(qw)); w["qzeros"].data.copy_(to_dev(qz)); w["scales"].data.copy_(to_dev(s))
        bias = torch.linspace(-1, 1, N, device="cuda").half()
        x3 = to_dev(x).reshape(2, M, K)
        y = method.apply_weights(w, x3, bias)
        y_again = method.apply_weights(w, x3, None)   # second call: already interleaved
        assert y.shape == (2, M, N)
        ea, er = gemm_err(y.reshape(-1, N) - bias, Yf)
        assert er <= 2 * TOL, (ea, er)
        ea, er = gemm_err(y_again.reshape(-1, N), Yf)
        assert er <= TOL, (ea, er)


def test_cuda_graph_capture(pkg, oracle, cuda_lib):
    x, qw, s, qz = oracle.make_inputs(16, 1024, 512, 128, seed=31)
    dx, ds, dqz = to_dev(x), to_dev(s), to_dev(qz)
    r = pkg.ops.quick_repack(to_dev(qw))
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        eager = pkg.ops.quick_gemm(dx, r, ds, dqz, 8)
        st.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=st):
            out = pkg.ops.quick_gemm(dx, r, ds, dqz, 8)
            out2 = pkg.ops.quick_gemm(out[:, :512].contiguous()[:, :1024 // 2].repeat(1, 2), r, ds, dqz, 8)
        graph.replay(); graph.replay()
        st.synchronize()
    assert torch.equal(out, eager)
    assert torch.isfinite(out2).all()


def test_cuda_graph_capture_of_cta_pair_launch(pkg, oracle, cuda_lib):
    """The 128-token kernel on CTA pairs is a cluster launch (cudaLaunchAttributeClusterDimension +
    programmatic stream serialization): it must capture into a CUDA graph and replay like the others."""
    M, K, N = 300, 5120, 37 * 128
    assert cuda_lib.b200_w4a16_gemm_cta_pair(M, K, N, 128) >= 1
    x, qw, s, qz, cols, Yf = sampled_case(oracle, M, K, N, 128, seed=77, n_words=8, edge_words=2)
    r = pkg.ops.quick_repack(qw)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        eager = pkg.ops.quick_gemm(x, r, s, qz, 8)
        st.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=st):
            out = pkg.ops.quick_gemm(x, r, s, qz, 8)
            out2 = pkg.ops.quick_gemm(x, r, s, qz, 8)          # back to back: programmatic dependent launch
        for _ in range(3):
            graph.replay()
        st.synchronize()
    assert torch.equal(out, eager) and torch.equal(out2, eager)
    gemm_check(out[:, cols], Yf, TOL, (M, K, N))


# ---- the exact shapes the driver-run benchmarks use (stream-K segment boundaries depend on
# (channel tiles, K blocks, 148): every new shape is a new split pattern)
LLAMA70B_TP = {  # per-rank (K, N) of qkv / o / gate_up / down, vLLM sharding (linear.py:312-552)
    1: [(8192, 10240), (8192, 8192), (8192, 57344), (28672, 8192)],
    2: [(8192, 5120), (4096, 8192), (8192, 28672), (14336, 8192)],
    4: [(8192, 2560), (2048, 8192), (8192, 14336), (7168, 8192)],
    8: [(8192, 1280), (1024, 8192), (8192, 7168), (3584, 8192)],
}


@pytest.mark.parametrize("tp", [1, 2, 4, 8])
@pytest.mark.parametrize("layer", [0, 1, 2, 3])
@pytest.mark.parametrize("M", [1, 16, 40])
def test_gemm_llama70b_tp_shards(pkg, oracle, cuda_lib, tp, layer, M):
    K, N = LLAMA70B_TP[tp][layer]
    x, qw, s, qz, cols, Yf = sampled_case(oracle, M, K, N, 128, seed=tp * 100 + layer * 10 + M)
    y = pkg.ops.quick_gemm(x, pkg.ops.quick_repack(qw), s, qz, 8)
    gemm_check(y[:, cols], Yf, TOL, (tp, layer, M))


@pytest.mark.parametrize("K,N", [(5120, 15360), (5120, 5120), (5120, 27648), (13824, 5120)])
@pytest.mark.parametrize("M", [1024, 8192])
def test_gemm_llama13b_prefill_shapes(pkg, oracle, cuda_lib, M, K, N):
    # BASELINE configs[2]: every token row, a sample of output channels, against the oracle
    x, qw, s, qz, cols, Yf = sampled_case(oracle, M, K, N, 128, seed=M + N, n_words=8)
    y = pkg.ops.quick_gemm(x, pkg.ops.quick_repack(qw), s, qz, 8)
    gemm_check(y[:, cols], Yf, TOL, (M, K, N))


@pytest.mark.parametrize("K,N", [(4096, 14336), (14336, 4096), (4096, 28672), (4096, 768), (512, 4096)])
@pytest.mark.parametrize("M", [3, 16, 64, 333])
def test_gemm_mixtral_shapes(pkg, oracle, cuda_lib, M, K, N):
    # BASELINE configs[4]: expert w1 / w3 (and their merge), w2, attention shards at TP = 8
    x, qw, s, qz, cols, Yf = sampled_case(oracle, M, K, N, 128, seed=M + K + N, n_words=16)
    y = pkg.ops.quick_gemm(x, pkg.ops.quick_repack(qw), s, qz, 8)
    gemm_check(y[:, cols], Yf, TOL, (M, K, N))


@pytest.mark.parametrize("M", [17, 24, 32, 33, 48, 64])
@pytest.mark.parametrize("K,N", [(4096, 4096), (4096, 22016), (11008, 4096), (1024, 8192), (256, 128), (4096, 192)])
def test_gemm_mid_batches(pkg, oracle, cuda_lib, M, K, N):
    """17..64 tokens: the tcgen05 kernel with 32- / 64-token tiles."""
    assert cuda_lib.b200_w4a16_gemm_variant(M, K, N, 128) == 0
    x, qw, s, qz, cols, Yf = sampled_case(oracle, M, K, N, 128, seed=M * 7 + N, n_words=min(16, N // 8))
    y = pkg.ops.quick_gemm(x, pkg.ops.quick_repack(qw), s, qz, 8)
    gemm_check(y[:, cols], Yf, TOL, (M, K, N))
    assert torch.equal(y, pkg.ops.quick_gemm(x, pkg.ops.quick_repack(qw), s, qz, 8))   # deterministic
    xg = torch.zeros_like(x)
    assert not bool(pkg.ops.quick_gemm(xg, pkg.ops.quick_repack(qw), s, qz, 8).any())   # y(0) == 0


@pytest.mark.parametrize("M,K,N,G", [
    (140, 8192, 19 * 128, 128),       # odd channel-tile count with pairs: the phantom tile is dropped
    (300, 5120, 37 * 128, 128),
    (130, 16384, 11 * 128 + 64, 128), # ragged last tile
    (520, 2048, 4096, 64),            # G < 128: two (scale, zero) rows per weight tile
    (200, 8192, 2560, 32),            # G = 32: four rows per tile
    (4100, 8192, 2560, 128),          # token-tile pairs with an odd token-tile count (33): phantom token tile
    (4096, 5120, 64 * 128, 128),      # token-tile pairs, wide layer
])
def test_gemm_channel_tile_pairs(pkg, oracle, cuda_lib, M, K, N, G):
    """Work units of two channel tiles per CTA (one activation tile feeds both) and of four per
    CTA pair (cta_group::2, the activation tile split between two SMs) on shapes where the
    grouping has an edge -- tile counts that are not multiples of 2 / 4, ragged N, G < 128: every
    token row, a sample of channels that always includes the first and the last ones."""
    assert cuda_lib.b200_w4a16_gemm_unit_tiles(M, K, N, G) in (2, 4)    # pairs within a CTA / CTA pairs
    x, qw, s, qz, cols, Yf = sampled_case(oracle, M, K, N, G, seed=M + N, n_words=12, edge_words=4)
    r = pkg.ops.quick_repack(qw)
    y = pkg.ops.quick_gemm(x, r, s, qz, 8)
    gemm_check(y[:, cols], Yf, TOL, (M, K, N, G))
    assert torch.isfinite(y).all()
    assert torch.equal(y, pkg.ops.quick_gemm(x, r, s, qz, 8))   # deterministic
