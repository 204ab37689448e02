"""CPU tests of the tensor-parallel linear layers (reference behaviour:
vllm/model_executor/layers/linear.py:74-552, parallel_utils/communication_op.py:17-70).

World size 1 runs in-process; world size 2 runs two ``gloo`` ranks on 127.0.0.1.  The
layers' GEMM is the CUDA op, which does not exist here, so the forward-path tests plug in a
test-only linear method that evaluates the same packed weights with the CPU oracle -- what is
under test is the sharding, the loaders and the collectives, not the kernel."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

G = 128


def _mods():
    pkg = importlib.import_module("vllm-quick_b200")
    pl = importlib.import_module("vllm-quick_b200.parallel_linear")
    pu = importlib.import_module("vllm-quick_b200.parallel_utils")
    from oracle import oracle
    return pkg, pl, pu, oracle


def _oracle_method(pkg, oracle):
    class OracleLinearMethod(pkg.AWQLinearMethod):
        """Test double: AWQ weights evaluated on the CPU oracle (fp32 accumulate)."""

        def apply_weights(self, weights, x, bias=None):
            W = oracle.dequantize(weights["qweight"].data.numpy(),
                                  weights["scales"].data.numpy().view(np.float16),
                                  weights["qzeros"].data.numpy())
            y = x.reshape(-1, x.shape[-1]).float() @ torch.from_numpy(W.astype(np.float32))
            if bias is not None:
                y = y + bias.float()
            return y.reshape(x.shape[:-1] + (W.shape[1],))

    return OracleLinearMethod(pkg.AWQConfig(4, G, True))


def _checkpoint(oracle, K, N, seed):
    _, qw, s, qz = oracle.make_inputs(1, K, N, G, seed=seed)
    return (torch.from_numpy(qw), torch.from_numpy(s.view(np.int16)).view(torch.float16),
            torch.from_numpy(qz))


def _load(layer, qw, s, qz, shard_id=None):
    args = () if shard_id is None else (shard_id,)
    layer.qweight.weight_loader(layer.qweight, qw, *args)
    layer.scales.weight_loader(layer.scales, s, *args)
    layer.qzeros.weight_loader(layer.qzeros, qz, *args)


# --------------------------------------------------------------------------- world size 1

def test_world_of_one_defaults():
    _, _, pu, _ = _mods()
    assert pu.get_tensor_model_parallel_world_size() == 1
    assert pu.get_tensor_model_parallel_rank() == 0
    x = torch.arange(6.0).reshape(2, 3)
    assert pu.tensor_model_parallel_all_reduce(x) is x
    assert pu.tensor_model_parallel_all_gather(x) is x


def test_layers_register_method_tensors_and_loader_attrs():
    pkg, pl, _, _ = _mods()
    m = pkg.QuickConfig(4, G, True).get_linear_method()
    col = pl.ColumnParallelLinear(256, 512, bias=True, linear_method=m, params_dtype=torch.float16)
    names = dict(col.named_parameters())
    assert set(names) == {"qweight", "qzeros", "scales", "bias"}
    assert col.qweight.shape == (256, 64) and col.qzeros.shape == (2, 64) and col.scales.shape == (2, 512)
    for p in (col.qweight, col.qzeros):
        assert (p.input_dim, p.output_dim, p.packed_dim, p.pack_factor) == (0, 1, 1, 8)
        assert p.weight_loader == col.weight_loader
    assert col.linear_weights["qweight_layout"] == "awq"      # non-tensor state is not a parameter
    rep = pl.ReplicatedLinear(256, 512, bias=False, linear_method=m, params_dtype=torch.float16)
    assert not hasattr(rep.qweight, "weight_loader") and rep.bias is None
    with pytest.raises(ValueError):
        pl.RowParallelLinear(256, 512, bias=True, reduce_results=False, linear_method=m)
    with pytest.raises(ValueError):
        pl.ColumnParallelLinear(256, 512, linear_method=None)
    with pytest.raises(ValueError):                            # K not a multiple of the group
        pl.RowParallelLinear(192, 512, bias=False, linear_method=m)


def test_merged_and_qkv_loaders_place_each_piece():
    pkg, pl, _, oracle = _mods()
    m = pkg.QuickConfig(4, G, True).get_linear_method()
    K = 256
    gate_up = pl.MergedColumnParallelLinear(K, [384, 384], bias=False, linear_method=m,
                                            params_dtype=torch.float16)
    g, u = _checkpoint(oracle, K, 384, 1), _checkpoint(oracle, K, 384, 2)
    _load(gate_up, *g, shard_id=0)
    _load(gate_up, *u, shard_id=1)
    assert torch.equal(gate_up.qweight.data, torch.cat([g[0], u[0]], 1))
    assert torch.equal(gate_up.scales.data, torch.cat([g[1], u[1]], 1))
    assert torch.equal(gate_up.qzeros.data, torch.cat([g[2], u[2]], 1))
    fused = pl.MergedColumnParallelLinear(K, [384, 384], bias=False, linear_method=m,
                                          params_dtype=torch.float16)
    _load(fused, torch.cat([g[0], u[0]], 1), torch.cat([g[1], u[1]], 1), torch.cat([g[2], u[2]], 1))
    assert torch.equal(fused.qweight.data, gate_up.qweight.data)
    assert torch.equal(fused.scales.data, gate_up.scales.data)

    qkv = pl.QKVParallelLinear(K, 64, 4, 2, bias=False, linear_method=m, params_dtype=torch.float16)
    assert qkv.output_size == (4 + 2 * 2) * 64
    q, k, v = _checkpoint(oracle, K, 256, 3), _checkpoint(oracle, K, 128, 4), _checkpoint(oracle, K, 128, 5)
    for sid, t in (("q", q), ("k", k), ("v", v)):
        _load(qkv, *t, shard_id=sid)
    assert torch.equal(qkv.qweight.data, torch.cat([q[0], k[0], v[0]], 1))
    assert torch.equal(qkv.scales.data, torch.cat([q[1], k[1], v[1]], 1))
    with pytest.raises(AssertionError):
        qkv.weight_loader(qkv.qweight, q[0], "x")
    with pytest.raises(AssertionError):
        qkv.weight_loader(qkv.qweight, k[0], "q")            # wrong piece size


# --------------------------------------------------------------------------- world size 2

def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _tiles(oracle, qw):
    """AWQ qweight [K, N/8] (torch) -> pre-interleaved checkpoint tensor [N/128, K/128, 2048]."""
    K, N = qw.shape[0], qw.shape[1] * 8
    return torch.from_numpy(oracle.np_b200_repack(qw.numpy()).reshape(N // 128, K // 128, 2048).copy())


def _pre_interleaved_checks(pkg, pl, oracle, rank, world):
    """a ``"version": "quick"`` checkpoint (qweight already in the B200 layout,
    [N/128, K/128, 2048]) loads shard by shard through the same loaders; what every rank ends
    up with is, word for word, the interleave of its own AWQ shard."""
    cfg = pkg.QuickConfig.from_config({"w_bit": 4, "q_group_size": G, "zero_point": True,
                                       "version": "QUICK"})
    assert cfg.pre_interleaved and not pkg.QuickConfig(4, G, True).pre_interleaved
    with pytest.raises(ValueError):
        pkg.QuickConfig(4, G, True, "gemv")
    method = cfg.get_linear_method()
    K, N = 512, 512
    qw, s, qz = _checkpoint(oracle, K, N, 21)
    nh, kh = N // world, K // world
    col = pl.ColumnParallelLinear(K, N, bias=False, linear_method=method, params_dtype=torch.float16)
    assert tuple(col.qweight.shape) == (nh // 128, K // 128, 2048)
    assert col.linear_weights["qweight_layout"] == "b200"
    _load(col, _tiles(oracle, qw), s, qz)
    assert torch.equal(col.qweight.data, _tiles(oracle, qw[:, rank * nh // 8:(rank + 1) * nh // 8].contiguous()))
    assert torch.equal(col.scales.data, s[:, rank * nh:(rank + 1) * nh])
    row = pl.RowParallelLinear(K, N, bias=False, linear_method=method, params_dtype=torch.float16)
    _load(row, _tiles(oracle, qw), s, qz)
    assert torch.equal(row.qweight.data, _tiles(oracle, qw[rank * kh:(rank + 1) * kh].contiguous()))
    assert torch.equal(row.qzeros.data, qz[rank * kh // G:(rank + 1) * kh // G])
    # merged gate_up from two checkpoint tensors, and from one already-fused tensor
    a, c = _checkpoint(oracle, K, 256, 22), _checkpoint(oracle, K, 256, 23)
    want = _tiles(oracle, torch.cat([a[0][:, rank * 128 // 8 * (2 // world):][:, :256 // world // 8],
                                     c[0][:, rank * 128 // 8 * (2 // world):][:, :256 // world // 8]], 1).contiguous())
    gu = pl.MergedColumnParallelLinear(K, [256, 256], bias=False, linear_method=method, params_dtype=torch.float16)
    for sid, t in ((0, a), (1, c)):
        _load(gu, _tiles(oracle, t[0]), t[1], t[2], shard_id=sid)
    assert torch.equal(gu.qweight.data, want)
    gu2 = pl.MergedColumnParallelLinear(K, [256, 256], bias=False, linear_method=method, params_dtype=torch.float16)
    _load(gu2, _tiles(oracle, torch.cat([a[0], c[0]], 1).contiguous()), torch.cat([a[1], c[1]], 1),
          torch.cat([a[2], c[2]], 1))
    assert torch.equal(gu2.qweight.data, want) and torch.equal(gu2.scales.data, gu.scales.data)
    # QKV, kv heads replicated when tp >= kv heads; head size 128 keeps pieces on tile borders
    qkv = pl.QKVParallelLinear(K, 128, 2, 1, bias=False, linear_method=method, params_dtype=torch.float16)
    qt, kt, vt = _checkpoint(oracle, K, 256, 24), _checkpoint(oracle, K, 128, 25), _checkpoint(oracle, K, 128, 26)
    for sid, t in (("q", qt), ("k", kt), ("v", vt)):
        _load(qkv, _tiles(oracle, t[0]), t[1], t[2], shard_id=sid)
    qn = 256 // world
    want = _tiles(oracle, torch.cat([qt[0][:, rank * qn // 8:(rank + 1) * qn // 8] if world > 1 else qt[0],
                                     kt[0], vt[0]], 1).contiguous())
    assert torch.equal(qkv.qweight.data, want)
    # pieces that do not fall on tile borders are refused, as is a misaligned partition
    bad = pl.QKVParallelLinear(K, 64, 4, 1, bias=False, linear_method=method, params_dtype=torch.float16) \
        if world == 1 else None
    with pytest.raises((ValueError, AssertionError)):
        if bad is None:
            pl.QKVParallelLinear(K, 64, 2, 2, bias=False, linear_method=method, params_dtype=torch.float16)
        else:
            bad.scales.weight_loader(bad.scales, _checkpoint(oracle, K, 64, 27)[1], "k")     # fine: per channel
            bad.qweight.weight_loader(bad.qweight, _tiles(oracle, _checkpoint(oracle, K, 128, 27)[0]), "k")
    with pytest.raises(ValueError):
        pl.ColumnParallelLinear(K, 192 * world, bias=False, linear_method=method, params_dtype=torch.float16)


def test_pre_interleaved_checkpoint_world_of_one():
    pkg, pl, pu, oracle = _mods()
    _pre_interleaved_checks(pkg, pl, oracle, 0, 1)


def _moe_checks(pkg, oracle, method, rank, world):
    """Mixture-of-experts block across ranks (mixtral_quant.py:113-118, :159): contiguous expert
    blocks per rank, parameters for the local experts only, partial sums all-reduced."""
    import types
    import torch.nn.functional as F
    hidden, ffn, experts, top_k = 256, 256, 8, 2
    cfg = types.SimpleNamespace(num_local_experts=experts, num_experts_per_tok=top_k,
                                hidden_size=hidden, intermediate_size=ffn)
    torch.set_default_dtype(torch.float16)
    try:
        moe = pkg.MixtralMoE(cfg, method)
    finally:
        torch.set_default_dtype(torch.float32)
    per = experts // world
    assert moe.expert_indicies == list(range(rank * per, (rank + 1) * per))
    owned = {int(n.split(".")[1]) for n, _ in moe.named_parameters() if n.startswith("experts.")}
    assert owned == set(moe.expert_indicies)
    dense = {}
    for e in range(experts):                                   # same checkpoint on every rank
        for j, name in enumerate(("w1", "w2", "w3")):
            K, N = (ffn, hidden) if name == "w2" else (hidden, ffn)
            _, qw, s, qz = oracle.make_inputs(1, K, N, G, seed=100 + 10 * e + j)
            dense[(e, name)] = torch.from_numpy(oracle.dequantize(qw, s, qz).astype(np.float32))
            if moe.experts[e] is not None:
                lin = getattr(moe.experts[e], name)
                lin.qweight.data.copy_(torch.from_numpy(qw))
                lin.scales.data.copy_(torch.from_numpy(s.view(np.int16)).view(torch.float16))
                lin.qzeros.data.copy_(torch.from_numpy(qz))
    g = torch.Generator().manual_seed(5)
    moe.gate.weight.data = torch.randn(experts, hidden, generator=g) * 0.5
    x = torch.randn(1, 6, hidden, generator=g)
    got = moe(x).reshape(6, hidden)                            # all-reduced inside
    xf = x.reshape(6, hidden)
    w = F.softmax(xf @ moe.gate.weight.data.t(), dim=1, dtype=torch.float)
    w, sel = torch.topk(w, top_k, dim=-1)
    w = w / w.sum(dim=-1, keepdim=True)
    want = torch.zeros(6, hidden)
    for e in range(experts):
        ew = (w * (sel == e)).sum(dim=-1, keepdim=True)
        want += ((F.silu(xf @ dense[(e, "w1")]) * (xf @ dense[(e, "w3")])) @ dense[(e, "w2")]) * ew
    assert torch.allclose(got.float(), want, rtol=1e-4, atol=1e-4)
    with pytest.raises(ValueError):                            # more ranks than experts
        pkg.MixtralMoE(types.SimpleNamespace(num_local_experts=1, num_experts_per_tok=1, hidden_size=hidden,
                                             intermediate_size=ffn), method)


def _worker(rank, world, port, q):
    try:
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        pkg, pl, pu, oracle = _mods()
        with pytest.raises(RuntimeError):
            pu.initialize_model_parallel(3)
        pu.initialize_model_parallel(world)
        assert pu.get_tensor_model_parallel_world_size() == world
        assert pu.get_tensor_model_parallel_rank() == rank
        method = _oracle_method(pkg, oracle)
        torch.manual_seed(0)
        K, N, M = 512, 256, 5
        x = torch.randn(M, K).half()

        # collectives
        t = torch.full((2, 3), float(rank + 1))
        assert torch.equal(pu.tensor_model_parallel_all_reduce(t), torch.full((2, 3), 3.0))
        gt = pu.tensor_model_parallel_all_gather(torch.full((2, 3), float(rank)), dim=-1)
        assert gt.shape == (2, 6) and torch.equal(gt[:, :3], torch.zeros(2, 3)) and torch.equal(gt[:, 3:], torch.ones(2, 3))
        g0 = pu.tensor_model_parallel_all_gather(torch.full((2, 3), float(rank)), dim=0)
        assert g0.shape == (4, 3) and torch.equal(g0[2:], torch.ones(2, 3))

        # column parallel: each rank holds its N/2 columns; gather_output restores the full product
        qw, s, qz = _checkpoint(oracle, K, N, 7)
        W = torch.from_numpy(oracle.dequantize(qw.numpy(), s.numpy().view(np.float16), qz.numpy()).astype(np.float32))
        full = x.float() @ W
        col = pl.ColumnParallelLinear(K, N, bias=False, gather_output=True, linear_method=method,
                                      params_dtype=torch.float16)
        _load(col, qw, s, qz)
        half = N // world
        assert torch.equal(col.qweight.data, qw[:, rank * half // 8:(rank + 1) * half // 8])
        assert torch.equal(col.scales.data, s[:, rank * half:(rank + 1) * half])
        y, b = col(x)
        assert b is None and torch.allclose(y, full, atol=1e-4)

        # row parallel: K rows split, partial products all-reduced; bias added once, after
        row = pl.RowParallelLinear(K, N, bias=True, input_is_parallel=False, linear_method=method,
                                   params_dtype=torch.float16)
        _load(row, qw, s, qz)
        bias = torch.linspace(-1, 1, N).half()
        row.bias.weight_loader(row.bias, bias)
        kh = K // world
        assert torch.equal(row.qweight.data, qw[rank * kh:(rank + 1) * kh])
        assert torch.equal(row.qzeros.data, qz[rank * kh // G:(rank + 1) * kh // G])
        y, b = row(x)
        assert b is None and torch.allclose(y, full + bias.float(), atol=1e-3)
        row_nr = pl.RowParallelLinear(K, N, bias=False, input_is_parallel=True, reduce_results=False,
                                      linear_method=method, params_dtype=torch.float16)
        _load(row_nr, qw, s, qz)
        part, _ = row_nr(x[:, rank * kh:(rank + 1) * kh].contiguous())
        dist.all_reduce(part)
        assert torch.allclose(part, full, atol=1e-3)

        # QKV with ONE kv head and two ranks: q split by head, k/v replicated on both ranks
        qkv = pl.QKVParallelLinear(K, 64, 4, 1, bias=False, linear_method=method, params_dtype=torch.float16)
        assert (qkv.num_heads, qkv.num_kv_heads, qkv.num_kv_head_replicas) == (2, 1, 2)
        qt, kt, vt = _checkpoint(oracle, K, 256, 8), _checkpoint(oracle, K, 64, 9), _checkpoint(oracle, K, 64, 10)
        for sid, tns in (("q", qt), ("k", kt), ("v", vt)):
            _load(qkv, *tns, shard_id=sid)
        assert torch.equal(qkv.scales.data[:, :128], qt[1][:, rank * 128:(rank + 1) * 128])
        assert torch.equal(qkv.scales.data[:, 128:192], kt[1]) and torch.equal(qkv.scales.data[:, 192:], vt[1])
        assert torch.equal(qkv.qweight.data[:, 16:24], kt[0])

        # merged column parallel, fused checkpoint tensor: every piece sharded on its own
        gu = pl.MergedColumnParallelLinear(K, [256, 256], bias=False, linear_method=method,
                                           params_dtype=torch.float16)
        a, c = _checkpoint(oracle, K, 256, 11), _checkpoint(oracle, K, 256, 12)
        _load(gu, torch.cat([a[0], c[0]], 1), torch.cat([a[1], c[1]], 1), torch.cat([a[2], c[2]], 1))
        assert torch.equal(gu.scales.data, torch.cat([a[1][:, rank * 128:(rank + 1) * 128],
                                                      c[1][:, rank * 128:(rank + 1) * 128]], 1))
        assert torch.equal(gu.qweight.data[:, 16:], c[0][:, rank * 16:(rank + 1) * 16])

        with pytest.raises(ValueError):
            pl.RowParallelLinear(128, N, bias=False, linear_method=method)  # K/tp not a multiple of G
        _pre_interleaved_checks(pkg, pl, oracle, rank, world)
        _moe_checks(pkg, oracle, method, rank, world)
        pu.destroy_model_parallel()
        assert pu.get_tensor_model_parallel_world_size() == 1
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except BaseException as e:  # noqa: BLE001 -- report to the parent, whatever it is
        import traceback
        q.put((rank, "".join(traceback.format_exception(type(e), e, e.__traceback__))))


@pytest.mark.timeout(180)
def test_two_rank_tensor_parallel_gloo():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=150) for _ in range(world)]
    for p in procs:
        p.join(30)
    for rank, msg in results:
        assert msg == "ok", f"rank {rank}: {msg}"
