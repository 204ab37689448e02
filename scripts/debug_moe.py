#!/usr/bin/env python
"""Debug aid: one Mixtral-shaped expert at a given batch, merged/fused path against the three
separate GEMMs + torch activation, stage by stage."""
import importlib, os, sys, types
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("vllm-quick_b200")
G = 128
T = int(sys.argv[1]) if len(sys.argv) > 1 else 64
H, F_ = 4096, 14336
g = torch.Generator(device="cuda").manual_seed(0)
def rnd(K, N):
    qw = torch.randint(-2**31, 2**31 - 1, (K, N // 8), dtype=torch.int32, device="cuda", generator=g)
    qz = torch.randint(-2**31, 2**31 - 1, (K // G, N // 8), dtype=torch.int32, device="cuda", generator=g)
    s = (torch.rand((K // G, N), device="cuda", generator=g) * 0.019 + 0.001).half()
    return qw, s, qz
torch.set_default_dtype(torch.float16)
method = pkg.QuickConfig(4, G, True).get_linear_method()
with torch.device("cuda"):
    mlp = pkg.MixtralMLP(1, H, F_, method)
torch.set_default_dtype(torch.float32)
ws = {}
for name, lin, (K, N) in (("w1", mlp.w1, (H, F_)), ("w2", mlp.w2, (F_, H)), ("w3", mlp.w3, (H, F_))):
    qw, s, qz = rnd(K, N)
    lin.qweight.data.copy_(qw); lin.scales.data.copy_(s); lin.qzeros.data.copy_(qz)
    ws[name] = (pkg.ops.quick_repack(qw), s, qz)
x = (torch.randn((T, H), device="cuda", generator=g) * 0.25).half()
g1 = pkg.ops.quick_gemm(x, *ws["w1"], 8)
g3 = pkg.ops.quick_gemm(x, *ws["w3"], 8)
sep = torch.empty((T, F_), dtype=torch.float16, device="cuda")
pkg.ops.silu_and_mul(sep, torch.cat([g1, g3], dim=1).contiguous())
y_sep = pkg.ops.quick_gemm(sep, *ws["w2"], 8)
mlp.process_weights_after_loading()
act = method.apply_weights_silu_and_mul(mlp._w13, x)
torch.cuda.synchronize()
print("T", T, "act equal", bool(torch.equal(act, sep)), "max diff", float((act.float() - sep.float()).abs().max()))
y = mlp(x)
print("y equal", bool(torch.equal(y, y_sep)), "max diff", float((y.float() - y_sep.float()).abs().max()), "max |y|", float(y_sep.float().abs().max()))
# full GEMM of the merged weight vs the paired activation kernel separately
full = pkg.ops.quick_gemm(x, mlp._w13["qweight"].data, mlp._w13["scales"].data, mlp._w13["qzeros"].data, 8)
o2 = torch.empty((T, F_), dtype=torch.float16, device="cuda"); pkg.ops.silu_and_mul_paired(o2, full)
print("paired kernel on full GEMM equal", bool(torch.equal(o2, sep)), float((o2.float() - sep.float()).abs().max()))
# un-pair the full GEMM output by hand: 16-channel blocks = 8 gate + 8 up
fb = full.reshape(T, -1, 2, 8)
print("gate cols equal", bool(torch.equal(fb[:, :, 0].reshape(T, -1), g1)), "up cols equal", bool(torch.equal(fb[:, :, 1].reshape(T, -1), g3)))
