#!/usr/bin/env python
"""Generate tests/golden/*.npz from the REFERENCE'S OWN kernels (run on a B200 box).

The reference has no golden vectors for the AWQ path and its kernels have no
CPU implementation, so the pin is produced here: oracle/build_ref.sh compiles
csrc/quantization/awq/gemm_kernels.cu unmodified for sm_100a into oracle/_ref/ (in the
development container, where /root/reference is mounted); this script loads that module on the
GPU box, runs ``awq_dequantize`` and ``awq_gemm`` on the seeded inputs of
``oracle.make_inputs`` and stores inputs' seeds + outputs.  Small cases are stored whole;
full-size cases as SHA-256 of the output bytes plus a strided sample.

  python tests/golden/make_golden.py gpurun_out/golden      # then copy *.npz to tests/golden/
"""
import glob
import hashlib
import importlib.util
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle  # noqa: E402

SMALL = [  # (M, K, N, G, seed)
    (16, 256, 128, 128, 0), (1, 512, 128, 128, 1), (7, 512, 192, 64, 2), (3, 256, 64, 32, 3),
    (16, 512, 256, 256, 4), (5, 1024, 256, 128, 5),
]
FULL = [(16, 4096, 4096, 128, 0), (1, 4096, 12288, 128, 1), (64, 11008, 4096, 128, 2)]
# RMSNorm / fused_add_rms_norm (csrc/layernorm_kernels.cu) and SiLU-and-multiply
# (csrc/activation_kernels.cu): (tokens, hidden | d, seed, input scale; None = the reference
# test's 1 / (2 hidden)).  Small cases whole, large ones as SHA-256 + strided sample.
NORM = [(1, 4096, 0, None), (7, 768, 1, None), (16, 8192, 2, 1.0), (83, 5120, 3, 1.0),
        (16, 4096, 4, 4.0), (256, 8192, 5, 1.0)]
NORM_EPS = 1e-6
ACT = [(1, 11008, 0, 1.0), (7, 512, 1, 1.0), (16, 14336, 2, 3.0), (83, 13824, 3, 1.0)]


def load_ref():
    paths = glob.glob(os.path.join(ROOT, "oracle", "_ref", "ref_awq*.so"))
    if not paths:
        raise SystemExit("oracle/_ref/ref_awq*.so is missing: run oracle/build_ref.sh where /root/reference is mounted")
    spec = importlib.util.spec_from_file_location("ref_awq", paths[0])
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def dev(a):
    if a.dtype == np.float16:
        return torch.from_numpy(a.view(np.int16).copy()).view(torch.float16).cuda()
    return torch.from_numpy(a).cuda()


def bits(t):
    return t.contiguous().cpu().view(torch.int16).numpy().view(np.uint16)


def put(dst, key, a):
    """Small outputs whole; large ones as SHA-256 of the bytes + a strided sample."""
    if a.size <= 65536:
        dst[key] = a
    else:
        dst[key + "_sha256"] = np.frombuffer(hashlib.sha256(a.tobytes()).digest(), dtype=np.uint8)
        dst[key + "_sample"] = a[::3, ::37].copy()


def main():
    out_dir = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(out_dir, exist_ok=True)
    ref = load_ref()
    small = {}
    for i, (M, K, N, G, seed) in enumerate(SMALL):
        x, qw, s, qz = oracle.make_inputs(M, K, N, G, seed=seed)
        W = ref.awq_dequantize(dev(qw), dev(s), dev(qz), 0, 0, 0)
        small[f"case{i}_shape"] = np.array([M, K, N, G, seed])
        small[f"case{i}_W"] = bits(W)
        if N % 64 == 0 and G % 32 == 0 and N % G == 0:      # the reference GEMM's own preconditions
            Y = ref.awq_gemm(dev(x), dev(qw), dev(s), dev(qz), 8)
            small[f"case{i}_Y"] = bits(Y)
    np.savez_compressed(os.path.join(out_dir, "ref_small.npz"), **small)
    full = {}
    for i, (M, K, N, G, seed) in enumerate(FULL):
        x, qw, s, qz = oracle.make_inputs(M, K, N, G, seed=seed)
        W = bits(ref.awq_dequantize(dev(qw), dev(s), dev(qz), 0, 0, 0))
        Y = bits(ref.awq_gemm(dev(x), dev(qw), dev(s), dev(qz), 8))
        full[f"case{i}_shape"] = np.array([M, K, N, G, seed])
        full[f"case{i}_W_sha256"] = np.frombuffer(hashlib.sha256(W.tobytes()).digest(), dtype=np.uint8)
        full[f"case{i}_W_sample"] = W[::97, ::89].copy()
        full[f"case{i}_Y_sample"] = Y[:, ::61].copy()
    np.savez_compressed(os.path.join(out_dir, "ref_full.npz"), **full)
    aux = {}
    for i, (tokens, hidden, seed, scale) in enumerate(NORM):
        x, r, w = oracle.make_norm_inputs(tokens, hidden, seed, scale)
        out = torch.empty_like(dev(x))
        ref.rms_norm(out, dev(x), dev(w), NORM_EPS)
        xi, ri = dev(x), dev(r)
        ref.fused_add_rms_norm(xi, ri, dev(w), NORM_EPS)      # in place: xi = normed, ri = x + r
        aux[f"norm{i}_shape"] = np.array([tokens, hidden, seed, -1 if scale is None else int(scale)])
        put(aux, f"norm{i}_out", bits(out))
        put(aux, f"norm{i}_fused_out", bits(xi))
        put(aux, f"norm{i}_fused_residual", bits(ri))
    for i, (tokens, d, seed, scale) in enumerate(ACT):
        a = oracle.make_act_inputs(tokens, d, seed, scale)
        out = torch.empty((tokens, d), dtype=torch.float16, device="cuda")
        ref.silu_and_mul(out, dev(a))
        aux[f"act{i}_shape"] = np.array([tokens, d, seed, int(scale)])
        put(aux, f"act{i}_out", bits(out))
    np.savez_compressed(os.path.join(out_dir, "ref_norm_act.npz"), **aux)
    torch.cuda.synchronize()
    print("golden vectors written to", out_dir, [f for f in os.listdir(out_dir)])


if __name__ == "__main__":
    main()
