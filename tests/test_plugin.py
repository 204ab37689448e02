"""CPU tests of the host-side mirror of the reference plugin
(vllm/model_executor/layers/quantization/awq.py:12-165, base_config.py, __init__.py)."""
import pytest
import torch


def test_registry(pkg):
    assert pkg.get_quantization_config("awq") is pkg.AWQConfig
    assert pkg.get_quantization_config("quick") is pkg.QuickConfig
    with pytest.raises(ValueError):
        pkg.get_quantization_config("gptq")


@pytest.mark.parametrize("cls_name,name", [("AWQConfig", "awq"), ("QuickConfig", "quick")])
def test_config(pkg, cls_name, name):
    cls = getattr(pkg, cls_name)
    cfg = cls.from_config({"w_bit": 4, "q_group_size": 128, "zero_point": True})
    assert (cfg.weight_bits, cfg.group_size, cfg.zero_point, cfg.pack_factor) == (4, 128, True, 8)
    cfg2 = cls.from_config({"bits": 4, "group_size": 64, "zero_point": False})
    assert cfg2.group_size == 64
    assert cfg.get_name() == name
    assert cfg.get_supported_act_dtypes() == [torch.half]
    assert cls.get_config_filenames() == ["quant_config.json", "quantize_config.json"]
    assert "gelu" in cfg.get_scaled_act_names()
    assert isinstance(cfg.get_linear_method(), pkg.LinearMethodBase)
    with pytest.raises(ValueError):
        cls(8, 128, True)
    with pytest.raises(ValueError):
        cls.from_config({"bits": 4})
    assert "group_size=128" in repr(cfg)


@pytest.mark.parametrize("cls_name", ["AWQConfig", "QuickConfig"])
def test_create_weights(pkg, cls_name):
    method = getattr(pkg, cls_name)(4, 128, True).get_linear_method()
    w = method.create_weights(4096, 11008, 4096, 22016, torch.float16)
    qweight, qzeros, scales = w["qweight"], w["qzeros"], w["scales"]
    assert qweight.shape == (4096, 11008 // 8) and qweight.dtype == torch.int32
    assert qzeros.shape == (32, 11008 // 8) and qzeros.dtype == torch.int32
    assert scales.shape == (32, 11008) and scales.dtype == torch.float16
    for p in (qweight, qzeros):
        assert (p.input_dim, p.output_dim, p.packed_dim, p.pack_factor) == (0, 1, 1, 8)
        assert not p.requires_grad
    assert (scales.input_dim, scales.output_dim) == (0, 1)
    assert not hasattr(scales, "packed_dim")
    # non-tensor entries are allowed in the dict (reference linear.py:110-112)
    tensors = {k for k, v in w.items() if isinstance(v, torch.Tensor)}
    assert tensors == {"qweight", "qzeros", "scales"}
    with pytest.raises(ValueError):
        method.create_weights(4096 + 64, 4096, 8192, 4096, torch.float16)
    with pytest.raises(ValueError):
        method.create_weights(4096, 4100, 4096, 8200, torch.float16)


def test_set_weight_attrs_refuses_overwrite(pkg):
    t = torch.nn.Parameter(torch.zeros(1), requires_grad=False)
    pkg.set_weight_attrs(t, {"input_dim": 0})
    with pytest.raises(AssertionError):
        pkg.set_weight_attrs(t, {"input_dim": 1})
    pkg.set_weight_attrs(t, None)


# ---- load-time selection of the method (reference config.py:151-182, weight_utils.py:86-122)

def test_resolve_quantization(pkg):
    r = pkg.resolve_quantization
    assert r(None) is None
    assert r("AWQ") == "awq"
    assert r(None, {"quant_method": "AWQ"}) == "awq"
    assert r("awq", {"quant_method": "awq"}) == "awq"
    assert r("quick", {"quant_method": "awq"}) == "quick"       # QUICK consumes AWQ checkpoints
    with pytest.raises(ValueError, match="does not match"):
        r("awq", {"quant_method": "gptq"})
    with pytest.raises(ValueError, match="Unknown quantization method"):
        r("squeezellm")


def test_get_quant_config_from_files(pkg, tmp_path):
    import json
    with pytest.raises(ValueError, match="Cannot find the config file"):
        pkg.get_quant_config("awq", str(tmp_path))
    (tmp_path / "config.json").write_text("{}")
    (tmp_path / "quant_config.json").write_text(json.dumps({"w_bit": 4, "q_group_size": 64, "zero_point": True}))
    cfg = pkg.get_quant_config("awq", str(tmp_path))
    assert isinstance(cfg, pkg.AWQConfig) and cfg.group_size == 64
    (tmp_path / "quantize_config.json").write_text(json.dumps({"bits": 4, "group_size": 128, "zero_point": True}))
    with pytest.raises(ValueError, match="multiple config files"):
        pkg.get_quant_config("awq", str(tmp_path))
    # the HF quantization_config wins over files
    cfg = pkg.get_quant_config("quick", str(tmp_path), {"bits": 4, "group_size": 128, "zero_point": True,
                                                        "version": "quick", "quant_method": "awq"})
    assert isinstance(cfg, pkg.QuickConfig) and cfg.pre_interleaved


def test_get_linear_method(pkg, tmp_path):
    import json
    assert pkg.get_linear_method(None) is None
    hf = {"quant_method": "awq", "bits": 4, "group_size": 128, "zero_point": True}
    m = pkg.get_linear_method(None, hf_quant_config=hf, capability=100)
    assert type(m) is pkg.AWQLinearMethod
    m = pkg.get_linear_method("quick", hf_quant_config=hf, capability=100)
    assert type(m) is pkg.QuickLinearMethod
    with pytest.raises(ValueError, match="not supported for the current GPU"):
        pkg.get_linear_method("awq", hf_quant_config=hf, capability=70)
    with pytest.raises(ValueError, match="is not supported for quantization method"):
        pkg.get_linear_method("awq", hf_quant_config=hf, dtype=torch.bfloat16, capability=100)
    (tmp_path / "quantize_config.json").write_text(json.dumps({"bits": 4, "group_size": 128, "zero_point": True}))
    assert isinstance(pkg.get_linear_method("awq", str(tmp_path), capability=100), pkg.LinearMethodBase)
