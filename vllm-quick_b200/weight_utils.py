"""Load-time selection of the quantized linear method.

Mirrors, for the two methods this package serves, what the reference does between the
engine arguments and the model constructor:

* ``resolve_quantization``   <- ``ModelConfig._verify_quantization`` (vllm/config.py:151-182)
* ``get_quant_config``       <- ``weight_utils.get_quant_config`` (vllm/model_executor/weight_utils.py:86-122)
* ``get_linear_method``      <- the three lines of ``get_model`` that turn the config into the
  ``LinearMethodBase`` handed to the model constructor (model_executor/model_loader.py:40-74)

Only local checkpoints are handled: hub download is the engine's business, not the path's.
"""
from __future__ import annotations

import glob
import json
import os
from typing import Any, Dict, Optional

import torch

from .linear import LinearMethodBase
from .quantization import QuantizationConfig, get_quantization_config

SUPPORTED_QUANTIZATION = ("awq", "quick")


def resolve_quantization(quantization: Optional[str],
                         hf_quant_config: Optional[Dict[str, Any]] = None) -> Optional[str]:
    """The method name the engine will use: the ``--quantization`` argument, lower-cased, checked
    against the ``quant_method`` a checkpoint declares in its HF ``quantization_config``.
    A ``"quick"`` request on an ``"awq"`` checkpoint is legal: QUICK consumes AWQ checkpoints and
    interleaves after loading."""
    name = quantization.lower() if quantization is not None else None
    if hf_quant_config is not None:
        declared = str(hf_quant_config["quant_method"]).lower()
        if name is None:
            name = declared
        elif name != declared and not (name == "quick" and declared == "awq"):
            raise ValueError(
                "Quantization method specified in the model config "
                f"({declared}) does not match the quantization "
                f"method specified in the `quantization` argument "
                f"({name}).")
    if name is not None and name not in SUPPORTED_QUANTIZATION:
        raise ValueError(f"Unknown quantization method: {name}. Must be one of "
                         f"{list(SUPPORTED_QUANTIZATION)}.")
    return name


def get_quant_config(quantization: str, model_dir: Optional[str] = None,
                     hf_quant_config: Optional[Dict[str, Any]] = None) -> QuantizationConfig:
    """Build the method's config object: from the HF ``quantization_config`` when the checkpoint
    carries one, else from the single ``quant_config.json`` / ``quantize_config.json`` in
    ``model_dir`` (file names: ``QuantizationConfig.get_config_filenames``)."""
    quant_cls = get_quantization_config(quantization)
    if hf_quant_config is not None:
        return quant_cls.from_config(hf_quant_config)
    if model_dir is None or not os.path.isdir(model_dir):
        raise ValueError(f"Cannot find the config file for {quantization}")
    wanted = tuple(quant_cls.get_config_filenames())
    found = sorted(f for f in glob.glob(os.path.join(model_dir, "*.json")) if f.endswith(wanted))
    if not found:
        raise ValueError(f"Cannot find the config file for {quantization}")
    if len(found) > 1:
        raise ValueError(f"Found multiple config files for {quantization}: {found}")
    with open(found[0], "r") as f:
        return quant_cls.from_config(json.load(f))


def get_linear_method(quantization: Optional[str], model_dir: Optional[str] = None,
                      hf_quant_config: Optional[Dict[str, Any]] = None,
                      dtype: torch.dtype = torch.float16,
                      capability: Optional[int] = None) -> Optional[LinearMethodBase]:
    """``None`` for an unquantized model, else the linear method to pass to the layers, after
    the reference's capability and activation-dtype checks (model_loader.py:46-66)."""
    name = resolve_quantization(quantization, hf_quant_config)
    if name is None:
        return None
    cfg = get_quant_config(name, model_dir, hf_quant_config)
    if capability is None and torch.cuda.is_available():
        major, minor = torch.cuda.get_device_capability()
        capability = major * 10 + minor
    if capability is not None and capability < cfg.get_min_capability():
        raise ValueError(
            f"The quantization method {name} is not supported for the current GPU. "
            f"Minimum capability: {cfg.get_min_capability()}. Current capability: {capability}.")
    supported = cfg.get_supported_act_dtypes()
    if dtype not in supported:
        raise ValueError(f"{dtype} is not supported for quantization method {name}. "
                         f"Supported dtypes: {supported}")
    return cfg.get_linear_method()
