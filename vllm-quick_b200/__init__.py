"""b200-native W4A16 (QUICK / AWQ) linear path: a drop-in for
``vllm._C.ops.awq_gemm`` / ``awq_dequantize`` and ``AWQLinearMethod`` of the reference
(vllm/model_executor/layers/quantization/awq.py), built for sm_100a only.

The directory name carries a hyphen (it mirrors the reference repository name); import it
with ``importlib.import_module("vllm-quick_b200")``.
"""
from . import _lib, ops  # noqa: F401
from .linear import LinearMethodBase, set_weight_attrs  # noqa: F401
from . import parallel_linear, parallel_utils  # noqa: F401
from .parallel_linear import (ColumnParallelLinear, MergedColumnParallelLinear,  # noqa: F401
                              QKVParallelLinear, ReplicatedLinear, RowParallelLinear)
from .layernorm import RMSNorm  # noqa: F401
from .activation import LlamaMLP, SiluAndMul  # noqa: F401
from .mixtral_moe import MixtralMLP, MixtralMoE  # noqa: F401
from .quantization import (AWQConfig, AWQLinearMethod, QuantizationConfig, QuickConfig,  # noqa: F401
                           QuickLinearMethod, get_quantization_config)
from . import weight_utils  # noqa: F401
from .weight_utils import get_linear_method, get_quant_config, resolve_quantization  # noqa: F401

__version__ = "0.1.0"
