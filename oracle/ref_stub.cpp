// oracle/ref_stub.cpp -- pybind11 stub that exposes the REFERENCE's own AWQ kernels
// (compiled unmodified from /root/reference/csrc/quantization/awq/gemm_kernels.cu, where the
// source lies) as a Python module.  TEST INFRASTRUCTURE ONLY: used to generate the golden
// vectors under tests/golden/ and as the same-box "kernel to beat" in profiles/.
// The prototypes restate csrc/ops.h:36-58 and :69-82 of the reference.
#include <torch/extension.h>

torch::Tensor awq_gemm(torch::Tensor _in_feats, torch::Tensor _kernel,
                       torch::Tensor _scaling_factors, torch::Tensor _zeros, int split_k_iters);

torch::Tensor awq_dequantize(torch::Tensor _kernel, torch::Tensor _scaling_factors,
                             torch::Tensor _zeros, int split_k_iters, int thx, int thy);

void rms_norm(torch::Tensor& out, torch::Tensor& input, torch::Tensor& weight, float epsilon);

void fused_add_rms_norm(torch::Tensor& input, torch::Tensor& residual, torch::Tensor& weight,
                        float epsilon);

void silu_and_mul(torch::Tensor& out, torch::Tensor& input);

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("awq_gemm", &awq_gemm, "reference AWQ GEMM (gemm_kernels.cu:389-446)");
  m.def("awq_dequantize", &awq_dequantize, "reference AWQ dequantize (gemm_kernels.cu:332-381)");
  m.def("rms_norm", &rms_norm, "reference RMSNorm (layernorm_kernels.cu:78-102)");
  m.def("fused_add_rms_norm", &fused_add_rms_norm,
        "reference residual add + RMSNorm, in place (layernorm_kernels.cu:104-129)");
  m.def("silu_and_mul", &silu_and_mul, "reference SiLU-and-multiply (activation_kernels.cu:30-52)");
}
