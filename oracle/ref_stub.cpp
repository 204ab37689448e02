// oracle/ref_stub.cpp -- pybind11 stub that exposes the REFERENCE's own AWQ kernels
// (compiled unmodified from /root/reference/csrc/quantization/awq/gemm_kernels.cu, where the
// source lies) as a Python module.  TEST INFRASTRUCTURE ONLY: used to generate the golden
// vectors under tests/golden/ and as the same-box "kernel to beat" in profiles/.
// The two prototypes restate csrc/ops.h:69-82 of the reference.
#include <torch/extension.h>

torch::Tensor awq_gemm(torch::Tensor _in_feats, torch::Tensor _kernel,
                       torch::Tensor _scaling_factors, torch::Tensor _zeros, int split_k_iters);

torch::Tensor awq_dequantize(torch::Tensor _kernel, torch::Tensor _scaling_factors,
                             torch::Tensor _zeros, int split_k_iters, int thx, int thy);

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("awq_gemm", &awq_gemm, "reference AWQ GEMM (gemm_kernels.cu:389-446)");
  m.def("awq_dequantize", &awq_dequantize, "reference AWQ dequantize (gemm_kernels.cu:332-381)");
}
