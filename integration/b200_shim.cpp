// integration/b200_shim.cpp -- the reference-side binding of INTEGRATION.md section A as a file that
// builds: the reference's C++ op names (csrc/ops.h:69-82, pybind.cpp:53-54) over libb200_w4a16.so.
// integration/build_shim.sh compiles it against this image's torch; tests/test_zx_integration_shim.py
// calls the reference's op signatures through it on a B200 and compares with the package's own ops.
// (The code between the markers is the snippet of INTEGRATION.md, verbatim; tests/test_plugin.py
// checks that the two stay identical.)
#include <map>
#include <stdexcept>
#include <string>
#include <cuda_runtime.h>
// ---- begin
// csrc/quantization/awq/b200_shim.cpp  (replaces gemm_kernels.cu; same symbols as csrc/ops.h:69-82)
#include <torch/extension.h>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include "b200_w4a16.h"

static torch::Tensor& workspace(const torch::Tensor& like, size_t bytes) {
  // One scratch per device, put into its idle state (all 0xFF: self-flagging split-K partials)
  // by b200_w4a16_workspace_init; the kernels hand it back in that state.  It must exist BEFORE
  // any CUDA-graph capture: the reference captures ~35 graphs on one capture stream
  // (worker/model_runner.py:781) and a scratch created inside capture #1 would be initialised
  // by a node of that graph only.  Create it at load time (e.g. from create_weights).
  static auto& ws = *new std::map<int, torch::Tensor>();   // never destroyed: no CUDA free at process exit
  const int dev = (int)like.get_device();
  auto it = ws.find(dev);
  if (it == ws.end() || (size_t)it->second.numel() < bytes) {
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(at::cuda::getCurrentCUDAStream().stream(), &cap);
    TORCH_CHECK(cap == cudaStreamCaptureStatusNone, "b200 workspace must be created before graph capture");
    auto t = torch::empty({(int64_t)std::max<size_t>(bytes, 148ull * 2 * 2 * 128 * 128 * 4)},
                          like.options().dtype(torch::kUInt8));
    b200_w4a16_workspace_init(t.data_ptr(), t.numel(), at::cuda::getCurrentCUDAStream().stream());
    it = ws.insert_or_assign(dev, t).first;
  }
  return it->second;
}

static void check(int st, const char* what) {
  if (st == B200_OK) return;
  if (st == B200_ERR_BAD_SHAPE) throw std::invalid_argument(std::string(what) + ": " + b200_w4a16_strerror(st));
  throw std::runtime_error(std::string(what) + ": " + b200_w4a16_strerror(st));
}

torch::Tensor awq_gemm(torch::Tensor in_feats, torch::Tensor kernel, torch::Tensor scaling_factors,
                       torch::Tensor zeros, int split_k_iters) {
  const at::cuda::OptionalCUDAGuard guard(device_of(in_feats));          // gemm_kernels.cu:398
  const int M = in_feats.size(0), K = in_feats.size(1), N = scaling_factors.size(1);
  const int G = K / scaling_factors.size(0);                             // gemm_kernels.cu:343
  auto out = torch::empty({M, N}, in_feats.options());                   // gemm_kernels.cu:401
  auto& ws = workspace(in_feats, b200_w4a16_workspace_bytes(M, K, N, B200_LAYOUT_AWQ));
  check(b200_w4a16_gemm((const uint16_t*)in_feats.data_ptr(), (const int32_t*)kernel.data_ptr(),
                        (const uint16_t*)scaling_factors.data_ptr(), (const int32_t*)zeros.data_ptr(),
                        (uint16_t*)out.data_ptr(), M, K, N, G, B200_LAYOUT_AWQ, split_k_iters,
                        ws.data_ptr(), ws.numel(), at::cuda::getCurrentCUDAStream().stream()),
        "awq_gemm");
  return out;
}

torch::Tensor awq_dequantize(torch::Tensor kernel, torch::Tensor scaling_factors, torch::Tensor zeros,
                             int /*split_k_iters*/, int /*thx*/, int /*thy*/) {
  const at::cuda::OptionalCUDAGuard guard(device_of(kernel));
  const int K = kernel.size(0), N = scaling_factors.size(1), G = K / scaling_factors.size(0);
  auto out = torch::empty({K, N}, scaling_factors.options());            // gemm_kernels.cu:366
  check(b200_w4a16_dequant((const int32_t*)kernel.data_ptr(), (const uint16_t*)scaling_factors.data_ptr(),
                           (const int32_t*)zeros.data_ptr(), (uint16_t*)out.data_ptr(), K, N, G,
                           B200_LAYOUT_AWQ, at::cuda::getCurrentCUDAStream().stream()),
        "awq_dequantize");
  return out;
}
// pybind.cpp:53-54 keeps:  ops.def("awq_gemm", &awq_gemm, ...); ops.def("awq_dequantize", &awq_dequantize, ...);
// and gains:               ops.def("quick_gemm", ...B200_LAYOUT_B200...); ops.def("quick_repack", ...b200_w4a16_repack...);
// ---- end

torch::Tensor quick_gemm(torch::Tensor in_feats, torch::Tensor kernel, torch::Tensor scaling_factors,
                         torch::Tensor zeros, int split_k_iters) {
  const at::cuda::OptionalCUDAGuard guard(device_of(in_feats));
  const int M = in_feats.size(0), K = in_feats.size(1), N = scaling_factors.size(1);
  const int G = K / scaling_factors.size(0);
  auto out = torch::empty({M, N}, in_feats.options());
  auto& ws = workspace(in_feats, b200_w4a16_workspace_bytes(M, K, N, B200_LAYOUT_B200));
  check(b200_w4a16_gemm((const uint16_t*)in_feats.data_ptr(), (const int32_t*)kernel.data_ptr(),
                        (const uint16_t*)scaling_factors.data_ptr(), (const int32_t*)zeros.data_ptr(),
                        (uint16_t*)out.data_ptr(), M, K, N, G, B200_LAYOUT_B200, split_k_iters,
                        ws.data_ptr(), ws.numel(), at::cuda::getCurrentCUDAStream().stream()),
        "quick_gemm");
  return out;
}

torch::Tensor quick_repack(torch::Tensor kernel) {
  const at::cuda::OptionalCUDAGuard guard(device_of(kernel));
  const int K = kernel.size(0), N = kernel.size(1) * 8;
  auto out = torch::empty({(int64_t)b200_w4a16_repacked_words(K, N)}, kernel.options());
  check(b200_w4a16_repack((const int32_t*)kernel.data_ptr(), (int32_t*)out.data_ptr(), K, N,
                          B200_REPACK_AWQ_TO_B200, at::cuda::getCurrentCUDAStream().stream()),
        "quick_repack");
  return out;
}

PYBIND11_MODULE(TORCH_EXTENSION_NAME, ops) {
  ops.def("awq_gemm", &awq_gemm, "Quantized GEMM for AWQ");              // pybind.cpp:53
  ops.def("awq_dequantize", &awq_dequantize, "Dequantization for AWQ");  // pybind.cpp:54
  ops.def("quick_gemm", &quick_gemm, "W4A16 GEMM on the pre-interleaved (B200) weight layout");
  ops.def("quick_repack", &quick_repack, "AWQ checkpoint layout -> B200 layout, once per weight");
}
