#include "cuda_api.hpp"
#include <mutex>

namespace smeb200 {

void check(cudaError_t e, const char *what) {
  if (e != cudaSuccess) {
    throw Error(std::string(what) + ": " + cudaGetErrorString(e));
  }
}

void check(CUresult e, const char *what) {
  if (e != CUDA_SUCCESS) {
    const char *s = nullptr;
    const auto &d = DriverApi::get();
    if (d.GetErrorString != nullptr) {
      d.GetErrorString(e, &s);
    }
    throw Error(std::string(what) + ": " + (s != nullptr ? s : "CUDA driver error"));
  }
}

void check(nvrtcResult e, const char *what) {
  if (e != NVRTC_SUCCESS) {
    throw Error(std::string(what) + ": " + nvrtcGetErrorString(e));
  }
}

template <typename F> static void resolve(const char *name, F &fn) {
  void *p = nullptr;
  cudaDriverEntryPointQueryResult status{};
  cudaError_t e = cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &status);
  if (e != cudaSuccess || p == nullptr ||
      status != cudaDriverEntryPointSuccess) {
    cudaGetLastError();
    throw Error(std::string("Failed to initialize the CUDA driver: missing "
                            "entry point ") +
                name);
  }
  fn = reinterpret_cast<F>(p);
}

const DriverApi &DriverApi::get() {
  static DriverApi api;
  static std::once_flag once;
  static std::string failure;
  std::call_once(once, [] {
    try {
      resolve("cuGetErrorString", api.GetErrorString);
      resolve("cuModuleLoadData", api.ModuleLoadData);
      resolve("cuModuleUnload", api.ModuleUnload);
      resolve("cuModuleGetFunction", api.ModuleGetFunction);
      resolve("cuLaunchKernel", api.LaunchKernel);
      resolve("cuLaunchKernelEx", api.LaunchKernelEx);
      resolve("cuFuncSetAttribute", api.FuncSetAttribute);
      resolve("cuOccupancyMaxActiveBlocksPerMultiprocessor",
              api.OccupancyMaxActiveBlocksPerMultiprocessor);
      resolve("cuTensorMapEncodeTiled", api.TensorMapEncodeTiled);
      resolve("cuStreamWriteValue32", api.StreamWriteValue32);
      resolve("cuStreamWaitValue32", api.StreamWaitValue32);
    } catch (const Error &e) {
      failure = e.what();
    }
  });
  if (!failure.empty()) {
    throw Error(failure);
  }
  return api;
}

std::string unavailableReason() {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e == cudaErrorInsufficientDriver || e == cudaErrorInitializationError ||
      e == cudaErrorSystemDriverMismatch ||
      e == cudaErrorCompatNotSupportedOnDevice) {
    cudaGetLastError();
    return std::string("Failed to initialize the CUDA driver: ") +
           cudaGetErrorString(e);
  }
  if (e == cudaErrorNoDevice || (e == cudaSuccess && n == 0)) {
    cudaGetLastError();
    return "No CUDA devices were found";
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    return std::string("Failed to query CUDA devices: ") +
           cudaGetErrorString(e);
  }
  return {};
}

} // namespace smeb200
