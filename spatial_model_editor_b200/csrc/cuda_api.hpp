// Thin access layer to the CUDA runtime (linked statically) and to the few
// driver-API entry points needed for NVRTC modules and TMA descriptors.  The
// driver symbols are resolved at first use through cudaGetDriverEntryPoint so
// that libsmeb200.so loads (and exports its C-ABI) on a machine without
// libcuda.so.1 — the role cuda_stubs.cpp:139-186 plays in the reference.
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <nvrtc.h>
#include <stdexcept>
#include <string>

namespace smeb200 {

struct Error : std::runtime_error {
  using std::runtime_error::runtime_error;
};

struct DriverApi {
  CUresult (*ModuleLoadData)(CUmodule *, const void *) = nullptr;
  CUresult (*ModuleUnload)(CUmodule) = nullptr;
  CUresult (*ModuleGetFunction)(CUfunction *, CUmodule, const char *) = nullptr;
  CUresult (*LaunchKernel)(CUfunction, unsigned, unsigned, unsigned, unsigned,
                           unsigned, unsigned, unsigned, CUstream, void **,
                           void **) = nullptr;
  CUresult (*LaunchKernelEx)(const CUlaunchConfig *, CUfunction, void **,
                             void **) = nullptr;
  CUresult (*FuncSetAttribute)(CUfunction, CUfunction_attribute, int) = nullptr;
  CUresult (*OccupancyMaxActiveBlocksPerMultiprocessor)(int *, CUfunction, int,
                                                        size_t) = nullptr;
  CUresult (*GetErrorString)(CUresult, const char **) = nullptr;
  CUresult (*TensorMapEncodeTiled)(
      CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *,
      const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
      const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
      CUtensorMapL2promotion, CUtensorMapFloatOOBfill) = nullptr;
  CUresult (*StreamWriteValue32)(CUstream, CUdeviceptr, cuuint32_t,
                                 unsigned int) = nullptr;
  CUresult (*StreamWaitValue32)(CUstream, CUdeviceptr, cuuint32_t,
                                unsigned int) = nullptr;
  // throws Error if the driver is not available
  static const DriverApi &get();
};

void check(cudaError_t e, const char *what);
void check(CUresult e, const char *what);
void check(nvrtcResult e, const char *what);

// reference-compatible "runtime unavailable" text (cudapixelsim_t.cpp:150-156
// skips on messages starting with these prefixes)
std::string unavailableReason();

} // namespace smeb200
