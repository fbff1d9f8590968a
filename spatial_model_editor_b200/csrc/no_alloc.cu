// smeb200_avg_min_max_no_alloc: the per-species {average, minimum, maximum}
// reduction of a compartment's state (device_kernels.cuh: smeb_stats +
// smeb_stats_finish, host side Sim::getAvgMinMax) for callers that must not
// allocate: code under CUDA graph capture, tight serving loops, harnesses that
// forbid cudaMalloc.  No counterpart in the
// reference.
//
// The block partials of the reduction live in ONE file-scope device array of a
// fixed, modest size instead of a per-simulator cudaMalloc'ed buffer.  Every GPU
// that loads the library holds the array (SMEB200_NO_ALLOC_SCRATCH_BYTES) until
// the process exits.  All calls on a device share it and it keeps the last
// call's contents: calls that overlap on the GPU (two streams, two threads) race
// and must be serialised by the caller.  Nothing in it has to start at zero:
// block b of the first kernel overwrites slot b before the finishing kernel of
// the same call reads slots [0, blocks) and nothing else.
//
// In host code `g_scratch` is a host-side stand-in for the symbol, not a device
// address, so the kernels get the address cudaGetSymbolAddress returns (once per
// device, with that device current).
#include "../../include/smeb200.h"
#include <cuda_runtime.h>
#include <cstdio>
#include <cstring>
#include <mutex>

namespace {

constexpr int BLOCK = 256;                       // SMEB_BLOCK
constexpr int MAX_SP = SMEB200_NO_ALLOC_MAX_SPECIES;
constexpr unsigned MAX_BLOCKS = SMEB200_NO_ALLOC_MAX_BLOCKS;
constexpr int MAX_DEVICES = 64;
constexpr double DMAX = 1.7976931348623157e308;

static_assert(sizeof(double) * 3 * MAX_SP * MAX_BLOCKS == SMEB200_NO_ALLOC_SCRATCH_BYTES,
              "scratch size constant out of date");

// [block][species][sum, min, max]
static __device__ __align__(256) double g_scratch[(size_t)MAX_BLOCKS * MAX_SP * 3];

// the summation order of smeb_stats<R>: per-thread grid-stride partials, xor
// shuffles inside a warp, the warps of a block in ascending order
__global__ void __launch_bounds__(BLOCK)
    stats_partial(const double *__restrict__ c, unsigned n, unsigned S, unsigned sp,
                  double *__restrict__ partial) {
  double sum[MAX_SP], mn[MAX_SP], mx[MAX_SP];
#pragma unroll
  for (int s = 0; s < MAX_SP; ++s) {
    sum[s] = 0.0;
    mn[s] = DMAX;
    mx[s] = 0.0;
  }
  for (unsigned i = blockIdx.x * BLOCK + threadIdx.x; i < n; i += gridDim.x * BLOCK) {
    const double *v = c + (size_t)i * S;
#pragma unroll
    for (int s = 0; s < MAX_SP; ++s) {
      if (s < (int)sp) {
        const double x = v[s];
        sum[s] += x;
        mn[s] = x < mn[s] ? x : mn[s];
        mx[s] = x > mx[s] ? x : mx[s];
      }
    }
  }
  __shared__ double ss[3][32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int s = 0; s < MAX_SP; ++s) {
    if (s >= (int)sp) {
      break; // uniform across the block
    }
    double a = sum[s], b = mn[s], d = mx[s];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      a += __shfl_xor_sync(0xffffffffu, a, o);
      const double b2 = __shfl_xor_sync(0xffffffffu, b, o);
      const double d2 = __shfl_xor_sync(0xffffffffu, d, o);
      b = b2 < b ? b2 : b;
      d = d2 > d ? d2 : d;
    }
    if (lane == 0) {
      ss[0][w] = a;
      ss[1][w] = b;
      ss[2][w] = d;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      double ta = 0.0, tb = DMAX, td = 0.0;
      for (int k = 0; k < BLOCK / 32; ++k) {
        ta += ss[0][k];
        tb = ss[1][k] < tb ? ss[1][k] : tb;
        td = ss[2][k] > td ? ss[2][k] : td;
      }
      double *dst = partial + ((size_t)blockIdx.x * sp + s) * 3;
      dst[0] = ta;
      dst[1] = tb;
      dst[2] = td;
    }
    __syncthreads();
  }
}

// out[3*s .. 3*s+2] = {average, min, max}: the fold of smeb_stats_finish
__global__ void __launch_bounds__(BLOCK)
    stats_finish(const double *__restrict__ partial, unsigned nBlocks, unsigned sp,
                 double nVoxels, double *__restrict__ out) {
  __shared__ double ss[3][BLOCK];
  for (unsigned s = 0; s < sp; ++s) {
    double a = 0.0, b = DMAX, d = 0.0;
    for (unsigned k = threadIdx.x; k < nBlocks; k += BLOCK) {
      const double *src = partial + ((size_t)k * sp + s) * 3;
      a += src[0];
      b = src[1] < b ? src[1] : b;
      d = src[2] > d ? src[2] : d;
    }
    ss[0][threadIdx.x] = a;
    ss[1][threadIdx.x] = b;
    ss[2][threadIdx.x] = d;
    __syncthreads();
    for (int o = BLOCK / 2; o > 0; o >>= 1) {
      if ((int)threadIdx.x < o) {
        ss[0][threadIdx.x] += ss[0][threadIdx.x + o];
        const double b2 = ss[1][threadIdx.x + o], d2 = ss[2][threadIdx.x + o];
        ss[1][threadIdx.x] = b2 < ss[1][threadIdx.x] ? b2 : ss[1][threadIdx.x];
        ss[2][threadIdx.x] = d2 > ss[2][threadIdx.x] ? d2 : ss[2][threadIdx.x];
      }
      __syncthreads();
    }
    if (threadIdx.x == 0) {
      out[3 * s] = ss[0][0] / nVoxels;
      out[3 * s + 1] = ss[1][0];
      out[3 * s + 2] = ss[2][0];
    }
    __syncthreads();
  }
}

struct DeviceSetup {
  bool ready{false};
  double *scratch{nullptr}; // cudaGetSymbolAddress(g_scratch) on this device
  unsigned blocks{0};       // min(8 per SM, MAX_BLOCKS): Sim::getAvgMinMax's grid cap
};
DeviceSetup g_setup[MAX_DEVICES];
std::mutex g_setupMutex;

int fail(char *err, size_t errLen, const char *what, cudaError_t e = cudaSuccess) {
  if (err != nullptr && errLen > 0) {
    if (e != cudaSuccess) {
      std::snprintf(err, errLen, "smeb200_avg_min_max_no_alloc: %s: %s", what,
                    cudaGetErrorString(e));
    } else {
      std::snprintf(err, errLen, "smeb200_avg_min_max_no_alloc: %s", what);
    }
  }
  return 1;
}

// once per device, with that device current
int setupDevice(int device, DeviceSetup **out, char *err, size_t errLen) {
  std::lock_guard<std::mutex> lock(g_setupMutex);
  DeviceSetup &d = g_setup[device];
  if (!d.ready) {
    void *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_scratch);
    if (e != cudaSuccess || p == nullptr) {
      return fail(err, errLen, "cudaGetSymbolAddress(scratch)", e);
    }
    int sm = 0;
    e = cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, device);
    if (e != cudaSuccess || sm <= 0) {
      return fail(err, errLen, "cudaDeviceGetAttribute(multiprocessor count)", e);
    }
    d.scratch = static_cast<double *>(p);
    const unsigned want = static_cast<unsigned>(sm) * 8u;
    d.blocks = want < MAX_BLOCKS ? want : MAX_BLOCKS;
    d.ready = true;
  }
  *out = &d;
  return 0;
}

} // namespace

extern "C" int smeb200_avg_min_max_no_alloc(const double *d_state, uint64_t n_voxels,
                                            uint32_t values_per_voxel,
                                            uint32_t n_primary_species, double *d_out,
                                            void *stream, char *err, size_t err_len) {
  if (err != nullptr && err_len > 0) {
    err[0] = '\0';
  }
  if (n_primary_species == 0 || n_primary_species > (uint32_t)MAX_SP) {
    char msg[128];
    std::snprintf(msg, sizeof msg,
                  "n_primary_species = %u is outside [1, %d] "
                  "(SMEB200_NO_ALLOC_MAX_SPECIES)",
                  n_primary_species, MAX_SP);
    return fail(err, err_len, msg);
  }
  if (values_per_voxel < n_primary_species ||
      values_per_voxel > (uint32_t)SMEB200_NO_ALLOC_MAX_VALUES_PER_VOXEL) {
    char msg[128];
    std::snprintf(msg, sizeof msg,
                  "values_per_voxel = %u is outside [n_primary_species, %d] "
                  "(SMEB200_NO_ALLOC_MAX_VALUES_PER_VOXEL)",
                  values_per_voxel, SMEB200_NO_ALLOC_MAX_VALUES_PER_VOXEL);
    return fail(err, err_len, msg);
  }
  if (n_voxels > (uint64_t)SMEB200_NO_ALLOC_MAX_VOXELS) {
    char msg[128];
    std::snprintf(msg, sizeof msg,
                  "n_voxels = %llu exceeds the maximum %llu "
                  "(SMEB200_NO_ALLOC_MAX_VOXELS)",
                  (unsigned long long)n_voxels,
                  (unsigned long long)SMEB200_NO_ALLOC_MAX_VOXELS);
    return fail(err, err_len, msg);
  }
  int device = -1;
  cudaError_t e = cudaGetDevice(&device);
  if (e != cudaSuccess) {
    return fail(err, err_len, "cudaGetDevice", e);
  }
  if (device < 0 || device >= MAX_DEVICES) {
    return fail(err, err_len, "device ordinal out of range");
  }
  DeviceSetup *d = nullptr;
  if (setupDevice(device, &d, err, err_len) != 0) {
    return 1;
  }
  if (n_voxels == 0) {
    return 0; // set-up only (do this once per device outside a stream capture)
  }
  if (d_state == nullptr || d_out == nullptr) {
    return fail(err, err_len, "d_state and d_out must be device pointers");
  }
  const unsigned n = static_cast<unsigned>(n_voxels);
  const uint64_t need = (n_voxels + BLOCK - 1) / BLOCK;
  const unsigned grid = static_cast<unsigned>(need < d->blocks ? need : d->blocks);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  stats_partial<<<grid, BLOCK, 0, st>>>(d_state, n, values_per_voxel,
                                        n_primary_species, d->scratch);
  e = cudaGetLastError();
  if (e != cudaSuccess) {
    return fail(err, err_len, "launch of the partial reduction", e);
  }
  stats_finish<<<1, BLOCK, 0, st>>>(d->scratch, grid, n_primary_species,
                                    static_cast<double>(n_voxels), d_out);
  e = cudaGetLastError();
  if (e != cudaSuccess) {
    return fail(err, err_len, "launch of the finishing reduction", e);
  }
  return 0;
}
