// BASELINE, NOT PRODUCT.  A restatement of the step structure of the reference's
// own GPU backend (SURVEY.md §2.3 / §8(d) "Reference-CUDA baseline"), written
// here from its description so that "the kernel to beat" can be timed on the
// same B200 beside ours.  The reference generates these kernels at run time
// (core/simulate/src/cudapixelsim.cpp:150-349) around a SymEngine-printed
// reaction function; SymEngine is not in this image and /root/reference does
// not exist on the GPU box, so the reaction function below is the hand-written
// gray-scott kinetics of the headline workload.  What is kept from the
// reference: one kernel per pass, nothing fused
//   reaction (1 thread / voxel, overwrites dcdt)      cudapixelsim.cpp:165-172
//   uniform diffusion via a uint32 nn[6N] gather       :174-204
//   rk212 substep 1 / 2 (1 thread / value)             :216-237
//   rk212 error (128-thread smem tree per block)       :239-273
//   error reduce (ONE block of 128 threads)            :275-301
//   clamp (1 thread / voxel)                           :336-348
// block size 128, grid = ceil(n / 128) (:40), separately stored dcdt / old /
// lower-order arrays, and one blocking download of the two error norms per step
// (:1425-1474), in the order of doRK212 + doRKAdaptive (:1549-1555, 1584-1616).
// Nothing in spatial_model_editor_b200/ links or loads this file; only
// `bench.py --impl refcuda-port` and tests/test_refcuda_port_gpu.py do.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>

namespace {

constexpr int kBlock = 128;
constexpr int S = 2; // gray-scott: U, V

struct Port {
  uint32_t nx{}, ny{}, nz{}, n{};
  double *conc{}, *dcdt{}, *old{}, *lower{}, *errAbs{}, *errRel{}, *errResult{};
  uint32_t *nn{};
  double D[3 * S]{}; // D_s / dx^2, D_s / dy^2, D_s / dz^2 per species
  double f{}, k{};
  uint32_t nErrBlocks{};
  cudaStream_t stream{};
  unsigned long long launches{};
};

// nn[6 i + {0..5}] = +x, -x, +y, -y, +z, -z neighbour, or i itself outside the
// compartment (zero-flux boundary); voxel order z-major, then x, then y
// (geometry.cpp:40-114)
__global__ void port_build_nn(uint32_t *nn, uint32_t nx, uint32_t ny, uint32_t nz) {
  const uint32_t n = nx * ny * nz;
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) {
    return;
  }
  const uint32_t y = i % ny, x = (i / ny) % nx, z = i / (ny * nx);
  nn[6 * i + 0] = x + 1 < nx ? i + ny : i;
  nn[6 * i + 1] = x > 0 ? i - ny : i;
  nn[6 * i + 2] = y + 1 < ny ? i + 1 : i;
  nn[6 * i + 3] = y > 0 ? i - 1 : i;
  nn[6 * i + 4] = z + 1 < nz ? i + nx * ny : i;
  nn[6 * i + 5] = z > 0 ? i - nx * ny : i;
}

__device__ void port_reaction_eval(double *out, const double *v, double f, double k) {
  const double U = v[0], V = v[1];
  out[0] = -(U * V * V) + f * (1.0 - U);
  out[1] = U * V * V - (f + k) * V;
}

__global__ void port_reaction(const double *conc, double *dcdt, double f, double k,
                              uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    port_reaction_eval(dcdt + i * S, conc + i * S, f, k);
  }
}

struct Dcoef {
  double d[3 * S];
};

__global__ void port_diffusion(const double *conc, double *dcdt, const uint32_t *nn,
                               Dcoef D, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) {
    return;
  }
  const uint32_t ux = nn[6 * i], dx = nn[6 * i + 1], uy = nn[6 * i + 2],
                 dy = nn[6 * i + 3], uz = nn[6 * i + 4], dz = nn[6 * i + 5];
  for (int s = 0; s < S; ++s) {
    const double c2 = 2.0 * conc[i * S + s];
    dcdt[i * S + s] += D.d[s] * (conc[ux * S + s] + conc[dx * S + s] - c2) +
                       D.d[S + s] * (conc[uy * S + s] + conc[dy * S + s] - c2) +
                       D.d[2 * S + s] * (conc[uz * S + s] + conc[dz * S + s] - c2);
  }
}

__global__ void port_sub1(double *conc, const double *dcdt, double *old, double dt,
                          uint32_t nv) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nv) {
    old[i] = conc[i];
    conc[i] += dt * dcdt[i];
  }
}

__global__ void port_sub2(double *conc, const double *dcdt, const double *old,
                          double *lower, double dt, uint32_t nv) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nv) {
    lower[i] = conc[i];
    conc[i] = 0.5 * old[i] + 0.5 * conc[i] + 0.5 * dt * dcdt[i];
  }
}

__global__ void port_error(const double *conc, const double *lower, const double *old,
                           double eps, double *errAbs, double *errRel, uint32_t nv) {
  __shared__ double sa[kBlock], sr[kBlock];
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  double a = 0.0, r = 0.0;
  if (i < nv) {
    a = fabs(conc[i] - lower[i]);
    r = a / (0.5 * (conc[i] + old[i] + eps));
  }
  sa[threadIdx.x] = a;
  sr[threadIdx.x] = r;
  __syncthreads();
  for (int w = kBlock / 2; w > 0; w /= 2) {
    if (threadIdx.x < w) {
      sa[threadIdx.x] = fmax(sa[threadIdx.x], sa[threadIdx.x + w]);
      sr[threadIdx.x] = fmax(sr[threadIdx.x], sr[threadIdx.x + w]);
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    errAbs[blockIdx.x] = sa[0];
    errRel[blockIdx.x] = sr[0];
  }
}

// one block strides over all per-block maxima
__global__ void port_error_reduce(const double *errAbs, const double *errRel,
                                  double *result, uint32_t nBlocks) {
  __shared__ double sa[kBlock], sr[kBlock];
  double a = 0.0, r = 0.0;
  for (uint32_t b = threadIdx.x; b < nBlocks; b += kBlock) {
    a = fmax(a, errAbs[b]);
    r = fmax(r, errRel[b]);
  }
  sa[threadIdx.x] = a;
  sr[threadIdx.x] = r;
  __syncthreads();
  for (int w = kBlock / 2; w > 0; w /= 2) {
    if (threadIdx.x < w) {
      sa[threadIdx.x] = fmax(sa[threadIdx.x], sa[threadIdx.x + w]);
      sr[threadIdx.x] = fmax(sr[threadIdx.x], sr[threadIdx.x + w]);
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    result[0] = sa[0];
    result[1] = sr[0];
  }
}

__global__ void port_clamp(double *conc, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    for (int s = 0; s < S; ++s) {
      if (conc[i * S + s] < 0.0) {
        conc[i * S + s] = 0.0;
      }
    }
  }
}

inline unsigned gridFor(uint64_t n) {
  return static_cast<unsigned>((n + kBlock - 1) / kBlock);
}

void evaluateDcdt(Port &p) {
  Dcoef D{};
  std::copy(p.D, p.D + 3 * S, D.d);
  port_reaction<<<gridFor(p.n), kBlock, 0, p.stream>>>(p.conc, p.dcdt, p.f, p.k, p.n);
  port_diffusion<<<gridFor(p.n), kBlock, 0, p.stream>>>(p.conc, p.dcdt, p.nn, D, p.n);
  p.launches += 2;
}

} // namespace

extern "C" {

// D: diffusion constants per species; h: voxel size; conc0: [n][2] host array
void *refport_create(int device, uint32_t nx, uint32_t ny, uint32_t nz, const double *D,
                     const double *h, double f, double k, const double *conc0) {
  if (cudaSetDevice(device) != cudaSuccess) {
    return nullptr;
  }
  auto *p = new Port();
  p->nx = nx;
  p->ny = ny;
  p->nz = nz;
  p->n = nx * ny * nz;
  p->f = f;
  p->k = k;
  for (int s = 0; s < S; ++s) {
    for (int d = 0; d < 3; ++d) {
      p->D[d * S + s] = D[s] / (h[d] * h[d]);
    }
  }
  const size_t nv = static_cast<size_t>(p->n) * S;
  p->nErrBlocks = gridFor(nv);
  bool ok = cudaStreamCreate(&p->stream) == cudaSuccess;
  ok = ok && cudaMalloc(&p->conc, nv * 8) == cudaSuccess;
  ok = ok && cudaMalloc(&p->dcdt, nv * 8) == cudaSuccess;
  ok = ok && cudaMalloc(&p->old, nv * 8) == cudaSuccess;
  ok = ok && cudaMalloc(&p->lower, nv * 8) == cudaSuccess;
  ok = ok && cudaMalloc(&p->nn, static_cast<size_t>(p->n) * 6 * 4) == cudaSuccess;
  ok = ok && cudaMalloc(&p->errAbs, p->nErrBlocks * 8ull) == cudaSuccess;
  ok = ok && cudaMalloc(&p->errRel, p->nErrBlocks * 8ull) == cudaSuccess;
  ok = ok && cudaMalloc(&p->errResult, 16) == cudaSuccess;
  ok = ok && cudaMemcpy(p->conc, conc0, nv * 8, cudaMemcpyHostToDevice) == cudaSuccess;
  if (ok) {
    port_build_nn<<<gridFor(p->n), kBlock, 0, p->stream>>>(p->nn, nx, ny, nz);
    ok = cudaStreamSynchronize(p->stream) == cudaSuccess;
  }
  if (!ok) {
    std::fprintf(stderr, "refport_create: %s\n", cudaGetErrorString(cudaGetLastError()));
    delete p;
    return nullptr;
  }
  return p;
}

// nSteps RK212 steps of size dt (fixed step: the error norms are computed and
// downloaded every step exactly as the adaptive loop does, never rejected);
// returns the device time in ms (CUDA events on the launch stream), or -1
double refport_steps(void *h, double dt, uint64_t nSteps, double *errOut) {
  auto &p = *static_cast<Port *>(h);
  const uint32_t nv = p.n * S;
  cudaEvent_t e0{}, e1{};
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaStreamSynchronize(p.stream);
  cudaEventRecord(e0, p.stream);
  double err[2] = {0.0, 0.0};
  for (uint64_t i = 0; i < nSteps; ++i) {
    evaluateDcdt(p);
    port_sub1<<<gridFor(nv), kBlock, 0, p.stream>>>(p.conc, p.dcdt, p.old, dt, nv);
    evaluateDcdt(p);
    port_sub2<<<gridFor(nv), kBlock, 0, p.stream>>>(p.conc, p.dcdt, p.old, p.lower, dt, nv);
    port_error<<<gridFor(nv), kBlock, 0, p.stream>>>(p.conc, p.lower, p.old, 1e-14,
                                                   p.errAbs, p.errRel, nv);
    port_error_reduce<<<1, kBlock, 0, p.stream>>>(p.errAbs, p.errRel, p.errResult,
                                                  p.nErrBlocks);
    cudaStreamSynchronize(p.stream);
    cudaMemcpy(err, p.errResult, 16, cudaMemcpyDeviceToHost);
    port_clamp<<<gridFor(p.n), kBlock, 0, p.stream>>>(p.conc, p.n);
    p.launches += 5;
  }
  cudaEventRecord(e1, p.stream);
  if (cudaEventSynchronize(e1) != cudaSuccess) {
    std::fprintf(stderr, "refport_steps: %s\n", cudaGetErrorString(cudaGetLastError()));
    return -1.0;
  }
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (errOut != nullptr) {
    errOut[0] = err[0];
    errOut[1] = err[1];
  }
  return ms;
}

int refport_get(void *h, double *out) {
  auto &p = *static_cast<Port *>(h);
  return cudaMemcpy(out, p.conc, static_cast<size_t>(p.n) * S * 8,
                    cudaMemcpyDeviceToHost) == cudaSuccess
             ? 0
             : 1;
}

int refport_set(void *h, const double *in) {
  auto &p = *static_cast<Port *>(h);
  return cudaMemcpy(p.conc, in, static_cast<size_t>(p.n) * S * 8,
                    cudaMemcpyHostToDevice) == cudaSuccess
             ? 0
             : 1;
}

unsigned long long refport_launches(void *h) { return static_cast<Port *>(h)->launches; }

void refport_destroy(void *h) {
  auto *p = static_cast<Port *>(h);
  if (p == nullptr) {
    return;
  }
  cudaFree(p->conc);
  cudaFree(p->dcdt);
  cudaFree(p->old);
  cudaFree(p->lower);
  cudaFree(p->nn);
  cudaFree(p->errAbs);
  cudaFree(p->errRel);
  cudaFree(p->errResult);
  cudaStreamDestroy(p->stream);
  delete p;
}

} // extern "C"
