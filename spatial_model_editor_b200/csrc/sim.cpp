#include "sim.hpp"
#include <algorithm>
#include <chrono>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <sstream>

namespace smeb200 {

// the kernel template text (device_kernels.cuh), generated at build time
extern const char *const kDeviceKernelSource;

namespace {

constexpr uint32_t MEMB_FLAG = 0x80000000u;
constexpr unsigned BLOCK = 256;

// NAME=0 in the environment switches an optimisation off (A/B measurements)
bool envFlagOff(const char *name) {
  const char *env = std::getenv(name);
  return env != nullptr && env[0] == '0';
}

// RK tableaux: pixelsim_common.hpp:138-171
namespace rk323 {
constexpr double g1[3]{1.0, 0.25, 0.666666666666666666666};
constexpr double g2[3]{0.0, 0.0, 0.0};
constexpr double g3[3]{0.0, 0.75, 0.333333333333333333333};
constexpr double beta[3]{1.0, 0.25, 0.6666666666666666666};
constexpr double delta[3]{0.0, 0.0, 1.0};
constexpr double finaliseFactors[3]{0.0, 2.0, -1.0};
} // namespace rk323
namespace rk435 {
constexpr double g1[5]{0.0, -0.497531095840104, 1.010070514199942,
                       -3.196559004608766, 1.717835630267259};
constexpr double g2[5]{1.0, 1.384996869124138, 3.878155713328178,
                       -2.324512951813145, -0.514633322274467};
constexpr double g3[5]{0.0, 0.0, 0.0, 1.642598936063715, 0.188295940828347};
constexpr double beta[5]{0.075152045700771, 0.211361016946069,
                         1.100713347634329, 0.728537814675568,
                         0.393172889823198};
constexpr double delta[7]{1.0,
                          0.081252332929194,
                          -1.083849060586449,
                          -1.096110881845602,
                          2.859440022030827,
                          -0.655568367959557,
                          -0.194421504490852};
constexpr double deltaSumReciprocal =
    1.0 / (delta[0] + delta[1] + delta[2] + delta[3] + delta[4] + delta[5] +
           delta[6]);
} // namespace rk435

template <typename T> T *devAlloc(std::size_t count, const char *what) {
  void *p = nullptr;
  check(cudaMalloc(&p, std::max<std::size_t>(count, 1) * sizeof(T)), what);
  return static_cast<T *>(p);
}

template <typename T>
T *devUpload(const std::vector<T> &v, cudaStream_t s, const char *what) {
  T *p = devAlloc<T>(v.size(), what);
  if (!v.empty()) {
    check(cudaMemcpyAsync(p, v.data(), v.size() * sizeof(T),
                          cudaMemcpyHostToDevice, s),
          what);
  }
  return p;
}

std::string functorSource(const std::string &header, uint32_t nVars,
                          const char *const *expressions, uint32_t nExpr,
                          uint32_t nParams) {
  std::ostringstream s;
  s << header;
  s << "  static __device__ __forceinline__ void eval(double* sme_out, const "
       "double* sme_values, const double* sme_params) {\n";
  for (uint32_t i = 0; i < nVars; ++i) {
    s << "    const double sme_var_" << i << " = sme_values[" << i << "];\n";
  }
  for (uint32_t i = 0; i < nParams; ++i) {
    s << "    const double sme_param_" << i << " = sme_params[" << i
      << "]; (void)sme_param_" << i << ";\n";
  }
  for (uint32_t i = 0; i < nExpr; ++i) {
    s << "    sme_out[" << i << "] = " << expressions[i] << ";\n";
  }
  s << "  }\n};\n\n";
  return s.str();
}

} // namespace


void Sim::bindDevice() const { check(cudaSetDevice(device), "cudaSetDevice"); }

uint32_t StageDesc::flags() const {
  uint32_t f = 0;
  if (useC0) {
    f |= F_USE_C0;
  }
  if (s2mode != 0) {
    f |= F_USE_S2;
    if (s2mode >= 2) {
      f |= F_USE_S2IN;
    }
    if (writeS2) {
      f |= F_WRITE_S2;
    }
  }
  if (final) {
    f |= F_FINAL;
  }
  if (lowerIsCin) {
    f |= F_LOWER_IS_CIN;
  }
  if (clampOut) {
    f |= F_CLAMP_OUT;
  }
  if (generic) {
    f |= F_GENERIC;
  }
  return f;
}

// ---- model module: generated functors + kernel template -> NVRTC -> cubin ---
void Sim::generateSource(const smeb200_problem &p) {
  std::ostringstream src;
  src << "#define SMEB_NCOMP " << std::max<uint32_t>(p.n_compartments, 1)
      << "\n";
  src << "#define SMEB_NPARAMS " << p.n_params << "\n";
  if (const char *env = std::getenv("SMEB200_C0_AHEAD");
      env != nullptr && env[0] >= '0' && env[0] <= '3') {
    src << "#define SMEB_C0_AHEAD " << env[0] << "\n"; // A/B measurements
  }
  if (const char *env = std::getenv("SMEB200_C0_SPLIT");
      env != nullptr && (env[0] == '0' || env[0] == '1')) {
    src << "#define SMEB_C0_SPLIT " << env[0] << "\n"; // A/B measurements
  }
  if (const char *env = std::getenv("SMEB200_NN_MIN_CTAS");
      env != nullptr && env[0] >= '0' && env[0] <= '8') {
    src << "#define SMEB_NN_MIN_CTAS " << env[0] << "\n"; // A/B measurements
  } else {
    // small models (<= 2 species, uniform diffusion): the voxel-list kernel is
    // held to 48 registers, five resident CTAs per SM; larger ones would spill
    bool small = p.n_compartments > 0;
    for (uint32_t c = 0; c < p.n_compartments; ++c) {
      const auto &pc = p.compartments[c];
      small = small && pc.n_species <= 2 && pc.diffusion_array == nullptr &&
              pc.n_cross_diffusion == 0;
    }
    if (small) {
      src << "#define SMEB_NN_MIN_CTAS 5\n";
    }
  }
  if (const char *env = std::getenv("SMEB200_SEG_MIN_CTAS");
      env != nullptr && env[0] >= '1' && env[0] <= '8') {
    src << "#define SMEB_SEG_MIN_CTAS " << env[0] << "\n"; // A/B measurements
  }
  if (const char *env = std::getenv("SMEB200_NN_PREFETCH");
      env != nullptr && env[0] >= '0' && env[0] <= '2') {
    src << "#define SMEB_NN_CODE_AHEAD " << env[0] << "\n"; // A/B measurements
  }
  src << kDeviceKernelSource << "\n";
  // SymEngine's code printer writes integer powers as pow(x, 2) (Symbolic::cudaCode,
  // symbolic.cpp:250-254), while the CPU solver's LLVM visitor multiplies
  // (llvm.powi).  Inside the generated functors pow with an INTEGER literal
  // exponent is therefore repeated multiplication — the same bits as the CPU
  // solver and no libdevice pow on the hot path; any other exponent is ::pow.
  src << "__device__ __forceinline__ double smeb_pow(double x, int n) {\n"
         "  return n == 2 ? x * x : n == 3 ? (x * x) * x : n == 4 ? (x * x) * (x * x)\n"
         "       : n == 1 ? x : n == 0 ? 1.0 : ::pow(x, (double)n);\n}\n"
         "__device__ __forceinline__ double smeb_pow(double x, double y) {\n"
         "  return ::pow(x, y);\n}\n"
         "#define pow(x, y) smeb_pow((x), (y))\n";
  for (uint32_t c = 0; c < p.n_compartments; ++c) {
    const auto &comp = p.compartments[c];
    if (comp.n_species == 0) {
      continue;
    }
    std::ostringstream h;
    h << "// compartment " << c << ": " << (comp.id ? comp.id : "") << "\n";
    const uint32_t sp =
        comp.n_primary_species != 0 ? comp.n_primary_species : comp.n_species;
    bool storage = false;
    for (uint32_t s = 0; comp.inv_storage != nullptr && s < comp.n_species;
         ++s) {
      storage = storage || comp.inv_storage[s] != 1.0;
    }
    h << "struct Reac" << c << " {\n  static constexpr int S = "
      << comp.n_species << ", SP = " << sp << ", NNS = " << comp.n_non_spatial
      << ", NX = " << comp.n_cross_diffusion
      << ";\n  static constexpr bool STORAGE = "
      << (storage ? "true" : "false")
      << ", NUD = " << (comp.diffusion_array != nullptr ? "true" : "false")
      << ";\n";
    if (comp.n_cross_diffusion > 0) {
      // coefficients of all terms at one voxel (symCrossDiffusion,
      // pixelsim_impl.cpp:372-389), then the face-averaged operator per term
      // (pixelsim_impl.cpp:545-589), unrolled
      const uint32_t nx = comp.n_cross_diffusion;
      h << "  static __device__ __forceinline__ void crossCoef(double* "
           "sme_out, const double* sme_values, const double* sme_params) {\n";
      for (uint32_t i = 0; i < comp.n_species; ++i) {
        h << "    const double sme_var_" << i << " = sme_values[" << i
          << "]; (void)sme_var_" << i << ";\n";
      }
      for (uint32_t i = 0; i < p.n_params; ++i) {
        h << "    const double sme_param_" << i << " = sme_params[" << i
          << "]; (void)sme_param_" << i << ";\n";
      }
      h << "    (void)sme_params;\n";
      for (uint32_t t = 0; t < nx; ++t) {
        h << "    sme_out[" << t << "] = " << comp.cross_expressions[t]
          << ";\n";
      }
      h << "  }\n";
      h << "  static __device__ __forceinline__ void cross(double* f, const "
           "double* c, const double* xp, const double* xm, const double* yp, "
           "const double* ym, const double* zp, const double* zm, const "
           "double* h2, const double* p) {\n";
      h << "    double kc[" << nx << "], k0[" << nx << "], k1[" << nx
        << "], k2[" << nx << "], k3[" << nx << "], k4[" << nx << "], k5["
        << nx << "];\n";
      h << "    crossCoef(kc, c, p); crossCoef(k0, xp, p); crossCoef(k1, xm, "
           "p); crossCoef(k2, yp, p); crossCoef(k3, ym, p); crossCoef(k4, zp, "
           "p); crossCoef(k5, zm, p);\n";
      // the terms read the state only, so their order matters per target
      // species, where it is the reference's (term order)
      for (uint32_t t = 0; t < nx; ++t) {
        const uint32_t tg = comp.cross_target[t], sr = comp.cross_source[t];
        h << "    {\n"
          << "      const double dxp = 0.5 * (kc[" << t << "] + k0[" << t
          << "]) / h2[0];\n"
          << "      const double dxm = 0.5 * (kc[" << t << "] + k1[" << t
          << "]) / h2[0];\n"
          << "      const double dyp = 0.5 * (kc[" << t << "] + k2[" << t
          << "]) / h2[1];\n"
          << "      const double dym = 0.5 * (kc[" << t << "] + k3[" << t
          << "]) / h2[1];\n"
          << "      const double dzp = 0.5 * (kc[" << t << "] + k4[" << t
          << "]) / h2[2];\n"
          << "      const double dzm = 0.5 * (kc[" << t << "] + k5[" << t
          << "]) / h2[2];\n"
          << "      f[" << tg << "] += dxp * (xp[" << sr << "] - c[" << sr
          << "]) - dxm * (c[" << sr << "] - xm[" << sr << "]) + dyp * (yp["
          << sr << "] - c[" << sr << "]) - dym * (c[" << sr << "] - ym[" << sr
          << "]) + dzp * (zp[" << sr << "] - c[" << sr << "]) - dzm * (c["
          << sr << "] - zm[" << sr << "]);\n    }\n";
      }
      h << "  }\n";
    }
    // dc/dt post-processing of the CPU solver (pixelsim.cpp:113-119)
    h << "  static __device__ __forceinline__ void nonSpatial(double* f, "
         "const double* av) {\n    (void)f; (void)av;\n";
    for (uint32_t j = 0; j < comp.n_non_spatial; ++j) {
      h << "    f[" << comp.non_spatial[j] << "] = av[" << j << "];\n";
    }
    h << "  }\n  static __device__ __forceinline__ void "
         "nonSpatialPick(const double* f, double* v) {\n    (void)f; "
         "(void)v;\n";
    for (uint32_t j = 0; j < comp.n_non_spatial; ++j) {
      h << "    v[" << j << "] = f[" << comp.non_spatial[j] << "];\n";
    }
    h << "  }\n  static __device__ __forceinline__ void storage(double* f) "
         "{\n    (void)f;\n";
    for (uint32_t s = 0; storage && s < comp.n_species; ++s) {
      if (comp.inv_storage[s] != 1.0) { // x * 1.0 is exact: skipped
        char buf[64];
        std::snprintf(buf, sizeof(buf), "%a", comp.inv_storage[s]);
        h << "    f[" << s << "] *= " << buf << ";\n";
      }
    }
    h << "  }\n";
    h << "  static __device__ __forceinline__ bool zeroStorage(int s) {\n"
         "    (void)s;\n    return false";
    for (uint32_t s = 0; comp.inv_storage != nullptr && s < comp.n_species;
         ++s) {
      if (comp.inv_storage[s] == 0.0) {
        h << " || s == " << s;
      }
    }
    h << ";\n  }\n";
    src << functorSource(h.str(), comp.n_species, comp.expressions,
                         comp.n_species, p.n_params);
  }
  for (uint32_t m = 0; m < p.n_membranes; ++m) {
    const auto &mb = p.membranes[m];
    auto stride = [&p](int32_t ci) -> uint32_t {
      return ci >= 0 ? p.compartments[ci].n_species : 0;
    };
    auto primary = [&p](int32_t ci) -> uint32_t {
      if (ci < 0) {
        return 0;
      }
      const auto &pc = p.compartments[ci];
      return pc.n_primary_species != 0 ? pc.n_primary_species : pc.n_species;
    };
    const uint32_t SA = stride(mb.comp_a), SB = stride(mb.comp_b);
    const uint32_t SAP = primary(mb.comp_a), SBP = primary(mb.comp_b);
    // padding (time, x, y, z) comes from B's voxel, else from A's
    // (pixelsim_impl.cpp:91-98)
    const uint32_t NE = mb.comp_b >= 0 ? SB - SBP : SA - SAP;
    std::ostringstream h;
    h << "// membrane " << m << ": " << (mb.id ? mb.id : "") << "\n";
    h << "struct Memb" << m << " {\n  static constexpr int SA = " << SA
      << ", SB = " << SB << ", SAP = " << SAP << ", SBP = " << SBP
      << ", NE = " << NE << ", CA = " << std::max(mb.comp_a, 0)
      << ", CB = " << std::max(mb.comp_b, 0) << ";\n";
    src << functorSource(h.str(), SAP + SBP + NE, mb.expressions, SAP + SBP,
                         p.n_params);
  }
  src << "#undef pow\n";
  src << "template <int S_OWN>\n__device__ __forceinline__ void\n"
         "smeb_memb_dispatch(smeb_u32 info, smeb_u32 partner, const double "
         "*cown,\n                   const SmebMembArgs &m, const double "
         "*params, double *f) {\n"
         "  switch (info & 0xffffu) {\n";
  for (uint32_t m = 0; m < p.n_membranes; ++m) {
    src << "  case " << m << ":\n    smeb_memb_term<Memb" << m
        << ", S_OWN>(info, partner, cown, m, params, f);\n    break;\n";
  }
  src << "  default:\n    break;\n  }\n}\n";
  kernelSource = src.str();
}

// name expression of the stage kernel of compartment ci for one flag set
std::string Sim::kernelName(std::size_t ci, uint32_t flags, bool memb) const {
  const auto &c = comps[ci];
  std::ostringstream s;
  if (memb) {
    s << "smeb_stage_memb<Reac" << ci << ", " << flags << "u>";
  } else if (c.geom == Geom::TILE) {
    s << "smeb_stage_tile<Reac" << ci << ", " << flags << "u, " << c.tile.tx
      << ", " << c.tile.ty << ", " << c.tile.nst << ", "
      << (c.tile.m3 ? "true" : "false") << ">";
  } else if (c.geom == Geom::BOX) {
    s << "smeb_stage_box<Reac" << ci << ", " << flags << "u>";
  } else if (c.geom == Geom::SEG) {
    s << "smeb_stage_seg<Reac" << ci << ", " << flags << "u, " << c.segLen << ", "
      << c.segNst << ", " << (c.segZ ? "true" : "false") << ">";
  } else {
    s << "smeb_stage_nn<Reac" << ci << ", " << flags << "u>";
  }
  return s.str();
}

// Process-wide cache of compiled modules: simulators with the same generated
// source and kernel list (ensemble members, optimizer candidates, the
// simulator Simulation re-creates after every event — simulate.cpp:265-267)
// reuse the cubin instead of running NVRTC again.
namespace {
struct CompiledModule {
  std::vector<char> cubin;
  std::map<std::string, std::string> lowered;
};
std::mutex gModuleCacheMutex;
std::map<std::string, std::shared_ptr<const CompiledModule>> gModuleCache;
} // namespace

Sim::Module Sim::compile(const std::vector<std::string> &names) {
  Module m;
  std::string key = kernelSource;
  for (const auto &n : names) {
    key += "\n//";
    key += n;
  }
  std::shared_ptr<const CompiledModule> compiled;
  {
    std::lock_guard<std::mutex> lock(gModuleCacheMutex);
    if (auto it = gModuleCache.find(key); it != gModuleCache.end()) {
      compiled = it->second;
      ++moduleCacheHits;
    }
  }
  if (!compiled) {
    nvrtcProgram prog{};
    check(nvrtcCreateProgram(&prog, kernelSource.c_str(), "smeb200_model.cu", 0,
                             nullptr, nullptr),
          "nvrtcCreateProgram");
    for (const auto &n : names) {
      nvrtcAddNameExpression(prog, n.c_str());
    }
    // sm_100a only; --fmad=false keeps every a*b+c as two IEEE operations so
    // the result matches the CPU solver's evaluation order bit for bit.
    const char *opts[] = {"--gpu-architecture=sm_100a", "--std=c++17",
                          "--fmad=false", "-lineinfo", "-default-device"};
    nvrtcResult r = nvrtcCompileProgram(prog, 5, opts);
    if (r != NVRTC_SUCCESS) {
      std::size_t n = 0;
      nvrtcGetProgramLogSize(prog, &n);
      std::string log(n, '\0');
      nvrtcGetProgramLog(prog, log.data());
      nvrtcDestroyProgram(&prog);
      throw Error("Failed to compile CUDA pixel kernels with NVRTC: " + log);
    }
    auto fresh = std::make_shared<CompiledModule>();
    std::size_t n = 0;
    check(nvrtcGetCUBINSize(prog, &n), "nvrtcGetCUBINSize");
    fresh->cubin.resize(n);
    check(nvrtcGetCUBIN(prog, fresh->cubin.data()), "nvrtcGetCUBIN");
    for (const auto &name : names) {
      const char *lowered = nullptr;
      check(nvrtcGetLoweredName(prog, name.c_str(), &lowered),
            "nvrtcGetLoweredName");
      fresh->lowered[name] = lowered;
    }
    nvrtcDestroyProgram(&prog);
    compiled = fresh;
    std::lock_guard<std::mutex> lock(gModuleCacheMutex);
    if (gModuleCache.size() >= 64) {
      gModuleCache.clear();
    }
    gModuleCache[key] = compiled;
  }
  cubinBytes += compiled->cubin.size();
  if (compileOnly) {
    return m;
  }
  const auto &drv = DriverApi::get();
  check(drv.ModuleLoadData(&m.mod, compiled->cubin.data()), "cuModuleLoadData");
  for (const auto &name : names) {
    CUfunction f{};
    check(drv.ModuleGetFunction(&f, m.mod, compiled->lowered.at(name).c_str()),
          "cuModuleGetFunction");
    m.fn[name] = f;
  }
  return m;
}

StageKernels Sim::lookup(const Module &m, std::size_t ci,
                         uint32_t flags) const {
  StageKernels k;
  const auto &drv = DriverApi::get();
  auto it = m.fn.find(kernelName(ci, flags, false));
  if (it != m.fn.end()) {
    k.main = it->second;
    if (comps[ci].geom == Geom::TILE || comps[ci].geom == Geom::SEG) {
      check(drv.FuncSetAttribute(
                k.main, CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES,
                static_cast<int>(comps[ci].geom == Geom::TILE
                                     ? comps[ci].tile.dynSmem
                                     : comps[ci].segSmem)),
            "cuFuncSetAttribute(max dynamic shared memory)");
    }
  }
  if (comps[ci].nMvox > 0) {
    it = m.fn.find(kernelName(ci, flags, true));
    if (it != m.fn.end()) {
      k.memb = it->second;
    }
  }
  return k;
}

// kernels that are only needed on demand (getDcdt, lower-order solution,
// error images) are compiled into an extra module at first use
StageKernels Sim::onDemandKernels(std::size_t ci, uint32_t flags) {
  const std::string name = kernelName(ci, flags, false);
  if (auxFn.find(name) == auxFn.end()) {
    std::vector<std::string> names{name};
    if (comps[ci].nMvox > 0) {
      names.push_back(kernelName(ci, flags, true));
    }
    auxModules.push_back(compile(names));
    for (const auto &[k, f] : auxModules.back().fn) {
      auxFn[k] = f;
    }
  }
  Module view;
  view.fn = auxFn;
  return lookup(view, ci, flags);
}

// ---- run-compressed neighbour codes (device_kernels.cuh: smeb_neighbours) -----
// nn: the reference's table (six indices per voxel, own index = no neighbour,
// bit 31 of the first entry = membrane flag).  A run is a maximal range of
// consecutive voxels whose existing +x, -x, +z, -z neighbours all sit at one
// index distance per direction; -y / +y neighbours must be i -/+ 1 (they are, in
// the reference's voxel order).  Returns false if the table does not have that
// structure or needs more than 2^25 runs: the caller keeps the plain table.
namespace {
bool compressNeighbours(const std::vector<uint32_t> &nn, uint64_t n,
                        std::vector<uint32_t> &codes, std::vector<int> &runs) {
  constexpr uint32_t runBits = 25;
  codes.assign(n, 0u);
  runs.clear();
  int64_t cur[4] = {0, 0, 0, 0};
  bool set[4] = {false, false, false, false};
  bool open = false;
  static const int dirOf[4] = {0, 1, 4, 5};
  for (uint64_t i = 0; i < n; ++i) {
    uint32_t nb[6];
    for (int k = 0; k < 6; ++k) {
      nb[k] = nn[6 * i + k];
    }
    const uint32_t memb = nb[0] & MEMB_FLAG;
    nb[0] &= ~MEMB_FLAG;
    uint32_t flags = 0;
    if (nb[2] != i) {
      if (nb[2] != i + 1) {
        return false;
      }
      flags |= 4u;
    }
    if (nb[3] != i) {
      if (nb[3] + 1 != i) {
        return false;
      }
      flags |= 8u;
    }
    int64_t d[4];
    bool exists[4];
    bool conflict = !open;
    for (int j = 0; j < 4; ++j) {
      exists[j] = nb[dirOf[j]] != i;
      d[j] = static_cast<int64_t>(nb[dirOf[j]]) - static_cast<int64_t>(i);
      if (exists[j] && (d[j] > INT32_MAX || d[j] < INT32_MIN)) {
        return false;
      }
      if (exists[j] && set[j] && cur[j] != d[j]) {
        conflict = true;
      }
    }
    if (conflict) {
      if (runs.size() / 4 >= (std::size_t{1} << runBits)) {
        return false;
      }
      runs.insert(runs.end(), {0, 0, 0, 0});
      std::fill(set, set + 4, false);
      open = true;
    }
    const std::size_t run = runs.size() / 4 - 1;
    for (int j = 0; j < 4; ++j) {
      if (exists[j]) {
        if (!set[j]) {
          set[j] = true;
          cur[j] = d[j];
          runs[4 * run + j] = static_cast<int>(d[j]);
        }
        flags |= (j == 0 ? 1u : j == 1 ? 2u : j == 2 ? 16u : 32u);
      }
    }
    codes[i] = static_cast<uint32_t>(run) | (flags << runBits) | memb;
  }
  return true;
}
} // namespace

bool compressNeighbourTable(const uint32_t *nn, uint64_t n,
                            std::vector<uint32_t> &codes,
                            std::vector<int> &runs) {
  return compressNeighbours(std::vector<uint32_t>(nn, nn + 6 * n), n, codes,
                            runs);
}

// ---- segment-staged voxel-list kernel (device_kernels.cuh: smeb_stage_seg) -------
// Decides per voxel-list compartment whether its stage kernels stage pieces of
// runs in shared memory, and cuts the list into those pieces.  Conditions: the
// run-compressed neighbour codes exist, voxels are 16-byte multiples (even
// species count: the bulk copies), uniform diffusion without cross terms (the
// staged operands are the seven states), one GPU (a z-slab updates a sub-range),
// and runs long enough that a piece fills most of a CTA.  variant 1 / 3 keep the
// direct kernel, SMEB200_SEG=0 does too (A/B).
void Sim::chooseSegments(int variant,
                         std::vector<std::vector<uint32_t>> &nnHost) {
  for (std::size_t ci = 0; ci < comps.size(); ++ci) {
    auto &c = comps[ci];
    if (c.box || c.S == 0 || c.n == 0 || compileOnly || nnHost[ci].empty() ||
        variant == 3) {
      continue;
    }
    if (!compressNeighbours(nnHost[ci], c.n, c.hostCodes, c.hostRuns)) {
      c.hostCodes.clear();
      c.hostRuns.clear();
      continue;
    }
    // Measured on B200 (profiles/r2_seg_ab.md): 8192^2 two-compartment brusselator
    // RK323 0.469 ms per stage launch against 0.470 ms of the direct kernel, the
    // 4096^2 disc 0.062 against 0.054 ms — both kernels sit at the same 76 % of the
    // HBM rate, limited by instruction issue and fp64 latency per voxel, not by
    // loads in flight.  The staged kernel therefore stays opt-in (kernel_variant 4
    // or SMEB200_SEG=1).
    const char *segEnv = std::getenv("SMEB200_SEG");
    const bool wanted = variant == 4 || (segEnv != nullptr && segEnv[0] == '1');
    if (!wanted || nRanks > 1 || c.S % 2 != 0 || c.nonUniform ||
        !c.crossTarget.empty()) {
      continue;
    }
    bool z = false;
    for (std::size_t r = 0; r < c.hostRuns.size() / 4; ++r) {
      z = z || c.hostRuns[4 * r + 2] != 0 || c.hostRuns[4 * r + 3] != 0;
    }
    // piece length: 512 voxels where four stages of (own + halo, +x, -x [, +z,
    // -z]) rows fit beside a second CTA, else 256
    const unsigned rows = z ? 5u : 3u;
    int seg = 0, nst = 2; // two stages of 512 voxels: four CTAs per SM
    int forcedLen = 0;
    if (const char *env = std::getenv("SMEB200_SEG_NST"); env != nullptr) {
      nst = std::max(2, std::min(8, std::atoi(env))); // A/B measurements
    }
    if (const char *env = std::getenv("SMEB200_SEG_LEN"); env != nullptr) {
      forcedLen = std::atoi(env);
    }
    for (int cand : {1024, 512, 256}) {
      if (cand == 1024 && forcedLen != 1024) {
        continue; // four voxels per thread: A/B measurements only
      }
      const unsigned stage =
          ((static_cast<unsigned>(cand) * rows + 2u) * c.S * 8u + 127u) & ~127u;
      if ((forcedLen == 0 || forcedLen == cand) &&
          stage * static_cast<unsigned>(nst) + 128u <= 200u * 1024u) {
        seg = cand;
        c.segSmem = stage * static_cast<unsigned>(nst) + 128u;
        break;
      }
    }
    if (seg == 0) {
      continue;
    }
    constexpr uint32_t runMask = (1u << 25) - 1u;
    std::vector<uint32_t> segs;
    uint64_t i = 0;
    while (i < c.n) {
      const uint32_t run = c.hostCodes[i] & runMask;
      uint64_t e = i + 1;
      while (e < c.n && e - i < static_cast<uint64_t>(seg) &&
             (c.hostCodes[e] & runMask) == run) {
        ++e;
      }
      segs.push_back(static_cast<uint32_t>(i));
      segs.push_back(static_cast<uint32_t>(e - i));
      i = e;
    }
    const double avg = static_cast<double>(c.n) / (segs.size() / 2);
    if (avg < 0.4 * seg) {
      continue; // short runs (cell outlines): the direct kernel
    }
    c.geom = Geom::SEG;
    c.segLen = seg;
    c.segNst = nst;
    c.segZ = z;
    c.nSeg = static_cast<uint32_t>(segs.size() / 2);
    c.hostSeg = std::move(segs);
  }
}

constexpr unsigned SMEB_TILE_CTAS_PER_SM = 2; // device_kernels.cuh SMEB_TILE_MIN_CTAS

// ---- geometry path per compartment -------------------------------------------
// Box compartments (implicit neighbours) use the TMA tile-marching kernel when
// the array can be described by a tensor map (16-byte row pitch) and the tile
// fits; otherwise the direct box kernel.  variant: 0 auto, 1 direct kernels
// only, 2 tile kernel required.
void Sim::chooseGeometry(int variant) {
  for (auto &c : comps) {
    if (!c.box) {
      continue; // Geom::NN, or Geom::SEG chosen by chooseSegments
    }
    c.geom = Geom::BOX;
    // per-voxel diffusion constants and cross-diffusion live in the direct
    // kernels only
    if (variant == 1 || c.S == 0 || c.n == 0 || c.nonUniform ||
        !c.crossTarget.empty()) {
      continue;
    }
    const bool pitchOk = (static_cast<uint64_t>(c.ny) * c.S) % 2 == 0;
    // TY: 64 voxels of y per tile, less for narrow images; the TMA box is
    // limited to 256 elements per dimension
    int ty = 64;
    while (ty > 8 && ty / 2 >= static_cast<int>(c.ny)) {
      ty /= 2;
    }
    // row of the TMA box in doubles; an odd species count is padded by one
    // double on each side so that the box starts 16-byte aligned
    auto boxRow = [&c](int t) {
      return (t + 2) * static_cast<int>(c.S) + 2 * static_cast<int>(c.S % 2);
    };
    while (ty > 8 && boxRow(ty) > 256) {
      ty /= 2;
    }
    const bool boxOk = boxRow(ty) <= 256;
    if (!pitchOk || !boxOk) {
      if (variant == 2) {
        throw Error("smeb200: the TMA tile kernel cannot take compartment '" +
                    c.id + "' (row pitch not a multiple of 16 bytes or too "
                           "many species)");
      }
      continue;
    }
    const int groups = static_cast<int>(BLOCK) / ty;
    int rpt = c.S <= 2 ? 4 : (c.S <= 4 ? 2 : 1);
    while (rpt > 1 && c.nx % rpt != 0) {
      rpt /= 2;
    }
    TileConfig t;
    t.ty = ty;
    t.tx = groups * rpt;
    t.m3 = c.nz > 1;
    const unsigned boxBytes = (t.tx + 2) * boxRow(t.ty) * 8;
    const unsigned slotBytes = (boxBytes + 127u) & ~127u;
    t.nst = t.m3 ? 5 : 3;
    const int nstMin = t.m3 ? 4 : 2;
    while (t.nst > nstMin && t.nst * slotBytes + 128u > 110u * 1024u) {
      --t.nst;
    }
    t.dynSmem = t.nst * slotBytes + 128u;
    if (t.dynSmem > 220u * 1024u) {
      if (variant == 2) {
        throw Error("smeb200: tile does not fit shared memory");
      }
      continue;
    }
    const unsigned tilesY = (c.ny + t.ty - 1) / t.ty;
    // Split of the march direction into chunks: the CTA count should fill whole
    // waves of resident CTAs (2 per SM) — a last wave that is 10 % full costs a
    // full CTA's time at a fraction of the bandwidth — while every chunk re-reads
    // its halo (2 planes in 3-D, 2 rows of a brick in 2-D).  Pick the chunk count
    // with the least (wave rounding x halo re-read) cost among those that give
    // at least two waves.
    const unsigned wave = static_cast<unsigned>(smCount) * SMEB_TILE_CTAS_PER_SM;
    // z-slabs: at least four chunks, so that the chunks holding the boundary
    // planes (scheduled first) are a small part of the launch and the halos
    // leave early
    const unsigned minChunks = nRanks > 1 ? 4u : 1u;
    double haloWeight = 0.5;
    if (const char *env = std::getenv("SMEB200_HALO_WEIGHT"); env != nullptr) {
      haloWeight = std::atof(env); // A/B measurements (round 1: 1.0)
    }
    auto pickChunks = [&](unsigned units, unsigned tiles, double haloUnits) {
      unsigned best = 1;
      double bestCost = 1e300;
      const unsigned maxChunks = std::min(units, 65535u);
      if (static_cast<double>(tiles) * units < 2.0 * wave) {
        // too small to fill the GPU whatever the split: chunks of ~8 units
        return std::max(1u, units / 8u);
      }
      for (unsigned chunks = std::min(minChunks, maxChunks); chunks <= maxChunks;
           ++chunks) {
        const unsigned chunk = (units + chunks - 1) / chunks;
        const unsigned used = (units + chunk - 1) / chunk; // non-empty chunks
        if (used != chunks) {
          continue;
        }
        const double total = static_cast<double>(tiles) * chunks;
        if (total < 2.0 * wave && chunks < maxChunks && chunk > 1) {
          continue;
        }
        const double waves = std::ceil(total / wave);
        // the halo units of a chunk are read again, but from L2 (ncu: DRAM reads
        // stay at 1.01 x the stage input down to 8-plane chunks): half weight
        const double cost =
            waves * wave / total * (chunk + haloWeight * haloUnits) / chunk;
        if (cost < bestCost - 1e-12) {
          bestCost = cost;
          best = chunks;
        }
        if (total > 64.0 * wave) {
          break;
        }
      }
      return best;
    };
    unsigned forced = 0;
    if (const char *env = std::getenv("SMEB200_TILE_CHUNKS"); env != nullptr) {
      forced = static_cast<unsigned>(std::atoi(env));
    }
    if (t.m3) {
      const unsigned tilesX = (c.nx + t.tx - 1) / t.tx;
      const unsigned owned = c.nz - c.zLoGhost - c.zHiGhost;
      const unsigned tiles = tilesX * tilesY;
      unsigned chunks = forced ? std::min(forced, owned) : pickChunks(owned, tiles, 2.0);
      unsigned chunk = (owned + chunks - 1) / chunks;
      chunks = (owned + chunk - 1) / chunk;
      t.chunk = chunk;
      t.grid[0] = tilesY;
      t.grid[1] = tilesX;
      t.grid[2] = chunks;
      // measured slower than the plain order (same box: 512^3 1.736 vs 1.678
      // ms/step, 512x512x64 0.2385 vs 0.2348): opt-in only
      const char *zz = std::getenv("SMEB200_ZIGZAG");
      t.zig = zz != nullptr && zz[0] == '1' && chunks > 1;
      zigzag = zigzag || t.zig;
    } else {
      const unsigned bricks = (c.nx + t.tx - 1) / t.tx;
      unsigned chunks = forced ? std::min(forced, bricks)
                               : pickChunks(bricks, tilesY, 2.0 / t.tx);
      unsigned chunk = (bricks + chunks - 1) / chunks;
      chunks = (bricks + chunk - 1) / chunk;
      t.chunk = chunk;
      t.grid[0] = tilesY;
      t.grid[1] = chunks;
      t.grid[2] = 1;
    }
    if (t.grid[1] > 65535u || t.grid[2] > 65535u) {
      if (variant == 2) {
        throw Error("smeb200: tile grid too large");
      }
      continue;
    }
    c.tile = t;
    c.geom = Geom::TILE;
  }
}

// one tensor map per concentration buffer: fp64 [nz][nx][ny*S], box
// [(TY+2)*S, TX+2, 1], no swizzle, out-of-bounds cells read as zero
void Sim::encodeTensorMap(const CompartmentState &c, double *buffer,
                          CUtensorMap *out) {
  const auto &drv = DriverApi::get();
  const cuuint64_t dims[3] = {static_cast<cuuint64_t>(c.ny) * c.S, c.nx, c.nz};
  const cuuint64_t strides[2] = {
      static_cast<cuuint64_t>(c.ny) * c.S * 8,
      static_cast<cuuint64_t>(c.nx) * c.ny * c.S * 8};
  const cuuint32_t box[3] = {
      static_cast<cuuint32_t>((c.tile.ty + 2) * c.S + 2 * (c.S % 2)),
      static_cast<cuuint32_t>(c.tile.tx + 2), 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  check(drv.TensorMapEncodeTiled(
            out, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, buffer, dims, strides, box,
            estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
            CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE),
        "cuTensorMapEncodeTiled");
}

void Sim::encodeTensorMaps(CompartmentState &c) {
  for (int b = 0; b < c.nbuf; ++b) {
    encodeTensorMap(c, c.buf[b], &c.tmap[b]);
  }
}

// ---- membranes: per-voxel face lists in the reference accumulation order ----
// (membranes in model order; within one membrane XP,XM,YP,YM,ZP,ZM; within one
// direction the pair order — pixelsim.cpp:105-111, pixelsim_impl.cpp:1012-1025)
void Sim::setupMembranes(const smeb200_problem &p,
                         std::vector<std::vector<uint32_t>> &nnHost) {
  nMembranes = p.n_membranes;
  fluxLen.assign(std::size_t{6} * std::max<uint32_t>(nMembranes, 1), 1.0);
  struct Face {
    uint32_t voxel;
    MembEntryHost e;
  };
  std::vector<std::vector<Face>> faces(comps.size());
  for (uint32_t m = 0; m < p.n_membranes; ++m) {
    const auto &mb = p.membranes[m];
    if (mb.comp_a >= static_cast<int32_t>(comps.size()) ||
        mb.comp_b >= static_cast<int32_t>(comps.size())) {
      throw Error("membrane refers to a compartment that does not exist");
    }
    for (uint32_t d = 0; d < 6; ++d) {
      fluxLen[6 * m + d] = mb.flux_length[d];
      for (uint64_t k = 0; k < mb.n_pairs[d]; ++k) {
        const uint32_t ixA = mb.pairs[d][2 * k];
        const uint32_t ixB = mb.pairs[d][2 * k + 1];
        // z-slabs: only owned voxels are updated; a partner may be a ghost
        auto owned = [this](int32_t ci, uint32_t ix) {
          return ix >= comps[ci].ownedBegin() && ix < comps[ci].ownedEnd();
        };
        if (mb.comp_a >= 0 && comps[mb.comp_a].S > 0 && owned(mb.comp_a, ixA)) {
          faces[mb.comp_a].push_back({ixA, {ixB, m | (0u << 16) | (d << 17)}});
        }
        if (mb.comp_b >= 0 && comps[mb.comp_b].S > 0 && owned(mb.comp_b, ixB)) {
          faces[mb.comp_b].push_back({ixB, {ixA, m | (1u << 16) | (d << 17)}});
        }
      }
    }
  }
  for (std::size_t c = 0; c < comps.size(); ++c) {
    auto &f = faces[c];
    if (f.empty()) {
      continue;
    }
    if (comps[c].box) {
      throw Error("a box compartment (implicit neighbours) cannot have "
                  "membrane faces: supply its neighbour list");
    }
    std::stable_sort(f.begin(), f.end(), [](const Face &a, const Face &b) {
      return a.voxel < b.voxel;
    });
    std::vector<uint32_t> mvox, mstart;
    std::vector<MembEntryHost> ent;
    ent.reserve(f.size());
    for (std::size_t k = 0; k < f.size(); ++k) {
      if (f[k].voxel >= comps[c].n) {
        throw Error("membrane voxel index out of range");
      }
      if (k == 0 || f[k].voxel != f[k - 1].voxel) {
        mvox.push_back(f[k].voxel);
        mstart.push_back(static_cast<uint32_t>(k));
      }
      ent.push_back(f[k].e);
    }
    mstart.push_back(static_cast<uint32_t>(f.size()));
    comps[c].nMvox = static_cast<uint32_t>(mvox.size());
    for (uint32_t v : mvox) {
      nnHost[c][std::size_t{6} * v] |= MEMB_FLAG;
    }
    comps[c].dMvox = devUpload(mvox, stream, "upload membrane voxels");
    comps[c].dMstart = devUpload(mstart, stream, "upload membrane offsets");
    comps[c].dEnt = devUpload(ent, stream, "upload membrane faces");
    check(cudaStreamSynchronize(stream), "upload membrane tables");
  }
  dFluxLen = devUpload(fluxLen, stream, "upload flux lengths");
  check(cudaStreamSynchronize(stream), "upload flux lengths");
}

// ---- the integrator as a list of fused stages over rotating buffers ---------
void Sim::makePlan() {
  plan.clear();
  int nbuf = 2;
  if (integrator == SMEB200_RK101) {
    // pixelsim.cpp:122-140
    StageDesc s;
    s.cin = CUR;
    s.cout = RA;
    s.clampOut = true;
    plan.push_back(s);
    nbuf = 2;
  } else if (integrator == SMEB200_RK212) {
    // pixelsim.cpp:142-164, pixelsim_impl.cpp:672-701
    StageDesc s1;
    s1.cin = CUR;
    s1.cout = RA;
    plan.push_back(s1);
    StageDesc s2;
    s2.cin = RA;
    s2.cout = RB;
    s2.useC0 = true;
    s2.g1 = 0.5;
    s2.g3 = 0.5;
    s2.beta = 0.5;
    s2.final = true;
    s2.lowerIsCin = true;
    s2.clampOut = true;
    plan.push_back(s2);
    nbuf = 3;
  } else if (integrator == SMEB200_RK323) {
    // pixelsim.cpp:166-178; s2 only ever holds delta[2]*c2, so it stays in
    // registers and the lower-order solution 2*c2 - c0 is formed in the
    // final stage (pixelsim_impl.cpp:737-743)
    const int cin[3] = {CUR, RA, RB};
    const int cout[3] = {RA, RB, RA};
    for (int i = 0; i < 3; ++i) {
      StageDesc s;
      s.generic = true;
      s.cin = cin[i];
      s.cout = cout[i];
      s.useC0 = rk323::g3[i] != 0.0;
      s.s2mode = rk323::delta[i] != 0.0 ? 1 : 0;
      s.g1 = rk323::g1[i];
      s.g2 = rk323::g2[i];
      s.g3 = rk323::g3[i];
      s.beta = rk323::beta[i];
      s.delta = rk323::delta[i];
      if (i == 2) {
        s.final = true;
        s.clampOut = true;
        s.useC0 = true;
        s.cF = rk323::finaliseFactors[0];
        s.s2F = rk323::finaliseFactors[1];
        s.s3F = rk323::finaliseFactors[2];
      }
      plan.push_back(s);
    }
    nbuf = 3;
  } else {
    // RK435, pixelsim.cpp:180-192: s2 is a true running sum.  After stage 0
    // it equals c0 exactly (0 + 1*c0), so stage 1 reads it from the c0 buffer.
    const int cin[5] = {CUR, RA, RB, RA, RB};
    const int cout[5] = {RA, RB, RA, RB, RA};
    for (int i = 0; i < 5; ++i) {
      StageDesc s;
      s.generic = true;
      s.cin = cin[i];
      s.cout = cout[i];
      s.useC0 = rk435::g3[i] != 0.0;
      s.s2mode = i == 0 ? 1 : (i == 1 ? 2 : 3);
      s.writeS2 = i >= 1 && i <= 3;
      s.g1 = rk435::g1[i];
      s.g2 = rk435::g2[i];
      s.g3 = rk435::g3[i];
      s.beta = rk435::beta[i];
      s.delta = rk435::delta[i];
      if (i == 4) {
        s.final = true;
        s.clampOut = true;
        s.useC0 = true;
        s.cF = rk435::deltaSumReciprocal * rk435::delta[5];
        s.s2F = rk435::deltaSumReciprocal;
        s.s3F = rk435::deltaSumReciprocal * rk435::delta[6];
      }
      plan.push_back(s);
    }
    nbuf = 4;
  }
  for (auto &c : comps) {
    c.nbuf = nbuf;
  }
}


// ---- constructor: mirrors PixelSim::PixelSim / CudaPixelSim::CudaPixelSim ---
Sim::Sim(const smeb200_problem &p, bool compileOnlyMode)
    : compileOnly{compileOnlyMode} {
  try {
    build(p);
    if (compileOnly) {
      ctorFailed = true; // not runnable
    }
  } catch (const std::exception &e) {
    // like the reference constructors: never throw, keep the message
    // (pixelsim.cpp:415-418, cudapixelsim.cpp:1156-1167)
    errorMessage = e.what();
    ctorFailed = true;
    abortSlabs(); // z-slabs: the other ranks' constructors are waiting
  }
}

void Sim::build(const smeb200_problem &p) {
  if (p.abi_version != SMEB200_ABI_VERSION) {
    throw Error("smeb200: ABI version mismatch");
  }
  integrator = p.integrator;
  errMaxAbs = p.max_err_abs;
  errMaxRel = p.max_err_rel;
  maxTimestep = p.max_timestep;
  imageNx = p.image_nx;
  imageNy = p.image_ny;
  imageNz = p.image_nz;
  if (integrator < SMEB200_RK101 || integrator > SMEB200_RK435) {
    throw Error("smeb200: unknown integrator");
  }
  // feature gates of the reference GPU backend, same texts
  // (cudapixelsim.cpp:789-809, pixelsim_common.hpp:80-121)
  if (p.unsupported & SMEB200_HAS_CROSS_DIFFUSION) {
    throw Error("CUDA pixel backend PoC does not yet support cross-diffusion");
  }
  if (p.unsupported & SMEB200_HAS_TIME_DEPENDENCE) {
    throw Error("CUDA pixel backend PoC does not yet support time-dependent "
                "reaction terms");
  }
  if (p.unsupported & SMEB200_HAS_SPACE_DEPENDENCE) {
    throw Error("CUDA pixel backend PoC does not yet support spatially "
                "dependent reaction terms");
  }
  if (p.unsupported & SMEB200_HAS_NON_SPATIAL_SPECIES) {
    throw Error("CUDA pixel backend PoC does not yet support non-spatial "
                "species");
  }
  if (p.unsupported & SMEB200_HAS_NON_UNIFORM_DIFFUSION) {
    throw Error("CUDA pixel backend PoC does not yet support non-uniform "
                "diffusion constants");
  }
  if (p.unsupported & SMEB200_HAS_NON_UNIT_STORAGE) {
    throw Error("CUDA pixel backend PoC currently requires unit storage for "
                "all species");
  }
  if (p.n_compartments > 32) {
    throw Error("smeb200: at most 32 compartments are supported");
  }
  if (p.n_params > 64) {
    throw Error("smeb200: at most 64 runtime parameters are supported");
  }
  if (p.n_params > 0) {
    params.assign(p.params, p.params + p.n_params);
  }
  if (!compileOnly) {
    // no CPU fallback: without a device the simulator carries an error
    if (auto reason = unavailableReason(); !reason.empty()) {
      throw Error(reason);
    }
    if (p.device >= 0) {
      device = p.device;
    } else {
      check(cudaGetDevice(&device), "Failed to get the primary CUDA device");
    }
    bindDevice();
    check(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking),
          "Failed to create the CUDA stream");
    DriverApi::get();
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) ==
            cudaSuccess &&
        sms > 0) {
      smCount = sms;
    }
  }

  comps.resize(p.n_compartments);
  std::vector<std::vector<uint32_t>> nnHost(p.n_compartments);
  for (uint32_t ci = 0; ci < p.n_compartments; ++ci) {
    const auto &pc = p.compartments[ci];
    auto &c = comps[ci];
    c.id = pc.id != nullptr ? pc.id : "";
    c.n = pc.n_voxels;
    c.S = pc.n_species;
    if (c.S > 64) {
      throw Error("smeb200: at most 64 species per compartment");
    }
    if (c.n >= (uint64_t{1} << 31)) {
      throw Error("smeb200: at most 2^31 voxels per compartment");
    }
    c.box = pc.nn == nullptr;
    if (c.box) {
      if (pc.box_nx <= 0 || pc.box_ny <= 0 || pc.box_nz <= 0 ||
          static_cast<uint64_t>(pc.box_nx) * pc.box_ny * pc.box_nz != c.n) {
        throw Error("smeb200: compartment '" + c.id +
                    "' has neither a neighbour list nor valid box dimensions");
      }
      c.nx = pc.box_nx;
      c.ny = pc.box_ny;
      c.nz = pc.box_nz;
    } else if (!compileOnly) {
      nnHost[ci].assign(pc.nn, pc.nn + 6 * c.n);
      for (uint32_t v : nnHost[ci]) {
        if (v >= c.n) {
          throw Error("smeb200: neighbour index out of range in '" + c.id +
                      "'");
        }
      }
    }
    if (pc.voxels != nullptr && !compileOnly) {
      c.voxels.assign(pc.voxels, pc.voxels + 3 * c.n);
    }
    c.D.assign(pc.diffusion, pc.diffusion + 3 * c.S);
    c.SP = pc.n_primary_species != 0 ? pc.n_primary_species : c.S;
    if (c.SP > c.S) {
      throw Error("smeb200: n_primary_species exceeds n_species in '" + c.id +
                  "'");
    }
    for (uint32_t k = 0; k < pc.n_non_spatial; ++k) {
      if (pc.non_spatial[k] >= c.S) {
        throw Error("smeb200: non-spatial species index out of range in '" +
                    c.id + "'");
      }
      c.nonSpatial.push_back(pc.non_spatial[k]);
    }
    if (!c.nonSpatial.empty() && p.n_ranks > 1) {
      if (c.nonSpatial.size() > 64) {
        throw Error("smeb200: too many non-spatial species for a z-slab "
                    "decomposed simulation");
      }
      nsInSlabs = true; // the spatial mean is summed over the ranks per stage
    }
    if (pc.inv_storage != nullptr) {
      c.invStorage.assign(pc.inv_storage, pc.inv_storage + c.S);
      for (double v : c.invStorage) {
        if (v == 0.0) {
          continue; // zero-storage (algebraic) species, see below
        }
        if (!(v > 0.0)) {
          throw Error("Invalid negative storage value for species in '" +
                      c.id + "'");
        }
      }
    }
    for (int k = 0; k < 3; ++k) {
      h2[k] = p.voxel_size[k] * p.voxel_size[k];
    }
    c.nonUniform = pc.diffusion_array != nullptr;
    if (pc.n_cross_diffusion > 0) {
      if (pc.diffusion_raw == nullptr || pc.cross_target == nullptr ||
          pc.cross_source == nullptr || pc.cross_expressions == nullptr) {
        throw Error("smeb200: cross-diffusion terms need targets, sources, "
                    "expressions and diffusion_raw");
      }
      if (pc.n_cross_diffusion > 64) {
        throw Error("smeb200: at most 64 cross-diffusion terms per "
                    "compartment");
      }
      for (uint32_t k = 0; k < pc.n_cross_diffusion; ++k) {
        if (pc.cross_target[k] >= c.SP || pc.cross_source[k] >= c.SP) {
          throw Error("smeb200: cross-diffusion species index out of range "
                      "in '" + c.id + "'");
        }
        c.crossTarget.push_back(pc.cross_target[k]);
        c.crossSource.push_back(pc.cross_source[k]);
      }
      c.diffusionRaw.assign(pc.diffusion_raw, pc.diffusion_raw + c.SP);
      anyCross = true;
    }
    if ((c.nonUniform || !c.crossTarget.empty()) &&
        !(h2[0] > 0.0 && h2[1] > 0.0 && h2[2] > 0.0)) {
      throw Error("smeb200: voxel_size is required with non-uniform or "
                  "cross-diffusion");
    }
    for (uint32_t s = 0; s < c.S; ++s) {
      c.speciesNames.emplace_back(
          pc.species_names != nullptr && pc.species_names[s] != nullptr
              ? pc.species_names[s]
              : "");
      // forwards-Euler stability bound, pixelsim_common.hpp:33-37 and
      // pixelsim_impl.cpp:249-255 (D * invStorage per direction)
      const double is = c.invStorage.empty() ? 1.0 : c.invStorage[s];
      if (is == 0.0) {
        // Zero storage: an algebraic species, relaxed in pseudo-time before
        // every stage (pixelsim.cpp:23-90).  Its diffusion bounds the
        // relaxation step, not the time step (pixelsim_impl.cpp:238-247).
        c.zeroStorage.push_back(s);
        anyZeroStorage = true;
        double dsum = c.D[3 * s] + c.D[3 * s + 1] + c.D[3 * s + 2];
        if (c.nonUniform) {
          double maxD = 0.0;
          for (uint64_t i = 0; i < c.n; ++i) {
            const double d = pc.diffusion_array[i * c.S + s];
            maxD = i == 0 ? d : std::max(maxD, d);
          }
          dsum = maxD / h2[0] + maxD / h2[1] + maxD / h2[2];
        }
        if (dsum > 0.0) {
          maxRelaxStableTimestep =
              std::min(maxRelaxStableTimestep, 1.0 / (2.0 * dsum));
        }
        continue;
      }
      double sum = c.D[3 * s] * is + c.D[3 * s + 1] * is + c.D[3 * s + 2] * is;
      if (c.nonUniform) {
        // per-voxel constants: bound from the species' maximum
        // (pixelsim_impl.cpp:257-290)
        double maxD = 0.0;
        for (uint64_t i = 0; i < c.n; ++i) {
          const double d = pc.diffusion_array[i * c.S + s];
          maxD = i == 0 ? d : std::max(maxD, d);
        }
        sum = maxD * is / h2[0] + maxD * is / h2[1] + maxD * is / h2[2];
      }
      c.maxStable = std::min(c.maxStable, 1.0 / (2.0 * sum));
      maxStableTimestep = std::min(maxStableTimestep, 1.0 / (2.0 * sum));
    }
  }
  if (anyZeroStorage &&
      maxRelaxStableTimestep == std::numeric_limits<double>::max()) {
    maxRelaxStableTimestep = 1.0; // every zero-storage species has D = 0
  }
  for (std::size_t ci = 0; ci < comps.size(); ++ci) {
    if (comps[ci].n * comps[ci].S >
        comps[dominantComp].n * comps[dominantComp].S) {
      dominantComp = ci;
    }
  }
  if (p.n_ranks > 1) {
    setupSlabs(p);
  }
  makePlan();
  if (compileOnly) {
    // no device: only mark which compartments need a membrane kernel
    for (uint32_t m = 0; m < p.n_membranes; ++m) {
      for (int32_t ci : {p.membranes[m].comp_a, p.membranes[m].comp_b}) {
        if (ci >= 0 && ci < static_cast<int32_t>(comps.size())) {
          comps[ci].nMvox = 1;
        }
      }
    }
  } else {
    setupMembranes(p, nnHost);
    chooseSegments(p.kernel_variant, nnHost);
  }
  chooseGeometry(p.kernel_variant);
  // z-slabs of box compartments on the TMA tile path exchange their halos inside
  // the stage kernels (device_kernels.cuh: "halo exchange inside the stage
  // kernel"); every other combination keeps the stream-ordered exchange
  // Measured on 2 B200s (profiles/r2_halo_ab.md), 64 planes per rank: 0.264 ms per
  // step against 0.243 ms with boundary-plane launches and copy-engine pushes on
  // the halo stream — the kernel variant alone costs 4 us per step and the
  // stores through the peer mapping 10 more — so the in-kernel exchange is opt-in
  // (SMEB200_FUSED_HALO=1).
  const char *fusedEnv = std::getenv("SMEB200_FUSED_HALO");
  if (nRanks > 1 && fusedEnv != nullptr && fusedEnv[0] == '1' && !nsInSlabs &&
      !anyZeroStorage) {
    bool any = false, all = true;
    for (const auto &c : comps) {
      if (c.S == 0 || c.n == 0) {
        continue;
      }
      any = true;
      all = all && c.geom == Geom::TILE && c.tile.m3;
    }
    fusedHalo = any && all;
  }

  // the model module: one kernel per compartment and distinct stage
  generateSource(p);
  std::vector<std::string> names;
  auto addName = [&names](const std::string &n) {
    if (std::find(names.begin(), names.end(), n) == names.end()) {
      names.push_back(n);
    }
  };
  for (std::size_t ci = 0; ci < comps.size(); ++ci) {
    if (comps[ci].S == 0) {
      continue;
    }
    for (const auto &sd : plan) {
      addName(kernelName(ci, stageFlags(sd), false));
      if (comps[ci].nMvox > 0) {
        addName(kernelName(ci, stageFlags(sd), true));
      }
    }
    if (!comps[ci].nonSpatial.empty()) {
      // pre-pass: block sums of dc/dt of the non-spatial species
      addName(kernelName(ci, F_SUM_NS, false));
      if (comps[ci].nMvox > 0) {
        addName(kernelName(ci, F_SUM_NS, true));
      }
      addName(kNsFinishKernel);
    }
    if (!comps[ci].crossTarget.empty()) {
      addName("smeb_xd_max<Reac" + std::to_string(ci) + ">");
    }
    if (!comps[ci].zeroStorage.empty()) {
      for (uint32_t f : {F_RELAX_A, F_RELAX_B | F_USE_C0}) {
        addName(kernelName(ci, f, false));
        if (comps[ci].nMvox > 0) {
          addName(kernelName(ci, f, true));
        }
      }
    }
  }
  if (fusedHalo) {
    names.push_back("smeb_halo_quiesce");
  }
  names.push_back("smeb_err_flush");
  mainModule = compile(names);
  if (compileOnly) {
    return;
  }
  kErrFlush = mainModule.fn.at("smeb_err_flush");
  if (fusedHalo) {
    kHaloQuiesce = mainModule.fn.at("smeb_halo_quiesce");
  }
  dominantKernelName = comps.empty() || comps[dominantComp].S == 0
                           ? std::string{}
                           : kernelName(dominantComp, stageFlags(plan.back()), false);

  for (uint32_t ci = 0; ci < p.n_compartments; ++ci) {
    auto &c = comps[ci];
    if (c.S == 0) {
      continue;
    }
    const std::size_t nv = c.n * c.S;
    for (int b = 0; b < c.nbuf; ++b) {
      c.buf[b] = devAlloc<double>(nv, "allocate concentration buffers");
      c.stateAllocs.push_back(c.buf[b]);
      c.role[b] = b;
    }
    check(cudaMemcpyAsync(c.buf[0], p.compartments[ci].conc0,
                          nv * sizeof(double), cudaMemcpyHostToDevice, stream),
          "upload initial concentrations");
    if (!c.box) {
      if (!c.hostCodes.empty()) {
        c.nRuns = static_cast<uint32_t>(c.hostRuns.size() / 4);
        c.dNn = devUpload(c.hostCodes, stream, "upload neighbour codes");
        c.dRuns = devUpload(c.hostRuns, stream, "upload neighbour runs");
        if (c.geom == Geom::SEG) {
          c.dSeg = devUpload(c.hostSeg, stream, "upload segment table");
        }
        check(cudaStreamSynchronize(stream), "upload neighbour codes");
        c.hostCodes = {};
        c.hostRuns = {};
        c.hostSeg = {};
      } else {
        c.dNn = devUpload(nnHost[ci], stream, "upload neighbour list");
      }
    }
    if (c.geom == Geom::TILE) {
      encodeTensorMaps(c);
    }
    for (const auto &sd : plan) {
      c.kStage.push_back(lookup(mainModule, ci, stageFlags(sd)));
      // occupancy queries up front: none while a step is being captured
      if (c.geom != Geom::TILE && c.kStage.back().main != nullptr) {
        persistentGrid(c.kStage.back().main, c.geom == Geom::SEG ? c.segSmem : 0u);
      }
      if (c.kStage.back().memb != nullptr) {
        persistentGrid(c.kStage.back().memb);
      }
    }
    if (c.nonUniform) {
      c.dDarr = devAlloc<double>(nv, "allocate diffusion constant array");
      check(cudaMemcpyAsync(c.dDarr, p.compartments[ci].diffusion_array,
                            nv * sizeof(double), cudaMemcpyHostToDevice,
                            stream),
            "upload diffusion constant array");
    }
    if (!c.zeroStorage.empty()) {
      c.kRelaxA = lookup(mainModule, ci, F_RELAX_A);
      c.kRelaxB = lookup(mainModule, ci, F_RELAX_B | F_USE_C0);
      for (const auto &k : {c.kRelaxA, c.kRelaxB}) {
        if (c.geom != Geom::TILE && k.main != nullptr) {
          persistentGrid(k.main, c.geom == Geom::SEG ? c.segSmem : 0u);
        }
        if (k.memb != nullptr) {
          persistentGrid(k.memb);
        }
      }
      c.pool.resize(3);
      for (auto &pb : c.pool) {
        pb.p = devAlloc<double>(nv, "allocate relaxation buffers");
        c.stateAllocs.push_back(pb.p);
        if (c.geom == Geom::TILE) {
          encodeTensorMap(c, pb.p, &pb.t);
        }
      }
    }
    if (!c.crossTarget.empty()) {
      c.dXdMax = devAlloc<unsigned long long>(c.crossTarget.size(),
                                              "allocate cross-diffusion maxima");
      c.kXdMax =
          mainModule.fn.at("smeb_xd_max<Reac" + std::to_string(ci) + ">");
    }
    if (!c.nonSpatial.empty()) {
      c.kSumNs = lookup(mainModule, ci, F_SUM_NS);
      if (c.geom != Geom::TILE && c.kSumNs.main != nullptr) {
        persistentGrid(c.kSumNs.main, c.geom == Geom::SEG ? c.segSmem : 0u);
      }
      if (c.kSumNs.memb != nullptr) {
        persistentGrid(c.kSumNs.memb);
      }
      const std::size_t nns = c.nonSpatial.size();
      const uint32_t mainBlocks =
          c.geom == Geom::TILE
              ? c.tile.grid[0] * c.tile.grid[1] * c.tile.grid[2]
              : static_cast<uint32_t>(smCount) * 8u;
      c.nsPartialBlocks = mainBlocks + static_cast<uint32_t>(smCount) * 8u;
      c.dNsPartial = devAlloc<double>(c.nsPartialBlocks * nns,
                                      "allocate non-spatial partial sums");
      c.dNsAv = devAlloc<double>(nns, "allocate non-spatial averages");
      check(cudaMemsetAsync(c.dNsAv, 0, nns * sizeof(double), stream),
            "clear non-spatial averages");
    }
    check(cudaMallocHost(reinterpret_cast<void **>(&c.hostConc),
                         std::max<std::size_t>(nv, 1) * sizeof(double)),
          "allocate pinned host mirror");
    // before the first run() the mirror is the initial state
    std::memcpy(c.hostConc, p.compartments[ci].conc0, nv * sizeof(double));
    c.hostStale = false;
    c.hostDcdt.assign(nv, 0.0);
    c.dcdtStale = false;
  }
  // error slot {abs, rel, (arrived, published)} + one scratch slot for on-demand
  // launches; the pinned ring the publishing CTA writes (smeb_err_arrive)
  dErr = devAlloc<unsigned long long>(8, "allocate error accumulators");
  dErrScratch = dErr + 4;
  check(cudaMemsetAsync(dErr, 0, 64, stream), "clear error accumulators");
  check(cudaHostAlloc(reinterpret_cast<void **>(&hErrRing), 32 * kErrRing,
                      cudaHostAllocMapped | cudaHostAllocPortable),
        "allocate pinned error ring");
  std::memset(hErrRing, 0, 32 * kErrRing);
  {
    void *dev = nullptr;
    check(cudaHostGetDevicePointer(&dev, hErrRing, 0),
          "cudaHostGetDevicePointer(error ring)");
    dErrRing = static_cast<unsigned long long *>(dev);
  }
  usePdl = !envFlagOff("SMEB200_PDL");
  check(cudaMallocHost(reinterpret_cast<void **>(&hDt), sizeof(double)),
        "allocate pinned step size");
  dDt = devAlloc<double>(1, "allocate step size");
  {
    // graphs pay off where a step is launch bound; SMEB200_GRAPHS=0/1 overrides
    uint64_t values = 0;
    for (const auto &c : comps) {
      values += c.n * c.S;
    }
    // (the zero-storage relaxation needs the host between the launches)
    useGraphs = !comm && !anyZeroStorage && values <= (uint64_t{1} << 22);
    if (const char *env = std::getenv("SMEB200_GRAPHS"); env != nullptr) {
      useGraphs = !comm && !anyZeroStorage && env[0] == '1';
    }
    std::size_t active = 0;
    for (const auto &c : comps) {
      active += c.S > 0 && c.n > 0 ? 1 : 0;
    }
    if (useGraphs && (active > 1 || nMembranes > 0)) {
      // capture-time fork streams: two per compartment (main, membrane)
      check(cudaEventCreateWithFlags(&forkEvent, cudaEventDisableTiming),
            "cudaEventCreate");
      for (std::size_t k = 0; k < 2 * comps.size(); ++k) {
        cudaStream_t s{};
        cudaEvent_t e{};
        check(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking),
              "cudaStreamCreate");
        check(cudaEventCreateWithFlags(&e, cudaEventDisableTiming),
              "cudaEventCreate");
        forkStreams.push_back(s);
        joinEvents.push_back(e);
      }
    }
  }
  check(cudaStreamSynchronize(stream), "initial upload");
  if (anyCross) {
    check(cudaMallocHost(reinterpret_cast<void **>(&hXdMax),
                         64 * sizeof(unsigned long long)),
          "allocate pinned cross-diffusion maxima");
    // SimCompartment ctor: coefficients of the initial state and their bound
    // (pixelsim_impl.cpp:425-428)
    updateCrossDiffusionMaxStableTimestep();
  }
  if (comm) {
    connectSlabs();
    slabsConnected = true;
    syncMaxStableTimestep(); // per-voxel diffusion constants differ per slab
  }
  if (nsInSlabs) {
    if (!comm) {
      throw Error("smeb200: non-spatial species in a z-slab decomposed "
                  "simulation need the shared-memory communicator (comm_name)");
    }
    check(cudaMallocHost(reinterpret_cast<void **>(&hNs), 64 * sizeof(double)),
          "allocate pinned non-spatial sums");
    for (auto &c : comps) { // every rank walks the compartments in the same order
      double v = static_cast<double>(c.ownedEnd() - c.ownedBegin());
      comm->allreduceSumOrdered(&v, 1);
      c.nGlobal = v;
    }
  }
}

// ---- z-slab decomposition -----------------------------------------------------
// rank r owns a contiguous range of z planes of every (box) compartment; the
// local arrays carry one ghost plane towards each neighbour
void Sim::setupSlabs(const smeb200_problem &p) {
  rank = p.rank;
  nRanks = p.n_ranks;
  if (rank < 0 || rank >= nRanks || nRanks > kCommMaxRanks) {
    throw Error("smeb200: invalid rank / n_ranks");
  }
  if (comps.size() > static_cast<std::size_t>(kCommMaxCompartments)) {
    throw Error("smeb200: too many compartments for z-slab decomposition");
  }
  for (std::size_t ci = 0; ci < comps.size(); ++ci) {
    auto &c = comps[ci];
    if (!c.box) {
      // voxel list: [ghosts of the plane below][owned][ghosts of the plane above]
      const auto &pc = p.compartments[ci];
      c.vGhostLo = pc.slab_ghost_lo;
      c.vGhostHi = pc.slab_ghost_hi;
      c.vFirstPlane = pc.slab_first_plane;
      c.vLastPlane = pc.slab_last_plane;
      const uint64_t owned = c.n - std::min<uint64_t>(c.n, uint64_t{c.vGhostLo} +
                                                                c.vGhostHi);
      if (uint64_t{c.vGhostLo} + c.vGhostHi > c.n || c.vFirstPlane > owned ||
          c.vLastPlane > owned || (rank == 0 && c.vGhostLo != 0) ||
          (rank + 1 == nRanks && c.vGhostHi != 0)) {
        throw Error("smeb200: inconsistent slab description of compartment '" +
                    c.id + "'");
      }
      continue;
    }
    c.zLoGhost = rank > 0 ? 1u : 0u;
    c.zHiGhost = rank + 1 < nRanks ? 1u : 0u;
    if (c.nz < c.zLoGhost + c.zHiGhost + 1u) {
      throw Error("smeb200: slab of compartment '" + c.id +
                  "' has no owned plane");
    }
  }
  if (!compileOnly) {
    if (p.comm_name == nullptr) {
      throw Error("smeb200: n_ranks > 1 needs comm_name");
    }
    comm = std::make_unique<ShmComm>(p.comm_name, rank, nRanks);
  }
}

// exchange CUDA IPC handles of all concentration buffers with the neighbours
// and register the handshake flags with this process's CUDA context
void Sim::connectSlabs() {
  auto *seg = comm->segment();
  auto &mine = seg->slot[rank];
  mine.device = device;
  for (std::size_t ci = 0; ci < comps.size(); ++ci) {
    auto &c = comps[ci];
    const uint64_t plane = uint64_t{c.nx} * c.ny * c.S;
    mine.ghostHiOff[ci] = c.box ? (c.nz - c.zHiGhost) * plane
                                : (c.n - c.vGhostHi) * c.S;
    mine.ghostLoCnt[ci] = c.box ? c.zLoGhost * plane : uint64_t{c.vGhostLo} * c.S;
    mine.ghostHiCnt[ci] = c.box ? c.zHiGhost * plane : uint64_t{c.vGhostHi} * c.S;
    for (int b = 0; b < c.nbuf && c.n > 0 && c.S > 0; ++b) {
      cudaIpcMemHandle_t h{};
      check(cudaIpcGetMemHandle(&h, c.buf[b]), "cudaIpcGetMemHandle");
      static_assert(sizeof(h) == 64, "CUDA IPC handle size");
      std::memcpy(mine.memHandle[ci][b], &h, sizeof(h));
    }
    // the relaxation buffers of zero-storage species take part in the exchange
    for (std::size_t k = 0; k < c.pool.size() && c.n > 0 && c.S > 0; ++k) {
      if (4 + k >= static_cast<std::size_t>(kCommMaxBuffers)) {
        throw Error("smeb200: internal error (relaxation buffers)");
      }
      cudaIpcMemHandle_t h{};
      check(cudaIpcGetMemHandle(&h, c.pool[k].p), "cudaIpcGetMemHandle");
      std::memcpy(mine.memHandle[ci][4 + k], &h, sizeof(h));
    }
  }
  mine.fusedHalo = fusedHalo ? 1 : 0;
  if (fusedHalo) {
    static_assert(2 * kCommMaxCompartments < kHaloErrWord, "halo words");
    dHalo = devAlloc<uint32_t>(kHaloWords, "allocate halo flags");
    check(cudaMemset(dHalo, 0, kHaloWords * sizeof(uint32_t)), "clear halo flags");
    cudaIpcMemHandle_t h{};
    check(cudaIpcGetMemHandle(&h, dHalo), "cudaIpcGetMemHandle");
    std::memcpy(mine.flagHandle, &h, sizeof(h));
  }
  mine.published.store(1, std::memory_order_release);
  comm->barrier();
  for (int side = 0; side < 2; ++side) {
    const int nb = side == 0 ? rank - 1 : rank + 1;
    if (nb < 0 || nb >= nRanks) {
      continue;
    }
    if ((seg->slot[nb].fusedHalo != 0) != fusedHalo) {
      throw Error("smeb200: neighbouring ranks disagree about the halo exchange "
                  "path");
    }
    if (fusedHalo) {
      cudaIpcMemHandle_t h{};
      std::memcpy(&h, seg->slot[nb].flagHandle, sizeof(h));
      void *ptr = nullptr;
      check(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess),
            "cudaIpcOpenMemHandle(halo flags)");
      peerHalo[side] = static_cast<uint32_t *>(ptr);
    }
    for (std::size_t ci = 0; ci < comps.size(); ++ci) {
      auto &c = comps[ci];
      // what this rank sends towards `side` must be what the neighbour holds
      // as ghosts on its facing side
      const uint64_t plane = uint64_t{c.nx} * c.ny * c.S;
      const uint64_t sendCnt =
          c.box ? plane
                : uint64_t{side == 0 ? c.vFirstPlane : c.vLastPlane} * c.S;
      const uint64_t peerCnt = side == 0 ? seg->slot[nb].ghostHiCnt[ci]
                                         : seg->slot[nb].ghostLoCnt[ci];
      if (sendCnt != peerCnt) {
        throw Error("smeb200: slab boundary of compartment '" + c.id +
                    "' does not match the neighbouring rank's ghost voxels");
      }
      if (peerCnt == 0) {
        continue;
      }
      for (int b = 0; b < c.nbuf; ++b) {
        cudaIpcMemHandle_t h{};
        std::memcpy(&h, seg->slot[nb].memHandle[ci][b], sizeof(h));
        void *ptr = nullptr;
        check(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess),
              "cudaIpcOpenMemHandle (is peer access between the GPUs of this "
              "node available?)");
        peerBuf[side][ci][b] = static_cast<double *>(ptr);
        ipcOpened.push_back(ptr);
      }
      for (std::size_t k = 0; k < c.pool.size(); ++k) {
        cudaIpcMemHandle_t h{};
        std::memcpy(&h, seg->slot[nb].memHandle[ci][4 + k], sizeof(h));
        void *ptr = nullptr;
        check(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess),
              "cudaIpcOpenMemHandle(relaxation buffer)");
        c.pool[k].peer[side] = static_cast<double *>(ptr);
        ipcOpened.push_back(ptr);
      }
    }
  }
  check(cudaHostRegister(comm->flagPages(), comm->flagBytes(),
                         cudaHostRegisterPortable | cudaHostRegisterMapped),
        "cudaHostRegister(halo flags)");
  flagsRegistered = true;
  void *dev = nullptr;
  check(cudaHostGetDevicePointer(&dev, comm->flagPages(), 0),
        "cudaHostGetDevicePointer(halo flags)");
  flagBaseDev = reinterpret_cast<CUdeviceptr>(dev);
  // halos are pushed on their own stream while the interior of the slab is
  // computed (SMEB200_HALO_OVERLAP=0: everything on the main stream)
  const char *ov = std::getenv("SMEB200_HALO_OVERLAP");
  if (!fusedHalo && (ov == nullptr || std::string(ov) != "0")) {
    int prLow = 0, prHigh = 0; // boundary CTAs are scheduled ahead of interior ones
    check(cudaDeviceGetStreamPriorityRange(&prLow, &prHigh), "stream priorities");
    check(cudaStreamCreateWithPriority(&haloStream, cudaStreamNonBlocking, prHigh),
          "cudaStreamCreate(halo)");
    check(cudaEventCreateWithFlags(&stageInputReady, cudaEventDisableTiming),
          "cudaEventCreate");
    check(cudaEventCreateWithFlags(&haloPushed, cudaEventDisableTiming),
          "cudaEventCreate");
    check(cudaEventCreateWithFlags(&boundaryDone, cudaEventDisableTiming),
          "cudaEventCreate");
  }
  comm->barrier();
}

// after the boundary planes of a stage wrote buffer `coutBuf`: push them into the
// neighbours' ghost planes of the same buffer and signal (on stream `s`).
// Everything is stream-ordered; the host does not wait.  (A neighbour can only
// overwrite a ghost plane of buffer B after this rank's earlier readers of B
// finished: it pushes stage k only after it saw this rank's flag of stage k-1,
// which this rank writes after the boundary-plane and membrane kernels of its
// stage k-1 -- the only kernels that read ghost voxels.)
void Sim::haloPush(int coutBuf, cudaStream_t s) {
  std::vector<const double *> local(comps.size(), nullptr);
  std::vector<std::array<double *, 2>> peer(comps.size(),
                                            std::array<double *, 2>{nullptr, nullptr});
  for (std::size_t ci = 0; ci < comps.size(); ++ci) {
    local[ci] = comps[ci].buf[coutBuf];
    peer[ci] = {peerBuf[0][ci][coutBuf], peerBuf[1][ci][coutBuf]};
  }
  haloPushPtrs(local, peer, s);
}

void Sim::haloPushPtrs(const std::vector<const double *> &local,
                       const std::vector<std::array<double *, 2>> &peer,
                       cudaStream_t s) {
  const auto &drv = DriverApi::get();
  auto *seg = comm->segment();
  ++haloSeq;
  auto flagDev = [&](int r, int side) {
    return flagBaseDev +
           static_cast<CUdeviceptr>(
               reinterpret_cast<unsigned char *>(comm->haloFlag(r, side)) -
               static_cast<unsigned char *>(comm->flagPages()));
  };
  for (std::size_t ci = 0; ci < comps.size(); ++ci) {
    auto &c = comps[ci];
    if (c.S == 0 || c.n == 0 || local[ci] == nullptr) {
      continue;
    }
    // box: whole planes; voxel list: the voxels of the first / last owned plane
    const std::size_t plane = static_cast<std::size_t>(c.nx) * c.ny * c.S;
    const std::size_t loOff =
        c.box ? c.zLoGhost * plane : std::size_t{c.vGhostLo} * c.S;
    const std::size_t loCnt = c.box ? plane : std::size_t{c.vFirstPlane} * c.S;
    const std::size_t hiCnt = c.box ? plane : std::size_t{c.vLastPlane} * c.S;
    const std::size_t hiOff =
        c.box ? (c.nz - 1 - c.zHiGhost) * plane
              : (c.n - c.vGhostHi) * c.S - hiCnt;
    if (rank > 0 && loCnt > 0) { // first owned plane -> HIGH ghosts of rank-1
      check(cudaMemcpyAsync(peer[ci][0] + seg->slot[rank - 1].ghostHiOff[ci],
                            local[ci] + loOff, loCnt * sizeof(double),
                            cudaMemcpyDefault, s),
            "halo push (low neighbour)");
    }
    if (rank + 1 < nRanks && hiCnt > 0) { // last owned plane -> LOW ghosts of rank+1
      check(cudaMemcpyAsync(peer[ci][1], local[ci] + hiOff,
                            hiCnt * sizeof(double), cudaMemcpyDefault, s),
            "halo push (high neighbour)");
    }
  }
  if (rank > 0) {
    check(drv.StreamWriteValue32(s, flagDev(rank - 1, 1), haloSeq,
                                 CU_STREAM_WRITE_VALUE_DEFAULT),
          "cuStreamWriteValue32");
  }
  if (rank + 1 < nRanks) {
    check(drv.StreamWriteValue32(s, flagDev(rank + 1, 0), haloSeq,
                                 CU_STREAM_WRITE_VALUE_DEFAULT),
          "cuStreamWriteValue32");
  }
  ++haloExchanges;
}

// make stream `s` wait for the neighbours' pushes of the latest stage
void Sim::haloWait(cudaStream_t s) {
  const auto &drv = DriverApi::get();
  auto flagDev = [&](int r, int side) {
    return flagBaseDev +
           static_cast<CUdeviceptr>(
               reinterpret_cast<unsigned char *>(comm->haloFlag(r, side)) -
               static_cast<unsigned char *>(comm->flagPages()));
  };
  if (rank > 0) {
    check(drv.StreamWaitValue32(s, flagDev(rank, 0), haloSeq,
                                CU_STREAM_WAIT_VALUE_GEQ),
          "cuStreamWaitValue32");
  }
  if (rank + 1 < nRanks) {
    check(drv.StreamWaitValue32(s, flagDev(rank, 1), haloSeq,
                                CU_STREAM_WAIT_VALUE_GEQ),
          "cuStreamWaitValue32");
  }
}

// Outside run() / benchSteps() nothing is in flight on the halo stream and all
// ghost voxels of the current state have arrived.
void Sim::haloQuiesce() {
  if (!haloPending) {
    return;
  }
  if (fusedHalo && std::getenv("SMEB200_HALO_DEBUG") != nullptr) {
    haloPending = false; // timing experiments: nothing was exchanged
    return;
  }
  if (fusedHalo) {
    // one thread per compartment waits for the neighbours' planes of the last
    // stage: every stage adds one count per tile of a boundary plane
    uint32_t mode = (rank > 0 ? 1u : 0u) | (rank + 1 < nRanks ? 2u : 0u);
    for (std::size_t ci = 0; ci < comps.size(); ++ci) {
      const auto &c = comps[ci];
      if (c.S == 0 || c.n == 0) {
        continue;
      }
      uint32_t want = haloSeq * (c.tile.grid[0] * c.tile.grid[1]);
      uint32_t *words = dHalo + 2 * ci;
      uint32_t *err = dHalo + kHaloErrWord;
      void *params[] = {&words, &mode, &want, &err};
      check(DriverApi::get().LaunchKernel(kHaloQuiesce, 1, 1, 1, 32, 1, 1, 0,
                                          stream, params, nullptr),
            "launch halo wait kernel");
    }
    haloPending = false;
    return;
  }
  if (haloStream == nullptr) {
    haloPending = false;
    return;
  }
  check(cudaEventRecord(haloPushed, haloStream), "halo event");
  haloWait(stream);
  check(cudaStreamWaitEvent(stream, haloPushed, 0), "halo pushed wait");
  haloPending = false;
}

// a bounded in-kernel wait that ran out raised halo[4] (the stream is idle)
void Sim::checkHaloError() {
  if (!fusedHalo) {
    return;
  }
  uint32_t flag = 0;
  check(cudaMemcpy(&flag, dHalo + kHaloErrWord, sizeof(flag),
                   cudaMemcpyDeviceToHost),
        "read halo status");
  if (flag != 0) {
    check(cudaMemset(dHalo + kHaloErrWord, 0, sizeof(flag)), "clear halo status");
    throw Error("smeb200: timed out waiting for the halo planes of the rank " +
                std::string(flag == 1 ? "below" : "above"));
  }
}

void Sim::abortSlabs() noexcept {
  if (!comm) {
    return;
  }
  try {
    comm->abort();
    if (!slabsConnected) {
      return;
    }
    // the waits compare cyclically: a sequence number far ahead satisfies every
    // wait the neighbours can still have queued
    const uint32_t far = haloSeq + 0x40000000u;
    if (fusedHalo) {
      for (std::size_t ci = 0; ci < comps.size(); ++ci) {
        const uint32_t farCount =
            haloSeq * (comps[ci].tile.grid[0] * comps[ci].tile.grid[1]) + 0x40000000u;
        if (rank > 0 && peerHalo[0] != nullptr) {
          cudaMemcpy(peerHalo[0] + 2 * ci + 1, &farCount, sizeof(farCount),
                     cudaMemcpyHostToDevice);
        }
        if (rank + 1 < nRanks && peerHalo[1] != nullptr) {
          cudaMemcpy(peerHalo[1] + 2 * ci, &farCount, sizeof(farCount),
                     cudaMemcpyHostToDevice);
        }
      }
      cudaGetLastError();
    } else {
      if (rank > 0) {
        __atomic_store_n(comm->haloFlag(rank - 1, 1), far, __ATOMIC_RELEASE);
      }
      if (rank + 1 < nRanks) {
        __atomic_store_n(comm->haloFlag(rank + 1, 0), far, __ATOMIC_RELEASE);
      }
    }
  } catch (...) {
  }
}

Sim::~Sim() {
  if (stream == nullptr) {
    return;
  }
  cudaSetDevice(device);
  cudaStreamSynchronize(stream);
  cudaFree(dGather);
  cudaFreeHost(hGather);
  if (haloStream != nullptr) {
    cudaStreamSynchronize(haloStream);
    cudaStreamDestroy(haloStream);
    cudaEventDestroy(stageInputReady);
    cudaEventDestroy(haloPushed);
    cudaEventDestroy(boundaryDone);
  }
  if (comm) {
    // nobody may still be pushing halos into this rank's buffers
    try {
      if (!comm->aborted()) {
        comm->barrier();
      }
    } catch (...) {
    }
    for (void *ptr : ipcOpened) {
      cudaIpcCloseMemHandle(ptr);
    }
    for (auto *ptr : peerHalo) {
      if (ptr != nullptr) {
        cudaIpcCloseMemHandle(ptr);
      }
    }
    if (flagsRegistered) {
      cudaHostUnregister(comm->flagPages());
    }
  }
  cudaFree(dHalo);
  if (xferStream != nullptr) {
    cudaStreamSynchronize(xferStream);
    cudaStreamDestroy(xferStream);
    cudaEventDestroy(downloadDone);
  }
  for (auto &c : comps) {
    for (void *b : c.stateAllocs) {
      cudaFree(b);
    }
    cudaFree(c.dNn);
    cudaFree(c.dRuns);
    cudaFree(c.dSeg);
    cudaFree(c.dMvox);
    cudaFree(c.dMstart);
    cudaFree(c.dEnt);
    cudaFree(c.scratch);
    cudaFree(c.dNsAv);
    cudaFree(c.dStats);
    cudaFree(c.dStatsPartial);
    cudaFree(c.dDarr);
    cudaFree(c.dXdMax);
    cudaFree(c.dNsPartial);
    if (c.userMirror) {
      cudaHostUnregister(c.hostConc);
    } else {
      cudaFreeHost(c.hostConc);
    }
  }
  cudaFree(dFluxLen);
  cudaFree(dErr);
  cudaFreeHost(hErrRing);
  cudaFreeHost(hNs);
  cudaFreeHost(hXdMax);
  cudaFreeHost(hDt);
  cudaFree(dDt);
  for (auto s2 : forkStreams) {
    cudaStreamDestroy(s2);
  }
  for (auto e : joinEvents) {
    cudaEventDestroy(e);
  }
  if (forkEvent != nullptr) {
    cudaEventDestroy(forkEvent);
  }
  for (auto &[k, g] : stepGraphs) {
    if (g.exec != nullptr) {
      cudaGraphExecDestroy(g.exec);
    }
  }
  if (runEv0 != nullptr) {
    cudaEventDestroy(runEv0);
    cudaEventDestroy(runEv1);
  }
  for (auto &[a, b] : timerEvents) {
    cudaEventDestroy(a);
    cudaEventDestroy(b);
  }
  try {
    const auto &drv = DriverApi::get();
    if (mainModule.mod != nullptr) {
      drv.ModuleUnload(mainModule.mod);
    }
    for (auto &m : auxModules) {
      if (m.mod != nullptr) {
        drv.ModuleUnload(m.mod);
      }
    }
  } catch (...) {
  }
  cudaStreamDestroy(stream);
}

StageArgsHost Sim::stageArgsTemplate(std::size_t ci,
                                     std::size_t stageIndex) const {
  const StageDesc &sd = plan[stageIndex];
  const auto &c = comps[ci];
  StageArgsHost a{};
  a.nn = c.dNn;
  a.runs = c.dRuns;
  a.seg = c.dSeg;
  a.nSeg = c.nSeg;
  a.err = dErr;
  a.errBase = dErr;
  a.errPairs = 1;
  a.errHost = dErrRing;
  a.n = static_cast<uint32_t>(c.n - c.vGhostHi);
  a.v0 = c.vGhostLo;
  a.nx = c.nx;
  a.ny = c.ny;
  a.nz = c.nz;
  a.z_begin = c.zLoGhost;
  a.z_end = c.nz - c.zHiGhost;
  a.g1 = sd.g1;
  a.g2 = sd.g2;
  a.g3 = sd.g3;
  a.beta = sd.beta;
  a.delta = sd.delta;
  a.cF = sd.cF;
  a.s2F = sd.s2F;
  a.s3F = sd.s3F;
  a.eps = epsilon;
  std::copy(c.D.begin(), c.D.end(), a.D);
  std::copy(params.begin(), params.end(), a.D + 3 * c.S);
  a.nsAv = c.dNsAv;
  a.nsPartial = c.dNsPartial;
  a.darr = c.dDarr;
  a.xdMax = c.dXdMax;
  std::copy(h2, h2 + 3, a.h2);
  return a;
}

// ---- one fused stage over all compartments ----------------------------------
// replaces calculateDcdt() + the RK substep (+ error norm, + clamp) of
// pixelsim.cpp:92-120,194-205
void Sim::launchStage(std::size_t stageIndex, double dt) {
  const StageDesc &sd = plan[stageIndex];
  const bool last = stageIndex + 1 == plan.size();
  // Captured step: within a stage the kernels of different compartments, and a
  // compartment's main and membrane kernels, touch disjoint voxels and only read
  // the stage input, so they become parallel branches of the graph (a step of a
  // small multi-compartment model is a chain of latencies otherwise).
  if (anyZeroStorage) {
    solveZeroStorageConstraints(stageIndex);
  }
  // stage inputs of all compartments as the membrane kernels see them (a
  // relaxed copy may stand in for the buffer of the stage's input role)
  membCinOverride.clear();
  if (anyZeroStorage) {
    for (auto &c : comps) {
      const double *pj = c.S > 0 ? c.buf[c.role[sd.cin]] : nullptr;
      if (c.haveOverride && stageIndex == 0) {
        pj = c.stageInOverride.p;
      }
      membCinOverride.push_back(pj);
    }
  }
  const bool fork = capturing && !forkStreams.empty();
  const bool overlapHalo = comm != nullptr && haloStream != nullptr;
  // pass 1 (host only): the arguments of every compartment's launches.  In a
  // z-slab run with the halo stream a compartment is launched in parts: the owned
  // boundary planes (lo, hi), the membrane voxels, the interior (mid).
  struct Item {
    std::size_t ci;
    StageArgsHost a, lo, hi, mid;
    bool hasLo{false}, hasHi{false}, hasMid{false};
  };
  std::vector<Item> items;
  if (fusedHalo) {
    ++haloSeq; // the planes this stage sends carry this number
  }
  for (std::size_t ci = 0; ci < comps.size(); ++ci) {
    auto &c = comps[ci];
    if (c.S == 0 || c.n == 0) {
      continue;
    }
    Item item;
    item.ci = ci;
    StageArgsHost &a = item.a;
    a = stageArgsTemplate(ci, stageIndex);
    a.cin = c.buf[c.role[sd.cin]];
    a.cout = c.buf[c.role[sd.cout]];
    a.c0 = c.buf[c.role[CUR]];
    if (c.haveOverride && stageIndex == 0) {
      // RK323 / RK435: the stage works on the relaxed copy, the pre-step state
      // c0 stays what doRKInit saved before the relaxation (pixelsim.cpp:166-171)
      a.cin = c.stageInOverride.p;
    }
    if (sd.s2mode == 2) {
      // s2 after stage 0 is exactly its input (0 + 1 * c): the relaxed copy if
      // there is one
      a.s2in = c.haveOverride ? c.stageInOverride.p : c.buf[c.role[CUR]];
    } else if (sd.s2mode == 3) {
      a.s2in = c.buf[c.role[RS]];
    }
    if (sd.writeS2) {
      a.s2out = c.buf[c.role[RS]];
    }
    a.bdt = sd.beta * dt;
    a.dtp = capturing ? dDt : nullptr;
    if (fusedHalo) {
      const int coutBuf = c.role[sd.cout];
      a.haloMode = (rank > 0 ? 1u : 0u) | (rank + 1 < nRanks ? 2u : 0u);
      a.halo = dHalo + 2 * ci;
      a.haloErr = dHalo + kHaloErrWord;
      a.haloSeq = haloSeq;
      if (rank > 0) {
        a.peerLo = peerBuf[0][ci][coutBuf] + comm->segment()->slot[rank - 1].ghostHiOff[ci];
        a.peerFlagLo = peerHalo[0] + 2 * ci + 1;
      }
      if (rank + 1 < nRanks) {
        a.peerHi = peerBuf[1][ci][coutBuf];
        a.peerFlagHi = peerHalo[1] + 2 * ci;
      }
      // timing experiments only (results are wrong): 1 = the boundary planes are
      // stored into local scratch memory instead of the neighbour's, nobody
      // waits; 2 = no boundary work at all in the halo-variant kernel
      static const int dbg = std::getenv("SMEB200_HALO_DEBUG") != nullptr
                                 ? std::atoi(std::getenv("SMEB200_HALO_DEBUG"))
                                 : 0;
      if (dbg == 1) {
        if (c.scratch == nullptr) {
          c.scratch = devAlloc<double>(c.n * c.S, "allocate scratch buffer");
        }
        a.peerLo = a.peerHi = c.scratch;
        a.peerFlagLo = a.peerFlagHi = dHalo + 12;
        a.haloSeq = 1;
      } else if (dbg == 2) {
        a.haloMode = 0;
      }
    }
    if (overlapHalo) {
      item.lo = item.hi = item.mid = a;
      if (c.box) {
        const uint32_t zb = a.z_begin, ze = a.z_end;
        const uint32_t loEnd = rank > 0 ? std::min(zb + 1, ze) : zb;
        const uint32_t hiBegin = rank + 1 < nRanks ? std::max(ze - 1, loEnd) : ze;
        item.lo.z_end = loEnd;
        item.hi.z_begin = hiBegin;
        item.mid.z_begin = loEnd;
        item.mid.z_end = hiBegin;
      } else {
        const uint32_t v0 = a.v0, v1 = a.n;
        const uint32_t loEnd = rank > 0 ? v0 + c.vFirstPlane : v0;
        const uint32_t hiBegin =
            rank + 1 < nRanks ? std::max(v1 - c.vLastPlane, loEnd) : v1;
        item.lo.n = loEnd;
        item.hi.v0 = hiBegin;
        item.mid.v0 = loEnd;
        item.mid.n = hiBegin;
      }
      const auto nonEmpty = [&](const StageArgsHost &r) {
        return c.box ? r.z_begin < r.z_end : r.v0 < r.n;
      };
      item.hasLo = nonEmpty(item.lo);
      item.hasHi = nonEmpty(item.hi);
      item.hasMid = nonEmpty(item.mid);
    }
    items.push_back(item);
  }
  // the CTAs that will report error norms in this stage: the one that completes
  // the count publishes the step's norms (smeb_err_arrive)
  errExpected = 0;
  uint32_t finalCtas = 0;
  if (sd.final) {
    for (const auto &it : items) {
      const auto &k = comps[it.ci].kStage[stageIndex];
      if (overlapHalo) {
        errExpected += it.hasLo ? mainGrid(it.ci, it.lo, k.main) : 0u;
        errExpected += it.hasHi ? mainGrid(it.ci, it.hi, k.main) : 0u;
        errExpected += it.hasMid ? mainGrid(it.ci, it.mid, k.main) : 0u;
      } else {
        errExpected += mainGrid(it.ci, it.a, k.main);
      }
      errExpected += membGrid(it.ci, k.memb);
    }
    finalCtas = errExpected;
    // Counting the CTAs costs each of them a fence and an atomic round trip while
    // it holds its SM slot.  Not needed when the host does not wait for these
    // norms before the next launch (lagged check), and for large grids a
    // one-warp flush kernel behind the stage is cheaper than the counting
    if (!capturing && !relaxing && (deferPublish || errExpected > 1024u)) {
      errExpected = 0;
    }
    for (auto &it : items) {
      it.a.errExpected = it.lo.errExpected = it.hi.errExpected =
          it.mid.errExpected = errExpected;
    }
  }
  std::vector<std::pair<std::size_t, StageArgsHost>> interior;
  std::size_t forksUsed = 0;
  if (fork) {
    check(cudaEventRecord(forkEvent, stream), "fork event");
  }
  if (overlapHalo) {
    // halo stream: the stage input is complete once the main stream got here
    // and the neighbours' halos of the previous stage have arrived
    check(cudaEventRecord(stageInputReady, stream), "stage input event");
    check(cudaStreamWaitEvent(haloStream, stageInputReady, 0), "halo stream wait");
    if (haloSeq > 0) {
      haloWait(haloStream);
      if (nsInSlabs) {
        // the pre-pass of the spatial mean runs on the main stream and
        // evaluates dc/dt of the boundary voxels too: it reads ghosts
        haloWait(stream);
      }
    }
  }
  for (auto &item : items) {
    const std::size_t ci = item.ci;
    auto &c = comps[ci];
    StageArgsHost &a = item.a;
    tmapOverride = nullptr;
    if (c.haveOverride && stageIndex == 0) {
      tmapOverride = &c.stageInOverride.t;
    }
    if (fork && forksUsed + 2 <= forkStreams.size()) {
      forkMain = forkStreams[forksUsed];
      // with a pre-pass the membrane kernels stay ordered with the main ones
      forkMemb = c.nonSpatial.empty() && c.nMvox > 0 ? forkStreams[forksUsed + 1]
                                                     : forkMain;
      check(cudaStreamWaitEvent(forkMain, forkEvent, 0), "fork wait");
      if (forkMemb != forkMain) {
        check(cudaStreamWaitEvent(forkMemb, forkEvent, 0), "fork wait");
      }
      forksUsed += 2;
    }
    if (!c.nonSpatial.empty()) {
      // SimCompartment::spatiallyAverageDcdt (pixelsim_impl.cpp:150-163): the
      // averages need dc/dt of every voxel before any voxel can be updated
      StageArgsHost pre = a;
      pre.errExpected = 0;
      launchCompartment(ci, pre, c.kSumNs, c.role[sd.cin], true);
      if (overlapHalo) {
        // the boundary and membrane kernels on the halo stream use the mean too
        check(cudaEventRecord(stageInputReady, stream), "mean ready event");
        check(cudaStreamWaitEvent(haloStream, stageInputReady, 0),
              "halo stream wait");
      }
    }
    if (overlapHalo) {
      // z-slab: the owned boundary planes (what the neighbours need) and the
      // membrane kernels go to the high-priority halo stream, which pushes the
      // halos as soon as they are done; the interior runs beside them on the
      // main stream (below)
      // The interior launch is enqueued first: the nine driver calls of the halo
      // stream must not sit between two interior launches on the host side; on
      // the device the halo stream's priority still gives the boundary CTAs
      // the next free SM slots.  (SMEB200_INTERIOR_FIRST=0: after the pushes.)
      static const bool interiorFirst = !envFlagOff("SMEB200_INTERIOR_FIRST");
      if (interiorFirst && item.hasMid) {
        launchCompartment(ci, item.mid, c.kStage[stageIndex], c.role[sd.cin], false, 1);
      }
      suppressTiming = true;
      forkMain = forkMemb = haloStream; // concurrent with the interior launch
      if (item.hasLo) {
        launchCompartment(ci, item.lo, c.kStage[stageIndex], c.role[sd.cin], false, 1);
      }
      if (item.hasHi) {
        launchCompartment(ci, item.hi, c.kStage[stageIndex], c.role[sd.cin], false, 1);
      }
      launchCompartment(ci, a, c.kStage[stageIndex], c.role[sd.cin], false, 2);
      forkMain = forkMemb = nullptr;
      suppressTiming = false;
      if (item.hasMid && !interiorFirst) {
        interior.emplace_back(ci, item.mid);
      }
    } else {
      launchCompartment(ci, a, c.kStage[stageIndex], c.role[sd.cin]);
    }
    tmapOverride = nullptr;
    if (forkMain != nullptr) { // join the branches of this compartment
      const bool two = forkMemb != forkMain;
      check(cudaEventRecord(joinEvents[forksUsed - 2], forkMain), "join event");
      check(cudaStreamWaitEvent(stream, joinEvents[forksUsed - 2], 0), "join");
      if (two) {
        check(cudaEventRecord(joinEvents[forksUsed - 1], forkMemb), "join event");
        check(cudaStreamWaitEvent(stream, joinEvents[forksUsed - 1], 0), "join");
      }
      forkMain = forkMemb = nullptr;
    }
    if (last) {
      c.last = a;
      c.last.errExpected = 0;
      c.last.haloMode = 0;
      c.lastCinBuf = c.role[sd.cin];
      c.lastFlags = sd.flags();
      c.haveLast = true;
    }
  }
  membCinOverride.clear();
  if (fusedHalo) {
    haloPending = true; // the stage kernels exchanged the planes themselves
    ++haloExchanges;
  } else if (comm && !comps.empty()) {
    const int coutBuf = comps[0].role[sd.cout];
    if (overlapHalo) {
      // the halos travel (and the neighbours' arrive) while the interior is
      // computed
      check(cudaEventRecord(boundaryDone, haloStream), "boundary event");
      haloPush(coutBuf, haloStream);
      for (auto &[ci, mid] : interior) {
        launchCompartment(ci, mid, comps[ci].kStage[stageIndex],
                          comps[ci].role[sd.cin], false, 1);
      }
      // the main stream never reads ghost voxels: it only needs this rank's own
      // boundary planes before the next stage (the flags of the neighbours are
      // awaited by the halo stream, and by haloQuiesce() when stepping ends)
      check(cudaStreamWaitEvent(stream, boundaryDone, 0), "boundary wait");
      haloPending = true;
    } else {
      haloPush(coutBuf, stream);
      haloWait(stream);
    }
  }
  if (sd.final && finalCtas > 0 && !capturing) {
    ++errSeq; // this step's norms will appear under this number (replays of a
              // captured step are counted by graphStep)
    if (errExpected == 0) {
      errUnpublished = true; // the next kernel launched, or flushErrors()
    }
  }
}

// the maxima of the last final stage leave the slot now (one warp, stream
// ordered behind that stage)
void Sim::flushErrors() {
  if (!errUnpublished) {
    return;
  }
  unsigned long long *base = dErr, *host = dErrRing;
  uint32_t pairs = 1;
  void *params[] = {&base, &pairs, &host};
  const auto &drv = DriverApi::get();
  check(drv.LaunchKernel(kErrFlush, 1, 1, 1, 32, 1, 1, 0, stream, params, nullptr),
        "launch error flush kernel");
  ++launchCount;
  errUnpublished = false;
}

// PixelSim::solveZeroStorageConstraints (pixelsim.cpp:23-90) before the stage
// `stageIndex`: adaptive RK212 iteration in pseudo-time on the zero-storage
// species of the stage input until the residual of 0 = R(c) + div(D grad c)
// meets the integrator's tolerances (at most 100 iterations).  One iteration =
// two fused launches per compartment (residual norm + first half; second half +
// error norm), each followed by a 16-byte read of the norms.  The reference
// updates conc in place; here an accepted iterate is a buffer that takes the
// place of the stage input (pointer swap with a pool of three scratch buffers).
void Sim::solveZeroStorageConstraints(std::size_t stageIndex) {
  const StageDesc &sd = plan[stageIndex];
  const bool protect =
      stageIndex == 0 &&
      (integrator == SMEB200_RK323 || integrator == SMEB200_RK435);
  using PoolBuf = CompartmentState::PoolBuf;
  std::vector<PoolBuf> in(comps.size());
  for (std::size_t ci = 0; ci < comps.size(); ++ci) {
    auto &c = comps[ci];
    if (stageIndex == 0) {
      c.haveOverride = false;
    }
    if (c.S == 0) {
      continue;
    }
    const int idx = c.role[sd.cin];
    in[ci].p = c.buf[idx];
    in[ci].t = c.tmap[idx];
    in[ci].peer[0] = peerBuf[0][ci][idx];
    in[ci].peer[1] = peerBuf[1][ci][idx];
    if (protect && !c.zeroStorage.empty() && c.n > 0) {
      // the pre-step state (s3) must keep the un-relaxed values: relax a copy
      check(cudaMemcpyAsync(c.pool[2].p, c.buf[idx], c.n * c.S * sizeof(double),
                            cudaMemcpyDeviceToDevice, stream),
            "copy stage input for relaxation");
      in[ci] = c.pool[2];
    }
  }
  auto launchHalf = [&](bool second, double dt) {
    std::vector<const double *> cins(comps.size(), nullptr);
    uint32_t expected = 0; // CTAs that report residual / error norms
    for (std::size_t ci = 0; ci < comps.size(); ++ci) {
      const auto &c = comps[ci];
      // second half: the compartments with zero-storage species are at c1
      cins[ci] = second && !c.zeroStorage.empty() && c.n > 0 ? c.pool[0].p
                                                              : in[ci].p;
      if (!c.zeroStorage.empty() && c.n > 0) {
        StageArgsHost r{};
        r.v0 = c.vGhostLo;
        r.n = static_cast<uint32_t>(c.n - c.vGhostHi);
        r.z_begin = c.zLoGhost;
        r.z_end = c.nz - c.zHiGhost;
        const auto &k = second ? c.kRelaxB : c.kRelaxA;
        expected += mainGrid(ci, r, k.main) + membGrid(ci, k.memb);
      }
    }
    for (std::size_t ci = 0; ci < comps.size(); ++ci) {
      auto &c = comps[ci];
      if (c.zeroStorage.empty() || c.n == 0) {
        continue;
      }
      StageArgsHost a{};
      a.cin = cins[ci];
      a.c0 = in[ci].p;
      a.cout = second ? c.pool[1].p : c.pool[0].p;
      a.nn = c.dNn;
      a.runs = c.dRuns;
      a.seg = c.dSeg;
      a.nSeg = c.nSeg;
      a.err = dErr;
      a.errBase = dErr;
      a.errPairs = 1;
    a.errBase = dErr;
    a.errPairs = 1;
      a.errHost = dErrRing;
      a.errExpected = expected;
      // z-slabs: the owned voxels; the iterate's ghosts come from the neighbours
      a.n = static_cast<uint32_t>(c.n - c.vGhostHi);
      a.v0 = c.vGhostLo;
      a.nx = c.nx;
      a.ny = c.ny;
      a.nz = c.nz;
      a.z_begin = c.zLoGhost;
      a.z_end = c.nz - c.zHiGhost;
      a.bdt = second ? 0.5 * dt : dt;
      a.eps = epsilon;
      std::copy(c.D.begin(), c.D.end(), a.D);
      std::copy(params.begin(), params.end(), a.D + 3 * c.S);
      a.nsAv = c.dNsAv;
      a.nsPartial = c.dNsPartial;
      a.darr = c.dDarr;
      a.xdMax = c.dXdMax;
      std::copy(h2, h2 + 3, a.h2);
      membCinOverride = cins;
      tmapOverride = second ? &c.pool[0].t : &in[ci].t;
      if (!c.nonSpatial.empty()) {
        StageArgsHost pre = a;
        pre.errExpected = 0;
        launchCompartment(ci, pre, c.kSumNs, 0, true);
      }
      launchCompartment(ci, a, second ? c.kRelaxB : c.kRelaxA, 0);
    }
    membCinOverride.clear();
    tmapOverride = nullptr;
    if (expected > 0) {
      ++errSeq;
    }
    if (comm && slabsConnected) {
      // z-slabs: the boundary planes of the new iterate go to the neighbours
      // before anybody reads it (every rank takes the same accept / reject
      // decisions from max-reduced norms, so the ranks' buffers swap in lock step
      // and a buffer's peer mapping stays its counterpart)
      std::vector<const double *> local(comps.size(), nullptr);
      std::vector<std::array<double *, 2>> peer(
          comps.size(), std::array<double *, 2>{nullptr, nullptr});
      for (std::size_t ci = 0; ci < comps.size(); ++ci) {
        const auto &c = comps[ci];
        if (c.zeroStorage.empty() || c.n == 0) {
          continue;
        }
        const auto &out = second ? c.pool[1] : c.pool[0];
        local[ci] = out.p;
        peer[ci] = {out.peer[0], out.peer[1]};
      }
      haloPushPtrs(local, peer, stream);
      haloWait(stream);
    }
  };
  if (comm && slabsConnected) {
    haloQuiesce(); // the stage input's ghosts have arrived, the halo stream is idle
  }
  relaxing = true;
  try {
    constexpr std::size_t maxIterations{100};
    double nextRelaxDt = maxRelaxStableTimestep;
    for (std::size_t iter = 0; iter < maxIterations; ++iter) {
      const double dt = nextRelaxDt;
      launchHalf(false, dt);
      double rabs = 0.0, rrel = 0.0;
      readError(rabs, rrel);
      if (rabs < errMaxAbs && rrel < errMaxRel) {
        break;
      }
      launchHalf(true, dt);
      double eabs = 0.0, erel = 0.0;
      readError(eabs, erel);
      double errFactor = std::min(errMaxAbs / (eabs + 1e-300),
                                  errMaxRel / (erel + 1e-300));
      errFactor = std::sqrt(errFactor);
      nextRelaxDt = std::min(0.95 * dt * errFactor, maxRelaxStableTimestep);
      if (!(eabs > errMaxAbs || erel > errMaxRel)) {
        for (std::size_t ci = 0; ci < comps.size(); ++ci) {
          if (!comps[ci].zeroStorage.empty() && comps[ci].n > 0) {
            std::swap(in[ci], comps[ci].pool[1]); // accepted: c2 is the state
          }
        }
      }
    }
  } catch (...) {
    relaxing = false;
    throw;
  }
  relaxing = false;
  for (std::size_t ci = 0; ci < comps.size(); ++ci) {
    auto &c = comps[ci];
    if (c.zeroStorage.empty() || c.n == 0) {
      continue;
    }
    if (protect) {
      c.pool[2] = in[ci];
      c.stageInOverride = in[ci];
      c.haveOverride = true;
    } else {
      const int idx = c.role[sd.cin];
      c.buf[idx] = in[ci].p;
      c.tmap[idx] = in[ci].t;
      peerBuf[0][ci][idx] = in[ci].peer[0];
      peerBuf[1][ci][idx] = in[ci].peer[1];
    }
  }
}

// cinBuf: index of the buffer every compartment's stage input lives in (the
// roles rotate in lock step, so one index serves all compartments)
void Sim::launchCompartment(std::size_t ci, StageArgsHost &a,
                            const StageKernels &k, int cinBuf,
                            bool nsPrepass, unsigned parts) {
  const auto &drv = DriverApi::get();
  auto &c = comps[ci];
  const bool doMain = (parts & 1u) != 0, doMemb = (parts & 2u) != 0;
  const bool timed = kernelTiming && ci == dominantComp && !nsPrepass &&
                     !relaxing && doMain && !suppressTiming &&
                     timingStep % kernelTimingStride == 0;
  uint32_t nsBlocks = 0; // blocks that wrote partial sums (pre-pass)
  // while a step is captured the kernels of one stage fork onto per-compartment
  // streams (parallel graph branches); otherwise everything is on `stream`
  cudaStream_t sMain = forkMain != nullptr ? forkMain : stream;
  cudaStream_t sMemb = forkMemb != nullptr ? forkMemb : stream;
  if (timed) {
    if (timerUsed == timerEvents.size()) {
      cudaEvent_t e0{}, e1{};
      check(cudaEventCreate(&e0), "cudaEventCreate");
      check(cudaEventCreate(&e1), "cudaEventCreate");
      timerEvents.emplace_back(e0, e1);
    }
    check(cudaEventRecord(timerEvents[timerUsed].first, stream),
          "cudaEventRecord");
  }
  // the first kernel launched behind a final stage that did not count its CTAs
  // publishes that stage's norms before it starts (every kernel of this stage is
  // stream-ordered behind all kernels of the previous one)
  const bool publishHere = errUnpublished && !capturing &&
                           (doMain || (c.nMvox > 0 && doMemb));
  if (publishHere) {
    a.errPublish = 1;
    errUnpublished = false;
  }
  if (!doMain) {
    // membrane kernel only
  } else if (c.geom == Geom::TILE) {
    uint32_t chunk = c.tile.chunk;
    unsigned gridZ = c.tile.grid[2];
    if (c.tile.m3) { // the planes of this launch (a z-slab may be launched in parts)
      gridZ = std::max(1u, (a.z_end - a.z_begin + chunk - 1) / chunk);
    }
    const CUtensorMap *tm = tmapOverride != nullptr ? tmapOverride : &c.tmap[cinBuf];
    a.zig = c.tile.m3 && c.tile.zig ? 1u : 0u;
    void *params[] = {&a, const_cast<CUtensorMap *>(tm), &chunk};
    if (a.zig != 0u) { // chunk index fastest: the chunks of a tile run together
      launchKernel(k.main, gridZ, c.tile.grid[0], c.tile.grid[1], c.tile.dynSmem,
                   sMain, params);
    } else {
      launchKernel(k.main, c.tile.grid[0], c.tile.grid[1],
                   c.tile.m3 ? gridZ : c.tile.grid[2], c.tile.dynSmem, sMain, params);
    }
    nsBlocks = c.tile.grid[0] * c.tile.grid[1] * c.tile.grid[2];
  } else {
    const unsigned grid = mainGrid(ci, a, k.main);
    void *params[] = {&a};
    launchKernel(k.main, grid, 1, 1, c.geom == Geom::SEG ? c.segSmem : 0u, sMain,
                 params);
    nsBlocks = grid;
  }
  if (doMain) {
    ++launchCount;
    a.errPublish = 0; // published by the main kernel: not again by the membrane one
  }
  if (timed) {
    check(cudaEventRecord(timerEvents[timerUsed].second, stream),
          "cudaEventRecord");
    ++timerUsed;
    if (timerUsed >= 4096) {
      flushTimers();
    }
  }
  if (c.nMvox > 0 && doMemb) {
    // SmebMembArgs: 4 pointers, SMEB_NCOMP stage inputs, SMEB_NCOMP member
    // tables (ensembles only), 2 x u32
    std::vector<unsigned long long> blob(4 + 2 * comps.size() + 1, 0ull);
    blob[0] = reinterpret_cast<unsigned long long>(c.dMvox);
    blob[1] = reinterpret_cast<unsigned long long>(c.dMstart);
    blob[2] = reinterpret_cast<unsigned long long>(c.dEnt);
    blob[3] = reinterpret_cast<unsigned long long>(dFluxLen);
    for (std::size_t j = 0; j < comps.size(); ++j) {
      const double *pj = comps[j].S > 0 ? comps[j].buf[cinBuf] : nullptr;
      if (!membCinOverride.empty()) {
        pj = membCinOverride[j];
      }
      blob[4 + j] = reinterpret_cast<unsigned long long>(pj);
    }
    blob[4 + 2 * comps.size()] = static_cast<unsigned long long>(c.nMvox);
    const unsigned mgrid = membGrid(ci, k.memb);
    StageArgsHost am = a;
    if (nsPrepass) { // the membrane voxels' partial sums follow the main ones
      am.nsPartial = a.nsPartial + nsBlocks * c.nonSpatial.size();
      nsBlocks += mgrid;
    }
    void *mparams[] = {&am, blob.data()};
    launchKernel(k.memb, mgrid, 1, 1, 0, sMemb, mparams);
    ++launchCount;
    a.errPublish = 0;
  }
  if (nsPrepass) {
    if (nsBlocks > c.nsPartialBlocks) {
      throw Error("smeb200: internal error (non-spatial partial sums)");
    }
    const double *partial = c.dNsPartial;
    uint32_t nns = static_cast<uint32_t>(c.nonSpatial.size());
    // z-slabs: this rank's sum over its owned voxels (divided by 1), completed
    // below; else the mean over the compartment
    double nVox = nsInSlabs ? 1.0 : static_cast<double>(c.n);
    double *av = c.dNsAv;
    void *fparams[] = {&partial, &nsBlocks, &nns, &nVox, &av};
    check(drv.LaunchKernel(mainModule.fn.at(kNsFinishKernel), 1, 1, 1, BLOCK, 1,
                           1, 0, sMain, fparams, nullptr),
          "launch non-spatial average kernel");
    ++launchCount;
    if (nsInSlabs) {
      // sum of the ranks' sums in rank order (the same bits on every rank),
      // divided by the compartment's voxel count: one small round trip per
      // stage — the stage kernels cannot start before the mean is known anyway
      check(cudaMemcpyAsync(hNs, c.dNsAv, nns * sizeof(double),
                            cudaMemcpyDeviceToHost, sMain),
            "download non-spatial sums");
      check(cudaStreamSynchronize(sMain), "cudaStreamSynchronize");
      for (uint32_t j = 0; j < nns; j += 4) {
        comm->allreduceSumOrdered(hNs + j, static_cast<int>(std::min(4u, nns - j)));
      }
      for (uint32_t j = 0; j < nns; ++j) {
        hNs[j] = hNs[j] / c.nGlobal;
      }
      check(cudaMemcpyAsync(c.dNsAv, hNs, nns * sizeof(double),
                            cudaMemcpyHostToDevice, sMain),
            "upload non-spatial averages");
    }
  }
}

// grid of the main kernel of one launch: the tiles of the launch's planes, or
// one resident wave at most for the grid-stride kernels
unsigned Sim::mainGrid(std::size_t ci, const StageArgsHost &a, CUfunction f) {
  const auto &c = comps[ci];
  if (f == nullptr) {
    return 0;
  }
  if (c.geom == Geom::TILE) {
    unsigned gridZ = c.tile.grid[2];
    if (c.tile.m3) {
      gridZ = std::max(1u, (a.z_end - a.z_begin + c.tile.chunk - 1) / c.tile.chunk);
    }
    return c.tile.grid[0] * c.tile.grid[1] * gridZ;
  }
  if (c.geom == Geom::SEG) {
    return std::max(1u, std::min(c.nSeg, persistentGrid(f, c.segSmem)));
  }
  uint64_t work = a.n > a.v0 ? a.n - a.v0 : 0;
  if (c.geom == Geom::BOX) {
    work = static_cast<uint64_t>(a.z_end - a.z_begin) * c.nx * c.ny;
  }
  return static_cast<unsigned>(std::max<uint64_t>(
      1, std::min<uint64_t>((work + BLOCK - 1) / BLOCK, persistentGrid(f))));
}

unsigned Sim::membGrid(std::size_t ci, CUfunction f) {
  const auto &c = comps[ci];
  if (c.nMvox == 0 || f == nullptr) {
    return 0;
  }
  return std::min((c.nMvox + BLOCK - 1) / BLOCK, persistentGrid(f));
}

void Sim::launchKernel(CUfunction f, unsigned gx, unsigned gy, unsigned gz,
                       unsigned smem, cudaStream_t s, void **params) {
  const auto &drv = DriverApi::get();
  if (!usePdl || capturing) {
    check(drv.LaunchKernel(f, gx, gy, gz, BLOCK, 1, 1, smem, s, params, nullptr),
          "launch stage kernel");
    return;
  }
  CUlaunchAttribute attr{};
  attr.id = CU_LAUNCH_ATTRIBUTE_PROGRAMMATIC_STREAM_SERIALIZATION;
  attr.value.programmaticStreamSerializationAllowed = 1;
  CUlaunchConfig cfg{};
  cfg.gridDimX = gx;
  cfg.gridDimY = gy;
  cfg.gridDimZ = gz;
  cfg.blockDimX = BLOCK;
  cfg.blockDimY = 1;
  cfg.blockDimZ = 1;
  cfg.sharedMemBytes = smem;
  cfg.hStream = s;
  cfg.attrs = &attr;
  cfg.numAttrs = 1;
  check(drv.LaunchKernelEx(&cfg, f, params, nullptr), "launch stage kernel");
}

// Grid-stride kernels are launched as exactly one wave of resident CTAs: with
// more CTAs than fit (e.g. 8 per SM when 6 are resident) the second, partial
// wave runs at a fraction of the occupancy while holding a full share of the
// work (ncu: 53 % achieved warps on smeb_stage_nn before this).
unsigned Sim::persistentGrid(CUfunction f, unsigned dynSmem) {
  if (auto it = persistentGridOf.find(f); it != persistentGridOf.end()) {
    return it->second;
  }
  int perSm = 0;
  const auto &drv = DriverApi::get();
  if (drv.OccupancyMaxActiveBlocksPerMultiprocessor(
          &perSm, f, static_cast<int>(BLOCK), dynSmem) != CUDA_SUCCESS ||
      perSm <= 0) {
    perSm = 4;
  }
  const unsigned g = static_cast<unsigned>(smCount) * static_cast<unsigned>(perSm);
  persistentGridOf[f] = g;
  return g;
}

void Sim::flushTimers() {
  if (timerUsed == 0) {
    return;
  }
  check(cudaStreamSynchronize(stream), "cudaStreamSynchronize");
  for (std::size_t i = 0; i < timerUsed; ++i) {
    float ms = 0;
    if (cudaEventElapsedTime(&ms, timerEvents[i].first,
                             timerEvents[i].second) == cudaSuccess) {
      kernelMs += ms;
      ++kernelLaunchesTimed;
    }
  }
  timerUsed = 0;
}

void Sim::setParams(const double *values, std::size_t n) {
  if (n != params.size()) {
    throw Error("set_params: wrong number of values");
  }
  params.assign(values, values + n);
  if (stream != nullptr && !stepGraphs.empty()) {
    bindDevice();
    check(cudaStreamSynchronize(stream), "cudaStreamSynchronize");
    for (auto &[k, g] : stepGraphs) {
      if (g.exec != nullptr) {
        cudaGraphExecDestroy(g.exec);
      }
    }
    stepGraphs.clear();
  }
}

void Sim::launchAllStages(double dt) {
  for (std::size_t i = 0; i < plan.size(); ++i) {
    launchStage(i, dt);
  }
}

// One cudaGraphLaunch per step instead of (stages x compartments x {main,
// membrane, pre-pass}) kernel launches plus the error-norm copies: what a step
// costs for the small models (ABtoC, liver-simplified ensembles) is launch
// latency, not bandwidth.
void Sim::graphStep(double dt) {
  uint32_t key = 0;
  if (!comps.empty()) {
    for (int b = 0; b < 4; ++b) {
      key = key * 4u + static_cast<uint32_t>(comps[0].role[b]);
    }
  }
  auto it = stepGraphs.find(key);
  if (it == stepGraphs.end()) {
    StepGraph g;
    const uint64_t launches0 = launchCount;
    cudaGraph_t graph{};
    check(cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal),
          "cudaStreamBeginCapture");
    capturing = true;
    try {
      check(cudaMemcpyAsync(dDt, hDt, sizeof(double), cudaMemcpyHostToDevice,
                            stream),
            "capture step size upload");
      launchAllStages(dt);
    } catch (...) {
      capturing = false;
      cudaStreamEndCapture(stream, &graph);
      if (graph != nullptr) {
        cudaGraphDestroy(graph);
      }
      throw;
    }
    capturing = false;
    check(cudaStreamEndCapture(stream, &graph), "cudaStreamEndCapture");
    g.launches = launchCount - launches0;
    launchCount = launches0;
    const cudaError_t e = cudaGraphInstantiate(&g.exec, graph, 0);
    cudaGraphDestroy(graph);
    check(e, "cudaGraphInstantiate");
    for (auto &c : comps) {
      g.last.push_back({c.last, c.lastCinBuf, c.lastFlags});
    }
    it = stepGraphs.emplace(key, std::move(g)).first;
  }
  if (dt != graphDt) {
    // the upload node of an earlier, still running replay reads hDt
    if (integrator == SMEB200_RK101) {
      check(cudaStreamSynchronize(stream), "cudaStreamSynchronize");
    }
    *hDt = dt;
    graphDt = dt;
  }
  check(cudaGraphLaunch(it->second.exec, stream), "cudaGraphLaunch");
  launchCount += it->second.launches;
  if (integrator != SMEB200_RK101) {
    ++errSeq; // the replayed step publishes its norms like a launched one
  }
  for (std::size_t ci = 0; ci < comps.size(); ++ci) {
    auto &c = comps[ci];
    if (c.S == 0 || c.n == 0) {
      continue;
    }
    c.last = it->second.last[ci].last;
    c.lastCinBuf = it->second.last[ci].lastCinBuf;
    c.lastFlags = it->second.last[ci].lastFlags;
    c.haveLast = true;
  }
}

void Sim::doStep(double dt) {
  if (useGraphs && !kernelTiming) {
    graphStep(dt);
  } else {
    launchAllStages(dt);
  }
  ++timingStep;
  ++stateVersion;
  for (auto &c : comps) {
    c.dcdtStale = true;
    c.lowerStale = true;
    c.hostStale = true;
  }
}

// accepted step: the final stage's output buffer becomes the state; the old
// state buffer takes its place in the rotation (no copy — the reference's
// undoRKStep / D2D restore, pixelsim_impl.cpp:759-763, cudapixelsim.cpp:
// 1476-1484, is simply "do not rotate").  The final stage already clamped its
// output (clampNegativeConcentrations, pixelsim.cpp:268).
void Sim::rotateAfterAccept() {
  const int outRole = plan.back().cout;
  for (auto &c : comps) {
    std::swap(c.role[CUR], c.role[outRole]);
  }
}

// Waits until the step with sequence number `seq` has published its error norms
// in the pinned ring.  A kernel that failed never publishes: the streams are
// looked at every few hundred polls.
void Sim::waitError(uint32_t seq, double v[2]) {
  volatile unsigned long long *h = hErrRing + 4 * (seq % kErrRing);
  const auto t0 = std::chrono::steady_clock::now();
  for (unsigned spins = 1;; ++spins) {
    if (static_cast<uint32_t>(h[2]) == seq) {
      break;
    }
    if ((spins & 0x3ffu) == 0) {
      bool idle = true;
      for (cudaStream_t st : {stream, haloStream}) {
        if (st == nullptr) {
          continue;
        }
        const cudaError_t e = cudaStreamQuery(st);
        if (e == cudaErrorNotReady) {
          idle = false;
        } else {
          check(e, "a stage kernel failed");
        }
      }
      if (idle && static_cast<uint32_t>(h[2]) != seq) {
        errDesync = true;
        throw Error("smeb200: internal error (the error norms of a step were "
                    "not published)");
      }
      if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0)
              .count() > 300.0) {
        errDesync = true;
        throw Error("smeb200: timed out waiting for the error norms of a step");
      }
    }
  }
  std::atomic_thread_fence(std::memory_order_acquire);
  unsigned long long bits[2] = {h[0], h[1]};
  std::memcpy(v, bits, 16);
}

// after a failed launch sequence: drain the device, zero the slot (norms, CTA
// count, published count) and the ring, start numbering from zero again
void Sim::resyncErrorSeq() {
  check(cudaStreamSynchronize(stream), "cudaStreamSynchronize");
  if (haloStream != nullptr) {
    check(cudaStreamSynchronize(haloStream), "cudaStreamSynchronize");
  }
  check(cudaMemsetAsync(dErr, 0, 64, stream), "clear error accumulators");
  check(cudaStreamSynchronize(stream), "cudaStreamSynchronize");
  std::memset(hErrRing, 0, 32 * kErrRing);
  errSeq = 0;
  errHead = errTail = 0;
  errDesync = false;
  errUnpublished = false;
}

void Sim::readError(double &eabs, double &erel) {
  double v[2];
  flushErrors();
  waitError(errSeq, v);
  if (allreduceMax != nullptr) {
    allreduceMax(v, 2, allreduceUser);
  } else if (comm) {
    comm->allreduceMax(v, 2);
  }
  eabs = v[0];
  erel = v[1];
}

// ---- controller: pixelsim.cpp:122-140, 207-270, 423-471 ----------------------
// RK101 with cross-diffusion: the forward-Euler bound depends on the state
// (calculateDcdt refreshes it before the step is taken, pixelsim.cpp:122-129,
// pixelsim_impl.cpp:506-543): one max-reduction kernel per compartment with
// cross-diffusion terms, one small D2H, then the reference's formula on the
// host.
void Sim::updateCrossDiffusionMaxStableTimestep() {
  const auto &drv = DriverApi::get();
  std::size_t off = 0;
  for (std::size_t ci = 0; ci < comps.size(); ++ci) {
    auto &c = comps[ci];
    if (c.crossTarget.empty() || c.n == 0) {
      continue;
    }
    const std::size_t nx = c.crossTarget.size();
    check(cudaMemsetAsync(c.dXdMax, 0, nx * sizeof(unsigned long long), stream),
          "clear cross-diffusion maxima");
    StageArgsHost a{};
    a.cin = c.buf[c.role[CUR]];
    // z-slabs: the owned voxels (the ghosts are other ranks' voxels, and their
    // halos may still be on the way)
    a.v0 = static_cast<uint32_t>(c.ownedBegin());
    a.n = static_cast<uint32_t>(c.ownedEnd());
    a.xdMax = c.dXdMax;
    std::copy(params.begin(), params.end(), a.D + 3 * c.S);
    const unsigned grid = static_cast<unsigned>(std::max<uint64_t>(
        1, std::min<uint64_t>((a.n - a.v0 + BLOCK - 1) / BLOCK,
                              static_cast<uint64_t>(smCount) * 8u)));
    void *kparams[] = {&a};
    check(drv.LaunchKernel(c.kXdMax, grid, 1, 1, BLOCK, 1, 1, 0, stream,
                           kparams, nullptr),
          "launch cross-diffusion maximum kernel");
    ++launchCount;
    check(cudaMemcpyAsync(hXdMax + off, c.dXdMax,
                          nx * sizeof(unsigned long long),
                          cudaMemcpyDeviceToHost, stream),
          "download cross-diffusion maxima");
    off += nx;
  }
  check(cudaStreamSynchronize(stream), "cross-diffusion maxima");
  maxStableTimestep = std::numeric_limits<double>::max();
  off = 0;
  for (auto &c : comps) {
    if (!c.crossTarget.empty() && c.n > 0) {
      auto maxEffectiveDiffusion = c.diffusionRaw;
      for (std::size_t t = 0; t < c.crossTarget.size(); ++t) {
        double m = 0.0;
        std::memcpy(&m, hXdMax + off + t, sizeof(double));
        maxEffectiveDiffusion[c.crossTarget[t]] += m;
      }
      off += c.crossTarget.size();
      c.maxStable = std::numeric_limits<double>::max();
      for (uint32_t s = 0; s < c.SP; ++s) {
        const double is = c.invStorage.empty() ? 1.0 : c.invStorage[s];
        const double d = maxEffectiveDiffusion[s];
        if (is == 0.0 || d <= 0.0) {
          continue;
        }
        c.maxStable = std::min(
            c.maxStable,
            1.0 / (2.0 * (d * is / h2[0] + d * is / h2[1] + d * is / h2[2])));
      }
    }
    if (c.S > 0) {
      maxStableTimestep = std::min(maxStableTimestep, c.maxStable);
    }
  }
  syncMaxStableTimestep();
}

// z-slabs: the forward-Euler bound is the minimum over the ranks' voxels
void Sim::syncMaxStableTimestep() {
  if (!comm || !slabsConnected) {
    return;
  }
  double v = -maxStableTimestep; // exact under negation
  comm->allreduceMax(&v, 1);
  maxStableTimestep = -v;
  if (anyZeroStorage) { // the pseudo-time step of the relaxation as well
    double r = -maxRelaxStableTimestep;
    comm->allreduceMax(&r, 1);
    maxRelaxStableTimestep = -r;
  }
}

double Sim::doRK101(double dt) {
  if (anyCross) {
    updateCrossDiffusionMaxStableTimestep();
  }
  dt = std::min(dt, maxStableTimestep);
  doStep(dt);
  rotateAfterAccept();
  return dt;
}

// The error norms of the step just launched will be published under errSeq;
// nothing is queued behind the step and nobody waits here.
void Sim::enqueueErrorRead(double dt, double dtMax) {
  errPending[errTail % kErrRing] = {dt, dtMax, errSeq};
  ++errTail;
}

// Checks queued error norms until at most `keep` are in flight.  Each one goes
// through the controller's own formula (pixelsim.cpp:231-233); with unbounded
// tolerances that yields "accept, next step = dtMax" for every finite norm, which
// is what was assumed when the following steps were launched.  A norm for which
// it does not (infinite: the state blew up) ends the run with the reference's
// message for a collapsed step size.
bool Sim::drainErrors(std::size_t keep) {
  const double errPower = integrator == SMEB200_RK323   ? 1.0 / 3.0
                          : integrator == SMEB200_RK435 ? 1.0 / 4.0
                                                        : 1.0 / 2.0;
  while (errTail - errHead > keep) {
    const std::size_t slot = errHead % kErrRing;
    double v[2];
    if (errPending[slot].seq == errSeq) {
      flushErrors(); // the newest step: no kernel behind it publishes its norms
    }
    waitError(errPending[slot].seq, v);
    ++errHead;
    if (allreduceMax != nullptr) {
      allreduceMax(v, 2, allreduceUser);
    } else if (comm) {
      comm->allreduceMax(v, 2);
    }
    const double dt = errPending[slot].dt, dtMax = errPending[slot].dtMax;
    double errFactor = std::min(errMaxAbs / v[0], errMaxRel / v[1]);
    errFactor = std::pow(errFactor, errPower);
    const double next = std::min(0.95 * dt * errFactor, dtMax);
    if (next != dtMax || v[0] > errMaxAbs || v[1] > errMaxRel) {
      // the steps launched behind the failed one are void
      errHead = errTail;
      check(cudaStreamSynchronize(stream), "cudaStreamSynchronize");
      nextTimestep = next;
      std::string problemSpecies{"unknown"};
      plotRKError(v[1], problemSpecies);
      errorMessage =
          "Failed to solve model to required accuracy. The largest "
          "relative integration error comes from species '" +
          problemSpecies +
          "'. The location "
          "of the pixels with the largest relative integration error are shown "
          "below in red:";
      return false;
    }
  }
  return true;
}

double Sim::doRKAdaptive(double dtMax) {
  double eabs = 0.0, erel = 0.0;
  double dt = 0.0;
  if (laggedError) {
    if (!drainErrors(errLag)) {
      return nextTimestep;
    }
    dt = std::min(nextTimestep, dtMax);
    doStep(dt);
    enqueueErrorRead(dt, dtMax);
    nextTimestep = dtMax; // min(0.95 dt (DBL_MAX / err)^p, dtMax), err finite
    rotateAfterAccept();
    return dt;
  }
  const double errPower = integrator == SMEB200_RK323   ? 1.0 / 3.0
                          : integrator == SMEB200_RK435 ? 1.0 / 4.0
                                                        : 1.0 / 2.0;
  do {
    dt = std::min(nextTimestep, dtMax);
    doStep(dt);
    readError(eabs, erel);
    double errFactor = std::min(errMaxAbs / eabs, errMaxRel / erel);
    errFactor = std::pow(errFactor, errPower);
    nextTimestep = std::min(0.95 * dt * errFactor, dtMax);
    if (nextTimestep / dtMax < 1e-20) {
      std::string problemSpecies{"unknown"};
      plotRKError(erel, problemSpecies);
      errorMessage =
          "Failed to solve model to required accuracy. The largest "
          "relative integration error comes from species '" +
          problemSpecies +
          "'. The location "
          "of the pixels with the largest relative integration error are shown "
          "below in red:";
      return nextTimestep;
    }
    if (eabs > errMaxAbs || erel > errMaxRel) {
      ++discardedSteps;
    }
  } while (eabs > errMaxAbs || erel > errMaxRel);
  rotateAfterAccept();
  return dt;
}

std::size_t Sim::run(double time, double timeoutMs, int (*stopCb)(void *),
                     void *user) {
  if (ctorFailed) {
    return 0; // cudapixelsim.cpp:1621-1623
  }
  errorMessage.clear();
  errorImages.clear();
  std::size_t steps = 0;
  try {
    bindDevice();
    if (runEv0 == nullptr) {
      check(cudaEventCreate(&runEv0), "cudaEventCreate");
      check(cudaEventCreate(&runEv1), "cudaEventCreate");
    }
    if (comm && slabsConnected) {
      // no rank sends halos of a new run before every rank has left the reads
      // of the last one (getters and on-demand kernels read ghost planes)
      comm->barrier();
    }
    awaitDownloadOnStream();
    check(cudaEventRecord(runEv0, stream), "cudaEventRecord");
    const auto t0 = std::chrono::steady_clock::now();
    double tNow = 0;
    discardedSteps = 0;
    constexpr double relativeTolerance = 1e-12;
    errHead = errTail; // nothing is in flight between runs
    if (errDesync) {
      resyncErrorSeq();
    }
    laggedError = integrator != SMEB200_RK101 && !useGraphs && !anyZeroStorage &&
                  errMaxAbs == std::numeric_limits<double>::max() &&
                  errMaxRel == std::numeric_limits<double>::max() &&
                  !envFlagOff("SMEB200_LAGGED_ERROR");
    deferPublish = laggedError && !envFlagOff("SMEB200_DEFER_PUBLISH");
    errLag = comm ? 4 : 2;
    if (const char *env = std::getenv("SMEB200_ERR_LAG");
        env != nullptr && env[0] >= '1' && env[0] <= '6') {
      errLag = static_cast<std::size_t>(env[0] - '0');
    }
    while (tNow + time * relativeTolerance < time) {
      const double maxDt = std::min(maxTimestep, time - tNow);
      if (integrator == SMEB200_RK101) {
        double timestep = std::min(maxDt, maxStableTimestep);
        tNow += doRK101(timestep);
      } else {
        tNow += doRKAdaptive(maxDt);
        if (!errorMessage.empty()) {
          haloQuiesce();
          flushTimers();
          return steps;
        }
      }
      ++steps;
      if (timeoutMs >= 0.0) {
        // the reference compares wall time after each step; with asynchronous
        // RK101 launches the queue is drained first so the clock means the
        // same thing
        if (integrator == SMEB200_RK101) {
          check(cudaStreamSynchronize(stream), "cudaStreamSynchronize");
        }
        const double elapsed =
            std::chrono::duration<double, std::milli>(
                std::chrono::steady_clock::now() - t0)
                .count();
        if (elapsed >= timeoutMs) {
          stopRequested.store(true);
        }
      }
      if (stopCb != nullptr && stopCb(user) != 0) {
        stopRequested.store(true);
      }
      if (comm && slabsConnected) {
        // z-slabs: deadline, callback and setStopRequested (another thread, one
        // rank) are per-rank events, the decision must be common — a rank that
        // stops alone leaves its neighbours waiting for halos that never come.
        // Every rank gets here after every step: one value through the
        // shared-memory segment
        double stop = stopRequested.load() ? 1.0 : 0.0;
        comm->allreduceMax(&stop, 1);
        if (stop > 0.0) {
          stopRequested.store(true);
        }
      }
      if (stopRequested.load()) {
        errorMessage = "Simulation stopped early";
        break;
      }
    }
    if (laggedError && !drainErrors(0)) {
      haloQuiesce();
      flushTimers();
      return steps;
    }
    haloQuiesce();
    check(cudaEventRecord(runEv1, stream), "cudaEventRecord");
    flushTimers();
    check(cudaStreamSynchronize(stream), "cudaStreamSynchronize");
    float ms = 0;
    check(cudaEventElapsedTime(&ms, runEv0, runEv1), "cudaEventElapsedTime");
    lastRunMs = ms;
    checkHaloError();
    if (asyncDownload && steps > 0) {
      startDownload();
    }
  } catch (const std::exception &e) {
    errorMessage = e.what();
    errDesync = true; // launches of an unfinished step may be missing
    abortSlabs();     // the other ranks must not wait for this one
  }
  return steps;
}

double Sim::benchSteps(double dt, std::size_t nSteps) {
  if (ctorFailed) {
    return -1.0;
  }
  try {
    bindDevice();
    cudaEvent_t e0{}, e1{};
    check(cudaEventCreate(&e0), "cudaEventCreate");
    check(cudaEventCreate(&e1), "cudaEventCreate");
    check(cudaStreamSynchronize(stream), "cudaStreamSynchronize");
    deferPublish = !useGraphs;
    check(cudaEventRecord(e0, stream), "cudaEventRecord");
    for (std::size_t i = 0; i < nSteps; ++i) {
      doStep(integrator == SMEB200_RK101 ? std::min(dt, maxStableTimestep)
                                         : dt);
      rotateAfterAccept();
    }
    flushErrors();
    haloQuiesce();
    check(cudaEventRecord(e1, stream), "cudaEventRecord");
    check(cudaEventSynchronize(e1), "cudaEventSynchronize");
    float ms = 0;
    check(cudaEventElapsedTime(&ms, e0, e1), "cudaEventElapsedTime");
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    flushTimers();
    return ms;
  } catch (const std::exception &e) {
    errorMessage = e.what();
    return -1.0;
  }
}

// ---- state getters: pixelsim.cpp:473-490 -------------------------------------
const double *Sim::getConcentrations(std::size_t comp, std::size_t *n) {
  if (comp >= comps.size()) {
    throw Error("compartment index out of range");
  }
  auto &c = comps[comp];
  const std::size_t nv = c.n * c.S;
  if (n != nullptr) {
    *n = nv;
  }
  if (c.downloadPending) {
    finishDownload(); // run() started this copy on the transfer stream
  }
  if (c.hostStale && nv > 0 && !ctorFailed) {
    bindDevice();
    check(cudaMemcpyAsync(c.hostConc, c.buf[c.role[CUR]], nv * sizeof(double),
                          cudaMemcpyDeviceToHost, stream),
          "download concentrations");
    check(cudaStreamSynchronize(stream), "download concentrations");
    c.hostStale = false;
  }
  return c.hostConc;
}

void Sim::finishDownload() {
  check(cudaEventSynchronize(downloadDone), "download concentrations");
  for (auto &o : comps) {
    if (o.downloadPending) {
      o.downloadPending = false;
      // the mirror is current unless steps were taken since the copy started
      o.hostStale = o.hostStale && downloadVersion != stateVersion;
    }
  }
}

// the state of every compartment starts its way to the host mirrors; the compute
// stream is idle (run() has just synchronised it)
void Sim::startDownload() {
  if (xferStream == nullptr) {
    check(cudaStreamCreateWithFlags(&xferStream, cudaStreamNonBlocking),
          "Failed to create the transfer stream");
    check(cudaEventCreateWithFlags(&downloadDone, cudaEventDisableTiming),
          "cudaEventCreate");
  }
  bool any = false;
  for (auto &c : comps) {
    const std::size_t nv = c.n * c.S;
    if (nv == 0 || !c.hostStale) {
      continue;
    }
    check(cudaMemcpyAsync(c.hostConc, c.buf[c.role[CUR]], nv * sizeof(double),
                          cudaMemcpyDeviceToHost, xferStream),
          "download concentrations");
    c.downloadPending = true;
    any = true;
  }
  if (any) {
    check(cudaEventRecord(downloadDone, xferStream), "cudaEventRecord");
    downloadVersion = stateVersion;
  }
}

// a download still in flight reads the current state buffer: whatever is queued
// next on the compute stream (a new run, an upload) starts after it
void Sim::awaitDownloadOnStream() {
  bool pending = false;
  for (const auto &c : comps) {
    pending = pending || c.downloadPending;
  }
  if (pending) {
    check(cudaStreamWaitEvent(stream, downloadDone, 0), "download wait");
  }
}

void Sim::setHostMirror(std::size_t comp, double *buffer, std::size_t n) {
  if (comp >= comps.size()) {
    throw Error("compartment index out of range");
  }
  auto &c = comps[comp];
  if (n != c.n * c.S || (buffer == nullptr && n > 0)) {
    throw Error("setHostMirror: wrong number of values");
  }
  if (ctorFailed || n == 0) {
    return;
  }
  bindDevice();
  if (c.downloadPending) {
    finishDownload();
  }
  check(cudaHostRegister(buffer, n * sizeof(double), cudaHostRegisterPortable),
        "cudaHostRegister(host mirror)");
  if (!c.hostStale) {
    std::memcpy(buffer, c.hostConc, n * sizeof(double));
  }
  if (c.userMirror) {
    cudaHostUnregister(c.hostConc);
  } else {
    cudaFreeHost(c.hostConc);
  }
  c.hostConc = buffer;
  c.userMirror = true;
}

void Sim::setConcentrations(std::size_t comp, const double *data,
                            std::size_t n) {
  if (comp >= comps.size()) {
    throw Error("compartment index out of range");
  }
  auto &c = comps[comp];
  if (n != c.n * c.S) {
    throw Error("setConcentrations: wrong number of values");
  }
  if (n == 0 || ctorFailed) {
    return;
  }
  bindDevice();
  if (c.downloadPending) { // the mirror is the destination of a copy in flight
    finishDownload();
  }
  ++stateVersion;
  // data may be the pinned mirror itself (filled in place by the caller):
  // then the upload is a single pinned H2D copy
  if (data != c.hostConc) {
    std::memcpy(c.hostConc, data, n * sizeof(double));
  }
  check(cudaMemcpyAsync(c.buf[c.role[CUR]], c.hostConc, n * sizeof(double),
                        cudaMemcpyHostToDevice, stream),
        "upload concentrations");
  check(cudaStreamSynchronize(stream), "upload concentrations");
  c.hostStale = false;
}

// re-run the last attempted step's final stage of compartment ci with an
// on-demand output (dc/dt, lower-order solution or unclamped new state)
// written to `out`; the stage's inputs are still intact in the rotation
void Sim::runOnDemand(std::size_t ci, uint32_t flags, double *out) {
  auto &c = comps[ci];
  const StageKernels k = onDemandKernels(ci, flags);
  StageArgsHost a = c.last;
  a.cout = out;
  a.s2out = nullptr;
  a.err = dErrScratch;
  a.errBase = dErrScratch;
  a.errPairs = 1;
  a.errExpected = 0; // the norms of an on-demand launch go nowhere
  const bool timing = kernelTiming;
  kernelTiming = false;
  launchCompartment(ci, a, k, c.lastCinBuf);
  kernelTiming = timing;
}

// Per-species {average, min, max} of the current state, reduced on the device
// (calculateAvgMinMax, simulate.cpp:272-290, runs on one CPU thread over the
// downloaded state after every image interval): 24 bytes per species cross
// PCIe instead of the whole state.  In a z-slab decomposed simulation the
// values cover this rank's owned planes.
void Sim::getAvgMinMax(std::size_t comp, double *out, std::size_t nOut) {
  if (comp >= comps.size()) {
    throw Error("compartment index out of range");
  }
  auto &c = comps[comp];
  if (nOut != 3 * static_cast<std::size_t>(c.SP) || out == nullptr) {
    throw Error("getAvgMinMax: output must hold 3 values per model species");
  }
  if (ctorFailed) {
    throw Error(errorMessage);
  }
  if (c.S == 0 || c.n == 0) {
    return;
  }
  bindDevice();
  const auto &drv = DriverApi::get();
  const std::string name = "smeb_stats<Reac" + std::to_string(comp) + ">";
  const std::string finish = "smeb_stats_finish";
  if (auxFn.find(name) == auxFn.end()) {
    auxModules.push_back(compile({name, finish}));
    for (const auto &[k, f] : auxModules.back().fn) {
      auxFn[k] = f;
    }
  }
  const unsigned blocks = static_cast<unsigned>(smCount) * 8u;
  if (c.dStats == nullptr) {
    c.dStatsPartial = devAlloc<double>(static_cast<std::size_t>(blocks) * 3 * c.SP,
                                       "allocate statistics partials");
    c.dStats = devAlloc<double>(3 * c.SP, "allocate statistics");
  }
  const double *state = c.buf[c.role[CUR]];
  uint32_t i0 = static_cast<uint32_t>(c.ownedBegin());
  uint32_t i1 = static_cast<uint32_t>(c.ownedEnd());
  uint64_t counted = i1 - i0;
  if (counted == 0) {
    return;
  }
  unsigned grid = static_cast<unsigned>(std::max<uint64_t>(
      1, std::min<uint64_t>((counted + BLOCK - 1) / BLOCK, blocks)));
  double *partial = c.dStatsPartial;
  void *p1[] = {&state, &i0, &i1, &partial};
  check(drv.LaunchKernel(auxFn.at(name), grid, 1, 1, BLOCK, 1, 1, 0, stream, p1,
                         nullptr),
        "launch statistics kernel");
  const double *cpartial = c.dStatsPartial;
  uint32_t sp = c.SP;
  double nVox = static_cast<double>(counted);
  double *dout = c.dStats;
  void *p2[] = {&cpartial, &grid, &sp, &nVox, &dout};
  check(drv.LaunchKernel(auxFn.at(finish), 1, 1, 1, BLOCK, 1, 1, 0, stream, p2,
                         nullptr),
        "launch statistics finish kernel");
  launchCount += 2;
  check(cudaMemcpyAsync(out, c.dStats, nOut * sizeof(double),
                        cudaMemcpyDeviceToHost, stream),
        "download statistics");
  check(cudaStreamSynchronize(stream), "download statistics");
}

// a decimated image of one species leaves the device instead
// of the whole state (Simulation::updateConcentrations, simulate.cpp:313-314)
const double *Sim::getSpeciesDecimated(std::size_t comp, std::size_t species,
                                       const uint32_t stride[3],
                                       std::size_t dims[3]) {
  if (comp >= comps.size()) {
    throw Error("compartment index out of range");
  }
  if (ctorFailed) {
    throw Error(errorMessage);
  }
  auto &c = comps[comp];
  if (species >= c.S) {
    throw Error("getSpeciesDecimated: species index out of range");
  }
  uint32_t sx = stride[0], sy = stride[1], sz = stride[2];
  if (sx == 0 || sy == 0 || sz == 0) {
    throw Error("getSpeciesDecimated: strides must be positive");
  }
  uint32_t nx = c.nx, ny = c.ny, z0 = c.zLoGhost;
  uint64_t nzOwned = c.nz - c.zLoGhost - c.zHiGhost;
  std::size_t voxelOffset = 0;
  if (!c.box) {
    // a voxel list is one row: every sx-th voxel of the owned range
    if (sy != 1 || sz != 1) {
      throw Error("getSpeciesDecimated: a voxel-list compartment takes one "
                  "stride (stride[1] = stride[2] = 1)");
    }
    voxelOffset = c.ownedBegin();
    ny = static_cast<uint32_t>(c.ownedEnd() - c.ownedBegin());
    nx = 1;
    nzOwned = ny > 0 ? 1 : 0;
    z0 = 0;
    sy = sx;
    sx = 1;
  }
  uint32_t mx = (nx + sx - 1) / sx, my = (ny + sy - 1) / sy;
  const uint64_t mz = (nzOwned + sz - 1) / sz;
  const uint64_t nOut = mz * mx * my;
  if (c.box) {
    dims[0] = mx;
    dims[1] = my;
    dims[2] = mz;
  } else {
    dims[0] = nOut;
    dims[1] = dims[2] = 1;
  }
  if (nOut == 0) {
    return hGather;
  }
  bindDevice();
  if (nOut > gatherCapacity) {
    check(cudaStreamSynchronize(stream), "cudaStreamSynchronize");
    cudaFree(dGather);
    cudaFreeHost(hGather);
    dGather = hGather = nullptr;
    gatherCapacity = 0;
    dGather = devAlloc<double>(nOut, "allocate decimated image");
    check(cudaMallocHost(reinterpret_cast<void **>(&hGather), nOut * sizeof(double)),
          "allocate pinned decimated image");
    gatherCapacity = nOut;
  }
  const std::string name = "smeb_gather_species";
  if (auxFn.find(name) == auxFn.end()) {
    auxModules.push_back(compile({name}));
    for (const auto &[k, f] : auxModules.back().fn) {
      auxFn[k] = f;
    }
  }
  const auto &drv = DriverApi::get();
  const double *state = c.buf[c.role[CUR]] + voxelOffset * c.S;
  double *dout = dGather;
  uint32_t S = c.S, s = static_cast<uint32_t>(species);
  unsigned long long n = nOut;
  const unsigned grid = static_cast<unsigned>(std::max<uint64_t>(
      1, std::min<uint64_t>((nOut + BLOCK - 1) / BLOCK,
                            static_cast<uint64_t>(smCount) * 8u)));
  void *params[] = {&state, &dout, &S, &s,  &nx, &ny, &z0,
                    &sx,    &sy,   &sz, &mx, &my, &n};
  check(drv.LaunchKernel(auxFn.at(name), grid, 1, 1, BLOCK, 1, 1, 0, stream, params,
                         nullptr),
        "launch gather kernel");
  ++launchCount;
  check(cudaMemcpyAsync(hGather, dGather, nOut * sizeof(double),
                        cudaMemcpyDeviceToHost, stream),
        "download decimated image");
  check(cudaStreamSynchronize(stream), "download decimated image");
  return hGather;
}

// dc/dt of the last evaluated stage (PixelSim::getDcdt).  The fused stages
// never store dc/dt; it is re-evaluated on demand from the retained input of
// the last stage.
void Sim::evaluateLastDcdt() {
  bindDevice();
  for (std::size_t ci = 0; ci < comps.size(); ++ci) {
    auto &c = comps[ci];
    if (c.S == 0 || c.n == 0 || !c.haveLast) {
      continue;
    }
    const std::size_t nv = c.n * c.S;
    if (c.scratch == nullptr) {
      c.scratch = devAlloc<double>(nv, "allocate scratch buffer");
    }
    runOnDemand(ci, F_OUT_DCDT, c.scratch);
    c.hostDcdt.resize(nv);
    check(cudaMemcpyAsync(c.hostDcdt.data(), c.scratch, nv * sizeof(double),
                          cudaMemcpyDeviceToHost, stream),
          "download dcdt");
    c.dcdtStale = false;
  }
  check(cudaStreamSynchronize(stream), "download dcdt");
}

const double *Sim::getDcdt(std::size_t comp, std::size_t *n) {
  if (comp >= comps.size()) {
    throw Error("compartment index out of range");
  }
  auto &c = comps[comp];
  if (n != nullptr) {
    *n = c.n * c.S;
  }
  if (c.dcdtStale && !ctorFailed) {
    evaluateLastDcdt();
  }
  return c.hostDcdt.data();
}

double Sim::getLowerOrder(std::size_t comp, std::size_t species,
                          std::size_t pixel) {
  if (comp >= comps.size()) {
    throw Error("compartment index out of range");
  }
  auto &c = comps[comp];
  // reference: s2 is empty until an adaptive step has run -> 0
  // (pixelsim_impl.hpp getLowerOrderConcentration)
  if (!c.haveLast || integrator == SMEB200_RK101 || ctorFailed) {
    return 0.0;
  }
  const std::size_t nv = c.n * c.S;
  if (species >= c.S || pixel >= c.n) {
    throw Error("getLowerOrderConcentration: index out of range");
  }
  if (c.lowerStale) {
    bindDevice();
    if (c.scratch == nullptr) {
      c.scratch = devAlloc<double>(nv, "allocate scratch buffer");
    }
    runOnDemand(comp, c.lastFlags | F_OUT_LOWER, c.scratch);
    c.hostLower.resize(nv);
    check(cudaMemcpyAsync(c.hostLower.data(), c.scratch, nv * sizeof(double),
                          cudaMemcpyDeviceToHost, stream),
          "download lower-order concentrations");
    check(cudaStreamSynchronize(stream), "download lower-order");
    c.lowerStale = false;
  }
  return c.hostLower[pixel * c.S + species];
}

// red-channel image of the relative error, pixelsim_impl.cpp:794-824
void Sim::plotRKError(double relMax, std::string &problemSpecies) {
  errorImages.assign(static_cast<std::size_t>(std::max(imageNx, 0)) *
                         std::max(imageNy, 0) * std::max(imageNz, 0),
                     0u);
  for (std::size_t ci = 0; ci < comps.size(); ++ci) {
    auto &c = comps[ci];
    if (c.S == 0 || c.n == 0 || !c.haveLast) {
      continue;
    }
    const std::size_t nv = c.n * c.S;
    // lower-order solution, unclamped new state and old state of the failed
    // step
    c.lowerStale = true;
    getLowerOrder(ci, 0, 0);
    std::vector<double> cnew(nv), cold(nv);
    runOnDemand(ci, (c.lastFlags & ~F_FINAL) | F_OUT_CNEW, c.scratch);
    check(cudaMemcpyAsync(cnew.data(), c.scratch, nv * sizeof(double),
                          cudaMemcpyDeviceToHost, stream),
          "download state");
    check(cudaMemcpyAsync(cold.data(), c.buf[c.role[CUR]], nv * sizeof(double),
                          cudaMemcpyDeviceToHost, stream),
          "download state");
    check(cudaStreamSynchronize(stream), "download state");
    std::size_t iSpecies = c.S + 1;
    for (std::size_t ix = 0; ix < c.n; ++ix) {
      for (std::size_t is = 0; is < c.S; ++is) {
        const std::size_t i = ix * c.S + is;
        const double localErr = std::abs(cnew[i] - c.hostLower[i]);
        const double localNorm = 0.5 * (cnew[i] + cold[i] + epsilon);
        const double pixelIntensity = localErr / localNorm / relMax;
        const int red = static_cast<int>(255.0 * pixelIntensity);
        int x = 0, y = 0, z = 0;
        if (!c.voxels.empty()) {
          x = c.voxels[3 * ix];
          y = c.voxels[3 * ix + 1];
          z = c.voxels[3 * ix + 2];
        } else if (c.box) {
          y = static_cast<int>(ix % c.ny);
          x = static_cast<int>((ix / c.ny) % c.nx);
          z = static_cast<int>(ix / (static_cast<std::size_t>(c.ny) * c.nx));
        } else {
          continue;
        }
        if (x < 0 || y < 0 || z < 0 || x >= imageNx || y >= imageNy ||
            z >= imageNz) {
          continue;
        }
        auto &px = errorImages[(static_cast<std::size_t>(z) * imageNy + y) *
                                   imageNx +
                               x];
        const int oldRed = static_cast<int>((px >> 16) & 0xffu);
        if (red > oldRed) {
          px = 0xff000000u | (static_cast<uint32_t>(red & 0xff) << 16);
          if (red > 254) {
            iSpecies = is;
          }
        }
      }
    }
    if (iSpecies < c.S && !c.speciesNames[iSpecies].empty()) {
      problemSpecies = c.speciesNames[iSpecies];
    }
  }
}

} // namespace smeb200
