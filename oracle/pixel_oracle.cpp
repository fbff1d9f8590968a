// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
//
// CPU restatement of the reference Pixel solver (spatial-model-editor
// core/simulate/src/pixelsim.cpp, pixelsim_impl.cpp, pixelsim_common.hpp,
// pixelsim_base.hpp) used as the parity oracle for the CUDA path and as the
// timed "port" CPU baseline of bench.py.  Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may load this.
//
// The reference itself cannot be built here (needs Qt6, SymEngine+LLVM,
// libSBML, oneTBB, ...), so this file restates its algorithm: the same passes
// over the same arrays (conc, dcdt, s2, s3 as AoS [voxel][species] fp64,
// size_t neighbour table), in the same floating point association order, with
// the same controller.  The JIT-compiled reaction function of the reference
// (Symbolic::eval -> LLVM) is replaced by the same expressions compiled by
// gcc -O2 -ffp-contract=off and dlopen'ed.  Threading: OpenMP static chunks
// stand in for TBB static_partitioner (pixelsim_impl.cpp:26-33).
//
// Parity pin: PINNED.  tests/test_golden_cpu.py checks this oracle against the
// reference's own known-answer tests (simulate_t.cpp:18-275, 301-324, 367-380,
// 882-994, 1596-1647; pixelsim_t.cpp:12-20) on fixtures extracted from the
// reference's SBML example models (tests/golden/, README.md there).
#include "../include/smeb200.h"
#include <algorithm>
#include <array>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <fstream>
#include <limits>
#include <memory>
#include <sstream>
#include <string>
#include <unistd.h>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

using ReacFn = void (*)(double *out, const double *in);

// ---- SimCompartment (pixelsim_impl.hpp:58-322) -------------------------------
struct OCompartment {
  std::string id;
  std::size_t nPixels{0};
  std::size_t nSpecies{0};
  std::vector<std::size_t> nn; // [6*nPixels]: +x,-x,+y,-y,+z,-z
  std::vector<std::array<double, 3>> diff; // diffConstantsUniform
  std::size_t nPrimarySpecies{0};          // pixelsim_impl.cpp:205
  std::vector<std::size_t> nonSpatialSpeciesIndices;
  std::vector<double> invStorage;
  bool hasNonUnitStorage{false};
  // zero-storage (algebraic) species, pixelsim_impl.cpp:218-224,826-891
  bool hasZeroStorageSpecies{false};
  std::vector<std::size_t> zeroStorageSpeciesIndices;
  std::vector<double> relaxOld, relaxFirstOrder;
  double maxRelaxStableTimestep{std::numeric_limits<double>::max()};
  // non-uniform diffusion: diffConstants[s][i] (pixelsim_impl.cpp:257-296)
  bool useUniformDiffusionOperator{true};
  std::vector<std::vector<double>> diffConstants;
  double dx2{1}, dy2{1}, dz2{1};
  // cross-diffusion (pixelsim_impl.cpp:350-389)
  struct CrossTerm {
    std::size_t targetSpeciesIndex, sourceSpeciesIndex;
  };
  std::vector<CrossTerm> crossDiffusionTerms;
  std::vector<double> crossDiffusionCoefficients; // [nPixels * nTerms]
  std::vector<double> maxPrimaryDiagonalDiffusion;
  ReacFn crossFn{nullptr};
  bool hasCrossDiffusion{false};
  std::vector<double> conc, dcdt, s2, s3;
  std::vector<int32_t> voxels;
  std::vector<std::string> speciesNames;
  ReacFn reac{nullptr};
  double maxStableTimestep{std::numeric_limits<double>::max()};

  // pixelsim_impl.cpp:488-492
  void evaluateReactions(std::size_t begin, std::size_t end) {
    for (std::size_t i = begin; i < end; ++i) {
      reac(dcdt.data() + i * nSpecies, conc.data() + i * nSpecies);
    }
  }
  // pixelsim_impl.cpp:453-485 (array branch)
  void evaluateDiffusionOperatorArray(std::size_t begin, std::size_t end) {
    for (std::size_t i = begin; i < end; ++i) {
      const std::size_t ix{i * nSpecies};
      const std::size_t ix_upx{nn[6 * i] * nSpecies};
      const std::size_t ix_dnx{nn[6 * i + 1] * nSpecies};
      const std::size_t ix_upy{nn[6 * i + 2] * nSpecies};
      const std::size_t ix_dny{nn[6 * i + 3] * nSpecies};
      const std::size_t ix_upz{nn[6 * i + 4] * nSpecies};
      const std::size_t ix_dnz{nn[6 * i + 5] * nSpecies};
      for (std::size_t is = 0; is < nSpecies; ++is) {
        const auto &d = diffConstants[is];
        double d_i = d[i];
        double d_upx = d[nn[6 * i]];
        double d_dnx = d[nn[6 * i + 1]];
        double d_upy = d[nn[6 * i + 2]];
        double d_dny = d[nn[6 * i + 3]];
        double d_upz = d[nn[6 * i + 4]];
        double d_dnz = d[nn[6 * i + 5]];
        double dxp = 0.5 * (d_i + d_upx) / dx2;
        double dxm = 0.5 * (d_i + d_dnx) / dx2;
        double dyp = 0.5 * (d_i + d_upy) / dy2;
        double dym = 0.5 * (d_i + d_dny) / dy2;
        double dzp = 0.5 * (d_i + d_upz) / dz2;
        double dzm = 0.5 * (d_i + d_dnz) / dz2;
        dcdt[ix + is] += dxp * (conc[ix_upx + is] - conc[ix + is]) -
                         dxm * (conc[ix + is] - conc[ix_dnx + is]) +
                         dyp * (conc[ix_upy + is] - conc[ix + is]) -
                         dym * (conc[ix + is] - conc[ix_dny + is]) +
                         dzp * (conc[ix_upz + is] - conc[ix + is]) -
                         dzm * (conc[ix + is] - conc[ix_dnz + is]);
      }
    }
  }
  // pixelsim_impl.cpp:494-504
  void evaluateCrossDiffusionCoefficients(std::size_t begin, std::size_t end) {
    const auto nTerms{crossDiffusionTerms.size()};
    for (std::size_t i = begin; i < end; ++i) {
      crossFn(crossDiffusionCoefficients.data() + i * nTerms,
              conc.data() + i * nSpecies);
    }
  }
  // pixelsim_impl.cpp:506-543
  void updateCrossDiffusionMaxStableTimestep() {
    if (!hasCrossDiffusion || crossDiffusionTerms.empty()) {
      return;
    }
    const auto nTerms{crossDiffusionTerms.size()};
    std::vector<double> maxAbsCrossDiffusionByTerm(nTerms, 0.0);
    for (std::size_t ix = 0; ix < nPixels; ++ix) {
      const auto iBase{ix * nTerms};
      for (std::size_t iTerm = 0; iTerm < nTerms; ++iTerm) {
        maxAbsCrossDiffusionByTerm[iTerm] =
            std::max(maxAbsCrossDiffusionByTerm[iTerm],
                     std::abs(crossDiffusionCoefficients[iBase + iTerm]));
      }
    }
    auto maxEffectiveDiffusion{maxPrimaryDiagonalDiffusion};
    for (std::size_t iTerm = 0; iTerm < nTerms; ++iTerm) {
      const auto &term{crossDiffusionTerms[iTerm]};
      maxEffectiveDiffusion[term.targetSpeciesIndex] +=
          maxAbsCrossDiffusionByTerm[iTerm];
    }
    maxStableTimestep = std::numeric_limits<double>::max();
    for (std::size_t iTarget = 0; iTarget < nPrimarySpecies; ++iTarget) {
      const double invStorageValue{invStorage[iTarget]};
      if (invStorageValue == 0.0) {
        continue;
      }
      const double maxEffectiveD{maxEffectiveDiffusion[iTarget]};
      if (maxEffectiveD <= 0.0) {
        continue;
      }
      maxStableTimestep = std::min(
          maxStableTimestep,
          1.0 / (2.0 * (maxEffectiveD * invStorageValue / dx2 +
                        maxEffectiveD * invStorageValue / dy2 +
                        maxEffectiveD * invStorageValue / dz2)));
    }
  }
  // pixelsim_impl.cpp:545-589
  void evaluateCrossDiffusionOperator(std::size_t begin, std::size_t end) {
    const auto nTerms{crossDiffusionTerms.size()};
    for (std::size_t i = begin; i < end; ++i) {
      const std::size_t ix{i * nSpecies};
      const std::size_t iupx{nn[6 * i]}, idnx{nn[6 * i + 1]};
      const std::size_t iupy{nn[6 * i + 2]}, idny{nn[6 * i + 3]};
      const std::size_t iupz{nn[6 * i + 4]}, idnz{nn[6 * i + 5]};
      const std::size_t ixUpx{iupx * nSpecies}, ixDnx{idnx * nSpecies};
      const std::size_t ixUpy{iupy * nSpecies}, ixDny{idny * nSpecies};
      const std::size_t ixUpz{iupz * nSpecies}, ixDnz{idnz * nSpecies};
      for (std::size_t iTerm = 0; iTerm < nTerms; ++iTerm) {
        const auto &term{crossDiffusionTerms[iTerm]};
        const std::size_t target{term.targetSpeciesIndex};
        const std::size_t source{term.sourceSpeciesIndex};
        const double dCenter{crossDiffusionCoefficients[i * nTerms + iTerm]};
        const double dUpx{crossDiffusionCoefficients[iupx * nTerms + iTerm]};
        const double dDnx{crossDiffusionCoefficients[idnx * nTerms + iTerm]};
        const double dUpy{crossDiffusionCoefficients[iupy * nTerms + iTerm]};
        const double dDny{crossDiffusionCoefficients[idny * nTerms + iTerm]};
        const double dUpz{crossDiffusionCoefficients[iupz * nTerms + iTerm]};
        const double dDnz{crossDiffusionCoefficients[idnz * nTerms + iTerm]};
        const double dxp{0.5 * (dCenter + dUpx) / dx2};
        const double dxm{0.5 * (dCenter + dDnx) / dx2};
        const double dyp{0.5 * (dCenter + dUpy) / dy2};
        const double dym{0.5 * (dCenter + dDny) / dy2};
        const double dzp{0.5 * (dCenter + dUpz) / dz2};
        const double dzm{0.5 * (dCenter + dDnz) / dz2};
        dcdt[ix + target] += dxp * (conc[ixUpx + source] - conc[ix + source]) -
                             dxm * (conc[ix + source] - conc[ixDnx + source]) +
                             dyp * (conc[ixUpy + source] - conc[ix + source]) -
                             dym * (conc[ix + source] - conc[ixDny + source]) +
                             dzp * (conc[ixUpz + source] - conc[ix + source]) -
                             dzm * (conc[ix + source] - conc[ixDnz + source]);
      }
    }
  }
  // pixelsim_impl.cpp:834-891
  void doRelaxSubstep1(double dt) {
    for (std::size_t is : zeroStorageSpeciesIndices) {
      for (std::size_t ix = 0; ix < nPixels; ++ix) {
        std::size_t i = ix * nSpecies + is;
        relaxOld[i] = conc[i];
        conc[i] += dt * dcdt[i];
      }
    }
  }
  void doRelaxSubstep2(double dt) {
    for (std::size_t is : zeroStorageSpeciesIndices) {
      for (std::size_t ix = 0; ix < nPixels; ++ix) {
        std::size_t i = ix * nSpecies + is;
        relaxFirstOrder[i] = conc[i];
        conc[i] = 0.5 * relaxOld[i] + 0.5 * conc[i] + 0.5 * dt * dcdt[i];
      }
    }
  }
  void undoRelaxStep() {
    for (std::size_t is : zeroStorageSpeciesIndices) {
      for (std::size_t ix = 0; ix < nPixels; ++ix) {
        std::size_t i = ix * nSpecies + is;
        conc[i] = relaxOld[i];
      }
    }
  }
  void calculateRelaxError(double epsilon, double &eabs, double &erel) const {
    for (std::size_t is : zeroStorageSpeciesIndices) {
      for (std::size_t ix = 0; ix < nPixels; ++ix) {
        std::size_t i = ix * nSpecies + is;
        double localErr = std::abs(conc[i] - relaxFirstOrder[i]);
        eabs = std::max(eabs, localErr);
        double localNorm = 0.5 * (conc[i] + relaxOld[i] + epsilon);
        erel = std::max(erel, localErr / localNorm);
      }
    }
  }
  void getZeroStorageResidual(double epsilon, double &rabs, double &rrel) const {
    for (std::size_t is : zeroStorageSpeciesIndices) {
      for (std::size_t ix = 0; ix < nPixels; ++ix) {
        std::size_t i = ix * nSpecies + is;
        double absRes = std::abs(dcdt[i]);
        rabs = std::max(rabs, absRes);
        double norm = std::abs(conc[i]) + epsilon;
        rrel = std::max(rrel, absRes / norm);
      }
    }
  }
  // pixelsim_impl.cpp:434-452 (uniform branch)
  void evaluateDiffusionOperator(std::size_t begin, std::size_t end) {
    if (!useUniformDiffusionOperator) {
      evaluateDiffusionOperatorArray(begin, end);
      return;
    }
    for (std::size_t i = begin; i < end; ++i) {
      const std::size_t ix = i * nSpecies;
      const std::size_t upx = nn[6 * i] * nSpecies;
      const std::size_t dnx = nn[6 * i + 1] * nSpecies;
      const std::size_t upy = nn[6 * i + 2] * nSpecies;
      const std::size_t dny = nn[6 * i + 3] * nSpecies;
      const std::size_t upz = nn[6 * i + 4] * nSpecies;
      const std::size_t dnz = nn[6 * i + 5] * nSpecies;
      for (std::size_t is = 0; is < nSpecies; ++is) {
        dcdt[ix + is] +=
            diff[is][0] * (conc[upx + is] + conc[dnx + is] - 2.0 * conc[ix + is]) +
            diff[is][1] * (conc[upy + is] + conc[dny + is] - 2.0 * conc[ix + is]) +
            diff[is][2] * (conc[upz + is] + conc[dnz + is] - 2.0 * conc[ix + is]);
      }
    }
  }
};

// pixelsim_impl.cpp:150-163 (serial in the reference)
inline void spatiallyAverageDcdt(OCompartment &c) {
  for (std::size_t is : c.nonSpatialSpeciesIndices) {
    double av = 0;
    for (std::size_t ix = 0; ix < c.nPixels; ++ix) {
      av += c.dcdt[ix * c.nSpecies + is];
    }
    av /= static_cast<double>(c.nPixels);
    for (std::size_t ix = 0; ix < c.nPixels; ++ix) {
      c.dcdt[ix * c.nSpecies + is] = av;
    }
  }
}

// pixelsim_impl.cpp:165-189
inline void applyStorage(OCompartment &c, std::size_t begin, std::size_t end) {
  for (std::size_t i = begin; i < end; ++i) {
    const std::size_t ix{i * c.nSpecies};
    for (std::size_t is = 0; is < c.nSpecies; ++is) {
      c.dcdt[ix + is] *= c.invStorage[is];
    }
  }
}

template <typename Body>
void parallelFor(bool threads, std::size_t n, const Body &body) {
  if (!threads) {
    body(0, n);
    return;
  }
#ifdef _OPENMP
#pragma omp parallel
  {
    const std::size_t nt = static_cast<std::size_t>(omp_get_num_threads());
    const std::size_t t = static_cast<std::size_t>(omp_get_thread_num());
    const std::size_t chunk = (n + nt - 1) / nt;
    const std::size_t b = std::min(n, t * chunk);
    const std::size_t e = std::min(n, b + chunk);
    if (b < e) {
      body(b, e);
    }
  }
#else
  body(0, n);
#endif
}

// ---- SimMembrane (pixelsim_impl.cpp:929-1047) --------------------------------
struct OMembrane {
  int compA{-1}, compB{-1};
  std::vector<std::pair<std::size_t, std::size_t>> pairs[6];
  double fluxLength[6]{1, 1, 1, 1, 1, 1};
  ReacFn reac{nullptr};
};

struct Oracle {
  std::vector<OCompartment> comps;
  std::vector<OMembrane> membs;
  int integrator{SMEB200_RK212};
  double errMaxAbs{std::numeric_limits<double>::max()}, errMaxRel{0.005};
  double maxTimestep{std::numeric_limits<double>::max()};
  // pixelsim_base.hpp:40-49
  double maxStableTimestep{std::numeric_limits<double>::max()};
  std::size_t discardedSteps{0};
  double nextTimestep{1e-7};
  double epsilon{1e-14};
  std::string errorMessage;
  std::vector<uint32_t> errorImages;
  std::atomic<bool> stopRequested{false};
  bool useThreads{false};
  int imageNx{0}, imageNy{0}, imageNz{0};
  void *dl{nullptr};
  std::string soPath;

  ~Oracle() {
    if (dl != nullptr) {
      dlclose(dl);
    }
    if (!soPath.empty()) {
      unlink(soPath.c_str());
    }
  }

  // pixelsim_impl.cpp:80-113
  void evaluateMembranePairRange(OMembrane &m, int dir, std::size_t begin,
                                 std::size_t end) {
    OCompartment *A = m.compA >= 0 ? &comps[m.compA] : nullptr;
    OCompartment *B = m.compB >= 0 ? &comps[m.compB] : nullptr;
    // strides include the time/space padding, nA / nB are the model species
    const std::size_t strideA = A != nullptr ? A->nSpecies : 0;
    const std::size_t strideB = B != nullptr ? B->nSpecies : 0;
    const std::size_t nA = A != nullptr ? A->nPrimarySpecies : 0;
    const std::size_t nB = B != nullptr ? B->nPrimarySpecies : 0;
    const std::size_t nExtraVars =
        B != nullptr ? strideB - nB : (A != nullptr ? strideA - nA : 0);
    std::vector<double> species(nA + nB + nExtraVars, 0.0),
        result(nA + nB + nExtraVars, 0.0);
    const double fluxLength = m.fluxLength[dir];
    for (std::size_t i = begin; i < end; ++i) {
      const auto [ixA, ixB] = m.pairs[dir][i];
      if (A != nullptr) {
        std::copy_n(&A->conc[ixA * strideA], nA, &species[0]);
      }
      if (B != nullptr) {
        std::copy_n(&B->conc[ixB * strideB], strideB, &species[nA]);
      } else if (A != nullptr) {
        std::copy_n(&A->conc[ixA * strideA + nA], nExtraVars, &species[nA]);
      }
      m.reac(result.data(), species.data());
      for (std::size_t is = 0; is < nA; ++is) {
        A->dcdt[ixA * strideA + is] += result[is] / fluxLength;
      }
      for (std::size_t is = 0; is < nB; ++is) {
        B->dcdt[ixB * strideB + is] += result[is + nA] / fluxLength;
      }
    }
  }

  bool hasAnyZeroStorageSpecies{false};
  double maxRelaxStableTimestep{std::numeric_limits<double>::max()};

  // pixelsim.cpp:23-90: pseudo-time relaxation of the zero-storage species to
  // 0 = R(c) + div(D grad c) with an adaptive RK212 iteration
  void solveZeroStorageConstraints() {
    if (!hasAnyZeroStorageSpecies) {
      return;
    }
    constexpr std::size_t maxIterations{100};
    double nextRelaxDt = maxRelaxStableTimestep;
    auto computeResidual = [this]() {
      evaluateCompartmentsAndMembranes();
      for (auto &c : comps) {
        spatiallyAverageDcdt(c);
      }
    };
    for (std::size_t iter = 0; iter < maxIterations; ++iter) {
      computeResidual();
      double rabs = 0.0, rrel = 0.0;
      for (const auto &c : comps) {
        c.getZeroStorageResidual(epsilon, rabs, rrel);
      }
      if (rabs < errMaxAbs && rrel < errMaxRel) {
        return;
      }
      double dt = nextRelaxDt;
      for (auto &c : comps) {
        c.doRelaxSubstep1(dt);
      }
      computeResidual();
      for (auto &c : comps) {
        c.doRelaxSubstep2(dt);
      }
      double eabs = 0.0, erel = 0.0;
      for (const auto &c : comps) {
        c.calculateRelaxError(epsilon, eabs, erel);
      }
      double errFactor = std::min(errMaxAbs / (eabs + 1e-300),
                                  errMaxRel / (erel + 1e-300));
      errFactor = std::sqrt(errFactor);
      nextRelaxDt = std::min(0.95 * dt * errFactor, maxRelaxStableTimestep);
      if (eabs > errMaxAbs || erel > errMaxRel) {
        for (auto &c : comps) {
          c.undoRelaxStep();
        }
      }
    }
  }

  // pixelsim.cpp:92-120
  void calculateDcdt() {
    evaluateCompartmentsAndMembranes();
    for (auto &c : comps) {
      spatiallyAverageDcdt(c);
      if (c.hasNonUnitStorage) {
        parallelFor(useThreads, c.nPixels, [&](std::size_t b, std::size_t e) {
          applyStorage(c, b, e);
        });
      }
    }
  }

  void evaluateCompartmentsAndMembranes() {
    maxStableTimestep = std::numeric_limits<double>::max();
    for (auto &c : comps) {
      parallelFor(useThreads, c.nPixels, [&](std::size_t b, std::size_t e) {
        c.evaluateReactions(b, e);
      });
      if (c.hasCrossDiffusion) { // pixelsim_impl.cpp:591-599
        parallelFor(useThreads, c.nPixels, [&](std::size_t b, std::size_t e) {
          c.evaluateCrossDiffusionCoefficients(b, e);
        });
        c.updateCrossDiffusionMaxStableTimestep();
      }
      parallelFor(useThreads, c.nPixels, [&](std::size_t b, std::size_t e) {
        c.evaluateDiffusionOperator(b, e);
      });
      if (c.hasCrossDiffusion) {
        parallelFor(useThreads, c.nPixels, [&](std::size_t b, std::size_t e) {
          c.evaluateCrossDiffusionOperator(b, e);
        });
      }
      maxStableTimestep = std::min(maxStableTimestep, c.maxStableTimestep);
    }
    for (auto &m : membs) {
      for (int dir = 0; dir < 6; ++dir) {
        // parallel only within one signed face direction
        // (pixelsim_impl.cpp:1027-1047)
        parallelFor(useThreads, m.pairs[dir].size(),
                    [&](std::size_t b, std::size_t e) {
                      evaluateMembranePairRange(m, dir, b, e);
                    });
      }
    }
  }

  template <typename F> void forValues(OCompartment &c, const F &f) {
    parallelFor(useThreads, c.conc.size(), [&](std::size_t b, std::size_t e) {
      for (std::size_t i = b; i < e; ++i) {
        f(i);
      }
    });
  }

  // pixelsim_impl.cpp:643-665
  void clampNegative(OCompartment &c) {
    parallelFor(useThreads, c.nPixels, [&](std::size_t b, std::size_t e) {
      for (std::size_t ix = b; ix < e; ++ix) {
        for (std::size_t is = 0; is < c.nPrimarySpecies; ++is) {
          double &v = c.conc[ix * c.nSpecies + is];
          if (v < 0.0) {
            v = 0.0;
          }
        }
      }
    });
  }

  // pixelsim.cpp:122-140
  double doRK101(double dt) {
    solveZeroStorageConstraints();
    calculateDcdt();
    dt = std::min(dt, maxStableTimestep);
    for (auto &c : comps) {
      forValues(c, [&](std::size_t i) { c.conc[i] += dt * c.dcdt[i]; });
      clampNegative(c);
    }
    return dt;
  }

  // pixelsim.cpp:142-164; pixelsim_impl.cpp:672-701
  void doRK212(double dt) {
    solveZeroStorageConstraints();
    calculateDcdt();
    for (auto &c : comps) {
      c.s2.resize(c.conc.size());
      c.s3.resize(c.conc.size());
      forValues(c, [&](std::size_t i) {
        c.s3[i] = c.conc[i];
        c.conc[i] += dt * c.dcdt[i];
      });
    }
    solveZeroStorageConstraints();
    calculateDcdt();
    for (auto &c : comps) {
      forValues(c, [&](std::size_t i) {
        c.s2[i] = c.conc[i];
        c.conc[i] = 0.5 * c.s3[i] + 0.5 * c.conc[i] + 0.5 * dt * c.dcdt[i];
      });
    }
  }

  // pixelsim.cpp:194-205; pixelsim_impl.cpp:714-721
  void doRKSubstep(double dt, double g1, double g2, double g3, double beta,
                   double delta) {
    solveZeroStorageConstraints(); // pixelsim.cpp:196
    calculateDcdt();
    for (auto &c : comps) {
      forValues(c, [&](std::size_t i) {
        c.s2[i] += delta * c.conc[i];
        c.conc[i] = g1 * c.conc[i] + g2 * c.s2[i] + g3 * c.s3[i] +
                    beta * dt * c.dcdt[i];
      });
    }
  }
  void doRKInit() {
    for (auto &c : comps) {
      c.s2.assign(c.conc.size(), 0.0);
      c.s3 = c.conc;
    }
  }
  void doRKFinalise(double cF, double s2F, double s3F) {
    for (auto &c : comps) {
      for (std::size_t i = 0; i < c.conc.size(); ++i) {
        c.s2[i] = cF * c.conc[i] + s2F * c.s2[i] + s3F * c.s3[i];
      }
    }
  }
  // pixelsim.cpp:166-192; tableaux pixelsim_common.hpp:138-171
  void doRK323(double dt) {
    static const double g1[3]{1.0, 0.25, 0.666666666666666666666};
    static const double g2[3]{0.0, 0.0, 0.0};
    static const double g3[3]{0.0, 0.75, 0.333333333333333333333};
    static const double beta[3]{1.0, 0.25, 0.6666666666666666666};
    static const double delta[3]{0.0, 0.0, 1.0};
    doRKInit();
    for (int i = 0; i < 3; ++i) {
      doRKSubstep(dt, g1[i], g2[i], g3[i], beta[i], delta[i]);
    }
    doRKFinalise(0.0, 2.0, -1.0);
  }
  void doRK435(double dt) {
    static const double g1[5]{0.0, -0.497531095840104, 1.010070514199942,
                              -3.196559004608766, 1.717835630267259};
    static const double g2[5]{1.0, 1.384996869124138, 3.878155713328178,
                              -2.324512951813145, -0.514633322274467};
    static const double g3[5]{0.0, 0.0, 0.0, 1.642598936063715,
                              0.188295940828347};
    static const double beta[5]{0.075152045700771, 0.211361016946069,
                                1.100713347634329, 0.728537814675568,
                                0.393172889823198};
    static const double delta[7]{1.0,
                                 0.081252332929194,
                                 -1.083849060586449,
                                 -1.096110881845602,
                                 2.859440022030827,
                                 -0.655568367959557,
                                 -0.194421504490852};
    const double dsr = 1.0 / (delta[0] + delta[1] + delta[2] + delta[3] +
                              delta[4] + delta[5] + delta[6]);
    doRKInit();
    for (int i = 0; i < 5; ++i) {
      doRKSubstep(dt, g1[i], g2[i], g3[i], beta[i], delta[i]);
    }
    doRKFinalise(dsr * delta[5], dsr, dsr * delta[6]);
  }

  // pixelsim_impl.cpp:774-792 — serial in the reference too
  void calculateRKError(const OCompartment &c, double &eabs, double &erel) const {
    for (std::size_t ix = 0; ix < c.nPixels; ++ix) {
      for (std::size_t is = 0; is < c.nSpecies; ++is) {
        if (c.invStorage[is] == 0.0) {
          continue;
        }
        const std::size_t i = ix * c.nSpecies + is;
        const double localErr = std::abs(c.conc[i] - c.s2[i]);
        eabs = std::max(eabs, localErr);
        const double localNorm = 0.5 * (c.conc[i] + c.s3[i] + epsilon);
        erel = std::max(erel, localErr / localNorm);
      }
    }
  }

  // pixelsim_impl.cpp:794-824 with a plain 0xffRRGGBB buffer for QImage
  std::string plotRKError(const OCompartment &c, double relMax) {
    std::size_t iSpecies = c.nSpecies + 1;
    for (std::size_t ix = 0; ix < c.nPixels; ++ix) {
      for (std::size_t is = 0; is < c.nSpecies; ++is) {
        const std::size_t i = ix * c.nSpecies + is;
        const double localErr = std::abs(c.conc[i] - c.s2[i]);
        const double localNorm = 0.5 * (c.conc[i] + c.s3[i] + epsilon);
        const double intensity = localErr / localNorm / relMax;
        const int red = static_cast<int>(255.0 * intensity);
        if (c.voxels.empty()) {
          continue;
        }
        const int x = c.voxels[3 * ix], y = c.voxels[3 * ix + 1],
                  z = c.voxels[3 * ix + 2];
        if (x < 0 || y < 0 || z < 0 || x >= imageNx || y >= imageNy ||
            z >= imageNz) {
          continue;
        }
        auto &px = errorImages[(static_cast<std::size_t>(z) * imageNy + y) *
                                   imageNx + x];
        const int oldRed = static_cast<int>((px >> 16) & 0xffu);
        if (red > oldRed) {
          px = 0xff000000u | (static_cast<uint32_t>(red & 0xff) << 16);
          if (red > 254) {
            iSpecies = is;
          }
        }
      }
    }
    if (iSpecies < c.nSpecies) {
      return c.speciesNames[iSpecies];
    }
    return {};
  }

  // pixelsim.cpp:207-270
  double doRKAdaptive(double dtMax) {
    double eabs = 0, erel = 0, dt = 0;
    const double errPower = integrator == SMEB200_RK323   ? 1.0 / 3.0
                            : integrator == SMEB200_RK435 ? 1.0 / 4.0
                                                          : 1.0 / 2.0;
    do {
      dt = std::min(nextTimestep, dtMax);
      if (integrator == SMEB200_RK212) {
        doRK212(dt);
      } else if (integrator == SMEB200_RK323) {
        doRK323(dt);
      } else {
        doRK435(dt);
      }
      eabs = 0;
      erel = 0;
      for (const auto &c : comps) {
        calculateRKError(c, eabs, erel);
      }
      double errFactor = std::min(errMaxAbs / eabs, errMaxRel / erel);
      errFactor = std::pow(errFactor, errPower);
      nextTimestep = std::min(0.95 * dt * errFactor, dtMax);
      if (nextTimestep / dtMax < 1e-20) {
        errorImages.assign(static_cast<std::size_t>(imageNx) * imageNy * imageNz,
                           0u);
        std::string problemSpecies{"unknown"};
        for (const auto &c : comps) {
          auto name = plotRKError(c, erel);
          if (!name.empty()) {
            problemSpecies = name;
          }
        }
        errorMessage =
            "Failed to solve model to required accuracy. The largest "
            "relative integration error comes from species '" +
            problemSpecies +
            "'. The location "
            "of the pixels with the largest relative integration error are "
            "shown "
            "below in red:";
        return nextTimestep;
      }
      if (eabs > errMaxAbs || erel > errMaxRel) {
        ++discardedSteps;
        for (auto &c : comps) {
          c.conc = c.s3; // undoRKStep, pixelsim_impl.cpp:759-763
        }
      }
    } while (eabs > errMaxAbs || erel > errMaxRel);
    for (auto &c : comps) {
      clampNegative(c);
    }
    return dt;
  }

  // pixelsim.cpp:423-471
  std::size_t run(double time, double timeoutMs, int (*stopCb)(void *),
                  void *user) {
    errorMessage.clear();
    const auto t0 = std::chrono::steady_clock::now();
    double tNow = 0;
    std::size_t steps = 0;
    discardedSteps = 0;
    constexpr double relativeTolerance = 1e-12;
    while (tNow + time * relativeTolerance < time) {
      const double maxDt = std::min(maxTimestep, time - tNow);
      if (integrator == SMEB200_RK101) {
        double timestep = std::min(maxDt, maxStableTimestep);
        timestep = doRK101(timestep);
        tNow += timestep;
      } else {
        tNow += doRKAdaptive(maxDt);
        if (!errorMessage.empty()) {
          return steps;
        }
      }
      ++steps;
      const double elapsed = std::chrono::duration<double, std::milli>(
                                 std::chrono::steady_clock::now() - t0)
                                 .count();
      if (timeoutMs >= 0.0 && elapsed >= timeoutMs) {
        stopRequested.store(true);
      }
      if (stopCb != nullptr && stopCb(user) != 0) {
        stopRequested.store(true);
      }
      if (stopRequested.load()) {
        errorMessage = "Simulation stopped early";
        return steps;
      }
    }
    return steps;
  }

  // stands in for Symbolic::parse + compile (LLVM JIT): the expressions are
  // compiled by the host C compiler without contraction or fast-math
  double maxStableTimestepAfterCtor{std::numeric_limits<double>::max()};
  void compileReactions(const smeb200_problem &p) {
    std::ostringstream src;
    src << "#include <math.h>\n";
    auto emit = [&](const std::string &name, uint32_t nVars, uint32_t nExpr,
                    const char *const *expr) {
      src << "void " << name
          << "(double* sme_out, const double* sme_values) {\n";
      for (uint32_t i = 0; i < nVars; ++i) {
        src << "  const double sme_var_" << i << " = sme_values[" << i
            << "]; (void)sme_var_" << i << ";\n";
      }
      // runtime parameters of the product are numeric constants here, as in
      // the reference (pde.cpp:65-77 inlines parameter values)
      for (uint32_t i = 0; i < p.n_params; ++i) {
        char buf[64];
        std::snprintf(buf, sizeof(buf), "%a", p.params[i]);
        src << "  const double sme_param_" << i << " = " << buf
            << "; (void)sme_param_" << i << ";\n";
      }
      for (uint32_t i = 0; i < nExpr; ++i) {
        src << "  sme_out[" << i << "] = " << expr[i] << ";\n";
      }
      src << "}\n";
    };
    for (uint32_t c = 0; c < p.n_compartments; ++c) {
      emit("reac" + std::to_string(c), p.compartments[c].n_species,
           p.compartments[c].n_species, p.compartments[c].expressions);
      if (p.compartments[c].n_cross_diffusion > 0) {
        emit("cross" + std::to_string(c), p.compartments[c].n_species,
             p.compartments[c].n_cross_diffusion,
             p.compartments[c].cross_expressions);
      }
    }
    for (uint32_t m = 0; m < p.n_membranes; ++m) {
      const auto &mb = p.membranes[m];
      auto primary = [&](int ci) -> uint32_t {
        if (ci < 0) {
          return 0;
        }
        const auto &pc = p.compartments[ci];
        return pc.n_primary_species != 0 ? pc.n_primary_species : pc.n_species;
      };
      auto extra = [&](int ci) -> uint32_t {
        return ci < 0 ? 0 : p.compartments[ci].n_species - primary(ci);
      };
      const uint32_t nA = primary(mb.comp_a), nB = primary(mb.comp_b);
      const uint32_t nE = mb.comp_b >= 0 ? extra(mb.comp_b) : extra(mb.comp_a);
      emit("memb" + std::to_string(m), nA + nB + nE, nA + nB, mb.expressions);
    }
    char tmpl[] = "/tmp/smeb200_oracle_XXXXXX";
    const int fd = mkstemp(tmpl);
    if (fd < 0) {
      throw std::runtime_error("oracle: mkstemp failed");
    }
    close(fd);
    const std::string base{tmpl};
    const std::string cPath = base + ".c";
    soPath = base + ".so";
    {
      std::ofstream f(cPath);
      f << src.str();
    }
    const char *cc = std::getenv("SMEB200_ORACLE_CC");
    const std::string cmd = std::string(cc != nullptr ? cc : "gcc") +
                            " -O2 -ffp-contract=off -fno-fast-math -shared "
                            "-fPIC -o " +
                            soPath + " " + cPath + " -lm 2> " + base + ".log";
    const int rc = std::system(cmd.c_str());
    unlink(cPath.c_str());
    unlink(tmpl);
    if (rc != 0) {
      std::ifstream lf(base + ".log");
      std::stringstream log;
      log << lf.rdbuf();
      unlink((base + ".log").c_str());
      throw std::runtime_error("oracle: failed to compile reactions: " +
                               log.str());
    }
    unlink((base + ".log").c_str());
    dl = dlopen(soPath.c_str(), RTLD_NOW | RTLD_LOCAL);
    if (dl == nullptr) {
      throw std::runtime_error(std::string("oracle: dlopen failed: ") +
                               dlerror());
    }
    for (uint32_t c = 0; c < p.n_compartments; ++c) {
      comps[c].reac = reinterpret_cast<ReacFn>(
          dlsym(dl, ("reac" + std::to_string(c)).c_str()));
      if (comps[c].hasCrossDiffusion) {
        comps[c].crossFn = reinterpret_cast<ReacFn>(
            dlsym(dl, ("cross" + std::to_string(c)).c_str()));
        // SimCompartment ctor, pixelsim_impl.cpp:425-428
        comps[c].evaluateCrossDiffusionCoefficients(0, comps[c].nPixels);
        comps[c].updateCrossDiffusionMaxStableTimestep();
        maxStableTimestepAfterCtor =
            std::min(maxStableTimestepAfterCtor, comps[c].maxStableTimestep);
      } else {
        maxStableTimestepAfterCtor =
            std::min(maxStableTimestepAfterCtor, comps[c].maxStableTimestep);
      }
    }
    for (uint32_t m = 0; m < p.n_membranes; ++m) {
      membs[m].reac = reinterpret_cast<ReacFn>(
          dlsym(dl, ("memb" + std::to_string(m)).c_str()));
    }
  }

  explicit Oracle(const smeb200_problem &p) {
    integrator = p.integrator;
    errMaxAbs = p.max_err_abs;
    errMaxRel = p.max_err_rel;
    maxTimestep = p.max_timestep;
    imageNx = p.image_nx;
    imageNy = p.image_ny;
    imageNz = p.image_nz;
    comps.resize(p.n_compartments);
    for (uint32_t ci = 0; ci < p.n_compartments; ++ci) {
      const auto &pc = p.compartments[ci];
      auto &c = comps[ci];
      c.id = pc.id != nullptr ? pc.id : "";
      c.nPixels = pc.n_voxels;
      c.nSpecies = pc.n_species;
      c.nn.resize(6 * c.nPixels);
      if (pc.nn != nullptr) {
        std::copy(pc.nn, pc.nn + 6 * c.nPixels, c.nn.begin());
      } else {
        // full box: Compartment ctor enumeration z, x, y
        // (geometry.cpp:55-72) and neighbour rule (geometry.cpp:98-110)
        const std::size_t nx = pc.box_nx, ny = pc.box_ny, nz = pc.box_nz;
        std::size_t i = 0;
        for (std::size_t z = 0; z < nz; ++z) {
          for (std::size_t x = 0; x < nx; ++x) {
            for (std::size_t y = 0; y < ny; ++y, ++i) {
              c.nn[6 * i] = x + 1 < nx ? i + ny : i;
              c.nn[6 * i + 1] = x > 0 ? i - ny : i;
              c.nn[6 * i + 2] = y + 1 < ny ? i + 1 : i;
              c.nn[6 * i + 3] = y > 0 ? i - 1 : i;
              c.nn[6 * i + 4] = z + 1 < nz ? i + nx * ny : i;
              c.nn[6 * i + 5] = z > 0 ? i - nx * ny : i;
            }
          }
        }
      }
      if (pc.voxels != nullptr) {
        c.voxels.assign(pc.voxels, pc.voxels + 3 * c.nPixels);
      } else if (pc.nn == nullptr) {
        c.voxels.resize(3 * c.nPixels);
        std::size_t i = 0;
        for (int z = 0; z < pc.box_nz; ++z) {
          for (int x = 0; x < pc.box_nx; ++x) {
            for (int y = 0; y < pc.box_ny; ++y, ++i) {
              c.voxels[3 * i] = x;
              c.voxels[3 * i + 1] = y;
              c.voxels[3 * i + 2] = z;
            }
          }
        }
      }
      c.nPrimarySpecies =
          pc.n_primary_species != 0 ? pc.n_primary_species : pc.n_species;
      c.dx2 = p.voxel_size[0] * p.voxel_size[0];
      c.dy2 = p.voxel_size[1] * p.voxel_size[1];
      c.dz2 = p.voxel_size[2] * p.voxel_size[2];
      c.useUniformDiffusionOperator = pc.diffusion_array == nullptr;
      for (uint32_t k = 0; k < pc.n_cross_diffusion; ++k) {
        c.crossDiffusionTerms.push_back({pc.cross_target[k], pc.cross_source[k]});
      }
      c.hasCrossDiffusion = pc.n_cross_diffusion > 0;
      if (c.hasCrossDiffusion) {
        if (pc.diffusion_raw == nullptr) {
          throw std::runtime_error("oracle: cross-diffusion needs diffusion_raw");
        }
        c.maxPrimaryDiagonalDiffusion.assign(
            pc.diffusion_raw, pc.diffusion_raw + c.nPrimarySpecies);
        c.crossDiffusionCoefficients.assign(
            c.nPixels * pc.n_cross_diffusion, 0.0);
      }
      for (uint32_t k = 0; k < pc.n_non_spatial; ++k) {
        c.nonSpatialSpeciesIndices.push_back(pc.non_spatial[k]);
      }
      for (uint32_t s = 0; s < pc.n_species; ++s) {
        c.diff.push_back({pc.diffusion[3 * s], pc.diffusion[3 * s + 1],
                          pc.diffusion[3 * s + 2]});
        c.invStorage.push_back(pc.inv_storage != nullptr ? pc.inv_storage[s]
                                                         : 1.0);
        if (c.invStorage.back() != 1.0) {
          c.hasNonUnitStorage = true;
        }
        if (c.invStorage.back() == 0.0) {
          // zero storage: algebraic species; the bound of the relaxation
          // iteration comes from its diffusion constant
          // (pixelsim_impl.cpp:218-224,238-247,273-282)
          c.hasZeroStorageSpecies = true;
          c.zeroStorageSpeciesIndices.push_back(s);
          double dsum = c.diff.back()[0] + c.diff.back()[1] + c.diff.back()[2];
          if (!c.useUniformDiffusionOperator) {
            double maxD = 0.0;
            for (std::size_t i = 0; i < c.nPixels; ++i) {
              const double d = pc.diffusion_array[i * pc.n_species + s];
              maxD = i == 0 ? d : std::max(maxD, d);
            }
            dsum = maxD / c.dx2 + maxD / c.dy2 + maxD / c.dz2;
            c.diffConstants.emplace_back(c.nPixels);
            for (std::size_t i = 0; i < c.nPixels; ++i) {
              c.diffConstants.back()[i] = pc.diffusion_array[i * pc.n_species + s];
            }
          }
          if (dsum > 0.0) {
            c.maxRelaxStableTimestep =
                std::min(c.maxRelaxStableTimestep, 1.0 / (2.0 * dsum));
          }
          c.speciesNames.emplace_back(
              pc.species_names != nullptr && pc.species_names[s] != nullptr
                  ? pc.species_names[s]
                  : "");
          continue;
        }
        if (c.useUniformDiffusionOperator) {
          // pixelsim_common.hpp:33-37, pixelsim_impl.cpp:249-255
          c.maxStableTimestep = std::min(
              c.maxStableTimestep,
              1.0 / (2.0 * (c.diff.back()[0] * c.invStorage.back() +
                            c.diff.back()[1] * c.invStorage.back() +
                            c.diff.back()[2] * c.invStorage.back())));
        } else {
          // pixelsim_impl.cpp:257-290: per-voxel constants, bound from max D
          c.diffConstants.emplace_back(c.nPixels);
          double maxD = 0.0;
          for (std::size_t i = 0; i < c.nPixels; ++i) {
            const double d = pc.diffusion_array[i * pc.n_species + s];
            c.diffConstants.back()[i] = d;
            maxD = i == 0 ? d : std::max(maxD, d);
          }
          c.maxStableTimestep = std::min(
              c.maxStableTimestep,
              1.0 / (2.0 * (maxD * c.invStorage.back() / c.dx2 +
                            maxD * c.invStorage.back() / c.dy2 +
                            maxD * c.invStorage.back() / c.dz2)));
        }
        c.speciesNames.emplace_back(
            pc.species_names != nullptr && pc.species_names[s] != nullptr
                ? pc.species_names[s]
                : "");
      }
      c.conc.assign(pc.conc0, pc.conc0 + c.nPixels * c.nSpecies);
      c.dcdt.assign(c.conc.size(), 0.0);
      if (c.hasZeroStorageSpecies) { // pixelsim.cpp:355-365
        c.relaxOld.resize(c.conc.size());
        c.relaxFirstOrder.resize(c.conc.size());
        hasAnyZeroStorageSpecies = true;
        maxRelaxStableTimestep =
            std::min(maxRelaxStableTimestep, c.maxRelaxStableTimestep);
      }
      maxStableTimestep = std::min(maxStableTimestep, c.maxStableTimestep);
    }
    membs.resize(p.n_membranes);
    for (uint32_t mi = 0; mi < p.n_membranes; ++mi) {
      const auto &pm = p.membranes[mi];
      auto &m = membs[mi];
      m.compA = pm.comp_a;
      m.compB = pm.comp_b;
      for (int d = 0; d < 6; ++d) {
        m.fluxLength[d] = pm.flux_length[d];
        for (uint64_t k = 0; k < pm.n_pairs[d]; ++k) {
          m.pairs[d].emplace_back(pm.pairs[d][2 * k], pm.pairs[d][2 * k + 1]);
        }
      }
    }
    if (hasAnyZeroStorageSpecies &&
        maxRelaxStableTimestep == std::numeric_limits<double>::max()) {
      maxRelaxStableTimestep = 1.0; // all zero-storage species have D = 0
    }
    compileReactions(p);
    if (p.n_compartments > 0) { // pixelsim.cpp:353-354
      maxStableTimestep = maxStableTimestepAfterCtor;
    }
  }
};

} // namespace

// ---- C interface (same shapes as include/smeb200.h, prefix oracle_) ----------
extern "C" {

struct oracle_sim {
  std::unique_ptr<Oracle> impl;
  std::string ctorError;
};

int oracle_create(const smeb200_problem *p, oracle_sim **out) {
  auto *s = new oracle_sim;
  try {
    s->impl = std::make_unique<Oracle>(*p);
  } catch (const std::exception &e) {
    s->ctorError = e.what();
  }
  *out = s;
  return s->ctorError.empty() ? 0 : 1;
}
void oracle_destroy(oracle_sim *s) { delete s; }
void oracle_set_threads(oracle_sim *s, int nThreads) {
  if (s->impl) {
    s->impl->useThreads = nThreads != 1;
#ifdef _OPENMP
    if (nThreads > 0) {
      omp_set_num_threads(nThreads);
    }
#endif
  }
}
int oracle_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
size_t oracle_run(oracle_sim *s, double time, double timeout_ms,
                  int (*stop_cb)(void *), void *user) {
  return s->impl ? s->impl->run(time, timeout_ms, stop_cb, user) : 0;
}
int oracle_get_concentrations(oracle_sim *s, size_t comp, const double **data,
                              size_t *n) {
  auto &c = s->impl->comps.at(comp);
  *data = c.conc.data();
  *n = c.conc.size();
  return 0;
}
int oracle_get_dcdt(oracle_sim *s, size_t comp, const double **data, size_t *n) {
  auto &c = s->impl->comps.at(comp);
  *data = c.dcdt.data();
  *n = c.dcdt.size();
  return 0;
}
int oracle_get_lower_order(oracle_sim *s, size_t comp, size_t species,
                           size_t pixel, double *value) {
  auto &c = s->impl->comps.at(comp);
  *value = c.s2.empty() ? 0.0 : c.s2[pixel * c.nSpecies + species];
  return 0;
}
int oracle_set_concentrations(oracle_sim *s, size_t comp, const double *data,
                              size_t n) {
  auto &c = s->impl->comps.at(comp);
  if (n != c.conc.size()) {
    return 1;
  }
  std::copy(data, data + n, c.conc.begin());
  return 0;
}
const char *oracle_error(const oracle_sim *s) {
  return s->impl ? s->impl->errorMessage.c_str() : s->ctorError.c_str();
}
int oracle_error_images(const oracle_sim *s, const uint32_t **data, size_t *n) {
  *data = s->impl->errorImages.data();
  *n = s->impl->errorImages.size();
  return 0;
}
void oracle_set_stop_requested(oracle_sim *s, int stop) {
  s->impl->stopRequested.store(stop != 0);
}
int oracle_get_stop_requested(const oracle_sim *s) {
  return s->impl->stopRequested.load() ? 1 : 0;
}
double oracle_get_next_timestep(const oracle_sim *s) {
  return s->impl->nextTimestep;
}
void oracle_set_next_timestep(oracle_sim *s, double dt) {
  s->impl->nextTimestep = dt;
}
size_t oracle_get_discarded_steps(const oracle_sim *s) {
  return s->impl->discardedSteps;
}
double oracle_get_max_stable_timestep(const oracle_sim *s) {
  return s->impl->maxStableTimestep;
}
}
