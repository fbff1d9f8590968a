// B200PixelSim — SME-side adapter for libsmeb200.so (include/smeb200.h).
//
// Drop this header into core/simulate/src/ of Spatial Model Editor and link
// libsmeb200.so.  The class has the constructor contract of PixelSim /
// CudaPixelSim (pixelsim.hpp:52-56, cudapixelsim.hpp:59-63), derives from
// PixelSimBase (pixelsim_base.hpp:17-50) and forwards every BaseSim virtual
// (basesim.hpp:24-54) to one C-ABI entry point.  It fills smeb200_problem from
// sme::model::Model exactly as CudaPixelSim's constructor does
// (cudapixelsim.cpp:770-1168): nn table as uint32, D/dx^2 table, AoS initial
// concentrations, ReacExpr -> Symbolic -> relabel -> cudaCode strings, per-face
// membrane pair lists, restart concentrations.
//
// It is written against the reference headers.  SME cannot be linked in this
// repository (Qt6 / SymEngine / libSBML are not installed here), but the header
// is type-checked against the reference's own headers by tests/test_adapter_cpu.py
// (g++ -fsyntax-only, declaration-only stand-ins for the third-party includes:
// tests/adapter_check/).  INTEGRATION.md shows the patch to simulate.cpp that
// selects it.
#pragma once

#include <functional> // basesim.hpp uses std::function without including it

#include "pixelsim_base.hpp"
#include "pixelsim_common.hpp"
#include "pixelsim_impl.hpp" // ReacExpr
#include "sme/geometry.hpp"
#include "sme/model.hpp"
#include "sme/simple_symbolic.hpp"
#include "sme/symbolic.hpp"
#include "sme/utils.hpp"
#include "smeb200.h"
#include <QImage>
#include <array>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace sme::simulate {

class B200PixelSim : public PixelSimBase {
private:
  struct SimDeleter {
    void operator()(smeb200_sim *s) const { smeb200_destroy(s); }
  };
  std::unique_ptr<smeb200_sim, SimDeleter> sim;
  const model::Model &doc;
  std::size_t nCompartments{0};
  std::size_t nExtraVars{0}; // time, x, y, z components behind the species
  // BaseSim returns references: host mirrors owned here
  mutable std::vector<std::vector<double>> conc;
  bool mirrored{true}; // conc[i] is the library's host mirror of compartment i
  mutable std::vector<std::vector<double>> dcdt;

  static std::vector<std::string> cudaExpressions(
      const std::vector<std::string> &variables,
      const std::vector<std::string> &expressions,
      const std::vector<std::pair<std::string, double>> &constants = {}) {
    // cudapixelsim.cpp:108-136
    std::vector<std::string> safe;
    for (std::size_t i = 0; i < variables.size(); ++i) {
      safe.push_back("sme_var_" + std::to_string(i));
    }
    common::Symbolic symbolic{expressions, variables, constants};
    if (!symbolic.isValid()) {
      throw std::runtime_error("Failed to convert reaction expression to CUDA "
                               "code: " +
                               symbolic.getErrorMessage());
    }
    symbolic.relabel(safe);
    std::vector<std::string> out;
    for (std::size_t i = 0; i < expressions.size(); ++i) {
      out.push_back(symbolic.cudaCode(i, false));
    }
    return out;
  }

  void syncErrorState() {
    currentErrorMessage = smeb200_error(sim.get());
    const std::uint32_t *px{nullptr};
    std::size_t n{0};
    smeb200_error_images(sim.get(), &px, &n);
    if (n > 0) {
      const auto size{doc.getGeometry().getImages().volume()};
      currentErrorImages = common::ImageStack(size, QImage::Format_ARGB32);
      std::size_t i{0};
      for (std::size_t z = 0; z < size.depth(); ++z) {
        for (int y = 0; y < size.height(); ++y) {
          for (int x = 0; x < size.width(); ++x) {
            currentErrorImages[z].setPixel(x, y, px[i++]);
          }
        }
      }
    }
  }

public:
  explicit B200PixelSim(
      const model::Model &sbmlDoc,
      const std::vector<std::string> &compartmentIds,
      const std::vector<std::vector<std::string>> &compartmentSpeciesIds,
      const std::map<std::string, double, std::less<>> &substitutions = {})
      : PixelSimBase{sbmlDoc.getSimulationSettings().options.pixel.integrator,
                     sbmlDoc.getSimulationSettings().options.pixel.maxErr,
                     sbmlDoc.getSimulationSettings().options.pixel.maxTimestep},
        doc{sbmlDoc} {
    try {
      smeb200_problem p{};
      p.abi_version = SMEB200_ABI_VERSION;
      p.integrator = static_cast<int32_t>(integrator);
      p.max_err_abs = errMax.abs;
      p.max_err_rel = errMax.rel;
      p.max_timestep = maxTimestep;
      p.device = -1;
      const auto imageSize{doc.getGeometry().getImages().volume()};
      p.image_nx = imageSize.width();
      p.image_ny = imageSize.height();
      p.image_nz = static_cast<int32_t>(imageSize.depth());
      // remaining feature gates (cudapixelsim.cpp:789-809): cross-diffusion
      // and non-uniform diffusion.  Time / space dependent reactions,
      // non-spatial species and non-unit storage are described to the library
      // the way PixelSim handles them (pixelsim.cpp:282-322).
      const auto &coords{doc.getParameters().getSpatialCoordinates()};
      // pixelsim.cpp:286-322: reactions or cross-diffusion coefficients that
      // use time / x, y, z
      const auto crossDependsOn = [&](const std::string &variableId) {
        for (const auto &speciesIds : compartmentSpeciesIds) {
          for (const auto &target : speciesIds) {
            for (const auto &source : speciesIds) {
              if (target == source) {
                continue;
              }
              const auto expression{doc.getSpecies().getCrossDiffusionConstant(
                  target.c_str(), source.c_str())};
              if (!expression.isEmpty() &&
                  common::SimpleSymbolic::contains(expression.toStdString(),
                                                   variableId)) {
                return true;
              }
            }
          }
        }
        return false;
      };
      const bool timeDependent{doc.getReactions().dependOnVariable("time") ||
                               crossDependsOn("time")};
      const bool spaceDependent{
          doc.getReactions().dependOnVariable(coords.x.id.c_str()) ||
          doc.getReactions().dependOnVariable(coords.y.id.c_str()) ||
          doc.getReactions().dependOnVariable(coords.z.id.c_str()) ||
          crossDependsOn(coords.x.id) || crossDependsOn(coords.y.id) ||
          crossDependsOn(coords.z.id)};
      // pixelsim.cpp:323-336: one non-uniform species switches every
      // compartment to the per-voxel (array) diffusion operator
      bool allUniformDiffusion{true};
      for (const auto &speciesIds : compartmentSpeciesIds) {
        for (const auto &speciesId : speciesIds) {
          if (doc.getSpecies().getIsConstant(speciesId.c_str())) {
            continue;
          }
          const auto *field{doc.getSpecies().getField(speciesId.c_str())};
          if (field == nullptr || !field->getIsUniformDiffusionConstant()) {
            allUniformDiffusion = false;
          }
        }
      }
      p.voxel_size[0] = doc.getGeometry().getVoxelSize().width();
      p.voxel_size[1] = doc.getGeometry().getVoxelSize().height();
      p.voxel_size[2] = doc.getGeometry().getVoxelSize().depth();
      nExtraVars = (timeDependent ? 1 : 0) + (spaceDependent ? 3 : 0);
      const auto origin{doc.getGeometry().getPhysicalOrigin()};
      const auto voxelSize{doc.getGeometry().getVoxelSize()};
      const double dx2 = voxelSize.width() * voxelSize.width();
      const double dy2 = voxelSize.height() * voxelSize.height();
      const double dz2 = voxelSize.depth() * voxelSize.depth();

      nCompartments = compartmentIds.size();
      std::vector<smeb200_compartment> comps(nCompartments);
      std::vector<std::vector<std::uint32_t>> nn(nCompartments);
      std::vector<std::vector<std::int32_t>> voxels(nCompartments);
      std::vector<std::vector<double>> diffusion(nCompartments);
      std::vector<std::vector<double>> invStorage(nCompartments);
      std::vector<std::vector<double>> diffusionArray(nCompartments);
      std::vector<std::vector<double>> diffusionRaw(nCompartments);
      std::vector<std::vector<std::uint32_t>> crossTarget(nCompartments);
      std::vector<std::vector<std::uint32_t>> crossSource(nCompartments);
      std::vector<std::vector<std::string>> crossExprs(nCompartments);
      std::vector<std::vector<const char *>> crossExprPtrs(nCompartments);
      std::vector<std::vector<std::uint32_t>> nonSpatial(nCompartments);
      std::vector<std::vector<std::string>> exprs(nCompartments);
      std::vector<std::vector<const char *>> exprPtrs(nCompartments);
      std::vector<std::vector<const char *>> namePtrs(nCompartments);
      std::map<std::string, std::size_t, std::less<>> compIndexById;
      conc.resize(nCompartments);
      dcdt.resize(nCompartments);
      for (std::size_t ci = 0; ci < nCompartments; ++ci) {
        const auto &speciesIds{compartmentSpeciesIds[ci]};
        const auto *comp{doc.getCompartments().getCompartment(
            compartmentIds[ci].c_str())};
        compIndexById[comp->getId()] = ci;
        const std::size_t n{comp->nVoxels()};
        const std::size_t nPrimary{speciesIds.size()};
        const std::size_t S{nPrimary + nExtraVars}; // + time, x, y, z padding
        auto &c{comps[ci]};
        c.id = comp->getId().c_str();
        c.n_voxels = n;
        c.n_species = static_cast<std::uint32_t>(S);
        c.n_primary_species = static_cast<std::uint32_t>(nPrimary);
        // a compartment that fills the whole image needs no neighbour table:
        // the library then uses the TMA tile kernel with implicit neighbours
        const bool fillsImage{n == static_cast<std::size_t>(p.image_nx) *
                                       p.image_ny * p.image_nz};
        voxels[ci].reserve(3 * n);
        for (std::size_t i = 0; i < n; ++i) {
          const auto &v{comp->getVoxel(i)};
          voxels[ci].push_back(v.p.x());
          voxels[ci].push_back(v.p.y());
          voxels[ci].push_back(static_cast<std::int32_t>(v.z));
        }
        c.voxels = voxels[ci].data();
        if (fillsImage) {
          c.box_nx = p.image_nx;
          c.box_ny = p.image_ny;
          c.box_nz = p.image_nz;
          c.nn = nullptr;
        } else {
          nn[ci].reserve(6 * n); // cudapixelsim.cpp:877-890
          for (std::size_t i = 0; i < n; ++i) {
            nn[ci].push_back(static_cast<std::uint32_t>(comp->up_x(i)));
            nn[ci].push_back(static_cast<std::uint32_t>(comp->dn_x(i)));
            nn[ci].push_back(static_cast<std::uint32_t>(comp->up_y(i)));
            nn[ci].push_back(static_cast<std::uint32_t>(comp->dn_y(i)));
            nn[ci].push_back(static_cast<std::uint32_t>(comp->up_z(i)));
            nn[ci].push_back(static_cast<std::uint32_t>(comp->dn_z(i)));
          }
          c.nn = nn[ci].data();
        }
        std::vector<const geometry::Field *> fields;
        for (const auto &speciesId : speciesIds) {
          const auto *field{doc.getSpecies().getField(speciesId.c_str())};
          if (field == nullptr) {
            throw std::runtime_error("could not find field for species '" +
                                     speciesId + "'");
          }
          if (!field->getIsSpatial()) { // pixelsim_impl.cpp:303-305
            nonSpatial[ci].push_back(
                static_cast<std::uint32_t>(fields.size()));
          }
          const bool isConstant{
              doc.getSpecies().getIsConstant(speciesId.c_str())};
          // pixelsim_impl.cpp:214-232; storage 0 (algebraic species) is
          // passed as 0: the library relaxes it like pixelsim.cpp:23-90
          const double storage{
              doc.getSpecies().getStorage(speciesId.c_str())};
          if (storage < 0.0) {
            throw std::runtime_error(
                "Invalid negative storage value for species '" + speciesId +
                "'");
          }
          invStorage[ci].push_back(storage == 0.0 ? 0.0 : 1.0 / storage);
          fields.push_back(field);
          const auto &d{field->getDiffusionConstant()};
          const double dc{isConstant || d.empty() ? 0.0 : d.front()};
          // maxPrimaryDiagonalDiffusion (pixelsim_impl.cpp:256,291)
          diffusionRaw[ci].push_back(
              allUniformDiffusion ? dc
                                  : (isConstant || d.empty() ? 0.0
                                                             : common::max(d)));
          diffusion[ci].push_back(dc / dx2);
          diffusion[ci].push_back(dc / dy2);
          diffusion[ci].push_back(dc / dz2);
          namePtrs[ci].push_back(speciesId.c_str());
        }
        // the padding components: zero diffusion, unit storage
        // (pixelsim_impl.cpp:312-349)
        static const std::array<std::string, 4> padNames{"time", "x", "y", "z"};
        for (std::size_t k = 0; k < nExtraVars; ++k) {
          diffusion[ci].insert(diffusion[ci].end(), {0.0, 0.0, 0.0});
          invStorage[ci].push_back(1.0);
          namePtrs[ci].push_back(
              padNames[(timeDependent ? 0 : 1) + k].c_str());
        }
        c.diffusion = diffusion[ci].data();
        c.species_names = namePtrs[ci].data();
        c.inv_storage = invStorage[ci].data();
        c.diffusion_raw = diffusionRaw[ci].data();
        if (!allUniformDiffusion) {
          // per-voxel constants, species-interleaved like the concentrations
          // (pixelsim_impl.cpp:257-266)
          diffusionArray[ci].assign(n * S, 0.0);
          for (std::size_t is = 0; is < nPrimary; ++is) {
            const auto &d{fields[is]->getDiffusionConstant()};
            if (doc.getSpecies().getIsConstant(speciesIds[is].c_str()) ||
                d.empty()) {
              continue;
            }
            for (std::size_t ix = 0; ix < n; ++ix) {
              diffusionArray[ci][ix * S + is] = d[ix];
            }
          }
          c.diffusion_array = diffusionArray[ci].data();
        }
        c.n_non_spatial = static_cast<std::uint32_t>(nonSpatial[ci].size());
        c.non_spatial = nonSpatial[ci].empty() ? nullptr : nonSpatial[ci].data();
        conc[ci].resize(n * S);
        auto it{conc[ci].begin()};
        const int imageNy{comp->getCompartmentImages()[0].height()};
        for (std::size_t ix = 0; ix < n; ++ix) {
          for (const auto *field : fields) {
            *it++ = field->getConcentration()[ix];
          }
          // pixelsim_impl.cpp:405-423
          if (timeDependent) {
            *it++ = 0.0;
          }
          if (spaceDependent) {
            const auto voxel{comp->getVoxel(ix)};
            *it++ = origin.p.x() +
                    (static_cast<double>(voxel.p.x()) + 0.5) * voxelSize.width();
            *it++ = origin.p.y() +
                    (static_cast<double>(imageNy - 1 - voxel.p.y()) + 0.5) *
                        voxelSize.height();
            *it++ = origin.z +
                    (static_cast<double>(voxel.z) + 0.5) * voxelSize.depth();
          }
        }
        c.conc0 = conc[ci].data();
        dcdt[ci].assign(n * S, 0.0);
        std::vector<std::string> reactionIDs;
        if (auto r{doc.getReactions().getIds(comp->getId().c_str())};
            !r.isEmpty()) {
          reactionIDs = common::toStdString(r);
        }
        const ReacExpr reacExpr(doc, speciesIds, reactionIDs, 1.0,
                                timeDependent, spaceDependent, substitutions);
        exprs[ci] = cudaExpressions(reacExpr.variables, reacExpr.expressions);
        for (const auto &e : exprs[ci]) {
          exprPtrs[ci].push_back(e.c_str());
        }
        c.expressions = exprPtrs[ci].data();
        // sparse cross-diffusion terms D_ij grad c_j, i != j
        // (pixelsim_impl.cpp:350-389)
        std::vector<std::string> crossDiffusionExpressions;
        for (std::size_t iTarget = 0; iTarget < nPrimary; ++iTarget) {
          const QString targetId{speciesIds[iTarget].c_str()};
          if (doc.getSpecies().getIsConstant(targetId)) {
            continue;
          }
          for (std::size_t iSource = 0; iSource < nPrimary; ++iSource) {
            if (iSource == iTarget) {
              continue;
            }
            const auto expression{doc.getSpecies().getCrossDiffusionConstant(
                targetId, speciesIds[iSource].c_str())};
            if (expression.isEmpty()) {
              continue;
            }
            crossDiffusionExpressions.push_back(
                doc.inlineExpr(expression.toStdString()));
            crossTarget[ci].push_back(static_cast<std::uint32_t>(iTarget));
            crossSource[ci].push_back(static_cast<std::uint32_t>(iSource));
          }
        }
        if (!crossDiffusionExpressions.empty()) {
          std::vector<std::pair<std::string, double>> constants;
          for (const auto &constant :
               doc.getParameters().getGlobalConstants()) {
            double value{constant.value};
            if (auto it{substitutions.find(constant.id)};
                it != substitutions.end()) {
              value = it->second;
            }
            constants.emplace_back(constant.id, value);
          }
          crossExprs[ci] = cudaExpressions(reacExpr.variables,
                                           crossDiffusionExpressions, constants);
          for (const auto &e : crossExprs[ci]) {
            crossExprPtrs[ci].push_back(e.c_str());
          }
          c.n_cross_diffusion =
              static_cast<std::uint32_t>(crossExprs[ci].size());
          c.cross_target = crossTarget[ci].data();
          c.cross_source = crossSource[ci].data();
          c.cross_expressions = crossExprPtrs[ci].data();
        }
      }
      // restart concentrations (pixelsim.cpp:399-406)
      const auto &data{doc.getSimulationData()};
      if (data.concentration.size() > 1 && !data.concentration.back().empty() &&
          data.concentration.back().size() == nCompartments) {
        for (std::size_t ci = 0; ci < nCompartments; ++ci) {
          if (data.concentration.back()[ci].size() == conc[ci].size()) {
            conc[ci] = data.concentration.back()[ci];
            comps[ci].conc0 = conc[ci].data();
          }
        }
      }
      // membranes with reactions (cudapixelsim.cpp:1020-1096)
      const auto &lengthUnit{doc.getUnits().getLength()};
      const auto &volumeUnit{doc.getUnits().getVolume()};
      const double volOverL3{model::getVolOverL3(lengthUnit, volumeUnit)};
      std::vector<smeb200_membrane> membs;
      std::vector<std::array<std::vector<std::uint32_t>, 6>> pairs;
      std::vector<std::vector<std::string>> mExprs;
      std::vector<std::vector<const char *>> mExprPtrs;
      const auto &membranes{doc.getMembranes().getMembranes()};
      pairs.reserve(membranes.size());
      mExprs.reserve(membranes.size());
      mExprPtrs.reserve(membranes.size());
      for (const auto &membrane : membranes) {
        auto reacs{doc.getReactions().getIds(membrane.getId().c_str())};
        if (reacs.isEmpty()) {
          continue;
        }
        smeb200_membrane m{};
        m.id = membrane.getId().c_str();
        m.comp_a = m.comp_b = -1;
        std::vector<std::string> speciesIds;
        if (auto it{compIndexById.find(membrane.getCompartmentA()->getId())};
            it != compIndexById.end()) {
          m.comp_a = static_cast<int32_t>(it->second);
          const auto &s{compartmentSpeciesIds[it->second]};
          speciesIds.insert(speciesIds.end(), s.begin(), s.end());
        }
        if (auto it{compIndexById.find(membrane.getCompartmentB()->getId())};
            it != compIndexById.end()) {
          m.comp_b = static_cast<int32_t>(it->second);
          const auto &s{compartmentSpeciesIds[it->second]};
          speciesIds.insert(speciesIds.end(), s.begin(), s.end());
        }
        if (speciesIds.empty()) {
          continue;
        }
        // variables: [A species | B species | time, x, y, z]; the library
        // uses the first SA + SB expressions (pixelsim_impl.cpp:960-1005)
        const ReacExpr reacExpr(doc, speciesIds, common::toStdString(reacs),
                                volOverL3, timeDependent, spaceDependent,
                                substitutions);
        mExprs.push_back(
            cudaExpressions(reacExpr.variables, reacExpr.expressions));
        mExprPtrs.emplace_back();
        for (const auto &e : mExprs.back()) {
          mExprPtrs.back().push_back(e.c_str());
        }
        m.expressions = mExprPtrs.back().data();
        pairs.emplace_back();
        for (std::size_t d = 0; d < 6; ++d) {
          const auto dir{static_cast<geometry::Membrane::FACE_DIRECTION>(d)};
          const auto &facePairs{membrane.getFaceIndexPairs(dir)};
          auto &flat{pairs.back()[d]};
          for (const auto &[ixA, ixB] : facePairs) {
            flat.push_back(static_cast<std::uint32_t>(ixA));
            flat.push_back(static_cast<std::uint32_t>(ixB));
          }
          m.n_pairs[d] = facePairs.size();
          m.pairs[d] = flat.data();
          m.flux_length[d] = detail::getFaceFluxLength(dir, voxelSize);
        }
        membs.push_back(m);
      }
      p.n_compartments = static_cast<std::uint32_t>(comps.size());
      p.compartments = comps.data();
      p.n_membranes = static_cast<std::uint32_t>(membs.size());
      p.membranes = membs.data();
      smeb200_sim *raw{nullptr};
      smeb200_create(&p, &raw);
      sim.reset(raw);
      if (!sim) {
        throw std::runtime_error("smeb200_create failed");
      }
      currentErrorMessage = smeb200_error(sim.get());
      maxStableTimestep = smeb200_get_max_stable_timestep(sim.get());
      if (currentErrorMessage.empty()) {
        // the vectors BaseSim::getConcentrations hands out are the library's host
        // mirrors (page-locked in place): one D2H, no second host copy; and the
        // download starts at the end of run() like the reference GPU backend's
        // (cudapixelsim.cpp:1486-1516) — Simulation reads the state after every
        // run (simulate.cpp:313-314)
        for (std::size_t ci = 0; ci < conc.size(); ++ci) {
          if (!conc[ci].empty()) {
            mirrored = smeb200_set_host_mirror(sim.get(), ci, conc[ci].data(),
                                               conc[ci].size()) == 0 &&
                       mirrored;
          }
        }
        smeb200_set_async_download(sim.get(), 1);
      }
    } catch (const std::exception &e) {
      // like the reference constructors: never throw (pixelsim.cpp:415-418)
      currentErrorMessage = e.what();
    }
  }

  std::size_t run(double time, double timeout_ms,
                  const std::function<bool()> &stopRunningCallback) override {
    if (!sim || !currentErrorMessage.empty()) {
      return 0; // cudapixelsim.cpp:1621-1623
    }
    smeb200_set_stop_requested(sim.get(), stopRequested.load() ? 1 : 0);
    auto trampoline = [](void *user) -> int {
      const auto &cb{*static_cast<const std::function<bool()> *>(user)};
      return cb && cb() ? 1 : 0;
    };
    const std::size_t steps{smeb200_run(
        sim.get(), time, timeout_ms, trampoline,
        const_cast<void *>(static_cast<const void *>(&stopRunningCallback)))};
    stopRequested.store(smeb200_get_stop_requested(sim.get()) != 0);
    discardedSteps = smeb200_get_discarded_steps(sim.get());
    nextTimestep = smeb200_get_next_timestep(sim.get());
    syncErrorState();
    return steps;
  }

  void setStopRequested(bool stop) override {
    stopRequested.store(stop);
    if (sim) {
      smeb200_set_stop_requested(sim.get(), stop ? 1 : 0); // thread-safe
    }
  }

  [[nodiscard]] const std::vector<double> &
  getConcentrations(std::size_t compartmentIndex) const override {
    if (sim) {
      const double *data{nullptr};
      std::size_t n{0};
      // waits for the download run() started; `data` is conc's own storage
      // when the mirror could be page-locked in place
      if (smeb200_get_concentrations(sim.get(), compartmentIndex, &data, &n) ==
              0 &&
          data != conc[compartmentIndex].data()) {
        conc[compartmentIndex].assign(data, data + n);
      }
    }
    return conc[compartmentIndex];
  }

  [[nodiscard]] std::size_t getConcentrationPadding() const override {
    return nExtraVars; // pixelsim.cpp:478
  }

  [[nodiscard]] const std::vector<double> &
  getDcdt(std::size_t compartmentIndex) const {
    if (sim) {
      const double *data{nullptr};
      std::size_t n{0};
      if (smeb200_get_dcdt(sim.get(), compartmentIndex, &data, &n) == 0) {
        dcdt[compartmentIndex].assign(data, data + n);
      }
    }
    return dcdt[compartmentIndex];
  }

  [[nodiscard]] double getLowerOrderConcentration(std::size_t compartmentIndex,
                                                  std::size_t speciesIndex,
                                                  std::size_t pixelIndex) const {
    double v{0.0};
    if (sim) {
      smeb200_get_lower_order(sim.get(), compartmentIndex, speciesIndex,
                              pixelIndex, &v);
    }
    return v;
  }
};

} // namespace sme::simulate
