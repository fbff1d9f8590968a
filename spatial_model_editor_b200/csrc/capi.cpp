// extern "C" surface of libsmeb200.so — see include/smeb200.h for the contract
// and the reference interface each entry point stands in for.
#include "comm.hpp"
#include "ensemble.hpp"
#include "sim.hpp"
#include <cstring>

using smeb200::Sim;

struct smeb200_sim {
  Sim impl;
  explicit smeb200_sim(const smeb200_problem &p, bool compileOnly = false)
      : impl(p, compileOnly) {}
};

namespace {
template <typename F> int guarded(smeb200_sim *sim, F &&f) {
  if (sim == nullptr) {
    return -1;
  }
  try {
    f();
    return 0;
  } catch (const std::exception &e) {
    sim->impl.errorMessage = e.what();
    return 1;
  }
}
} // namespace

extern "C" {

int smeb200_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

int smeb200_create(const smeb200_problem *p, smeb200_sim **out) {
  if (out == nullptr) {
    return -1;
  }
  *out = nullptr;
  if (p == nullptr) {
    return -1;
  }
  try {
    *out = new smeb200_sim(*p);
  } catch (...) {
    return -1;
  }
  return (*out)->impl.errorMessage.empty() ? 0 : 1;
}

void smeb200_destroy(smeb200_sim *sim) { delete sim; }

int smeb200_compile_only(const smeb200_problem *p, smeb200_sim **out) {
  if (out == nullptr || p == nullptr) {
    return -1;
  }
  *out = nullptr;
  try {
    *out = new smeb200_sim(*p, true);
  } catch (...) {
    return -1;
  }
  return (*out)->impl.errorMessage.empty() && (*out)->impl.cubinBytes > 0 ? 0
                                                                         : 1;
}

size_t smeb200_run(smeb200_sim *sim, double time, double timeout_ms,
                   int (*stop_cb)(void *user), void *user) {
  if (sim == nullptr) {
    return 0;
  }
  return sim->impl.run(time, timeout_ms, stop_cb, user);
}

int smeb200_get_concentrations(smeb200_sim *sim, size_t compartment,
                               const double **data, size_t *n) {
  return guarded(sim, [&] {
    const double *p = sim->impl.getConcentrations(compartment, n);
    if (data != nullptr) {
      *data = p;
    }
  });
}

int smeb200_get_dcdt(smeb200_sim *sim, size_t compartment, const double **data,
                     size_t *n) {
  return guarded(sim, [&] {
    const double *p = sim->impl.getDcdt(compartment, n);
    if (data != nullptr) {
      *data = p;
    }
  });
}

int smeb200_get_lower_order(smeb200_sim *sim, size_t compartment,
                            size_t species, size_t pixel, double *value) {
  return guarded(sim, [&] {
    const double v = sim->impl.getLowerOrder(compartment, species, pixel);
    if (value != nullptr) {
      *value = v;
    }
  });
}

int smeb200_get_avg_min_max(smeb200_sim *sim, size_t compartment, double *out,
                            size_t n_out) {
  return guarded(sim,
                 [&] { sim->impl.getAvgMinMax(compartment, out, n_out); });
}

int smeb200_get_species_decimated(smeb200_sim *sim, size_t compartment,
                                  size_t species, const uint32_t stride[3],
                                  const double **data, size_t dims[3]) {
  return guarded(sim, [&] {
    if (stride == nullptr || data == nullptr || dims == nullptr) {
      throw smeb200::Error("getSpeciesDecimated: null argument");
    }
    *data = sim->impl.getSpeciesDecimated(compartment, species, stride, dims);
  });
}

size_t smeb200_neighbour_bytes_per_voxel(const smeb200_sim *s, size_t comp) {
  return s != nullptr && comp < s->impl.nCompartments()
             ? s->impl.neighbourBytesPerVoxel(comp)
             : 0;
}

size_t smeb200_neighbour_runs(const smeb200_sim *s, size_t comp) {
  return s != nullptr && comp < s->impl.nCompartments()
             ? s->impl.neighbourRuns(comp)
             : 0;
}

int smeb200_compress_neighbours(const uint32_t *nn, uint64_t n_voxels,
                                uint32_t *codes, int32_t *runs,
                                uint64_t *n_runs) {
  try {
    std::vector<uint32_t> c;
    std::vector<int> r;
    if (nn == nullptr || n_runs == nullptr ||
        !smeb200::compressNeighbourTable(nn, n_voxels, c, r)) {
      return 1;
    }
    if (codes != nullptr) {
      std::copy(c.begin(), c.end(), codes);
    }
    if (runs != nullptr && r.size() / 4 <= *n_runs) {
      std::copy(r.begin(), r.end(), runs);
    }
    *n_runs = r.size() / 4;
    return 0;
  } catch (...) {
    return 2;
  }
}

size_t smeb200_get_concentration_padding(const smeb200_sim *s) {
  return s != nullptr ? s->impl.concentrationPadding() : 0;
}

int smeb200_set_concentrations(smeb200_sim *sim, size_t compartment,
                               const double *data, size_t n) {
  return guarded(sim,
                 [&] { sim->impl.setConcentrations(compartment, data, n); });
}

void smeb200_set_async_download(smeb200_sim *sim, int enable) {
  if (sim != nullptr) {
    sim->impl.setAsyncDownload(enable != 0);
  }
}

int smeb200_set_host_mirror(smeb200_sim *sim, size_t compartment, double *buffer,
                            size_t n) {
  if (sim == nullptr) {
    return -1;
  }
  try {
    sim->impl.setHostMirror(compartment, buffer, n);
  } catch (const std::exception &e) {
    sim->impl.errorMessage = e.what();
    return 1;
  }
  return 0;
}

int smeb200_set_params(smeb200_sim *sim, const double *values, size_t n) {
  return guarded(sim, [&] { sim->impl.setParams(values, n); });
}

const char *smeb200_error(const smeb200_sim *sim) {
  return sim != nullptr ? sim->impl.errorMessage.c_str() : "null simulator";
}

void smeb200_set_error(smeb200_sim *sim, const char *msg) {
  if (sim != nullptr) {
    sim->impl.errorMessage = msg != nullptr ? msg : "";
  }
}

int smeb200_error_images(const smeb200_sim *sim, const uint32_t **data,
                         size_t *n) {
  if (sim == nullptr) {
    return -1;
  }
  if (data != nullptr) {
    *data = sim->impl.errorImages.data();
  }
  if (n != nullptr) {
    *n = sim->impl.errorImages.size();
  }
  return 0;
}

int smeb200_get_stop_requested(const smeb200_sim *sim) {
  return sim != nullptr && sim->impl.stopRequested.load() ? 1 : 0;
}

void smeb200_set_stop_requested(smeb200_sim *sim, int stop) {
  if (sim != nullptr) {
    sim->impl.stopRequested.store(stop != 0);
  }
}

double smeb200_get_next_timestep(const smeb200_sim *sim) {
  return sim != nullptr ? sim->impl.nextTimestep : 0.0;
}

void smeb200_set_next_timestep(smeb200_sim *sim, double dt) {
  if (sim != nullptr) {
    sim->impl.nextTimestep = dt;
  }
}

size_t smeb200_get_discarded_steps(const smeb200_sim *sim) {
  return sim != nullptr ? sim->impl.discardedSteps : 0;
}

double smeb200_get_max_stable_timestep(const smeb200_sim *sim) {
  return sim != nullptr ? sim->impl.maxStableTimestep : 0.0;
}

double smeb200_bench_steps(smeb200_sim *sim, double dt, size_t n_steps) {
  return sim != nullptr ? sim->impl.benchSteps(dt, n_steps) : -1.0;
}

double smeb200_last_run_ms(const smeb200_sim *sim) {
  return sim != nullptr ? sim->impl.lastRunMs : 0.0;
}

uint64_t smeb200_launch_count(const smeb200_sim *sim) {
  return sim != nullptr ? sim->impl.launchCount : 0;
}

void smeb200_set_kernel_timing(smeb200_sim *sim, int enable) {
  if (sim != nullptr) {
    sim->impl.kernelTiming = enable != 0;
    sim->impl.kernelTimingStride = enable > 1 ? static_cast<unsigned>(enable) : 1u;
    sim->impl.timingStep = 0;
    sim->impl.kernelMs = 0;
    sim->impl.kernelLaunchesTimed = 0;
  }
}

int smeb200_get_kernel_timing(const smeb200_sim *sim, double *total_ms,
                              uint64_t *launches) {
  if (sim == nullptr) {
    return -1;
  }
  if (total_ms != nullptr) {
    *total_ms = sim->impl.kernelMs;
  }
  if (launches != nullptr) {
    *launches = sim->impl.kernelLaunchesTimed;
  }
  return 0;
}

const char *smeb200_kernel_source(const smeb200_sim *sim) {
  return sim != nullptr ? sim->impl.kernelSource.c_str() : "";
}

const char *smeb200_dominant_kernel(const smeb200_sim *sim) {
  return sim != nullptr ? sim->impl.dominantKernelName.c_str() : "";
}

struct smeb200_comm {
  smeb200::ShmComm impl;
  std::string error;
  smeb200_comm(const char *name, int rank, int n) : impl(name, rank, n) {}
};

int smeb200_comm_create(const char *name, int rank, int n_ranks,
                        smeb200_comm **out) {
  if (out == nullptr || name == nullptr) {
    return -1;
  }
  *out = nullptr;
  try {
    *out = new smeb200_comm(name, rank, n_ranks);
  } catch (...) {
    return 1;
  }
  return 0;
}

int smeb200_comm_allreduce_sum(smeb200_comm *comm, double *values, int n) {
  if (comm == nullptr || values == nullptr) {
    return -1;
  }
  try {
    comm->impl.allreduceSumOrdered(values, n);
  } catch (const std::exception &e) {
    comm->error = e.what();
    return 1;
  }
  return 0;
}

int smeb200_comm_allreduce_max(smeb200_comm *comm, double *values, int n) {
  if (comm == nullptr || values == nullptr) {
    return -1;
  }
  try {
    comm->impl.allreduceMax(values, n);
  } catch (const std::exception &e) {
    comm->error = e.what();
    return 1;
  }
  return 0;
}

void smeb200_comm_abort(smeb200_comm *comm) {
  if (comm != nullptr) {
    comm->impl.abort();
  }
}

int smeb200_comm_barrier(smeb200_comm *comm) {
  if (comm == nullptr) {
    return -1;
  }
  try {
    comm->impl.barrier();
  } catch (const std::exception &e) {
    comm->error = e.what();
    return 1;
  }
  return 0;
}

void smeb200_comm_destroy(smeb200_comm *comm) { delete comm; }

size_t smeb200_timed_launch_planes(const smeb200_sim *sim) {
  try {
    return sim != nullptr ? sim->impl.timedLaunchPlanes() : 0;
  } catch (...) {
    return 0;
  }
}

uint64_t smeb200_halo_exchanges(const smeb200_sim *sim) {
  return sim != nullptr ? sim->impl.haloExchanges : 0;
}

void smeb200_set_allreduce_max(smeb200_sim *sim, smeb200_allreduce_max_fn fn,
                               void *user) {
  if (sim != nullptr) {
    sim->impl.allreduceMax = fn;
    sim->impl.allreduceUser = user;
  }
}


/* ---- member-batched ensembles ---------------------------------------------- */
struct smeb200_ensemble {
  smeb200::EnsembleSim impl;
  smeb200_ensemble(const smeb200_problem &p, uint32_t n, const double *params)
      : impl(p, n, params) {}
};

int smeb200_ensemble_create(const smeb200_problem *p, uint32_t n_members,
                            const double *member_params, smeb200_ensemble **out) {
  if (out == nullptr) {
    return -1;
  }
  *out = nullptr;
  if (p == nullptr) {
    return -1;
  }
  try {
    *out = new smeb200_ensemble(*p, n_members, member_params);
  } catch (...) {
    return -1;
  }
  return (*out)->impl.errorMessage.empty() ? 0 : 1;
}

void smeb200_ensemble_destroy(smeb200_ensemble *ens) { delete ens; }

size_t smeb200_ensemble_run(smeb200_ensemble *ens, double time, double timeout_ms,
                            int (*stop_cb)(void *user), void *user) {
  if (ens == nullptr) {
    return 0;
  }
  return ens->impl.run(time, timeout_ms, stop_cb, user);
}

int smeb200_ensemble_get_concentrations(smeb200_ensemble *ens, uint32_t member,
                                        size_t compartment, const double **data,
                                        size_t *n) {
  if (ens == nullptr) {
    return -1;
  }
  try {
    const double *p = ens->impl.getConcentrations(member, compartment, n);
    if (data != nullptr) {
      *data = p;
    }
  } catch (const std::exception &e) {
    ens->impl.errorMessage = e.what();
    return 1;
  }
  return 0;
}

int smeb200_ensemble_set_concentrations(smeb200_ensemble *ens, uint32_t member,
                                        size_t compartment, const double *data,
                                        size_t n) {
  if (ens == nullptr) {
    return -1;
  }
  try {
    ens->impl.setConcentrations(member, compartment, data, n);
  } catch (const std::exception &e) {
    ens->impl.errorMessage = e.what();
    return 1;
  }
  return 0;
}

size_t smeb200_ensemble_get_steps(const smeb200_ensemble *ens, uint32_t member) {
  return ens != nullptr && member < ens->impl.nMembers() ? ens->impl.steps[member]
                                                         : 0;
}

size_t smeb200_ensemble_get_discarded_steps(const smeb200_ensemble *ens,
                                            uint32_t member) {
  return ens != nullptr && member < ens->impl.nMembers()
             ? ens->impl.discarded[member]
             : 0;
}

double smeb200_ensemble_get_next_timestep(const smeb200_ensemble *ens,
                                          uint32_t member) {
  return ens != nullptr && member < ens->impl.nMembers()
             ? ens->impl.nextTimestep[member]
             : 0.0;
}

void smeb200_ensemble_set_next_timestep(smeb200_ensemble *ens, uint32_t member,
                                        double dt) {
  if (ens != nullptr && member < ens->impl.nMembers()) {
    ens->impl.nextTimestep[member] = dt;
  }
}

const char *smeb200_ensemble_error(const smeb200_ensemble *ens, int member) {
  if (ens == nullptr) {
    return "";
  }
  if (member < 0 || static_cast<uint32_t>(member) >= ens->impl.nMembers()) {
    return ens->impl.errorMessage.c_str();
  }
  return ens->impl.memberError[member].c_str();
}

void smeb200_ensemble_set_stop_requested(smeb200_ensemble *ens, int stop) {
  if (ens != nullptr) {
    ens->impl.stopRequested.store(stop != 0);
  }
}

uint64_t smeb200_ensemble_launch_count(const smeb200_ensemble *ens) {
  return ens != nullptr ? ens->impl.launchCount : 0;
}

uint64_t smeb200_ensemble_batched_steps(const smeb200_ensemble *ens) {
  return ens != nullptr ? ens->impl.batchedSteps : 0;
}

double smeb200_ensemble_last_run_ms(const smeb200_ensemble *ens) {
  return ens != nullptr ? ens->impl.lastRunMs : 0.0;
}

} // extern "C"
