#include "ensemble.hpp"
#include <array>
#include <chrono>
#include <cmath>
#include <cstring>
#include <limits>

namespace smeb200 {

namespace {
constexpr unsigned BLOCK = 256;
constexpr std::size_t kRing = 8; // SMEB_ERR_RING

template <typename T> T *devAllocE(std::size_t count, const char *what) {
  void *p = nullptr;
  check(cudaMalloc(&p, std::max<std::size_t>(count, 1) * sizeof(T)), what);
  return static_cast<T *>(p);
}
} // namespace

EnsembleSim::EnsembleSim(const smeb200_problem &p, uint32_t nMembers,
                         const double *memberParams)
    : M{nMembers} {
  try {
    build(p, memberParams);
  } catch (const std::exception &e) {
    // like the simulator constructors: never throw, keep the message
    errorMessage = e.what();
    ctorFailed = true;
  }
}

void EnsembleSim::build(const smeb200_problem &p, const double *memberParams) {
  if (M == 0) {
    throw Error("smeb200: an ensemble needs at least one member");
  }
  if (p.n_params > 0 && memberParams == nullptr) {
    throw Error("smeb200: an ensemble needs the members' parameter values");
  }
  if (p.n_ranks > 1) {
    throw Error("smeb200: ensembles are sharded by member, not z-slab "
                "decomposed (one ensemble per GPU)");
  }
  // the prototype: geometry and membrane tables, plan, compiled module.  Box
  // compartments use the direct kernels (a tensor map describes one buffer).
  smeb200_problem q = p;
  q.kernel_variant = 1;
  q.params = memberParams; // member 0's values
  proto = std::make_unique<Sim>(q);
  if (proto->ctorFailed) {
    throw Error(proto->errorMessage);
  }
  if (proto->anyZeroStorage || proto->anyCross) {
    throw Error("smeb200: zero-storage species and cross-diffusion are not "
                "supported in a member-batched ensemble");
  }
  for (const auto &c : proto->comps) {
    if (!c.nonSpatial.empty()) {
      throw Error("smeb200: non-spatial species are not supported in a "
                  "member-batched ensemble");
    }
  }
  proto->bindDevice();
  cudaStream_t stream = proto->stream;
  const std::size_t nc = proto->comps.size();
  cd.resize(nc);
  role.assign(M, std::array<int, 4>{0, 1, 2, 3});
  memberError.assign(M, std::string{});
  steps.assign(M, 0);
  discarded.assign(M, 0);
  nextTimestep.assign(M, 1e-7); // pixelsim_base.hpp:45
  for (std::size_t ci = 0; ci < nc; ++ci) {
    const auto &c = proto->comps[ci];
    auto &d = cd[ci];
    const std::size_t nv = c.n * c.S;
    if (nv == 0) {
      continue;
    }
    for (int b = 0; b < c.nbuf; ++b) {
      d.buf[b] = devAllocE<double>(nv * M, "allocate ensemble buffers");
    }
    // every member starts from the model's initial state (the prototype's
    // host mirror holds it)
    for (uint32_t m = 0; m < M; ++m) {
      check(cudaMemcpyAsync(d.buf[0] + m * nv, c.hostConc, nv * sizeof(double),
                            cudaMemcpyHostToDevice, stream),
            "upload initial concentrations");
    }
    check(cudaMallocHost(reinterpret_cast<void **>(&d.hTable),
                         sizeof(MemberRec) * M * kTableRing),
          "allocate pinned member table");
    d.dTable = devAllocE<MemberRec>(M, "allocate member table");
    check(cudaMallocHost(reinterpret_cast<void **>(&d.hostConc),
                         nv * M * sizeof(double)),
          "allocate pinned ensemble mirror");
    for (uint32_t m = 0; m < M; ++m) {
      std::memcpy(d.hostConc + m * nv, c.hostConc, nv * sizeof(double));
    }
    d.hostStale.assign(M, 0);
    for (const auto &sd : proto->plan) {
      d.kStage.push_back(
          proto->onDemandKernels(ci, proto->stageFlags(sd) | F_MEMBERS));
      const auto &k = d.kStage.back();
      if (k.main != nullptr) {
        proto->persistentGrid(k.main, c.geom == Geom::SEG ? c.segSmem : 0u);
      }
      if (k.memb != nullptr) {
        proto->persistentGrid(k.memb);
      }
    }
  }
  const std::size_t np = proto->params.size();
  dParams = devAllocE<double>(static_cast<std::size_t>(M) * std::max<std::size_t>(np, 1),
                              "allocate member parameters");
  if (np > 0) {
    check(cudaMemcpyAsync(dParams, memberParams, sizeof(double) * M * np,
                          cudaMemcpyHostToDevice, stream),
          "upload member parameters");
  }
  dErr = devAllocE<unsigned long long>(2 * M + 2, "allocate error accumulators");
  check(cudaMemsetAsync(dErr, 0, sizeof(unsigned long long) * (2 * M + 2), stream),
        "clear error accumulators");
  const std::size_t ringBytes = sizeof(unsigned long long) * (2 * M + 2) * kRing;
  check(cudaHostAlloc(reinterpret_cast<void **>(&hErrRing), ringBytes,
                      cudaHostAllocMapped | cudaHostAllocPortable),
        "allocate pinned error ring");
  std::memset(hErrRing, 0, ringBytes);
  {
    void *dev = nullptr;
    check(cudaHostGetDevicePointer(&dev, hErrRing, 0),
          "cudaHostGetDevicePointer(error ring)");
    dErrRing = static_cast<unsigned long long *>(dev);
  }
  for (auto &e : tableEvents) {
    check(cudaEventCreateWithFlags(&e, cudaEventDisableTiming), "cudaEventCreate");
  }
  if (const char *env = std::getenv("SMEB200_GRAPHS"); env != nullptr) {
    useGraph = env[0] != '0';
  }
  check(cudaEventCreateWithFlags(&forkEvent, cudaEventDisableTiming),
        "cudaEventCreate");
  for (std::size_t k = 0; k < 2 * nc; ++k) {
    cudaStream_t st{};
    cudaEvent_t e{};
    check(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking), "cudaStreamCreate");
    check(cudaEventCreateWithFlags(&e, cudaEventDisableTiming), "cudaEventCreate");
    forkStreams.push_back(st);
    joinEvents.push_back(e);
  }
  check(cudaEventCreate(&ev0), "cudaEventCreate");
  check(cudaEventCreate(&ev1), "cudaEventCreate");
  check(cudaStreamSynchronize(stream), "initial upload");
}

EnsembleSim::~EnsembleSim() {
  if (!proto || proto->stream == nullptr) {
    return;
  }
  cudaSetDevice(proto->device);
  cudaStreamSynchronize(proto->stream);
  for (auto &d : cd) {
    for (double *b : d.buf) {
      cudaFree(b);
    }
    cudaFreeHost(d.hTable);
    cudaFree(d.dTable);
    cudaFreeHost(d.hostConc);
  }
  cudaFree(dParams);
  cudaFree(dErr);
  cudaFreeHost(hErrRing);
  for (auto e : tableEvents) {
    if (e != nullptr) {
      cudaEventDestroy(e);
    }
  }
  if (stepGraph != nullptr) {
    cudaGraphExecDestroy(stepGraph);
  }
  for (auto st : forkStreams) {
    cudaStreamDestroy(st);
  }
  for (auto e : joinEvents) {
    cudaEventDestroy(e);
  }
  if (forkEvent != nullptr) {
    cudaEventDestroy(forkEvent);
  }
  if (ev0 != nullptr) {
    cudaEventDestroy(ev0);
    cudaEventDestroy(ev1);
  }
}

// one attempted step of every active member: the member tables are rewritten in
// pinned memory, then the step's launches are replayed (or issued)
void EnsembleSim::batchedStep(const std::vector<double> &dt,
                              const std::vector<char> &active) {
  Sim &s = *proto;
  cudaStream_t stream = s.stream;
  for (std::size_t ci = 0; ci < cd.size(); ++ci) {
    const auto &c = s.comps[ci];
    auto &d = cd[ci];
    const std::size_t nv = c.n * c.S;
    if (nv == 0) {
      continue;
    }
    MemberRec *t = d.hTable;
    for (uint32_t m = 0; m < M; ++m) {
      for (int r = 0; r < 4; ++r) {
        const int b = role[m][r];
        t[m].buf[r] = d.buf[b] != nullptr ? d.buf[b] + m * nv : nullptr;
      }
      t[m].dt = dt[m];
      t[m].active = active[m] ? 1u : 0u;
      t[m].pad_ = 0;
      if (active[m]) {
        d.hostStale[m] = 1;
      }
    }
  }
  if (useGraph) {
    if (stepGraph == nullptr) {
      cudaGraph_t graph{};
      const uint64_t launches0 = launchCount;
      const uint32_t seq0 = errSeq;
      check(cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal),
            "cudaStreamBeginCapture");
      s.capturing = true;
      try {
        enqueueStep(true);
      } catch (...) {
        s.capturing = false;
        cudaStreamEndCapture(stream, &graph);
        if (graph != nullptr) {
          cudaGraphDestroy(graph);
        }
        throw;
      }
      s.capturing = false;
      check(cudaStreamEndCapture(stream, &graph), "cudaStreamEndCapture");
      launchesPerStep = launchCount - launches0;
      launchCount = launches0;
      errSeq = seq0;
      const cudaError_t e = cudaGraphInstantiate(&stepGraph, graph, 0);
      cudaGraphDestroy(graph);
      check(e, "cudaGraphInstantiate");
    }
    check(cudaGraphLaunch(stepGraph, stream), "cudaGraphLaunch");
    launchCount += launchesPerStep;
    if (s.integrator != SMEB200_RK101) {
      ++errSeq;
    }
  } else {
    enqueueStep(false);
  }
  ++batchedSteps;
}

void EnsembleSim::enqueueStep(bool fork) {
  Sim &s = *proto;
  cudaStream_t stream = s.stream;
  for (std::size_t ci = 0; ci < cd.size(); ++ci) {
    const auto &c = s.comps[ci];
    if (c.n * c.S == 0) {
      continue;
    }
    check(cudaMemcpyAsync(cd[ci].dTable, cd[ci].hTable, sizeof(MemberRec) * M,
                          cudaMemcpyHostToDevice, stream),
          "upload member table");
  }
  for (std::size_t si = 0; si < s.plan.size(); ++si) {
    const StageDesc &sd = s.plan[si];
    uint32_t expected = 0;
    std::vector<StageArgsHost> args(cd.size());
    for (std::size_t ci = 0; ci < cd.size(); ++ci) {
      const auto &c = s.comps[ci];
      if (c.n * c.S == 0) {
        continue;
      }
      StageArgsHost a = s.stageArgsTemplate(ci, si);
      a.members = cd[ci].dTable;
      a.memberParams = dParams;
      a.err = dErr;
      a.errBase = dErr;
      a.errPairs = M;
      a.errHost = dErrRing;
      const uint32_t s2in = sd.s2mode == 2 ? 0u : (sd.s2mode == 3 ? 3u : 15u);
      a.roles = static_cast<uint32_t>(sd.cin) |
                (static_cast<uint32_t>(sd.cout) << 4) | (s2in << 8) |
                ((sd.writeS2 ? 3u : 15u) << 12);
      args[ci] = a;
      if (sd.final) {
        const auto &k = cd[ci].kStage[si];
        expected += (s.mainGrid(ci, a, k.main) + s.membGrid(ci, k.memb)) * M;
      }
    }
    // within a stage the kernels of different compartments, and a compartment's
    // main and membrane kernels, write disjoint voxels and only read the stage
    // input: parallel branches of the captured step
    if (fork) {
      check(cudaEventRecord(forkEvent, stream), "fork event");
    }
    std::size_t used = 0;
    for (std::size_t ci = 0; ci < cd.size(); ++ci) {
      const auto &c = s.comps[ci];
      if (c.n * c.S == 0) {
        continue;
      }
      StageArgsHost &a = args[ci];
      a.errExpected = expected;
      const auto &k = cd[ci].kStage[si];
      cudaStream_t sMain = stream, sMemb = stream;
      if (fork) {
        sMain = forkStreams[2 * ci];
        sMemb = forkStreams[2 * ci + 1];
        check(cudaStreamWaitEvent(sMain, forkEvent, 0), "fork wait");
      }
      {
        const unsigned grid = s.mainGrid(ci, a, k.main);
        void *params[] = {&a};
        s.launchKernel(k.main, grid, M, 1, c.geom == Geom::SEG ? c.segSmem : 0u,
                       sMain, params);
        ++launchCount;
      }
      if (fork) {
        check(cudaEventRecord(joinEvents[2 * ci], sMain), "join event");
        check(cudaStreamWaitEvent(stream, joinEvents[2 * ci], 0), "join");
      }
      if (c.nMvox > 0 && k.memb != nullptr) {
        // SmebMembArgs (sim.cpp: launchCompartment): the stage inputs of the
        // partner compartments come from the member tables
        const std::size_t ncomp = cd.size();
        std::vector<unsigned long long> blob(4 + 2 * ncomp + 1, 0ull);
        blob[0] = reinterpret_cast<unsigned long long>(c.dMvox);
        blob[1] = reinterpret_cast<unsigned long long>(c.dMstart);
        blob[2] = reinterpret_cast<unsigned long long>(c.dEnt);
        blob[3] = reinterpret_cast<unsigned long long>(s.dFluxLen);
        for (std::size_t j = 0; j < ncomp; ++j) {
          blob[4 + ncomp + j] = reinterpret_cast<unsigned long long>(cd[j].dTable);
        }
        blob[4 + 2 * ncomp] = static_cast<unsigned long long>(c.nMvox);
        void *mparams[] = {&a, blob.data()};
        if (fork) {
          check(cudaStreamWaitEvent(sMemb, forkEvent, 0), "fork wait");
        }
        s.launchKernel(k.memb, s.membGrid(ci, k.memb), M, 1, 0, sMemb, mparams);
        ++launchCount;
        if (fork) {
          check(cudaEventRecord(joinEvents[2 * ci + 1], sMemb), "join event");
          check(cudaStreamWaitEvent(stream, joinEvents[2 * ci + 1], 0), "join");
        }
      }
      ++used;
    }
    (void)used;
    if (sd.final && expected > 0) {
      ++errSeq;
    }
  }
}

void EnsembleSim::waitErrors(std::vector<double> &eabs, std::vector<double> &erel) {
  const std::size_t stride = 2 * M + 2;
  volatile unsigned long long *h = hErrRing + stride * (errSeq % kRing);
  const auto t0 = std::chrono::steady_clock::now();
  for (unsigned spins = 1;; ++spins) {
    if (static_cast<uint32_t>(h[2 * M]) == errSeq) {
      break;
    }
    if ((spins & 0x3ffu) == 0) {
      const cudaError_t e = cudaStreamQuery(proto->stream);
      if (e != cudaErrorNotReady) {
        check(e, "a stage kernel of the ensemble failed");
        if (static_cast<uint32_t>(h[2 * M]) != errSeq) {
          throw Error("smeb200: internal error (the error norms of an ensemble "
                      "step were not published)");
        }
      }
      if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0)
              .count() > 300.0) {
        throw Error("smeb200: timed out waiting for the error norms of an "
                    "ensemble step");
      }
    }
  }
  std::atomic_thread_fence(std::memory_order_acquire);
  for (uint32_t m = 0; m < M; ++m) {
    const unsigned long long a = h[2 * m], r = h[2 * m + 1];
    std::memcpy(&eabs[m], &a, 8);
    std::memcpy(&erel[m], &r, 8);
  }
}

// PixelSim::run and doRKAdaptive (pixelsim.cpp:423-471, 207-270) for every member,
// one batched launch sequence per round of attempts
std::size_t EnsembleSim::run(double time, double timeoutMs, int (*stopCb)(void *),
                             void *user) {
  if (ctorFailed) {
    return 0;
  }
  errorMessage.clear();
  std::size_t total = 0;
  try {
    Sim &s = *proto;
    s.bindDevice();
    check(cudaEventRecord(ev0, s.stream), "cudaEventRecord");
    const auto t0 = std::chrono::steady_clock::now();
    constexpr double relativeTolerance = 1e-12;
    std::vector<double> tNow(M, 0.0), dt(M, 0.0), dtMax(M, 0.0), eabs(M, 0.0),
        erel(M, 0.0);
    std::vector<char> active(M, 0);
    for (uint32_t m = 0; m < M; ++m) {
      steps[m] = 0;
      discarded[m] = 0;
      memberError[m].clear();
    }
    const double errPower = s.integrator == SMEB200_RK323   ? 1.0 / 3.0
                            : s.integrator == SMEB200_RK435 ? 1.0 / 4.0
                                                            : 1.0 / 2.0;
    const int outRole = s.plan.back().cout;
    for (;;) {
      bool any = false;
      for (uint32_t m = 0; m < M; ++m) {
        active[m] = memberError[m].empty() &&
                    tNow[m] + time * relativeTolerance < time;
        dt[m] = 0.0;
        if (!active[m]) {
          continue;
        }
        any = true;
        dtMax[m] = std::min(s.maxTimestep, time - tNow[m]);
        dt[m] = s.integrator == SMEB200_RK101
                    ? std::min(dtMax[m], s.maxStableTimestep)
                    : std::min(nextTimestep[m], dtMax[m]);
      }
      if (!any) {
        break;
      }
      batchedStep(dt, active);
      if (s.integrator != SMEB200_RK101) {
        waitErrors(eabs, erel);
      } else {
        // no norms to wait for: the uploads of this step must have read the
        // pinned tables before the next step rewrites them
        check(cudaEventRecord(tableEvents[0], s.stream), "cudaEventRecord");
        check(cudaEventSynchronize(tableEvents[0]), "cudaEventSynchronize");
      }
      for (uint32_t m = 0; m < M; ++m) {
        if (!active[m]) {
          continue;
        }
        bool accept = true;
        if (s.integrator != SMEB200_RK101) {
          double errFactor = std::min(s.errMaxAbs / eabs[m], s.errMaxRel / erel[m]);
          errFactor = std::pow(errFactor, errPower);
          nextTimestep[m] = std::min(0.95 * dt[m] * errFactor, dtMax[m]);
          if (nextTimestep[m] / dtMax[m] < 1e-20) {
            memberError[m] =
                "Failed to solve model to required accuracy. The largest "
                "relative integration error comes from species 'unknown'. The "
                "location of the pixels with the largest relative integration "
                "error are shown below in red:";
            continue;
          }
          accept = !(eabs[m] > s.errMaxAbs || erel[m] > s.errMaxRel);
        }
        if (accept) {
          std::swap(role[m][CUR], role[m][outRole]);
          tNow[m] += dt[m];
          ++steps[m];
          ++total;
        } else {
          ++discarded[m];
        }
      }
      if (timeoutMs >= 0.0) {
        if (s.integrator == SMEB200_RK101) {
          check(cudaStreamSynchronize(s.stream), "cudaStreamSynchronize");
        }
        const double elapsed = std::chrono::duration<double, std::milli>(
                                   std::chrono::steady_clock::now() - t0)
                                   .count();
        if (elapsed >= timeoutMs) {
          stopRequested.store(true);
        }
      }
      if (stopCb != nullptr && stopCb(user) != 0) {
        stopRequested.store(true);
      }
      if (stopRequested.load()) {
        errorMessage = "Simulation stopped early";
        break;
      }
    }
    check(cudaEventRecord(ev1, s.stream), "cudaEventRecord");
    check(cudaStreamSynchronize(s.stream), "cudaStreamSynchronize");
    float ms = 0;
    check(cudaEventElapsedTime(&ms, ev0, ev1), "cudaEventElapsedTime");
    lastRunMs = ms;
  } catch (const std::exception &e) {
    errorMessage = e.what();
  }
  return total;
}

const double *EnsembleSim::getConcentrations(uint32_t member, std::size_t comp,
                                             std::size_t *n) {
  if (ctorFailed) {
    throw Error(errorMessage);
  }
  if (member >= M || comp >= cd.size()) {
    throw Error("member or compartment index out of range");
  }
  const auto &c = proto->comps[comp];
  auto &d = cd[comp];
  const std::size_t nv = c.n * c.S;
  if (n != nullptr) {
    *n = nv;
  }
  if (nv > 0 && d.hostStale[member]) {
    proto->bindDevice();
    check(cudaMemcpyAsync(d.hostConc + member * nv,
                          d.buf[role[member][CUR]] + member * nv,
                          nv * sizeof(double), cudaMemcpyDeviceToHost,
                          proto->stream),
          "download concentrations");
    check(cudaStreamSynchronize(proto->stream), "download concentrations");
    d.hostStale[member] = 0;
  }
  return d.hostConc + member * nv;
}

void EnsembleSim::setConcentrations(uint32_t member, std::size_t comp,
                                    const double *data, std::size_t n) {
  if (ctorFailed) {
    throw Error(errorMessage);
  }
  if (member >= M || comp >= cd.size()) {
    throw Error("member or compartment index out of range");
  }
  const auto &c = proto->comps[comp];
  auto &d = cd[comp];
  const std::size_t nv = c.n * c.S;
  if (n != nv) {
    throw Error("setConcentrations: wrong number of values");
  }
  if (nv == 0) {
    return;
  }
  proto->bindDevice();
  std::memcpy(d.hostConc + member * nv, data, nv * sizeof(double));
  check(cudaMemcpyAsync(d.buf[role[member][CUR]] + member * nv,
                        d.hostConc + member * nv, nv * sizeof(double),
                        cudaMemcpyHostToDevice, proto->stream),
        "upload concentrations");
  check(cudaStreamSynchronize(proto->stream), "upload concentrations");
  d.hostStale[member] = 0;
}

} // namespace smeb200
