// Member-batched parameter-scan ensembles (SURVEY.md section 8(e)/(f) rank 4,
// BASELINE config 5).  The reference has no such thing: a scan is a loop that
// sets a parameter and simulates again (sme_model.cpp:541-617), the optimizer
// evaluates its candidates one simulation at a time (optimize_impl.cpp:150-186).
//
// An EnsembleSim holds M simulations of ONE model that differ in the values of
// the runtime parameters.  Geometry tables, membrane tables and the compiled
// module are those of a prototype simulator; the state is [member][voxel][species]
// per rotating buffer; a stage of all members of a compartment is one launch with
// gridDim.y = M (device_kernels.cuh: SMEB_F_MEMBERS).  Every member has its own
// step-size controller — the loops of PixelSim::run / doRKAdaptive
// (pixelsim.cpp:207-270, 423-471), evaluated on the host for all members after
// each batched step from one 16 M byte publication of the error norms — so a
// member takes exactly the steps it would take alone.
#pragma once

#include "sim.hpp"
#include <string>
#include <vector>

namespace smeb200 {

class EnsembleSim {
public:
  EnsembleSim(const smeb200_problem &p, uint32_t nMembers,
              const double *memberParams);
  ~EnsembleSim();
  EnsembleSim(const EnsembleSim &) = delete;
  EnsembleSim &operator=(const EnsembleSim &) = delete;

  // every member integrates [0, time]; returns the accepted steps of all members
  std::size_t run(double time, double timeoutMs, int (*stopCb)(void *),
                  void *user);
  const double *getConcentrations(uint32_t member, std::size_t comp,
                                  std::size_t *n);
  void setConcentrations(uint32_t member, std::size_t comp, const double *data,
                         std::size_t n);

  uint32_t nMembers() const { return M; }
  std::string errorMessage;              // the ensemble as a whole
  std::vector<std::string> memberError;  // per member (adaptive failure)
  std::vector<std::size_t> steps, discarded; // of the last run()
  std::vector<double> nextTimestep;      // controller state per member
  bool ctorFailed{false};
  uint64_t launchCount{0};
  uint64_t batchedSteps{0};
  double lastRunMs{0};
  std::atomic<bool> stopRequested{false};

private:
  struct MemberRec { // device_kernels.cuh: SmebMember
    double *buf[4];
    double dt;
    uint32_t active;
    uint32_t pad_;
  };
  struct CompData {
    double *buf[4]{nullptr, nullptr, nullptr, nullptr}; // [M][n][S] each
    MemberRec *hTable{nullptr}; // pinned, kTableRing x M
    MemberRec *dTable{nullptr}; // M
    double *hostConc{nullptr};  // pinned [M][n][S]
    std::vector<char> hostStale;
    std::vector<StageKernels> kStage;
  };
  static constexpr int kTableRing = 1; // refilled only after the step before it has
                                       // reported in (or the stream was drained)
  void build(const smeb200_problem &p, const double *memberParams);
  void batchedStep(const std::vector<double> &dt, const std::vector<char> &active);
  // the launches of one batched step: table uploads, then per stage the kernels of
  // all compartments as parallel branches (captured once into a CUDA graph: a
  // step is launch-latency bound)
  void enqueueStep(bool fork);
  cudaGraphExec_t stepGraph{nullptr};
  uint64_t launchesPerStep{0};
  bool useGraph{true};
  std::vector<cudaStream_t> forkStreams;
  std::vector<cudaEvent_t> joinEvents;
  cudaEvent_t forkEvent{nullptr};
  void waitErrors(std::vector<double> &eabs, std::vector<double> &erel);

  std::unique_ptr<Sim> proto;
  uint32_t M{0};
  std::vector<CompData> cd;
  std::vector<std::array<int, 4>> role; // per member: role -> buffer index
  double *dParams{nullptr};
  unsigned long long *dErr{nullptr};     // 2 M pairs + count/seq
  unsigned long long *hErrRing{nullptr}; // kErrRing x (2 M + 2), pinned + mapped
  unsigned long long *dErrRing{nullptr};
  uint32_t errSeq{0};
  int tableSlot{0};
  cudaEvent_t tableEvents[kTableRing]{};
  cudaEvent_t ev0{nullptr}, ev1{nullptr};
};

} // namespace smeb200
