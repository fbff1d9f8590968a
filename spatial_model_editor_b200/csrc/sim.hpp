// B200PixelSim host side: the controller and device-state owner behind the
// C-ABI of include/smeb200.h.  Mirrors the member layout and control flow of
// sme::simulate::PixelSim / PixelSimBase (reference pixelsim.cpp:423-471,
// 207-270, 122-205; pixelsim_base.hpp:40-49) — the loops are the reference's,
// the per-stage work is one fused kernel launch per compartment.
#pragma once

#include "../../include/smeb200.h"
#include "comm.hpp"
#include "cuda_api.hpp"
#include <algorithm>
#include <array>
#include <atomic>
#include <cstdint>
#include <limits>
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace smeb200 {

// stage flags: compile-time template argument of the stage kernels
// (device_kernels.cuh SMEB_F_*)
constexpr uint32_t F_USE_C0 = 1u, F_USE_S2 = 2u, F_USE_S2IN = 4u,
                   F_WRITE_S2 = 8u, F_FINAL = 16u, F_LOWER_IS_CIN = 32u,
                   F_CLAMP_OUT = 64u, F_GENERIC = 128u, F_OUT_DCDT = 256u,
                   F_OUT_LOWER = 512u, F_OUT_CNEW = 1024u, F_SUM_NS = 2048u,
                   F_RELAX_A = 4096u, F_RELAX_B = 8192u, F_HALO = 16384u,
                   F_MEMBERS = 32768u, F_ZIG = 65536u;

// host mirror of the device-side SmebStageArgs<S> (device_kernels.cuh); the
// D[3*S] table is the tail.
struct StageArgsHost {
  const double *cin;
  const double *c0;
  const double *s2in;
  double *s2out;
  double *cout;
  const uint32_t *nn;
  const int *runs;
  const uint32_t *seg;
  unsigned long long *err;
  unsigned long long *errBase;
  unsigned long long *errHost;
  const void *members;
  const double *memberParams;
  const double *nsAv;
  double *nsPartial;
  const double *darr;
  unsigned long long *xdMax;
  double h2[3];
  uint32_t n;
  uint32_t v0;
  uint32_t nx, ny, nz;
  uint32_t z_begin, z_end;
  uint32_t errExpected;
  uint32_t haloMode;
  double *peerLo;
  double *peerHi;
  uint32_t *halo;
  uint32_t *peerFlagLo;
  uint32_t *peerFlagHi;
  uint32_t *haloErr;
  uint32_t haloSeq;
  uint32_t nSeg;
  uint32_t errPairs;
  uint32_t roles;
  uint32_t zig;
  uint32_t errPublish;
  double g1, g2, g3, bdt, delta;
  double beta;
  const double *dtp;
  double cF, s2F, s3F, eps;
  double D[3 * 64 + 64]; // + runtime parameters
};

struct MembEntryHost {
  uint32_t partner;
  uint32_t info;
};

// one Runge-Kutta stage in terms of buffer roles
enum Role : int { CUR = 0, RA = 1, RB = 2, RS = 3 };
struct StageDesc {
  int cin{CUR}, cout{RA};
  bool useC0{false};
  int s2mode{0}; // 0 none, 1 starts at zero, 2 read from CUR, 3 read from RS
  bool writeS2{false};
  bool generic{false};
  double g1{1}, g2{0}, g3{0}, beta{1}, delta{0};
  bool final{false}, lowerIsCin{false}, clampOut{false};
  double cF{0}, s2F{0}, s3F{0};
  uint32_t flags() const;
};

enum class Geom { NN, BOX, TILE, SEG };

// kernel pair for one compartment and one set of stage flags
struct StageKernels {
  CUfunction main{nullptr};
  CUfunction memb{nullptr};
};
constexpr const char *kNsFinishKernel = "smeb_ns_finish";

struct TileConfig {
  int tx{0}, ty{0}, nst{0};
  bool m3{true};
  bool zig{false}; // 3-D: chunk index fastest in the grid, alternating directions
  uint32_t chunk{1};
  unsigned grid[3]{1, 1, 1};
  unsigned dynSmem{0};
};

struct CompartmentState {
  std::string id;
  uint64_t n{0};
  uint32_t S{0};
  bool box{false};
  Geom geom{Geom::NN};
  uint32_t nx{0}, ny{0}, nz{0};
  uint32_t zLoGhost{0}, zHiGhost{0}; // z-slab ghost planes (read, not updated)
  // z-slab of a voxel list: ghost voxels at both ends of the local list and the
  // sizes of the first / last owned plane (what the neighbours hold as ghosts)
  uint32_t vGhostLo{0}, vGhostHi{0}, vFirstPlane{0}, vLastPlane{0};
  uint64_t ownedBegin() const { return box ? uint64_t{zLoGhost} * nx * ny : vGhostLo; }
  uint64_t ownedEnd() const {
    return box ? uint64_t{nz - zHiGhost} * nx * ny : n - vGhostHi;
  }
  TileConfig tile;
  // segment-staged voxel-list kernel (Geom::SEG): pieces of runs of at most
  // segLen voxels, staged in shared memory by 1-D bulk copies
  int segLen{0}, segNst{0};
  bool segZ{false};
  unsigned segSmem{0};
  uint32_t nSeg{0};
  uint32_t *dSeg{nullptr};
  std::vector<uint32_t> hostSeg, hostCodes; // until the upload
  std::vector<int> hostRuns;
  std::vector<std::string> speciesNames;
  std::vector<double> D;
  // CPU-solver semantics beyond the reference GPU backend (smeb200.h)
  uint32_t SP{0};                    // model species; S - SP = t,x,y,z padding
  std::vector<uint32_t> nonSpatial;  // species whose dc/dt is averaged
  std::vector<double> invStorage;    // [S], empty = unit storage
  double *dNsAv{nullptr};            // [nonSpatial.size()]
  double nGlobal{0.0};               // voxels of the compartment over all z-slabs
  double *dNsPartial{nullptr};       // block sums of the pre-pass
  uint32_t nsPartialBlocks{0};       // capacity of dNsPartial in blocks
  StageKernels kSumNs;               // pre-pass kernels (F_SUM_NS)
  double *dDarr{nullptr};            // [n][S] per-voxel diffusion constants
  bool nonUniform{false};
  std::vector<uint32_t> crossTarget, crossSource; // cross-diffusion terms
  std::vector<double> diffusionRaw;  // maxPrimaryDiagonalDiffusion
  unsigned long long *dXdMax{nullptr};
  CUfunction kXdMax{nullptr};
  double maxStable{std::numeric_limits<double>::max()};
  // zero-storage (algebraic) species: relaxation kernels, scratch buffers and
  // the relaxed copy that replaces the stage-0 input when the pre-step state
  // must survive (RK323 / RK435)
  struct PoolBuf {
    double *p{nullptr};
    CUtensorMap t{};
    // z-slabs: the same buffer of the rank below / above (CUDA IPC mapping); the
    // ranks swap their buffers in lock step, so the pairing survives the swaps
    double *peer[2]{nullptr, nullptr};
  };
  std::vector<uint32_t> zeroStorage;
  std::vector<PoolBuf> pool;
  StageKernels kRelaxA, kRelaxB;
  bool haveOverride{false};
  PoolBuf stageInOverride;
  double *dStatsPartial{nullptr};    // [blocks][3*SP]
  double *dStats{nullptr};           // [3*SP] avg, min, max per species
  std::vector<int32_t> voxels;
  int nbuf{0};
  double *buf[4]{nullptr, nullptr, nullptr, nullptr};
  std::vector<void *> stateAllocs; // what buf[] / pool[] point into (freed once)
  CUtensorMap tmap[4]{};
  int role[4]{0, 1, 2, 3};
  uint32_t *dNn{nullptr};
  int *dRuns{nullptr};               // run table of the compressed neighbour codes
  uint32_t nRuns{0};
  uint32_t nMvox{0};
  uint32_t *dMvox{nullptr};
  uint32_t *dMstart{nullptr};
  MembEntryHost *dEnt{nullptr};
  std::vector<StageKernels> kStage; // per plan stage
  // host mirrors
  double *hostConc{nullptr}; // pinned; owned unless userMirror
  bool userMirror{false};    // hostConc is the caller's buffer, page-locked here
  bool hostStale{true};
  bool downloadPending{false}; // a D2H of the state is in flight (xferStream)
  std::vector<double> hostDcdt;
  std::vector<double> hostLower;
  bool dcdtStale{true}, lowerStale{true};
  double *scratch{nullptr};
  // record of the last attempted step's final stage (getDcdt / lower order)
  bool haveLast{false};
  StageArgsHost last{};
  int lastCinBuf{0};
  uint32_t lastFlags{0};
};

// run-compressed neighbour codes of a voxel-list compartment (sim.cpp); false if
// the table does not have the reference's voxel-order structure
bool compressNeighbourTable(const uint32_t *nn, uint64_t n,
                            std::vector<uint32_t> &codes,
                            std::vector<int> &runs);

class EnsembleSim;

class Sim {
  friend class EnsembleSim;

public:
  explicit Sim(const smeb200_problem &p, bool compileOnlyMode = false);
  ~Sim();
  Sim(const Sim &) = delete;
  Sim &operator=(const Sim &) = delete;

  std::size_t run(double time, double timeoutMs, int (*stopCb)(void *),
                  void *user);
  const double *getConcentrations(std::size_t comp, std::size_t *n);
  const double *getDcdt(std::size_t comp, std::size_t *n);
  double getLowerOrder(std::size_t comp, std::size_t species, std::size_t pixel);
  void setConcentrations(std::size_t comp, const double *data, std::size_t n);
  // asynchronous state download: run() ends by queueing the D2H of every
  // compartment's state on a transfer stream; getConcentrations waits for it
  // (cudapixelsim.cpp:1486-1516,1657,1665-1669)
  void setAsyncDownload(bool enable) { asyncDownload = enable; }
  // the caller's buffer becomes the host mirror of a compartment (page-locked by
  // the library): getConcentrations returns it, no second copy on the caller's side
  void setHostMirror(std::size_t comp, double *buffer, std::size_t n);
  void getAvgMinMax(std::size_t comp, double *out, std::size_t nOut);
  const double *getSpeciesDecimated(std::size_t comp, std::size_t species,
                                    const uint32_t stride[3], std::size_t dims[3]);
  // new values of the runtime parameters: captured step graphs carry the old
  // ones in their kernel arguments and are dropped
  void setParams(const double *values, std::size_t n);
  double benchSteps(double dt, std::size_t nSteps);

  std::string errorMessage;
  bool ctorFailed{false};
  std::vector<uint32_t> errorImages;
  std::atomic<bool> stopRequested{false};
  // PixelSimBase state (pixelsim_base.hpp:40-49)
  double maxStableTimestep{std::numeric_limits<double>::max()};
  std::size_t discardedSteps{0};
  int integrator{SMEB200_RK212};
  double errMaxAbs{std::numeric_limits<double>::max()};
  double errMaxRel{0.005};
  double maxTimestep{std::numeric_limits<double>::max()};
  double nextTimestep{1e-7};
  double epsilon{1e-14};

  uint64_t launchCount{0};
  double lastRunMs{0}; // device time of the last run() (CUDA events)
  cudaEvent_t runEv0{nullptr}, runEv1{nullptr};
  bool kernelTiming{false};
  // every n-th step has its dominant-kernel launches bracketed by events (1 =
  // every step); the events cost a few microseconds of stream time per launch
  unsigned kernelTimingStride{1};
  std::size_t timingStep{0};
  double kernelMs{0};
  uint64_t kernelLaunchesTimed{0};
  std::string kernelSource;
  std::string dominantKernelName;
  bool compileOnly{false};
  std::size_t cubinBytes{0};
  std::size_t nCompartments() const { return comps.size(); }
  // bytes of neighbour-index data the stage kernels read per voxel: 0 (box,
  // implicit), 4 (run-compressed codes) or 24 (plain table)
  std::size_t neighbourBytesPerVoxel(std::size_t comp) const {
    const auto &c = comps.at(comp);
    return c.box ? 0 : (c.dRuns != nullptr ? 4 : 24);
  }
  std::size_t neighbourRuns(std::size_t comp) const { return comps.at(comp).nRuns; }
  // planes of the dominant (box) compartment the timed launch covers: all owned
  // planes, or the slab interior when the boundary planes are launched apart
  std::size_t timedLaunchPlanes() const {
    const auto &c = comps.at(dominantComp);
    std::size_t owned = c.nz - c.zLoGhost - c.zHiGhost;
    if (comm && haloStream != nullptr && !fusedHalo) {
      owned -= std::min<std::size_t>(owned, c.zLoGhost + c.zHiGhost);
    }
    return owned;
  }
  std::size_t concentrationPadding() const {
    std::size_t pad = 0;
    for (const auto &c : comps) {
      pad = std::max<std::size_t>(pad, c.S - c.SP);
    }
    return pad;
  }
  smeb200_allreduce_max_fn allreduceMax{nullptr};
  void *allreduceUser{nullptr};
  uint64_t haloExchanges{0};
  std::vector<double> params; // runtime model parameters (sme_param_k)
  uint64_t moduleCacheHits{0};

private:
  struct Module {
    CUmodule mod{nullptr};
    std::map<std::string, CUfunction> fn;
  };
  void build(const smeb200_problem &p);
  void generateSource(const smeb200_problem &p);
  Module compile(const std::vector<std::string> &names);
  std::string kernelName(std::size_t ci, uint32_t flags, bool memb) const;
  StageKernels lookup(const Module &m, std::size_t ci, uint32_t flags) const;
  StageKernels onDemandKernels(std::size_t ci, uint32_t flags);
  void chooseGeometry(int variant);
  void chooseSegments(int variant, std::vector<std::vector<uint32_t>> &nnHost);
  void encodeTensorMaps(CompartmentState &c);
  void setupMembranes(const smeb200_problem &p,
                      std::vector<std::vector<uint32_t>> &nnHost);
  void makePlan();
  void bindDevice() const;
  void doStep(double dt); // all stages of one RK step
  void launchAllStages(double dt);
  void graphStep(double dt);
  // CUDA-graph replay of a whole step (all stage launches of all compartments,
  // the error-norm download): for small models a step is launch-latency bound.
  // One graph per rotation state of the buffers; the step size is read from
  // device memory (dDt), refreshed from pinned hDt by the graph's first node.
  struct LastRecord {
    StageArgsHost last;
    int lastCinBuf;
    uint32_t lastFlags;
  };
  struct StepGraph {
    cudaGraphExec_t exec{nullptr};
    uint64_t launches{0};
    std::vector<LastRecord> last;
  };
  std::map<uint32_t, StepGraph> stepGraphs;
  bool useGraphs{false};
  bool capturing{false};
  double *hDt{nullptr};       // pinned
  double *dDt{nullptr};
  double graphDt{-1.0};
  std::vector<cudaStream_t> forkStreams; // capture-time parallel branches
  std::vector<cudaEvent_t> joinEvents;
  cudaEvent_t forkEvent{nullptr};
  cudaStream_t forkMain{nullptr}, forkMemb{nullptr};
  void launchStage(std::size_t stageIndex, double dt);
  // the arguments of compartment ci's kernels in plan stage `stageIndex` that do
  // not depend on the step: geometry tables, stage coefficients, constants
  StageArgsHost stageArgsTemplate(std::size_t ci, std::size_t stageIndex) const;
  // parts: 1 = main kernel over the range in `a` (z_begin..z_end / v0..n),
  // 2 = membrane kernel, 3 = both
  void launchCompartment(std::size_t ci, StageArgsHost &a,
                         const StageKernels &k, int cinBuf,
                         bool nsPrepass = false, unsigned parts = 3);
  // grid of the main / membrane kernel of one launch (host-side, no side effects)
  unsigned mainGrid(std::size_t ci, const StageArgsHost &a, CUfunction f);
  unsigned membGrid(std::size_t ci, CUfunction f);
  // kernel launch with the programmatic-stream-serialization attribute: the
  // kernel may become resident while its predecessor in the stream drains (the
  // stage kernels wait with griddepcontrol.wait before touching memory)
  void launchKernel(CUfunction f, unsigned gx, unsigned gy, unsigned gz,
                    unsigned smem, cudaStream_t s, void **params);
  bool usePdl{true};
  void rotateAfterAccept();
  void readError(double &eabs, double &erel);
  // unbounded tolerances (maxErr = {DBL_MAX, DBL_MAX}, the reference benchmark's
  // fixed-step configuration): the controller's decision does not depend on a
  // finite error norm, so the norms are checked kErrLag steps late instead of
  // stalling the stream after every step
  static constexpr std::size_t kErrRing = 8; // SMEB_ERR_RING
  // steps the host may run ahead of the norms it has looked at: 2 on one GPU; 4 in
  // a z-slab run, where a step is a few dozen driver calls on every rank and the
  // ranks meet in two host reductions per step (SMEB200_ERR_LAG = 1 .. 6)
  std::size_t errLag{2};
  struct PendingError {
    double dt, dtMax;
    uint32_t seq;
  };
  // z-slabs with non-spatial species: the spatial mean is summed over the ranks
  bool nsInSlabs{false};
  double *hNs{nullptr}; // pinned, 64 doubles
  bool laggedError{false};
  // Error norms of a step: the last CTA of the step's final-stage kernels writes
  // {abs, rel, sequence number} into this pinned, device-mapped ring
  // (device_kernels.cuh: smeb_err_arrive); nothing is queued behind the kernels
  unsigned long long *hErrRing{nullptr}; // 4 x kErrRing
  unsigned long long *dErrRing{nullptr}; // device address of hErrRing
  uint32_t errSeq{0};      // sequence number of the last publishing step launched
  bool errDesync{false};   // a failed launch sequence left the device counter behind
  uint32_t errExpected{0}; // CTAs of the error-norm launches of the current step
  // deferred publication (device_kernels.cuh: smeb_err_publish_previous): the
  // final stage of a step only leaves its maxima in the slot and the first kernel
  // launched after it — or smeb_err_flush — publishes them
  bool deferPublish{false};   // this run's final stages do not count their CTAs
  bool errUnpublished{false}; // a final stage's maxima wait in the slot
  CUfunction kErrFlush{nullptr};
  void flushErrors();
  void waitError(uint32_t seq, double v[2]);
  void resyncErrorSeq();
  PendingError errPending[kErrRing]{};
  std::size_t errHead{0}, errTail{0}; // slots [errHead, errTail) are in flight
  void enqueueErrorRead(double dt, double dtMax);
  bool drainErrors(std::size_t keep);
  double doRK101(double dt);
  double doRKAdaptive(double dtMax);
  void plotRKError(double relMax, std::string &problemSpecies);
  void runOnDemand(std::size_t ci, uint32_t flags, double *out);
  void evaluateLastDcdt();
  void flushTimers();
  void setupSlabs(const smeb200_problem &p);
  void connectSlabs();
  void haloPush(int coutBuf, cudaStream_t s);
  // boundary planes of arbitrary per-compartment buffers (nullptr: skip) into the
  // neighbours' ghost planes of the paired buffers, then the flags
  void haloPushPtrs(const std::vector<const double *> &local,
                    const std::vector<std::array<double *, 2>> &peer,
                    cudaStream_t s);
  void haloWait(cudaStream_t s);
  void haloQuiesce();
  void syncMaxStableTimestep();
  bool slabsConnected{false};
  bool haloPending{false};
  // halo exchange inside the tile kernel (all compartments box / TMA tile path):
  // own flag words in device memory, the neighbours' through CUDA IPC
  bool fusedHalo{false};
  bool zigzag{false}; // SMEB200_ZIGZAG=1: the tile kernels carry the zig-zag order
  // [2 ci + side] counts of boundary-plane tiles arrived from below / above per
  // compartment; [kHaloErrWord] raised by a wait that ran out
  static constexpr int kHaloWords = 16, kHaloErrWord = 15;
  uint32_t *dHalo{nullptr};
  uint32_t *peerHalo[2]{nullptr, nullptr};
  CUfunction kHaloQuiesce{nullptr};
  void checkHaloError();
  // this rank gives up: the other ranks' host waits throw and their device-side
  // halo waits are released
  void abortSlabs() noexcept;
  // template flags of a plan stage's kernels (on-demand variants never carry
  // F_HALO)
  uint32_t stageFlags(const StageDesc &sd) const {
    return sd.flags() | (fusedHalo ? F_HALO : 0u) | (zigzag ? F_ZIG : 0u);
  }
  cudaStream_t haloStream{nullptr};
  cudaEvent_t stageInputReady{nullptr}, haloPushed{nullptr}, boundaryDone{nullptr};
  bool suppressTiming{false};
  void updateCrossDiffusionMaxStableTimestep();
  void solveZeroStorageConstraints(std::size_t stageIndex);
  void encodeTensorMap(const CompartmentState &c, double *buffer,
                       CUtensorMap *out);
  bool anyZeroStorage{false};
  bool relaxing{false};
  double maxRelaxStableTimestep{std::numeric_limits<double>::max()};
  // per-launch overrides of launchCompartment (zero-storage relaxation)
  const CUtensorMap *tmapOverride{nullptr};
  std::vector<const double *> membCinOverride;
  // one full wave of resident CTAs for a grid-stride kernel: SMs x the CTAs of
  // this kernel that fit on one SM (cached per function)
  unsigned persistentGrid(CUfunction f, unsigned dynSmem = 0);
  std::map<CUfunction, unsigned> persistentGridOf;
  double h2[3]{1.0, 1.0, 1.0}; // voxel width^2, height^2, depth^2
  bool anyCross{false};
  unsigned long long *hXdMax{nullptr}; // pinned

  // z-slab decomposition (n_ranks > 1): communicator, neighbours' buffers
  // (CUDA IPC), device addresses of the stream-ordered handshake flags
  int rank{0}, nRanks{1};
  std::unique_ptr<ShmComm> comm;
  double *peerBuf[2][kCommMaxCompartments][kCommMaxBuffers]{};
  std::vector<void *> ipcOpened; // every mapping opened (closed once, at the end)
  bool flagsRegistered{false};
  CUdeviceptr flagBaseDev{0};
  uint32_t haloSeq{0};

  bool asyncDownload{false};
  cudaStream_t xferStream{nullptr};
  cudaEvent_t downloadDone{nullptr};
  uint64_t stateVersion{0}, downloadVersion{0}; // steps taken / state in flight
  void finishDownload();     // wait for the copy in flight
  void startDownload();      // end of run()
  void awaitDownloadOnStream(); // before anything overwrites a state buffer
  int device{0};
  cudaStream_t stream{nullptr};
  Module mainModule;
  double *dGather{nullptr}, *hGather{nullptr}; // decimated species image
  std::size_t gatherCapacity{0};
  std::vector<Module> auxModules;
  std::map<std::string, CUfunction> auxFn;
  std::vector<CompartmentState> comps;
  std::vector<StageDesc> plan;
  std::vector<double> fluxLen;
  double *dFluxLen{nullptr};
  uint32_t nMembranes{0};
  unsigned long long *dErr{nullptr};
  unsigned long long *dErrScratch{nullptr};
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> timerEvents;
  std::size_t timerUsed{0};
  int imageNx{0}, imageNy{0}, imageNz{0};
  int smCount{148};
  std::size_t dominantComp{0};
};

} // namespace smeb200
