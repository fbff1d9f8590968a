/*
 * smeb200.h — C-ABI of the B200-native Pixel simulator (libsmeb200.so).
 *
 * This is the drop-in boundary for ONE path of Spatial Model Editor: the Pixel
 * (finite-difference, explicit Runge-Kutta) simulator that lives behind
 * sme::simulate::BaseSim (reference core/simulate/src/basesim.hpp:15-55) and is
 * implemented there by PixelSim (pixelsim.cpp) and CudaPixelSim
 * (cudapixelsim.cpp).  An SME-side adapter class (adapter/b200pixelsim.hpp)
 * fills `smeb200_problem` from sme::model::Model exactly as
 * cudapixelsim.cpp:857-1150 does and forwards every BaseSim virtual to one of
 * the entry points below.  Plain pointers and sizes only: no C++ types, no
 * exceptions, no torch types.  Errors are int codes + smeb200_error().
 *
 * All floating point data is fp64.  Concentrations are species-interleaved
 * ("AoS"): value of species s at voxel i is conc[i * n_species + s], voxels in
 * the order of geometry::Compartment (z outermost, then x, then y;
 * geometry.cpp:55-72).
 */
#ifndef SMEB200_H
#define SMEB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SMEB200_ABI_VERSION 3

/* mirrors sme::simulate::PixelIntegratorType (simulate_options.hpp:113-119) */
enum smeb200_integrator {
  SMEB200_RK101 = 0,
  SMEB200_RK212 = 1,
  SMEB200_RK323 = 2,
  SMEB200_RK435 = 3
};

/* membrane face directions, geometry::Membrane::FACE_DIRECTION
 * (geometry.cpp:18-36): sign of (voxel B - voxel A), x first, then y, then z */
enum smeb200_face { SMEB200_XP = 0, SMEB200_XM, SMEB200_YP, SMEB200_YM,
                    SMEB200_ZP, SMEB200_ZM };

/* bits of smeb200_problem.unsupported: features of the CPU PixelSim that this
 * library does not compute.  The adapter sets them; smeb200_create then fails
 * with the message text of the reference GPU backend (cudapixelsim.cpp:789-809,
 * pixelsim_common.hpp:80-121).  Time / space dependent reactions, non-spatial
 * species and non-unit storage ARE computed (describe them through
 * smeb200_compartment::n_primary_species / non_spatial / inv_storage); their
 * bits remain for adapters that prefer to refuse such models. */
enum smeb200_unsupported {
  SMEB200_HAS_CROSS_DIFFUSION = 1,
  SMEB200_HAS_TIME_DEPENDENCE = 2,
  SMEB200_HAS_NON_SPATIAL_SPECIES = 4,
  SMEB200_HAS_NON_UNIFORM_DIFFUSION = 8,
  SMEB200_HAS_NON_UNIT_STORAGE = 16,
  SMEB200_HAS_SPACE_DEPENDENCE = 32
};

/* One geometry::Compartment + SimCompartment (pixelsim_impl.hpp:58-322). */
typedef struct smeb200_compartment {
  const char *id;          /* compartment id (diagnostics only) */
  uint64_t n_voxels;       /* Compartment::nVoxels() */
  uint32_t n_species;      /* simulated species in this compartment */
  /* Geometry, one of:
   *  - box: nn == NULL and box_nx*box_ny*box_nz == n_voxels: the compartment
   *    fills its whole image; neighbours are implicit (zero-flux Neumann at
   *    the faces, i.e. missing neighbour = own index, geometry.cpp:98-110);
   *  - voxel list: nn[6*i + k] = index of the +x,-x,+y,-y,+z,-z neighbour of
   *    voxel i (own index if outside the compartment), as uint32 like the
   *    reference GPU backend (cudapixelsim.cpp:877-890). */
  int32_t box_nx, box_ny, box_nz;
  const uint32_t *nn;
  /* optional [3*n_voxels] image coordinates (x, y, z) of each voxel, used for
   * errorImages (pixelsim_impl.cpp:794-824) and for z-slab partitioning;
   * may be NULL (implied for a box) */
  const int32_t *voxels;
  /* [3*n_species] dimensionless diffusion D/dx^2, D/dy^2, D/dz^2 per species
   * (pixelsim_impl.cpp:236-237) */
  const double *diffusion;
  /* [n_voxels*n_species] initial (or restart) concentrations */
  const double *conc0;
  /* n_species CUDA-C expressions for d(species)/dt from reactions, written in
   * terms of sme_var_0 .. sme_var_{n_species-1}: exactly the strings
   * Symbolic::cudaCode(i) returns after Symbolic::relabel
   * (cudapixelsim.cpp:108-136) */
  const char *const *expressions;
  const char *const *species_names; /* optional, for the adaptive-failure text */
  /* --- features of the CPU PixelSim beyond the reference GPU backend --------
   * n_primary_species: the first n_primary_species of the n_species state
   * components are model species; the remaining ones are the "time, x, y, z"
   * padding the CPU solver appends when reactions depend on time / space
   * (pixelsim_impl.cpp:312-349,405-423): ordinary state components with zero
   * diffusion whose expressions are "1" (time) and "0" (x, y, z), initialised
   * by the adapter to 0 and to the voxel-centre coordinates.  They are not
   * clamped (pixelsim_impl.cpp:643-654) and are what getConcentrationPadding
   * reports (pixelsim.cpp:478).  0 means n_species (no padding). */
  uint32_t n_primary_species;
  /* species whose dc/dt is replaced by its spatial average before the update
   * (SimCompartment::spatiallyAverageDcdt, pixelsim_impl.cpp:150-163) */
  uint32_t n_non_spatial;
  const uint32_t *non_spatial;  /* [n_non_spatial] species indices */
  /* [n_species] 1/storage per state component (pixelsim_impl.cpp:165-189,
   * 214-232); NULL = unit storage.  A zero entry marks a zero-storage species
   * (algebraic constraint 0 = R(c) + div(D grad c)): it is relaxed in
   * pseudo-time before every Runge-Kutta stage
   * (PixelSim::solveZeroStorageConstraints, pixelsim.cpp:23-90), its dc/dt is
   * zero and it is left out of the error norm (pixelsim_impl.cpp:778-781).
   * In a z-slab decomposed simulation the iterates exchange their boundary
   * planes and the norms are max-reduced over the ranks. */
  const double *inv_storage;
  /* Non-uniform diffusion (SimCompartment::evaluateDiffusionOperator, array
   * branch, pixelsim_impl.cpp:453-485): [n_voxels*n_species] diffusion
   * constant of species s at voxel i at diffusion_array[i*n_species + s] (NOT
   * divided by the voxel size; the face-averaged stencil divides by
   * smeb200_problem.voxel_size^2 per face).  The reference uses this branch
   * for EVERY compartment as soon as one species of the model has a
   * non-uniform constant (pixelsim.cpp:323-336), so the adapter passes arrays
   * for all compartments or for none.  NULL = uniform (`diffusion` is used). */
  const double *diffusion_array;
  /* Cross-diffusion terms D_ts(c) grad c_s added to species t
   * (pixelsim_impl.cpp:350-389,494-589): per term the target and source
   * species index and a CUDA-C expression of the coefficient in
   * sme_var_0..sme_var_{n_species-1}, evaluated at each voxel's own state and
   * face-averaged.  With cross-diffusion `diffusion_raw` ([n_species] diffusion
   * constants, or their per-species maxima, not divided by the voxel size;
   * maxPrimaryDiagonalDiffusion of pixelsim_impl.cpp:256,291) is required: the
   * forward-Euler bound is then re-evaluated every step
   * (updateCrossDiffusionMaxStableTimestep, pixelsim_impl.cpp:506-543). */
  uint32_t n_cross_diffusion;
  const uint32_t *cross_target;
  const uint32_t *cross_source;
  const char *const *cross_expressions;
  const double *diffusion_raw;
  /* z-slab decomposition of a VOXEL-LIST compartment (n_ranks > 1, nn != NULL).
   * The voxel order is z-major, so the voxels of the planes a rank owns are a
   * contiguous range of the list; the local list is
   *   [slab_ghost_lo voxels of the plane below][owned voxels][slab_ghost_hi
   *    voxels of the plane above],
   * nn holds LOCAL indices (rows of ghost voxels are ignored), membrane pairs
   * hold local indices and list every pair with at least one owned voxel.
   * slab_first_plane / slab_last_plane = number of owned voxels in the first /
   * last owned plane: they are what the lower / upper neighbour rank holds as
   * its slab_ghost_hi / slab_ghost_lo voxels. */
  uint32_t slab_ghost_lo, slab_ghost_hi, slab_first_plane, slab_last_plane;
} smeb200_compartment;

/* One geometry::Membrane + SimMembrane (pixelsim_impl.cpp:929-1047). */
typedef struct smeb200_membrane {
  const char *id;
  int32_t comp_a, comp_b;  /* indices into compartments[], -1 if not simulated */
  /* per face direction: n_pairs[d] (ixA, ixB) uint32 pairs, in the order of
   * Membrane::getFaceIndexPairs(d) (geometry.cpp:165-172) */
  uint64_t n_pairs[6];
  const uint32_t *pairs[6];
  /* voxel length in the flux direction per face (pixelsim_common.hpp:39-54) */
  double flux_length[6];
  /* (SA + SB) CUDA-C expressions in sme_var_0..sme_var_{SA+SB+E-1}: variables
   * are [primary species of A | primary species of B | the E padding
   * components (time, x, y, z) of B's voxel, or of A's if B is not simulated],
   * results [0,SA) go to A and [SA,SA+SB) to B (pixelsim_impl.cpp:960-973,
   * 84-111); already scaled by volOverL3.  SA, SB = n_primary_species. */
  const char *const *expressions;
} smeb200_membrane;

typedef struct smeb200_problem {
  uint32_t abi_version;    /* SMEB200_ABI_VERSION */
  uint32_t n_compartments;
  const smeb200_compartment *compartments;
  uint32_t n_membranes;
  const smeb200_membrane *membranes;
  /* PixelOptions (simulate_options.hpp:160-200) */
  int32_t integrator;      /* enum smeb200_integrator */
  double max_err_abs;      /* PixelIntegratorError::abs */
  double max_err_rel;      /* PixelIntegratorError::rel */
  double max_timestep;
  /* geometry image size, for errorImages */
  int32_t image_nx, image_ny, image_nz;
  uint32_t unsupported;    /* enum smeb200_unsupported bits */
  int32_t device;          /* CUDA device ordinal; -1 = current device */
  /* z-slab decomposition over `n_ranks` processes on one node, one GPU each
   * (0/1 = single GPU).  Rank r then describes only ITS slab.  Box
   * compartments: box_nz = owned planes + one ghost plane below (r > 0) + one
   * above (r < n_ranks-1), conc0 including the ghost planes.  Voxel-list
   * compartments (and membranes between them): see
   * smeb200_compartment::slab_ghost_lo.  See "multi-GPU". */
  int32_t rank, n_ranks;
  /* kernel variant: 0 = auto (TMA tile-marching kernel for box compartments
   * where the shape allows it), 1 = direct-gather kernels only, 2 = require
   * the TMA tile kernel for box compartments (error if it cannot be used),
   * 3 = keep the plain six-index neighbour table of voxel-list compartments
   * (default: run-compressed to one 32-bit code per voxel where possible),
   * 4 = voxel-list compartments use the segment-staged kernel where their runs
   * allow it (pieces of runs staged in shared memory by bulk copies; measured
   * level with the direct kernel, kept for comparison) */
  int32_t kernel_variant;
  /* runtime model parameters: expressions may refer to sme_param_0 ..
   * sme_param_{n_params-1}; their values are kernel arguments, so simulators
   * that differ only in these values (parameter scans, optimizer candidates)
   * share one compiled module, and smeb200_set_params changes them between
   * runs.  The reference inlines parameter values into the expression text
   * and re-JITs per member (pde.cpp:65-77). */
  uint32_t n_params;
  /* name of the node-local communicator (POSIX shared memory segment) the
   * n_ranks processes of one decomposed simulation rendezvous on; must be the
   * same string on all ranks and unique per simulation; NULL if n_ranks <= 1 */
  const char *comm_name;
  const double *params;    /* [n_params] */
  /* voxel width, height, depth (model units): needed with diffusion_array or
   * cross-diffusion only */
  double voxel_size[3];
  uint32_t reserved[2];
} smeb200_problem;

typedef struct smeb200_sim smeb200_sim;

/* number of usable CUDA devices, 0 if none; never fails */
int smeb200_device_count(void);

/* Constructor (PixelSim::PixelSim / CudaPixelSim::CudaPixelSim).  Like the
 * reference constructors it never throws: on failure a simulator object is
 * still returned whose smeb200_error() is non-empty and whose run() returns 0
 * (cudapixelsim.cpp:1156-1167,1621-1623).  Returns 0 on success, non-zero if
 * the simulator carries an error. *out is NULL only if p or out is NULL. */
int smeb200_create(const smeb200_problem *p, smeb200_sim **out);
void smeb200_destroy(smeb200_sim *sim);
/* Generate the model's CUDA module and compile it with NVRTC for sm_100a
 * without touching a device (works on a machine with no GPU).  The returned
 * object only serves smeb200_kernel_source / smeb200_error / smeb200_destroy;
 * its run() returns 0.  Returns 0 if the module compiled. */
int smeb200_compile_only(const smeb200_problem *p, smeb200_sim **out);

/* BaseSim::run (basesim.hpp:24-27; pixelsim.cpp:423-471): integrate for
 * `time`, returns the number of accepted steps.  timeout_ms < 0 = none.
 * stop_cb (may be NULL) is polled once per step; non-zero = stop. */
size_t smeb200_run(smeb200_sim *sim, double time, double timeout_ms,
                   int (*stop_cb)(void *user), void *user);

/* BaseSim::getConcentrations (host mirror, valid until the next run/destroy) */
int smeb200_get_concentrations(smeb200_sim *sim, size_t compartment,
                               const double **data, size_t *n);
/* PixelSim::getDcdt (pixelsim.cpp:481-484): last evaluated dc/dt */
int smeb200_get_dcdt(smeb200_sim *sim, size_t compartment, const double **data,
                     size_t *n);
/* PixelSim::getLowerOrderConcentration (pixelsim.cpp:486-490) */
int smeb200_get_lower_order(smeb200_sim *sim, size_t compartment,
                            size_t species, size_t pixel, double *value);
/* Per-species {average, minimum, maximum} of a compartment's current state,
 * reduced on the device: what Simulation::updateConcentrations computes from
 * the downloaded state after every image interval (calculateAvgMinMax,
 * simulate.cpp:272-290; minimum starts at DBL_MAX and maximum at 0 like
 * sme::simulate::AvgMinMax).  out[3*s .. 3*s+2] for the n_primary_species model
 * species; n_out must be 3 * n_primary_species. */
int smeb200_get_avg_min_max(smeb200_sim *sim, size_t compartment, double *out,
                            size_t n_out);
/* The same reduction for callers that cannot allocate (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart
 * in the reference.
 *
 * d_state: DEVICE pointer to n_voxels x values_per_voxel doubles, the layout of
 * smeb200_get_concentrations (the first n_primary_species values of a voxel are
 * the model species).  d_out: DEVICE pointer to 3 * n_primary_species doubles,
 * out[3*s .. 3*s+2] = {average, minimum, maximum} as above, written by two
 * kernels enqueued on `stream` (a cudaStream_t; NULL = the default stream) of
 * the CURRENT device; the call does not synchronise.  It calls no cudaMalloc,
 * cudaFree or allocating library routine: the block partials go into one
 * file-scope device array of SMEB200_NO_ALLOC_SCRATCH_BYTES that every GPU
 * loading the library holds until the process exits.  All calls on a device
 * share that array, which keeps the last call's contents (nothing in it has to
 * start at zero: a call overwrites every slot it later reads); calls that
 * overlap on the GPU race, so serialise them (one stream, or events).
 * The first call on a device resolves the array's address
 * (cudaGetSymbolAddress); with n_voxels = 0 a call does only that, which is the
 * way to prepare a device before a stream capture begins.
 * Returns 0, or 1 with a message in err[0 .. err_len) (may be NULL) when a size
 * exceeds the limits below, a pointer is NULL or a CUDA call fails. */
#define SMEB200_NO_ALLOC_MAX_SPECIES 16
#define SMEB200_NO_ALLOC_MAX_VALUES_PER_VOXEL 64
#define SMEB200_NO_ALLOC_MAX_VOXELS 2147483648ull /* 32-bit voxel indices */
#define SMEB200_NO_ALLOC_MAX_BLOCKS 2048
#define SMEB200_NO_ALLOC_SCRATCH_BYTES 786432 /* 2048 blocks x 16 species x 3 x 8 */
int smeb200_avg_min_max_no_alloc(const double *d_state, uint64_t n_voxels,
                                 uint32_t values_per_voxel,
                                 uint32_t n_primary_species, double *d_out,
                                 void *stream, char *err, size_t err_len);
/* One species of every (stride[0], stride[1], stride[2]) = (x, y, z)-th voxel of a
 * compartment's current state, gathered on the device and downloaded into a
 * pinned buffer owned by the simulator (valid until the next call on `sim`):
 * the decimated image a front end draws after an image interval instead of the
 * whole state Simulation::updateConcentrations copies (simulate.cpp:313-314,
 * SURVEY 8(f) rank 2).  Box compartments: data[(zi * mx + xi) * my + yi] with
 * mx = ceil(nx / stride[0]) etc., dims[0..2] = {mx, my, mz} (z-slabs: the owned
 * planes).  Voxel-list compartments: every stride[0]-th voxel of the (owned)
 * list, stride[1] = stride[2] = 1, dims = {n_out, 1, 1}.  `species` counts the
 * stored values of a voxel (padding included). */
int smeb200_get_species_decimated(smeb200_sim *sim, size_t compartment,
                                  size_t species, const uint32_t stride[3],
                                  const double **data, size_t dims[3]);
/* The device layout of a voxel-list compartment's neighbour table, exposed for
 * inspection and tests (no device needed): nn[6*i+k] as in
 * smeb200_compartment -> one code per voxel (bits 0-24 run id, bits 25-30
 * "neighbour exists" for +x,-x,+y,-y,+z,-z, bit 31 membrane flag) and four index
 * deltas (+x,-x,+z,-z) per run.  On entry *n_runs is the capacity of `runs` in
 * runs (codes needs n_voxels entries; either may be NULL), on return the number
 * of runs.  Returns 0, or 1 if the table cannot be compressed. */
int smeb200_compress_neighbours(const uint32_t *nn, uint64_t n_voxels,
                                uint32_t *codes, int32_t *runs,
                                uint64_t *n_runs);
/* neighbour-index bytes the stage kernels read per voxel of a compartment: 0
 * (box, implicit neighbours), 4 (run-compressed codes) or 24 (plain table); and
 * the number of runs of the compressed layout */
size_t smeb200_neighbour_bytes_per_voxel(const smeb200_sim *sim, size_t compartment);
size_t smeb200_neighbour_runs(const smeb200_sim *sim, size_t compartment);
/* BaseSim::getConcentrationPadding (pixelsim.cpp:478): number of time / x, y, z
 * components behind the model species of every voxel */
size_t smeb200_get_concentration_padding(const smeb200_sim *sim);
/* restart: overwrite the state of one compartment (SimCompartment::
 * setConcentrations, used with SimulationData, pixelsim.cpp:399-406) */
int smeb200_set_concentrations(smeb200_sim *sim, size_t compartment,
                               const double *data, size_t n);

/* Asynchronous state download, as the reference GPU backend does it
 * (cudapixelsim.cpp:1486-1516,1657,1665-1669): with enable != 0 every
 * smeb200_run ends by queueing the device-to-host copy of every compartment's
 * state on a transfer stream and returns; smeb200_get_concentrations waits for
 * that copy instead of starting one.  A following run / set_concentrations is
 * ordered after the copy on the device.  Off by default (a caller that never
 * reads the state should not pay for the copy); the SME adapter switches it on,
 * Simulation::updateConcentrations reads the state after every run. */
void smeb200_set_async_download(smeb200_sim *sim, int enable);
/* The caller's buffer (n = n_voxels * n_species doubles, alive until
 * smeb200_destroy or the next call for this compartment) becomes the host
 * mirror of the compartment: the library page-locks it, downloads into it and
 * smeb200_get_concentrations returns it — the adapter's std::vector behind
 * BaseSim::getConcentrations is filled without a second host copy. */
int smeb200_set_host_mirror(smeb200_sim *sim, size_t compartment, double *buffer,
                            size_t n);

/* new values for the runtime model parameters (takes effect at the next run) */
int smeb200_set_params(smeb200_sim *sim, const double *values, size_t n);

/* BaseSim::errorMessage / setCurrentErrormessage */
const char *smeb200_error(const smeb200_sim *sim);
void smeb200_set_error(smeb200_sim *sim, const char *msg);
/* BaseSim::errorImages: image_nz RGB32 images of image_nx*image_ny pixels
 * (0x00RRGGBB, row-major, y = 0 at the top like QImage), empty (n = 0) unless
 * the adaptive controller failed (pixelsim.cpp:236-253) */
int smeb200_error_images(const smeb200_sim *sim, const uint32_t **data,
                         size_t *n);
/* BaseSim::getStopRequested / setStopRequested (thread-safe) */
int smeb200_get_stop_requested(const smeb200_sim *sim);
void smeb200_set_stop_requested(smeb200_sim *sim, int stop);

/* controller state (PixelSimBase, pixelsim_base.hpp:40-49) */
double smeb200_get_next_timestep(const smeb200_sim *sim);
void smeb200_set_next_timestep(smeb200_sim *sim, double dt);
size_t smeb200_get_discarded_steps(const smeb200_sim *sim);
double smeb200_get_max_stable_timestep(const smeb200_sim *sim);

/* ---- measurement hooks (bench.py; not part of BaseSim) ------------------ */
/* Take `n_steps` steps of fixed size dt without host-side control: no
 * download, state stays resident in HBM.  Returns device milliseconds for
 * the n_steps (CUDA events on the launching stream), <0 on error. */
double smeb200_bench_steps(smeb200_sim *sim, double dt, size_t n_steps);
/* device time (ms, CUDA events on the launching stream) of the last
 * smeb200_run call */
double smeb200_last_run_ms(const smeb200_sim *sim);
/* kernels launched since construction */
uint64_t smeb200_launch_count(const smeb200_sim *sim);
/* device time of the dominant stage kernel over the last bench/run window:
 * sum of ms and number of launches (CUDA events, only if profiling enabled
 * with smeb200_set_kernel_timing(sim, n): n = 1 brackets the launches of every
 * step, n > 1 those of every n-th step — the two event records cost a few
 * microseconds of stream time per launch, which shows on 100 us kernels) */
void smeb200_set_kernel_timing(smeb200_sim *sim, int enable);
int smeb200_get_kernel_timing(const smeb200_sim *sim, double *total_ms,
                              uint64_t *launches);
/* generated CUDA source of the model module (for inspection / tests) */
const char *smeb200_kernel_source(const smeb200_sim *sim);
/* name expression of the stage kernel that set_kernel_timing times: the final
 * stage of the largest compartment, e.g.
 * "smeb_stage_tile<Reac0, 113u, 16, 64, 5, true>" */
const char *smeb200_dominant_kernel(const smeb200_sim *sim);

/* ---- multi-GPU ---------------------------------------------------------- */
/* z-slab decomposition (no counterpart in the reference, which has no
 * distributed layer): with n_ranks > 1 every rank's simulator
 *  - updates only its owned planes; in every Runge-Kutta stage its two boundary
 *    planes also go into the neighbours' ghost planes over NVLink (CUDA IPC peer
 *    memory).  Box compartments on the TMA tile path do that inside the stage
 *    kernel: the CTAs of a boundary plane store their results a second time
 *    through the peer mapping and the last of them releases a sequence number
 *    into a flag word in the neighbour's memory, which the neighbour's CTAs
 *    acquire before they load the ghost plane — one kernel per stage, no copy
 *    engine, no stream waits.  Voxel-list compartments and membranes: peer
 *    copies and stream-ordered 32-bit flags (cuStreamWriteValue32 /
 *    cuStreamWaitValue32) on a high-priority stream beside the slab interior.
 *    No host involvement either way;
 *  - max-reduces the two error norms over ranks once per adaptive step through
 *    the shared-memory communicator, so that all ranks take identical
 *    step-size decisions.
 * smeb200_get_concentrations returns the local slab including ghost planes.
 *
 * The communicator on its own (usable without a GPU; CPU tests): */
typedef struct smeb200_comm smeb200_comm;
int smeb200_comm_create(const char *name, int rank, int n_ranks,
                        smeb200_comm **out);
int smeb200_comm_allreduce_max(smeb200_comm *comm, double *values, int n);
/* in-place sum over the ranks of n doubles, added in rank order on every
 * rank (all ranks get the same bits): the spatial mean of non-spatial species in
 * a z-slab decomposed simulation (SimCompartment::spatiallyAverageDcdt,
 * pixelsim_impl.cpp:150-163, across slabs) */
int smeb200_comm_allreduce_sum(smeb200_comm *comm, double *values, int n);
int smeb200_comm_barrier(smeb200_comm *comm);
/* this rank gives up (its run failed): every wait of the other ranks — barrier,
 * reductions — returns an error from now on instead of spinning to its timeout */
void smeb200_comm_abort(smeb200_comm *comm);
void smeb200_comm_destroy(smeb200_comm *comm);
/* z planes of the largest (box) compartment that the launch timed by
 * smeb200_set_kernel_timing covers: all owned planes, or the slab interior when
 * the boundary planes are launched apart (bench.py: roofline of a slab) */
size_t smeb200_timed_launch_planes(const smeb200_sim *sim);
/* halo exchanges performed since construction (0 on a single GPU) */
uint64_t smeb200_halo_exchanges(const smeb200_sim *sim);

/* alternative error-norm exchange for the adaptive controller: the host
 * supplies an allreduce-max over ranks for 2 doubles (NCCL/gloo in the caller);
 * used instead of the communicator's when set */
typedef void (*smeb200_allreduce_max_fn)(double *values, int n, void *user);
void smeb200_set_allreduce_max(smeb200_sim *sim, smeb200_allreduce_max_fn fn,
                               void *user);

/* ---- member-batched ensembles -------------------------------------------------
 * A parameter scan / a population of optimizer candidates: n_members simulations
 * of ONE model that differ only in the values of the runtime parameters
 * (smeb200_problem::n_params).  The reference has no counterpart — a scan is a
 * loop that sets a parameter and simulates again (sme_model.cpp:541-617), the
 * optimizer evaluates its candidates one simulation at a time
 * (optimize_impl.cpp:150-186) — and every member re-compiles the reactions
 * because parameter values are inlined into the expression text
 * (pde.cpp:65-77).  Here the members share geometry tables and one compiled
 * module, the state is [member][voxel][species], a Runge-Kutta stage of all
 * members of a compartment is ONE kernel launch (members = gridDim.y), and every
 * member has its own step-size controller (PixelSim::run / doRKAdaptive per
 * member, evaluated after each batched step), so a member takes exactly the
 * steps it would take alone.  member_params: [n_members][n_params], row m = the
 * values of member m (p->params is ignored).  Not available with z-slab
 * decomposition (an ensemble lives on one GPU; shard members over GPUs by
 * creating one ensemble per rank), zero-storage / non-spatial species or
 * cross-diffusion.  Error conventions as for smeb200_create. */
typedef struct smeb200_ensemble smeb200_ensemble;
int smeb200_ensemble_create(const smeb200_problem *p, uint32_t n_members,
                            const double *member_params, smeb200_ensemble **out);
void smeb200_ensemble_destroy(smeb200_ensemble *ens);
/* every member integrates [0, time] (BaseSim::run per member); returns the sum of
 * the members' accepted steps.  timeout / stop_cb are polled once per batched
 * step and stop all members ("Simulation stopped early"). */
size_t smeb200_ensemble_run(smeb200_ensemble *ens, double time, double timeout_ms,
                            int (*stop_cb)(void *user), void *user);
/* BaseSim::getConcentrations of one member (pinned mirror, valid until the next
 * run) and the restart path (SimCompartment::setConcentrations) */
int smeb200_ensemble_get_concentrations(smeb200_ensemble *ens, uint32_t member,
                                        size_t compartment, const double **data,
                                        size_t *n);
int smeb200_ensemble_set_concentrations(smeb200_ensemble *ens, uint32_t member,
                                        size_t compartment, const double *data,
                                        size_t n);
/* per member: accepted / discarded steps of the last run, controller state */
size_t smeb200_ensemble_get_steps(const smeb200_ensemble *ens, uint32_t member);
size_t smeb200_ensemble_get_discarded_steps(const smeb200_ensemble *ens,
                                            uint32_t member);
double smeb200_ensemble_get_next_timestep(const smeb200_ensemble *ens,
                                          uint32_t member);
void smeb200_ensemble_set_next_timestep(smeb200_ensemble *ens, uint32_t member,
                                        double dt);
/* member >= 0: that member's error (adaptive failure), empty = OK; member < 0: the
 * ensemble's own error (construction, stop, device failure) */
const char *smeb200_ensemble_error(const smeb200_ensemble *ens, int member);
void smeb200_ensemble_set_stop_requested(smeb200_ensemble *ens, int stop);
/* measurement hooks */
uint64_t smeb200_ensemble_launch_count(const smeb200_ensemble *ens);
uint64_t smeb200_ensemble_batched_steps(const smeb200_ensemble *ens);
double smeb200_ensemble_last_run_ms(const smeb200_ensemble *ens);

#ifdef __cplusplus
}
#endif
#endif /* SMEB200_H */
