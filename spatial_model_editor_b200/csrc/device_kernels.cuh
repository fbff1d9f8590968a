// smeb200 device kernels for the Pixel simulator hot path (sm_100a, fp64).
//
// This file is the *kernel template* of the library: it is embedded verbatim
// into libsmeb200.so and compiled once per model by NVRTC together with the
// model's generated reaction functors (struct Reac<k>, struct Memb<k>), with
// --gpu-architecture=sm_100a --fmad=false.  It is also compiled at build time
// by nvcc with a default model (default_model.cu) so that ptxas/SASS can be
// inspected without a GPU.  No #include: NVRTC sees exactly this text.
//
// What the kernels fuse, per Runge-Kutta stage and per voxel (one pass over
// HBM instead of the reference's 4-6 passes):
//   dcdt  = R(c)                                   pixelsim_impl.cpp:488-492
//         + Dx(c[+x]+c[-x]-2c)+Dy(..)+Dz(..)       pixelsim_impl.cpp:434-452
//         + sum of membrane fluxes / L_face        pixelsim_impl.cpp:80-113
//   s2   += delta*c ; c' = g1 c + g2 s2 + g3 s3 + beta dt dcdt
//                                                  pixelsim_impl.cpp:625-721
//   err   = max|c'-lower| , max|c'-lower|/(0.5(c'+s3+eps))
//                                                  pixelsim_impl.cpp:774-792
//   clamp = max(c',0) at the store of the last stage (the state is only used
//           if the step is accepted)               pixelsim_impl.cpp:643-665
// The floating point association order of every expression follows the cited
// reference lines so that fixed-step results agree to the last bits.
//
// Stage flags are a compile-time template argument: the host instantiates one
// kernel per distinct stage of the chosen integrator (2 for RK212).

typedef unsigned long long smeb_u64;
typedef unsigned int smeb_u32;

#define SMEB_F_USE_C0 1u         // update reads the pre-step state c0 (s3)
#define SMEB_F_USE_S2 2u         // running sum s2 is live (RK323/RK435)
#define SMEB_F_USE_S2IN 4u       // s2 read from memory (else starts at 0)
#define SMEB_F_WRITE_S2 8u       // s2 written back
#define SMEB_F_FINAL 16u         // last stage: fused error norm
#define SMEB_F_LOWER_IS_CIN 32u  // RK212: lower order solution = stage input
#define SMEB_F_CLAMP_OUT 64u     // clamp output to >= 0 at the store
#define SMEB_F_GENERIC 128u      // update = g1 c + g2 s2 + g3 c0 + bdt f
                                 // (else the two-term forms of RK101/RK212)
#define SMEB_F_OUT_DCDT 256u     // on demand: store dc/dt, nothing else
#define SMEB_F_OUT_LOWER 512u    // on demand: store the lower-order solution
#define SMEB_F_OUT_CNEW 1024u    // on demand: store the unclamped new state
#define SMEB_F_SUM_NS 2048u      // pre-pass: block sums of dc/dt of the
                                 // non-spatial species, nothing else
#define SMEB_F_RELAX_A 4096u     // zero-storage relaxation, first half: residual
                                 // norms, out = c + dt f for those species
#define SMEB_F_RELAX_B 8192u     // second half: out = c0/2 + c/2 + dt/2 f, error
#define SMEB_F_HALO 16384u       // z-slabs, tile kernel: halo exchange in the kernel
#define SMEB_F_MEMBERS 32768u    // member-batched ensemble: gridDim.y = members
#define SMEB_F_ZIG 65536u        // tile kernel, 3-D: zig-zag chunk order (opt-in)

#define SMEB_MEMB_FLAG 0x80000000u
#define SMEB_BLOCK 256
#ifndef SMEB_NPARAMS
#define SMEB_NPARAMS 0 // runtime model parameters (sme_param_k), per module
#endif
#ifndef SMEB_C0_AHEAD
// planes the register prefetch of c0 / s2 runs ahead of their use; 0 = chosen per
// kernel: 2 when a stage reads c0 and not s2 (B200, 512^3 RK212 stage 2: 87 % ->
// 97 % of the HBM copy rate — one plane of lead, about 2 us, is less than the
// loaded DRAM latency), 1 when it also streams s2 (RK435: a deeper second queue
// costs 16 more registers and the kernels get slower)
#define SMEB_C0_AHEAD 0
#endif
#ifndef SMEB_C0_SPLIT
#define SMEB_C0_SPLIT 0 // 1: c0 two planes ahead also in stages that stream s2 (A/B)
#endif
#ifndef SMEB_NN_CODE_AHEAD
#define SMEB_NN_CODE_AHEAD 2 // voxel-list kernel: what is fetched one iteration early (see smeb_stage_nn)
#endif
#ifndef SMEB_TILE_MIN_CTAS
#define SMEB_TILE_MIN_CTAS 2 // resident CTAs per SM the tile kernel is built for
#endif

template <int S> struct SmebStageArgs {
  const double *cin;   // stage input: RHS is evaluated on this state
  const double *c0;    // pre-step state (reference: s3)
  const double *s2in;  // low-storage running sum in (reference: s2)
  double *s2out;       // running sum out (may alias s2in)
  double *cout;        // stage output (or the on-demand output)
  const smeb_u32 *nn;  // voxel-list geometry: [6 n] neighbour indices, or (runs
                       // != 0) [n] run-compressed neighbour codes
  const int *runs;     // [4 n_runs] index deltas of the +x, -x, +z, -z neighbours
  const smeb_u32 *seg; // segment-staged voxel-list kernel: [2 nSeg] (first voxel,
                       // length) of pieces of runs, at most SEG voxels each
  smeb_u64 *err;       // [2] bit patterns of max abs / max rel error of this
                       // simulation (ensembles: of this member, errBase + 2 m)
  smeb_u64 *errBase;   // error slot: errPairs pairs, then two u32: CTAs arrived,
                       // steps published
  smeb_u64 *errHost;   // pinned host ring the last CTA of a step publishes into
  // member-batched ensembles (SMEB_F_MEMBERS): gridDim.y = members; a member's
  // buffers, step size and parameter values come from these tables
  const struct SmebMember *members; // [nMembers] of this compartment
  const double *memberParams;       // [nMembers][SMEB_NPARAMS]
  const double *nsAv;  // [R::NNS] spatial averages of dc/dt (non-spatial species)
  double *nsPartial;   // [blocks][R::NNS] block sums written by the pre-pass
  const double *darr;  // [n][S] per-voxel diffusion constants (R::NUD)
  smeb_u64 *xdMax;     // [R::NX] max |cross-diffusion coefficient| (bit patterns)
  double h2[3];        // voxel width^2, height^2, depth^2 (R::NUD / R::NX)
  smeb_u32 n;          // voxel list: one past the last voxel updated
  smeb_u32 v0;         // voxel list: first voxel updated (z-slab ghosts before it)
  smeb_u32 nx, ny, nz; // box dims; voxel = (z*nx + x)*ny + y
  smeb_u32 z_begin, z_end; // planes updated by this launch (ghosts excluded)
  smeb_u32 errExpected;    // CTAs of all error-norm launches of this step (0: the
                           // norms stay in the slot, nothing is published)
  // z-slabs, tile kernel: halo exchange inside the stage kernel (see
  // smeb_stage_tile).  haloMode bit 0 / 1: there is a rank below / above.
  smeb_u32 haloMode;
  double *peerLo;          // rank below: its HIGH ghost plane of the output buffer
  double *peerHi;          // rank above: its LOW ghost plane of the output buffer
  smeb_u32 *halo;          // own words: [0] / [1] count of boundary-plane tiles that
                           // have arrived from below / above since the start (every
                           // stage adds one per tile of the plane)
  smeb_u32 *peerFlagLo;    // rank below: its halo[1]
  smeb_u32 *peerFlagHi;    // rank above: its halo[0]
  smeb_u32 *haloErr;       // own word: raised when a bounded wait ran out
  smeb_u32 haloSeq;        // number of this stage: its inputs are complete when the
                           // counts have reached (haloSeq - 1) * tiles
  smeb_u32 nSeg;           // segments in `seg`
  smeb_u32 errPairs;       // pairs in the error slot: 1, or the members of an ensemble
  smeb_u32 roles;          // ensembles: buffer roles (CUR, RA, RB, RS = 0..3; 15 =
                           // none) of cin | cout << 4 | s2in << 8 | s2out << 12
  smeb_u32 zig;            // tile kernel, 3-D: grid = (chunks, tiles y, tiles x) and
                           // the chunks of a tile march in alternating directions
  smeb_u32 errPublish;     // the first CTA publishes the error slot before it starts:
                           // the norms of the step BEFORE this kernel, whose final
                           // stage only left its maxima in the slot (errExpected 0)
  double g1, g2, g3, bdt, delta; // stage coefficients (bdt = beta*dt)
  // CUDA-graph replay of a step: the step size lives in device memory (a
  // captured launch cannot carry a new bdt); bdt = beta * *dtp when dtp != 0
  double beta;
  const double *dtp;
  double cF, s2F, s3F, eps;      // finalise factors, error epsilon
  // D/dx^2, D/dy^2, D/dz^2 per species, then the runtime model parameters
  // (values of sme_param_k: scanned parameters of an ensemble stay symbolic so
  // that one compiled module serves every member)
  double D[3 * S + SMEB_NPARAMS];
};

// one member of an ensemble, per compartment; rewritten by the host before every
// step (the roles of a member's buffers rotate when ITS step is accepted)
struct SmebMember {
  double *buf[4];  // concentration buffers by role: CUR, RA, RB, RS
  double dt;       // step size of this attempt
  smeb_u32 active; // 0: the member has reached the end time (its CTAs only report in)
  smeb_u32 pad_;
};

struct SmebMembEntry {
  smeb_u32 partner; // voxel index in the other compartment
  smeb_u32 info;    // membrane id | side (bit 16: own voxel is in B) | dir<<17
};

#ifndef SMEB_NCOMP
#define SMEB_NCOMP 1
#endif

struct SmebMembArgs {
  const smeb_u32 *mvox;       // [n_mvox] voxels with at least one membrane face
  const smeb_u32 *mstart;     // [n_mvox+1] CSR offsets into ent
  const SmebMembEntry *ent;   // faces in reference order (membrane, direction)
  const double *flux_len;     // [n_membranes*6]
  const double *comp_cin[SMEB_NCOMP]; // stage input of every compartment
  const SmebMember *comp_members[SMEB_NCOMP]; // ensembles: member tables
  smeb_u32 n_mvox;
  smeb_u32 pad_;
};

template <int S>
__device__ __forceinline__ void smeb_load_voxel(const double *__restrict__ p,
                                                size_t i, double *v) {
  if constexpr (S % 2 == 0) {
    const double2 *q = reinterpret_cast<const double2 *>(p) + i * (S / 2);
#pragma unroll
    for (int k = 0; k < S / 2; ++k) {
      double2 t = q[k];
      v[2 * k] = t.x;
      v[2 * k + 1] = t.y;
    }
  } else {
    const double *q = p + i * S;
#pragma unroll
    for (int k = 0; k < S; ++k) {
      v[k] = q[k];
    }
  }
}

template <int S>
__device__ __forceinline__ void smeb_store_voxel(double *__restrict__ p,
                                                 size_t i, const double *v) {
  if constexpr (S % 2 == 0) {
    double2 *q = reinterpret_cast<double2 *>(p) + i * (S / 2);
#pragma unroll
    for (int k = 0; k < S / 2; ++k) {
      q[k] = make_double2(v[2 * k], v[2 * k + 1]);
    }
  } else {
    double *q = p + i * S;
#pragma unroll
    for (int k = 0; k < S; ++k) {
      q[k] = v[k];
    }
  }
}

// Programmatic dependent launch (sm_90+).  Every stage kernel first waits until
// the kernel in front of it in the stream has completed and flushed
// (griddepcontrol.wait) and only then lets the kernel behind it become resident
// (griddepcontrol.launch_dependents): a kernel's prologue therefore overlaps the
// tail of its immediate predecessor and nothing older, and what it does before
// the wait (barrier setup, index arithmetic, loads of data that the predecessor
// does not write — the pre-step state c0) is safe.  Both are no-ops for launches
// without the stream-serialization attribute.
__device__ __forceinline__ void smeb_pdl_trigger() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void smeb_pdl_wait() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
}

#define SMEB_ERR_RING 8 // entries of the pinned host ring (sim.hpp kErrRing)

// One CTA of an error-norm launch is done: returns true to the thread of the CTA
// that completes the step's count.  That CTA takes the maxima out of the slot
// (leaving it zeroed for the next step) and publishes them with the step's
// sequence number into the pinned host ring (smeb_err_publish): the host polls
// its own memory instead of queueing a copy, a memset and an event behind every
// step (which also kept two kernels from overlapping their tails).  `pairs` pairs
// of maxima (ensembles: one per member); a ring entry is 2 pairs + 2 words, the
// sequence number at [2 pairs].  Called by one thread per CTA, after its
// atomicMax.
__device__ __forceinline__ bool smeb_err_count(smeb_u64 *errBase, smeb_u32 pairs,
                                               smeb_u32 expected) {
  if (expected == 0u) {
    return false;
  }
  unsigned int *cnt = reinterpret_cast<unsigned int *>(errBase + 2u * pairs);
  __threadfence();
  const unsigned int prev = atomicAdd(cnt, 1u);
  if (prev + 1u == expected) {
    __threadfence();
    return true;
  }
  return false;
}

// the publishing CTA: lanes [0, nLanes) of one warp (or one thread, nLanes = 1)
// move the maxima to the host ring entry of the next sequence number
__device__ __forceinline__ void smeb_err_publish(smeb_u64 *errBase, smeb_u32 pairs,
                                                 smeb_u64 *errHost, int lane,
                                                 int nLanes) {
  unsigned int *cnt = reinterpret_cast<unsigned int *>(errBase + 2u * pairs);
  const unsigned int seq = *reinterpret_cast<volatile unsigned int *>(cnt + 1) + 1u;
  volatile smeb_u64 *h =
      errHost + (size_t)(2u * pairs + 2u) * (seq % SMEB_ERR_RING);
  for (smeb_u32 k = lane; k < 2u * pairs; k += nLanes) {
    h[k] = atomicExch(errBase + k, 0ull);
  }
  __threadfence_system();
  if (nLanes > 1) {
    __syncwarp();
  }
  if (lane == 0) {
    atomicExch(cnt, 0u);
    atomicExch(cnt + 1, seq);
    __threadfence_system();
    h[2u * pairs] = (smeb_u64)seq;
    __threadfence_system();
  }
}

// a CTA that only reports in (an empty tile CTA, a finished ensemble member)
__device__ __forceinline__ void smeb_err_arrive(smeb_u64 *errBase, smeb_u32 pairs,
                                                smeb_u64 *errHost,
                                                smeb_u32 expected) {
  if (smeb_err_count(errBase, pairs, expected)) {
    smeb_err_publish(errBase, pairs, errHost, 0, 1);
  }
}

// max-reduction of the two error norms over the block, then one atomicMax per
// block on the ordered bit patterns (values are >= 0 and never NaN, because
// like std::max in pixelsim_impl.cpp:783-788 a NaN or negative candidate is
// never taken).  Called once per thread at the end of a kernel.
__device__ __forceinline__ void smeb_block_max2(double eabs, double erel,
                                                smeb_u64 *err, smeb_u64 *errBase,
                                                smeb_u32 pairs, smeb_u64 *errHost,
                                                smeb_u32 expected) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    double a = __shfl_xor_sync(0xffffffffu, eabs, o);
    double r = __shfl_xor_sync(0xffffffffu, erel, o);
    eabs = a > eabs ? a : eabs;
    erel = r > erel ? r : erel;
  }
  __shared__ double sa[32], sr[32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int nw = (blockDim.x + 31) >> 5;
  if (lane == 0) {
    sa[w] = eabs;
    sr[w] = erel;
  }
  __syncthreads();
  if (w == 0) {
    eabs = lane < nw ? sa[lane] : 0.0;
    erel = lane < nw ? sr[lane] : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      double a = __shfl_xor_sync(0xffffffffu, eabs, o);
      double r = __shfl_xor_sync(0xffffffffu, erel, o);
      eabs = a > eabs ? a : eabs;
      erel = r > erel ? r : erel;
    }
    int publisher = 0;
    if (lane == 0) {
      if (eabs > 0.0) {
        atomicMax(err, (smeb_u64)__double_as_longlong(eabs));
      }
      if (erel > 0.0) {
        atomicMax(err + 1, (smeb_u64)__double_as_longlong(erel));
      }
      publisher = smeb_err_count(errBase, pairs, expected) ? 1 : 0;
    }
    publisher = __shfl_sync(0xffffffffu, publisher, 0);
    if (publisher != 0) { // the whole warp moves the (2 x members) maxima
      smeb_err_publish(errBase, pairs, errHost, lane, 32);
    }
  }
}

// Deferred publication.  Counting the CTAs of a final stage costs every one of
// them a fence and an atomic round trip at its end, while it still holds its
// shared memory and registers: with 2048 CTAs on 296 slots that was 14 us per
// final-stage launch (24 of 151 us on a 64-plane slab; ncu, profiles/r2_*).  Where
// the host does not wait for a step's norms before it launches the next step
// (unbounded tolerances: they are checked two steps late), the final stage only
// leaves its maxima in the slot, and the first CTA of the NEXT kernel — which
// starts after the final stage has completed (griddepcontrol.wait / stream order)
// — moves them to the host ring.  smeb_err_flush does the same for the last step
// of a run, and behind final stages of large grids when the host does wait.
template <int S>
__device__ __forceinline__ void smeb_err_publish_previous(const SmebStageArgs<S> &a) {
  if (a.errPublish != 0u && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 &&
      threadIdx.x < 32) {
    smeb_err_publish(a.errBase, a.errPairs, a.errHost, (int)threadIdx.x, 32);
  }
}

__global__ void smeb_err_flush(smeb_u64 *errBase, smeb_u32 pairs, smeb_u64 *errHost) {
  smeb_pdl_wait();
  if (blockIdx.x == 0 && threadIdx.x < 32) {
    smeb_err_publish(errBase, pairs, errHost, (int)threadIdx.x, 32);
  }
}

// deterministic block sums of N per-thread accumulators (fixed shuffle tree,
// then warp 0 over the warp results): dst[j] = sum over the block of acc[j]
template <int N>
__device__ __forceinline__ void smeb_block_sum(const double *acc, double *dst) {
  __shared__ double ss[32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int nw = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int j = 0; j < N; ++j) {
    double v = acc[j];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      v += __shfl_xor_sync(0xffffffffu, v, o);
    }
    if (lane == 0) {
      ss[w] = v;
    }
    __syncthreads();
    if (w == 0) {
      v = lane < nw ? ss[lane] : 0.0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        v += __shfl_xor_sync(0xffffffffu, v, o);
      }
      if (lane == 0) {
        dst[j] = v;
      }
    }
    __syncthreads();
  }
}

// RK stage update + error norm for one voxel, all in registers.
//   c    stage input, f complete dc/dt, c0v pre-step state (if USE_C0),
//   s2v  in: running sum from memory or 0; out: updated running sum,
//   out  the value the caller stores to a.cout.
// Association order: pixelsim_impl.cpp:627,675,699,718-719,741,783-788.
//
// The model functor R carries what the CPU solver does to dc/dt after the
// membranes (pixelsim.cpp:113-119): R::nonSpatial replaces dc/dt of the
// non-spatial species by their spatial averages (computed by the SUM_NS
// pre-pass), R::storage multiplies by 1/storage; R::SP is the number of model
// species — the time/space padding components behind them are not clamped
// (pixelsim_impl.cpp:643-654).
template <class R, unsigned F>
__device__ __forceinline__ void
smeb_update(const SmebStageArgs<R::S> &a, const double *c, const double *fin,
            const double *c0v, double *s2v, double *out, double &eabs,
            double &erel) {
  constexpr int S = R::S;
  double f[S];
#pragma unroll
  for (int s = 0; s < S; ++s) {
    f[s] = fin[s];
  }
  if constexpr (R::NNS > 0) {
    R::nonSpatial(f, a.nsAv);
  }
  // Pseudo-time relaxation of the zero-storage (algebraic) species towards
  // 0 = R(c) + div(D grad c): one adaptive RK212 iteration of
  // PixelSim::solveZeroStorageConstraints (pixelsim.cpp:23-90) is two launches.
  // The residual is dc/dt before the storage factor; the other species are
  // copied.  a.bdt = dt (A) or dt/2 (B), a.c0 = the state before the iteration.
  if constexpr ((F & SMEB_F_RELAX_A) != 0) {
#pragma unroll
    for (int s = 0; s < S; ++s) {
      out[s] = c[s];
      if (R::zeroStorage(s)) {
        // getZeroStorageResidual, pixelsim_impl.cpp:877-891
        const double r = fabs(f[s]);
        eabs = eabs < r ? r : eabs;
        const double q = r / (fabs(c[s]) + a.eps);
        erel = erel < q ? q : erel;
        out[s] = c[s] + a.bdt * f[s]; // doRelaxSubstep1, :834-842
      }
    }
    return;
  }
  if constexpr ((F & SMEB_F_RELAX_B) != 0) {
#pragma unroll
    for (int s = 0; s < S; ++s) {
      out[s] = c[s];
      if (R::zeroStorage(s)) {
        // doRelaxSubstep2 (:844-852) and calculateRelaxError (:864-876)
        const double cn = 0.5 * c0v[s] + 0.5 * c[s] + a.bdt * f[s];
        const double e = fabs(cn - c[s]);
        eabs = eabs < e ? e : eabs;
        const double q = e / (0.5 * (cn + c0v[s] + a.eps));
        erel = erel < q ? q : erel;
        out[s] = cn;
      }
    }
    return;
  }
  if constexpr (R::STORAGE) {
    R::storage(f);
  }
  if constexpr ((F & SMEB_F_OUT_DCDT) != 0) {
#pragma unroll
    for (int s = 0; s < S; ++s) {
      out[s] = f[s];
    }
    return;
  }
  // same product as the host's sd.beta * dt
  const double bdt = a.dtp != nullptr ? a.beta * (*a.dtp) : a.bdt;
  double cn[S];
#pragma unroll
  for (int s = 0; s < S; ++s) {
    if constexpr ((F & SMEB_F_USE_S2) != 0) {
      s2v[s] = s2v[s] + a.delta * c[s];
    }
    if constexpr ((F & SMEB_F_GENERIC) != 0) {
      cn[s] = a.g1 * c[s] + a.g2 * s2v[s] + a.g3 * c0v[s] + bdt * f[s];
    } else if constexpr ((F & SMEB_F_USE_C0) != 0) {
      cn[s] = a.g3 * c0v[s] + a.g1 * c[s] + bdt * f[s];
    } else {
      cn[s] = c[s] + bdt * f[s];
    }
  }
  if constexpr ((F & SMEB_F_FINAL) != 0) {
#pragma unroll
    for (int s = 0; s < S; ++s) {
      double lower;
      if constexpr ((F & SMEB_F_LOWER_IS_CIN) != 0) {
        lower = c[s];
      } else {
        lower = a.cF * cn[s] + a.s2F * s2v[s] + a.s3F * c0v[s];
      }
      if constexpr ((F & SMEB_F_OUT_LOWER) != 0) {
        out[s] = lower;
      } else if (!R::zeroStorage(s)) { // pixelsim_impl.cpp:778-781
        const double e = fabs(cn[s] - lower);
        eabs = eabs < e ? e : eabs;
        const double norm = 0.5 * (cn[s] + c0v[s] + a.eps);
        // e / norm can only raise the running maximum if it exceeds erel;
        // the exact division is done only for candidates within 0.1% of
        // that (or with a non-positive / NaN norm), so the result is the
        // same value std::max over all quotients gives
        if (!(e < 0.999 * erel * norm)) {
          const double r = e / norm;
          erel = erel < r ? r : erel;
        }
      }
    }
  }
  if constexpr ((F & SMEB_F_OUT_LOWER) == 0) {
#pragma unroll
    for (int s = 0; s < S; ++s) {
      double v = cn[s];
      if constexpr ((F & SMEB_F_CLAMP_OUT) != 0 && (F & SMEB_F_OUT_CNEW) == 0) {
        if (s < R::SP) {
          v = v < 0.0 ? 0.0 : v;
        }
      }
      out[s] = v;
    }
  }
}

// complete per-voxel epilogue for the direct kernels: loads of c0 / s2,
// update, stores
// c0pre: the voxel's c0 if the caller already holds it (nullptr: loaded here)
template <class R, unsigned F>
__device__ __forceinline__ void
smeb_finish_voxel(const SmebStageArgs<R::S> &a, size_t i, const double *c,
                  const double *f, double &eabs, double &erel, double *nsacc,
                  const double *c0pre = nullptr) {
  constexpr int S = R::S;
  if constexpr ((F & SMEB_F_SUM_NS) != 0) {
    // pre-pass: only the sums of dc/dt of the non-spatial species are wanted
    double v[R::NNS > 0 ? R::NNS : 1];
    R::nonSpatialPick(f, v);
#pragma unroll
    for (int j = 0; j < R::NNS; ++j) {
      nsacc[j] += v[j];
    }
    return;
  }
  double c0v[S], s2v[S], out[S];
#pragma unroll
  for (int s = 0; s < S; ++s) {
    c0v[s] = 0.0;
    s2v[s] = 0.0;
  }
  if constexpr ((F & SMEB_F_OUT_DCDT) == 0) {
    if constexpr ((F & SMEB_F_USE_C0) != 0) {
      if (c0pre != nullptr) {
#pragma unroll
        for (int s = 0; s < S; ++s) {
          c0v[s] = c0pre[s];
        }
      } else {
        smeb_load_voxel<S>(a.c0, i, c0v);
      }
    }
    if constexpr ((F & SMEB_F_USE_S2IN) != 0) {
      smeb_load_voxel<S>(a.s2in, i, s2v);
    }
  }
  smeb_update<R, F>(a, c, f, c0v, s2v, out, eabs, erel);
  constexpr bool onDemand =
      (F & (SMEB_F_OUT_DCDT | SMEB_F_OUT_LOWER | SMEB_F_OUT_CNEW)) != 0;
  if constexpr ((F & SMEB_F_WRITE_S2) != 0 && !onDemand) {
    smeb_store_voxel<S>(a.s2out, i, s2v);
  }
  smeb_store_voxel<S>(a.cout, i, out);
}

// reaction + uniform diffusion given the voxel and its six neighbours
// (pixelsim_impl.cpp:444-450: x, y, z terms summed left to right)
template <class R>
__device__ __forceinline__ void
smeb_rhs_values(const SmebStageArgs<R::S> &a, const double *c,
                const double *xp, const double *xm, const double *yp,
                const double *ym, const double *zp, const double *zm,
                double *f) {
  constexpr int S = R::S;
  R::eval(f, c, a.D + 3 * S);
#pragma unroll
  for (int s = 0; s < S; ++s) {
    f[s] += a.D[3 * s] * (xp[s] + xm[s] - 2.0 * c[s]) +
            a.D[3 * s + 1] * (yp[s] + ym[s] - 2.0 * c[s]) +
            a.D[3 * s + 2] * (zp[s] + zm[s] - 2.0 * c[s]);
  }
}

// non-uniform diffusion: face-averaged constants, conservative form
// (pixelsim_impl.cpp:453-485, same association order)
template <int S>
__device__ __forceinline__ void
smeb_diffusion_array(const double *h2, const double *d, const double *dxp_,
                     const double *dxm_, const double *dyp_, const double *dym_,
                     const double *dzp_, const double *dzm_, const double *c,
                     const double *xp, const double *xm, const double *yp,
                     const double *ym, const double *zp, const double *zm,
                     double *f) {
#pragma unroll
  for (int s = 0; s < S; ++s) {
    const double dxp = 0.5 * (d[s] + dxp_[s]) / h2[0];
    const double dxm = 0.5 * (d[s] + dxm_[s]) / h2[0];
    const double dyp = 0.5 * (d[s] + dyp_[s]) / h2[1];
    const double dym = 0.5 * (d[s] + dym_[s]) / h2[1];
    const double dzp = 0.5 * (d[s] + dzp_[s]) / h2[2];
    const double dzm = 0.5 * (d[s] + dzm_[s]) / h2[2];
    f[s] += dxp * (xp[s] - c[s]) - dxm * (c[s] - xm[s]) +
            dyp * (yp[s] - c[s]) - dym * (c[s] - ym[s]) +
            dzp * (zp[s] - c[s]) - dzm * (c[s] - zm[s]);
  }
}

// Direct kernels: the voxel and its six neighbours are gathered from global
// memory (neighbour lines come from L1/L2).  Besides the uniform stencil this
// path carries the CPU solver's other diffusion operators: per-voxel diffusion
// constants (R::NUD) and cross-diffusion (R::NX terms).  The cross-diffusion
// coefficient of a neighbour is a function of that neighbour's state, which is
// already in registers, so it is re-evaluated instead of being stored per voxel
// as the CPU solver does (pixelsim_impl.cpp:494-504,545-589).
template <class R, bool HAVE_C = false>
__device__ __forceinline__ void
smeb_rhs_gather(const SmebStageArgs<R::S> &a, smeb_u32 i, const smeb_u32 *nb,
                double *c, double *f) {
  constexpr int S = R::S;
  double xp[S], xm[S], yp[S], ym[S], zp[S], zm[S];
  if constexpr (!HAVE_C) { // else the caller loaded the voxel itself
    smeb_load_voxel<S>(a.cin, i, c);
  }
  smeb_load_voxel<S>(a.cin, nb[0], xp);
  smeb_load_voxel<S>(a.cin, nb[1], xm);
  smeb_load_voxel<S>(a.cin, nb[2], yp);
  smeb_load_voxel<S>(a.cin, nb[3], ym);
  smeb_load_voxel<S>(a.cin, nb[4], zp);
  smeb_load_voxel<S>(a.cin, nb[5], zm);
  if constexpr (R::NUD) {
    double d[S], d0[S], d1[S], d2[S], d3[S], d4[S], d5[S];
    smeb_load_voxel<S>(a.darr, i, d);
    smeb_load_voxel<S>(a.darr, nb[0], d0);
    smeb_load_voxel<S>(a.darr, nb[1], d1);
    smeb_load_voxel<S>(a.darr, nb[2], d2);
    smeb_load_voxel<S>(a.darr, nb[3], d3);
    smeb_load_voxel<S>(a.darr, nb[4], d4);
    smeb_load_voxel<S>(a.darr, nb[5], d5);
    R::eval(f, c, a.D + 3 * S);
    smeb_diffusion_array<S>(a.h2, d, d0, d1, d2, d3, d4, d5, c, xp, xm, yp, ym,
                            zp, zm, f);
  } else {
    smeb_rhs_values<R>(a, c, xp, xm, yp, ym, zp, zm, f);
  }
  if constexpr (R::NX > 0) {
    R::cross(f, c, xp, xm, yp, ym, zp, zm, a.h2, a.D + 3 * S);
  }
}

// max |cross-diffusion coefficient| per term over the voxels of a compartment:
// the input of the forward-Euler bound RK101 re-evaluates every step
// (updateCrossDiffusionMaxStableTimestep, pixelsim_impl.cpp:506-543)
template <class R>
__global__ void __launch_bounds__(SMEB_BLOCK)
    smeb_xd_max(const __grid_constant__ SmebStageArgs<R::S> a) {
  constexpr int S = R::S;
  constexpr int NX = R::NX > 0 ? R::NX : 1;
  double mx[NX] = {};
  for (smeb_u32 i = a.v0 + blockIdx.x * SMEB_BLOCK + threadIdx.x; i < a.n;
       i += gridDim.x * SMEB_BLOCK) {
    double c[S], k[NX];
    smeb_load_voxel<S>(a.cin, i, c);
    R::crossCoef(k, c, a.D + 3 * S);
#pragma unroll
    for (int t = 0; t < R::NX; ++t) {
      const double v = fabs(k[t]);
      mx[t] = mx[t] < v ? v : mx[t];
    }
  }
#pragma unroll
  for (int t = 0; t < R::NX; ++t) {
    double v = mx[t];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double w = __shfl_xor_sync(0xffffffffu, v, o);
      v = w > v ? w : v;
    }
    if ((threadIdx.x & 31) == 0 && v > 0.0) {
      atomicMax(a.xdMax + t, (smeb_u64)__double_as_longlong(v));
    }
  }
}

// ---- member-batched ensembles (SMEB_F_MEMBERS) -----------------------------------
// A parameter scan is many simulations of one model that differ in the values of
// runtime parameters (BASELINE config 5: 64 x liver-simplified).  Each of them is
// launch bound on its own; batched, the members are gridDim.y of ONE launch per
// stage and compartment.  Geometry tables, coefficients and stage constants are
// shared (the kernel arguments); what differs per member comes from a table the
// host rewrites before every step: the member's buffers by role (a member whose
// step was rejected keeps its roles, the others rotate), its step size and
// whether it is still running, and its parameter values.  The kernel patches a
// private copy of its arguments and runs the ordinary body.
template <class R, unsigned F>
__device__ __forceinline__ bool smeb_member_select(SmebStageArgs<R::S> &a) {
  constexpr int S = R::S;
  const smeb_u32 mIdx = blockIdx.y;
  const SmebMember mb = a.members[mIdx];
  const smeb_u32 rCin = a.roles & 15u, rCout = (a.roles >> 4) & 15u,
                 rS2in = (a.roles >> 8) & 15u, rS2out = (a.roles >> 12) & 15u;
  a.cin = mb.buf[rCin];
  a.cout = mb.buf[rCout];
  a.c0 = mb.buf[0];
  a.s2in = rS2in < 4u ? mb.buf[rS2in] : nullptr;
  a.s2out = rS2out < 4u ? mb.buf[rS2out] : nullptr;
  a.bdt = a.beta * mb.dt; // the product the host forms for a single simulation
  a.dtp = nullptr;
  a.err = a.errBase + 2u * mIdx;
#pragma unroll
  for (int k = 0; k < SMEB_NPARAMS; ++k) {
    a.D[3 * S + k] = a.memberParams[(size_t)mIdx * SMEB_NPARAMS + k];
  }
  if (mb.active == 0u) {
    // a finished member: its CTAs still count towards the step's error norms
    smeb_pdl_wait();
    smeb_pdl_trigger();
    smeb_err_publish_previous<R::S>(a);
    if constexpr (((F & SMEB_F_FINAL) != 0 && (F & SMEB_F_OUT_LOWER) == 0)) {
      if (threadIdx.x == 0) {
        smeb_err_arrive(a.errBase, a.errPairs, a.errHost, a.errExpected);
      }
    }
    return false;
  }
  return true;
}

// ---- direct-gather stage kernels: grid-stride loop, one voxel at a time -----
// voxel-list geometry.  The neighbour table of the reference (six indices per
// voxel, geometry.cpp:98-110; 24 B per voxel as uint32) is stored run-compressed
// where the host could build it: the voxel order is z, x, y with y fastest
// (geometry.cpp:55-72), so the -y / +y neighbours are i -/+ 1 and along a run of
// consecutive voxels of one column the index distance to the neighbouring
// column / plane is constant.  One 32-bit code per voxel — run id (25 bits), six
// "neighbour exists" bits (missing neighbour = own index, the zero-flux
// boundary), membrane flag — plus a table of four deltas per run that stays in
// L1/L2: 4 B of index traffic per voxel instead of 24.  Voxels flagged
// SMEB_MEMB_FLAG are left to smeb_stage_memb.
#define SMEB_RUN_BITS 25
// `code` = nn[i] of the run-compressed layout, loaded by the caller (the stage
// kernel fetches it one iteration early; ignored for the plain table)
__device__ __forceinline__ bool
smeb_neighbours(const smeb_u32 *__restrict__ nn, const int *__restrict__ runs,
                smeb_u32 i, smeb_u32 code, smeb_u32 *nb) {
  if (runs != nullptr) {
    const int4 d = reinterpret_cast<const int4 *>(
        runs)[code & ((1u << SMEB_RUN_BITS) - 1u)];
    const smeb_u32 fl = code >> SMEB_RUN_BITS;
    nb[0] = (fl & 1u) ? i + d.x : i;
    nb[1] = (fl & 2u) ? i + d.y : i;
    nb[2] = (fl & 4u) ? i + 1u : i;
    nb[3] = (fl & 8u) ? i - 1u : i;
    nb[4] = (fl & 16u) ? i + d.z : i;
    nb[5] = (fl & 32u) ? i + d.w : i;
    return (code & SMEB_MEMB_FLAG) != 0;
  }
  const uint2 *q = reinterpret_cast<const uint2 *>(nn) + (size_t)i * 3;
  const uint2 n0 = q[0], n1 = q[1], n2 = q[2];
  nb[0] = n0.x & ~SMEB_MEMB_FLAG;
  nb[1] = n0.y;
  nb[2] = n1.x;
  nb[3] = n1.y;
  nb[4] = n2.x;
  nb[5] = n2.y;
  return (n0.x & SMEB_MEMB_FLAG) != 0;
}

#ifndef SMEB_NN_MIN_CTAS
#define SMEB_NN_MIN_CTAS 0 // resident CTAs per SM the register allocation aims at (0: compiler's choice)
#endif
#if SMEB_NN_MIN_CTAS > 0
#define SMEB_NN_BOUNDS __launch_bounds__(SMEB_BLOCK, SMEB_NN_MIN_CTAS)
#else
#define SMEB_NN_BOUNDS __launch_bounds__(SMEB_BLOCK)
#endif
template <class R, unsigned F>
__device__ __forceinline__ void smeb_stage_nn_body(const SmebStageArgs<R::S> &a) {
  constexpr int S = R::S;
  double eabs = 0.0, erel = 0.0;
  double nsacc[R::NNS > 0 ? R::NNS : 1] = {};
  // neighbour code -> run deltas -> gathers is a chain of dependent loads: the
  // code of the next voxel of this thread is fetched one iteration early
  // (SMEB_NN_CODE_AHEAD >= 1), and so are its own state and c0 (>= 2): with six
  // resident CTAs the loads of one iteration alone do not keep enough bytes in
  // flight to cover the DRAM latency
  const smeb_u32 stride = gridDim.x * SMEB_BLOCK;
  const bool coded = a.runs != nullptr;
  constexpr bool PRE = SMEB_NN_CODE_AHEAD >= 2;
  constexpr bool PRE_C0 = PRE && (F & SMEB_F_USE_C0) != 0 &&
                          (F & (SMEB_F_OUT_DCDT | SMEB_F_SUM_NS)) == 0;
  smeb_u32 i = a.v0 + blockIdx.x * SMEB_BLOCK + threadIdx.x;
  smeb_pdl_wait();
  smeb_pdl_trigger();
  smeb_err_publish_previous<R::S>(a);
  smeb_u32 codeNext = coded && i < a.n ? a.nn[i] : 0u;
  double cNext[S], c0Next[S];
#pragma unroll
  for (int s = 0; s < S; ++s) {
    cNext[s] = 0.0;
    c0Next[s] = 0.0;
  }
  if (PRE && i < a.n) {
    smeb_load_voxel<S>(a.cin, i, cNext);
    if constexpr (PRE_C0) {
      smeb_load_voxel<S>(a.c0, i, c0Next);
    }
  }
  for (; i < a.n; i += stride) {
    const smeb_u32 code =
        SMEB_NN_CODE_AHEAD != 0 ? codeNext : (coded ? a.nn[i] : 0u);
    double c[S], c0v[S], f[S];
#pragma unroll
    for (int s = 0; s < S; ++s) {
      c[s] = cNext[s];
      c0v[s] = c0Next[s];
    }
    if (stride < a.n - i) {
      if constexpr (SMEB_NN_CODE_AHEAD != 0) {
        if (coded) {
          codeNext = a.nn[i + stride];
        }
      }
      if constexpr (PRE) {
        smeb_load_voxel<S>(a.cin, i + stride, cNext);
      }
      if constexpr (PRE_C0) {
        smeb_load_voxel<S>(a.c0, i + stride, c0Next);
      }
    }
    smeb_u32 nb[6];
    if (smeb_neighbours(a.nn, a.runs, i, code, nb)) {
      continue;
    }
    smeb_rhs_gather<R, PRE>(a, i, nb, c, f);
    smeb_finish_voxel<R, F>(a, i, c, f, eabs, erel, nsacc,
                            PRE_C0 ? c0v : nullptr);
  }
  if constexpr ((F & SMEB_F_SUM_NS) != 0) {
    smeb_block_sum<R::NNS>(nsacc, a.nsPartial + (size_t)blockIdx.x * R::NNS);
  } else if constexpr (((F & SMEB_F_FINAL) != 0 && (F & SMEB_F_OUT_LOWER) == 0) ||
                       (F & (SMEB_F_RELAX_A | SMEB_F_RELAX_B)) != 0) {
    smeb_block_max2(eabs, erel, a.err, a.errBase, a.errPairs, a.errHost,
                    a.errExpected);
  }
}

template <class R, unsigned F>
__global__ void SMEB_NN_BOUNDS
    smeb_stage_nn(const __grid_constant__ SmebStageArgs<R::S> ain) {
  if constexpr ((F & SMEB_F_MEMBERS) != 0) {
    SmebStageArgs<R::S> a = ain;
    if (smeb_member_select<R, F>(a)) {
      smeb_stage_nn_body<R, F>(a);
    }
  } else {
    smeb_stage_nn_body<R, F>(ain);
  }
}

// box geometry without tiling (fallback for shapes the TMA kernel does not
// take): neighbours implicit, zero-flux Neumann at the faces (missing
// neighbour = own index, geometry.cpp:98-110); only planes
// [z_begin, z_end) are updated (the others are z-slab ghost planes).
template <class R, unsigned F>
__device__ __forceinline__ void smeb_stage_box_body(const SmebStageArgs<R::S> &a) {
  constexpr int S = R::S;
  double eabs = 0.0, erel = 0.0;
  double nsacc[R::NNS > 0 ? R::NNS : 1] = {};
  const smeb_u32 sx = a.ny, sz = a.nx * a.ny;
  const smeb_u32 iEnd = a.z_end * sz;
  smeb_pdl_wait();
  smeb_pdl_trigger();
  smeb_err_publish_previous<R::S>(a);
  for (smeb_u32 i = a.z_begin * sz + blockIdx.x * SMEB_BLOCK + threadIdx.x;
       i < iEnd; i += gridDim.x * SMEB_BLOCK) {
    const smeb_u32 y = i % a.ny;
    const smeb_u32 t = i / a.ny;
    const smeb_u32 x = t % a.nx;
    const smeb_u32 z = t / a.nx;
    smeb_u32 nb[6];
    nb[0] = x + 1 < a.nx ? i + sx : i;
    nb[1] = x > 0 ? i - sx : i;
    nb[2] = y + 1 < a.ny ? i + 1 : i;
    nb[3] = y > 0 ? i - 1 : i;
    nb[4] = z + 1 < a.nz ? i + sz : i;
    nb[5] = z > 0 ? i - sz : i;
    double c[S], f[S];
    smeb_rhs_gather<R>(a, i, nb, c, f);
    smeb_finish_voxel<R, F>(a, i, c, f, eabs, erel, nsacc);
  }
  if constexpr ((F & SMEB_F_SUM_NS) != 0) {
    smeb_block_sum<R::NNS>(nsacc, a.nsPartial + (size_t)blockIdx.x * R::NNS);
  } else if constexpr (((F & SMEB_F_FINAL) != 0 && (F & SMEB_F_OUT_LOWER) == 0) ||
                       (F & (SMEB_F_RELAX_A | SMEB_F_RELAX_B)) != 0) {
    smeb_block_max2(eabs, erel, a.err, a.errBase, a.errPairs, a.errHost,
                    a.errExpected);
  }
}

template <class R, unsigned F>
__global__ void __launch_bounds__(SMEB_BLOCK)
    smeb_stage_box(const __grid_constant__ SmebStageArgs<R::S> ain) {
  if constexpr ((F & SMEB_F_MEMBERS) != 0) {
    SmebStageArgs<R::S> a = ain;
    if (smeb_member_select<R, F>(a)) {
      smeb_stage_box_body<R, F>(a);
    }
  } else {
    smeb_stage_box_body<R, F>(ain);
  }
}

// ---- membrane-adjacent voxels ----------------------------------------------
// One thread per voxel that has >= 1 membrane face.  The voxel is processed
// completely here (reaction, diffusion, its membrane fluxes in the reference's
// accumulation order: membranes in model order, then XP,XM,YP,YM,ZP,ZM —
// pixelsim.cpp:105-111, pixelsim_impl.cpp:1012-1025), so no atomics are needed
// and the sum order is deterministic.
template <class M, int S_OWN>
__device__ __forceinline__ void
smeb_memb_term(smeb_u32 info, smeb_u32 partner, const double *cown,
               const SmebMembArgs &m, const double *params, double *f) {
  // SA / SB: state components per voxel of A / B (model species + NE padding
  // components), SAP / SBP: model species.  Variables of the membrane
  // expression: [A species | B species | padding of B's voxel, or of A's when
  // B is not simulated] (pixelsim_impl.cpp:84-98).
  constexpr int SA = M::SA, SB = M::SB, SAP = M::SAP, SBP = M::SBP,
                NE = M::NE;
  double in[SAP + SBP + NE + 1], out[SAP + SBP + NE + 1];
  const bool ownIsB = (info >> 16) & 1u;
  const smeb_u32 dir = (info >> 17) & 7u;
  const double L = m.flux_len[(info & 0xffffu) * 6 + dir];
  if (!ownIsB) {
    if constexpr (S_OWN == SA && SA > 0) {
#pragma unroll
      for (int s = 0; s < SAP; ++s) {
        in[s] = cown[s];
      }
      if constexpr (SB > 0) {
        double b[SB];
        smeb_load_voxel<SB>(m.comp_cin[M::CB], partner, b);
#pragma unroll
        for (int s = 0; s < SB; ++s) {
          in[SAP + s] = b[s];
        }
      } else {
#pragma unroll
        for (int s = 0; s < NE; ++s) {
          in[SAP + s] = cown[SAP + s];
        }
      }
      M::eval(out, in, params);
#pragma unroll
      for (int s = 0; s < SAP; ++s) {
        f[s] += out[s] / L;
      }
    }
  } else {
    if constexpr (S_OWN == SB && SB > 0) {
#pragma unroll
      for (int s = 0; s < SB; ++s) {
        in[SAP + s] = cown[s];
      }
      if constexpr (SA > 0) {
        double av[SA];
        smeb_load_voxel<SA>(m.comp_cin[M::CA], partner, av);
#pragma unroll
        for (int s = 0; s < SAP; ++s) {
          in[s] = av[s];
        }
      }
      M::eval(out, in, params);
#pragma unroll
      for (int s = 0; s < SBP; ++s) {
        f[s] += out[SAP + s] / L;
      }
    }
  }
}

// generated per model: switch over the membrane id -> smeb_memb_term<Memb_k>
template <int S_OWN>
__device__ __forceinline__ void
smeb_memb_dispatch(smeb_u32 info, smeb_u32 partner, const double *cown,
                   const SmebMembArgs &m, const double *params, double *f);

template <class R, unsigned F>
__device__ __forceinline__ void smeb_stage_memb_body(const SmebStageArgs<R::S> &a,
                                                     const SmebMembArgs &m) {
  constexpr int S = R::S;
  double eabs = 0.0, erel = 0.0;
  double nsacc[R::NNS > 0 ? R::NNS : 1] = {};
  smeb_pdl_wait();
  smeb_pdl_trigger();
  smeb_err_publish_previous<R::S>(a);
  for (smeb_u32 k = blockIdx.x * SMEB_BLOCK + threadIdx.x; k < m.n_mvox;
       k += gridDim.x * SMEB_BLOCK) {
    const smeb_u32 i = m.mvox[k];
    smeb_u32 nb[6];
    smeb_neighbours(a.nn, a.runs, i, a.runs != nullptr ? a.nn[i] : 0u, nb);
    double c[S], f[S];
    smeb_rhs_gather<R>(a, i, nb, c, f);
    const smeb_u32 e1 = m.mstart[k + 1];
    for (smeb_u32 e = m.mstart[k]; e < e1; ++e) {
      const SmebMembEntry ent = m.ent[e];
      smeb_memb_dispatch<S>(ent.info, ent.partner, c, m, a.D + 3 * S, f);
    }
    smeb_finish_voxel<R, F>(a, i, c, f, eabs, erel, nsacc);
  }
  if constexpr ((F & SMEB_F_SUM_NS) != 0) {
    smeb_block_sum<R::NNS>(nsacc, a.nsPartial + (size_t)blockIdx.x * R::NNS);
  } else if constexpr (((F & SMEB_F_FINAL) != 0 && (F & SMEB_F_OUT_LOWER) == 0) ||
                       (F & (SMEB_F_RELAX_A | SMEB_F_RELAX_B)) != 0) {
    smeb_block_max2(eabs, erel, a.err, a.errBase, a.errPairs, a.errHost,
                    a.errExpected);
  }
}

template <class R, unsigned F>
__global__ void __launch_bounds__(SMEB_BLOCK)
    smeb_stage_memb(const __grid_constant__ SmebStageArgs<R::S> ain,
                    const __grid_constant__ SmebMembArgs min) {
  if constexpr ((F & SMEB_F_MEMBERS) != 0) {
    SmebStageArgs<R::S> a = ain;
    SmebMembArgs m = min;
    // the partner voxels are read from the same member's stage input
#pragma unroll
    for (int j = 0; j < SMEB_NCOMP; ++j) {
      if (m.comp_members[j] != nullptr) {
        m.comp_cin[j] = m.comp_members[j][blockIdx.y].buf[a.roles & 15u];
      }
    }
    if (smeb_member_select<R, F>(a)) {
      smeb_stage_memb_body<R, F>(a, m);
    }
  } else {
    smeb_stage_memb_body<R, F>(ain, min);
  }
}

// ---- TMA-staged tile-marching stage kernel for box compartments -------------
// The compartment fills its image, so the voxel list is a dense
// [nz][nx][ny*S] fp64 array (y contiguous).  A CTA owns a TX x TY tile of the
// (x, y) plane and marches along z (M3) — or, for 2-D images (nz == 1), owns a
// TY-wide strip and marches along x in bricks of TX rows.  Every plane/brick of
// the stage input is fetched ONCE by TMA (cp.async.bulk.tensor, halo of one
// voxel in x and y, out-of-image cells zero-filled and never used) into a ring
// of NST shared-memory slots, NST-3 (M3) or NST-1 loads ahead of the compute;
// the seven stencil operands are then read from shared memory, the pre-step
// state / running sum are register-prefetched one plane ahead with coalesced
// 16-byte loads, and the result is written with coalesced 16-byte stores.
// Zero-flux Neumann faces (missing neighbour = own voxel) are address selects
// computed once per thread, not per voxel.  Host guarantees nx % RPT == 0.
struct alignas(64) SmebTensorMap {
  smeb_u64 opaque[16];
};

__device__ __forceinline__ smeb_u32 smeb_smem_addr(const void *p) {
  smeb_u64 a;
  asm volatile("cvta.to.shared.u64 %0, %1;" : "=l"(a) : "l"(p));
  return (smeb_u32)a;
}

__device__ __forceinline__ void smeb_mbar_init(smeb_u32 bar, smeb_u32 count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count)
               : "memory");
}

__device__ __forceinline__ void smeb_mbar_expect_tx(smeb_u32 bar,
                                                    smeb_u32 bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(
                   bar),
               "r"(bytes)
               : "memory");
}

__device__ __forceinline__ void smeb_mbar_wait(smeb_u32 bar, smeb_u32 parity) {
  asm volatile("{\n"
               ".reg .pred P1;\n"
               "SMEB_WAIT:\n"
               "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
               "@P1 bra SMEB_DONE;\n"
               "bra SMEB_WAIT;\n"
               "SMEB_DONE:\n"
               "}" ::"r"(bar),
               "r"(parity)
               : "memory");
}

__device__ __forceinline__ void smeb_tma_load_3d(smeb_u32 dst,
                                                 const SmebTensorMap *tmap,
                                                 int c0, int c1, int c2,
                                                 smeb_u32 bar) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::"
               "complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
               "l"((smeb_u64)tmap), "r"(c0), "r"(c1), "r"(c2), "r"(bar)
               : "memory");
}

template <int S>
__device__ __forceinline__ void smeb_lds(const double *p, double *v) {
  if constexpr (S % 2 == 0) {
    const double2 *q = reinterpret_cast<const double2 *>(p);
#pragma unroll
    for (int k = 0; k < S / 2; ++k) {
      double2 t = q[k];
      v[2 * k] = t.x;
      v[2 * k + 1] = t.y;
    }
  } else {
#pragma unroll
    for (int k = 0; k < S; ++k) {
      v[k] = p[k];
    }
  }
}

extern __shared__ __align__(128) unsigned char smeb_dyn_smem[];

// ---- halo exchange inside the stage kernel (z-slabs over NVLink) ---------------
// A rank's boundary planes are written twice: into its own output buffer and,
// with plain stores through the peer mapping, into the neighbour's ghost plane of
// the same buffer.  A CTA that has stored its tile of a boundary plane adds one,
// with release semantics at system scope, to a counter word in the NEIGHBOUR's
// memory (fire and forget: nothing comes back, nobody on this side waits for the
// other tiles); the neighbour's CTAs that read that ghost plane in the next stage
// acquire the counter before their TMA loads: (stage - 1) x tiles means every
// tile of every earlier stage has landed.  No copy engine, no extra launches, no
// stream waits: one kernel per stage, and the stages chain with programmatic
// dependent launch.  The wait is bounded: after about 10 s it raises haloErr and
// goes on (the host reports it), it never hangs the GPU.
__device__ __forceinline__ smeb_u32 smeb_ld_acquire_sys(const smeb_u32 *p) {
  smeb_u32 v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void smeb_red_release_sys(smeb_u32 *p, smeb_u32 v) {
  asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void smeb_halo_acquire(smeb_u32 *halo, int side,
                                                  smeb_u32 want, smeb_u32 *err) {
  long long t0 = 0;
  unsigned spins = 0;
  while ((int)(smeb_ld_acquire_sys(halo + side) - want) < 0) {
    if (++spins == 64u) {
      t0 = clock64();
    } else if (spins > 64u && (spins & 1023u) == 0u &&
               clock64() - t0 > 20000000000ll) {
      atomicExch(err, 1u + (smeb_u32)side);
      break;
    }
  }
  // the plane was written through the generic proxy (by the neighbour's SMs);
  // the TMA load that follows reads it through the async proxy
  asm volatile("fence.proxy.async.global;" ::: "memory");
}

// waits (one thread) until the planes of stage `seq` have arrived from both
// neighbours: the end of run(), before getters read ghost planes
__global__ void smeb_halo_quiesce(smeb_u32 *halo, smeb_u32 mode, smeb_u32 want,
                                  smeb_u32 *err) {
  // want: the count the words must have reached (stages x tiles of a plane)
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    if (mode & 1u) {
      smeb_halo_acquire(halo, 0, want, err);
    }
    if (mode & 2u) {
      smeb_halo_acquire(halo, 1, want, err);
    }
  }
}

template <class R, unsigned F, int TX, int TY, int NST, bool M3>
__global__ void __launch_bounds__(SMEB_BLOCK, SMEB_TILE_MIN_CTAS)
    smeb_stage_tile(const __grid_constant__ SmebStageArgs<R::S> a,
                    const __grid_constant__ SmebTensorMap tmap,
                    const smeb_u32 chunk) {
  constexpr int S = R::S;
  constexpr int G = SMEB_BLOCK / TY;   // thread groups along x
  constexpr int RPT = TX / G;          // consecutive rows per thread
  // TMA needs the box start to be 16-byte aligned in the row: for an odd
  // number of species the box starts one double early (and ends one late)
  constexpr int PAD = S % 2;
  constexpr int PITCH = (TY + 2) * S + 2 * PAD; // doubles per smem row
  constexpr smeb_u32 BOX_BYTES = (TX + 2) * PITCH * 8;
  constexpr smeb_u32 SLOT_BYTES = (BOX_BYTES + 127u) & ~127u;
  constexpr int LOOK = M3 ? 2 : 0;     // planes needed beyond the current one
  constexpr bool USE_C0 = (F & SMEB_F_USE_C0) != 0 &&
                          (F & (SMEB_F_OUT_DCDT | SMEB_F_SUM_NS)) == 0;
  constexpr bool USE_S2IN = (F & SMEB_F_USE_S2IN) != 0 &&
                            (F & (SMEB_F_OUT_DCDT | SMEB_F_SUM_NS)) == 0;
  constexpr bool ON_DEMAND =
      (F & (SMEB_F_OUT_DCDT | SMEB_F_OUT_LOWER | SMEB_F_OUT_CNEW)) != 0;
  static_assert(SMEB_BLOCK % TY == 0 && TX % G == 0, "tile shape");
  static_assert(NST >= LOOK + 2, "ring too short");

  __shared__ smeb_u64 fullBar[NST];
  unsigned char *ring = smeb_dyn_smem;
  {
    const smeb_u32 mis = smeb_smem_addr(ring) & 127u;
    ring += (128u - mis) & 127u;
  }
  const smeb_u32 ringAddr = smeb_smem_addr(ring);
  const smeb_u32 barAddr = smeb_smem_addr(fullBar);

  const int tid = threadIdx.x;
  const int ty = tid % TY, tg = tid / TY;
  const int r0 = tg * RPT; // first tile row of this thread
  // march range [m0, m1): planes (M3) or bricks of TX rows (2-D)
  int m0, m1, x0, y0;
  // Zig-zag (3-D): the chunks of one tile are adjacent in the launch order (chunk
  // index fastest), so they run together, and they march in alternating
  // directions — even chunks down from their top plane, odd chunks up from their
  // bottom plane.  Two neighbouring chunks then start at their common boundary,
  // or arrive at it, at the same time, and the two halo planes each of them reads
  // beyond its own range are the planes the other one is reading just then: L2
  // hits instead of a second trip to HBM ((chunk + 2) / chunk of the stage input
  // otherwise: 25 % at 8 planes per chunk).
  bool down = false;
  smeb_u32 tilesPerPlane = gridDim.x * gridDim.y;
  constexpr bool ZIG = M3 && (F & SMEB_F_ZIG) != 0; // compiled in only on request
  if constexpr (M3) {
    int zc;
    if (ZIG && a.zig != 0u) {
      zc = blockIdx.x;
      y0 = blockIdx.y * TY;
      x0 = blockIdx.z * TX;
      down = (zc & 1) == 0;
      tilesPerPlane = gridDim.y * gridDim.z;
    } else {
      zc = blockIdx.z;
      y0 = blockIdx.x * TY;
      x0 = blockIdx.y * TX;
    }
    m0 = a.z_begin + zc * chunk;
    m1 = min(m0 + (int)chunk, (int)a.z_end);
  } else {
    y0 = blockIdx.x * TY;
    x0 = 0;
    m0 = blockIdx.y * chunk;
    m1 = min(m0 + (int)chunk, (int)((a.nx + TX - 1) / TX));
  }
  const int gy = y0 + ty;
  const int dir = down ? -1 : 1;
  const int nIter = m1 - m0;
  if (nIter <= 0) {
    smeb_pdl_wait();
    smeb_pdl_trigger();
    smeb_err_publish_previous<R::S>(a);
    if constexpr ((F & SMEB_F_SUM_NS) != 0) {
      const size_t blk =
          ((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x +
          blockIdx.x;
      if (threadIdx.x < R::NNS) {
        a.nsPartial[blk * R::NNS + threadIdx.x] = 0.0;
      }
    } else if constexpr (((F & SMEB_F_FINAL) != 0 && (F & SMEB_F_OUT_LOWER) == 0) ||
                         (F & (SMEB_F_RELAX_A | SMEB_F_RELAX_B)) != 0) {
      if (threadIdx.x == 0) { // an empty CTA still counts towards the step's norms
        smeb_err_arrive(a.errBase, a.errPairs, a.errHost, a.errExpected);
      }
    }
    return;
  }
  const int nLoads = nIter + LOOK;
  // halo exchange in the kernel: does this CTA read a ghost plane / own a
  // boundary plane?
  constexpr bool HALO = M3 && (F & SMEB_F_HALO) != 0 &&
                        (F & (SMEB_F_OUT_DCDT | SMEB_F_OUT_LOWER |
                              SMEB_F_OUT_CNEW | SMEB_F_SUM_NS | SMEB_F_RELAX_A |
                              SMEB_F_RELAX_B)) == 0;
  bool haloLo = false, haloHi = false;
  if constexpr (HALO) {
    haloLo = (a.haloMode & 1u) != 0u && m0 == (int)a.z_begin;
    haloHi = (a.haloMode & 2u) != 0u && m1 == (int)a.z_end;
  }
  // load j fetches plane m0 - 1 + j, or m1 - j marching down (M3) / brick m0 + j
  // (2-D)
  auto issue = [&](int j, int slot) {
    const smeb_u32 bar = barAddr + 8u * slot;
    smeb_mbar_expect_tx(bar, BOX_BYTES);
    if constexpr (M3) {
      // planes outside the array are never read (Neumann select below);
      // the coordinate is clamped so that every box overlaps the tensor
      const int plane =
          max(0, min((int)a.nz - 1, down ? m1 - j : m0 - 1 + j));
      smeb_tma_load_3d(ringAddr + slot * SLOT_BYTES, &tmap,
                       (y0 - 1) * S - PAD, x0 - 1, plane, bar);
    } else {
      smeb_tma_load_3d(ringAddr + slot * SLOT_BYTES, &tmap,
                       (y0 - 1) * S - PAD, (m0 + j) * TX - 1, 0, bar);
    }
  };
  // per-thread geometry: Neumann faces as address selects
  const int offYm = gy == 0 ? 0 : -S;
  const int offYp = gy + 1 >= (int)a.ny ? 0 : S;
  const int ownOff = (r0 + 1) * PITCH + (ty + 1) * S + PAD; // row r0, own col
  int gx0 = x0 + r0;
  bool active = gy < (int)a.ny && (M3 ? gx0 < (int)a.nx : true);
  int offLo = (M3 && gx0 == 0) ? 0 : -PITCH;               // row r0 - 1
  int offHi = (M3 && gx0 + RPT >= (int)a.nx) ? (RPT - 1) * PITCH
                                              : RPT * PITCH; // row r0 + RPT
  const size_t rowStride = (size_t)a.ny;               // voxels per x row
  // voxels per iteration (negative marching down; size_t arithmetic wraps)
  const size_t marchStride =
      M3 ? (size_t)((long long)dir * (long long)((size_t)a.nx * a.ny))
         : (size_t)TX * a.ny;
  size_t g = M3 ? ((size_t)(down ? m1 - 1 : m0) * a.nx + gx0) * a.ny + gy
                : ((size_t)m0 * TX + r0) * a.ny + gy;

  // register prefetch queues of c0 / s2: entry q holds the values of iteration
  // it + q; entry 0 is consumed, the last entry is loaded that many iterations
  // early.  AHC / AHS: depth of the c0 / s2 queue
  constexpr int AHC = SMEB_C0_AHEAD > 0 ? SMEB_C0_AHEAD
                      : (USE_C0 && (!USE_S2IN || SMEB_C0_SPLIT != 0)) ? 2
                                                                      : 1;
  constexpr int AHS = SMEB_C0_AHEAD > 0 ? SMEB_C0_AHEAD : 1;
  double c0q[AHC + 1][RPT][S], s2q[AHS + 1][RPT][S];
#pragma unroll
  for (int k = 0; k < RPT; ++k) {
#pragma unroll
    for (int s = 0; s < S; ++s) {
#pragma unroll
      for (int q = 0; q <= AHC; ++q) {
        c0q[q][k][s] = 0.0;
      }
#pragma unroll
      for (int q = 0; q <= AHS; ++q) {
        s2q[q][k][s] = 0.0;
      }
    }
  }
  if constexpr (!M3) {
    active = gy < (int)a.ny && m0 * TX + r0 < (int)a.nx;
  }
  // whether this thread's voxels exist in iteration `it` (2-D: the brick moves)
  auto activeAt = [&](int it) {
    if constexpr (M3) {
      return active;
    } else {
      return gy < (int)a.ny && (m0 + it) * TX + r0 < (int)a.nx;
    }
  };
  // The pre-step state c0 is not written by any kernel of this step, so its first
  // planes are requested before the wait for the kernel in front of this one:
  // they travel while that kernel drains.
#pragma unroll
  for (int q = 0; q < AHC; ++q) {
    if constexpr (USE_C0) {
      if (q < nIter && activeAt(q)) {
#pragma unroll
        for (int k = 0; k < RPT; ++k) {
          smeb_load_voxel<S>(a.c0, g + q * marchStride + k * rowStride,
                             c0q[q][k]);
        }
      }
    }
  }
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < NST; ++s) {
      smeb_mbar_init(barAddr + 8u * s, 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  // everything above is independent of the kernels before this one
  smeb_pdl_wait();
  smeb_pdl_trigger();
  smeb_err_publish_previous<R::S>(a);
  if constexpr (HALO) {
    // a boundary chunk reads a ghost plane (its first / last load): the
    // neighbour's plane of the previous stage must have arrived.  Both waits sit
    // here, outside the march loop, which stays the single-GPU loop
    if (tid == 0) {
      const smeb_u32 want = (a.haloSeq - 1u) * tilesPerPlane;
      if (haloLo) {
        smeb_halo_acquire(a.halo, 0, want, a.haloErr);
      }
      if (haloHi) {
        smeb_halo_acquire(a.halo, 1, want, a.haloErr);
      }
    }
  }
  if (tid == 0) {
    for (int j = 0; j < NST && j < nLoads; ++j) {
      issue(j, j);
    }
  }
  __syncthreads();

  // the running sum was written by the stage before this one: after the wait
#pragma unroll
  for (int q = 0; q < AHS; ++q) {
    if constexpr (USE_S2IN) {
      if (q < nIter && activeAt(q)) {
#pragma unroll
        for (int k = 0; k < RPT; ++k) {
          smeb_load_voxel<S>(a.s2in, g + q * marchStride + k * rowStride,
                             s2q[q][k]);
        }
      }
    }
  }

  double eabs = 0.0, erel = 0.0;
  double nsacc[R::NNS > 0 ? R::NNS : 1] = {};
  // ring bookkeeping: slot / parity of load j = it + LOOK (the newest plane
  // this iteration needs), and slots of the current and previous plane
  int slotW = LOOK % NST, parW = (LOOK / NST) & 1;
  int slotC = M3 ? 1 % NST : 0, slotM = 0;
  if constexpr (M3) {
    smeb_mbar_wait(barAddr, 0);
    smeb_mbar_wait(barAddr + 8u * (1 % NST), 0);
  }
  for (int it = 0; it < nIter; ++it) {
    const int m = down ? m1 - 1 - it : m0 + it;
    // 2-D: the brick's x range changes every iteration
    if constexpr (!M3) {
      gx0 = m * TX + r0;
      active = gy < (int)a.ny && gx0 < (int)a.nx;
      offLo = gx0 == 0 ? 0 : -PITCH;
      offHi = gx0 + RPT >= (int)a.nx ? (RPT - 1) * PITCH : RPT * PITCH;
    }
    // register prefetch of c0 / s2, AHC / AHS iterations ahead
#pragma unroll
    for (int k = 0; k < RPT; ++k) {
#pragma unroll
      for (int s = 0; s < S; ++s) {
        c0q[AHC][k][s] = 0.0;
        s2q[AHS][k][s] = 0.0;
      }
    }
    if constexpr (USE_C0) {
      if (it + AHC < nIter && activeAt(it + AHC)) {
#pragma unroll
        for (int k = 0; k < RPT; ++k) {
          smeb_load_voxel<S>(a.c0, g + AHC * marchStride + k * rowStride,
                             c0q[AHC][k]);
        }
      }
    }
    if constexpr (USE_S2IN) {
      if (it + AHS < nIter && activeAt(it + AHS)) {
#pragma unroll
        for (int k = 0; k < RPT; ++k) {
          smeb_load_voxel<S>(a.s2in, g + AHS * marchStride + k * rowStride,
                             s2q[AHS][k]);
        }
      }
    }
    smeb_mbar_wait(barAddr + 8u * slotW, parW);
    if (active) {
      const double *sc =
          reinterpret_cast<const double *>(ring + slotC * SLOT_BYTES) + ownOff;
      const double *sm = sc, *sp = sc;
      if constexpr (M3) {
        // the plane loaded before the current one is z - 1 marching up, z + 1
        // marching down; the newest plane is the other neighbour
        const int slotBelow = down ? slotW : slotM, slotAbove = down ? slotM : slotW;
        if (m > 0) {
          sm = reinterpret_cast<const double *>(ring + slotBelow * SLOT_BYTES) +
               ownOff;
        }
        if (m + 1 < (int)a.nz) {
          sp = reinterpret_cast<const double *>(ring + slotAbove * SLOT_BYTES) +
               ownOff;
        }
      }
      double col[RPT + 2][S];
      smeb_lds<S>(sc + offLo, col[0]);
#pragma unroll
      for (int k = 0; k < RPT; ++k) {
        smeb_lds<S>(sc + k * PITCH, col[k + 1]);
      }
      smeb_lds<S>(sc + offHi, col[RPT + 1]);
#pragma unroll
      for (int k = 0; k < RPT; ++k) {
        double yp[S], ym[S], zp[S], zm[S], f[S], out[S];
        smeb_lds<S>(sc + k * PITCH + offYp, yp);
        smeb_lds<S>(sc + k * PITCH + offYm, ym);
        smeb_lds<S>(sp + k * PITCH, zp);
        smeb_lds<S>(sm + k * PITCH, zm);
        smeb_rhs_values<R>(a, col[k + 1], col[k + 2], col[k], yp, ym, zp, zm,
                           f);
        if constexpr ((F & SMEB_F_SUM_NS) != 0) {
          double v[R::NNS > 0 ? R::NNS : 1];
          R::nonSpatialPick(f, v);
#pragma unroll
          for (int j = 0; j < R::NNS; ++j) {
            nsacc[j] += v[j];
          }
          continue;
        }
        smeb_update<R, F>(a, col[k + 1], f, c0q[0][k], s2q[0][k], out, eabs,
                          erel);
        if constexpr ((F & SMEB_F_WRITE_S2) != 0 && !ON_DEMAND) {
          smeb_store_voxel<S>(a.s2out, g + k * rowStride, s2q[0][k]);
        }
        smeb_store_voxel<S>(a.cout, g + k * rowStride, out);
        if constexpr (HALO) {
          // a boundary plane also goes into the neighbour's ghost plane: plain
          // stores through the peer mapping, they drain while the march goes on
          if ((haloLo && m == m0) || (haloHi && m == m1 - 1)) {
            const size_t inPlane = (size_t)(gx0 + k) * a.ny + gy;
            if (haloLo && m == m0) {
              smeb_store_voxel<S>(a.peerLo, inPlane, out);
            }
            if (haloHi && m == m1 - 1) {
              smeb_store_voxel<S>(a.peerHi, inPlane, out);
            }
          }
        }
      }
    }
#pragma unroll
    for (int k = 0; k < RPT; ++k) {
#pragma unroll
      for (int s = 0; s < S; ++s) {
#pragma unroll
        for (int q = 0; q < AHC; ++q) {
          c0q[q][k][s] = c0q[q + 1][k][s];
        }
#pragma unroll
        for (int q = 0; q < AHS; ++q) {
          s2q[q][k][s] = s2q[q + 1][k][s];
        }
      }
    }
    g += marchStride;
    __syncthreads(); // every thread is done with the oldest slot
    const int slotFree = M3 ? slotM : slotC;
    if (tid == 0 && it + NST < nLoads) {
      issue(it + NST, slotFree);
    }
    slotM = slotC;
    slotC = (slotC + 1 == NST) ? 0 : slotC + 1;
    slotW = slotW + 1;
    if (slotW == NST) {
      slotW = 0;
      parW ^= 1;
    }
  }
  if constexpr (HALO) {
    // this CTA's tile of a boundary plane is in the neighbour's ghost plane (all
    // threads are past the barrier that ended the last iteration): count it,
    // with release semantics, in the neighbour's memory
    if (tid == 0) {
      if (haloLo) {
        smeb_red_release_sys(a.peerFlagLo, 1u);
      }
      if (haloHi) {
        smeb_red_release_sys(a.peerFlagHi, 1u);
      }
    }
  }
  if constexpr ((F & SMEB_F_SUM_NS) != 0) {
    const size_t blk =
        ((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    smeb_block_sum<R::NNS>(nsacc, a.nsPartial + blk * R::NNS);
  } else if constexpr (((F & SMEB_F_FINAL) != 0 && (F & SMEB_F_OUT_LOWER) == 0) ||
                       (F & (SMEB_F_RELAX_A | SMEB_F_RELAX_B)) != 0) {
    smeb_block_max2(eabs, erel, a.err, a.errBase, a.errPairs, a.errHost,
                    a.errExpected);
  }
}

// ---- segment-staged stage kernel for voxel-list compartments -------------------
// The voxel list is z-major with y fastest (geometry.cpp:55-72), so along a piece
// of a run (consecutive voxels with constant index distances to the neighbouring
// columns / planes, see smeb_neighbours) the -y / +y neighbours are the voxels
// before / after, and the +x, -x (+z, -z) neighbours of the piece are CONTIGUOUS
// ranges of the state array: [i0 + d, i0 + d + L).  A CTA therefore stages, per
// segment, three (3-D: five) rows of the stage input in shared memory with 1-D
// bulk copies (cp.async.bulk + mbarrier complete_tx), NST segments deep, and all
// seven stencil operands come from shared memory.  The bytes in flight no longer
// depend on how many gathers a thread can have outstanding: the direct kernel
// (smeb_stage_nn) waited on its code -> run deltas -> gather chain with 55-70 %
// long-scoreboard stalls at 5 resident CTAs.  Segments are handed out round robin
// (segment = block + j * grid), so the CTAs that run together work on a
// contiguous window of the list and the rows they share meet in L2.  A voxel
// whose neighbour does not exist (code bit off) uses its own value; the rows are
// clamped to the array, so such a voxel's slot may hold anything and is never
// read.  Voxels flagged SMEB_MEMB_FLAG are left to smeb_stage_memb.
__device__ __forceinline__ void smeb_bulk_load(smeb_u32 dst, const void *src,
                                               smeb_u32 bytes, smeb_u32 bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes "
               "[%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

#ifndef SMEB_SEG_MIN_CTAS
#define SMEB_SEG_MIN_CTAS 4 // resident CTAs per SM the register allocation aims at
#endif
template <class R, unsigned F, int SEG, int NST, bool Z>
__device__ __forceinline__ void smeb_stage_seg_body(const SmebStageArgs<R::S> &a) {
  constexpr int S = R::S;
  static_assert(S % 2 == 0, "bulk copies need 16-byte voxels");
  static_assert(SEG % SMEB_BLOCK == 0, "segment length");
  constexpr int VPT = SEG / SMEB_BLOCK; // voxels per thread and segment
  constexpr int NROW = Z ? 5 : 3;       // own (+ halo), +x, -x (, +z, -z)
  constexpr smeb_u32 VB = S * 8;        // bytes per voxel
  constexpr smeb_u32 OWN_BYTES = (SEG + 2) * VB;
  constexpr smeb_u32 ROW_BYTES = SEG * VB;
  constexpr smeb_u32 STAGE_BYTES =
      (OWN_BYTES + (NROW - 1) * ROW_BYTES + 127u) & ~127u;
  constexpr bool USE_C0 = (F & SMEB_F_USE_C0) != 0 &&
                          (F & (SMEB_F_OUT_DCDT | SMEB_F_SUM_NS)) == 0;
  __shared__ smeb_u64 fullBar[NST];
  unsigned char *ring = smeb_dyn_smem;
  {
    const smeb_u32 mis = smeb_smem_addr(ring) & 127u;
    ring += (128u - mis) & 127u;
  }
  const smeb_u32 ringAddr = smeb_smem_addr(ring);
  const smeb_u32 barAddr = smeb_smem_addr(fullBar);
  const int tid = threadIdx.x;
  const smeb_u32 G = gridDim.x;
  const int nMine =
      blockIdx.x < a.nSeg ? (int)((a.nSeg - 1u - blockIdx.x) / G) + 1 : 0;
  const uint2 *segs = reinterpret_cast<const uint2 *>(a.seg);
  const int4 *runs = reinterpret_cast<const int4 *>(a.runs);

  // thread 0: bulk copies of segment j of this CTA into ring slot j % NST
  auto issue = [&](int j) {
    const uint2 sd = segs[blockIdx.x + (smeb_u32)j * G];
    const long long i0 = sd.x, L = sd.y, n = a.n;
    const int4 d = runs[a.nn[sd.x] & ((1u << SMEB_RUN_BITS) - 1u)];
    const int slot = j % NST;
    const smeb_u32 base = ringAddr + slot * STAGE_BYTES;
    const smeb_u32 bar = barAddr + 8u * slot;
    long long src[NROW], cnt[NROW];
    smeb_u32 dst[NROW];
    // own row with one voxel of halo on each side, clamped to the array
    {
      const long long lo = i0 > 0 ? i0 - 1 : i0;
      const long long hi = i0 + L + 1 < n ? i0 + L + 1 : n;
      src[0] = lo;
      cnt[0] = hi - lo;
      dst[0] = base + (smeb_u32)(lo - (i0 - 1)) * VB;
    }
    const int delta[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
    for (int r = 1; r < NROW; ++r) {
      const long long s0 = i0 + delta[r - 1];
      const long long lo = s0 > 0 ? s0 : 0;
      const long long hi = s0 + L < n ? s0 + L : n;
      // delta 0: no voxel of the run has a neighbour in this direction
      const bool used = delta[r - 1] != 0 && hi > lo;
      src[r] = lo;
      cnt[r] = used ? hi - lo : 0;
      dst[r] = base + OWN_BYTES + (smeb_u32)(r - 1) * ROW_BYTES +
               (smeb_u32)(lo - s0) * VB;
    }
    smeb_u32 total = 0;
#pragma unroll
    for (int r = 0; r < NROW; ++r) {
      total += (smeb_u32)cnt[r] * VB;
    }
    smeb_mbar_expect_tx(bar, total);
#pragma unroll
    for (int r = 0; r < NROW; ++r) {
      if (cnt[r] > 0) {
        smeb_bulk_load(dst[r], a.cin + src[r] * S, (smeb_u32)cnt[r] * VB, bar);
      }
    }
  };

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < NST; ++s) {
      smeb_mbar_init(barAddr + 8u * s, 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  smeb_pdl_wait();
  smeb_pdl_trigger();
  smeb_err_publish_previous<R::S>(a);
  if (tid == 0) {
    for (int j = 0; j < NST && j < nMine; ++j) {
      issue(j);
    }
  }
  __syncthreads();

  double eabs = 0.0, erel = 0.0;
  double nsacc[R::NNS > 0 ? R::NNS : 1] = {};
  // neighbour codes and the pre-step state of a thread's voxels are fetched one
  // segment early with plain coalesced loads
  smeb_u32 codeN[VPT];
  double c0N[VPT][S];
  auto prefetch = [&](int j) {
#pragma unroll
    for (int v = 0; v < VPT; ++v) {
      codeN[v] = 0u;
#pragma unroll
      for (int s = 0; s < S; ++s) {
        c0N[v][s] = 0.0;
      }
    }
    if (j < nMine) {
      const uint2 sd = segs[blockIdx.x + (smeb_u32)j * G];
#pragma unroll
      for (int v = 0; v < VPT; ++v) {
        const smeb_u32 t = tid + v * SMEB_BLOCK;
        if (t < sd.y) {
          codeN[v] = a.nn[sd.x + t];
          if constexpr (USE_C0) {
            smeb_load_voxel<S>(a.c0, sd.x + t, c0N[v]);
          }
        }
      }
    }
  };
  prefetch(0);
  for (int j = 0; j < nMine; ++j) {
    const uint2 sd = segs[blockIdx.x + (smeb_u32)j * G];
    smeb_u32 code[VPT];
    double c0v[VPT][S];
#pragma unroll
    for (int v = 0; v < VPT; ++v) {
      code[v] = codeN[v];
#pragma unroll
      for (int s = 0; s < S; ++s) {
        c0v[v][s] = c0N[v][s];
      }
    }
    prefetch(j + 1);
    const int slot = j % NST;
    smeb_mbar_wait(barAddr + 8u * slot, (smeb_u32)(j / NST) & 1u);
    const double *own =
        reinterpret_cast<const double *>(ring + slot * STAGE_BYTES);
    const double *rows = own + (SEG + 2) * S;
    // The VPT voxels of a thread are independent: the stencil and reaction
    // arithmetic of all of them is written without branches so that the compiler
    // interleaves their fp64 dependency chains; only the global loads and stores
    // are predicated (a slot beyond a short segment holds stale data that is
    // computed on and dropped).
    bool valid[VPT];
    double c[VPT][S], f[VPT][S];
#pragma unroll
    for (int v = 0; v < VPT; ++v) {
      const smeb_u32 t = tid + v * SMEB_BLOCK;
      valid[v] = t < sd.y && (code[v] & SMEB_MEMB_FLAG) == 0u;
      const smeb_u32 fl = code[v] >> SMEB_RUN_BITS;
      const double *o = own + (t + 1) * S;
      double xp[S], xm[S], yp[S], ym[S], zp[S], zm[S];
      smeb_lds<S>(o, c[v]);
      smeb_lds<S>((fl & 1u) ? rows + t * S : o, xp);
      smeb_lds<S>((fl & 2u) ? rows + (SEG + t) * S : o, xm);
      smeb_lds<S>((fl & 4u) ? o + S : o, yp);
      smeb_lds<S>((fl & 8u) ? o - S : o, ym);
      if constexpr (Z) {
        smeb_lds<S>((fl & 16u) ? rows + (2 * SEG + t) * S : o, zp);
        smeb_lds<S>((fl & 32u) ? rows + (3 * SEG + t) * S : o, zm);
      } else {
#pragma unroll
        for (int s = 0; s < S; ++s) {
          zp[s] = c[v][s];
          zm[s] = c[v][s];
        }
      }
      smeb_rhs_values<R>(a, c[v], xp, xm, yp, ym, zp, zm, f[v]);
    }
#pragma unroll
    for (int v = 0; v < VPT; ++v) {
      const size_t i = (size_t)sd.x + tid + v * SMEB_BLOCK;
      if constexpr ((F & SMEB_F_SUM_NS) != 0) {
        double pick[R::NNS > 0 ? R::NNS : 1];
        R::nonSpatialPick(f[v], pick);
#pragma unroll
        for (int q = 0; q < R::NNS; ++q) {
          nsacc[q] += valid[v] ? pick[q] : 0.0;
        }
      } else {
        constexpr bool onDemand =
            (F & (SMEB_F_OUT_DCDT | SMEB_F_OUT_LOWER | SMEB_F_OUT_CNEW)) != 0;
        double s2v[S], out[S];
#pragma unroll
        for (int s = 0; s < S; ++s) {
          s2v[s] = 0.0;
        }
        if constexpr ((F & SMEB_F_USE_S2IN) != 0 && (F & SMEB_F_OUT_DCDT) == 0) {
          if (valid[v]) {
            smeb_load_voxel<S>(a.s2in, i, s2v);
          }
        }
        double ea = eabs, er = erel;
        smeb_update<R, F>(a, c[v], f[v], c0v[v], s2v, out, ea, er);
        if (valid[v]) {
          eabs = ea;
          erel = er;
          if constexpr ((F & SMEB_F_WRITE_S2) != 0 && !onDemand) {
            smeb_store_voxel<S>(a.s2out, i, s2v);
          }
          smeb_store_voxel<S>(a.cout, i, out);
        }
      }
    }
    __syncthreads(); // every thread is done with this slot
    if (tid == 0 && j + NST < nMine) {
      issue(j + NST);
    }
  }
  if constexpr ((F & SMEB_F_SUM_NS) != 0) {
    smeb_block_sum<R::NNS>(nsacc, a.nsPartial + (size_t)blockIdx.x * R::NNS);
  } else if constexpr (((F & SMEB_F_FINAL) != 0 && (F & SMEB_F_OUT_LOWER) == 0) ||
                       (F & (SMEB_F_RELAX_A | SMEB_F_RELAX_B)) != 0) {
    smeb_block_max2(eabs, erel, a.err, a.errBase, a.errPairs, a.errHost,
                    a.errExpected);
  }
}

template <class R, unsigned F, int SEG, int NST, bool Z>
__global__ void __launch_bounds__(SMEB_BLOCK, SMEB_SEG_MIN_CTAS)
    smeb_stage_seg(const __grid_constant__ SmebStageArgs<R::S> ain) {
  if constexpr ((F & SMEB_F_MEMBERS) != 0) {
    SmebStageArgs<R::S> a = ain;
    if (smeb_member_select<R, F>(a)) {
      smeb_stage_seg_body<R, F, SEG, NST, Z>(a);
    }
  } else {
    smeb_stage_seg_body<R, F, SEG, NST, Z>(ain);
  }
}

// ---- spatial average of dc/dt of the non-spatial species -----------------------
// av[j] = (sum of the block sums of the pre-pass, in a fixed order) / n
// (SimCompartment::spatiallyAverageDcdt, pixelsim_impl.cpp:150-163: the CPU sums
// serially; the association order differs, the result agrees to rounding)
__global__ void __launch_bounds__(SMEB_BLOCK)
    smeb_ns_finish(const double *__restrict__ partial, smeb_u32 nBlocks,
                   smeb_u32 nns, double n, double *__restrict__ av) {
  for (smeb_u32 j = 0; j < nns; ++j) {
    double acc = 0.0;
    for (smeb_u32 b = threadIdx.x; b < nBlocks; b += SMEB_BLOCK) {
      acc += partial[(size_t)b * nns + j];
    }
    smeb_block_sum<1>(&acc, av + j);
    if (threadIdx.x == 0) {
      av[j] = av[j] / n;
    }
    __syncthreads();
  }
}

// ---- per-species average / minimum / maximum of the state -------------------------
// What Simulation::updateConcentrations computes on one CPU thread from the
// downloaded state after every image interval (calculateAvgMinMax,
// simulate.cpp:272-290), done where the state lives: block partials
// [blocks][3*SP] = {sum, min, max} per model species (padding skipped), then one
// block folds them in a fixed order; min starts at DBL_MAX, max at 0 like
// sme::simulate::AvgMinMax (simulate_options.hpp:226-238).
template <class R>
__global__ void __launch_bounds__(SMEB_BLOCK)
    smeb_stats(const double *__restrict__ c, smeb_u32 i0, smeb_u32 i1,
               double *__restrict__ partial) {
  constexpr int S = R::S, SP = R::SP;
  double sum[SP], mn[SP], mx[SP];
#pragma unroll
  for (int s = 0; s < SP; ++s) {
    sum[s] = 0.0;
    mn[s] = 1.7976931348623157e308;
    mx[s] = 0.0;
  }
  // [i0, i1): the owned voxels (z-slab ghosts are skipped)
  for (smeb_u32 i = i0 + blockIdx.x * SMEB_BLOCK + threadIdx.x; i < i1;
       i += gridDim.x * SMEB_BLOCK) {
    double v[S];
    smeb_load_voxel<S>(c, i, v);
#pragma unroll
    for (int s = 0; s < SP; ++s) {
      sum[s] += v[s];
      mn[s] = v[s] < mn[s] ? v[s] : mn[s];
      mx[s] = v[s] > mx[s] ? v[s] : mx[s];
    }
  }
  __shared__ double ss[3][32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int s = 0; s < SP; ++s) {
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
      double ta = 0.0, tb = 1.7976931348623157e308, td = 0.0;
      for (int k = 0; k < SMEB_BLOCK / 32; ++k) {
        ta += ss[0][k];
        tb = ss[1][k] < tb ? ss[1][k] : tb;
        td = ss[2][k] > td ? ss[2][k] : td;
      }
      double *dst = partial + ((size_t)blockIdx.x * SP + s) * 3;
      dst[0] = ta;
      dst[1] = tb;
      dst[2] = td;
    }
    __syncthreads();
  }
}

// One species of every (sx, sy, sz)-th voxel of a z,x,y-ordered block of voxels:
// out[(zi * mx + xi) * my + yi] = c[(((z0 + zi*sz) * nx + xi*sx) * ny + yi*sy) * S + s].
// A voxel list is the block nx = nz = 1, ny = n.  The decimated image of a
// species is all that leaves the device (Simulation::updateConcentrations copies
// the whole state per image interval, simulate.cpp:313-314).
__global__ void __launch_bounds__(SMEB_BLOCK)
    smeb_gather_species(const double *__restrict__ c, double *__restrict__ out,
                        smeb_u32 S, smeb_u32 s, smeb_u32 nx, smeb_u32 ny,
                        smeb_u32 z0, smeb_u32 sx, smeb_u32 sy, smeb_u32 sz,
                        smeb_u32 mx, smeb_u32 my, unsigned long long nOut) {
  for (unsigned long long j = (unsigned long long)blockIdx.x * SMEB_BLOCK + threadIdx.x;
       j < nOut; j += (unsigned long long)gridDim.x * SMEB_BLOCK) {
    const unsigned long long yi = j % my, xi = (j / my) % mx, zi = j / ((unsigned long long)my * mx);
    const unsigned long long v = ((z0 + zi * sz) * nx + xi * sx) * ny + yi * sy;
    out[j] = c[v * S + s];
  }
}

// out[3*s .. 3*s+2] = {average, min, max} of species s
__global__ void __launch_bounds__(SMEB_BLOCK)
    smeb_stats_finish(const double *__restrict__ partial, smeb_u32 nBlocks,
                      smeb_u32 sp, double nVoxels, double *__restrict__ out) {
  __shared__ double ss[3][SMEB_BLOCK];
  for (smeb_u32 s = 0; s < sp; ++s) {
    double a = 0.0, b = 1.7976931348623157e308, d = 0.0;
    for (smeb_u32 k = threadIdx.x; k < nBlocks; k += SMEB_BLOCK) {
      const double *src = partial + ((size_t)k * sp + s) * 3;
      a += src[0];
      b = src[1] < b ? src[1] : b;
      d = src[2] > d ? src[2] : d;
    }
    ss[0][threadIdx.x] = a;
    ss[1][threadIdx.x] = b;
    ss[2][threadIdx.x] = d;
    __syncthreads();
    for (int o = SMEB_BLOCK / 2; o > 0; o >>= 1) {
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

// ---- elementwise helper -------------------------------------------------------
// c = max(c, 0) in place: SimCompartment::clampNegativeConcentrations for a
// state injected from the host (pixelsim_impl.cpp:643-665); the stage kernels
// clamp at their own store
__global__ void __launch_bounds__(SMEB_BLOCK)
    smeb_clamp(double *__restrict__ c, smeb_u64 n) {
  for (smeb_u64 i = (smeb_u64)blockIdx.x * SMEB_BLOCK + threadIdx.x; i < n;
       i += (smeb_u64)gridDim.x * SMEB_BLOCK) {
    const double v = c[i];
    if (v < 0.0) {
      c[i] = 0.0;
    }
  }
}
