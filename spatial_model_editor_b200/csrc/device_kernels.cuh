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
//   clamp = max(c,0) applied lazily on the next load
//                                                  pixelsim_impl.cpp:643-665
// The floating point association order of every expression follows the cited
// reference lines so that fixed-step results agree to the last bits.

typedef unsigned long long smeb_u64;
typedef unsigned int smeb_u32;

#define SMEB_F_CLAMP_CIN 1u      // clamp stage input to >= 0 on load
#define SMEB_F_CLAMP_C0 2u       // clamp pre-step state to >= 0 on load
#define SMEB_F_CLAMP_OUT 4u      // clamp output (RK101: no accept/reject)
#define SMEB_F_USE_C0 8u         // load c0 (s3)
#define SMEB_F_USE_S2 16u        // running sum s2 is live (RK323/RK435)
#define SMEB_F_USE_S2IN 32u      // s2 read from memory (else starts at 0)
#define SMEB_F_WRITE_S2 64u      // s2 written back
#define SMEB_F_FINAL 128u        // last stage: fused error norm
#define SMEB_F_LOWER_IS_CIN 256u // RK212: lower order solution = stage input
#define SMEB_F_DCDT_ONLY 512u    // only evaluate dc/dt into dcdt_out
#define SMEB_F_CLAMP_S2IN 1024u  // s2 is read from the (lazily clamped) c0

#define SMEB_GEOM_NN 0
#define SMEB_GEOM_BOX 1
#define SMEB_MEMB_FLAG 0x80000000u
#define SMEB_BLOCK 256

template <int S> struct SmebStageArgs {
  const double *cin;   // stage input: RHS is evaluated on this state
  const double *c0;    // pre-step state (reference: s3)
  const double *s2in;  // low-storage running sum in (reference: s2)
  double *s2out;       // running sum out (may alias s2in)
  double *cout;        // stage output
  double *dcdt_out;    // SMEB_F_DCDT_ONLY target
  const smeb_u32 *nn;  // [6 n] neighbour list (SMEB_GEOM_NN)
  smeb_u64 *err;       // [2] bit patterns of max abs / max rel error
  smeb_u32 n;          // voxels in this launch's index space
  smeb_u32 nx, ny, nz; // box dims (SMEB_GEOM_BOX); voxel = (z*nx + x)*ny + y
  smeb_u32 z_lo_ghost, z_hi_ghost; // 1 if plane 0 / nz-1 is a ghost plane
  smeb_u32 flags;
  double g1, g2, g3, bdt, delta; // stage coefficients (bdt = beta*dt)
  double cF, s2F, s3F, eps;      // finalise factors, error epsilon
  double D[3 * S];               // D/dx^2, D/dy^2, D/dz^2 per species
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
  smeb_u32 n_mvox;
  smeb_u32 clamp_mask;        // bit c: clamp comp_cin[c] on load
};

template <int S>
__device__ __forceinline__ void smeb_load_voxel(const double *__restrict__ p,
                                                smeb_u32 i, bool clamp,
                                                double *v) {
  if constexpr (S % 2 == 0) {
    const double2 *q = reinterpret_cast<const double2 *>(p) + (size_t)i * (S / 2);
#pragma unroll
    for (int k = 0; k < S / 2; ++k) {
      double2 t = q[k];
      v[2 * k] = t.x;
      v[2 * k + 1] = t.y;
    }
  } else {
    const double *q = p + (size_t)i * S;
#pragma unroll
    for (int k = 0; k < S; ++k) {
      v[k] = q[k];
    }
  }
  if (clamp) {
#pragma unroll
    for (int k = 0; k < S; ++k) {
      v[k] = v[k] < 0.0 ? 0.0 : v[k];
    }
  }
}

template <int S>
__device__ __forceinline__ void smeb_store_voxel(double *__restrict__ p,
                                                 smeb_u32 i, const double *v) {
  if constexpr (S % 2 == 0) {
    double2 *q = reinterpret_cast<double2 *>(p) + (size_t)i * (S / 2);
#pragma unroll
    for (int k = 0; k < S / 2; ++k) {
      q[k] = make_double2(v[2 * k], v[2 * k + 1]);
    }
  } else {
    double *q = p + (size_t)i * S;
#pragma unroll
    for (int k = 0; k < S; ++k) {
      q[k] = v[k];
    }
  }
}

// max-reduction of the two error norms over the block, then one atomicMax per
// block on the ordered bit patterns (values are >= 0 and never NaN, because
// like std::max in pixelsim_impl.cpp:783-788 a NaN or negative candidate is
// never taken).
__device__ __forceinline__ void smeb_block_max2(double eabs, double erel,
                                                smeb_u64 *err) {
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
    if (lane == 0) {
      if (eabs > 0.0) {
        atomicMax(err, (smeb_u64)__double_as_longlong(eabs));
      }
      if (erel > 0.0) {
        atomicMax(err + 1, (smeb_u64)__double_as_longlong(erel));
      }
    }
  }
}

// RK stage update + error norm for one voxel, given c (stage input) and the
// complete dc/dt in f.  Association order: pixelsim_impl.cpp:627,675,699,
// 718-719,741,783-788.
template <int S>
__device__ __forceinline__ void
smeb_stage_update(const SmebStageArgs<S> &a, smeb_u32 i, const double *c,
                  const double *f, double &eabs, double &erel) {
  const smeb_u32 flags = a.flags;
  if (flags & SMEB_F_DCDT_ONLY) {
    smeb_store_voxel<S>(a.dcdt_out, i, f);
    return;
  }
  double c0v[S], s2v[S], cn[S];
#pragma unroll
  for (int s = 0; s < S; ++s) {
    c0v[s] = 0.0;
    s2v[s] = 0.0;
  }
  if (flags & SMEB_F_USE_C0) {
    smeb_load_voxel<S>(a.c0, i, flags & SMEB_F_CLAMP_C0, c0v);
  }
  if (flags & SMEB_F_USE_S2) {
    if (flags & SMEB_F_USE_S2IN) {
      smeb_load_voxel<S>(a.s2in, i, flags & SMEB_F_CLAMP_S2IN, s2v);
    }
#pragma unroll
    for (int s = 0; s < S; ++s) {
      s2v[s] = s2v[s] + a.delta * c[s];
    }
    if (flags & SMEB_F_WRITE_S2) {
      smeb_store_voxel<S>(a.s2out, i, s2v);
    }
  }
#pragma unroll
  for (int s = 0; s < S; ++s) {
    double v = a.g1 * c[s] + a.g2 * s2v[s] + a.g3 * c0v[s] + a.bdt * f[s];
    if (flags & SMEB_F_CLAMP_OUT) {
      v = v < 0.0 ? 0.0 : v;
    }
    cn[s] = v;
  }
  smeb_store_voxel<S>(a.cout, i, cn);
  if (flags & SMEB_F_FINAL) {
#pragma unroll
    for (int s = 0; s < S; ++s) {
      const double lower = (flags & SMEB_F_LOWER_IS_CIN)
                               ? c[s]
                               : a.cF * cn[s] + a.s2F * s2v[s] + a.s3F * c0v[s];
      const double e = fabs(cn[s] - lower);
      eabs = e > eabs ? e : eabs;
      const double norm = 0.5 * (cn[s] + c0v[s] + a.eps);
      const double r = e / norm;
      erel = r > erel ? r : erel;
    }
  }
}

// reaction + uniform diffusion for one voxel whose six neighbour indices are
// known; cin values of the voxel itself are returned in c.
template <class R>
__device__ __forceinline__ void
smeb_rhs(const SmebStageArgs<R::S> &a, smeb_u32 i, const smeb_u32 *nb,
         double *c, double *f) {
  constexpr int S = R::S;
  const bool clamp = a.flags & SMEB_F_CLAMP_CIN;
  double xp[S], xm[S], yp[S], ym[S], zp[S], zm[S];
  smeb_load_voxel<S>(a.cin, i, clamp, c);
  smeb_load_voxel<S>(a.cin, nb[0], clamp, xp);
  smeb_load_voxel<S>(a.cin, nb[1], clamp, xm);
  smeb_load_voxel<S>(a.cin, nb[2], clamp, yp);
  smeb_load_voxel<S>(a.cin, nb[3], clamp, ym);
  smeb_load_voxel<S>(a.cin, nb[4], clamp, zp);
  smeb_load_voxel<S>(a.cin, nb[5], clamp, zm);
  R::eval(f, c);
#pragma unroll
  for (int s = 0; s < S; ++s) {
    f[s] += a.D[3 * s] * (xp[s] + xm[s] - 2.0 * c[s]) +
            a.D[3 * s + 1] * (yp[s] + ym[s] - 2.0 * c[s]) +
            a.D[3 * s + 2] * (zp[s] + zm[s] - 2.0 * c[s]);
  }
}

// ---- direct-gather stage kernel: one thread per voxel -----------------------
// GEOM_NN: neighbour indices from the uint32 list (24 B per voxel); voxels
// flagged SMEB_MEMB_FLAG are left to smeb_stage_memb.
// GEOM_BOX: neighbours implicit; planes marked as ghost are read but not
// updated (z-slab decomposition).
template <class R, int GEOM>
__global__ void __launch_bounds__(SMEB_BLOCK)
    smeb_stage(const __grid_constant__ SmebStageArgs<R::S> a) {
  constexpr int S = R::S;
  const smeb_u32 i = blockIdx.x * SMEB_BLOCK + threadIdx.x;
  double eabs = 0.0, erel = 0.0;
  if (i < a.n) {
    smeb_u32 nb[6];
    bool active = true;
    if constexpr (GEOM == SMEB_GEOM_BOX) {
      const smeb_u32 y = i % a.ny;
      const smeb_u32 t = i / a.ny;
      const smeb_u32 x = t % a.nx;
      const smeb_u32 z = t / a.nx;
      const smeb_u32 sx = a.ny, sz = a.nx * a.ny;
      nb[0] = x + 1 < a.nx ? i + sx : i;
      nb[1] = x > 0 ? i - sx : i;
      nb[2] = y + 1 < a.ny ? i + 1 : i;
      nb[3] = y > 0 ? i - 1 : i;
      nb[4] = z + 1 < a.nz ? i + sz : i;
      nb[5] = z > 0 ? i - sz : i;
      active = !((z == 0 && a.z_lo_ghost) || (z + 1 == a.nz && a.z_hi_ghost));
    } else {
      const uint2 *q = reinterpret_cast<const uint2 *>(a.nn) + (size_t)i * 3;
      const uint2 n0 = q[0], n1 = q[1], n2 = q[2];
      active = !(n0.x & SMEB_MEMB_FLAG);
      nb[0] = n0.x & ~SMEB_MEMB_FLAG;
      nb[1] = n0.y;
      nb[2] = n1.x;
      nb[3] = n1.y;
      nb[4] = n2.x;
      nb[5] = n2.y;
    }
    if (active) {
      double c[S], f[S];
      smeb_rhs<R>(a, i, nb, c, f);
      smeb_stage_update<S>(a, i, c, f, eabs, erel);
    }
  }
  if (a.flags & SMEB_F_FINAL) {
    smeb_block_max2(eabs, erel, a.err);
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
               const SmebMembArgs &m, double *f) {
  constexpr int SA = M::SA, SB = M::SB;
  double in[SA + SB + 1], out[SA + SB + 1];
  const bool ownIsB = (info >> 16) & 1u;
  const smeb_u32 dir = (info >> 17) & 7u;
  const double L = m.flux_len[(info & 0xffffu) * 6 + dir];
  if (!ownIsB) {
    if constexpr (S_OWN == SA) {
#pragma unroll
      for (int s = 0; s < SA; ++s) {
        in[s] = cown[s];
      }
      if constexpr (SB > 0) {
        smeb_load_voxel<SB>(m.comp_cin[M::CB], partner,
                            (m.clamp_mask >> M::CB) & 1u, in + SA);
      }
      M::eval(out, in);
#pragma unroll
      for (int s = 0; s < SA; ++s) {
        f[s] += out[s] / L;
      }
    }
  } else {
    if constexpr (S_OWN == SB) {
#pragma unroll
      for (int s = 0; s < SB; ++s) {
        in[SA + s] = cown[s];
      }
      if constexpr (SA > 0) {
        smeb_load_voxel<SA>(m.comp_cin[M::CA], partner,
                            (m.clamp_mask >> M::CA) & 1u, in);
      }
      M::eval(out, in);
#pragma unroll
      for (int s = 0; s < SB; ++s) {
        f[s] += out[SA + s] / L;
      }
    }
  }
}

// generated per model: switch over the membrane id -> smeb_memb_term<Memb_k>
template <int S_OWN>
__device__ __forceinline__ void
smeb_memb_dispatch(smeb_u32 info, smeb_u32 partner, const double *cown,
                   const SmebMembArgs &m, double *f);

template <class R>
__global__ void __launch_bounds__(SMEB_BLOCK)
    smeb_stage_memb(const __grid_constant__ SmebStageArgs<R::S> a,
                    const __grid_constant__ SmebMembArgs m) {
  constexpr int S = R::S;
  const smeb_u32 k = blockIdx.x * SMEB_BLOCK + threadIdx.x;
  double eabs = 0.0, erel = 0.0;
  if (k < m.n_mvox) {
    const smeb_u32 i = m.mvox[k];
    const uint2 *q = reinterpret_cast<const uint2 *>(a.nn) + (size_t)i * 3;
    const uint2 n0 = q[0], n1 = q[1], n2 = q[2];
    smeb_u32 nb[6] = {n0.x & ~SMEB_MEMB_FLAG, n0.y, n1.x, n1.y, n2.x, n2.y};
    double c[S], f[S];
    smeb_rhs<R>(a, i, nb, c, f);
    const smeb_u32 e1 = m.mstart[k + 1];
    for (smeb_u32 e = m.mstart[k]; e < e1; ++e) {
      const SmebMembEntry ent = m.ent[e];
      smeb_memb_dispatch<S>(ent.info, ent.partner, c, m, f);
    }
    smeb_stage_update<S>(a, i, c, f, eabs, erel);
  }
  if (a.flags & SMEB_F_FINAL) {
    smeb_block_max2(eabs, erel, a.err);
  }
}

// ---- on-demand elementwise kernels ------------------------------------------
// lower-order solution (pixelsim_impl.cpp:737-743; RK212: s2 = stage-2 input)
struct SmebLowerArgs {
  const double *cin, *c0, *s2in, *cnew;
  double *out;
  smeb_u64 n; // values
  smeb_u32 flags;
  double delta, cF, s2F, s3F;
};

__global__ void __launch_bounds__(SMEB_BLOCK)
    smeb_lower_order(const __grid_constant__ SmebLowerArgs a) {
  const smeb_u64 i = (smeb_u64)blockIdx.x * SMEB_BLOCK + threadIdx.x;
  if (i >= a.n) {
    return;
  }
  const double c = a.cin[i];
  if (a.flags & SMEB_F_LOWER_IS_CIN) {
    a.out[i] = c;
    return;
  }
  double c0 = a.c0[i];
  if (a.flags & SMEB_F_CLAMP_C0) {
    c0 = c0 < 0.0 ? 0.0 : c0;
  }
  double s2 = (a.flags & SMEB_F_USE_S2IN) ? a.s2in[i] : 0.0;
  if (a.flags & SMEB_F_CLAMP_S2IN) {
    s2 = s2 < 0.0 ? 0.0 : s2;
  }
  s2 = s2 + a.delta * c;
  a.out[i] = a.cF * a.cnew[i] + a.s2F * s2 + a.s3F * c0;
}

// materialise the lazy clamp into a contiguous copy (download path)
__global__ void __launch_bounds__(SMEB_BLOCK)
    smeb_clamp_copy(const double *__restrict__ in, double *__restrict__ out,
                    smeb_u64 n, int clamp) {
  const smeb_u64 i = (smeb_u64)blockIdx.x * SMEB_BLOCK + threadIdx.x;
  if (i < n) {
    double v = in[i];
    if (clamp) {
      v = v < 0.0 ? 0.0 : v;
    }
    out[i] = v;
  }
}
