// Build-time instantiation of the kernel template with a default model
// (gray-scott kinetics, S = 2, plus a 2+2 species linear membrane) so that
// nvcc/ptxas check the template for sm_100a and SASS can be inspected without
// a GPU (cuobjdump -sass build/default_model.cubin).  At run time the same
// template text is compiled per model by NVRTC (sim.cpp: compileModule).
#define SMEB_NCOMP 2
#include "device_kernels.cuh"

struct Reac0 {
  static constexpr int S = 2, SP = 2, NNS = 0, NX = 0;
  static constexpr bool STORAGE = false, NUD = false;
  static __device__ __forceinline__ void nonSpatial(double *, const double *) {}
  static __device__ __forceinline__ void nonSpatialPick(const double *, double *) {}
  static __device__ __forceinline__ void storage(double *) {}
  static __device__ __forceinline__ bool zeroStorage(int) { return false; }
  static __device__ __forceinline__ void eval(double *sme_out,
                                              const double *sme_values,
                                              const double *) {
    const double sme_var_0 = sme_values[0];
    const double sme_var_1 = sme_values[1];
    sme_out[0] = -sme_var_0 * sme_var_1 * sme_var_1 + 0.037 * (1.0 - sme_var_0);
    sme_out[1] = sme_var_0 * sme_var_1 * sme_var_1 - 0.097 * sme_var_1;
  }
};
struct Reac1 : Reac0 {};
// a model using every post-processing hook: 2 species + time + x,y,z padding,
// species 1 non-spatial, storage 2 for species 0
struct Reac2 {
  static constexpr int S = 6, SP = 2, NNS = 1, NX = 1;
  static constexpr bool STORAGE = true, NUD = true;
  // one cross-diffusion term: target 0, source 1, coefficient 0.1 * c0
  static __device__ __forceinline__ void crossCoef(double *k, const double *v,
                                                   const double *) {
    k[0] = 0.1 * v[0];
  }
  static __device__ __forceinline__ void
  cross(double *f, const double *c, const double *xp, const double *xm,
        const double *yp, const double *ym, const double *zp, const double *zm,
        const double *h2, const double *p) {
    double kc[1], k0[1], k1[1], k2[1], k3[1], k4[1], k5[1];
    crossCoef(kc, c, p);
    crossCoef(k0, xp, p);
    crossCoef(k1, xm, p);
    crossCoef(k2, yp, p);
    crossCoef(k3, ym, p);
    crossCoef(k4, zp, p);
    crossCoef(k5, zm, p);
    const double dxp = 0.5 * (kc[0] + k0[0]) / h2[0];
    const double dxm = 0.5 * (kc[0] + k1[0]) / h2[0];
    const double dyp = 0.5 * (kc[0] + k2[0]) / h2[1];
    const double dym = 0.5 * (kc[0] + k3[0]) / h2[1];
    const double dzp = 0.5 * (kc[0] + k4[0]) / h2[2];
    const double dzm = 0.5 * (kc[0] + k5[0]) / h2[2];
    f[0] += dxp * (xp[1] - c[1]) - dxm * (c[1] - xm[1]) +
            dyp * (yp[1] - c[1]) - dym * (c[1] - ym[1]) +
            dzp * (zp[1] - c[1]) - dzm * (c[1] - zm[1]);
  }
  static __device__ __forceinline__ void nonSpatial(double *f, const double *av) {
    f[1] = av[0];
  }
  static __device__ __forceinline__ void nonSpatialPick(const double *f, double *v) {
    v[0] = f[1];
  }
  static __device__ __forceinline__ void storage(double *f) {
    f[0] *= 0.5;
    f[1] *= 0.0;
  }
  static __device__ __forceinline__ bool zeroStorage(int s) { return s == 1; }
  static __device__ __forceinline__ void eval(double *sme_out,
                                              const double *sme_values,
                                              const double *) {
    sme_out[0] = sme_values[2] * sme_values[3] - sme_values[0];
    sme_out[1] = sme_values[0] - sme_values[1] * sme_values[4];
    sme_out[2] = 1.0;
    sme_out[3] = 0.0;
    sme_out[4] = 0.0;
    sme_out[5] = 0.0;
  }
};
struct Memb0 {
  static constexpr int SA = 2, SB = 2, SAP = 2, SBP = 2, NE = 0, CA = 0, CB = 1;
  static __device__ __forceinline__ void eval(double *sme_out,
                                              const double *sme_values,
                                              const double *) {
    sme_out[0] = 0.1 * (sme_values[2] - sme_values[0]);
    sme_out[1] = 0.0;
    sme_out[2] = -sme_out[0];
    sme_out[3] = 0.0;
  }
};
template <int S_OWN>
__device__ __forceinline__ void
smeb_memb_dispatch(smeb_u32 info, smeb_u32 partner, const double *cown,
                   const SmebMembArgs &m, const double *params, double *f) {
  switch (info & 0xffffu) {
  case 0:
    smeb_memb_term<Memb0, S_OWN>(info, partner, cown, m, params, f);
    break;
  default:
    break;
  }
}

// RK212: stage 1 (Euler) and stage 2 (final, fused error norm, clamp)
#define F1 0u
#define F2 (SMEB_F_USE_C0 | SMEB_F_FINAL | SMEB_F_LOWER_IS_CIN | SMEB_F_CLAMP_OUT)
// RK435 stage 3: generic low-storage update with the running sum in memory
#define F3 (SMEB_F_GENERIC | SMEB_F_USE_S2 | SMEB_F_USE_S2IN | SMEB_F_WRITE_S2)
template __global__ void
smeb_stage_tile<Reac0, F1, 16, 64, 5, true>(const __grid_constant__ SmebStageArgs<2>,
                                            const __grid_constant__ SmebTensorMap,
                                            const smeb_u32);
template __global__ void
smeb_stage_tile<Reac0, F2, 16, 64, 5, true>(const __grid_constant__ SmebStageArgs<2>,
                                            const __grid_constant__ SmebTensorMap,
                                            const smeb_u32);
template __global__ void
smeb_stage_tile<Reac0, F3, 16, 64, 5, true>(const __grid_constant__ SmebStageArgs<2>,
                                            const __grid_constant__ SmebTensorMap,
                                            const smeb_u32);
// voxel lists staged in shared memory piece by piece (2-D and 3-D)
template __global__ void
smeb_stage_seg<Reac0, F1, 512, 2, false>(const __grid_constant__ SmebStageArgs<2>);
template __global__ void
smeb_stage_seg<Reac0, F2, 512, 2, false>(const __grid_constant__ SmebStageArgs<2>);
template __global__ void
smeb_stage_seg<Reac0, F2, 256, 2, true>(const __grid_constant__ SmebStageArgs<2>);
// member-batched ensemble variants (gridDim.y = members)
template __global__ void
smeb_stage_nn<Reac0, F2 | SMEB_F_MEMBERS>(const __grid_constant__ SmebStageArgs<2>);
template __global__ void
smeb_stage_memb<Reac0, F2 | SMEB_F_MEMBERS>(const __grid_constant__ SmebStageArgs<2>,
                                            const __grid_constant__ SmebMembArgs);
template __global__ void
smeb_stage_seg<Reac0, F2 | SMEB_F_MEMBERS, 512, 2, false>(
    const __grid_constant__ SmebStageArgs<2>);
// z-slab variants: halo exchange inside the kernel
template __global__ void
smeb_stage_tile<Reac0, F1 | SMEB_F_HALO, 16, 64, 5, true>(
    const __grid_constant__ SmebStageArgs<2>, const __grid_constant__ SmebTensorMap,
    const smeb_u32);
template __global__ void
smeb_stage_tile<Reac0, F2 | SMEB_F_HALO, 16, 64, 5, true>(
    const __grid_constant__ SmebStageArgs<2>, const __grid_constant__ SmebTensorMap,
    const smeb_u32);
template __global__ void
smeb_stage_tile<Reac0, F2, 16, 64, 3, false>(const __grid_constant__ SmebStageArgs<2>,
                                             const __grid_constant__ SmebTensorMap,
                                             const smeb_u32);
template __global__ void
smeb_stage_box<Reac0, F2>(const __grid_constant__ SmebStageArgs<2>);
template __global__ void
smeb_stage_nn<Reac0, F2>(const __grid_constant__ SmebStageArgs<2>);
template __global__ void
smeb_stage_memb<Reac0, F2>(const __grid_constant__ SmebStageArgs<2>,
                           const __grid_constant__ SmebMembArgs);
// pre-pass and stage kernels of the model with non-spatial species / storage /
// time-space padding
#define F4 (SMEB_F_SUM_NS)
template __global__ void
smeb_stage_nn<Reac2, F4>(const __grid_constant__ SmebStageArgs<6>);
template __global__ void
smeb_stage_nn<Reac2, F2>(const __grid_constant__ SmebStageArgs<6>);
template __global__ void
smeb_stage_tile<Reac2, F4, 8, 32, 4, true>(const __grid_constant__ SmebStageArgs<6>,
                                            const __grid_constant__ SmebTensorMap,
                                            const smeb_u32);
template __global__ void
smeb_stage_tile<Reac2, F2, 8, 32, 4, true>(const __grid_constant__ SmebStageArgs<6>,
                                            const __grid_constant__ SmebTensorMap,
                                            const smeb_u32);
template __global__ void
smeb_stage_box<Reac2, F2>(const __grid_constant__ SmebStageArgs<6>);
template __global__ void smeb_xd_max<Reac2>(const __grid_constant__ SmebStageArgs<6>);
template __global__ void smeb_stats<Reac0>(const double *, smeb_u32, smeb_u32, double *);
template __global__ void smeb_stats<Reac2>(const double *, smeb_u32, smeb_u32, double *);
// zero-storage relaxation kernels
#define F5 (SMEB_F_RELAX_A)
#define F6 (SMEB_F_RELAX_B | SMEB_F_USE_C0)
template __global__ void
smeb_stage_nn<Reac2, F5>(const __grid_constant__ SmebStageArgs<6>);
template __global__ void
smeb_stage_nn<Reac2, F6>(const __grid_constant__ SmebStageArgs<6>);
template __global__ void
smeb_stage_tile<Reac2, F6, 8, 32, 4, true>(const __grid_constant__ SmebStageArgs<6>,
                                           const __grid_constant__ SmebTensorMap,
                                           const smeb_u32);
