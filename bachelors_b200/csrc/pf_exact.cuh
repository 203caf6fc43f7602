// pf_exact.cuh - the analytic ("exact") mode of the reference: exact.h:17-96, simulation.cu:179-184, 1069-1089,
// main.cpp:104-110, config.h:493-509. A radially symmetric Stefan-like solution used to check convergence order:
//   * do_exact adds the source term fu = exact_fu(iter*dt) to every dT/dt (host-computed constant of a step);
//   * solver = exact writes the analytic Phi, T at params.time instead of stepping;
//   * the initial conditions become the analytic profile.
// The reference's exact Dirichlet values (simulation.cu:60-66) sit behind COMPILE_EXACT, which its build leaves
// undefined (simulation.h:17 "do not use slightly broken!"): not reproduced, as in the reference binary.
// Host + device, Real arithmetic as in the reference; exp/erf/hypot are the CUDA math library on the device.
#pragma once
#include <math.h>

namespace pf {

template <typename Real> struct ExactParams {                  // get_static_exact_params, exact.h:90-96
    static constexpr Real lambda = (Real) 0.5, R_ini = (Real) 0.25, epsilon = (Real) 0.001;
};

template <typename Real> __host__ __device__ inline Real ex_sqrt(Real v);
template <> __host__ __device__ inline double ex_sqrt<double>(double v) { return sqrt(v); }
template <> __host__ __device__ inline float  ex_sqrt<float>(float v)   { return sqrtf(v); }
__host__ __device__ inline double ex_exp(double v) { return exp(v); }
__host__ __device__ inline float  ex_exp(float v)  { return expf(v); }
__host__ __device__ inline double ex_erf(double v) { return erf(v); }
__host__ __device__ inline float  ex_erf(float v)  { return erff(v); }
__host__ __device__ inline double ex_hypot(double a, double b) { return hypot(a, b); }
__host__ __device__ inline float  ex_hypot(float a, float b)   { return hypotf(a, b); }

template <typename Real> __host__ __device__ inline Real exact_R(Real t)                 // exact.h:17-20
{
    using P = ExactParams<Real>;
    return ex_sqrt<Real>(P::R_ini * P::R_ini + 2 * P::lambda * t);
}
template <typename Real> __host__ __device__ inline Real exact_U(Real Rt)                // exact.h:22-25
{
    using P = ExactParams<Real>;
    return -P::epsilon * (P::lambda + 2) / Rt;
}
template <typename Real> __host__ __device__ inline Real exact_T(Real s)                 // exact.h:27-35
{
    const Real l = ExactParams<Real>::lambda;
    const Real pi = (Real) 3.14159265359;
    const Real sqrtl2 = ex_sqrt<Real>(l / 2);
    const Real integral = ex_exp(-l / 2) - 1 / s * ex_exp(-l / 2 * s * s) + sqrtl2 * pi * (ex_erf(sqrtl2) - ex_erf(s * sqrtl2));
    return -l * ex_exp(l / 2) * integral;
}
template <typename Real> __host__ __device__ inline Real exact_fu(Real t)                // exact.h:37-42
{
    using P = ExactParams<Real>;
    const Real Rt = exact_R<Real>(t);
    const Real l = P::lambda;
    return P::epsilon * l * (l + 2) / (Rt * Rt * Rt);
}
template <typename Real> __host__ __device__ inline Real exact_u(Real t, Real r)         // exact.h:44-53
{
    const Real Rt = exact_R<Real>(t);
    Real ut = exact_U<Real>(Rt);
    if(r > Rt) ut += exact_T<Real>(r / Rt);
    return ut;
}
template <typename Real> __host__ __device__ inline Real exact_phi(Real t, Real r)       // exact.h:55-58
{
    return r <= exact_R<Real>(t) ? (Real) 1 : (Real) 0;
}
template <typename Real> __host__ __device__ inline Real exact_phi_ini(Real r, Real xi)  // exact_corresponing_phi_ini, exact.h:70-88
{
    const Real fade = 1;
    const Real lo = ExactParams<Real>::R_ini - fade * xi / 2;
    const Real hi = ExactParams<Real>::R_ini + fade * xi / 2;
    if(r < lo) return 1;
    if(r > hi) return 0;
    return 1 - (r - lo) / (hi - lo);
}

#ifdef __CUDACC__
// exact_solver_step: reference simulation.cu:1069-1089. Write-only, one thread per cell.
template <typename Real>
__global__ void __launch_bounds__(256)
k_exact_fields(int nx, int ny, Real dx, Real dy, Real L0, Real t, Real* __restrict__ next_F, Real* __restrict__ next_U)
{
    const int xi = blockIdx.x * blockDim.x + threadIdx.x, yi = blockIdx.y;
    if(xi >= nx || yi >= ny) return;
    const Real x = (Real)(((Real) xi + 0.5) * dx - L0 / 2);            // `+ 0.5` promotes to double for Real = float, as in the reference
    const Real y = (Real)(((Real) yi + 0.5) * dy - L0 / 2);
    const Real r = ex_hypot(x, y);
    const size_t i = (size_t) xi + (size_t) yi * (size_t) nx;
    next_U[i] = exact_u<Real>(t, r);
    next_F[i] = exact_phi<Real>(t, r);
}
#endif

} // namespace pf
