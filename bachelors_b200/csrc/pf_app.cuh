// pf_app.cuh - device kernels of the rows around the stepping path:
//   * k_init_fields : sim_make_initial_conditions (reference main.cpp:93-136) on the device;
//   * k_render_rgba : the viewer's colour-map fragment shader (reference gl.h:221-283) as a CUDA kernel.
// Both are one pass, one thread per cell, HBM-bound (write 2 reals / read 1 real + write 4 bytes per cell).
#pragma once
#include "pf_common.cuh"

namespace pf {

struct InitArgs {
    double inside_phi, inside_T, outside_phi, outside_T;
    double cx, cy, radius, fade;
    double sq_x0, sq_y0, sq_x1, sq_y1;
    double L0, xi;
    int nx, ny_global;
    int row0, rows;          // global row of local row 0, rows to fill (ghost rows included)
    int do_exact;            // analytic initial profile (main.cpp:104-110) instead of the circle / square blend
    int wrap;                // ghost rows outside [0, ny) hold the periodic image (else the clamped row; never read)
};

// hypot(a, b) in double, rounded like a correctly rounded hypot in all but a vanishing fraction of inputs: the square
// root of the fused sum, then one correction from the exactly computed residual a^2 + b^2 - h^2 (error-free products via
// fma). The host's libm hypot (what the reference calls, main.cpp:116) is correctly rounded or within 1 ulp of it.
__device__ __forceinline__ double hypot_cr(double a, double b)
{
    a = fabs(a); b = fabs(b);
    const double hi = fmax(a, b), lo = fmin(a, b);
    if(!(hi > 0) || hi > 1e150 || lo < 1e-150) return ::hypot(a, b);       // zeros, huge/tiny scales, non-finite: library path
    const double p = hi * hi, pe = ::fma(hi, hi, -p);
    const double q = lo * lo, qe = ::fma(lo, lo, -q);
    const double h = sqrt(p + q);
    const double hh = h * h, he = ::fma(h, h, -hh);
    const double resid = ((p - hh) + q) + ((pe + qe) - he);
    return h + resid / (h + h);
}

template <typename Real>
__global__ void __launch_bounds__(256)
k_init_fields(const InitArgs a, Real* __restrict__ F, Real* __restrict__ U)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int ly = blockIdx.y;
    if(x >= a.nx || ly >= a.rows) return;
    int gy = a.row0 + ly;
    if(a.wrap) gy = ((gy % a.ny_global) + a.ny_global) % a.ny_global;
    else       gy = gy < 0 ? 0 : (gy >= a.ny_global ? a.ny_global - 1 : gy);

    // main.cpp:101: ((x+0.5)) / nx * L0, left to right, in double
    const double px = ((double) x + 0.5) / (double) a.nx * a.L0;
    const double py = ((double) gy + 0.5) / (double) a.ny_global * a.L0;
    const size_t i = (size_t) ly * (size_t) a.nx + (size_t) x;
    if(a.do_exact)
    {
        const Real r = (Real) hypot_cr(px - a.L0 / 2, py - a.L0 / 2);                  // main.cpp:106
        F[i] = exact_phi_ini<Real>(r, (Real) a.xi);
        U[i] = exact_u<Real>((Real) 0, r);
        return;
    }
    // main.cpp:113-114: lo/hi are Real in the reference; for Real = float they are rounded to float first
    const Real lo_r = (Real)(a.radius - a.xi * a.fade / 2);
    const Real hi_r = (Real)(a.radius + a.xi * a.fade / 2);
    const double lo = (double) lo_r, hi = (double) hi_r;
    const double r = hypot_cr(a.cx - px, a.cy - py);
    const bool within = (a.sq_x0 <= px && px < a.sq_x1) && (a.sq_y0 <= py && py < a.sq_y1);
    double circle = 1 - (r - lo) / (hi - lo);
    if(circle > 1) circle = 1;
    if(circle < 0) circle = 0;
    const double cube = within ? 1 : 0;
    const double factor = cube > circle ? cube : circle;
    F[i] = (Real)(factor * a.inside_phi + (1 - factor) * a.outside_phi);
    U[i] = (Real)(factor * a.inside_T + (1 - factor) * a.outside_T);
}

// gl.h:236-281, float arithmetic as in GLSL. Colour channels in [0, 1] -> UNORM8 by round-to-nearest (what the GL
// framebuffer conversion does).
__device__ __forceinline__ unsigned char unorm8(float c)
{
    c = c < 0.0f ? 0.0f : (c > 1.0f ? 1.0f : c);
    return (unsigned char)(int) floorf(c * 255.0f + 0.5f);
}

template <typename Real>
__global__ void __launch_bounds__(256)
k_render_rgba(size_t N, const Real* __restrict__ field, float minVal, float maxVal, uchar4* __restrict__ out)
{
    const float PI = 3.14159265359f;
    for(size_t i = blockIdx.x * (size_t) blockDim.x + threadIdx.x; i < N; i += (size_t) gridDim.x * blockDim.x)
    {
        float val = (float) field[i];                      // the viewer downloads as float (gl.h:396)
        float r, g, b;
        if(val != val)          { r = 1; g = 0; b = 1; }
        else if(val < minVal)   { float d = (1 - atanf(minVal - val) / PI * 2) * 0.3f; r = g = b = d; }
        else if(val > maxVal)   { float d = atanf(val - minVal) / PI * 2 * 0.3f + 0.7f; r = g = b = d; }
        else
        {
            val = fminf(fmaxf(val, minVal), maxVal - 0.0001f);
            const float d = maxVal - minVal;
            val = d == 0.0f ? 0.5f : (val - minVal) / d;
            const float m = 0.25f;
            const float num = floorf(val / m);
            const float s = (val - num * m) / m;
            r = 0; g = 0; b = 0;
            switch((int) num)
            {
                case 0: r = 0.0f; g = s; b = 1.0f; break;
                case 1: r = 0.0f; g = 1.0f; b = 1.0f - s; break;
                case 2: r = s; g = 1.0f; b = 0.0f; break;
                case 3: r = 1.0f; g = 1.0f - s; b = 0.0f; break;
            }
        }
        out[i] = make_uchar4(unorm8(r), unorm8(g), unorm8(b), 255);
    }
}

} // namespace pf
