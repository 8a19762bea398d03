// v and w of the truncate factor, shared by the round
// and the streaming rating kernels.
//
// win:  x = t - eps, v = pdf(x)/cdf(x), w = v (v + x).  On |x| < 6 -- everything
// a real ladder produces -- v comes from a piecewise degree-10 polynomial table
// (vtable.inc, generated at 60 digits; max relative error 4e-16 in double)
// evaluated with Estrin's scheme: ~6 dependent FMAs and one L1-resident table
// row instead of the ~100-deep chain of erfcx + division.  The v/w chain sits on
// the critical path of every dependency hop, so its latency, not its
// throughput, is what counts.  Outside the table: v = sqrt(2/pi)/erfcx(-x/sqrt2).
// draw: pdf/cdf differences at +-eps - |t|, or the analytic limit v = -t, w = 1
// when draw_probability = 0.
// Returns false where the trueskill package raises FloatingPointError.
#pragma once
#include "vtable.inc"

namespace tsvw {

constexpr double kInvSqrt2 = 0.70710678118654752440;
constexpr double kSqrt2OverPi = 0.79788456080286535588;   // sqrt(2/pi)
constexpr double kInvSqrt2Pi = 0.39894228040143267794;

__device__ __forceinline__ double npdf(double x) { return __dmul_rn(kInvSqrt2Pi, exp(__dmul_rn(__dmul_rn(-0.5, x), x))); }
__device__ __forceinline__ double ncdf(double x) { return 0.5 * erfc(-x * kInvSqrt2); }

__device__ __forceinline__ double v_win_table(double x)
{
    static_assert(TS_VTAB_DEG == 10, "Estrin scheme below is written for degree 10");
    const double tpos = (x + TS_VTAB_XMAX) * (TS_VTAB_NSEG / (2.0 * TS_VTAB_XMAX));
    const int i = (int)tpos;
    const double u = fma(2.0, tpos - (double)i, -1.0);
    const double *c = g_vtab + i * (TS_VTAB_DEG + 1);
    const double c0 = __ldg(c), c1 = __ldg(c + 1), c2 = __ldg(c + 2), c3 = __ldg(c + 3), c4 = __ldg(c + 4),
                 c5 = __ldg(c + 5), c6 = __ldg(c + 6), c7 = __ldg(c + 7), c8 = __ldg(c + 8), c9 = __ldg(c + 9),
                 c10 = __ldg(c + 10);
    const double u2 = u * u, u4 = u2 * u2, u8 = u4 * u4;
    const double p01 = fma(c1, u, c0), p23 = fma(c3, u, c2), p45 = fma(c5, u, c4), p67 = fma(c7, u, c6),
                 p89 = fma(c9, u, c8);
    const double q0 = fma(p23, u2, p01), q1 = fma(p67, u2, p45), q2 = fma(c10, u2, p89);
    return fma(q2, u8, fma(q1, u4, q0));
}

__device__ __forceinline__ bool trunc_vw(double t, double eps, bool is_draw, int draw_is_zero, double &v, double &w)
{
    if (!is_draw) {
        const double x = t - eps;
        if (fabs(x) < TS_VTAB_XMAX) {
            v = v_win_table(x);
        } else {
            // no cancellation in either tail
            v = kSqrt2OverPi / erfcx(-x * kInvSqrt2);
        }
        w = __dmul_rn(v, __dadd_rn(v, x));
        if (w > 0.0 && w < 1.0) return true;
        // far favourite-wins tail: erfcx overflows where the 50-digit reference
        // still sees 0 < w << 1e-300; the update is then the tau-only one.
        if (x > 30.0 && w >= 0.0) { v = 0.0; w = 0.0; return true; }
        return false;
    }
    if (draw_is_zero == 2) return false;   // the integrator asked for a draw at draw_probability = 0 to fail like the float backend
    if (draw_is_zero) {   // draw_probability = 0: limit of v_draw/w_draw for eps -> 0
        v = -t;
        w = 1.0;
        return true;
    }
    const double at = fabs(t);
    const double a = eps - at, b = -eps - at;
    const double mass = ncdf(a) - ncdf(b);
    if (mass == 0.0) return false;
    const double pa = npdf(a), pb = npdf(b);
    const double vv = __ddiv_rn(__dadd_rn(pb, -pa), mass);
    w = __fma_rn(vv, vv, __ddiv_rn(__fma_rn(a, pa, -__dmul_rn(b, pb)), mass));   // explicit rounding: same bits in every kernel
    v = t < 0.0 ? -vv : vv;
    return true;
}

}  // namespace tsvw
