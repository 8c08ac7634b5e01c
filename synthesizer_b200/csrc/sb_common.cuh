// Shared host/device helpers of libsynth_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/synth_b200.h"

namespace sb {

void set_error(const char *fmt, ...);
void count_launch(int n = 1);

#define SB_CUDA_CHECK(expr)                                                          \
    do {                                                                             \
        cudaError_t _e = (expr);                                                     \
        if (_e != cudaSuccess) {                                                     \
            sb::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),    \
                          __FILE__, __LINE__);                                       \
            return SB_ECUDA;                                                         \
        }                                                                            \
    } while (0)

#define SB_LAUNCH_CHECK()                                                            \
    do {                                                                             \
        cudaError_t _e = cudaGetLastError();                                         \
        if (_e != cudaSuccess) {                                                     \
            sb::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e),\
                          __FILE__, __LINE__);                                       \
            return SB_ECUDA;                                                         \
        }                                                                            \
        sb::count_launch();                                                          \
    } while (0)

#define SB_REQUIRE(cond, msg)                                                        \
    do {                                                                             \
        if (!(cond)) { sb::set_error("%s (%s)", msg, #cond); return SB_EINVAL; }     \
    } while (0)

constexpr int kMaxSegs = SB_TIMEBASE_MAX_SEGS;

// ------------------------------------------------------------------ time base (device)
// Segment cursor: finds the segment that holds n and walks forward as n grows.
struct TimeCursor {
    const sb_time_seg *segs;
    int nsegs;
    int idx;
    int64_t next_start;
    int64_t n_start;
    double t_start, step;

    __device__ __forceinline__ void load(int i)
    {
        idx = i;
        n_start = segs[i].n_start;
        t_start = segs[i].t_start;
        step = segs[i].step;
        next_start = (i + 1 < nsegs) ? segs[i + 1].n_start : INT64_MAX;
    }
    __device__ __forceinline__ void seek(const sb_time_seg *s, int ns, int64_t n)
    {
        segs = s;
        nsegs = ns;
        int lo = 0, hi = ns - 1;            // last segment with n_start <= n
        while (lo < hi) {
            int mid = (lo + hi + 1) >> 1;
            if (s[mid].n_start <= n) lo = mid; else hi = mid - 1;
        }
        load(lo);
    }
    // exact value of the serially accumulated t after n additions; n must not decrease
    __device__ __forceinline__ double at(int64_t n)
    {
        while (n >= next_start) load(idx + 1);
        return __dadd_rn(t_start, __dmul_rn((double)(n - n_start), step));
    }
    __device__ __forceinline__ double step_at(int64_t n)
    {
        while (n >= next_start) load(idx + 1);
        return step;
    }
};

// ------------------------------------------------------------------ waveforms (device)
// All phase arithmetic that decides which side of a discontinuity a sample falls on is done
// with explicitly rounded fp64 operations (no FMA contraction) in the order upstream's
// Python expressions evaluate, so the decisions are bit-identical to CPython's.

constexpr double kTwoPi = 6.283185307179586476925286766559;      // == 2*math.pi in fp64
constexpr double kInvTwoPi = 0.15915494309189533576888376337251;
// 2*pi split for Cody-Waite reduction: kTwoPiHi carries 29 significant bits, so k*kTwoPiHi is
// exact for |k| < 2^24 turns (20 kHz for 10 minutes is 1.2e7) and kTwoPiLo = 2*pi - kTwoPiHi.
constexpr double kTwoPiHi = 6.283185303211212;        // 0x1.921fb54000000p+2
constexpr double kTwoPiLo = 3.968374318722162e-09;

__device__ __forceinline__ float sin_of_radians(double x)
{
    // reduce to [-pi, pi] in fp64, evaluate in fp32
    double k = rint(x * kInvTwoPi);
    double r = fma(-k, kTwoPiHi, x);
    r = fma(-k, kTwoPiLo, r);
    return sinf((float)r);
}

__device__ __forceinline__ double py_mod1(double x)
{
    // Python's x % 1.0 : result in [0, 1)
    double m = x - trunc(x);          // == fmod(x, 1.0): exact
    if (m < 0.0) m = __dadd_rn(m, 1.0);
    return m;
}

// value of a waveform at fp64 phase ph (radians for sine, cycles otherwise), before
// amplitude/bias: returns the unit-amplitude shape in [-1, 1] as float.
__device__ __forceinline__ float wave_shape(int wave, double ph, double pw)
{
    switch (wave) {
    case SB_WAVE_SINE:
        return sin_of_radians(ph);
    case SB_WAVE_TRIANGLE: {
        double f = py_mod1(__dadd_rn(ph, 0.75));
        return (float)(4.0 * (fabs(f - 0.5) - 0.25));
    }
    case SB_WAVE_SQUARE: {
        double x2 = __dmul_rn(ph, 2.0);
        long long i = (long long)x2;               // int(): toward zero
        return (i & 1) ? -1.0f : 1.0f;             // Python's i % 2 is the low bit for either sign
    }
    case SB_WAVE_SAWTOOTH: {
        double f = floor(__dadd_rn(0.5, ph));
        return (float)(2.0 * (ph - f));
    }
    case SB_WAVE_PULSE:
        return (py_mod1(ph) < pw) ? 1.0f : -1.0f;
    case SB_WAVE_SEMICIRCLE: {
        // half circle over each cycle: ft = ph - floor(ph) - 0.5, sqrt(0.25 - ft*ft) * (2*amplitude)
        const double ft = __dadd_rn(__dadd_rn(ph, -floor(ph)), -0.5);
        return (float)(2.0 * sqrt(__dadd_rn(0.25, -__dmul_rn(ft, ft))));
    }
    case SB_WAVE_POINTY: {
        // 'inverted cosine': vv = 1 - |cos(ph)|, value = copysign(vv*vv, sin(ph)); ph in radians
        double k = rint(ph * kInvTwoPi);
        double r = fma(-k, kTwoPiHi, ph);
        r = fma(-k, kTwoPiLo, r);
        const double vv = 1.0 - fabs(cos(r));
        return (float)copysign(vv * vv, sin(r));
    }
    }
    return 0.0f;
}

// The same shapes in fp64 (for an oscillator rendered as somebody's FM modulator: the running sum
// of its values over minutes keeps float32 rounding as a systematic phase error of the carrier).
__device__ __forceinline__ double wave_shape64(int wave, double ph, double pw)
{
    switch (wave) {
    case SB_WAVE_SINE: {
        double k = rint(ph * kInvTwoPi);
        double r = fma(-k, kTwoPiHi, ph);
        r = fma(-k, kTwoPiLo, r);
        return sin(r);
    }
    case SB_WAVE_TRIANGLE: {
        double f = py_mod1(__dadd_rn(ph, 0.75));
        return 4.0 * (fabs(f - 0.5) - 0.25);
    }
    case SB_WAVE_SQUARE: {
        double x2 = __dmul_rn(ph, 2.0);
        long long i = (long long)x2;
        return (i & 1) ? -1.0 : 1.0;
    }
    case SB_WAVE_SAWTOOTH:
        return 2.0 * (ph - floor(__dadd_rn(0.5, ph)));
    case SB_WAVE_PULSE:
        return (py_mod1(ph) < pw) ? 1.0 : -1.0;
    case SB_WAVE_SEMICIRCLE: {
        const double ft = __dadd_rn(__dadd_rn(ph, -floor(ph)), -0.5);
        return 2.0 * sqrt(__dadd_rn(0.25, -__dmul_rn(ft, ft)));
    }
    case SB_WAVE_POINTY: {
        double k = rint(ph * kInvTwoPi);
        double r = fma(-k, kTwoPiHi, ph);
        r = fma(-k, kTwoPiLo, r);
        const double vv = 1.0 - fabs(cos(r));
        return copysign(vv * vv, sin(r));
    }
    }
    return 0.0;
}

__device__ __forceinline__ double clamp_pulsewidth(double pw)
{
    const double eps = 2.220446049250313e-16;
    if (pw <= 0.0) return eps;
    if (pw >= 1.0) return 1.0 - eps;
    return pw;
}

// phase offset upstream starts from: radians for Sine, cycles otherwise
__host__ __device__ __forceinline__ double phase0_of(int wave, double phase)
{
#ifdef __CUDA_ARCH__
    return (wave == SB_WAVE_SINE || wave == SB_WAVE_POINTY) ? __dmul_rn(__dmul_rn(phase, 2.0), 3.141592653589793) : phase;
#else
    return (wave == SB_WAVE_SINE || wave == SB_WAVE_POINTY) ? phase * 2 * 3.141592653589793 : phase;
#endif
}

// ADSR factor at sample index n (EnvelopeFilter); 0 after the release ends.
__device__ __forceinline__ float envelope_at(const sb_envelope &e, int64_t n)
{
    if (!e.enabled) return 1.0f;
    if (n < e.n_a) return (float)((double)n * e.a_slope);
    if (n < e.n_d) return (float)(1.0 + (double)(n - e.n_a) * e.d_slope);
    if (n < e.n_s) return (float)e.sustain_level;
    if (n < e.n_r) return (float)(e.sustain_level + (double)(n - e.n_s) * e.r_slope);
    if (n == e.n_r && e.extra) return (float)e.amp_extra;
    return 0.0f;
}

// ... in fp64 (an enveloped oscillator as an FM modulator)
__device__ __forceinline__ double envelope_at64(const sb_envelope &e, int64_t n)
{
    if (!e.enabled) return 1.0;
    if (n < e.n_a) return (double)n * e.a_slope;
    if (n < e.n_d) return 1.0 + (double)(n - e.n_a) * e.d_slope;
    if (n < e.n_s) return e.sustain_level;
    if (n < e.n_r) return e.sustain_level + (double)(n - e.n_s) * e.r_slope;
    if (n == e.n_r && e.extra) return e.amp_extra;
    return 0.0;
}

// int <-> fp64 without conversion instructions (I2F / F2I / FRND run on the 16-lane XU pipe and
// bound the integer PCM kernels; these are one integer and one fp64-pipe operation each):
// exact for |x| < 2^31, and for |v| < 2^31 respectively.
__device__ __forceinline__ double i32_to_f64(int x)
{
    // bits of 2^52 + 2^31 + x, minus that offset
    return __dadd_rn(__hiloint2double(0x43300000, (int)((unsigned)x ^ 0x80000000u)), -4503601774854144.0);
}
__device__ __forceinline__ int f64_floor_to_i32(double v)
{
    // v + 1.5*2^52 rounded toward -inf: the low word holds floor(v) in two's complement
    return __double2loint(__dadd_rd(v, 6755399441055744.0));
}
__device__ __forceinline__ int f64_trunc_to_i32(double v)
{
    const int m = __double2loint(__dadd_rd(fabs(v), 6755399441055744.0));      // floor(|v|)
    return (__double2hiint(v) < 0) ? -m : m;
}

// IEEE a/b for a divisor known in advance: rcp_b = RN(1/b) once, then a product and one fused
// correction step per quotient instead of the ~40-instruction fp64 divide.  By Markstein's
// theorem the result is the correctly rounded quotient for normal a/b (b's significand not all
// ones, which an integer sample count never is); tools/fma_division_check.py replays it in
// exact arithmetic.
__device__ __forceinline__ double div_by_known(double a, double b, double rcp_b)
{
    const double q = __dmul_rn(a, rcp_b);
    return __fma_rn(__fma_rn(-b, q, a), rcp_b, q);
}

inline int grid_for(int64_t work_items, int block, int max_blocks)
{
    int64_t b = (work_items + block - 1) / block;
    if (b < 1) b = 1;
    if (b > max_blocks) b = max_blocks;
    return (int)b;
}

}  // namespace sb
