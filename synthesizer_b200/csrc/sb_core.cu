// Host-side core of libsynth_b200: error state, launch counter, the exact time base and the
// ADSR boundary computation.  No device code here.
#include <stdarg.h>
#include <string.h>
#include <math.h>
#include <atomic>
#include <cmath>
#include "sb_common.cuh"

namespace sb {

static thread_local char g_error[512] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

}  // namespace sb

extern "C" int sb_abi_version(void) { return 2; }
extern "C" const char *sb_last_error(void) { return sb::g_error; }
extern "C" int64_t sb_launch_count(void) { return sb::g_launches.load(); }

// The two additions below must be real IEEE double additions in program order: the table is
// built by observing them.  `volatile` keeps the compiler from folding or reassociating.
static inline double fp_add(double a, double b)
{
    volatile double r = a + b;
    return r;
}

extern "C" int sb_timebase_build(double increment, int64_t max_n, sb_time_seg *segs, int max_segs)
{
    return sb_timebase_build_from(0.0, increment, max_n, segs, max_segs);
}

extern "C" int sb_timebase_build_from(double t0, double increment, int64_t max_n, sb_time_seg *segs, int max_segs)
{
    if (!(increment > 0.0) || !isfinite(increment) || !(t0 >= 0.0) || !isfinite(t0) || max_n < 0 || !segs ||
        max_segs < 4) {
        sb::set_error("sb_timebase_build: bad arguments");
        return SB_EINVAL;
    }
    int ns = 0;
    auto emit = [&](int64_t n, double t, double step) -> bool {
        if (ns > 0) {
            const sb_time_seg &p = segs[ns - 1];
            // same line as the previous segment: merge
            if (p.step == step && p.t_start + (double)(n - p.n_start) * p.step == t) return true;
        }
        if (ns >= max_segs) return false;
        segs[ns].n_start = n;
        segs[ns].t_start = t;
        segs[ns].step = step;
        ns++;
        return true;
    };
    int64_t n = 0;
    double t = t0;
    while (n <= max_n) {
        double t1 = fp_add(t, increment);
        double s1 = t1 - t;
        double t2 = fp_add(t1, increment);
        double s2 = t2 - t1;
        if (s1 == 0.0) {                         // stagnation: t no longer advances
            if (!emit(n, t, 0.0)) goto overflow;
            break;
        }
        if (s1 != s2) {                          // step still changing: one-sample segment
            if (!emit(n, t, s1)) goto overflow;
            n += 1;
            t = t1;
            continue;
        }
        {
            // constant step s1 while the sums stay inside the binade of t1
            int e;
            frexp(t1, &e);                        // t1 in [2^(e-1), 2^e)
            double limit = ldexp(1.0, e);
            double room = limit - t1;             // exact
            double kf = floor(room / s1);
            int64_t K = (kf > 4.0e18) ? (int64_t)4e18 : (int64_t)kf;
            K -= 2;                               // stay clear of the binade edge
            if (K < 0) K = 0;
            while (K > 0 && t1 + (double)K * s1 >= limit) K--;
            if (!emit(n, t, s1)) goto overflow;
            n += 1 + K;
            t = t + (double)(1 + K) * s1;          // exact: multiple of ulp inside the binade
        }
    }
    return ns;
overflow:
    sb::set_error("sb_timebase_build: more than %d segments", max_segs);
    return SB_ERANGE;
}

static int seg_index(const sb_time_seg *segs, int nsegs, int64_t n)
{
    int lo = 0, hi = nsegs - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (segs[mid].n_start <= n) lo = mid; else hi = mid - 1;
    }
    return lo;
}

extern "C" double sb_timebase_at(const sb_time_seg *segs, int nsegs, int64_t n)
{
    int i = seg_index(segs, nsegs, n);
    return segs[i].t_start + (double)(n - segs[i].n_start) * segs[i].step;
}

extern "C" int64_t sb_timebase_first_ge(const sb_time_seg *segs, int nsegs, int64_t n_from, double limit)
{
    // t(n) is non-decreasing: binary search on n over a bracket found by doubling
    if (sb_timebase_at(segs, nsegs, n_from) >= limit) return n_from;
    int64_t lo = n_from, hi = n_from + 1;
    const int64_t cap = (int64_t)1 << 60;
    while (sb_timebase_at(segs, nsegs, hi) < limit) {
        lo = hi;
        hi = (hi - n_from) * 2 + n_from + 1;
        if (hi > cap) return cap;                 // never reached (stagnated time base)
    }
    while (hi - lo > 1) {                          // invariant: t(lo) < limit <= t(hi)
        int64_t mid = lo + (hi - lo) / 2;
        if (sb_timebase_at(segs, nsegs, mid) >= limit) hi = mid; else lo = mid;
    }
    return hi;
}

// x after `count` serial fp64 additions of c (x = x + c, rounded each time), without doing them
// one by one.  While x stays inside one binade every addition rounds on the same grid, so the
// step actually taken, RN(x + c) - x, is the same each time (a tie of round-half-even settles
// after one step): two real additions establish it, then the remaining steps inside the
// binade are one exact multiply-add; a step that leaves the binade (or touches zero) is taken
// for real.  A 10 s release at 48 kHz is ~50 iterations instead of 480 000 additions.
static double serial_add_repeat(double x, double c, int64_t count)
{
    while (count > 0) {
        volatile double a1v = x + c;
        const double a1 = a1v;
        count -= 1;
        if (count == 0 || a1 == 0.0 || x == 0.0 || !std::isfinite(a1) || std::ilogb(a1) != std::ilogb(x) ||
            std::signbit(a1) != std::signbit(x)) {
            x = a1;
            continue;
        }
        volatile double a2v = a1 + c;
        const double a2 = a2v;
        const double s1 = a1 - x, s2 = a2 - a1;          // exact: neighbours in one binade
        if (s1 != s2 || s2 == 0.0 || a2 == 0.0 || std::ilogb(a2) != std::ilogb(a1) ||
            std::signbit(a2) != std::signbit(a1)) {
            x = a1;                                       // not settled, or about to leave: single steps
            if (s2 == 0.0 && s1 == 0.0) return x;        // c vanishes against x: nothing changes any more
            continue;
        }
        // from a1, how many more steps of s2 stay inside the binade [2^e, 2^(e+1)) of |a1|
        const int e = std::ilogb(a1);
        const double lo = std::ldexp(1.0, e), mag = std::fabs(a1), step = std::fabs(s2);
        const bool away = (std::signbit(s2) == std::signbit(a1));          // magnitude grows
        double room = away ? (std::ldexp(1.0, e + 1) - mag) : (mag - lo);  // exact
        int64_t k = (int64_t)std::floor(room / step);
        if (away && (double)k * step >= room && k > 0) k -= 1;             // stay strictly below 2^(e+1)
        if (k > count) k = count;
        if (k > 1) k -= 1;                                                  // (one step of slack for the quotient's rounding)
        else k = 0;
        x = a1 + (double)k * s2;                                            // exact: all values on the binade's grid
        count -= k;
    }
    return x;
}

extern "C" int sb_envelope_build(double samplerate, double attack, double decay, double sustain,
                                 double sustain_level, double release, sb_envelope *out)
{
    if (!out || !(samplerate > 1) || attack < 0 || decay < 0 || sustain < 0 || release < 0 ||
        sustain_level < 0 || sustain_level > 1) {
        sb::set_error("sb_envelope_build: bad arguments");
        return SB_EINVAL;
    }
    double increment = 1.0 / samplerate;
    double end_time_decay = attack + decay;
    double end_time_sustain = end_time_decay + sustain;
    double end_time_release = end_time_sustain + release;
    // time base long enough for the whole envelope
    int64_t need = (int64_t)((end_time_release + 1.0) * samplerate) + 16;
    sb_time_seg segs[SB_TIMEBASE_MAX_SEGS];
    int ns = sb_timebase_build(increment, need, segs, SB_TIMEBASE_MAX_SEGS);
    if (ns < 0) return ns;
    memset(out, 0, sizeof(*out));
    out->enabled = 1;
    out->sustain_level = sustain_level;
    int64_t n = 0;
    if (attack != 0.0) {
        out->a_slope = 1.0 / attack * increment;
        n = sb_timebase_first_ge(segs, ns, n, attack);
    }
    out->n_a = n;
    if (decay != 0.0) {
        out->d_slope = (sustain_level - 1.0) / decay * increment;
        n = sb_timebase_first_ge(segs, ns, n, end_time_decay);
    }
    out->n_d = n;
    n = sb_timebase_first_ge(segs, ns, n, end_time_sustain);
    out->n_s = n;
    if (release != 0.0) {
        out->r_slope = (-sustain_level) / release * increment;
        n = sb_timebase_first_ge(segs, ns, n, end_time_release);
        // the serially accumulated release amplitude decides whether one more sample follows
        const double amp = serial_add_repeat(sustain_level, out->r_slope, n - out->n_s);
        if (amp > 0.0) {
            out->extra = 1;
            out->amp_extra = amp;
        }
    }
    out->n_r = n;
    return SB_OK;
}

// CUDA events for callers that have no CUDA runtime binding of their own (bench.py times the
// dominant kernel with them through sb_bank_profile_events)
extern "C" int sb_event_create(void **event)
{
    SB_REQUIRE(event, "sb_event_create: null pointer");
    cudaEvent_t e;
    SB_CUDA_CHECK(cudaEventCreate(&e));
    *event = (void *)e;
    return SB_OK;
}

extern "C" int sb_event_elapsed_ms(void *start_event, void *stop_event, float *ms)
{
    SB_REQUIRE(start_event && stop_event && ms, "sb_event_elapsed_ms: null pointer");
    SB_CUDA_CHECK(cudaEventElapsedTime(ms, (cudaEvent_t)start_event, (cudaEvent_t)stop_event));
    return SB_OK;
}

extern "C" int sb_event_destroy(void *event)
{
    if (event) SB_CUDA_CHECK(cudaEventDestroy((cudaEvent_t)event));
    return SB_OK;
}
