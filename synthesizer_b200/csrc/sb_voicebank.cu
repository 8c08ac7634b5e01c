// The fused hot path: a bank of FM voices with ADSR, mixed to stereo.
//
// Decomposition: work unit = (group of voices) x (span of consecutive frames), one WARP per
// unit; warps never synchronise with each other.  A warp walks its span in sub-tiles of 512
// frames (16 consecutive frames per lane) and, for every voice of its group, evaluates
// modulator -> FM prefix sum -> carrier -> ADSR -> pan, accumulating the voice sum in
// registers; coalesced 128-bit stores write the group's partial mix, and a fold kernel sums
// the groups.
//
// Voice classes (voice_class / sb_voice_is_fast) share that decomposition:
//
//  FAST (1) any of the five carriers with no FM or a Sine modulator of moderate gain.
//          Everything that needs fp64 is done once per lane per 16-frame run: the exact time
//          base, the carrier's phase base reduced mod one turn, and the FM sum ahead of the
//          run, which has a closed form for a Sine modulator (sum of sin over an arithmetic
//          progression).  That closed form needs cos/sin of the modulator's phase at the
//          run's first frame; they come from a per-(voice, sub-tile) fp64 anchor rotated by
//          a per-(voice, lane) fp64 table, four fp64 multiply-adds instead of a range
//          reduction and a polynomial.  The 16 samples of the run then advance in fp32: the
//          modulator by the two-term recurrence s += d, d -= k*s (no SFU work), the carrier
//          by the SFU sine, in packed fp32x2 over frame pairs (q, q+8).  No pre-pass over
//          the samples, no running carry, and fp32 error cannot build up over more than 16
//          frames.  Square / Sawtooth / Pulse carriers (bank_fast_kernel<1>) walk an integer
//          phase and re-evaluate the few samples within a guard band of an edge exactly in fp64.
//          Sine / Triangle carriers under strong FM (gain above kFastMaxFmGain, up to
//          kFastHighFmGain) take the same run with the phase reduced to one turn after every step.
//
//  FAST (2) the same carriers with a Triangle / Square / Sawtooth / Pulse modulator
//          (bank_fast_kernel<2>).  Between its edges and kinks such a modulator is a straight line:
//          a run of 16 frames that stays clear of them (tested on an integer phase, with a guard
//          band) needs one exactly formed phase - increments, half-run sum and the running sum
//          ahead of the lane are closed forms; a sub-tile that is one line end to end (found by the
//          span's set-up) evaluates nothing of the modulator in its runs.  A run with one edge
//          takes two lines around a settled break (lfo_run_break); only runs with more evaluate
//          the 16 values exactly as upstream does (lfo_run_slow); then the lanes' totals are
//          scanned in fp64.  The FM sum
//          ahead of a sub-tile is a per-512-frame tile prefix from a pre-pass (bank_lfo_tile_kernel:
//          a tile as one line, or two around a single edge settled by exact values, else by runs;
//          bank_lfo_prefix_kernel).
//
//  GENERIC (0) anything else (FM gain above kFastHighFmGain - kFastMaxFmGain for edged carriers -,
//          carriers or modulators with a negative start phase or frequency).  Phase arithmetic per sample in explicitly
//          rounded fp64, FM prefix by fp64 warp scan with a running carry that starts from
//          per-(voice, span) sums computed by a pre-pass.
//
// Order of kernels in a render: bank_prepare_kernel (per-voice constants and tables), for banks
// of more than one full group bank_key_hist / scan / scatter (the voices sorted by kind across
// the bank, so that a group - a slice of 64 of that order - holds one or two kinds),
// bank_sort_kernel (each group's voices by kernel part and kind), the tile pre-pass if any
// class-2 voice, then bank_fast_kernel<0>, <1>, <2> and the generic kernels as far as the bank has
// such voices (per group, the lowest launched part with voices in it writes the partial mix and the
// others add into it), bank_fold_kernel.  <0> maps its 2-D grid to (block of spans, group); the CTAs
// of <1> and <2> draw their (group, block of spans) unit from a queue per kind, chosen by the SM
// they run on, so that an SM runs one kind's loop at a time.
#include "sb_common.cuh"

namespace sb {

constexpr int kItems = 8;
constexpr int kSub = 32 * kItems;            // frames per warp step (generic kernel; plan granularity)
#ifndef FAST_ITEMS
#define FAST_ITEMS 16          /* measured on B200: 8 frames per lane 2.95 ms, 16 frames per lane 2.63 ms */
#endif
constexpr int kFItems = FAST_ITEMS;          // fast kernels: consecutive frames per lane
constexpr int kFSub = 32 * kFItems;          // fast kernels: frames per warp step (a multiple of kSub)
static_assert(kFSub % kSub == 0, "the fast sub-tile must be a multiple of the plan granularity");
constexpr int kMaxVoicesPerGroup = 64;

constexpr int kGenWarps = 8;                 // generic kernel: warps per CTA
constexpr int kFastWarps = 4;                // fast kernel: warps per CTA
#ifndef FAST_MAX_SUB
#define FAST_MAX_SUB 4
#endif
constexpr int kFastMaxSub = FAST_MAX_SUB;    // fast kernel: sub-tiles per span (anchors in smem)
#ifndef FAST_MIN_BLOCKS
#define FAST_MIN_BLOCKS 5      /* CTAs per SM (register cap 96; 39 KB of shared memory each).  Measured on B200:
                                  4/5 CTAs -> 2.20/2.19 ms; 6 needs 80 registers (spills) and a shorter span: 2.56 ms */
#endif
#ifndef FAST_MAX_FM_GAIN
#define FAST_MAX_FM_GAIN 1.5
#endif
#ifndef FAST_PAIR_VOICES
#define FAST_PAIR_VOICES 1     /* measured on B200: 2.42 ms without, 2.19 ms with; three per iteration 2.22 ms (4 CTAs/SM,
                                  126 registers) or 2.37 ms (5 CTAs/SM, spills); separate code per (flat, ramp) mix
                                  of a pair 2.29 ms (instruction cache) */
#endif

enum : int {
    FV_FAST = 1,
    FV_HAS_MOD = 2,
    FV_TRIANGLE = 4,
    FV_ENV = 8,
    FV_ENV_EXTRA = 16,
    FV_BIAS = 32,
    FV_SCAN = 64,            // Triangle/Square/Sawtooth/Pulse modulator: FM sum from the tile pre-pass + warp scan
    FV_HIGH = 128,           // FM gain above kFastMaxFmGain (Sine / Triangle carrier): the run's phase is
                             // brought back into one turn after every step
    FV_WAVE_SHIFT = 8,       // bits 8..10: the carrier's waveform id
    FV_MWAVE_SHIFT = 12,     // bits 12..14: the modulator's waveform id (FV_SCAN voices)
};
constexpr double kFastMaxFmGain = FAST_MAX_FM_GAIN;
// ... and up to this gain for Sine / Triangle carriers (any modulator), whose fp32 run then
// reduces its phase after every step (what limits the plain run is the resolution of an fp32 phase
// that has grown to 16 steps' worth of gain: measured at full scale, tools/accuracy_sweep.py and the
// sweep in test_gpu_voicebank.py: plain run 0.44 LSB at 1.5 rad, 0.92 at 6; reduced 0.13 and 0.39)
constexpr double kFastHighFmGain = 6.0;
constexpr int kFastKinds = 15;              // kind = MOD | (carrier waveform << 1) for no / Sine modulators,
                                             // 10 + carrier waveform for the other modulators
__host__ __device__ inline int fast_kind_of(int flags)
{
    const int wave = (flags >> FV_WAVE_SHIFT) & 7;
    return (flags & FV_SCAN) ? 10 + wave : (((flags & FV_HAS_MOD) ? 1 : 0) | (wave << 1));
}
// kinds 0..3 (Sine / Triangle carrier, no or Sine modulator) run in bank_fast_kernel<0>, 4..9
// (Square / Sawtooth / Pulse carrier) in <1>, 10..14 (piecewise-linear modulators) in <2>
__host__ __device__ inline int fast_kind_part(int kind) { return kind >= 10 ? 2 : (kind >= 4 ? 1 : 0); }

// Per-voice constants of the fast path, derived once per render by bank_prepare_kernel.
// The first six 128-bit words are everything the per-run set-up and the per-sample loop need
// (values used by packed fp32x2 instructions are stored as pairs); the rest is used by the
// per-span precompute and by the rare sub-tile that straddles an envelope corner.
// H = kFItems/2 frames is half a run; w the modulator's phase step per frame; "scaled" means
// multiplied by gd, the carrier phase per unit of the modulator's running sum.
struct alignas(16) FastVoice {
    double2 car;             // f0b, pc0: carrier frequency * (1 + modulator bias), start phase
    int flags;
    int tab_row;             // voice index * nsegs_rad: the voice's row of the closed-form FM table
    int lt_row;              // voice index * kLaneTab: the voice's row of the lane rotation table
    float Hwc;               // H carrier steps
    float4 stepk;            // wc, wc, -k, -k: carrier phase step; k = 4 sin^2(w/2)
    float4 gain;             // aL, aL, aR, aR: c_amp*gain
    float4 ini;              // k/2, k/2, sin w, sin w: first backward difference of the modulator recurrence
    float4 half;             // cos(H w), sin(H w), P, Q: second half-run start; sum of the first H sines = s0*P + c0*Q
    double2 mod;             // fm, pcm: modulator frequency and start phase (radians); fm is read per run
    double gd;               // carrier phase per unit of sum(sin(theta_k)): f0 * inc_c * depth
    double pw64;             // Pulse carriers: the (clamped) pulse width in fp64, for the exact re-evaluation
    long long n_a, n_d, n_s, n_r;
    float a_slope, d_slope, r_slope, sl;
    float amp_extra;
    float bL, bR;            // c_bias*gain
    float low_limit;         // edged carriers: a run whose first frame's unreduced phase (cycles) is not above
                             // this is evaluated exactly throughout (its phases might be negative)
};
static_assert(sizeof(FastVoice) == 192, "FastVoice layout");
// FV_SCAN voices have no use for `ini` and `half`; a Pulse modulator's (clamped) pulse width is
// kept there in fp64
__host__ __device__ inline double scan_m_pw64(const FastVoice &F) { return *reinterpret_cast<const double *>(&F.half); }
// ... then the size of the modulator's slope per frame at the nominal time step (Triangle 4w, Sawtooth 2w, else 0)
__host__ __device__ inline double scan_m_slope(const FastVoice &F) { return *reinterpret_cast<const double *>(&F.half.z); }
// ... and in ini: the pulse width and the phase advance over a run of kFItems frames as integers
// (2^32 = one cycle), then (two floats' room) the modulator's step per frame in those units as fp64
__host__ __device__ inline double scan_m_step32(const FastVoice &F) { return *reinterpret_cast<const double *>(&F.ini.z); }

// Closed-form FM table entry per (voice, modulator time-base segment b), step w_b = fm*step_b:
//   gd * sum_{k<n} sin(theta_k) = A - Cc*(gd cos theta_n) - (gd sin theta_n)/2      (n in segment b)
// with Cc = cot(w_b/2)/2 and A = gd*(sum over earlier segments + C_b cos(theta_b - w_b/2)),
// C_b = 1/(2 sin(w_b/2)).
struct alignas(16) FmSeg {
    double A, Cc;
};

// Per-(voice, sub-tile) records precomputed at the start of a span:
//   anchor = gd * (cos, sin)(theta) at the sub-tile's first frame, fp64;
//   env    = the ADSR line over the sub-tile (e(i) = e_sub + i*de), or e_sub = kEnvSilent / kEnvSlow.
constexpr float kEnvSilent = -1.0f;          // after the release: the voice adds nothing here
constexpr float kEnvSlow = -2.0f;            // an envelope corner falls inside: per-sample ADSR

constexpr int kLaneTab = 33;     // entries per voice of the lane rotation table: lanes 0..31, then one whole sub-tile

struct BankArgs {
    const sb_voice *voices;
    int nvoices, voices_per_group, ngroups, nspans;
    int64_t frames, span_len, n0;
    const sb_time_seg *tb_rad;
    int nsegs_rad;
    const sb_time_seg *tb_cyc;
    int nsegs_cyc;
    double samplerate;
    double *carry;        // [nvoices][nprefix + nspans] per-span LFO sums of generic voices (pre-pass), or null
    int nprefix;          // spans of span_len frames covering [0, n0): the FM sum before this render
    FastVoice *fastv;     // [nvoices]
    FmSeg *tab;           // [nvoices][nsegs_rad]: closed-form FM table per modulator segment
    double2 *lanetab;     // [nvoices][kLaneTab]: (cos, sin)(l * kFItems * w), l = 0..32: the rotation from a
                          // sub-tile's anchor to lane l's first frame; l = 32 is the next sub-tile's anchor
    double *lfo_prefix;   // [nvoices][lfo_rows]: FV_SCAN voices: sum of the modulator's shape over all frames
                          // before tile `row` (tiles of kFSub frames: lfo_nprefix over [0, n0), then the render's)
    int lfo_rows, lfo_nprefix;
    int *perm;            // [nvoices]: the voices sorted by kind across the whole bank (stable); group g is the
                          // slice perm[g*voices_per_group ...], so a group holds one or two kinds and its CTAs
                          // run one or two instantiations of the voice loop (a bank of mixed kinds otherwise
                          // walks all of the kernel's code in every CTA and falls out of the instruction cache)
    int *key_hist;        // [kSortKeys][ceil(nvoices / kSortBlock)] counts, then their exclusive scan
    int *part_kinds;      // [64]: [0..2] how many kinds of voices each kernel part holds (bank_key_scan_kernel; 0
                          // when the bank is not sorted), then the unit queues of parts 1 and 2, one per kind
                          // (kQ* below): a kind's groups, and a counter its CTAs draw (group, span block) units from
    int sm_count;
    int parts_mask;       // bit p: bank_fast_kernel<p> is launched in this render
    int *order;           // [3][nvoices]: per kernel part, each group's voices of that part sorted by kind
                          // (bank_sort_kernel); entry = voice index, the group's slice starts at its first voice
    int *kind_end;        // [ngroups][3][kFastKinds]: end position of each kind in the group's sorted slice
    int *scan_list;       // [2 + nvoices]: count of FV_SCAN voices, count of voices left to the generic kernels,
                          // then the indices of the FV_SCAN voices (any order)
    int scan_cap;         // rows of the pre-pass grid: how many list entries the launch covers
    int fast_nspans;
    int64_t fast_span_len;
    int first_part;       // the first of the three fast kernels that is launched (it writes, the others add)
    float *out;           // [ngroups][stride] partial mixes (or the mix itself if ngroups==1)
    int64_t out_stride;   // floats between groups: 2*frames rounded up to a multiple of 4
};

// ------------------------------------------------------------------ classification
// 0 = generic, 1 = fast (no FM, or a Sine modulator: closed-form FM sum), 2 = fast with a
// Triangle / Square / Sawtooth / Pulse modulator (FM sum from the tile pre-pass + a warp scan)
// (inv_sr = 1/samplerate and w1 = 2*pi/samplerate are passed in: a caller classifying a whole
// bank on the host - sb_host_bank_render_pcm16, inside its end-to-end time - pays the two
// divisions once; host and device evaluate the same expressions, so they always agree)
__host__ __device__ inline int voice_class(const sb_voice &V, double samplerate, double inv_sr, double w1)
{
    if (V.c_wave < SB_WAVE_SINE || V.c_wave > SB_WAVE_PULSE) return 0;
    const double inc_c = (V.c_wave == SB_WAVE_SINE ? w1 : inv_sr);
    const double turns = (V.c_wave == SB_WAVE_SINE ? 6.283185307179586 : 1.0);
    // carrier advance per 8-sample run must stay small enough for fp32 increments
    if (!(fabs(V.c_freq) * inc_c * 8.0 < 64.0 * turns)) return 0;
    if (V.c_wave >= SB_WAVE_SQUARE) {
        // discontinuous carriers: the fp32 run must resolve the phase to ~1e-6 cycles so that only
        // samples in a narrow guard band around an edge need the exact fp64 evaluation
        if (!(V.c_freq >= 0.0 && V.c_freq * 2.0 <= samplerate && V.c_phase >= 0.0)) return 0;
    }
    if (V.m_wave == SB_WAVE_NONE) return 1;
    if (V.m_wave < SB_WAVE_SINE || V.m_wave > SB_WAVE_PULSE) return 0;
    // (depth and bias only need to be sane numbers: what bounds the fp32 run is the FM gain below)
    if (!(fabs((double)V.m_amp) <= 64.0 && fabs((double)V.m_bias) <= 64.0)) return 0;
    // FM gain: carrier phase (radians) per unit of the modulator's running sum.  The fast path
    // anchors that sum exactly (fp64) at the first frame of every 16-frame run and adds fp32
    // values only inside the run; measured against the serial fp64 computation
    // (tools/accuracy_sweep.py) the error stays below half an LSB of full scale up to a gain
    // of 1.5 (it was 3.5 LSB per radian with an fp32 warp scan over the 512-frame sub-tile);
    // stronger FM takes the generic fp64 path.
    const double gain_rad = fabs(V.c_freq) * w1 * fabs((double)V.m_amp);
    if (!(gain_rad <= kFastMaxFmGain) &&
        !(gain_rad <= kFastHighFmGain && V.c_wave < SB_WAVE_SQUARE)) return 0;
    if (V.m_wave == SB_WAVE_SINE) {
        const double w = V.m_freq * w1;                              // modulator phase step
        if (!(fabs(w) > 1e-7 && fabs(w) < 3.0)) return 0;            // closed form needs sin(w/2) != 0
        return 1;
    }
    // piecewise-linear modulators: the periodic rules below hold for non-negative phases, and the
    // fp32 run must resolve the modulator phase like an edged carrier's
    if (!(V.m_freq >= 0.0 && V.m_freq * 2.0 <= samplerate && V.m_phase >= 0.0)) return 0;
    return 2;
}
__host__ __device__ inline int voice_class(const sb_voice &V, double samplerate)
{
    return voice_class(V, samplerate, 1.0 / samplerate, 6.283185307179586 / samplerate);
}

// cos(r) for |r| <= pi + small, fp64 Taylor to r^26 (error < 1e-13)
__device__ __forceinline__ double cos_reduced(double r)
{
    const double x = r * r;
    double p = -1.0 / 403291461126605635584000000.0;           // -1/26!
    p = fma(p, x, 1.0 / 620448401733239439360000.0);           // 1/24!
    p = fma(p, x, -1.0 / 1124000727777607680000.0);            // -1/22!
    p = fma(p, x, 1.0 / 2432902008176640000.0);                // 1/20!
    p = fma(p, x, -1.0 / 6402373705728000.0);                  // -1/18!
    p = fma(p, x, 1.0 / 20922789888000.0);                     // 1/16!
    p = fma(p, x, -1.0 / 87178291200.0);                       // -1/14!
    p = fma(p, x, 1.0 / 479001600.0);                          // 1/12!
    p = fma(p, x, -1.0 / 3628800.0);                           // -1/10!
    p = fma(p, x, 1.0 / 40320.0);                              // 1/8!
    p = fma(p, x, -1.0 / 720.0);                               // -1/6!
    p = fma(p, x, 1.0 / 24.0);                                 // 1/4!
    p = fma(p, x, -0.5);
    return fma(p, x, 1.0);
}

// sin(r) for |r| <= pi + small, fp64 Taylor to r^27 (error < 1e-13)
__device__ __forceinline__ double sin_reduced(double r)
{
    const double x = r * r;
    double p = -1.0 / 10888869450418352160768000000.0;         // -1/27!
    p = fma(p, x, 1.0 / 15511210043330985984000000.0);         // 1/25!
    p = fma(p, x, -1.0 / 25852016738884976640000.0);           // -1/23!
    p = fma(p, x, 1.0 / 51090942171709440000.0);               // 1/21!
    p = fma(p, x, -1.0 / 121645100408832000.0);                // -1/19!
    p = fma(p, x, 1.0 / 355687428096000.0);                    // 1/17!
    p = fma(p, x, -1.0 / 1307674368000.0);                     // -1/15!
    p = fma(p, x, 1.0 / 6227020800.0);                         // 1/13!
    p = fma(p, x, -1.0 / 39916800.0);                          // -1/11!
    p = fma(p, x, 1.0 / 362880.0);                             // 1/9!
    p = fma(p, x, -1.0 / 5040.0);                              // -1/7!
    p = fma(p, x, 1.0 / 120.0);                                // 1/5!
    p = fma(p, x, -1.0 / 6.0);
    return fma(p * x, r, r);
}

__device__ __forceinline__ double reduce_two_pi(double x)
{
    double k = rint(x * kInvTwoPi);
    double r = fma(-k, kTwoPiHi, x);
    return fma(-k, kTwoPiLo, r);
}

// ------------------------------------------------------------------ prepare (one warp per voice)
// Lane 0 derives the voice's constants; all lanes share the trigonometry of the FM tables
// (one time-base segment and one lane-table entry each).
__global__ void bank_prepare_kernel(BankArgs a)
{
    const int v = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const int lane = threadIdx.x & 31;
    if (v >= a.nvoices) return;                          // the whole warp leaves
    const sb_voice V = a.voices[v];
    constexpr int H = kFItems / 2;
    const int cls = voice_class(V, a.samplerate);
    const bool fast = cls != 0, scan = cls == 2;
    const bool tri = (V.c_wave != SB_WAVE_SINE);         // phase counted in cycles (all but Sine)
    const double inc_rad = 6.283185307179586 / a.samplerate;
    const double inc_c = tri ? 1.0 / a.samplerate : inc_rad;
    const bool has_mod = (V.m_wave != SB_WAVE_NONE);
    const double mbias = has_mod ? (double)V.m_bias : 0.0;
    const double fm = V.m_freq, pcm = scan ? V.m_phase : phase0_of(SB_WAVE_SINE, V.m_phase);
    const double f0b = V.c_freq * (1.0 + mbias);
    const double gd = has_mod ? V.c_freq * inc_c * (double)V.m_amp : 0.0;
    const double w = fm * inc_rad;                        // nominal modulator phase step
    if (lane == 0) {
        FastVoice F;
        F.flags = 0;
        F.tab_row = v * a.nsegs_rad;
        F.lt_row = v * kLaneTab;
        // the phase cannot fall by more than this over a run: kFItems FM increments of at most |gd| each
        // (unit-amplitude modulator shapes), and the carrier's own steps should its frequency be negative
        F.low_limit = (float)((double)kFItems * (fabs(gd) + fmax(0.0, -f0b * inc_c)) + 1.0 / 512.0);
        // ... and never falls at all where the FM index is below one (every frame's phase increment
        // f0b*inc + gd*value is positive: the phase stays at or above the start phase, itself >= 0)
        if (fabs(gd) <= 0.999 * f0b * inc_c) F.low_limit = -1.0f;
        F.pw64 = clamp_pulsewidth((double)V.c_pw);
        F.mod = make_double2(fm, pcm);
        F.car = make_double2(f0b, phase0_of(V.c_wave, V.c_phase));
        F.gd = gd;
        F.n_a = V.env.n_a; F.n_d = V.env.n_d; F.n_s = V.env.n_s; F.n_r = V.env.n_r;
        const float wc = (float)(f0b * inc_c);
        const double sh = sin(0.5 * w);
        const float k = (float)(4.0 * sh * sh);
        F.stepk = make_float4(wc, wc, -k, -k);
        if (V.c_wave >= SB_WAVE_SQUARE) {
            // edged carriers walk an integer phase (2^32 = one cycle): the step and the pulse
            // width as such integers, in the slots of the fp32 step
            const double cyc = f0b * inc_c;
            const double wfrac = cyc - floor(cyc);
            F.stepk.x = __uint_as_float((uint32_t)(unsigned long long)llrint(wfrac * 4294967296.0));
            // (a pulse width clamped to just below 1 rounds to 2^32: keep it in 32 bits)
            const unsigned long long pwll = (unsigned long long)llrint(clamp_pulsewidth((double)V.c_pw) * 4294967296.0);
            F.stepk.y = __uint_as_float(pwll > 0xffffffffull ? 0xffffffffu : (uint32_t)pwll);
        }
        F.Hwc = (float)((double)H * f0b * inc_c);
        const float hk = (float)(2.0 * sh * sh), sw = (float)sin(w);
        F.ini = make_float4(hk, hk, sw, sw);
        // sum_{j<H} sin(theta + j w) = sin(theta + (H-1) w/2) * sin(H w/2)/sin(w/2)
        const double R = (fabs(sh) > 1e-300) ? sin(0.5 * (double)H * w) / sh : (double)H;
        const double mid = 0.5 * (double)(H - 1) * w;
        F.half = make_float4((float)cos((double)H * w), (float)sin((double)H * w), (float)(cos(mid) * R), (float)(sin(mid) * R));
        bool demote = false;
        if (scan) {
            // slot in the list = row block of the modulator prefix table (kept in tab_row, unused
            // here).  A voice beyond the capacity the caller's scan_voices hint allowed for is
            // left to the generic kernels.
            const int slot = a.scan_cap > 0 ? atomicAdd(a.scan_list, 1) : 0;
            if (slot < a.scan_cap) {
                a.scan_list[2 + slot] = v;
                F.tab_row = slot;
                *reinterpret_cast<double *>(&F.half) = clamp_pulsewidth((double)V.m_pw);
                *reinterpret_cast<double *>(&F.half.z) =
                    (V.m_wave == SB_WAVE_TRIANGLE ? 4.0 : (V.m_wave == SB_WAVE_SAWTOOTH ? 2.0 : 0.0)) * (fm / a.samplerate);
                F.flags |= FV_SCAN | ((V.m_wave & 7) << FV_MWAVE_SHIFT);
                // the modulator on an integer phase (2^32 = one cycle), for the test whether a run
                // of kFItems frames passes an edge or a kink of it: the width of a pulse, and how
                // far the phase moves over the run (all ones: more than the test can tell)
                const double wm = fm / a.samplerate;                       // cycles per frame, in [0, 0.5]
                // (up to the next run's first frame: when every lane's run is clear, no edge lies
                // between two runs either, and the whole sub-tile is one straight line)
                const double reach = (double)kFItems * wm * 4294967296.0;
                const unsigned long long pwll = (unsigned long long)llrint(clamp_pulsewidth((double)V.m_pw) * 4294967296.0);
                F.ini.x = __uint_as_float(pwll > 0xffffffffull ? 0xffffffffu : (uint32_t)pwll);
                F.ini.y = __uint_as_float(reach < 3.0e9 ? (uint32_t)(unsigned long long)ceil(reach) : 0xffffffffu);
                *reinterpret_cast<double *>(&F.ini.z) = wm * 4294967296.0;
            } else {
                demote = true;
            }
        }
        F.bL = V.c_bias * V.gain_l;
        F.bR = V.c_bias * V.gain_r;
        F.gain = make_float4(V.c_amp * V.gain_l, V.c_amp * V.gain_l, V.c_amp * V.gain_r, V.c_amp * V.gain_r);
        F.a_slope = (float)V.env.a_slope; F.d_slope = (float)V.env.d_slope; F.r_slope = (float)V.env.r_slope;
        F.sl = (float)V.env.sustain_level; F.amp_extra = (float)V.env.amp_extra;
        if (fast && !demote) F.flags |= FV_FAST;
        else atomicAdd(a.scan_list + 1, 1);              // sb_bank_render_check() reads this
        if (has_mod) F.flags |= FV_HAS_MOD;
        if (tri) F.flags |= FV_TRIANGLE;
        F.flags |= (V.c_wave & 7) << FV_WAVE_SHIFT;
        if (V.env.enabled) F.flags |= FV_ENV;
        if (V.env.enabled && V.env.extra) F.flags |= FV_ENV_EXTRA;
        if (V.c_bias != 0.0f) F.flags |= FV_BIAS;
        if (has_mod && fabs(gd) * (tri ? 6.283185307179586 : 1.0) > kFastMaxFmGain) F.flags |= FV_HIGH;
        a.fastv[v] = F;
        a.perm[v] = v;                                   // (a bank of one group keeps this order)
    }
    if (!(cls == 1 && has_mod)) return;

    // Closed form of sum_{k<n} sin(theta_k), theta_k = fm*t(k) + pcm, t piecewise linear:
    // inside segment b (start n_b, t_b, step_b): theta_k = theta_b + (k-n_b)*w_b and
    //   sum_{k=n_b}^{n-1} sin(theta_k) = [cos(theta_b - w_b/2) - cos(theta_n - w_b/2)] / (2 sin(w_b/2)),
    // so with C_b = 1/(2 sin(w_b/2)) and S_b = (sum over earlier segments) + C_b cos(theta_b - w_b/2)
    // the sum up to n in segment b is S_b - C_b cos(theta_n - w_b/2)
    //                               = S_b - C_b cos(w_b/2) cos(theta_n) - sin(theta_n)/2.
    // A lane takes one segment; the sums over earlier segments come from a warp scan.
    FmSeg *row = a.tab + (int64_t)v * a.nsegs_rad;
    double carry = 0.0;
    for (int b0 = 0; b0 < a.nsegs_rad; b0 += 32) {
        const int b = b0 + lane;
        double whole = 0.0, C = 0.0, c_lo = 0.0, half = 0.0;
        if (b < a.nsegs_rad) {
            const sb_time_seg s = a.tb_rad[b];
            half = 0.5 * fm * s.step;
            const double sh = sin(half);
            C = (fabs(sh) > 1e-300) ? 0.5 / sh : 0.0;
            c_lo = cos(fm * (s.t_start - 0.5 * s.step) + pcm);
            if (b + 1 < a.nsegs_rad) {                    // this segment's whole sum, for the ones after it
                const double len = (double)(a.tb_rad[b + 1].n_start - s.n_start);
                const double th_hi = fm * (s.t_start + (len - 0.5) * s.step) + pcm;
                whole = (C != 0.0) ? C * (c_lo - cos(th_hi)) : len * sin(fm * s.t_start + pcm);
            }
        }
        double incl = whole;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const double o = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += o;
        }
        if (b < a.nsegs_rad) {
            FmSeg e;
            e.A = gd * (carry + (incl - whole) + C * c_lo);
            e.Cc = C * cos(half);
            row[b] = e;
        }
        carry += __shfl_sync(0xffffffffu, incl, 31);
    }
    // rotation from a sub-tile's first frame to each lane's first frame, at the nominal step
    const double ang = reduce_two_pi((double)(lane * kFItems) * w);
    a.lanetab[(int64_t)v * kLaneTab + lane] = make_double2(cos(ang), sin(ang));
    if (lane == 0) {
        const double ang32 = reduce_two_pi((double)(32 * kFItems) * w);
        a.lanetab[(int64_t)v * kLaneTab + 32] = make_double2(cos(ang32), sin(ang32));
    }
}

__device__ __forceinline__ float tri_shape(float cycles)
{
    // 4*(|((tt+0.75) % 1.0) - 0.5| - 0.25)
    float t = cycles + 0.75f;
    float f = t - floorf(t);
    return 4.0f * (fabsf(f - 0.5f) - 0.25f);
}

// ------------------------------------------------------------------ fast kernel
// The ADSR segment frame k lies in, as a line: the factor at k, its step per frame, and how many
// frames from k the next segment starts (a sub-tile that holds an envelope corner walks these lines,
// switching at the corner, instead of looking every sample up: a real call, made twice per run and
// once more in the one lane whose run holds the corner)
struct EnvLine {
    float v, dv;
    int left;
};
__device__ __noinline__ EnvLine fast_env_line(const FastVoice *Fp, int64_t k)
{
    const FastVoice &F = *Fp;
    EnvLine r;
    int64_t left = 0x7fffffff;
    r.dv = 0.0f;
    if (k < F.n_a) { r.v = (float)k * F.a_slope; r.dv = F.a_slope; left = F.n_a - k; }
    else if (k < F.n_d) { r.v = 1.0f + (float)(k - F.n_a) * F.d_slope; r.dv = F.d_slope; left = F.n_d - k; }
    else if (k < F.n_s) { r.v = F.sl; left = F.n_s - k; }
    else if (k < F.n_r) { r.v = F.sl + (float)(k - F.n_s) * F.r_slope; r.dv = F.r_slope; left = F.n_r - k; }
    else if (k == F.n_r && (F.flags & FV_ENV_EXTRA)) { r.v = F.amp_extra; left = 1; }
    else r.v = 0.0f;
    r.left = left > 0x7fffffff ? 0x7fffffff : (int)left;
    return r;
}

// x mod 2*pi into [-pi, pi] without the FRND (XU pipe) that rint() costs: magic-number rounding
__device__ __forceinline__ double reduce_two_pi_fast(double x)
{
    const double M = 6755399441055744.0;                       // 1.5 * 2^52
    const double k = __dadd_rn(__fma_rn(x, kInvTwoPi, M), -M);
    const double r = fma(-k, kTwoPiHi, x);
    return fma(-k, kTwoPiLo, r);
}

// Exact unit-amplitude value of a Triangle / Square / Sawtooth / Pulse modulator at time t
// (cycles time base), operation for operation what upstream's oscillator computes.  The phase
// is non-negative here (voice_class admits only such modulators), so Python's int() and % 1.0
// are floors, taken with one round-down add of 1.5*2^52 instead of a conversion.
__device__ __forceinline__ double floor_nonneg(double x)
{
    return __dadd_rn(__dadd_rd(x, 6755399441055744.0), -6755399441055744.0);
}
__device__ __forceinline__ double lfo_shape_exact(int mwave, double t, double fm, double phm, double pw)
{
    const double ph = __dadd_rn(__dmul_rn(t, fm), phm);
    switch (mwave) {
    case SB_WAVE_TRIANGLE: {
        const double x = __dadd_rn(ph, 0.75);
        return 4.0 * (fabs((x - floor_nonneg(x)) - 0.5) - 0.25);
    }
    case SB_WAVE_SQUARE:
        return (__double2loint(__dadd_rd(__dmul_rn(ph, 2.0), 6755399441055744.0)) & 1) ? -1.0 : 1.0;
    case SB_WAVE_SAWTOOTH:
        return 2.0 * (ph - floor_nonneg(__dadd_rn(0.5, ph)));
    default:
        return ((ph - floor_nonneg(ph)) < pw) ? 1.0 : -1.0;
    }
}

// exact serial time of frame k (cycles time base), as a real call for rare paths
__device__ __noinline__ double exact_time_at(const sb_time_seg *segs, int nsegs, int64_t k)
{
    TimeCursor c;
    c.seek(segs, nsegs, k);
    return c.at(k);
}

#ifdef SB_CTA_TRACE
// debugging build only (tools/cta_trace.py): start / end time, SM and unit of every warp of the fast
// kernels, and how many times it called the edged carriers' fix-up pass / took it for a whole run
constexpr int kTraceWarps = 1 << 16;
__device__ unsigned long long g_cta_trace[3][kTraceWarps][4];
__device__ unsigned int g_cta_fix[kTraceWarps][2];
#endif

struct FastCtx;

struct FastCtx {
    const sb_time_seg *s_cyc;
    int nsegs_cyc;
    double cstep;            // cycles time base: step of the segment holding this lane's first frame
    bool cstraddle;          // ... and whether the lane's run crosses into the next segment
    const FmSeg *tab_seg;    // closed-form FM table, offset to this lane's time-base segment: row stride nsegs_rad
    const double2 *lt_lane;  // anchor -> lane rotation table, offset to this lane: row stride 32
};

// The 16 values of a piecewise-linear modulator over a lane's run, exactly as upstream computes
// them (v[] = the FM increments gd*value in fp32, frame pairs (q, q+H); tot = the sum of the values),
// as a real call (one body, the waveform chosen at run time): only runs that pass an edge or a kink
// of the modulator come here, and inlined into every carrier's loop the four waveforms' copies made
// up most of the part-2 kernel's code.
struct LfoRun {
    float2 v[kFItems / 2];
    double tot;
};

__device__ __noinline__ LfoRun lfo_run_slow(const int mwave, const sb_time_seg *s_cyc, const int nsegs_cyc,
                                            const double cstep, const bool cstraddle, const double Tc0, const int64_t n,
                                            const double fm, const double phm, const double mpw, const double gd)
{
    // (a rolled loop on purpose: this is the uncommon path, and its code must stay small - a bank of
    // many kinds of voices otherwise keeps more code in flight than the instruction cache holds)
    constexpr int H = kFItems / 2;
    LfoRun r;
    r.tot = 0.0;
    if (mwave == SB_WAVE_TRIANGLE && !cstraddle) {
        // A triangle has no edges: its value does not hinge on the last bits of the phase, so only
        // the run's first phase is formed the way upstream forms it and the other 15 step from it
        // (phase fraction += step, wrapped), which halves the fp64 work per sample.  The difference
        // from upstream's own rounding of each phase is ~1e-16 of the value.
        const double x0 = __dadd_rn(__dadd_rn(__dmul_rn(Tc0, fm), phm), 0.75);
        double f = x0 - floor_nonneg(x0);                                   // in [0, 1)
        const double w = fm * cstep;                                        // cycles per frame (< 1)
#pragma unroll
        for (int j = 0; j < kFItems; j++) {
            const double v = fma(4.0, fabs(f - 0.5), -1.0);
            r.tot += v;
            const float inc = (float)(gd * v);
            if (j < H) r.v[j].x = inc; else r.v[j - H].y = inc;
            f += w;
            if (f >= 1.0) f -= 1.0;
        }
        return r;
    }
#pragma unroll 1
    for (int j = 0; j < kFItems; j++) {
        const double t = cstraddle ? exact_time_at(s_cyc, nsegs_cyc, n + j)
                                   : __dadd_rn(Tc0, __dmul_rn((double)j, cstep));   // exact inside a time-base segment
        const double v = lfo_shape_exact(mwave, t, fm, phm, mpw);
        r.tot += v;
        const float inc = (float)(gd * v);
        if (j < H) r.v[j].x = inc; else r.v[j - H].y = inc;
    }
    return r;
}

// The FM increments gd*v_j (fp32, frame pairs (q, q+H)) of a class-2 modulator over this lane's
// run, the sum of the first half-run's increments, and the sum of the v_j of the lanes before this
// one (the modulator's running sum from the sub-tile's first frame to the run's): the straight
// line if every lane's run is clear - then that sum has a closed form too and no scan is needed -
// else (the whole warp) the per-sample evaluation and an fp64 warp scan of the lanes' totals.
__device__ __forceinline__ void lfo_run_any(const FastVoice &F, const FastCtx &ctx, const double sub_line, const double Tc0,
                                            const int64_t n, float2 (&sv)[kFItems / 2], float &half_sum, double &before);

// ---- class-2 modulators: the run as a straight line wherever it is one
// A Triangle / Square / Sawtooth / Pulse modulator is linear (or constant) between its edges and
// kinks.  For a run of kFItems frames that stays clear of them by a guard band - at LFO rates
// nearly every run - the 16 values follow from the exactly formed phase of the run's first frame:
// v_j = v_0 + j*dv.  Whether the run is clear is tested on an integer phase (2^32 = one cycle);
// the guard (1e-6 cycles) covers what separates the straight line from upstream's own 16
// roundings, so a clear run has every sample on the side of every edge where upstream has it.
// Only runs that are not clear take the per-sample exact evaluation (lfo_run above).
constexpr uint32_t kGuardLfo = 4096u;

struct LfoLine {
    double v0, dv;       // value at the run's first frame and its step per frame (unit amplitude)
    bool clear;          // no edge or kink within the run (else the line only holds up to the first one)
    uint32_t d1, d2;     // phase distance (2^32 = one cycle) from the first frame to the next edge / kink of each kind
};

__device__ __forceinline__ LfoLine lfo_line(const int mwave, const double Tc0, const double fm, const double phm,
                                            const double cstep, const uint32_t pwi, const uint32_t reach)
{
    LfoLine L;
    const double ph = __dadd_rn(__dmul_rn(Tc0, fm), phm);                 // as upstream forms it; >= 0
    const double fr = ph - floor_nonneg(ph);                              // exact, in [0, 1)
    const uint32_t p = (uint32_t)__double2loint(__dadd_rd(fr * 4294967296.0, 6755399441055744.0));
    const double w = fm * cstep;                                          // cycles per frame
    // edges / kinks of the waveform on the integer phase
    uint32_t e1, e2;
    switch (mwave) {
    case SB_WAVE_TRIANGLE: e1 = 0x40000000u; e2 = 0xC0000000u; break;     // (ph + 0.75) % 1 crosses 0 and 0.5
    case SB_WAVE_SQUARE: e1 = 0u; e2 = 0x80000000u; break;
    case SB_WAVE_SAWTOOTH: e1 = e2 = 0x80000000u; break;                  // 0.5 + ph crosses an integer
    default: e1 = 0u; e2 = pwi; break;
    }
    // the run covers [p, p + reach]: clear if neither edge lies in [p - g, p + reach + g] (mod 2^32)
    const uint32_t lim = reach + 2u * kGuardLfo;
    L.d1 = e1 - p;
    L.d2 = e2 - p;
    L.clear = reach < 0xf0000000u && (L.d1 + kGuardLfo) > lim && (L.d2 + kGuardLfo) > lim;
    switch (mwave) {
    case SB_WAVE_TRIANGLE: {
        const double x0 = __dadd_rn(fr, 0.75);
        const double x = x0 >= 1.0 ? x0 - 1.0 : x0;
        L.v0 = fma(4.0, fabs(x - 0.5), -1.0);
        L.dv = x < 0.5 ? -4.0 * w : 4.0 * w;
        break;
    }
    case SB_WAVE_SQUARE:
        L.v0 = fr < 0.5 ? 1.0 : -1.0;
        L.dv = 0.0;
        break;
    case SB_WAVE_SAWTOOTH:
        L.v0 = fr < 0.5 ? 2.0 * fr : fma(2.0, fr, -2.0);
        L.dv = 2.0 * w;
        break;
    default:
        L.v0 = p < pwi ? 1.0 : -1.0;
        L.dv = 0.0;
        break;
    }
    return L;
}

// A run that is not clear usually holds exactly one edge (or kink): two straight lines.  Where the
// second one starts is estimated from the phase distance to the edge and settled by the exact values
// of the two frames around it (an edge makes the value jump by 2; a Triangle's kink needs no
// settling: the value is continuous there).  A real call: only the lanes whose run holds an edge
// come here, a few per cent of the runs at LFO rates.
struct LfoBreak {
    double v1, dv1;      // the second line: value at frame js and step per frame
    int js;              // first frame (1 .. kFItems-1) of the second line; 0: not one clean break
};
__device__ __noinline__ LfoBreak lfo_run_break(const FastVoice *Fp, const int mwave, const double Tc0, const double cstep,
                                               const double v0, const double dv, const uint32_t d1, const uint32_t d2)
{
    const FastVoice &F = *Fp;
    LfoBreak r;
    r.js = 0;
    r.v1 = r.dv1 = 0.0;
    const double step32 = scan_m_step32(F), fm = F.mod.x, phm = F.mod.y;
    const double reach = (double)(kFItems - 1) * step32;
    const uint32_t dn = min(d1 - 1u, d2 - 1u) + 1u;                   // nearest edge strictly ahead (0 = a whole cycle away)
    const uint32_t df = max(d1 - 1u, d2 - 1u) + 1u;
    if (!(dn != 0u && step32 > 0.0 && reach < 2.0e9 && (d1 == d2 || (double)df > reach + 3.0 * kGuardLfo + step32))) return r;
    int js = (int)ceil((double)dn / step32);                          // first frame at or past the edge
    js = max(1, min(kFItems - 1, js));
    bool ok = (mwave == SB_WAVE_TRIANGLE);
    const double mpw = scan_m_pw64(F);
#pragma unroll 1
    for (int tries = 0; tries < 4 && !ok; tries++) {
        const double ta = __dadd_rn(Tc0, __dmul_rn((double)(js - 1), cstep)), tb = __dadd_rn(Tc0, __dmul_rn((double)js, cstep));
        const bool before_a = fabs(lfo_shape_exact(mwave, ta, fm, phm, mpw) - fma((double)(js - 1), dv, v0)) < 0.5;
        const bool before_b = fabs(lfo_shape_exact(mwave, tb, fm, phm, mpw) - fma((double)js, dv, v0)) < 0.5;
        if (before_a && !before_b) ok = true;
        else if (!before_a) js--;                                      // the edge came earlier
        else js++;                                                     // not yet
        if (js < 1 || js > kFItems - 1) break;
    }
    if (!ok) return r;
    const int rest = kFItems - js;
    const double tj = __dadd_rn(Tc0, __dmul_rn((double)js, cstep));
    const LfoLine L2 = lfo_line(mwave, tj, fm, phm, cstep, __float_as_uint(F.ini.x),
                                (uint32_t)(long long)ceil((double)(rest - 1) * step32));
    if (!L2.clear) return r;
    r.js = js;
    r.v1 = L2.v0;
    r.dv1 = L2.dv;
    return r;
}

__device__ __forceinline__ void lfo_run_any(const FastVoice &F, const FastCtx &ctx, const double sub_line, const double Tc0,
                                            const int64_t n, float2 (&sv)[kFItems / 2], float &half_sum, double &before)
{
    constexpr int H = kFItems / 2;
    const int mwave = (F.flags >> FV_MWAVE_SHIFT) & 7;
    const double fm = F.mod.x, phm = F.mod.y, gd = F.gd;
    const int lane = threadIdx.x & 31;
    if (sub_line == sub_line) {
        // The whole sub-tile is one straight line (the span's set-up found it clear of the modulator's
        // edges and kinks - at LFO rates nearly every sub-tile is): sub_line is its value at the
        // sub-tile's first frame, plus 4 for a rising Triangle.  Nothing of the modulator is evaluated
        // per run: v_i = vs + i*dv, sum_{i<m} v_i = m*vs + dv*m(m-1)/2, m = this lane's first frame.
        const bool rising = sub_line > 2.0;
        const double vs = rising ? sub_line - 4.0 : sub_line;
        const double slope = scan_m_slope(F);
        const double dv = (mwave == SB_WAVE_TRIANGLE && !rising) ? -slope : slope;
        const double m = (double)(lane * kFItems);
        before = fma(0.5 * dv, m * (m - 1.0), m * vs);
        const float s0 = (float)(gd * fma(m, dv, vs)), ds = (float)(gd * dv);
        const float2 s2 = make_float2(s0, s0), d2 = make_float2(ds, ds);
#pragma unroll
        for (int q = 0; q < H; q++) sv[q] = __ffma2_rn(make_float2((float)q, (float)(q + H)), d2, s2);
        half_sum = fmaf((float)(H * (H - 1) / 2), ds, (float)H * s0);
        return;
    }
    const LfoLine L = lfo_line(mwave, Tc0, fm, phm, ctx.cstep, __float_as_uint(F.ini.x), __float_as_uint(F.ini.y));
    // A run that crosses into the next segment of the time base keeps the line too: the step changes
    // by at most half an ulp of the time, which over a sub-tile moves the phase by < 1e-7 cycles while
    // the time is below 128 s (the guard is 1e-6).  This matters at the start of a render, where the
    // time doubles - a new segment - every few frames: the first span's warps otherwise take the exact
    // path in half their sub-tiles (2.8 ms for one warp, against 0.34 ms for the others).
    const bool seam = ctx.cstraddle && !(Tc0 < 128.0);
    if (__all_sync(0xffffffffu, L.clear && !seam)) {
        // the whole sub-tile is one straight line v_i = v0 + (i - m)*dv, m = this lane's first frame:
        // sum_{i<m} v_i = m*v0 - dv*m(m+1)/2
        const double m = (double)(lane * kFItems);
        before = fma(-L.dv, 0.5 * m * (m + 1.0), m * L.v0);
        const float s0 = (float)(gd * L.v0), ds = (float)(gd * L.dv);
        const float2 s2 = make_float2(s0, s0), d2 = make_float2(ds, ds);
#pragma unroll
        for (int q = 0; q < H; q++) sv[q] = __ffma2_rn(make_float2((float)q, (float)(q + H)), d2, s2);
        half_sum = fmaf((float)(H * (H - 1) / 2), ds, (float)H * s0);
    } else {
        // some lane's run holds an edge: each lane its own line, or two lines with one break
        const bool mine = L.clear && !seam;
        LfoBreak bk;
        bk.js = kFItems;
        bk.v1 = bk.dv1 = 0.0;
        if (!mine) {
            bk.js = 0;
            if (!ctx.cstraddle) bk = lfo_run_break(&F, mwave, Tc0, ctx.cstep, L.v0, L.dv, L.d1, L.d2);
        }
        double tot;
        if (__all_sync(0xffffffffu, bk.js != 0)) {
            const int js = bk.js, rest = kFItems - js;
            const float s0 = (float)(gd * L.v0), ds = (float)(gd * L.dv);
            const float s1 = (float)(gd * bk.v1), ds1 = (float)(gd * bk.dv1);
            half_sum = 0.0f;
#pragma unroll
            for (int q = 0; q < H; q++) {
                sv[q].x = q < js ? fmaf((float)q, ds, s0) : fmaf((float)(q - js), ds1, s1);
                sv[q].y = q + H < js ? fmaf((float)(q + H), ds, s0) : fmaf((float)(q + H - js), ds1, s1);
                half_sum += sv[q].x;
            }
            tot = fma(0.5 * (double)js * (double)(js - 1), L.dv, (double)js * L.v0) +
                  fma(0.5 * (double)rest * (double)(rest - 1), bk.dv1, (double)rest * bk.v1);
        } else {
            const LfoRun r = lfo_run_slow(mwave, ctx.s_cyc, ctx.nsegs_cyc, ctx.cstep, ctx.cstraddle, Tc0, n, fm, phm,
                                          scan_m_pw64(F), gd);
            half_sum = 0.0f;
#pragma unroll
            for (int q = 0; q < H; q++) {
                sv[q] = r.v[q];
                half_sum += r.v[q].x;
            }
            tot = r.tot;
        }
        double incl = tot;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const double o = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += o;
        }
        before = incl - tot;
    }
}

// Exact value of a discontinuous waveform at absolute frame k, with the fp64 phase formed the
// way upstream forms it (no FM: fl(fl(t*f) + phase), bit for bit).  Called only for samples
// whose run phase is within the guard band of an edge, or while the phase may still be negative.
__device__ __noinline__ float exact_shape(int wave, FastCtx ctx, int64_t k, double f0b, double pc0,
                                          double fm_term, int has_mod, double pw)
{
    TimeCursor c;
    c.seek(ctx.s_cyc, ctx.nsegs_cyc, k);
    const double t = c.at(k);
    const double tt = has_mod ? __dadd_rn(pc0, __dadd_rn(__dmul_rn(f0b, t), fm_term))
                              : __dadd_rn(__dmul_rn(t, f0b), pc0);
    return wave_shape(wave, tt, pw);
}

// One voice over one lane's run of kFItems frames.  WAVE: the carrier (Sine or Triangle).  RARE: per-sample envelope and/or carrier bias (warp-uniform, uncommon).
//
// Frame pairing: the packed fp32x2 operations (Blackwell FFMA2/FADD2/FMUL2) work on frames
// (q, q+H) of the lane's run, H = kFItems/2.  accL2[q] = left-channel sums of frames (q, q+H),
// accR2[q] the right-channel ones.
//
// Per pair and step, with P the carrier phase (within a turn, fp32), s = gd*sin(theta) the FM
// increment and d its backward difference:
//     value = shape(P);   acc += value * e * gain;
//     P += wc + s;        d -= k*s;   s += d;        e += de
// The recurrence for s is sin(theta+w) = 2cos(w) sin(theta) - sin(theta-w) written with
// k = 2 - 2cos(w) = 4 sin^2(w/2) so that slow modulators lose nothing to cancellation.
//
// FLAT: the envelope is constant over the sub-tile (sustain, or no envelope): it is folded
// into the pan gains once per run, which takes two packed operations out of every step.
//
// Set-up (fp64, once per run): gd*(cos, sin)(theta) at the lane's first frame by rotating the
// sub-tile's anchor with the lane table (plus the first-order term for the time-base segment's
// own step: dstep = (frames from the span start) * (segment step - nominal step)); the FM sum ahead of the run from the closed form; the carrier phase base.
//
// (Square / Sawtooth / Pulse carriers: edged_voice_run below.)
// MODC: 0 = no modulator, 1 = Sine modulator (closed form + recurrence), 2 = Triangle / Square /
// Sawtooth / Pulse modulator: its 16 values are evaluated exactly (fp64, as upstream does) and
// the FM sum ahead of the run is the tile pre-pass's prefix plus an fp64 warp scan of the lanes'
// run totals.
template <int MODC, int WAVE, bool RARE, bool FLAT>
__device__ __forceinline__ void fast_voice_run(const FastVoice &F, const double2 anc, const float2 env,
                                               const double Tr0, const double Tc0, const double dstep,
                                               const float lane8, const int64_t n,
                                               const FastCtx ctx, float2 (&accL2)[kFItems / 2],
                                               float2 (&accR2)[kFItems / 2])
{
    constexpr int H = kFItems / 2;
    constexpr bool MOD = MODC != 0;
    constexpr bool CYC = (WAVE != SB_WAVE_SINE);            // phase counted in cycles
    static_assert(WAVE == SB_WAVE_SINE || WAVE == SB_WAVE_TRIANGLE, "edged carriers go through edged_voice_run");
    const double2 car = F.car;
    const float4 stepk = F.stepk;           // wc, wc, -k, -k
    const float4 gain = F.gain;             // aL, aL, aR, aR
    double B = fma(car.x, CYC ? Tc0 : Tr0, car.y);
    double gbefore = 0.0;                   // gd * (FM sum ahead of this lane's run)
    float2 sn = make_float2(0.0f, 0.0f), dl = make_float2(0.0f, 0.0f);
    float half_sum = 0.0f;                  // gd * (sum of the first half-run's sines)
    float2 sv[kFItems / 2];                 // MODC 2: the FM increments gd*lfo of frames (q, q+H)
    if (MODC == 2) {
        double before;
        lfo_run_any(F, ctx, anc.y, Tc0, n, sv, half_sum, before);
        gbefore = F.gd * (anc.x + before);                 // anc.x: the modulator's sum before this sub-tile (tile pre-pass)
        B += gbefore;
    }
    if (MODC == 1) {
        const double2 L = ctx.lt_lane[F.lt_row];
        const FmSeg fs = ctx.tab_seg[F.tab_row];
        // gd*(cos, sin) of the modulator phase at this lane's first frame
        const double cl = fma(-anc.y, L.y, anc.x * L.x);
        const double sl = fma(anc.x, L.y, anc.y * L.x);
        // closed form, with theta advanced to first order by the small angle fm*dstep: how far
        // this segment's time step has carried the lane from where the nominal step would have
        double u = fma(-fs.Cc, cl, fs.A);
        u = fma(-0.5, sl, u);
        const double vv = fma(fs.Cc, sl, -0.5 * cl);
        gbefore = fma(F.mod.x * dstep, vv, u);
        B += gbefore;
        const float s0 = (float)sl, c0 = (float)cl;
        const float4 hf = F.half;           // cos(H w), sin(H w), P, Q
        const float4 ini = F.ini;           // k/2, k/2, sin w, sin w
        const float sH = fmaf(c0, hf.y, s0 * hf.x);
        const float cH = fmaf(-s0, hf.y, c0 * hf.x);
        sn = make_float2(s0, sH);
        dl = __ffma2_rn(make_float2(c0, cH), make_float2(ini.z, ini.w), __fmul2_rn(sn, make_float2(ini.x, ini.y)));
        half_sum = fmaf(s0, hf.z, c0 * hf.w);
    }
    // carrier: phase base reduced to one turn in fp64, fp32 from there
    float b0;
    if (CYC) {
        const double M = 6755399441055744.0;
        b0 = (float)(B - __dadd_rn(__dadd_rn(B, M), -M));
    } else {
        b0 = (float)reduce_two_pi_fast(B);
    }
    float2 P = make_float2(b0, b0 + F.Hwc + half_sum);
    const float e0 = fmaf(lane8, env.y, env.x);
    float2 e = make_float2(e0, fmaf((float)H, env.y, e0));
    const float2 cstep = make_float2(stepk.x, stepk.y), nk2 = make_float2(stepk.z, stepk.w);
    float2 estep = make_float2(env.y, env.y);
    int swx = 0x7fffffff, swy = 0x7fffffff;                // RARE: the step at which frame q / q+H enters a new ADSR segment
    const bool high = RARE && (F.flags & FV_HIGH);          // warp-uniform
    if (RARE && env.x == kEnvSlow) {
        const EnvLine lx = fast_env_line(&F, n), ly = fast_env_line(&F, n + H);
        e = make_float2(lx.v, ly.v);
        estep = make_float2(lx.dv, ly.dv);
        swx = lx.left;
        swy = ly.left;
    }
    float2 aL2 = make_float2(gain.x, gain.y), aR2 = make_float2(gain.z, gain.w);
    if (FLAT) {
        const float2 ec = make_float2(env.x, env.x);
        aL2 = __fmul2_rn(aL2, ec);
        aR2 = __fmul2_rn(aR2, ec);
    }
#pragma unroll
    for (int q = 0; q < H; q++) {
        float2 w;
        if (WAVE == SB_WAVE_SINE) {
            w.x = __sinf(P.x);
            w.y = __sinf(P.y);
        } else {
            w.x = tri_shape(P.x);
            w.y = tri_shape(P.y);
        }
        const float2 ee = e;
        if (RARE) {
            accL2[q] = __ffma2_rn(ee, make_float2(F.bL, F.bL), accL2[q]);
            accR2[q] = __ffma2_rn(ee, make_float2(F.bR, F.bR), accR2[q]);
        }
        const float2 we = FLAT ? w : __fmul2_rn(w, ee);
        accL2[q] = __ffma2_rn(we, aL2, accL2[q]);
        accR2[q] = __ffma2_rn(we, aR2, accR2[q]);
        if (q + 1 < H) {
            P = __fadd2_rn(P, cstep);
            if (MODC == 2) sn = sv[q];
            if (MOD) P = __fadd2_rn(P, sn);
            if (MODC == 1) {
                dl = __ffma2_rn(nk2, sn, dl);
                sn = __fadd2_rn(sn, dl);
            }
            if (!FLAT) e = __fadd2_rn(e, estep);
            if (RARE && MOD && high) {
                // strong FM: the phase back into one turn (cycles: exactly; radians: two-constant 2*pi)
                const float2 big = make_float2(12582912.0f, 12582912.0f), nbig = make_float2(-12582912.0f, -12582912.0f);
                if (CYC) {
                    const float2 t = __fadd2_rn(__fadd2_rn(P, big), nbig);
                    P = __fadd2_rn(P, make_float2(-t.x, -t.y));
                } else {
                    const float2 t = __fadd2_rn(__ffma2_rn(P, make_float2(0.15915494f, 0.15915494f), big), nbig);
                    P = __ffma2_rn(t, make_float2(-6.28318548f, -6.28318548f), P);
                    P = __ffma2_rn(t, make_float2(1.74845553e-7f, 1.74845553e-7f), P);
                }
            }
            if (RARE) {
                if (q + 1 == swx) {
                    const EnvLine l = fast_env_line(&F, n + q + 1);
                    e.x = l.v;
                    estep.x = l.dv;
                    swx = l.left > 0x7fffffff - H ? 0x7fffffff : q + 1 + l.left;
                }
                if (q + 1 == swy) {
                    const EnvLine l = fast_env_line(&F, n + q + 1 + H);
                    e.y = l.v;
                    estep.y = l.dv;
                    swy = l.left > 0x7fffffff - H ? 0x7fffffff : q + 1 + l.left;
                }
            }
        }
    }
}

// ------------------------------------------------------------------ edged carriers, integer phase
// Square / Sawtooth / Pulse carriers: their value is exact (+-1, or a continuous ramp) unless the
// phase sits within a guard band of a discontinuity; those few samples are redone in fp64
// exactly as upstream forms the phase, so every edge lands on the sample upstream puts it on.
// They keep their phase as a 32-bit fraction of a cycle (2^32 = one
// cycle): the carrier step is an integer add that wraps for free, the waveform value and the
// "within the guard band of an edge" test are two or three integer operations each, and the
// resolution (2^-32 cycles, against fp32's 6e-8) lets the guard band shrink to 2e-6 cycles for
// voices without FM.  The FM part of the phase stays an fp32 sum of the increments (cycles),
// converted per sample at 2^-20 cycles resolution by a magic-number add.
constexpr uint32_t kGuardIntFm = 85899u;     // 2e-5 cycles: fp32 FM sum + its 2^-20 quantisation
constexpr uint32_t kGuardIntPlain = 8192u;   // 1.9e-6 cycles: 16 steps of a 2^-33-rounded step, plus the drift of the
                                             // time base's step from the nominal one (1e-8 relative an hour into a render)

__device__ __forceinline__ uint32_t cycles_to_phase(float phi)          // |phi| < 4 cycles
{
    return (uint32_t)(__float_as_int(__fmaf_rn(phi, 1048576.0f, 12582912.0f)) - 0x4B400000) << 12;
}

template <int WAVE>
__device__ __forceinline__ float edged_value(uint32_t p, uint32_t pwi, uint32_t g, bool &near)
{
    if (WAVE == SB_WAVE_SQUARE) {                          // int(tt*2) % 2: +1 on the first half cycle
        near = ((p + g) << 1) < (g << 2);                  // edges at 0 and 2^31
        return __uint_as_float(0x3f800000u | (p & 0x80000000u));
    }
    if (WAVE == SB_WAVE_SAWTOOTH) {                        // 2*(tt - floor(0.5 + tt)) = 2*frac(tt + 0.5) - 1
        const uint32_t q = p + 0x80000000u;
        near = (q + g) < (g << 1);                         // wraps where 0.5 + tt crosses an integer
        return fmaf(__uint_as_float(0x3f800000u | (q >> 9)), 2.0f, -3.0f);
    }
    near = ((p + g) < (g << 1)) | ((p - pwi + g) < (g << 1));          // tt % 1.0 < pulsewidth: edges at 0 and pw
    return p < pwi ? 1.0f : -1.0f;
}

// One pass over a lane's run.  FIX = false: the values from the integer phase, returns whether
// any sample sat in a guard band.  FIX = true (rare, per lane): the same walk, adding for every
// such sample the difference between the exact fp64 value (phase formed as upstream forms it)
// and the value the first pass used.
struct RunIncrements {                      // MODC 2: the FM increments gd*lfo of frames (q, q+H), by value
    float2 v[kFItems / 2];
};

template <int MODC, int WAVE, bool RARE, bool FLAT, bool FIX>
__device__ __forceinline__ bool edged_pass(const FastVoice &F, const float2 env, const float lane8, const int64_t n,
                                           const FastCtx ctx, const uint32_t p0, const double gbefore,
                                           float2 sn, float2 dl, const RunIncrements sv, const float half_sum,
                                           const bool low, float2 (&accL2)[kFItems / 2], float2 (&accR2)[kFItems / 2])
{
    constexpr int H = kFItems / 2;
    constexpr bool MOD = MODC != 0;
    constexpr uint32_t g = MOD ? kGuardIntFm : kGuardIntPlain;
    const float4 stepk = F.stepk;           // carrier step and pulse width as phase integers, -k, -k
    const float4 gain = F.gain;
    const uint32_t wci = __float_as_uint(stepk.x), pwi = __float_as_uint(stepk.y);
    const float2 nk2 = make_float2(stepk.z, stepk.w);
    uint32_t bx = p0, by = p0 + (uint32_t)H * wci;
    float2 run = make_float2(0.0f, half_sum);               // FM phase (cycles) ahead of frames (q, q+H)
    const float e0 = fmaf(lane8, env.y, env.x);
    float2 e = make_float2(e0, fmaf((float)H, env.y, e0));
    float2 estep = make_float2(env.y, env.y);
    int swx = 0x7fffffff, swy = 0x7fffffff;                // RARE: the step at which frame q / q+H enters a new ADSR segment
    if (RARE && env.x == kEnvSlow) {
        const EnvLine lx = fast_env_line(&F, n), ly = fast_env_line(&F, n + H);
        e = make_float2(lx.v, ly.v);
        estep = make_float2(lx.dv, ly.dv);
        swx = lx.left;
        swy = ly.left;
    }
    float2 aL2 = make_float2(gain.x, gain.y), aR2 = make_float2(gain.z, gain.w);
    if (FLAT) {                                             // constant envelope: folded into the gains
        const float2 ec = make_float2(env.x, env.x);
        aL2 = __fmul2_rn(aL2, ec);
        aR2 = __fmul2_rn(aR2, ec);
    }
    bool any = false;
#pragma unroll
    for (int q = 0; q < H; q++) {
        if (MODC == 2) sn = sv.v[q];
        const uint32_t px = MOD ? bx + cycles_to_phase(run.x) : bx;
        const uint32_t py = MOD ? by + cycles_to_phase(run.y) : by;
        bool nx, ny;
        float2 w = make_float2(edged_value<WAVE>(px, pwi, g, nx), edged_value<WAVE>(py, pwi, g, ny));
        if (!FIX) {
            any |= nx | ny;
        } else {
            float2 d = make_float2(0.0f, 0.0f);
            if (nx || low)
                d.x = exact_shape(WAVE, ctx, n + q, F.car.x, F.car.y, MOD ? gbefore + (double)run.x : 0.0, MOD, F.pw64) - w.x;
            if (ny || low)
                d.y = exact_shape(WAVE, ctx, n + q + H, F.car.x, F.car.y, MOD ? gbefore + (double)run.y : 0.0, MOD, F.pw64) - w.y;
            w = d;
        }
        const float2 ee = e;
        if (RARE) {
            if (!FIX) {
                accL2[q] = __ffma2_rn(ee, make_float2(F.bL, F.bL), accL2[q]);
                accR2[q] = __ffma2_rn(ee, make_float2(F.bR, F.bR), accR2[q]);
            }
        }
        const float2 we = FLAT ? w : __fmul2_rn(w, ee);
        accL2[q] = __ffma2_rn(we, aL2, accL2[q]);
        accR2[q] = __ffma2_rn(we, aR2, accR2[q]);
        if (q + 1 < H) {
            bx += wci;
            by += wci;
            if (MOD) run = __fadd2_rn(run, sn);
            if (MODC == 1) {
                dl = __ffma2_rn(nk2, sn, dl);
                sn = __fadd2_rn(sn, dl);
            }
            if (!FLAT) e = __fadd2_rn(e, estep);
            if (RARE) {
                if (q + 1 == swx) {
                    const EnvLine l = fast_env_line(&F, n + q + 1);
                    e.x = l.v;
                    estep.x = l.dv;
                    swx = l.left > 0x7fffffff - H ? 0x7fffffff : q + 1 + l.left;
                }
                if (q + 1 == swy) {
                    const EnvLine l = fast_env_line(&F, n + q + 1 + H);
                    e.y = l.v;
                    estep.y = l.dv;
                    swy = l.left > 0x7fffffff - H ? 0x7fffffff : q + 1 + l.left;
                }
            }
        }
    }
    return any;
}

// The fix-up pass as a real call: it is rare (a sample in a guard band, or the first cycles of a
// voice), it contains the calls of the exact evaluation, and kept inline it makes the compiler
// spill the accumulators around every one of them on the main path too.  It returns the
// corrections of the run's 16 frames through local memory.
struct RunFix {
    float2 L[kFItems / 2], R[kFItems / 2];
};

template <int MODC, int WAVE, bool RARE, bool FLAT>
__device__ __noinline__ RunFix edged_fix(const FastVoice *Fp, const float2 env, const float lane8, const int64_t n,
                                         const FastCtx ctx, const uint32_t p0, const double gbefore, const float2 sn,
                                         const float2 dl, const RunIncrements sv, const float half_sum, const bool low)
{
    RunFix r = {};
    edged_pass<MODC, WAVE, RARE, FLAT, true>(*Fp, env, lane8, n, ctx, p0, gbefore, sn, dl, sv, half_sum, low, r.L, r.R);
    return r;
}

template <int MODC, int WAVE, bool RARE, bool FLAT>
__device__ __forceinline__ void edged_voice_run(const FastVoice &F, const double2 anc, const float2 env,
                                                const double Tc0, const double dstep,
                                                const float lane8, const int64_t n,
                                                const FastCtx ctx, float2 (&accL2)[kFItems / 2],
                                                float2 (&accR2)[kFItems / 2])
{
    constexpr int H = kFItems / 2;
    double B = fma(F.car.x, Tc0, F.car.y);
    double gbefore = 0.0;
    float2 sn = make_float2(0.0f, 0.0f), dl = make_float2(0.0f, 0.0f);
    float half_sum = 0.0f;
    RunIncrements sv = {};
    if (MODC == 2) {
        double before;
        lfo_run_any(F, ctx, anc.y, Tc0, n, sv.v, half_sum, before);
        gbefore = F.gd * (anc.x + before);                 // anc.x: the modulator's sum before this sub-tile (tile pre-pass)
        B += gbefore;
    }
    if (MODC == 1) {
        const double2 L = ctx.lt_lane[F.lt_row];
        const FmSeg fs = ctx.tab_seg[F.tab_row];
        const double cl = fma(-anc.y, L.y, anc.x * L.x);
        const double sl = fma(anc.x, L.y, anc.y * L.x);
        double u = fma(-fs.Cc, cl, fs.A);
        u = fma(-0.5, sl, u);
        const double vv = fma(fs.Cc, sl, -0.5 * cl);
        gbefore = fma(F.mod.x * dstep, vv, u);
        B += gbefore;
        const float s0 = (float)sl, c0 = (float)cl;
        const float4 hf = F.half;           // cos(H w), sin(H w), P, Q
        const float4 ini = F.ini;           // k/2, k/2, sin w, sin w
        const float sH = fmaf(c0, hf.y, s0 * hf.x);
        const float cH = fmaf(-s0, hf.y, c0 * hf.x);
        sn = make_float2(s0, sH);
        dl = __ffma2_rn(make_float2(c0, cH), make_float2(ini.z, ini.w), __fmul2_rn(sn, make_float2(ini.x, ini.y)));
        half_sum = fmaf(s0, hf.z, c0 * hf.w);
    }
    // phase base: its fraction of a cycle as a 32-bit integer (two magic-number adds, no conversion)
    const double M = 6755399441055744.0;
    const double fr = B - __dadd_rn(__dadd_rn(B, M), -M);                      // [-0.5, 0.5]
    const uint32_t p0 = (uint32_t)__double2loint(__dadd_rn(fr * 4294967296.0, M));
    // while the (unreduced) phase may still be negative, Square's int() truncation is not the
    // periodic rule: take the exact path for the whole run (only the first frames of a voice:
    // this used to be "the first 16 cycles", which made the first span of a low voice - all of
    // its frames on the exact path - the kernel's critical path: 1.7 ms for one warp of a
    // 3.3 ms kernel)
    const bool low = !(B > (double)F.low_limit);
    const bool any = edged_pass<MODC, WAVE, RARE, FLAT, false>(F, env, lane8, n, ctx, p0, gbefore, sn, dl, sv, half_sum, low, accL2, accR2);
    if (any || low) {
#ifdef SB_CTA_TRACE
        {
            const unsigned w = (blockIdx.y * gridDim.x + blockIdx.x) * kFastWarps + (threadIdx.x >> 5);
            if (w < (unsigned)kTraceWarps) {
                const unsigned act = __activemask();
                if ((threadIdx.x & 31) == __ffs(act) - 1) atomicAdd(&g_cta_fix[w][0], 1u);
                if (low) atomicAdd(&g_cta_fix[w][1], 1u);
            }
        }
#endif
        const RunFix r = edged_fix<MODC, WAVE, RARE, FLAT>(&F, env, lane8, n, ctx, p0, gbefore, sn, dl, sv, half_sum, low);
#pragma unroll
        for (int q = 0; q < H; q++) {
            accL2[q] = __fadd2_rn(accL2[q], r.L[q]);
            accR2[q] = __fadd2_rn(accR2[q], r.R[q]);
        }
    }
}

// One instantiation per kind (bit0 = MOD, bits 1.. = carrier waveform); voices are stored
// sorted by kind, so each loop is a branch-free run.
template <int KIND>
__device__ __forceinline__ void fast_voice_loop(const FastVoice *s_fv, const double2 *s_anc, const float2 *s_env,
                                                int begin, int end, int s, const double Tr0, const double Tc0,
                                                const double dstep, const float lane8,
                                                const int64_t n, const FastCtx ctx,
                                                float2 (&accL2)[kFItems / 2], float2 (&accR2)[kFItems / 2])
{
    constexpr int MOD = KIND >= 10 ? 2 : (KIND & 1);         // modulator class (see fast_voice_run)
    constexpr int WAVE = KIND >= 10 ? KIND - 10 : (KIND >> 1);
    const double2 none = make_double2(0.0, 0.0);
    int v = begin;
#if FAST_PAIR_VOICES
    // two voices per iteration in one basic block: their (independent) dependency chains give
    // the scheduler twice the instruction-level parallelism to hide MUFU / fp64 latency
    if constexpr (WAVE < SB_WAVE_SQUARE && MOD != 2) {      // (pairing edged carriers was measured slower)
        for (; v + 1 < end; v += 2) {
            const FastVoice &F0 = s_fv[v], &F1 = s_fv[v + 1];
            const float2 e0 = s_env[v * kFastMaxSub + s], e1 = s_env[(v + 1) * kFastMaxSub + s];
            // both envelopes plain lines (sign bits clear: not kEnvSilent / kEnvSlow), no carrier bias
            const bool plain = (__float_as_int(e0.x) | __float_as_int(e1.x)) >= 0 && !((F0.flags | F1.flags) & (FV_BIAS | FV_HIGH));
            if (!plain) break;                                                  // warp-uniform
            const double2 a0 = MOD == 1 ? s_anc[v * kFastMaxSub + s] : none;
            const double2 a1 = MOD == 1 ? s_anc[(v + 1) * kFastMaxSub + s] : none;
            if (((__float_as_int(e0.y) | __float_as_int(e1.y)) << 1) == 0) {     // both slopes +-0: warp-uniform
                fast_voice_run<MOD, WAVE, false, true>(F0, a0, e0, Tr0, Tc0, dstep, lane8, n, ctx, accL2, accR2);
                fast_voice_run<MOD, WAVE, false, true>(F1, a1, e1, Tr0, Tc0, dstep, lane8, n, ctx, accL2, accR2);
            } else {
                fast_voice_run<MOD, WAVE, false, false>(F0, a0, e0, Tr0, Tc0, dstep, lane8, n, ctx, accL2, accR2);
                fast_voice_run<MOD, WAVE, false, false>(F1, a1, e1, Tr0, Tc0, dstep, lane8, n, ctx, accL2, accR2);
            }
        }
    }
#endif
    for (; v < end; v++) {
        const FastVoice &F = s_fv[v];
        const float2 env = s_env[v * kFastMaxSub + s];
        if (env.x == kEnvSilent) continue;                                      // warp-uniform
        const double2 anc = MOD != 0 ? s_anc[v * kFastMaxSub + s] : none;
        if constexpr (WAVE >= SB_WAVE_SQUARE) {
            if (env.x == kEnvSlow || (F.flags & FV_BIAS))
                edged_voice_run<MOD, WAVE, true, false>(F, anc, env, Tc0, dstep, lane8, n, ctx, accL2, accR2);
            else if (MOD != 2 && (__float_as_int(env.y) << 1) == 0)              // slope +-0: warp-uniform
                edged_voice_run<MOD, WAVE, false, MOD != 2>(F, anc, env, Tc0, dstep, lane8, n, ctx, accL2, accR2);
            else
                edged_voice_run<MOD, WAVE, false, false>(F, anc, env, Tc0, dstep, lane8, n, ctx, accL2, accR2);
        } else {
            if (env.x == kEnvSlow || (F.flags & (FV_BIAS | FV_HIGH)))
                fast_voice_run<MOD, WAVE, true, false>(F, anc, env, Tr0, Tc0, dstep, lane8, n, ctx, accL2, accR2);
            else
                fast_voice_run<MOD, WAVE, false, false>(F, anc, env, Tr0, Tc0, dstep, lane8, n, ctx, accL2, accR2);
        }
    }
}

// ---- the bank's voices sorted by kind (counting sort, stable, deterministic): histogram per block
// of 256 voices, one scan over (key, block), scatter by stable rank within the block.
constexpr int kSortKeys = kFastKinds + 1;          // the fast kinds, then the generic voices
constexpr int kSortBlock = 256;
// unit queues in part_kinds[]: per part the total of (group, span block) units, per kind the first
// group it owns, how many, and the draw counter.  A group belongs to the kind of its first voice of
// the part (the bank is sorted by kind, so a kind's groups are consecutive).
constexpr int kQUnits = 4, kQFirst = 8, kQGroups = 24, kQDrawn = 40;
__device__ __forceinline__ int sort_key_of(int flags) { return (flags & FV_FAST) ? fast_kind_of(flags) : kFastKinds; }

__global__ void __launch_bounds__(kSortBlock)
bank_key_hist_kernel(BankArgs a, int nblocks)
{
    __shared__ int s_count[kSortKeys];
    if (threadIdx.x < kSortKeys) s_count[threadIdx.x] = 0;
    __syncthreads();
    const int v = blockIdx.x * kSortBlock + threadIdx.x;
    if (v < a.nvoices) atomicAdd(&s_count[sort_key_of(a.fastv[v].flags)], 1);
    __syncthreads();
    if (threadIdx.x < kSortKeys) a.key_hist[threadIdx.x * nblocks + blockIdx.x] = s_count[threadIdx.x];
}

// exclusive scan of key_hist in place, key-major: where each (key, block) bucket starts in perm
__global__ void __launch_bounds__(1024)
bank_key_scan_kernel(BankArgs a, int nblocks)
{
    __shared__ int s_warp[32];
    __shared__ int s_carry;
    const int n = kSortKeys * nblocks;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_carry = 0;
    if (warp == 0) {
        // kinds present per kernel part (a lane per kind: is any of its per-block counts non-zero?)
        int present = 0;
        if (lane < kFastKinds)
            for (int b = 0; b < nblocks && !present; b++) present = a.key_hist[lane * nblocks + b] != 0;
        const unsigned m = __ballot_sync(0xffffffffu, present);
        if (lane < 3) {
            int c = 0;
            for (int k = 0; k < kFastKinds; k++) c += ((m >> k) & 1u) && fast_kind_part(k) == lane;
            a.part_kinds[lane] = c;
        }
    }
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int i = base + threadIdx.x;
        const int x = i < n ? a.key_hist[i] : 0;
        int inc = x;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int o = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += o;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            int w = s_warp[lane], winc = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int o = __shfl_up_sync(0xffffffffu, winc, d);
                if (lane >= d) winc += o;
            }
            s_warp[lane] = winc - w;
        }
        __syncthreads();
        const int carry = s_carry;
        if (i < n) a.key_hist[i] = carry + s_warp[warp] + inc - x;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = carry + s_warp[warp] + inc;
        __syncthreads();
    }
    // the unit queues: key_hist[k * nblocks] is now where kind k starts in the sorted bank
    if (warp == 0) {
        const int k = lane, vpg = a.voices_per_group;
        const int nblk = (a.fast_nspans + kFastWarps - 1) / kFastWarps;
        int units = 0, part = -1;
        if (k < kFastKinds) {
            part = fast_kind_part(k);
            const int pfirst = part == 0 ? 0 : (part == 1 ? 4 : 10);
            const int ps = a.key_hist[pfirst * nblocks];
            const int lo = a.key_hist[k * nblocks], hi = a.key_hist[(k + 1) * nblocks];   // k + 1 <= kFastKinds: the generic key
            int g_lo = 0, ng = 0;
            if (hi > lo) {
                g_lo = (lo == ps) ? ps / vpg : (lo + vpg - 1) / vpg;
                ng = (hi + vpg - 1) / vpg - g_lo;
            }
            // the part that writes the partial mixes covers every group of the bank (zeros where it has no voice)
            const unsigned ne = __ballot_sync(0x7fffu, hi > lo) & (part == 1 ? 0x3f0u : (part == 2 ? 0x7c00u : 0xfu));
            if (part == a.first_part) {
                if (ne == 0) {
                    if (k == pfirst) ng = a.ngroups;
                } else {
                    if (k == __ffs(ne) - 1) { ng += g_lo; g_lo = 0; }
                    if (k == 31 - __clz(ne)) ng = a.ngroups - g_lo;
                }
            }
            a.part_kinds[kQFirst + k] = g_lo;
            a.part_kinds[kQGroups + k] = ng;
            a.part_kinds[kQDrawn + k] = 0;
            units = ng * nblk;
        }
        for (int pt = 0; pt < 3; pt++) {
            int u = part == pt ? units : 0;
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) u += __shfl_xor_sync(0xffffffffu, u, d);
            if (lane == 0) a.part_kinds[kQUnits + pt] = u;
        }
    }
}

__global__ void __launch_bounds__(kSortBlock)
bank_key_scatter_kernel(BankArgs a, int nblocks)
{
    __shared__ int s_warp_count[kSortBlock / 32][kSortKeys];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < (kSortBlock / 32) * kSortKeys; i += kSortBlock) (&s_warp_count[0][0])[i] = 0;
    __syncthreads();
    const int v = blockIdx.x * kSortBlock + threadIdx.x;
    const int key = v < a.nvoices ? sort_key_of(a.fastv[v].flags) : -1;
    // lanes of this warp with the same key; the voice's rank among them by lane (= by voice index)
    const unsigned same = __match_any_sync(0xffffffffu, key);
    const int rank_in_warp = __popc(same & ((1u << lane) - 1));
    if (key >= 0 && rank_in_warp == 0) s_warp_count[warp][key] = __popc(same);
    __syncthreads();
    if (key >= 0) {
        int before = 0;
        for (int w = 0; w < warp; w++) before += s_warp_count[w][key];
        a.perm[a.key_hist[key * nblocks + blockIdx.x] + before + rank_in_warp] = v;
    }
}

// One CTA per voice group, once per render: for each of the three kernel parts, the group's
// voices of that part sorted by kind (stable), and where each kind ends.  The fast kernels'
// CTAs - there are thousands per render - then stage their voices with a plain indexed copy
// instead of each repeating this O(n^2) ranking against flags in global memory.
__global__ void __launch_bounds__(kMaxVoicesPerGroup)
bank_sort_kernel(BankArgs a)
{
    __shared__ int s_flags[kMaxVoicesPerGroup];
    const int group = blockIdx.x;
    const int v_begin = group * a.voices_per_group;
    int nv = a.nvoices - v_begin;
    if (nv > a.voices_per_group) nv = a.voices_per_group;
    const int v = threadIdx.x;
    if (v < nv) s_flags[v] = a.fastv[a.perm[v_begin + v]].flags;
    __syncthreads();
    for (int part = 0; part < 3; part++) {
        if (v < nv) {
            const int fl = s_flags[v];
            const int kind = fast_kind_of(fl);
            if ((fl & FV_FAST) && fast_kind_part(kind) == part) {
                int pos = 0;
                for (int u = 0; u < nv; u++) {
                    const int fu = s_flags[u];
                    const int ku = fast_kind_of(fu);
                    if (!(fu & FV_FAST) || fast_kind_part(ku) != part) continue;
                    pos += (ku < kind || (ku == kind && u < v)) ? 1 : 0;
                }
                a.order[(int64_t)part * a.nvoices + v_begin + pos] = a.perm[v_begin + v];
            }
        }
        if (v < kFastKinds) {
            int c = 0;
            for (int u = 0; u < nv; u++) {
                const int fu = s_flags[u];
                const int ku = fast_kind_of(fu);
                if (!(fu & FV_FAST) || fast_kind_part(ku) != part) continue;
                c += (ku <= v) ? 1 : 0;
            }
            a.kind_end[((int64_t)group * 3 + part) * kFastKinds + v] = c;
        }
    }
}

// PART 0: Sine / Triangle carriers with no or a Sine modulator.
// PART 1: Square / Sawtooth / Pulse carriers with no or a Sine modulator.
// PART 2: any carrier with a Triangle / Square / Sawtooth / Pulse modulator.
// Launched in this order; per group, the lowest part with voices in it writes the partial mix and
// the others add to it.  Separate kernels rather than one so that the exact-fallback call and its
// register needs stay out of the smooth carriers' loop (the headline path).
template <int PART>
__global__ void __launch_bounds__(kFastWarps * 32, PART != 0 ? 4 : FAST_MIN_BLOCKS)   // parts 1, 2: 128 registers
bank_fast_kernel(BankArgs a)
{
    extern __shared__ __align__(16) unsigned char smem[];
    FastVoice *s_fv = reinterpret_cast<FastVoice *>(smem);
    const size_t nrec = (size_t)kFastWarps * a.voices_per_group * kFastMaxSub;
    double2 *s_anc_all = reinterpret_cast<double2 *>(s_fv + a.voices_per_group);
    float2 *s_env_all = reinterpret_cast<float2 *>(s_anc_all + nrec);
    sb_time_seg *s_rad = reinterpret_cast<sb_time_seg *>(s_env_all + nrec);
    sb_time_seg *s_cyc = s_rad + a.nsegs_rad;
    __shared__ int s_kind_end[kFastKinds];     // voices are stored sorted by kind; end index of each kind

    // Grid over (voice group, block of spans).  The smooth carriers' kernel walks the span
    // blocks of a group first.  The other two go group-first when their voices are of one or two
    // kinds (measured on 8192 voices x 10 s: Sawtooth x Sine 4.5 -> 3.6 ms, Sine x Triangle LFO
    // 5.6 -> 5.0 ms), but not in a bank of many kinds, where that order would put every kind's code
    // in flight at once (instruction cache: 14 -> 21 ms for the mixed 5x5 bank), and not in the
    // smooth kernel (2.13 -> 2.29 ms).
    // (the smooth kernel keeps its 2-D grid, x = block of spans, y = group; the other two are launched 1-D)
    int group, span_block;
    if (PART == 0) {
        group = blockIdx.y;
        span_block = blockIdx.x;
    } else if (a.nvoices <= kMaxVoicesPerGroup) {
        const int nblk = (a.fast_nspans + kFastWarps - 1) / kFastWarps;
        group = (int)(blockIdx.x / (unsigned)nblk);
        span_block = (int)(blockIdx.x % (unsigned)nblk);
    } else {
        // a sorted bank: the CTA draws its unit from the queue of the kind its SM is given (the SMs
        // are shared out over the part's kinds in proportion to their units), moving on to the next
        // kind when that one is drained - so an SM runs one kind's loop at a time, whatever the
        // mix of kinds in the bank.  Units of a kind are drawn group-first.
        __shared__ int s_unit[2];
        if (threadIdx.x == 0) {
            const int *q = a.part_kinds;
            const int total = q[kQUnits + PART];
            int g = -1, sb = 0;
            if ((int)blockIdx.x < total) {
                constexpr int kfirst = PART == 1 ? 4 : 10, nk = PART == 1 ? 6 : 5;
                const int nblk = (a.fast_nspans + kFastWarps - 1) / kFastWarps;
                unsigned smid;
                asm("mov.u32 %0, %%smid;" : "=r"(smid));
                const int64_t target = ((int64_t)(smid % (unsigned)a.sm_count) * 2 + 1) * total / (2 * a.sm_count);
                int start = 0, cum = 0;
                for (int t = 0; t < nk; t++) {
                    const int u = q[kQGroups + kfirst + t] * nblk;
                    if (target >= cum && target < cum + u) start = t;
                    cum += u;
                }
                for (int t = 0; t < nk && g < 0; t++) {
                    const int k = kfirst + (start + t) % nk, ng = q[kQGroups + k];
                    if (ng == 0) continue;
                    const int u = atomicAdd(a.part_kinds + kQDrawn + k, 1);
                    if (u < ng * nblk) {
                        g = q[kQFirst + k] + u % ng;
                        sb = u / ng;
                    }
                }
            }
            s_unit[0] = g;
            s_unit[1] = sb;
        }
        __syncthreads();
        group = s_unit[0];
        span_block = s_unit[1];
        if (group < 0) return;                             // block-uniform: the grid covers every group
    }
    const int v_begin = group * a.voices_per_group;
    int nv = a.nvoices - v_begin;
    if (nv > a.voices_per_group) nv = a.voices_per_group;

    // this part's voices of the group, already ranked by kind (bank_sort_kernel): where each kind ends
    const int *kend = a.kind_end + ((int64_t)group * 3 + PART) * kFastKinds;
    nv = kend[kFastKinds - 1];
    // Who writes the group's partial mix and who adds to it: the lowest part that has voices in the
    // group writes (the parts are launched in order), the others add; a group without any voice
    // of the fast classes is written as zeros by the first part launched (the generic kernel adds
    // into it).  In a sorted bank a group's voices are nearly always of one part, so the partial
    // mixes are written once and read by the fold alone.
    const int *kall = a.kind_end + (int64_t)group * 3 * kFastKinds;
    // (only the parts this render launches count: the plan's hints may have ruled one out)
    const int n0 = (a.parts_mask & 1) ? kall[kFastKinds - 1] : 0, n1 = (a.parts_mask & 2) ? kall[2 * kFastKinds - 1] : 0;
    const int n2 = (a.parts_mask & 4) ? kall[3 * kFastKinds - 1] : 0;
    const int n_lower = (PART >= 1 ? n0 : 0) + (PART == 2 ? n1 : 0);
    const int n_fast = n0 + n1 + n2;
    const bool add = n_lower > 0;
    if (nv == 0 && (n_fast > 0 || a.first_part != PART)) return;       // block-uniform
    if (threadIdx.x < kFastKinds) s_kind_end[threadIdx.x] = kend[threadIdx.x];
    for (int i = threadIdx.x; i < a.nsegs_rad; i += blockDim.x) s_rad[i] = a.tb_rad[i];
    for (int i = threadIdx.x; i < a.nsegs_cyc; i += blockDim.x) s_cyc[i] = a.tb_cyc[i];
    // stage them in shared memory in that order, so that the inner loops are branch-free runs
    // of one instantiation each
    {
        constexpr int kWords = (int)(sizeof(FastVoice) / 16);
        const int *ord = a.order + (int64_t)PART * a.nvoices + v_begin;
        for (int i = threadIdx.x; i < nv * kWords; i += blockDim.x) {
            const int pos = i / kWords, w = i - pos * kWords;
            reinterpret_cast<int4 *>(s_fv + pos)[w] = reinterpret_cast<const int4 *>(a.fastv + ord[pos])[w];
        }
    }
    __syncthreads();
    const int *ke = s_kind_end;                        // read from shared memory where needed: no registers held
    const bool any_cyc = PART != 0 || nv > ke[1];

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int span = span_block * kFastWarps + warp;
    if (span >= a.fast_nspans) return;                 // no block-wide sync after this point
#ifdef SB_CTA_TRACE
    unsigned long long trace_t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(trace_t0));
#endif
    double2 *s_anc = s_anc_all + (size_t)warp * a.voices_per_group * kFastMaxSub;
    float2 *s_env = s_env_all + (size_t)warp * a.voices_per_group * kFastMaxSub;

    const int64_t span_begin = (int64_t)span * a.fast_span_len;
    int64_t span_end = span_begin + a.fast_span_len;
    if (span_end > a.frames) span_end = a.frames;
    const int nsub = (int)((span_end - span_begin + kFSub - 1) / kFSub);

    // ---- per-(voice, sub-tile) records: the modulator's (cos, sin) at the sub-tile's first
    //      frame, scaled by gd, and the ADSR line across the sub-tile.  The anchor of the span's
    //      first sub-tile is evaluated exactly (range reduction + fp64 polynomials); those of the
    //      following ones are that anchor rotated by one sub-tile's worth of modulator steps
    //      (four fp64 multiply-adds; the lanes' first-order step correction counts from the
    //      span's first frame, so it covers the rotation too).
    for (int v = lane; v < nv; v += 32) {
        const FastVoice &F = s_fv[v];
        const bool sine_mod = (F.flags & (FV_HAS_MOD | FV_SCAN)) == FV_HAS_MOD;
        double2 anc = make_double2(0.0, 0.0), rot = make_double2(1.0, 0.0);
        if (sine_mod) {
            const int64_t lo = a.n0 + span_begin;
            TimeCursor c;
            c.seek(s_rad, a.nsegs_rad, lo);
            const double th = reduce_two_pi(fma(F.mod.x, c.at(lo), F.mod.y));
            anc = make_double2(F.gd * cos_reduced(th), F.gd * sin_reduced(th));
            rot = a.lanetab[F.lt_row + 32];
        }
        for (int s = 0; s < nsub; s++) {
            const int64_t lo = a.n0 + span_begin + (int64_t)s * kFSub, hi = lo + kFSub;
            float2 r = make_float2(1.0f, 0.0f);                 // e_sub, de
            if (F.flags & FV_ENV) {
                if (lo > F.n_r) r.x = kEnvSilent;
                else if (hi <= F.n_a) { r.x = (float)lo * F.a_slope; r.y = F.a_slope; }
                else if (lo >= F.n_a && hi <= F.n_d) { r.x = 1.0f + (float)(lo - F.n_a) * F.d_slope; r.y = F.d_slope; }
                else if (lo >= F.n_d && hi <= F.n_s) { r.x = F.sl; }
                else if (lo >= F.n_s && hi <= F.n_r) { r.x = F.sl + (float)(lo - F.n_s) * F.r_slope; r.y = F.r_slope; }
                else r.x = kEnvSlow;
            }
            // class-2 voices: the modulator's sum over all frames before this sub-tile, from the tile
            // pre-pass's prefix table (read here, once per span, rather than from global memory in
            // the middle of every run)
            // ... and, where the whole sub-tile is clear of the modulator's edges and kinks, its value at
            // the sub-tile's first frame (plus 4 for a rising Triangle; NaN: not clear - the runs look
            // for themselves).  A seam of the time base inside the sub-tile is harmless while the time
            // is small (see lfo_run_any).
            if (PART == 2) {
                anc.x = a.lfo_prefix[(int64_t)F.tab_row * a.lfo_rows + a.lfo_nprefix + span_begin / kFSub + s];
                const int mwave = (F.flags >> FV_MWAVE_SHIFT) & 7;
                TimeCursor c;
                c.seek(s_cyc, a.nsegs_cyc, lo);
                const double Ts = c.at(lo);
                const bool seam = (hi >= c.next_start) && !(Ts < 128.0);
                const double reach = (double)(kFSub - 1) * scan_m_step32(F);
                const LfoLine L = lfo_line(mwave, Ts, F.mod.x, F.mod.y, c.step, __float_as_uint(F.ini.x),
                                           reach < 3.0e9 ? (uint32_t)(long long)ceil(reach) : 0xffffffffu);
                anc.y = (L.clear && !seam) ? ((mwave == SB_WAVE_TRIANGLE && L.dv > 0.0) ? L.v0 + 4.0 : L.v0)
                                           : __longlong_as_double(0x7ff8000000000000ll);
            }
            s_anc[v * kFastMaxSub + s] = anc;
            s_env[v * kFastMaxSub + s] = r;
            if (sine_mod) anc = make_double2(fma(-anc.y, rot.y, anc.x * rot.x), fma(anc.x, rot.y, anc.y * rot.x));
        }
    }
    __syncwarp();

    float *out = a.out + (int64_t)group * a.out_stride;
    const float lane8 = (float)(lane * kFItems);
    const double inc_rad = 6.283185307179586 / a.samplerate;

    for (int s = 0; s < nsub; s++) {
        const int64_t i0 = span_begin + (int64_t)s * kFSub + lane * kFItems;
        const int64_t n = a.n0 + i0;
        // exact serial time at this lane's first frame in both bases (a fresh look-up per
        // sub-tile costs ~40 instructions per 64 voices and keeps no cursor state in registers)
        double Tr0, Tc0 = 0.0, dstep;
        FastCtx ctx;
        ctx.s_cyc = s_cyc;
        ctx.nsegs_cyc = a.nsegs_cyc;
        ctx.lt_lane = a.lanetab + lane;
        {
            TimeCursor c;
            c.seek(s_rad, a.nsegs_rad, n);
            Tr0 = c.at(n);
            ctx.tab_seg = a.tab + c.idx;
            dstep = (double)(s * kFSub + lane * kFItems) * (c.step - inc_rad);      // frames from the span's first one
            ctx.cstep = 0.0;
            ctx.cstraddle = false;
            if (any_cyc) {                              // carriers whose phase counts in cycles present
                c.seek(s_cyc, a.nsegs_cyc, n);
                Tc0 = c.at(n);
                if (PART == 2) {                        // (modulators evaluated per sample live in this part only)
                    ctx.cstep = c.step;
                    ctx.cstraddle = (n + kFItems >= c.next_start);
                }
            }
        }
        // keep both in registers for the voice loops (the compiler would otherwise recompute
        // the int64 -> fp64 conversion per voice)
        asm volatile("" : "+d"(Tr0), "+d"(Tc0), "+d"(dstep));
        constexpr int H = kFItems / 2;
        float2 accL2[H], accR2[H];
#pragma unroll
        for (int q = 0; q < H; q++) accL2[q] = accR2[q] = make_float2(0.0f, 0.0f);

        if (PART == 0) {
            fast_voice_loop<0>(s_fv, s_anc, s_env, 0, ke[0], s, Tr0, Tc0, dstep, lane8, n, ctx, accL2, accR2);
            fast_voice_loop<1>(s_fv, s_anc, s_env, ke[0], ke[1], s, Tr0, Tc0, dstep, lane8, n, ctx, accL2, accR2);
            fast_voice_loop<2>(s_fv, s_anc, s_env, ke[1], ke[2], s, Tr0, Tc0, dstep, lane8, n, ctx, accL2, accR2);
            fast_voice_loop<3>(s_fv, s_anc, s_env, ke[2], ke[3], s, Tr0, Tc0, dstep, lane8, n, ctx, accL2, accR2);
        } else if (PART == 1) {
            fast_voice_loop<4>(s_fv, s_anc, s_env, 0, ke[4], s, Tr0, Tc0, dstep, lane8, n, ctx, accL2, accR2);
            fast_voice_loop<5>(s_fv, s_anc, s_env, ke[4], ke[5], s, Tr0, Tc0, dstep, lane8, n, ctx, accL2, accR2);
            fast_voice_loop<6>(s_fv, s_anc, s_env, ke[5], ke[6], s, Tr0, Tc0, dstep, lane8, n, ctx, accL2, accR2);
            fast_voice_loop<7>(s_fv, s_anc, s_env, ke[6], ke[7], s, Tr0, Tc0, dstep, lane8, n, ctx, accL2, accR2);
            fast_voice_loop<8>(s_fv, s_anc, s_env, ke[7], ke[8], s, Tr0, Tc0, dstep, lane8, n, ctx, accL2, accR2);
            fast_voice_loop<9>(s_fv, s_anc, s_env, ke[8], ke[9], s, Tr0, Tc0, dstep, lane8, n, ctx, accL2, accR2);
        } else {
            fast_voice_loop<10>(s_fv, s_anc, s_env, 0, ke[10], s, Tr0, Tc0, dstep, lane8, n, ctx, accL2, accR2);
            fast_voice_loop<11>(s_fv, s_anc, s_env, ke[10], ke[11], s, Tr0, Tc0, dstep, lane8, n, ctx, accL2, accR2);
            fast_voice_loop<12>(s_fv, s_anc, s_env, ke[11], ke[12], s, Tr0, Tc0, dstep, lane8, n, ctx, accL2, accR2);
            fast_voice_loop<13>(s_fv, s_anc, s_env, ke[12], ke[13], s, Tr0, Tc0, dstep, lane8, n, ctx, accL2, accR2);
            fast_voice_loop<14>(s_fv, s_anc, s_env, ke[13], ke[14], s, Tr0, Tc0, dstep, lane8, n, ctx, accL2, accR2);
        }

        // frame j of the run is accL2[j].x for j < H and accL2[j-H].y above; two frames per float4
        if (i0 + kFItems <= span_end) {
            float4 *dst = reinterpret_cast<float4 *>(out + i0 * 2);
#pragma unroll
            for (int k = 0; k < H; k++) {               // float4 k holds frames 2k, 2k+1
                const int fa = 2 * k, fb = 2 * k + 1;
                float4 o = make_float4(fa < H ? accL2[fa].x : accL2[fa - H].y, fa < H ? accR2[fa].x : accR2[fa - H].y,
                                       fb < H ? accL2[fb].x : accL2[fb - H].y, fb < H ? accR2[fb].x : accR2[fb - H].y);
                if (add) {                              // add to what the earlier kernel parts wrote
                    const float4 p = dst[k];
                    o = make_float4(o.x + p.x, o.y + p.y, o.z + p.z, o.w + p.w);
                }
                dst[k] = o;
            }
        } else {
#pragma unroll
            for (int q = 0; q < H; q++) {
                if (i0 + q < span_end) {
                    out[(i0 + q) * 2] = accL2[q].x + (add ? out[(i0 + q) * 2] : 0.0f);
                    out[(i0 + q) * 2 + 1] = accR2[q].x + (add ? out[(i0 + q) * 2 + 1] : 0.0f);
                }
                if (i0 + q + H < span_end) {
                    out[(i0 + q + H) * 2] = accL2[q].y + (add ? out[(i0 + q + H) * 2] : 0.0f);
                    out[(i0 + q + H) * 2 + 1] = accR2[q].y + (add ? out[(i0 + q + H) * 2 + 1] : 0.0f);
                }
            }
        }
    }
#ifdef SB_CTA_TRACE
    {
        unsigned long long trace_t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(trace_t1));
        unsigned smid;
        asm("mov.u32 %0, %%smid;" : "=r"(smid));
        const unsigned w = (blockIdx.y * gridDim.x + blockIdx.x) * kFastWarps + warp;
        if (lane == 0 && w < (unsigned)kTraceWarps) {
            g_cta_trace[PART][w][0] = trace_t0;
            g_cta_trace[PART][w][1] = trace_t1;
            g_cta_trace[PART][w][2] = ((unsigned long long)group << 32) | (unsigned)span;
            g_cta_trace[PART][w][3] = ((unsigned long long)nv << 32) | ((unsigned long long)(g_cta_fix[w][0] & 0xfffu) << 20) |
                                      ((unsigned long long)(g_cta_fix[w][1] & 0xfffu) << 8) | (smid & 0xffu);
            g_cta_fix[w][0] = g_cta_fix[w][1] = 0;
        }
    }
#endif
}

__device__ __forceinline__ double warp_scan_incl(double v, int lane)
{
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        double o = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += o;
    }
    return v;
}

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}

// ------------------------------------------------------------------ modulator tile pre-pass
// FV_SCAN voices (Triangle / Square / Sawtooth / Pulse modulators): the sum of the modulator's
// unit-amplitude values over every tile of kFSub frames, evaluated exactly (the same function
// the fast kernel uses per sample), then turned into an exclusive prefix per voice.  Rows:
// lfo_nprefix tiles over [0, n0) (the FM sum accumulated before this render starts; the last
// one may be short), then the render's own tiles from n0.  One warp per (voice, tile).
// The sum of a class-2 modulator's values over the `cnt` frames from n on (cnt <= kFSub, one
// sub-tile), by the whole warp: a lane takes the run of kFItems frames it takes in the render
// kernel - the straight line where the run is clear, else the same per-sample evaluation the
// render kernel falls back to, so that tile sums and run sums are sums of the same numbers.
__device__ __forceinline__ double lfo_tile_by_runs(const FastVoice &F, const sb_time_seg *s_cyc, const int nsegs_cyc,
                                                   const int64_t begin, const int tile_cnt, const int lane)
{
    const int mwave = (F.flags >> FV_MWAVE_SHIFT) & 7;
    const double fm = F.mod.x, phm = F.mod.y, mpw = scan_m_pw64(F);
    const int64_t n = begin + (int64_t)lane * kFItems;
    const int cnt = max(0, min(kFItems, tile_cnt - lane * kFItems));               // < kFItems only on a last, short tile
    double Tc0 = 0.0, cstep = 0.0;
    bool straddle = false;
    if (cnt > 0) {
        TimeCursor c;
        c.seek(s_cyc, nsegs_cyc, n);
        Tc0 = c.at(n);
        cstep = c.step;
        straddle = (n + kFItems >= c.next_start);
    }
    const LfoLine L = lfo_line(mwave, Tc0, fm, phm, cstep, __float_as_uint(F.ini.x), __float_as_uint(F.ini.y));
    // (per lane, as the render kernel's runs do it: the line; two lines around a single edge; else the
    // per-sample evaluation - in the lanes that need it only, so a tile with an edge in each of two
    // runs costs two settled breaks and not a pass of the exact path by the whole warp)
    double acc = 0.0;
    if (cnt == kFItems && L.clear && !straddle) {
        acc = fma((double)(kFItems * (kFItems - 1) / 2), L.dv, (double)kFItems * L.v0);
    } else if (cnt == kFItems) {
        LfoBreak bk;
        bk.js = 0;
        if (!straddle) bk = lfo_run_break(&F, mwave, Tc0, cstep, L.v0, L.dv, L.d1, L.d2);
        if (bk.js) {
            const int js = bk.js, rest = kFItems - js;
            acc = fma(0.5 * (double)js * (double)(js - 1), L.dv, (double)js * L.v0) +
                  fma(0.5 * (double)rest * (double)(rest - 1), bk.dv1, (double)rest * bk.v1);
        } else {
            acc = lfo_run_slow(mwave, s_cyc, nsegs_cyc, cstep, straddle, Tc0, n, fm, phm, mpw, 1.0).tot;
        }
    } else {
        for (int j = 0; j < cnt; j++) acc += lfo_shape_exact(mwave, exact_time_at(s_cyc, nsegs_cyc, n + j), fm, phm, mpw);
    }
    return warp_sum(acc);
}

// One warp per (voice, 32 consecutive tiles): a lane first tries its whole tile as one straight
// line (one exact phase, a handful of operations: at LFO rates most tiles are clear of the
// modulator's edges and kinks); the tiles that are not are then summed run by run by the warp.
__global__ void __launch_bounds__(kGenWarps * 32)
bank_lfo_tile_kernel(BankArgs a)
{
    extern __shared__ __align__(16) unsigned char smem[];
    sb_time_seg *s_cyc = reinterpret_cast<sb_time_seg *>(smem);
    for (int i = threadIdx.x; i < a.nsegs_cyc; i += blockDim.x) s_cyc[i] = a.tb_cyc[i];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int chunk = blockIdx.x * kGenWarps + (threadIdx.x >> 5);               // 32 rows each
    if (chunk * 32 >= a.lfo_rows) return;
    const int row = chunk * 32 + lane;
    int64_t begin = 0, end = 0;
    if (row < a.lfo_nprefix) {
        begin = (int64_t)row * kFSub;
        end = min(begin + kFSub, a.n0);
    } else if (row < a.lfo_rows) {
        begin = a.n0 + (int64_t)(row - a.lfo_nprefix) * kFSub;
        end = min(begin + kFSub, a.n0 + a.frames);
    }
    const int cnt = (int)(end - begin);                                           // 0 for lanes past the last row
    double T0 = 0.0, cstep = 0.0;
    bool straddle = false;
    if (cnt > 0) {
        TimeCursor c;
        c.seek(s_cyc, a.nsegs_cyc, begin);
        T0 = c.at(begin);
        cstep = c.step;
        straddle = (begin + cnt >= c.next_start);
    }
    const int nscan = min(a.scan_list[0], a.scan_cap);
    for (int e = blockIdx.y; e < nscan; e += gridDim.y) {
        const FastVoice &F = a.fastv[a.scan_list[2 + e]];
        const int mwave = (F.flags >> FV_MWAVE_SHIFT) & 7;
        // how far the phase moves over the tile, on the integer phase (all ones: too far to tell)
        const double reach = (double)(cnt - 1) * scan_m_step32(F);
        const LfoLine L = lfo_line(mwave, T0, F.mod.x, F.mod.y, cstep, __float_as_uint(F.ini.x),
                                   reach < 3.0e9 ? (uint32_t)(long long)ceil(reach) : 0xffffffffu);
        bool whole = cnt > 0 && L.clear && !straddle;
        double acc = 0.0;
        if (whole) acc = fma(0.5 * (double)cnt * (double)(cnt - 1), L.dv, (double)cnt * L.v0);
        if (cnt > 0 && !whole && !straddle && reach < 2.0e9) {
            // One edge (or kink) inside the tile - the usual case at LFO rates: two straight lines.
            // Where the second one starts is estimated from the phase distance to the edge and then
            // settled by the exact values of the two frames around it (an edge makes the value jump
            // by 2; a Triangle's kink needs no settling: the value is continuous there).
            const double step32 = scan_m_step32(F);
            const uint32_t dn = min(L.d1 - 1u, L.d2 - 1u) + 1u;               // nearest edge strictly ahead (0 = a whole cycle away)
            const uint32_t df = max(L.d1 - 1u, L.d2 - 1u) + 1u;
            const bool one = dn != 0u && step32 > 0.0 && (L.d1 == L.d2 || (double)df > reach + 3.0 * kGuardLfo + step32);
            if (one) {
                int js = (int)ceil((double)dn / step32);                      // first frame at or past the edge
                js = max(1, min(cnt - 1, js));
                bool ok = (mwave == SB_WAVE_TRIANGLE);
                const double mpw = scan_m_pw64(F);
                for (int tries = 0; tries < 4 && !ok; tries++) {
                    const double ta = __dadd_rn(T0, __dmul_rn((double)(js - 1), cstep)), tb2 = __dadd_rn(T0, __dmul_rn((double)js, cstep));
                    const bool before_a = fabs(lfo_shape_exact(mwave, ta, F.mod.x, F.mod.y, mpw) - fma((double)(js - 1), L.dv, L.v0)) < 0.5;
                    const bool before_b = fabs(lfo_shape_exact(mwave, tb2, F.mod.x, F.mod.y, mpw) - fma((double)js, L.dv, L.v0)) < 0.5;
                    if (before_a && !before_b) ok = true;
                    else if (!before_a) js--;                                  // the edge came earlier
                    else js++;                                                 // not yet
                    if (js < 1 || js > cnt - 1) break;
                }
                if (ok) {
                    const int rest = cnt - js;
                    const double tj = __dadd_rn(T0, __dmul_rn((double)js, cstep));
                    const double reach2 = (double)(rest - 1) * step32;
                    const LfoLine L2 = lfo_line(mwave, tj, F.mod.x, F.mod.y, cstep, __float_as_uint(F.ini.x),
                                                (uint32_t)(long long)ceil(reach2));
                    if (L2.clear) {
                        acc = fma(0.5 * (double)js * (double)(js - 1), L.dv, (double)js * L.v0) +
                              fma(0.5 * (double)rest * (double)(rest - 1), L2.dv, (double)rest * L2.v0);
                        whole = true;
                    }
                }
            }
        }
        unsigned todo = __ballot_sync(0xffffffffu, cnt > 0 && !whole);
        while (todo) {
            const int t = __ffs(todo) - 1;
            todo &= todo - 1;
            const int64_t tb = __shfl_sync(0xffffffffu, begin, t);
            const int tc = __shfl_sync(0xffffffffu, cnt, t);
            const double sum = lfo_tile_by_runs(F, s_cyc, a.nsegs_cyc, tb, tc, lane);
            if (lane == t) acc = sum;
        }
        if (cnt > 0) a.lfo_prefix[(int64_t)e * a.lfo_rows + row] = acc;
    }
}

// in-place exclusive prefix over the rows of each FV_SCAN voice: one warp per voice
__global__ void __launch_bounds__(kGenWarps * 32)
bank_lfo_prefix_kernel(BankArgs a)
{
    const int lane = threadIdx.x & 31;
    const int nscan = min(a.scan_list[0], a.scan_cap);
    for (int e = blockIdx.x * kGenWarps + (threadIdx.x >> 5); e < nscan; e += gridDim.x * kGenWarps) {
        double *rowp = a.lfo_prefix + (int64_t)e * a.lfo_rows;
        double carry = 0.0;
        for (int r0 = 0; r0 < a.lfo_rows; r0 += 32) {
            const int r = r0 + lane;
            const double v = (r < a.lfo_rows) ? rowp[r] : 0.0;
            const double incl = warp_scan_incl(v, lane);
            if (r < a.lfo_rows) rowp[r] = carry + (incl - v);
            carry += __shfl_sync(0xffffffffu, incl, 31);
        }
    }
}

// ------------------------------------------------------------------ generic path
// Base value and step of a time base for a run of kItems+1 consecutive indices.
struct TimeRun {
    double tb, step;
    bool straddles;       // a segment boundary falls inside the run: use the cursor per sample
};

__device__ __forceinline__ TimeRun time_run(TimeCursor &cur, int64_t n)
{
    TimeRun r;
    r.tb = cur.at(n);
    r.step = cur.step;
    r.straddles = (n + kItems >= cur.next_start);
    return r;
}

__device__ __forceinline__ double run_at(const TimeRun &r, const sb_time_seg *segs, int nsegs, int64_t n, int j)
{
    if (!r.straddles) return __dadd_rn(r.tb, __dmul_rn((double)j, r.step));
    TimeCursor c;
    c.seek(segs, nsegs, n + j);
    return c.at(n + j);
}

// Modulator value in fp64, as upstream computes it.  fp32 would not do here: the FM sum adds
// these values over the whole render, and an fp32 sine's error is not zero-mean over a stretch
// of a slow LFO — measured 87 LSB off on a 10 kHz carrier with a 0.1 Hz, depth-2 modulator
// (tools/accuracy_sweep.py) before this was fp64.
__device__ __forceinline__ double lfo_value(const sb_voice &V, double t)
{
    const double ph = __dadd_rn(__dmul_rn(t, V.m_freq), phase0_of(V.m_wave, V.m_phase));
    double s;
    switch (V.m_wave) {
    case SB_WAVE_SINE:
        s = sin(ph);
        break;
    case SB_WAVE_TRIANGLE:
        s = 4.0 * (fabs(py_mod1(__dadd_rn(ph, 0.75)) - 0.5) - 0.25);
        break;
    case SB_WAVE_SQUARE:
        s = (((long long)__dmul_rn(ph, 2.0)) & 1) ? -1.0 : 1.0;
        break;
    case SB_WAVE_SAWTOOTH:
        s = 2.0 * (ph - floor(__dadd_rn(0.5, ph)));
        break;
    default:
        s = (py_mod1(ph) < clamp_pulsewidth((double)V.m_pw)) ? 1.0 : -1.0;
        break;
    }
    return __dadd_rn(__dmul_rn(s, (double)V.m_amp), (double)V.m_bias);
}

// pre-pass: per-(voice, span) LFO sums of the generic FM voices
__global__ void __launch_bounds__(kGenWarps * 32)
bank_carry_kernel(BankArgs a)
{
    extern __shared__ __align__(16) unsigned char smem[];
    sb_time_seg *s_rad = reinterpret_cast<sb_time_seg *>(smem);
    sb_time_seg *s_cyc = s_rad + a.nsegs_rad;
    for (int i = threadIdx.x; i < a.nsegs_rad; i += blockDim.x) s_rad[i] = a.tb_rad[i];
    for (int i = threadIdx.x; i < a.nsegs_cyc; i += blockDim.x) s_cyc[i] = a.tb_cyc[i];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int64_t unit = ((int64_t)blockIdx.x * kGenWarps + (threadIdx.x >> 5));
    const int nrow = a.nprefix + a.nspans;
    const int64_t nunits = (int64_t)a.nvoices * nrow;
    if (unit >= nunits) return;
    const int v = (int)(unit / nrow), row = (int)(unit % nrow);
    if (a.fastv[v].flags & FV_FAST) return;            // closed form, no pre-pass
    const sb_voice V = a.voices[v];
    double acc = 0.0;
    if (V.m_wave != SB_WAVE_NONE) {
        const bool c_rad = (V.c_wave == SB_WAVE_SINE), m_rad = (V.m_wave == SB_WAVE_SINE);
        const sb_time_seg *cs = c_rad ? s_rad : s_cyc, *ms = m_rad ? s_rad : s_cyc;
        const int cn = c_rad ? a.nsegs_rad : a.nsegs_cyc, mn = m_rad ? a.nsegs_rad : a.nsegs_cyc;
        // absolute frame range of this row: a piece of [0, n0) (the FM sum accumulated before
        // this render starts) or a span of the render itself
        int64_t begin, end;
        if (row < a.nprefix) {
            begin = (int64_t)row * a.span_len;
            end = begin + a.span_len;
            if (end > a.n0) end = a.n0;
        } else {
            begin = a.n0 + (int64_t)(row - a.nprefix) * a.span_len;
            end = begin + a.span_len;
            if (end > a.n0 + a.frames) end = a.n0 + a.frames;
        }
        TimeCursor cc, mc;
        cc.seek(cs, cn, begin + lane);
        mc.seek(ms, mn, begin + lane);
        for (int64_t n = begin + lane; n < end; n += 32) {
            double t_n = cc.at(n);                               // cursor only moves forward
            double step = __dadd_rn(cc.at(n + 1), -t_n);         // exact: both on the table
            acc += __dmul_rn(lfo_value(V, mc.at(n)), step);
        }
    }
    acc = warp_sum(acc);
    if (lane == 0) a.carry[(int64_t)v * nrow + row] = acc;
}

__global__ void __launch_bounds__(kGenWarps * 32)
bank_generic_kernel(BankArgs a)
{
    extern __shared__ __align__(16) unsigned char smem[];
    sb_time_seg *s_rad = reinterpret_cast<sb_time_seg *>(smem);
    sb_time_seg *s_cyc = s_rad + a.nsegs_rad;
    sb_voice *s_voices = reinterpret_cast<sb_voice *>(s_cyc + a.nsegs_cyc);
    double *s_carry = reinterpret_cast<double *>(s_voices + a.voices_per_group);
    int *s_vidx = reinterpret_cast<int *>(s_carry + kGenWarps * a.voices_per_group);
    __shared__ int s_ngeneric;

    const int group = blockIdx.y;
    const int v_begin = group * a.voices_per_group;
    int nvg = a.nvoices - v_begin;
    if (nvg > a.voices_per_group) nvg = a.voices_per_group;

    // compact the generic voices of this group into shared memory (in voice order)
    if (threadIdx.x == 0) {
        int c = 0;
        for (int v = 0; v < nvg; v++) {
            const int vi = a.perm[v_begin + v];
            if (!(a.fastv[vi].flags & FV_FAST)) s_vidx[c++] = vi;
        }
        s_ngeneric = c;
    }
    __syncthreads();
    const int nv = s_ngeneric;
    if (nv == 0) return;                               // block-uniform

    for (int i = threadIdx.x; i < a.nsegs_rad; i += blockDim.x) s_rad[i] = a.tb_rad[i];
    for (int i = threadIdx.x; i < a.nsegs_cyc; i += blockDim.x) s_cyc[i] = a.tb_cyc[i];
    {
        const int words = (int)(sizeof(sb_voice) / 4);
        for (int i = threadIdx.x; i < nv * words; i += blockDim.x) {
            const int v = i / words, w = i - v * words;
            reinterpret_cast<uint32_t *>(s_voices)[i] = reinterpret_cast<const uint32_t *>(a.voices + s_vidx[v])[w];
        }
    }
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int span = blockIdx.x * kGenWarps + warp;
    if (span >= a.nspans) return;                      // no block-wide sync after this point
    double *carry = s_carry + warp * a.voices_per_group;

    // FM running sums at the start of this span: sum of the earlier spans' totals
    for (int v = 0; v < nv; v++) {
        double c = 0.0;
        if (a.carry && s_voices[v].m_wave != SB_WAVE_NONE) {
            const double *row = a.carry + (int64_t)s_vidx[v] * (a.nprefix + a.nspans);
            for (int s = lane; s < a.nprefix + span; s += 32) c += row[s];
            c = warp_sum(c);
        }
        if (lane == 0) carry[v] = c;
    }
    __syncwarp();

    const int64_t span_begin = (int64_t)span * a.span_len;
    int64_t span_end = span_begin + a.span_len;
    if (span_end > a.frames) span_end = a.frames;
    float *out = a.out + (int64_t)group * a.out_stride;

    TimeCursor cur_rad, cur_cyc;
    cur_rad.seek(s_rad, a.nsegs_rad, a.n0 + span_begin + lane * kItems);
    cur_cyc.seek(s_cyc, a.nsegs_cyc, a.n0 + span_begin + lane * kItems);

    for (int64_t sub = span_begin; sub < span_end; sub += kSub) {
        const int64_t i0 = sub + lane * kItems;        // first frame of this lane (relative)
        const int64_t n = a.n0 + i0;                   // absolute sample index
        const TimeRun rr = time_run(cur_rad, n);
        const TimeRun rc = time_run(cur_cyc, n);
        float accL[kItems], accR[kItems];
#pragma unroll
        for (int j = 0; j < kItems; j++) accL[j] = accR[j] = 0.0f;

        for (int v = 0; v < nv; v++) {
            const sb_voice &V = s_voices[v];
            // EnvelopeFilter yields 0.0 after the release without touching its source
            if (V.env.enabled && a.n0 + sub > V.env.n_r) continue;
            const bool c_rad = (V.c_wave == SB_WAVE_SINE);
            const TimeRun &cr = c_rad ? rr : rc;
            const sb_time_seg *cs = c_rad ? s_rad : s_cyc;
            const int cn = c_rad ? a.nsegs_rad : a.nsegs_cyc;
            const double pc0 = phase0_of(V.c_wave, V.c_phase);
            const double pw = (V.c_wave == SB_WAVE_PULSE) ? clamp_pulsewidth((double)V.c_pw) : 0.0;

            double ph[kItems];
            if (V.m_wave != SB_WAVE_NONE) {
                const bool m_rad = (V.m_wave == SB_WAVE_SINE);
                const TimeRun &mr = m_rad ? rr : rc;
                const sb_time_seg *ms = m_rad ? s_rad : s_cyc;
                const int mn = m_rad ? a.nsegs_rad : a.nsegs_cyc;
                double w[kItems];
                double local = 0.0;
                double t_prev = run_at(cr, cs, cn, n, 0);
                double tc[kItems];
#pragma unroll
                for (int j = 0; j < kItems; j++) {
                    double t_next = run_at(cr, cs, cn, n, j + 1);
                    tc[j] = t_prev;
                    w[j] = __dmul_rn(lfo_value(V, run_at(mr, ms, mn, n, j)), __dadd_rn(t_next, -t_prev));
                    local += w[j];
                    t_prev = t_next;
                }
                double incl = warp_scan_incl(local, lane);
                double base = carry[v] + (incl - local);
                double total = __shfl_sync(0xffffffffu, incl, 31);
                __syncwarp();
                if (lane == 0) carry[v] += total;
                __syncwarp();
#pragma unroll
                for (int j = 0; j < kItems; j++) {
                    ph[j] = __dadd_rn(pc0, __dmul_rn(V.c_freq, __dadd_rn(tc[j], base)));
                    base += w[j];
                }
            } else {
#pragma unroll
                for (int j = 0; j < kItems; j++)
                    ph[j] = __dadd_rn(__dmul_rn(run_at(cr, cs, cn, n, j), V.c_freq), pc0);
            }
#pragma unroll
            for (int j = 0; j < kItems; j++) {
                float s = wave_shape(V.c_wave, ph[j], pw);
                float val = __fadd_rn(__fmul_rn(s, V.c_amp), V.c_bias);
                val = __fmul_rn(val, envelope_at(V.env, n + j));
                accL[j] = fmaf(val, V.gain_l, accL[j]);
                accR[j] = fmaf(val, V.gain_r, accR[j]);
            }
        }

        // add to what the fast kernel wrote: 8 frames x 2 channels = 16 consecutive floats per lane
        if (i0 + kItems <= span_end) {
            float4 *dst = reinterpret_cast<float4 *>(out + i0 * 2);
#pragma unroll
            for (int q = 0; q < kItems / 2; q++) {
                float4 o = dst[q];
                dst[q] = make_float4(o.x + accL[2 * q], o.y + accR[2 * q], o.z + accL[2 * q + 1], o.w + accR[2 * q + 1]);
            }
        } else {
#pragma unroll
            for (int j = 0; j < kItems; j++)
                if (i0 + j < span_end) {
                    out[(i0 + j) * 2] += accL[j];
                    out[(i0 + j) * 2 + 1] += accR[j];
                }
        }
    }
}

// mix[f] = sum over groups of partial[g][f]; float4 = two stereo frames
__global__ void bank_fold_kernel(const float *__restrict__ partial, int ngroups, int64_t nfloats,
                                 int64_t gstride, float *__restrict__ mix)
{
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t nvec = nfloats / 4;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
        float4 acc = reinterpret_cast<const float4 *>(partial)[i];
        int g = 1;
        // eight groups' loads in flight per thread; the additions stay in group order
        for (; g + 8 <= ngroups; g += 8) {
            float4 p[8];
#pragma unroll
            for (int k = 0; k < 8; k++) p[k] = __ldcs(reinterpret_cast<const float4 *>(partial + (int64_t)(g + k) * gstride) + i);
#pragma unroll
            for (int k = 0; k < 8; k++) { acc.x += p[k].x; acc.y += p[k].y; acc.z += p[k].z; acc.w += p[k].w; }
        }
        for (; g < ngroups; g++) {
            float4 p = reinterpret_cast<const float4 *>(partial + (int64_t)g * gstride)[i];
            acc.x += p.x; acc.y += p.y; acc.z += p.z; acc.w += p.w;
        }
        reinterpret_cast<float4 *>(mix)[i] = acc;
    }
    for (int64_t i = nvec * 4 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nfloats; i += stride) {
        float acc = partial[i];
        for (int g = 1; g < ngroups; g++) acc += partial[(int64_t)g * gstride + i];
        mix[i] = acc;
    }
}

static inline int64_t align_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }

// Scratch layout after the span sums and FastVoice array (plan->off_tab onwards): closed-form FM
// table, lane rotation table, list of FV_SCAN voices, the per-part voice order and kind ends, and last - so that its size can follow
// the scan_voices hint - the modulator tile prefix table [scan slots][rows].
static inline int64_t bank_off_lanetab(const sb_bank_plan *p)
{
    return align_up(p->off_tab + (int64_t)p->nvoices * p->nsegs_rad * (int64_t)sizeof(FmSeg), 256);
}
static inline int64_t bank_off_scanlist(const sb_bank_plan *p)
{
    return bank_off_lanetab(p) + align_up((int64_t)p->nvoices * kLaneTab * (int64_t)sizeof(double2), 256);
}
static inline int64_t bank_off_order(const sb_bank_plan *p)
{
    return bank_off_scanlist(p) + align_up((int64_t)(p->nvoices + 2) * (int64_t)sizeof(int), 256);
}
static inline int64_t bank_off_kindend(const sb_bank_plan *p)
{
    return bank_off_order(p) + align_up(3 * (int64_t)p->nvoices * (int64_t)sizeof(int), 256);
}
static inline int64_t bank_off_perm(const sb_bank_plan *p)
{
    return bank_off_kindend(p) + align_up((int64_t)p->ngroups * 3 * kFastKinds * (int64_t)sizeof(int), 256);
}
static inline int64_t bank_sort_blocks(const sb_bank_plan *p) { return (p->nvoices + kSortBlock - 1) / kSortBlock; }
static inline int64_t bank_off_keyhist(const sb_bank_plan *p)
{
    return bank_off_perm(p) + align_up((int64_t)p->nvoices * (int64_t)sizeof(int), 256);
}
static inline int64_t bank_off_partkinds(const sb_bank_plan *p)
{
    return bank_off_keyhist(p) + align_up(kSortKeys * bank_sort_blocks(p) * (int64_t)sizeof(int), 256);
}
static inline int64_t bank_off_lfo(const sb_bank_plan *p) { return bank_off_partkinds(p) + 256; }
static inline int64_t bank_lfo_nprefix(const sb_bank_plan *p) { return (p->n0 + kFSub - 1) / kFSub; }
static inline int64_t bank_lfo_rows(const sb_bank_plan *p) { return bank_lfo_nprefix(p) + (p->frames + kFSub - 1) / kFSub; }
static inline int64_t bank_scratch_bytes(const sb_bank_plan *p)
{
    const int64_t slots = p->scan_voices < 0 ? p->nvoices : p->scan_voices;
    return bank_off_lfo(p) + align_up(slots * bank_lfo_rows(p) * (int64_t)sizeof(double), 256);
}

}  // namespace sb

using namespace sb;

namespace sb {
// measurement hook: CUDA events recorded on the render's stream right before and after the
// launch of bank_fast_kernel<0> (bench.py times the dominant kernel alone with them)
static cudaEvent_t g_prof_start = nullptr, g_prof_stop = nullptr;

// the hint counts of a whole bank, for host-side callers (one pass, divisions hoisted)
// generic = class 0; smooth / edged = class 1 with a Sine, Triangle / Square, Sawtooth, Pulse carrier;
// scan = class 2: the voices each kernel of a render works on
void bank_count_classes(const sb_voice *voices, int nvoices, double samplerate, int *generic, int *smooth, int *edged,
                        int *scan)
{
    const double inv_sr = 1.0 / samplerate, w1 = 6.283185307179586 / samplerate;
    int g = 0, sm = 0, e = 0, sc = 0;
    for (int v = 0; v < nvoices; v++) {
        const int cls = voice_class(voices[v], samplerate, inv_sr, w1);
        if (!cls) g++;
        else if (cls == 2) sc++;
        else if (voices[v].c_wave >= SB_WAVE_SQUARE) e++;
        else sm++;
    }
    *generic = g;
    *smooth = sm;
    *edged = e;
    *scan = sc;
}
}  // namespace sb

namespace sb {
// SMs of the current device (asked once per device)
static int bank_sm_count()
{
    static int cached[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return -1;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n < 1) return -1;
        cached[dev] = n;
    }
    return cached[dev];
}
}  // namespace sb

#ifdef SB_CTA_TRACE
extern "C" int sb_debug_cta_trace(unsigned long long *host, int clear)
{
    if (clear) {
        void *p = nullptr;
        SB_CUDA_CHECK(cudaGetSymbolAddress(&p, g_cta_trace));
        SB_CUDA_CHECK(cudaMemset(p, 0, sizeof(g_cta_trace)));
        return SB_OK;
    }
    SB_CUDA_CHECK(cudaDeviceSynchronize());
    SB_CUDA_CHECK(cudaMemcpyFromSymbol(host, g_cta_trace, sizeof(g_cta_trace)));
    return SB_OK;
}
#endif

extern "C" int sb_bank_profile_events(void *start_event, void *stop_event)
{
    g_prof_start = (cudaEvent_t)start_event;
    g_prof_stop = (cudaEvent_t)stop_event;
    return SB_OK;
}

extern "C" int sb_bank_count_classes(const sb_voice *voices, int nvoices, double samplerate, int *counts4)
{
    SB_REQUIRE(voices && counts4 && nvoices >= 1 && samplerate > 1, "sb_bank_count_classes: bad arguments");
    bank_count_classes(voices, nvoices, samplerate, counts4 + 0, counts4 + 1, counts4 + 2, counts4 + 3);
    return SB_OK;
}

extern "C" int sb_voice_is_fast(const sb_voice *voice, double samplerate)
{
    if (!voice || !(samplerate > 1)) return 0;
    return voice_class(*voice, samplerate);
}

extern "C" int sb_bank_plan_make(int nvoices, int64_t frames, int64_t n0, int sm_count, int nsegs_rad, sb_bank_plan *plan)
{
    SB_REQUIRE(plan && nvoices >= 1 && frames >= 1 && n0 >= 0 && nsegs_rad >= 1 && nsegs_rad <= kMaxSegs,
               "sb_bank_plan_make: bad arguments");
    if (sm_count <= 0) sm_count = 148;
    // target: several waves of warps per SM (each warp is an independent unit, so many short
    // units balance the tail), but spans of at least 4 sub-tiles so that the per-span set-up
    // stays ~1% of the work.
    const int64_t target_warps = (int64_t)sm_count * 192;
    int vpg = kMaxVoicesPerGroup;
    if (nvoices < vpg) vpg = nvoices;
    int64_t nsub = (frames + kSub - 1) / kSub;
    int ngroups = (nvoices + vpg - 1) / vpg;
    int64_t nspans = (target_warps + ngroups - 1) / ngroups;
    const int64_t max_spans = (nsub + 3) / 4;
    if (nspans > max_spans) {
        // not enough time parallelism: split voices finer
        nspans = max_spans;
        int64_t want_groups = (target_warps + nspans - 1) / nspans;
        if (want_groups > nvoices) want_groups = nvoices;
        if (want_groups > 256) want_groups = 256;       // partial buffers: ngroups * frames * 8 B
        if (want_groups < 1) want_groups = 1;
        vpg = (int)((nvoices + want_groups - 1) / want_groups);
        ngroups = (nvoices + vpg - 1) / vpg;
    }
    if (nspans < 1) nspans = 1;
    int64_t span_len = ((nsub + nspans - 1) / nspans) * kSub;
    nspans = (frames + span_len - 1) / span_len;
    // fast path: spans of at most kFastMaxSub sub-tiles (their anchors sit in shared memory)
    int64_t fast_sub = span_len / kFSub;
    if (fast_sub > kFastMaxSub) fast_sub = kFastMaxSub;
    if (fast_sub < 1) fast_sub = 1;
    plan->nvoices = nvoices;
    plan->voices_per_group = vpg;
    plan->ngroups = ngroups;
    plan->nspans = (int32_t)nspans;
    plan->frames = frames;
    plan->span_len = span_len;
    plan->fast_span_len = fast_sub * kFSub;
    plan->fast_nspans = (int32_t)((frames + plan->fast_span_len - 1) / plan->fast_span_len);
    plan->nsegs_rad = nsegs_rad;
    plan->generic_voices = -1;
    plan->smooth_voices = -1;
    plan->edged_voices = -1;
    plan->scan_voices = -1;
    plan->n0 = n0;
    const int64_t nprefix = (n0 + span_len - 1) / span_len;
    SB_REQUIRE(nprefix < 0x7fffffff, "sb_bank_plan_make: start offset too large");
    plan->nprefix = (int32_t)nprefix;
    plan->off_fastv = align_up((int64_t)nvoices * (nprefix + nspans) * (int64_t)sizeof(double), 256);
    plan->off_tab = align_up(plan->off_fastv + (int64_t)nvoices * (int64_t)sizeof(FastVoice), 256);
    SB_REQUIRE(bank_lfo_rows(plan) < 0x7fffffff, "sb_bank_plan_make: too many frames");
    plan->carry_bytes = bank_scratch_bytes(plan);          // worst case until sb_bank_plan_hint() says otherwise
    plan->partial_bytes = (ngroups > 1) ? (int64_t)ngroups * ((frames * 2 + 3) / 4 * 4) * (int64_t)sizeof(float) : 0;
    return SB_OK;
}

extern "C" int sb_bank_plan_hint(sb_bank_plan *plan, int generic_voices, int smooth_voices, int edged_voices,
                                 int scan_voices)
{
    SB_REQUIRE(plan && plan->nvoices >= 1 && generic_voices >= -1 && edged_voices >= -1 && scan_voices >= -1 &&
                   smooth_voices >= -1 && generic_voices <= plan->nvoices && edged_voices <= plan->nvoices &&
                   scan_voices <= plan->nvoices && smooth_voices <= plan->nvoices,
               "sb_bank_plan_hint: bad arguments");
    plan->generic_voices = generic_voices;
    plan->smooth_voices = smooth_voices;
    plan->edged_voices = edged_voices;
    plan->scan_voices = scan_voices;
    plan->carry_bytes = bank_scratch_bytes(plan);
    return SB_OK;
}

extern "C" int sb_bank_render_f32(const sb_voice *voices_dev, const sb_bank_plan *plan,
                                  const sb_time_seg *tb_rad_dev, int nsegs_rad,
                                  const sb_time_seg *tb_cyc_dev, int nsegs_cyc,
                                  int64_t n0, double samplerate, void *scratch_dev, float *partial_dev,
                                  float *mix_dev, void *stream)
{
    SB_REQUIRE(voices_dev && plan && tb_rad_dev && tb_cyc_dev && mix_dev && scratch_dev,
               "sb_bank_render_f32: null pointer");
    SB_REQUIRE(nsegs_rad > 0 && nsegs_rad <= kMaxSegs && nsegs_cyc > 0 && nsegs_cyc <= kMaxSegs &&
                   nsegs_rad == plan->nsegs_rad && samplerate > 1,
               "sb_bank_render_f32: bad time base");
    SB_REQUIRE(plan->voices_per_group >= 1 && plan->voices_per_group <= kMaxVoicesPerGroup &&
                   plan->span_len % kSub == 0 && plan->nspans >= 1 && plan->ngroups >= 1 && n0 >= 0 &&
                   plan->fast_span_len % kFSub == 0 && plan->fast_span_len / kFSub <= kFastMaxSub &&
                   plan->fast_nspans >= 1,
               "sb_bank_render_f32: bad plan");
    SB_REQUIRE(plan->ngroups == 1 || partial_dev, "sb_bank_render_f32: partial scratch missing");
    SB_REQUIRE(plan->ngroups <= 65535, "sb_bank_render_f32: too many voice groups");
    cudaStream_t s = (cudaStream_t)stream;
    unsigned char *scratch = (unsigned char *)scratch_dev;
    BankArgs a;
    a.voices = voices_dev;
    a.nvoices = plan->nvoices;
    a.voices_per_group = plan->voices_per_group;
    a.ngroups = plan->ngroups;
    a.nspans = plan->nspans;
    a.frames = plan->frames;
    a.span_len = plan->span_len;
    a.n0 = n0;
    a.tb_rad = tb_rad_dev;
    a.nsegs_rad = nsegs_rad;
    a.tb_cyc = tb_cyc_dev;
    a.nsegs_cyc = nsegs_cyc;
    a.samplerate = samplerate;
    SB_REQUIRE(plan->n0 == n0, "sb_bank_render_f32: the plan was made for another start offset");
    a.nprefix = plan->nprefix;
    a.carry = (plan->nspans + plan->nprefix > 1) ? (double *)scratch : nullptr;
    a.fastv = (FastVoice *)(scratch + plan->off_fastv);
    a.tab = (FmSeg *)(scratch + plan->off_tab);
    a.lanetab = (double2 *)(scratch + bank_off_lanetab(plan));
    a.scan_list = (int *)(scratch + bank_off_scanlist(plan));
    a.order = (int *)(scratch + bank_off_order(plan));
    a.perm = (int *)(scratch + bank_off_perm(plan));
    a.key_hist = (int *)(scratch + bank_off_keyhist(plan));
    a.part_kinds = (int *)(scratch + bank_off_partkinds(plan));
    a.sm_count = bank_sm_count();
    SB_REQUIRE(a.sm_count >= 1, "sb_bank_render_f32: no device");
    a.kind_end = (int *)(scratch + bank_off_kindend(plan));
    a.lfo_prefix = (double *)(scratch + bank_off_lfo(plan));
    a.lfo_nprefix = (int)bank_lfo_nprefix(plan);
    a.lfo_rows = (int)bank_lfo_rows(plan);
    a.scan_cap = plan->scan_voices < 0 ? plan->nvoices : plan->scan_voices;
    SB_CUDA_CHECK(cudaMemsetAsync(a.scan_list, 0, 2 * sizeof(int), s));
    a.fast_nspans = plan->fast_nspans;
    a.fast_span_len = plan->fast_span_len;
    a.out = (plan->ngroups > 1) ? partial_dev : mix_dev;
    a.out_stride = (plan->frames * 2 + 3) / 4 * 4;

    // a part no voice belongs to is not launched; the first one that is writes the partial mixes
    const bool run0 = plan->smooth_voices != 0 || (plan->edged_voices == 0 && plan->scan_voices == 0);
    a.first_part = run0 ? 0 : (plan->edged_voices != 0 ? 1 : 2);
    a.parts_mask = (run0 ? 1 : 0) | (plan->edged_voices != 0 ? 2 : 0) | (plan->scan_voices != 0 ? 4 : 0);

    bank_prepare_kernel<<<(a.nvoices + 3) / 4, 128, 0, s>>>(a);       // one warp per voice
    SB_LAUNCH_CHECK();
    if (a.nvoices > kMaxVoicesPerGroup) {
        // the voices by kind across the whole bank, then each group's (a slice of that order) by part
        // (a bank that fits one full group is ranked by bank_sort_kernel alone)
        const int nb = (int)bank_sort_blocks(plan);
        bank_key_hist_kernel<<<nb, kSortBlock, 0, s>>>(a, nb);
        SB_LAUNCH_CHECK();
        bank_key_scan_kernel<<<1, 1024, 0, s>>>(a, nb);
        SB_LAUNCH_CHECK();
        bank_key_scatter_kernel<<<nb, kSortBlock, 0, s>>>(a, nb);
        SB_LAUNCH_CHECK();
    }
    bank_sort_kernel<<<a.ngroups, kMaxVoicesPerGroup, 0, s>>>(a);     // one CTA per voice group
    SB_LAUNCH_CHECK();
    if (plan->scan_voices != 0) {
        // modulator tile sums of the FV_SCAN voices, then their prefix (the fast kernel reads it)
        // a CTA keeps its frames' times in registers and walks the voices in steps of grid.y
        const int64_t tx = ((a.lfo_rows + 31) / 32 + kGenWarps - 1) / kGenWarps;   // a warp per 32 tiles
        // (each iteration of a warp's voice loop starts with dependent global loads - the list entry, then
        // the voice record - so many short warps beat few long ones: per 8192 voices x 10 s, 592 CTAs
        // 2.5 ms, 4736 0.76 ms, 18944 0.6 ms)
        int64_t ty = (18944 + tx - 1) / tx;
        if (ty > a.scan_cap) ty = a.scan_cap;
        if (ty < 1) ty = 1;
        dim3 tgrid((unsigned)tx, (unsigned)ty);
        bank_lfo_tile_kernel<<<tgrid, kGenWarps * 32, (size_t)nsegs_cyc * sizeof(sb_time_seg), s>>>(a);
        SB_LAUNCH_CHECK();
        const int pgrid = (a.scan_cap + kGenWarps - 1) / kGenWarps;
        bank_lfo_prefix_kernel<<<pgrid < 1184 ? pgrid : 1184, kGenWarps * 32, 0, s>>>(a);
        SB_LAUNCH_CHECK();
    }

    const size_t tb_bytes = (size_t)(nsegs_rad + nsegs_cyc) * sizeof(sb_time_seg);
    {
        size_t smem = tb_bytes + (size_t)a.voices_per_group * sizeof(FastVoice) +
                      (size_t)kFastWarps * a.voices_per_group * kFastMaxSub * (sizeof(double2) + sizeof(float2));
        if (smem > 48 * 1024)
        {
            SB_CUDA_CHECK(cudaFuncSetAttribute(bank_fast_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            SB_CUDA_CHECK(cudaFuncSetAttribute(bank_fast_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            SB_CUDA_CHECK(cudaFuncSetAttribute(bank_fast_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        }
        const int64_t nctas = (int64_t)((a.fast_nspans + kFastWarps - 1) / kFastWarps) * a.ngroups;
        SB_REQUIRE(nctas < 0x7fffffff, "sb_bank_render_f32: too many (group, span) units");
        const dim3 grid0((unsigned)((a.fast_nspans + kFastWarps - 1) / kFastWarps), (unsigned)a.ngroups);
        const unsigned grid = (unsigned)nctas;
        if (run0) {
            if (g_prof_start) SB_CUDA_CHECK(cudaEventRecord(g_prof_start, s));
            bank_fast_kernel<0><<<grid0, kFastWarps * 32, smem, s>>>(a);
            SB_LAUNCH_CHECK();
            if (g_prof_stop) SB_CUDA_CHECK(cudaEventRecord(g_prof_stop, s));
        }
        if (plan->edged_voices != 0) {
            bank_fast_kernel<1><<<grid, kFastWarps * 32, smem, s>>>(a);
            SB_LAUNCH_CHECK();
        }
        if (plan->scan_voices != 0) {
            bank_fast_kernel<2><<<grid, kFastWarps * 32, smem, s>>>(a);
            SB_LAUNCH_CHECK();
        }
    }
    if (plan->generic_voices != 0) {
        if (a.carry) {
            int64_t nunits = (int64_t)a.nvoices * (a.nprefix + a.nspans);
            int64_t blocks = (nunits + kGenWarps - 1) / kGenWarps;
            SB_REQUIRE(blocks < 0x7fffffff, "sb_bank_render_f32: too many (voice, span) units");
            bank_carry_kernel<<<(unsigned)blocks, kGenWarps * 32, tb_bytes, s>>>(a);
            SB_LAUNCH_CHECK();
        }
        size_t smem = tb_bytes + (size_t)a.voices_per_group * sizeof(sb_voice) +
                      (size_t)kGenWarps * a.voices_per_group * sizeof(double) +
                      (size_t)a.voices_per_group * sizeof(int);
        dim3 grid((unsigned)((a.nspans + kGenWarps - 1) / kGenWarps), (unsigned)a.ngroups);
        bank_generic_kernel<<<grid, kGenWarps * 32, smem, s>>>(a);
        SB_LAUNCH_CHECK();
    }
    if (a.ngroups > 1) {
        int64_t nfloats = a.frames * 2;
        bank_fold_kernel<<<grid_for(nfloats / 4 + 1, 256, 148 * 8), 256, 0, s>>>(partial_dev, a.ngroups,
                                                                               nfloats, a.out_stride, mix_dev);
        SB_LAUNCH_CHECK();
    }
    return SB_OK;
}

extern "C" int sb_bank_render_check(const sb_bank_plan *plan, const void *scratch_dev, int *generic_needed, void *stream)
{
    SB_REQUIRE(plan && scratch_dev, "sb_bank_render_check: null pointer");
    int words[2] = {0, 0};
    const unsigned char *scratch = (const unsigned char *)scratch_dev;
    SB_CUDA_CHECK(cudaMemcpyAsync(words, scratch + bank_off_scanlist(plan), sizeof(words), cudaMemcpyDeviceToHost,
                                  (cudaStream_t)stream));
    SB_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
    if (generic_needed) *generic_needed = words[1];
    if (plan->generic_voices == 0 && words[1] != 0) {
        set_error("sb_bank_render_f32 left %d voice(s) out of the mix: the plan's hint said no voice needs the generic "
                  "kernels (generic_voices = 0), but %d do (class-0 voices, or class-2 voices beyond scan_voices)",
                  words[1], words[1]);
        return SB_EINVAL;
    }
    return SB_OK;
}
