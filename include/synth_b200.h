/*
 * synth_b200.h — C ABI of libsynth_b200.so: the B200 (sm_100a) render backend for the
 * data-parallel path of irmen/synthesizer's `synthplayer` package.
 *
 * Every entry point takes plain pointers and sizes; no torch types.  Pointers named *_dev are
 * device pointers, everything else is host memory.  `stream` is a cudaStream_t passed as
 * void* (NULL = the legacy default stream).  All functions return 0 on success or a negative
 * SB_E* code; sb_last_error() gives the message for the calling thread.
 *
 * REFERENCE CITATIONS.  /root/reference holds only README.md (lines 1-2: "removed from
 * Github, now hosted at Codeberg"); the synthplayer sources are absent offline, so no
 * file:line can be cited.  Each entry point instead names the upstream class/method it stands
 * in for, as published in the `synthplayer` package (modules oscillators.py, sample.py,
 * synth.py), and — where the arithmetic is CPython's stdlib audioop (Modules/audioop.c,
 * CPython 3.12.3) — the audioop routine.  See DESIGN.md "Parity status".
 */
#ifndef SYNTH_B200_H
#define SYNTH_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SB_OK 0
#define SB_EINVAL (-1)     /* bad argument */
#define SB_ECUDA (-2)      /* CUDA runtime error, see sb_last_error() */
#define SB_ENOMEM (-3)
#define SB_ERANGE (-4)     /* a size exceeds a documented limit */

/* waveform ids (synthplayer.oscillators class names) */
#define SB_WAVE_NONE (-1)
#define SB_WAVE_SINE 0
#define SB_WAVE_TRIANGLE 1
#define SB_WAVE_SQUARE 2
#define SB_WAVE_SAWTOOTH 3
#define SB_WAVE_PULSE 4
/* general render path only (sb_osc_render_f32); not admitted as bank voices */
#define SB_WAVE_SEMICIRCLE 5
#define SB_WAVE_POINTY 6

int sb_abi_version(void);
const char *sb_last_error(void);
/* number of kernel launches made by this library in this process (bench.py's gpu_launches) */
int64_t sb_launch_count(void);

/* ------------------------------------------------------------------ time base ---------
 * Upstream oscillators advance `t += increment` serially in fp64 (increment = 2*pi/sr for
 * Sine, 1/sr for the others, and EnvelopeFilter's `time`).  That sum drifts in a way that is
 * fully determined by IEEE rounding: inside one binade the rounded step is constant.  The time
 * base is the list of (n_start, t_start, step) segments such that the serially accumulated
 * value after n additions is EXACTLY  t_start + (n - n_start) * step  for the segment holding
 * n.  Kernels evaluate it per sample index, so a render is parallel and still bit-identical
 * to the serial accumulation.  (stands in for: the `t += increment` loops of
 * synthplayer.oscillators.*.blocks and EnvelopeFilter.generator)
 */
typedef struct sb_time_seg {
    int64_t n_start;
    double t_start;
    double step;
} sb_time_seg;

#define SB_TIMEBASE_MAX_SEGS 256

/* Fills segs[] for n in [0, max_n]; returns the number of segments or a negative error. */
int sb_timebase_build(double increment, int64_t max_n, sb_time_seg *segs, int max_segs);
/* The same for an accumulation that starts at t0 >= 0 instead of 0 (upstream's Fast*
 * oscillators start their t at phase*2*pi resp. phase/frequency). */
int sb_timebase_build_from(double t0, double increment, int64_t max_n, sb_time_seg *segs, int max_segs);
/* The accumulated value after n additions (host helper; same formula as the kernels). */
double sb_timebase_at(const sb_time_seg *segs, int nsegs, int64_t n);
/* Smallest n >= n_from with t(n) >= limit, i.e. where `while time < limit` stops. */
int64_t sb_timebase_first_ge(const sb_time_seg *segs, int nsegs, int64_t n_from, double limit);

/* ADSR boundaries of synthplayer.oscillators.EnvelopeFilter.generator, as sample indices:
 * attack [0,n_a) decay [n_a,n_d) sustain [n_d,n_s) release [n_s,n_r) (+1 sample when the
 * serially accumulated release amplitude is still > 0, which `extra` reports). */
typedef struct sb_envelope {
    int64_t n_a, n_d, n_s, n_r;
    int32_t extra;          /* 1: one more sample at index n_r with amplitude amp_extra */
    int32_t enabled;        /* 0: no envelope (factor 1 everywhere) */
    double a_slope, d_slope, r_slope;   /* amp_change per sample in each ramp */
    double sustain_level;
    double amp_extra;
} sb_envelope;

int sb_envelope_build(double samplerate, double attack, double decay, double sustain,
                      double sustain_level, double release, sb_envelope *out);

/* ------------------------------------------------------------------ oscillators -------
 * sb_osc_render_f32: samples [n0, n0+count) of one oscillator into out_dev (float32).
 * stands in for synthplayer.oscillators.{Sine,Triangle,Square,Sawtooth,Pulse}.blocks.
 *   fmsum_dev : NULL, or the exclusive weighted prefix sum S[i] = sum_{k<n0+i} lfo[k]*step[k]
 *               of the FM LFO (from sb_fm_prefix_f64), which is what upstream's
 *               `phase_correction += (freq_previous-freq)*t` recurrence amounts to.
 *   pwm_dev   : NULL, or the pulse-width LFO values for the same range (Pulse only).
 *   tb_dev    : device copy of the time base for this waveform's increment.
 */
typedef struct sb_osc_params {
    int32_t wave;
    int32_t reserved;
    double frequency, amplitude, phase, bias, pulsewidth, samplerate;
} sb_osc_params;

int sb_osc_render_f32(const sb_osc_params *p, const sb_time_seg *tb_dev, int nsegs,
                      int64_t n0, int64_t count, const double *fmsum_dev, const float *pwm_dev,
                      float *out_dev, void *stream);
/* The same oscillator with its values in fp64 (shape, amplitude and bias as upstream's Python
 * floats): for an oscillator rendered as another one's fm_lfo.  The modulator's values are summed
 * over the whole render; rounded to float32 they leave a systematic phase error in the carrier
 * (measured: 200 - 1200 LSB after 60 s under a slow, deep LFO that itself has an LFO). */
int sb_osc_render_f64(const sb_osc_params *p, const sb_time_seg *tb_dev, int nsegs,
                      int64_t n0, int64_t count, const double *fmsum_dev, const float *pwm_dev,
                      double *out_dev, void *stream);

/* synthplayer.oscillators.Harmonics.blocks (and SquareH / SawtoothH built on it): the sum over
 * harmonics k of sin(k*q)*amp_k with q = t*frequency + phase_correction in radians, times
 * p->amplitude plus p->bias; FM as for Sine (p->wave is ignored).  harm_n / harm_amp are host
 * arrays of nharm entries, already filtered to the harmonics below Nyquist. */
#define SB_HARMONICS_MAX 128
int sb_harmonics_render_f32(const sb_osc_params *p, const int32_t *harm_n, const float *harm_amp,
                            int nharm, const sb_time_seg *tb_dev, int nsegs, int64_t n0,
                            int64_t count, const double *fmsum_dev, float *out_dev, void *stream);

/* Exclusive prefix sum of lfo[i]*step(n0+i) in fp64 with a running carry (carry_dev: one
 * device double, read at entry, advanced at exit) so that consecutive chunks continue the
 * sum.  scratch_dev must hold sb_fm_prefix_scratch_bytes(count) bytes. */
int64_t sb_fm_prefix_scratch_bytes(int64_t count);
int sb_fm_prefix_f64(const float *lfo_dev, const sb_time_seg *tb_dev, int nsegs, int64_t n0,
                     int64_t count, double *carry_dev, double *out_dev, void *scratch_dev,
                     void *stream);
/* ... of fp64 modulator values (sb_osc_render_f64) */
int sb_fm_prefix_f64_from_f64(const double *lfo_dev, const sb_time_seg *tb_dev, int nsegs, int64_t n0,
                              int64_t count, double *carry_dev, double *out_dev, void *scratch_dev,
                              void *stream);

/* sb_fm_prefix_f64_noalloc — no counterpart in the reference; an addition of this library.
 *
 * The same FM prefix sum as sb_fm_prefix_f64, for callers that are not allowed to allocate
 * device memory at all: a real-time audio callback (cudaMalloc/cudaFree synchronise the device
 * and can stall the audio thread for milliseconds), a plugin host or engine that hands each
 * component a fixed memory budget up front, or a process whose allocator already owns the
 * whole GPU.  Such callers cannot provide the scratch buffer sb_fm_prefix_f64 asks for, so this
 * variant keeps its scratch inside the library: a fixed 64 KiB array that is part of the
 * loaded module on every GPU and lives until the process exits.  It allocates nothing — no
 * cudaMalloc, no cudaFree, no library routine that would.  Inputs and outputs are still yours.
 *
 *  - Limit: count <= sb_fm_prefix_noalloc_max_count() (16,777,216 samples = 8192 tiles of
 *    2048).  A larger count is refused with SB_ERANGE and a message; split the range into
 *    consecutive calls (the carry continues the sum) or use sb_fm_prefix_f64.
 *  - The first call on each device looks up the device address of the built-in array (a cheap
 *    one-time runtime query, made with that device current).  Make that first call during
 *    set-up, before you begin capturing a stream.
 *  - All calls on a device share the one array, and it keeps whatever the previous call left
 *    in it.  That is harmless for a single call (every slot is written before it is read,
 *    nothing needs to start at zero), but two calls whose kernels overlap on the GPU — on
 *    different streams, or from different threads — would overwrite each other's tile sums.
 *    Keep calls to this entry point on one stream, or order them yourself.
 */
int64_t sb_fm_prefix_noalloc_max_count(void);
int sb_fm_prefix_f64_noalloc(const float *lfo_dev, const sb_time_seg *tb_dev, int nsegs, int64_t n0,
                             int64_t count, double *carry_dev, double *out_dev, void *stream);

/* synthplayer.oscillators.WhiteNoise.blocks: value = (-amp + 2*amp*u) + bias, each draw held
 * for `cycles` = int(samplerate/frequency) samples.  u_dev[j] is draw number j (j counted
 * from draw0); parity tests inject the same draws on both sides. */
int sb_noise_render_f32(const float *u_dev, int64_t draw0, int64_t cycles, double amplitude,
                        double bias, int64_t n0, int64_t count, float *out_dev, void *stream);
/* synthplayer.oscillators.Linear: upstream's serial value = clamp(value + increment): the first
 * value is `start` itself, unclamped; the following ones ramp from clamp(start + increment). */
int sb_linear_render_f32(double start, double increment, double min_value, double max_value,
                         int64_t n0, int64_t count, float *out_dev, void *stream);
/* ... its values in fp64 (a Linear as an FM modulator: a pitch slide; see sb_osc_render_f64) */
int sb_linear_render_f64(double start, double increment, double min_value, double max_value,
                         int64_t n0, int64_t count, double *out_dev, void *stream);
/* Counter-based uniform draws (Philox-4x32-10) for product use when no buffer is injected. */
int sb_uniform_f32(uint64_t seed, int64_t first_draw, int64_t count, float *out_dev, void *stream);

/* ------------------------------------------------------------------ filters -----------
 * All work on float32 device buffers holding samples [n0, n0+count) of their sources. */

/* synthplayer.oscillators.EnvelopeFilter.generator */
int sb_envelope_apply_f32(const float *in_dev, const sb_envelope *env, int64_t n0, int64_t count,
                          float *out_dev, void *stream);
/* ... on fp64 values (an enveloped LFO as an FM modulator: vibrato that fades in) */
int sb_envelope_apply_f64(const double *in_dev, const sb_envelope *env, int64_t n0, int64_t count,
                          double *out_dev, void *stream);
/* AmpModulationFilter: out = a*b */
int sb_mul_f32(const float *a_dev, const float *b_dev, int64_t count, float *out_dev, void *stream);
/* MixingFilter: out = sum of nsrc buffers (src_ptrs_dev: device array of device pointers) */
int sb_mix_f32(const float *const *src_ptrs_dev, int nsrc, int64_t count, float *out_dev, void *stream);
/* ClipFilter / AbsFilter / scaling */
int sb_clip_f32(const float *in_dev, float lo, float hi, int64_t count, float *out_dev, void *stream);
int sb_abs_f32(const float *in_dev, int64_t count, float *out_dev, void *stream);
int sb_scale_bias_f32(const float *in_dev, float scale, float bias, int64_t count, float *out_dev, void *stream);
/* One channel of an interleaved float buffer (the mono mix out of sb_bank_render_f32 when a
 * MixingFilter is rendered through the voice bank). */
int sb_channel_f32(const float *in_dev, int nchannels, int channel, int64_t frames, float *out_dev,
                   void *stream);
/* EchoFilter: out[n] = src[n] + sum_k gain[k]*src[n - delay[k]] for n >= n_after, taking
 * src[m] = 0 for m < n_after in the delayed copies; out[n] = src[n] before n_after.
 * src_dev[0] is sample index src_n0 (history the caller keeps); out covers [n0, n0+count). */
#define SB_ECHO_MAX 64
int sb_echo_f32(const float *src_dev, int64_t src_n0, int64_t n_after, int ntaps,
                const int64_t *delays, const float *gains, int64_t n0, int64_t count,
                float *out_dev, void *stream);

/* float -> PCM (WaveSynth / Sample.from_osc_block): int(v*scale), truncation toward zero.
 * Values outside the integer range are counted into *overflow_dev (upstream raises
 * OverflowError; the host wrapper does too) and saturated in the output. */
int sb_f32_to_pcm16(const float *in_dev, int64_t count, float scale, int16_t *out_dev,
                    unsigned long long *overflow_dev, void *stream);
int sb_f32_to_pcm32(const float *in_dev, int64_t count, double scale, int32_t *out_dev,
                    unsigned long long *overflow_dev, void *stream);
/* two mono float mixes -> interleaved stereo int16 */
int sb_f32x2_to_pcm16_stereo(const float *left_dev, const float *right_dev, int64_t frames,
                             float scale, int16_t *out_dev, unsigned long long *overflow_dev,
                             void *stream);

/* ------------------------------------------------------------------ voice bank --------
 * The fused hot path: a bank of voices, each
 *     EnvelopeFilter( Carrier(fm_lfo = Modulator) ) * (gain_left, gain_right)
 * summed over voices into a stereo float mix (MixingFilter per channel).  One kernel
 * evaluates modulator, FM prefix, carrier, ADSR and the voice sum; a second one folds the
 * per-group partial mixes.  (stands in for: Sine/.../Pulse.blocks with fm_lfo,
 * EnvelopeFilter.generator, AmpModulationFilter and MixingFilter, composed)
 */
typedef struct sb_voice {
    int32_t c_wave, m_wave;              /* m_wave = SB_WAVE_NONE: no FM */
    double c_freq, c_phase;              /* phase as passed upstream (fraction of a cycle) */
    double m_freq, m_phase;
    /* Everything that decides WHERE an edge falls, or that is summed over the whole render,
     * is fp64 as upstream has it: the modulator's amplitude and bias (they scale the FM sum) and
     * both pulse widths (a float32 width moves an LFO edge by one sample whenever the phase
     * hits the width exactly, e.g. 5.3 Hz, phase 0.1, width 0.4 at t = 1 s).  The carrier's
     * amplitude, bias and the pan gains only scale the output value: float32. */
    double m_amp, m_bias;
    double c_pw, m_pw;
    float c_amp, c_bias;
    float gain_l, gain_r;
    sb_envelope env;
} sb_voice;

typedef struct sb_bank_plan {
    int32_t nvoices, voices_per_group, ngroups, nspans;
    int64_t frames, span_len;            /* generic path: span_len is a multiple of 256 */
    int64_t fast_span_len;               /* fast path: a multiple of 256, at most 8*256 */
    int32_t fast_nspans;
    int32_t nsegs_rad;                   /* width of the closed-form FM table */
    int32_t generic_voices;              /* hint: 0 = every voice is fast (skip the generic kernels),
                                            -1 = unknown (run them; they exit early per group) */
    int32_t nprefix;                     /* generic path: spans covering [0, n0), see below */
    int32_t edged_voices;                /* hint: 0 = no fast voice has a Square/Sawtooth/Pulse carrier
                                            (skip that kernel), -1 = unknown */
    int32_t scan_voices;                 /* hint: 0 = no voice has a Triangle/Square/Sawtooth/Pulse modulator on the
                                            fast path (skip their pre-pass), -1 = unknown (run it over every voice) */
    int64_t n0;                          /* first frame of the render the plan was made for */
    int64_t off_fastv, off_tab;          /* layout of the scratch buffer */
    int64_t carry_bytes;                 /* size of scratch_dev: span sums + per-voice constants + FM tables */
    int64_t partial_bytes;               /* size of partial_dev: ngroups * (2*frames rounded up to 4) floats */
    int32_t smooth_voices;               /* hint: 0 = no class-1 voice has a Sine/Triangle carrier (skip that kernel), -1 = unknown */
} sb_bank_plan;

/* Chooses the (voice group) x (time span) decomposition for `nvoices` x `frames` starting at
 * frame n0.  nsegs_rad = segment count of the 2*pi/samplerate time base.
 * A render that starts at n0 > 0 is exact for every voice: fast voices evaluate their FM sum
 * in closed form at any frame; generic FM voices get theirs from a pre-pass that also covers
 * [0, n0) (cost proportional to n0 for those voices only). */
int sb_bank_plan_make(int nvoices, int64_t frames, int64_t n0, int sm_count, int nsegs_rad, sb_bank_plan *plan);

/* Optional, between sb_bank_plan_make and allocating scratch_dev: how many voices of the bank
 * are generic (sb_voice_is_fast == 0), of class 1 with a Sine/Triangle carrier (smooth), of class 1
 * with a Square/Sawtooth/Pulse carrier (edged), and of class 2 (scan); -1 = unknown;
 * sb_bank_count_classes() counts them.  Kernels no voice needs are not launched, and
 * plan->carry_bytes shrinks to what the class-2 voices need (without the hint it is sized as if
 * every voice were one: nvoices * (n0+frames)/512 * 8 bytes).  Counts must not be too small:
 * a voice whose kernel was hinted away is not rendered, and a class-2 voice beyond scan_voices is
 * handed to the generic kernels, which a generic_voices of 0 would have skipped
 * (sb_bank_render_check reports the latter). */
int sb_bank_plan_hint(sb_bank_plan *plan, int generic_voices, int smooth_voices, int edged_voices, int scan_voices);
/* counts4 = {generic, smooth, edged, scan} of a HOST array of voices, as sb_bank_plan_hint takes them */
int sb_bank_count_classes(const sb_voice *voices, int nvoices, double samplerate, int *counts4);

/* Which path the voice takes:
 *   1  fast: any of the five carriers, with no FM or a Sine modulator of moderate gain: fp32
 *      per-sample arithmetic off one fp64 base per 16 samples, SFU sine, closed-form FM sums;
 *      Square/Sawtooth/Pulse carriers re-evaluate in exact fp64 the few samples whose phase is
 *      within 2e-5 cycles of an edge, so edges fall where upstream puts them;
 *   2  fast with a Triangle/Square/Sawtooth/Pulse modulator: the same, with the FM sum taken
 *      from a pre-pass (exact fp64 sums per 512 frames) plus a warp scan inside them;
 *   0  generic (stronger FM, negative phases or frequencies on edged carriers or modulators):
 *      per-sample fp64 phase. */
int sb_voice_is_fast(const sb_voice *voice, double samplerate);

/* Renders frames [n0, n0+plan->frames) of all voices.  tb_rad_dev / tb_cyc_dev: time bases
 * for increment 2*pi/sr and 1/sr.  scratch_dev (plan->carry_bytes) and partial_dev
 * (plan->partial_bytes, may be NULL when ngroups == 1) are caller-provided scratch.
 * mix_dev receives plan->frames interleaved (left,right) float pairs. */
int sb_bank_render_f32(const sb_voice *voices_dev, const sb_bank_plan *plan,
                       const sb_time_seg *tb_rad_dev, int nsegs_rad,
                       const sb_time_seg *tb_cyc_dev, int nsegs_cyc,
                       int64_t n0, double samplerate, void *scratch_dev, float *partial_dev,
                       float *mix_dev, void *stream);
/* After sb_bank_render_f32 on the same stream: waits for it and reads back how many voices were
 * left to the generic kernels (class 0, or class-2 voices beyond the scan_voices hint) into
 * *generic_needed (may be NULL).  Returns SB_EINVAL when the plan's hint said generic_voices = 0
 * and that count is not zero: those voices are missing from the mix (a hint was too small).
 * Callers that pass exact hints (sb_host_bank_render_pcm16, the Python VoiceBank) need not call it. */
int sb_bank_render_check(const sb_bank_plan *plan, const void *scratch_dev, int *generic_needed, void *stream);

/* Measurement hook (no upstream counterpart): from now on every sb_bank_render_f32 records the two
 * CUDA events (cudaEvent_t, made by the caller with timing enabled) on its stream right before
 * and right after the launch of the smooth-carrier kernel, bank_fast_kernel<0>; pass NULL, NULL to
 * stop.  bench.py times the dominant kernel alone with them. */
int sb_bank_profile_events(void *start_event, void *stop_event);
/* CUDA events (timing enabled) for callers without a CUDA runtime binding of their own */
int sb_event_create(void **event);
int sb_event_elapsed_ms(void *start_event, void *stop_event, float *ms);
int sb_event_destroy(void *event);

/* ------------------------------------------------------------------ integer PCM -------
 * Bit-exact restatements of the CPython audioop routines synthplayer's Sample calls. */

/* audioop.add (Sample.mix): saturating add */
int sb_pcm16_add(const int16_t *a_dev, const int16_t *b_dev, int64_t count, int16_t *out_dev, void *stream);
/* audioop.mul (Sample.amplify): floor(v*factor) clipped (audioop's fbound) */
int sb_pcm16_mul(const int16_t *in_dev, int64_t count, double factor, int16_t *out_dev, void *stream);
/* audioop.bias (Sample.bias): wrapping add */
int sb_pcm16_bias(const int16_t *in_dev, int64_t count, int bias, int16_t *out_dev, void *stream);
/* audioop.reverse */
int sb_pcm16_reverse(const int16_t *in_dev, int64_t count, int16_t *out_dev, void *stream);
/* audioop.max: max |v| as unsigned, into *out_dev */
int sb_pcm16_maxabs(const int16_t *in_dev, int64_t count, unsigned int *out_dev, void *stream);
/* audioop.tomono / tostereo (Sample.mono / stereo / pan) */
int sb_pcm16_tomono(const int16_t *in_dev, int64_t frames, double lfactor, double rfactor, int16_t *out_dev, void *stream);
int sb_pcm16_tostereo(const int16_t *in_dev, int64_t frames, double lfactor, double rfactor, int16_t *out_dev, void *stream);
/* audioop.lin2lin 2<->4 (Sample.make_32bit / make_16bit) */
int sb_pcm16_to_pcm32(const int16_t *in_dev, int64_t count, int32_t *out_dev, void *stream);
int sb_pcm32_to_pcm16(const int32_t *in_dev, int64_t count, int16_t *out_dev, void *stream);
/* Sample.fadeout / fadein: int(v * (1 - k*decrease/numsamples)) resp. (k*increase/numsamples + start) */
int sb_pcm16_fadeout(int16_t *buf_dev, int64_t nsamples, double target_volume, void *stream);
int sb_pcm16_fadein(int16_t *buf_dev, int64_t nsamples, double start_volume, void *stream);
/* Sample.modulate_amp: int(v * m) */
int sb_pcm16_modulate(int16_t *buf_dev, const float *mod_dev, int64_t count, void *stream);
/* Sample.modulate_amp with a Sample / list / array modulator (upstream cycles it and divides by
   its largest magnitude in fp64): int(v * mod[i % mod_len]), factors in fp64 */
int sb_pcm16_modulate_cycle_f64(int16_t *buf_dev, const double *mod_dev, int64_t mod_len, int64_t count, void *stream);
int sb_pcm32_modulate_cycle_f64(int32_t *buf_dev, const double *mod_dev, int64_t mod_len, int64_t count, void *stream);
/* audioop.ratecv(frames, 2, nchannels, inrate, outrate, None) (Sample.resample / speed).
 * sb_pcm16_ratecv_frames gives the number of output frames. */
int64_t sb_pcm16_ratecv_frames(int64_t in_frames, int64_t inrate, int64_t outrate);
int sb_pcm16_ratecv(const int16_t *in_dev, int64_t in_frames, int nchannels, int64_t inrate,
                    int64_t outrate, int16_t *out_dev, void *stream);
/* The same for many (short) samples in ONE launch: source s is in_pool[in_off[s] ..) with
 * in_frames[s] frames at reduced rate inrate_red[s] -> outrate_red[s] (both divided by their
 * gcd, below 2^21), written to out_pool[out_off[s] ..) with out_frames[s] =
 * sb_pcm16_ratecv_frames(...) frames.  All arrays are device arrays of nsrc int64 (offsets in
 * int16 units); max_out_frames = the largest out_frames[s]. */
int sb_pcm16_ratecv_many(const int16_t *in_pool_dev, const int64_t *in_off_dev, const int64_t *in_frames_dev,
                         const int64_t *inrate_red_dev, const int64_t *outrate_red_dev,
                         const int64_t *out_off_dev, const int64_t *out_frames_dev, int64_t nsrc,
                         int nchannels, int64_t max_out_frames, int16_t *out_pool_dev, void *stream);
/* Sample.mix_at for a list of sources applied IN ORDER: each step is audioop.add on the
 * overlapped span (saturation per step).  Source s is src_pool_dev[src_off[s] ..
 * src_off[s]+src_len[s]) (int16 sample units) mixed at dst sample index dst_off[s]; the
 * three arrays are device arrays of nsrc int64.  dst must already be long enough.
 * scratch_dev (16-byte aligned; sb_pcm16_mix_at_many_scratch_bytes; total_src_samples = sum of
 * src_len) holds the per-tile source lists (32 bytes per (source, tile) pair); pass NULL to run without an index (each 2048-sample tile then scans
 * the whole source list — fine for a few sources, slow for thousands).
 * Bad input cannot make the kernels write outside their buffers: a source with a negative dst_off
 * is left out, and per-tile lists that do not fit the scratch given are refused (those tiles
 * stay unmixed); the first four ints of scratch_dev then read {list cursor, sources left out,
 * 1 if a list did not fit, 0} - all zero but the cursor after a call with valid arguments. */
int64_t sb_pcm16_mix_at_many_scratch_bytes(int64_t dst_count, int64_t nsrc, int64_t total_src_samples);
int sb_pcm16_mix_at_many(int16_t *dst_dev, int64_t dst_count, const int16_t *src_pool_dev,
                         const int64_t *src_off_dev, const int64_t *src_len_dev,
                         const int64_t *dst_off_dev, int64_t nsrc, void *scratch_dev,
                         int64_t scratch_bytes, void *stream);

/* The same routines for sample width 4 (audioop with size == 4); a Sample of 32-bit frames
 * (Sample.make_32bit, WaveSynth(samplewidth=4)) uses these. */
int sb_pcm32_add(const int32_t *a_dev, const int32_t *b_dev, int64_t count, int32_t *out_dev, void *stream);
int sb_pcm32_mul(const int32_t *in_dev, int64_t count, double factor, int32_t *out_dev, void *stream);
int sb_pcm32_bias(const int32_t *in_dev, int64_t count, int bias, int32_t *out_dev, void *stream);
int sb_pcm32_reverse(const int32_t *in_dev, int64_t count, int32_t *out_dev, void *stream);
int sb_pcm32_maxabs(const int32_t *in_dev, int64_t count, unsigned int *out_dev, void *stream);
int sb_pcm32_tomono(const int32_t *in_dev, int64_t frames, double lfactor, double rfactor, int32_t *out_dev, void *stream);
int sb_pcm32_tostereo(const int32_t *in_dev, int64_t frames, double lfactor, double rfactor, int32_t *out_dev, void *stream);
int sb_pcm32_fadeout(int32_t *buf_dev, int64_t nsamples, double target_volume, void *stream);
int sb_pcm32_fadein(int32_t *buf_dev, int64_t nsamples, double start_volume, void *stream);
int sb_pcm32_modulate(int32_t *buf_dev, const float *mod_dev, int64_t count, void *stream);
int sb_pcm32_ratecv(const int32_t *in_dev, int64_t in_frames, int nchannels, int64_t inrate,
                    int64_t outrate, int32_t *out_dev, void *stream);

/* 24-bit frames (audioop with size == 3): a Sample of 3-byte frames keeps sign-extended int32
 * values on the device and runs the size-4 routines above; these supply what differs: packing,
 * the 24-bit clip of audioop's fbound / overflow, and the <<8 / >>8 of GETSAMPLE32 and lin2lin
 * (sb_pcm32_shift: shift > 0 left, < 0 arithmetic right). */
int sb_pcm24_unpack(const uint8_t *bytes_dev, int64_t count, int32_t *out_dev, void *stream);
int sb_pcm24_pack(const int32_t *in_dev, int64_t count, uint8_t *bytes_dev, void *stream);
int sb_pcm24_clamp(int32_t *buf_dev, int64_t count, void *stream);
int sb_pcm32_shift(int32_t *buf_dev, int64_t count, int shift, void *stream);

/* ------------------------------------------------------------------ host-buffer calls -
 * The same path with HOST buffers (copies included), for callers without device memory of
 * their own; bench.py's e2e figure goes through these.
 * State: device workspaces (grow-only) and a pinned word are kept PER DEVICE, keyed by the device
 * that is current at the call; a process-wide mutex serialises the calls, so they may come from
 * several host threads and for several devices (their GPU work does not overlap).  The plain
 * variants run on the legacy default stream; the _on_stream ones take a cudaStream_t. */
int sb_host_bank_render_pcm16(const sb_voice *voices, int nvoices, int64_t frames,
                              double samplerate, float scale, int16_t *out_stereo,
                              unsigned long long *overflow);
int sb_host_bank_render_pcm16_on_stream(const sb_voice *voices, int nvoices, int64_t frames,
                                        double samplerate, float scale, int16_t *out_stereo,
                                        unsigned long long *overflow, void *stream);
/* The same render up to the float mix, left in DEVICE memory the caller owns (2*frames floats):
 * the per-GPU half of a voice-sharded render, whose partial mixes the caller then reduces
 * (NCCL) and converts once.  Asynchronous on `stream`; the library's workspace is in use until
 * that work has run, so make the next host call on this device on the same stream or after a
 * synchronisation. */
int sb_host_bank_render_mix_f32(const sb_voice *voices, int nvoices, int64_t frames, double samplerate,
                                float *mix_dev, void *stream);
/* Grows this device's workspaces to what a bank of nvoices x frames can need at most, so that
 * later calls of that size or smaller allocate nothing (growing frees and re-allocates, which
 * synchronises the device). */
int sb_host_reserve(int nvoices, int64_t frames, double samplerate);
int sb_host_pcm16_mix_at_many(int16_t *dst, int64_t dst_count, const int16_t *src_pool,
                              int64_t pool_count, const int64_t *src_off, const int64_t *src_len,
                              const int64_t *dst_off, int64_t nsrc);
int sb_host_pcm16_mix_at_many_on_stream(int16_t *dst, int64_t dst_count, const int16_t *src_pool,
                                        int64_t pool_count, const int64_t *src_off, const int64_t *src_len,
                                        const int64_t *dst_off, int64_t nsrc, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* SYNTH_B200_H */
