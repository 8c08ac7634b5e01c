/*
 * ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product.
 *
 * C restatement of oracle/synth_oracle.py and oracle/audioop_port.py: the same serial,
 * one-sample-at-a-time fp64 loops, compiled so that parity checks at BASELINE.json's full
 * sizes (60 s, 10 min, thousands of voices) finish in seconds and so that bench.py has a
 * CPU baseline.  Built with -O2 -ffp-contract=off (no FMA contraction) so that every value
 * is bit-identical to the pure-Python oracle (same IEEE double operations, same glibc sin).
 * tests/test_oracle_c.py checks that bit-identity.
 *
 * PARITY UNPINNED for the oscillator/filter part: the reference source is absent from
 * /root/reference (README.md:1-2 only).  Each function names the upstream
 * synthplayer class it restates (recalled, no line numbers possible).
 * The audioop part restates CPython 3.12.3 Modules/audioop.c and is pinned against the
 * real stdlib module by the tests.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may load this.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>

#define ORC_SINE 0
#define ORC_TRIANGLE 1
#define ORC_SQUARE 2
#define ORC_SAWTOOTH 3
#define ORC_PULSE 4

static double py_mod1(double x)            /* Python's  x % 1.0  */
{
    double m = fmod(x, 1.0);
    if (m != 0.0) { if (m < 0.0) m += 1.0; } else m = 0.0;
    return m;
}

static double wave_value(int wave, double ph, double amplitude, double bias, double pw)
{
    switch (wave) {
    case ORC_SINE:     return sin(ph) * amplitude + bias;
    case ORC_TRIANGLE: return 4.0 * amplitude * (fabs(py_mod1(ph + 0.75) - 0.5) - 0.25) + bias;
    case ORC_SQUARE: {
        double x = ph * 2.0;
        long long i = (long long)x;            /* int(): truncation toward zero */
        long long r = i % 2; if (r < 0) r += 2; /* Python % is non-negative for modulus 2 */
        return (r ? -amplitude : amplitude) + bias;
    }
    case ORC_SAWTOOTH: return bias + amplitude * 2.0 * (ph - floor(0.5 + ph));
    case ORC_PULSE:    return (py_mod1(ph) < pw ? amplitude : -amplitude) + bias;
    }
    return 0.0;
}

/* synthplayer.oscillators.{Sine,Triangle,Square,Sawtooth,Pulse}.blocks (recalled).
 * fm / pwm may be NULL.  Serial: t += increment; phase_correction += (freq_prev-freq)*t. */
void orc_osc(int wave, long long n, double samplerate, double frequency, double amplitude,
             double phase, double bias, const double *fm, double pulsewidth, const double *pwm,
             double *out)
{
    double phase_correction = (wave == ORC_SINE) ? phase * 2 * M_PI : phase;
    double increment = (wave == ORC_SINE) ? 2.0 * M_PI / samplerate : 1.0 / samplerate;
    double freq_previous = frequency;
    double t = 0.0;
    for (long long i = 0; i < n; i++) {
        double freq = frequency * (1.0 + (fm ? fm[i] : 0.0));
        phase_correction += (freq_previous - freq) * t;
        freq_previous = freq;
        double ph = t * freq + phase_correction;
        double pw = pulsewidth;
        if (wave == ORC_PULSE) {
            if (pwm) pw = pwm[i];
            if (pw <= 0.0) pw = DBL_EPSILON; else if (pw >= 1.0) pw = 1.0 - DBL_EPSILON;
        }
        out[i] = wave_value(wave, ph, amplitude, bias, pw);
        t += increment;
    }
}

/* synthplayer.oscillators.WhiteNoise (recalled) with injected uniform draws u[] in [0,1). */
void orc_noise(long long n, double samplerate, double frequency, double amplitude, double bias,
               const double *u, double *out)
{
    long long cycles = (long long)(samplerate / frequency);
    if (cycles < 1) cycles = 1;
    for (long long i = 0; i < n; i++) {
        double value = (-amplitude + (amplitude - (-amplitude)) * u[i / cycles]) + bias;
        out[i] = value;
    }
}

/* synthplayer.oscillators.EnvelopeFilter.generator (recalled).  Returns the number of
 * samples produced (== n unless stop_at_end cuts the stream). */
long long orc_envelope(const double *in, long long n, double samplerate, double attack, double decay,
                       double sustain, double sustain_level, double release, int stop_at_end,
                       double *out)
{
    double time = 0.0;
    double end_time_decay = attack + decay;
    double end_time_sustain = end_time_decay + sustain;
    double end_time_release = end_time_sustain + release;
    double increment = 1.0 / samplerate;
    long long i = 0;
    double amp, amp_change;
    if (attack != 0.0) {
        amp_change = 1.0 / attack * increment;
        amp = 0.0;
        while (time < attack) {
            if (i >= n) return i;
            out[i] = in[i] * amp; i++;
            amp += amp_change; time += increment;
        }
    }
    if (decay != 0.0) {
        amp = 1.0;
        amp_change = (sustain_level - 1.0) / decay * increment;
        while (time < end_time_decay) {
            if (i >= n) return i;
            out[i] = in[i] * amp; i++;
            amp += amp_change; time += increment;
        }
    }
    while (time < end_time_sustain) {
        if (i >= n) return i;
        out[i] = in[i] * sustain_level; i++;
        time += increment;
    }
    if (release != 0.0) {
        amp = sustain_level;
        amp_change = (-sustain_level) / release * increment;
        while (time < end_time_release) {
            if (i >= n) return i;
            out[i] = in[i] * amp; i++;
            amp += amp_change; time += increment;
        }
        if (amp > 0.0) {
            if (i >= n) return i;
            out[i] = in[i] * amp; i++;
        }
    }
    if (stop_at_end) return i;
    for (; i < n; i++) out[i] = 0.0;
    return n;
}

/* Python >= 3.12 builtin sum() over floats is Neumaier-compensated (Python/bltinmodule.c,
 * builtin_sum_impl).  MixingFilter and EchoFilter call sum(), so the restatement does too;
 * the effect is at the 1e-16 level, far below one PCM LSB. */
static void py_sum_step(double *f_result, double *c, double x)
{
    double t = *f_result + x;
    if (fabs(*f_result) >= fabs(x)) *c += (*f_result - t) + x;
    else *c += (x - t) + *f_result;
    *f_result = t;
}
static double py_sum_end(double f_result, double c)
{
    if (c != 0.0 && isfinite(c)) f_result += c;
    return f_result;
}

/* synthplayer.oscillators.EchoFilter.generator (recalled): dry + sum_k decay^k * delayed copy,
 * summed in the order dry, echo 1, echo 2, ... like Python's sum() of the list. */
void orc_echo(const double *in, long long n, double samplerate, double after, int amount,
              double delay, double decay, double *out)
{
    if (decay < 1) {
        double lim = log(0.000001) / log(decay);
        if ((double)amount > lim) amount = (int)lim;
    }
    long long nafter = (long long)(samplerate * after);
    long long i;
    for (i = 0; i < n && i < nafter; i++) out[i] = in[i];
    for (; i < n; i++) {
        double acc = 0.0 + in[i];  /* Python sum() starts from int 0: 0 + x */
        double comp = 0.0;
        double amp = decay, echo_delay = delay;
        for (int k = 0; k < amount; k++) {
            long long nz = (long long)(samplerate * echo_delay);
            long long m = i - nafter - nz;
            double v = (m >= 0) ? in[nafter + m] * amp : 0.0;
            py_sum_step(&acc, &comp, v);
            echo_delay += delay;
            amp *= decay;
        }
        out[i] = py_sum_end(acc, comp);
    }
}

/* Python's math.log(x, base) is log(x)/log(base): orc_echo uses the same expression. */

/* float -> PCM: int(v*scale); returns the number of out-of-range values (OverflowError upstream). */
long long orc_to_pcm16(const double *in, long long n, double scale, int16_t *out)
{
    long long bad = 0;
    for (long long i = 0; i < n; i++) {
        double x = (scale != 1.0) ? in[i] * scale : in[i];
        long long v = (long long)x;
        if (v < -32768 || v > 32767) { bad++; v = v < 0 ? -32768 : 32767; }
        out[i] = (int16_t)v;
    }
    return bad;
}

/* ---- a bank of FM voices mixed to stereo: config 3 / config 5 of BASELINE.json ----------
 * One voice = EnvelopeFilter(Carrier(fm_lfo=Modulator)) scaled into a left/right mix, i.e.
 * MixingFilter over voices evaluated per channel.  Parameters per voice (doubles):
 *  0 carrier_wave 1 carrier_freq 2 carrier_amp 3 carrier_phase 4 carrier_bias 5 pulsewidth
 *  6 mod_wave (-1 = no FM) 7 mod_freq 8 mod_amp(depth) 9 mod_phase 10 mod_bias
 *  11 attack 12 decay 13 sustain 14 sustain_level 15 release (attack<0 = no envelope)
 *  16 left_gain 17 right_gain
 * Voices are summed in index order the way Python's sum() does (Neumaier-compensated).
 */
#define ORC_VOICE_STRIDE 18
void orc_voice_bank(const double *params, long long nvoices, long long nframes, double samplerate,
                    double *mix_left, double *mix_right)
{
    double *m = (double *)malloc(sizeof(double) * (size_t)nframes);
    double *c = (double *)malloc(sizeof(double) * (size_t)nframes);
    double *e = (double *)malloc(sizeof(double) * (size_t)nframes);
    double *cl = (double *)calloc((size_t)nframes, sizeof(double));
    double *cr = (double *)calloc((size_t)nframes, sizeof(double));
    memset(mix_left, 0, sizeof(double) * (size_t)nframes);
    memset(mix_right, 0, sizeof(double) * (size_t)nframes);
    for (long long v = 0; v < nvoices; v++) {
        const double *p = params + v * ORC_VOICE_STRIDE;
        const double *fm = NULL;
        if (p[6] >= 0) {
            orc_osc((int)p[6], nframes, samplerate, p[7], p[8], p[9], p[10], NULL, 0.1, NULL, m);
            fm = m;
        }
        orc_osc((int)p[0], nframes, samplerate, p[1], p[2], p[3], p[4], fm, p[5], NULL, c);
        const double *src = c;
        if (p[11] >= 0) {
            orc_envelope(c, nframes, samplerate, p[11], p[12], p[13], p[14], p[15], 0, e);
            src = e;
        }
        for (long long i = 0; i < nframes; i++) {
            if (v == 0) {                       /* int 0 + first float */
                mix_left[i] = 0.0 + src[i] * p[16];
                mix_right[i] = 0.0 + src[i] * p[17];
            } else {
                py_sum_step(&mix_left[i], &cl[i], src[i] * p[16]);
                py_sum_step(&mix_right[i], &cr[i], src[i] * p[17]);
            }
        }
    }
    for (long long i = 0; i < nframes; i++) {
        mix_left[i] = py_sum_end(mix_left[i], cl[i]);
        mix_right[i] = py_sum_end(mix_right[i], cr[i]);
    }
    free(m); free(c); free(e); free(cl); free(cr);
}

/* ---------------------------------------------------------------- audioop (CPython 3.12.3) */

static int fbound16(double val)
{
    if (val > 32767.0) val = 32767.0;
    else if (val < -32768.0 + 1.0) val = -32768.0;
    val = floor(val);
    return (int)val;
}

void orc_pcm16_add(const int16_t *a, const int16_t *b, long long n, int16_t *out)
{
    for (long long i = 0; i < n; i++) {
        int v = (int)a[i] + (int)b[i];
        if (v > 32767) v = 32767; else if (v < -32768) v = -32768;
        out[i] = (int16_t)v;
    }
}

void orc_pcm16_mul(const int16_t *a, long long n, double factor, int16_t *out)
{
    for (long long i = 0; i < n; i++) out[i] = (int16_t)fbound16((double)a[i] * factor);
}

/* Sample.mix_at (recalled) applied for a list of (offset, source) pairs in order: each is
 * audioop.add on the overlapped span, i.e. a saturating add per step, not one wide sum. */
void orc_pcm16_mix_at_many(int16_t *dst, long long ndst, const int16_t *src_pool,
                           const long long *src_off, const long long *src_len,
                           const long long *dst_off, long long nsrc)
{
    for (long long s = 0; s < nsrc; s++) {
        const int16_t *src = src_pool + src_off[s];
        long long o = dst_off[s];
        for (long long i = 0; i < src_len[s] && o + i < ndst; i++) {
            int v = (int)dst[o + i] + (int)src[i];
            if (v > 32767) v = 32767; else if (v < -32768) v = -32768;
            dst[o + i] = (int16_t)v;
        }
    }
}

/* audioop.ratecv(frames, 2, nchannels, inrate, outrate, None)[0]; returns frames written. */
long long orc_pcm16_ratecv(const int16_t *in, long long nframes, int nchannels, long long inrate,
                           long long outrate, int16_t *out, long long out_capacity_frames)
{
    long long a = inrate, b = outrate;
    while (b) { long long t = a % b; a = b; b = t; }
    inrate /= a; outrate /= a;
    int prev_i[2] = {0, 0}, cur_i[2] = {0, 0};
    long long d = -outrate, k = 0, j = 0;
    for (;;) {
        while (d < 0) {
            if (k == nframes) return j;
            for (int c = 0; c < nchannels; c++) {
                prev_i[c] = cur_i[c];
                cur_i[c] = (int)((uint32_t)(int32_t)in[k * nchannels + c] << 16);
                cur_i[c] = (int)((1.0 * (double)cur_i[c] + 0.0 * (double)prev_i[c]) / (1.0 + 0.0));
            }
            k++;
            d += outrate;
        }
        while (d >= 0) {
            if (j >= out_capacity_frames) return -1;
            for (int c = 0; c < nchannels; c++) {
                int cur_o = (int)(((double)prev_i[c] * (double)d + (double)cur_i[c] * (double)(outrate - d))
                                  / (double)outrate);
                out[j * nchannels + c] = (int16_t)(cur_o >> 16);
            }
            j++;
            d -= inrate;
        }
    }
}

/* Sample.fadeout / fadein (recalled): int(sample * factor(k)), truncation toward zero. */
void orc_pcm16_fadeout(int16_t *a, long long nsamples, double target_volume)
{
    double numsamples = (double)nsamples, decrease = 1.0 - target_volume;
    for (long long k = 0; k < nsamples; k++)
        a[k] = (int16_t)(long long)((double)a[k] * (1.0 - (double)k * decrease / numsamples));
}

void orc_pcm16_fadein(int16_t *a, long long nsamples, double start_volume)
{
    double numsamples = (double)nsamples, increase = 1.0 - start_volume;
    for (long long k = 0; k < nsamples; k++)
        a[k] = (int16_t)(long long)((double)a[k] * ((double)k * increase / numsamples + start_volume));
}
