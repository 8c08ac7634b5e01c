// The fused hot path: a bank of FM voices with ADSR, mixed to stereo.
//
// Decomposition: work unit = (group of voices) x (span of consecutive frames), one WARP per
// unit.  A warp walks its span in sub-tiles of 256 frames (8 consecutive frames per lane) and,
// for every voice of its group, evaluates modulator -> weighted prefix sum (the FM recurrence
// in closed form) -> carrier -> ADSR -> pan, accumulating the voice sum in registers.  The
// only cross-lane step is one fp64 warp scan per voice per sub-tile; warps never synchronise
// with each other.  The running FM sum of each voice lives in shared memory (one fp64 per
// voice per warp) and starts from a per-(voice, span) carry computed by a pre-pass.
#include "sb_common.cuh"

namespace sb {

constexpr int kWarpsPerCta = 8;
constexpr int kItems = 8;
constexpr int kSub = 32 * kItems;            // frames per warp step
constexpr int kMaxVoicesPerGroup = 64;

struct BankArgs {
    const sb_voice *voices;
    int nvoices, voices_per_group, ngroups, nspans;
    int64_t frames, span_len, n0;
    const sb_time_seg *tb_rad;
    int nsegs_rad;
    const sb_time_seg *tb_cyc;
    int nsegs_cyc;
    double *carry;        // [nvoices][nspans] per-span LFO sums (pre-pass), or null
    float *out;           // [ngroups][stride] partial mixes (or the mix itself if ngroups==1)
    int64_t out_stride;   // floats between groups: 2*frames rounded up to a multiple of 4
};

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

// modulator value at fp64 phase; amplitude/bias applied in fp32
__device__ __forceinline__ float lfo_value(const sb_voice &V, double t)
{
    double ph = __dadd_rn(__dmul_rn(t, V.m_freq), phase0_of(V.m_wave, V.m_phase));
    double pw = (V.m_wave == SB_WAVE_PULSE) ? clamp_pulsewidth((double)V.m_pw) : 0.0;
    float s = wave_shape(V.m_wave, ph, pw);
    return __fadd_rn(__fmul_rn(s, V.m_amp), V.m_bias);
}

// ---------------------------------------------------------------- pre-pass: per-span LFO sums
__global__ void __launch_bounds__(kWarpsPerCta * 32)
bank_carry_kernel(BankArgs a)
{
    extern __shared__ __align__(16) unsigned char smem[];
    sb_time_seg *s_rad = reinterpret_cast<sb_time_seg *>(smem);
    sb_time_seg *s_cyc = s_rad + a.nsegs_rad;
    for (int i = threadIdx.x; i < a.nsegs_rad; i += blockDim.x) s_rad[i] = a.tb_rad[i];
    for (int i = threadIdx.x; i < a.nsegs_cyc; i += blockDim.x) s_cyc[i] = a.tb_cyc[i];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int64_t unit = ((int64_t)blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5));
    const int64_t nunits = (int64_t)a.nvoices * a.nspans;
    if (unit >= nunits) return;
    const int v = (int)(unit / a.nspans), span = (int)(unit % a.nspans);
    const sb_voice V = a.voices[v];
    double acc = 0.0;
    if (V.m_wave != SB_WAVE_NONE) {
        const bool c_rad = (V.c_wave == SB_WAVE_SINE), m_rad = (V.m_wave == SB_WAVE_SINE);
        const sb_time_seg *cs = c_rad ? s_rad : s_cyc, *ms = m_rad ? s_rad : s_cyc;
        const int cn = c_rad ? a.nsegs_rad : a.nsegs_cyc, mn = m_rad ? a.nsegs_rad : a.nsegs_cyc;
        int64_t begin = (int64_t)span * a.span_len, end = begin + a.span_len;
        if (end > a.frames) end = a.frames;
        TimeCursor cc, mc;
        cc.seek(cs, cn, a.n0 + begin + lane);
        mc.seek(ms, mn, a.n0 + begin + lane);
        for (int64_t i = begin + lane; i < end; i += 32) {
            int64_t n = a.n0 + i;
            double t_n = cc.at(n);                               // cursor only moves forward
            double step = __dadd_rn(cc.at(n + 1), -t_n);         // exact: both on the table
            float m = lfo_value(V, mc.at(n));
            acc += __dmul_rn((double)m, step);
        }
    }
    acc = warp_sum(acc);
    if (lane == 0) a.carry[(int64_t)v * a.nspans + span] = acc;
}

// ---------------------------------------------------------------- main kernel
__global__ void __launch_bounds__(kWarpsPerCta * 32)
bank_render_kernel(BankArgs a)
{
    extern __shared__ __align__(16) unsigned char smem[];
    sb_time_seg *s_rad = reinterpret_cast<sb_time_seg *>(smem);
    sb_time_seg *s_cyc = s_rad + a.nsegs_rad;
    sb_voice *s_voices = reinterpret_cast<sb_voice *>(s_cyc + a.nsegs_cyc);
    double *s_carry = reinterpret_cast<double *>(s_voices + a.voices_per_group);

    const int group = blockIdx.y;
    const int v_begin = group * a.voices_per_group;
    int nv = a.nvoices - v_begin;
    if (nv > a.voices_per_group) nv = a.voices_per_group;

    for (int i = threadIdx.x; i < a.nsegs_rad; i += blockDim.x) s_rad[i] = a.tb_rad[i];
    for (int i = threadIdx.x; i < a.nsegs_cyc; i += blockDim.x) s_cyc[i] = a.tb_cyc[i];
    {
        const int words = (int)(sizeof(sb_voice) / 4);
        const uint32_t *src = reinterpret_cast<const uint32_t *>(a.voices + v_begin);
        uint32_t *dst = reinterpret_cast<uint32_t *>(s_voices);
        for (int i = threadIdx.x; i < nv * words; i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int span = blockIdx.x * kWarpsPerCta + warp;
    if (span >= a.nspans) return;                      // no block-wide sync after this point
    double *carry = s_carry + warp * a.voices_per_group;

    // FM running sums at the start of this span: sum of the earlier spans' totals
    for (int v = 0; v < nv; v++) {
        double c = 0.0;
        if (a.carry && s_voices[v].m_wave != SB_WAVE_NONE) {
            const double *row = a.carry + (int64_t)(v_begin + v) * a.nspans;
            for (int s = lane; s < span; s += 32) c += row[s];
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
                    float m = lfo_value(V, run_at(mr, ms, mn, n, j));
                    w[j] = __dmul_rn((double)m, __dadd_rn(t_next, -t_prev));
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

        // 8 frames x 2 channels = 16 consecutive floats per lane
        if (i0 + kItems <= span_end) {
            float4 *dst = reinterpret_cast<float4 *>(out + i0 * 2);
#pragma unroll
            for (int q = 0; q < kItems / 2; q++)
                dst[q] = make_float4(accL[2 * q], accR[2 * q], accL[2 * q + 1], accR[2 * q + 1]);
        } else {
#pragma unroll
            for (int j = 0; j < kItems; j++)
                if (i0 + j < span_end) {
                    out[(i0 + j) * 2] = accL[j];
                    out[(i0 + j) * 2 + 1] = accR[j];
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
        for (int g = 1; g < ngroups; g++) {
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

}  // namespace sb

using namespace sb;

extern "C" int sb_bank_plan_make(int nvoices, int64_t frames, int sm_count, sb_bank_plan *plan)
{
    SB_REQUIRE(plan && nvoices >= 1 && frames >= 1, "sb_bank_plan_make: bad arguments");
    if (sm_count <= 0) sm_count = 148;
    // target: several waves of warps per SM (each warp is an independent unit, so many short
    // units balance the tail), but spans of at least 4 sub-tiles so that the per-span set-up
    // (summing the carry row) stays ~1% of the work.
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
    plan->nvoices = nvoices;
    plan->voices_per_group = vpg;
    plan->ngroups = ngroups;
    plan->nspans = (int32_t)nspans;
    plan->frames = frames;
    plan->span_len = span_len;
    plan->carry_bytes = (int64_t)nvoices * nspans * (int64_t)sizeof(double);
    plan->partial_bytes = (ngroups > 1) ? (int64_t)ngroups * ((frames * 2 + 3) / 4 * 4) * (int64_t)sizeof(float) : 0;
    return SB_OK;
}

extern "C" int sb_bank_render_f32(const sb_voice *voices_dev, const sb_bank_plan *plan,
                                  const sb_time_seg *tb_rad_dev, int nsegs_rad,
                                  const sb_time_seg *tb_cyc_dev, int nsegs_cyc,
                                  int64_t n0, double samplerate, double *carry_dev, float *partial_dev,
                                  float *mix_dev, void *stream)
{
    (void)samplerate;
    SB_REQUIRE(voices_dev && plan && tb_rad_dev && tb_cyc_dev && mix_dev, "sb_bank_render_f32: null pointer");
    SB_REQUIRE(nsegs_rad > 0 && nsegs_rad <= kMaxSegs && nsegs_cyc > 0 && nsegs_cyc <= kMaxSegs,
               "sb_bank_render_f32: bad time base");
    SB_REQUIRE(plan->voices_per_group >= 1 && plan->voices_per_group <= kMaxVoicesPerGroup &&
                   plan->span_len % kSub == 0 && plan->nspans >= 1 && plan->ngroups >= 1 && n0 >= 0,
               "sb_bank_render_f32: bad plan");
    SB_REQUIRE(plan->nspans == 1 || carry_dev, "sb_bank_render_f32: carry scratch missing");
    SB_REQUIRE(plan->ngroups == 1 || partial_dev, "sb_bank_render_f32: partial scratch missing");
    cudaStream_t s = (cudaStream_t)stream;
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
    a.carry = (plan->nspans > 1) ? carry_dev : nullptr;
    a.out = (plan->ngroups > 1) ? partial_dev : mix_dev;
    a.out_stride = (plan->frames * 2 + 3) / 4 * 4;

    size_t tb_bytes = (size_t)(nsegs_rad + nsegs_cyc) * sizeof(sb_time_seg);
    if (a.carry) {
        int64_t nunits = (int64_t)a.nvoices * a.nspans;
        int64_t blocks = (nunits + kWarpsPerCta - 1) / kWarpsPerCta;
        SB_REQUIRE(blocks < 0x7fffffff, "sb_bank_render_f32: too many (voice, span) units");
        bank_carry_kernel<<<(unsigned)blocks, kWarpsPerCta * 32, tb_bytes, s>>>(a);
        SB_LAUNCH_CHECK();
    }
    size_t smem = tb_bytes + (size_t)a.voices_per_group * sizeof(sb_voice) +
                  (size_t)kWarpsPerCta * a.voices_per_group * sizeof(double);
    dim3 grid((unsigned)((a.nspans + kWarpsPerCta - 1) / kWarpsPerCta), (unsigned)a.ngroups);
    SB_REQUIRE(a.ngroups <= 65535, "sb_bank_render_f32: too many voice groups");
    bank_render_kernel<<<grid, kWarpsPerCta * 32, smem, s>>>(a);
    SB_LAUNCH_CHECK();
    if (a.ngroups > 1) {
        int64_t nfloats = a.frames * 2;
        bank_fold_kernel<<<grid_for(nfloats / 4 + 1, 256, 148 * 8), 256, 0, s>>>(partial_dev, a.ngroups,
                                                                               nfloats, a.out_stride, mix_dev);
        SB_LAUNCH_CHECK();
    }
    return SB_OK;
}
