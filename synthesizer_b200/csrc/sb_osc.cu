// Single-oscillator and filter kernels: the general (one node per launch) render path.
// Every kernel maps one thread to a run of consecutive sample indices, so the exact time
// base cursor moves forward only, and global accesses are coalesced 128-bit where the
// element type allows.
#include <mutex>
#include "sb_common.cuh"

namespace sb {

constexpr int kBlock = 256;
constexpr int kMaxBlocks = 148 * 16;

// ------------------------------------------------------------------ oscillator
// ITEMS consecutive samples per thread: thread i of the grid owns [i*ITEMS, i*ITEMS+ITEMS).
// OUT = float: the node values of the render graph.  OUT = double: the oscillator as an FM modulator
// (sb_osc_render_f64): shape, amplitude and bias in fp64, as upstream's Python floats are.
template <int ITEMS, typename OUT>
__global__ void __launch_bounds__(kBlock)
osc_render_kernel(sb_osc_params p, const sb_time_seg *__restrict__ tb, int nsegs, int64_t n0,
                  int64_t count, const double *__restrict__ fmsum, const float *__restrict__ pwm,
                  OUT *__restrict__ out)
{
    __shared__ sb_time_seg s_tb[kMaxSegs];
    for (int i = threadIdx.x; i < nsegs; i += blockDim.x) s_tb[i] = tb[i];
    __syncthreads();

    const double pc0 = phase0_of(p.wave, p.phase);
    const float amp = (float)p.amplitude, bias = (float)p.bias;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * ITEMS;
    for (int64_t base = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * ITEMS; base < count;
         base += stride) {
        TimeCursor cur;
        cur.seek(s_tb, nsegs, n0 + base);
#pragma unroll
        for (int j = 0; j < ITEMS; j++) {
            int64_t i = base + j;
            if (i >= count) break;
            double t = cur.at(n0 + i);
            double ph;
            if (fmsum) {
                // phase = pc0 + f0*(t_n + S_n): the closed form of upstream's running
                // phase_correction (see DESIGN.md "FM as a prefix sum")
                ph = __dadd_rn(pc0, __dmul_rn(p.frequency, __dadd_rn(t, fmsum[i])));
            } else {
                ph = __dadd_rn(__dmul_rn(t, p.frequency), pc0);   // bit-identical to upstream
            }
            double pw = p.pulsewidth;
            if (p.wave == SB_WAVE_PULSE) {
                if (pwm) pw = (double)pwm[i];
                pw = clamp_pulsewidth(pw);
            }
            if constexpr (sizeof(OUT) == 8) {
                const double shape = wave_shape64(p.wave, ph, pw);
                out[i] = (p.wave == SB_WAVE_SAWTOOTH) ? __dadd_rn(p.bias, __dmul_rn(p.amplitude, shape))
                                                      : __dadd_rn(__dmul_rn(shape, p.amplitude), p.bias);
            } else {
                float shape = wave_shape(p.wave, ph, pw);
                out[i] = (p.wave == SB_WAVE_SAWTOOTH) ? fmaf(amp, shape, bias)
                                                      : __fadd_rn(__fmul_rn(shape, amp), bias);
            }
        }
    }
}

// ------------------------------------------------------------------ harmonics
struct HarmonicSet {
    int n;
    int k[SB_HARMONICS_MAX];
    float amp[SB_HARMONICS_MAX];
};

__global__ void __launch_bounds__(kBlock)
harmonics_kernel(sb_osc_params p, const __grid_constant__ HarmonicSet hs, const sb_time_seg *__restrict__ tb, int nsegs,
                 int64_t n0, int64_t count, const double *__restrict__ fmsum, float *__restrict__ out)
{
    __shared__ sb_time_seg s_tb[kMaxSegs];
    for (int i = threadIdx.x; i < nsegs; i += blockDim.x) s_tb[i] = tb[i];
    __syncthreads();
    const double pc0 = phase0_of(SB_WAVE_SINE, p.phase);
    const float amp = (float)p.amplitude, bias = (float)p.bias;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
        TimeCursor cur;
        cur.seek(s_tb, nsegs, n0 + i);
        const double t = cur.at(n0 + i);
        const double q = fmsum ? __dadd_rn(pc0, __dmul_rn(p.frequency, __dadd_rn(t, fmsum[i])))
                               : __dadd_rn(__dmul_rn(t, p.frequency), pc0);
        float h = 0.0f;
        for (int j = 0; j < hs.n; j++) {
            // upstream evaluates sin(q*k) in fp64: form the product in fp64, reduce it to one
            // turn there, then the SFU sine on the reduced argument
            const double x = __dmul_rn(q, (double)hs.k[j]);
            const double kk = rint(x * kInvTwoPi);
            double r = fma(-kk, kTwoPiHi, x);
            r = fma(-kk, kTwoPiLo, r);
            h = fmaf(__sinf((float)r), hs.amp[j], h);
        }
        out[i] = fmaf(h, amp, bias);
    }
}

// ------------------------------------------------------------------ FM prefix sum
// Three phases over tiles of kScanTile samples: tile sums -> scan of tile sums (one block)
// -> per-tile exclusive scan with the tile offset.  All sums in fp64.
constexpr int kScanItems = 8;
constexpr int kScanTile = kBlock * kScanItems;   // 2048

__device__ __forceinline__ double warp_inclusive_scan(double v)
{
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        double o = __shfl_up_sync(0xffffffffu, v, d);
        if ((threadIdx.x & 31) >= d) v += o;
    }
    return v;
}

// exclusive scan of one value per thread across the block; returns the exclusive prefix and
// writes the block total to *total
__device__ __forceinline__ double block_exclusive_scan(double v, double *total)
{
    __shared__ double s_warp[kBlock / 32];
    __shared__ double s_total;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double inc = warp_inclusive_scan(v);
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        double w = (lane < kBlock / 32) ? s_warp[lane] : 0.0;
        double winc = warp_inclusive_scan(w);
        if (lane < kBlock / 32) s_warp[lane] = winc - w;     // exclusive warp offsets
        if (lane == kBlock / 32 - 1) s_total = winc;
    }
    __syncthreads();
    double r = s_warp[warp] + (inc - v);
    *total = s_total;
    __syncthreads();
    return r;
}

template <typename LFO>
__global__ void __launch_bounds__(kBlock)
fm_tile_sums_kernel(const LFO *__restrict__ lfo, const sb_time_seg *__restrict__ tb, int nsegs,
                    int64_t n0, int64_t count, double *__restrict__ tile_sums)
{
    __shared__ sb_time_seg s_tb[kMaxSegs];
    for (int i = threadIdx.x; i < nsegs; i += blockDim.x) s_tb[i] = tb[i];
    __syncthreads();
    const int64_t ntiles = (count + kScanTile - 1) / kScanTile;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        int64_t base = tile * kScanTile + (int64_t)threadIdx.x * kScanItems;
        double acc = 0.0;
        if (base < count) {
            TimeCursor cur;
            cur.seek(s_tb, nsegs, n0 + base);
#pragma unroll
            for (int j = 0; j < kScanItems; j++) {
                int64_t i = base + j;
                if (i < count) acc += __dmul_rn((double)lfo[i], cur.step_at(n0 + i));
            }
        }
        double total;
        block_exclusive_scan(acc, &total);
        if (threadIdx.x == 0) tile_sums[tile] = total;
    }
}

// one block: exclusive scan of the tile sums in place, starting from *carry; advances *carry
__global__ void __launch_bounds__(kBlock)
fm_scan_tiles_kernel(double *__restrict__ tile_sums, int64_t ntiles, double *__restrict__ carry)
{
    double running = *carry;
    __syncthreads();
    for (int64_t base = 0; base < ntiles; base += kBlock) {
        int64_t i = base + threadIdx.x;
        double v = (i < ntiles) ? tile_sums[i] : 0.0;
        double total;
        double ex = block_exclusive_scan(v, &total);
        if (i < ntiles) tile_sums[i] = running + ex;
        running += total;
    }
    if (threadIdx.x == 0) *carry = running;
}

template <typename LFO>
__global__ void __launch_bounds__(kBlock)
fm_scan_apply_kernel(const LFO *__restrict__ lfo, const sb_time_seg *__restrict__ tb, int nsegs,
                     int64_t n0, int64_t count, const double *__restrict__ tile_offsets,
                     double *__restrict__ out)
{
    __shared__ sb_time_seg s_tb[kMaxSegs];
    for (int i = threadIdx.x; i < nsegs; i += blockDim.x) s_tb[i] = tb[i];
    __syncthreads();
    const int64_t ntiles = (count + kScanTile - 1) / kScanTile;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        int64_t base = tile * kScanTile + (int64_t)threadIdx.x * kScanItems;
        double w[kScanItems];
        double acc = 0.0;
        if (base < count) {
            TimeCursor cur;
            cur.seek(s_tb, nsegs, n0 + base);
#pragma unroll
            for (int j = 0; j < kScanItems; j++) {
                int64_t i = base + j;
                w[j] = (i < count) ? __dmul_rn((double)lfo[i], cur.step_at(n0 + i)) : 0.0;
                acc += w[j];
            }
        }
        double total;
        double ex = block_exclusive_scan(acc, &total) + tile_offsets[tile];
#pragma unroll
        for (int j = 0; j < kScanItems; j++) {
            int64_t i = base + j;
            if (i < count) out[i] = ex;
            ex += w[j];
        }
    }
}

// ------------------------------------------------------------------ noise
__global__ void noise_kernel(const float *__restrict__ u, int64_t draw0, int64_t cycles, float amp,
                             float bias, int64_t n0, int64_t count, float *__restrict__ out)
{
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
        int64_t draw = (n0 + i) / cycles - draw0;
        float v = __fadd_rn(-amp, __fmul_rn(__fsub_rn(amp, -amp), u[draw]));
        out[i] = __fadd_rn(v, bias);
    }
}

template <typename OUT>
__global__ void linear_kernel(double start, double inc, double lo, double hi, int64_t n0, int64_t count,
                              OUT *__restrict__ out)
{
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
        // upstream's serial value = clamp(value + inc): the first value is the start level as
        // given; from the second on the ramp runs from v1 = clamp(start + inc), clamped
        // (a start level outside [lo, hi] is pulled in by the first step, not by the n-th)
        int64_t n = n0 + i;
        double v = start;
        if (n > 0) {
            const double v1 = fmin(hi, fmax(lo, start + inc));
            v = fmin(hi, fmax(lo, v1 + (double)(n - 1) * inc));
        }
        out[i] = (OUT)v;
    }
}

// Philox-4x32-10 (Salmon et al., SC'11), one counter per four draws
__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1)
{
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    uint32_t hi0 = __umulhi(M0, c[0]), lo0 = M0 * c[0];
    uint32_t hi1 = __umulhi(M1, c[2]), lo1 = M1 * c[2];
    uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
}

__global__ void uniform_kernel(uint64_t seed, int64_t first, int64_t count, float *__restrict__ out)
{
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t q0 = first >> 2, q1 = (first + count + 3) >> 2;
    for (int64_t q = q0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < q1; q += stride) {
        uint32_t c[4] = {(uint32_t)q, (uint32_t)((uint64_t)q >> 32), 0u, 0u};
        uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
        for (int r = 0; r < 10; r++) {
            philox_round(c, k0, k1);
            k0 += 0x9E3779B9u;
            k1 += 0xBB67AE85u;
        }
#pragma unroll
        for (int j = 0; j < 4; j++) {
            int64_t d = q * 4 + j - first;
            if (d >= 0 && d < count) out[d] = (float)(c[j] >> 8) * (1.0f / 16777216.0f);
        }
    }
}

// ------------------------------------------------------------------ filters
__global__ void envelope_kernel(const float *__restrict__ in, sb_envelope e, int64_t n0, int64_t count,
                                float *__restrict__ out)
{
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
        out[i] = __fmul_rn(in[i], envelope_at(e, n0 + i));
}

__global__ void envelope64_kernel(const double *__restrict__ in, sb_envelope e, int64_t n0, int64_t count,
                                  double *__restrict__ out)
{
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
        out[i] = __dmul_rn(in[i], envelope_at64(e, n0 + i));
}

template <int OP>
__global__ void ew_kernel(const float *__restrict__ a, const float *__restrict__ b, float p0, float p1,
                          int64_t count, float *__restrict__ out)
{
    // 128-bit path over the aligned bulk, scalar tail
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    auto f = [&](float x, float y) -> float {
        if (OP == 0) return __fmul_rn(x, y);                    // AmpModulationFilter
        if (OP == 1) return fmaxf(fminf(x, p1), p0);            // ClipFilter: max(min(v,max),min)
        if (OP == 2) return fabsf(x);                           // AbsFilter
        return fmaf(x, p0, p1);                                 // scale + bias
    };
    bool aligned = ((((uintptr_t)a) | ((uintptr_t)out) | (b ? (uintptr_t)b : 0)) & 15) == 0;
    int64_t nvec = aligned ? count / 4 : 0;
    for (int64_t v = tid; v < nvec; v += stride) {
        float4 x = reinterpret_cast<const float4 *>(a)[v];
        float4 y = b ? reinterpret_cast<const float4 *>(b)[v] : make_float4(0, 0, 0, 0);
        float4 r = make_float4(f(x.x, y.x), f(x.y, y.y), f(x.z, y.z), f(x.w, y.w));
        reinterpret_cast<float4 *>(out)[v] = r;
    }
    for (int64_t i = nvec * 4 + tid; i < count; i += stride) out[i] = f(a[i], b ? b[i] : 0.0f);
}

__global__ void mix_kernel(const float *const *__restrict__ srcs, int nsrc, int64_t count,
                           float *__restrict__ out)
{
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
        float acc = 0.0f;
        for (int s = 0; s < nsrc; s++) acc = __fadd_rn(acc, srcs[s][i]);
        out[i] = acc;
    }
}

// out[i] = in[i*stride + offset]: one channel of an interleaved buffer
__global__ void deinterleave_kernel(const float *__restrict__ in, int stride_elems, int offset, int64_t count,
                                    float *__restrict__ out)
{
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
        out[i] = in[i * stride_elems + offset];
}

struct EchoTaps {
    int ntaps;
    int64_t delay[SB_ECHO_MAX];
    float gain[SB_ECHO_MAX];
};

__global__ void echo_kernel(const float *__restrict__ src, int64_t src_n0, int64_t n_after, EchoTaps taps,
                            int64_t n0, int64_t count, float *__restrict__ out)
{
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
        int64_t n = n0 + i;
        float acc = src[n - src_n0];
        if (n >= n_after) {
            for (int k = 0; k < taps.ntaps; k++) {
                int64_t m = n - taps.delay[k];
                if (m >= n_after) acc = __fadd_rn(acc, __fmul_rn(src[m - src_n0], taps.gain[k]));
            }
        }
        out[i] = acc;
    }
}

// ------------------------------------------------------------------ float -> PCM
__device__ __forceinline__ int trunc_to_int_sat(float x, int lo, int hi, unsigned int *bad)
{
    // int(x): toward zero.  NaN counts as out of range.
    if (!(x > (float)lo - 1.0f && x < (float)hi + 1.0f)) {
        (*bad)++;
        return (x < 0.0f) ? lo : hi;
    }
    return __float2int_rz(x);
}

__global__ void to_pcm16_kernel(const float *__restrict__ in, int64_t count, float scale,
                                int16_t *__restrict__ out, unsigned long long *overflow)
{
    // 8 samples per thread step: two 128-bit loads, one 128-bit store (scalar when unaligned / tail)
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned int bad = 0;
    const bool aligned = ((((uintptr_t)in) | ((uintptr_t)out)) & 15) == 0;
    const int64_t nvec = aligned ? count / 8 : 0;
    for (int64_t v = tid; v < nvec; v += stride) {
        const float4 x0 = reinterpret_cast<const float4 *>(in)[2 * v];
        const float4 x1 = reinterpret_cast<const float4 *>(in)[2 * v + 1];
        const float x[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
        union { int4 q; short s[8]; } r;
#pragma unroll
        for (int k = 0; k < 8; k++) r.s[k] = (short)trunc_to_int_sat(__fmul_rn(x[k], scale), -32768, 32767, &bad);
        reinterpret_cast<int4 *>(out)[v] = r.q;
    }
    for (int64_t i = nvec * 8 + tid; i < count; i += stride)
        out[i] = (int16_t)trunc_to_int_sat(__fmul_rn(in[i], scale), -32768, 32767, &bad);
    if (bad && overflow) atomicAdd(overflow, (unsigned long long)bad);
}

__global__ void to_pcm32_kernel(const float *__restrict__ in, int64_t count, double scale,
                                int32_t *__restrict__ out, unsigned long long *overflow)
{
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    unsigned int bad = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
        double x = __dmul_rn((double)in[i], scale);
        if (!(x > -2147483649.0 && x < 2147483648.0)) {
            bad++;
            out[i] = x < 0 ? INT32_MIN : INT32_MAX;
        } else {
            out[i] = __double2int_rz(x);
        }
    }
    if (bad && overflow) atomicAdd(overflow, (unsigned long long)bad);
}

__global__ void to_pcm16_stereo_kernel(const float *__restrict__ left, const float *__restrict__ right,
                                       int64_t frames, float scale, int16_t *__restrict__ out,
                                       unsigned long long *overflow)
{
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    unsigned int bad = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < frames; i += stride) {
        short2 v;
        v.x = (short)trunc_to_int_sat(__fmul_rn(left[i], scale), -32768, 32767, &bad);
        v.y = (short)trunc_to_int_sat(__fmul_rn(right[i], scale), -32768, 32767, &bad);
        reinterpret_cast<short2 *>(out)[i] = v;
    }
    if (bad && overflow) atomicAdd(overflow, (unsigned long long)bad);
}

}  // namespace sb

using namespace sb;

extern "C" int sb_osc_render_f32(const sb_osc_params *p, const sb_time_seg *tb_dev, int nsegs,
                                 int64_t n0, int64_t count, const double *fmsum_dev,
                                 const float *pwm_dev, float *out_dev, void *stream)
{
    SB_REQUIRE(p && tb_dev && out_dev, "sb_osc_render_f32: null pointer");
    SB_REQUIRE(p->wave >= SB_WAVE_SINE && p->wave <= SB_WAVE_POINTY, "sb_osc_render_f32: unknown waveform");
    SB_REQUIRE(nsegs > 0 && nsegs <= kMaxSegs && n0 >= 0 && count >= 0, "sb_osc_render_f32: bad range");
    if (count == 0) return SB_OK;
    constexpr int ITEMS = 4;
    int grid = grid_for((count + ITEMS - 1) / ITEMS, kBlock, kMaxBlocks);
    osc_render_kernel<ITEMS, float><<<grid, kBlock, 0, (cudaStream_t)stream>>>(*p, tb_dev, nsegs, n0, count,
                                                                              fmsum_dev, pwm_dev, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_osc_render_f64(const sb_osc_params *p, const sb_time_seg *tb_dev, int nsegs,
                                 int64_t n0, int64_t count, const double *fmsum_dev,
                                 const float *pwm_dev, double *out_dev, void *stream)
{
    SB_REQUIRE(p && tb_dev && out_dev, "sb_osc_render_f64: null pointer");
    SB_REQUIRE(p->wave >= SB_WAVE_SINE && p->wave <= SB_WAVE_POINTY, "sb_osc_render_f64: unknown waveform");
    SB_REQUIRE(nsegs > 0 && nsegs <= kMaxSegs && n0 >= 0 && count >= 0, "sb_osc_render_f64: bad range");
    if (count == 0) return SB_OK;
    constexpr int ITEMS = 4;
    int grid = grid_for((count + ITEMS - 1) / ITEMS, kBlock, kMaxBlocks);
    osc_render_kernel<ITEMS, double><<<grid, kBlock, 0, (cudaStream_t)stream>>>(*p, tb_dev, nsegs, n0, count,
                                                                               fmsum_dev, pwm_dev, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_harmonics_render_f32(const sb_osc_params *p, const int32_t *harm_n, const float *harm_amp,
                                       int nharm, const sb_time_seg *tb_dev, int nsegs, int64_t n0,
                                       int64_t count, const double *fmsum_dev, float *out_dev, void *stream)
{
    SB_REQUIRE(p && tb_dev && out_dev, "sb_harmonics_render_f32: null pointer");
    SB_REQUIRE(nharm >= 0 && nharm <= SB_HARMONICS_MAX && (nharm == 0 || (harm_n && harm_amp)),
               "sb_harmonics_render_f32: between 0 and 128 harmonics");
    SB_REQUIRE(nsegs > 0 && nsegs <= kMaxSegs && n0 >= 0 && count >= 0, "sb_harmonics_render_f32: bad range");
    if (count == 0) return SB_OK;
    HarmonicSet hs;
    hs.n = nharm;
    for (int j = 0; j < nharm; j++) {
        hs.k[j] = harm_n[j];
        hs.amp[j] = harm_amp[j];
    }
    harmonics_kernel<<<grid_for(count, kBlock, kMaxBlocks), kBlock, 0, (cudaStream_t)stream>>>(
        *p, hs, tb_dev, nsegs, n0, count, fmsum_dev, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int64_t sb_fm_prefix_scratch_bytes(int64_t count)
{
    int64_t ntiles = (count + kScanTile - 1) / kScanTile;
    return (ntiles + 1) * (int64_t)sizeof(double);
}

extern "C" int sb_fm_prefix_f64(const float *lfo_dev, const sb_time_seg *tb_dev, int nsegs, int64_t n0,
                                int64_t count, double *carry_dev, double *out_dev, void *scratch_dev,
                                void *stream)
{
    SB_REQUIRE(lfo_dev && tb_dev && carry_dev && out_dev && scratch_dev, "sb_fm_prefix_f64: null pointer");
    SB_REQUIRE(nsegs > 0 && nsegs <= kMaxSegs && n0 >= 0 && count >= 0, "sb_fm_prefix_f64: bad range");
    if (count == 0) return SB_OK;
    int64_t ntiles = (count + kScanTile - 1) / kScanTile;
    double *tiles = (double *)scratch_dev;
    int grid = grid_for(ntiles, 1, kMaxBlocks);
    cudaStream_t s = (cudaStream_t)stream;
    fm_tile_sums_kernel<float><<<grid, kBlock, 0, s>>>(lfo_dev, tb_dev, nsegs, n0, count, tiles);
    SB_LAUNCH_CHECK();
    fm_scan_tiles_kernel<<<1, kBlock, 0, s>>>(tiles, ntiles, carry_dev);
    SB_LAUNCH_CHECK();
    fm_scan_apply_kernel<float><<<grid, kBlock, 0, s>>>(lfo_dev, tb_dev, nsegs, n0, count, tiles, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_fm_prefix_f64_from_f64(const double *lfo_dev, const sb_time_seg *tb_dev, int nsegs, int64_t n0,
                                         int64_t count, double *carry_dev, double *out_dev, void *scratch_dev,
                                         void *stream)
{
    SB_REQUIRE(lfo_dev && tb_dev && carry_dev && out_dev && scratch_dev, "sb_fm_prefix_f64_from_f64: null pointer");
    SB_REQUIRE(nsegs > 0 && nsegs <= kMaxSegs && n0 >= 0 && count >= 0, "sb_fm_prefix_f64_from_f64: bad range");
    if (count == 0) return SB_OK;
    int64_t ntiles = (count + kScanTile - 1) / kScanTile;
    double *tiles = (double *)scratch_dev;
    int grid = grid_for(ntiles, 1, kMaxBlocks);
    cudaStream_t s = (cudaStream_t)stream;
    fm_tile_sums_kernel<double><<<grid, kBlock, 0, s>>>(lfo_dev, tb_dev, nsegs, n0, count, tiles);
    SB_LAUNCH_CHECK();
    fm_scan_tiles_kernel<<<1, kBlock, 0, s>>>(tiles, ntiles, carry_dev);
    SB_LAUNCH_CHECK();
    fm_scan_apply_kernel<double><<<grid, kBlock, 0, s>>>(lfo_dev, tb_dev, nsegs, n0, count, tiles, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

// ------------------------------------------------------------------ FM prefix sum, no allocation
// The same three kernels as sb_fm_prefix_f64, but the tile-sum scratch is this fixed array in
// the library's device image instead of a buffer the caller has to allocate.  It exists once
// per GPU that loads the library and stays resident until the process exits (64 KiB + 8 B).
// One slot per 2048-sample tile plus one spare, hence the per-call limit below.
#define SB_NOALLOC_MAX_TILES 8192
static __device__ __align__(256) double g_noalloc_tile_sums[SB_NOALLOC_MAX_TILES + 1];

namespace sb {
constexpr int kNoallocMaxDevices = 64;
static std::mutex g_noalloc_mu;
static double *g_noalloc_ptr[kNoallocMaxDevices] = {nullptr};

// `g_noalloc_tile_sums` as written in host code is only a handle for the runtime; the address
// kernels can dereference has to be asked for, per device, with that device current.
static int noalloc_scratch(double **out)
{
    int dev = -1;
    SB_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kNoallocMaxDevices) {
        set_error("sb_fm_prefix_f64_noalloc: device ordinal %d is outside 0..%d", dev, kNoallocMaxDevices - 1);
        return SB_ERANGE;
    }
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        void *p = nullptr;
        SB_CUDA_CHECK(cudaGetSymbolAddress(&p, g_noalloc_tile_sums));
        g_noalloc_ptr[dev] = static_cast<double *>(p);
    }
    *out = g_noalloc_ptr[dev];
    return SB_OK;
}
}  // namespace sb

extern "C" int64_t sb_fm_prefix_noalloc_max_count(void)
{
    return (int64_t)SB_NOALLOC_MAX_TILES * kScanTile;
}

extern "C" int sb_fm_prefix_f64_noalloc(const float *lfo_dev, const sb_time_seg *tb_dev, int nsegs, int64_t n0,
                                        int64_t count, double *carry_dev, double *out_dev, void *stream)
{
    SB_REQUIRE(lfo_dev && tb_dev && carry_dev && out_dev, "sb_fm_prefix_f64_noalloc: null pointer");
    SB_REQUIRE(nsegs > 0 && nsegs <= kMaxSegs && n0 >= 0 && count >= 0, "sb_fm_prefix_f64_noalloc: bad range");
    const int64_t max_count = (int64_t)SB_NOALLOC_MAX_TILES * kScanTile;
    if (count > max_count) {
        set_error("sb_fm_prefix_f64_noalloc: count %lld exceeds the built-in scratch (at most %lld samples per "
                  "call); split the range or use sb_fm_prefix_f64 with caller-provided scratch",
                  (long long)count, (long long)max_count);
        return SB_ERANGE;
    }
    if (count == 0) return SB_OK;
    double *tiles = nullptr;
    int rc = noalloc_scratch(&tiles);
    if (rc) return rc;
    // Nothing in the scratch has to start at zero: phase 1 writes slot [0, ntiles) before
    // phase 2 reads exactly those slots, so whatever an earlier call left behind is never seen.
    int64_t ntiles = (count + kScanTile - 1) / kScanTile;
    int grid = grid_for(ntiles, 1, kMaxBlocks);
    cudaStream_t s = (cudaStream_t)stream;
    fm_tile_sums_kernel<float><<<grid, kBlock, 0, s>>>(lfo_dev, tb_dev, nsegs, n0, count, tiles);
    SB_LAUNCH_CHECK();
    fm_scan_tiles_kernel<<<1, kBlock, 0, s>>>(tiles, ntiles, carry_dev);
    SB_LAUNCH_CHECK();
    fm_scan_apply_kernel<float><<<grid, kBlock, 0, s>>>(lfo_dev, tb_dev, nsegs, n0, count, tiles, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_noise_render_f32(const float *u_dev, int64_t draw0, int64_t cycles, double amplitude,
                                   double bias, int64_t n0, int64_t count, float *out_dev, void *stream)
{
    SB_REQUIRE(u_dev && out_dev, "sb_noise_render_f32: null pointer");
    SB_REQUIRE(cycles >= 1 && n0 >= 0 && count >= 0 && draw0 >= 0 && n0 / cycles >= draw0,
               "sb_noise_render_f32: bad range");
    if (count == 0) return SB_OK;
    noise_kernel<<<grid_for(count, kBlock, kMaxBlocks), kBlock, 0, (cudaStream_t)stream>>>(
        u_dev, draw0, cycles, (float)amplitude, (float)bias, n0, count, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_linear_render_f32(double start, double increment, double min_value, double max_value,
                                    int64_t n0, int64_t count, float *out_dev, void *stream)
{
    SB_REQUIRE(out_dev && n0 >= 0 && count >= 0, "sb_linear_render_f32: bad arguments");
    if (count == 0) return SB_OK;
    linear_kernel<float><<<grid_for(count, kBlock, kMaxBlocks), kBlock, 0, (cudaStream_t)stream>>>(
        start, increment, min_value, max_value, n0, count, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_linear_render_f64(double start, double increment, double min_value, double max_value,
                                    int64_t n0, int64_t count, double *out_dev, void *stream)
{
    SB_REQUIRE(out_dev && n0 >= 0 && count >= 0, "sb_linear_render_f64: bad arguments");
    if (count == 0) return SB_OK;
    linear_kernel<double><<<grid_for(count, kBlock, kMaxBlocks), kBlock, 0, (cudaStream_t)stream>>>(
        start, increment, min_value, max_value, n0, count, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_uniform_f32(uint64_t seed, int64_t first_draw, int64_t count, float *out_dev, void *stream)
{
    SB_REQUIRE(out_dev && first_draw >= 0 && count >= 0, "sb_uniform_f32: bad arguments");
    if (count == 0) return SB_OK;
    uniform_kernel<<<grid_for(count / 4 + 1, kBlock, kMaxBlocks), kBlock, 0, (cudaStream_t)stream>>>(
        seed, first_draw, count, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_envelope_apply_f32(const float *in_dev, const sb_envelope *env, int64_t n0, int64_t count,
                                     float *out_dev, void *stream)
{
    SB_REQUIRE(in_dev && env && out_dev && n0 >= 0 && count >= 0, "sb_envelope_apply_f32: bad arguments");
    if (count == 0) return SB_OK;
    envelope_kernel<<<grid_for(count, kBlock, kMaxBlocks), kBlock, 0, (cudaStream_t)stream>>>(
        in_dev, *env, n0, count, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_envelope_apply_f64(const double *in_dev, const sb_envelope *env, int64_t n0, int64_t count,
                                     double *out_dev, void *stream)
{
    SB_REQUIRE(in_dev && env && out_dev && n0 >= 0 && count >= 0, "sb_envelope_apply_f64: bad arguments");
    if (count == 0) return SB_OK;
    envelope64_kernel<<<grid_for(count, kBlock, kMaxBlocks), kBlock, 0, (cudaStream_t)stream>>>(
        in_dev, *env, n0, count, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

template <int OP>
static int launch_ew(const float *a, const float *b, float p0, float p1, int64_t count, float *out, void *stream)
{
    if (count == 0) return SB_OK;
    ew_kernel<OP><<<grid_for(count / 4 + 1, kBlock, kMaxBlocks), kBlock, 0, (cudaStream_t)stream>>>(
        a, b, p0, p1, count, out);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_mul_f32(const float *a_dev, const float *b_dev, int64_t count, float *out_dev, void *stream)
{
    SB_REQUIRE(a_dev && b_dev && out_dev && count >= 0, "sb_mul_f32: bad arguments");
    return launch_ew<0>(a_dev, b_dev, 0, 0, count, out_dev, stream);
}

extern "C" int sb_clip_f32(const float *in_dev, float lo, float hi, int64_t count, float *out_dev, void *stream)
{
    SB_REQUIRE(in_dev && out_dev && count >= 0, "sb_clip_f32: bad arguments");
    return launch_ew<1>(in_dev, nullptr, lo, hi, count, out_dev, stream);
}

extern "C" int sb_abs_f32(const float *in_dev, int64_t count, float *out_dev, void *stream)
{
    SB_REQUIRE(in_dev && out_dev && count >= 0, "sb_abs_f32: bad arguments");
    return launch_ew<2>(in_dev, nullptr, 0, 0, count, out_dev, stream);
}

extern "C" int sb_scale_bias_f32(const float *in_dev, float scale, float bias, int64_t count, float *out_dev,
                                 void *stream)
{
    SB_REQUIRE(in_dev && out_dev && count >= 0, "sb_scale_bias_f32: bad arguments");
    return launch_ew<3>(in_dev, nullptr, scale, bias, count, out_dev, stream);
}

extern "C" int sb_mix_f32(const float *const *src_ptrs_dev, int nsrc, int64_t count, float *out_dev, void *stream)
{
    SB_REQUIRE(src_ptrs_dev && out_dev && nsrc >= 1 && count >= 0, "sb_mix_f32: bad arguments");
    if (count == 0) return SB_OK;
    mix_kernel<<<grid_for(count, kBlock, kMaxBlocks), kBlock, 0, (cudaStream_t)stream>>>(src_ptrs_dev, nsrc,
                                                                                        count, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_channel_f32(const float *in_dev, int nchannels, int channel, int64_t frames, float *out_dev,
                              void *stream)
{
    SB_REQUIRE(frames >= 0 && nchannels >= 1 && channel >= 0 && channel < nchannels && (frames == 0 || (in_dev && out_dev)),
               "sb_channel_f32: bad arguments");
    if (frames == 0) return SB_OK;
    deinterleave_kernel<<<grid_for(frames, kBlock, kMaxBlocks), kBlock, 0, (cudaStream_t)stream>>>(
        in_dev, nchannels, channel, frames, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_echo_f32(const float *src_dev, int64_t src_n0, int64_t n_after, int ntaps,
                           const int64_t *delays, const float *gains, int64_t n0, int64_t count,
                           float *out_dev, void *stream)
{
    SB_REQUIRE(src_dev && out_dev && n0 >= src_n0 && count >= 0, "sb_echo_f32: bad arguments");
    SB_REQUIRE(ntaps >= 0 && ntaps <= SB_ECHO_MAX, "sb_echo_f32: too many echoes");
    SB_REQUIRE(ntaps == 0 || (delays && gains), "sb_echo_f32: null taps");
    if (count == 0) return SB_OK;
    EchoTaps taps;
    taps.ntaps = ntaps;
    for (int k = 0; k < ntaps; k++) {
        SB_REQUIRE(delays[k] >= 0, "sb_echo_f32: negative delay");
        // the oldest sample a tap can read is max(n_after, n0 - delay): it must be in src
        int64_t oldest = n0 - delays[k];
        if (oldest < n_after) oldest = n_after;
        SB_REQUIRE(oldest >= src_n0 || n0 + count <= n_after + delays[k], "sb_echo_f32: source history too short");
        taps.delay[k] = delays[k];
        taps.gain[k] = gains[k];
    }
    echo_kernel<<<grid_for(count, kBlock, kMaxBlocks), kBlock, 0, (cudaStream_t)stream>>>(
        src_dev, src_n0, n_after, taps, n0, count, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_f32_to_pcm16(const float *in_dev, int64_t count, float scale, int16_t *out_dev,
                               unsigned long long *overflow_dev, void *stream)
{
    SB_REQUIRE(count >= 0 && (count == 0 || (in_dev && out_dev)), "sb_f32_to_pcm16: bad arguments");
    if (count == 0) return SB_OK;
    to_pcm16_kernel<<<grid_for(count / 8 + 1, kBlock, kMaxBlocks), kBlock, 0, (cudaStream_t)stream>>>(
        in_dev, count, scale, out_dev, overflow_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_f32_to_pcm32(const float *in_dev, int64_t count, double scale, int32_t *out_dev,
                               unsigned long long *overflow_dev, void *stream)
{
    SB_REQUIRE(in_dev && out_dev && count >= 0, "sb_f32_to_pcm32: bad arguments");
    if (count == 0) return SB_OK;
    to_pcm32_kernel<<<grid_for(count, kBlock, kMaxBlocks), kBlock, 0, (cudaStream_t)stream>>>(
        in_dev, count, scale, out_dev, overflow_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_f32x2_to_pcm16_stereo(const float *left_dev, const float *right_dev, int64_t frames,
                                        float scale, int16_t *out_dev, unsigned long long *overflow_dev,
                                        void *stream)
{
    SB_REQUIRE(left_dev && right_dev && out_dev && frames >= 0, "sb_f32x2_to_pcm16_stereo: bad arguments");
    if (frames == 0) return SB_OK;
    to_pcm16_stereo_kernel<<<grid_for(frames, kBlock, kMaxBlocks), kBlock, 0, (cudaStream_t)stream>>>(
        left_dev, right_dev, frames, scale, out_dev, overflow_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}
