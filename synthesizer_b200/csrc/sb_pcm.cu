// Integer PCM kernels: bit-exact restatements of the CPython audioop routines that
// synthplayer's Sample delegates to (Modules/audioop.c, CPython 3.12.3), plus Sample's own
// fade / modulate formulas.  All are HBM-bound streaming kernels: int16 data moves as
// 128-bit vectors (8 samples) over the 16-byte aligned bulk, scalar at the ragged edges.
#include "sb_common.cuh"

namespace sb {

constexpr int kPBlock = 256;
constexpr int kPMaxBlocks = 148 * 16;

__device__ __forceinline__ int sat16(int v) { return max(-32768, min(32767, v)); }

// audioop.c fbound(): clip (note: val < minval + 1.0 -> minval), floor, cast
__device__ __forceinline__ int fbound16(double val)
{
    if (val > 32767.0) val = 32767.0;
    else if (val < -32767.0) val = -32768.0;
    return (int)floor(val);
}

// the same for |val| < 2^31 (callers check their factors on the host): no conversion instructions
__device__ __forceinline__ int fbound16_small(double val)
{
    return max(-32768, min(32767, f64_floor_to_i32(val)));
}

union Vec8 {
    int4 v;
    short s[8];
};

// generic unary/binary streaming map over int16 with a 128-bit fast path
template <class F>
// (no __restrict__: the in-place callers pass the same buffer as `a` and `out`; every element is
// read and written by the same thread, loads before the store)
__device__ __forceinline__ void map16(const int16_t *a, const int16_t *b, int16_t *out, int64_t count, F f)
{
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    // head: until `out` is 16-byte aligned; the fast path needs a (and b) aligned the same way
    int64_t head = ((16 - ((uintptr_t)out & 15)) & 15) / 2;
    if (head > count) head = count;
    bool same = (((uintptr_t)(a + head)) & 15) == 0 && (!b || (((uintptr_t)(b + head)) & 15) == 0);
    if (!same) {
        for (int64_t i = tid; i < count; i += stride) out[i] = (int16_t)f(i, a[i], b ? b[i] : 0);
        return;
    }
    for (int64_t i = tid; i < head; i += stride) out[i] = (int16_t)f(i, a[i], b ? b[i] : 0);
    const int64_t nvec = (count - head) / 8;
    const int4 *av = reinterpret_cast<const int4 *>(a + head);
    const int4 *bv = b ? reinterpret_cast<const int4 *>(b + head) : nullptr;
    int4 *ov = reinterpret_cast<int4 *>(out + head);
    // (two vectors in flight per thread were measured slower: add 0.83 -> 0.78, fadeout 0.64 -> 0.61 of peak)
    for (int64_t v = tid; v < nvec; v += stride) {
        Vec8 x, y, r;
        x.v = av[v];
        if (bv) y.v = bv[v]; else y.v = make_int4(0, 0, 0, 0);
#pragma unroll
        for (int k = 0; k < 8; k++) r.s[k] = (short)f(head + v * 8 + k, x.s[k], y.s[k]);
        ov[v] = r.v;
    }
    for (int64_t i = head + nvec * 8 + tid; i < count; i += stride) out[i] = (int16_t)f(i, a[i], b ? b[i] : 0);
}

__global__ void pcm16_add_kernel(const int16_t *a, const int16_t *b, int64_t count, int16_t *out)
{
    map16(a, b, out, count, [](int64_t, int x, int y) { return sat16(x + y); });
}

// SMALL: |factor| <= 16384 (checked on the host), so every product is below 2^31 in magnitude
template <bool SMALL>
__global__ void pcm16_mul_kernel(const int16_t *a, int64_t count, double factor, int16_t *out)
{
    map16(a, nullptr, out, count, [factor](int64_t, int x, int) {
        return SMALL ? fbound16_small(__dmul_rn(i32_to_f64(x), factor)) : fbound16(__dmul_rn((double)x, factor));
    });
}

__global__ void pcm16_bias_kernel(const int16_t *a, int64_t count, int bias, int16_t *out)
{
    map16(a, nullptr, out, count,
          [bias](int64_t, int x, int) { return (int)(int16_t)(uint16_t)((uint32_t)(uint16_t)x + (uint32_t)bias); });
}

// SMALL: fewer than 2^31 samples and volumes of at most 16384 in magnitude (checked on the host)
// RN(k * factor) for 0 <= k < 2^31 in ONE fused operation, without converting k: the bit pattern
// 0x43300000:k is the double 2^52 + k, and (2^52 + k)*factor - 2^52*factor is formed exactly
// inside the FMA (2^52*factor is a power-of-two scaling, exact) and rounded once.
__device__ __forceinline__ double index_times(int k, double factor, double neg_factor_2p52)
{
    return __fma_rn(__hiloint2double(0x43300000, k), factor, neg_factor_2p52);
}

template <bool SMALL>
__global__ void pcm16_fadeout_kernel(int16_t *buf, int64_t nsamples, double decrease)
{
    const double numsamples = (double)nsamples, rcp_n = __drcp_rn(numsamples);
    const double dec52 = -decrease * 4503599627370496.0;
    map16(buf, nullptr, buf, nsamples, [=](int64_t k, int x, int) {
        // int(sample * (1.0 - k*decrease/numsamples)), every operation rounded like CPython's
        const double kdec = SMALL ? index_times((int)k, decrease, dec52) : __dmul_rn((double)k, decrease);
        double f = __dsub_rn(1.0, div_by_known(kdec, numsamples, rcp_n));
        return SMALL ? f64_trunc_to_i32(__dmul_rn(i32_to_f64(x), f)) : __double2int_rz(__dmul_rn((double)x, f));
    });
}

template <bool SMALL>
__global__ void pcm16_fadein_kernel(int16_t *buf, int64_t nsamples, double increase, double start_volume)
{
    const double numsamples = (double)nsamples, rcp_n = __drcp_rn(numsamples);
    const double inc52 = -increase * 4503599627370496.0;
    map16(buf, nullptr, buf, nsamples, [=](int64_t k, int x, int) {
        const double kinc = SMALL ? index_times((int)k, increase, inc52) : __dmul_rn((double)k, increase);
        double f = __dadd_rn(div_by_known(kinc, numsamples, rcp_n), start_volume);
        return SMALL ? f64_trunc_to_i32(__dmul_rn(i32_to_f64(x), f)) : __double2int_rz(__dmul_rn((double)x, f));
    });
}

__global__ void pcm16_modulate_kernel(int16_t *buf, const float *__restrict__ mod, int64_t count)
{
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
        buf[i] = (int16_t)__double2int_rz(__dmul_rn((double)buf[i], (double)mod[i]));
}

// the cycled, normalised modulator of Sample.modulate_amp(list / array / Sample): the factors
// stay in fp64, as upstream's v/biggest are (float32 factors give 6 where 70/100*10 gives 7)
template <class T>
__global__ void pcm_modulate_cycle_kernel(T *buf, const double *__restrict__ mod, int64_t mod_len, int64_t count)
{
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
        buf[i] = (T)__double2int_rz(__dmul_rn((double)buf[i], mod[i % mod_len]));
}

__global__ void pcm16_reverse_kernel(const int16_t *__restrict__ in, int64_t count, int16_t *__restrict__ out)
{
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
        out[i] = in[count - 1 - i];
}

__global__ void pcm16_maxabs_kernel(const int16_t *__restrict__ in, int64_t count, unsigned int *out)
{
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    unsigned int m = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
        int v = in[i];
        unsigned int a = v < 0 ? (unsigned int)(-v) : (unsigned int)v;
        m = max(m, a);
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, d));
    if ((threadIdx.x & 31) == 0 && m) atomicMax(out, m);
}

// SMALL: |lf|, |rf| <= 16384 (checked on the host)
template <bool SMALL>
__device__ __forceinline__ int mono_of(int l, int r, double lf, double rf)
{
    if (SMALL) return fbound16_small(__dadd_rn(__dmul_rn(i32_to_f64(l), lf), __dmul_rn(i32_to_f64(r), rf)));
    return fbound16(__dadd_rn(__dmul_rn((double)l, lf), __dmul_rn((double)r, rf)));
}

// four frames per thread over the aligned bulk: 16 B in, 8 B out
template <bool SMALL>
__global__ void pcm16_tomono_kernel(const int16_t *__restrict__ in, int64_t frames, double lf, double rf,
                                    int16_t *__restrict__ out)
{
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (int64_t)gridDim.x * blockDim.x;
    const bool aligned = ((((uintptr_t)in) & 15) | (((uintptr_t)out) & 7)) == 0;
    const int64_t nvec = aligned ? frames / 4 : 0;
    for (int64_t v = tid; v < nvec; v += stride) {
        Vec8 x;
        x.v = reinterpret_cast<const int4 *>(in)[v];
        short4 o;
        o.x = (short)mono_of<SMALL>(x.s[0], x.s[1], lf, rf);
        o.y = (short)mono_of<SMALL>(x.s[2], x.s[3], lf, rf);
        o.z = (short)mono_of<SMALL>(x.s[4], x.s[5], lf, rf);
        o.w = (short)mono_of<SMALL>(x.s[6], x.s[7], lf, rf);
        reinterpret_cast<short4 *>(out)[v] = o;
    }
    for (int64_t i = nvec * 4 + tid; i < frames; i += stride)
        out[i] = (int16_t)mono_of<SMALL>(in[2 * i], in[2 * i + 1], lf, rf);
}

template <bool SMALL>
__device__ __forceinline__ short2 stereo_of(int x, double lf, double rf)
{
    short2 o;
    if (SMALL) {
        const double v = i32_to_f64(x);
        o.x = (short)fbound16_small(__dmul_rn(v, lf));
        o.y = (short)fbound16_small(__dmul_rn(v, rf));
    } else {
        const double v = (double)x;
        o.x = (short)fbound16(__dmul_rn(v, lf));
        o.y = (short)fbound16(__dmul_rn(v, rf));
    }
    return o;
}

// four frames per thread over the aligned bulk: 8 B in, 16 B out
template <bool SMALL>
__global__ void pcm16_tostereo_kernel(const int16_t *__restrict__ in, int64_t frames, double lf, double rf,
                                      int16_t *__restrict__ out)
{
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (int64_t)gridDim.x * blockDim.x;
    const bool aligned = ((((uintptr_t)in) & 7) | (((uintptr_t)out) & 15)) == 0;
    const int64_t nvec = aligned ? frames / 4 : 0;
    for (int64_t v = tid; v < nvec; v += stride) {
        const short4 x = reinterpret_cast<const short4 *>(in)[v];
        const short2 a = stereo_of<SMALL>(x.x, lf, rf), b = stereo_of<SMALL>(x.y, lf, rf);
        const short2 c = stereo_of<SMALL>(x.z, lf, rf), d = stereo_of<SMALL>(x.w, lf, rf);
        Vec8 o;
        o.s[0] = a.x, o.s[1] = a.y, o.s[2] = b.x, o.s[3] = b.y;
        o.s[4] = c.x, o.s[5] = c.y, o.s[6] = d.x, o.s[7] = d.y;
        reinterpret_cast<int4 *>(out)[v] = o.v;
    }
    for (int64_t i = nvec * 4 + tid; i < frames; i += stride) {
        const short2 o = stereo_of<SMALL>(in[i], lf, rf);
        out[2 * i] = o.x;
        out[2 * i + 1] = o.y;
    }
}

__global__ void pcm16_to_32_kernel(const int16_t *__restrict__ in, int64_t count, int32_t *__restrict__ out)
{
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
        out[i] = (int32_t)((uint32_t)(int32_t)in[i] << 16);
}

__global__ void pcm32_to_16_kernel(const int32_t *__restrict__ in, int64_t count, int16_t *__restrict__ out)
{
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
        out[i] = (int16_t)(in[i] >> 16);
}

// audioop.ratecv, state=None, weightA=1, weightB=0, in closed form per output frame j:
//   c = ceil(j*inrate/outrate) is the newest input frame read when j is emitted,
//   d = c*outrate - j*inrate in [0, outrate) the interpolation counter,
//   out = (int)((prev*d + cur*(outrate-d)) / outrate) on samples widened to 32 bits (<<16),
//   then >>16.
//
// The per-sample arithmetic in 32-bit integers.  With p, q the 16-bit samples and M = p*d +
// q*(outrate-d) = q*outrate + (p-q)*d, audioop's numerator is 65536*M: below 2^53 for reduced
// rates under 2^21, so its two fp64 products and their sum are exact, and the correctly rounded
// quotient truncates like the exact one (a non-integer quotient is at least 1/outrate from the
// next integer, far more than half an ulp).  Hence the value stored is
//   trunc(65536*M / outrate) >> 16  =  floor(M / outrate)                       for M >= 0
//                                   =  floor(M / outrate) + [0 < outrate - rem < outrate/65536]   for M < 0
// (rem = M mod outrate; the second case is C's truncation toward zero followed by an arithmetic
// shift), and floor(M / outrate) = q + floor((p-q)*d / outrate), a 17-bit quotient: one fp32
// estimate (error < 0.02) and a remainder correction in wrapping 32-bit arithmetic.
// tools/ratecv_formula_check.py replays exactly this against CPython's audioop.
struct RcvRate {
    uint32_t outr;       // reduced output rate (< 2^21)
    uint32_t Qk, Rk;     // 32*inrate = Qk*outr + Rk: what 32 output frames advance (c, d) by
    uint32_t lim;        // outr - ceil(outr/65536): remainders above it trigger the negative fix-up
    float rcp;           // 1/outr
    // outr <= 16384 (every usual pair of sample rates, reduced): floor(n/outr) for 0 <= n < 2^31 as one
    // widening multiply and a shift, floor(n/outr) = (n*mg) >> msh with msh = 31 + ceil(log2 outr),
    // mg = ceil(2^msh/outr) (Granlund & Montgomery; tools/ratecv_formula_check.py replays it)
    uint32_t mg;
    int msh;
    uint32_t moff;       // 65536*outr: makes the numerator (prev-cur)*d non-negative, and shifts the quotient by 65536
};

__device__ __forceinline__ RcvRate rcv_rate(int64_t inrate, int64_t outrate)
{
    RcvRate rt;
    rt.outr = (uint32_t)outrate;
    const uint32_t K = 32u * (uint32_t)inrate;
    rt.Qk = K / rt.outr;
    rt.Rk = K - rt.Qk * rt.outr;
    rt.lim = rt.outr - ((rt.outr + 65535u) >> 16);
    rt.rcp = __frcp_rn((float)rt.outr);
    rt.mg = 0;
    rt.msh = 0;
    rt.moff = 0;
    if (rt.outr <= 16384u) {
        int L = 0;
        while ((1u << L) < rt.outr) L++;
        rt.msh = 31 + L;
        rt.mg = (uint32_t)(((1ull << rt.msh) + rt.outr - 1) / rt.outr);
        rt.moff = 65536u * rt.outr;
    }
    return rt;
}

// int -> float and floor(float) -> int for magnitudes below 2^22, by adding 1.5*2^23: one integer
// and one FMA-pipe operation each instead of a conversion on the 16-lane XU pipe
__device__ __forceinline__ float small_int_to_float(int x)
{
    return __fadd_rn(__int_as_float(0x4B400000 + x), -12582912.0f);
}
__device__ __forceinline__ int small_floor_to_int(float x)
{
    return __float_as_int(__fadd_rd(x, 12582912.0f)) - 0x4B400000;
}

// FIXUP = false for outr <= 65536 (every real sample-rate pair): then ceil(outr/65536) = 1, a
// remainder can never exceed outr - 1 = lim, and the negative fix-up above never applies.
// cur + floor((prev-cur)*d/outr) by the invariant multiplication (outr <= 16384): five integer operations
__device__ __forceinline__ int rcv_sample_magic(int prev, int cur, uint32_t d, const RcvRate &rt)
{
    const uint32_t n = (uint32_t)((prev - cur) * (int)d) + rt.moff;             // in [0, 2^31)
    const unsigned long long wide = (unsigned long long)n * rt.mg;
    return cur + (int)(uint32_t)(wide >> rt.msh) - 65536;
}

template <bool FIXUP>
__device__ __forceinline__ int rcv_sample(int prev, int cur, uint32_t d, float w, const RcvRate &rt)
{
    const int diff = prev - cur;
    // w = d/outr; the estimate is biased down by more than its error, so it is floor or floor-1
    int est = small_floor_to_int(__fmaf_rn(small_int_to_float(diff), w, -0.015625f));
    uint32_t r = (uint32_t)diff * d - (uint32_t)est * rt.outr;       // in [0, 2*outr)
    const bool hi = r >= rt.outr;
    est += hi ? 1 : 0;
    const int qf = cur + est;
    if (!FIXUP) return qf;
    r -= hi ? rt.outr : 0u;
    return qf + ((qf < 0 && r > rt.lim) ? 1 : 0);
}

// A lane's ITERS frames j0, j0+32, ...: all loads first (2*ITERS in flight per lane), then the
// arithmetic and the stores.  `pin` points at the lane's newest input frame for j0, `pout` at
// output frame j0; everything inside is 32-bit offsets from those.  FULL = every frame exists.
template <int NCH, bool VEC, int ITERS, bool FULL, bool FIXUP, bool MAGIC>
__device__ __forceinline__ void rcv_batch(const int16_t *__restrict__ pin, int16_t *__restrict__ pout, int nleft,
                                          bool first_is_frame0, uint32_t &crel, uint32_t &d, const RcvRate &rt)
{
    uint32_t dd[ITERS];
    uint32_t wc[ITERS], wp[ITERS];              // raw current / previous frames (one or two int16)
#pragma unroll
    for (int it = 0; it < ITERS; it++) {
        dd[it] = d;
        wc[it] = wp[it] = 0;
        if (FULL || it * 32 < nleft) {
            // frame 0 of the output has no previous frame; its weight d is 0 there, any value does
            const int back = (it == 0 && first_is_frame0) ? 0 : 1;
            if (NCH == 1) {
                wc[it] = (uint16_t)pin[crel];
                wp[it] = (uint16_t)pin[(int)crel - back];
            } else if (VEC) {
                wc[it] = reinterpret_cast<const uint32_t *>(pin)[crel];
                wp[it] = reinterpret_cast<const uint32_t *>(pin)[(int)crel - back];
            } else {
                wc[it] = (uint32_t)(uint16_t)pin[2 * crel] | ((uint32_t)(uint16_t)pin[2 * crel + 1] << 16);
                const int pb = 2 * ((int)crel - back);
                wp[it] = (uint32_t)(uint16_t)pin[pb] | ((uint32_t)(uint16_t)pin[pb + 1] << 16);
            }
        }
        const bool inc = rt.Rk > d;
        crel += rt.Qk + (inc ? 1u : 0u);
        d = d - rt.Rk + (inc ? rt.outr : 0u);
    }
#pragma unroll
    for (int it = 0; it < ITERS; it++) {
        if (FULL || it * 32 < nleft) {
            const float w = MAGIC ? 0.0f : __fmul_rn(small_int_to_float((int)dd[it]), rt.rcp);
            const int o0 = MAGIC ? rcv_sample_magic((int)(int16_t)(wp[it] & 0xffffu), (int)(int16_t)(wc[it] & 0xffffu), dd[it], rt)
                                 : rcv_sample<FIXUP>((int)(int16_t)(wp[it] & 0xffffu), (int)(int16_t)(wc[it] & 0xffffu), dd[it], w, rt);
            if (NCH == 1) {
                pout[it * 32] = (int16_t)o0;
            } else {
                const int o1 = MAGIC ? rcv_sample_magic((int)wp[it] >> 16, (int)wc[it] >> 16, dd[it], rt)
                                     : rcv_sample<FIXUP>((int)wp[it] >> 16, (int)wc[it] >> 16, dd[it], w, rt);
                if (VEC) reinterpret_cast<uint32_t *>(pout)[it * 32] = ((uint32_t)o0 & 0xffffu) | ((uint32_t)o1 << 16);
                else pout[it * 64] = (int16_t)o0, pout[it * 64 + 1] = (int16_t)o1;
            }
        }
    }
}

// A warp renders the 32*ITERS*BATCHES consecutive output frames starting at jc; lane l owns
// frames jc + l + 32*k.  A lane's first frame is placed by one 64-bit division (fp64 estimate
// plus correction); each further frame is 32 outputs on, which moves (c, d) by constants.
// Lanes map to consecutive output frames: coalesced stores, near-coalesced loads.
template <int NCH, bool VEC, int ITERS, int BATCHES, bool FIXUP, bool MAGIC>
__device__ __forceinline__ void rcv_chunk(const int16_t *__restrict__ in, int16_t *__restrict__ out, int64_t jc,
                                          int64_t out_frames, int64_t inrate, int64_t outrate, const RcvRate &rt,
                                          double rcp64)
{
    const int64_t j0 = jc + (threadIdx.x & 31);
    if (j0 >= out_frames) return;
    const int64_t num = j0 * inrate;
    int64_t q = __double2ll_rd(__dmul_rn((double)num, rcp64));     // floor(num/outrate) +-1
    int64_t r = num - q * outrate;
    if (r < 0) { q -= 1; r += outrate; }
    else if (r >= outrate) { q += 1; r -= outrate; }
    const int64_t c = q + (r > 0 ? 1 : 0);                         // ceil
    uint32_t d = r > 0 ? (uint32_t)(outrate - r) : 0u;
    const int16_t *pin = in + c * NCH;
    int16_t *pout = out + j0 * NCH;
    uint32_t crel = 0;
    const bool full = jc + 32 * ITERS * BATCHES <= out_frames;     // warp-uniform
    if (full) {
#pragma unroll 1
        for (int b = 0; b < BATCHES; b++)
            rcv_batch<NCH, VEC, ITERS, true, FIXUP, MAGIC>(pin, pout + (int64_t)b * 32 * ITERS * NCH, 0, b == 0 && j0 == 0, crel, d, rt);
    } else {
        const int64_t left = out_frames - j0;
        for (int b = 0; b < BATCHES && (int64_t)b * 32 * ITERS < left; b++)
            rcv_batch<NCH, VEC, ITERS, false, FIXUP, MAGIC>(pin, pout + (int64_t)b * 32 * ITERS * NCH,
                                              (int)min(left - (int64_t)b * 32 * ITERS, (int64_t)(32 * ITERS)),
                                              b == 0 && j0 == 0, crel, d, rt);
    }
}

#ifndef RCV_ITERS
#define RCV_ITERS 8
#endif
#ifndef RCV_BATCHES
#define RCV_BATCHES 4
#endif
#ifndef RCV_MIN_BLOCKS
#define RCV_MIN_BLOCKS 4
#endif
constexpr int kRcvIters = RCV_ITERS;     // frames a lane has in flight
constexpr int kRcvBatches = RCV_BATCHES; // a warp owns 32*kRcvIters*kRcvBatches consecutive output frames

template <int NCH, bool VEC>
__global__ void __launch_bounds__(kPBlock, RCV_MIN_BLOCKS)
pcm16_ratecv_kernel(const int16_t *__restrict__ in, int64_t inrate, int64_t outrate, int64_t out_frames,
                    int16_t *__restrict__ out)
{
    const int64_t warp_id = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const RcvRate rt = rcv_rate(inrate, outrate);
    const double rcp64 = 1.0 / (double)outrate;
    const int64_t chunk = 32 * kRcvIters * kRcvBatches;
    if (outrate <= 16384) {
        for (int64_t jc = warp_id * chunk; jc < out_frames; jc += nwarps * chunk)
            rcv_chunk<NCH, VEC, kRcvIters, kRcvBatches, false, true>(in, out, jc, out_frames, inrate, outrate, rt, rcp64);
    } else if (outrate <= 65536) {
        for (int64_t jc = warp_id * chunk; jc < out_frames; jc += nwarps * chunk)
            rcv_chunk<NCH, VEC, kRcvIters, kRcvBatches, false, false>(in, out, jc, out_frames, inrate, outrate, rt, rcp64);
    } else {
        for (int64_t jc = warp_id * chunk; jc < out_frames; jc += nwarps * chunk)
            rcv_chunk<NCH, VEC, kRcvIters, kRcvBatches, true, false>(in, out, jc, out_frames, inrate, outrate, rt, rcp64);
    }
}

// Sample.mix_at for many sources, in order.  Thread <-> destination sample; the sources that
// overlap a destination tile are found through a per-tile candidate list built on the fly:
// sources are visited in index order and each applies one saturating add, exactly the
// sequence of audioop.add calls upstream would make.  A block first filters, cooperatively
// and in order, the sources that touch its tile into shared memory.
constexpr int kMixTile = 2048;           // destination samples per block step
constexpr int kMixList = 1024;           // candidate list chunk

__global__ void __launch_bounds__(kPBlock)
pcm16_mix_at_many_kernel(int16_t *__restrict__ dst, int64_t dst_count, const int16_t *__restrict__ pool,
                         const int64_t *__restrict__ src_off, const int64_t *__restrict__ src_len,
                         const int64_t *__restrict__ dst_off, int64_t nsrc)
{
    __shared__ int s_list[kMixList];
    __shared__ int s_count;
    __shared__ int s_warp_counts[kPBlock / 32];
    const int64_t ntiles = (dst_count + kMixTile - 1) / kMixTile;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t t0 = tile * kMixTile;
        const int64_t t1 = min(t0 + (int64_t)kMixTile, dst_count);
        int acc[kMixTile / kPBlock];
#pragma unroll
        for (int k = 0; k < kMixTile / kPBlock; k++) {
            int64_t i = t0 + threadIdx.x + (int64_t)k * kPBlock;
            acc[k] = (i < t1) ? dst[i] : 0;
        }
        bool touched = false;
        for (int64_t sbase = 0; sbase < nsrc; sbase += kMixList) {
            // ordered compaction of the sources in [sbase, sbase+kMixList) that overlap the tile
            if (threadIdx.x == 0) s_count = 0;
            __syncthreads();
            for (int64_t c0 = sbase; c0 < min(sbase + (int64_t)kMixList, nsrc); c0 += kPBlock) {
                int64_t s = c0 + threadIdx.x;
                bool hit = false;
                if (s < nsrc && s < sbase + kMixList) {
                    int64_t o = dst_off[s], l = src_len[s];
                    hit = (l > 0) && (o >= 0) && (o < t1) && (o + l > t0);
                }
                unsigned ballot = __ballot_sync(0xffffffffu, hit);
                if (lane == 0) s_warp_counts[warp] = __popc(ballot);
                __syncthreads();
                int before = s_count;
                for (int w = 0; w < warp; w++) before += s_warp_counts[w];
                if (hit) s_list[before + __popc(ballot & ((1u << lane) - 1))] = (int)(s - sbase);
                __syncthreads();
                if (threadIdx.x == 0) {
                    int tot = 0;
                    for (int w = 0; w < kPBlock / 32; w++) tot += s_warp_counts[w];
                    s_count += tot;
                }
                __syncthreads();
            }
            const int nlist = s_count;
            for (int q = 0; q < nlist; q++) {
                const int64_t s = sbase + s_list[q];
                const int64_t o = dst_off[s], l = src_len[s];
                const int16_t *src = pool + src_off[s];
#pragma unroll
                for (int k = 0; k < kMixTile / kPBlock; k++) {
                    int64_t i = t0 + threadIdx.x + (int64_t)k * kPBlock;
                    int64_t r = i - o;
                    if (i < t1 && r >= 0 && r < l) acc[k] = sat16(acc[k] + (int)src[r]);
                }
            }
            touched |= (nlist > 0);
            __syncthreads();
        }
        if (touched) {
#pragma unroll
            for (int k = 0; k < kMixTile / kPBlock; k++) {
                int64_t i = t0 + threadIdx.x + (int64_t)k * kPBlock;
                if (i < t1) dst[i] = (int16_t)acc[k];
            }
        }
    }
}

// ---- indexed variant: per-tile source lists built on the device (count, allocate, fill), so that a
// tile only visits the sources that touch it.  The lists come out of the atomics unordered;
// the apply kernel rank-sorts each (short) list in shared memory to restore source order.
// (a warp per source, a lane per tile it touches: the atomics of one source are all in flight at once)
// (a source with a negative destination offset is left out and counted in status word 1)
__global__ void mix_count_kernel(const int64_t *__restrict__ src_len, const int64_t *__restrict__ dst_off,
                                 int64_t nsrc, int64_t dst_count, int *__restrict__ tile_count, int *__restrict__ status)
{
    const int64_t s = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (s >= nsrc) return;
    const int64_t o = dst_off[s], l = src_len[s];
    if (o < 0) {
        if ((threadIdx.x & 31) == 0 && l > 0) atomicAdd(status + 1, 1);
        return;
    }
    if (l <= 0 || o >= dst_count) return;
    const int64_t last = min(o + l, dst_count) - 1;
    for (int64_t t = o / kMixTile + (threadIdx.x & 31); t <= last / kMixTile; t += 32) atomicAdd(&tile_count[t], 1);
}

// Gives every tile a private range of the list: ranges need not be in tile order (each list is
// rank-sorted when it is used), so a warp-aggregated atomic on one cursor replaces a scan.
// (list_cap = entries the caller's scratch has room for: a range that would pass it is refused -
// the tile's count is set to 0, so nothing is mixed there - and status word 2 is raised)
__global__ void mix_alloc_kernel(int *__restrict__ tile_count, int64_t ntiles, int *__restrict__ cursor,
                                 int *__restrict__ tile_start, int *__restrict__ tile_fill, int64_t list_cap)
{
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const int v = (t < ntiles) ? tile_count[t] : 0;
    int inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    int base = 0;
    if (lane == 31 && inc) base = atomicAdd(cursor, inc);
    base = __shfl_sync(0xffffffffu, base, 31);
    if (t < ntiles) {
        const int64_t first = (int64_t)base + inc - v;
        const bool fits = first + v <= list_cap;
        tile_start[t] = fits ? (int)first : 0;
        tile_fill[t] = fits ? 0 : 0x40000000;              // the fill kernel skips refused tiles
        if (!fits) {
            tile_count[t] = 0;
            if (v) atomicOr(cursor + 2, 1);
        }
    }
}

// A list entry carries the source's placement with its index: the apply kernel then goes from a
// tile's list straight to the samples (one dependent round trip to global memory less per tile; the
// kernel is bound by that chain of latencies, not by bandwidth).
struct alignas(16) MixEntry {
    int64_t o, l;            // destination offset and length of the source
    const int16_t *p;        // its first sample
    int s, pad;              // its index (sources apply in index order)
};
static_assert(sizeof(MixEntry) == 32, "MixEntry layout");

__global__ void mix_fill_kernel(const int16_t *__restrict__ pool, const int64_t *__restrict__ src_off,
                                const int64_t *__restrict__ src_len, const int64_t *__restrict__ dst_off,
                                int64_t nsrc, int64_t dst_count, const int *__restrict__ tile_start,
                                int *__restrict__ tile_fill, MixEntry *__restrict__ list)
{
    const int64_t s = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (s >= nsrc) return;
    const int64_t o = dst_off[s], l = src_len[s];
    if (l <= 0 || o < 0 || o >= dst_count) return;
    MixEntry e;
    e.o = o; e.l = l; e.p = pool + src_off[s]; e.s = (int)s; e.pad = 0;
    const int64_t last = min(o + l, dst_count) - 1;
    for (int64_t t = o / kMixTile + (threadIdx.x & 31); t <= last / kMixTile; t += 32) {
        const int slot = atomicAdd(&tile_fill[t], 1);
        if (slot < 0x40000000) list[tile_start[t] + slot] = e;
    }
}

constexpr int kMixMeta = 128;            // sources per list chunk whose placement is staged in shared memory
__global__ void __launch_bounds__(kPBlock, 8)
mix_apply_kernel(int16_t *__restrict__ dst, int64_t dst_count, const int16_t *__restrict__ pool,
                 const int64_t *__restrict__ src_off, const int64_t *__restrict__ src_len,
                 const int64_t *__restrict__ dst_off, const int *__restrict__ tile_count,
                 const int *__restrict__ tile_start, const MixEntry *__restrict__ list)
{
    __shared__ int s_raw[kMixList];
    __shared__ int s_sorted[kMixList];
    __shared__ int64_t s_o[kMixMeta], s_l[kMixMeta];
    __shared__ const int16_t *s_p[kMixMeta];
    const int64_t ntiles = (dst_count + kMixTile - 1) / kMixTile;
    constexpr int kPer = kMixTile / kPBlock;           // 8 consecutive destination samples per thread
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int n = tile_count[tile];
        if (n == 0) continue;                          // block-uniform: nothing mixes into this tile
        const int begin = tile_start[tile];
        const int64_t t0 = tile * kMixTile;
        const int64_t t1 = min(t0 + (int64_t)kMixTile, dst_count);
        const int64_t i0 = t0 + (int64_t)threadIdx.x * kPer;
        const int live = (int)max((int64_t)0, min((int64_t)kPer, t1 - i0));    // this thread's samples inside the buffer
        int acc[kPer];
        const bool full = (live == kPer) && ((((uintptr_t)(dst + i0)) & 15) == 0);
        if (full) {
            Vec8 v;
            v.v = *reinterpret_cast<const int4 *>(dst + i0);
#pragma unroll
            for (int k = 0; k < kPer; k++) acc[k] = v.s[k];
        } else {
#pragma unroll
            for (int k = 0; k < kPer; k++) acc[k] = (k < live) ? dst[i0 + k] : 0;
        }
        // the first mm staged sources, in order, onto this thread's run.  All per-sample index arithmetic
        // is 32-bit: the 64-bit part is done once per (thread, source).
        // (Loading several sources' samples before adding any was tried and is slower: empty batch
        // slots cost instructions.)
        auto apply_staged = [&](const int mm) {
            for (int q = 0; q < mm; q++) {
                const int64_t r0 = i0 - s_o[q];                 // source index of this thread's first sample
                // samples k in [klo, khi) of the run are covered by the source (none if khi <= klo)
                const int klo = (int)max((int64_t)0, min((int64_t)kPer, -r0));
                const int khi = (int)max((int64_t)0, min((int64_t)live, s_l[q] - r0));
                if (khi <= klo) continue;
                const int16_t *p0 = s_p[q] + r0;
                if (klo == 0 && khi == kPer) {
                    // the source covers the whole run (all but the threads at its two ends):
                    // no per-sample tests
#pragma unroll
                    for (int k = 0; k < kPer; k++) acc[k] = sat16(acc[k] + (int)p0[k]);
                } else {
#pragma unroll
                    for (int k = 0; k < kPer; k++)
                        if (k >= klo && k < khi) acc[k] = sat16(acc[k] + (int)p0[k]);
                }
            }
        };
        if (n <= kMixMeta) {
            // the usual tile (a few sources): entries to shared memory in index order, one pass
            __syncthreads();
            MixEntry e = {};
            if ((int)threadIdx.x < n) {
                e = list[begin + threadIdx.x];
                s_raw[threadIdx.x] = e.s;
            }
            __syncthreads();
            if ((int)threadIdx.x < n) {
                int rank = 0;
                for (int u = 0; u < n; u++) rank += (s_raw[u] < e.s) ? 1 : 0;
                s_o[rank] = e.o;
                s_l[rank] = e.l;
                s_p[rank] = e.p;
            }
            __syncthreads();
            apply_staged(n);
        } else
        // the list in chunks of kMixList, each rank-sorted so that sources apply in index order;
        // chunks themselves are ordered by splitting on the source index range
        for (int lo = 0; lo < n; lo += kMixList) {
            // (n > kMixList is rare: then every chunk pass selects the next kMixList smallest)
            __syncthreads();
            const int m = min(kMixList, n - lo);
            if (n <= kMixList) {
                for (int i = threadIdx.x; i < m; i += kPBlock) s_raw[i] = list[begin + i].s;
                __syncthreads();
                for (int i = threadIdx.x; i < m; i += kPBlock) {
                    const int mine = s_raw[i];
                    int rank = 0;
                    for (int u = 0; u < m; u++) rank += (s_raw[u] < mine) ? 1 : 0;
                    s_sorted[rank] = mine;
                }
            } else {
                // rank over the whole list in global memory; keep ranks [lo, lo+m)
                for (int i = threadIdx.x; i < n; i += kPBlock) {
                    const int mine = list[begin + i].s;
                    int rank = 0;
                    for (int u = 0; u < n; u++) rank += (list[begin + u].s < mine) ? 1 : 0;
                    if (rank >= lo && rank < lo + m) s_sorted[rank - lo] = mine;
                }
            }
            __syncthreads();
            // Stage the placements of the chunk's sources with one round trip (instead of three
            // dependent loads per source), then apply the sources in order.  All per-sample index
            // arithmetic is 32-bit: the 64-bit part is done once per (thread, source).
            // (Loading several sources' samples before adding any was tried and is slower: the
            // kernel is bound by instruction issue, and empty batch slots cost instructions.)
            for (int mb = 0; mb < m; mb += kMixMeta) {
                const int mm = min(kMixMeta, m - mb);
                if (mb) __syncthreads();
                for (int i = threadIdx.x; i < mm; i += kPBlock) {
                    const int64_t s = s_sorted[mb + i];
                    s_o[i] = dst_off[s];
                    s_l[i] = src_len[s];
                    s_p[i] = pool + src_off[s];
                }
                __syncthreads();
                apply_staged(mm);
            }
        }
        if (full) {
            Vec8 v;
#pragma unroll
            for (int k = 0; k < kPer; k++) v.s[k] = (short)acc[k];
            *reinterpret_cast<int4 *>(dst + i0) = v.v;
        } else {
#pragma unroll
            for (int k = 0; k < kPer; k++)
                if (k < live) dst[i0 + k] = (int16_t)acc[k];
        }
    }
}

}  // namespace sb

using namespace sb;

#define PCM_GRID(count) grid_for((count) / 8 + 1, kPBlock, kPMaxBlocks)

extern "C" int sb_pcm16_add(const int16_t *a_dev, const int16_t *b_dev, int64_t count, int16_t *out_dev, void *stream)
{
    SB_REQUIRE(count >= 0 && (count == 0 || (a_dev && b_dev && out_dev)), "sb_pcm16_add: bad arguments");
    if (!count) return SB_OK;
    pcm16_add_kernel<<<PCM_GRID(count), kPBlock, 0, (cudaStream_t)stream>>>(a_dev, b_dev, count, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm16_mul(const int16_t *in_dev, int64_t count, double factor, int16_t *out_dev, void *stream)
{
    SB_REQUIRE(count >= 0 && (count == 0 || (in_dev && out_dev)), "sb_pcm16_mul: bad arguments");
    if (!count) return SB_OK;
    if (fabs(factor) <= 16384.0)
        pcm16_mul_kernel<true><<<PCM_GRID(count), kPBlock, 0, (cudaStream_t)stream>>>(in_dev, count, factor, out_dev);
    else
        pcm16_mul_kernel<false><<<PCM_GRID(count), kPBlock, 0, (cudaStream_t)stream>>>(in_dev, count, factor, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm16_bias(const int16_t *in_dev, int64_t count, int bias, int16_t *out_dev, void *stream)
{
    SB_REQUIRE(in_dev && out_dev && count >= 0, "sb_pcm16_bias: bad arguments");
    if (!count) return SB_OK;
    pcm16_bias_kernel<<<PCM_GRID(count), kPBlock, 0, (cudaStream_t)stream>>>(in_dev, count, bias, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm16_reverse(const int16_t *in_dev, int64_t count, int16_t *out_dev, void *stream)
{
    SB_REQUIRE(in_dev && out_dev && in_dev != out_dev && count >= 0, "sb_pcm16_reverse: bad arguments");
    if (!count) return SB_OK;
    pcm16_reverse_kernel<<<grid_for(count, kPBlock, kPMaxBlocks), kPBlock, 0, (cudaStream_t)stream>>>(in_dev, count, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm16_maxabs(const int16_t *in_dev, int64_t count, unsigned int *out_dev, void *stream)
{
    SB_REQUIRE(out_dev && count >= 0 && (in_dev || !count), "sb_pcm16_maxabs: bad arguments");
    SB_CUDA_CHECK(cudaMemsetAsync(out_dev, 0, sizeof(unsigned int), (cudaStream_t)stream));
    if (!count) return SB_OK;
    pcm16_maxabs_kernel<<<grid_for(count, kPBlock, 148 * 4), kPBlock, 0, (cudaStream_t)stream>>>(in_dev, count, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm16_tomono(const int16_t *in_dev, int64_t frames, double lfactor, double rfactor,
                               int16_t *out_dev, void *stream)
{
    SB_REQUIRE(in_dev && out_dev && frames >= 0, "sb_pcm16_tomono: bad arguments");
    if (!frames) return SB_OK;
    const int grid = grid_for(frames / 4 + 1, kPBlock, kPMaxBlocks);
    if (fabs(lfactor) <= 16384.0 && fabs(rfactor) <= 16384.0)
        pcm16_tomono_kernel<true><<<grid, kPBlock, 0, (cudaStream_t)stream>>>(in_dev, frames, lfactor, rfactor, out_dev);
    else
        pcm16_tomono_kernel<false><<<grid, kPBlock, 0, (cudaStream_t)stream>>>(in_dev, frames, lfactor, rfactor, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm16_tostereo(const int16_t *in_dev, int64_t frames, double lfactor, double rfactor,
                                 int16_t *out_dev, void *stream)
{
    SB_REQUIRE(in_dev && out_dev && frames >= 0, "sb_pcm16_tostereo: bad arguments");
    if (!frames) return SB_OK;
    const int grid = grid_for(frames / 4 + 1, kPBlock, kPMaxBlocks);
    if (fabs(lfactor) <= 16384.0 && fabs(rfactor) <= 16384.0)
        pcm16_tostereo_kernel<true><<<grid, kPBlock, 0, (cudaStream_t)stream>>>(in_dev, frames, lfactor, rfactor, out_dev);
    else
        pcm16_tostereo_kernel<false><<<grid, kPBlock, 0, (cudaStream_t)stream>>>(in_dev, frames, lfactor, rfactor, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm16_to_pcm32(const int16_t *in_dev, int64_t count, int32_t *out_dev, void *stream)
{
    SB_REQUIRE(in_dev && out_dev && count >= 0, "sb_pcm16_to_pcm32: bad arguments");
    if (!count) return SB_OK;
    pcm16_to_32_kernel<<<grid_for(count, kPBlock, kPMaxBlocks), kPBlock, 0, (cudaStream_t)stream>>>(in_dev, count, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm32_to_pcm16(const int32_t *in_dev, int64_t count, int16_t *out_dev, void *stream)
{
    SB_REQUIRE(in_dev && out_dev && count >= 0, "sb_pcm32_to_pcm16: bad arguments");
    if (!count) return SB_OK;
    pcm32_to_16_kernel<<<grid_for(count, kPBlock, kPMaxBlocks), kPBlock, 0, (cudaStream_t)stream>>>(in_dev, count, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm16_fadeout(int16_t *buf_dev, int64_t nsamples, double target_volume, void *stream)
{
    SB_REQUIRE(nsamples >= 0 && (nsamples == 0 || buf_dev), "sb_pcm16_fadeout: bad arguments");
    if (!nsamples) return SB_OK;
    if (nsamples < (1LL << 31) && fabs(target_volume) <= 16384.0)
        pcm16_fadeout_kernel<true><<<PCM_GRID(nsamples), kPBlock, 0, (cudaStream_t)stream>>>(buf_dev, nsamples, 1.0 - target_volume);
    else
        pcm16_fadeout_kernel<false><<<PCM_GRID(nsamples), kPBlock, 0, (cudaStream_t)stream>>>(buf_dev, nsamples, 1.0 - target_volume);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm16_fadein(int16_t *buf_dev, int64_t nsamples, double start_volume, void *stream)
{
    SB_REQUIRE(buf_dev && nsamples >= 0, "sb_pcm16_fadein: bad arguments");
    if (!nsamples) return SB_OK;
    if (nsamples < (1LL << 31) && fabs(start_volume) <= 16384.0)
        pcm16_fadein_kernel<true><<<PCM_GRID(nsamples), kPBlock, 0, (cudaStream_t)stream>>>(buf_dev, nsamples, 1.0 - start_volume,
                                                                                          start_volume);
    else
        pcm16_fadein_kernel<false><<<PCM_GRID(nsamples), kPBlock, 0, (cudaStream_t)stream>>>(buf_dev, nsamples, 1.0 - start_volume,
                                                                                           start_volume);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm16_modulate_cycle_f64(int16_t *buf_dev, const double *mod_dev, int64_t mod_len, int64_t count,
                                           void *stream)
{
    SB_REQUIRE(buf_dev && mod_dev && mod_len >= 1 && count >= 0, "sb_pcm16_modulate_cycle_f64: bad arguments");
    if (!count) return SB_OK;
    pcm_modulate_cycle_kernel<int16_t><<<grid_for(count, kPBlock, kPMaxBlocks), kPBlock, 0, (cudaStream_t)stream>>>(
        buf_dev, mod_dev, mod_len, count);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm32_modulate_cycle_f64(int32_t *buf_dev, const double *mod_dev, int64_t mod_len, int64_t count,
                                           void *stream)
{
    SB_REQUIRE(buf_dev && mod_dev && mod_len >= 1 && count >= 0, "sb_pcm32_modulate_cycle_f64: bad arguments");
    if (!count) return SB_OK;
    pcm_modulate_cycle_kernel<int32_t><<<grid_for(count, kPBlock, kPMaxBlocks), kPBlock, 0, (cudaStream_t)stream>>>(
        buf_dev, mod_dev, mod_len, count);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm16_modulate(int16_t *buf_dev, const float *mod_dev, int64_t count, void *stream)
{
    SB_REQUIRE(buf_dev && mod_dev && count >= 0, "sb_pcm16_modulate: bad arguments");
    if (!count) return SB_OK;
    pcm16_modulate_kernel<<<grid_for(count, kPBlock, kPMaxBlocks), kPBlock, 0, (cudaStream_t)stream>>>(buf_dev, mod_dev, count);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

static int64_t gcd64(int64_t a, int64_t b)
{
    while (b) { int64_t t = a % b; a = b; b = t; }
    return a;
}

extern "C" int64_t sb_pcm16_ratecv_frames(int64_t in_frames, int64_t inrate, int64_t outrate)
{
    if (in_frames <= 0 || inrate <= 0 || outrate <= 0) return 0;
    int64_t g = gcd64(inrate, outrate);
    inrate /= g;
    outrate /= g;
    return ((in_frames - 1) * outrate) / inrate + 1;
}

extern "C" int sb_pcm16_ratecv(const int16_t *in_dev, int64_t in_frames, int nchannels, int64_t inrate,
                               int64_t outrate, int16_t *out_dev, void *stream)
{
    SB_REQUIRE(in_dev && out_dev && in_frames >= 0 && (nchannels == 1 || nchannels == 2) && inrate > 0 && outrate > 0,
               "sb_pcm16_ratecv: bad arguments");
    if (!in_frames) return SB_OK;
    int64_t g = gcd64(inrate, outrate);
    inrate /= g;
    outrate /= g;
    // prev*d + cur*(outrate-d) must stay exact in fp64: |sample<<16| < 2^31, so outrate < 2^21
    SB_REQUIRE(outrate < (1 << 21) && inrate < (1 << 21), "sb_pcm16_ratecv: reduced rates must be below 2^21");
    int64_t out_frames = ((in_frames - 1) * outrate) / inrate + 1;
    const int per_lane = kRcvIters * kRcvBatches;
    const int grid = grid_for((out_frames + per_lane - 1) / per_lane, kPBlock, kPMaxBlocks);
    const bool vec = (((uintptr_t)in_dev | (uintptr_t)out_dev) & 3) == 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (nchannels == 1)
        pcm16_ratecv_kernel<1, false><<<grid, kPBlock, 0, st>>>(in_dev, inrate, outrate, out_frames, out_dev);
    else if (vec)
        pcm16_ratecv_kernel<2, true><<<grid, kPBlock, 0, st>>>(in_dev, inrate, outrate, out_frames, out_dev);
    else
        pcm16_ratecv_kernel<2, false><<<grid, kPBlock, 0, st>>>(in_dev, inrate, outrate, out_frames, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

// ---- many short samples resampled to one rate in ONE launch (config 4: thousands of drum hits).
// Source s is in_pool[in_off[s] .. + in_frames[s]*nch) at rate inrate[s]; its output goes to
// out_pool[out_off[s] .. + out_frames[s]*nch).  blockIdx.y walks the sources, blockIdx.x
// (grid-stride) the output frames of a source.  Same arithmetic as pcm16_ratecv_kernel.
constexpr int kRcvManyIters = 4;

__global__ void __launch_bounds__(kPBlock)
pcm16_ratecv_many_kernel(const int16_t *__restrict__ in_pool, const int64_t *__restrict__ in_off,
                         const int64_t *__restrict__ in_frames, const int64_t *__restrict__ inrate_red,
                         const int64_t *__restrict__ outrate_red, const int64_t *__restrict__ out_off,
                         const int64_t *__restrict__ out_frames, int64_t nsrc, int nch, int16_t *__restrict__ out_pool)
{
    const bool pools_aligned = (((uintptr_t)in_pool | (uintptr_t)out_pool) & 3) == 0;
    for (int64_t s = blockIdx.y; s < nsrc; s += gridDim.y) {
        const int16_t *in = in_pool + in_off[s];
        int16_t *out = out_pool + out_off[s];
        const int64_t nout = out_frames[s], inr = inrate_red[s], outr = outrate_red[s];
        if (inr == outr) {                                  // same rate: audioop.ratecv copies
            for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nout * nch; i += (int64_t)gridDim.x * blockDim.x)
                out[i] = in[i];
            continue;
        }
        // the block's warps take 32*kRcvManyIters-frame chunks of this source in turn
        const RcvRate rt = rcv_rate(inr, outr);
        const double rcp64 = 1.0 / (double)outr;
        const int64_t chunk = 32 * kRcvManyIters;
        const int64_t warp_id = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
        const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
        const bool vec = ((in_off[s] | out_off[s]) & 1) == 0 && pools_aligned;
        for (int64_t jc = warp_id * chunk; jc < nout; jc += nwarps * chunk) {
            if (nch == 1) rcv_chunk<1, false, kRcvManyIters, 1, true, false>(in, out, jc, nout, inr, outr, rt, rcp64);
            else if (vec) rcv_chunk<2, true, kRcvManyIters, 1, true, false>(in, out, jc, nout, inr, outr, rt, rcp64);
            else rcv_chunk<2, false, kRcvManyIters, 1, true, false>(in, out, jc, nout, inr, outr, rt, rcp64);
        }
    }
}

extern "C" int sb_pcm16_ratecv_many(const int16_t *in_pool_dev, const int64_t *in_off_dev, const int64_t *in_frames_dev,
                                    const int64_t *inrate_red_dev, const int64_t *outrate_red_dev,
                                    const int64_t *out_off_dev, const int64_t *out_frames_dev, int64_t nsrc,
                                    int nchannels, int64_t max_out_frames, int16_t *out_pool_dev, void *stream)
{
    SB_REQUIRE(nsrc >= 0 && (nchannels == 1 || nchannels == 2) && max_out_frames >= 0, "sb_pcm16_ratecv_many: bad arguments");
    if (!nsrc || !max_out_frames) return SB_OK;
    SB_REQUIRE(in_pool_dev && in_off_dev && in_frames_dev && inrate_red_dev && outrate_red_dev && out_off_dev &&
                   out_frames_dev && out_pool_dev, "sb_pcm16_ratecv_many: null pointer");
    dim3 grid((unsigned)grid_for((max_out_frames + kRcvManyIters - 1) / kRcvManyIters, kPBlock, 64), (unsigned)(nsrc < 65535 ? nsrc : 65535));
    pcm16_ratecv_many_kernel<<<grid, kPBlock, 0, (cudaStream_t)stream>>>(
        in_pool_dev, in_off_dev, in_frames_dev, inrate_red_dev, outrate_red_dev, out_off_dev, out_frames_dev, nsrc,
        nchannels, out_pool_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

// scratch layout: cursor (4 ints: list cursor, sources left out for a negative offset, list overflow
// flag, spare), tile_count, tile_start, tile_fill (ntiles ints each), then the list entries
static inline int64_t mix_list_offset(int64_t ntiles) { return ((3 * ntiles + 4) * (int64_t)sizeof(int) + 15) / 16 * 16; }

extern "C" int64_t sb_pcm16_mix_at_many_scratch_bytes(int64_t dst_count, int64_t nsrc, int64_t total_src_samples)
{
    if (dst_count <= 0 || nsrc <= 0) return 0;
    int64_t ntiles = (dst_count + kMixTile - 1) / kMixTile;
    int64_t cap = total_src_samples / kMixTile + 2 * nsrc + 1;     // every source touches <= len/tile + 2 tiles
    return mix_list_offset(ntiles) + cap * (int64_t)sizeof(MixEntry);
}

extern "C" int sb_pcm16_mix_at_many(int16_t *dst_dev, int64_t dst_count, const int16_t *src_pool_dev,
                                    const int64_t *src_off_dev, const int64_t *src_len_dev,
                                    const int64_t *dst_off_dev, int64_t nsrc, void *scratch_dev,
                                    int64_t scratch_bytes, void *stream)
{
    SB_REQUIRE(dst_dev && dst_count >= 0 && nsrc >= 0, "sb_pcm16_mix_at_many: bad arguments");
    if (!nsrc || !dst_count) return SB_OK;
    SB_REQUIRE(src_pool_dev && src_off_dev && src_len_dev && dst_off_dev, "sb_pcm16_mix_at_many: null pointer");
    SB_REQUIRE(nsrc < 0x7fffffff, "sb_pcm16_mix_at_many: too many sources");
    int64_t ntiles = (dst_count + kMixTile - 1) / kMixTile;
    cudaStream_t s = (cudaStream_t)stream;
    if (!scratch_dev) {
        // no scratch: every tile scans the whole source list
        pcm16_mix_at_many_kernel<<<grid_for(ntiles, 1, kPMaxBlocks), kPBlock, 0, s>>>(
            dst_dev, dst_count, src_pool_dev, src_off_dev, src_len_dev, dst_off_dev, nsrc);
        SB_LAUNCH_CHECK();
        return SB_OK;
    }
    SB_REQUIRE(scratch_bytes >= mix_list_offset(ntiles) + (int64_t)sizeof(MixEntry) && ((uintptr_t)scratch_dev & 15) == 0,
               "sb_pcm16_mix_at_many: scratch too small or not 16-byte aligned");
    int64_t list_cap = (scratch_bytes - mix_list_offset(ntiles)) / (int64_t)sizeof(MixEntry);
    if (list_cap > 0x3fffffff) list_cap = 0x3fffffff;
    int *cursor = (int *)scratch_dev;
    int *tile_count = cursor + 4;
    int *tile_start = tile_count + ntiles;
    int *tile_fill = tile_start + ntiles;
    MixEntry *list = (MixEntry *)((unsigned char *)scratch_dev + mix_list_offset(ntiles));
    SB_CUDA_CHECK(cudaMemsetAsync(cursor, 0, (size_t)(ntiles + 4) * sizeof(int), s));
    int sgrid = (int)((nsrc * 32 + 255) / 256);             // a warp per source
    mix_count_kernel<<<sgrid, 256, 0, s>>>(src_len_dev, dst_off_dev, nsrc, dst_count, tile_count, cursor);
    SB_LAUNCH_CHECK();
    mix_alloc_kernel<<<(unsigned)((ntiles + 255) / 256), 256, 0, s>>>(tile_count, ntiles, cursor, tile_start, tile_fill, list_cap);
    SB_LAUNCH_CHECK();
    mix_fill_kernel<<<sgrid, 256, 0, s>>>(src_pool_dev, src_off_dev, src_len_dev, dst_off_dev, nsrc, dst_count, tile_start,
                                          tile_fill, list);
    SB_LAUNCH_CHECK();
    mix_apply_kernel<<<grid_for(ntiles, 1, kPMaxBlocks * 4), kPBlock, 0, s>>>(
        dst_dev, dst_count, src_pool_dev, src_off_dev, src_len_dev, dst_off_dev, tile_count, tile_start, list);
    SB_LAUNCH_CHECK();
    return SB_OK;
}
