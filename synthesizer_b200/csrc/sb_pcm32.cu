// 32-bit integer PCM: the same audioop routines as sb_pcm.cu for sample width 4
// (CPython Modules/audioop.c with size == 4: maxval 0x7fffffff, minval -0x80000000; add goes
// through fp64 and fbound, ratecv works on the raw 32-bit samples).  Samples are 4 bytes, so a
// warp's plain loads and stores are already full 128-byte lines; kernels are grid-stride maps.
#include "sb_common.cuh"

namespace sb {

constexpr int kBlock32 = 256;
constexpr int kMaxBlocks32 = 148 * 16;

// audioop.c fbound() for size 4
__device__ __forceinline__ int fbound32(double val)
{
    if (val > 2147483647.0) val = 2147483647.0;
    else if (val < -2147483647.0) val = -2147483648.0;
    return (int)floor(val);
}

template <class F>
__global__ void map32_kernel(const int32_t *a, const int32_t *b, int32_t *out,        // (in-place callers alias a and out)
                             int64_t count, F f)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
        out[i] = f(i, a[i], b ? b[i] : 0);
}

__global__ void pcm32_reverse_kernel(const int32_t *__restrict__ in, int64_t count, int32_t *__restrict__ out)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) out[i] = in[count - 1 - i];
}

__global__ void pcm32_maxabs_kernel(const int32_t *__restrict__ in, int64_t count, unsigned int *out)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    unsigned int m = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
        const int v = in[i];
        const unsigned int a = v < 0 ? (unsigned int)(-(long long)v) : (unsigned int)v;
        m = max(m, a);
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, d));
    if ((threadIdx.x & 31) == 0 && m) atomicMax(out, m);
}

__global__ void pcm32_tomono_kernel(const int32_t *__restrict__ in, int64_t frames, double lf, double rf,
                                    int32_t *__restrict__ out)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < frames; i += stride) {
        const int2 v = reinterpret_cast<const int2 *>(in)[i];
        out[i] = fbound32(__dadd_rn(__dmul_rn((double)v.x, lf), __dmul_rn((double)v.y, rf)));
    }
}

__global__ void pcm32_tostereo_kernel(const int32_t *__restrict__ in, int64_t frames, double lf, double rf,
                                      int32_t *__restrict__ out)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < frames; i += stride) {
        const double v = (double)in[i];
        reinterpret_cast<int2 *>(out)[i] = make_int2(fbound32(__dmul_rn(v, lf)), fbound32(__dmul_rn(v, rf)));
    }
}

__global__ void pcm32_modulate_kernel(int32_t *buf, const float *__restrict__ mod, int64_t count)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
        buf[i] = __double2int_rz(__dmul_rn((double)buf[i], (double)mod[i]));
}

// audioop.ratecv for width 4: see pcm16_ratecv_kernel; samples are already 32-bit wide
__global__ void pcm32_ratecv_kernel(const int32_t *__restrict__ in, int64_t in_frames, int nch, int64_t inrate,
                                    int64_t outrate, int64_t out_frames, int32_t *__restrict__ out)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const double rcp = 1.0 / (double)outrate;
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < out_frames; j += stride) {
        const int64_t num = j * inrate;
        const int64_t c = (num + outrate - 1) / outrate;
        const int64_t d = c * outrate - num;
        for (int ch = 0; ch < nch; ch++) {
            const long long cur = in[c * nch + ch];
            const long long prev = (c > 0) ? in[(c - 1) * nch + ch] : 0;
            // exact integer quotient == truncation of audioop's correctly rounded fp64 quotient
            // (|sum| < 2^52, outrate < 2^21; see the 16-bit kernel)
            const long long sum = prev * d + cur * (outrate - d);
            const long long mag = sum < 0 ? -sum : sum;
            long long q = (long long)__double2ll_rz(__dmul_rn((double)mag, rcp));
            const long long rem = mag - q * outrate;
            if (rem < 0) q -= 1;
            else if (rem >= outrate) q += 1;
            out[j * nch + ch] = (int32_t)(sum < 0 ? -q : q);
        }
    }
}

}  // namespace sb

using namespace sb;

#define GRID32(n) grid_for((n), kBlock32, kMaxBlocks32)
#define LAUNCH_MAP32(a, b, out, count, stream, lambda)                                              \
    do {                                                                                            \
        auto _f = lambda;                                                                           \
        map32_kernel<<<GRID32(count), kBlock32, 0, (cudaStream_t)(stream)>>>(a, b, out, count, _f); \
        SB_LAUNCH_CHECK();                                                                          \
    } while (0)

extern "C" int sb_pcm32_add(const int32_t *a_dev, const int32_t *b_dev, int64_t count, int32_t *out_dev, void *stream)
{
    SB_REQUIRE(a_dev && b_dev && out_dev && count >= 0, "sb_pcm32_add: bad arguments");
    if (!count) return SB_OK;
    LAUNCH_MAP32(a_dev, b_dev, out_dev, count, stream,
                 [] __device__(int64_t, int x, int y) { return fbound32(__dadd_rn((double)x, (double)y)); });
    return SB_OK;
}

extern "C" int sb_pcm32_mul(const int32_t *in_dev, int64_t count, double factor, int32_t *out_dev, void *stream)
{
    SB_REQUIRE(in_dev && out_dev && count >= 0, "sb_pcm32_mul: bad arguments");
    if (!count) return SB_OK;
    LAUNCH_MAP32(in_dev, (const int32_t *)nullptr, out_dev, count, stream,
                 [factor] __device__(int64_t, int x, int) { return fbound32(__dmul_rn((double)x, factor)); });
    return SB_OK;
}

extern "C" int sb_pcm32_bias(const int32_t *in_dev, int64_t count, int bias, int32_t *out_dev, void *stream)
{
    SB_REQUIRE(in_dev && out_dev && count >= 0, "sb_pcm32_bias: bad arguments");
    if (!count) return SB_OK;
    LAUNCH_MAP32(in_dev, (const int32_t *)nullptr, out_dev, count, stream,
                 [bias] __device__(int64_t, int x, int) { return (int)((unsigned int)x + (unsigned int)bias); });
    return SB_OK;
}

extern "C" int sb_pcm32_fadeout(int32_t *buf_dev, int64_t nsamples, double target_volume, void *stream)
{
    SB_REQUIRE(buf_dev && nsamples >= 0, "sb_pcm32_fadeout: bad arguments");
    if (!nsamples) return SB_OK;
    const double decrease = 1.0 - target_volume, numsamples = (double)nsamples, rcp_n = 1.0 / numsamples;
    LAUNCH_MAP32((const int32_t *)buf_dev, (const int32_t *)nullptr, buf_dev, nsamples, stream,
                 [=] __device__(int64_t k, int x, int) {
                     const double f = __dsub_rn(1.0, div_by_known(__dmul_rn((double)k, decrease), numsamples, rcp_n));
                     return __double2int_rz(__dmul_rn((double)x, f));
                 });
    return SB_OK;
}

extern "C" int sb_pcm32_fadein(int32_t *buf_dev, int64_t nsamples, double start_volume, void *stream)
{
    SB_REQUIRE(buf_dev && nsamples >= 0, "sb_pcm32_fadein: bad arguments");
    if (!nsamples) return SB_OK;
    const double increase = 1.0 - start_volume, numsamples = (double)nsamples, rcp_n = 1.0 / numsamples;
    LAUNCH_MAP32((const int32_t *)buf_dev, (const int32_t *)nullptr, buf_dev, nsamples, stream,
                 [=] __device__(int64_t k, int x, int) {
                     const double f = __dadd_rn(div_by_known(__dmul_rn((double)k, increase), numsamples, rcp_n), start_volume);
                     return __double2int_rz(__dmul_rn((double)x, f));
                 });
    return SB_OK;
}

extern "C" int sb_pcm32_reverse(const int32_t *in_dev, int64_t count, int32_t *out_dev, void *stream)
{
    SB_REQUIRE(in_dev && out_dev && in_dev != out_dev && count >= 0, "sb_pcm32_reverse: bad arguments");
    if (!count) return SB_OK;
    pcm32_reverse_kernel<<<GRID32(count), kBlock32, 0, (cudaStream_t)stream>>>(in_dev, count, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm32_maxabs(const int32_t *in_dev, int64_t count, unsigned int *out_dev, void *stream)
{
    SB_REQUIRE(out_dev && count >= 0 && (in_dev || !count), "sb_pcm32_maxabs: bad arguments");
    SB_CUDA_CHECK(cudaMemsetAsync(out_dev, 0, sizeof(unsigned int), (cudaStream_t)stream));
    if (!count) return SB_OK;
    pcm32_maxabs_kernel<<<grid_for(count, kBlock32, 148 * 4), kBlock32, 0, (cudaStream_t)stream>>>(in_dev, count, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm32_tomono(const int32_t *in_dev, int64_t frames, double lfactor, double rfactor,
                               int32_t *out_dev, void *stream)
{
    SB_REQUIRE(in_dev && out_dev && frames >= 0, "sb_pcm32_tomono: bad arguments");
    if (!frames) return SB_OK;
    pcm32_tomono_kernel<<<GRID32(frames), kBlock32, 0, (cudaStream_t)stream>>>(in_dev, frames, lfactor, rfactor, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm32_tostereo(const int32_t *in_dev, int64_t frames, double lfactor, double rfactor,
                                 int32_t *out_dev, void *stream)
{
    SB_REQUIRE(in_dev && out_dev && frames >= 0, "sb_pcm32_tostereo: bad arguments");
    if (!frames) return SB_OK;
    pcm32_tostereo_kernel<<<GRID32(frames), kBlock32, 0, (cudaStream_t)stream>>>(in_dev, frames, lfactor, rfactor, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm32_modulate(int32_t *buf_dev, const float *mod_dev, int64_t count, void *stream)
{
    SB_REQUIRE(buf_dev && mod_dev && count >= 0, "sb_pcm32_modulate: bad arguments");
    if (!count) return SB_OK;
    pcm32_modulate_kernel<<<GRID32(count), kBlock32, 0, (cudaStream_t)stream>>>(buf_dev, mod_dev, count);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm32_ratecv(const int32_t *in_dev, int64_t in_frames, int nchannels, int64_t inrate,
                               int64_t outrate, int32_t *out_dev, void *stream)
{
    SB_REQUIRE(in_dev && out_dev && in_frames >= 0 && (nchannels == 1 || nchannels == 2) && inrate > 0 && outrate > 0,
               "sb_pcm32_ratecv: bad arguments");
    if (!in_frames) return SB_OK;
    int64_t a = inrate, b = outrate;
    while (b) { int64_t t = a % b; a = b; b = t; }
    inrate /= a;
    outrate /= a;
    SB_REQUIRE(outrate < (1 << 21) && inrate < (1 << 21), "sb_pcm32_ratecv: reduced rates must be below 2^21");
    const int64_t out_frames = ((in_frames - 1) * outrate) / inrate + 1;
    pcm32_ratecv_kernel<<<GRID32(out_frames), kBlock32, 0, (cudaStream_t)stream>>>(in_dev, in_frames, nchannels, inrate,
                                                                                  outrate, out_frames, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

// ------------------------------------------------------------------ 24-bit frames (audioop size 3)
// A Sample of 3-byte frames keeps its values as int32 (sign-extended 24-bit) on the device and
// runs the size-4 kernels above; what differs from size 4 is done by these: packing, the
// 24-bit clip of audioop's fbound / overflow, and the <<8 / >>8 of GETSAMPLE32 / lin2lin.
namespace sb {
__global__ void pcm24_unpack_kernel(const uint8_t *__restrict__ in, int64_t count, int32_t *__restrict__ out)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
        const uint32_t v = (uint32_t)in[3 * i] | ((uint32_t)in[3 * i + 1] << 8) | ((uint32_t)in[3 * i + 2] << 16);
        out[i] = (int32_t)(v << 8) >> 8;
    }
}

__global__ void pcm24_pack_kernel(const int32_t *__restrict__ in, int64_t count, uint8_t *__restrict__ out)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
        const uint32_t v = (uint32_t)in[i];
        out[3 * i] = (uint8_t)v;
        out[3 * i + 1] = (uint8_t)(v >> 8);
        out[3 * i + 2] = (uint8_t)(v >> 16);
    }
}

__global__ void pcm24_clamp_kernel(int32_t *buf, int64_t count)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
        buf[i] = max(-0x800000, min(0x7fffff, buf[i]));
}

// shift > 0: left (wrapping), shift < 0: arithmetic right
__global__ void pcm32_shift_kernel(int32_t *buf, int64_t count, int shift)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
        buf[i] = shift >= 0 ? (int32_t)((uint32_t)buf[i] << shift) : (buf[i] >> (-shift));
}
}  // namespace sb

extern "C" int sb_pcm24_unpack(const uint8_t *bytes_dev, int64_t count, int32_t *out_dev, void *stream)
{
    SB_REQUIRE(count >= 0 && (count == 0 || (bytes_dev && out_dev)), "sb_pcm24_unpack: bad arguments");
    if (!count) return SB_OK;
    pcm24_unpack_kernel<<<GRID32(count), kBlock32, 0, (cudaStream_t)stream>>>(bytes_dev, count, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm24_pack(const int32_t *in_dev, int64_t count, uint8_t *bytes_dev, void *stream)
{
    SB_REQUIRE(count >= 0 && (count == 0 || (bytes_dev && in_dev)), "sb_pcm24_pack: bad arguments");
    if (!count) return SB_OK;
    pcm24_pack_kernel<<<GRID32(count), kBlock32, 0, (cudaStream_t)stream>>>(in_dev, count, bytes_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm24_clamp(int32_t *buf_dev, int64_t count, void *stream)
{
    SB_REQUIRE(count >= 0 && (count == 0 || buf_dev), "sb_pcm24_clamp: bad arguments");
    if (!count) return SB_OK;
    pcm24_clamp_kernel<<<GRID32(count), kBlock32, 0, (cudaStream_t)stream>>>(buf_dev, count);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm32_shift(int32_t *buf_dev, int64_t count, int shift, void *stream)
{
    SB_REQUIRE(count >= 0 && (count == 0 || buf_dev) && shift > -32 && shift < 32, "sb_pcm32_shift: bad arguments");
    if (!count || !shift) return SB_OK;
    pcm32_shift_kernel<<<GRID32(count), kBlock32, 0, (cudaStream_t)stream>>>(buf_dev, count, shift);
    SB_LAUNCH_CHECK();
    return SB_OK;
}
