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

union Vec8 {
    int4 v;
    short s[8];
};

// generic unary/binary streaming map over int16 with a 128-bit fast path
template <class F>
__device__ __forceinline__ void map16(const int16_t *__restrict__ a, const int16_t *__restrict__ b,
                                      int16_t *__restrict__ out, int64_t count, F f)
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

__global__ void pcm16_mul_kernel(const int16_t *a, int64_t count, double factor, int16_t *out)
{
    map16(a, nullptr, out, count, [factor](int64_t, int x, int) { return fbound16(__dmul_rn((double)x, factor)); });
}

__global__ void pcm16_bias_kernel(const int16_t *a, int64_t count, int bias, int16_t *out)
{
    map16(a, nullptr, out, count,
          [bias](int64_t, int x, int) { return (int)(int16_t)(uint16_t)((uint32_t)(uint16_t)x + (uint32_t)bias); });
}

__global__ void pcm16_fadeout_kernel(int16_t *buf, int64_t nsamples, double decrease)
{
    const double numsamples = (double)nsamples;
    map16(buf, nullptr, buf, nsamples, [=](int64_t k, int x, int) {
        // int(sample * (1.0 - k*decrease/numsamples)), every operation rounded like CPython's
        double f = __dsub_rn(1.0, __ddiv_rn(__dmul_rn((double)k, decrease), numsamples));
        return __double2int_rz(__dmul_rn((double)x, f));
    });
}

__global__ void pcm16_fadein_kernel(int16_t *buf, int64_t nsamples, double increase, double start_volume)
{
    const double numsamples = (double)nsamples;
    map16(buf, nullptr, buf, nsamples, [=](int64_t k, int x, int) {
        double f = __dadd_rn(__ddiv_rn(__dmul_rn((double)k, increase), numsamples), start_volume);
        return __double2int_rz(__dmul_rn((double)x, f));
    });
}

__global__ void pcm16_modulate_kernel(int16_t *buf, const float *__restrict__ mod, int64_t count)
{
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
        buf[i] = (int16_t)__double2int_rz(__dmul_rn((double)buf[i], (double)mod[i]));
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

__global__ void pcm16_tomono_kernel(const int16_t *__restrict__ in, int64_t frames, double lf, double rf,
                                    int16_t *__restrict__ out)
{
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < frames; i += stride) {
        short2 v = reinterpret_cast<const short2 *>(in)[i];
        double val = __dadd_rn(__dmul_rn((double)v.x, lf), __dmul_rn((double)v.y, rf));
        out[i] = (int16_t)fbound16(val);
    }
}

__global__ void pcm16_tostereo_kernel(const int16_t *__restrict__ in, int64_t frames, double lf, double rf,
                                      int16_t *__restrict__ out)
{
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < frames; i += stride) {
        double v = (double)in[i];
        short2 o;
        o.x = (short)fbound16(__dmul_rn(v, lf));
        o.y = (short)fbound16(__dmul_rn(v, rf));
        reinterpret_cast<short2 *>(out)[i] = o;
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
//   then >>16.  The fp64 expression is evaluated exactly as audioop.c writes it.
__global__ void pcm16_ratecv_kernel(const int16_t *__restrict__ in, int64_t in_frames, int nch, int64_t inrate,
                                    int64_t outrate, int64_t out_frames, int16_t *__restrict__ out)
{
    int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const double outr = (double)outrate;
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < out_frames; j += stride) {
        int64_t num = j * inrate;
        int64_t c = (num + outrate - 1) / outrate;
        int64_t d = c * outrate - num;
        for (int ch = 0; ch < nch; ch++) {
            int cur = (int)((uint32_t)(int32_t)in[c * nch + ch] << 16);
            int prev = (c > 0) ? (int)((uint32_t)(int32_t)in[(c - 1) * nch + ch] << 16) : 0;
            double v = __ddiv_rn(__dadd_rn(__dmul_rn((double)prev, (double)d),
                                           __dmul_rn((double)cur, (double)(outrate - d))), outr);
            int o = __double2int_rz(v);
            out[j * nch + ch] = (int16_t)(o >> 16);
        }
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
                    hit = (l > 0) && (o < t1) && (o + l > t0);
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

}  // namespace sb

using namespace sb;

#define PCM_GRID(count) grid_for((count) / 8 + 1, kPBlock, kPMaxBlocks)

extern "C" int sb_pcm16_add(const int16_t *a_dev, const int16_t *b_dev, int64_t count, int16_t *out_dev, void *stream)
{
    SB_REQUIRE(a_dev && b_dev && out_dev && count >= 0, "sb_pcm16_add: bad arguments");
    if (!count) return SB_OK;
    pcm16_add_kernel<<<PCM_GRID(count), kPBlock, 0, (cudaStream_t)stream>>>(a_dev, b_dev, count, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm16_mul(const int16_t *in_dev, int64_t count, double factor, int16_t *out_dev, void *stream)
{
    SB_REQUIRE(in_dev && out_dev && count >= 0, "sb_pcm16_mul: bad arguments");
    if (!count) return SB_OK;
    pcm16_mul_kernel<<<PCM_GRID(count), kPBlock, 0, (cudaStream_t)stream>>>(in_dev, count, factor, out_dev);
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
    pcm16_tomono_kernel<<<grid_for(frames, kPBlock, kPMaxBlocks), kPBlock, 0, (cudaStream_t)stream>>>(
        in_dev, frames, lfactor, rfactor, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm16_tostereo(const int16_t *in_dev, int64_t frames, double lfactor, double rfactor,
                                 int16_t *out_dev, void *stream)
{
    SB_REQUIRE(in_dev && out_dev && frames >= 0, "sb_pcm16_tostereo: bad arguments");
    if (!frames) return SB_OK;
    pcm16_tostereo_kernel<<<grid_for(frames, kPBlock, kPMaxBlocks), kPBlock, 0, (cudaStream_t)stream>>>(
        in_dev, frames, lfactor, rfactor, out_dev);
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
    SB_REQUIRE(buf_dev && nsamples >= 0, "sb_pcm16_fadeout: bad arguments");
    if (!nsamples) return SB_OK;
    pcm16_fadeout_kernel<<<PCM_GRID(nsamples), kPBlock, 0, (cudaStream_t)stream>>>(buf_dev, nsamples, 1.0 - target_volume);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm16_fadein(int16_t *buf_dev, int64_t nsamples, double start_volume, void *stream)
{
    SB_REQUIRE(buf_dev && nsamples >= 0, "sb_pcm16_fadein: bad arguments");
    if (!nsamples) return SB_OK;
    pcm16_fadein_kernel<<<PCM_GRID(nsamples), kPBlock, 0, (cudaStream_t)stream>>>(buf_dev, nsamples, 1.0 - start_volume,
                                                                                   start_volume);
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
    pcm16_ratecv_kernel<<<grid_for(out_frames, kPBlock, kPMaxBlocks), kPBlock, 0, (cudaStream_t)stream>>>(
        in_dev, in_frames, nchannels, inrate, outrate, out_frames, out_dev);
    SB_LAUNCH_CHECK();
    return SB_OK;
}

extern "C" int sb_pcm16_mix_at_many(int16_t *dst_dev, int64_t dst_count, const int16_t *src_pool_dev,
                                    const int64_t *src_off_dev, const int64_t *src_len_dev,
                                    const int64_t *dst_off_dev, int64_t nsrc, void *stream)
{
    SB_REQUIRE(dst_dev && dst_count >= 0 && nsrc >= 0, "sb_pcm16_mix_at_many: bad arguments");
    if (!nsrc || !dst_count) return SB_OK;
    SB_REQUIRE(src_pool_dev && src_off_dev && src_len_dev && dst_off_dev, "sb_pcm16_mix_at_many: null pointer");
    int64_t ntiles = (dst_count + kMixTile - 1) / kMixTile;
    pcm16_mix_at_many_kernel<<<grid_for(ntiles, 1, kPMaxBlocks), kPBlock, 0, (cudaStream_t)stream>>>(
        dst_dev, dst_count, src_pool_dev, src_off_dev, src_len_dev, dst_off_dev, nsrc);
    SB_LAUNCH_CHECK();
    return SB_OK;
}
