// Host-buffer entry points: the same kernels for callers that own no device memory.
// Copies in and out are part of the call (bench.py's e2e figure is measured through these).
// Device workspaces are grow-only and cached per process, so steady-state calls do not
// allocate.
#include <mutex>
#include <vector>
#include <string.h>
#include "sb_common.cuh"

namespace sb {

struct Workspace {
    void *ptr = nullptr;
    size_t cap = 0;
    int ensure(size_t bytes)
    {
        if (bytes <= cap) return SB_OK;
        if (ptr) cudaFree(ptr);
        ptr = nullptr;
        cap = 0;
        size_t want = bytes + bytes / 4 + 256;
        cudaError_t e = cudaMalloc(&ptr, want);
        if (e != cudaSuccess) {
            set_error("cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e));
            return SB_ENOMEM;
        }
        cap = want;
        return SB_OK;
    }
};

void bank_count_classes(const sb_voice *voices, int nvoices, double samplerate, int *generic, int *edged, int *scan);

static std::mutex g_mu;
static Workspace g_ws[8];

struct TimebaseCache {
    double samplerate = 0;
    int64_t max_n = 0;
    std::vector<sb_time_seg> rad, cyc;
};
static TimebaseCache g_tb;

static int host_timebases(double samplerate, int64_t max_n)
{
    if (g_tb.samplerate == samplerate && g_tb.max_n >= max_n) return SB_OK;
    g_tb.rad.resize(SB_TIMEBASE_MAX_SEGS);
    g_tb.cyc.resize(SB_TIMEBASE_MAX_SEGS);
    int64_t cover = max_n * 2 + 1024;
    int nr = sb_timebase_build(2.0 * 3.141592653589793 / samplerate, cover, g_tb.rad.data(), SB_TIMEBASE_MAX_SEGS);
    if (nr < 0) return nr;
    int nc = sb_timebase_build(1.0 / samplerate, cover, g_tb.cyc.data(), SB_TIMEBASE_MAX_SEGS);
    if (nc < 0) return nc;
    g_tb.rad.resize(nr);
    g_tb.cyc.resize(nc);
    g_tb.samplerate = samplerate;
    g_tb.max_n = cover;
    return SB_OK;
}

static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

// one pinned word for the overflow counter, so that its device->host copy is asynchronous too
static unsigned long long *g_pinned_ovf = nullptr;

}  // namespace sb

using namespace sb;

extern "C" int sb_host_bank_render_pcm16(const sb_voice *voices, int nvoices, int64_t frames,
                                         double samplerate, float scale, int16_t *out_stereo,
                                         unsigned long long *overflow)
{
    SB_REQUIRE(voices && out_stereo && nvoices >= 1 && frames >= 1 && samplerate > 1,
               "sb_host_bank_render_pcm16: bad arguments");
    std::lock_guard<std::mutex> lock(g_mu);
    int rc = host_timebases(samplerate, frames);
    if (rc) return rc;
    int dev = 0, sms = 148;
    SB_CUDA_CHECK(cudaGetDevice(&dev));
    SB_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    sb_bank_plan plan;
    rc = sb_bank_plan_make(nvoices, frames, 0, sms, (int)g_tb.rad.size(), &plan);
    if (rc) return rc;
    if (!g_pinned_ovf) SB_CUDA_CHECK(cudaHostAlloc((void **)&g_pinned_ovf, sizeof(unsigned long long), cudaHostAllocDefault));

    // the voices go to their own workspace first, so that their copy can run while the host
    // counts which kernels the bank needs (the count decides how much scratch the render takes)
    const size_t b_voices = align256((size_t)nvoices * sizeof(sb_voice));
    rc = g_ws[2].ensure(b_voices);
    if (rc) return rc;
    sb_voice *d_voices = (sb_voice *)g_ws[2].ptr;
    cudaStream_t s = 0;
    SB_CUDA_CHECK(cudaMemcpyAsync(d_voices, voices, (size_t)nvoices * sizeof(sb_voice), cudaMemcpyHostToDevice, s));
    int n_generic = 0, n_edged = 0, n_scan = 0;
    bank_count_classes(voices, nvoices, samplerate, &n_generic, &n_edged, &n_scan);
    rc = sb_bank_plan_hint(&plan, n_generic, n_edged, n_scan);
    if (rc) return rc;

    const size_t b_rad = align256(g_tb.rad.size() * sizeof(sb_time_seg));
    const size_t b_cyc = align256(g_tb.cyc.size() * sizeof(sb_time_seg));
    const size_t b_carry = align256((size_t)plan.carry_bytes);
    const size_t b_partial = align256((size_t)plan.partial_bytes);
    const size_t b_mix = align256((size_t)frames * 2 * sizeof(float));
    const size_t b_pcm = align256((size_t)frames * 2 * sizeof(int16_t));
    const size_t b_ovf = 256;
    rc = g_ws[0].ensure(b_rad + b_cyc + b_carry + b_partial + b_mix + b_pcm + b_ovf);
    if (rc) return rc;
    unsigned char *p = (unsigned char *)g_ws[0].ptr;
    sb_time_seg *d_rad = (sb_time_seg *)p; p += b_rad;
    sb_time_seg *d_cyc = (sb_time_seg *)p; p += b_cyc;
    void *d_carry = (void *)p; p += b_carry;
    float *d_partial = (float *)p; p += b_partial;
    float *d_mix = (float *)p; p += b_mix;
    int16_t *d_pcm = (int16_t *)p; p += b_pcm;
    unsigned long long *d_ovf = (unsigned long long *)p;

    SB_CUDA_CHECK(cudaMemcpyAsync(d_rad, g_tb.rad.data(), g_tb.rad.size() * sizeof(sb_time_seg), cudaMemcpyHostToDevice, s));
    SB_CUDA_CHECK(cudaMemcpyAsync(d_cyc, g_tb.cyc.data(), g_tb.cyc.size() * sizeof(sb_time_seg), cudaMemcpyHostToDevice, s));
    SB_CUDA_CHECK(cudaMemsetAsync(d_ovf, 0, sizeof(unsigned long long), s));
    rc = sb_bank_render_f32(d_voices, &plan, d_rad, (int)g_tb.rad.size(), d_cyc, (int)g_tb.cyc.size(), 0,
                            samplerate, d_carry, d_partial, d_mix, s);
    if (rc) return rc;
    rc = sb_f32_to_pcm16(d_mix, frames * 2, scale, d_pcm, d_ovf, s);
    if (rc) return rc;
    SB_CUDA_CHECK(cudaMemcpyAsync(out_stereo, d_pcm, (size_t)frames * 2 * sizeof(int16_t), cudaMemcpyDeviceToHost, s));
    SB_CUDA_CHECK(cudaMemcpyAsync(g_pinned_ovf, d_ovf, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    SB_CUDA_CHECK(cudaStreamSynchronize(s));
    if (overflow) *overflow = *g_pinned_ovf;
    return SB_OK;
}

extern "C" int sb_host_pcm16_mix_at_many(int16_t *dst, int64_t dst_count, const int16_t *src_pool,
                                         int64_t pool_count, const int64_t *src_off, const int64_t *src_len,
                                         const int64_t *dst_off, int64_t nsrc)
{
    SB_REQUIRE(dst && dst_count >= 0 && nsrc >= 0 && pool_count >= 0, "sb_host_pcm16_mix_at_many: bad arguments");
    if (!nsrc || !dst_count) return SB_OK;
    SB_REQUIRE(src_pool && src_off && src_len && dst_off, "sb_host_pcm16_mix_at_many: null pointer");
    for (int64_t s = 0; s < nsrc; s++) {
        SB_REQUIRE(src_off[s] >= 0 && src_len[s] >= 0 && src_off[s] + src_len[s] <= pool_count,
                   "sb_host_pcm16_mix_at_many: source outside the pool");
        SB_REQUIRE(dst_off[s] >= 0 && dst_off[s] + src_len[s] <= dst_count,
                   "sb_host_pcm16_mix_at_many: source does not fit the destination");
    }
    std::lock_guard<std::mutex> lock(g_mu);
    const size_t b_dst = align256((size_t)dst_count * 2), b_pool = align256((size_t)pool_count * 2);
    const size_t b_idx = align256((size_t)nsrc * 8);
    int64_t total_len = 0;
    for (int64_t s = 0; s < nsrc; s++) total_len += src_len[s];
    const int64_t n_scratch = sb_pcm16_mix_at_many_scratch_bytes(dst_count, nsrc, total_len);
    int rc = g_ws[1].ensure(b_dst + b_pool + 3 * b_idx + align256((size_t)n_scratch));
    if (rc) return rc;
    unsigned char *p = (unsigned char *)g_ws[1].ptr;
    int16_t *d_dst = (int16_t *)p; p += b_dst;
    int16_t *d_pool = (int16_t *)p; p += b_pool;
    int64_t *d_so = (int64_t *)p; p += b_idx;
    int64_t *d_sl = (int64_t *)p; p += b_idx;
    int64_t *d_do = (int64_t *)p; p += b_idx;
    void *d_scratch = (void *)p;
    cudaStream_t s = 0;
    SB_CUDA_CHECK(cudaMemcpyAsync(d_dst, dst, (size_t)dst_count * 2, cudaMemcpyHostToDevice, s));
    SB_CUDA_CHECK(cudaMemcpyAsync(d_pool, src_pool, (size_t)pool_count * 2, cudaMemcpyHostToDevice, s));
    SB_CUDA_CHECK(cudaMemcpyAsync(d_so, src_off, (size_t)nsrc * 8, cudaMemcpyHostToDevice, s));
    SB_CUDA_CHECK(cudaMemcpyAsync(d_sl, src_len, (size_t)nsrc * 8, cudaMemcpyHostToDevice, s));
    SB_CUDA_CHECK(cudaMemcpyAsync(d_do, dst_off, (size_t)nsrc * 8, cudaMemcpyHostToDevice, s));
    rc = sb_pcm16_mix_at_many(d_dst, dst_count, d_pool, d_so, d_sl, d_do, nsrc, d_scratch, n_scratch, s);
    if (rc) return rc;
    SB_CUDA_CHECK(cudaMemcpyAsync(dst, d_dst, (size_t)dst_count * 2, cudaMemcpyDeviceToHost, s));
    SB_CUDA_CHECK(cudaStreamSynchronize(s));
    return SB_OK;
}
