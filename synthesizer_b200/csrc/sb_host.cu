// Host-buffer entry points: the same kernels for callers that own no device memory.
// Copies in and out are part of the call (bench.py's e2e figure is measured through these).
//
// State: one set of grow-only device workspaces PER DEVICE (keyed by the ordinal that is
// current when the call is made), a host-side cache of the time base per sample rate, and one
// pinned word per device for the overflow counter.  A mutex serialises the calls: two host
// threads may call concurrently, on the same or on different devices, and get the same
// bytes they would get alone; their GPU work does not overlap.  Workspaces only ever grow
// (cudaFree + cudaMalloc, a device-wide synchronisation, when a larger request arrives):
// sb_host_reserve() takes them to a high-water mark up front so that steady-state calls
// allocate nothing.
#include <map>
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

void bank_count_classes(const sb_voice *voices, int nvoices, double samplerate, int *generic, int *smooth, int *edged,
                        int *scan);

struct DeviceState {
    Workspace ws[3];                      // 0: bank render, 1: mix_at_many, 2: voices
    unsigned long long *pinned_ovf = nullptr;   // so that the overflow counter's device->host copy is asynchronous too
    int sms = 0;
};

static std::mutex g_mu;
static std::map<int, DeviceState> g_dev;        // by device ordinal; entries live until the process exits

static int device_state(DeviceState **out)
{
    int dev = 0;
    SB_CUDA_CHECK(cudaGetDevice(&dev));
    DeviceState &d = g_dev[dev];
    if (!d.sms) SB_CUDA_CHECK(cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, dev));
    if (!d.pinned_ovf) SB_CUDA_CHECK(cudaHostAlloc((void **)&d.pinned_ovf, sizeof(unsigned long long), cudaHostAllocDefault));
    *out = &d;
    return SB_OK;
}

struct TimebaseCache {
    double samplerate = 0;
    int64_t max_n = 0;
    std::vector<sb_time_seg> rad, cyc;
};
static TimebaseCache g_tb;                      // host memory: the same for every device

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

// Layout of workspace 0 for one bank render (everything but the voices)
struct BankLayout {
    size_t b_rad, b_cyc, b_carry, b_partial, b_mix, b_pcm, b_ovf;
    size_t total() const { return b_rad + b_cyc + b_carry + b_partial + b_mix + b_pcm + b_ovf; }
};

static BankLayout bank_layout(const sb_bank_plan &plan, int64_t frames, bool own_mix)
{
    BankLayout l;
    l.b_rad = align256(g_tb.rad.size() * sizeof(sb_time_seg));
    l.b_cyc = align256(g_tb.cyc.size() * sizeof(sb_time_seg));
    l.b_carry = align256((size_t)plan.carry_bytes);
    l.b_partial = align256((size_t)plan.partial_bytes);
    l.b_mix = own_mix ? align256((size_t)frames * 2 * sizeof(float)) : 0;
    l.b_pcm = own_mix ? align256((size_t)frames * 2 * sizeof(int16_t)) : 0;
    l.b_ovf = 256;
    return l;
}

// voices in host memory -> the float mix in device memory (mix_dev given, or inside the workspace);
// *mix_out / *pcm_out / *ovf_out point into the workspace when it owns them.  Caller holds g_mu.
static int bank_render_locked(DeviceState &d, const sb_voice *voices, int nvoices, int64_t frames, double samplerate,
                              float *mix_dev, cudaStream_t s, float **mix_out, int16_t **pcm_out,
                              unsigned long long **ovf_out)
{
    int rc = host_timebases(samplerate, frames);
    if (rc) return rc;
    sb_bank_plan plan;
    rc = sb_bank_plan_make(nvoices, frames, 0, d.sms, (int)g_tb.rad.size(), &plan);
    if (rc) return rc;
    // the voices go to their own workspace first, so that their copy can run while the host
    // counts which kernels the bank needs (the count decides how much scratch the render takes)
    rc = d.ws[2].ensure(align256((size_t)nvoices * sizeof(sb_voice)));
    if (rc) return rc;
    sb_voice *d_voices = (sb_voice *)d.ws[2].ptr;
    SB_CUDA_CHECK(cudaMemcpyAsync(d_voices, voices, (size_t)nvoices * sizeof(sb_voice), cudaMemcpyHostToDevice, s));
    int n_generic = 0, n_smooth = 0, n_edged = 0, n_scan = 0;
    bank_count_classes(voices, nvoices, samplerate, &n_generic, &n_smooth, &n_edged, &n_scan);
    rc = sb_bank_plan_hint(&plan, n_generic, n_smooth, n_edged, n_scan);
    if (rc) return rc;

    const BankLayout l = bank_layout(plan, frames, mix_dev == nullptr);
    rc = d.ws[0].ensure(l.total());
    if (rc) return rc;
    unsigned char *p = (unsigned char *)d.ws[0].ptr;
    sb_time_seg *d_rad = (sb_time_seg *)p; p += l.b_rad;
    sb_time_seg *d_cyc = (sb_time_seg *)p; p += l.b_cyc;
    void *d_carry = (void *)p; p += l.b_carry;
    float *d_partial = (float *)p; p += l.b_partial;
    float *d_mix = mix_dev ? mix_dev : (float *)p; p += l.b_mix;
    if (pcm_out) *pcm_out = (int16_t *)p;
    p += l.b_pcm;
    if (ovf_out) *ovf_out = (unsigned long long *)p;
    if (mix_out) *mix_out = d_mix;

    SB_CUDA_CHECK(cudaMemcpyAsync(d_rad, g_tb.rad.data(), g_tb.rad.size() * sizeof(sb_time_seg), cudaMemcpyHostToDevice, s));
    SB_CUDA_CHECK(cudaMemcpyAsync(d_cyc, g_tb.cyc.data(), g_tb.cyc.size() * sizeof(sb_time_seg), cudaMemcpyHostToDevice, s));
    return sb_bank_render_f32(d_voices, &plan, d_rad, (int)g_tb.rad.size(), d_cyc, (int)g_tb.cyc.size(), 0,
                              samplerate, d_carry, d_partial, d_mix, s);
}

}  // namespace sb

using namespace sb;

extern "C" int sb_host_bank_render_pcm16_on_stream(const sb_voice *voices, int nvoices, int64_t frames,
                                                   double samplerate, float scale, int16_t *out_stereo,
                                                   unsigned long long *overflow, void *stream)
{
    SB_REQUIRE(voices && out_stereo && nvoices >= 1 && frames >= 1 && samplerate > 1,
               "sb_host_bank_render_pcm16: bad arguments");
    std::lock_guard<std::mutex> lock(g_mu);
    DeviceState *d = nullptr;
    int rc = device_state(&d);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    float *d_mix = nullptr;
    int16_t *d_pcm = nullptr;
    unsigned long long *d_ovf = nullptr;
    rc = bank_render_locked(*d, voices, nvoices, frames, samplerate, nullptr, s, &d_mix, &d_pcm, &d_ovf);
    if (rc) return rc;
    SB_CUDA_CHECK(cudaMemsetAsync(d_ovf, 0, sizeof(unsigned long long), s));
    rc = sb_f32_to_pcm16(d_mix, frames * 2, scale, d_pcm, d_ovf, s);
    if (rc) return rc;
    SB_CUDA_CHECK(cudaMemcpyAsync(out_stereo, d_pcm, (size_t)frames * 2 * sizeof(int16_t), cudaMemcpyDeviceToHost, s));
    SB_CUDA_CHECK(cudaMemcpyAsync(d->pinned_ovf, d_ovf, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    SB_CUDA_CHECK(cudaStreamSynchronize(s));
    if (overflow) *overflow = *d->pinned_ovf;
    return SB_OK;
}

extern "C" int sb_host_bank_render_pcm16(const sb_voice *voices, int nvoices, int64_t frames,
                                         double samplerate, float scale, int16_t *out_stereo,
                                         unsigned long long *overflow)
{
    return sb_host_bank_render_pcm16_on_stream(voices, nvoices, frames, samplerate, scale, out_stereo, overflow, nullptr);
}

extern "C" int sb_host_bank_render_mix_f32(const sb_voice *voices, int nvoices, int64_t frames, double samplerate,
                                           float *mix_dev, void *stream)
{
    SB_REQUIRE(voices && mix_dev && nvoices >= 1 && frames >= 1 && samplerate > 1,
               "sb_host_bank_render_mix_f32: bad arguments");
    std::lock_guard<std::mutex> lock(g_mu);
    DeviceState *d = nullptr;
    int rc = device_state(&d);
    if (rc) return rc;
    // (asynchronous: the caller's stream orders what follows, e.g. the reduce of the partial mixes;
    // the next host call on this device reuses the workspace and is stream-ordered behind this one
    // only if it uses the same stream)
    return bank_render_locked(*d, voices, nvoices, frames, samplerate, mix_dev, (cudaStream_t)stream, nullptr, nullptr, nullptr);
}

extern "C" int sb_host_reserve(int nvoices, int64_t frames, double samplerate)
{
    SB_REQUIRE(nvoices >= 1 && frames >= 1 && samplerate > 1, "sb_host_reserve: bad arguments");
    std::lock_guard<std::mutex> lock(g_mu);
    DeviceState *d = nullptr;
    int rc = device_state(&d);
    if (rc) return rc;
    rc = host_timebases(samplerate, frames);
    if (rc) return rc;
    sb_bank_plan plan;
    rc = sb_bank_plan_make(nvoices, frames, 0, d->sms, (int)g_tb.rad.size(), &plan);      // unhinted = the worst case
    if (rc) return rc;
    rc = d->ws[2].ensure(align256((size_t)nvoices * sizeof(sb_voice)));
    if (rc) return rc;
    return d->ws[0].ensure(bank_layout(plan, frames, true).total());
}

extern "C" int sb_host_pcm16_mix_at_many_on_stream(int16_t *dst, int64_t dst_count, const int16_t *src_pool,
                                                   int64_t pool_count, const int64_t *src_off, const int64_t *src_len,
                                                   const int64_t *dst_off, int64_t nsrc, void *stream)
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
    DeviceState *d = nullptr;
    int rc = device_state(&d);
    if (rc) return rc;
    const size_t b_dst = align256((size_t)dst_count * 2), b_pool = align256((size_t)pool_count * 2);
    const size_t b_idx = align256((size_t)nsrc * 8);
    int64_t total_len = 0;
    for (int64_t s = 0; s < nsrc; s++) total_len += src_len[s];
    const int64_t n_scratch = sb_pcm16_mix_at_many_scratch_bytes(dst_count, nsrc, total_len);
    rc = d->ws[1].ensure(b_dst + b_pool + 3 * b_idx + align256((size_t)n_scratch));
    if (rc) return rc;
    unsigned char *p = (unsigned char *)d->ws[1].ptr;
    int16_t *d_dst = (int16_t *)p; p += b_dst;
    int16_t *d_pool = (int16_t *)p; p += b_pool;
    int64_t *d_so = (int64_t *)p; p += b_idx;
    int64_t *d_sl = (int64_t *)p; p += b_idx;
    int64_t *d_do = (int64_t *)p; p += b_idx;
    void *d_scratch = (void *)p;
    cudaStream_t s = (cudaStream_t)stream;
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

extern "C" int sb_host_pcm16_mix_at_many(int16_t *dst, int64_t dst_count, const int16_t *src_pool,
                                         int64_t pool_count, const int64_t *src_off, const int64_t *src_len,
                                         const int64_t *dst_off, int64_t nsrc)
{
    return sb_host_pcm16_mix_at_many_on_stream(dst, dst_count, src_pool, pool_count, src_off, src_len, dst_off, nsrc, nullptr);
}
