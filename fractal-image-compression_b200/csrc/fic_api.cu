// C ABI of libfic_b200.so (include/fic.h): handle management, staging of host buffers, dispatch.
#include <stdarg.h>

#include "fic_common.cuh"

int bf_prepare8(fic_handle* h, const uint8_t* d_img, int H, int W, bool with_stats, cudaStream_t st);
int bf_screen_run(fic_handle* h, const uint8_t* d_img, int H, int W, int mode, int range_begin, int n_ranges, int32_t* d_idx,
                  double* d_so, double* d_R, unsigned flags, float* probe, cudaStream_t st);
// device-side counter totals of the brute-force pipeline: waits for the handle's last call, copies {survivors, ranges
// screened, ranges re-scanned, fault code} to t[4]; t == nullptr clears them instead
int bf_totals_read(fic_handle* h, unsigned long long* t);

static thread_local char g_err[512] = "";

void fic_set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int fic_copy_in(fic_handle* h, void* dst_dev, const void* src, size_t bytes, bool src_is_dev, cudaStream_t st)
{
    if (bytes == 0) return FIC_OK;
    FIC_CUDA(cudaMemcpyAsync(dst_dev, src, bytes, src_is_dev ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
    if (!src_is_dev) h->counters[CNT_H2D] += (int64_t)bytes;
    return FIC_OK;
}

int fic_sync(fic_handle* h, cudaStream_t st)
{
    FIC_CUDA(cudaStreamSynchronize(st));
    for (const fic_handle::PendingOut& po : h->pending_out) memcpy(po.dst, h->pin_stage.as<char>() + po.at, po.bytes);
    h->pending_out.clear();
    h->stage_used = 0;
    return FIC_OK;
}

int fic_copy_out(fic_handle* h, void* dst, const void* src_dev, size_t bytes, bool dst_is_dev, cudaStream_t st)
{
    if (bytes == 0) return FIC_OK;
    if (dst_is_dev) {
        FIC_CUDA(cudaMemcpyAsync(dst, src_dev, bytes, cudaMemcpyDeviceToDevice, st));
        return FIC_OK;
    }
    if (bytes > ((size_t)128 << 20)) {   // not worth holding that much pinned memory: let the driver stage it
        FIC_CUDA(cudaMemcpyAsync(dst, src_dev, bytes, cudaMemcpyDeviceToHost, st));
        h->counters[CNT_D2H] += (int64_t)bytes;
        return FIC_OK;
    }
    const size_t at = (h->stage_used + 255) & ~(size_t)255;
    if (at + bytes > h->pin_stage.cap) {
        FIC_TRY(fic_sync(h, st));   // the staging area moves: deliver what is in it first
        FIC_TRY(h->pin_stage.reserve(bytes + bytes / 2 + (1 << 20)));
        return fic_copy_out(h, dst, src_dev, bytes, false, st);
    }
    FIC_CUDA(cudaMemcpyAsync(h->pin_stage.as<char>() + at, src_dev, bytes, cudaMemcpyDeviceToHost, st));
    h->pending_out.push_back(fic_handle::PendingOut{dst, at, bytes});
    h->stage_used = at + bytes;
    h->counters[CNT_D2H] += (int64_t)bytes;
    return FIC_OK;
}

extern "C" {

const char* fic_version(void) { return "fic_b200 0.1 (sm_100a)"; }
const char* fic_last_error(void) { return g_err; }

int fic_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int fic_create(int device, fic_handle** out)
{
    if (!out) { fic_set_error("fic_create: out is NULL"); return FIC_EINVAL; }
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        fic_set_error("fic_create: no CUDA device (%s); this library has no CPU fallback",
                      e == cudaSuccess ? "count is 0" : cudaGetErrorString(e));
        return FIC_ENOGPU;
    }
    if (device < 0 || device >= n) { fic_set_error("fic_create: device %d out of range [0,%d)", device, n); return FIC_EINVAL; }
    cudaDeviceProp prop;
    FIC_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        fic_set_error("fic_create: device %d is sm_%d%d; the kernels are built for sm_100a only", device, prop.major, prop.minor);
        return FIC_ENOGPU;
    }
    DeviceScope scope(device);
    FIC_CUDA(scope.err);
    fic_handle* h = new fic_handle();
    h->device = device;
    h->num_sms = prop.multiProcessorCount;
    e = cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->tail, cudaEventDisableTiming);
    if (e != cudaSuccess) {
        fic_set_error("stream / event creation failed: %s", cudaGetErrorString(e));
        if (h->own_stream) cudaStreamDestroy(h->own_stream);
        delete h;
        return FIC_ECUDA;
    }
    *out = h;
    return FIC_OK;
}

int fic_destroy(fic_handle* h)
{
    if (!h) return FIC_OK;
    DeviceScope scope(h->device);
    if (h->tail_valid) cudaEventSynchronize(h->tail);   // work of the last call, whatever stream it went to
    cudaStreamSynchronize(h->own_stream);
    DevBuf* bufs[] = {&h->d_img, &h->d_out_a, &h->d_out_b, &h->d_out_c, &h->d_in_a, &h->d_in_b, &h->bf_boxsum, &h->bf_stats,
                      &h->bf_partial, &h->bf_ranges, &h->bf_pool, &h->bf_cand, &h->bf_A, &h->bf_thr,
                      &h->bf_surv, &h->bf_ctrl, &h->bf_best, &h->bf_list, &h->dec_img, &h->dec_next, &h->dec_map,
                      &h->dec_sched, &h->dec_red, &h->t_int1, &h->t_int2, &h->t_cls, &h->t_lists, &h->t_work, &h->t_res,
                      &h->t_aux, &h->t_hvpool, &h->t_part, &h->dec_keys, &h->dec_roots, &h->bf_done, &h->t_nodes, &h->t_count, &h->t_rcls, &h->t_need};
    for (DevBuf* b : bufs) b->release();
    h->pin_a.release();
    h->pin_b.release();
    h->pin_out.release();
    h->pin_stage.release();
    h->pin_tree.release();
    for (cudaEvent_t e : h->prof_events) cudaEventDestroy(e);
    cudaEventDestroy(h->tail);
    cudaStreamDestroy(h->own_stream);
    delete h;
    return FIC_OK;
}

int fic_get_counters(fic_handle* h, int64_t out[8])
{
    if (!h || !out) { fic_set_error("fic_get_counters: NULL argument"); return FIC_EINVAL; }
    for (int k = 0; k < 8; ++k) out[k] = h->counters[k];
    if (h->bf_totals_live) {
        // The brute-force pipeline counts on the device (its calls do not wait for the GPU): fetch the totals now.
        // This is one of the library's host-visible points, so a pipeline fault recorded by an earlier call surfaces here.
        DeviceScope scope(h->device);
        FIC_CUDA(scope.err);
        unsigned long long t[4] = {0, 0, 0, 0};
        FIC_TRY(bf_totals_read(h, t));
        out[CNT_SURVIVORS] += (int64_t)t[0];
        out[CNT_SCREEN_RANGES] += (int64_t)t[1];
        out[CNT_RESCAN_RANGES] += (int64_t)t[2];
        if (t[3]) { fic_set_error("bf_screen_kernel: pipeline wait %llu timed out in an earlier call", t[3]); return FIC_ECUDA; }
    }
    return FIC_OK;
}

int fic_reset_counters(fic_handle* h)
{
    if (!h) { fic_set_error("fic_reset_counters: NULL handle"); return FIC_EINVAL; }
    for (int k = 0; k < 8; ++k) h->counters[k] = 0;
    if (h->bf_totals_live) {
        DeviceScope scope(h->device);
        FIC_CUDA(scope.err);
        FIC_TRY(bf_totals_read(h, nullptr));
    }
    return FIC_OK;
}

int fic_synchronize(fic_handle* h)
{
    if (!h) { fic_set_error("fic_synchronize: NULL handle"); return FIC_EINVAL; }
    DeviceScope scope(h->device);
    FIC_CUDA(scope.err);
    if (h->tail_valid) FIC_CUDA(cudaEventSynchronize(h->tail));   // the last call's work, whatever stream it was given
    FIC_TRY(fic_sync(h, h->own_stream));
    return FIC_OK;
}

int fic_profile(fic_handle* h, int enable)
{
    if (!h) { fic_set_error("fic_profile: NULL handle"); return FIC_EINVAL; }
    h->prof_on = enable != 0;
    return FIC_OK;
}

int fic_profile_read(fic_handle* h, double* total_ms, int64_t* launches)
{
    if (!h || !total_ms || !launches) { fic_set_error("fic_profile_read: NULL argument"); return FIC_EINVAL; }
    DeviceScope scope(h->device);
    FIC_CUDA(scope.err);
    double tot = 0.0;
    int64_t n = 0;
    for (size_t k = 0; k + 1 < h->prof_events.size(); k += 2) {
        float ms = 0.f;
        FIC_CUDA(cudaEventSynchronize(h->prof_events[k + 1]));
        FIC_CUDA(cudaEventElapsedTime(&ms, h->prof_events[k], h->prof_events[k + 1]));
        tot += ms;
        ++n;
    }
    for (cudaEvent_t e : h->prof_events) cudaEventDestroy(e);
    h->prof_events.clear();
    *total_ms = tot;
    *launches = n;
    return FIC_OK;
}

// ---------------------------------------------------------------------------------------------
int fic_bf_encode(fic_handle* h, const uint8_t* img, int H, int W, int range, int domain, int stride_i, int stride_j,
                  int mode, int range_begin, int range_end, int32_t* out_idx, double* out_so, double* out_R,
                  unsigned flags, void* stream)
{
    FIC_REQUIRE(h && img && out_idx && out_so, "fic_bf_encode: NULL argument");
    FIC_REQUIRE(range > 0 && domain == 2 * range,
                "fic_bf_encode: domain_size must be 2*range_size (the reference cannot multiply a %dx%d pooled block "
                "with a %dx%d range)", domain / 2, domain / 2, range, range);
    FIC_REQUIRE(H > 0 && W > 0 && H % range == 0 && W % range == 0, "fic_bf_encode: image %dx%d is not a multiple of range %d", H, W, range);
    FIC_REQUIRE(stride_i > 0 && stride_j > 0, "fic_bf_encode: strides must be positive");
    FIC_REQUIRE(mode == FIC_BF_MODE_NUMPY || mode == FIC_BF_MODE_JIT, "fic_bf_encode: unknown mode %d", mode);
    const int total = (H / range) * (W / range);
    FIC_REQUIRE(0 <= range_begin && range_begin <= range_end && range_end <= total,
                "fic_bf_encode: range slice [%d,%d) outside [0,%d)", range_begin, range_end, total);
    FIC_ENTER(h, stream);
    const bool dev = flags & FIC_DEVICE_PTRS;
    const int n = range_end - range_begin;
    if (n == 0) return FIC_OK;
    const uint8_t* d_img = img;
    int32_t* d_idx = out_idx;
    double* d_so = out_so;
    double* d_R = out_R;
    if (!dev) {
        FIC_TRY(h->d_img.reserve((size_t)H * W));
        FIC_TRY(h->d_out_a.reserve((size_t)n * 4 * sizeof(int32_t)));
        FIC_TRY(h->d_out_b.reserve((size_t)n * 2 * sizeof(double)));
        FIC_TRY(h->d_out_c.reserve((size_t)n * sizeof(double)));
        FIC_TRY(fic_copy_in(h, h->d_img.p, img, (size_t)H * W, false, st));
        d_img = h->d_img.as<uint8_t>();
        d_idx = h->d_out_a.as<int32_t>();
        d_so = h->d_out_b.as<double>();
        d_R = h->d_out_c.as<double>();
    } else if (!d_R) {
        FIC_TRY(h->d_out_c.reserve((size_t)n * sizeof(double)));
        d_R = h->d_out_c.as<double>();
    }
    const size_t b_idx = (size_t)n * 4 * sizeof(int32_t), b_so = (size_t)n * 2 * sizeof(double),
                 b_R = out_R ? (size_t)n * sizeof(double) : 0;
    if (!dev) {
        FIC_TRY(h->pin_out.reserve(b_idx + b_so + b_R));
        h->eager.armed = true;
        h->eager.done = false;
        h->eager.src[0] = d_idx; h->eager.bytes[0] = b_idx; h->eager.at[0] = 0;
        h->eager.src[1] = d_so;  h->eager.bytes[1] = b_so;  h->eager.at[1] = b_idx;
        h->eager.src[2] = d_R;   h->eager.bytes[2] = b_R;   h->eager.at[2] = b_idx + b_so;
    }
    const int rc = bf_encode_impl(h, d_img, H, W, range, stride_i, stride_j, mode, range_begin, range_end, d_idx, d_so, d_R,
                                  flags, st);
    h->eager.armed = false;
    FIC_TRY(rc);
    if (!dev) {
        if (!h->eager.done) {   // paths without the in-pipeline copy (exact scan, fallback ranges)
            FIC_CUDA(cudaMemcpyAsync(h->pin_out.as<char>(), d_idx, b_idx, cudaMemcpyDeviceToHost, st));
            FIC_CUDA(cudaMemcpyAsync(h->pin_out.as<char>() + b_idx, d_so, b_so, cudaMemcpyDeviceToHost, st));
            if (b_R) FIC_CUDA(cudaMemcpyAsync(h->pin_out.as<char>() + b_idx + b_so, d_R, b_R, cudaMemcpyDeviceToHost, st));
            FIC_TRY(fic_sync(h, st));
        }
        h->counters[CNT_D2H] += (int64_t)(b_idx + b_so + b_R);
        memcpy(out_idx, h->pin_out.as<char>(), b_idx);
        memcpy(out_so, h->pin_out.as<char>() + b_idx, b_so);
        if (b_R) memcpy(out_R, h->pin_out.as<char>() + b_idx + b_so, b_R);
    }
    return FIC_OK;
}

int fic_bf_screen_probe(fic_handle* h, const uint8_t* img, int H, int W, int range_begin, int n_ranges, float* out_u,
                        unsigned flags, void* stream)
{
    FIC_REQUIRE(h && img && out_u, "fic_bf_screen_probe: NULL argument");
    FIC_REQUIRE(H > 16 && W > 16 && H % 8 == 0 && W % 8 == 0, "fic_bf_screen_probe: image %dx%d unsupported", H, W);
    const int total = (H / 8) * (W / 8);
    FIC_REQUIRE(n_ranges >= 1 && n_ranges <= 32 && range_begin >= 0 && range_begin + n_ranges <= total,
                "fic_bf_screen_probe: at most 32 ranges inside [0,%d)", total);
    FIC_ENTER(h, stream);
    const bool dev = flags & FIC_DEVICE_PTRS;
    const size_t NP = (size_t)(H - 16) * (W - 16);
    const size_t rows = (NP + 127) / 128 * 128;
    const uint8_t* d_img = img;
    float* d_out = out_u;
    if (!dev) {
        FIC_TRY(h->d_img.reserve((size_t)H * W));
        FIC_TRY(fic_copy_in(h, h->d_img.p, img, (size_t)H * W, false, st));
        d_img = h->d_img.as<uint8_t>();
        FIC_TRY(h->d_out_a.reserve(rows * 256 * sizeof(float)));
        d_out = h->d_out_a.as<float>();
    }
    FIC_CUDA(cudaMemsetAsync(d_out, 0, rows * 256 * sizeof(float), st));
    FIC_TRY(bf_prepare8(h, d_img, H, W, false, st));
    FIC_TRY(bf_screen_run(h, d_img, H, W, FIC_BF_MODE_JIT, range_begin, n_ranges, nullptr, nullptr, nullptr, flags, d_out, st));
    if (!dev) {
        FIC_TRY(fic_copy_out(h, out_u, d_out, rows * 256 * sizeof(float), false, st));
        FIC_TRY(fic_sync(h, st));
    }
    return FIC_OK;
}

int fic_bf_encode_batch(fic_handle* h, const uint8_t* imgs, int B, int H, int W, int range, int domain, int mode,
                        int32_t* out_idx, double* out_so, double* out_R, unsigned flags, void* stream)
{
    FIC_REQUIRE(h && imgs && out_idx && out_so, "fic_bf_encode_batch: NULL argument");
    FIC_REQUIRE(B >= 0, "fic_bf_encode_batch: negative batch");
    FIC_REQUIRE(range > 0 && H > 0 && W > 0 && H % range == 0 && W % range == 0, "fic_bf_encode_batch: bad geometry");
    const size_t n = (size_t)(H / range) * (W / range);
    for (int b = 0; b < B; ++b) {
        int rc = fic_bf_encode(h, imgs + (size_t)b * H * W, H, W, range, domain, 1, 1, mode, 0, (int)n, out_idx + b * n * 4,
                               out_so + b * n * 2, out_R ? out_R + b * n : nullptr, flags, stream);
        if (rc < 0) return rc;
    }
    return FIC_OK;
}

static int bf_decode_common(const char* who, fic_handle* h, int planes, const int32_t* idx, const double* so, int n, int width,
                            int height, int iterations, uint8_t* out, unsigned flags, void* stream)
{
    FIC_REQUIRE(h && idx && so && out, "%s: NULL argument", who);
    FIC_REQUIRE(planes >= 1 && n >= 0 && width > 0 && height > 0 && iterations >= 0, "%s: bad sizes", who);
    FIC_REQUIRE(width % 8 == 0 && height % 8 == 0, "%s: sides must be multiples of the hard-coded range size 8", who);
    FIC_ENTER(h, stream);
    const bool dev = flags & FIC_DEVICE_PTRS;
    // the execution schedule is built on the host from the transforms, so host copies are needed either way
    const size_t n_all = (size_t)planes * n;
    const int32_t* h_idx = idx;
    const double* h_so = so;
    uint8_t* d_out = out;
    if (dev) {
        FIC_TRY(h->pin_a.reserve(n_all * (4 * sizeof(int32_t) + 2 * sizeof(double)) + 64));
        double* ps = h->pin_a.as<double>();
        int32_t* pi = reinterpret_cast<int32_t*>(ps + n_all * 2);
        FIC_CUDA(cudaMemcpyAsync(pi, idx, n_all * 4 * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        FIC_CUDA(cudaMemcpyAsync(ps, so, n_all * 2 * sizeof(double), cudaMemcpyDeviceToHost, st));
        FIC_TRY(fic_sync(h, st));
        h->counters[CNT_D2H] += (int64_t)(n_all * (4 * sizeof(int32_t) + 2 * sizeof(double)));
        h_idx = pi;
        h_so = ps;
    } else {
        FIC_TRY(h->d_out_a.reserve((size_t)planes * width * height));
        d_out = h->d_out_a.as<uint8_t>();
    }
    for (size_t q = 0; q < n_all; ++q) {
        const int k = (int)(q % (size_t)(n ? n : 1));
        const int32_t* r = h_idx + 4 * q;
        FIC_REQUIRE(r[0] >= 0 && r[1] >= 0 && r[0] + 16 <= height && r[1] + 16 <= width && (unsigned)r[2] < 2u && (unsigned)r[3] < 4u,
                    "%s: transform %d (%d,%d,%d,%d) is outside the image", who, k, r[0], r[1], r[2], r[3]);
        // decoder.py:19-20 mixes width and height; numpy raises when the 8x8 target slice leaves the image
        const int range_i = (k / (width / 8)) * 8, range_j = (k % (height / 8)) * 8;
        FIC_REQUIRE(range_i + 8 <= height && range_j + 8 <= width,
                    "%s: transform %d maps to range block (%d,%d) outside the %dx%d image "
                    "(the reference's index arithmetic, bruteforce/decoder.py:19-20)", who, k, range_i, range_j, height, width);
    }
    FIC_TRY(bf_decode_impl(h, planes, h_idx, h_so, n, width, height, iterations, d_out, st));
    if (!dev) {
        FIC_TRY(fic_copy_out(h, out, d_out, (size_t)planes * width * height, false, st));
        FIC_TRY(fic_sync(h, st));
    }
    return FIC_OK;
}

int fic_bf_decode(fic_handle* h, const int32_t* idx, const double* so, int n, int width, int height, int iterations,
                  uint8_t* out, unsigned flags, void* stream)
{
    return bf_decode_common("fic_bf_decode", h, 1, idx, so, n, width, height, iterations, out, flags, stream);
}

int fic_bf_decode_batch(fic_handle* h, int planes, const int32_t* idx, const double* so, int n, int width, int height,
                        int iterations, uint8_t* out, unsigned flags, void* stream)
{
    return bf_decode_common("fic_bf_decode_batch", h, planes, idx, so, n, width, height, iterations, out, flags, stream);
}

int fic_average_subsample(fic_handle* h, const double* arr, int h2, int w2, int jit, double* out, unsigned flags, void* stream)
{
    FIC_REQUIRE(h && arr && out, "fic_average_subsample: NULL argument");
    FIC_REQUIRE(h2 > 0 && w2 > 0 && h2 % 2 == 0 && w2 % 2 == 0, "fic_average_subsample: shape (%d,%d) must be even", h2, w2);
    FIC_ENTER(h, stream);
    const bool dev = flags & FIC_DEVICE_PTRS;
    const size_t nin = (size_t)h2 * w2 * sizeof(double), nout = nin / 4;
    const double* d_arr = arr;
    double* d_out = out;
    if (!dev) {
        FIC_TRY(h->d_in_a.reserve(nin));
        FIC_TRY(h->d_out_a.reserve(nout));
        FIC_TRY(fic_copy_in(h, h->d_in_a.p, arr, nin, false, st));
        d_arr = h->d_in_a.as<double>();
        d_out = h->d_out_a.as<double>();
    }
    FIC_TRY(subsample_impl(h, d_arr, h2, w2, jit, d_out, st));
    if (!dev) {
        FIC_TRY(fic_copy_out(h, out, d_out, nout, false, st));
        FIC_TRY(fic_sync(h, st));
    }
    return FIC_OK;
}

static int stage_image(fic_handle* h, const uint8_t* img, int H, int W, bool dev, cudaStream_t st, const uint8_t** d_img)
{
    if (dev) { *d_img = img; return FIC_OK; }
    FIC_TRY(h->d_img.reserve((size_t)H * W));
    FIC_TRY(fic_copy_in(h, h->d_img.p, img, (size_t)H * W, false, st));
    *d_img = h->d_img.as<uint8_t>();
    return FIC_OK;
}

int fic_qt_classify(fic_handle* h, const uint8_t* img, int H, int W, const int32_t* blocks, int count, int32_t* out,
                    unsigned flags, void* stream)
{
    FIC_REQUIRE(h && img && blocks && out, "fic_qt_classify: NULL argument");
    FIC_REQUIRE(H > 0 && W > 0 && count >= 0, "fic_qt_classify: bad sizes");
    FIC_ENTER(h, stream);
    const bool dev = flags & FIC_DEVICE_PTRS;
    const uint8_t* d_img;
    FIC_TRY(stage_image(h, img, H, W, dev, st, &d_img));
    const int32_t* d_blocks = blocks;
    int32_t* d_out = out;
    const size_t nb = (size_t)count * 4 * sizeof(int32_t);
    if (!dev) {
        for (int k = 0; k < count; ++k) {
            const int32_t* b = blocks + 4 * k;
            FIC_REQUIRE(b[2] >= 2 && (b[2] & (b[2] - 1)) == 0 && b[0] >= 0 && b[1] >= 0 && b[0] + b[2] <= H && b[1] + b[2] <= W,
                        "fic_qt_classify: block %d (%d,%d,side %d) invalid", k, b[0], b[1], b[2]);
        }
        FIC_TRY(h->d_in_a.reserve(nb + 16));
        FIC_TRY(h->d_out_a.reserve(nb + 16));
        FIC_TRY(fic_copy_in(h, h->d_in_a.p, blocks, nb, false, st));
        d_blocks = h->d_in_a.as<int32_t>();
        d_out = h->d_out_a.as<int32_t>();
    }
    FIC_TRY(qt_classify_impl(h, d_img, H, W, d_blocks, count, d_out, st));
    if (!dev) {
        FIC_TRY(fic_copy_out(h, out, d_out, nb, false, st));
        FIC_TRY(fic_sync(h, st));
    }
    return FIC_OK;
}

int fic_qt_encode(fic_handle* h, const uint8_t* img, int H, int W, int scale_bits, int offset_bits, double max_scale,
                  int full_class_search, int only_positive_scale, int min_range_side, int max_range_side,
                  double error_tolerance, int domain_hstep, int domain_vstep, int32_t* nodes, double* vals, int cap,
                  int32_t* n_roots, unsigned flags, void* stream)
{
    FIC_REQUIRE(h && img && nodes && vals, "fic_qt_encode: NULL argument");
    FIC_REQUIRE(H > 0 && W > 0, "fic_qt_encode: empty image");
    FIC_REQUIRE(min_range_side >= 1 && max_range_side >= 1, "fic_qt_encode: range sides must be >= 1");
    FIC_REQUIRE(scale_bits >= 0 && scale_bits <= 30 && offset_bits >= 0 && offset_bits <= 30, "fic_qt_encode: bit widths out of range");
    FIC_REQUIRE((domain_hstep == -1 || domain_hstep > 0) && (domain_vstep == -1 || domain_vstep > 0), "fic_qt_encode: domain steps must be -1 or positive");
    FIC_ENTER(h, stream);
    const uint8_t* d_img;
    FIC_TRY(stage_image(h, img, H, W, flags & FIC_DEVICE_PTRS, st, &d_img));
    return qt_encode_impl(h, d_img, H, W, scale_bits, offset_bits, max_scale, full_class_search, only_positive_scale,
                          min_range_side, max_range_side, error_tolerance, domain_hstep, domain_vstep, nodes, vals, cap,
                          n_roots, st);
}

int fic_hv_search(fic_handle* h, const uint8_t* img, int H, int W, const int32_t* rects, int count, int scale_bits,
                  int offset_bits, double max_scale, int32_t* rec, double* vals, unsigned flags, void* stream)
{
    FIC_REQUIRE(h && img && rects && rec && vals, "fic_hv_search: NULL argument");
    FIC_REQUIRE(H > 0 && W > 0 && count >= 0, "fic_hv_search: bad sizes");
    FIC_ENTER(h, stream);
    const bool dev = flags & FIC_DEVICE_PTRS;
    const uint8_t* d_img;
    FIC_TRY(stage_image(h, img, H, W, dev, st, &d_img));
    const int32_t* d_rects = rects;
    int32_t* d_rec = rec;
    double* d_vals = vals;
    if (!dev) {
        for (int k = 0; k < count; ++k) {
            const int32_t* r = rects + 4 * k;
            FIC_REQUIRE(r[0] >= 0 && r[1] >= 0 && r[2] >= 1 && r[3] >= 1 && r[0] + r[3] <= H && r[1] + r[2] <= W,
                        "fic_hv_search: rect %d (i=%d,j=%d,w=%d,h=%d) invalid", k, r[0], r[1], r[2], r[3]);
        }
        FIC_TRY(h->d_in_a.reserve((size_t)count * 16 + 16));
        FIC_TRY(h->d_out_a.reserve((size_t)count * 24 + 16));
        FIC_TRY(h->d_out_b.reserve((size_t)count * 24 + 16));
        FIC_TRY(fic_copy_in(h, h->d_in_a.p, rects, (size_t)count * 16, false, st));
        d_rects = h->d_in_a.as<int32_t>();
        d_rec = h->d_out_a.as<int32_t>();
        d_vals = h->d_out_b.as<double>();
    }
    FIC_TRY(hv_search_impl(h, d_img, H, W, d_rects, count, scale_bits, offset_bits, max_scale, d_rec, d_vals, st));
    if (!dev) {
        FIC_TRY(fic_copy_out(h, rec, d_rec, (size_t)count * 24, false, st));
        FIC_TRY(fic_copy_out(h, vals, d_vals, (size_t)count * 24, false, st));
        FIC_TRY(fic_sync(h, st));
    }
    return FIC_OK;
}

int fic_hv_encode(fic_handle* h, const uint8_t* img, int H, int W, int min_range_height, int min_range_width,
                  int max_range_height, int max_range_width, int scale_bits, int offset_bits, double err_threshold,
                  double max_scale, int32_t* nodes, double* vals, int cap, unsigned flags, void* stream)
{
    FIC_REQUIRE(h && img && nodes && vals, "fic_hv_encode: NULL argument");
    FIC_REQUIRE(H > 0 && W > 0, "fic_hv_encode: empty image");
    FIC_REQUIRE(min_range_height >= 1 && min_range_width >= 1, "fic_hv_encode: minimum range sides must be >= 1");
    FIC_REQUIRE(scale_bits >= 0 && scale_bits <= 30 && offset_bits >= 0 && offset_bits <= 30, "fic_hv_encode: bit widths out of range");
    FIC_ENTER(h, stream);
    const uint8_t* d_img;
    FIC_TRY(stage_image(h, img, H, W, flags & FIC_DEVICE_PTRS, st, &d_img));
    return hv_encode_impl(h, d_img, H, W, min_range_height, min_range_width, max_range_height, max_range_width, scale_bits,
                          offset_bits, err_threshold, max_scale, nodes, vals, cap, st);
}

static int check_leaves(fic_handle* h, const char* who, int kind, const int32_t* leaves, int n, int H, int W, bool domains)
{
    for (int k = 0; k < n; ++k) {
        const int32_t* L = leaves + 8 * k;
        FIC_REQUIRE(L[0] >= 0 && L[1] >= 0 && L[2] >= 1 && L[3] >= 1 && L[0] + L[2] <= H && L[1] + L[3] <= W,
                    "%s: leaf %d range (%d,%d,%dx%d) outside the image", who, k, L[0], L[1], L[2], L[3]);
        FIC_REQUIRE(!domains || (L[4] >= 0 && L[5] >= 0 && L[4] + 2 * L[2] <= H && L[5] + 2 * L[3] <= W),
                    "%s: leaf %d domain (%d,%d) does not fit the image", who, k, L[4], L[5]);
        FIC_REQUIRE(kind == FIC_TREE_HV || L[2] == L[3], "%s: quadtree leaf %d is not square", who, k);
    }
    return FIC_OK;
}

int fic_tree_decode(fic_handle* h, int kind, const int32_t* leaves, const double* so, int n, int H, int W, int iterations,
                    double stop_on_relative_error, int min_iterations, int max_iterations, int f32_affine, uint8_t* out,
                    int32_t* iters_done, unsigned flags, void* stream)
{
    FIC_REQUIRE(h && leaves && so && out, "fic_tree_decode: NULL argument");
    FIC_REQUIRE(kind == FIC_TREE_QUADTREE || kind == FIC_TREE_HV, "fic_tree_decode: unknown tree kind %d", kind);
    FIC_REQUIRE(n >= 0 && H > 0 && W > 0 && (iterations >= 0 || iterations == FIC_ITER_AUTO), "fic_tree_decode: bad sizes");
    FIC_REQUIRE(!(flags & FIC_QT_DEBLOCK) || kind == FIC_TREE_QUADTREE,
                "fic_tree_decode: the deblocking post-processor exists for the quadtree codec only");
    FIC_ENTER(h, stream);
    const bool dev = flags & FIC_DEVICE_PTRS;
    const unsigned post = flags & (FIC_QT_DEBLOCK | FIC_TREE_OUT_F64);
    const size_t out_bytes = (size_t)H * W * ((flags & FIC_TREE_OUT_F64) ? sizeof(double) : 1);
    const int32_t* d_leaves = leaves;
    const double* d_so = so;
    void* d_out = out;
    if (!dev) {
        FIC_TRY(check_leaves(h, "fic_tree_decode", kind, leaves, n, H, W, true));
        FIC_TRY(h->d_in_a.reserve((size_t)n * 32 + 16));
        FIC_TRY(h->d_in_b.reserve((size_t)n * 16 + 16));
        FIC_TRY(h->d_out_a.reserve(out_bytes));
        FIC_TRY(fic_copy_in(h, h->d_in_a.p, leaves, (size_t)n * 32, false, st));
        FIC_TRY(fic_copy_in(h, h->d_in_b.p, so, (size_t)n * 16, false, st));
        d_leaves = h->d_in_a.as<int32_t>();
        d_so = h->d_in_b.as<double>();
        d_out = h->d_out_a.p;
    }
    FIC_TRY(tree_decode_impl(h, kind, d_leaves, d_so, n, H, W, iterations, stop_on_relative_error, min_iterations,
                             max_iterations, f32_affine, post, d_out, iters_done, st));
    if (!dev) {
        FIC_TRY(fic_copy_out(h, out, d_out, out_bytes, false, st));
        FIC_TRY(fic_sync(h, st));
    }
    return FIC_OK;
}

int fic_qt_deblock(fic_handle* h, const double* img, const int32_t* leaves, int n, int H, int W, const double* weights,
                   int max_depth_range_size, double* out, unsigned flags, void* stream)
{
    FIC_REQUIRE(h && img && leaves && out, "fic_qt_deblock: NULL argument");
    FIC_REQUIRE(n >= 0 && H > 0 && W > 0, "fic_qt_deblock: bad sizes");
    FIC_ENTER(h, stream);
    const bool dev = flags & FIC_DEVICE_PTRS;
    const size_t bytes = (size_t)H * W * sizeof(double);
    const double* d_img = img;
    const int32_t* d_leaves = leaves;
    double* d_out = out;
    if (!dev) {
        FIC_TRY(check_leaves(h, "fic_qt_deblock", FIC_TREE_QUADTREE, leaves, n, H, W, false));
        FIC_TRY(h->d_in_a.reserve((size_t)n * 32 + 16));
        FIC_TRY(h->d_in_b.reserve(bytes));
        FIC_TRY(h->d_out_a.reserve(bytes));
        FIC_TRY(fic_copy_in(h, h->d_in_a.p, leaves, (size_t)n * 32, false, st));
        FIC_TRY(fic_copy_in(h, h->d_in_b.p, img, bytes, false, st));
        d_leaves = h->d_in_a.as<int32_t>();
        d_img = h->d_in_b.as<double>();
        d_out = h->d_out_a.as<double>();
    }
    FIC_TRY(qt_deblock_impl(h, d_img, d_leaves, n, H, W, weights, max_depth_range_size, d_out, st));
    if (!dev) {
        FIC_TRY(fic_copy_out(h, out, d_out, bytes, false, st));
        FIC_TRY(fic_sync(h, st));
    }
    return FIC_OK;
}

int fic_rgb_to_ycbcr(fic_handle* h, const uint8_t* rgb, int N, int M, int subsample, uint8_t* y, uint8_t* cb, uint8_t* cr,
                     unsigned flags, void* stream)
{
    FIC_REQUIRE(h && rgb && y && cb && cr, "fic_rgb_to_ycbcr: NULL argument");
    FIC_REQUIRE(N > 0 && M > 0 && (subsample == 0 || subsample == 1), "fic_rgb_to_ycbcr: bad sizes or subsampling mode");
    FIC_REQUIRE(!subsample || N % 2 == 0,
                "fic_rgb_to_ycbcr: 4:2:0 with an odd height %d (the reference raises or leaves the last chroma row unset)", N);
    FIC_ENTER(h, stream);
    const bool dev = flags & FIC_DEVICE_PTRS;
    const size_t npx = (size_t)N * M, nc = subsample ? (size_t)(N / 2) * ((M + 1) / 2) : npx;
    const uint8_t* d_rgb = rgb;
    uint8_t *d_y = y, *d_cb = cb, *d_cr = cr;
    if (!dev) {
        FIC_TRY(h->d_in_a.reserve(3 * npx));
        FIC_TRY(h->d_out_a.reserve(npx));
        FIC_TRY(h->d_out_b.reserve(nc));
        FIC_TRY(h->d_out_c.reserve(nc));
        FIC_TRY(fic_copy_in(h, h->d_in_a.p, rgb, 3 * npx, false, st));
        d_rgb = h->d_in_a.as<uint8_t>();
        d_y = h->d_out_a.as<uint8_t>();
        d_cb = h->d_out_b.as<uint8_t>();
        d_cr = h->d_out_c.as<uint8_t>();
    }
    FIC_TRY(rgb_to_ycbcr_impl(h, d_rgb, N, M, subsample, d_y, d_cb, d_cr, st));
    if (!dev) {
        FIC_TRY(fic_copy_out(h, y, d_y, npx, false, st));
        FIC_TRY(fic_copy_out(h, cb, d_cb, nc, false, st));
        FIC_TRY(fic_copy_out(h, cr, d_cr, nc, false, st));
        FIC_TRY(fic_sync(h, st));
    }
    return FIC_OK;
}

int fic_ycbcr_to_rgb(fic_handle* h, const uint8_t* y, const uint8_t* cb, const uint8_t* cr, int N, int M, int subsample,
                     uint8_t* rgb, unsigned flags, void* stream)
{
    FIC_REQUIRE(h && rgb && y && cb && cr, "fic_ycbcr_to_rgb: NULL argument");
    FIC_REQUIRE(N > 0 && M > 0 && (subsample == 0 || subsample == 1), "fic_ycbcr_to_rgb: bad sizes or subsampling mode");
    FIC_REQUIRE(!subsample || N % 4 == 0,
                "fic_ycbcr_to_rgb: 4:2:0 with height %d not a multiple of 4 (the reference raises, utils.py:73-76)", N);
    FIC_ENTER(h, stream);
    const bool dev = flags & FIC_DEVICE_PTRS;
    const size_t npx = (size_t)N * M, nc = subsample ? (size_t)(N / 2) * ((M + 1) / 2) : npx;
    const uint8_t *d_y = y, *d_cb = cb, *d_cr = cr;
    uint8_t* d_rgb = rgb;
    if (!dev) {
        FIC_TRY(h->d_in_a.reserve(npx));
        FIC_TRY(h->d_in_b.reserve(nc));
        FIC_TRY(h->d_img.reserve(nc));
        FIC_TRY(h->d_out_a.reserve(3 * npx));
        FIC_TRY(fic_copy_in(h, h->d_in_a.p, y, npx, false, st));
        FIC_TRY(fic_copy_in(h, h->d_in_b.p, cb, nc, false, st));
        FIC_TRY(fic_copy_in(h, h->d_img.p, cr, nc, false, st));
        d_y = h->d_in_a.as<uint8_t>();
        d_cb = h->d_in_b.as<uint8_t>();
        d_cr = h->d_img.as<uint8_t>();
        d_rgb = h->d_out_a.as<uint8_t>();
    }
    FIC_TRY(ycbcr_to_rgb_impl(h, d_y, d_cb, d_cr, N, M, subsample, d_rgb, st));
    if (!dev) {
        FIC_TRY(fic_copy_out(h, rgb, d_rgb, 3 * npx, false, st));
        FIC_TRY(fic_sync(h, st));
    }
    return FIC_OK;
}

}  // extern "C"
