// Internal shared definitions of libfic_b200.so (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/fic.h"

// thread-local error text behind fic_last_error()
void fic_set_error(const char* fmt, ...);

#define FIC_CUDA(call)                                                                            \
    do {                                                                                          \
        cudaError_t _e = (call);                                                                  \
        if (_e != cudaSuccess) {                                                                  \
            fic_set_error("%s failed at %s:%d: %s", #call, __FILE__, __LINE__, cudaGetErrorString(_e)); \
            return FIC_ECUDA;                                                                     \
        }                                                                                         \
    } while (0)

#define FIC_REQUIRE(cond, ...)       \
    do {                             \
        if (!(cond)) {               \
            fic_set_error(__VA_ARGS__); \
            return FIC_EINVAL;       \
        }                            \
    } while (0)

// A grow-only device buffer cached in the handle (the library's only allocations).
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes)
    {
        if (bytes <= cap) return FIC_OK;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = bytes + bytes / 4 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) {
            fic_set_error("cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e));
            return FIC_ENOMEM;
        }
        cap = want;
        return FIC_OK;
    }
    void release()
    {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <class T>
    T* as() const { return reinterpret_cast<T*>(p); }
};

struct PinnedBuf {
    void* p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes)
    {
        if (bytes <= cap) return FIC_OK;
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
        cudaError_t e = cudaMallocHost(&p, bytes + 256);
        if (e != cudaSuccess) {
            fic_set_error("cudaMallocHost(%zu) failed: %s", bytes, cudaGetErrorString(e));
            return FIC_ENOMEM;
        }
        cap = bytes + 256;
        return FIC_OK;
    }
    void release()
    {
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
    }
    template <class T>
    T* as() const { return reinterpret_cast<T*>(p); }
};

enum {
    CNT_LAUNCHES = 0,
    CNT_SCREEN_RANGES = 1,
    CNT_RESCAN_RANGES = 2,
    CNT_SURVIVORS = 3,
    CNT_H2D = 4,
    CNT_D2H = 5,
};

struct fic_handle {
    int device = 0;
    int num_sms = 0;
    cudaStream_t own_stream = nullptr;
    int64_t counters[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    // Order between calls: every call records `tail` on its stream when it returns; a call on a DIFFERENT stream first
    // makes its stream wait for that event, so the shared workspaces are never touched by two streams at once.
    cudaEvent_t tail = nullptr;
    cudaStream_t tail_stream = nullptr;
    bool tail_valid = false;
    // optional event timing of the dominant kernel (fic_profile)
    bool prof_on = false;
    bool prof_suppress = false;
    std::vector<cudaEvent_t> prof_events;  // start/stop pairs
    // generic staging
    DevBuf d_img, d_out_a, d_out_b, d_out_c, d_in_a, d_in_b;
    PinnedBuf pin_a, pin_b, pin_out, pin_stage, pin_tree;
    // Device -> host results are staged in pinned memory (a copy into pageable memory is a blocking, driver-staged transfer)
    // and handed to the caller's buffers at the next fic_sync() of the stream.
    struct PendingOut { void* dst; size_t at, bytes; };
    std::vector<PendingOut> pending_out;
    size_t stage_used = 0;
    // Host-buffer encodes: the records are copied to pinned memory by the pipeline itself, in front of the one host
    // synchronisation it needs anyway (control-block read-back), instead of by the entry point after it.
    struct EagerOut {
        bool armed = false, done = false;
        const void* src[3] = {nullptr, nullptr, nullptr};
        size_t bytes[3] = {0, 0, 0};
        size_t at[3] = {0, 0, 0};   // offsets into pin_out
    } eager;
    // brute force
    DevBuf bf_boxsum, bf_stats, bf_partial, bf_ranges;
    DevBuf bf_pool, bf_cand, bf_A, bf_thr, bf_surv, bf_ctrl, bf_best, bf_list, bf_done;
    bool qt_screen_attr_ready = false;
    bool bf4_attr_ready = false;        // bf_screen_kernel<8, 16> (4x4 brute force) has its dynamic shared memory limit raised
    bool hv_group_attr_ready = false;   // hv_search_grouped_kernel has its dynamic shared memory limit raised on this device
    bool screen_attr_ready = false;  // the screen / rescoring kernels have their dynamic shared memory limit raised on this device
    bool bf_totals_live = false;     // bf_ctrl holds device-side counter totals that fic_get_counters has to fetch
    // decoders
    DevBuf dec_img, dec_next, dec_map, dec_sched, dec_red, dec_keys, dec_roots;
    bool dec_shadow_ready = false;   // bf_decode_kernel<true> has its dynamic shared memory limit raised on this device
    // quadtree / hv
    DevBuf t_int1, t_int2, t_cls, t_lists, t_work, t_res, t_aux, t_hvpool, t_part, t_nodes, t_count, t_rcls, t_need;
};

static inline cudaStream_t fic_stream(fic_handle* h, void* s) { return s ? (cudaStream_t)s : h->own_stream; }

// Makes the handle's device current for the duration of an entry point and puts the caller's device back afterwards.
struct DeviceScope {
    int prev = -1;
    bool switched = false;
    cudaError_t err = cudaSuccess;
    explicit DeviceScope(int device)
    {
        if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
        if (prev != device) { err = cudaSetDevice(device); switched = err == cudaSuccess && prev >= 0; }
    }
    ~DeviceScope() { if (switched) cudaSetDevice(prev); }
};

// See fic_handle::tail.
struct StreamOrder {
    fic_handle* h;
    cudaStream_t st;
    StreamOrder(fic_handle* h_, cudaStream_t st_) : h(h_), st(st_)
    {
        if (h->tail_valid && h->tail_stream != st) cudaStreamWaitEvent(st, h->tail, 0);
    }
    ~StreamOrder()
    {
        if (cudaEventRecord(h->tail, st) == cudaSuccess) { h->tail_stream = st; h->tail_valid = true; }
        else cudaGetLastError();
    }
};

// First lines of every entry point that works on a handle: device, stream, ordering behind the previous call.
#define FIC_ENTER(h, stream)                                  \
    DeviceScope _fic_dev((h)->device);                        \
    FIC_CUDA(_fic_dev.err);                                   \
    cudaStream_t st = fic_stream((h), (stream));              \
    StreamOrder _fic_order((h), st)

// copy helpers that account traffic; `dev` selects device-to-device when the caller passed device pointers
int fic_copy_in(fic_handle* h, void* dst_dev, const void* src, size_t bytes, bool src_is_dev, cudaStream_t st);
int fic_copy_out(fic_handle* h, void* dst, const void* src_dev, size_t bytes, bool dst_is_dev, cudaStream_t st);
// cudaStreamSynchronize + delivery of the staged results of fic_copy_out; every host wait of the library goes through it
int fic_sync(fic_handle* h, cudaStream_t st);

#define FIC_LAUNCHED(h) ((h)->counters[CNT_LAUNCHES]++)

// RAII bracket: records a start event now and a stop event on destruction when profiling is enabled
struct ProfScope {
    fic_handle* h;
    cudaStream_t st;
    ProfScope(fic_handle* h_, cudaStream_t st_) : h(h_), st(st_)
    {
        if (!h->prof_on || h->prof_suppress) { h = nullptr; return; }
        cudaEvent_t a, b;
        if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) { h = nullptr; return; }
        h->prof_events.push_back(a);
        h->prof_events.push_back(b);
        cudaEventRecord(a, st);
    }
    ~ProfScope()
    {
        if (h && h->prof_on) cudaEventRecord(h->prof_events.back(), st);
    }
};

static inline int fic_check_launch(const char* what)
{
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        fic_set_error("launch of %s failed: %s", what, cudaGetErrorString(e));
        return FIC_ECUDA;
    }
    return FIC_OK;
}
#define FIC_CHECK_LAUNCH(h, what)                 \
    do {                                          \
        FIC_LAUNCHED(h);                          \
        int _rc = fic_check_launch(what);         \
        if (_rc != FIC_OK) return _rc;            \
    } while (0)
#define FIC_TRY(expr)                  \
    do {                               \
        int _rc = (expr);              \
        if (_rc < 0) return _rc;       \
    } while (0)

// ---- entry points implemented in the per-codec translation units -------------------------
int bf_encode_impl(fic_handle* h, const uint8_t* d_img, int H, int W, int range, int si, int sj, int mode,
                   int range_begin, int range_end, int32_t* d_idx, double* d_so, double* d_R, unsigned flags,
                   cudaStream_t st);
int bf_decode_impl(fic_handle* h, int planes, const int32_t* h_idx, const double* h_so, int n, int width, int height,
                   int iterations, uint8_t* d_out, cudaStream_t st);
int tree_decode_impl(fic_handle* h, int kind, const int32_t* d_leaves, const double* d_so, int n, int H, int W,
                     int iterations, double stop, int min_it, int max_it, int f32_affine, unsigned post, void* d_out,
                     int32_t* iters_done, cudaStream_t st);
int qt_deblock_impl(fic_handle* h, const double* d_img, const int32_t* d_leaves, int n, int H, int W, const double* w4,
                    int small_side, double* d_res, cudaStream_t st);
int rgb_to_ycbcr_impl(fic_handle* h, const uint8_t* d_rgb, int N, int M, int subsample, uint8_t* d_y, uint8_t* d_cb,
                      uint8_t* d_cr, cudaStream_t st);
int ycbcr_to_rgb_impl(fic_handle* h, const uint8_t* d_y, const uint8_t* d_cb, const uint8_t* d_cr, int N, int M,
                      int subsample, uint8_t* d_rgb, cudaStream_t st);
int subsample_impl(fic_handle* h, const double* d_arr, int h2, int w2, int jit, double* d_out, cudaStream_t st);
int qt_classify_impl(fic_handle* h, const uint8_t* d_img, int H, int W, const int32_t* d_blocks, int count,
                     int32_t* d_out, cudaStream_t st);
int qt_encode_impl(fic_handle* h, const uint8_t* d_img, int H, int W, int scale_bits, int offset_bits, double max_scale,
                   int full, int onlypos, int min_side, int max_side, double tol, int hstep, int vstep,
                   int32_t* nodes, double* vals, int cap, int32_t* n_roots, cudaStream_t st);
int hv_search_impl(fic_handle* h, const uint8_t* d_img, int H, int W, const int32_t* d_rects, int count, int scale_bits,
                   int offset_bits, double max_scale, int32_t* d_rec, double* d_vals, cudaStream_t st);
int hv_encode_impl(fic_handle* h, const uint8_t* d_img, int H, int W, int min_h, int min_w, int max_h, int max_w,
                   int scale_bits, int offset_bits, double thr, double max_scale, int32_t* nodes, double* vals,
                   int cap, cudaStream_t st);
