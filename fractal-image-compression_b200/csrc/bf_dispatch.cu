// Brute-force encode dispatch: chooses between the tensor-core screen pipeline (bf_screen.cu) and the exact
// CUDA-core scan (bf_exact.cu).  Both produce the reference's records bit for bit; the screen is only a filter.
#include "fic_common.cuh"

int bf_prepare8(fic_handle* h, const uint8_t* d_img, int H, int W, bool with_stats, cudaStream_t st);
int bf_exact_scan8(fic_handle* h, const uint8_t* d_img, int H, int W, int si, int sj, int mode, int range_begin,
                   int n_ranges, const int32_t* d_range_list, int32_t* d_idx, double* d_so, double* d_R,
                   cudaStream_t st);
int bf_exact_generic(fic_handle* h, const uint8_t* d_img, int H, int W, int range, int si, int sj, int mode,
                     int range_begin, int n_ranges, int32_t* d_idx, double* d_so, double* d_R, cudaStream_t st);
int bf_screen_encode8(fic_handle* h, const uint8_t* d_img, int H, int W, int mode, int range_begin, int n_ranges, int32_t* d_idx,
                      double* d_so, double* d_R, unsigned flags, cudaStream_t st);
bool bf_screen_applicable(int H, int W);
bool bf4_tc_applicable(int H, int W, int si, int sj, int mode);
int bf4_tc_encode(fic_handle* h, const uint8_t* d_img, int H, int W, int si, int sj, int mode, int range_begin, int n_ranges,
                  int32_t* d_idx, double* d_so, double* d_R, cudaStream_t st);

int bf_encode_impl(fic_handle* h, const uint8_t* d_img, int H, int W, int range, int si, int sj, int mode,
                   int range_begin, int range_end, int32_t* d_idx, double* d_so, double* d_R, unsigned flags,
                   cudaStream_t st)
{
    const int n = range_end - range_begin;
    // 4x4 ranges: exact pass on the tensor cores (K = 16, every sum of either encoder is exact); every other size: the generic scan
    if (range == 4 && !(flags & FIC_BF_EXACT_ONLY) && bf4_tc_applicable(H, W, si, sj, mode))
        return bf4_tc_encode(h, d_img, H, W, si, sj, mode, range_begin, n, d_idx, d_so, d_R, st);
    if (range != 8) return bf_exact_generic(h, d_img, H, W, range, si, sj, mode, range_begin, n, d_idx, d_so, d_R, st);
    const long long nq = (long long)(H - 16 > 0 ? (H - 16 + si - 1) / si : 0) * (W - 16 > 0 ? (W - 16 + sj - 1) / sj : 0);
    FIC_REQUIRE(nq * 8 < 0xffffffffll, "fic_bf_encode: %lld candidates per range exceed the 32-bit scan index", nq * 8);
    const bool screen = si == 1 && sj == 1 && !(flags & FIC_BF_EXACT_ONLY) &&
                        bf_screen_applicable(H, W);
    FIC_TRY(bf_prepare8(h, d_img, H, W, !screen, st));   // the screen's pool kernel computes the position statistics itself
    if (screen) return bf_screen_encode8(h, d_img, H, W, mode, range_begin, n, d_idx, d_so, d_R, flags, st);
    return bf_exact_scan8(h, d_img, H, W, si, sj, mode, range_begin, n, nullptr, d_idx, d_so, d_R, st);
}
