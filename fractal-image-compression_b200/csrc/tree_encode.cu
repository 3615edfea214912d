// Quadtree and HV encoders: split decisions batched per tree level on the device.
//
// quadtree/encoder.py:67-384 -- classification (_classify) from integral images in exact integer arithmetic,
//   72-class domain pools, one CTA per pending range scanning its candidate lists with the float32-sequential
//   replay of _find_min_err_domain / _find_optimal_mean_square_error / _calc_mean_square_error.
// hvv2/encoder.py:24-188 -- one CTA per pending rectangle over its domain lattice x 4 isometries
//   (_run_domain_search), one warp per rectangle for the biased-difference split (_compute_biased_sum_deltas).
//
// Level-synchronous driver: all pending ranges of one size (quadtree) / one depth (HV) are searched by a single
// launch; the host reads back one error per range, decides leaf vs split exactly as the reference does, and
// queues the children.  The preorder node stream of include/fic.h is emitted at the end.
//
// HBM layout: I1, I2 int64 [(H+1),(W+1)] integral images of v and v^2; pools as CSR (offsets[73], rows[.][4]).
#include <math.h>

#include <algorithm>

#include "fic_common.cuh"
#include "scorer.cuh"

#define TE_THREADS 128
#define HV_MINB 4        // resident CTAs per SM the HV search is compiled for (128 registers; 5 spills and loses, 3 gains nothing)
// A rectangle's candidate list is cut into S chunks, one CTA each.  The work of a chunk varies with the rectangle (small
// rectangles have many positions) and a level's CTAs run in several waves, so fine chunks balance far better than "a few
// CTAs per SM": same box, lena 512x512, S aimed at 4 / 8 / 16 / 32 / 64 / 128 CTAs per SM: 70.9 / 62.8 / 54.6 / 44.7 / 42.1 /
// 44.3 ms per encode.
#define HV_SMULT 64
#define HV_SCAP 256

// ---- integral images ----------------------------------------------------------------------------
// rows: one warp per image row, 32 pixels per step, warp inclusive scan plus a running carry (coalesced reads and writes)
__global__ void integral_rows_kernel(const uint8_t* __restrict__ img, int H, int W, long long* __restrict__ I1,
                                     long long* __restrict__ I2)
{
    const int y = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (y > H) return;
    const int W1 = W + 1;
    long long* r1 = I1 + (size_t)y * W1;
    long long* r2 = I2 + (size_t)y * W1;
    if (y == 0) {
        for (int x = lane; x <= W; x += 32) { r1[x] = 0; r2[x] = 0; }
        return;
    }
    if (lane == 0) { r1[0] = 0; r2[0] = 0; }
    const uint8_t* src = img + (size_t)(y - 1) * W;
    long long cs = 0, cq = 0;   // carries: sums of everything left of this step
    for (int x0 = 0; x0 < W; x0 += 32) {
        const int x = x0 + lane;
        const long long v = x < W ? src[x] : 0;
        long long s = v, q = v * v;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const long long ts = __shfl_up_sync(0xffffffffu, s, off), tq = __shfl_up_sync(0xffffffffu, q, off);
            if (lane >= off) { s += ts; q += tq; }
        }
        if (x < W) { r1[x + 1] = cs + s; r2[x + 1] = cq + q; }
        cs += __shfl_sync(0xffffffffu, s, 31);
        cq += __shfl_sync(0xffffffffu, q, 31);
    }
}
// columns: a CTA takes 32 columns, its sixteen warps sixteen bands of rows: every thread sums its band of its column,
// the band totals are scanned in shared memory, then the band is walked again (L1/L2) with the running sum and stored.
// (One thread per column over all rows was 56 us at 512x512 -- 513 threads on the whole GPU.)
#define INTEGRAL_BANDS 16
__global__ void __launch_bounds__(32 * INTEGRAL_BANDS)
integral_cols_kernel(int H, int W, long long* __restrict__ I1, long long* __restrict__ I2)
{
    __shared__ long long tot1[INTEGRAL_BANDS][32], tot2[INTEGRAL_BANDS][32];
    const int lane = threadIdx.x & 31, band = threadIdx.x >> 5;
    const int x = blockIdx.x * 32 + lane;
    const int W1 = W + 1;
    const int per = (H + INTEGRAL_BANDS - 1) / INTEGRAL_BANDS;
    const int y0 = 1 + band * per, y1 = y0 + per <= H + 1 ? y0 + per : H + 1;    // rows [y0, y1)
    const bool live = x <= W;
    long long s = 0, q = 0;
    if (live)
        for (int y = y0; y < y1; ++y) {
            s += I1[(size_t)y * W1 + x];
            q += I2[(size_t)y * W1 + x];
        }
    tot1[band][lane] = s;
    tot2[band][lane] = q;
    __syncthreads();
    s = 0; q = 0;
    for (int b = 0; b < band; ++b) { s += tot1[b][lane]; q += tot2[b][lane]; }
    if (!live) return;
    constexpr int U = 8;     // rows loaded before any is stored, so that the loads are not serialised behind the in-place stores
    for (int yb = y0; yb < y1; yb += U) {
        long long a[U], c[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const bool in = yb + u < y1;
            a[u] = in ? I1[(size_t)(yb + u) * W1 + x] : 0;
            c[u] = in ? I2[(size_t)(yb + u) * W1 + x] : 0;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (yb + u < y1) {
                s += a[u];
                q += c[u];
                I1[(size_t)(yb + u) * W1 + x] = s;
                I2[(size_t)(yb + u) * W1 + x] = q;
            }
        }
    }
}
__device__ __forceinline__ long long box(const long long* __restrict__ I, int W1, int y0, int x0, int y1, int x1)
{
    return I[(size_t)y1 * W1 + x1] - I[(size_t)y0 * W1 + x1] - I[(size_t)y1 * W1 + x0] + I[(size_t)y0 * W1 + x0];
}

static int build_integrals(fic_handle* h, const uint8_t* d_img, int H, int W, cudaStream_t st)
{
    size_t n = (size_t)(H + 1) * (W + 1);
    FIC_TRY(h->t_int1.reserve(n * sizeof(long long)));
    FIC_TRY(h->t_int2.reserve(n * sizeof(long long)));
    integral_rows_kernel<<<((H + 1) * 32 + 127) / 128, 128, 0, st>>>(d_img, H, W, h->t_int1.as<long long>(), h->t_int2.as<long long>());
    FIC_CHECK_LAUNCH(h, "integral_rows_kernel");
    integral_cols_kernel<<<(W + 1 + 31) / 32, 32 * INTEGRAL_BANDS, 0, st>>>(H, W, h->t_int1.as<long long>(), h->t_int2.as<long long>());
    FIC_CHECK_LAUNCH(h, "integral_cols_kernel");
    return FIC_OK;
}

// ---- _classify (quadtree/encoder.py:195-259), integer-exact ------------------------
__device__ void classify_block(const long long* __restrict__ I1, const long long* __restrict__ I2, int W1, int si, int sj,
                               int side, int neg, int& omc, int& ovc, int& orot, int& oorient)
{
    const int hs = side >> 1;
    const long long m = (long long)hs * hs;
    long long S[4], V[4];
    const int oi[4] = {0, 0, hs, hs}, oj[4] = {0, hs, 0, hs};  // a1 TL, a2 TR, a3 BL, a4 BR
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        long long s = box(I1, W1, si + oi[q], sj + oj[q], si + oi[q] + hs, sj + oj[q] + hs);
        long long s2 = box(I2, W1, si + oi[q], sj + oj[q], si + oi[q] + hs, sj + oj[q] + hs);
        S[q] = neg ? -s : s;
        V[q] = m * s2 - s * s;
    }
    long long means[4] = {S[0], S[1], S[3], S[2]};  // [a1,a2,a4,a3], :213
    long long vars[4] = {V[0], V[1], V[3], V[2]};
    int am = 0;
#pragma unroll
    for (int k = 1; k < 4; ++k)
        if (means[k] > means[am]) am = k;  // np.argmax: first maximum
    const int rotation = (4 - am) & 3;
    long long mr[4], vr[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {  // np.roll(x, rotation)
        mr[(k + rotation) & 3] = means[k];
        vr[(k + rotation) & 3] = vars[k];
    }
    const int orientation = mr[1] < mr[3] ? 1 : 0;
    if (orientation) {
        long long t = mr[1]; mr[1] = mr[3]; mr[3] = t;
        t = vr[1]; vr[1] = vr[3]; vr[3] = t;
    }
    const long long A2 = mr[1], A3 = mr[3], A4 = mr[2];
    long long ov[4] = {vr[0], vr[1], vr[3], vr[2]};
    int oid[4] = {0, 1, 2, 3};
    int mc;
    if (A2 >= A3 && A3 >= A4) mc = 0;
    else if (A2 >= A4 && A4 >= A3) mc = 1;
    else mc = 2;
    for (int i = 1; i < 4; ++i)
        for (int j = i; j > 0; --j) {
            if (ov[j] > ov[j - 1]) {
                long long t = ov[j]; ov[j] = ov[j - 1]; ov[j - 1] = t;
                int ti = oid[j]; oid[j] = oid[j - 1]; oid[j - 1] = ti;
            } else break;
        }
    int vc = 0;
    for (int i = 2; i >= 0; --i)
        for (int j = 0; j <= i; ++j)
            if (oid[j] > oid[j + 1]) {
                int t = oid[j]; oid[j] = oid[j + 1]; oid[j + 1] = t;
                if (oid[j] == 0 || oid[j + 1] == 0) vc += 6;
                else if (oid[j] == 1 || oid[j + 1] == 1) vc += 2;
                else if (oid[j] == 2 || oid[j + 1] == 2) vc += 1;
            }
    omc = mc; ovc = vc; orot = rotation; oorient = orientation;
}

__global__ void classify_blocks_kernel(const long long* __restrict__ I1, const long long* __restrict__ I2, int W1,
                                       const int32_t* __restrict__ blocks, int count, int32_t* __restrict__ out)
{
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= count) return;
    int mc, vc, rot, orient;
    classify_block(I1, I2, W1, blocks[4 * k], blocks[4 * k + 1], blocks[4 * k + 2], blocks[4 * k + 3], mc, vc, rot, orient);
    out[4 * k] = mc; out[4 * k + 1] = vc; out[4 * k + 2] = rot; out[4 * k + 3] = orient;
}

// all grid domains of one pool level (quadtree/encoder.py:261-283): out[pos] = {mc*24+vc, rot, orient}
__global__ void classify_grid_kernel(const long long* __restrict__ I1, const long long* __restrict__ I2, int W1, int side,
                                     int nv, int nh, int vstep, int hstep, int32_t* __restrict__ out)
{
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nv * nh) return;
    int i = k / nh, j = k - i * nh;
    int mc, vc, rot, orient;
    classify_block(I1, I2, W1, i * vstep, j * hstep, side, 0, mc, vc, rot, orient);
    out[3 * (size_t)k] = mc * 24 + vc;
    out[3 * (size_t)k + 1] = rot;
    out[3 * (size_t)k + 2] = orient;
}

// ---- 72-class domain pools on the device (quadtree/encoder.py:261-283) ------------------------------------------
// A stable counting sort of the grid positions by class: the reference appends the domains of a pool size to their
// class lists in row-major grid order, so inside a class the order is the position index.  Three launches per pool
// size and no host round trip: (1) per-CTA class histograms over contiguous chunks of positions, (2) one CTA turns them
// into per-chunk start offsets and the CSR offsets[73] -- a class without domains gets one placeholder row (0,0,0,0),
// encoder.py:279-280 -- (3) every CTA scatters its chunk in order (ranks from warp match masks + running class counters).
#define QP_THREADS 256
#define QP_CHUNK 2048     // positions per CTA

__global__ void __launch_bounds__(QP_THREADS)
qt_pool_count_kernel(const int32_t* __restrict__ cls3, int cnt, int32_t* __restrict__ blockhist /*[nblocks][72]*/)
{
    __shared__ int hist[72];
    if (threadIdx.x < 72) hist[threadIdx.x] = 0;
    __syncthreads();
    const int p0 = blockIdx.x * QP_CHUNK;
    for (int p = p0 + threadIdx.x; p < min(p0 + QP_CHUNK, cnt); p += QP_THREADS) atomicAdd(&hist[cls3[3 * (size_t)p]], 1);
    __syncthreads();
    if (threadIdx.x < 72) blockhist[(size_t)blockIdx.x * 72 + threadIdx.x] = hist[threadIdx.x];
}

__global__ void qt_pool_scan_kernel(int32_t* __restrict__ blockhist, int nblocks, int32_t* __restrict__ offsets /*[73]*/,
                                    int32_t* __restrict__ rows)
{
    __shared__ int total[72], off[73];
    const int c = threadIdx.x;
    if (c < 72) {
        int acc = 0;
        for (int b = 0; b < nblocks; ++b) {   // exclusive prefix over the chunks, in place
            const int v = blockhist[(size_t)b * 72 + c];
            blockhist[(size_t)b * 72 + c] = acc;
            acc += v;
        }
        total[c] = acc;
    }
    __syncthreads();
    if (c == 0) {
        int acc = 0;
        for (int k = 0; k < 72; ++k) { off[k] = acc; acc += total[k] > 0 ? total[k] : 1; }
        off[72] = acc;
    }
    __syncthreads();
    if (c < 73) offsets[c] = off[c];
    if (c < 72 && total[c] == 0) {   // placeholder row of an empty class
        int32_t* r = rows + 4 * (size_t)off[c];
        r[0] = r[1] = r[2] = r[3] = 0;
    }
}

__global__ void __launch_bounds__(QP_THREADS)
qt_pool_scatter_kernel(const int32_t* __restrict__ cls3, int cnt, int nh, int vstep, int hstep,
                       const int32_t* __restrict__ blockhist, const int32_t* __restrict__ offsets, int32_t* __restrict__ rows)
{
    __shared__ int run[72];                         // next free row of every class for this chunk
    __shared__ int wcnt[QP_THREADS / 32][72];       // this pass: members of every class per warp
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid < 72) run[tid] = offsets[tid] + blockhist[(size_t)blockIdx.x * 72 + tid];
    const int p0 = blockIdx.x * QP_CHUNK, p1 = min(p0 + QP_CHUNK, cnt);
    for (int base = p0; base < p1; base += QP_THREADS) {
        for (int e = tid; e < (QP_THREADS / 32) * 72; e += QP_THREADS) (&wcnt[0][0])[e] = 0;
        __syncthreads();
        const int p = base + tid;
        const bool live = p < p1;
        int c = 0, rot = 0, orient = 0, rank = 0;
        if (live) { c = cls3[3 * (size_t)p]; rot = cls3[3 * (size_t)p + 1]; orient = cls3[3 * (size_t)p + 2]; }
        const unsigned act = __ballot_sync(0xffffffffu, live);
        if (live) {
            const unsigned same = __match_any_sync(act, c);
            rank = __popc(same & ((1u << lane) - 1u));
            if (rank == 0) wcnt[warp][c] = __popc(same);
        }
        __syncthreads();
        if (live) {
            int at = run[c] + rank;
            for (int w = 0; w < warp; ++w) at += wcnt[w][c];
            int32_t* r = rows + 4 * (size_t)at;
            r[0] = (p / nh) * vstep; r[1] = (p % nh) * hstep; r[2] = rot; r[3] = orient;
        }
        __syncthreads();
        if (tid < 72) {
            int add = 0;
            for (int w = 0; w < QP_THREADS / 32; ++w) add += wcnt[w][tid];
            run[tid] += add;
        }
        __syncthreads();
    }
}

int qt_classify_impl(fic_handle* h, const uint8_t* d_img, int H, int W, const int32_t* d_blocks, int count, int32_t* d_out,
                     cudaStream_t st)
{
    if (count == 0) return FIC_OK;
    FIC_TRY(build_integrals(h, d_img, H, W, st));
    classify_blocks_kernel<<<(count + 127) / 128, 128, 0, st>>>(h->t_int1.as<long long>(), h->t_int2.as<long long>(), W + 1,
                                                                d_blocks, count, d_out);
    FIC_CHECK_LAUNCH(h, "classify_blocks_kernel");
    return FIC_OK;
}

// ---- shared CTA reduction: first strict minimum in scan order -----------------------------------
struct TeBest {
    double R, s, o;
    unsigned long long seq;
    int di, dj, orient, rot, qs, qo;
};
__device__ __forceinline__ void te_init(TeBest& b)
{
    b.R = __longlong_as_double(0x7ff0000000000000ll);
    b.seq = ~0ull;
    b.s = 0.0; b.o = 0.0;
    b.di = b.dj = b.orient = b.rot = b.qs = b.qo = 0;
}
__device__ void te_block_reduce(TeBest& b, int32_t* __restrict__ rec6, double* __restrict__ vals3,
                                unsigned long long* __restrict__ seq_out = nullptr)
{
    __shared__ double sR[TE_THREADS / 32];
    __shared__ unsigned long long sSeq[TE_THREADS / 32];
    __shared__ int winner;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    double R = b.R;
    unsigned long long seq = b.seq;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        double oR = __shfl_xor_sync(0xffffffffu, R, off);
        unsigned long long oS = __shfl_xor_sync(0xffffffffu, seq, off);
        if (fic_better(oR, oS, R, seq)) { R = oR; seq = oS; }
    }
    if (lane == 0) { sR[warp] = R; sSeq[warp] = seq; }
    if (tid == 0) winner = -1;
    __syncthreads();
    double bR = sR[0];
    unsigned long long bS = sSeq[0];
    for (int w = 1; w < TE_THREADS / 32; ++w)
        if (fic_better(sR[w], sSeq[w], bR, bS)) { bR = sR[w]; bS = sSeq[w]; }
    if (bS != ~0ull && b.seq == bS) winner = tid;  // seq values are unique per CTA
    __syncthreads();
    if (winner < 0) {
        if (tid == 0) {
            for (int k = 0; k < 6; ++k) rec6[k] = 0;
            vals3[0] = 0.0; vals3[1] = 0.0;
            vals3[2] = __longlong_as_double(0x7ff0000000000000ll);
            if (seq_out) *seq_out = ~0ull;
        }
    } else if (tid == winner) {
        rec6[0] = b.di; rec6[1] = b.dj; rec6[2] = b.orient; rec6[3] = b.rot; rec6[4] = b.qs; rec6[5] = b.qo;
        vals3[0] = b.s; vals3[1] = b.o; vals3[2] = b.R;
        if (seq_out) *seq_out = b.seq;
    }
}

// ---- quadtree range search (quadtree/encoder.py:148-193, 285-322) ---------------------------------
struct QtPoolDev {
    const int32_t* offsets;  // [73]
    const int32_t* rows;     // [offsets[72]][4] = (i, j, rot, orient)
};

#define QT_RANGE_SMEM 1024

__global__ void __launch_bounds__(TE_THREADS)
qt_search_kernel(const uint8_t* __restrict__ img, const float* __restrict__ Dp, int W, const long long* __restrict__ I1,
                 const long long* __restrict__ I2,
                 const int32_t* __restrict__ ranges /*[n][4] = si,sj,node,-*/, int side, QtPoolDev pool, int full, int nmodes,
                 double max_scale, int scale_bits, int offset_bits, int32_t* __restrict__ rec /*[n][6]*/,
                 double* __restrict__ vals /*[n][3]*/,
                 const int32_t* __restrict__ need /* optional: only ranges with need[k] != 0 are searched (the others were decided by
                                                     the tensor-core screen, qt_tc_search) */)
{
    if (need && !need[blockIdx.x]) return;
    __shared__ float s_rs, s_sqr;
    __shared__ int seg_begin[2 * 72 + 1];  // prefix of candidate counts over (mode, class) segments
    __shared__ int seg_list[2 * 72];       // class index of the segment
    __shared__ int seg_rot[2], seg_orient[2];
    __shared__ int n_seg;
    __shared__ unsigned long long s_bound;   // bit pattern of the best score found so far by this CTA (+inf at start)
    __shared__ float s_range[QT_RANGE_SMEM];  // the range block as floats (sides up to 32)
    const int tid = threadIdx.x;
    const int si = ranges[4 * (size_t)blockIdx.x], sj = ranges[4 * (size_t)blockIdx.x + 1];
    const int nn = side * side;
    const bool in_smem = nn <= QT_RANGE_SMEM;
    if (in_smem)
        for (int t = tid; t < nn; t += TE_THREADS) s_range[t] = (float)img[(size_t)(si + t / side) * W + sj + t % side];
    if (tid == 0) {
        s_bound = 0x7ff0000000000000ull;
        float rs = 0.f, sq = 0.f;
        for (int t = 0; t < nn; ++t) rs = __fadd_rn(rs, (float)img[(size_t)(si + t / side) * W + sj + t % side]);
        for (int t = 0; t < nn; ++t) {
            float v = (float)img[(size_t)(si + t / side) * W + sj + t % side];
            sq = __fadd_rn(sq, __fmul_rn(v, v));
        }
        s_rs = rs;
        s_sqr = sq;
        int ns = 0, acc = 0;
        for (int mode = 0; mode < nmodes; ++mode) {
            int mc, vc, rot, orient;
            classify_block(I1, I2, W + 1, si, sj, side, mode, mc, vc, rot, orient);
            seg_rot[mode] = rot;
            seg_orient[mode] = orient;
            int c0 = full ? 0 : mc * 24 + vc, c1 = full ? 72 : c0 + 1;
            for (int c = c0; c < c1; ++c) {
                seg_begin[ns] = acc;
                seg_list[ns] = c | (mode << 8);
                acc += pool.offsets[c + 1] - pool.offsets[c];
                ++ns;
            }
        }
        seg_begin[ns] = acc;
        n_seg = ns;
    }
    __syncthreads();
    const float rs = s_rs, sqr = s_sqr;
    const double dn = (double)nn;
    const double slack_a = DS((double)sqr, DD(DM((double)rs, (double)rs), dn));   // sqr - rs^2 / n
    const int total = seg_begin[n_seg];
    TeBest best;
    te_init(best);
    for (int c = tid; c < total; c += TE_THREADS) {
        int lo = 0, hi = n_seg - 1;  // last segment with seg_begin <= c
        while (lo < hi) {
            int mid = (lo + hi + 1) >> 1;
            if (seg_begin[mid] <= c) lo = mid; else hi = mid - 1;
        }
        const int cls = seg_list[lo] & 255, mode = seg_list[lo] >> 8;
        const int32_t* row = pool.rows + 4 * (size_t)(pool.offsets[cls] + (c - seg_begin[lo]));
        const int di = row[0], dj = row[1], drot = row[2], dorient = row[3];
        const int rrot = seg_rot[mode], rorient = seg_orient[mode];
        int d_rot, d_orient;
        if (dorient == rorient) { d_rot = (drot - rrot) & 3; d_orient = 0; }
        else { d_rot = (drot - rrot + 2) & 3; d_orient = 1; }
        // The isometry is affine in (i, j): the pooled sample of range pixel (i, j) sits at  o00 + i * step_i + j * step_j
        // in the pooled image (2x2 means at every pixel offset, exact in float32 -- hv_pool_kernel), so the map is
        // evaluated at three corners per candidate instead of once per pixel.
        int a00, b00, a10, b10, a01, b01;
        fic_iso_qt_enc(d_orient, d_rot, side, 0, 0, a00, b00);
        fic_iso_qt_enc(d_orient, d_rot, side, 1, 0, a10, b10);
        fic_iso_qt_enc(d_orient, d_rot, side, 0, 1, a01, b01);
        const int PW = W - 1;
        const int step_i = 2 * ((a10 - a00) * PW + (b10 - b00)), step_j = 2 * ((a01 - a00) * PW + (b01 - b00));
        const float* drow = Dp + (size_t)(di + 2 * a00) * PW + dj + 2 * b00;
        float ds = 0.f, sds = 0.f, ab = 0.f;
        int t = 0;
        for (int i = 0; i < side; ++i, drow += step_i) {
            const float* q = drow;
            for (int j = 0; j < side; ++j, q += step_j, ++t) {
                const float d = *q;
                const float r = in_smem ? s_range[t] : (float)img[(size_t)(si + i) * W + sj + j];
                ds = __fadd_rn(ds, d);
                sds = __fadd_rn(sds, __fmul_rn(d, d));
                ab = __fadd_rn(ab, __fmul_rn(d, r));
            }
        }
        {   // exact lower bound of the score (minimum of its quadratic form, see hv_search_kernel) against the CTA's best
            const double bound = __longlong_as_double((long long)*(volatile unsigned long long*)&s_bound);
            const double room = DS(slack_a, DM(dn, bound));
            if (room > 0.0) {
                const double dsd = (double)ds;
                const double den = DS(DM(dn, (double)sds), DM(dsd, dsd));
                const double num = DS(DM(dn, (double)ab), DM((double)rs, dsd));
                if (den > 0.0 && DM(num, num) < DM(DM(DM(room, dn), den), 0.999999999)) continue;
            }
        }
        QScore sc = fic_q_score(nn, rs, sqr, ds, sds, ab, max_scale, scale_bits, offset_bits);
        if (fic_better(sc.R, (unsigned long long)c, best.R, best.seq)) {
            best.R = sc.R; best.seq = c; best.s = sc.s; best.o = sc.o; best.qs = sc.qs; best.qo = sc.qo;
            best.di = di; best.dj = dj; best.orient = d_orient; best.rot = d_rot;
            atomicMin(&s_bound, (unsigned long long)__double_as_longlong(sc.R > 0.0 ? sc.R : 0.0));
        }
    }
    te_block_reduce(best, rec + 6 * (size_t)blockIdx.x, vals + 3 * (size_t)blockIdx.x);
}

// The same search with a WARP per range, for the class-restricted case (full_class_search off) and blocks of at most 64
// pixels: a range then meets the one or two class lists of its own classification -- a hundred candidates or so in a default
// pool -- and a CTA per range spends its time in the serial prologue of one thread and in block barriers (ncu launch list of
// the default 512x512 encode: 278 us for 15 916 CTAs, 10 us per CTA).  Here the lanes load the block, two lanes classify the
// two scale signs at the same time, the candidates are strided over the lanes and the winner is found by shuffles: no
// block barrier at all, four ranges per CTA.  Scoring, order and tie-break are qt_search_kernel's.
#define QT_WARP_RANGES 4
__global__ void __launch_bounds__(QT_WARP_RANGES * 32)
qt_search_warp_kernel(const uint8_t* __restrict__ img, const float* __restrict__ Dp, int W, const long long* __restrict__ I1,
                      const long long* __restrict__ I2, const int32_t* __restrict__ ranges /*[n][4] = si,sj,node,-*/, int n, int side,
                      QtPoolDev pool, int nmodes, double max_scale, int scale_bits, int offset_bits,
                      int32_t* __restrict__ rec /*[n][6]*/, double* __restrict__ vals /*[n][3]*/)
{
    __shared__ float s_range[QT_WARP_RANGES][64];
    __shared__ unsigned long long s_bound[QT_WARP_RANGES];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = blockIdx.x * QT_WARP_RANGES + warp;
    if (q >= n) return;
    const int si = ranges[4 * (size_t)q], sj = ranges[4 * (size_t)q + 1];
    const int nn = side * side;
    for (int t = lane; t < nn; t += 32) s_range[warp][t] = (float)img[(size_t)(si + t / side) * W + sj + t % side];
    if (lane == 0) s_bound[warp] = 0x7ff0000000000000ull;
    int mc = 0, vc = 0, rot = 0, orient = 0;
    if (lane < nmodes) classify_block(I1, I2, W + 1, si, sj, side, lane, mc, vc, rot, orient);
    __syncwarp();
    float rs = 0.f, sqr = 0.f;
    if (lane == 0) {   // the reference's sequential float32 sums
        for (int t = 0; t < nn; ++t) rs = __fadd_rn(rs, s_range[warp][t]);
        for (int t = 0; t < nn; ++t) sqr = __fadd_rn(sqr, __fmul_rn(s_range[warp][t], s_range[warp][t]));
    }
    rs = __shfl_sync(0xffffffffu, rs, 0);
    sqr = __shfl_sync(0xffffffffu, sqr, 0);
    const int cls = mc * 24 + vc;
    const int cls0 = __shfl_sync(0xffffffffu, cls, 0), cls1 = __shfl_sync(0xffffffffu, cls, 1);
    const int rot0 = __shfl_sync(0xffffffffu, rot, 0), rot1 = __shfl_sync(0xffffffffu, rot, 1);
    const int ori0 = __shfl_sync(0xffffffffu, orient, 0), ori1 = __shfl_sync(0xffffffffu, orient, 1);
    const int b0 = pool.offsets[cls0], n0 = pool.offsets[cls0 + 1] - b0;
    const int b1 = nmodes > 1 ? pool.offsets[cls1] : 0, n1 = nmodes > 1 ? pool.offsets[cls1 + 1] - b1 : 0;
    const int total = n0 + n1;
    const double dn = (double)nn;
    const double slack_a = DS((double)sqr, DD(DM((double)rs, (double)rs), dn));   // sqr - rs^2 / n
    const int PW = W - 1;
    TeBest best;
    te_init(best);
    for (int c = lane; c < total; c += 32) {
        const bool second = c >= n0;
        const int32_t* row = pool.rows + 4 * (size_t)(second ? b1 + (c - n0) : b0 + c);
        const int di = row[0], dj = row[1], drot = row[2], dorient = row[3];
        const int rrot = second ? rot1 : rot0, rorient = second ? ori1 : ori0;
        int d_rot, d_orient;
        if (dorient == rorient) { d_rot = (drot - rrot) & 3; d_orient = 0; }
        else { d_rot = (drot - rrot + 2) & 3; d_orient = 1; }
        int a00, b00, a10, b10, a01, b01;      // the isometry is affine in (i, j): three corners instead of every pixel
        fic_iso_qt_enc(d_orient, d_rot, side, 0, 0, a00, b00);
        fic_iso_qt_enc(d_orient, d_rot, side, 1, 0, a10, b10);
        fic_iso_qt_enc(d_orient, d_rot, side, 0, 1, a01, b01);
        const int step_i = 2 * ((a10 - a00) * PW + (b10 - b00)), step_j = 2 * ((a01 - a00) * PW + (b01 - b00));
        const float* drow = Dp + (size_t)(di + 2 * a00) * PW + dj + 2 * b00;
        float ds = 0.f, sds = 0.f, ab = 0.f;
        int t = 0;
        for (int i = 0; i < side; ++i, drow += step_i) {
            const float* qd = drow;
            for (int j = 0; j < side; ++j, qd += step_j, ++t) {
                const float d = *qd;
                ds = __fadd_rn(ds, d);
                sds = __fadd_rn(sds, __fmul_rn(d, d));
                ab = __fadd_rn(ab, __fmul_rn(d, s_range[warp][t]));
            }
        }
        {   // exact lower bound of the score (see hv_search_kernel) against the warp's best
            const double bound = __longlong_as_double((long long)*(volatile unsigned long long*)&s_bound[warp]);
            const double room = DS(slack_a, DM(dn, bound));
            if (room > 0.0) {
                const double dsd = (double)ds;
                const double den = DS(DM(dn, (double)sds), DM(dsd, dsd));
                const double num = DS(DM(dn, (double)ab), DM((double)rs, dsd));
                if (den > 0.0 && DM(num, num) < DM(DM(DM(room, dn), den), 0.999999999)) continue;
            }
        }
        const QScore sc = fic_q_score(nn, rs, sqr, ds, sds, ab, max_scale, scale_bits, offset_bits);
        if (fic_better(sc.R, (unsigned long long)c, best.R, best.seq)) {
            best.R = sc.R; best.seq = c; best.s = sc.s; best.o = sc.o; best.qs = sc.qs; best.qo = sc.qo;
            best.di = di; best.dj = dj; best.orient = d_orient; best.rot = d_rot;
            atomicMin(&s_bound[warp], (unsigned long long)__double_as_longlong(sc.R > 0.0 ? sc.R : 0.0));
        }
    }
    __syncwarp();
    // first strict minimum in scan order over the lanes (te_block_reduce's rule)
    double R = best.R;
    unsigned long long seq = best.seq;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        const double oR = __shfl_xor_sync(0xffffffffu, R, off);
        const unsigned long long oS = __shfl_xor_sync(0xffffffffu, seq, off);
        if (fic_better(oR, oS, R, seq)) { R = oR; seq = oS; }
    }
    int32_t* rec6 = rec + 6 * (size_t)q;
    double* vals3 = vals + 3 * (size_t)q;
    if (seq == ~0ull) {
        if (lane == 0) {
            for (int k = 0; k < 6; ++k) rec6[k] = 0;
            vals3[0] = 0.0; vals3[1] = 0.0;
            vals3[2] = __longlong_as_double(0x7ff0000000000000ll);
        }
    } else if (best.seq == seq) {   // seq values are unique per range
        rec6[0] = best.di; rec6[1] = best.dj; rec6[2] = best.orient; rec6[3] = best.rot; rec6[4] = best.qs; rec6[5] = best.qo;
        vals3[0] = best.s; vals3[1] = best.o; vals3[2] = best.R;
    }
}

// ---- HV domain search (hvv2/encoder.py:97-126) -----------------------------------------------------
// 2x2-pooled image at every pixel offset: Dp[y][x] = (img[y][x] + img[y][x+1] + img[y+1][x] + img[y+1][x+1]) * 0.25,
// exact in float32 (the reference's average_subsample of a domain block reads every second entry of it).
__global__ void hv_pool_kernel(const uint8_t* __restrict__ img, int H, int W, float* __restrict__ Dp)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= W - 1 || y >= H - 1) return;
    const uint8_t* p = img + (size_t)y * W + x;
    Dp[(size_t)y * (W - 1) + x] = __fmul_rn((float)((int)p[0] + p[1] + p[W] + p[W + 1]), 0.25f);
}

// One rectangle per blockIdx.x; its candidate list (domain position x 4 isometries, scan order = seq) is cut into
// gridDim.y contiguous chunks so that the deep, narrow levels of the tree (a handful of rectangles) still fill the
// GPU.  Each CTA writes the first minimum of its chunk; hv_pick_kernel merges the chunks with the same rule.
// The three sums run in the reference's order (row-major over the oriented block, float32, no FMA).
__global__ void __launch_bounds__(TE_THREADS, HV_MINB)
hv_search_kernel(const uint8_t* __restrict__ img, const float* __restrict__ Dp, int H, int W,
                 const int32_t* __restrict__ rects /*[n][4] = i,j,w,h*/, double max_scale, int scale_bits, int offset_bits,
                 int32_t* __restrict__ rec, double* __restrict__ vals, unsigned long long* __restrict__ seqs, int smem_px,
                 const int32_t* __restrict__ slist /* optional: CTA x works on rectangle slist[x] (results stay indexed by x) */)
{
    extern __shared__ float s_range[];   // the rectangle's pixels (when they fit)
    const int rq = slist ? slist[blockIdx.x] : (int)blockIdx.x;
    __shared__ float s_rs, s_sqr;
    __shared__ unsigned long long s_bound;   // bit pattern of the best score found so far by this CTA (+inf at start)
    const int tid = threadIdx.x;
    if (tid == 0) s_bound = 0x7ff0000000000000ull;
    const int ri = rects[4 * (size_t)rq], rj = rects[4 * (size_t)rq + 1], w = rects[4 * (size_t)rq + 2],
              hgt = rects[4 * (size_t)rq + 3];
    const int nn = w * hgt;
    const bool in_smem = nn <= smem_px;
    if (in_smem)
        for (int t = tid; t < nn; t += TE_THREADS) s_range[t] = (float)img[(size_t)(ri + t / w) * W + rj + t % w];
    __syncthreads();
    if (tid == 0) {
        float rs = 0.f, sq = 0.f;
        for (int t = 0; t < nn; ++t) {
            const float v = in_smem ? s_range[t] : (float)img[(size_t)(ri + t / w) * W + rj + t % w];
            rs = __fadd_rn(rs, v);
            sq = __fadd_rn(sq, __fmul_rn(v, v));
        }
        s_rs = rs;
        s_sqr = sq;
    }
    __syncthreads();
    const float rs = s_rs, sqr = s_sqr;
    const double dn = (double)nn;
    const double slack_a = DS((double)sqr, DD(DM((double)rs, (double)rs), dn));   // sqr - rs^2 / n
    const int ndi = H - 2 * hgt > 0 ? (H - 2 * hgt + hgt - 1) / hgt : 0;  // range(0, H-2h, h): exclusive
    const int ndj = W - 2 * w > 0 ? (W - 2 * w + w - 1) / w : 0;
    const long long npos = (long long)ndi * ndj;
    const long long per = (npos + gridDim.y - 1) / gridDim.y;
    const long long p0 = per * blockIdx.y, p1 = p0 + per < npos ? p0 + per : npos;
    const int PW = W - 1;
    TeBest best;
    te_init(best);
    // One thread takes the four isometries of a domain position together: candidate index c = pos * 4 + iso with
    // iso = orient * 2 + (rot == 2); the four traversals read the block from its four corners (fic_iso_hv: rot180
    // flips rows and columns, flipud flips the rows again), which gives twelve independent float chains per thread.
    for (long long pos = p0 + tid; pos < p1; pos += TE_THREADS) {
        const int di = (int)(pos / ndj) * hgt, dj = (int)(pos % ndj) * w;
        const float* top = Dp + (size_t)di * PW + dj;                       // row 0 of the pooled block
        const float* bot = top + (size_t)(2 * (hgt - 1)) * PW;              // its last row
        const int last = 2 * (w - 1);
        float ds0 = 0.f, sq0 = 0.f, ab0 = 0.f;   // iso 0: identity               rows down, columns forward
        float ds1 = 0.f, sq1 = 0.f, ab1 = 0.f;   // iso 1: rot180                 rows up,   columns backward
        float ds2 = 0.f, sq2 = 0.f, ab2 = 0.f;   // iso 2: flipud                 rows up,   columns forward
        float ds3 = 0.f, sq3 = 0.f, ab3 = 0.f;   // iso 3: flipud after rot180    rows down, columns backward
        int t = 0;
        for (int i = 0; i < hgt; ++i) {
            const float* rd = top + (size_t)(2 * i) * PW;
            const float* ru = bot - (size_t)(2 * i) * PW;
            for (int j = 0; j < w; ++j, ++t) {
                const float r = in_smem ? s_range[t] : (float)img[(size_t)(ri + i) * W + rj + j];
                const float d0 = rd[2 * j], d1 = ru[last - 2 * j], d2 = ru[2 * j], d3 = rd[last - 2 * j];
                ds0 = __fadd_rn(ds0, d0); sq0 = __fadd_rn(sq0, __fmul_rn(d0, d0)); ab0 = __fadd_rn(ab0, __fmul_rn(d0, r));
                ds1 = __fadd_rn(ds1, d1); sq1 = __fadd_rn(sq1, __fmul_rn(d1, d1)); ab1 = __fadd_rn(ab1, __fmul_rn(d1, r));
                ds2 = __fadd_rn(ds2, d2); sq2 = __fadd_rn(sq2, __fmul_rn(d2, d2)); ab2 = __fadd_rn(ab2, __fmul_rn(d2, r));
                ds3 = __fadd_rn(ds3, d3); sq3 = __fadd_rn(sq3, __fmul_rn(d3, d3)); ab3 = __fadd_rn(ab3, __fmul_rn(d3, r));
            }
        }
        const float dsv[4] = {ds0, ds1, ds2, ds3}, sqv[4] = {sq0, sq1, sq2, sq3}, abv[4] = {ab0, ab1, ab2, ab3};
        // Lower bound before the expensive quantising scorer.  The score is the quadratic form
        //   R(s, o) = (sqr + s^2 sds - 2 s ab + 2 s o ds + n o^2 - 2 o rs) / n
        // of the five sums evaluated at the QUANTISED (s, o); for n sds - ds^2 > 0 its minimum over all real (s, o) is
        //   (sqr - rs^2/n - num^2 / (n den)) / n,   num = n ab - rs ds,  den = n sds - ds^2,
        // so a candidate with  num^2 < (sqr - rs^2/n - n B) n den  cannot score below B, the best score any thread of
        // the CTA has found so far.  The 1e-9 relative slack is six orders above the rounding of either side.
        const double bound = __longlong_as_double((long long)*(volatile unsigned long long*)&s_bound);
        const double room = DS(slack_a, DM(dn, bound));   // sqr - rs^2/n - n B
#pragma unroll
        for (int iso = 0; iso < 4; ++iso) {
            const unsigned long long c = (unsigned long long)pos * 4ull + (unsigned)iso;
            if (room > 0.0) {
                const double dsd = (double)dsv[iso];
                const double den = DS(DM(dn, (double)sqv[iso]), DM(dsd, dsd));
                const double num = DS(DM(dn, (double)abv[iso]), DM((double)rs, dsd));
                if (den > 0.0 && DM(num, num) < DM(DM(DM(room, dn), den), 0.999999999)) continue;
            }
            QScore sc = fic_q_score(nn, rs, sqr, dsv[iso], sqv[iso], abv[iso], max_scale, scale_bits, offset_bits);
            if (fic_better(sc.R, c, best.R, best.seq)) {
                best.R = sc.R; best.seq = c; best.s = sc.s; best.o = sc.o; best.qs = sc.qs; best.qo = sc.qo;
                best.di = di; best.dj = dj; best.orient = iso >> 1; best.rot = (iso & 1) * 2;
                // publish as a bound for the whole CTA (bit pattern order = numeric order for non-negative doubles)
                atomicMin(&s_bound, (unsigned long long)__double_as_longlong(sc.R > 0.0 ? sc.R : 0.0));
            }
        }
    }
    const size_t slot = (size_t)blockIdx.x * gridDim.y + blockIdx.y;
    te_block_reduce(best, rec + 6 * slot, vals + 3 * slot, seqs ? seqs + slot : nullptr);
}

// ---- HV search, rectangles of equal shape together --------------------------------------------------------------------
// The search of a rectangle costs 4 * positions * pixels ~ 4 * H * W multiply-adds whatever its shape, and the rectangles
// of a level that share a shape (lena 512x512: 78 % of them sit in groups of eight or more, profiles/hv_shape_stats.py)
// walk the SAME candidate blocks.  A CTA takes four consecutive entries of the level's search list, which
// hv_group_kernel has sorted by shape, and searches each run of equal shape in one sweep: a thread loads the four pooled
// samples of a pixel once, keeps sum(d) and sum(d^2) once per isometry, and one sum(d*r) chain per rectangle and isometry
// (the reference's order and rounding: row-major, float32, multiply then add).  Per multiply-add of sum(d*r) that is 2.75
// arithmetic instructions and 0.3 loads instead of 5 and 1.  The per-thread winners live in shared memory (they change
// rarely); everything after the sums is hv_search_kernel's.
#define HV_GROUP 4
#define HV_GROUP_MAX_PX 1024                 /* range pixels of the four rectangles in shared memory: 16 KB */
struct HvGroupBest {                         // per thread and rectangle of the group, structure of arrays in shared memory
    double R[HV_GROUP][TE_THREADS], s[HV_GROUP][TE_THREADS], o[HV_GROUP][TE_THREADS];
    unsigned long long seq[HV_GROUP][TE_THREADS];
    int qs[HV_GROUP][TE_THREADS], qo[HV_GROUP][TE_THREADS];
};

template <int G>
__device__ __forceinline__ void hv_group_body(const uint8_t* __restrict__ img, const float* __restrict__ Dp, int H, int W,
                                              const int32_t* __restrict__ rects, const int* rq, int x0, double max_scale,
                                              int scale_bits, int offset_bits, int32_t* __restrict__ rec, double* __restrict__ vals,
                                              unsigned long long* __restrict__ seqs, unsigned long long* __restrict__ gbound,
                                              const float2* __restrict__ rsums, float* s_range, HvGroupBest* sb,
                                              int g_real /* rq[g_real..G) repeat rq[0] and are ignored */)
{
    __shared__ float s_rs[HV_GROUP], s_sqr[HV_GROUP];
    __shared__ double s_slack[HV_GROUP];
    __shared__ unsigned long long s_bound[HV_GROUP];
    const int tid = threadIdx.x;
    const int w = rects[4 * (size_t)rq[0] + 2], hgt = rects[4 * (size_t)rq[0] + 3];
    const int nn = w * hgt;
    for (int t = tid; t < nn; t += TE_THREADS) {
        const int i = t / w, j = t - i * w;
#pragma unroll
        for (int k = 0; k < G; ++k)
            s_range[t * G + k] = (float)img[(size_t)(rects[4 * (size_t)rq[k]] + i) * W + rects[4 * (size_t)rq[k] + 1] + j];
    }
    // The best score any CTA has found so far for the rectangle (all-ones = none yet: a NaN, which drops nothing).  The
    // chunks of a rectangle are blockIdx.y, the slow grid dimension, so all but the first few start with a bound that
    // means something instead of scoring every candidate until they have found one of their own.
    if (tid < HV_GROUP) s_bound[tid] = tid < g_real ? __ldcg(&gbound[x0 + tid]) : ~0ull;
#pragma unroll
    for (int k = 0; k < G; ++k) { sb->R[k][tid] = __longlong_as_double(0x7ff0000000000000ll); sb->seq[k][tid] = ~0ull; }
    __syncthreads();
    if (tid < G) {   // the rectangle's two sums: sequential float32 chains, computed once per level by hv_group_kernel
        const float2 rr = rsums[x0 + (tid < g_real ? tid : 0)];
        s_rs[tid] = rr.x;
        s_sqr[tid] = rr.y;
        s_slack[tid] = DS((double)rr.y, DD(DM((double)rr.x, (double)rr.x), (double)nn));   // sqr - rs^2 / n
    }
    __syncthreads();
    const double dn = (double)nn;
    const int ndi = H - 2 * hgt > 0 ? (H - 2 * hgt + hgt - 1) / hgt : 0;  // range(0, H-2h, h): exclusive
    const int ndj = W - 2 * w > 0 ? (W - 2 * w + w - 1) / w : 0;
    const int npos = ndi * ndj;                                           // (at most H * W / 4 < 2^31, hv_encode_impl)
    const int per = (npos + (int)gridDim.y - 1) / (int)gridDim.y;
    const int p0 = per * (int)blockIdx.y, p1 = p0 + per < npos ? p0 + per : npos;
    const int PW = W - 1;
    for (int pos = p0 + tid; pos < p1; pos += TE_THREADS) {
        const int pi = pos / ndj;
        const int di = pi * hgt, dj = (pos - pi * ndj) * w;
        const float* top = Dp + (size_t)di * PW + dj;                       // row 0 of the pooled block
        const float* bot = top + (size_t)(2 * (hgt - 1)) * PW;              // its last row
        const int last = 2 * (w - 1);
        float ds[4] = {0.f, 0.f, 0.f, 0.f}, sq[4] = {0.f, 0.f, 0.f, 0.f};    // per isometry (see hv_search_kernel)
        float ab[G][4];
#pragma unroll
        for (int k = 0; k < G; ++k)
#pragma unroll
            for (int e = 0; e < 4; ++e) ab[k][e] = 0.f;
        const float* rp = s_range;
        for (int i = 0; i < hgt; ++i) {
            const float* rd = top + (size_t)(2 * i) * PW;
            const float* ru = bot - (size_t)(2 * i) * PW;
            for (int j = 0; j < w; ++j, rp += G) {
                const float d[4] = {rd[2 * j], ru[last - 2 * j], ru[2 * j], rd[last - 2 * j]};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    ds[e] = __fadd_rn(ds[e], d[e]);
                    sq[e] = __fadd_rn(sq[e], __fmul_rn(d[e], d[e]));
                }
#pragma unroll
                for (int k = 0; k < G; ++k) {
                    const float r = rp[k];
#pragma unroll
                    for (int e = 0; e < 4; ++e) ab[k][e] = __fadd_rn(ab[k][e], __fmul_rn(d[e], r));
                }
            }
        }
        // Lower bound of the score first (see hv_search_kernel), for all G x 4 candidates of this position; the few that pass
        // go through ONE copy of the quantising scorer (sixteen inlined copies would not fit the instruction cache).
        double nden[4], dsd[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            dsd[e] = (double)ds[e];
            const double den = DS(DM(dn, (double)sq[e]), DM(dsd[e], dsd[e]));
            nden[e] = den > 0.0 ? DM(DM(dn, den), 0.999999999) : -1.0;     // -1: never dropped
        }
        unsigned pass = 0u;
#pragma unroll
        for (int k = 0; k < G; ++k) {
            const double rsd = (double)s_rs[k];
            const double bound = __longlong_as_double((long long)*(volatile unsigned long long*)&s_bound[k]);
            const double room = DS(s_slack[k], DM(dn, bound));   // sqr - rs^2/n - n B
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const double num = DS(DM(dn, (double)ab[k][e]), DM(rsd, dsd[e]));
                const bool drop = room > 0.0 && nden[e] > 0.0 && DM(num, num) < DM(room, nden[e]);
                if (!drop && k < g_real) pass |= 1u << (k * 4 + e);
            }
        }
        while (pass) {
            const int idx = __ffs((int)pass) - 1;
            pass &= pass - 1u;
            const int k = idx >> 2, iso = idx & 3;
            const float dsv = iso == 0 ? ds[0] : iso == 1 ? ds[1] : iso == 2 ? ds[2] : ds[3];
            const float sqv = iso == 0 ? sq[0] : iso == 1 ? sq[1] : iso == 2 ? sq[2] : sq[3];
            float abv = ab[0][0];
#pragma unroll
            for (int kk = 0; kk < G; ++kk)
#pragma unroll
                for (int e = 0; e < 4; ++e)
                    if (idx == kk * 4 + e) abv = ab[kk][e];
            const unsigned long long c = (unsigned long long)pos * 4ull + (unsigned)iso;
            const QScore sc = fic_q_score(nn, s_rs[k], s_sqr[k], dsv, sqv, abv, max_scale, scale_bits, offset_bits);
            if (fic_better(sc.R, c, sb->R[k][tid], sb->seq[k][tid])) {
                sb->R[k][tid] = sc.R; sb->seq[k][tid] = c; sb->s[k][tid] = sc.s; sb->o[k][tid] = sc.o;
                sb->qs[k][tid] = sc.qs; sb->qo[k][tid] = sc.qo;
                const unsigned long long bits = (unsigned long long)__double_as_longlong(sc.R > 0.0 ? sc.R : 0.0);
                if (bits < atomicMin(&s_bound[k], bits)) atomicMin(&gbound[x0 + k], bits);
            }
        }
    }
#pragma unroll
    for (int k = 0; k < G; ++k) {
        if (k >= g_real) break;
        TeBest b;
        b.R = sb->R[k][tid]; b.seq = sb->seq[k][tid];
        b.s = b.o = 0.0; b.di = b.dj = b.orient = b.rot = b.qs = b.qo = 0;
        if (b.seq != ~0ull) {
            b.s = sb->s[k][tid]; b.o = sb->o[k][tid]; b.qs = sb->qs[k][tid]; b.qo = sb->qo[k][tid];
            const int pos = (int)(b.seq >> 2);
            const int iso = (int)(b.seq & 3ull);
            b.di = (pos / ndj) * hgt; b.dj = (pos % ndj) * w; b.orient = iso >> 1; b.rot = (iso & 1) * 2;
        }
        const size_t slot = (size_t)(x0 + k) * gridDim.y + blockIdx.y;
        __syncthreads();      // (te_block_reduce's shared scratch is reused from call to call)
        te_block_reduce(b, rec + 6 * slot, vals + 3 * slot, seqs + slot);
    }
    __syncthreads();
}

__global__ void __launch_bounds__(TE_THREADS, HV_MINB)
hv_search_grouped_kernel(const uint8_t* __restrict__ img, const float* __restrict__ Dp, int H, int W,
                         const int32_t* __restrict__ rects /*[n][4] = i,j,w,h*/, double max_scale, int scale_bits, int offset_bits,
                         int32_t* __restrict__ rec, double* __restrict__ vals, unsigned long long* __restrict__ seqs,
                         const int32_t* __restrict__ slist /*sorted by shape*/, int count, unsigned long long* __restrict__ gbound,
                         const float2* __restrict__ rsums /*[count] = (sum r, sum r^2) of the entries of slist*/)
{
    extern __shared__ __align__(16) unsigned char s_dyn[];
    float* s_range = reinterpret_cast<float*>(s_dyn);
    HvGroupBest* sb = reinterpret_cast<HvGroupBest*>(s_dyn + HV_GROUP * HV_GROUP_MAX_PX * sizeof(float));
    const int x0 = blockIdx.x * HV_GROUP;
    const int m = count - x0 < HV_GROUP ? count - x0 : HV_GROUP;
    int rq[HV_GROUP];
#pragma unroll
    for (int k = 0; k < HV_GROUP; ++k) rq[k] = slist[x0 + (k < m ? k : 0)];
    int a = 0;
    while (a < m) {
        int b = a + 1;
        while (b < m && rects[4 * (size_t)rq[b] + 2] == rects[4 * (size_t)rq[a] + 2] && rects[4 * (size_t)rq[b] + 3] == rects[4 * (size_t)rq[a] + 3]) ++b;
        int run[HV_GROUP];
#pragma unroll
        for (int k = 0; k < HV_GROUP; ++k) run[k] = rq[k < b - a ? a + k : a];
        // Two bodies only -- a single rectangle, or up to four with the unused slots repeating the first: different CTAs of an
        // SM run different bodies at the same time and the kernel's code has to stay in the instruction cache (with four
        // bodies a third of the stall samples were waits for instructions).
        if (b - a == 1) hv_group_body<1>(img, Dp, H, W, rects, run, x0 + a, max_scale, scale_bits, offset_bits, rec, vals, seqs, gbound, rsums, s_range, sb, 1);
        else hv_group_body<HV_GROUP>(img, Dp, H, W, rects, run, x0 + a, max_scale, scale_bits, offset_bits, rec, vals, seqs, gbound, rsums, s_range, sb, b - a);
        a = b;
    }
}

// Sorts the search list of a level by shape (counting sort, one CTA): slist_sorted[k] = k-th rectangle, sidx[rect] = k.
// The order inside a shape is not specified (atomic counters); every rectangle's result is independent of it.
#define HV_SORT_BINS 4096
__global__ void __launch_bounds__(1024)
hv_group_kernel(const int32_t* __restrict__ rects, const int32_t* __restrict__ slist, const int* __restrict__ count_ptr,
                int32_t* __restrict__ slist_sorted, int32_t* __restrict__ sidx, unsigned long long* __restrict__ gbound,
                const uint8_t* __restrict__ img, int W, float2* __restrict__ rsums)
{
    __shared__ int hist[HV_SORT_BINS];
    __shared__ int warp_sum[32];
    const int tid = threadIdx.x, n = *count_ptr;
    for (int b = tid; b < HV_SORT_BINS; b += 1024) hist[b] = 0;
    __syncthreads();
    auto key = [&](int r) {
        const int w = rects[4 * (size_t)r + 2], hgt = rects[4 * (size_t)r + 3];
        return (w < 63 ? w : 63) * 64 + (hgt < 63 ? hgt : 63);
    };
    for (int e = tid; e < n; e += 1024) atomicAdd(&hist[key(slist[e])], 1);
    __syncthreads();
    // exclusive scan of the 4096 counters: four per thread, warp scan, scan of the warp totals
    int v[4], sum = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) { v[k] = hist[tid * 4 + k]; sum += v[k]; }
    int incl = sum;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const int o = __shfl_up_sync(0xffffffffu, incl, off);
        if ((tid & 31) >= off) incl += o;
    }
    if ((tid & 31) == 31) warp_sum[tid >> 5] = incl;
    __syncthreads();
    if (tid < 32) {
        int ws = warp_sum[tid], wi = ws;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int o = __shfl_up_sync(0xffffffffu, wi, off);
            if (tid >= off) wi += o;
        }
        warp_sum[tid] = wi - ws;
    }
    __syncthreads();
    int base = warp_sum[tid >> 5] + incl - sum;
#pragma unroll
    for (int k = 0; k < 4; ++k) { hist[tid * 4 + k] = base; base += v[k]; }
    __syncthreads();
    for (int e = tid; e < n; e += 1024) {
        const int r = slist[e];
        const int at = atomicAdd(&hist[key(r)], 1);
        slist_sorted[at] = r;
        sidx[r] = at;
        gbound[e] = ~0ull;     // no score yet for the e-th rectangle of the search list
    }
    __syncthreads();
    // sum r and sum r^2 of every rectangle, in the reference's order (row-major, float32): two serial chains per rectangle,
    // done here once instead of by one thread of every chunk CTA with the other 127 waiting at the barrier
    for (int e = tid; e < n; e += 1024) {
        const int r = slist_sorted[e];
        const int ri = rects[4 * (size_t)r], rj = rects[4 * (size_t)r + 1], w = rects[4 * (size_t)r + 2], hgt = rects[4 * (size_t)r + 3];
        float rs = 0.f, sq = 0.f;
        for (int i = 0; i < hgt; ++i)
            for (int j = 0; j < w; ++j) {
                const float v = (float)img[(size_t)(ri + i) * W + rj + j];
                rs = __fadd_rn(rs, v);
                sq = __fadd_rn(sq, __fmul_rn(v, v));
            }
        rsums[e] = make_float2(rs, sq);
    }
}

// merge the chunk winners of every rectangle: first strict minimum in scan order
__global__ void hv_pick_kernel(const int32_t* __restrict__ prec, const double* __restrict__ pvals,
                               const unsigned long long* __restrict__ pseq, int count, int S, int32_t* __restrict__ rec,
                               double* __restrict__ vals)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= count) return;
    int bs = 0;
    double bR = pvals[3 * ((size_t)k * S) + 2];
    unsigned long long bq = pseq[(size_t)k * S];
    for (int s = 1; s < S; ++s) {
        const double R = pvals[3 * ((size_t)k * S + s) + 2];
        const unsigned long long q = pseq[(size_t)k * S + s];
        if (fic_better(R, q, bR, bq)) { bR = R; bq = q; bs = s; }
    }
    const size_t src = (size_t)k * S + bs;
    for (int t = 0; t < 6; ++t) rec[6 * (size_t)k + t] = prec[6 * src + t];
    for (int t = 0; t < 3; ++t) vals[3 * (size_t)k + t] = pvals[3 * src + t];
}

// launches the search for `count` rectangles already on the device; results in d_rec / d_vals
static int hv_search_launch(fic_handle* h, const uint8_t* d_img, int H, int W, const int32_t* d_rects, int count, int max_px,
                            int scale_bits, int offset_bits, double max_scale, int32_t* d_rec, double* d_vals, cudaStream_t st)
{
    if (count == 0) return FIC_OK;
    const float* Dp = h->t_hvpool.as<float>();
    int smem_px = max_px <= 10240 ? max_px : 0;      // range pixels in shared memory when they fit (40 KB)
    int S = (HV_SMULT * h->num_sms + count - 1) / count;      // see HV_SMULT
    if (S > HV_SCAP) S = HV_SCAP;
    if (S < 1) S = 1;
    if (S == 1) {
        hv_search_kernel<<<dim3(count, 1), TE_THREADS, smem_px * sizeof(float), st>>>(d_img, Dp, H, W, d_rects, max_scale, scale_bits,
                                                                                     offset_bits, d_rec, d_vals, nullptr, smem_px, nullptr);
        FIC_CHECK_LAUNCH(h, "hv_search_kernel");
        return FIC_OK;
    }
    const size_t np_ = (size_t)count * S;
    FIC_TRY(h->t_part.reserve(np_ * (6 * sizeof(int32_t) + 3 * sizeof(double) + sizeof(unsigned long long)) + 64));
    double* pvals = h->t_part.as<double>();
    unsigned long long* pseq = reinterpret_cast<unsigned long long*>(pvals + 3 * np_);
    int32_t* prec = reinterpret_cast<int32_t*>(pseq + np_);
    hv_search_kernel<<<dim3(count, S), TE_THREADS, smem_px * sizeof(float), st>>>(d_img, Dp, H, W, d_rects, max_scale, scale_bits,
                                                                                 offset_bits, prec, pvals, pseq, smem_px, nullptr);
    FIC_CHECK_LAUNCH(h, "hv_search_kernel");
    hv_pick_kernel<<<(count + 127) / 128, 128, 0, st>>>(prec, pvals, pseq, count, S, d_rec, d_vals);
    FIC_CHECK_LAUNCH(h, "hv_pick_kernel");
    return FIC_OK;
}

static int hv_build_pool(fic_handle* h, const uint8_t* d_img, int H, int W, cudaStream_t st)
{
    FIC_TRY(h->t_hvpool.reserve((size_t)(H - 1) * (W - 1) * sizeof(float) + 64));
    hv_pool_kernel<<<dim3((W - 1 + 255) / 256, H - 1), 256, 0, st>>>(d_img, H, W, h->t_hvpool.as<float>());
    FIC_CHECK_LAUNCH(h, "hv_pool_kernel");
    return FIC_OK;
}

int hv_search_impl(fic_handle* h, const uint8_t* d_img, int H, int W, const int32_t* d_rects, int count, int scale_bits,
                   int offset_bits, double max_scale, int32_t* d_rec, double* d_vals, cudaStream_t st)
{
    if (count == 0) return FIC_OK;
    FIC_TRY(hv_build_pool(h, d_img, H, W, st));
    // the rectangles may live on the device only: 4096 pixels of shared memory cover every rectangle up to 64x64, larger
    // ones read the image directly
    return hv_search_launch(h, d_img, H, W, d_rects, count, 4096, scale_bits, offset_bits, max_scale, d_rec, d_vals, st);
}

// ---- host drivers -----------------------------------------------------------------------------------
static int ilog2_floor(int v)
{
    int e = 0;
    while ((1 << (e + 1)) <= v) ++e;
    return e;
}

// quadtree/encoder.py:81-95
static int qt_smallest_range(int width, int height)
{
    if (width == 0) return height;
    if (height == 0) return width;
    if (width == 1 || height == 1) return 1;
    int ms = 1 << ilog2_floor(std::min(width, height));
    if (ms == width && ms == height) return ms;
    int fx = width / ms, fy = height / ms;
    int min_right = qt_smallest_range(width - fx * ms, height);
    int min_bottom = qt_smallest_range(fx * ms, height - fy * ms);
    return std::min(min_bottom, min_right);
}

// classification of the queued ranges for positive and negative scales: out[k][mode] = {class, rotation, orientation}
__global__ void qt_range_class_kernel(const long long* __restrict__ I1, const long long* __restrict__ I2, int W1,
                                      const int32_t* __restrict__ queue, int n, int side, int32_t* __restrict__ out)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= 2 * n) return;
    const int q = k >> 1, mode = k & 1;
    int mc, vc, rot, orient;
    classify_block(I1, I2, W1, queue[4 * (size_t)q], queue[4 * (size_t)q + 1], side, mode, mc, vc, rot, orient);
    out[3 * (size_t)k] = mc * 24 + vc;
    out[3 * (size_t)k + 1] = rot;
    out[3 * (size_t)k + 2] = orient;
}

// bf_screen.cu / bf_exact.cu: the dense all-classes search on the tensor-core screen
bool qt_tc_applicable(int H, int W, int e, double max_scale, int hs, int vs);
int qt_tc_search(fic_handle* h, const uint8_t* d_img, int H, int W, int e, const int32_t* d_queue, int n_ranges,
                 const int32_t* d_rcls, const int32_t* d_dcls, int nmodes, double max_scale, int scale_bits, int offset_bits,
                 int32_t* d_rec, double* d_vals, int32_t* d_need_exact, int hs, int vs, const int32_t* d_pool_offsets,
                 const int32_t* d_pool_rows, cudaStream_t st);
int bf_prepare8(fic_handle* h, const uint8_t* d_img, int H, int W, bool with_stats, cudaStream_t st);

// ---- quadtree partition on the device (quadtree/encoder.py:127-146) ----------------------------------------------
// Node table and one work queue per size exponent live in device memory.  A level = all pending ranges of one size:
// one search launch, then qt_decide_kernel applies `rms > tolerance and exponent > min_exponent` (encoder.py:134),
// stores the leaf records or appends the four children (TL, TR, BL, BR) to the node table and to the queue of the next
// smaller size.  The host reads back one integer per level (how many ranges the next launch has) and, at the end, the
// node table, from which it emits the preorder stream.
struct QtNodeDev {
    int32_t si, sj, exp, kind;    // kind: -1 pending, 0 internal, 1 leaf
    int32_t child0, pad_;         // the four children are consecutive nodes
    int32_t rec[6];
    double vals[3];
};
static_assert(sizeof(QtNodeDev) == 72, "QtNodeDev layout");

// mode 0: decide from the search results; 1: split unconditionally (sizes above max_range_side); 2: side-1 dummy leaves
__global__ void qt_decide_kernel(const int32_t* __restrict__ queue /*[n][4]*/, int n, int e, int mode, double tol, int min_exp,
                                 const int32_t* __restrict__ rec, const double* __restrict__ vals, QtNodeDev* __restrict__ nodes,
                                 int* __restrict__ counts /*[0] = nodes, [1 + e] = queue length of exponent e*/,
                                 int32_t* __restrict__ next_queue, int node_cap, int next_cap)
{
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    const int si = queue[4 * (size_t)q], sj = queue[4 * (size_t)q + 1], id = queue[4 * (size_t)q + 2];
    QtNodeDev& nd = nodes[id];
    bool split = mode == 1;
    if (mode == 0) {
        const double rms = sqrt(vals[3 * (size_t)q + 2]);   // NaN for a negative error -> comparison false -> leaf
        split = rms > tol && e > min_exp;
    }
    if (!split) {
        nd.kind = 1;
        if (mode == 2) {   // side 1: dummy leaf with rms 1e-4 (encoder.py:150-152)
            for (int k = 0; k < 6; ++k) nd.rec[k] = 0;
            nd.vals[0] = 0.0; nd.vals[1] = 0.0; nd.vals[2] = 1e-8;
        } else {
            for (int k = 0; k < 6; ++k) nd.rec[k] = rec[6 * (size_t)q + k];
            for (int k = 0; k < 3; ++k) nd.vals[k] = vals[3 * (size_t)q + k];
        }
        return;
    }
    const int base = atomicAdd(&counts[0], 4);
    const int at = atomicAdd(&counts[1 + (e - 1)], 4);
    if (base + 4 > node_cap || at + 4 > next_cap) { counts[40] = 1; return; }   // cannot happen with the host's bounds; flagged
    nd.kind = 0;
    nd.child0 = base;
    const int ns = 1 << (e - 1);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        QtNodeDev ch;
        ch.si = si + (c >> 1) * ns; ch.sj = sj + (c & 1) * ns; ch.exp = e - 1; ch.kind = -1; ch.child0 = -1; ch.pad_ = 0;
        for (int k = 0; k < 6; ++k) ch.rec[k] = 0;
        ch.vals[0] = ch.vals[1] = ch.vals[2] = 0.0;
        nodes[base + c] = ch;
        int32_t* qe = next_queue + 4 * (size_t)(at + c);
        qe[0] = ch.si; qe[1] = ch.sj; qe[2] = base + c; qe[3] = 0;
    }
}

// Sizes above max_range_side are split without a search (encoder.py:131-133), so the tree down to that size is known
// in advance: one launch builds it for every root.  A root of exponent E with D = E - max_exp forced levels owns
// (4^(D+1) - 4) / 3 descendants, numbered level by level behind `sub_base`: node m of level d (its path in base 4, first
// split in the top digit) sits at sub_base + (4^d - 4)/3 + m and its children at level d + 1, 4m .. 4m + 3 (TL, TR, BL, BR).
// The nodes of the last level go to the queue of exponent max_exp at q_base + m.
struct QtForcedRoot { int32_t node, si, sj, exp, depth, sub_base, q_base, first_thread; };

__global__ void qt_forced_kernel(const QtForcedRoot* __restrict__ roots, int n_roots, int total_threads, QtNodeDev* __restrict__ nodes,
                                 int32_t* __restrict__ queue)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= total_threads) return;
    int r = 0;
    while (r + 1 < n_roots && roots[r + 1].first_thread <= t) ++r;     // a handful of roots
    const QtForcedRoot R = roots[r];
    int local = t - R.first_thread;     // index among the root's descendants, level order
    int d = 1, level_start = 0, level_size = 4;
    while (local >= level_start + level_size) { level_start += level_size; level_size *= 4; ++d; }
    const int m = local - level_start;
    int si = R.si, sj = R.sj;
    for (int k = d - 1; k >= 0; --k) {          // digits of m, first split first
        const int c = (m >> (2 * k)) & 3;
        const int half = 1 << (R.exp - (d - k));
        si += (c >> 1) * half;
        sj += (c & 1) * half;
    }
    QtNodeDev nd;
    nd.si = si; nd.sj = sj; nd.exp = R.exp - d; nd.pad_ = 0;
    for (int k = 0; k < 6; ++k) nd.rec[k] = 0;
    nd.vals[0] = nd.vals[1] = nd.vals[2] = 0.0;
    if (d < R.depth) {
        nd.kind = 0;
        nd.child0 = R.sub_base + level_start + level_size + 4 * m;
    } else {
        nd.kind = -1;
        nd.child0 = -1;
        int32_t* qe = queue + 4 * (size_t)(R.q_base + m);
        qe[0] = si; qe[1] = sj; qe[2] = R.sub_base + local; qe[3] = 0;
    }
    nodes[R.sub_base + local] = nd;
}

// quadtree/encoder.py:111-125: roots in _partition_image order
static void qt_roots(std::vector<QtNodeDev>& nodes, int si, int sj, int width, int height)
{
    int e = ilog2_floor(std::min(width, height));
    int side = 1 << e;
    QtNodeDev nd{};
    nd.si = si; nd.sj = sj; nd.exp = e; nd.kind = -1; nd.child0 = -1;
    nodes.push_back(nd);
    if (side < width) qt_roots(nodes, si, sj + side, width - side, height);
    if (side < height) qt_roots(nodes, si + side, sj, side, height - side);
}

// The preorder stream of include/fic.h from the node table, on the device (one CTA; the host used to walk the table with a
// stack, a fifth of a default 512x512 encode).  A node's children are four consecutive nodes one size exponent below it, so
// the nodes are first bucketed by exponent (counting sort), subtree sizes are added up from the smallest exponent to the
// largest, preorder positions handed down from the largest to the smallest -- roots one after the other, a child behind
// its parent and its elder siblings' subtrees -- and every node writes its record at its position.
#define QT_EMIT_EXPS 32
#define QT_EMIT_ILP 4          /* nodes a thread works on at once: the kernel is one CTA and nothing but dependent loads */
__global__ void __launch_bounds__(1024)
qt_emit_kernel(const QtNodeDev* __restrict__ nodes, int n_nodes, int n_root, int32_t* __restrict__ order /*[n]*/,
               int32_t* __restrict__ sz /*[n]*/, int32_t* __restrict__ pre /*[n]*/, int32_t* __restrict__ out_nodes /*[n][8]*/,
               double* __restrict__ out_vals /*[n][3]*/)
{
    __shared__ int cnt[QT_EMIT_EXPS], first[QT_EMIT_EXPS + 1], fill[QT_EMIT_EXPS];
    constexpr int U = QT_EMIT_ILP, STEP = 1024 * QT_EMIT_ILP;
    const int tid = threadIdx.x;
    if (tid < QT_EMIT_EXPS) cnt[tid] = 0;
    __syncthreads();
    for (int base = tid; base < n_nodes; base += STEP) {
        int e[U];
#pragma unroll
        for (int u = 0; u < U; ++u) e[u] = base + u * 1024 < n_nodes ? nodes[base + u * 1024].exp & (QT_EMIT_EXPS - 1) : -1;
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (e[u] >= 0) atomicAdd(&cnt[e[u]], 1);
    }
    __syncthreads();
    if (tid == 0) {
        int acc = 0;
        for (int e = 0; e < QT_EMIT_EXPS; ++e) { first[e] = acc; fill[e] = acc; acc += cnt[e]; }
        first[QT_EMIT_EXPS] = acc;
    }
    __syncthreads();
    for (int base = tid; base < n_nodes; base += STEP) {
        int e[U];
#pragma unroll
        for (int u = 0; u < U; ++u) e[u] = base + u * 1024 < n_nodes ? nodes[base + u * 1024].exp & (QT_EMIT_EXPS - 1) : -1;
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (e[u] >= 0) order[atomicAdd(&fill[e[u]], 1)] = base + u * 1024;
    }
    __syncthreads();
    for (int e = 0; e < QT_EMIT_EXPS; ++e) {          // subtree sizes, small blocks first
        const int k1 = first[e + 1];
        if (first[e] == k1) continue;                  // (uniform: no barrier needed for an empty exponent)
        for (int base = first[e] + tid; base < k1; base += STEP) {
            int id[U], kind[U], c0[U], s[U];
#pragma unroll
            for (int u = 0; u < U; ++u) id[u] = base + u * 1024 < k1 ? order[base + u * 1024] : -1;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                kind[u] = id[u] >= 0 ? nodes[id[u]].kind : 1;
                c0[u] = id[u] >= 0 ? nodes[id[u]].child0 : 0;
            }
#pragma unroll
            for (int u = 0; u < U; ++u) s[u] = kind[u] == 0 ? 1 + sz[c0[u]] + sz[c0[u] + 1] + sz[c0[u] + 2] + sz[c0[u] + 3] : 1;
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (id[u] >= 0) sz[id[u]] = s[u];
        }
        __syncthreads();
    }
    if (tid == 0) {
        int acc = 0;
        for (int r = 0; r < n_root; ++r) { pre[r] = acc; acc += sz[r]; }
    }
    __syncthreads();
    for (int e = QT_EMIT_EXPS - 1; e >= 0; --e) {     // positions, large blocks first
        const int k1 = first[e + 1];
        if (first[e] == k1) continue;
        for (int base = first[e] + tid; base < k1; base += STEP) {
            int id[U], at[U];
#pragma unroll
            for (int u = 0; u < U; ++u) id[u] = base + u * 1024 < k1 ? order[base + u * 1024] : -1;
#pragma unroll
            for (int u = 0; u < U; ++u) at[u] = id[u] >= 0 ? pre[id[u]] : 0;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (id[u] < 0) continue;
                const QtNodeDev nd = nodes[id[u]];
                int32_t* o = out_nodes + 8 * (size_t)at[u];
                double* v = out_vals + 3 * (size_t)at[u];
                const bool leaf = nd.kind == 1;
                reinterpret_cast<int4*>(o)[0] = make_int4(nd.kind, leaf ? nd.rec[0] : 0, leaf ? nd.rec[1] : 0, leaf ? nd.rec[2] : 0);
                reinterpret_cast<int4*>(o)[1] = make_int4(leaf ? nd.rec[3] : 0, leaf ? nd.rec[4] : 0, leaf ? nd.rec[5] : 0, nd.exp);
                v[0] = leaf ? nd.vals[0] : 0.0; v[1] = leaf ? nd.vals[1] : 0.0; v[2] = leaf ? nd.vals[2] : 0.0;
                if (nd.kind == 0) {
                    const int s0 = sz[nd.child0], s1 = sz[nd.child0 + 1], s2 = sz[nd.child0 + 2];
                    pre[nd.child0] = at[u] + 1;
                    pre[nd.child0 + 1] = at[u] + 1 + s0;
                    pre[nd.child0 + 2] = at[u] + 1 + s0 + s1;
                    pre[nd.child0 + 3] = at[u] + 1 + s0 + s1 + s2;
                }
            }
        }
        __syncthreads();
    }
}

int qt_encode_impl(fic_handle* h, const uint8_t* d_img, int H, int W, int scale_bits, int offset_bits, double max_scale,
                   int full, int onlypos, int min_side, int max_side, double tol, int hstep_in, int vstep_in,
                   int32_t* out_nodes, double* out_vals, int cap, int32_t* n_roots, cudaStream_t st)
{
    const int min_exp = ilog2_floor(min_side), max_exp = ilog2_floor(max_side);  // int(np.log2(.)), encoder.py:56-57
    int e_small = ilog2_floor(qt_smallest_range(W, H));
    if (e_small < 1) e_small = 1;
    const int min_needed = std::min(min_exp, e_small);  // encoder.py:73-74
    {
        // the reference raises inside numba as soon as it searches a range whose 2x domain does not fit
        int mn = std::min(H, W);
        int e_top = std::min(ilog2_floor(mn), max_exp);
        FIC_REQUIRE(!(e_top >= 1 && (2 << e_top) > mn),
                    "fic_qt_encode: a %dx%d range needs a %dx%d domain, which does not fit the %dx%d image "
                    "(the reference fails with 'unable to broadcast')", 1 << e_top, 1 << e_top, 2 << e_top, 2 << e_top, H, W);
    }
    FIC_TRY(build_integrals(h, d_img, H, W, st));
    FIC_TRY(hv_build_pool(h, d_img, H, W, st));   // 2x2 means at every pixel offset: one load per pooled sample in the search
    const long long* I1 = h->t_int1.as<long long>();
    const long long* I2 = h->t_int2.as<long long>();

    // ---- domain pools (encoder.py:97-109, 261-283), built on the device: [offsets(73) | rows] per pool size ----
    const int npools = std::max(0, max_exp + 1 - (min_needed + 1) + 1);
    std::vector<size_t> pool_base(npools);
    size_t tot = 0, max_cnt = 0, cls_tot = 0;
    struct PoolGeom { int side, hs, vs, nv, nh; size_t cnt, cls_off; };
    std::vector<PoolGeom> geom(npools);
    for (int k = 0; k < npools; ++k) {
        PoolGeom& g = geom[k];
        g.side = 1 << (min_needed + 1 + k);
        g.hs = hstep_in == -1 ? g.side : hstep_in;
        g.vs = vstep_in == -1 ? g.side : vstep_in;
        g.nv = H - g.side >= 0 ? 1 + (H - g.side) / g.vs : 0;
        g.nh = W - g.side >= 0 ? 1 + (W - g.side) / g.hs : 0;
        g.cnt = (size_t)g.nv * g.nh;
        FIC_REQUIRE(g.cnt < ((size_t)1 << 30), "fic_qt_encode: %zu domains in one pool", g.cnt);
        pool_base[k] = tot;
        tot += 73 + 4 * (g.cnt + 72);          // every class may need its placeholder row
        tot = (tot + 3) & ~(size_t)3;
        max_cnt = std::max(max_cnt, g.cnt);
        g.cls_off = cls_tot;                   // the classification of every pool is kept (the tensor-core search reads it)
        cls_tot += (g.cnt * 3 + 3) & ~(size_t)3;
    }
    FIC_TRY(h->t_lists.reserve((tot + 4) * sizeof(int32_t)));
    if (max_cnt) {
        const size_t max_blocks = (max_cnt + QP_CHUNK - 1) / QP_CHUNK;
        FIC_TRY(h->t_cls.reserve(cls_tot * sizeof(int32_t) + max_blocks * 72 * sizeof(int32_t) + 64));
    }
    for (int k = 0; k < npools; ++k) {
        const PoolGeom& g = geom[k];
        int32_t* offsets = h->t_lists.as<int32_t>() + pool_base[k];
        int32_t* rows = offsets + 73;
        int32_t* cls3 = h->t_cls.as<int32_t>() + g.cls_off;
        const int nblocks = (int)((g.cnt + QP_CHUNK - 1) / QP_CHUNK);
        int32_t* blockhist = h->t_cls.as<int32_t>() + cls_tot;
        if (g.cnt) {
            classify_grid_kernel<<<(unsigned)((g.cnt + 127) / 128), 128, 0, st>>>(I1, I2, W + 1, g.side, g.nv, g.nh, g.vs, g.hs, cls3);
            FIC_CHECK_LAUNCH(h, "classify_grid_kernel");
            qt_pool_count_kernel<<<nblocks, QP_THREADS, 0, st>>>(cls3, (int)g.cnt, blockhist);
            FIC_CHECK_LAUNCH(h, "qt_pool_count_kernel");
        }
        qt_pool_scan_kernel<<<1, 96, 0, st>>>(blockhist, nblocks, offsets, rows);
        FIC_CHECK_LAUNCH(h, "qt_pool_scan_kernel");
        if (g.cnt) {
            qt_pool_scatter_kernel<<<nblocks, QP_THREADS, 0, st>>>(cls3, (int)g.cnt, g.nh, g.vs, g.hs, blockhist, offsets, rows);
            FIC_CHECK_LAUNCH(h, "qt_pool_scatter_kernel");
        }
    }

    // ---- level-synchronous partition (encoder.py:127-146), decisions on the device ----
    std::vector<QtNodeDev> nodes;
    qt_roots(nodes, 0, 0, W, H);
    const int n_root = (int)nodes.size();
    int top = 0;
    for (const QtNodeDev& r : nodes) top = std::max(top, r.exp);
    FIC_REQUIRE(top < 32, "fic_qt_encode: image too large");
    // capacity: a node of exponent e exists at most once per aligned 2^e cell of its root; ragged images add the roots
    std::vector<size_t> qcap(top + 1), qoff(top + 2, 0);
    size_t node_cap = n_root;
    const int low = std::max(0, std::min(min_exp, max_exp));   // splits stop here (forced ones at max_exp, the others at min_exp)
    for (int e = top; e >= 0; --e) {
        // roots of a ragged image can have any exponent down to 0; split children exist from `low` upwards
        qcap[e] = (e >= low ? (size_t)((H >> e) + 1) * ((W >> e) + 1) : 0) + n_root + 4;
        node_cap += e < top && e >= low ? qcap[e] : 0;
    }
    for (int e = 0; e <= top; ++e) qoff[e + 1] = qoff[e] + qcap[e] * 4;
    FIC_REQUIRE(node_cap < ((size_t)1 << 30), "fic_qt_encode: image too large");
    FIC_TRY(h->t_work.reserve(qoff[top + 1] * sizeof(int32_t) + 64));
    FIC_TRY(h->t_nodes.reserve(node_cap * sizeof(QtNodeDev) + 64));
    FIC_TRY(h->t_count.reserve(64 * sizeof(int)));
    FIC_TRY(h->pin_a.reserve(256 + 64 * sizeof(int)));
    int32_t* d_queue = h->t_work.as<int32_t>();
    QtNodeDev* d_nodes = h->t_nodes.as<QtNodeDev>();
    int* d_counts = h->t_count.as<int>();
    int* h_counts = h->pin_a.as<int>() + 64;     // pinned: [0] nodes, [1 + e] queue lengths, [40] overflow flag
    // ---- initial state: the roots; the levels above max_range_side, which split without a search, in one launch ----
    int start_exp = top;      // first exponent whose queue needs the level loop
    {
        std::vector<QtForcedRoot> forced;
        std::vector<int> qn(top + 1, 0);
        std::vector<std::vector<int32_t>> init(top + 1);
        int n_nodes0 = n_root, threads = 0;
        const int stop = std::max(max_exp, 0);
        for (int r = 0; r < n_root; ++r) {
            const int e = nodes[r].exp;
            if (e > stop) {
                const int D = e - stop;
                FIC_REQUIRE(D < 15, "fic_qt_encode: image too large");
                const int desc = (int)((((long long)1 << (2 * (D + 1))) - 4) / 3);
                QtForcedRoot fr{r, nodes[r].si, nodes[r].sj, e, D, n_nodes0, qn[stop], threads};
                forced.push_back(fr);
                nodes[r].kind = 0;
                nodes[r].child0 = n_nodes0;
                n_nodes0 += desc;
                threads += desc;
                qn[stop] += 1 << (2 * D);
            } else {
                init[e].insert(init[e].end(), {nodes[r].si, nodes[r].sj, r, 0});
            }
        }
        FIC_REQUIRE((size_t)n_nodes0 <= node_cap, "fic_qt_encode: internal: node table too small for the forced levels");
        // roots that start at or below max_exp sit behind the forced entries of their queue
        std::vector<int> q_first(top + 1, 0);
        for (int e = 0; e <= top; ++e) { q_first[e] = qn[e]; qn[e] += (int)init[e].size() / 4; }
        // one pinned staging block: counts | roots | forced descriptors | queue entries
        size_t bytes = 64 * sizeof(int) + (size_t)n_root * sizeof(QtNodeDev) + forced.size() * sizeof(QtForcedRoot);
        for (int e = 0; e <= top; ++e) bytes += init[e].size() * sizeof(int32_t);
        // (the block belongs to the tree encoders alone and every call of theirs ends with a host wait, so it is free here)
        FIC_TRY(h->pin_tree.reserve(bytes + 256));
        FIC_TRY(h->t_aux.reserve(forced.size() * sizeof(QtForcedRoot) + 64));
        char* stage = h->pin_tree.as<char>();
        int* c0 = reinterpret_cast<int*>(stage);
        memset(c0, 0, 64 * sizeof(int));
        c0[0] = n_nodes0;
        for (int e = 0; e <= top; ++e) c0[1 + e] = qn[e];
        memcpy(h_counts, c0, 64 * sizeof(int));
        size_t at = 64 * sizeof(int);
        FIC_CUDA(cudaMemcpyAsync(d_counts, c0, 64 * sizeof(int), cudaMemcpyHostToDevice, st));
        memcpy(stage + at, nodes.data(), (size_t)n_root * sizeof(QtNodeDev));
        FIC_CUDA(cudaMemcpyAsync(d_nodes, stage + at, (size_t)n_root * sizeof(QtNodeDev), cudaMemcpyHostToDevice, st));
        at += (size_t)n_root * sizeof(QtNodeDev);
        if (!forced.empty()) {
            memcpy(stage + at, forced.data(), forced.size() * sizeof(QtForcedRoot));
            FIC_CUDA(cudaMemcpyAsync(h->t_aux.p, stage + at, forced.size() * sizeof(QtForcedRoot), cudaMemcpyHostToDevice, st));
            at += forced.size() * sizeof(QtForcedRoot);
        }
        for (int e = 0; e <= top; ++e)
            if (!init[e].empty()) {
                memcpy(stage + at, init[e].data(), init[e].size() * sizeof(int32_t));
                FIC_CUDA(cudaMemcpyAsync(d_queue + qoff[e] + 4 * (size_t)q_first[e], stage + at, init[e].size() * sizeof(int32_t),
                                         cudaMemcpyHostToDevice, st));
                at += init[e].size() * sizeof(int32_t);
            }
        h->counters[CNT_H2D] += (int64_t)at;
        if (threads) {
            FIC_REQUIRE((size_t)qn[stop] <= qcap[stop], "fic_qt_encode: internal: queue of exponent %d overflows", stop);
            qt_forced_kernel<<<(threads + 127) / 128, 128, 0, st>>>(h->t_aux.as<QtForcedRoot>(), (int)forced.size(), threads, d_nodes,
                                                                    d_queue + qoff[stop]);
            FIC_CHECK_LAUNCH(h, "qt_forced_kernel");
            start_exp = stop;
        }
    }
    const int nmodes = onlypos ? 1 : 2;
    bool boxsum_ready = false;     // 2x2 cell sums of the image, made once if a level goes through the tensor-core search
    bool counts_fresh = true;      // h_counts mirrors the device counters
    for (int e = start_exp; e >= 0; --e) {
        const int n = h_counts[1 + e];
        if (n == 0) continue;
        FIC_REQUIRE((size_t)n <= qcap[e], "fic_qt_encode: internal: queue of exponent %d overflows", e);
        const int32_t* queue = d_queue + qoff[e];
        int32_t* next_queue = e > 0 ? d_queue + qoff[e - 1] : nullptr;
        const int next_cap = e > 0 ? (int)qcap[e - 1] : 0;
        int mode = 0;
        if (e > max_exp) mode = 1;
        else if (e == 0) mode = 2;
        if (mode == 0) {
            const int pk = e - min_needed;
            FIC_REQUIRE(pk >= 0 && pk < npools, "fic_qt_encode: internal: no pool for exponent %d", e);
            FIC_TRY(h->t_res.reserve((size_t)n * 6 * sizeof(int32_t) + 16));
            FIC_TRY(h->t_part.reserve((size_t)n * 3 * sizeof(double) + 16));
            QtPoolDev pool;
            pool.offsets = h->t_lists.as<int32_t>() + pool_base[pk];
            pool.rows = pool.offsets + 73;
            // A pool searched over all classes is the brute-force contraction: on the tensor cores (qt_tc_search), the CUDA-core
            // search only for the ranges it hands back.
            const int32_t* need = nullptr;
            if (full && n >= 32 && qt_tc_applicable(H, W, e, max_scale, geom[pk].hs, geom[pk].vs)) {
                if (!boxsum_ready) {
                    FIC_TRY(bf_prepare8(h, d_img, H, W, false, st));
                    boxsum_ready = true;
                }
                FIC_TRY(h->t_rcls.reserve((size_t)n * 6 * sizeof(int32_t) + 64));
                FIC_TRY(h->t_need.reserve((size_t)n * sizeof(int32_t) + 64));
                qt_range_class_kernel<<<(2 * n + 127) / 128, 128, 0, st>>>(I1, I2, W + 1, queue, n, 1 << e, h->t_rcls.as<int32_t>());
                FIC_CHECK_LAUNCH(h, "qt_range_class_kernel");
                FIC_TRY(qt_tc_search(h, d_img, H, W, e, queue, n, h->t_rcls.as<int32_t>(), h->t_cls.as<int32_t>() + geom[pk].cls_off, nmodes,
                                     max_scale, scale_bits, offset_bits, h->t_res.as<int32_t>(), h->t_part.as<double>(),
                                     h->t_need.as<int32_t>(), geom[pk].hs, geom[pk].vs, pool.offsets, pool.rows, st));
                need = h->t_need.as<int32_t>();
#ifdef FIC_TUNING
                {
                    std::vector<int32_t> hn(n);
                    cudaStreamSynchronize(st);
                    cudaMemcpy(hn.data(), need, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost);
                    long long f = 0;
                    for (int v : hn) f += v != 0;
                    fprintf(stderr, "quadtree level e=%d: %d ranges, %lld left to the exact search\n", e, n, f);
                }
#endif
            }
            // (a warp per range while the class lists are short -- 256 domains per class on average; beyond that a CTA per range)
            if (!full && !need && (1 << e) * (1 << e) <= 64 && geom[pk].cnt <= (size_t)72 * 256) {
                qt_search_warp_kernel<<<(n + QT_WARP_RANGES - 1) / QT_WARP_RANGES, QT_WARP_RANGES * 32, 0, st>>>(
                    d_img, h->t_hvpool.as<float>(), W, I1, I2, queue, n, 1 << e, pool, nmodes, max_scale, scale_bits, offset_bits,
                    h->t_res.as<int32_t>(), h->t_part.as<double>());
                FIC_CHECK_LAUNCH(h, "qt_search_warp_kernel");
            } else {
                qt_search_kernel<<<n, TE_THREADS, 0, st>>>(d_img, h->t_hvpool.as<float>(), W, I1, I2, queue, 1 << e, pool, full, nmodes,
                                                           max_scale, scale_bits, offset_bits, h->t_res.as<int32_t>(),
                                                           h->t_part.as<double>(), need);
                FIC_CHECK_LAUNCH(h, "qt_search_kernel");
            }
        }
        qt_decide_kernel<<<(n + 127) / 128, 128, 0, st>>>(queue, n, e, mode, tol, min_exp, h->t_res.as<int32_t>(), h->t_part.as<double>(),
                                                          d_nodes, d_counts, next_queue, (int)node_cap, next_cap);
        FIC_CHECK_LAUNCH(h, "qt_decide_kernel");
        // Can this level have split anything?  Only then does the host need the new queue length (and node count).
        const bool may_split = mode == 1 || (mode == 0 && e > min_exp);
        counts_fresh = !may_split;
        if (may_split && e > 0) {
            FIC_CUDA(cudaMemcpyAsync(h_counts, d_counts, 64 * sizeof(int), cudaMemcpyDeviceToHost, st));
            FIC_TRY(fic_sync(h, st));
            h->counters[CNT_D2H] += 64 * sizeof(int);
            FIC_REQUIRE(h_counts[40] == 0, "fic_qt_encode: internal: node table overflow");
            counts_fresh = true;
        }
    }
    // ---- node table back, preorder emission ----
    if (!counts_fresh) {
        FIC_CUDA(cudaMemcpyAsync(h_counts, d_counts, 64 * sizeof(int), cudaMemcpyDeviceToHost, st));
        FIC_TRY(fic_sync(h, st));
    }
    const int n_nodes = h_counts[0];
    if (n_nodes > cap) { fic_set_error("fic_qt_encode: node capacity %d too small", cap); return FIC_ECAP; }
    FIC_REQUIRE(top < QT_EMIT_EXPS, "fic_qt_encode: image too large");
    FIC_TRY(h->d_in_a.reserve((size_t)n_nodes * 3 * sizeof(int32_t) + 64));
    FIC_TRY(h->d_out_a.reserve((size_t)n_nodes * 8 * sizeof(int32_t) + 64));
    FIC_TRY(h->d_out_b.reserve((size_t)n_nodes * 3 * sizeof(double) + 64));
    int32_t* d_order = h->d_in_a.as<int32_t>();
    qt_emit_kernel<<<1, 1024, 0, st>>>(d_nodes, n_nodes, n_root, d_order, d_order + n_nodes, d_order + 2 * (size_t)n_nodes,
                                       h->d_out_a.as<int32_t>(), h->d_out_b.as<double>());
    FIC_CHECK_LAUNCH(h, "qt_emit_kernel");
    FIC_TRY(fic_copy_out(h, out_nodes, h->d_out_a.p, (size_t)n_nodes * 8 * sizeof(int32_t), false, st));
    FIC_TRY(fic_copy_out(h, out_vals, h->d_out_b.p, (size_t)n_nodes * 3 * sizeof(double), false, st));
    FIC_TRY(fic_sync(h, st));
    const int count = n_nodes;
    if (n_roots) *n_roots = (int32_t)n_root;
    return count;
}

// ---- HV partition on the device (hvv2/encoder.py:35-82) -------------------------------------------------------------
// One level = all rectangles of one tree depth.  The level's rectangles, node ids and flags (bit 0: small enough to be
// searched, bit 1: large enough to be split) sit in device arrays; after the search and split launches hv_decide_kernel
// applies `err < threshold` / "cannot be split any further" (encoder.py:36-43), stores leaf records in the node table or
// appends the two children (child 1 keeps split + 1 columns / rows) to the node table and to the next level's arrays.
// The host reads back one integer per level -- the size of the next one -- and the node table at the end.
struct HvNodeDev {
    int32_t i, j, w, h;
    int32_t kind;           // -1 pending, 0 internal, 1 leaf
    int32_t vertical, split, c1;   // children c1, c1 + 1
    int32_t rec[6];
    double vals[3];
};
static_assert(sizeof(HvNodeDev) == 80, "HvNodeDev layout");

__host__ __device__ __forceinline__ int hv_flags(int w, int hgt, int min_w, int min_h, int max_w, int max_h)
{
    const bool small = w <= min_w && hgt <= min_h;              // must become a leaf (searched whatever the maximum says)
    const bool fits = w <= max_w && hgt <= max_h;               // searched
    return (fits || small ? 1 : 0) | (!small ? 2 : 0);
}

// One warp per rectangle of the level: merges the chunk winners of its search (hv_pick_kernel's rule), finds the split
// (the biased-difference rule, hvv2/encoder.py:45-63) and decides.  counts: [0] nodes, [1 + parity] size of the next level (the other one is the
// current level's and is cleared here for the level after next), [3] error flag (1 = a rectangle that must be split cannot
// be, 2 = capacity), [4 + parity] number of rectangles of the next level that will be searched.
// Level arrays (level_cap entries each): rects[4], ids, flags, sidx (index into the search results, -1 = not searched),
// slist (the rectangles to search, compacted).
__global__ void hv_decide_kernel(const long long* __restrict__ I1, int W1, const int32_t* __restrict__ rects,
                                 const int32_t* __restrict__ ids, const int32_t* __restrict__ flags,
                                 const int32_t* __restrict__ sidx, int n,
                                 const int32_t* __restrict__ prec, const double* __restrict__ pvals,
                                 const unsigned long long* __restrict__ pseq, int S, double thr, int min_w, int min_h, int max_w,
                                 int max_h, HvNodeDev* __restrict__ nodes, int* __restrict__ counts, int parity,
                                 int32_t* __restrict__ next_rects, int32_t* __restrict__ next_ids, int32_t* __restrict__ next_flags,
                                 int32_t* __restrict__ next_sidx, int32_t* __restrict__ next_slist, int node_cap, int level_cap)
{
    const int q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (blockIdx.x == 0 && threadIdx.x == 0) {   // nobody reads or writes these during this launch
        counts[1 + (parity ^ 1)] = 0;
        counts[4 + (parity ^ 1)] = 0;
    }
    if (q >= n) return;
    const int id = ids[q], f = flags[q];
    const int ri = rects[4 * (size_t)q], rj = rects[4 * (size_t)q + 1], w = rects[4 * (size_t)q + 2], hgt = rects[4 * (size_t)q + 3];
    HvNodeDev& nd = nodes[id];
    if (f & 1) {
        // first strict minimum in scan order over the S chunk winners
        const size_t k0 = (size_t)sidx[q] * S;
        int bs = lane < S ? lane : 0;
        double bR = pvals[3 * (k0 + bs) + 2];
        unsigned long long bq = pseq[k0 + bs];
        for (int c = lane + 32; c < S; c += 32) {
            const double R = pvals[3 * (k0 + c) + 2];
            const unsigned long long sq = pseq[k0 + c];
            if (fic_better(R, sq, bR, bq)) { bR = R; bq = sq; bs = c; }
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            const double oR = __shfl_xor_sync(0xffffffffu, bR, off);
            const unsigned long long oq = __shfl_xor_sync(0xffffffffu, bq, off);
            const int os = __shfl_xor_sync(0xffffffffu, bs, off);
            if (fic_better(oR, oq, bR, bq) || (oq == bq && os < bs)) { bR = oR; bq = oq; bs = os; }
        }
        bs = __shfl_sync(0xffffffffu, bs, 0);
        const size_t src = k0 + bs;
        const double err = pvals[3 * src + 2];
        if (!(f & 2) || err < thr) {   // hvv2/encoder.py:38,42: raw MSE against the threshold
            if (lane == 0) {
                nd.kind = 1;
                for (int k = 0; k < 6; ++k) nd.rec[k] = prec[6 * src + k];
                for (int k = 0; k < 3; ++k) nd.vals[k] = pvals[3 * src + k];
            }
            return;
        }
    }
    // biased-difference split (hvv2/encoder.py:45-63, 77-82)
    bool have[2];
    int best_k[2];
    double best_v[2];
    for (int axis = 0; axis < 2; ++axis) {
        have[axis] = axis == 0 ? w > min_w : hgt > min_h;
        const int L = axis == 0 ? w : hgt;
        double bv = -__longlong_as_double(0x7ff0000000000000ll);
        int bk = 0x7fffffff;
        if (have[axis]) {
            for (int t = lane; t < L - 1; t += 32) {
                long long s0, s1;
                if (axis == 0) {  // column sums (np.sum(axis=0))
                    s0 = box(I1, W1, ri, rj + t, ri + hgt, rj + t + 1);
                    s1 = box(I1, W1, ri, rj + t + 1, ri + hgt, rj + t + 2);
                } else {
                    s0 = box(I1, W1, ri + t, rj, ri + t + 1, rj + w);
                    s1 = box(I1, W1, ri + t + 1, rj, ri + t + 2, rj + w);
                }
                double bt = (double)t, other = DS((double)(L - 1), bt);
                double bias = DD(bt < other ? bt : other, (double)(L - 1));
                double v = DM((double)(s0 - s1), bias);
                if (v > bv || (v == bv && t < bk)) { bv = v; bk = t; }
            }
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) {
                double ov = __shfl_xor_sync(0xffffffffu, bv, off);
                int ok = __shfl_xor_sync(0xffffffffu, bk, off);
                if (ov > bv || (ov == bv && ok < bk)) { bv = ov; bk = ok; }
            }
        }
        best_k[axis] = have[axis] ? bk : 0;
        best_v[axis] = have[axis] ? bv : 0.0;
    }
    if (lane != 0) return;
    if (!have[0] && !have[1]) { counts[3] = 1; return; }
    bool vertical;
    if (!have[0]) vertical = false;
    else if (!have[1]) vertical = true;
    else vertical = fabs(best_v[0]) > fabs(best_v[1]);  // encoder.py:63
    const int s = vertical ? best_k[0] : best_k[1];
    const int base = atomicAdd(&counts[0], 2);
    const int at = atomicAdd(&counts[1 + parity], 2);
    if (base + 2 > node_cap || at + 2 > level_cap) { counts[3] = 2; return; }
    nd.kind = 0;
    nd.vertical = vertical ? 1 : 0;
    nd.split = s;
    nd.c1 = base;
    HvNodeDev a, b;
    if (vertical) {
        a.i = ri; a.j = rj; a.w = s + 1; a.h = hgt;
        b.i = ri; b.j = rj + s + 1; b.w = w - s - 1; b.h = hgt;
    } else {
        a.i = ri; a.j = rj; a.w = w; a.h = s + 1;
        b.i = ri + s + 1; b.j = rj; b.w = w; b.h = hgt - s - 1;
    }
    a.kind = b.kind = -1;
    a.vertical = b.vertical = a.split = b.split = 0;
    a.c1 = b.c1 = -1;
    for (int k = 0; k < 6; ++k) a.rec[k] = b.rec[k] = 0;
    for (int k = 0; k < 3; ++k) a.vals[k] = b.vals[k] = 0.0;
    nodes[base] = a;
    nodes[base + 1] = b;
    const HvNodeDev* ch[2] = {&a, &b};
#pragma unroll
    for (int c = 0; c < 2; ++c) {
        int32_t* r = next_rects + 4 * (size_t)(at + c);
        r[0] = ch[c]->i; r[1] = ch[c]->j; r[2] = ch[c]->w; r[3] = ch[c]->h;
        next_ids[at + c] = base + c;
        const int cf = hv_flags(ch[c]->w, ch[c]->h, min_w, min_h, max_w, max_h);
        next_flags[at + c] = cf;
        int k = -1;
        if (cf & 1) {
            k = atomicAdd(&counts[4 + parity], 1);
            next_slist[k] = at + c;
        }
        next_sidx[at + c] = k;
    }
}

// the search of one level with the chunk winners left in the handle's partial buffers (hv_decide_kernel merges them)
static int hv_search_level(fic_handle* h, const uint8_t* d_img, int H, int W, const int32_t* d_rects, const int32_t* d_slist,
                           const int32_t* d_slist_sorted /* list sorted by shape (hv_group_kernel) or nullptr */,
                           unsigned long long* d_gbound /* with the sorted list: one bound per entry, all-ones */,
                           const float2* d_rsums /* ... and (sum r, sum r^2) per entry */, int count, int max_px, int scale_bits, int offset_bits, double max_scale, int* S_out,
                           const int32_t** prec_out, const double** pvals_out, const unsigned long long** pseq_out, cudaStream_t st)
{
    const float* Dp = h->t_hvpool.as<float>();
    const int smem_px = max_px <= 10240 ? max_px : 0;
    const bool grouped = d_slist_sorted != nullptr;
    const int n_cta = grouped ? (count + HV_GROUP - 1) / HV_GROUP : count;
    int S = (HV_SMULT * h->num_sms + n_cta - 1) / n_cta;      // see HV_SMULT
    if (S > HV_SCAP) S = HV_SCAP;
    if (S < 1) S = 1;
    const size_t np_ = (size_t)count * S;
    FIC_TRY(h->t_part.reserve(np_ * (6 * sizeof(int32_t) + 3 * sizeof(double) + sizeof(unsigned long long)) + 64));
    double* pvals = h->t_part.as<double>();
    unsigned long long* pseq = reinterpret_cast<unsigned long long*>(pvals + 3 * np_);
    int32_t* prec = reinterpret_cast<int32_t*>(pseq + np_);
    if (grouped) {
        const size_t smem = HV_GROUP * HV_GROUP_MAX_PX * sizeof(float) + sizeof(HvGroupBest);
        if (!h->hv_group_attr_ready) {
            FIC_CUDA(cudaFuncSetAttribute(hv_search_grouped_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            h->hv_group_attr_ready = true;
        }
        hv_search_grouped_kernel<<<dim3(n_cta, S), TE_THREADS, smem, st>>>(d_img, Dp, H, W, d_rects, max_scale, scale_bits, offset_bits,
                                                                          prec, pvals, pseq, d_slist_sorted, count, d_gbound, d_rsums);
        FIC_CHECK_LAUNCH(h, "hv_search_grouped_kernel");
        *S_out = S; *prec_out = prec; *pvals_out = pvals; *pseq_out = pseq;
        return FIC_OK;
    }
    hv_search_kernel<<<dim3(count, S), TE_THREADS, smem_px * sizeof(float), st>>>(d_img, Dp, H, W, d_rects, max_scale, scale_bits,
                                                                                 offset_bits, prec, pvals, pseq, smem_px, d_slist);
    FIC_CHECK_LAUNCH(h, "hv_search_kernel");
    *S_out = S; *prec_out = prec; *pvals_out = pvals; *pseq_out = pseq;
    return FIC_OK;
}

int hv_encode_impl(fic_handle* h, const uint8_t* d_img, int H, int W, int min_h, int min_w, int max_h, int max_w,
                   int scale_bits, int offset_bits, double thr, double max_scale, int32_t* out_nodes, double* out_vals,
                   int cap, cudaStream_t st)
{
    FIC_TRY(build_integrals(h, d_img, H, W, st));
    FIC_TRY(hv_build_pool(h, d_img, H, W, st));
    const long long* I1 = h->t_int1.as<long long>();
    // Capacity: a tree over H*W pixels has at most 2*H*W - 1 nodes and H*W rectangles per depth, far more than real
    // images produce; start from an estimate and run again with more room if the device reports an overflow.
    const size_t hard = 2 * (size_t)H * W + 64;
    size_t node_cap = std::min(hard, std::max((size_t)65536, (size_t)H * W / 4));
    std::vector<HvNodeDev> nodes;
    for (;;) {
        const size_t level_cap = std::min((size_t)H * W + 2, node_cap);
        FIC_REQUIRE(node_cap < ((size_t)1 << 30), "fic_hv_encode: image too large");
        FIC_TRY(h->t_nodes.reserve(node_cap * sizeof(HvNodeDev) + 64));
        FIC_TRY(h->t_work.reserve(2 * level_cap * 9 * sizeof(int32_t) + level_cap * (sizeof(unsigned long long) + sizeof(float2)) + 128));     // two levels x (rect[4], id, flags, sidx, slist, slist sorted by shape) + bounds + range sums
        FIC_TRY(h->t_count.reserve(64 * sizeof(int)));
        FIC_TRY(h->pin_a.reserve(256 + 64 * sizeof(int)));
        HvNodeDev* d_nodes = h->t_nodes.as<HvNodeDev>();
        int* d_counts = h->t_count.as<int>();
        int* h_counts = h->pin_a.as<int>() + 64;
        int n_search = 0;
        int32_t* lvl[2] = {h->t_work.as<int32_t>(), h->t_work.as<int32_t>() + level_cap * 9};
        unsigned long long* d_gbound = reinterpret_cast<unsigned long long*>((reinterpret_cast<uintptr_t>(lvl[1] + level_cap * 9) + 15) & ~(uintptr_t)15);
        float2* d_rsums = reinterpret_cast<float2*>(d_gbound + level_cap);
        {
            HvNodeDev root{};
            root.i = 0; root.j = 0; root.w = W; root.h = H; root.kind = -1; root.c1 = -1;
            const int f = hv_flags(W, H, min_w, min_h, max_w, max_h);
            int32_t first[8] = {0, 0, W, H, 0, f, (f & 1) ? 0 : -1, 0};
            memset(h_counts, 0, 8 * sizeof(int));
            h_counts[0] = 1;
            h_counts[6] = (f & 1) ? 1 : 0;      // length of the root level's search list (read by hv_group_kernel below)
            FIC_CUDA(cudaMemcpyAsync(d_counts, h_counts, 8 * sizeof(int), cudaMemcpyHostToDevice, st));
            FIC_CUDA(cudaMemcpyAsync(d_nodes, &root, sizeof(root), cudaMemcpyHostToDevice, st));
            FIC_CUDA(cudaMemcpyAsync(lvl[0], first, 4 * sizeof(int32_t), cudaMemcpyHostToDevice, st));
            FIC_CUDA(cudaMemcpyAsync(lvl[0] + level_cap * 4, first + 4, sizeof(int32_t), cudaMemcpyHostToDevice, st));
            FIC_CUDA(cudaMemcpyAsync(lvl[0] + level_cap * 5, first + 5, sizeof(int32_t), cudaMemcpyHostToDevice, st));
            FIC_CUDA(cudaMemcpyAsync(lvl[0] + level_cap * 6, first + 6, sizeof(int32_t), cudaMemcpyHostToDevice, st));
            FIC_CUDA(cudaMemcpyAsync(lvl[0] + level_cap * 7, first + 7, sizeof(int32_t), cudaMemcpyHostToDevice, st));
            FIC_CUDA(cudaMemcpyAsync(lvl[0] + level_cap * 8, first + 7, sizeof(int32_t), cudaMemcpyHostToDevice, st));
            if (std::max(max_w * max_h, min_w * min_h) <= HV_GROUP_MAX_PX) {
                hv_group_kernel<<<1, 1024, 0, st>>>(lvl[0], lvl[0] + level_cap * 7, d_counts + 6, lvl[0] + level_cap * 8, lvl[0] + level_cap * 6,
                                                    d_gbound, d_img, W, d_rsums);
                FIC_CHECK_LAUNCH(h, "hv_group_kernel");
            }
            FIC_TRY(fic_sync(h, st));
            n_search = (f & 1) ? 1 : 0;
        }
        int n = 1, cur = 0, fault = 0, parity = 0;
        const int max_px = max_w * max_h > 0 ? max_w * max_h : 1;   // searched rectangles are at most max_w x max_h
        const int small_px = min_w * min_h > 0 ? min_w * min_h : 1;  // ... or at most min_w x min_h (hv_flags)
        const bool grouped = std::max(max_px, small_px) <= HV_GROUP_MAX_PX;   // rectangles of equal shape searched together
        while (n > 0) {
            const int32_t* rects = lvl[cur];
            const int32_t* ids = lvl[cur] + level_cap * 4;
            const int32_t* flags = lvl[cur] + level_cap * 5;
            const int32_t* sidx = lvl[cur] + level_cap * 6;
            const int32_t* slist = lvl[cur] + level_cap * 7;
            int32_t* nrects = lvl[cur ^ 1];
            int S = 1;
            const int32_t* prec = nullptr;
            const double* pvals = nullptr;
            const unsigned long long* pseq = nullptr;
            if (n_search)
                FIC_TRY(hv_search_level(h, d_img, H, W, rects, slist, grouped ? lvl[cur] + level_cap * 8 : nullptr, d_gbound, d_rsums, n_search, max_px,
                                        scale_bits, offset_bits, max_scale, &S, &prec, &pvals, &pseq, st));
            hv_decide_kernel<<<(n * 32 + 127) / 128, 128, 0, st>>>(I1, W + 1, rects, ids, flags, sidx, n, prec, pvals, pseq, S, thr,
                                                                   min_w, min_h, max_w, max_h, d_nodes, d_counts, parity, nrects,
                                                                   nrects + level_cap * 4, nrects + level_cap * 5,
                                                                   nrects + level_cap * 6, nrects + level_cap * 7, (int)node_cap,
                                                                   (int)level_cap);
            FIC_CHECK_LAUNCH(h, "hv_decide_kernel");
            if (grouped) {   // the next level's search list sorted by shape; its length stays on the device
                hv_group_kernel<<<1, 1024, 0, st>>>(nrects, nrects + level_cap * 7, d_counts + 4 + parity, nrects + level_cap * 8,
                                                    nrects + level_cap * 6, d_gbound, d_img, W, d_rsums);
                FIC_CHECK_LAUNCH(h, "hv_group_kernel");
            }
            // the two numbers the host needs: the size of the next level and how many of its rectangles are searched
            FIC_CUDA(cudaMemcpyAsync(h_counts, d_counts, 8 * sizeof(int), cudaMemcpyDeviceToHost, st));
            FIC_TRY(fic_sync(h, st));
            h->counters[CNT_D2H] += 8 * sizeof(int);
            fault = h_counts[3];
            if (fault) break;
            n = h_counts[1 + parity];
            n_search = h_counts[4 + parity];
            parity ^= 1;
            cur ^= 1;
        }
        if (fault == 2 && node_cap < hard) { node_cap = std::min(hard, node_cap * 4); continue; }
        FIC_REQUIRE(fault == 0, fault == 1 ? "fic_hv_encode: internal: a rectangle that must be split cannot be"
                                           : "fic_hv_encode: internal: node table overflow");
        nodes.resize(h_counts[0]);
        FIC_TRY(fic_copy_out(h, nodes.data(), d_nodes, nodes.size() * sizeof(HvNodeDev), false, st));
        FIC_TRY(fic_sync(h, st));
        break;
    }
    int count = 0;
    std::vector<int> stack{0};
    while (!stack.empty()) {
        int id = stack.back();
        stack.pop_back();
        const HvNodeDev& nd = nodes[id];
        if (count >= cap) { fic_set_error("fic_hv_encode: node capacity %d too small", cap); return FIC_ECAP; }
        int32_t* o = out_nodes + 8 * (size_t)count;
        double* v = out_vals + 3 * (size_t)count;
        memset(o, 0, 8 * sizeof(int32_t));
        v[0] = v[1] = v[2] = 0.0;
        if (nd.kind == 1) {
            o[0] = 1;
            for (int k = 0; k < 6; ++k) o[1 + k] = nd.rec[k];
            for (int k = 0; k < 3; ++k) v[k] = nd.vals[k];
        } else {
            o[0] = 0; o[1] = nd.vertical; o[2] = nd.split;
            stack.push_back(nd.c1 + 1);
            stack.push_back(nd.c1);
        }
        ++count;
    }
    return count;
}
