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
// columns: one thread per column (coalesced across the warp), sixteen rows loaded before any is stored so that the
// loads are not serialised behind the in-place stores
__global__ void integral_cols_kernel(int H, int W, long long* __restrict__ I1, long long* __restrict__ I2)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x > W) return;
    const int W1 = W + 1;
    constexpr int U = 16;
    long long s = 0, q = 0;
    for (int y0 = 1; y0 <= H; y0 += U) {
        long long a[U], b[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const bool in = y0 + u <= H;
            a[u] = in ? I1[(size_t)(y0 + u) * W1 + x] : 0;
            b[u] = in ? I2[(size_t)(y0 + u) * W1 + x] : 0;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (y0 + u <= H) {
                s += a[u];
                q += b[u];
                I1[(size_t)(y0 + u) * W1 + x] = s;
                I2[(size_t)(y0 + u) * W1 + x] = q;
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
    integral_cols_kernel<<<(W + 1 + 31) / 32, 32, 0, st>>>(H, W, h->t_int1.as<long long>(), h->t_int2.as<long long>());
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
                 const int32_t* __restrict__ ranges /*[n][2] = si,sj*/, int side, QtPoolDev pool, int full, int nmodes,
                 double max_scale, int scale_bits, int offset_bits, int32_t* __restrict__ rec /*[n][6]*/,
                 double* __restrict__ vals /*[n][3]*/)
{
    __shared__ float s_rs, s_sqr;
    __shared__ int seg_begin[2 * 72 + 1];  // prefix of candidate counts over (mode, class) segments
    __shared__ int seg_list[2 * 72];       // class index of the segment
    __shared__ int seg_rot[2], seg_orient[2];
    __shared__ int n_seg;
    __shared__ unsigned long long s_bound;   // bit pattern of the best score found so far by this CTA (+inf at start)
    __shared__ float s_range[QT_RANGE_SMEM];  // the range block as floats (sides up to 32)
    const int tid = threadIdx.x;
    const int si = ranges[2 * blockIdx.x], sj = ranges[2 * blockIdx.x + 1];
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
__global__ void __launch_bounds__(TE_THREADS, 4)
hv_search_kernel(const uint8_t* __restrict__ img, const float* __restrict__ Dp, int H, int W,
                 const int32_t* __restrict__ rects /*[n][4] = i,j,w,h*/, double max_scale, int scale_bits, int offset_bits,
                 int32_t* __restrict__ rec, double* __restrict__ vals, unsigned long long* __restrict__ seqs, int smem_px)
{
    extern __shared__ float s_range[];   // the rectangle's pixels (when they fit)
    __shared__ float s_rs, s_sqr;
    __shared__ unsigned long long s_bound;   // bit pattern of the best score found so far by this CTA (+inf at start)
    const int tid = threadIdx.x;
    if (tid == 0) s_bound = 0x7ff0000000000000ull;
    const int ri = rects[4 * blockIdx.x], rj = rects[4 * blockIdx.x + 1], w = rects[4 * blockIdx.x + 2],
              hgt = rects[4 * blockIdx.x + 3];
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
    int S = (4 * h->num_sms + count - 1) / count;      // aim at >= 4 CTAs per SM
    if (S > 64) S = 64;
    if (S < 1) S = 1;
    if (S == 1) {
        hv_search_kernel<<<dim3(count, 1), TE_THREADS, smem_px * sizeof(float), st>>>(d_img, Dp, H, W, d_rects, max_scale, scale_bits,
                                                                                     offset_bits, d_rec, d_vals, nullptr, smem_px);
        FIC_CHECK_LAUNCH(h, "hv_search_kernel");
        return FIC_OK;
    }
    const size_t np_ = (size_t)count * S;
    FIC_TRY(h->t_part.reserve(np_ * (6 * sizeof(int32_t) + 3 * sizeof(double) + sizeof(unsigned long long)) + 64));
    double* pvals = h->t_part.as<double>();
    unsigned long long* pseq = reinterpret_cast<unsigned long long*>(pvals + 3 * np_);
    int32_t* prec = reinterpret_cast<int32_t*>(pseq + np_);
    hv_search_kernel<<<dim3(count, S), TE_THREADS, smem_px * sizeof(float), st>>>(d_img, Dp, H, W, d_rects, max_scale, scale_bits,
                                                                                 offset_bits, prec, pvals, pseq, smem_px);
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

// ---- HV split (hvv2/encoder.py:45-63, 77-82): one warp per rectangle --------------------------------
// out[k] = {have_v, best_v, have_h, best_h}; outv[k] = {value_v, value_h}
__global__ void hv_split_kernel(const long long* __restrict__ I1, int W1, const int32_t* __restrict__ rects, int count,
                                int min_w, int min_h, int32_t* __restrict__ out, double* __restrict__ outv)
{
    const int k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (k >= count) return;
    const int ri = rects[4 * k], rj = rects[4 * k + 1], w = rects[4 * k + 2], hgt = rects[4 * k + 3];
    for (int axis = 0; axis < 2; ++axis) {
        const bool have = axis == 0 ? w > min_w : hgt > min_h;
        const int L = axis == 0 ? w : hgt;
        double bv = -__longlong_as_double(0x7ff0000000000000ll);
        int bk = 0x7fffffff;
        if (have) {
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
        if (lane == 0) {
            out[4 * (size_t)k + 2 * axis] = have ? 1 : 0;
            out[4 * (size_t)k + 2 * axis + 1] = have ? bk : 0;
            outv[2 * (size_t)k + axis] = have ? bv : 0.0;
        }
    }
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

struct QtNode {
    int si, sj, exp;
    int kind;       // 0 internal, 1 leaf
    int child[4];
    int32_t rec[6];
    double vals[3];
};

// quadtree/encoder.py:111-125: roots in _partition_image order
static void qt_roots(std::vector<QtNode>& nodes, std::vector<int>& roots, int si, int sj, int width, int height)
{
    int e = ilog2_floor(std::min(width, height));
    int side = 1 << e;
    QtNode nd{};
    nd.si = si; nd.sj = sj; nd.exp = e; nd.kind = -1;
    roots.push_back((int)nodes.size());
    nodes.push_back(nd);
    if (side < width) qt_roots(nodes, roots, si, sj + side, width - side, height);
    if (side < height) qt_roots(nodes, roots, si + side, sj, side, height - side);
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

    // ---- domain pools (encoder.py:97-109, 261-283): classify on device, stable bucket on host ----
    const int npools = std::max(0, max_exp + 1 - (min_needed + 1) + 1);
    std::vector<std::vector<int32_t>> pool_off(npools), pool_rows(npools);
    for (int k = 0; k < npools; ++k) {
        const int side = 1 << (min_needed + 1 + k);
        const int hs = hstep_in == -1 ? side : hstep_in, vs = vstep_in == -1 ? side : vstep_in;
        const int nv = H - side >= 0 ? 1 + (H - side) / vs : 0, nh = W - side >= 0 ? 1 + (W - side) / hs : 0;
        const size_t cnt = (size_t)nv * nh;
        std::vector<int32_t> cls(cnt * 3);
        if (cnt) {
            FIC_TRY(h->t_cls.reserve(cnt * 3 * sizeof(int32_t)));
            classify_grid_kernel<<<(unsigned)((cnt + 127) / 128), 128, 0, st>>>(I1, I2, W + 1, side, nv, nh, vs, hs, h->t_cls.as<int32_t>());
            FIC_CHECK_LAUNCH(h, "classify_grid_kernel");
            FIC_TRY(fic_copy_out(h, cls.data(), h->t_cls.p, cnt * 3 * sizeof(int32_t), false, st));
            FIC_TRY(fic_sync(h, st));
        }
        std::vector<int32_t>& off = pool_off[k];
        off.assign(73, 0);
        for (size_t p = 0; p < cnt; ++p) off[cls[3 * p] + 1]++;
        for (int c = 0; c < 72; ++c)
            if (off[c + 1] == 0) off[c + 1] = 1;  // placeholder row (0,0,0,0), encoder.py:279-280
        for (int c = 0; c < 72; ++c) off[c + 1] += off[c];
        std::vector<int32_t>& rows = pool_rows[k];
        rows.assign((size_t)off[72] * 4, 0);
        std::vector<int32_t> cur(off.begin(), off.begin() + 72);
        for (size_t p = 0; p < cnt; ++p) {
            int c = cls[3 * p];
            int32_t* r = rows.data() + 4 * (size_t)cur[c]++;
            r[0] = (int)(p / nh) * vs; r[1] = (int)(p % nh) * hs; r[2] = cls[3 * p + 1]; r[3] = cls[3 * p + 2];
        }
    }
    // upload pools: [off0(73) rows0 | off1(73) rows1 | ...]
    std::vector<size_t> pool_base(npools);
    size_t tot = 0;
    for (int k = 0; k < npools; ++k) { pool_base[k] = tot; tot += 73 + pool_rows[k].size(); tot = (tot + 3) & ~(size_t)3; }
    FIC_TRY(h->t_lists.reserve((tot + 4) * sizeof(int32_t)));
    {
        std::vector<int32_t> flat(tot + 4, 0);
        for (int k = 0; k < npools; ++k) {
            std::copy(pool_off[k].begin(), pool_off[k].end(), flat.begin() + pool_base[k]);
            std::copy(pool_rows[k].begin(), pool_rows[k].end(), flat.begin() + pool_base[k] + 73);
        }
        FIC_CUDA(cudaMemcpyAsync(h->t_lists.p, flat.data(), flat.size() * sizeof(int32_t), cudaMemcpyHostToDevice, st));
        h->counters[CNT_H2D] += (int64_t)(flat.size() * sizeof(int32_t));
        FIC_TRY(fic_sync(h, st));
    }

    // ---- level-synchronous partition (encoder.py:127-146) ----
    std::vector<QtNode> nodes;
    std::vector<int> roots;
    qt_roots(nodes, roots, 0, 0, W, H);
    int top = 0;
    for (int r : roots) top = std::max(top, nodes[r].exp);
    std::vector<std::vector<int>> pending(top + 1);
    for (int r : roots) pending[nodes[r].exp].push_back(r);
    const int nmodes = onlypos ? 1 : 2;
    for (int e = top; e >= 0; --e) {
        std::vector<int>& cur = pending[e];
        if (cur.empty()) continue;
        std::vector<int> to_split;
        if (e > max_exp) {
            to_split = cur;
        } else if (e == 0) {  // side 1: dummy leaf, rms 1e-4 (encoder.py:150-152)
            for (int id : cur) {
                QtNode& nd = nodes[id];
                nd.kind = 1;
                memset(nd.rec, 0, sizeof(nd.rec));
                nd.vals[0] = nd.vals[1] = 0.0;
                nd.vals[2] = 1e-8;
                // rms 1e-4 > tol could still ask for a split, but exp 0 > min_exp never holds
            }
        } else {
            const int n = (int)cur.size();
            const int pk = e - min_needed;
            FIC_REQUIRE(pk >= 0 && pk < npools, "fic_qt_encode: internal: no pool for exponent %d", e);
            std::vector<int32_t> rng((size_t)n * 2);
            for (int q = 0; q < n; ++q) { rng[2 * q] = nodes[cur[q]].si; rng[2 * q + 1] = nodes[cur[q]].sj; }
            FIC_TRY(h->t_work.reserve((size_t)n * 2 * sizeof(int32_t) + 16));
            FIC_TRY(h->t_res.reserve((size_t)n * 6 * sizeof(int32_t) + 16));
            FIC_TRY(h->t_aux.reserve((size_t)n * 3 * sizeof(double) + 16));
            FIC_CUDA(cudaMemcpyAsync(h->t_work.p, rng.data(), rng.size() * sizeof(int32_t), cudaMemcpyHostToDevice, st));
            h->counters[CNT_H2D] += (int64_t)(rng.size() * sizeof(int32_t));
            QtPoolDev pool;
            pool.offsets = h->t_lists.as<int32_t>() + pool_base[pk];
            pool.rows = pool.offsets + 73;
            qt_search_kernel<<<n, TE_THREADS, 0, st>>>(d_img, h->t_hvpool.as<float>(), W, I1, I2, h->t_work.as<int32_t>(), 1 << e, pool, full, nmodes,
                                                       max_scale, scale_bits, offset_bits, h->t_res.as<int32_t>(),
                                                       h->t_aux.as<double>());
            FIC_CHECK_LAUNCH(h, "qt_search_kernel");
            std::vector<int32_t> rec((size_t)n * 6);
            std::vector<double> vals((size_t)n * 3);
            FIC_TRY(fic_copy_out(h, rec.data(), h->t_res.p, rec.size() * sizeof(int32_t), false, st));
            FIC_TRY(fic_copy_out(h, vals.data(), h->t_aux.p, vals.size() * sizeof(double), false, st));
            FIC_TRY(fic_sync(h, st));  // also keeps `rng` alive until the upload is consumed
            for (int q = 0; q < n; ++q) {
                QtNode& nd = nodes[cur[q]];
                const double rms = sqrt(vals[3 * q + 2]);  // NaN for a negative error -> comparison false -> leaf
                if (rms > tol && e > min_exp) {
                    to_split.push_back(cur[q]);
                } else {
                    nd.kind = 1;
                    memcpy(nd.rec, &rec[6 * q], sizeof(nd.rec));
                    memcpy(nd.vals, &vals[3 * q], sizeof(nd.vals));
                }
            }
        }
        for (int id : to_split) {
            const int ns = 1 << (e - 1);
            const int si = nodes[id].si, sj = nodes[id].sj;
            nodes[id].kind = 0;
            const int ci[4] = {si, si, si + ns, si + ns}, cj[4] = {sj, sj + ns, sj, sj + ns};  // TL,TR,BL,BR
            for (int c = 0; c < 4; ++c) {
                QtNode ch{};
                ch.si = ci[c]; ch.sj = cj[c]; ch.exp = e - 1; ch.kind = -1;
                nodes[id].child[c] = (int)nodes.size();
                pending[e - 1].push_back((int)nodes.size());
                nodes.push_back(ch);
            }
        }
    }
    // ---- preorder emission ----
    int count = 0;
    std::vector<int> stack;
    for (int r : roots) {
        stack.push_back(r);
        while (!stack.empty()) {
            int id = stack.back();
            stack.pop_back();
            const QtNode& nd = nodes[id];
            if (count >= cap) { fic_set_error("fic_qt_encode: node capacity %d too small", cap); return FIC_ECAP; }
            int32_t* o = out_nodes + 8 * (size_t)count;
            double* v = out_vals + 3 * (size_t)count;
            o[0] = nd.kind; o[7] = nd.exp;
            for (int k = 0; k < 6; ++k) o[1 + k] = nd.kind == 1 ? nd.rec[k] : 0;
            for (int k = 0; k < 3; ++k) v[k] = nd.kind == 1 ? nd.vals[k] : 0.0;
            ++count;
            if (nd.kind == 0)
                for (int c = 3; c >= 0; --c) stack.push_back(nd.child[c]);
        }
    }
    if (n_roots) *n_roots = (int32_t)roots.size();
    return count;
}

struct HvNode {
    int i, j, w, h;
    int kind;  // 0 internal, 1 leaf
    int vertical, split;
    int c1, c2;
    int32_t rec[6];
    double vals[3];
};

int hv_encode_impl(fic_handle* h, const uint8_t* d_img, int H, int W, int min_h, int min_w, int max_h, int max_w,
                   int scale_bits, int offset_bits, double thr, double max_scale, int32_t* out_nodes, double* out_vals,
                   int cap, cudaStream_t st)
{
    FIC_TRY(build_integrals(h, d_img, H, W, st));
    FIC_TRY(hv_build_pool(h, d_img, H, W, st));
    const long long* I1 = h->t_int1.as<long long>();
    std::vector<HvNode> nodes;
    {
        HvNode root{};
        root.i = 0; root.j = 0; root.w = W; root.h = H; root.kind = -1;
        nodes.push_back(root);
    }
    std::vector<int> level{0};
    while (!level.empty()) {
        // classify the rectangles of this depth (hvv2/encoder.py:36-43)
        std::vector<int> search, split;           // node ids
        std::vector<char> must_leaf;              // parallel to `search`
        for (int id : level) {
            const HvNode& nd = nodes[id];
            if (nd.w <= min_w && nd.h <= min_h) { search.push_back(id); must_leaf.push_back(1); }
            else if (nd.w <= max_w && nd.h <= max_h) { search.push_back(id); must_leaf.push_back(0); split.push_back(id); }
            else split.push_back(id);
        }
        const int ns = (int)search.size(), np = (int)split.size();
        std::vector<int32_t> rects((size_t)(ns + np) * 4);
        for (int q = 0; q < ns; ++q) { const HvNode& nd = nodes[search[q]]; int32_t* r = &rects[4 * (size_t)q]; r[0] = nd.i; r[1] = nd.j; r[2] = nd.w; r[3] = nd.h; }
        for (int q = 0; q < np; ++q) { const HvNode& nd = nodes[split[q]]; int32_t* r = &rects[4 * (size_t)(ns + q)]; r[0] = nd.i; r[1] = nd.j; r[2] = nd.w; r[3] = nd.h; }
        FIC_TRY(h->t_work.reserve(rects.size() * sizeof(int32_t) + 16));
        FIC_TRY(h->t_res.reserve((size_t)(ns * 6 + np * 4) * sizeof(int32_t) + 16));
        FIC_TRY(h->t_aux.reserve((size_t)(ns * 3 + np * 2) * sizeof(double) + 16));
        FIC_CUDA(cudaMemcpyAsync(h->t_work.p, rects.data(), rects.size() * sizeof(int32_t), cudaMemcpyHostToDevice, st));
        h->counters[CNT_H2D] += (int64_t)(rects.size() * sizeof(int32_t));
        int32_t* d_rects = h->t_work.as<int32_t>();
        int32_t* d_rec = h->t_res.as<int32_t>();
        int32_t* d_split = d_rec + (size_t)ns * 6;
        double* d_vals = h->t_aux.as<double>();
        double* d_splitv = d_vals + (size_t)ns * 3;
        if (ns) {
            int max_px = 1;
            for (int q = 0; q < ns; ++q) { const HvNode& nd = nodes[search[q]]; if (nd.w * nd.h > max_px) max_px = nd.w * nd.h; }
            FIC_TRY(hv_search_launch(h, d_img, H, W, d_rects, ns, max_px, scale_bits, offset_bits, max_scale, d_rec, d_vals, st));
        }
        if (np) {
            hv_split_kernel<<<(np * 32 + 127) / 128, 128, 0, st>>>(I1, W + 1, d_rects + (size_t)ns * 4, np, min_w, min_h, d_split, d_splitv);
            FIC_CHECK_LAUNCH(h, "hv_split_kernel");
        }
        std::vector<int32_t> rec((size_t)ns * 6 + (size_t)np * 4);
        std::vector<double> vals((size_t)ns * 3 + (size_t)np * 2);
        FIC_TRY(fic_copy_out(h, rec.data(), d_rec, rec.size() * sizeof(int32_t), false, st));
        FIC_TRY(fic_copy_out(h, vals.data(), d_vals, vals.size() * sizeof(double), false, st));
        FIC_TRY(fic_sync(h, st));
        std::vector<char> is_leaf(nodes.size(), 0);
        for (int q = 0; q < ns; ++q) {
            HvNode& nd = nodes[search[q]];
            const double err = vals[3 * (size_t)q + 2];
            if (must_leaf[q] || err < thr) {  // hvv2/encoder.py:38,42: raw MSE against the threshold
                nd.kind = 1;
                memcpy(nd.rec, &rec[6 * (size_t)q], sizeof(nd.rec));
                memcpy(nd.vals, &vals[3 * (size_t)q], sizeof(nd.vals));
                is_leaf[search[q]] = 1;
            }
        }
        std::vector<int> next;
        for (int q = 0; q < np; ++q) {
            const int id = split[q];
            if (is_leaf[id]) continue;
            const int32_t* sp = &rec[(size_t)ns * 6 + 4 * (size_t)q];
            const double* sv = &vals[(size_t)ns * 3 + 2 * (size_t)q];
            const bool have_v = sp[0] != 0, have_h = sp[2] != 0;
            FIC_REQUIRE(have_v || have_h, "fic_hv_encode: internal: rectangle %dx%d cannot be split", nodes[id].w, nodes[id].h);
            bool vertical;
            if (!have_v) vertical = false;
            else if (!have_h) vertical = true;
            else vertical = fabs(sv[0]) > fabs(sv[1]);  // encoder.py:63
            const int s = vertical ? sp[1] : sp[3];
            HvNode a{}, b{};
            const HvNode nd = nodes[id];
            if (vertical) {
                a.i = nd.i; a.j = nd.j; a.w = s + 1; a.h = nd.h;
                b.i = nd.i; b.j = nd.j + s + 1; b.w = nd.w - s - 1; b.h = nd.h;
            } else {
                a.i = nd.i; a.j = nd.j; a.w = nd.w; a.h = s + 1;
                b.i = nd.i + s + 1; b.j = nd.j; b.w = nd.w; b.h = nd.h - s - 1;
            }
            a.kind = b.kind = -1;
            nodes[id].kind = 0;
            nodes[id].vertical = vertical ? 1 : 0;
            nodes[id].split = s;
            nodes[id].c1 = (int)nodes.size();
            nodes.push_back(a);
            nodes[id].c2 = (int)nodes.size();
            nodes.push_back(b);
            next.push_back(nodes[id].c1);
            next.push_back(nodes[id].c2);
        }
        level.swap(next);
    }
    int count = 0;
    std::vector<int> stack{0};
    while (!stack.empty()) {
        int id = stack.back();
        stack.pop_back();
        const HvNode& nd = nodes[id];
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
            stack.push_back(nd.c2);
            stack.push_back(nd.c1);
        }
        ++count;
    }
    return count;
}
