// Exact device restatement of the reference's scalar scoring arithmetic.
// Every operation is spelled with an explicit round-to-nearest intrinsic so nvcc can neither fuse
// a multiply-add nor reassociate; the translation units are also compiled with -fmad=false.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define DM(a, b) __dmul_rn((a), (b))
#define DA(a, b) __dadd_rn((a), (b))
#define DS(a, b) __dsub_rn((a), (b))
#define DD(a, b) __ddiv_rn((a), (b))

// R = 1./n * (sqr + s*(s*sds - 2*ab + 2*o*ds) + o*(n*o - 2*rs))
// bruteforce/jit_bfencoder.py:93-95, bf_encoder.py:69, quadtree/encoder.py:378-382, hvv2/encoder.py:182-186
__device__ __forceinline__ double fic_R(double inv_n, double n, double sqr, double s, double sds, double ab, double o,
                                        double ds, double rs)
{
    double inner = DA(DS(DM(s, sds), DM(2.0, ab)), DM(DM(2.0, o), ds));
    double t = DA(DA(sqr, DM(s, inner)), DM(o, DS(DM(n, o), DM(2.0, rs))));
    return DM(inv_n, t);
}

struct BfScore {
    double R, s, o;
};

// bruteforce/jit_bfencoder.py:64-91.  All inputs are the float64 widenings of the reference's values:
// rs = uint64 range sum, sqr = int64 sum of squares, ds/sds/ab = float32 domain sum / squared sum / dot,
// dsds = f32(ds*ds).
__device__ __forceinline__ BfScore fic_bfjit_score(double n, double inv_n, double rs, double sqr, double ds, double sds,
                                                   double dsds, double ab)
{
    BfScore r;
    double den = DS(DM(n, sds), dsds);
    double s, o;
    if (fabs(den) < 1e-4) {
        s = 0.0;
        o = DM(inv_n, rs);
    } else {
        s = DD(DS(DM(n, ab), DM(rs, ds)), den);
        o = DM(inv_n, DS(rs, DM(s, ds)));
    }
    if (s < -1.0 || s > 1.0) {
        double o_left = DM(inv_n, DS(rs, DM(-1.0, ds)));
        double R_left = fic_R(inv_n, n, sqr, -1.0, sds, ab, o_left, ds, rs);
        double o_right = DM(inv_n, DS(rs, DM(1.0, ds)));
        double R_right = fic_R(inv_n, n, sqr, 1.0, sds, ab, o_right, ds, rs);
        if (R_right < R_left) {
            r.R = R_right; r.s = 1.0; r.o = o_right;
        } else {
            r.R = R_left; r.s = -1.0; r.o = o_left;
        }
        return r;
    }
    r.R = fic_R(inv_n, n, sqr, s, sds, ab, o, ds, rs);
    r.s = s;
    r.o = o;
    return r;
}

// bruteforce/bf_encoder.py:56-71: float64 throughout, threshold 1e-4, no clamp.  sqr is the
// uint8-wrapped sum of squares (bf_encoder.py:33).
__device__ __forceinline__ BfScore fic_bfnp_score(double n, double inv_n, double rs, double sqr, double ds, double sds,
                                                  double ab)
{
    BfScore r;
    double den = DS(DM(n, sds), DM(ds, ds));
    double s, o;
    if (fabs(den) < 1e-4) {
        s = 0.0;
        o = DM(inv_n, rs);
    } else {
        s = DD(DS(DM(n, ab), DM(rs, ds)), den);
        o = DM(inv_n, DS(rs, DM(s, ds)));
    }
    r.R = fic_R(inv_n, n, sqr, s, sds, ab, o, ds, rs);
    r.s = s;
    r.o = o;
    return r;
}

struct QScore {
    double R, s, o;
    int qs, qo;
};

// quadtree/encoder.py:357-384 == hvv2/encoder.py:161-188
__device__ __forceinline__ double fic_q_mse(double n, double inv_n, double sqr, double s, double sds, double ab, double o,
                                            double ds, double rs, double max_scale, int scale_bits, int offset_bits,
                                            int* oqs, int* oqo)
{
    // Divisions by a power of two are exact multiplications by its reciprocal (same IEEE result, a fraction of the
    // instructions of a float64 division): 2^scale_bits and 2^offset_bits always, 2*max_scale when it is one.
    const double sb2 = (double)(1ll << scale_bits), ob2 = (double)(1ll << offset_bits);
    const double inv_sb2 = __longlong_as_double((long long)(1023 - scale_bits) << 52);
    const double inv_ob2 = __longlong_as_double((long long)(1023 - offset_bits) << 52);
    const double two_ms = DM(2.0, max_scale);
    const long long tm_bits = __double_as_longlong(two_ms);
    const bool tm_pow2 = (tm_bits & 0x000fffffffffffffll) == 0 && tm_bits > (2ll << 52) && tm_bits < (2044ll << 52);
    const double inv_tm = __longlong_as_double((2046ll << 52) - tm_bits);
    const double sn = DA(s, max_scale);
    long long qs = (long long)DM(tm_pow2 ? DM(sn, inv_tm) : DD(sn, two_ms), sb2);  // int(): truncation toward zero
    if (qs < 0) qs = 0;
    if (qs > (1ll << scale_bits) - 1) qs = (1ll << scale_bits) - 1;
    s = DS(DM(DM((double)qs, two_ms), inv_sb2), max_scale);
    double coef = 255.0;
    if (s < 0.0) {
        coef = DM(coef, DA(1.0, s));
        if (s == -1.0) coef = 1e-6;
    }
    long long qo = (long long)DM(DD(DA(o, DM(255.0, s)), coef), ob2);
    if (qo < 0) qo = 0;
    if (qo > (1ll << offset_bits) - 1) qo = (1ll << offset_bits) - 1;
    o = DS(DM(DM((double)qo, inv_ob2), coef), DM(255.0, s));
    *oqs = (int)qs;
    *oqo = (int)qo;
    return fic_R(inv_n, n, sqr, s, sds, ab, o, ds, rs);
}

// quadtree/encoder.py:324-355 == hvv2/encoder.py:128-159.  rs, sqr, ds, sds, ab are the float32 sequential sums.
__device__ __forceinline__ QScore fic_q_score(int nn, float rs_f, float sqr_f, float ds_f, float sds_f, float ab_f,
                                              double max_scale, int scale_bits, int offset_bits)
{
    QScore r;
    const double n = (double)nn;
    const double inv_n = DD(1.0, n);
    const double rs = (double)rs_f, sqr = (double)sqr_f, ds = (double)ds_f, sds = (double)sds_f, ab = (double)ab_f;
    const float dsds = __fmul_rn(ds_f, ds_f);
    double den = DS(DM(n, sds), (double)dsds);
    double s, o;
    if (fabs(den) < 1e-6) {
        s = 0.0;
        o = DM(inv_n, rs);
    } else {
        const float rsds = __fmul_rn(rs_f, ds_f);
        s = DD(DS(DM(n, ab), (double)rsds), den);
        o = DM(inv_n, DS(rs, DM(s, ds)));
    }
    if (s < -max_scale || s > max_scale) {
        int qsl, qol, qsr, qor;
        double sl = -max_scale;
        double o_left = DM(inv_n, DS(rs, DM(sl, ds)));
        double R_left = fic_q_mse(n, inv_n, sqr, sl, sds, ab, o_left, ds, rs, max_scale, scale_bits, offset_bits, &qsl, &qol);
        double o_right = DM(inv_n, DS(rs, DM(max_scale, ds)));
        double R_right = fic_q_mse(n, inv_n, sqr, max_scale, sds, ab, o_right, ds, rs, max_scale, scale_bits, offset_bits, &qsr, &qor);
        if (R_right < R_left) {
            r.R = R_right; r.s = max_scale; r.o = o_right; r.qs = qsr; r.qo = qor;
        } else {
            r.R = R_left; r.s = sl; r.o = o_left; r.qs = qsl; r.qo = qol;
        }
        return r;
    }
    r.R = fic_q_mse(n, inv_n, sqr, s, sds, ab, o, ds, rs, max_scale, scale_bits, offset_bits, &r.qs, &r.qo);
    r.s = s;
    r.o = o;
    return r;
}

// ---- isometry index maps --------------------------------------------------------
// brute force enc+dec and quadtree DECODER: Y = rot90(A or A.T, k)
__device__ __forceinline__ void fic_iso_bf(int orient, int rot, int n, int i, int j, int& si, int& sj)
{
    int a, b;
    switch (rot & 3) {
    case 0: a = i; b = j; break;
    case 1: a = j; b = n - 1 - i; break;
    case 2: a = n - 1 - i; b = n - 1 - j; break;
    default: a = n - 1 - j; b = i; break;
    }
    si = orient ? b : a;
    sj = orient ? a : b;
}
// quadtree ENCODER: rot90 first, then .T (quadtree/encoder.py:307-309)
__device__ __forceinline__ void fic_iso_qt_enc(int orient, int rot, int n, int i, int j, int& si, int& sj)
{
    if (orient) { int t = i; i = j; j = t; }
    switch (rot & 3) {
    case 0: si = i; sj = j; break;
    case 1: si = j; sj = n - 1 - i; break;
    case 2: si = n - 1 - i; sj = n - 1 - j; break;
    default: si = n - 1 - j; sj = i; break;
    }
}
// HV: orientation 1 = flipud, rotation 2 = rot180 (hvv2/encoder.py:111-113)
__device__ __forceinline__ void fic_iso_hv(int orient, int rot, int h, int w, int i, int j, int& si, int& sj)
{
    if (rot == 2) { i = h - 1 - i; j = w - 1 - j; }
    if (orient) i = h - 1 - i;
    si = i;
    sj = j;
}

// ---- argmin key: first strict minimum in scan order, NaN never wins ---------------------------
struct BestRec {
    double R;
    unsigned long long seq;  // scan-order index, ~0ull = nothing yet
};
__device__ __forceinline__ bool fic_better(double R, unsigned long long seq, double bR, unsigned long long bseq)
{
    // (R < bR) || (R == bR && seq < bseq); a NaN R fails both; "nothing yet" has bR = +inf and bseq = max
    // +inf never wins either: the reference starts from best = inf and compares with strict <
    return (R < bR) || (R == bR && seq < bseq && R < __longlong_as_double(0x7ff0000000000000ll));
}
