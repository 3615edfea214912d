// Brute-force encoder, exact scan path (CUDA cores).
//
// Replaces the scan of bruteforce/jit_bfencoder.py:32-95 (mode 1) and bruteforce/bf_encoder.py:29-71
// (mode 0): every range block against every lattice position under the 8 isometries, the reference's
// scalar arithmetic replayed bit for bit (scorer.cuh), first strict minimum in scan order.
//
// This path is (a) the parity baseline for the tensor-core screen in bf_screen.cu, (b) the producer of
// the per-range upper bounds the screen needs, (c) the fallback for ranges the screen cannot decide and
// (d) the only path for NumPy semantics, strides != 1 and range sizes != 8.
//
// Data layout in HBM (8x8 ranges):
//   boxsum  uint16 [(H-1),(W-1)]   P[y][x] = sum of the 2x2 pixel cell at (y,x); the pooled domain at
//                                  (di,dj) is D4[a][b] = P[di+2a][dj+2b] = 4*d  (integers <= 1020)
//   stats   per stride-1 position p = di*(W-16)+dj, structure of arrays:
//             sd4  uint32 [NP]     sum of D4  (4 * float32 domain_sum, exact)
//             sq4  uint32 [NP]     sum of D4^2 (16 * exact sum of d^2, NumPy mode)
//             sdd  float  [8][NP]  float32 SEQUENTIAL sum of d*d in the row-major order of each
//                                  isometry -- the one order-dependent quantity (jit_bfencoder.py:54)
//   partial BestRec [ranges][chunks]  per-CTA winners, reduced by bf_finalize_kernel
//
// For 8x8 blocks everything else is exact: sum d*r < 2^24 quarter-units, so float32 sequential == integer.
#include "fic_common.cuh"
#include "scorer.cuh"

#define BF_THREADS 256
#define BF_G 8  // ranges per CTA
#define BF_SCAN_DYN_SMEM (BF_G * BF_THREADS * (sizeof(double) + sizeof(uint32_t)))

struct BfGeom {
    int H, W;        // image
    int npi, npj;    // lattice size
    int si, sj;      // lattice strides
    int PW;          // W-16 : positions per row at stride 1
    int BW;          // W-1  : boxsum row pitch
};

// ---- prep: box sums -------------------------------------------------------------------------
__global__ void bf_boxsum_kernel(const uint8_t* __restrict__ img, int H, int W, uint16_t* __restrict__ P)
{
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y;
    if (x >= W - 1 || y >= H - 1) return;
    const uint8_t* p = img + (size_t)y * W + x;
    P[(size_t)y * (W - 1) + x] = (uint16_t)((int)p[0] + p[1] + p[W] + p[W + 1]);
}

// ---- prep: per-position statistics (stride-1 positions) ---------------------------------------
__global__ void bf_stats_kernel(const uint16_t* __restrict__ P, int H, int W, uint32_t* __restrict__ sd4,
                                uint32_t* __restrict__ sq4, float* __restrict__ sdd)
{
    const int PW = W - 16, PH = H - 16, BW = W - 1;
    const int NP = PW * PH;
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= NP) return;
    int di = p / PW, dj = p - di * PW;
    float sq[64];
    uint32_t s = 0, s2 = 0;
#pragma unroll
    for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            uint32_t v = P[(size_t)(di + 2 * a) * BW + dj + 2 * b];
            s += v;
            s2 += v * v;
            float d = __fmul_rn((float)v, 0.25f);
            sq[a * 8 + b] = __fmul_rn(d, d);  // exact: v^2/16 < 2^24 significant bits
        }
    sd4[p] = s;
    sq4[p] = s2;
    // fully unrolled: the isometry index maps become compile-time constants and sq[] stays in registers
#pragma unroll
    for (int t = 0; t < 8; ++t) {
        float acc = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                int sa, sb;
                fic_iso_bf(t >> 2, t & 3, 8, i, j, sa, sb);
                acc = __fadd_rn(acc, sq[sa * 8 + sb]);
            }
        sdd[(size_t)t * NP + p] = acc;
    }
}

// ---- the scan ---------------------------------------------------------------------------------
// grid = (range groups, chunks).  Each thread walks lattice positions q = chunk*chunk_len + tid + k*256,
// keeps the 64 D4 values of its position in registers and scores the CTA's BF_G ranges x 8 isometries.
template <int MODE>
__device__ __forceinline__ void bf_scan8_body(const uint8_t* __restrict__ img, const uint16_t* __restrict__ P,
                                              const uint32_t* __restrict__ sd4, const uint32_t* __restrict__ sq4,
                                              const float* __restrict__ sdd, const BfGeom& g, int range_begin, int n_ranges,
                                              const int32_t* __restrict__ range_list, int chunk_len, int n_chunks,
                                              BestRec* __restrict__ partial, int grp, int chunk)
{
    __shared__ __align__(16) float rv[BF_G][8][64];  // inverse-transformed range blocks
    __shared__ double r_sum[BF_G], r_sq[BF_G];
    __shared__ BestRec red[BF_G][BF_THREADS / 32];

    const int tid = threadIdx.x;
    const int rw = g.W / 8;
    const int NP = g.PW * (g.H - 16);

    // stage the ranges: rv[gi][t][sa*8+sb] = r[i][j] where (sa,sb) = iso_t(i,j)
    for (int e = tid; e < BF_G * 8 * 64; e += BF_THREADS) {
        int gi = e / 512, t = (e >> 6) & 7, k = e & 63;
        int slot = grp * BF_G + gi;
        float v = 0.f;
        int sa = 0, sb = 0;
        fic_iso_bf(t >> 2, t & 3, 8, k >> 3, k & 7, sa, sb);
        if (slot < n_ranges) {
            int ridx = range_list ? range_list[slot] : range_begin + slot;
            int ri = (ridx / rw) * 8, rj = (ridx % rw) * 8;
            v = (float)img[(size_t)(ri + (k >> 3)) * g.W + rj + (k & 7)];
        }
        rv[gi][t][sa * 8 + sb] = v;
    }
    if (tid < BF_G) {
        int slot = grp * BF_G + tid;
        unsigned long long s = 0, q = 0;
        if (slot < n_ranges) {
            int ridx = range_list ? range_list[slot] : range_begin + slot;
            int ri = (ridx / rw) * 8, rj = (ridx % rw) * 8;
            for (int k = 0; k < 64; ++k) {
                unsigned v = img[(size_t)(ri + (k >> 3)) * g.W + rj + (k & 7)];
                s += v;
                q += MODE == FIC_BF_MODE_JIT ? v * v : ((v * v) & 255u);  // bf_encoder.py:33 wraps in uint8
            }
        }
        r_sum[tid] = (double)s;
        r_sq[tid] = (double)q;
    }
    __syncthreads();

    // per-thread running winners live in shared memory ([range][thread], conflict-free) so the range loop can
    // stay rolled; they are pulled into registers once per (position, range) and written back on change only
    extern __shared__ __align__(16) unsigned char dyn_smem[];
    double* sBestR = reinterpret_cast<double*>(dyn_smem);                           // [BF_G][BF_THREADS]
    uint32_t* sBestSeq = reinterpret_cast<uint32_t*>(sBestR + BF_G * BF_THREADS);   // [BF_G][BF_THREADS]
#pragma unroll
    for (int gi = 0; gi < BF_G; ++gi) {
        sBestR[gi * BF_THREADS + tid] = __longlong_as_double(0x7ff0000000000000ll);
        sBestSeq[gi * BF_THREADS + tid] = 0xffffffffu;
    }

    const long long nq = (long long)g.npi * g.npj;
    const long long q0 = (long long)chunk * chunk_len;
    long long q1 = q0 + chunk_len;
    if (q1 > nq) q1 = nq;

    for (long long q = q0 + tid; q < q1; q += BF_THREADS) {
        int pi = (int)(q / g.npj), pj = (int)(q - (long long)pi * g.npj);
        int di = pi * g.si, dj = pj * g.sj;
        size_t p = (size_t)di * g.PW + dj;
        float D[64];
#pragma unroll
        for (int a = 0; a < 8; ++a)
#pragma unroll
            for (int b = 0; b < 8; ++b) D[a * 8 + b] = (float)P[(size_t)(di + 2 * a) * g.BW + dj + 2 * b];
        const float dsf = __fmul_rn((float)sd4[p], 0.25f);
        const double ds = (double)dsf;
        float sds_t[8];
        double dsds = 0.0, sds_np = 0.0;
        if (MODE == FIC_BF_MODE_JIT) {
#pragma unroll
            for (int t = 0; t < 8; ++t) sds_t[t] = sdd[(size_t)t * NP + p];
            dsds = (double)__fmul_rn(dsf, dsf);
        } else {
            sds_np = DM((double)sq4[p], 0.0625);
        }
#pragma unroll 1
        for (int gi = 0; gi < BF_G; ++gi) {
            const double rs = r_sum[gi], sqr = r_sq[gi];
            double bR = sBestR[gi * BF_THREADS + tid];
            uint32_t bSeq = sBestSeq[gi * BF_THREADS + tid];
            bool changed = false;
#pragma unroll
            for (int t = 0; t < 8; ++t) {
                const float4* rp = reinterpret_cast<const float4*>(&rv[gi][t][0]);
                float x0 = 0.f, x1 = 0.f, x2 = 0.f, x3 = 0.f;
#pragma unroll
                for (int k = 0; k < 16; ++k) {
                    float4 r4 = rp[k];
                    x0 = fmaf(r4.x, D[4 * k + 0], x0);  // integers < 2^24: every partial sum is exact
                    x1 = fmaf(r4.y, D[4 * k + 1], x1);
                    x2 = fmaf(r4.z, D[4 * k + 2], x2);
                    x3 = fmaf(r4.w, D[4 * k + 3], x3);
                }
                const double ab = DM((double)((x0 + x1) + (x2 + x3)), 0.25);
                BfScore sc;
                if (MODE == FIC_BF_MODE_JIT) sc = fic_bfjit_score(64.0, 0.015625, rs, sqr, ds, (double)sds_t[t], dsds, ab);
                else sc = fic_bfnp_score(64.0, 0.015625, rs, sqr, ds, sds_np, ab);
                uint32_t seq = (uint32_t)q * 8u + t;
                if (fic_better(sc.R, seq, bR, bSeq)) {
                    bR = sc.R;
                    bSeq = seq;
                    changed = true;
                }
            }
            if (changed) {
                sBestR[gi * BF_THREADS + tid] = bR;
                sBestSeq[gi * BF_THREADS + tid] = bSeq;
            }
        }
    }

    // CTA reduction per range
    const int lane = tid & 31, warp = tid >> 5;
#pragma unroll 1
    for (int gi = 0; gi < BF_G; ++gi) {
        double R = sBestR[gi * BF_THREADS + tid];
        uint32_t sq_ = sBestSeq[gi * BF_THREADS + tid];
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            double oR = __shfl_xor_sync(0xffffffffu, R, off);
            uint32_t oS = __shfl_xor_sync(0xffffffffu, sq_, off);
            if (fic_better(oR, oS, R, sq_)) { R = oR; sq_ = oS; }
        }
        if (lane == 0) { red[gi][warp].R = R; red[gi][warp].seq = sq_ == 0xffffffffu ? ~0ull : (unsigned long long)sq_; }
    }
    __syncthreads();
    if (tid < BF_G) {
        int slot = grp * BF_G + tid;
        if (slot < n_ranges) {
            BestRec b = red[tid][0];
            for (int w = 1; w < BF_THREADS / 32; ++w)
                if (fic_better(red[tid][w].R, red[tid][w].seq, b.R, b.seq)) b = red[tid][w];
            partial[(size_t)slot * n_chunks + chunk] = b;
        }
    }
}

template <int MODE>
__global__ void __launch_bounds__(BF_THREADS)
bf_scan8_kernel(const uint8_t* __restrict__ img, const uint16_t* __restrict__ P, const uint32_t* __restrict__ sd4,
                const uint32_t* __restrict__ sq4, const float* __restrict__ sdd, BfGeom g, int range_begin, int n_ranges,
                const int32_t* __restrict__ range_list, int chunk_len, int n_chunks, BestRec* __restrict__ partial)
{
    bf_scan8_body<MODE>(img, P, sd4, sq4, sdd, g, range_begin, n_ranges, range_list, chunk_len, n_chunks, partial,
                        blockIdx.x, blockIdx.y);
}

// L2 load of a partial winner: the records may have been written by other SMs of the same launch (bf_scan8_dyn_kernel)
__device__ __forceinline__ BestRec ld_rec_cg(const BestRec* p)
{
    const ulonglong2 v = __ldcg(reinterpret_cast<const ulonglong2*>(p));
    BestRec b;
    b.R = __longlong_as_double((long long)v.x);
    b.seq = v.y;
    return b;
}

// One thread per range: reduce the chunk winners, decode the scan index, re-score the winner for (s, o).
template <int MODE>
__device__ __forceinline__ void bf_finalize8_slot(const uint8_t* __restrict__ img, const uint16_t* __restrict__ P,
                                                  const uint32_t* __restrict__ sd4, const uint32_t* __restrict__ sq4,
                                                  const float* __restrict__ sdd, const BfGeom& g, int range_begin,
                                                  const int32_t* __restrict__ range_list, int n_chunks,
                                                  const BestRec* __restrict__ partial, int32_t* __restrict__ out_idx,
                                                  double* __restrict__ out_so, double* __restrict__ out_R, int slot)
{
    BestRec b = ld_rec_cg(partial + (size_t)slot * n_chunks);
    for (int c = 1; c < n_chunks; ++c) {
        BestRec o = ld_rec_cg(partial + (size_t)slot * n_chunks + c);
        if (fic_better(o.R, o.seq, b.R, b.seq)) b = o;
    }
    int ridx = range_list ? range_list[slot] : range_begin + slot;
    // with a range list the outputs are indexed by absolute range number minus range_begin
    size_t oslot = range_list ? (size_t)(ridx - range_begin) : (size_t)slot;
    if (b.seq == ~0ull) {  // nothing scanned / nothing finite: jit_bfencoder.py:35 keeps the zero tuple
        for (int k = 0; k < 4; ++k) out_idx[4 * oslot + k] = 0;
        out_so[2 * oslot] = 0.0;
        out_so[2 * oslot + 1] = 0.0;
        if (out_R) out_R[oslot] = 0.0;
        return;
    }
    int t = (int)(b.seq & 7ull);
    long long q = (long long)(b.seq >> 3);
    int pi = (int)(q / g.npj), pj = (int)(q - (long long)pi * g.npj);
    int di = pi * g.si, dj = pj * g.sj;
    size_t p = (size_t)di * g.PW + dj;
    const int NP = g.PW * (g.H - 16);
    const int rw = g.W / 8;
    int ri = (ridx / rw) * 8, rj = (ridx % rw) * 8;
    unsigned long long rs = 0, sq = 0;
    long long x = 0;
    for (int i = 0; i < 8; ++i)
        for (int j = 0; j < 8; ++j) {
            unsigned v = img[(size_t)(ri + i) * g.W + rj + j];
            rs += v;
            sq += MODE == FIC_BF_MODE_JIT ? v * v : ((v * v) & 255u);
            int sa, sb;
            fic_iso_bf(t >> 2, t & 3, 8, i, j, sa, sb);
            x += (long long)v * P[(size_t)(di + 2 * sa) * g.BW + dj + 2 * sb];
        }
    const double ds = (double)__fmul_rn((float)sd4[p], 0.25f);
    const double ab = DM((double)x, 0.25);
    BfScore sc;
    if (MODE == FIC_BF_MODE_JIT) {
        float dsf = __fmul_rn((float)sd4[p], 0.25f);
        sc = fic_bfjit_score(64.0, 0.015625, (double)rs, (double)sq, ds, (double)sdd[(size_t)t * NP + p],
                             (double)__fmul_rn(dsf, dsf), ab);
    } else {
        sc = fic_bfnp_score(64.0, 0.015625, (double)rs, (double)sq, ds, DM((double)sq4[p], 0.0625), ab);
    }
    out_idx[4 * oslot + 0] = di;
    out_idx[4 * oslot + 1] = dj;
    out_idx[4 * oslot + 2] = t >> 2;
    out_idx[4 * oslot + 3] = t & 3;
    out_so[2 * oslot] = sc.s;
    out_so[2 * oslot + 1] = sc.o;
    if (out_R) out_R[oslot] = sc.R;
}

template <int MODE>
__global__ void bf_finalize8_kernel(const uint8_t* __restrict__ img, const uint16_t* __restrict__ P,
                                    const uint32_t* __restrict__ sd4, const uint32_t* __restrict__ sq4,
                                    const float* __restrict__ sdd, BfGeom g, int range_begin, int n_ranges,
                                    const int32_t* __restrict__ range_list, int n_chunks, const BestRec* __restrict__ partial,
                                    int32_t* __restrict__ out_idx, double* __restrict__ out_so, double* __restrict__ out_R)
{
    int slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= n_ranges) return;
    bf_finalize8_slot<MODE>(img, P, sd4, sq4, sdd, g, range_begin, range_list, n_chunks, partial, out_idx, out_so, out_R, slot);
}

// ---- the scan with a range count that only the device knows ----------------------------------------------------
// The tensor-core pipeline (bf_screen.cu) leaves the few ranges it cannot decide in a device-side list whose length
// is ctrl[1].  Reading that length on the host would cost a stream synchronisation per encode, so this kernel is
// launched unconditionally with a fixed grid: every CTA derives the (group, chunk) decomposition from the count
// itself -- the formula of bf_exact_scan8 -- walks its share of it, and the CTA that delivers the last chunk of a
// group reduces that group's winners and writes the records (what bf_finalize8_kernel does).  With an empty list the
// launch costs a few microseconds.  Thread 0 of CTA 0 also folds the call's counters into the handle's device totals.
//   done[group]: chunk arrival counters, zero on entry, set back to zero by the finalising CTA.
//   totals[4] (u64): survivors rescored, ranges decided by the screen, ranges re-scanned, fault code.
template <int MODE>
__global__ void __launch_bounds__(BF_THREADS)
bf_scan8_dyn_kernel(const uint8_t* __restrict__ img, const uint16_t* __restrict__ P, const uint32_t* __restrict__ sd4,
                    const uint32_t* __restrict__ sq4, const float* __restrict__ sdd, BfGeom g, int range_begin,
                    const int* __restrict__ ctrl, int n_total, const int32_t* __restrict__ range_list, int want_ctas,
                    BestRec* __restrict__ partial, unsigned int* __restrict__ done, unsigned long long* __restrict__ totals,
                    int32_t* __restrict__ out_idx, double* __restrict__ out_so, double* __restrict__ out_R)
{
    __shared__ int s_last;
    const int n_ranges = ctrl[1];
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        unsigned long long surv;
        memcpy(&surv, ctrl + 8, sizeof(surv));
        atomicAdd(&totals[0], surv);
        atomicAdd(&totals[1], (unsigned long long)(n_total - n_ranges));
        atomicAdd(&totals[2], (unsigned long long)n_ranges);
        if (ctrl[2]) atomicMax(&totals[3], (unsigned long long)ctrl[2]);
    }
    if (n_ranges <= 0) return;
    const long long nq = (long long)g.npi * g.npj;
    const int groups = (n_ranges + BF_G - 1) / BF_G;
    int n_chunks = 1, chunk_len = BF_THREADS;
    if (nq > 0) {   // bf_exact_scan8's decomposition
        n_chunks = (want_ctas + groups - 1) / groups;
        const int max_chunks = (int)((nq + BF_THREADS * 4 - 1) / (BF_THREADS * 4));
        if (n_chunks > max_chunks) n_chunks = max_chunks;
        if (n_chunks < 1) n_chunks = 1;
        chunk_len = (int)((nq + n_chunks - 1) / n_chunks);
        chunk_len = (chunk_len + BF_THREADS - 1) / BF_THREADS * BF_THREADS;
        n_chunks = (int)((nq + chunk_len - 1) / chunk_len);
    }
    const long long items = (long long)groups * n_chunks;
    for (long long item = blockIdx.x; item < items; item += gridDim.x) {
        const int grp = (int)(item / n_chunks), chunk = (int)(item - (long long)grp * n_chunks);
        __syncthreads();   // the previous item's shared data is dead
        bf_scan8_body<MODE>(img, P, sd4, sq4, sdd, g, range_begin, n_ranges, range_list, chunk_len, n_chunks, partial, grp, chunk);
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();   // this CTA's partial records before its arrival
            const unsigned int arrived = atomicAdd(&done[grp], 1u);
            s_last = arrived + 1u == (unsigned)n_chunks;
            if (s_last) { done[grp] = 0u; __threadfence(); }
        }
        __syncthreads();
        if (s_last && threadIdx.x < BF_G) {
            const int slot = grp * BF_G + threadIdx.x;
            if (slot < n_ranges)
                bf_finalize8_slot<MODE>(img, P, sd4, sq4, sdd, g, range_begin, range_list, n_chunks, partial, out_idx, out_so, out_R, slot);
        }
    }
}

// ---- generic range size: sequential replay -----------------------------------------------------
// One CTA per range, threads over lattice positions, the 8 isometries inside.  float32 sequential sums in
// the row-major order of the oriented pooled block (jit) or exact integer sums (numpy).  The pooled block is read from
// the 2x2 cell sums (bf_boxsum_kernel: one load per sample instead of four bytes), the range block sits in shared
// memory as floats, and the isometry -- affine in the pixel index -- is evaluated at three corners per candidate instead
// of once per pixel: 674 -> 224 ms for the 16x16 ranges of a 512x512 image.
template <int MODE>
__global__ void __launch_bounds__(BF_THREADS)
bf_scan_generic_kernel(const uint8_t* __restrict__ img, const uint16_t* __restrict__ P, int H, int W, int R, int si, int sj,
                       int npi, int npj, int range_begin, int32_t* __restrict__ out_idx, double* __restrict__ out_so,
                       double* __restrict__ out_R, const int32_t* __restrict__ need /* optional: only slots with need[slot] != 0 */,
                       int smem_px)
{
    if (need && !need[blockIdx.x]) return;
    extern __shared__ float s_rng[];   // the range block (when it fits)
    __shared__ BestRec red[BF_THREADS / 32];
    __shared__ double red_s[BF_THREADS / 32], red_o[BF_THREADS / 32];
    const int tid = threadIdx.x;
    const int slot = blockIdx.x;
    const int ridx = range_begin + slot;
    const int rw = W / R;
    const int ri = (ridx / rw) * R, rj = (ridx % rw) * R;
    const int nn = R * R;
    const bool in_smem = nn <= smem_px;
    if (in_smem) {
        for (int k = tid; k < nn; k += BF_THREADS) s_rng[k] = (float)img[(size_t)(ri + k / R) * W + rj + k % R];
        __syncthreads();
    }
    const double n = (double)nn, inv_n = DD(1.0, n);
    unsigned long long rs = 0, sq = 0;
    for (int k = 0; k < nn; ++k) {
        unsigned v = in_smem ? (unsigned)s_rng[k] : img[(size_t)(ri + k / R) * W + rj + k % R];
        rs += v;
        sq += MODE == FIC_BF_MODE_JIT ? v * v : ((v * v) & 255u);
    }
    double bR = __longlong_as_double(0x7ff0000000000000ll), bs = 0.0, bo = 0.0;
    unsigned long long bseq = ~0ull;
    const long long nq = (long long)npi * npj;
    const int BW = W - 1;
    for (long long q = tid; q < nq; q += BF_THREADS) {
        int pi = (int)(q / npj), pj = (int)(q - (long long)pi * npj);
        int di = pi * si, dj = pj * sj;
        // the eight isometries side by side: 24 independent float chains (or integer sums) per thread instead of three
        int step_i[8], step_j[8];
        const uint16_t* drow[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            int a00, b00, a10, b10, a01, b01;
            fic_iso_bf(t >> 2, t & 3, R, 0, 0, a00, b00);
            fic_iso_bf(t >> 2, t & 3, R, 1, 0, a10, b10);
            fic_iso_bf(t >> 2, t & 3, R, 0, 1, a01, b01);
            step_i[t] = 2 * ((a10 - a00) * BW + (b10 - b00));
            step_j[t] = 2 * ((a01 - a00) * BW + (b01 - b00));
            drow[t] = P + (size_t)(di + 2 * a00) * BW + dj + 2 * b00;
        }
        float dsf[8], sdsf[8], abf[8];
        long long isd[8], isq[8], iab[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) { dsf[t] = sdsf[t] = abf[t] = 0.f; isd[t] = isq[t] = iab[t] = 0; }
        int k = 0;
        for (int i = 0; i < R; ++i) {
            for (int j = 0; j < R; ++j, ++k) {
                const float rf = in_smem ? s_rng[k] : (float)img[(size_t)(ri + i) * W + rj + j];
#pragma unroll
                for (int t = 0; t < 8; ++t) {
                    const int d4 = drow[t][j * step_j[t]];
                    if (MODE == FIC_BF_MODE_JIT) {
                        const float d = __fmul_rn((float)d4, 0.25f);
                        dsf[t] = __fadd_rn(dsf[t], d);
                        sdsf[t] = __fadd_rn(sdsf[t], __fmul_rn(d, d));
                        abf[t] = __fadd_rn(abf[t], __fmul_rn(d, rf));
                    } else {
                        isd[t] += d4;
                        isq[t] += (long long)d4 * d4;
                        iab[t] += (long long)d4 * (long long)(unsigned)rf;
                    }
                }
            }
#pragma unroll
            for (int t = 0; t < 8; ++t) drow[t] += step_i[t];
        }
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            BfScore sc;
            if (MODE == FIC_BF_MODE_JIT)
                sc = fic_bfjit_score(n, inv_n, (double)rs, (double)sq, (double)dsf[t], (double)sdsf[t],
                                     (double)__fmul_rn(dsf[t], dsf[t]), (double)abf[t]);
            else
                sc = fic_bfnp_score(n, inv_n, (double)rs, (double)sq, DM((double)isd[t], 0.25), DM((double)isq[t], 0.0625),
                                    DM((double)iab[t], 0.25));
            unsigned long long seq = (unsigned long long)q * 8ull + t;
            if (fic_better(sc.R, seq, bR, bseq)) { bR = sc.R; bseq = seq; bs = sc.s; bo = sc.o; }
        }
    }
    const int lane = tid & 31, warp = tid >> 5;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        double oR = __shfl_xor_sync(0xffffffffu, bR, off);
        unsigned long long oS = __shfl_xor_sync(0xffffffffu, bseq, off);
        double os = __shfl_xor_sync(0xffffffffu, bs, off), oo = __shfl_xor_sync(0xffffffffu, bo, off);
        if (fic_better(oR, oS, bR, bseq)) { bR = oR; bseq = oS; bs = os; bo = oo; }
    }
    if (lane == 0) { red[warp].R = bR; red[warp].seq = bseq; red_s[warp] = bs; red_o[warp] = bo; }
    __syncthreads();
    if (tid == 0) {
        BestRec b = red[0];
        double s = red_s[0], o = red_o[0];
        for (int w = 1; w < BF_THREADS / 32; ++w)
            if (fic_better(red[w].R, red[w].seq, b.R, b.seq)) { b = red[w]; s = red_s[w]; o = red_o[w]; }
        if (b.seq == ~0ull) {
            for (int k = 0; k < 4; ++k) out_idx[4 * (size_t)slot + k] = 0;
            out_so[2 * (size_t)slot] = 0.0; out_so[2 * (size_t)slot + 1] = 0.0;
            if (out_R) out_R[slot] = 0.0;
        } else {
            int t = (int)(b.seq & 7ull);
            long long q = (long long)(b.seq >> 3);
            int pi = (int)(q / npj), pj = (int)(q - (long long)pi * npj);
            out_idx[4 * (size_t)slot + 0] = pi * si;
            out_idx[4 * (size_t)slot + 1] = pj * sj;
            out_idx[4 * (size_t)slot + 2] = t >> 2;
            out_idx[4 * (size_t)slot + 3] = t & 3;
            out_so[2 * (size_t)slot] = s;
            out_so[2 * (size_t)slot + 1] = o;
            if (out_R) out_R[slot] = b.R;
        }
    }
}

// ---- generic range size, jit arithmetic, four ranges per CTA ----------------------------------------------------------
// Two of the three sums of a candidate -- sum d and sum d^2 in the order of its isometry -- do not depend on the range:
// bf_gen_stats_kernel computes them once per (position, isometry).  The scan then gives a CTA four consecutive ranges: a
// thread walks the pooled block of its position once per isometry for all four (one load and one conversion per sample
// instead of four, one float4 of range pixels from shared memory), keeping 32 sum(d r) chains, each in the reference's
// order and rounding.  Winners are kept as (score, scan index) per thread in shared memory and the winning candidate of
// a range is scored once more at the end for (s, o).  Per (range, isometry, pixel): 2.8 instructions instead of 9.
#define BF_GEN_G 4
// (NumPy semantics: the sums are integers and do not depend on the isometry -- isd_tab / isq_tab [NP] instead)
template <int MODE>
__global__ void bf_gen_stats_kernel(const uint16_t* __restrict__ P, int W, int R, int si, int sj, int npj, int NP,
                                    float* __restrict__ ds_tab /*[8][NP]*/, float* __restrict__ sds_tab /*[8][NP]*/,
                                    long long* __restrict__ isd_tab /*[NP]*/, long long* __restrict__ isq_tab /*[NP]*/)
{
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= NP) return;
    const int BW = W - 1;
    const int di = (q / npj) * si, dj = (q % npj) * sj;
    if (MODE == FIC_BF_MODE_NUMPY) {
        long long sd = 0, sq = 0;
        for (int i = 0; i < R; ++i)
            for (int j = 0; j < R; ++j) {
                const long long d4 = P[(size_t)(di + 2 * i) * BW + dj + 2 * j];
                sd += d4;
                sq += d4 * d4;
            }
        isd_tab[q] = sd;
        isq_tab[q] = sq;
        return;
    }
    int step_i[8], step_j[8];
    const uint16_t* drow[8];
#pragma unroll
    for (int t = 0; t < 8; ++t) {
        int a00, b00, a10, b10, a01, b01;
        fic_iso_bf(t >> 2, t & 3, R, 0, 0, a00, b00);
        fic_iso_bf(t >> 2, t & 3, R, 1, 0, a10, b10);
        fic_iso_bf(t >> 2, t & 3, R, 0, 1, a01, b01);
        step_i[t] = 2 * ((a10 - a00) * BW + (b10 - b00));
        step_j[t] = 2 * ((a01 - a00) * BW + (b01 - b00));
        drow[t] = P + (size_t)(di + 2 * a00) * BW + dj + 2 * b00;
    }
    float dsf[8], sdsf[8];
#pragma unroll
    for (int t = 0; t < 8; ++t) dsf[t] = sdsf[t] = 0.f;
    for (int i = 0; i < R; ++i) {
        for (int j = 0; j < R; ++j) {
#pragma unroll
            for (int t = 0; t < 8; ++t) {
                const float d = __fmul_rn((float)drow[t][j * step_j[t]], 0.25f);
                dsf[t] = __fadd_rn(dsf[t], d);
                sdsf[t] = __fadd_rn(sdsf[t], __fmul_rn(d, d));
            }
        }
#pragma unroll
        for (int t = 0; t < 8; ++t) drow[t] += step_i[t];
    }
#pragma unroll
    for (int t = 0; t < 8; ++t) {
        ds_tab[(size_t)t * NP + q] = dsf[t];
        sds_tab[(size_t)t * NP + q] = sdsf[t];
    }
}

template <int MODE>
__global__ void __launch_bounds__(BF_THREADS)
bf_scan_generic4_kernel(const uint8_t* __restrict__ img, const uint16_t* __restrict__ P, int H, int W, int R, int si, int sj,
                        int npi, int npj, int range_begin, int n_ranges, const float* __restrict__ ds_tab,
                        const float* __restrict__ sds_tab, const long long* __restrict__ isd_tab,
                        const long long* __restrict__ isq_tab, BestRec* __restrict__ partial /*[n_ranges][gridDim.y]*/)
{
    extern __shared__ __align__(16) unsigned char s_dyn4[];
    const int nn = R * R;
    float4* s_rng = reinterpret_cast<float4*>(s_dyn4);                                   // [nn]: pixel k of the four ranges
    double* s_bR = reinterpret_cast<double*>(s_dyn4 + (size_t)nn * sizeof(float4));       // [BF_GEN_G][BF_THREADS]
    unsigned long long* s_bseq = reinterpret_cast<unsigned long long*>(s_bR + BF_GEN_G * BF_THREADS);
    __shared__ BestRec red[BF_GEN_G][BF_THREADS / 32];
    __shared__ double s_rs[BF_GEN_G], s_sq[BF_GEN_G];
    const int tid = threadIdx.x;
    const int slot0 = blockIdx.x * BF_GEN_G;
    const int g_real = n_ranges - slot0 < BF_GEN_G ? n_ranges - slot0 : BF_GEN_G;
    const int rw = W / R;
    int ri[BF_GEN_G], rj[BF_GEN_G];
#pragma unroll
    for (int g = 0; g < BF_GEN_G; ++g) {
        const int ridx = range_begin + slot0 + (g < g_real ? g : 0);
        ri[g] = (ridx / rw) * R; rj[g] = (ridx % rw) * R;
    }
    for (int k = tid; k < nn; k += BF_THREADS) {
        const int i = k / R, j = k % R;
        s_rng[k] = make_float4((float)img[(size_t)(ri[0] + i) * W + rj[0] + j], (float)img[(size_t)(ri[1] + i) * W + rj[1] + j],
                               (float)img[(size_t)(ri[2] + i) * W + rj[2] + j], (float)img[(size_t)(ri[3] + i) * W + rj[3] + j]);
    }
#pragma unroll
    for (int g = 0; g < BF_GEN_G; ++g) { s_bR[g * BF_THREADS + tid] = __longlong_as_double(0x7ff0000000000000ll); s_bseq[g * BF_THREADS + tid] = ~0ull; }
    __syncthreads();
    if (tid < BF_GEN_G) {
        unsigned long long rs = 0, sq = 0;
        for (int k = 0; k < nn; ++k) {
            const float4 v4 = s_rng[k];
            const unsigned v = (unsigned)(tid == 0 ? v4.x : tid == 1 ? v4.y : tid == 2 ? v4.z : v4.w);
            rs += v;
            sq += MODE == FIC_BF_MODE_JIT ? v * v : ((v * v) & 255u);      // bf_encoder.py:33 squares in uint8
        }
        s_rs[tid] = (double)rs;
        s_sq[tid] = (double)sq;
    }
    __syncthreads();
    const double n = (double)nn, inv_n = DD(1.0, n);
    const int NP = npi * npj;
    const int BW = W - 1;
    // the positions are cut into gridDim.y contiguous chunks, one CTA each: few ranges still fill the GPU
    const int per = (NP + (int)gridDim.y - 1) / (int)gridDim.y;
    const int q0 = per * (int)blockIdx.y, q1 = q0 + per < NP ? q0 + per : NP;
    for (int q = q0 + tid; q < q1; q += BF_THREADS) {
        const int di = (q / npj) * si, dj = (q % npj) * sj;
        int step_i[8], step_j[8];
        const uint16_t* drow[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            int a00, b00, a10, b10, a01, b01;
            fic_iso_bf(t >> 2, t & 3, R, 0, 0, a00, b00);
            fic_iso_bf(t >> 2, t & 3, R, 1, 0, a10, b10);
            fic_iso_bf(t >> 2, t & 3, R, 0, 1, a01, b01);
            step_i[t] = 2 * ((a10 - a00) * BW + (b10 - b00));
            step_j[t] = 2 * ((a01 - a00) * BW + (b01 - b00));
            drow[t] = P + (size_t)(di + 2 * a00) * BW + dj + 2 * b00;
        }
        float ab[BF_GEN_G][8];
        int iab[BF_GEN_G][8];        // NumPy semantics: integer sums (nn <= 1024: below 2^31)
#pragma unroll
        for (int g = 0; g < BF_GEN_G; ++g)
#pragma unroll
            for (int t = 0; t < 8; ++t) { ab[g][t] = 0.f; iab[g][t] = 0; }
        int k = 0;
        for (int i = 0; i < R; ++i) {
            for (int j = 0; j < R; ++j, ++k) {
                const float4 r4 = s_rng[k];
#pragma unroll
                for (int t = 0; t < 8; ++t) {
                    if (MODE == FIC_BF_MODE_JIT) {
                        const float d = __fmul_rn((float)drow[t][j * step_j[t]], 0.25f);
                        ab[0][t] = __fadd_rn(ab[0][t], __fmul_rn(d, r4.x));
                        ab[1][t] = __fadd_rn(ab[1][t], __fmul_rn(d, r4.y));
                        ab[2][t] = __fadd_rn(ab[2][t], __fmul_rn(d, r4.z));
                        ab[3][t] = __fadd_rn(ab[3][t], __fmul_rn(d, r4.w));
                    } else {
                        const int d4 = drow[t][j * step_j[t]];
                        iab[0][t] += d4 * (int)r4.x;
                        iab[1][t] += d4 * (int)r4.y;
                        iab[2][t] += d4 * (int)r4.z;
                        iab[3][t] += d4 * (int)r4.w;
                    }
                }
            }
#pragma unroll
            for (int t = 0; t < 8; ++t) drow[t] += step_i[t];
        }
        // one copy of the scorer: the 32 candidates of this position in turn (scan order inside a range: isometry)
#pragma unroll 1
        for (int idx = 0; idx < BF_GEN_G * 8; ++idx) {
            const int g = idx >> 3, t = idx & 7;
            if (g >= g_real) break;
            float abv = 0.f;
            int iabv = 0;
#pragma unroll
            for (int gg = 0; gg < BF_GEN_G; ++gg)
#pragma unroll
                for (int tt = 0; tt < 8; ++tt)
                    if (idx == gg * 8 + tt) { abv = ab[gg][tt]; iabv = iab[gg][tt]; }
            BfScore sc;
            if (MODE == FIC_BF_MODE_JIT) {
                const float dsf = ds_tab[(size_t)t * NP + q], sdsf = sds_tab[(size_t)t * NP + q];
                sc = fic_bfjit_score(n, inv_n, s_rs[g], s_sq[g], (double)dsf, (double)sdsf, (double)__fmul_rn(dsf, dsf), (double)abv);
            } else {
                sc = fic_bfnp_score(n, inv_n, s_rs[g], s_sq[g], DM((double)isd_tab[q], 0.25), DM((double)isq_tab[q], 0.0625),
                                    DM((double)iabv, 0.25));
            }
            const unsigned long long seq = (unsigned long long)q * 8ull + (unsigned)t;
            if (fic_better(sc.R, seq, s_bR[g * BF_THREADS + tid], s_bseq[g * BF_THREADS + tid])) {
                s_bR[g * BF_THREADS + tid] = sc.R;
                s_bseq[g * BF_THREADS + tid] = seq;
            }
        }
    }
    const int lane = tid & 31, warp = tid >> 5;
#pragma unroll
    for (int g = 0; g < BF_GEN_G; ++g) {
        double bR = s_bR[g * BF_THREADS + tid];
        unsigned long long bseq = s_bseq[g * BF_THREADS + tid];
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            const double oR = __shfl_xor_sync(0xffffffffu, bR, off);
            const unsigned long long oS = __shfl_xor_sync(0xffffffffu, bseq, off);
            if (fic_better(oR, oS, bR, bseq)) { bR = oR; bseq = oS; }
        }
        if (lane == 0) { red[g][warp].R = bR; red[g][warp].seq = bseq; }
    }
    __syncthreads();
    if (tid < g_real) {
        const int g = tid;
        BestRec b = red[g][0];
        for (int w = 1; w < BF_THREADS / 32; ++w)
            if (fic_better(red[g][w].R, red[g][w].seq, b.R, b.seq)) b = red[g][w];
        partial[(size_t)(slot0 + g) * gridDim.y + blockIdx.y] = b;
    }
}

// merges the chunk winners of a range (first strict minimum in scan order) and scores the winner once more for its (s, o):
// the same sums in the same order
template <int MODE>
__global__ void bf_gen_finalize_kernel(const uint8_t* __restrict__ img, const uint16_t* __restrict__ P, int W, int R, int si, int sj,
                                       int npj, int NP, int range_begin, int n_ranges, int n_chunks,
                                       const BestRec* __restrict__ partial, const float* __restrict__ ds_tab,
                                       const float* __restrict__ sds_tab, const long long* __restrict__ isd_tab,
                                       const long long* __restrict__ isq_tab, int32_t* __restrict__ out_idx,
                                       double* __restrict__ out_so, double* __restrict__ out_R)
{
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= n_ranges) return;
    BestRec b = partial[(size_t)slot * n_chunks];
    for (int c = 1; c < n_chunks; ++c) {
        const BestRec o = partial[(size_t)slot * n_chunks + c];
        if (fic_better(o.R, o.seq, b.R, b.seq)) b = o;
    }
    if (b.seq == ~0ull) {
        for (int k = 0; k < 4; ++k) out_idx[4 * (size_t)slot + k] = 0;
        out_so[2 * (size_t)slot] = 0.0; out_so[2 * (size_t)slot + 1] = 0.0;
        if (out_R) out_R[slot] = 0.0;
        return;
    }
    const int nn = R * R, BW = W - 1, rw = W / R;
    const int ridx = range_begin + slot;
    const int ri = (ridx / rw) * R, rj = (ridx % rw) * R;
    unsigned long long rs = 0, sq = 0;
    for (int k = 0; k < nn; ++k) {
        const unsigned v = img[(size_t)(ri + k / R) * W + rj + k % R];
        rs += v;
        sq += MODE == FIC_BF_MODE_JIT ? v * v : ((v * v) & 255u);
    }
    const int t = (int)(b.seq & 7ull), q = (int)(b.seq >> 3);
    const int di = (q / npj) * si, dj = (q % npj) * sj;
    float abf = 0.f;
    long long iab = 0;
    for (int i = 0; i < R; ++i)
        for (int j = 0; j < R; ++j) {
            int sa, sb;
            fic_iso_bf(t >> 2, t & 3, R, i, j, sa, sb);
            const int d4 = P[(size_t)(di + 2 * sa) * BW + dj + 2 * sb];
            const int rv = img[(size_t)(ri + i) * W + rj + j];
            abf = __fadd_rn(abf, __fmul_rn(__fmul_rn((float)d4, 0.25f), (float)rv));
            iab += (long long)d4 * rv;
        }
    const double n = (double)nn, inv_n = DD(1.0, n);
    BfScore sc;
    if (MODE == FIC_BF_MODE_JIT) {
        const float dsf = ds_tab[(size_t)t * NP + q], sdsf = sds_tab[(size_t)t * NP + q];
        sc = fic_bfjit_score(n, inv_n, (double)rs, (double)sq, (double)dsf, (double)sdsf, (double)__fmul_rn(dsf, dsf), (double)abf);
    } else {
        sc = fic_bfnp_score(n, inv_n, (double)rs, (double)sq, DM((double)isd_tab[q], 0.25), DM((double)isq_tab[q], 0.0625), DM((double)iab, 0.25));
    }
    out_idx[4 * (size_t)slot + 0] = di;
    out_idx[4 * (size_t)slot + 1] = dj;
    out_idx[4 * (size_t)slot + 2] = t >> 2;
    out_idx[4 * (size_t)slot + 3] = t & 3;
    out_so[2 * (size_t)slot] = sc.s;
    out_so[2 * (size_t)slot + 1] = sc.o;
    if (out_R) out_R[slot] = sc.R;
}

// ---- host side -------------------------------------------------------------------------------
static inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

// Builds boxsum + stats for an 8x8 encode (shared with the screen path).
// with_stats = false: the caller's own pool kernel fills sd4 / sq4 / sdd (bf_screen.cu), only the buffers are reserved
int bf_prepare8(fic_handle* h, const uint8_t* d_img, int H, int W, bool with_stats, cudaStream_t st)
{
    const int PW = W - 16, PH = H - 16;
    FIC_TRY(h->bf_boxsum.reserve((size_t)(H - 1) * (W - 1) * sizeof(uint16_t)));
    dim3 gb(ceil_div(W - 1, 256), H - 1);
    bf_boxsum_kernel<<<gb, 256, 0, st>>>(d_img, H, W, h->bf_boxsum.as<uint16_t>());
    FIC_CHECK_LAUNCH(h, "bf_boxsum_kernel");
    if (PW > 0 && PH > 0) {
        size_t NP = (size_t)PW * PH;
        FIC_TRY(h->bf_stats.reserve(NP * (2 * sizeof(uint32_t) + 8 * sizeof(float))));
        uint32_t* sd4 = h->bf_stats.as<uint32_t>();
        uint32_t* sq4 = sd4 + NP;
        float* sdd = reinterpret_cast<float*>(sq4 + NP);
        if (with_stats) {
            bf_stats_kernel<<<ceil_div(NP, 128), 128, 0, st>>>(h->bf_boxsum.as<uint16_t>(), H, W, sd4, sq4, sdd);
            FIC_CHECK_LAUNCH(h, "bf_stats_kernel");
        }
    }
    return FIC_OK;
}

// Exact scan of `n_ranges` ranges (contiguous from range_begin, or the explicit device list) over the lattice
// with strides (si,sj).  Outputs are written at index (range - range_begin).  Requires bf_prepare8.
int bf_exact_scan8(fic_handle* h, const uint8_t* d_img, int H, int W, int si, int sj, int mode, int range_begin,
                   int n_ranges, const int32_t* d_range_list, int32_t* d_idx, double* d_so, double* d_R,
                   cudaStream_t st)
{
    if (n_ranges <= 0) return FIC_OK;
    BfGeom g;
    g.H = H; g.W = W; g.si = si; g.sj = sj; g.PW = W - 16; g.BW = W - 1;
    g.npi = H - 16 > 0 ? ceil_div(H - 16, si) : 0;
    g.npj = W - 16 > 0 ? ceil_div(W - 16, sj) : 0;
    const long long nq = (long long)g.npi * g.npj;
    const int groups = ceil_div(n_ranges, BF_G);
    // enough CTAs for ~8 waves of 148 SMs x 2, but never less than 4 positions per thread
    int n_chunks = 1;
    if (nq > 0) {
        long long want = (long long)h->num_sms * 16;
        n_chunks = (int)((want + groups - 1) / groups);
        int max_chunks = ceil_div(nq, BF_THREADS * 4);
        if (n_chunks > max_chunks) n_chunks = max_chunks;
        if (n_chunks < 1) n_chunks = 1;
        if (n_chunks > 65535) n_chunks = 65535;
    }
    int chunk_len = nq > 0 ? ceil_div(nq, n_chunks) : 1;
    chunk_len = ceil_div(chunk_len, BF_THREADS) * BF_THREADS;
    n_chunks = nq > 0 ? ceil_div(nq, chunk_len) : 1;
    FIC_TRY(h->bf_partial.reserve((size_t)n_ranges * n_chunks * sizeof(BestRec)));
    const size_t NP = (size_t)(W - 16 > 0 ? W - 16 : 0) * (size_t)(H - 16 > 0 ? H - 16 : 0);
    const uint16_t* P = h->bf_boxsum.as<uint16_t>();
    const uint32_t* sd4 = h->bf_stats.as<uint32_t>();
    const uint32_t* sq4 = sd4 + NP;
    const float* sdd = reinterpret_cast<const float*>(sq4 + NP);
    BestRec* partial = h->bf_partial.as<BestRec>();
    dim3 grid(groups, n_chunks);
    if (mode == FIC_BF_MODE_JIT) {
        {
            ProfScope prof(h, st);
            bf_scan8_kernel<FIC_BF_MODE_JIT><<<grid, BF_THREADS, BF_SCAN_DYN_SMEM, st>>>(d_img, P, sd4, sq4, sdd, g, range_begin, n_ranges,
                                                                      d_range_list, chunk_len, n_chunks, partial);
        }
        FIC_CHECK_LAUNCH(h, "bf_scan8_kernel<jit>");
        bf_finalize8_kernel<FIC_BF_MODE_JIT><<<ceil_div(n_ranges, 128), 128, 0, st>>>(
            d_img, P, sd4, sq4, sdd, g, range_begin, n_ranges, d_range_list, n_chunks, partial, d_idx, d_so, d_R);
        FIC_CHECK_LAUNCH(h, "bf_finalize8_kernel<jit>");
    } else {
        bf_scan8_kernel<FIC_BF_MODE_NUMPY><<<grid, BF_THREADS, BF_SCAN_DYN_SMEM, st>>>(d_img, P, sd4, sq4, sdd, g, range_begin, n_ranges,
                                                                        d_range_list, chunk_len, n_chunks, partial);
        FIC_CHECK_LAUNCH(h, "bf_scan8_kernel<numpy>");
        bf_finalize8_kernel<FIC_BF_MODE_NUMPY><<<ceil_div(n_ranges, 128), 128, 0, st>>>(
            d_img, P, sd4, sq4, sdd, g, range_begin, n_ranges, d_range_list, n_chunks, partial, d_idx, d_so, d_R);
        FIC_CHECK_LAUNCH(h, "bf_finalize8_kernel<numpy>");
    }
    return FIC_OK;
}

// The scan of the ranges listed in d_range_list[0 .. ctrl[1]) -- a count only the device knows (see
// bf_scan8_dyn_kernel).  n_total = ranges of the whole call (for the counters).  No host synchronisation.
int bf_exact_scan8_dyn(fic_handle* h, const uint8_t* d_img, int H, int W, int mode, int range_begin, int n_total,
                       const int* d_ctrl, const int32_t* d_range_list, unsigned long long* d_totals, int32_t* d_idx,
                       double* d_so, double* d_R, cudaStream_t st)
{
    BfGeom g;
    g.H = H; g.W = W; g.si = 1; g.sj = 1; g.PW = W - 16; g.BW = W - 1;
    g.npi = H - 16; g.npj = W - 16;
    const int want = h->num_sms * 16;
    // groups * n_chunks <= want + groups, eight ranges per group
    FIC_TRY(h->bf_partial.reserve(((size_t)want + (size_t)n_total / BF_G + 2) * BF_G * sizeof(BestRec)));
    const size_t done_bytes = ((size_t)n_total / BF_G + 2) * sizeof(unsigned int);
    if (done_bytes > h->bf_done.cap) {
        FIC_TRY(h->bf_done.reserve(done_bytes));
        FIC_CUDA(cudaMemsetAsync(h->bf_done.p, 0, h->bf_done.cap, st));   // the kernel leaves the counters at zero
    }
    const size_t NP = (size_t)g.npi * g.npj;
    const uint16_t* P = h->bf_boxsum.as<uint16_t>();
    const uint32_t* sd4 = h->bf_stats.as<uint32_t>();
    const uint32_t* sq4 = sd4 + NP;
    const float* sdd = reinterpret_cast<const float*>(sq4 + NP);
    const int grid = h->num_sms * 4;
    if (mode == FIC_BF_MODE_JIT)
        bf_scan8_dyn_kernel<FIC_BF_MODE_JIT><<<grid, BF_THREADS, BF_SCAN_DYN_SMEM, st>>>(
            d_img, P, sd4, sq4, sdd, g, range_begin, d_ctrl, n_total, d_range_list, want, h->bf_partial.as<BestRec>(),
            h->bf_done.as<unsigned int>(), d_totals, d_idx, d_so, d_R);
    else
        bf_scan8_dyn_kernel<FIC_BF_MODE_NUMPY><<<grid, BF_THREADS, BF_SCAN_DYN_SMEM, st>>>(
            d_img, P, sd4, sq4, sdd, g, range_begin, d_ctrl, n_total, d_range_list, want, h->bf_partial.as<BestRec>(),
            h->bf_done.as<unsigned int>(), d_totals, d_idx, d_so, d_R);
    FIC_CHECK_LAUNCH(h, "bf_scan8_dyn_kernel");
    return FIC_OK;
}

// the scan for `n_ranges` ranges from range_begin; with d_need, only for the flagged ones (the 4x4 tensor-core path hands
// back what it cannot decide)
int bf_exact_generic_listed(fic_handle* h, const uint8_t* d_img, int H, int W, int range, int si, int sj, int mode,
                            int range_begin, int n_ranges, const int32_t* d_need, int32_t* d_idx, double* d_so, double* d_R,
                            cudaStream_t st)
{
    if (n_ranges <= 0) return FIC_OK;
    const int D = 2 * range;
    int npi = H - D > 0 ? ceil_div(H - D, si) : 0, npj = W - D > 0 ? ceil_div(W - D, sj) : 0;
    FIC_TRY(h->bf_boxsum.reserve((size_t)(H > 1 ? H - 1 : 1) * (W > 1 ? W - 1 : 1) * sizeof(uint16_t)));
    if (H > 1 && W > 1) {
        dim3 gb(ceil_div(W - 1, 256), H - 1);
        bf_boxsum_kernel<<<gb, 256, 0, st>>>(d_img, H, W, h->bf_boxsum.as<uint16_t>());
        FIC_CHECK_LAUNCH(h, "bf_boxsum_kernel");
    }
    const int nn = range * range;
    const long long NPq = (long long)npi * npj;
    if (!d_need && NPq > 0 && NPq < (1ll << 28) && nn <= 1024 && H > 1 && W > 1) {
        // four ranges per CTA over per-position tables of sum d / sum d^2 (see bf_scan_generic4_kernel)
        FIC_TRY(h->bf_stats.reserve((size_t)NPq * 16 * sizeof(float) + 256));
        float* ds_tab = h->bf_stats.as<float>();
        float* sds_tab = ds_tab + (size_t)8 * NPq;
        long long* isd_tab = h->bf_stats.as<long long>();      // (NumPy semantics: two integer tables in the same buffer)
        long long* isq_tab = isd_tab + NPq;
        const uint16_t* P = h->bf_boxsum.as<uint16_t>();
        const size_t smem = (size_t)nn * sizeof(float4) + BF_GEN_G * BF_THREADS * (sizeof(double) + sizeof(unsigned long long));
        const int n_cta = ceil_div(n_ranges, BF_GEN_G);
        int n_chunks = (8 * h->num_sms + n_cta - 1) / n_cta;       // about eight CTAs per SM in all
        const int max_chunks = (int)std::max<long long>(1, NPq / BF_THREADS);
        if (n_chunks > max_chunks) n_chunks = max_chunks;
        if (n_chunks < 1) n_chunks = 1;
        FIC_TRY(h->bf_partial.reserve((size_t)n_ranges * n_chunks * sizeof(BestRec) + 64));
        BestRec* partial = h->bf_partial.as<BestRec>();
        if (mode == FIC_BF_MODE_JIT) {
            bf_gen_stats_kernel<FIC_BF_MODE_JIT><<<ceil_div(NPq, 128), 128, 0, st>>>(P, W, range, si, sj, npj, (int)NPq, ds_tab, sds_tab, isd_tab, isq_tab);
            FIC_CHECK_LAUNCH(h, "bf_gen_stats_kernel");
            bf_scan_generic4_kernel<FIC_BF_MODE_JIT><<<dim3(n_cta, n_chunks), BF_THREADS, smem, st>>>(
                d_img, P, H, W, range, si, sj, npi, npj, range_begin, n_ranges, ds_tab, sds_tab, isd_tab, isq_tab, partial);
            FIC_CHECK_LAUNCH(h, "bf_scan_generic4_kernel");
            bf_gen_finalize_kernel<FIC_BF_MODE_JIT><<<ceil_div(n_ranges, 128), 128, 0, st>>>(
                d_img, P, W, range, si, sj, npj, (int)NPq, range_begin, n_ranges, n_chunks, partial, ds_tab, sds_tab, isd_tab, isq_tab, d_idx, d_so, d_R);
        } else {
            bf_gen_stats_kernel<FIC_BF_MODE_NUMPY><<<ceil_div(NPq, 128), 128, 0, st>>>(P, W, range, si, sj, npj, (int)NPq, ds_tab, sds_tab, isd_tab, isq_tab);
            FIC_CHECK_LAUNCH(h, "bf_gen_stats_kernel");
            bf_scan_generic4_kernel<FIC_BF_MODE_NUMPY><<<dim3(n_cta, n_chunks), BF_THREADS, smem, st>>>(
                d_img, P, H, W, range, si, sj, npi, npj, range_begin, n_ranges, ds_tab, sds_tab, isd_tab, isq_tab, partial);
            FIC_CHECK_LAUNCH(h, "bf_scan_generic4_kernel");
            bf_gen_finalize_kernel<FIC_BF_MODE_NUMPY><<<ceil_div(n_ranges, 128), 128, 0, st>>>(
                d_img, P, W, range, si, sj, npj, (int)NPq, range_begin, n_ranges, n_chunks, partial, ds_tab, sds_tab, isd_tab, isq_tab, d_idx, d_so, d_R);
        }
        FIC_CHECK_LAUNCH(h, "bf_gen_finalize_kernel");
        return FIC_OK;
    }
    const int smem_px = nn <= 10240 ? nn : 0;      // range block in shared memory when it fits 40 KB
    if (mode == FIC_BF_MODE_JIT)
        bf_scan_generic_kernel<FIC_BF_MODE_JIT><<<n_ranges, BF_THREADS, smem_px * sizeof(float), st>>>(
            d_img, h->bf_boxsum.as<uint16_t>(), H, W, range, si, sj, npi, npj, range_begin, d_idx, d_so, d_R, d_need, smem_px);
    else
        bf_scan_generic_kernel<FIC_BF_MODE_NUMPY><<<n_ranges, BF_THREADS, smem_px * sizeof(float), st>>>(
            d_img, h->bf_boxsum.as<uint16_t>(), H, W, range, si, sj, npi, npj, range_begin, d_idx, d_so, d_R, d_need, smem_px);
    FIC_CHECK_LAUNCH(h, "bf_scan_generic_kernel");
    return FIC_OK;
}

int bf_exact_generic(fic_handle* h, const uint8_t* d_img, int H, int W, int range, int si, int sj, int mode,
                     int range_begin, int n_ranges, int32_t* d_idx, double* d_so, double* d_R, cudaStream_t st)
{
    return bf_exact_generic_listed(h, d_img, H, W, range, si, sj, mode, range_begin, n_ranges, nullptr, d_idx, d_so, d_R, st);
}

// ---- fic_bf_encode_static: the exact 8x8 scan with no allocation at all ------------------------
// Some hosts of this library may not allocate device memory when they call it: a codec plugin running inside an
// application that handed the whole GPU memory budget to its own pool allocator, a capture or inference server
// whose steady-state loop must not touch the driver's allocator (cudaMalloc / cudaFree synchronise the device and
// can fail under fragmentation), or a sandboxed worker that is only given pre-allocated buffers.  For them this
// entry point runs the same four kernels as the exact scan above (box sums, per-position statistics, scan,
// finalize) but takes every byte of scratch from one array that is part of the loaded module.
//
// Limits and rules (also in include/fic.h):
//   * image at most FIC_STATIC_MAX_SIDE x FIC_STATIC_MAX_SIDE pixels (8x8 ranges, 16x16 domains); larger images
//     are rejected with FIC_EINVAL -- nothing is truncated;
//   * the scratch costs FIC_STATIC_SCRATCH_BYTES of device memory on every GPU that loads the library, from load
//     until process exit, whether or not this entry point is ever called;
//   * all calls on one GPU share that array.  Nothing in it has to start at zero: every element a kernel reads was
//     written earlier in the same call (box sums and statistics cover exactly the positions the scan visits, and
//     every (range, chunk) slot of the partial-winner table is written by the CTA that owns it), so the previous
//     call's leftovers are harmless.  But two calls that overlap in time on the same GPU (different streams or
//     host threads) would overwrite each other's scratch: run them one after the other;
//   * the first call on each device looks up the array's device address.  That lookup may initialise the CUDA
//     context / load the module, so make the first call before starting any stream capture, and make it from one
//     host thread: the per-device slot below is filled without a lock.
#define FIC_STATIC_MAX_SIDE 256
#define FIC_STATIC_BOX_BYTES 131072u                                   /* u16 255*255 rounded up   */
#define FIC_STATIC_NP_MAX ((FIC_STATIC_MAX_SIDE - 16) * (FIC_STATIC_MAX_SIDE - 16))
#define FIC_STATIC_STATS_BYTES (FIC_STATIC_NP_MAX * 40u)               /* sd4 + sq4 + 8 x sdd      */
#define FIC_STATIC_PARTIAL_RECS 65536u                                 /* BestRec, 16 bytes each   */
#define FIC_STATIC_SCRATCH_BYTES (FIC_STATIC_BOX_BYTES + FIC_STATIC_STATS_BYTES + FIC_STATIC_PARTIAL_RECS * 16u)

static __device__ __align__(256) unsigned char g_static_scratch[FIC_STATIC_SCRATCH_BYTES];

// Host-side bookkeeping, one slot per device ordinal: the scratch's DEVICE address (the symbol name itself is only
// a host-side handle and must never be handed to a kernel) and the SM count used to size the grid.
#define FIC_STATIC_MAX_DEVICES 64
static unsigned char* g_static_ptr[FIC_STATIC_MAX_DEVICES];
static int g_static_sms[FIC_STATIC_MAX_DEVICES];

extern "C" int fic_bf_encode_static(const uint8_t* d_img, int H, int W, int stride_i, int stride_j, int mode,
                                    int range_begin, int range_end, int32_t* d_idx, double* d_so, double* d_R,
                                    void* stream)
{
    FIC_REQUIRE(d_img && d_idx && d_so && d_R, "fic_bf_encode_static: NULL argument (all buffers are device pointers)");
    FIC_REQUIRE(H > 0 && W > 0 && H % 8 == 0 && W % 8 == 0, "fic_bf_encode_static: image %dx%d is not a multiple of 8", H, W);
    FIC_REQUIRE(H <= FIC_STATIC_MAX_SIDE && W <= FIC_STATIC_MAX_SIDE,
                "fic_bf_encode_static: image %dx%d exceeds the built-in scratch (at most %dx%d); use fic_bf_encode",
                H, W, FIC_STATIC_MAX_SIDE, FIC_STATIC_MAX_SIDE);
    FIC_REQUIRE(stride_i > 0 && stride_j > 0, "fic_bf_encode_static: strides must be positive");
    FIC_REQUIRE(mode == FIC_BF_MODE_NUMPY || mode == FIC_BF_MODE_JIT, "fic_bf_encode_static: unknown mode %d", mode);
    const int total = (H / 8) * (W / 8);
    FIC_REQUIRE(0 <= range_begin && range_begin <= range_end && range_end <= total,
                "fic_bf_encode_static: range slice [%d,%d) outside [0,%d)", range_begin, range_end, total);
    const int n_ranges = range_end - range_begin;
    if (n_ranges == 0) return FIC_OK;

    int dev = -1;
    FIC_CUDA(cudaGetDevice(&dev));
    FIC_REQUIRE(dev >= 0 && dev < FIC_STATIC_MAX_DEVICES, "fic_bf_encode_static: device ordinal %d not supported", dev);
    if (!g_static_ptr[dev]) {  // once per device, with that device current
        unsigned char* p = nullptr;
        FIC_CUDA(cudaGetSymbolAddress((void**)&p, g_static_scratch));
        int sms = 0;
        FIC_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        g_static_sms[dev] = sms;
        g_static_ptr[dev] = p;
    }
    unsigned char* base = g_static_ptr[dev];
    cudaStream_t st = (cudaStream_t)stream;

    uint16_t* P = reinterpret_cast<uint16_t*>(base);
    const size_t NP = (size_t)(W - 16 > 0 ? W - 16 : 0) * (size_t)(H - 16 > 0 ? H - 16 : 0);
    uint32_t* sd4 = reinterpret_cast<uint32_t*>(base + FIC_STATIC_BOX_BYTES);
    uint32_t* sq4 = sd4 + NP;
    float* sdd = reinterpret_cast<float*>(sq4 + NP);
    BestRec* partial = reinterpret_cast<BestRec*>(base + FIC_STATIC_BOX_BYTES + FIC_STATIC_STATS_BYTES);

    dim3 gb(ceil_div(W - 1, 256), H - 1);
    bf_boxsum_kernel<<<gb, 256, 0, st>>>(d_img, H, W, P);
    FIC_TRY(fic_check_launch("bf_boxsum_kernel (static)"));
    if (NP) {
        bf_stats_kernel<<<ceil_div(NP, 128), 128, 0, st>>>(P, H, W, sd4, sq4, sdd);
        FIC_TRY(fic_check_launch("bf_stats_kernel (static)"));
    }
    BfGeom g;
    g.H = H; g.W = W; g.si = stride_i; g.sj = stride_j; g.PW = W - 16; g.BW = W - 1;
    g.npi = H - 16 > 0 ? ceil_div(H - 16, stride_i) : 0;
    g.npj = W - 16 > 0 ? ceil_div(W - 16, stride_j) : 0;
    const long long nq = (long long)g.npi * g.npj;
    const int groups = ceil_div(n_ranges, BF_G);
    int n_chunks = 1;
    if (nq > 0) {
        long long want = (long long)g_static_sms[dev] * 16;
        n_chunks = (int)((want + groups - 1) / groups);
        int max_chunks = ceil_div(nq, BF_THREADS * 4);
        if (n_chunks > max_chunks) n_chunks = max_chunks;
        int fit = (int)(FIC_STATIC_PARTIAL_RECS / (unsigned)n_ranges);  // the partial table is fixed-size here
        if (n_chunks > fit) n_chunks = fit;
        if (n_chunks < 1) n_chunks = 1;
    }
    int chunk_len = nq > 0 ? ceil_div(nq, n_chunks) : 1;
    chunk_len = ceil_div(chunk_len, BF_THREADS) * BF_THREADS;
    n_chunks = nq > 0 ? ceil_div(nq, chunk_len) : 1;
    dim3 grid(groups, n_chunks);
    if (mode == FIC_BF_MODE_JIT) {
        bf_scan8_kernel<FIC_BF_MODE_JIT><<<grid, BF_THREADS, BF_SCAN_DYN_SMEM, st>>>(d_img, P, sd4, sq4, sdd, g, range_begin, n_ranges,
                                                                                     nullptr, chunk_len, n_chunks, partial);
        FIC_TRY(fic_check_launch("bf_scan8_kernel<jit> (static)"));
        bf_finalize8_kernel<FIC_BF_MODE_JIT><<<ceil_div(n_ranges, 128), 128, 0, st>>>(d_img, P, sd4, sq4, sdd, g, range_begin, n_ranges,
                                                                                      nullptr, n_chunks, partial, d_idx, d_so, d_R);
        FIC_TRY(fic_check_launch("bf_finalize8_kernel<jit> (static)"));
    } else {
        bf_scan8_kernel<FIC_BF_MODE_NUMPY><<<grid, BF_THREADS, BF_SCAN_DYN_SMEM, st>>>(d_img, P, sd4, sq4, sdd, g, range_begin, n_ranges,
                                                                                       nullptr, chunk_len, n_chunks, partial);
        FIC_TRY(fic_check_launch("bf_scan8_kernel<numpy> (static)"));
        bf_finalize8_kernel<FIC_BF_MODE_NUMPY><<<ceil_div(n_ranges, 128), 128, 0, st>>>(d_img, P, sd4, sq4, sdd, g, range_begin, n_ranges,
                                                                                        nullptr, n_chunks, partial, d_idx, d_so, d_R);
        FIC_TRY(fic_check_launch("bf_finalize8_kernel<numpy> (static)"));
    }
    return FIC_OK;
}
