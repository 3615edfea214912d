// Brute-force encoder, tensor-core path: pool kernel -> tcgen05 correlation GEMM with a fused screening
// epilogue -> exact rescoring of the survivors.  8x8 ranges, 16x16 domains, stride 1, JIT semantics
// (bruteforce/jit_bfencoder.py:32-95).  The GEMM never decides anything on its own: it only discards candidates
// that provably cannot win, everything that survives is scored by the exact arithmetic of scorer.cuh, so the
// records are bit-identical to the exact scan (bf_exact.cu) and to the reference.
//
// Math.  With D4 = 4*d (integers), X = sum r_k*D4_k, SD4 = sum D4, the reference's least-squares numerator is
//     n*ab - sum(r)*sum(d) = (64*X - sum(r)*SD4)/4 = 16*u,        u := X - sum(r)*SD4/64 .
// Centre both operands by integers: r' = r - m_r, D4' = D4 - m_c (m = round(mean)), rho = sum r', eps = sum D4':
//     u = sum r'_k*D4'_k - rho*eps/64                              (exact identity)
// The first term is one K=64 dot product of fp16-exact integer operands (|r'|<=255, |D4'|<=1020, every partial sum
// < 2^24: exact in the fp32 accumulator); the second is a rank-1 term of magnitude at most 32*32/64 = 16 that the GEMM
// does not compute: it is covered by the slack DELTA_U of the threshold (|u| >= |X'| - 16 with X' the accumulator).
// (An earlier version carried it in a K=16 extension slab: 5 instead of 4 tcgen05.mma per tile, 160 instead of
// 128 bytes per pool row, for a threshold that was 14 units of u tighter.)
//
// Bound.  For ANY (s,o) the exact-arithmetic collage error obeys R >= R_opt = (A_r - 64*u^2/den16)/64 with
// A_r = sum r^2 - (sum r)^2/64, den16 = 64*sum(D4^2) - SD4^2 (integers).  The reference's R differs from exact
// arithmetic only through its float32 sequential sum of d^2 (every other input is exact for 8x8 blocks):
//     R_ref(c) >= R_opt(c) - kappa_c * s_ref^2,   kappa_c = 1.02 * 63 * 2^-24 / 64 * sum d^2   (<= 0.25)
// and |s_ref| <= 1 (clamp).  For a well-conditioned domain (den16 >= 5e-4 * sum D4^2, i.e. the float32 noise in the
// denominator is below half of it) |s_ref| <= 2|s_opt| = 512|u|/den16, so the guard is proportional to u^2 and folds
// into the per-candidate scale:  candidate c can win or tie only if
//     u^2 * h_c^2 >= A_r - 64*UB_r,     h_c^2 = 64/den16 * (1 + 1e-3 * sum D4^2 / den16)          (tau_A)
// For the (rare) ill-conditioned domains the constant guard is used:  u^2 * 64/den16 >= A_r - 64*(UB_r + 0.26)  (tau_B).
// UB_r is the exact R of some candidate (any exactly scored candidate bounds the winner from above).
// The epilogue evaluates  (|X'| + DELTA)*h_c >= tau  as  |X'| >= theta := tau/h_c - DELTA  (DELTA = 16 covers the
// rank-1 centring term); theta is one FFMA per (thread, range, tile), the 8 isometry columns of a range are
// max-reduced first, so the common case is ~5 instructions per 8 outputs.
//
// Rounds: a seed round ranks a coarse lattice of candidates (~NP/100 positions) per range with the same GEMM and a
// ranking epilogue; the best one and position 0 are scored exactly and give the first bounds.  Round 1 then screens a
// sparse subset of the candidate tiles (every 8th for 512x512, every 256th for 4096x4096), its survivors are rescored
// exactly and give much tighter bounds for the next, ~16x larger, round, and so on until every tile has been screened once.
// Ranges the screen cannot decide (bound not below the range's own variance, survivor list overflow) go to the
// exact scan; constant ranges are decided by the seed list (every candidate ties at R = 0, the first one wins).
// The NumPy encoder's semantics (bf_encoder.py) go through the same screen: see bf_init_kernel.
//
// GEMM shape (per CTA): M = 128 candidates (TMEM lanes, streamed by TMA through a ring of stages), N = 256 columns = 32
// ranges x 8 isometries (resident in shared memory), K = 64.  tcgen05.mma cta_group::1 kind::f16, fp32 accumulators in
// TMEM, 2 accumulator stages x 256 columns = all 512 TMEM columns, so the MMA of tile t+1 overlaps the epilogue of
// tile t.  Warp roles: warps 0..15 epilogue, warp 16 TMA producer, warp 17 MMA issuer + TMEM allocator (both run in
// lock step with one elected lane issuing); epilogue in two groups of eight, one group per accumulator stage (two warps
// per TMEM lane quadrant, 128 columns each).  Thread = candidate, so h_c is a per-thread register and the 8 isometry
// columns of a range share one tau.  The (group, tile) space is cut into equal contiguous spans, one per CTA.
//
// HBM/L2 layout written by bf_pool_kernel ("TMA-friendly domain-candidate pool"):
//   poolA   fp16 [NP][64]  K-major rows (128 B) -> TMA box 64x128, SWIZZLE_128B
//   hk      float2 [NP]    (+-1/h_c, den16); a negative .x marks an ill-conditioned domain, .y only ranks seed candidates
// and by bf_init_kernel:  rngB fp16 [Ncols][64], column = slot*8 + isometry.
#include <algorithm>
#include <vector>
#include <cuda.h>
#include <cuda_fp16.h>
#include <stdlib.h>

#include "fic_common.cuh"
#include "scorer.cuh"

int bf_exact_scan8_dyn(fic_handle* h, const uint8_t* d_img, int H, int W, int mode, int range_begin, int n_total,
                       const int* d_ctrl, const int32_t* d_range_list, unsigned long long* d_totals, int32_t* d_idx,
                       double* d_so, double* d_R, cudaStream_t st);

// Measurement knobs (environment variables read per encode, kernel variants whose records are wrong on purpose) exist
// only in builds made with `make TUNING=1` (-DFIC_TUNING); the shipped library ignores the environment.
#ifdef FIC_TUNING
static const char* tuning_env(const char* name) { return getenv(name); }
#else
static const char* tuning_env(const char*) { return nullptr; }
#endif

namespace {

constexpr int BM = 128;      // candidates per tile
constexpr int BN = 256;      // columns per unit (32 ranges x 8 isometries)
constexpr int KMAIN = 64;
#ifndef FIC_SCREEN_STAGES
#define FIC_SCREEN_STAGES 6
#endif
constexpr int STAGES = FIC_SCREEN_STAGES;
constexpr int WARPS_PER_STAGE = 8;                    // epilogue warps per accumulator stage (two per TMEM lane quadrant)
constexpr int EPI_WARPS = 2 * WARPS_PER_STAGE;
constexpr int RANGES_PER_WARP = 16;                   // ranges (groups of 8 columns) one epilogue warp looks at per tile
// the scheduler favours high warp ids: the two single-thread roles sit ABOVE the epilogue warps so that the busy
// epilogue can never delay a TMA issue or an MMA issue
constexpr int TMA_WARP = EPI_WARPS, MMA_WARP = EPI_WARPS + 1;
constexpr int SCREEN_THREADS = 64 + EPI_WARPS * 32;   // registers are allocated per four warps: 576 threads -> 96 each
constexpr uint32_t A_MAIN_BYTES = BM * KMAIN * 2;   // 16384
constexpr uint32_t B_MAIN_BYTES = BN * KMAIN * 2;   // 32768
constexpr uint32_t OFF_B_MAIN = 0;
constexpr uint32_t OFF_A_MAIN = OFF_B_MAIN + B_MAIN_BYTES;
constexpr uint32_t OFF_BARS = OFF_A_MAIN + STAGES * A_MAIN_BYTES;
constexpr uint32_t OFF_TAU = OFF_BARS + 256;                        // per epilogue warp: 16 x float2 thresholds
constexpr uint32_t OFF_QUEUE = OFF_TAU + EPI_WARPS * 16 * 8;          // per epilogue warp: survivor queue
constexpr int QFLUSH = 128;                                          // a queue is flushed between tiles once it holds this many entries
constexpr int QCAP = QFLUSH + RANGES_PER_WARP * 32;                  // ... and a tile adds at most one entry per (range, lane)
constexpr uint32_t QUEUE_BYTES = QCAP * 8;                           // sizeof(WarpQueue), checked below
constexpr uint32_t SCREEN_SMEM = OFF_QUEUE + EPI_WARPS * QUEUE_BYTES + 1024;  // + slack to align the base to 1024
static_assert(SCREEN_SMEM <= 232448, "shared memory of the screen kernel exceeds the 227 KB a CTA can have");

constexpr int BF_CTRL_TOTALS = 16;   // int index inside the control block where the four u64 totals start

constexpr double GUARD_ILL = 0.26;   // constant guard (R units) for ill-conditioned domains: kappa_c <= 0.25
constexpr float DELTA_U = 16.0f;     // |u - X'| = |rho*eps|/64 <= 32*32/64: the rank-1 centring term the GEMM leaves out

// ---- PTX wrappers ----------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred P1;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, P1;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded wait: a broken pipeline must not hang the GPU.  ~4 s at 2 GHz, then flag + trap.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, int* err_flag, int code)
{
    if (mbar_try_wait(bar, parity)) return;
    const long long t0 = clock64();
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if ((++spins & 1023u) == 0 && clock64() - t0 > 8000000000ll) {   // look at the clock only now and then
            if (err_flag) atomicExch(err_flag, code);
            __threadfence_system();
            __trap();
        }
    }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
        "l"(map), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ bool elect_one()
{
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_commit(uint32_t bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// K-major operand tile in shared memory (cute::UMMA::SmemDescriptor, mma_sm100_desc.hpp): start address >> 4 in
// bits [0,14), LBO (unused for swizzled K-major, = 1) in [16,30), SBO >> 4 in [32,46), version 1 in [46,48),
// layout type in [61,64): 2 = SWIZZLE_128B (rows of 128 B, 8-row groups 1024 B apart), 6 = SWIZZLE_32B (rows of 32 B,
// 8-row groups 256 B apart).
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t sbo_bytes, uint32_t layout_type)
{
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)layout_type << 61;
    return d;
}
// cute::UMMA::InstrDescriptor: c_format F32 (1) at bit 4, a/b format F16 (0), K-major both, N>>3 at bit 17, M>>4 at bit 24
constexpr uint32_t IDESC = (1u << 4) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32])
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
          "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
          "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr));
}
// wait for the outstanding TMEM loads; the registers are tied to the asm so no consumer can be scheduled above it
__device__ __forceinline__ void tmem_ld_wait(uint32_t (&v)[32])
{
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]),
                   "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]), "+r"(v[16]),
                   "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]), "+r"(v[21]), "+r"(v[22]), "+r"(v[23]), "+r"(v[24]),
                   "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]), "+r"(v[29]), "+r"(v[30]), "+r"(v[31])
                 :
                 : "memory");
}

// Per-range constants of the scoring loop, one 32-byte record (written by qt_tcinit_kernel).
struct __align__(16) QtcRange {
    int rho;                  // sum r - N * round(sum r / N): the centring remainder of the range
    int ok;                   // 0: not searched here (past the end, or handed to the exact search)
    uint32_t tw0, tw1;        // isometry by domain type, eight 4-bit entries, for positive / negative scales
    double slack;             // sum r^2 - (sum r)^2 / N
    float rs, sqr;
};

struct ScreenParams {
    const float2* hk;        // [NP]
    const float2* tau;       // [slots padded to 32] = (tau_A, tau_B)
    const int* n_elig;       // device scalar
    uint32_t* cnt;           // [slots] survivors found this round
    uint32_t* list;          // [slots][cap] entries = position*8 + isometry
    uint32_t cap;
    int NP;
    int n_tiles;             // candidate tiles in the whole pool
    int m_cur;               // this round visits the tiles t with t % m_cur == 0 ...
    int r_prev;              // ... that were not visited before: (t / m_cur) % r_prev != 0   (0 = first round: all of them)
    int n_round_tiles;       // tiles visited in this round
    int window_tiles;        // round tiles per L2-sized window: all CTAs sweep one window before the next
    float* dump;             // probe mode: [n_tiles*128][256] raw accumulators of column group 0 (else nullptr)
    int* err_flag;
    int debug_mode;          // 0 = normal; 1/2/3 = timing experiments (FIC_SCREEN_DEBUG), results invalid
    unsigned long long* seed_best;   // seed mode: [slots] max of (gain bits << 32 | ~local candidate)
    // MODE 7 (quadtree exact pass): the bound test runs in the epilogue, survivors are scored from the warp's queue
    const uint8_t* dtyp;     // [NP] domain type = rotation | orientation << 2
    int grp_begin, grp_count;        // column groups of this launch
    const QtcRange* rngc;    // [slots]
    unsigned long long* bound;       // [slots] best score found so far (bits of a non-negative double)
    const uint32_t* sd4;     // [NP] sum D4
    const uint32_t* sq4;     // [NP] sum D4^2
    const float* sds8;       // [8][NP], 8x8 blocks: float32 sequential sum of d^2 per isometry
    const double* den;       // [NP] N * sds - ds^2 (8x8 blocks: the smallest of the eight)
    double max_scale;
    int scale_bits, offset_bits, nmodes, seed_pass;
    // MODE 8 (brute force, 4x4 ranges): survivors are rescored from the pixels
    const uint8_t* img;      // the image, row pitch img_W
    const uint16_t* boxsum;  // 2x2 cell sums at every pixel offset, (H-1) x (W-1)
    int img_W, lat_si, lat_sj, lat_npj;      // lattice of domain positions: (q / lat_npj * lat_si, q % lat_npj * lat_sj)
    int bf_mode;             // FIC_BF_MODE_JIT or FIC_BF_MODE_NUMPY: which scorer the survivors get
};

// j-th tile of the round, in increasing order.  Rounds use nested strides m_1 > m_2 > ... > m_last = 1 (each divides
// the previous one): round k takes the multiples of m_k that are not multiples of m_{k-1}.
__device__ __forceinline__ int tile_of(const ScreenParams& p, int j)
{
    if (p.r_prev == 0) return j * p.m_cur;
    const int r1 = p.r_prev - 1;
    const int q = j / r1;
    return (q * p.r_prev + (j - q * r1) + 1) * p.m_cur;
}

// Work split.  The round's tiles are cut into windows (a window's slice of the pool fits L2); inside a window the
// linearised (group, tile) space is cut into equal contiguous spans, one per CTA, so the load is balanced to one tile
// and all CTAs read the same part of the pool at about the same time.
struct Span {
    long long lin, end;
    int w0, wn;            // current window: first round tile, number of round tiles
    int n_round_tiles, window_tiles, n_groups;
    __device__ Span(const ScreenParams& p, int n_groups_)
    {
        n_round_tiles = p.n_round_tiles; window_tiles = p.window_tiles; n_groups = n_groups_;
        w0 = -window_tiles; wn = 0; lin = 0; end = 0;
    }
    __device__ bool next(int& grp, int& j0, int& j1)
    {
        while (lin >= end) {   // open the next window
            w0 += window_tiles;
            if (w0 >= n_round_tiles || n_groups == 0) return false;
            wn = min(window_tiles, n_round_tiles - w0);
            const long long total = (long long)n_groups * wn;
            const long long per = (total + gridDim.x - 1) / gridDim.x;
            lin = per * blockIdx.x;
            end = lin + per < total ? lin + per : total;
        }
        grp = (int)(lin / wn);
        const int o = (int)(lin - (long long)grp * wn);
        const long long room = end - lin;
        const int o1 = (long long)wn - o < room ? wn : o + (int)room;
        j0 = w0 + o;
        j1 = w0 + o1;
        lin += o1 - o;
        return true;
    }
};

// Screen one 32-column chunk held in registers.  theta[g] = tau_g/h_c - DELTA is this thread's threshold on |u| for
// the g-th range of the chunk:  (|u| + DELTA)*h_c >= tau  <=>  |u| >= theta.  The 8 isometry columns of a range are
// reduced with max first, so the common case costs ~5 instructions per 8 outputs.
// rare path, kept out of line so the hot loop stays small: one reservation for all passing isometries of a
// (candidate, range) pair, then the entries
__device__ __noinline__ void append_hits(uint32_t* __restrict__ cnt, uint32_t* __restrict__ list, uint32_t cap, int slot,
                                         uint32_t cand, unsigned m8)
{
    uint32_t at = atomicAdd(&cnt[slot], (uint32_t)__popc(m8));
    while (m8) {
        const int i = __ffs(m8) - 1;
        m8 &= m8 - 1;
        if (at < cap) list[(size_t)slot * cap + at] = cand * 8u + (uint32_t)i;
        ++at;
    }
}

// Per-warp survivor queue in shared memory.  The fill level lives in a (warp-uniform) register and hits are placed
// with a ballot + prefix count, so a hit costs one vote and one store -- no atomics, no divergent lanes.  The global
// reservations (one atomicAdd per entry, a full round trip to L2) are issued 32 at a time when the queue is flushed
// between tiles.  A tile can add at most RANGES_PER_WARP x 32 lanes entries per warp, and the queue is flushed
// whenever a tile ends with QFLUSH or more entries, so QCAP can never overflow.
struct WarpQueue {
    uint2 e[QCAP];   // x = (slot << 8) | isometry mask, y = candidate position
};
static_assert(sizeof(WarpQueue) == QUEUE_BYTES, "QUEUE_BYTES must match WarpQueue");

__device__ __noinline__ void queue_flush(WarpQueue* q, const ScreenParams& p, int lane, uint32_t n)
{
    __syncwarp();
    for (uint32_t k = lane; k < n; k += 32) {
        const uint2 ent = q->e[k];
        append_hits(p.cnt, p.list, p.cap, (int)(ent.x >> 8), ent.y, ent.x & 0xffu);
    }
    __syncwarp();
}

// Quadtree exact pass: queue of the candidates the bound test did not drop, one per (position, range, scale sign), scored
// 32 at a time -- one entry per lane, so the scorer never runs in a divergent warp.  Exact score from the exact dot product
// (ab), the position's integer sums (ds, and sds of a 4x4 block) and the table of sequential float32 sums (sds of an 8x8
// block): bit for bit the inputs qt_search_kernel's pixel loop produces.  A candidate at or below the best score of its
// range lowers the bound and -- outside the seed pass -- enters the range's list; the rescoring kernel picks the first
// minimum in scan order among the listed.
constexpr int QT_COLS = 4;                    // columns per range: (scale sign) x (orientation of the domain's class)
constexpr int QT_RPW = 128 / QT_COLS;         // ranges one epilogue warp looks at per tile (its 128 columns)
constexpr int QT_RPG = BN / QT_COLS;          // ranges per column group
constexpr int QT_QCAP = QUEUE_BYTES / 8;
constexpr int QT_QROOM = 8 * 2 * 32;          // a 32-column chunk adds at most 8 ranges x 2 scale signs x 32 lanes entries
static_assert(QT_QCAP >= QT_QROOM + 32, "queue of the quadtree pass too small");
struct QtQueue {
    uint2 e[QT_QCAP];        // x = the accumulator (float bits, an integer), y = position << 6 | range of the warp's 32 << 1 | scale sign
};
static_assert(sizeof(QtQueue) <= QUEUE_BYTES, "QtQueue must fit a warp's queue slice");

// The isometry that pairs a domain type with a range type (quadtree/encoder.py:168-176) factors, for each orientation of
// the domain's class, into a permutation of the domain block that depends on the domain type alone and one of the range
// block that depends on the range type alone:   sum_k r[k] d[iso_t(k)] = sum_k r[beta(k)] d[alpha(k)] .
// So the pool holds every block already permuted by its own alpha, the range operand holds two permuted copies per scale
// sign (one per domain orientation), and the pair's exact dot product is ONE accumulator, picked by the domain's orientation
// bit -- not one of eight picked by the full isometry.  The tables give alpha / beta as isometry numbers
// (orientation * 4 + rotation, fic_iso_qt_enc); indices: domain type = rotation | orientation << 2, same for ranges.
__device__ __constant__ unsigned char QT_ALPHA[8] = {0, 1, 2, 3, 6, 7, 4, 5};
__device__ __constant__ unsigned char QT_BETA[2][8] = {{0, 1, 2, 3, 6, 5, 4, 7}, {0, 3, 2, 1, 6, 7, 4, 5}};

template <int N>
__device__ __noinline__ void qt_queue_flush(const QtQueue* q, const ScreenParams& p, int slot0, int lane, uint32_t n)
{
    __syncwarp();
    for (uint32_t k = lane; k < n; k += 32) {
        const uint2 en = q->e[k];
        const int slot = slot0 + (int)((en.y >> 1) & 31u), cand = (int)(en.y >> 6), Xi = (int)__uint_as_float(en.x);
        const QtcRange rc = p.rngc[slot];
        if (!rc.ok) continue;     // (a range that is not searched here can still meet a candidate that is never dropped)
        const int t = (int)((((en.y & 1u) ? rc.tw1 : rc.tw0) >> (4u * (p.dtyp[cand] & 7u))) & 7u);   // the pair's isometry
        const unsigned sd = p.sd4[cand], sq = p.sq4[cand];
        const int mc = (int)((sd + (unsigned)(N / 2)) / (unsigned)N), eps = (int)sd - N * mc;
        const int mr = ((int)rc.rs - rc.rho) / N;
        const int dot = Xi + mr * eps + mc * rc.rho + N * mr * mc;     // sum r D4, exact
        const float ab = __fmul_rn((float)dot, 0.25f);
        const float ds = __fmul_rn((float)sd, 0.25f);
        const float sds = N == 64 ? p.sds8[(size_t)t * p.NP + cand] : __fmul_rn((float)sq, 0.0625f);
        const QScore sc = fic_q_score(N, rc.rs, rc.sqr, ds, sds, ab, p.max_scale, p.scale_bits, p.offset_bits);
        const double B = __longlong_as_double((long long)__ldcg(&p.bound[slot]));
        if (sc.R <= B && sc.R < __longlong_as_double(0x7ff0000000000000ll)) {
            if (sc.R < B) atomicMin(&p.bound[slot], (unsigned long long)__double_as_longlong(sc.R > 0.0 ? sc.R : 0.0));
            if (!p.seed_pass) {
                const uint32_t at = atomicAdd(&p.cnt[slot], 1u);
                if (at < p.cap) p.list[(size_t)slot * p.cap + at] = (uint32_t)cand * 8u + (uint32_t)t;
            }
        }
    }
    __syncwarp();
}

// ---- brute force with 4x4 ranges (MODE 8) -----------------------------------------------------------------------------
// Sixteen pixels per block: every float32 sum of the reference (jit_bfencoder.py:40-62) is exact -- sum d^2 needs 24 bits,
// sum d r 22 -- so the integer sums ARE the reference's sums and, as in the quadtree pass, the GEMM's accumulator gives the
// exact dot product.  All eight isometries of a (range, domain) pair are candidates (eight columns per range, one
// accumulator each); the epilogue tests the scorer's lower bound against the best score of the range and queues what it
// cannot drop as (position, range, isometry) -- four bytes; the survivor's sums are rebuilt from the pixels when it is scored.
constexpr int BF4_QCAP = QUEUE_BYTES / 4;
static_assert(BF4_QCAP >= 4 * 8 * 32 + 32, "queue of the 4x4 brute-force pass too small");   // a 32-column chunk adds at most 1024 entries
struct Bf4Queue {
    uint32_t e[BF4_QCAP];     // position << 7 | range of the warp's 16 << 3 | isometry
};

// score of range `rc` (origin in tw0 / tw1) against the domain at lattice position q under isometry t, the reference's way
__device__ __forceinline__ BfScore bf4_score_pair(const ScreenParams& p, const QtcRange& rc, int q, int t)
{
    const int di = (q / p.lat_npj) * p.lat_si, dj = (q % p.lat_npj) * p.lat_sj;
    const int BW = p.img_W - 1;
    const uint8_t* rp = p.img + (size_t)rc.tw0 * p.img_W + rc.tw1;
    const uint16_t* dp = p.boxsum + (size_t)di * BW + dj;
    int sd = 0, sq = 0, ab = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int a, b;
            fic_iso_bf(t >> 2, t & 3, 4, i, j, a, b);
            const int d4 = dp[(size_t)(2 * a) * BW + 2 * b];
            sd += d4;
            sq += d4 * d4;
            ab += d4 * (int)rp[(size_t)i * p.img_W + j];
        }
    if (p.bf_mode == FIC_BF_MODE_NUMPY)      // bf_encoder.py:56-71: float64 sums (exact), no clamp; rc.sqr is the uint8-wrapped sum
        return fic_bfnp_score(16.0, 0.0625, (double)rc.rs, (double)rc.sqr, DM((double)sd, 0.25), DM((double)sq, 0.0625), DM((double)ab, 0.25));
    const float ds = __fmul_rn((float)sd, 0.25f);
    return fic_bfjit_score(16.0, 0.0625, (double)rc.rs, (double)rc.sqr, (double)ds, (double)__fmul_rn((float)sq, 0.0625f),
                           (double)__fmul_rn(ds, ds), (double)__fmul_rn((float)ab, 0.25f));
}

__device__ __noinline__ void bf4_queue_flush(const Bf4Queue* q, const ScreenParams& p, int slot0, int lane, uint32_t n)
{
    __syncwarp();
    for (uint32_t k = lane; k < n; k += 32) {
        const uint32_t en = q->e[k];
        const int slot = slot0 + (int)((en >> 3) & 15u), cand = (int)(en >> 7), t = (int)(en & 7u);
        const QtcRange rc = p.rngc[slot];
        if (!(rc.ok & 1)) continue;
        const BfScore sc = bf4_score_pair(p, rc, cand, t);
        const double B = __longlong_as_double((long long)__ldcg(&p.bound[slot]));
        if (sc.R <= B && sc.R < __longlong_as_double(0x7ff0000000000000ll)) {
            if (sc.R < B) atomicMin(&p.bound[slot], (unsigned long long)__double_as_longlong(sc.R > 0.0 ? sc.R : 0.0));
            if (!p.seed_pass) {
                const uint32_t at = atomicAdd(&p.cnt[slot], 1u);
                if (at < p.cap) p.list[(size_t)slot * p.cap + at] = (uint32_t)cand * 8u + (uint32_t)t;
            }
        }
    }
    __syncwarp();
}

template <bool APPEND>
__device__ __forceinline__ void screen_chunk(const uint32_t (&v)[32], const float* theta, const ScreenParams& p, int slot0,
                                             int my_cand, WarpQueue* q, uint32_t& qn, int lane)
{
    float m[4];
    bool any = false;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        m[g] = fabsf(__uint_as_float(v[g * 8]));
#pragma unroll
        for (int i = 1; i < 8; ++i) m[g] = fmaxf(m[g], fabsf(__uint_as_float(v[g * 8 + i])));
        any |= m[g] >= theta[g];
    }
    if (!APPEND) {  // timing experiment: the screening math without the survivor path
        if (any && m[0] == 1.2345e-30f) p.cnt[0] = 1;
        return;
    }
    if (__any_sync(0xffffffffu, any)) {   // warp-uniform from here on
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            const bool hit = m[g] >= theta[g];
            const unsigned mask = __ballot_sync(0xffffffffu, hit);
            if (mask) {
                if (hit) {
                    unsigned m8 = 0;
#pragma unroll
                    for (int i = 0; i < 8; ++i) m8 |= (fabsf(__uint_as_float(v[g * 8 + i])) >= theta[g] ? 1u : 0u) << i;
                    q->e[qn + __popc(mask & ((1u << lane) - 1u))] = make_uint2(((uint32_t)(slot0 + g) << 8) | m8, (uint32_t)my_cand);
                }
                qn += __popc(mask);
            }
        }
    }
}

// Two chunks at once (both already in registers, no load depends on them): eight independent reductions, one vote.
template <bool APPEND>
__device__ __forceinline__ void screen_pair(const uint32_t (&va)[32], const uint32_t (&vb)[32], const float* theta,
                                            const ScreenParams& p, int slot0, int my_cand, WarpQueue* q, uint32_t& qn, int lane)
{
    float m[8];
    bool any = false;
#pragma unroll
    for (int g = 0; g < 8; ++g) {
        const uint32_t* v = g < 4 ? va + g * 8 : vb + (g - 4) * 8;
        m[g] = fabsf(__uint_as_float(v[0]));
#pragma unroll
        for (int i = 1; i < 8; ++i) m[g] = fmaxf(m[g], fabsf(__uint_as_float(v[i])));
        any |= m[g] >= theta[g];
    }
    if (!APPEND) {
        if (any && m[0] == 1.2345e-30f) p.cnt[0] = 1;
        return;
    }
    if (__any_sync(0xffffffffu, any)) {   // warp-uniform from here on
#pragma unroll
        for (int g = 0; g < 8; ++g) {
            const bool hit = m[g] >= theta[g];
            const unsigned mask = __ballot_sync(0xffffffffu, hit);
            if (mask) {
                if (hit) {
                    const uint32_t* v = g < 4 ? va + g * 8 : vb + (g - 4) * 8;
                    unsigned m8 = 0;
#pragma unroll
                    for (int i = 0; i < 8; ++i) m8 |= (fabsf(__uint_as_float(v[i])) >= theta[g] ? 1u : 0u) << i;
                    q->e[qn + __popc(mask & ((1u << lane) - 1u))] = make_uint2(((uint32_t)(slot0 + g) << 8) | m8, (uint32_t)my_cand);
                }
                qn += __popc(mask);
            }
        }
    }
}

// ---- the screen ---------------------------------------------------------------------------------------------
// MODE 0 = production; MODE 1 = probe (dump raw accumulators of group 0); MODE 2 / 3 = timing experiments selected by
// FIC_SCREEN_DEBUG (2: TMEM loads but no screening, 3: no epilogue work at all) -- their results are wrong on purpose.
// KT = K of the contraction = pixels per block: 64 (8x8 blocks: 128-byte operand rows, SWIZZLE_128B, four MMAs per tile) or
// 16 (4x4 blocks of the quadtree's smallest ranges: 32-byte rows, SWIZZLE_32B, one MMA per tile).  The shared-memory layout
// is the K = 64 one in both cases.
template <int MODE, int KT = KMAIN>
__global__ void __launch_bounds__(SCREEN_THREADS, 1)
bf_screen_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, const __grid_constant__ ScreenParams p)
{
    static_assert(KT == 64 || KT == 16, "operand rows of 128 or 32 bytes");
    constexpr uint32_t A_BYTES = BM * KT * 2, B_BYTES = BN * KT * 2;
    constexpr uint32_t DESC_SBO = KT == 64 ? 1024 : 256, DESC_LAYOUT = KT == 64 ? 2 : 6;   // 8-row group pitch, swizzle mode
    extern __shared__ unsigned char smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;  // SWIZZLE_128B tiles want 1024-byte alignment
    unsigned char* gbase = smem_raw + (base - raw);
    const uint32_t bars = base + OFF_BARS;
    // barrier slots (8 bytes each): full[STAGES], empty[STAGES], b_full, b_empty, tmem_full[2], tmem_empty[2]
    auto full_bar = [&](int s) { return bars + 8u * s; };
    auto empty_bar = [&](int s) { return bars + 8u * (STAGES + s); };
    const uint32_t b_full = bars + 8u * (2 * STAGES), b_empty = bars + 8u * (2 * STAGES + 1);
    auto tfull_bar = [&](int s) { return bars + 8u * (2 * STAGES + 2 + s); };
    auto tempty_bar = [&](int s) { return bars + 8u * (2 * STAGES + 6 + s); };
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gbase + OFF_BARS + 8 * (2 * STAGES + 10));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_elig = *p.n_elig;
    const int grp_off = MODE == 7 || MODE == 8 ? p.grp_begin : 0;            // the quadtree modes name their column groups
    const int n_groups = MODE == 7 || MODE == 8 ? p.grp_count : (n_elig + 31) / 32;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        mbar_init(b_full, 1);
        mbar_init(b_empty, 1);
        for (int s = 0; s < 2; ++s) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), WARPS_PER_STAGE); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == MMA_WARP) {  // TMEM: all 512 columns (one CTA per SM by construction: > 113 KB of shared memory)
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32((const void*)tmem_slot)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == TMA_WARP) {
        // ===== TMA producer =====  (whole warp in lock step, one elected lane issues -- see the MMA issuer)
        int s = 0;
        uint32_t ph = 0, bph = 0;
        const bool issuer = elect_one();
        Span span(p, n_groups);
        int grp, j0, j1;
        while (span.next(grp, j0, j1)) {
            grp += grp_off;
            mbar_wait(b_empty, bph ^ 1, p.err_flag, 1);  // previous unit's MMAs have finished reading B
            if (issuer) {
                mbar_expect_tx(b_full, B_BYTES);
                tma_load_2d(base + OFF_B_MAIN, &mapB, b_full, 0, grp * BN);
            }
            bph ^= 1;
            // tile index without a division per tile (same walk as the epilogue's)
            int t_cur = tile_of(p, j0);
            const int r1 = p.r_prev > 0 ? p.r_prev - 1 : 0x7fffffff;
            int rem = p.r_prev > 0 ? j0 % r1 : 0;
            for (int j = j0; j < j1; ++j) {
                mbar_wait(empty_bar(s), ph ^ 1, p.err_flag, 2);
                if (issuer) {
                    mbar_expect_tx(full_bar(s), A_BYTES);
                    tma_load_2d(base + OFF_A_MAIN + s * A_MAIN_BYTES, &mapA, full_bar(s), 0, t_cur * BM);
                }
                __syncwarp();
                int step = 1;
                if (++rem >= r1) { rem = 0; step = 2; }
                t_cur += step * p.m_cur;
                if (++s == STAGES) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp == MMA_WARP) {
        // ===== MMA issuer =====
        // The issue loop is the pacemaker of the whole CTA: every cycle it spends between "accumulator free" and the
        // last tcgen05.mma is a cycle the tensor pipe idles.  All 32 lanes run it in lock step (so ptxas keeps
        // descriptors and barrier addresses in uniform registers instead of broadcasting them out of a divergent lane
        // before every instruction), one elected lane issues, and the descriptors are advanced by additions.
        int s = 0, as = 0;
        uint32_t ph = 0, aph = 0, bph = 0;
        const uint64_t bdesc = make_desc(base + OFF_B_MAIN, DESC_SBO, DESC_LAYOUT);
        const uint64_t adesc0 = make_desc(base + OFF_A_MAIN, DESC_SBO, DESC_LAYOUT);
        const bool issuer = elect_one();
        Span span(p, n_groups);
        int grp, j0, j1;
        while (span.next(grp, j0, j1)) {
            mbar_wait(b_full, bph, p.err_flag, 3);
            bph ^= 1;
            for (int j = j0; j < j1; ++j) {
                const uint64_t adesc = adesc0 + (uint64_t)(s * (A_MAIN_BYTES >> 4));
                const uint32_t d_tmem = tmem_base + (uint32_t)as * BN;
                mbar_wait(tempty_bar(as), aph ^ 1, p.err_flag, 4);  // epilogue drained this accumulator stage
                mbar_wait(full_bar(s), ph, p.err_flag, 5);          // TMA bytes have landed
                tcgen05_fence_after();
                if (issuer) {
#pragma unroll
                    for (int k = 0; k < KT / 16; ++k)  // +32 bytes per K=16 step inside the 128-byte swizzle atom
                        umma_f16(d_tmem, adesc + 2 * k, bdesc + 2 * k, IDESC, k > 0);
                    tcgen05_commit(empty_bar(s));   // frees the A stage once the MMAs above have completed
                    tcgen05_commit(tfull_bar(as));  // accumulator ready for the epilogue
                }
                __syncwarp();
                if (++s == STAGES) { s = 0; ph ^= 1; }
                if (++as == 2) { as = 0; aph ^= 1; }
            }
            if (issuer) tcgen05_commit(b_empty);
            __syncwarp();
        }
    } else if (warp < EPI_WARPS) {
        // ===== epilogue: TMEM -> registers -> screen -> (rare) survivor append =====
        // Two warp groups of WARPS_PER_STAGE warps; group g owns accumulator stage g, i.e. every second tile of the CTA's
        // span, so the two groups work on different tiles at the same time and only ever wait on their own stage.  Inside
        // a group two warps share each TMEM lane quadrant and take 128 columns (16 ranges) each.
        const int ew = warp;
        const int stage = ew / WARPS_PER_STAGE;              // the accumulator stage this warp owns
        const int quad = warp & 3;                           // TMEM lane quadrant this warp may touch
        const int part = (ew % WARPS_PER_STAGE) >> 2;        // which 128-column half of the 256 columns
        const int lane_off = quad * 32 + lane;
        // address of this warp's slice of its accumulator stage; pinned so it is not rebuilt from %tid after the wait
        uint32_t taddr0 = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(stage * BN + part * 128);
        asm volatile("" : "+r"(taddr0));
        const uint32_t my_tfull = tfull_bar(stage), my_tempty = tempty_bar(stage);
        uint32_t aph = 0;
        int n = 0;  // tiles issued so far by this CTA (all units): tile n uses stage n & 1
        float* my_tau = reinterpret_cast<float*>(gbase + OFF_TAU) + ew * 32;   // [0,16) = tau_A, [16,32) = tau_B
        WarpQueue* queue = reinterpret_cast<WarpQueue*>(gbase + OFF_QUEUE) + ew;
        uint32_t qn = 0;   // queue fill, warp-uniform
        int qt_slot0 = 0;  // MODE 7: first range of the unit the queued entries belong to
        Span span(p, n_groups);
        int grp, j0, j1;
        while (span.next(grp, j0, j1)) {
            grp += grp_off;
            if (MODE == 0 && qn) { queue_flush(queue, p, lane, qn); qn = 0; }
            if (MODE == 7 && qn) { qt_queue_flush<KT>(reinterpret_cast<QtQueue*>(queue), p, qt_slot0, lane, qn); qn = 0; }   // (entries name ranges relative to their unit)
            if (MODE == 8 && qn) { bf4_queue_flush(reinterpret_cast<Bf4Queue*>(queue), p, qt_slot0, lane, qn); qn = 0; }
            // this warp's (tau_A, tau_B) pairs live in a private slice of shared memory; broadcast reads, no bank conflicts
            __syncwarp();
            if (MODE != 7 && MODE != 8 && lane < 16) {
                const float2 tt = p.tau[grp * 32 + part * 16 + lane];
                my_tau[lane] = tt.x;
                my_tau[16 + lane] = tt.y;
            }
            __syncwarp();
            const int slot0 = MODE == 7 ? grp * QT_RPG + part * QT_RPW : grp * 32 + part * 16;
            qt_slot0 = slot0;
            float my_rho = 0.f;
            double my_slack = -1.0;
            if (MODE == 7) {   // lane r keeps what the test needs of the r-th range of this warp's 128 columns
                const QtcRange rc = p.rngc[grp * QT_RPG + part * QT_RPW + lane];
                my_rho = (float)rc.rho;
                my_slack = rc.ok ? rc.slack : -1.0;
            }
            bool my_ok = false;
            if (MODE == 8) {   // sixteen ranges of eight columns: lanes r and r + 16 keep range r
                const QtcRange rc = p.rngc[slot0 + (lane & 15)];
                my_rho = (float)rc.rho;
                my_slack = rc.slack;       // (may be negative: the NumPy encoder's sum of squares wraps at 256)
                my_ok = (rc.ok & 1) != 0;
            }
            int j = j0 + ((n ^ stage) & 1);   // first tile of this unit that is mine
            n += j1 - j0;
            // tile index of j without a division per tile: t = (j + j / (r_prev - 1) + 1) * m_cur, advanced by 2
            // The loop carries this lane's candidate index (tile * 128 + quadrant * 32 + lane) rather than the tile:
            // the lane is then a mask away from a live register where it is needed (release, queue slot) and the
            // compiler has less reason to re-read %tid.x -- a ~50-cycle S2R -- in the middle of a tile.
            int cand = (j < j1 ? tile_of(p, j) : 0) * BM + lane_off;
            const int r1 = p.r_prev > 0 ? p.r_prev - 1 : 0x7fffffff;
            int rem = p.r_prev > 0 && j < j1 ? j % r1 : 0;
            // +-1/h_c of this lane's candidate, fetched one tile ahead; sign = conditioning class; +inf (threshold never
            // met) for the padding rows of the last tile
            const float hk_pad = __int_as_float(0x7f800000);
            float hq_next = (MODE != 7 && MODE != 8 && j < j1 && cand < p.NP) ? p.hk[cand].x : hk_pad;
            // the sixteen tau of the next tile (class picked by the sign of its hq) are fetched at the END of the current
            // tile, when both chunk buffers are dead, so the shared-memory latency is paid under the index arithmetic
            float4 tq0, tq1, tq2, tq3;
            {
                const float4* tau4 = reinterpret_cast<const float4*>(my_tau + (hq_next < 0.f ? 16 : 0));
                tq0 = tau4[0]; tq1 = tau4[1]; tq2 = tau4[2]; tq3 = tau4[3];
            }
            float seed_gain[16];   // seed mode only: this lane's best candidate per range over the unit's tiles
            int seed_cand[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) { seed_gain[i] = 0.f; seed_cand[i] = 0; }
            for (; j < j1; j += 2) {
                // ---- everything up to the wait is off the accumulator's critical path (MMA -> epilogue -> MMA) ----
                const int my_cand = cand;
                {   // advance to tile j + 2:  t = (j + j / (r_prev - 1) + 1) * m_cur without a division per tile
                    rem += 2;
                    int wraps = 0;
                    if (rem >= r1) { rem -= r1; wraps = 1; }
                    if (rem >= r1) { rem -= r1; wraps = 2; }   // r1 == 1 (r_prev == 2)
                    cand += (2 + wraps) * p.m_cur * BM;
                }
                const float hq = hq_next;
                hq_next = (MODE != 7 && MODE != 8 && j + 2 < j1 && cand < p.NP) ? p.hk[cand].x : hk_pad;
                const int lane = my_cand & 31;            // shadows the outer lane on purpose
                const uint32_t lane0 = lane == 0;
                if (MODE == 5) {
                    // seed mode: no thresholds yet.  Rank the (few) candidates of the seed pool by the collage-error
                    // reduction the exact-arithmetic optimum would give (contrast clamped to |s| <= 1) and keep the best
                    // one per range; it is scored exactly afterwards and provides the first upper bound.  Every lane keeps
                    // its own best per range over the tiles of the unit; the warps merge them once, after the last tile.
                    const float den = my_cand < p.NP ? p.hk[my_cand].y : 0.f;
                    const float inv64 = den > 0.f ? 64.f / den : 0.f, den1024 = den * (1.0f / 1024.0f);
                    uint32_t va[32], vb[32];
                    auto rank4 = [&](const uint32_t (&v)[32], int c4) {
#pragma unroll
                        for (int g = 0; g < 4; ++g) {
                            float m = fabsf(__uint_as_float(v[g * 8]));
#pragma unroll
                            for (int i = 1; i < 8; ++i) m = fmaxf(m, fabsf(__uint_as_float(v[g * 8 + i])));
                            const float gain = 256.f * m <= den ? m * m * inv64 : 0.5f * m - den1024;   // den == 0: m == 0
                            if (gain > seed_gain[c4 * 4 + g]) { seed_gain[c4 * 4 + g] = gain; seed_cand[c4 * 4 + g] = my_cand; }
                        }
                    };
                    mbar_wait(my_tfull, aph, p.err_flag, 6);
                    tcgen05_fence_after();
                    tmem_ld32(taddr0, va);
                    tmem_ld32(taddr0 + 32, vb);
                    tmem_ld_wait(va);
                    tmem_ld_wait(vb);
                    rank4(va, 0);
                    tmem_ld32(taddr0 + 64, va);
                    rank4(vb, 1);
                    tmem_ld32(taddr0 + 96, vb);
                    tmem_ld_wait(va);
                    tmem_ld_wait(vb);
                    tcgen05_fence_before();
                    __syncwarp();
                    if (lane0) mbar_arrive(my_tempty);
                    rank4(va, 2);
                    rank4(vb, 3);
                    aph ^= 1;
                    continue;
                }
                if (MODE == 8) {
                    // Brute force, 4x4 ranges: the eight accumulators of a (candidate, range) pair are the exact dot products
                    // of the eight isometries.  Same test as the quadtree pass (integer numerator N * X - rho * eps, float32
                    // comparison inside the 1e-6 the right-hand side is shrunk by), here for every column.  Masks first, for the
                    // four ranges of a 32-column chunk; the queue is filled from the masks afterwards, so that no accumulator
                    // is live across the (out of line) flush.
                    const bool have = my_cand < p.NP;
                    unsigned sdc = 0;
                    float denf = 0.f;            // 0: this candidate is never dropped (degenerate block)
                    if (have) {
                        sdc = p.sd4[my_cand];
                        const double den = p.den[my_cand];
                        denf = den > 0.0 ? (float)den : 0.f;
                    }
                    const float neps = -(float)((int)sdc - KT * (int)((sdc + (unsigned)(KT / 2)) / (unsigned)KT));
                    float my_rhs;
                    {
                        const double B = __longlong_as_double((long long)__ldcg(&p.bound[slot0 + (lane & 15)]));
                        const double rhs_g = DM(DS(my_slack, DM((double)KT, B)), 16.0 * KT * 0.999999);
                        my_rhs = !my_ok ? __int_as_float(0x7f800000) : (rhs_g > 0.0 ? (float)rhs_g : -1.0f);
                    }
                    Bf4Queue* qq = reinterpret_cast<Bf4Queue*>(queue);
                    const unsigned lt = (1u << lane) - 1u;
                    const uint32_t ent_cand = (uint32_t)my_cand << 7;
                    mbar_wait(my_tfull, aph, p.err_flag, 6);
                    tcgen05_fence_after();
#pragma unroll 1
                    for (int c4 = 0; c4 < 4; ++c4) {
                        unsigned m8[4];
                        {
                            uint32_t v[32];
                            tmem_ld32(taddr0 + 32 * c4, v);
                            tmem_ld_wait(v);
                            if (c4 == 3) {     // the whole slice has been read: hand the accumulator back
                                tcgen05_fence_before();
                                __syncwarp();
                                if (lane0) mbar_arrive(my_tempty);
                            }
#pragma unroll
                            for (int g = 0; g < 4; ++g) {
                                const int r = c4 * 4 + g;
                                const float rhs = __fmul_rn(__shfl_sync(0xffffffffu, my_rhs, r), denf);
                                const float nre = __fmul_rn(__shfl_sync(0xffffffffu, my_rho, r), neps);
                                // the eight numerators; the largest magnitude decides for all of them in the common case
                                float na[8];
#pragma unroll
                                for (int t = 0; t < 8; ++t) na[t] = fmaf((float)KT, __uint_as_float(v[g * 8 + t]), nre);
                                const float mx = fmaxf(fmaxf(fmaxf(fabsf(na[0]), fabsf(na[1])), fmaxf(fabsf(na[2]), fabsf(na[3]))),
                                                       fmaxf(fmaxf(fabsf(na[4]), fabsf(na[5])), fmaxf(fabsf(na[6]), fabsf(na[7]))));
                                unsigned m = 0u;
                                if (have && !(__fmul_rn(mx, mx) < rhs)) {
#pragma unroll
                                    for (int t = 0; t < 8; ++t) m |= (__fmul_rn(na[t], na[t]) < rhs ? 0u : 1u) << t;
                                }
                                m8[g] = m;
                            }
                        }
#pragma unroll
                        for (int g = 0; g < 4; ++g) {
                            if (!__any_sync(0xffffffffu, m8[g] != 0u)) continue;
                            const uint32_t ent = ent_cand | (uint32_t)((c4 * 4 + g) << 3);
#pragma unroll
                            for (int t = 0; t < 8; ++t) {
                                const bool s1 = (m8[g] >> t) & 1u;
                                const unsigned mb = __ballot_sync(0xffffffffu, s1);
                                if (s1) qq->e[qn + __popc(mb & lt)] = ent | (uint32_t)t;
                                qn += __popc(mb);
                            }
                        }
                        if (qn > (uint32_t)(BF4_QCAP - 4 * 8 * 32)) { bf4_queue_flush(qq, p, slot0, lane, qn); qn = 0; }
                    }
                    aph ^= 1;
                    continue;
                }
                if (MODE == 7) {
                    // Exact pass of the quadtree search.  The accumulator of a (candidate, range, scale sign) pair -- the exact
                    // integer sum X of the centred range times the centred block under the pair's isometry -- sits in the
                    // column its scale sign and the orientation bit of the candidate's class select (see QT_ALPHA).  With it
                    // 4 * (N * ab - rs * ds) = N * X - rho * eps is an integer, and the pair is dropped when its square is
                    // below 16 * N * den * (slack - N * B), the scorer's lower bound against the best score B of the range
                    // with both sides multiplied out and the right one shrunk by 1e-6.  The comparison runs in float32: X is
                    // an integer below 2^24 and N a power of two, so fmaf(N, X, -rho * eps) is the numerator rounded once;
                    // its square, the two converted factors of the right-hand side and their product add five more
                    // roundings, 4e-7 relative in all.  Everything not dropped is queued and scored exactly.
                    const bool have = my_cand < p.NP;
                    unsigned dty = 0, sdc = 0;
                    float denf = 0.f;            // 0: this candidate is never dropped (degenerate block)
                    if (have) {
                        dty = p.dtyp[my_cand]; sdc = p.sd4[my_cand];
                        const double den = p.den[my_cand];
                        denf = den > 0.0 ? (float)den : 0.f;
                    }
                    const float neps = -(float)((int)sdc - KT * (int)((sdc + (unsigned)(KT / 2)) / (unsigned)KT));
                    const bool dori = (dty & 4u) != 0u;
                    const bool two = p.nmodes == 2;
                    // lane r: the right-hand side of range r without the candidate's factor; +inf = drop everything (range not
                    // searched here), -1 = drop nothing (bound not yet below the range's variance)
                    float my_rhs;
                    {
                        const double B = __longlong_as_double((long long)__ldcg(&p.bound[slot0 + lane]));
                        const double rhs_g = DM(DS(my_slack, DM((double)KT, B)), 16.0 * KT * 0.999999);
                        my_rhs = my_slack < 0.0 ? __int_as_float(0x7f800000) : (rhs_g > 0.0 ? (float)rhs_g : -1.0f);
                    }
                    QtQueue* qq = reinterpret_cast<QtQueue*>(queue);
                    const unsigned lt = (1u << lane) - 1u;
                    const uint32_t ent_cand = (uint32_t)my_cand << 6;
                    mbar_wait(my_tfull, aph, p.err_flag, 6);
                    tcgen05_fence_after();
                    // One 32-column chunk (eight ranges) in registers at a time, in a rolled loop: the queue is emptied
                    // between chunks, where no accumulator is live across the call, and from one call site.
#pragma unroll 1
                    for (int c8 = 0; c8 < 4; ++c8) {
                        uint32_t v[32];
                        tmem_ld32(taddr0 + 32 * c8, v);
                        tmem_ld_wait(v);
                        if (c8 == 3) {     // the whole slice has been read: hand the accumulator back
                            tcgen05_fence_before();
                            __syncwarp();
                            if (lane0) mbar_arrive(my_tempty);
                        }
#pragma unroll
                        for (int g = 0; g < 8; ++g) {
                            const int r = c8 * 8 + g;
                            const float rhs = __fmul_rn(__shfl_sync(0xffffffffu, my_rhs, r), denf);
                            const float nre = __fmul_rn(__shfl_sync(0xffffffffu, my_rho, r), neps);
                            const uint32_t xa = dori ? v[g * 4 + 1] : v[g * 4 + 0], xb = dori ? v[g * 4 + 3] : v[g * 4 + 2];
                            const float na = fmaf((float)KT, __uint_as_float(xa), nre), nb = fmaf((float)KT, __uint_as_float(xb), nre);
                            const bool sa = have && !(__fmul_rn(na, na) < rhs), sb = two && have && !(__fmul_rn(nb, nb) < rhs);
                            if (__any_sync(0xffffffffu, sa || sb)) {
                                const unsigned ma = __ballot_sync(0xffffffffu, sa);
                                if (sa) qq->e[qn + __popc(ma & lt)] = make_uint2(xa, ent_cand | (uint32_t)(r << 1));
                                qn += __popc(ma);
                                const unsigned mb = __ballot_sync(0xffffffffu, sb);
                                if (sb) qq->e[qn + __popc(mb & lt)] = make_uint2(xb, ent_cand | (uint32_t)(r << 1) | 1u);
                                qn += __popc(mb);
                            }
                        }
                        if (qn > (uint32_t)(QT_QCAP - QT_QROOM)) { qt_queue_flush<KT>(qq, p, slot0, lane, qn); qn = 0; }
                    }
                    aph ^= 1;
                    continue;
                }
                const float qx = fabsf(hq);
                // thresholds on |u| for the 16 ranges: theta = tau/h_c - DELTA, tau picked by the candidate's class
                float theta[16];
                theta[0] = fmaf(tq0.x, qx, -DELTA_U); theta[1] = fmaf(tq0.y, qx, -DELTA_U);
                theta[2] = fmaf(tq0.z, qx, -DELTA_U); theta[3] = fmaf(tq0.w, qx, -DELTA_U);
                theta[4] = fmaf(tq1.x, qx, -DELTA_U); theta[5] = fmaf(tq1.y, qx, -DELTA_U);
                theta[6] = fmaf(tq1.z, qx, -DELTA_U); theta[7] = fmaf(tq1.w, qx, -DELTA_U);
                theta[8] = fmaf(tq2.x, qx, -DELTA_U); theta[9] = fmaf(tq2.y, qx, -DELTA_U);
                theta[10] = fmaf(tq2.z, qx, -DELTA_U); theta[11] = fmaf(tq2.w, qx, -DELTA_U);
                theta[12] = fmaf(tq3.x, qx, -DELTA_U); theta[13] = fmaf(tq3.y, qx, -DELTA_U);
                theta[14] = fmaf(tq3.z, qx, -DELTA_U); theta[15] = fmaf(tq3.w, qx, -DELTA_U);
#pragma unroll
                for (int i = 0; i < 16; ++i) asm volatile("" : "+f"(theta[i]));   // materialise before the wait
                mbar_wait(my_tfull, aph, p.err_flag, 6);
                tcgen05_fence_after();
                uint32_t va[32], vb[32];
                if (MODE == 1) {
                    for (int c4 = 0; c4 < 4; ++c4) {
                        tmem_ld32(taddr0 + c4 * 32, va);
                        tmem_ld_wait(va);
                        if (grp == 0) {
#pragma unroll
                            for (int q = 0; q < 32; ++q)
                                p.dump[(size_t)my_cand * BN + part * 128 + c4 * 32 + q] = __uint_as_float(va[q]);
                        }
                    }
                    tcgen05_fence_before();
                    __syncwarp();
                    if (lane0) mbar_arrive(my_tempty);
                } else if (MODE == 3) {
                    tcgen05_fence_before();
                    __syncwarp();
                    if (lane0) mbar_arrive(my_tempty);
                } else {
                    constexpr bool SCREEN = MODE == 0 || MODE == 4;
                    tmem_ld32(taddr0, va);
                    tmem_ld32(taddr0 + 32, vb);
                    tmem_ld_wait(va);
                    tmem_ld_wait(vb);
                    if (SCREEN) screen_chunk<MODE == 0>(va, theta + 0, p, slot0, my_cand, queue, qn, lane);
                    tmem_ld32(taddr0 + 64, va);        // in flight while chunk 1 is screened
                    if (SCREEN) screen_chunk<MODE == 0>(vb, theta + 4, p, slot0 + 4, my_cand, queue, qn, lane);
                    tmem_ld32(taddr0 + 96, vb);
                    tmem_ld_wait(va);
                    tmem_ld_wait(vb);
                    // the whole slice is in registers now: hand the accumulator back before screening the second half
                    tcgen05_fence_before();
                    __syncwarp();
                    if (lane0) mbar_arrive(my_tempty);
                    if (SCREEN) screen_pair<MODE == 0>(va, vb, theta + 8, p, slot0 + 8, my_cand, queue, qn, lane);
                    if (!SCREEN && (va[0] == 0x7fc12345u || vb[0] == 0x7fc12345u)) p.cnt[0] = 1;
                }
                aph ^= 1;
                {   // next tile's tau (see above); harmless after the last tile
                    const float4* tau4 = reinterpret_cast<const float4*>(my_tau + (hq_next < 0.f ? 16 : 0));
                    tq0 = tau4[0]; tq1 = tau4[1]; tq2 = tau4[2]; tq3 = tau4[3];
                }
                if (MODE == 0 && qn >= QFLUSH) { queue_flush(queue, p, lane, qn); qn = 0; }
            }
            if (MODE == 5) {   // merge the lanes' seed candidates of this unit: warp maximum, then one atomic per range
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const uint32_t bits = seed_gain[i] > 0.f ? __float_as_uint(seed_gain[i]) : 0u;
                    const uint32_t wmax = __reduce_max_sync(0xffffffffu, bits);
                    if (wmax != 0u && bits == wmax)   // ties (rare): every tied lane offers its candidate, the smaller index wins
                        atomicMax(&p.seed_best[slot0 + i], ((unsigned long long)bits << 32) | (0xffffffffu - (uint32_t)seed_cand[i]));
                }
            }
        }
        if (MODE == 0 && qn) queue_flush(queue, p, lane, qn);
        if (MODE == 7 && qn) qt_queue_flush<KT>(reinterpret_cast<QtQueue*>(queue), p, qt_slot0, lane, qn);
        if (MODE == 8 && qn) bf4_queue_flush(reinterpret_cast<Bf4Queue*>(queue), p, qt_slot0, lane, qn);
        }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == MMA_WARP) {
        tcgen05_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
    }
}

// ---- pool kernel: 2x pooling + centring + statistics, one thread per stride-1 position ------------------------
// Also the per-position statistics of the exact scorer (the same values bf_stats_kernel of bf_exact.cu produces: this
// path skips that launch): sd4 = sum D4, sq4 = sum D4^2, sdd[t] = float32 sequential sum of d^2 in the order of isometry t.
// HBM-bound: 176 bytes written per position (128 B pool row, 8 B hk, 8 B integer sums, 32 B float sums) against 2 bytes of
// box sums read.  A CTA covers POOL_TILE consecutive positions of one lattice row: the eight box-sum rows it needs are
// staged in shared memory once (each value is used by up to eight positions), and the pool rows -- 128 B per thread,
// which would leave every store instruction touching 32 different lines -- go through a swizzled shared-memory tile and
// leave the SM as one contiguous 16 KB block, 512 B per warp instruction.
constexpr int POOL_TILE = 128;
constexpr int POOL_SP_W = POOL_TILE + 14;      // box-sum columns a tile of positions touches

__global__ void __launch_bounds__(POOL_TILE)
bf_pool_kernel(const uint16_t* __restrict__ P, int H, int W, __half* __restrict__ poolA, float2* __restrict__ hk,
               uint32_t* __restrict__ sd4, uint32_t* __restrict__ sq4, float* __restrict__ sdd, int seed_S, int seed_gx,
               int n_seed, int n_seed_pad, __half* __restrict__ seedA, float2* __restrict__ seedhk)
{
    __shared__ uint16_t sP[8][POOL_SP_W + 2];
    __shared__ __align__(16) uint4 sRow[POOL_TILE * 8];     // 128 pool rows of 8 x 16 B, chunk q of row t at q ^ (t & 7)
    const int PW = W - 16, PH = H - 16, BW = W - 1;
    const int NP = PW * PH;
    const int tid = threadIdx.x;
    const int di = blockIdx.y, dj0 = blockIdx.x * POOL_TILE, dj = dj0 + tid;
    const bool live = dj < PW;
    const int p = di * PW + dj;
    for (int e = tid; e < 8 * POOL_SP_W; e += POOL_TILE) {
        const int a = e / POOL_SP_W, c = e - a * POOL_SP_W, x = dj0 + c;
        sP[a][c] = x < BW ? P[(size_t)(di + 2 * a) * BW + x] : (uint16_t)0;
    }
    // seed pool: the lattice nodes (every seed_S-th row and column) get a second copy in a small pool of their own; seed
    // candidate k sits at node (k / gx, k % gx); rows n_seed .. n_seed_pad are padding (zero rows, den16 = 0)
    int seed_k = -1;
    if (seedA) {
        if (live && di % seed_S == 0 && dj % seed_S == 0) seed_k = (di / seed_S) * seed_gx + dj / seed_S;
        const int lin = (blockIdx.y * gridDim.x + blockIdx.x) * POOL_TILE + tid;
        if (lin < n_seed_pad - n_seed) {
            uint4* za = reinterpret_cast<uint4*>(seedA + (size_t)(n_seed + lin) * 64);
#pragma unroll
            for (int q = 0; q < 8; ++q) za[q] = make_uint4(0u, 0u, 0u, 0u);
            seedhk[n_seed + lin] = make_float2(__int_as_float(0x7f800000), 0.f);
        }
    }
    __syncthreads();
    int v[64];
    int s = 0;
    long long s2 = 0;
#pragma unroll
    for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            int x = sP[a][tid + 2 * b];
            v[a * 8 + b] = x;
            s += x;
            s2 += x * x;
        }
    const int m = (s + 32) >> 6;  // round(SD4 / 64); eps = s - 64 m in [-32, 31] is what DELTA_U covers
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        __half2 h0 = __halves2half2(__int2half_rn(v[8 * q + 0] - m), __int2half_rn(v[8 * q + 1] - m));
        __half2 h1 = __halves2half2(__int2half_rn(v[8 * q + 2] - m), __int2half_rn(v[8 * q + 3] - m));
        __half2 h2 = __halves2half2(__int2half_rn(v[8 * q + 4] - m), __int2half_rn(v[8 * q + 5] - m));
        __half2 h3 = __halves2half2(__int2half_rn(v[8 * q + 6] - m), __int2half_rn(v[8 * q + 7] - m));
        uint4 o;
        o.x = *reinterpret_cast<uint32_t*>(&h0);
        o.y = *reinterpret_cast<uint32_t*>(&h1);
        o.z = *reinterpret_cast<uint32_t*>(&h2);
        o.w = *reinterpret_cast<uint32_t*>(&h3);
        sRow[tid * 8 + (q ^ (tid & 7))] = o;
        if (seed_k >= 0) reinterpret_cast<uint4*>(seedA + (size_t)seed_k * 64)[q] = o;
    }
    __syncthreads();
    {   // the tile's rows are contiguous in the pool: positions di*PW + dj0 .. + n_live - 1
        const int n_live = min(POOL_TILE, PW - dj0);
        uint4* dst = reinterpret_cast<uint4*>(poolA + (size_t)(di * PW + dj0) * 64);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int e = k * POOL_TILE + tid, r = e >> 3, c = e & 7;
            if (r < n_live) dst[e] = sRow[r * 8 + (c ^ (r & 7))];
        }
    }
    if (!live) return;
    const long long den16 = 64ll * s2 - (long long)s * s;  // >= 0, exact
    // q_c = 1/h_c, rounded DOWN a little so that theta = tau*q - DELTA never over-estimates; +inf for a constant domain
    // (den16 == 0: u == 0 exactly, s_ref == 0, R_ref == A_r/64 without noise -> it can never beat tau_A > 0)
    float hc = __int_as_float(0x7f800000), kc = 0.f;
    if (den16 > 0) {
        const double d = (double)den16, q = (double)s2;
        const bool well = d >= 5e-4 * q;
        const double h2 = well ? 64.0 / d * (1.0 + 1.0e-3 * q / d) : 64.0 / d;
        hc = (float)(1.0 / sqrt(h2)) * (1.0f - 1.0f / 1048576.0f);
        if (!well) hc = -hc;                                   // sign flags the ill-conditioned class
        kc = (float)d;                                         // den16, only used to rank seed candidates
    }
    hk[p] = make_float2(hc, kc);
    if (seed_k >= 0) seedhk[seed_k] = make_float2(hc, kc);
    sd4[p] = (uint32_t)s;
    sq4[p] = (uint32_t)s2;
    float sq[64];
#pragma unroll
    for (int k = 0; k < 64; ++k) {
        const float d = __fmul_rn((float)v[k], 0.25f);
        sq[k] = __fmul_rn(d, d);  // exact: v^2/16 < 2^24 significant bits
    }
    // Eight sequential float32 sums of 64 terms, one per isometry, each in its own order: 512 dependent additions are what
    // this kernel spends its time on (it is bound by them, not by HBM).  Fully unrolled with the eight chains advancing
    // together: the index maps are compile-time constants, sq[] stays in registers, eight independent additions per step.
    float acc[8];
#pragma unroll
    for (int t = 0; t < 8; ++t) acc[t] = 0.f;
#pragma unroll
    for (int k = 0; k < 64; ++k)
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            int sa, sb;
            fic_iso_bf(t >> 2, t & 3, 8, k >> 3, k & 7, sa, sb);
            acc[t] = __fadd_rn(acc[t], sq[sa * 8 + sb]);
        }
#pragma unroll
    for (int t = 0; t < 8; ++t) sdd[(size_t)t * NP + p] = acc[t];
}

// ---- per-range state --------------------------------------------------------------------------------------------
// The running winner of every range (exact R, scan index, s, o), seeded by the coarse pre-scan, improved by every
// rescoring round; plus the bound ingredients.
struct RangeState {
    double R, s, o, TA0;           // TA0 = A_r - 64e-9*sum r^2  (threshold numerator before the bound is applied)
    unsigned long long seq;        // position*8 + isometry of the running winner
    int ridx;                      // absolute range index
    int flags;                     // 1 = sent to the exact scan, 2 = constant range (decided)
    int rsum, pad_;                // sum of the range's 64 pixels
};

// ctrl[0] = n_elig, ctrl[1] = n_fallback, ctrl[2] = error flag, ctrl[8..9] = survivors rescored (u64)
// One warp per slot (the slots are padded to whole column groups).  Every range of the slice gets a slot in natural
// order; its sum and sums of squares come from a warp reduction over the 64 pixels.  Nothing is known about the winners
// yet: R = +inf, no candidate, thresholds +inf (nothing passes), empty survivor list; constant ranges are flagged, they
// are decided by the seed list alone (every candidate scores R = 0 exactly, the first in scan order -- position 0,
// isometry 0 -- wins).
// NumPy mode: bf_encoder.py:33 squares the uint8 pixels in uint8, every term wraps mod 256, so the reference's R carries
// the wrapped sum of squares, R_ref = R_true - (q - qw)/64 for every candidate of the range alike; the bound is on
// R_true, so the shift goes into the threshold numerator once.
__global__ void bf_init_kernel(const uint8_t* __restrict__ img, int W, int mode, int range_begin, int n_ranges,
                               RangeState* __restrict__ state, float2* __restrict__ tau, uint32_t* __restrict__ cnt,
                               unsigned long long* __restrict__ seed_best, int* __restrict__ ctrl, int tau_pad,
                               __half* __restrict__ rngB)
{
    const int slot = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (slot >= tau_pad) return;
    if (slot >= n_ranges) {   // padding columns of the last group: zero operand
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            rngB[((size_t)slot * 8 + t) * 64 + lane] = __float2half_rn(0.f);
            rngB[((size_t)slot * 8 + t) * 64 + 32 + lane] = __float2half_rn(0.f);
        }
    }
    if (slot == 0 && lane < 16) ctrl[lane] = lane == 0 ? n_ranges : 0;
    if (lane == 0) {
        tau[slot] = make_float2(__int_as_float(0x7f800000), __int_as_float(0x7f800000));
        seed_best[slot] = 0ull;
    }
    if (slot >= n_ranges) return;
    const int rw = W / 8, ridx = range_begin + slot;
    const int ri = (ridx / rw) * 8, rj = (ridx % rw) * 8;
    const unsigned r0 = img[(size_t)(ri + (lane >> 3)) * W + rj + (lane & 7)];
    const unsigned r1 = img[(size_t)(ri + 4 + (lane >> 3)) * W + rj + (lane & 7)];
    unsigned sm = r0 + r1, q32 = r0 * r0 + r1 * r1, qw = ((r0 * r0) & 255u) + ((r1 * r1) & 255u);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        sm += __shfl_xor_sync(0xffffffffu, sm, off);
        q32 += __shfl_xor_sync(0xffffffffu, q32, off);
        qw += __shfl_xor_sync(0xffffffffu, qw, off);
    }
    {
        // range operand (rngB fp16 [cols][64], column = slot * 8 + isometry): eight inverse-transformed copies of the
        // range centred by the integer m = round(mean) (rho = sum - 64 m in [-32, 31] is what DELTA_U covers).
        // Entry k = (sa,sb) of column t holds r'[i][j] with (sa,sb) = iso_t(i,j): invert fic_iso_bf.
        const int m = ((int)sm + 32) >> 6;
#pragma unroll
        for (int t = 0; t < 8; ++t) {
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                const int k = lane + 32 * half, sa = k >> 3, sb = k & 7;
                const int a = (t >> 2) ? sb : sa, b = (t >> 2) ? sa : sb;  // undo the transpose
                int i, j;
                switch (t & 3) {
                case 0: i = a; j = b; break;
                case 1: j = a; i = 7 - b; break;          // a = j, b = 7 - i
                case 2: i = 7 - a; j = 7 - b; break;
                default: j = 7 - a; i = b; break;         // a = 7 - j, b = i
                }
                rngB[((size_t)slot * 8 + t) * 64 + k] = __int2half_rn((int)img[(size_t)(ri + i) * W + rj + j] - m);
            }
        }
    }
    if (lane == 0) {
        const long long s = sm, q = q32;
        const long long A64 = 64 * q - s * s;                             // 64 * A_r, exact
        RangeState st;
        st.R = __longlong_as_double(0x7ff0000000000000ll); st.s = 0.0; st.o = 0.0;
        st.TA0 = (double)A64 * 0.015625 - 64e-9 * (double)q;
        if (mode == FIC_BF_MODE_NUMPY) st.TA0 -= (double)(q - (long long)qw);
        st.seq = ~0ull; st.ridx = ridx; st.flags = A64 == 0 ? 2 : 0;
        st.rsum = (int)s; st.pad_ = 0;
        state[slot] = st;
        cnt[slot] = 0;
    }
}

// ---- exact rescoring of the survivors: one warp per eligible range, merges into the range state -------------------
// The survivor's pooled block is read back from the TMA pool (one 128-byte row, D4 = D4' + m_c), the dot product is
// an exact integer sum in fp32, the score is the float64 restatement of jit_bfencoder.py:64-95 (scorer.cuh).
// Ranges with a short list (the common case) are handled by their own warp; the few with a long list are then
// worked through by the whole CTA one after the other, so a single crowded range cannot become a 256-iteration tail.
#define RESCORE_WARPS 16    // ranges per CTA (512 threads, two CTAs per SM): same-box A/B against 32 / 1 and 8 / 4: -0.5 % / +0.4 % of the step
#define RESCORE_MINB 2
#define RESCORE_SMEM (RESCORE_WARPS * 8 * 65 * 4)   // dynamic: the oriented range copies

struct RescoreBest {
    double R, s, o;
    unsigned long long seq;
};

// order-preserving map of a double onto uint64 (R may be negative with the NumPy encoder's wrapped sum of squares);
// -0.0 and +0.0 compare equal in the reference, so they share a key
__device__ __forceinline__ unsigned long long rescore_key(double R)
{
    const long long bits = __double_as_longlong(R + 0.0);
    return bits < 0 ? ~(unsigned long long)bits : (unsigned long long)bits | 0x8000000000000000ull;
}

__device__ __forceinline__ void rescore_entry(uint32_t ent, const float (*rvw)[65], double rs, double sqr,
                                              const __half* __restrict__ poolA, const uint32_t* __restrict__ sd4,
                                              const float* __restrict__ sdd, int NP, int mode, RescoreBest& b)
{
    const int t = ent & 7;
    const int p = ent >> 3;
    const uint32_t sd = sd4[p];
    const float mc = (float)((sd + 32u) >> 6);          // the pool stores D4 - m_c
    const uint4* row = reinterpret_cast<const uint4*>(poolA + (size_t)p * 64);
    const float* rvt = rvw[t];
    float x = 0.f;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const uint4 u = row[c];
        const __half2* h2 = reinterpret_cast<const __half2*>(&u);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float2 f = __half22float2(h2[k]);
            x = fmaf(rvt[c * 8 + 2 * k], f.x + mc, x);        // exact: integers, every partial sum < 2^24
            x = fmaf(rvt[c * 8 + 2 * k + 1], f.y + mc, x);
        }
    }
    const float dsf = __fmul_rn((float)sd, 0.25f);
    BfScore sc;
    if (mode == FIC_BF_MODE_JIT)
        sc = fic_bfjit_score(64.0, 0.015625, rs, sqr, (double)dsf, (double)sdd[(size_t)t * NP + p],
                             (double)__fmul_rn(dsf, dsf), DM((double)x, 0.25));
    else   // bf_encoder.py:56-71: float64 throughout, exact sums (sq4 = sum of D4^2 sits behind sd4)
        sc = fic_bfnp_score(64.0, 0.015625, rs, sqr, DM((double)sd, 0.25), DM((double)sd4[(size_t)NP + p], 0.0625),
                            DM((double)x, 0.25));
    if (fic_better(sc.R, (unsigned long long)ent, b.R, b.seq)) { b.R = sc.R; b.seq = ent; b.s = sc.s; b.o = sc.o; }
}

__device__ __forceinline__ void rescore_warp_reduce(RescoreBest& b)
{
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        double oR = __shfl_xor_sync(0xffffffffu, b.R, off);
        unsigned long long oS = __shfl_xor_sync(0xffffffffu, b.seq, off);
        double os = __shfl_xor_sync(0xffffffffu, b.s, off), oo = __shfl_xor_sync(0xffffffffu, b.o, off);
        if (fic_better(oR, oS, b.R, b.seq)) { b.R = oR; b.seq = oS; b.s = os; b.o = oo; }
    }
}

// one thread: store the winner of `slot` (and emit the record in the last round)
__device__ void rescore_commit(int slot, int ridx, const RescoreBest& b, RangeState* __restrict__ state, int* __restrict__ ctrl,
                               uint32_t* __restrict__ cnt, int32_t* __restrict__ fallback, float2* __restrict__ tau, int emit,
                               int range_begin, int PW, int32_t* __restrict__ out_idx, double* __restrict__ out_so,
                               double* __restrict__ out_R)
{
    const float never = __int_as_float(0x7f800000);
    cnt[slot] = 0;  // the next round starts from an empty list
    if (b.seq == ~0ull) {  // nothing finite at all: exact scan decides
        int at = atomicAdd(&ctrl[1], 1);
        fallback[at] = ridx;
        state[slot].flags = 1;
        tau[slot] = make_float2(never, never);
        return;
    }
    state[slot].R = b.R; state[slot].seq = b.seq; state[slot].s = b.s; state[slot].o = b.o;
    if (emit) {
        const size_t os_ = (size_t)(ridx - range_begin);
        const int t = (int)(b.seq & 7ull), p = (int)(b.seq >> 3);
        out_idx[4 * os_ + 0] = p / PW;
        out_idx[4 * os_ + 1] = p % PW;
        out_idx[4 * os_ + 2] = t >> 2;
        out_idx[4 * os_ + 3] = t & 3;
        out_so[2 * os_] = b.s;
        out_so[2 * os_ + 1] = b.o;
        out_R[os_] = b.R;
        return;
    }
    // Thresholds of the next round from the new winner.  A range whose bound is not clearly below its own variance
    // cannot be screened: it is handed to the exact scan and never passes anything.
    float2 t = make_float2(never, never);
    if (state[slot].flags == 0) {
        const double TA = state[slot].TA0 - 64.0 * b.R, TB = TA - 64.0 * GUARD_ILL;
        if (b.R == b.R && b.R < 1e300 && TA > 1e-3) {
            t.x = (float)sqrt(TA) * (1.0f - 1.0f / 1048576.0f);     // rounded down a little: never over-estimate
            t.y = TB > 0.0 ? (float)sqrt(TB) * (1.0f - 1.0f / 1048576.0f) : 0.f;
        } else {
            state[slot].flags = 1;
            fallback[atomicAdd(&ctrl[1], 1)] = ridx;
        }
    }
    tau[slot] = t;
}

__global__ void __launch_bounds__(RESCORE_WARPS * 32, RESCORE_MINB)   // 512 threads x 64 registers: a long list is walked 512 entries at a time
bf_rescore_kernel(const uint8_t* __restrict__ img, const __half* __restrict__ poolA, const uint32_t* __restrict__ sd4,
                  const float* __restrict__ sdd, int H, int W, int range_begin, RangeState* __restrict__ state,
                  int* __restrict__ ctrl, uint32_t* __restrict__ cnt, uint32_t* list /* written in the seed round */, uint32_t cap,
                  int32_t* __restrict__ fallback, float2* __restrict__ tau, int emit, int mode, int32_t* __restrict__ out_idx,
                  double* __restrict__ out_so, double* __restrict__ out_R, const unsigned long long* __restrict__ seed_best,
                  int seed_S, int seed_gx)
{
    extern __shared__ float rv_raw[];
    float (*rv)[8][65] = reinterpret_cast<float (*)[8][65]>(rv_raw);   // inverse-transformed range per isometry, padded against bank conflicts
    __shared__ double s_rs[RESCORE_WARPS], s_sq[RESCORE_WARPS];
    __shared__ double s_R[RESCORE_WARPS], s_s[RESCORE_WARPS], s_o[RESCORE_WARPS];   // running winner of each range
    __shared__ unsigned long long s_key[RESCORE_WARPS], s_prev[RESCORE_WARPS], s_seq[RESCORE_WARPS];
    __shared__ uint32_t s_n[RESCORE_WARPS + 1];   // list lengths, then their exclusive prefix
    __shared__ int s_ridx[RESCORE_WARPS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, tid = threadIdx.x;
    const int slot = blockIdx.x * RESCORE_WARPS + warp;
    const int n_elig = ctrl[0];
    const int rw = W / 8, PW = W - 16;
    const int NP = PW * (H - 16);
    // ---- phase A: warp w prepares range w of the CTA (pixels, sums, the eight oriented copies, its running winner) ----
    if (lane == 0) { s_n[warp] = 0; s_ridx[warp] = -1; }
    if (slot < n_elig) {
        const RangeState st = state[slot];
        uint32_t n;
        if (seed_best) {
            // seed round: the list is written here -- position 0 / isometry 0 (decides constant ranges, and the exact scan
            // visits it first anyway) plus the eight isometries of the best-ranked seed candidate (lattice node k / gx,
            // k % gx); same-CTA readers see it after the barrier below
            const unsigned long long key = seed_best[slot];
            n = key != 0ull ? 9u : 1u;
            if (lane < n) {
                uint32_t ent = 0u;
                if (lane > 0) {
                    const uint32_t k = 0xffffffffu - (uint32_t)(key & 0xffffffffull);
                    const uint32_t pos = (k / (uint32_t)seed_gx) * (uint32_t)seed_S * (uint32_t)PW + (k % (uint32_t)seed_gx) * (uint32_t)seed_S;
                    ent = pos * 8u + (uint32_t)(lane - 1);
                }
                list[(size_t)slot * cap + lane] = ent;
            }
        } else {
            n = cnt[slot];
        }
        const int ridx = st.ridx;
        if (st.flags & 1) {
            // already handed to the exact scan
        } else if (n > cap) {  // survivor list overflowed: let the exact scan decide this range
            if (lane == 0) {
                int at = atomicAdd(&ctrl[1], 1);
                fallback[at] = ridx;
                state[slot].flags = 1;
                cnt[slot] = 0;
                tau[slot] = make_float2(__int_as_float(0x7f800000), __int_as_float(0x7f800000));
            }
        } else {
            const int ri = (ridx / rw) * 8, rj = (ridx % rw) * 8;
            // lane holds range pixels k = lane and k = lane + 32
            const unsigned r0 = img[(size_t)(ri + (lane >> 3)) * W + rj + (lane & 7)];
            const unsigned r1 = img[(size_t)(ri + 4 + (lane >> 3)) * W + rj + (lane & 7)];
            unsigned s = r0 + r1;
            unsigned q = mode == FIC_BF_MODE_JIT ? r0 * r0 + r1 * r1 : ((r0 * r0) & 255u) + ((r1 * r1) & 255u);  // bf_encoder.py:33
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, off); q += __shfl_xor_sync(0xffffffffu, q, off); }
#pragma unroll
            for (int t = 0; t < 8; ++t) {
                int sa, sb;
                fic_iso_bf(t >> 2, t & 3, 8, lane >> 3, lane & 7, sa, sb);
                rv[warp][t][sa * 8 + sb] = (float)r0;
                fic_iso_bf(t >> 2, t & 3, 8, 4 + (lane >> 3), lane & 7, sa, sb);
                rv[warp][t][sa * 8 + sb] = (float)r1;
            }
            if (lane == 0) {
                s_rs[warp] = (double)s; s_sq[warp] = (double)q;
                s_n[warp] = n; s_ridx[warp] = ridx;
                s_R[warp] = st.R; s_s[warp] = st.s; s_o[warp] = st.o; s_seq[warp] = st.seq;   // the running winner competes too
                s_key[warp] = s_prev[warp] = rescore_key(st.R);
            }
        }
    }
    __syncthreads();
    // ---- phase B: the CTA's lists laid end to end, one entry per thread and pass --------------------------------------
    // (a warp per range would leave most lanes idle on the short lists and make one long list a serial tail.)
    // Winner of a range = first strict minimum in scan order: minimum of the order-preserving key of R, then, among
    // the entries that reach it, minimum of the scan index.  Non-finite R never takes part (fic_better).
    if (tid == 0) {
        uint32_t acc = 0;
        for (int r = 0; r < RESCORE_WARPS; ++r) { const uint32_t n = s_n[r]; s_n[r] = acc; acc += n; }
        s_n[RESCORE_WARPS] = acc;
        if (acc) atomicAdd(reinterpret_cast<unsigned long long*>(ctrl + 8), (unsigned long long)acc);
    }
    __syncthreads();
    const uint32_t total = s_n[RESCORE_WARPS];
    for (uint32_t base = 0; base < total; base += RESCORE_WARPS * 32) {
        const uint32_t e = base + tid;
        int r = -1;
        RescoreBest b;
        b.R = __longlong_as_double(0x7ff0000000000000ll); b.s = 0.0; b.o = 0.0; b.seq = ~0ull;
        unsigned long long key = ~0ull;
        if (e < total) {
            r = 0;   // last range whose offset is <= e (binary search over the 32 offsets; empty ranges share an offset)
#pragma unroll
            for (int step = RESCORE_WARPS / 2; step > 0; step >>= 1)
                if (e >= s_n[r + step]) r += step;
            const int rslot = blockIdx.x * RESCORE_WARPS + r;
            rescore_entry(list[(size_t)rslot * cap + (e - s_n[r])], rv[r], s_rs[r], s_sq[r], poolA, sd4, sdd, NP, mode, b);
            if (b.seq != ~0ull) {   // finite R
                key = rescore_key(b.R);
                atomicMin(&s_key[r], key);
            }
        }
        __syncthreads();
        if (tid < RESCORE_WARPS && s_key[tid] != s_prev[tid]) { s_prev[tid] = s_key[tid]; s_seq[tid] = ~0ull; }  // a better R: forget the old index
        __syncthreads();
        const bool contender = r >= 0 && b.seq != ~0ull && key == s_key[r];
        if (contender) atomicMin(&s_seq[r], b.seq);
        __syncthreads();
        if (contender && s_seq[r] == b.seq) { s_R[r] = b.R; s_s[r] = b.s; s_o[r] = b.o; }   // duplicates write the same values
        __syncthreads();
    }
    if (lane == 0 && s_ridx[warp] >= 0) {
        RescoreBest w;
        w.R = s_R[warp]; w.s = s_s[warp]; w.o = s_o[warp]; w.seq = s_seq[warp];
        rescore_commit(slot, s_ridx[warp], w, state, ctrl, cnt, fallback, tau, emit, range_begin, PW, out_idx, out_so, out_R);
    }
}

// ---- host ---------------------------------------------------------------------------------------------------
typedef CUresult (*PFN_tmapEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                        const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                        CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

PFN_tmapEncodeTiled get_encode()
{
    static PFN_tmapEncodeTiled fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = (PFN_tmapEncodeTiled)p;
    }
    return fn;
}

// fp16 matrix [rows][cols] row-major, box = cols x box_rows
int make_map(CUtensorMap* m, void* base, uint64_t rows, uint32_t cols, uint32_t box_rows, CUtensorMapSwizzle sw)
{
    PFN_tmapEncodeTiled enc = get_encode();
    if (!enc) { fic_set_error("cuTensorMapEncodeTiled is not available from this driver"); return FIC_ECUDA; }
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
    cuuint32_t box[2] = {cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { fic_set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows %llu cols %u)", (int)r, (unsigned long long)rows, cols); return FIC_ECUDA; }
    return FIC_OK;
}

inline int cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }


// =================================================================================================================
// Quadtree: the all-classes search on the tensor cores (quadtree/encoder.py:148-193, 285-384 with full_class_search; any
// pool lattice, dense or not).  Every range of a level is matched against every domain of the pool, which is the
// brute-force contraction with two differences:
//   * one isometry per (range, domain) pair and scale sign, derived from the two classifications -- it factors into a
//     permutation of the domain block and one of the range block (QT_ALPHA / QT_BETA), so the range operand has four
//     columns (scale sign x orientation of the domain's class) and the epilogue picks the pair's own with one select;
//   * the score is the QUANTISED one (fic_q_score), the winner the first minimum in the reference's scan order
//     (scale sign, class, position).
// The GEMM's accumulators are exact integers (centred pixel sums in fp16, fp32 accumulation well below 2^24), so the
// picked column gives the pair's exact dot product and with it all five inputs of the scorer bit for bit -- nothing here
// is a screen that could be wrong.  What the epilogue tests is the scorer's own lower bound
//     n E >= A_r - num^2 / (n den) ,   num = n sum(r d) - sum r sum d ,   den = n sum(d^2) - (sum d)^2 ,
// against the best score found so far for the range: a pair whose bound is above it cannot win and is dropped, everything
// else is scored exactly and, when at or below the best score at that moment, listed.  qt_tcpick_kernel then takes the
// first minimum in scan order among the listed.  Constant ranges (every candidate ties, the scan order alone decides) and
// ranges whose list overflowed are flagged and searched by qt_search_kernel.
struct QtcState {
    int32_t flags;                // 1 = handed to the exact search
    int32_t si, sj;
    int32_t rot[2], orient[2];    // the range's classification for positive / negative scales
    float rs, sqr;                // sum r, sum r^2 (integers below 2^24)
};

template <int S>
__global__ void qt_tcpool_kernel(const uint16_t* __restrict__ P, int H, int W, const int32_t* __restrict__ dcls,
                                 __half* __restrict__ poolA, uint32_t* __restrict__ sd4, uint32_t* __restrict__ sq4,
                                 float* __restrict__ sds8, uint8_t* __restrict__ dtyp, double* __restrict__ den, int hs, int vs)
{
    constexpr int N = S * S;
    const int PW = (W - 2 * S) / hs + 1, PH = (H - 2 * S) / vs + 1, BW = W - 1;   // the pool's lattice (quadtree/encoder.py:262-283)
    const int NP = PW * PH;
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= NP) return;
    const int di = (p / PW) * vs, dj = (p % PW) * hs;
    int v[N];
    int s = 0;
    long long s2 = 0;
#pragma unroll
    for (int a = 0; a < S; ++a)
#pragma unroll
        for (int b = 0; b < S; ++b) {
            const int x = P[(size_t)(di + 2 * a) * BW + dj + 2 * b];
            v[a * S + b] = x;
            s += x;
            s2 += x * x;
        }
    const int m = (s + N / 2) / N;      // round(SD4 / n); eps = s - n m in [-n/2, n/2)
    const unsigned dty = (unsigned)((dcls[3 * (size_t)p + 1] & 3) | ((dcls[3 * (size_t)p + 2] & 1) << 2));
    // the pool row is the block permuted by the alpha of its own type (QT_ALPHA): element k = block[alpha(k)], read again
    // from the box sums (L1) rather than through a runtime index into v[]
    const int al = QT_ALPHA[dty];
    uint4* dst = reinterpret_cast<uint4*>(poolA + (size_t)p * N);
#pragma unroll
    for (int q = 0; q < N / 8; ++q) {
        int w[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            int a, b;
            fic_iso_qt_enc(al >> 2, al & 3, S, (8 * q + i) / S, (8 * q + i) % S, a, b);
            w[i] = (int)P[(size_t)(di + 2 * a) * BW + dj + 2 * b] - m;
        }
        __half2 h0 = __halves2half2(__int2half_rn(w[0]), __int2half_rn(w[1]));
        __half2 h1 = __halves2half2(__int2half_rn(w[2]), __int2half_rn(w[3]));
        __half2 h2 = __halves2half2(__int2half_rn(w[4]), __int2half_rn(w[5]));
        __half2 h3 = __halves2half2(__int2half_rn(w[6]), __int2half_rn(w[7]));
        uint4 o;
        o.x = *reinterpret_cast<uint32_t*>(&h0);
        o.y = *reinterpret_cast<uint32_t*>(&h1);
        o.z = *reinterpret_cast<uint32_t*>(&h2);
        o.w = *reinterpret_cast<uint32_t*>(&h3);
        dst[q] = o;
    }
    sd4[p] = (uint32_t)s;
    sq4[p] = (uint32_t)s2;
    dtyp[p] = (uint8_t)dty;
    if (N == 64) {
        // the one order-dependent input of the scorer for 8x8 blocks: the float32 sequential sum of d^2 in the order the
        // reference walks the oriented block (row-major over the range pixels, quadtree/encoder.py:302-322), per isometry
        float sq[N];
#pragma unroll
        for (int k = 0; k < N; ++k) {
            const float d = __fmul_rn((float)v[k], 0.25f);
            sq[k] = __fmul_rn(d, d);
        }
        float acc[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) acc[t] = 0.f;
#pragma unroll
        for (int k = 0; k < N; ++k)
#pragma unroll
            for (int t = 0; t < 8; ++t) {
                int a, b;
                fic_iso_qt_enc(t >> 2, t & 3, S, k / S, k % S, a, b);
                acc[t] = __fadd_rn(acc[t], sq[a * S + b]);
            }
#pragma unroll
        for (int t = 0; t < 8; ++t) sds8[(size_t)t * NP + p] = acc[t];
        // the eight sums differ by rounding only; the bound test of the fused pass uses the smallest denominator (drops less)
        float lo = acc[0];
#pragma unroll
        for (int t = 1; t < 8; ++t) lo = fminf(lo, acc[t]);
        const double dsd = (double)__fmul_rn((float)s, 0.25f);
        den[p] = DS(DM((double)N, (double)lo), DM(dsd, dsd));
    } else {
        const double dsd = (double)__fmul_rn((float)s, 0.25f);
        den[p] = DS(DM((double)N, (double)__fmul_rn((float)(uint32_t)s2, 0.0625f)), DM(dsd, dsd));
    }
}

// one warp per slot: range sums, state, the record of the bound test, the range operand (eight isometric copies, centred)
template <int S>
__global__ void qt_tcinit_kernel(const uint8_t* __restrict__ img, int W, const int32_t* __restrict__ queue /*[n][4]*/,
                                 const int32_t* __restrict__ rcls /*[n][2][3] = class, rot, orient per scale sign*/, int n_ranges,
                                 QtcState* __restrict__ state, uint32_t* __restrict__ cnt, unsigned long long* __restrict__ bound,
                                 int* __restrict__ ctrl, int slots_pad, __half* __restrict__ rngB, int32_t* __restrict__ need_exact,
                                 QtcRange* __restrict__ rngc)
{
    constexpr int N = S * S;
    const int slot = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (slot >= slots_pad) return;
    if (slot == 0 && lane < 16) ctrl[lane] = lane == 0 ? n_ranges : 0;
    if (lane == 0) bound[slot] = 0x7ff0000000000000ull;      // best score found so far: +inf
    if (slot >= n_ranges) {   // padding columns of the last group: zero operand, record that says "skip"
        for (int e = lane; e < QT_COLS * N; e += 32) rngB[(size_t)slot * QT_COLS * N + e] = __float2half_rn(0.f);
        if (lane == 0) {
            QtcRange rc;
            rc.rho = 0; rc.ok = 0; rc.tw0 = rc.tw1 = 0u; rc.slack = 0.0; rc.rs = rc.sqr = 0.f;
            rngc[slot] = rc;
        }
        return;
    }
    const int si = queue[4 * (size_t)slot], sj = queue[4 * (size_t)slot + 1];
    int px[2] = {0, 0};
    unsigned sm = 0, q32 = 0;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        const int k = lane + 32 * half;
        if (k < N) {
            px[half] = img[(size_t)(si + k / S) * W + sj + k % S];
            sm += px[half];
            q32 += px[half] * px[half];
        }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        sm += __shfl_xor_sync(0xffffffffu, sm, off);
        q32 += __shfl_xor_sync(0xffffffffu, q32, off);
    }
    const int m = ((int)sm + N / 2) / N;
    // column c = scale sign * 2 + orientation of the domain's class: element k = range[beta(k)] (QT_BETA), centred
#pragma unroll
    for (int c = 0; c < QT_COLS; ++c) {
        const int md = c >> 1;
        const int rty = (rcls[6 * (size_t)slot + 3 * md + 1] & 3) | ((rcls[6 * (size_t)slot + 3 * md + 2] & 1) << 2);
        const int be = QT_BETA[c & 1][rty];
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const int k = lane + 32 * half;
            if (k < N) {
                int a, b;
                fic_iso_qt_enc(be >> 2, be & 3, S, k / S, k % S, a, b);
                rngB[((size_t)slot * QT_COLS + c) * N + k] = __int2half_rn((int)img[(size_t)(si + a) * W + sj + b] - m);
            }
        }
    }
    if (lane == 0) {
        const long long sl = sm, ql = q32;
        QtcState st;
        st.flags = (long long)N * ql - sl * sl == 0 ? 1 : 0;      // constant block: every candidate ties, the scan order decides -> exact search
        st.si = si; st.sj = sj;
        for (int md = 0; md < 2; ++md) { st.rot[md] = rcls[6 * (size_t)slot + 3 * md + 1]; st.orient[md] = rcls[6 * (size_t)slot + 3 * md + 2]; }
        st.rs = (float)sl; st.sqr = (float)ql;     // exact: <= 64 * 255^2 < 2^24
        state[slot] = st;
        cnt[slot] = 0;
        need_exact[slot] = st.flags;
        QtcRange rc;
        rc.rho = (int)sm - N * m;
        rc.ok = !st.flags;
        rc.rs = st.rs; rc.sqr = st.sqr;
        rc.slack = DS((double)st.sqr, DD(DM((double)st.rs, (double)st.rs), (double)N));
        for (int md = 0; md < 2; ++md) {
            const int rrot = st.rot[md] & 3, rori = st.orient[md] & 1;
            uint32_t w = 0;
            for (int dty = 0; dty < 8; ++dty) {   // the pair's isometry (quadtree/encoder.py:168-176) for each of the eight domain types
                const int drot = dty & 3, dori = dty >> 2;
                const bool same = dori == rori;
                w |= (uint32_t)(((drot - rrot + (same ? 0 : 2)) & 3) | (same ? 0 : 4)) << (4 * dty);
            }
            if (md) rc.tw1 = w; else rc.tw0 = w;
        }
        rngc[slot] = rc;
    }
}

// ---- the winner of every range: first minimum in scan order among the listed candidates -------------------------------
// QTC_WARPS ranges per CTA; their lists are laid end to end and every thread rescores one entry per pass from the pixels
// (the reference's own float32 sums, row-major over the range block), so a long list is shared by the whole CTA.
#define QTC_WARPS 32
template <int S>
__global__ void __launch_bounds__(QTC_WARPS * 32, 1)
qt_tcpick_kernel(const uint8_t* __restrict__ img, const uint16_t* __restrict__ P /*2x2 box sums, (H-1) x (W-1)*/,
                 const int32_t* __restrict__ dcls /*[NP][3] = class, rot, orient*/, int H, int W, QtcState* __restrict__ state,
                 const int* __restrict__ ctrl, const uint32_t* __restrict__ cnt, const uint32_t* __restrict__ list, uint32_t cap,
                 int nmodes, double max_scale, int scale_bits, int offset_bits, int32_t* __restrict__ out_rec,
                 double* __restrict__ out_vals, int32_t* __restrict__ need_exact, int hs, int vs,
                 const int32_t* __restrict__ pool_offsets /*[73]*/, const int32_t* __restrict__ pool_rows /*[..][4]*/)
{
    constexpr int N = S * S;
    __shared__ float rv[QTC_WARPS][N];            // the range blocks as floats
    __shared__ float s_rs[QTC_WARPS], s_sqr[QTC_WARPS];
    __shared__ int s_rot[QTC_WARPS][2], s_orient[QTC_WARPS][2];
    __shared__ unsigned long long s_key[QTC_WARPS], s_prev[QTC_WARPS], s_tie[QTC_WARPS];
    __shared__ double s_E[QTC_WARPS], s_s[QTC_WARPS], s_o[QTC_WARPS];
    __shared__ int s_rec[QTC_WARPS][6];
    __shared__ uint32_t s_n[QTC_WARPS + 1];
    __shared__ int s_live[QTC_WARPS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, tid = threadIdx.x;
    const int slot = blockIdx.x * QTC_WARPS + warp;
    const int n_elig = ctrl[0];
    const int PW = (W - 2 * S) / hs + 1;
    // ---- phase A: warp w prepares range w of the CTA ----
    if (lane == 0) { s_n[warp] = 0; s_live[warp] = 0; }
    if (slot < n_elig) {
        const QtcState st = state[slot];
        const uint32_t n = cnt[slot];
        if (st.flags & 1) {
            // already handed to the exact search
        } else if (n > cap) {
            if (lane == 0) {
                state[slot].flags = 1;
                need_exact[slot] = 1;
            }
        } else {
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                const int k = lane + 32 * half;
                if (k < N) rv[warp][k] = (float)img[(size_t)(st.si + k / S) * W + st.sj + k % S];
            }
            if (lane == 0) {
                s_rs[warp] = st.rs; s_sqr[warp] = st.sqr;
                for (int md = 0; md < 2; ++md) { s_rot[warp][md] = st.rot[md]; s_orient[warp][md] = st.orient[md]; }
                s_n[warp] = n; s_live[warp] = 1;
                s_E[warp] = __longlong_as_double(0x7ff0000000000000ll); s_s[warp] = 0.0; s_o[warp] = 0.0; s_tie[warp] = ~0ull;
                for (int k = 0; k < 6; ++k) s_rec[warp][k] = 0;
                s_key[warp] = s_prev[warp] = rescore_key(s_E[warp]);
            }
        }
    }
    __syncthreads();
    if (tid == 0) {
        uint32_t acc = 0;
        for (int r = 0; r < QTC_WARPS; ++r) { const uint32_t n = s_n[r]; s_n[r] = acc; acc += n; }
        s_n[QTC_WARPS] = acc;
    }
    __syncthreads();
    // ---- phase B: the CTA's lists laid end to end, one entry per thread and pass ----
    const uint32_t total = s_n[QTC_WARPS];
    for (uint32_t base = 0; base < total; base += QTC_WARPS * 32) {
        const uint32_t e = base + tid;
        int r = -1;
        double bE = __longlong_as_double(0x7ff0000000000000ll), bs = 0.0, bo = 0.0;
        unsigned long long btie = ~0ull, key = ~0ull;
        int brec[6] = {0, 0, 0, 0, 0, 0};
        if (e < total) {
            r = 0;
#pragma unroll
            for (int step = QTC_WARPS / 2; step > 0; step >>= 1)
                if (e >= s_n[r + step]) r += step;
            const int rslot = blockIdx.x * QTC_WARPS + r;
            const uint32_t ent = list[(size_t)rslot * cap + (e - s_n[r])];
            const int t = ent & 7, p = ent >> 3;
            const int dclass = dcls[3 * (size_t)p], drot = dcls[3 * (size_t)p + 1], dorient = dcls[3 * (size_t)p + 2];
            const int di = (p / PW) * vs, dj = (p % PW) * hs;
            for (int md = 0; md < nmodes; ++md) {
                const int rrot = s_rot[r][md], rorient = s_orient[r][md];
                int d_rot, d_orient;
                if (dorient == rorient) { d_rot = (drot - rrot) & 3; d_orient = 0; }
                else { d_rot = (drot - rrot + 2) & 3; d_orient = 1; }
                if (d_orient * 4 + d_rot != t) continue;      // the entry belongs to the other scale sign
                // the three float32 sums in the reference's order: row-major over the range block (encoder.py:302-322)
                const uint16_t* blk = P + (size_t)di * (W - 1) + dj;
                float ds = 0.f, sds = 0.f, ab = 0.f;
#pragma unroll 4
                for (int k = 0; k < N; ++k) {
                    int a, b;
                    fic_iso_qt_enc(d_orient, d_rot, S, k / S, k % S, a, b);
                    const float d = __fmul_rn((float)blk[(size_t)(2 * a) * (W - 1) + 2 * b], 0.25f);
                    ds = __fadd_rn(ds, d);
                    sds = __fadd_rn(sds, __fmul_rn(d, d));
                    ab = __fadd_rn(ab, __fmul_rn(d, rv[r][k]));
                }
                const QScore sc = fic_q_score(N, s_rs[r], s_sqr[r], ds, sds, ab, max_scale, scale_bits, offset_bits);
                const unsigned long long tie = ((unsigned long long)(md * 72 + dclass) << 32) | (unsigned)p;
                if (fic_better(sc.R, tie, bE, btie)) {
                    bE = sc.R; btie = tie; bs = sc.s; bo = sc.o;
                    brec[0] = di; brec[1] = dj; brec[2] = d_orient; brec[3] = d_rot; brec[4] = sc.qs; brec[5] = sc.qo;
                }
            }
            if (btie != ~0ull) {
                key = rescore_key(bE);
                atomicMin(&s_key[r], key);
            }
        }
        __syncthreads();
        if (tid < QTC_WARPS && s_key[tid] != s_prev[tid]) { s_prev[tid] = s_key[tid]; s_tie[tid] = ~0ull; }   // a better score: forget the old tie key
        __syncthreads();
        const bool contender = r >= 0 && btie != ~0ull && key == s_key[r];
        if (contender) atomicMin(&s_tie[r], btie);
        __syncthreads();
        if (contender && s_tie[r] == btie) {
            s_E[r] = bE; s_s[r] = bs; s_o[r] = bo;
            for (int k = 0; k < 6; ++k) s_rec[r][k] = brec[k];
        }
        __syncthreads();
    }
    // ---- the placeholder rows: a class without a domain is represented by the row (0, 0, rotation 0, orientation 0)
    // (quadtree/encoder.py:279-280), which the reference's loop over all classes scores like any other domain -- the block at
    // the origin, with this class's place in the scan order.  (A class whose only domain IS that row gives the same candidate.)
    if (s_live[warp]) {
        double bE = __longlong_as_double(0x7ff0000000000000ll), bs = 0.0, bo = 0.0;
        unsigned long long btie = ~0ull;
        int brec[6] = {0, 0, 0, 0, 0, 0};
        for (int c = lane; c < 72; c += 32) {
            const int b0 = pool_offsets[c];
            if (pool_offsets[c + 1] - b0 != 1) continue;
            const int32_t* row = pool_rows + 4 * (size_t)b0;
            if (row[0] | row[1] | row[2] | row[3]) continue;
            for (int md = 0; md < nmodes; ++md) {
                const int rrot = s_rot[warp][md], rorient = s_orient[warp][md];
                const int d_orient = rorient == 0 ? 0 : 1;
                const int d_rot = rorient == 0 ? ((0 - rrot) & 3) : ((0 - rrot + 2) & 3);
                float ds = 0.f, sds = 0.f, ab = 0.f;
                for (int k = 0; k < N; ++k) {
                    int a, b;
                    fic_iso_qt_enc(d_orient, d_rot, S, k / S, k % S, a, b);
                    const float d = __fmul_rn((float)P[(size_t)(2 * a) * (W - 1) + 2 * b], 0.25f);
                    ds = __fadd_rn(ds, d);
                    sds = __fadd_rn(sds, __fmul_rn(d, d));
                    ab = __fadd_rn(ab, __fmul_rn(d, rv[warp][k]));
                }
                const QScore sc = fic_q_score(N, s_rs[warp], s_sqr[warp], ds, sds, ab, max_scale, scale_bits, offset_bits);
                const unsigned long long tie = (unsigned long long)(md * 72 + c) << 32;
                if (fic_better(sc.R, tie, bE, btie)) {
                    bE = sc.R; btie = tie; bs = sc.s; bo = sc.o;
                    brec[0] = 0; brec[1] = 0; brec[2] = d_orient; brec[3] = d_rot; brec[4] = sc.qs; brec[5] = sc.qo;
                }
            }
        }
        // the warp's best placeholder candidate against the winner among the listed
        double wE = bE;
        unsigned long long wt = btie;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            const double oE = __shfl_xor_sync(0xffffffffu, wE, off);
            const unsigned long long ot = __shfl_xor_sync(0xffffffffu, wt, off);
            if (fic_better(oE, ot, wE, wt)) { wE = oE; wt = ot; }
        }
        if (wt != ~0ull && btie == wt && fic_better(wE, wt, s_E[warp], s_tie[warp])) {   // tie keys are unique: one lane
            s_E[warp] = bE; s_s[warp] = bs; s_o[warp] = bo; s_tie[warp] = btie;
            for (int k = 0; k < 6; ++k) s_rec[warp][k] = brec[k];
        }
        __syncwarp();
    }
    // ---- commit ----
    if (lane == 0 && s_live[warp]) {
        if (s_tie[warp] == ~0ull) {      // nothing finite at all
            state[slot].flags = 1;
            need_exact[slot] = 1;
        } else {
            for (int k = 0; k < 6; ++k) out_rec[6 * (size_t)slot + k] = s_rec[warp][k];
            out_vals[3 * (size_t)slot] = s_s[warp]; out_vals[3 * (size_t)slot + 1] = s_o[warp]; out_vals[3 * (size_t)slot + 2] = s_E[warp];
        }
    }
}

// ---- brute force, 4x4 ranges: pool, range operand, winners ------------------------------------------------------------
// pool row = the pooled 4x4 block (cell sums, centred by their rounded mean) in plain row-major order; the isometries sit on
// the range side.  Lattice position q = (q / npj * si, q % npj * sj), jit_bfencoder.py:27-30.
__global__ void bf4_pool_kernel(const uint16_t* __restrict__ P, int W, int si, int sj, int npj, int NP, __half* __restrict__ poolA,
                                uint32_t* __restrict__ sd4, uint32_t* __restrict__ sq4, double* __restrict__ den)
{
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= NP) return;
    const int BW = W - 1;
    const uint16_t* dp = P + (size_t)(q / npj) * si * BW + (q % npj) * sj;
    int v[16], s = 0, s2 = 0;
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int x = dp[(size_t)(2 * a) * BW + 2 * b];
            v[a * 4 + b] = x;
            s += x;
            s2 += x * x;
        }
    const int m = (s + 8) / 16;
    uint4* dst = reinterpret_cast<uint4*>(poolA + (size_t)q * 16);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        __half2 h0 = __halves2half2(__int2half_rn(v[8 * h + 0] - m), __int2half_rn(v[8 * h + 1] - m));
        __half2 h1 = __halves2half2(__int2half_rn(v[8 * h + 2] - m), __int2half_rn(v[8 * h + 3] - m));
        __half2 h2 = __halves2half2(__int2half_rn(v[8 * h + 4] - m), __int2half_rn(v[8 * h + 5] - m));
        __half2 h3 = __halves2half2(__int2half_rn(v[8 * h + 6] - m), __int2half_rn(v[8 * h + 7] - m));
        uint4 o;
        o.x = *reinterpret_cast<uint32_t*>(&h0);
        o.y = *reinterpret_cast<uint32_t*>(&h1);
        o.z = *reinterpret_cast<uint32_t*>(&h2);
        o.w = *reinterpret_cast<uint32_t*>(&h3);
        dst[h] = o;
    }
    sd4[q] = (uint32_t)s;
    sq4[q] = (uint32_t)s2;
    // n sum(d^2) - (sum d)^2 with d = cell / 4: exact in float64
    den[q] = DS((double)s2, DM(DM((double)s, (double)s), 0.0625));
}

// one warp per range slot: sums, the record of the bound test (range origin in tw0 / tw1), eight isometric copies of the
// centred block as operand columns -- column t holds range pixel k at the place isometry t reads for it (fic_iso_bf)
__global__ void bf4_init_kernel(const uint8_t* __restrict__ img, int W, int mode, int range_begin, int n_ranges, int slots_pad,
                                __half* __restrict__ rngB, QtcRange* __restrict__ rngc, unsigned long long* __restrict__ bound,
                                uint32_t* __restrict__ cnt, int* __restrict__ ctrl, int32_t* __restrict__ need_exact)
{
    const int slot = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (slot >= slots_pad) return;
    if (slot == 0 && lane < 16) ctrl[lane] = lane == 0 ? n_ranges : 0;
    if (lane == 0) bound[slot] = 0x7ff0000000000000ull;
    if (slot >= n_ranges) {
        for (int e = lane; e < 8 * 16; e += 32) rngB[(size_t)slot * 128 + e] = __float2half_rn(0.f);
        if (lane == 0) {
            QtcRange rc;
            rc.rho = 0; rc.ok = 0; rc.tw0 = rc.tw1 = 0u; rc.slack = 0.0; rc.rs = rc.sqr = 0.f;
            rngc[slot] = rc;
        }
        return;
    }
    const int ridx = range_begin + slot, rw = W / 4;
    const int si = (ridx / rw) * 4, sj = (ridx % rw) * 4;
    const int k = lane & 15;
    const int px = img[(size_t)(si + k / 4) * W + sj + k % 4];
    int sm = lane < 16 ? px : 0, q2 = lane < 16 ? px * px : 0, qw = lane < 16 ? (px * px) & 255 : 0;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        sm += __shfl_xor_sync(0xffffffffu, sm, off);
        q2 += __shfl_xor_sync(0xffffffffu, q2, off);
        qw += __shfl_xor_sync(0xffffffffu, qw, off);
    }
    const int m = (sm + 8) / 16;
#pragma unroll
    for (int tt = 0; tt < 4; ++tt) {        // lanes 0-15 write isometries 0..3, lanes 16-31 isometries 4..7
        const int t = tt + (lane >> 4) * 4;
        int a, b;
        fic_iso_bf(t >> 2, t & 3, 4, k / 4, k % 4, a, b);
        rngB[((size_t)slot * 8 + t) * 16 + a * 4 + b] = __int2half_rn(px - m);
    }
    if (lane == 0) {
        QtcRange rc;
        const bool flat = 16 * q2 - sm * sm == 0;       // constant block: every candidate scores exactly 0 (bf4_pick_kernel)
        rc.rho = sm - 16 * m;
        rc.ok = flat ? 0 : 1;
        rc.tw0 = (uint32_t)si; rc.tw1 = (uint32_t)sj;
        const int sqr = mode == FIC_BF_MODE_NUMPY ? qw : q2;      // bf_encoder.py:33 squares the uint8 pixels in uint8
        rc.rs = (float)sm; rc.sqr = (float)sqr;
        rc.slack = DS((double)sqr, DM(DM((double)sm, (double)sm), 0.0625));
        rngc[slot] = rc;
        cnt[slot] = 0;
        need_exact[slot] = 0;
    }
}

// A first bound for every range before any tile is looked at: the score of one candidate (a position in the middle of the
// lattice, identity).  With the bounds at +inf the first tile of the seed pass would queue -- and score -- all of its 1024
// candidates per range.
__global__ void bf4_first_bound_kernel(const ScreenParams p, int n_ranges)
{
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= n_ranges) return;
    const QtcRange rc = p.rngc[slot];
    if (!(rc.ok & 1)) return;
    const BfScore sc = bf4_score_pair(p, rc, p.NP / 2, 0);
    if (sc.R < __longlong_as_double(0x7ff0000000000000ll))
        p.bound[slot] = (unsigned long long)__double_as_longlong(sc.R > 0.0 ? sc.R : 0.0);
}

// one warp per range: the first minimum in scan order (position * 8 + isometry) among the listed candidates
__global__ void bf4_pick_kernel(const ScreenParams p, int n_ranges, int32_t* __restrict__ out_idx, double* __restrict__ out_so,
                                double* __restrict__ out_R, int32_t* __restrict__ need_exact)
{
    const int slot = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (slot >= n_ranges) return;
    const QtcRange rc = p.rngc[slot];
    if (!(rc.ok & 1)) {
        // A constant range: num = n * ab - rs * ds is exactly zero for every candidate (ab = c * ds without rounding), so the
        // contrast is zero, the offset the gray level and the score exactly +0 whatever the domain -- the first candidate
        // of the scan wins (jit_bfencoder.py keeps the first strict minimum), and its record is the scorer's.
        if (lane == 0) {
            const BfScore sc = bf4_score_pair(p, rc, 0, 0);
            for (int k = 0; k < 4; ++k) out_idx[4 * (size_t)slot + k] = 0;
            out_so[2 * (size_t)slot] = sc.s;
            out_so[2 * (size_t)slot + 1] = sc.o;
            if (out_R) out_R[slot] = sc.R;
        }
        return;
    }
    const uint32_t n = p.cnt[slot];
    if (n > p.cap || n == 0) {
        if (lane == 0) need_exact[slot] = 1;
        return;
    }
    double bR = __longlong_as_double(0x7ff0000000000000ll), bs = 0.0, bo = 0.0;
    unsigned long long bseq = ~0ull;
    for (uint32_t e = lane; e < n; e += 32) {
        const uint32_t ent = p.list[(size_t)slot * p.cap + e];
        const BfScore sc = bf4_score_pair(p, rc, (int)(ent >> 3), (int)(ent & 7u));
        if (fic_better(sc.R, (unsigned long long)ent, bR, bseq)) { bR = sc.R; bseq = ent; bs = sc.s; bo = sc.o; }
    }
    double wR = bR;
    unsigned long long ws = bseq;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        const double oR = __shfl_xor_sync(0xffffffffu, wR, off);
        const unsigned long long oS = __shfl_xor_sync(0xffffffffu, ws, off);
        if (fic_better(oR, oS, wR, ws)) { wR = oR; ws = oS; }
    }
    if (ws == ~0ull) {
        if (lane == 0) need_exact[slot] = 1;
    } else if (bseq == ws) {      // entries are unique: one lane
        const int q = (int)(ws >> 3), t = (int)(ws & 7ull);
        out_idx[4 * (size_t)slot + 0] = (q / p.lat_npj) * p.lat_si;
        out_idx[4 * (size_t)slot + 1] = (q % p.lat_npj) * p.lat_sj;
        out_idx[4 * (size_t)slot + 2] = t >> 2;
        out_idx[4 * (size_t)slot + 3] = t & 3;
        out_so[2 * (size_t)slot] = bs;
        out_so[2 * (size_t)slot + 1] = bo;
        if (out_R) out_R[slot] = bR;
    }
}

}  // namespace

bool bf_screen_applicable(int H, int W)
{
    // below ~32 tiles of candidates the fixed costs of the pipeline exceed the exact scan
    return H > 16 && W > 16 && (long long)(H - 16) * (W - 16) >= 4096;
}

// `probe` != nullptr: debugging/validation mode -- one column group is run over all tiles with infinite thresholds and
// the raw accumulators are written to probe[n_tiles*128][256].
int bf_screen_run(fic_handle* h, const uint8_t* d_img, int H, int W, int mode, int range_begin, int n_ranges, int32_t* d_idx,
                  double* d_so, double* d_R, unsigned flags, float* probe, cudaStream_t st)
{
    const int PW = W - 16, PH = H - 16;
    const int NP = PW * PH;
    const int n_tiles = cdiv(NP, BM);
    const int n_cols_pad = cdiv((long long)n_ranges * 8, BN) * BN;
    const int tau_pad = n_cols_pad / 8;
    const uint16_t* P = h->bf_boxsum.as<uint16_t>();
    uint32_t* sd4 = h->bf_stats.as<uint32_t>();
    float* sdd = reinterpret_cast<float*>(sd4 + 2 * (size_t)NP);

    // survivor capacity per range: bounded by a 1 GiB list
    uint32_t cap = 8192;
    while (cap > 256 && (size_t)n_ranges * cap * 4 > ((size_t)1 << 30)) cap >>= 1;

    FIC_TRY(h->bf_pool.reserve((size_t)NP * 64 * 2 + 256));
    FIC_TRY(h->bf_cand.reserve((size_t)NP * sizeof(float2) + 256));
    FIC_TRY(h->bf_A.reserve((size_t)n_cols_pad * 64 * 2 + 256));
    FIC_TRY(h->bf_thr.reserve((size_t)tau_pad * sizeof(float2) + (size_t)n_ranges * (2 * sizeof(int32_t) + sizeof(uint2) + sizeof(RangeState)) + 512));
    FIC_TRY(h->bf_surv.reserve((size_t)n_ranges * 4 + 64));
    FIC_TRY(h->bf_list.reserve((size_t)n_ranges * cap * 4 + 256));
    if (!h->bf_ctrl.p) {   // control block of the call + the totals, which start at zero and live as long as the handle
        FIC_TRY(h->bf_ctrl.reserve(256));
        FIC_CUDA(cudaMemsetAsync(h->bf_ctrl.p, 0, 256, st));
    }
    // seed lattice: about NP/50 positions (at least 256, at most 8192), a whole number of candidate tiles
    int seed_S, seed_gx, n_seed, n_seed_pad;
    {
        double want = (double)NP / 50.0;       // same-box sweeps on lena 512x512: 5 000 positions beat 2 500 and 7 500
        if (want < 256.0) want = 256.0;
        if (want > 8192.0) want = 8192.0;
        if (const char* e = tuning_env("FIC_SEED_POSITIONS")) want = atof(e) > 1.0 ? atof(e) : want;  // tuning knob
        seed_S = (int)floor(sqrt((double)NP / want) + 0.5);
        if (seed_S < 1) seed_S = 1;
        seed_gx = cdiv(PW, seed_S);
        n_seed = seed_gx * cdiv(PH, seed_S);
        n_seed_pad = cdiv(n_seed, BM) * BM;
    }
    const size_t off_seedhk = ((size_t)tau_pad * 8 + 255) & ~(size_t)255;
    const size_t off_seedA = (off_seedhk + (size_t)n_seed_pad * 8 + 255) & ~(size_t)255;
    FIC_TRY(h->bf_best.reserve(off_seedA + (size_t)n_seed_pad * 128 + 256));
    FIC_TRY(h->pin_a.reserve(256));

    __half* poolA = h->bf_pool.as<__half>();
    float2* hk = h->bf_cand.as<float2>();
    __half* rngB = h->bf_A.as<__half>();
    RangeState* state = h->bf_thr.as<RangeState>();                                   // 8-byte aligned members first
    float2* tau = reinterpret_cast<float2*>(state + n_ranges);
    int32_t* fallback = reinterpret_cast<int32_t*>(tau + tau_pad);
    uint32_t* cnt = h->bf_surv.as<uint32_t>();
    uint32_t* list = h->bf_list.as<uint32_t>();
    int* ctrl = h->bf_ctrl.as<int>();
    unsigned char* seed_raw = h->bf_best.as<unsigned char>();
    unsigned long long* seed_best = reinterpret_cast<unsigned long long*>(seed_raw);
    float2* seedhk = reinterpret_cast<float2*>(seed_raw + off_seedhk);
    __half* seedA = reinterpret_cast<__half*>(seed_raw + off_seedA);

    // 1. pool (independent of the ranges)
    bf_pool_kernel<<<dim3(cdiv(PW, POOL_TILE), PH), POOL_TILE, 0, st>>>(P, H, W, poolA, hk, sd4, sd4 + NP, sdd, seed_S, seed_gx, n_seed,
                                                                        n_seed_pad, probe ? nullptr : seedA, seedhk);
    FIC_CHECK_LAUNCH(h, "bf_pool_kernel");

    // 2. range states and the range operand: all eligible in natural order, nothing known yet (also clears the control block and the lists;
    //    the probe keeps these thresholds: nothing ever passes)
    bf_init_kernel<<<cdiv((long long)tau_pad * 32, 256), 256, 0, st>>>(d_img, W, mode, range_begin, n_ranges, state, tau, cnt, seed_best, ctrl, tau_pad, rngB);
    FIC_CHECK_LAUNCH(h, "bf_init_kernel");

    // 4. the screen, in two rounds
    CUtensorMap mapA, mapB;
    FIC_TRY(make_map(&mapA, poolA, (uint64_t)NP, 64, BM, CU_TENSOR_MAP_SWIZZLE_128B));
    FIC_TRY(make_map(&mapB, rngB, (uint64_t)n_cols_pad, 64, BN, CU_TENSOR_MAP_SWIZZLE_128B));
    if (!h->screen_attr_ready) {   // the opt-in is per device: kept in the handle, set with the handle's device current
        FIC_CUDA(cudaFuncSetAttribute(bf_screen_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SCREEN_SMEM));
        FIC_CUDA(cudaFuncSetAttribute(bf_screen_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SCREEN_SMEM));
#ifdef FIC_TUNING
        FIC_CUDA(cudaFuncSetAttribute(bf_screen_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SCREEN_SMEM));
        FIC_CUDA(cudaFuncSetAttribute(bf_screen_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SCREEN_SMEM));
        FIC_CUDA(cudaFuncSetAttribute(bf_screen_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SCREEN_SMEM));
#endif
        FIC_CUDA(cudaFuncSetAttribute(bf_screen_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SCREEN_SMEM));
        FIC_CUDA(cudaFuncSetAttribute(bf_rescore_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)RESCORE_SMEM));
        h->screen_attr_ready = true;
    }
    ScreenParams sp;
    sp.hk = hk; sp.tau = tau; sp.n_elig = ctrl; sp.cnt = cnt; sp.list = list; sp.cap = cap; sp.NP = NP; sp.n_tiles = n_tiles;
    sp.dump = probe;
    sp.seed_best = nullptr;
    sp.err_flag = ctrl + 2;
    {
        const char* dbg = tuning_env("FIC_SCREEN_DEBUG");  // kernel timing experiments only; records are wrong when set
        sp.debug_mode = dbg ? atoi(dbg) : 0;
    }
    // 4a. seed round: the GEMM ranks the lattice candidates per range (no thresholds exist yet), the best one and
    // position 0 are scored exactly and give every range its first upper bound
    if (!probe) {
        CUtensorMap mapS;
        FIC_TRY(make_map(&mapS, seedA, (uint64_t)n_seed_pad, 64, BM, CU_TENSOR_MAP_SWIZZLE_128B));
        ScreenParams ss = sp;
        ss.hk = seedhk; ss.NP = n_seed; ss.n_tiles = n_seed_pad / BM; ss.m_cur = 1; ss.r_prev = 0;
        ss.n_round_tiles = ss.n_tiles; ss.window_tiles = ss.n_tiles; ss.seed_best = seed_best;
        {
            ProfScope prof(h, st);
            bf_screen_kernel<5><<<h->num_sms, SCREEN_THREADS, SCREEN_SMEM, st>>>(mapS, mapB, ss);
        }
        FIC_CHECK_LAUNCH(h, "bf_screen_kernel(seed)");
        bf_rescore_kernel<<<cdiv(n_ranges, RESCORE_WARPS), RESCORE_WARPS * 32, RESCORE_SMEM, st>>>(d_img, poolA, sd4, sdd, H, W, range_begin, state, ctrl, cnt, list, cap, fallback, tau, 0, mode,
                                                    d_idx, d_so, d_R, seed_best, seed_S, seed_gx);
        FIC_CHECK_LAUNCH(h, "bf_rescore_kernel(seed)");
    }
    const bool prof_rescore = tuning_env("FIC_PROFILE_RESCORE") != nullptr;   // measurement knob: time the rescoring kernels instead
    // Round plan: nested strides m_1 > m_2 > ... > 1 over the candidate tiles.  A round's survivor count per range is
    // about (fraction screened now) / (fraction scored exactly before), so each round may look at ~128x more candidates
    // than everything before it; the first one is sized against the pre-scan sample.
    std::vector<int> strides;
    if (probe || n_tiles < 64) {
        strides.push_back(1);
    } else {
        double f_prev = (double)n_seed / (double)NP;
        int growth = 16;
        if (const char* e = tuning_env("FIC_ROUND_GROWTH")) growth = atoi(e) >= 2 ? atoi(e) : growth;  // tuning knob
        int m = 1;
        while (1.0 / m > (double)growth * f_prev && 2 * m * 8 <= n_tiles) m *= 2;
        for (;;) {
            strides.push_back(m);
            if (m == 1) break;
            m = m / growth > 1 ? m / growth : 1;
        }
    }
    sp.window_tiles = (64 << 20) / (BM * KMAIN * 2);   // 64 MiB of pool per window: stays in the 126 MB L2
    for (size_t r = 0; r < strides.size(); ++r) {
        sp.m_cur = strides[r];
        sp.r_prev = r == 0 ? 0 : strides[r - 1] / strides[r];
        sp.n_round_tiles = cdiv(n_tiles, strides[r]) - (r == 0 ? 0 : cdiv(n_tiles, strides[r - 1]));
        int grid = probe ? 1 : h->num_sms;
        {
            h->prof_suppress = prof_rescore;
            ProfScope prof(h, st);
            if (probe) bf_screen_kernel<1><<<grid, SCREEN_THREADS, SCREEN_SMEM, st>>>(mapA, mapB, sp);
#ifdef FIC_TUNING
            else if (sp.debug_mode == 1) bf_screen_kernel<2><<<grid, SCREEN_THREADS, SCREEN_SMEM, st>>>(mapA, mapB, sp);
            else if (sp.debug_mode == 2) bf_screen_kernel<3><<<grid, SCREEN_THREADS, SCREEN_SMEM, st>>>(mapA, mapB, sp);
            else if (sp.debug_mode == 3) bf_screen_kernel<4><<<grid, SCREEN_THREADS, SCREEN_SMEM, st>>>(mapA, mapB, sp);
#endif
            else bf_screen_kernel<0><<<grid, SCREEN_THREADS, SCREEN_SMEM, st>>>(mapA, mapB, sp);
        }
        FIC_CHECK_LAUNCH(h, "bf_screen_kernel");
        if (probe) return FIC_OK;
        // 5. exact rescoring of this round's survivors, merged into the range states
        const int emit = r + 1 == strides.size();
        if (tuning_env("FIC_DBG_LISTS")) {   // measurement knob: survivor list lengths before this rescoring
            std::vector<uint32_t> hc_(n_ranges);
            cudaStreamSynchronize(st);
            cudaMemcpy(hc_.data(), cnt, (size_t)n_ranges * 4, cudaMemcpyDeviceToHost);
            std::sort(hc_.begin(), hc_.end());
            unsigned long long sum = 0;
            for (uint32_t v : hc_) sum += v;
            fprintf(stderr, "round %zu lists: total %llu median %u p90 %u p99 %u max %u (cap %u)\n", r, sum, hc_[n_ranges / 2],
                    hc_[(size_t)n_ranges * 9 / 10], hc_[(size_t)n_ranges * 99 / 100], hc_.back(), cap);
        }
        h->prof_suppress = !prof_rescore;
        ProfScope prof_rs(h, st);
        bf_rescore_kernel<<<cdiv(n_ranges, RESCORE_WARPS), RESCORE_WARPS * 32, RESCORE_SMEM, st>>>(d_img, poolA, sd4, sdd, H, W, range_begin, state, ctrl, cnt, list, cap, fallback, tau, emit, mode,
                                                    d_idx, d_so, d_R, nullptr, 0, 0);
        FIC_CHECK_LAUNCH(h, "bf_rescore_kernel");
    }
    h->prof_suppress = false;

    // 6. whatever the screen could not decide goes through the exact scan.  The number of such ranges is known to the
    // device only (ctrl[1]); the scan is launched unconditionally and sizes itself (bf_scan8_dyn_kernel), so a call with
    // device pointers returns without waiting for the GPU and the next call's launches queue up behind this one's.  The
    // counters are folded into device-side totals by the same kernel (fic_get_counters fetches them).
    unsigned long long* totals = reinterpret_cast<unsigned long long*>(ctrl + BF_CTRL_TOTALS);
    h->bf_totals_live = true;
    if (!(flags & FIC_BF_NO_FALLBACK)) {
        h->prof_suppress = true;
        int rc_fb = bf_exact_scan8_dyn(h, d_img, H, W, mode, range_begin, n_ranges, ctrl, fallback, totals, d_idx, d_so, d_R, st);
        h->prof_suppress = false;
        FIC_TRY(rc_fb);
    }
    if (!h->eager.armed && !(flags & FIC_BF_NO_FALLBACK)) return FIC_OK;
    // Host-buffer call (the records go to pinned memory in front of the one wait the call needs anyway), or the testing
    // flag that wants to hear about undecided ranges.
    if (h->eager.armed) {
        for (int k = 0; k < 3; ++k)
            if (h->eager.bytes[k])
                FIC_CUDA(cudaMemcpyAsync(h->pin_out.as<char>() + h->eager.at[k], h->eager.src[k], h->eager.bytes[k],
                                         cudaMemcpyDeviceToHost, st));
        h->eager.done = true;
    }
    int* hc = h->pin_a.as<int>();
    FIC_CUDA(cudaMemcpyAsync(hc, ctrl, 64, cudaMemcpyDeviceToHost, st));
    FIC_TRY(fic_sync(h, st));
    h->counters[CNT_D2H] += 64;
    if (hc[2] != 0) { fic_set_error("bf_screen_kernel: pipeline wait %d timed out", hc[2]); return FIC_ECUDA; }
    if (flags & FIC_BF_NO_FALLBACK) {
        const int n_fb = hc[1];
        unsigned long long st_total;
        memcpy(&st_total, hc + 8, sizeof(st_total));
        h->counters[CNT_SURVIVORS] += (int64_t)st_total;
        h->counters[CNT_SCREEN_RANGES] += n_ranges - n_fb;
        h->counters[CNT_RESCAN_RANGES] += n_fb;
        if (n_fb > 0) { fic_set_error("screen left %d of %d ranges undecided", n_fb, n_ranges); return FIC_ECAP; }
    }
    return FIC_OK;
}

// Device totals of the pipeline (see fic_api.cu: fic_get_counters / fic_reset_counters).
int bf_totals_read(fic_handle* h, unsigned long long* t)
{
    if (!h->bf_ctrl.p) return FIC_OK;
    if (h->tail_valid) FIC_CUDA(cudaEventSynchronize(h->tail));
    unsigned long long* totals = reinterpret_cast<unsigned long long*>(h->bf_ctrl.as<int>() + BF_CTRL_TOTALS);
    if (t) {
        FIC_CUDA(cudaMemcpy(t, totals, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    } else {
        FIC_CUDA(cudaMemsetAsync(totals, 0, 4 * sizeof(unsigned long long), h->own_stream));
        FIC_CUDA(cudaStreamSynchronize(h->own_stream));
    }
    return FIC_OK;
}

int bf_screen_encode8(fic_handle* h, const uint8_t* d_img, int H, int W, int mode, int range_begin, int n_ranges, int32_t* d_idx,
                      double* d_so, double* d_R, unsigned flags, cudaStream_t st)
{
    return bf_screen_run(h, d_img, H, W, mode, range_begin, n_ranges, d_idx, d_so, d_R, flags, nullptr, st);
}

// ---- host driver of the quadtree search on the tensor cores -----------------------------------------------------------------
// One level of the quadtree partition: the n ranges of side 2^e in d_queue ([n][4] = si, sj, node, -) against the
// ((H - 2^(e+1)) / vs + 1) x ((W - 2^(e+1)) / hs + 1) domains of the pool's lattice (steps vs, hs).  d_rcls [n][2][3] = (class,
// rotation, orientation) of every range for positive / negative scales, d_dcls [NP][3] the same for every domain (row-major
// lattice), d_pool_offsets / d_pool_rows the pool's class lists (for the placeholder rows of empty classes).  Writes rec
// [n][6] / vals [n][3] for the ranges it decides and need_exact[n] = 1 for the others.  e = 3 (8x8 ranges, K = 64) or e = 2
// (4x4, K = 16).
template <int S>
static int qt_tc_search_impl(fic_handle* h, const uint8_t* d_img, int H, int W, const int32_t* d_queue, int n_ranges,
                             const int32_t* d_rcls, const int32_t* d_dcls, int nmodes, double max_scale, int scale_bits,
                             int offset_bits, int32_t* d_rec, double* d_vals, int32_t* d_need_exact, int hs, int vs,
                             const int32_t* d_pool_offsets, const int32_t* d_pool_rows, cudaStream_t st)
{
    constexpr int N = S * S;
    const int PW = (W - 2 * S) / hs + 1, PH = (H - 2 * S) / vs + 1;
    const int NP = PW * PH;
    const int n_tiles = cdiv(NP, BM);
    const int np_pad = n_tiles * BM;
    const int n_cols_pad = cdiv((long long)n_ranges * QT_COLS, BN) * BN;
    const int slots_pad = n_cols_pad / QT_COLS;
    const int n_groups = slots_pad / QT_RPG;
    const uint16_t* P = h->bf_boxsum.as<uint16_t>();
    // list entries per range: with the bounds seeded, the longest lists measured are around a hundred entries
    uint32_t cap = 2048;
    while (cap > 256 && (size_t)n_ranges * cap * 4 > ((size_t)1 << 30)) cap >>= 1;
    FIC_TRY(h->bf_pool.reserve((size_t)np_pad * N * 2 + 256));
    FIC_TRY(h->bf_stats.reserve((size_t)NP * (2 * sizeof(uint32_t) + 8 * sizeof(float) + 1 + sizeof(double)) + 1024));
    FIC_TRY(h->bf_A.reserve((size_t)n_cols_pad * N * 2 + 256));
    FIC_TRY(h->bf_thr.reserve((size_t)n_ranges * sizeof(QtcState) + 512));
    FIC_TRY(h->bf_surv.reserve((size_t)n_ranges * 4 + 64));
    FIC_TRY(h->bf_list.reserve((size_t)n_ranges * cap * 4 + 256));
    FIC_TRY(h->bf_best.reserve((size_t)slots_pad * (8 + sizeof(QtcRange)) + 512));
    if (!h->bf_ctrl.p) {
        FIC_TRY(h->bf_ctrl.reserve(256));
        FIC_CUDA(cudaMemsetAsync(h->bf_ctrl.p, 0, 256, st));
    }
    __half* poolA = h->bf_pool.as<__half>();
    uint32_t* sd4 = h->bf_stats.as<uint32_t>();
    uint32_t* sq4 = sd4 + NP;
    float* sds8 = reinterpret_cast<float*>(sq4 + NP);
    double* den = reinterpret_cast<double*>((reinterpret_cast<uintptr_t>(sds8 + (size_t)8 * NP) + 15) & ~(uintptr_t)15);
    uint8_t* dtyp = reinterpret_cast<uint8_t*>(den + NP);
    __half* rngB = h->bf_A.as<__half>();
    QtcState* state = h->bf_thr.as<QtcState>();
    uint32_t* cnt = h->bf_surv.as<uint32_t>();
    uint32_t* list = h->bf_list.as<uint32_t>();
    int* ctrl = h->bf_ctrl.as<int>();
    unsigned long long* bound = h->bf_best.as<unsigned long long>();
    QtcRange* rngc = reinterpret_cast<QtcRange*>((reinterpret_cast<uintptr_t>(bound + slots_pad) + 15) & ~(uintptr_t)15);

    qt_tcpool_kernel<S><<<cdiv(NP, 128), 128, 0, st>>>(P, H, W, d_dcls, poolA, sd4, sq4, sds8, dtyp, den, hs, vs);
    FIC_CHECK_LAUNCH(h, "qt_tcpool_kernel");
    qt_tcinit_kernel<S><<<cdiv((long long)slots_pad * 32, 256), 256, 0, st>>>(d_img, W, d_queue, d_rcls, n_ranges, state, cnt, bound, ctrl, slots_pad, rngB,
                                                                              d_need_exact, rngc);
    FIC_CHECK_LAUNCH(h, "qt_tcinit_kernel");

    const CUtensorMapSwizzle sw = N == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_32B;
    CUtensorMap mapA, mapB;
    FIC_TRY(make_map(&mapA, poolA, (uint64_t)NP, N, BM, sw));
    FIC_TRY(make_map(&mapB, rngB, (uint64_t)n_cols_pad, N, BN, sw));
    if (!h->qt_screen_attr_ready) {
        FIC_CUDA(cudaFuncSetAttribute(bf_screen_kernel<7, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SCREEN_SMEM));
        FIC_CUDA(cudaFuncSetAttribute(bf_screen_kernel<7, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SCREEN_SMEM));
        h->qt_screen_attr_ready = true;
    }
    ScreenParams sp;
    sp.hk = nullptr; sp.tau = nullptr; sp.n_elig = ctrl; sp.cnt = cnt; sp.list = list; sp.cap = cap; sp.NP = NP; sp.n_tiles = n_tiles;
    sp.seed_best = nullptr; sp.err_flag = ctrl + 2; sp.debug_mode = 0;
    sp.r_prev = 0;
    sp.window_tiles = (64 << 20) / (BM * N * 2);
    sp.dump = nullptr; sp.dtyp = dtyp;
    sp.grp_begin = 0; sp.grp_count = n_groups;
    sp.rngc = rngc; sp.bound = bound; sp.sd4 = sd4; sp.sq4 = sq4; sp.sds8 = sds8; sp.den = den;
    sp.max_scale = max_scale; sp.scale_bits = scale_bits; sp.offset_bits = offset_bits; sp.nmodes = nmodes;
    // Seed pass: every stride-th tile of 128 positions is looked at first and only lowers the bounds -- with the bounds at
    // +inf, and every CTA on a different range, nearly all of its candidates would be listed.  The full pass then lists
    // from all positions against bounds that mean something.
    const int stride = NP >= 64 * 1024 ? 61 : (NP >= 8192 ? 13 : (NP >= 1024 ? 5 : 1));
    if (stride > 1) {
        sp.m_cur = stride; sp.n_round_tiles = cdiv(n_tiles, stride); sp.seed_pass = 1;
        bf_screen_kernel<7, N><<<h->num_sms, SCREEN_THREADS, SCREEN_SMEM, st>>>(mapA, mapB, sp);
        FIC_CHECK_LAUNCH(h, "bf_screen_kernel(quadtree exact pass, seed)");
    }
    sp.m_cur = 1; sp.n_round_tiles = n_tiles; sp.seed_pass = 0;
    bf_screen_kernel<7, N><<<h->num_sms, SCREEN_THREADS, SCREEN_SMEM, st>>>(mapA, mapB, sp);
    FIC_CHECK_LAUNCH(h, "bf_screen_kernel(quadtree exact pass)");
#ifdef FIC_QTC_DEBUG
    {
        std::vector<uint32_t> c(n_ranges);
        cudaStreamSynchronize(st);
        cudaMemcpy(c.data(), cnt, (size_t)n_ranges * 4, cudaMemcpyDeviceToHost);
        unsigned long long sum = 0; uint32_t mx = 0;
        for (uint32_t v : c) { sum += v; mx = std::max(mx, v); }
        fprintf(stderr, "[qtc S=%d] ranges %d, listed %llu, longest %u (cap %u)\n", S, n_ranges, sum, mx, cap);
    }
#endif
    // the lists hold every candidate that was at or below its range's best score when it was looked at
    qt_tcpick_kernel<S><<<cdiv(n_ranges, QTC_WARPS), QTC_WARPS * 32, 0, st>>>(d_img, P, d_dcls, H, W, state, ctrl, cnt, list, cap, nmodes, max_scale,
                                                                             scale_bits, offset_bits, d_rec, d_vals, d_need_exact, hs, vs,
                                                                             d_pool_offsets, d_pool_rows);
    FIC_CHECK_LAUNCH(h, "qt_tcpick_kernel");
    return FIC_OK;
}

bool qt_tc_applicable(int H, int W, int e, double max_scale, int hs, int vs)
{
    if (e != 2 && e != 3) return false;
    const int S = 1 << e;
    if (H < 2 * S || W < 2 * S || hs < 1 || vs < 1 || !(max_scale > 0.0 && max_scale <= 1.0)) return false;
    const long long np = (long long)((H - 2 * S) / vs + 1) * ((W - 2 * S) / hs + 1);
    return np >= 512 && np < (1ll << 26);      // below that the fixed costs of the pipeline dominate; queue entries hold 26 bits of position
}

int qt_tc_search(fic_handle* h, const uint8_t* d_img, int H, int W, int e, const int32_t* d_queue, int n_ranges,
                 const int32_t* d_rcls, const int32_t* d_dcls, int nmodes, double max_scale, int scale_bits, int offset_bits,
                 int32_t* d_rec, double* d_vals, int32_t* d_need_exact, int hs, int vs, const int32_t* d_pool_offsets,
                 const int32_t* d_pool_rows, cudaStream_t st)
{
    if (e == 3)
        return qt_tc_search_impl<8>(h, d_img, H, W, d_queue, n_ranges, d_rcls, d_dcls, nmodes, max_scale, scale_bits, offset_bits, d_rec,
                                    d_vals, d_need_exact, hs, vs, d_pool_offsets, d_pool_rows, st);
    return qt_tc_search_impl<4>(h, d_img, H, W, d_queue, n_ranges, d_rcls, d_dcls, nmodes, max_scale, scale_bits, offset_bits, d_rec,
                                d_vals, d_need_exact, hs, vs, d_pool_offsets, d_pool_rows, st);
}

// ---- brute force with 4x4 ranges on the tensor cores (both encoders' arithmetic) --------------------------------
int bf_exact_generic_listed(fic_handle* h, const uint8_t* d_img, int H, int W, int range, int si, int sj, int mode,
                            int range_begin, int n_ranges, const int32_t* d_need, int32_t* d_idx, double* d_so, double* d_R,
                            cudaStream_t st);
int bf_prepare8(fic_handle* h, const uint8_t* d_img, int H, int W, bool with_stats, cudaStream_t st);

bool bf4_tc_applicable(int H, int W, int si, int sj, int mode)
{
    if ((mode != FIC_BF_MODE_JIT && mode != FIC_BF_MODE_NUMPY) || H < 16 || W < 16 || H % 4 || W % 4 || si < 1 || sj < 1) return false;
    const long long np = (long long)((H - 8 + si - 1) / si) * ((W - 8 + sj - 1) / sj);
    return np >= 4096 && np < (1ll << 25);      // (queue entries hold 25 bits of position)
}

int bf4_tc_encode(fic_handle* h, const uint8_t* d_img, int H, int W, int si, int sj, int mode, int range_begin, int n_ranges,
                  int32_t* d_idx, double* d_so, double* d_R, cudaStream_t st)
{
    if (n_ranges <= 0) return FIC_OK;
    constexpr int N = 16;
    const int npi = (H - 8 + si - 1) / si, npj = (W - 8 + sj - 1) / sj;
    const int NP = npi * npj;
    const int n_tiles = cdiv(NP, BM);
    const int n_cols_pad = cdiv((long long)n_ranges * 8, BN) * BN;
    const int slots_pad = n_cols_pad / 8;
    const int n_groups = slots_pad / 32;
    uint32_t cap = 1024;
    while (cap > 128 && (size_t)n_ranges * cap * 4 > ((size_t)1 << 30)) cap >>= 1;
    FIC_TRY(h->bf_boxsum.reserve((size_t)(H - 1) * (W - 1) * sizeof(uint16_t)));
    FIC_TRY(h->bf_pool.reserve((size_t)n_tiles * BM * N * 2 + 256));
    FIC_TRY(h->bf_stats.reserve((size_t)NP * (2 * sizeof(uint32_t) + sizeof(double)) + 1024));
    FIC_TRY(h->bf_A.reserve((size_t)n_cols_pad * N * 2 + 256));
    FIC_TRY(h->bf_surv.reserve((size_t)n_ranges * 4 + 64));
    FIC_TRY(h->bf_list.reserve((size_t)n_ranges * cap * 4 + 256));
    FIC_TRY(h->bf_best.reserve((size_t)slots_pad * (8 + sizeof(QtcRange)) + 512));
    FIC_TRY(h->t_need.reserve((size_t)n_ranges * sizeof(int32_t) + 64));
    if (!h->bf_ctrl.p) {
        FIC_TRY(h->bf_ctrl.reserve(256));
        FIC_CUDA(cudaMemsetAsync(h->bf_ctrl.p, 0, 256, st));
    }
    uint16_t* P = h->bf_boxsum.as<uint16_t>();
    __half* poolA = h->bf_pool.as<__half>();
    double* den = h->bf_stats.as<double>();
    uint32_t* sd4 = reinterpret_cast<uint32_t*>(den + NP);
    uint32_t* sq4 = sd4 + NP;
    __half* rngB = h->bf_A.as<__half>();
    uint32_t* cnt = h->bf_surv.as<uint32_t>();
    uint32_t* list = h->bf_list.as<uint32_t>();
    int* ctrl = h->bf_ctrl.as<int>();
    unsigned long long* bound = h->bf_best.as<unsigned long long>();
    QtcRange* rngc = reinterpret_cast<QtcRange*>((reinterpret_cast<uintptr_t>(bound + slots_pad) + 15) & ~(uintptr_t)15);
    int32_t* need = h->t_need.as<int32_t>();

    FIC_TRY(bf_prepare8(h, d_img, H, W, false, st));      // (the 2x2 cell sums; its statistics buffers are not used here)
    P = h->bf_boxsum.as<uint16_t>();
    den = h->bf_stats.as<double>();
    sd4 = reinterpret_cast<uint32_t*>(den + NP);
    sq4 = sd4 + NP;
    bf4_pool_kernel<<<cdiv(NP, 128), 128, 0, st>>>(P, W, si, sj, npj, NP, poolA, sd4, sq4, den);
    FIC_CHECK_LAUNCH(h, "bf4_pool_kernel");
    bf4_init_kernel<<<cdiv((long long)slots_pad * 32, 256), 256, 0, st>>>(d_img, W, mode, range_begin, n_ranges, slots_pad, rngB, rngc, bound,
                                                                        cnt, ctrl, need);
    FIC_CHECK_LAUNCH(h, "bf4_init_kernel");
    CUtensorMap mapA, mapB;
    FIC_TRY(make_map(&mapA, poolA, (uint64_t)NP, N, BM, CU_TENSOR_MAP_SWIZZLE_32B));
    FIC_TRY(make_map(&mapB, rngB, (uint64_t)n_cols_pad, N, BN, CU_TENSOR_MAP_SWIZZLE_32B));
    if (!h->bf4_attr_ready) {
        FIC_CUDA(cudaFuncSetAttribute(bf_screen_kernel<8, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SCREEN_SMEM));
        h->bf4_attr_ready = true;
    }
    ScreenParams sp;
    memset(&sp, 0, sizeof(sp));
    sp.n_elig = ctrl; sp.cnt = cnt; sp.list = list; sp.cap = cap; sp.NP = NP; sp.n_tiles = n_tiles;
    sp.err_flag = ctrl + 2;
    sp.r_prev = 0;
    sp.window_tiles = (64 << 20) / (BM * N * 2);
    sp.grp_begin = 0; sp.grp_count = n_groups;
    sp.rngc = rngc; sp.bound = bound; sp.sd4 = sd4; sp.sq4 = sq4; sp.den = den;
    sp.img = d_img; sp.boxsum = P; sp.img_W = W; sp.lat_si = si; sp.lat_sj = sj; sp.lat_npj = npj; sp.bf_mode = mode;
    // seed pass over every stride-th tile (only lowers the bounds), then the full pass (see qt_tc_search_impl)
    const int stride = NP >= 64 * 1024 ? 61 : (NP >= 8192 ? 13 : 5);
    bf4_first_bound_kernel<<<cdiv(n_ranges, 128), 128, 0, st>>>(sp, n_ranges);
    FIC_CHECK_LAUNCH(h, "bf4_first_bound_kernel");
    sp.m_cur = stride; sp.n_round_tiles = cdiv(n_tiles, stride); sp.seed_pass = 1;
    bf_screen_kernel<8, 16><<<h->num_sms, SCREEN_THREADS, SCREEN_SMEM, st>>>(mapA, mapB, sp);
    FIC_CHECK_LAUNCH(h, "bf_screen_kernel(4x4 ranges, seed)");
    sp.m_cur = 1; sp.n_round_tiles = n_tiles; sp.seed_pass = 0;
    bf_screen_kernel<8, 16><<<h->num_sms, SCREEN_THREADS, SCREEN_SMEM, st>>>(mapA, mapB, sp);
    FIC_CHECK_LAUNCH(h, "bf_screen_kernel(4x4 ranges)");
    bf4_pick_kernel<<<cdiv((long long)n_ranges * 32, 256), 256, 0, st>>>(sp, n_ranges, d_idx, d_so, d_R, need);
    FIC_CHECK_LAUNCH(h, "bf4_pick_kernel");
    // empty or overflowed lists (neither seen so far): the exact scan, for those ranges only
    return bf_exact_generic_listed(h, d_img, H, W, 4, si, sj, mode, range_begin, n_ranges, need, d_idx, d_so, d_R, st);
}
