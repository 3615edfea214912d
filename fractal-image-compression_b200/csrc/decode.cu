// Decoders and the 2x2 pooling primitive.
//
//  * bf_decode_*   bruteforce/decoder.py:12-32 -- float32 image, sweep 1 from zeros, then IN PLACE in range
//                  order (the reference aliases img and next_img at decoder.py:30), int-truncating pool
//                  (utils.py:22), unclipped uint8 cast.  The in-place order is reproduced exactly with a level
//                  schedule: range k may run once every earlier range it reads from (RAW), every earlier range
//                  that reads its target block (WAR) and every earlier writer of the same block (WAW) is done.
//  * tree_decode_* quadtree/decoder.py:34-116 and hvv2/decoder.py:29-89 -- true Jacobi sweeps in float64, one
//                  thread per output pixel through a pixel->leaf map, coalesced row-major writes.
//  * subsample     utils.py:13-34.
//
// HBM traffic per Jacobi sweep: H*W*(4 gathered reads + 1 write)*8 B + 4 B leaf id + 48 B per leaf record.
#include <algorithm>

#include "fic_common.cuh"
#include "scorer.cuh"

// ------------------------------------------------------------------------------------------------
// brute-force decoder
// ------------------------------------------------------------------------------------------------
#define BFD_THREADS 1024
#define BFD_RANGES_PER_STEP (BFD_THREADS / 32)
#define BFD_SHADOW_PIXELS 102400   // 200 KB of int16 next to the CTA's static shared memory (256 x 256, 320 x 320, ...)

// One transform in execution order; a step is BFD_RANGES_PER_STEP consecutive records (levels are padded with dst = -1).
struct BfdRec {
    int32_t di, dj, iso, dst;   // iso = orient * 4 + rot; dst = pixel index of the target block's corner, -1 = padding
    double s, o;
};

// One CTA per plane, one warp per range (two pixels per lane), one CTA-wide barrier per step.  A range may read its own
// target block, so its reads are separated from its writes by a warp barrier; no other range of the same level touches
// what it writes (RAW / WAR / WAW edges are all cross-level), so the CTA barrier is only needed between steps.  The
// record of the next step is fetched while the current one is gathered.
//
// SHADOW: a step is one dependent round trip to the image (gather -> affine -> scatter), ~1.9 us through L2.  The gather
// only ever uses the int32-truncated pixel (utils.py:22), so for images of up to BFD_SHADOW_PIXELS pixels the CTA keeps a
// truncated int16 copy in shared memory and gathers from there; the float32 values are only needed from the last sweep
// (every range is rewritten in every sweep) and go to global memory then.  A value outside int16 raises a flag and the
// plane is decoded again from zeros through global memory, in the same launch.
template <bool SHADOW>
__device__ __forceinline__ void bfd_sweeps(const BfdRec* __restrict__ recs, int n_steps, int width, int iterations,
                                           float* __restrict__ buf, int16_t* __restrict__ shadow, int* overflow)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    BfdRec cur = recs[warp];
    for (int it = 0; it < iterations; ++it) {
        const bool last = it + 1 == iterations;
        for (int st = 0; st < n_steps; ++st) {
            const int nx = st + 1 < n_steps ? st + 1 : 0;
            const BfdRec nxt = recs[(size_t)nx * BFD_RANGES_PER_STEP + warp];
            if (cur.dst >= 0) {
                float val[2];
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    const int px = lane + 32 * half, a = px >> 3, b = px & 7;
                    int acc = 0;
                    if (it > 0) {  // sweep 1 reads the all-zero image
                        // an isometry of the 16x16 block maps 2x2 cells onto 2x2 cells and the pool is an integer sum,
                        // so the cell of output pixel (a, b) is located once, on the 8x8 grid of cells
                        int ca, cb;
                        fic_iso_bf(cur.iso >> 2, cur.iso & 3, 8, a, b, ca, cb);
                        const size_t at = (size_t)(cur.di + 2 * ca) * width + cur.dj + 2 * cb;
#pragma unroll
                        for (int u = 0; u < 2; ++u)
#pragma unroll
                            for (int v = 0; v < 2; ++v) {
                                const size_t q = at + (size_t)u * width + v;
                                acc += SHADOW ? (int)shadow[q] : __float2int_rz(buf[q]);  // astype(int32), utils.py:22
                            }
                    }
                    val[half] = __double2float_rn(DA(DM(DM((double)acc, 0.25), cur.s), cur.o));
                }
                __syncwarp();  // every read of this range happens before any of its writes
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    const int px = lane + 32 * half;
                    const int at = cur.dst + (px >> 3) * width + (px & 7);
                    if (!SHADOW || last) buf[at] = val[half];
                    if (SHADOW) {
                        if (!(fabsf(val[half]) < 32768.0f)) *overflow = 1;
                        shadow[at] = (int16_t)__float2int_rz(val[half]);
                    }
                }
            }
            __syncthreads();
            cur = nxt;
        }
    }
}

template <bool SHADOW>
__global__ void __launch_bounds__(BFD_THREADS)
bf_decode_kernel(const BfdRec* __restrict__ recs_all, const long long* __restrict__ step_off, int width, int height,
                 int iterations, float* __restrict__ buf_all, uint8_t* __restrict__ out_all)
{
    extern __shared__ int16_t bfd_shadow[];
    __shared__ int s_overflow;
    const int tid = threadIdx.x;
    const BfdRec* recs = recs_all + step_off[blockIdx.x] * BFD_RANGES_PER_STEP;
    const int n_steps = (int)(step_off[blockIdx.x + 1] - step_off[blockIdx.x]);
    const int npx = width * height;
    float* buf = buf_all + (size_t)blockIdx.x * npx;
    uint8_t* out = out_all + (size_t)blockIdx.x * npx;
    if (n_steps > 0) {
        if (SHADOW) {
            for (int p = tid; p < npx; p += BFD_THREADS) bfd_shadow[p] = 0;
            if (tid == 0) s_overflow = 0;
            __syncthreads();
            bfd_sweeps<true>(recs, n_steps, width, iterations, buf, bfd_shadow, &s_overflow);
            __syncthreads();
            if (s_overflow) {   // some pixel left the int16 range: decode this plane again through global memory
                for (int p = tid; p < npx; p += BFD_THREADS) buf[p] = 0.0f;
                __syncthreads();
                bfd_sweeps<false>(recs, n_steps, width, iterations, buf, nullptr, nullptr);
            }
        } else {
            bfd_sweeps<false>(recs, n_steps, width, iterations, buf, nullptr, nullptr);
        }
    }
    __syncthreads();
    for (int p = tid; p < npx; p += BFD_THREADS) out[p] = (uint8_t)__float2int_rz(buf[p]);
}

// h_idx [planes][n][4], h_so [planes][n][2] on the host; d_out [planes][height*width] on the device
int bf_decode_impl(fic_handle* h, int planes, const int32_t* h_idx, const double* h_so, int n, int width, int height,
                   int iterations, uint8_t* d_out, cudaStream_t st)
{
    const int bw = width / 8, bh = height / 8;
    std::vector<BfdRec> recs;
    std::vector<long long> step_off(planes + 1, 0);
    std::vector<int> last_write, last_read, level(n), count, cursor;
    for (int pl = 0; pl < planes; ++pl) {
        const int32_t* idx = h_idx + (size_t)pl * n * 4;
        const double* so = h_so + (size_t)pl * n * 2;
        // level schedule over 8x8 blocks
        last_write.assign((size_t)bw * bh, -1);
        last_read.assign((size_t)bw * bh, -1);
        int n_levels = 0;
        for (int k = 0; k < n; ++k) {
            const int wi = k / bw, wj = k % bh;  // block coordinates of the target
            const int di = idx[4 * k], dj = idx[4 * k + 1];
            const int b0i = di / 8, b1i = (di + 15) / 8, b0j = dj / 8, b1j = (dj + 15) / 8;
            int lv = 0;
            for (int bi = b0i; bi <= b1i; ++bi)
                for (int bj = b0j; bj <= b1j; ++bj) lv = std::max(lv, last_write[(size_t)bi * bw + bj] + 1);  // RAW
            const size_t wb = (size_t)wi * bw + wj;
            lv = std::max(lv, last_read[wb] + 1);   // WAR (reads of this very range do not count: registered below)
            lv = std::max(lv, last_write[wb] + 1);  // WAW
            level[k] = lv;
            for (int bi = b0i; bi <= b1i; ++bi)
                for (int bj = b0j; bj <= b1j; ++bj) last_read[(size_t)bi * bw + bj] = std::max(last_read[(size_t)bi * bw + bj], lv);
            last_write[wb] = lv;
            n_levels = std::max(n_levels, lv + 1);
        }
        // records in (level, range) order, every level padded to whole steps
        count.assign(n_levels, 0);
        for (int k = 0; k < n; ++k) count[level[k]]++;
        cursor.assign(n_levels, 0);
        long long total = 0;
        for (int l = 0; l < n_levels; ++l) {
            cursor[l] = (int)total;
            total += (long long)((count[l] + BFD_RANGES_PER_STEP - 1) / BFD_RANGES_PER_STEP) * BFD_RANGES_PER_STEP;
        }
        const size_t base = recs.size();
        BfdRec pad{};
        pad.dst = -1;
        recs.resize(base + (size_t)total, pad);
        for (int k = 0; k < n; ++k) {
            BfdRec& r = recs[base + (size_t)cursor[level[k]]++];
            r.di = idx[4 * k]; r.dj = idx[4 * k + 1]; r.iso = idx[4 * k + 2] * 4 + idx[4 * k + 3];
            r.dst = (k / bw) * 8 * width + (k % bh) * 8;   // decoder.py:19-20 (sic)
            r.s = so[2 * k]; r.o = so[2 * k + 1];
        }
        step_off[pl + 1] = step_off[pl] + total / BFD_RANGES_PER_STEP;
    }
    const size_t rec_bytes = recs.size() * sizeof(BfdRec), off_bytes = step_off.size() * sizeof(long long);
    FIC_TRY(h->dec_sched.reserve(rec_bytes + off_bytes + 64));
    FIC_TRY(h->pin_b.reserve(rec_bytes + off_bytes + 64));
    memcpy(h->pin_b.p, step_off.data(), off_bytes);
    const size_t rec_at = (off_bytes + 31) & ~(size_t)31;
    if (rec_bytes) memcpy((char*)h->pin_b.p + rec_at, recs.data(), rec_bytes);
    FIC_TRY(fic_copy_in(h, h->dec_sched.p, h->pin_b.p, rec_at + rec_bytes, false, st));
    FIC_TRY(h->dec_img.reserve((size_t)planes * width * height * sizeof(float)));
    FIC_CUDA(cudaMemsetAsync(h->dec_img.p, 0, (size_t)planes * width * height * sizeof(float), st));
    const long long* d_off = h->dec_sched.as<long long>();
    const BfdRec* d_recs = reinterpret_cast<const BfdRec*>(h->dec_sched.as<char>() + rec_at);
    const size_t shadow_bytes = (size_t)width * height * sizeof(int16_t);
    if ((size_t)width * height <= BFD_SHADOW_PIXELS) {
        if (!h->dec_shadow_ready) {     // once per handle, i.e. per device
            FIC_CUDA(cudaFuncSetAttribute(bf_decode_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)(BFD_SHADOW_PIXELS * sizeof(int16_t))));
            h->dec_shadow_ready = true;
        }
        bf_decode_kernel<true><<<planes, BFD_THREADS, shadow_bytes, st>>>(d_recs, d_off, width, height, iterations,
                                                                        h->dec_img.as<float>(), d_out);
    } else {
        bf_decode_kernel<false><<<planes, BFD_THREADS, 0, st>>>(d_recs, d_off, width, height, iterations,
                                                               h->dec_img.as<float>(), d_out);
    }
    FIC_CHECK_LAUNCH(h, "bf_decode_kernel");
    // pin_b is reused by the next call: make sure the upload has been consumed
    FIC_TRY(fic_sync(h, st));
    return FIC_OK;
}

// ------------------------------------------------------------------------------------------------
// quadtree / HV Jacobi decoder
// ------------------------------------------------------------------------------------------------
// One warp per leaf (grid-stride over the leaves): most leaves of a real tree are 4x4 .. 16x16 blocks, for which a CTA per
// leaf is almost all launch overhead (262 144 CTAs of 64 threads for the 8x8 leaves of a 4096x4096 image: 0.14 ms against
// 0.18 ms for a whole Jacobi sweep).  Lanes walk the leaf row-major, so a row of 8+ pixels is one contiguous store.
#define TP_THREADS 256
__global__ void __launch_bounds__(TP_THREADS)
tree_paint_kernel(const int32_t* __restrict__ leaves, int n, int W, int32_t* __restrict__ leaf_map)
{
    const int lane = threadIdx.x & 31;
    const int warps = gridDim.x * (TP_THREADS / 32);
    for (int k = blockIdx.x * (TP_THREADS / 32) + (threadIdx.x >> 5); k < n; k += warps) {
        const int4 L = *reinterpret_cast<const int4*>(leaves + 8 * (size_t)k);
        const int ri = L.x, rj = L.y, hh = L.z, ww = L.w;
        for (int e = lane; e < hh * ww; e += 32) {
            const int a = e / ww, b = e - a * ww;
            leaf_map[(size_t)(ri + a) * W + rj + b] = k;
        }
    }
}

#define TD_THREADS 256

template <int KIND, bool REDUCE>
__global__ void __launch_bounds__(TD_THREADS)
tree_sweep_kernel(const int32_t* __restrict__ leaves, const double* __restrict__ so, const int32_t* __restrict__ leaf_map,
                  int H, int W, int f32_affine, const double* __restrict__ img, double* __restrict__ next,
                  double* __restrict__ red)
{
    const size_t p = (size_t)blockIdx.x * TD_THREADS + threadIdx.x;
    double e2 = 0.0, s2 = 0.0;
    if (p < (size_t)H * W) {
        const int k = leaf_map[p];
        if (k >= 0) {
            const int y = (int)(p / W), x = (int)(p - (size_t)y * W);
            const int4 L0 = *reinterpret_cast<const int4*>(leaves + 8 * (size_t)k);
            const int4 L1 = *reinterpret_cast<const int4*>(leaves + 8 * (size_t)k + 4);
            const int a = y - L0.x, b = x - L0.y, hh = L0.z, ww = L0.w, di = L1.x, dj = L1.y, orient = L1.z, rot = L1.w;
            const double scale = so[2 * (size_t)k], offset = so[2 * (size_t)k + 1];
            double val;
            if (KIND == FIC_TREE_QUADTREE) {
                // .T then rot90 (decoder.py:102-104) maps 2x2 cells onto 2x2 cells and the pool is an integer sum
                // (utils.py:22), so the cell of this output pixel is located once, on the hh x hh grid of cells
                int acc = 0, ca, cb;
                fic_iso_bf(orient, rot, hh, a, b, ca, cb);
                const double* cell = img + (size_t)(di + 2 * ca) * W + dj + 2 * cb;
#pragma unroll
                for (int u = 0; u < 2; ++u)
#pragma unroll
                    for (int v = 0; v < 2; ++v) acc += __double2int_rz(cell[(size_t)u * W + v]);
                val = DA(DM(DM((double)acc, 0.25), scale), offset);
            } else {
                float q[4];
#pragma unroll
                for (int u = 0; u < 2; ++u)
#pragma unroll
                    for (int v = 0; v < 2; ++v) {
                        int sa, sb;
                        fic_iso_hv(orient, rot, 2 * hh, 2 * ww, 2 * a + u, 2 * b + v, sa, sb);
                        q[u * 2 + v] = __double2float_rn(img[(size_t)(di + sa) * W + dj + sb]);
                    }
                float pooled = __fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(q[0], q[1]), q[2]), q[3]), 0.25f);  // utils.py:32-33
                if (f32_affine)
                    val = (double)__fadd_rn(__fmul_rn(pooled, __double2float_rn(scale)), __double2float_rn(offset));
                else
                    val = DA(DM((double)pooled, scale), offset);
            }
            next[p] = val;
            if (REDUCE) {
                double dlt = DS(val, img[p]);
                e2 = DM(dlt, dlt);
                s2 = DM(val, val);
            }
        } else if (REDUCE) {
            // uncovered pixel: next stays what it was (zero here)
            double v = next[p], dlt = DS(v, img[p]);
            e2 = DM(dlt, dlt);
            s2 = DM(v, v);
        }
    }
    if (REDUCE) {
        __shared__ double se[TD_THREADS / 32], ss[TD_THREADS / 32];
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            e2 = DA(e2, __shfl_xor_sync(0xffffffffu, e2, off));
            s2 = DA(s2, __shfl_xor_sync(0xffffffffu, s2, off));
        }
        if ((threadIdx.x & 31) == 0) { se[threadIdx.x >> 5] = e2; ss[threadIdx.x >> 5] = s2; }
        __syncthreads();
        if (threadIdx.x == 0) {
            double a = 0.0, b = 0.0;
            for (int w = 0; w < TD_THREADS / 32; ++w) { a = DA(a, se[w]); b = DA(b, ss[w]); }
            red[2 * (size_t)blockIdx.x] = a;
            red[2 * (size_t)blockIdx.x + 1] = b;
        }
    }
}

// fixed-order final sum of the per-CTA partials (deterministic run to run)
__global__ void tree_reduce_kernel(const double* __restrict__ red, int nblocks, double* __restrict__ out2)
{
    __shared__ double se[256], ss[256];
    double a = 0.0, b = 0.0;
    for (int k = threadIdx.x; k < nblocks; k += 256) { a = DA(a, red[2 * (size_t)k]); b = DA(b, red[2 * (size_t)k + 1]); }
    se[threadIdx.x] = a;
    ss[threadIdx.x] = b;
    __syncthreads();
    for (int off = 128; off > 0; off >>= 1) {
        if (threadIdx.x < off) { se[threadIdx.x] = DA(se[threadIdx.x], se[threadIdx.x + off]); ss[threadIdx.x] = DA(ss[threadIdx.x], ss[threadIdx.x + off]); }
        __syncthreads();
    }
    if (threadIdx.x == 0) { out2[0] = se[0]; out2[1] = ss[0]; }
}

// clip(0., 255.) and the uint8 cast (decoder.py:61): 8 B read, 1 B written per pixel; four pixels per thread so that the
// bytes leave as one 32-bit store (the image base comes from cudaMalloc, so groups of four stay aligned)
__global__ void tree_finish_kernel(const double* __restrict__ img, size_t npx, uint8_t* __restrict__ out)
{
    const size_t q = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (q >= npx) return;
    auto clip = [](double v) -> unsigned {
        v = v < 0.0 ? 0.0 : (v > 255.0 ? 255.0 : v);
        return (unsigned)__double2int_rz(v);
    };
    if (q + 4 <= npx && (reinterpret_cast<uintptr_t>(out) & 3u) == 0 && (reinterpret_cast<uintptr_t>(img) & 15u) == 0) {
        const double2 a = *reinterpret_cast<const double2*>(img + q), b = *reinterpret_cast<const double2*>(img + q + 2);
        *reinterpret_cast<uint32_t*>(out + q) = clip(a.x) | (clip(a.y) << 8) | (clip(b.x) << 16) | (clip(b.y) << 24);
    } else {
        for (size_t p = q; p < npx && p < q + 4; ++p) out[p] = (uint8_t)clip(img[p]);
    }
}

// ------------------------------------------------------------------------------------------------
// DeblockingQuadtreePostprocessor (quadtree/postprocessing.py:15-71)
// ------------------------------------------------------------------------------------------------
// The reference visits the roots in _de_partition order and each root breadth first; every leaf overwrites four lines
// of the result from the untouched input (its first row and column, and the row and column just past it), so where
// lines of two leaves coincide the LAST leaf visited wins.  Here every pixel pulls instead: at most four writers can
// reach it (its own leaf through its first row / first column, the leaf above through the row past its end, the leaf
// to the left through the column past its end) and the one with the largest visiting key is applied.
// Key = root index | depth | Morton code of the leaf's corner inside its root (row bit above column bit: the children
// are queued TL, TR, BL, BR, so one level of the breadth-first order is the Morton order) | line (0 top, 1 past-bottom,
// 2 left, 3 past-right: the order of the four assignments, postprocessing.py:59-66).
struct DeblockWeights {
    double w[4];      // { own, neighbour } for sides <= small_side, then for larger blocks (postprocessing.py:17-21)
    int small_side;
};

__device__ __forceinline__ unsigned fic_spread16(unsigned v)
{
    v &= 0xffffu;
    v = (v | (v << 8)) & 0x00ff00ffu;
    v = (v | (v << 4)) & 0x0f0f0f0fu;
    v = (v | (v << 2)) & 0x33333333u;
    v = (v | (v << 1)) & 0x55555555u;
    return v;
}

__global__ void qt_leaf_key_kernel(const int32_t* __restrict__ leaves, int n, const int32_t* __restrict__ roots, int n_roots,
                                   unsigned long long* __restrict__ keys)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const int ri = leaves[8 * (size_t)k], rj = leaves[8 * (size_t)k + 1], size = leaves[8 * (size_t)k + 2];
    unsigned long long key = 0;
    for (int r = 0; r < n_roots; ++r) {
        const int oi = roots[3 * r], oj = roots[3 * r + 1], side = roots[3 * r + 2];
        if (ri >= oi && ri < oi + side && rj >= oj && rj < oj + side) {
            const int depth = (31 - __clz(side)) - (31 - __clz(size));
            const unsigned morton = (fic_spread16((unsigned)(ri - oi)) << 1) | fic_spread16((unsigned)(rj - oj));
            key = ((unsigned long long)(r + 1) << 39) | ((unsigned long long)depth << 34) | ((unsigned long long)morton << 2);
            break;
        }
    }
    keys[k] = key;
}

template <bool TO_U8>
__global__ void qt_deblock_kernel(const double* __restrict__ img, const int32_t* __restrict__ leaves,
                                  const unsigned long long* __restrict__ keys, const int32_t* __restrict__ leaf_map, int H,
                                  int W, DeblockWeights dw, double* __restrict__ res, uint8_t* __restrict__ out)
{
    const size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= (size_t)H * W) return;
    const int y = (int)(p / W), x = (int)(p - (size_t)y * W);
    const double own = img[p];
    double v = own;
    unsigned long long best = 0;
    auto offer = [&](unsigned long long key, int size, size_t q) {
        if (key > best) {
            best = key;
            const int c = size <= dw.small_side ? 0 : 2;
            v = DA(DM(dw.w[c], own), DM(dw.w[c + 1], img[q]));
        }
    };
    const int k = leaf_map[p];
    if (k >= 0) {
        const int ri = leaves[8 * (size_t)k], rj = leaves[8 * (size_t)k + 1], size = leaves[8 * (size_t)k + 2];
        if (y == ri && ri >= 1) offer(keys[k] | 0ull, size, p - W);
        if (x == rj && rj >= 1) offer(keys[k] | 2ull, size, p - 1);
    }
    if (y >= 1 && y + 1 < H) {
        const int a = leaf_map[p - W];
        if (a >= 0 && a != k) {
            const int size = leaves[8 * (size_t)a + 2];
            if (leaves[8 * (size_t)a] + size == y) offer(keys[a] | 1ull, size, p + W);
        }
    }
    if (x >= 1 && x + 1 < W) {
        const int b = leaf_map[p - 1];
        if (b >= 0 && b != k) {
            const int size = leaves[8 * (size_t)b + 2];
            if (leaves[8 * (size_t)b + 1] + size == x) offer(keys[b] | 3ull, size, p + 1);
        }
    }
    if (TO_U8) {
        v = v < 0.0 ? 0.0 : (v > 255.0 ? 255.0 : v);  // clip(0., 255.), decoder.py:61
        out[p] = (uint8_t)__double2int_rz(v);
    } else {
        res[p] = v;
    }
}

// _de_partition (postprocessing.py:29-41): the largest power-of-two square at the corner, then the strip to its right
// (full height), then the strip below it (square width)
static void qt_root_squares(int i, int j, int w, int h, std::vector<int32_t>& out)
{
    int e = 0;
    while ((2 << e) <= (w < h ? w : h)) ++e;
    const int side = 1 << e;
    out.push_back(i);
    out.push_back(j);
    out.push_back(side);
    if (side < w) qt_root_squares(i, j + side, w - side, h, out);
    if (side < h) qt_root_squares(i + side, j, side, h - side, out);
}

// d_img float64 [H,W] -> d_res float64 (or, d_out8 != nullptr, the clipped uint8 image).  leaf_map must already hold
// the leaf index of every pixel (tree_paint_kernel).
static int qt_deblock_run(fic_handle* h, const double* d_img, const int32_t* d_leaves, int n, int H, int W,
                          const DeblockWeights& dw, const int32_t* leaf_map, double* d_res, uint8_t* d_out8, cudaStream_t st)
{
    std::vector<int32_t> roots;
    qt_root_squares(0, 0, W, H, roots);
    const int n_roots = (int)(roots.size() / 3);
    FIC_TRY(h->dec_roots.reserve(roots.size() * sizeof(int32_t)));
    FIC_TRY(h->dec_keys.reserve(((size_t)n + 1) * sizeof(unsigned long long)));
    FIC_TRY(fic_copy_in(h, h->dec_roots.p, roots.data(), roots.size() * sizeof(int32_t), false, st));
    FIC_TRY(fic_sync(h, st));  // `roots` is pageable and goes out of scope
    if (n > 0) {
        qt_leaf_key_kernel<<<(n + 255) / 256, 256, 0, st>>>(d_leaves, n, h->dec_roots.as<int32_t>(), n_roots,
                                                            h->dec_keys.as<unsigned long long>());
        FIC_CHECK_LAUNCH(h, "qt_leaf_key_kernel");
    }
    const size_t npx = (size_t)H * W;
    const unsigned grid = (unsigned)((npx + 255) / 256);
    if (d_out8)
        qt_deblock_kernel<true><<<grid, 256, 0, st>>>(d_img, d_leaves, h->dec_keys.as<unsigned long long>(), leaf_map, H, W, dw,
                                                      nullptr, d_out8);
    else
        qt_deblock_kernel<false><<<grid, 256, 0, st>>>(d_img, d_leaves, h->dec_keys.as<unsigned long long>(), leaf_map, H, W, dw,
                                                       d_res, nullptr);
    FIC_CHECK_LAUNCH(h, "qt_deblock_kernel");
    return FIC_OK;
}

static DeblockWeights qt_default_weights()
{
    DeblockWeights dw;
    dw.w[0] = 5. / 6.;           // own_weight_max_depth
    dw.w[1] = 1. - 5. / 6.;      // neigh_weight_max_depth
    dw.w[2] = 2. / 3.;           // own_weight_not_max_depth
    dw.w[3] = 1. - 2. / 3.;      // neigh_weight_not_max_depth
    dw.small_side = 4;           // max_depth_range_size
    return dw;
}

int qt_deblock_impl(fic_handle* h, const double* d_img, const int32_t* d_leaves, int n, int H, int W, const double* w4,
                    int small_side, double* d_res, cudaStream_t st)
{
    const size_t npx = (size_t)H * W;
    FIC_TRY(h->dec_map.reserve(npx * sizeof(int32_t)));
    FIC_CUDA(cudaMemsetAsync(h->dec_map.p, 0xff, npx * sizeof(int32_t), st));
    if (n > 0) {
        tree_paint_kernel<<<std::min((n + TP_THREADS / 32 - 1) / (TP_THREADS / 32), h->num_sms * 32), TP_THREADS, 0, st>>>(d_leaves, n, W, h->dec_map.as<int32_t>());
        FIC_CHECK_LAUNCH(h, "tree_paint_kernel");
    }
    DeblockWeights dw = qt_default_weights();
    if (w4) for (int c = 0; c < 4; ++c) dw.w[c] = w4[c];
    if (small_side >= 0) dw.small_side = small_side;
    return qt_deblock_run(h, d_img, d_leaves, n, H, W, dw, h->dec_map.as<int32_t>(), d_res, nullptr, st);
}

template <int KIND>
static int launch_sweep(fic_handle* h, bool reduce, const int32_t* leaves, const double* so, const int32_t* map, int H,
                        int W, int f32_affine, const double* img, double* next, double* red, int nblocks, cudaStream_t st)
{
    if (reduce)
        tree_sweep_kernel<KIND, true><<<nblocks, TD_THREADS, 0, st>>>(leaves, so, map, H, W, f32_affine, img, next, red);
    else
        tree_sweep_kernel<KIND, false><<<nblocks, TD_THREADS, 0, st>>>(leaves, so, map, H, W, f32_affine, img, next, red);
    FIC_CHECK_LAUNCH(h, "tree_sweep_kernel");
    return FIC_OK;
}

// post & FIC_QT_DEBLOCK: the float image goes through the deblocking filter before the clip (decoder.py:58-61);
// post & FIC_TREE_OUT_F64: d_out receives the float64 image (filtered or not) instead of the clipped bytes.
int tree_decode_impl(fic_handle* h, int kind, const int32_t* d_leaves, const double* d_so, int n, int H, int W,
                     int iterations, double stop, int min_it, int max_it, int f32_affine, unsigned post, void* d_out,
                     int32_t* iters_done, cudaStream_t st)
{
    const size_t npx = (size_t)H * W;
    const int nblocks = (int)((npx + TD_THREADS - 1) / TD_THREADS);
    FIC_TRY(h->dec_img.reserve(npx * sizeof(double)));
    FIC_TRY(h->dec_next.reserve(npx * sizeof(double)));
    FIC_TRY(h->dec_map.reserve(npx * sizeof(int32_t)));
    FIC_TRY(h->dec_red.reserve(((size_t)nblocks * 2 + 2) * sizeof(double)));
    FIC_TRY(h->pin_b.reserve(64));
    FIC_CUDA(cudaMemsetAsync(h->dec_img.p, 0, npx * sizeof(double), st));
    FIC_CUDA(cudaMemsetAsync(h->dec_next.p, 0, npx * sizeof(double), st));
    FIC_CUDA(cudaMemsetAsync(h->dec_map.p, 0xff, npx * sizeof(int32_t), st));
    if (n > 0) {
        tree_paint_kernel<<<std::min((n + TP_THREADS / 32 - 1) / (TP_THREADS / 32), h->num_sms * 32), TP_THREADS, 0, st>>>(d_leaves, n, W, h->dec_map.as<int32_t>());
        FIC_CHECK_LAUNCH(h, "tree_paint_kernel");
    }
    double* img = h->dec_img.as<double>();
    double* next = h->dec_next.as<double>();
    double* red = h->dec_red.as<double>();
    double* red_out = red + (size_t)nblocks * 2;
    const int32_t* map = h->dec_map.as<int32_t>();
    int done = 0;
    auto sweep = [&](bool reduce) -> int {
        int rc = kind == FIC_TREE_QUADTREE
                     ? launch_sweep<FIC_TREE_QUADTREE>(h, reduce, d_leaves, d_so, map, H, W, f32_affine, img, next, red, nblocks, st)
                     : launch_sweep<FIC_TREE_HV>(h, reduce, d_leaves, d_so, map, H, W, f32_affine, img, next, red, nblocks, st);
        if (rc < 0) return rc;
        std::swap(img, next);  // img[:] = next_img[:]; every covered pixel is rewritten each sweep
        ++done;
        return FIC_OK;
    };
    if (iterations == FIC_ITER_AUTO) {
        for (int it = 0; it < min_it; ++it) FIC_TRY(sweep(false));
        for (int it = min_it; it < max_it; ++it) {
            FIC_TRY(sweep(true));
            tree_reduce_kernel<<<1, 256, 0, st>>>(red, nblocks, red_out);
            FIC_CHECK_LAUNCH(h, "tree_reduce_kernel");
            double* hp = h->pin_b.as<double>();
            FIC_TRY(fic_copy_out(h, hp, red_out, 2 * sizeof(double), false, st));
            FIC_TRY(fic_sync(h, st));
            if (sqrt(hp[0]) / sqrt(hp[1]) < stop) break;  // decoder.py:45-49
        }
    } else {
        for (int it = 0; it < iterations; ++it) FIC_TRY(sweep(false));
    }
    if (iters_done) *iters_done = done;
    const bool f64 = post & FIC_TREE_OUT_F64;
    if (post & FIC_QT_DEBLOCK)
        return qt_deblock_run(h, img, d_leaves, n, H, W, qt_default_weights(), map, f64 ? (double*)d_out : nullptr,
                              f64 ? nullptr : (uint8_t*)d_out, st);
    if (f64) return fic_copy_out(h, d_out, img, npx * sizeof(double), true, st);
    tree_finish_kernel<<<(unsigned)((npx + 1023) / 1024), 256, 0, st>>>(img, npx, (uint8_t*)d_out);
    FIC_CHECK_LAUNCH(h, "tree_finish_kernel");
    return FIC_OK;
}

// ------------------------------------------------------------------------------------------------
// utils.average_subsample / average_subsample_jit
// ------------------------------------------------------------------------------------------------
__global__ void subsample_kernel(const double* __restrict__ arr, int h2, int w2, int jit, double* __restrict__ out)
{
    const int ow = w2 / 2, oh = h2 / 2;
    int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= ow * oh) return;
    int a = e / ow, b = e - a * ow;
    const double* p = arr + (size_t)(2 * a) * w2 + 2 * b;
    if (jit) {
        float s4 = __fadd_rn(__fadd_rn(__fadd_rn(__double2float_rn(p[0]), __double2float_rn(p[1])), __double2float_rn(p[w2])),
                             __double2float_rn(p[w2 + 1]));
        out[e] = (double)__fmul_rn(s4, 0.25f);
    } else {
        int acc = __double2int_rz(p[0]) + __double2int_rz(p[w2]) + __double2int_rz(p[1]) + __double2int_rz(p[w2 + 1]);
        out[e] = DM((double)acc, 0.25);
    }
}

int subsample_impl(fic_handle* h, const double* d_arr, int h2, int w2, int jit, double* d_out, cudaStream_t st)
{
    int n = (h2 / 2) * (w2 / 2);
    subsample_kernel<<<(n + 255) / 256, 256, 0, st>>>(d_arr, h2, w2, jit, d_out);
    FIC_CHECK_LAUNCH(h, "subsample_kernel");
    return FIC_OK;
}
