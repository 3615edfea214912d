/*
 * fic.h -- C ABI of libfic_b200.so: the B200 (sm_100a) implementation of the
 * range->domain block-matching hot path of Fl0k3n/fractal-image-compression.
 *
 * The reference exposes no FFI of its own (it is pure Python + numba); the drop-in
 * boundary is its Python class API.  Each entry point below is what a
 * ctypes stub inside the reference's class would bind; the reference interface it
 * replaces is cited as file:line relative to the reference's src/ directory.
 * INTEGRATION.md shows the stubs.
 *
 * Conventions
 *  - Plain pointers and sizes only.  No torch / numpy types.
 *  - Every function returns >= 0 on success and a negative FIC_E* code on failure;
 *    fic_last_error() returns a human-readable message for the calling thread.
 *  - Buffers are caller-allocated.  By default pointers are HOST pointers and the
 *    library performs the host<->device copies itself; pass FIC_DEVICE_PTRS in
 *    `flags` when every buffer argument of the call is a device pointer on the
 *    handle's GPU.  The library owns only internal workspaces (cached in the handle).
 *  - `stream` is a cudaStream_t passed as void* (NULL = the handle's own stream).
 *    With host pointers the call returns after the results are in the host buffers;
 *    with device pointers the work is enqueued on `stream` and the call returns
 *    without synchronising unless it needs a size from the device (tree encoders,
 *    the brute-force decoder's schedule).  fic_bf_encode with device pointers never
 *    waits: its counters stay on the device until fic_get_counters asks for them, and
 *    a pipeline fault of an earlier call is reported there or by fic_synchronize.
 *  - A handle is used by one host thread at a time; one handle per GPU.  Batches and
 *    range slices on several GPUs use one handle (and one process or thread) per GPU:
 *    no entry point takes a device list.
 *  - All calls on a handle share its cached workspaces.  The library orders them itself:
 *    every call leaves an event on its stream and a call on a different stream first
 *    waits for the previous call's event, so device-pointer calls on several streams are
 *    safe but run one after the other.  fic_synchronize waits for the last call
 *    whatever stream it was given.
 *  - Every entry point makes the handle's device current while it runs and puts the
 *    caller's current device back before it returns.
 *  - There is NO CPU fallback: without a usable sm_100 GPU fic_create fails.
 */
#ifndef FIC_H_
#define FIC_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FIC_OK 0
#define FIC_EINVAL (-1)   /* bad argument (the reference would raise / crash on it) */
#define FIC_ECAP (-2)     /* caller-provided output capacity too small */
#define FIC_ECUDA (-3)    /* CUDA runtime / driver error, see fic_last_error() */
#define FIC_ENOGPU (-4)   /* no usable sm_100 device */
#define FIC_ENOMEM (-5)

#define FIC_DEVICE_PTRS 1u  /* all buffer arguments are device pointers */
#define FIC_BF_EXACT_ONLY 2u /* brute force: skip the tensor-core screen, run the exact scan only */
#define FIC_BF_NO_FALLBACK 4u /* brute force (testing): fail instead of re-scanning ranges the screen could not decide */
#define FIC_QT_DEBLOCK 8u     /* fic_tree_decode, quadtree: DeblockingQuadtreePostprocessor before the clip (decoder.py:58-59) */
#define FIC_TREE_OUT_F64 16u  /* fic_tree_decode: `out` is double[H*W], the float image decoder.py:61 clips (not clipped here) */

#define FIC_BF_MODE_NUMPY 0 /* BruteForceEncoder semantics  (bruteforce/bf_encoder.py)      */
#define FIC_BF_MODE_JIT 1   /* JitedBruteForceEncoder semantics (bruteforce/jit_bfencoder.py) */

#define FIC_TREE_QUADTREE 0
#define FIC_TREE_HV 1
#define FIC_ITER_AUTO (-1) /* quadtree/decoder.py:11 */

typedef struct fic_handle fic_handle;

/* ---- library / handle ------------------------------------------------------------------ */
const char* fic_version(void);
const char* fic_last_error(void);
int fic_device_count(void);
/* Creates a handle bound to CUDA device `device`.  Fails with FIC_ENOGPU when the device is
 * missing or is not compute capability 10.x. */
int fic_create(int device, fic_handle** out);
int fic_destroy(fic_handle* h);
/* Counters since creation / last reset: [0] kernels launched by this library,
 * [1] brute-force ranges decided by the tensor-core screen, [2] ranges re-scanned exactly,
 * [3] screen survivors rescored, [4] bytes copied H2D, [5] bytes copied D2H, [6..7] reserved. */
int fic_get_counters(fic_handle* h, int64_t out[8]);
int fic_reset_counters(fic_handle* h);
int fic_synchronize(fic_handle* h);
/* Device-side timing of the dominant kernel of each path (the brute-force scan / screen kernel): when enabled,
 * every launch of it is bracketed by CUDA events on the launching stream.  fic_profile_read synchronises those
 * events, returns their summed duration and launch count since the last read, and clears them.  Used by
 * bench.py for the roofline figure; off by default (two event records per launch). */
int fic_profile(fic_handle* h, int enable);
int fic_profile_read(fic_handle* h, double* total_ms, int64_t* launches);

/* ---- brute force ----------------------------------------------------------------------- */
/*
 * Replaces BruteForceEncoder._find_best_domain + _calc_rms_error over all ranges
 *   (bruteforce/bf_encoder.py:18-71, mode FIC_BF_MODE_NUMPY) and
 * JitedBruteForceEncoder.encode -> _find_best_domain / _calc_rms_error / _calc_R_error
 *   (bruteforce/jit_bfencoder.py:18-95, mode FIC_BF_MODE_JIT).
 * img: uint8 [H,W] row-major; H, W multiples of `range`; domain must equal 2*range.
 * Ranges are numbered row-major (jit_bfencoder.py:22-23); the call encodes ranges
 * [range_begin, range_end) -- the slice a rank owns when range blocks are sharded.
 * out_idx [n,4] int32 = (start_i, start_j, orientation, rotation); out_so [n,2] float64 =
 * (s, o); out_R [n] float64 = collage error R of the winner (may be NULL).
 */
int fic_bf_encode(fic_handle* h, const uint8_t* img, int H, int W, int range, int domain, int stride_i,
                  int stride_j, int mode, int range_begin, int range_end, int32_t* out_idx, double* out_so,
                  double* out_R, unsigned flags, void* stream);

/*
 * Batch form: imgs uint8 [B,H,W]; all ranges of every image; outputs [B,n,...].
 * This is the image-parallel path: one call per rank (= per GPU, per handle) on its share of the batch.  There is
 * no device-list argument -- a handle is bound to one GPU; fic_b200/parallel.py shows the sharding over ranks.
 */
int fic_bf_encode_batch(fic_handle* h, const uint8_t* imgs, int B, int H, int W, int range, int domain,
                        int mode, int32_t* out_idx, double* out_so, double* out_R, unsigned flags,
                        void* stream);

/*
 * Diagnostic hook for the tensor-core path of fic_bf_encode (no counterpart in the reference; the quantity it
 * exposes is the cross-correlation numerator of bruteforce/jit_bfencoder.py:76): runs the pool kernel, the range
 * operand kernel and the tcgen05 correlation GEMM for the n_ranges <= 32 ranges starting at range_begin and returns
 * the raw fp32 accumulators  X' = sum_k r'_k*D4'_k  (range and 4x-pooled domain centred by their rounded means;
 * (64*ab - sum(r)*sum(d))/4 in the reference's terms equals X' - rho*eps/64 with |rho*eps/64| <= 16):
 * out_u float32 [ceil(NP/128)*128][256], row = stride-1 domain position di*(W-16)+dj, column = range*8 + orientation*4
 * + rotation.  Tests use it to check TMA / descriptor / MMA plumbing and the accumulator's exactness.
 */
int fic_bf_screen_probe(fic_handle* h, const uint8_t* img, int H, int W, int range_begin, int n_ranges, float* out_u,
                        unsigned flags, void* stream);

/*
 * No counterpart in the reference (the reference never manages device memory).
 *
 * The same exact 8x8 / 16x16 brute-force encode as fic_bf_encode, for callers that are not allowed to allocate
 * device memory around the call: plugins inside an application whose own pool allocator owns the GPU's memory,
 * serving or capture loops that must stay away from cudaMalloc/cudaFree (both synchronise the device and can fail
 * late), sandboxed workers that only ever see pre-allocated buffers.  It needs no fic_handle and calls neither
 * cudaMalloc nor cudaFree nor any allocating library routine; its scratch (box sums, per-position statistics,
 * per-CTA winners) lives in a fixed array inside the loaded module.
 *   - every buffer is a DEVICE pointer on the current device: d_img uint8 [H,W]; outputs as in fic_bf_encode,
 *     sized for range_end - range_begin ranges; d_R is mandatory.  Work is enqueued on `stream` (NULL = the
 *     default stream); the call does not synchronise.
 *   - limit: H, W <= 256 (multiples of 8).  Anything larger returns FIC_EINVAL with a message; use fic_bf_encode.
 *   - cost: about 3.4 MB of device memory on every GPU that loads the library, held until the process exits.
 *   - the array is shared by all calls on a GPU and keeps the previous call's contents; nothing in it needs to
 *     start at zero (each call writes everything it later reads).  Calls that overlap on the GPU -- two streams,
 *     two host threads -- race on it: issue them one after the other.
 *   - the first call on each device resolves the array's device address; do that first call before starting a
 *     stream capture, and from one host thread (later calls may come from any thread, one at a time).
 */
int fic_bf_encode_static(const uint8_t* d_img, int H, int W, int stride_i, int stride_j, int mode, int range_begin,
                         int range_end, int32_t* d_idx, double* d_so, double* d_R, void* stream);

/*
 * Replaces Decoder.decode (bruteforce/decoder.py:12-32), including its buffer aliasing after the
 * first sweep, int-truncating pool and unclipped uint8 cast.  idx [n,4], so [n,2] as above, n =
 * number of transforms in range order; out uint8 [height,width].  Range 8 / domain 16 are
 * hard-coded as in the reference (decoder.py:9-10).
 */
int fic_bf_decode(fic_handle* h, const int32_t* idx, const double* so, int n, int width, int height,
                  int iterations, uint8_t* out, unsigned flags, void* stream);

/*
 * `planes` independent decodes of equal shape in one launch (one CTA per plane): the colour fan-out of
 * quadtree/rgb.py-style per-channel coding and image batches (BASELINE.json configs[4]).  No single-call counterpart
 * in the reference, which loops over Decoder.decode.  idx [planes,n,4], so [planes,n,2], out uint8 [planes,height,width].
 */
int fic_bf_decode_batch(fic_handle* h, int planes, const int32_t* idx, const double* so, int n, int width, int height,
                        int iterations, uint8_t* out, unsigned flags, void* stream);

/* ---- pooling --------------------------------------------------------------------------- */
/* utils.average_subsample (utils.py:13-25, jit=0: int32-truncating, float64 result) and
 * utils.average_subsample_jit (utils.py:27-34, jit=1: float32 adds).  arr float64 [h2,w2] ->
 * out float64 [h2/2,w2/2] (jit results are float32 values widened). */
int fic_average_subsample(fic_handle* h, const double* arr, int h2, int w2, int jit, double* out,
                          unsigned flags, void* stream);

/* ---- quadtree -------------------------------------------------------------------------- */
/*
 * Replaces QuadtreeEncoder.encode and everything under it (quadtree/encoder.py:67-384):
 * domain classification and 72-class pools, per-level batched range search, quantised scorer,
 * split decisions.  Output is the forest as a PREORDER node stream, roots concatenated in
 * _partition_image order (encoder.py:111-125), children TL,TR,BL,BR (encoder.py:141-146):
 *   nodes[k] = { kind (0 internal, 1 leaf), start_i, start_j, orientation, rotation,
 *                quantized_scale, quantized_offset, size_exponent }
 *   vals[k]  = { scale, offset, best_err }      (pre-quantisation s, o; MSE of the winner)
 * Returns the node count, FIC_ECAP if cap is too small (2*H*W+64 always suffices), FIC_EINVAL
 * where the reference raises (a searched range whose 2x domain does not fit the image).
 * nodes / vals / n_roots are HOST pointers even with FIC_DEVICE_PTRS (img may be on device).
 */
int fic_qt_encode(fic_handle* h, const uint8_t* img, int H, int W, int scale_bits, int offset_bits,
                  double max_scale, int full_class_search, int only_positive_scale, int min_range_side,
                  int max_range_side, double error_tolerance, int domain_hstep, int domain_vstep,
                  int32_t* nodes, double* vals, int cap, int32_t* n_roots, unsigned flags, void* stream);

/* _classify (quadtree/encoder.py:195-259) for `count` blocks: blocks[k] = {start_i, start_j, side, negate};
 * out[k] = {mean_class, variance_class, rotation, orientation}.  blocks/out follow `flags`. */
int fic_qt_classify(fic_handle* h, const uint8_t* img, int H, int W, const int32_t* blocks, int count,
                    int32_t* out, unsigned flags, void* stream);

/* ---- HV -------------------------------------------------------------------------------- */
/*
 * Replaces HVEncoder.encode (hvv2/encoder.py:24-188).  Preorder stream of the binary tree:
 *   leaf     nodes[k] = { 1, start_i, start_j, orientation, rotation, qs, qo, 0 }
 *   internal nodes[k] = { 0, vertical_split, split_idx, 0, 0, 0, 0, 0 }  then child 1, child 2
 *   vals[k] = { scale, offset, err }
 * Host output pointers as for fic_qt_encode.
 */
int fic_hv_encode(fic_handle* h, const uint8_t* img, int H, int W, int min_range_height, int min_range_width,
                  int max_range_height, int max_range_width, int scale_bits, int offset_bits,
                  double err_threshold, double max_scale, int32_t* nodes, double* vals, int cap,
                  unsigned flags, void* stream);

/* One _run_domain_search (hvv2/encoder.py:97-126) per rect: rects[k] = {i, j, w, h};
 * rec[k] = {start_i, start_j, orientation, rotation, qs, qo}; vals[k] = {scale, offset, err}. */
int fic_hv_search(fic_handle* h, const uint8_t* img, int H, int W, const int32_t* rects, int count,
                  int scale_bits, int offset_bits, double max_scale, int32_t* rec, double* vals,
                  unsigned flags, void* stream);

/* ---- tree decoders --------------------------------------------------------------------- */
/*
 * Replaces QuadtreeDecoder.decode (quadtree/decoder.py:34-116, kind FIC_TREE_QUADTREE) and
 * HVDecoder.decode (hvv2/decoder.py:29-89, kind FIC_TREE_HV) after the host has flattened the
 * tree: leaves[k] = { range_i, range_j, height, width, start_i, start_j, orientation, rotation },
 * so[k] = { scale, offset } (already de-quantised by the caller, decoder.py:90-98).
 * iterations = FIC_ITER_AUTO reproduces the 2 + <=8 sweep rule with the Frobenius stop test.
 * f32_affine != 0: HV only, `pooled*scale+offset` evaluated in float32 (python-float scale, i.e.
 * using_not_quantized_values()).  out uint8 [H,W]; *iters_done (host, may be NULL) = sweeps run.
 * flags & FIC_QT_DEBLOCK (quadtree only): decode(qimg, DeblockingQuadtreePostprocessor()) -- the filter of
 * fic_qt_deblock with the reference's default weights runs on the float image before the clip
 * (decoder.py:58-61).  flags & FIC_TREE_OUT_F64: out is double[H*W] and receives that float image unclipped,
 * which is what a caller-side QuadtreePostprocessor subclass is handed by the reference (decoder.py:59).
 */
int fic_tree_decode(fic_handle* h, int kind, const int32_t* leaves, const double* so, int n, int H, int W,
                    int iterations, double stop_on_relative_error, int min_iterations, int max_iterations,
                    int f32_affine, uint8_t* out, int32_t* iters_done, unsigned flags, void* stream);

/* ---- quadtree post-processing ------------------------------------------------------------ */
/*
 * Replaces DeblockingQuadtreePostprocessor.postprocess (quadtree/postprocessing.py:24-71): img float64 [H,W] ->
 * out float64 [H,W].  leaves as for fic_tree_decode (only range_i, range_j, size are read) in ANY order: the
 * order-dependent overwrites of the reference (roots in _de_partition order, breadth first inside a root, four
 * line assignments per leaf) are reproduced from the geometry.  weights = { own, neighbour } for blocks of side
 * <= max_depth_range_size, then { own, neighbour } for larger ones (postprocessing.py:17-21); NULL and a negative
 * max_depth_range_size select the reference's defaults { 5/6, 1-5/6, 2/3, 1-2/3 } and 4.
 */
int fic_qt_deblock(fic_handle* h, const double* img, const int32_t* leaves, int n, int H, int W, const double* weights,
                   int max_depth_range_size, double* out, unsigned flags, void* stream);

/* ---- colour planes ---------------------------------------------------------------------- */
/*
 * Replace utils.rgb_to_ycbcr (utils.py:36-63) and utils.ycbcr_to_rgb (utils.py:65-97), the plane fan-out either
 * side of RGBQuadtreeEncoder / RGBQuadtreeDecoder (quadtree/rgb.py:24-27, 35-48).  rgb uint8 [N,M,3] interleaved;
 * y uint8 [N,M]; subsample 0 = "4:4:4" (cb, cr [N,M]), 1 = "4:2:0" (cb, cr [N/2, (M+1)/2]).
 * "4:2:0" is defined where the reference is: forward for even N (an odd N makes utils.py:58 raise or leave the last
 * chroma row uninitialised), backward for N % 4 == 0 (otherwise utils.py:73-76 raises); FIC_EINVAL elsewhere.
 */
int fic_rgb_to_ycbcr(fic_handle* h, const uint8_t* rgb, int N, int M, int subsample, uint8_t* y, uint8_t* cb, uint8_t* cr,
                     unsigned flags, void* stream);
int fic_ycbcr_to_rgb(fic_handle* h, const uint8_t* y, const uint8_t* cb, const uint8_t* cr, int N, int M, int subsample,
                     uint8_t* rgb, unsigned flags, void* stream);

/* ---- container files (host only, no device work) ------------------------------------------ */
/*
 * Replaces the node loops of QuadtreeDeserializer._read_data (quadtree/serialization.py:182-223) and
 * HVDeserializer._read_data (hvv2/serialization.py:186-224): reads the node section of a `.fqt` / HV stream -- per
 * root a breadth-first list of `1` + row, col, orientation (1 bit), rotation (quadtree 2 bits; HV 1 bit, doubled),
 * scale, offset for a leaf, `0` (HV: + vertical flag + 16-bit split index) for an inner node -- and returns it as
 * the PREORDER stream fic_qt_encode / fic_hv_encode produce (nodes[k][8]; quadtree: size exponent in column 7).
 * data/nbytes: the file from any word boundary on; bit_pos: first bit of the node section, bits counted MSB first
 * (bitbuffer.py:39-56).  Fields are coded per CodingDimension (entropy/processor.py:11-16: row, col, scale, offset):
 * huff_root[d] < 0 = raw_width[d] plain bits, else the index of the root of a Huffman tree in huff_table[m][3] =
 * { left, right, value }, left < 0 marking a leaf (entropy/huffman.py:100-124).  root_exps: the quadtree roots' size
 * exponents in _de_partition order (HV: n_roots = 1, root_exps unused).  *end_bit = first bit after the section.
 * huff_rows = rows of huff_table.  The stream is untrusted input: a raw width outside 0..32 (the bit buffer's
 * MAX_BIT_WIDTH), a root or child index outside the table, or a child that does not follow its parent is FIC_EINVAL.
 * Returns the node count, FIC_ECAP, or FIC_EINVAL when the data ends early (the reference's "buffer returned EOF").
 */
int fic_tree_read_nodes(int kind, const uint8_t* data, int64_t nbytes, int64_t bit_pos, const int32_t* root_exps,
                        int n_roots, const int32_t* raw_width, const int32_t* huff_root, const int32_t* huff_table,
                        int huff_rows, int32_t* nodes, int cap, int64_t* end_bit);

/* Replaces HuffmanDeserializer.deserialize (entropy/huffman.py:100-124): reads one breadth-first tree dump (`1` + value
 * in val_bits bits for a leaf, `0` for an inner node) into table[m][3] = { left, right, value } rows in stream order
 * (row 0 = root, left < 0 = leaf) -- the form fic_tree_read_nodes takes.  Returns the node count. */
int fic_huffman_read_tree(const uint8_t* data, int64_t nbytes, int64_t bit_pos, int val_bits, int32_t* table, int cap,
                          int64_t* end_bit);

/* BitBuffer.write (bitbuffer.py:18-37) for n fields at once: the low widths[k] bits of values[k], most significant bit
 * first, appended to `pending_bits` (< 32) pending bits; every completed big-endian 32-bit word goes to `out`.
 * Returns the bytes written (a multiple of 4); *rest / *rest_bits = the bits still pending afterwards. */
int64_t fic_pack_fields(const uint64_t* values, const int64_t* widths, int64_t n, uint32_t pending, int pending_bits,
                        uint8_t* out, int64_t out_cap, uint32_t* rest, int* rest_bits);

/* Rectangle rects[k] = { i, j, w, h } and depth of every node of an HV preorder stream over an H x W image: the split
 * walk of HVDecoder (hvv2/decoder.py:59-88; child 1 keeps split_idx + 1 columns / rows). */
int fic_hv_layout(const int32_t* nodes, int n, int H, int W, int32_t* rects, int32_t* depth);

#ifdef __cplusplus
}
#endif
#endif /* FIC_H_ */
