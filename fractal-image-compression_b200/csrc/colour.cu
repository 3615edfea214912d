// colour.cu -- the plane fan-out either side of the colour quadtree codec, on the device:
//   utils.rgb_to_ycbcr (utils.py:36-63) and utils.ycbcr_to_rgb (utils.py:65-97), called by RGBQuadtreeEncoder.encode
//   (quadtree/rgb.py:24-27) and RGBQuadtreeDecoder.decode (quadtree/rgb.py:35-48).
// Byte work bound by HBM (3 B in, 3 B out per pixel at 4:4:4; 3 -> 1.5 at 4:2:0); one thread per pixel (or per 2x2 chroma
// cell), consecutive threads on consecutive pixels.  The arithmetic is the reference's float64 expression evaluated left
// to right, every operation rounded on its own (the file is built with -fmad=false and spells the operations out).
#include "fic_common.cuh"
#include "scorer.cuh"

namespace {

struct Ycc { uint8_t y, cb, cr; };

__device__ __forceinline__ uint8_t clip_u8(double v)
{
    v = v < 0.0 ? 0.0 : (v > 255.0 ? 255.0 : v);   // .clip(0., 255.).astype(np.uint8)
    return (uint8_t)__double2int_rz(v);
}

// utils.py:39-44: x = rgb / 256.;  Y = 16 + 65.738 R + 129.057 G + 25.064 B;  Cb = 128 - 37.945 R - 74.494 G + 112.439 B;
// Cr = 128 + 112.439 R - 94.154 G - 18.285 B
__device__ __forceinline__ Ycc forward_pixel(const uint8_t* __restrict__ px)
{
    const double R = DM((double)px[0], 0.00390625), G = DM((double)px[1], 0.00390625), B = DM((double)px[2], 0.00390625);
    Ycc o;
    o.y = clip_u8(DA(DA(DA(16.0, DM(65.738, R)), DM(129.057, G)), DM(25.064, B)));
    o.cb = clip_u8(DA(DS(DS(128.0, DM(37.945, R)), DM(74.494, G)), DM(112.439, B)));
    o.cr = clip_u8(DS(DS(DA(128.0, DM(112.439, R)), DM(94.154, G)), DM(18.285, B)));
    return o;
}

__global__ void rgb_to_ycbcr444_kernel(const uint8_t* __restrict__ rgb, size_t npx, uint8_t* __restrict__ y,
                                       uint8_t* __restrict__ cb, uint8_t* __restrict__ cr)
{
    const size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npx) return;
    const Ycc o = forward_pixel(rgb + 3 * p);
    y[p] = o.y;
    cb[p] = o.cb;
    cr[p] = o.cr;
}

// one thread per chroma sample: a 2x2 cell (average_subsample, utils.py:13-25: integer sum / 4, truncated by the uint8
// store), or the 2x1 cell of an odd last column (utils.py:53-56: (a + b) / 2 truncated)
__global__ void rgb_to_ycbcr420_kernel(const uint8_t* __restrict__ rgb, int N, int M, uint8_t* __restrict__ y,
                                       uint8_t* __restrict__ cb, uint8_t* __restrict__ cr)
{
    const int Ms = (M + 1) / 2;
    const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (size_t)(N / 2) * Ms) return;
    const int ci = (int)(e / Ms), cj = (int)(e - (size_t)ci * Ms);
    const int wide = (2 * cj + 1 < M) ? 2 : 1;
    int sb = 0, sr = 0;
#pragma unroll
    for (int u = 0; u < 2; ++u)
        for (int v = 0; v < wide; ++v) {
            const size_t p = (size_t)(2 * ci + u) * M + 2 * cj + v;
            const Ycc o = forward_pixel(rgb + 3 * p);
            y[p] = o.y;
            sb += o.cb;
            sr += o.cr;
        }
    cb[e] = (uint8_t)(wide == 2 ? sb >> 2 : sb >> 1);
    cr[e] = (uint8_t)(wide == 2 ? sr >> 2 : sr >> 1);
}

// utils.py:91-96: Y, Cb, Cr = planes / 256.;  R = 298.082 Y + 408.583 Cr - 222.921;
// G = 298.082 Y - 100.291 Cb - 208.120 Cr + 135.576;  B = 298.082 Y + 516.412 Cb - 276.836.
// SUB: the chroma sample of the pixel's 2x2 cell (utils.py:73-80; an odd last column reads the last chroma column).
template <bool SUB>
__global__ void ycbcr_to_rgb_kernel(const uint8_t* __restrict__ yp, const uint8_t* __restrict__ cbp,
                                    const uint8_t* __restrict__ crp, int N, int M, uint8_t* __restrict__ rgb)
{
    const size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= (size_t)N * M) return;
    size_t q = p;
    if (SUB) {
        const int i = (int)(p / M), j = (int)(p - (size_t)i * M);
        q = (size_t)(i / 2) * ((M + 1) / 2) + j / 2;
    }
    const double Y = DM((double)yp[p], 0.00390625), Cb = DM((double)cbp[q], 0.00390625), Cr = DM((double)crp[q], 0.00390625);
    const double ly = DM(298.082, Y);
    rgb[3 * p] = clip_u8(DS(DA(ly, DM(408.583, Cr)), 222.921));
    rgb[3 * p + 1] = clip_u8(DA(DS(DS(ly, DM(100.291, Cb)), DM(208.120, Cr)), 135.576));
    rgb[3 * p + 2] = clip_u8(DS(DA(ly, DM(516.412, Cb)), 276.836));
}

}  // namespace

int rgb_to_ycbcr_impl(fic_handle* h, const uint8_t* d_rgb, int N, int M, int subsample, uint8_t* d_y, uint8_t* d_cb,
                      uint8_t* d_cr, cudaStream_t st)
{
    if (subsample) {
        const size_t cells = (size_t)(N / 2) * ((M + 1) / 2);
        rgb_to_ycbcr420_kernel<<<(unsigned)((cells + 255) / 256), 256, 0, st>>>(d_rgb, N, M, d_y, d_cb, d_cr);
    } else {
        const size_t npx = (size_t)N * M;
        rgb_to_ycbcr444_kernel<<<(unsigned)((npx + 255) / 256), 256, 0, st>>>(d_rgb, npx, d_y, d_cb, d_cr);
    }
    FIC_CHECK_LAUNCH(h, "rgb_to_ycbcr_kernel");
    return FIC_OK;
}

int ycbcr_to_rgb_impl(fic_handle* h, const uint8_t* d_y, const uint8_t* d_cb, const uint8_t* d_cr, int N, int M,
                      int subsample, uint8_t* d_rgb, cudaStream_t st)
{
    const size_t npx = (size_t)N * M;
    const unsigned grid = (unsigned)((npx + 255) / 256);
    if (subsample)
        ycbcr_to_rgb_kernel<true><<<grid, 256, 0, st>>>(d_y, d_cb, d_cr, N, M, d_rgb);
    else
        ycbcr_to_rgb_kernel<false><<<grid, 256, 0, st>>>(d_y, d_cb, d_cr, N, M, d_rgb);
    FIC_CHECK_LAUNCH(h, "ycbcr_to_rgb_kernel");
    return FIC_OK;
}
