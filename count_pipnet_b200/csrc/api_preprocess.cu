// C ABI: the reference's evaluation-time input pipeline on the GPU.
//
// Replaces `transform_no_augment = Compose([Resize(size=(S, S)), ToTensor(), Normalize(mean, std)])` of util/data.py:381-385
// (the same three steps at :265-275, :296-307, :465-480, :522-531), which the reference applies image by image in DataLoader
// worker processes (util/data.py:596-617) -- at more than 15 k images/s per GPU that host path is the bottleneck.
//
// Bit-exact with the reference's dependencies:
//   * Pillow `Image.resize(size, BILINEAR)` (src/libImaging/Resample.c): triangle filter whose support widens with the
//     down-scaling factor; coefficients in double precision (precompute_coeffs: here with __d*_rn so that nothing is
//     contracted into an FMA), rounded to 22-bit fixed point (normalize_coeffs_8bpc); a horizontal pass rounded and
//     clipped to uint8, then a vertical pass rounded and clipped to uint8;
//   * torchvision ToTensor (true fp32 division by 255) and Normalize ((x - mean) / std in fp32): one of 3 x 256 values, so
//     a table per call.
// Layout: src [B, Hs, Ws, 3] uint8 (decoded RGB, channels last, what PIL / nvJPEG deliver) -> out [B, 3, Ho, Wo] fp32 NCHW
// (what net.forward takes). The coefficient tables live in a caller-provided workspace: the library allocates nothing.
#include "api_common.h"

using namespace cpn_api;

namespace {

constexpr int PP_BITS = 22;       // Resample.c PRECISION_BITS = 32 - 8 - 2
constexpr int PP_MAX_IN = 16384;  // source side
constexpr int PP_MAX_KSIZE = 131; // down-scaling factors up to 64

__host__ __device__ inline int pp_ksize(int in_size, int out_size) {
    double fs = static_cast<double>(in_size) / out_size;
    if (fs < 1.0) fs = 1.0;
    const int c = static_cast<int>(fs);
    return ((static_cast<double>(c) < fs) ? c + 1 : c) * 2 + 1;  // (int)ceil(support) * 2 + 1
}

// One thread per output index of one axis: bounds[i] = (first input index, count), coeffs[i][ksize] (unused entries 0).
__global__ void pp_coeffs_kernel(int in_size, int out_size, int ksize, int* __restrict__ bounds, int* __restrict__ coeffs) {
    const int xx = blockIdx.x * blockDim.x + threadIdx.x;
    if (xx >= out_size) return;
    const double scale = __ddiv_rn(static_cast<double>(in_size), static_cast<double>(out_size));
    const double filterscale = scale < 1.0 ? 1.0 : scale;
    const double support = filterscale;  // triangle filter: support 1.0 * filterscale
    const double ss = __ddiv_rn(1.0, filterscale);
    const double center = __dadd_rn(0.0, __dmul_rn(__dadd_rn(static_cast<double>(xx), 0.5), scale));
    int xmin = static_cast<int>(__dadd_rn(__dsub_rn(center, support), 0.5));
    if (xmin < 0) xmin = 0;
    int xmax = static_cast<int>(__dadd_rn(__dadd_rn(center, support), 0.5));
    if (xmax > in_size) xmax = in_size;
    xmax -= xmin;
    auto weight = [&](int x) {
        double a = __dmul_rn(__dadd_rn(__dsub_rn(static_cast<double>(x + xmin), center), 0.5), ss);
        if (a < 0.0) a = -a;
        return a < 1.0 ? __dsub_rn(1.0, a) : 0.0;
    };
    double ww = 0.0;
    for (int x = 0; x < xmax; ++x) ww = __dadd_rn(ww, weight(x));
    int* k = coeffs + static_cast<size_t>(xx) * ksize;
    for (int x = 0; x < ksize; ++x) {
        double w = x < xmax ? weight(x) : 0.0;
        if (x < xmax && ww != 0.0) w = __ddiv_rn(w, ww);
        const double v = __dmul_rn(w, static_cast<double>(1 << PP_BITS));
        k[x] = w < 0.0 ? static_cast<int>(__dadd_rn(-0.5, v)) : static_cast<int>(__dadd_rn(0.5, v));
    }
    bounds[2 * xx] = xmin;
    bounds[2 * xx + 1] = xmax;
}

__device__ __forceinline__ int pp_clip8(int acc) {
    const int v = acc >> PP_BITS;
    return v < 0 ? 0 : (v > 255 ? 255 : v);
}

// ToTensor + Normalize of each of the 3 x 256 possible values, once per call (true fp32 divisions: ~30 instructions each --
// far too many to repeat in every CTA of the resize kernel).
__global__ void pp_lut_kernel(float* __restrict__ lut, float m0, float m1, float m2, float s0, float s1, float s2) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 768) return;
    const int c = i >> 8, v = i & 255;
    const float m = c == 0 ? m0 : (c == 1 ? m1 : m2), s = c == 0 ? s0 : (c == 1 ? s1 : s2);
    lut[i] = __fdiv_rn(__fsub_rn(__fdiv_rn(static_cast<float>(v), 255.0f), m), s);
}

// grid = (ceil(Wo / 128), Ho, B), 128 threads: one output pixel (3 channels) per thread.
__global__ void __launch_bounds__(128)
pp_resize_normalize_kernel(const uint8_t* __restrict__ src, float* __restrict__ out, const int* __restrict__ xb,
                           const int* __restrict__ xk, const int* __restrict__ yb, const int* __restrict__ yk,
                           const float* __restrict__ lut_g, int Hs, int Ws, int Ho, int Wo, int kx, int ky) {
    __shared__ float lut[3][256];
    for (int i = threadIdx.x; i < 768; i += blockDim.x) (&lut[0][0])[i] = __ldg(lut_g + i);
    __syncthreads();
    const int ox = blockIdx.x * blockDim.x + threadIdx.x;
    const int oy = blockIdx.y, b = blockIdx.z;
    if (ox >= Wo) return;
    const int x0 = __ldg(xb + 2 * ox), nx = __ldg(xb + 2 * ox + 1);
    const int y0 = __ldg(yb + 2 * oy), ny = __ldg(yb + 2 * oy + 1);
    const int* cx = xk + static_cast<size_t>(ox) * kx;
    const int* cy = yk + static_cast<size_t>(oy) * ky;
    const uint8_t* img = src + static_cast<size_t>(b) * Hs * Ws * 3;
    constexpr int HALF = 1 << (PP_BITS - 1);
    int v0 = HALF, v1 = HALF, v2 = HALF;
    for (int y = 0; y < ny; ++y) {
        const uint8_t* row = img + (static_cast<size_t>(y0 + y) * Ws + x0) * 3;
        int h0 = HALF, h1 = HALF, h2 = HALF;
        for (int x = 0; x < nx; ++x) {
            const int k = __ldg(cx + x);
            h0 += static_cast<int>(__ldg(row + 3 * x)) * k;
            h1 += static_cast<int>(__ldg(row + 3 * x + 1)) * k;
            h2 += static_cast<int>(__ldg(row + 3 * x + 2)) * k;
        }
        const int k = __ldg(cy + y);  // the horizontal pass of Pillow ends in uint8: clip before the vertical pass
        v0 += pp_clip8(h0) * k;
        v1 += pp_clip8(h1) * k;
        v2 += pp_clip8(h2) * k;
    }
    const size_t plane = static_cast<size_t>(Ho) * Wo;
    float* o = out + static_cast<size_t>(b) * 3 * plane + static_cast<size_t>(oy) * Wo + ox;
    o[0] = lut[0][pp_clip8(v0)];
    o[plane] = lut[1][pp_clip8(v1)];
    o[2 * plane] = lut[2][pp_clip8(v2)];
}

}  // namespace

extern "C" {

long long cpn_preprocess_ws_ints(int Hs, int Ws, int Ho, int Wo) {
    if (Hs <= 0 || Ws <= 0 || Ho <= 0 || Wo <= 0) return 0;
    return 2LL * Wo + static_cast<long long>(Wo) * pp_ksize(Ws, Wo) + 2LL * Ho + static_cast<long long>(Ho) * pp_ksize(Hs, Ho) +
           768;  // + the 3 x 256 fp32 table of ToTensor + Normalize
}

int cpn_preprocess_u8(cpn_plan* plan, const unsigned char* src, float* out, int* ws, int B, int Hs, int Ws, int Ho, int Wo,
                      const float* mean3, const float* std3, void* stream) {
    if (src == nullptr || out == nullptr || ws == nullptr || mean3 == nullptr || std3 == nullptr)
        return fail("cpn_preprocess_u8: missing operand");
    if (B <= 0 || B > 65535 || Ho <= 0 || Ho > 65535 || Wo <= 0 || Hs <= 0 || Ws <= 0 || Hs > PP_MAX_IN || Ws > PP_MAX_IN)
        return fail("cpn_preprocess_u8: bad shape B=%d source %dx%d target %dx%d", B, Hs, Ws, Ho, Wo);
    const int kx = pp_ksize(Ws, Wo), ky = pp_ksize(Hs, Ho);
    if (kx > PP_MAX_KSIZE || ky > PP_MAX_KSIZE)
        return fail("cpn_preprocess_u8: down-scaling by more than 64 is not supported (%dx%d -> %dx%d)", Hs, Ws, Ho, Wo);
    int* xb = ws;
    int* xk = xb + 2 * Wo;
    int* yb = xk + static_cast<size_t>(Wo) * kx;
    int* yk = yb + 2 * Ho;
    float* lut = reinterpret_cast<float*>(yk + static_cast<size_t>(Ho) * ky);
    const float m0 = mean3[0], m1 = mean3[1], m2 = mean3[2], s0 = std3[0], s1 = std3[1], s2 = std3[2];  // host values, read now
    set_info("preprocess_u8", 0.0, 3.0 * B * Hs * Ws + 12.0 * B * Ho * Wo, "B=%d %dx%d -> %dx%d taps=%dx%d", B, Hs, Ws, Ho, Wo, ky, kx);
    const dim3 grid((Wo + 127) / 128, Ho, B);
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        pp_coeffs_kernel<<<(Wo + 127) / 128, 128, 0, st>>>(Ws, Wo, kx, xb, xk);
        pp_coeffs_kernel<<<(Ho + 127) / 128, 128, 0, st>>>(Hs, Ho, ky, yb, yk);
        pp_lut_kernel<<<3, 256, 0, st>>>(lut, m0, m1, m2, s0, s1, s2);
        pp_resize_normalize_kernel<<<grid, 128, 0, st>>>(src, out, xb, xk, yb, yk, lut, Hs, Ws, Ho, Wo, kx, ky);
        CPN_CUDA(cudaGetLastError());
        return 0;
    });
}

}  // extern "C"
