// C ABI: depthwise 7x7 conv of the CNBlocks (include/cpn_b200.h, cpn_dwconv7_stats / cpn_dwconv7_stats_split).
#include <cstdlib>
#include <cstring>

#include "api_common.h"
#include "dwconv_v2.cuh"
#include "dwconv_v3.cuh"

using namespace cpn_api;

namespace {

// Job geometry of dwconv7_v2_kernel (see dwconv_v2.cuh): a job = (32-channel group, image group, column block, row
// segment); a CTA's consumer warps hold two units = (image, 4-column group) each.
int dw2_geometry(cpn::Dw2Params& p, int B, int H, int W, int C) {
    constexpr int TW = cpn::DW2_TW, MAXU = 2 * cpn::DW2_MAX_WARPS;
    const int ng = (W + TW - 1) / TW;
    if (ng <= MAXU) {
        p.ncb = 1;
        p.ngb = ng;
        // several images per job when one image's columns do not fill the warps; prefer an even number of units
        int best = 1, best_score = -1;
        for (int nimg = 1; nimg <= 8; nimg *= 2) {
            const int units = nimg * ng;
            if (units > MAXU || (nimg > 1 && nimg > B)) break;
            const int score = units * 2 - (units & 1) * 3;  // idle half warp costs more than it brings
            if (score > best_score) { best_score = score; best = nimg; }
        }
        p.nimg = best;
    } else {
        p.ncb = (ng + MAXU - 1) / MAXU;
        p.ngb = (ng + p.ncb - 1) / p.ncb;
        p.ngb += p.ngb & 1;
        if (p.ngb > MAXU) p.ngb = MAXU;
        p.ncb = (ng + p.ngb - 1) / p.ngb;
        p.nimg = 1;
    }
    p.nip = (B + p.nimg - 1) / p.nimg;
    p.units = p.nimg * p.ngb;
    p.nwc = (p.units + 1) / 2;
    const int ncols = p.ngb * TW + 6;
    if (ncols > 256) return fail("cpn_dwconv7_stats: column block of %d columns exceeds a TMA box", ncols);
    p.box_tx = ncols * 128;
    p.box_stride = (p.box_tx + 1023) / 1024 * 1024;
    p.stage_bytes = p.nimg * p.box_stride;
    // row segments: enough jobs that the contiguous per-CTA ranges differ by at most ~1/4 (jobs >= 3.5 x CTAs), but at
    // least 7 output rows per segment (each segment re-reads 6 halo rows from L2)
    const long long base_jobs = static_cast<long long>(C / 32) * p.nip * p.ncb;
    const long long slots = 2LL * num_sms();
    int nseg = 1;
    while (base_jobs * nseg < (7 * slots) / 2 && (H + nseg) / (nseg + 1) >= 7) ++nseg;
    if (const char* e = getenv("CPN_DW2_NSEG")) nseg = atoi(e) > 0 ? atoi(e) : nseg;  // tuning knob (tools/run_op.py)
    p.rows_per_seg = (H + nseg - 1) / nseg;
    p.nseg = (H + p.rows_per_seg - 1) / p.rows_per_seg;
    p.njobs = static_cast<int>(base_jobs * p.nseg);
    return 0;
}

int dwconv7_v2(cpn_plan* plan, const float* x, const float* w49, const float* bias, void* out_hi, void* out_lo,
               float* stats, int B, int H, int W, int C, void* stream) {
    cpn::Dw2Params p;
    memset(&p, 0, sizeof(p));
    p.w49 = w49; p.bias = bias;
    p.out_hi = static_cast<__nv_bfloat16*>(out_hi);
    p.out_lo = static_cast<__nv_bfloat16*>(out_lo);
    p.stats = stats;
    p.B = B; p.H = H; p.W = W; p.C = C;
    if (dw2_geometry(p, B, H, W, C) != 0) return -1;
    CUtensorMap tm;
    const cuuint64_t dims[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H),
                                static_cast<cuuint64_t>(B)};
    const cuuint64_t strides[3] = {static_cast<cuuint64_t>(C) * 4, static_cast<cuuint64_t>(W) * C * 4,
                                   static_cast<cuuint64_t>(H) * W * C * 4};
    const cuuint32_t box[4] = {32, static_cast<cuuint32_t>(p.ngb * cpn::DW2_TW + 6), 1, 1};
    if (make_tmap_tiled(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, x, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B) != 0)
        return -1;
    const size_t smem = static_cast<size_t>(cpn::DW2_NSTAGE) * p.stage_bytes + cpn::DW2_WBYTES +
                        2 * cpn::DW2_NSTAGE * 8 + 1024;
    const long long slots = 2LL * num_sms();
    const int grid = static_cast<int>(p.njobs < slots ? p.njobs : slots);
    const int threads = 32 * (p.nwc + 1);
    const bool split = out_lo != nullptr;
    const double tok = static_cast<double>(B) * H * W;
    set_info(split ? "dwconv7_split" : "dwconv7_stats", 2.0 * 49 * tok * C,
             tok * C * (4 + 2 + (split ? 2 : 0)) + tok * (C / 32) * 8,
             "B=%d H=%d W=%d C=%d v2 nimg=%d ngb=%d ncb=%d warps=%d rows/seg=%d jobs=%d grid=%d", B, H, W, C, p.nimg, p.ngb,
             p.ncb, p.nwc, p.rows_per_seg, p.njobs, grid);
    // The seven-fold unrolled row loop (slots renamed instead of moved) is the faster one (0.094 vs 0.106 ms at 64 x 27 x 27
    // x 384, profiles/r02_dwconv.md); CPN_DW2_ROLL=1 selects the compact loop (A/B only)
    static const bool unrolled = [] { const char* e = getenv("CPN_DW2_ROLL"); return e == nullptr || atoi(e) == 0; }();
    return submit(plan, stream, [=](cudaStream_t st) -> int {
#define CPN_DW2_LAUNCH(SP, RL)                                                                                        \
    do {                                                                                                              \
        CPN_CUDA(ensure_smem(cpn::dwconv7_v2_kernel<SP, RL>, smem));                                                  \
        CPN_CUDA(launch_pdl(cpn::dwconv7_v2_kernel<SP, RL>, dim3(grid), dim3(threads), smem, st, 1, tm, p));          \
    } while (0)
        if (split) {
            if (unrolled) CPN_DW2_LAUNCH(true, false);
            else CPN_DW2_LAUNCH(true, true);
        } else {
            if (unrolled) CPN_DW2_LAUNCH(false, false);
            else CPN_DW2_LAUNCH(false, true);
        }
#undef CPN_DW2_LAUNCH
        CPN_CUDA(cudaGetLastError());
        return 0;
    });
}

// Tile geometry of dwconv7_v3_kernel (dwconv_v3.cuh): a tile = 2 row blocks (14 output rows) x NS <= 7 column strips (<= 28
// columns), one warp per (strip, row block). Returns false when the map does not suit the kernel (too small to give a
// CTA 10 warps, or mostly padding): the caller then uses one of the row-marching kernels.
bool dw3_geometry(cpn::Dw3Params& p, size_t& smem, double& useful, int B, int H, int W, int C) {
    constexpr int RB = cpn::DW3_RB, TW = cpn::DW3_TW, NB = cpn::DW3_NB;
    const int strips = (W + TW - 1) / TW;
    p.nct = (strips + cpn::DW3_MAX_NS - 1) / cpn::DW3_MAX_NS;
    p.NS = (strips + p.nct - 1) / p.nct;
    if (p.NS * NB < 10) return false;
    p.wbox = TW * p.NS + 6;
    p.rows_in = NB * RB + 6;
    p.nrt = (H + NB * RB - 1) / (NB * RB);
    p.nitems = (C / 32) * B * p.nrt * p.nct;
    p.stage_bytes = p.rows_in * p.wbox * 32 * 4;  // a multiple of 128
    smem = 2 * static_cast<size_t>(p.stage_bytes) + static_cast<size_t>(cpn::DW3_WARPS) * cpn::DW3_SCRATCH * 4 + 4 * 8;
    if (smem > 227 * 1024) return false;
    useful = static_cast<double>(H) * W / (static_cast<double>(p.nrt) * NB * RB * p.nct * p.NS * TW);  // computed outputs kept
    return useful >= 0.6;
}

int dwconv7_v3(cpn_plan* plan, const cpn::Dw3Params& geo, size_t smem, const float* x, const float* w49, const float* bias,
               void* out_hi, void* out_lo, float* stats, int B, int H, int W, int C, void* stream) {
    cpn::Dw3Params p = geo;
    p.w49 = w49; p.bias = bias;
    p.out_hi = static_cast<__nv_bfloat16*>(out_hi);
    p.out_lo = static_cast<__nv_bfloat16*>(out_lo);
    p.stats = stats;
    p.B = B; p.H = H; p.W = W; p.C = C;
    CUtensorMap tm;
    const cuuint64_t dims[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H),
                                static_cast<cuuint64_t>(B)};
    const cuuint64_t strides[3] = {static_cast<cuuint64_t>(C) * 4, static_cast<cuuint64_t>(W) * C * 4,
                                   static_cast<cuuint64_t>(H) * W * C * 4};
    const cuuint32_t box[4] = {32, static_cast<cuuint32_t>(p.wbox), static_cast<cuuint32_t>(p.rows_in), 1};
    if (make_tmap_tiled(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, x, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_NONE) != 0)
        return -1;
    const int sms = num_sms();
    const int grid = p.nitems < sms ? p.nitems : sms;
    const int threads = 32 * cpn::DW3_WARPS;
    const bool split = out_lo != nullptr;
    const double tok = static_cast<double>(B) * H * W;
    set_info(split ? "dwconv7_split" : "dwconv7_stats", 2.0 * 49 * tok * C,
             tok * C * (4 + 2 + (split ? 2 : 0)) + tok * (C / 32) * 8,
             "B=%d H=%d W=%d C=%d v3 NS=%d items=%d (%dx%d per image) threads=%d grid=%d smem=%zu", B, H, W, C, p.NS,
             p.nitems, p.nrt, p.nct, threads, grid, smem);
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        if (split) {
            CPN_CUDA(ensure_smem(cpn::dwconv7_v3_kernel<true>, smem));
            CPN_CUDA(launch_pdl(cpn::dwconv7_v3_kernel<true>, dim3(grid), dim3(threads), smem, st, 1, tm, p));
        } else {
            CPN_CUDA(ensure_smem(cpn::dwconv7_v3_kernel<false>, smem));
            CPN_CUDA(launch_pdl(cpn::dwconv7_v3_kernel<false>, dim3(grid), dim3(threads), smem, st, 1, tm, p));
        }
        CPN_CUDA(cudaGetLastError());
        return 0;
    });
}

// CPN_DW_IMPL = v1 | v2 | v3 forces a kernel (A/B runs); unset = the default choice below.
const char* dw_impl() {
    static const char* impl = [] { const char* e = getenv("CPN_DW_IMPL"); return e != nullptr ? e : ""; }();
    return impl;
}

// Which kernel: dwconv7_v3_kernel where its tiles fit the map (dwconv7_any below); else the taps-in-registers kernel
// (dwconv_stats.cuh) for the four ConvNeXt-tiny widths, dwconv7_v2_kernel for every other width (any multiple of 32) or
// when CPN_DW_IMPL=v2 asks for it (A/B runs).
bool use_v2(int C) {
    return strcmp(dw_impl(), "v2") == 0 || !(C == 96 || C == 192 || C == 384 || C == 768);
}

int dwconv7_any(cpn_plan* plan, const float* x, const float* w49, const float* bias, void* out_hi, void* out_lo,
                float* stats, int B, int H, int W, int C, void* stream) {
    const char* impl = dw_impl();
    const bool aligned32 = ((reinterpret_cast<uintptr_t>(out_hi) | reinterpret_cast<uintptr_t>(out_lo)) & 31) == 0;  // 256-bit stores
    if (aligned32 && strcmp(impl, "v1") != 0 && strcmp(impl, "v2") != 0) {
        cpn::Dw3Params geo;
        memset(&geo, 0, sizeof(geo));
        size_t smem = 0;
        double useful = 0.0;
        // Measured on B200 against the row-marching kernel (profiles/r02_dwconv_v3.md): 0.79-0.83 of its time on maps whose
        // tiles are full (56, 28, 112, 55 wide), 0.90-0.92 at 27 / 26 / 54, slower when a tile has 12 units (48, 24, 23 wide)
        if (dw3_geometry(geo, smem, useful, B, H, W, C) &&
            (strcmp(impl, "v3") == 0 || (geo.NS == cpn::DW3_MAX_NS && useful >= 0.85)))
            return dwconv7_v3(plan, geo, smem, x, w49, bias, out_hi, out_lo, stats, B, H, W, C, stream);
    }
    if (use_v2(C)) return dwconv7_v2(plan, x, w49, bias, out_hi, out_lo, stats, B, H, W, C, stream);
    return dwconv7_stats_v1(plan, x, w49, bias, out_hi, out_lo, stats, B, H, W, C, stream);
}

int check_dw(const char* who, const void* x, int B, int H, int W, int C) {
    if (C <= 0 || C % 32 != 0) return fail("%s: C=%d must be a multiple of 32", who, C);
    if (B <= 0 || H <= 0 || W <= 0) return fail("%s: bad shape B=%d H=%d W=%d", who, B, H, W);
    if ((reinterpret_cast<uintptr_t>(x) & 15) != 0) return fail("%s: x not 16-byte aligned", who);
    return 0;
}

}  // namespace

extern "C" {

int cpn_dwconv7_stats(cpn_plan* plan, const float* x, const float* w49, const float* bias, void* out_bf16,
                      float* stats, int B, int H, int W, int C, void* stream) {
    if (check_dw("cpn_dwconv7_stats", x, B, H, W, C) != 0) return -1;
    return dwconv7_any(plan, x, w49, bias, out_bf16, nullptr, stats, B, H, W, C, stream);
}

int cpn_dwconv7_stats_split(cpn_plan* plan, const float* x, const float* w49, const float* bias, void* out_hi,
                            void* out_lo, float* stats, int B, int H, int W, int C, void* stream) {
    if (check_dw("cpn_dwconv7_stats_split", x, B, H, W, C) != 0) return -1;
    if (out_lo == nullptr) return fail("cpn_dwconv7_stats_split: the lo plane is required");
    return dwconv7_any(plan, x, w49, bias, out_hi, out_lo, stats, B, H, W, C, stream);
}

}  // extern "C"
