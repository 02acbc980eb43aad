// C ABI: ResNet convolutions with folded BatchNorm as tcgen05 GEMMs with implicit 3x3 taps (conv_gemm.cuh).
#include <cstdlib>

#include "api_common.h"
#include "conv_gemm.cuh"

using namespace cpn_api;

namespace {

struct ConvLaunch {
    CUtensorMap ta, tb, tres, tout, tout_bf16;
    cpn::ConvGemmParams p;
    int grid;
};

template <int BN, int EPI, int CG>
int launch_conv_inst(const ConvLaunch& g, cudaStream_t st) {
    using S = cpn::ConvGemmCfg<BN, EPI, CG>;
    auto kern = cpn::conv_gemm_kernel<BN, EPI, CG>;
    CPN_CUDA(ensure_smem(kern, S::TOTAL));
    CPN_CUDA(launch_pdl(kern, dim3(g.grid), dim3(S::THREADS), S::TOTAL, st, CG, g.ta, g.tb, g.tres, g.tout, g.tout_bf16, g.p));
    CPN_CUDA(cudaGetLastError());
    return 0;
}

template <int EPI>
int launch_conv_bn(const ConvLaunch& g, int bn, int cg, cudaStream_t st) {
    if (cg == 2) {
        if (bn == 256) return launch_conv_inst<256, EPI, 2>(g, st);
        if (bn == 128) return launch_conv_inst<128, EPI, 2>(g, st);
    } else {
        if (bn == 256) return launch_conv_inst<256, EPI, 1>(g, st);
        if (bn == 128) return launch_conv_inst<128, EPI, 1>(g, st);
        if (bn == 64) return launch_conv_inst<64, EPI, 1>(g, st);
    }
    return fail("cpn_conv_bn_act: no instantiation for BN=%d cg=%d", bn, cg);
}

const char* epi_name(int epi, int relu) {
    switch (epi) {
        case cpn::CEPI_RELU_BF16: return relu ? "relu_bf16" : "bias_bf16";
        case cpn::CEPI_RES_RELU: return "res_relu";
        case cpn::CEPI_BIAS_F32: return "bias_f32";
        case cpn::CEPI_RESB_RELU: return "resb_relu";
        default: return "resb_relu_f32";
    }
}

int conv_bn_act(cpn_plan* plan, const void* x_bf16, const void* w_bf16, const float* bias, const float* res_f32,
                const void* res_bf16, float* out_f32, void* out_bf16, int B, int H, int W, int Cin, int Cout, int ksize,
                int relu, void* stream);

}  // namespace

extern "C" {

int cpn_conv3x3_tile_rows(int H, int W) {
    // image rows per M tile of the implicit 3x3 conv: the most rows with Ht * W <= 128, H % Ht == 0, (Ht * W) % 16 == 0
    for (int ht = 128 / (W > 0 ? W : 1); ht >= 1; --ht)
        if (H % ht == 0 && (ht * W) % 16 == 0) return ht;
    return 0;
}

int cpn_conv_bn_act(cpn_plan* plan, const void* x_bf16, const void* w_bf16, const float* bias, const float* res_f32,
                    float* out_f32, void* out_bf16, int B, int H, int W, int Cin, int Cout, int ksize, int relu,
                    void* stream) {
    return conv_bn_act(plan, x_bf16, w_bf16, bias, res_f32, nullptr, out_f32, out_bf16, B, H, W, Cin, Cout, ksize, relu, stream);
}

int cpn_conv_bn_act_resb(cpn_plan* plan, const void* x_bf16, const void* w_bf16, const float* bias, const void* res_bf16,
                         float* out_f32, void* out_bf16, int B, int H, int W, int Cin, int Cout, int ksize, void* stream) {
    if (res_bf16 == nullptr) return fail("cpn_conv_bn_act_resb: the bf16 residual is required");
    return conv_bn_act(plan, x_bf16, w_bf16, bias, nullptr, res_bf16, out_f32, out_bf16, B, H, W, Cin, Cout, ksize, 1, stream);
}

}  // extern "C"

namespace {

int conv_bn_act(cpn_plan* plan, const void* x_bf16, const void* w_bf16, const float* bias, const float* res_f32,
                const void* res_bf16, float* out_f32, void* out_bf16, int B, int H, int W, int Cin, int Cout, int ksize,
                int relu, void* stream) {
    if (x_bf16 == nullptr || w_bf16 == nullptr || bias == nullptr) return fail("cpn_conv_bn_act: missing operand");
    if (B <= 0 || H <= 0 || W <= 0) return fail("cpn_conv_bn_act: bad shape B=%d H=%d W=%d", B, H, W);
    if (ksize != 1 && ksize != 3) return fail("cpn_conv_bn_act: kernel size must be 1 or 3 (stride 1)");
    int epi;
    if (res_bf16 != nullptr) {
        if (out_bf16 != nullptr && out_f32 == nullptr) epi = cpn::CEPI_RESB_RELU;
        else if (out_bf16 == nullptr && out_f32 != nullptr) epi = cpn::CEPI_RESB_RELU_F32;
        else return fail("cpn_conv_bn_act_resb: exactly one of out_f32 / out_bf16");
    }
    else if (res_f32 != nullptr && out_f32 != nullptr && out_bf16 != nullptr && relu) epi = cpn::CEPI_RES_RELU;
    else if (res_f32 == nullptr && out_f32 == nullptr && out_bf16 != nullptr) epi = cpn::CEPI_RELU_BF16;  // relu = 0: bias only
    else if (res_f32 == nullptr && out_f32 != nullptr && out_bf16 == nullptr && !relu) epi = cpn::CEPI_BIAS_F32;
    else return fail("cpn_conv_bn_act: unsupported output combination");
    const int bn_cands[3] = {256, 128, 64};
    const long long M = static_cast<long long>(B) * H * W;
    if (M >= (1LL << 31)) return fail("cpn_conv_bn_act: too many rows");
    ConvLaunch g;
    memset(&g, 0, sizeof(g));
    g.p.M = static_cast<int>(M); g.p.N = Cout; g.p.bias = bias; g.p.relu = relu;
    const int K = ksize * ksize * Cin;
    if (ksize == 3) {
        if (Cin % 64 != 0) return fail("cpn_conv_bn_act: the implicit 3x3 conv needs Cin %% 64 == 0 (Cin=%d)", Cin);
        const int ht = cpn_conv3x3_tile_rows(H, W);
        if (ht == 0) return fail("cpn_conv_bn_act: no tile shape for a %dx%d map (see cpn_conv3x3_tile_rows)", H, W);
        g.p.implicit = 1; g.p.ht = ht; g.p.tile_rows = ht * W; g.p.tiles_per_img = H / ht;
        g.p.cblocks = Cin / 64; g.p.ksize = 3; g.p.pad = 1;
        g.p.num_m_tiles = B * g.p.tiles_per_img;
        g.p.num_k_blocks = 9 * g.p.cblocks;
    } else {
        if (Cin % 8 != 0) return fail("cpn_conv_bn_act: Cin must be a multiple of 8 (Cin=%d)", Cin);
        g.p.implicit = 0; g.p.tile_rows = cpn::GEMM_BM;
        g.p.num_m_tiles = static_cast<int>((M + cpn::GEMM_BM - 1) / cpn::GEMM_BM);
        g.p.num_k_blocks = (Cin + cpn::GEMM_BK - 1) / cpn::GEMM_BK;
    }
    int bn = 0;
    for (int c : bn_cands) {
        if (Cout % c != 0) continue;
        if (bn == 0) bn = c;
        if (static_cast<long long>(g.p.num_m_tiles) * (Cout / c) >= num_sms()) { bn = c; break; }
        bn = c;  // no width fills the machine yet: keep narrowing
    }
    if (bn == 0) return fail("cpn_conv_bn_act: Cout=%d must be a multiple of 64", Cout);
    static const int allow = [] { const char* e = getenv("CPN_GEMM_2SM"); return e == nullptr ? 1 : atoi(e); }();
    const long long pair_tiles = static_cast<long long>((g.p.num_m_tiles + 1) / 2) * (Cout / bn);
    const int cg = (allow != 0 && bn >= 128 && (pair_tiles >= num_sms() / 2 || allow == 2)) ? 2 : 1;
    g.p.num_n_blocks = Cout / bn;
    // operand maps
    if (g.p.implicit) {
        const cuuint64_t dims[4] = {static_cast<cuuint64_t>(Cin), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H),
                                    static_cast<cuuint64_t>(B)};
        const cuuint64_t strides[3] = {static_cast<cuuint64_t>(Cin) * 2, static_cast<cuuint64_t>(W) * Cin * 2,
                                       static_cast<cuuint64_t>(H) * W * Cin * 2};
        const cuuint32_t box[4] = {64, static_cast<cuuint32_t>(W), static_cast<cuuint32_t>(g.p.ht), 1};
        if (make_tmap_tiled(&g.ta, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, x_bf16, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B) != 0)
            return -1;
    } else if (make_tmap(&g.ta, x_bf16, M, Cin, cpn::GEMM_BM) != 0) {
        return -1;
    }
    if (make_tmap(&g.tb, w_bf16, Cout, K, bn / cg) != 0) return -1;
    const cuuint64_t odims[2] = {static_cast<cuuint64_t>(Cout), static_cast<cuuint64_t>(M)};
    const cuuint32_t obox[2] = {32, 16};
    const cuuint64_t f32_stride[1] = {static_cast<cuuint64_t>(Cout) * 4}, bf16_stride[1] = {static_cast<cuuint64_t>(Cout) * 2};
    if (epi == cpn::CEPI_RELU_BF16 || epi == cpn::CEPI_RESB_RELU) {
        if (make_tmap_tiled(&g.tout, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, out_bf16, odims, bf16_stride, obox, CU_TENSOR_MAP_SWIZZLE_64B) != 0)
            return -1;
        g.tres = g.tout; g.tout_bf16 = g.tout;
        if (epi == cpn::CEPI_RESB_RELU &&
            make_tmap_tiled(&g.tres, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, res_bf16, odims, bf16_stride, obox, CU_TENSOR_MAP_SWIZZLE_64B) != 0)
            return -1;
    } else {
        if (make_tmap_tiled(&g.tout, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, out_f32, odims, f32_stride, obox, CU_TENSOR_MAP_SWIZZLE_128B) != 0)
            return -1;
        g.tres = g.tout; g.tout_bf16 = g.tout;
        if (epi == cpn::CEPI_RESB_RELU_F32 &&
            make_tmap_tiled(&g.tres, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, res_bf16, odims, bf16_stride, obox, CU_TENSOR_MAP_SWIZZLE_64B) != 0)
            return -1;
        if (epi == cpn::CEPI_RES_RELU) {
            if (make_tmap_tiled(&g.tres, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, res_f32, odims, f32_stride, obox, CU_TENSOR_MAP_SWIZZLE_128B) != 0 ||
                make_tmap_tiled(&g.tout_bf16, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, out_bf16, odims, bf16_stride, obox, CU_TENSOR_MAP_SWIZZLE_64B) != 0)
                return -1;
        }
    }
    const int groups = (g.p.num_m_tiles + cg - 1) / cg;
    const long long tiles = static_cast<long long>(groups) * g.p.num_n_blocks;
    const int slots = num_sms() / cg;
    g.grid = cg * static_cast<int>(tiles < slots ? tiles : slots);
    const double mn = static_cast<double>(M) * Cout;
    set_info(ksize == 3 ? "conv_3x3" : "conv_1x1", 2.0 * mn * K,
             2.0 * (static_cast<double>(M) * Cin + static_cast<double>(Cout) * K) +
                 mn * (epi == cpn::CEPI_RELU_BF16 ? 2.0 : epi == cpn::CEPI_RES_RELU ? 10.0 : epi == cpn::CEPI_RESB_RELU ? 4.0
                       : epi == cpn::CEPI_RESB_RELU_F32 ? 6.0 : 4.0),
             "B=%d H=%d W=%d Cin=%d Cout=%d k=%d BN=%d cg=%d rows/tile=%d epi=%s", B, H, W, Cin, Cout, ksize, bn, cg,
             g.p.tile_rows, epi_name(epi, relu));
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        switch (epi) {
            case cpn::CEPI_RELU_BF16: return launch_conv_bn<cpn::CEPI_RELU_BF16>(g, bn, cg, st);
            case cpn::CEPI_RES_RELU: return launch_conv_bn<cpn::CEPI_RES_RELU>(g, bn, cg, st);
            case cpn::CEPI_RESB_RELU: return launch_conv_bn<cpn::CEPI_RESB_RELU>(g, bn, cg, st);
            case cpn::CEPI_RESB_RELU_F32: return launch_conv_bn<cpn::CEPI_RESB_RELU_F32>(g, bn, cg, st);
            default: return launch_conv_bn<cpn::CEPI_BIAS_F32>(g, bn, cg, st);
        }
    });
}

}  // namespace
