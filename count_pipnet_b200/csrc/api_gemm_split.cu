// C ABI: the fp32-mode (three-pass bf16 hi/lo) GEMMs with the TMA epilogue (gemm_split_epi.cuh).
#include <cstdlib>

#include "api_common.h"
#include "gemm_split_epi.cuh"

using namespace cpn_api;

namespace {

struct SplitLaunch {
    CUtensorMap ta, ta_lo, tb, tb_lo, tout, tout_lo;
    cpn::Gemm2Params p;
    int grid;
};

template <int BN, int EPI, int CG>
int launch_split_inst(const SplitLaunch& g, cudaStream_t st) {
    using S = cpn::GemmSplitCfg<BN, EPI, CG>;
    auto kern = cpn::gemm_split_epi_kernel<BN, EPI, CG>;
    CPN_CUDA(ensure_smem(kern, S::TOTAL));
    CPN_CUDA(launch_pdl(kern, dim3(g.grid), dim3(S::THREADS), S::TOTAL, st, CG, g.ta, g.ta_lo, g.tb, g.tb_lo, g.tout,
                        g.tout_lo, g.p));
    CPN_CUDA(cudaGetLastError());
    return 0;
}

template <int EPI>
int launch_split_bn(const SplitLaunch& g, int bn, int cg, cudaStream_t st) {
    if (cg == 2) {
        switch (bn) {
            case 256: return launch_split_inst<256, EPI, 2>(g, st);
            case 128: return launch_split_inst<128, EPI, 2>(g, st);
        }
        if (!cpn::epi_out_bf16(EPI) && bn == 192) return launch_split_inst<192, cpn::epi_out_bf16(EPI) ? cpn::EPI_BIAS_F32 : EPI, 2>(g, st);
        return fail("no CTA-pair split GEMM for BN=%d", bn);
    }
    switch (bn) {
        case 256: return launch_split_inst<256, EPI, 1>(g, st);
        case 128: return launch_split_inst<128, EPI, 1>(g, st);
    }
    if (!cpn::epi_out_bf16(EPI)) {
        constexpr int E = cpn::epi_out_bf16(EPI) ? cpn::EPI_BIAS_F32 : EPI;
        if (bn == 192) return launch_split_inst<192, E, 1>(g, st);
        if (bn == 96) return launch_split_inst<96, E, 1>(g, st);
    }
    return fail("no split GEMM instantiation for BN=%d", bn);
}

// Tile width: the widest of 256 / 192 / 128 / 96 that divides N (fp32 outputs), 256 / 128 for the bf16-plane outputs.
int pick_split_bn(int N, int epi) {
    if (N % 256 == 0) return 256;
    if (!cpn::epi_out_bf16(epi) && N % 192 == 0) return 192;
    if (N % 128 == 0) return 128;
    if (!cpn::epi_out_bf16(epi) && N % 96 == 0) return 96;
    return 0;
}

}  // namespace

bool cpn_api::gemm_split_supported(int N, int K, int epi) { return K % 8 == 0 && pick_split_bn(N, epi) != 0; }

int cpn_api::gemm_split(cpn_plan* plan, int epi, const void* A, const void* A_lo, const void* Bw, const void* Bw_lo,
                        int M, int N, int K, const float* bias, const float* scale, const float* stats, int nparts,
                        const float* colsum, float eps, void* out, void* out_lo, void* stream) {
    const int bn = pick_split_bn(N, epi);
    if (bn == 0) return fail("split GEMM: no tile width divides N=%d", N);
    if (K % 8 != 0) return fail("split GEMM: K must be a multiple of 8, K=%d", K);
    SplitLaunch g;
    memset(&g, 0, sizeof(g));
    static const int allow = [] { const char* e = getenv("CPN_GEMM_2SM"); return e == nullptr ? 1 : atoi(e); }();
    const int pairs = num_sms() / 2;
    const long long pair_tiles = static_cast<long long>((M + 2 * cpn::GEMM_BM - 1) / (2 * cpn::GEMM_BM)) * (N / bn);
    const int cg = (allow != 0 && bn != 96 && (pair_tiles >= pairs || allow == 2)) ? 2 : 1;
    if (make_tmap(&g.ta, A, M, K, cpn::GEMM_BM) != 0 || make_tmap(&g.ta_lo, A_lo, M, K, cpn::GEMM_BM) != 0) return -1;
    if (make_tmap(&g.tb, Bw, N, K, bn / cg) != 0 || make_tmap(&g.tb_lo, Bw_lo, N, K, bn / cg) != 0) return -1;
    if (cpn::epi_out_bf16(epi)) {
        const cuuint64_t dims[2] = {static_cast<cuuint64_t>(N), static_cast<cuuint64_t>(M)};
        const cuuint64_t strides[1] = {static_cast<cuuint64_t>(N) * 2};
        const cuuint32_t box[2] = {32, 32};
        if (make_tmap_tiled(&g.tout, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, out, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_64B) != 0 ||
            make_tmap_tiled(&g.tout_lo, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, out_lo, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_64B) != 0)
            return -1;
    } else {
        if (make_tmap_out(&g.tout, out, M, N, false) != 0) return -1;
        g.tout_lo = g.tout;
    }
    g.p.M = M; g.p.N = N; g.p.K = K;
    g.p.num_m_blocks = (M + cg * cpn::GEMM_BM - 1) / (cg * cpn::GEMM_BM);
    g.p.num_n_blocks = N / bn;
    g.p.num_k_blocks = (K + cpn::GEMM_BK - 1) / cpn::GEMM_BK;
    g.p.bias = bias; g.p.scale = scale;
    g.p.stats = stats; g.p.colsum = colsum; g.p.nparts = nparts; g.p.inv_k = 1.0f / static_cast<float>(K); g.p.eps = eps;
    g.p.m_fastest = 0; g.p.dbg = 0;
    const int tiles = g.p.num_m_blocks * g.p.num_n_blocks;
    const int slots = num_sms() / cg;
    g.grid = cg * (tiles < slots ? tiles : slots);
    const double mn = static_cast<double>(M) * N;
    const double obytes = cpn::epi_out_bf16(epi) ? 4.0 : (epi == cpn::EPI_SCALE_RES_F32 ? 8.0 : 4.0);
    set_info("gemm_bf16x3", 6.0 * mn * K, 4.0 * (static_cast<double>(M) * K + static_cast<double>(N) * K) + obytes * mn,
             "M=%d N=%d K=%d BN=%d cg=%d epi=%s tma", M, N, K, bn, cg,
             epi == cpn::EPI_LN_GELU_BF16 ? "ln_gelu_split" : (epi == cpn::EPI_SCALE_RES_F32 ? "scale_res_f32" : "bias_f32"));
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        switch (epi) {
            case cpn::EPI_LN_GELU_BF16: return launch_split_bn<cpn::EPI_LN_GELU_BF16>(g, bn, cg, st);
            case cpn::EPI_SCALE_RES_F32: return launch_split_bn<cpn::EPI_SCALE_RES_F32>(g, bn, cg, st);
            default: return launch_split_bn<cpn::EPI_BIAS_F32>(g, bn, cg, st);
        }
    });
}

extern "C" {

int cpn_gemm_ln_gelu_split(cpn_plan* plan, const void* A_hi, const void* A_lo, const void* Bw_hi, const void* Bw_lo, int M,
                           int N, int K, const float* stats, int nparts, const float* colsum, const float* bias, float eps,
                           void* out_hi, void* out_lo, void* stream) {
    if (M <= 0 || N <= 0 || K <= 0) return fail("cpn_gemm_ln_gelu_split: bad shape M=%d N=%d K=%d", M, N, K);
    if (nparts <= 0 || nparts > cpn::LN_MAX_PARTS) return fail("cpn_gemm_ln_gelu_split: nparts=%d out of range", nparts);
    if (A_hi == nullptr || A_lo == nullptr || Bw_hi == nullptr || Bw_lo == nullptr || stats == nullptr ||
        colsum == nullptr || bias == nullptr || out_hi == nullptr || out_lo == nullptr)
        return fail("cpn_gemm_ln_gelu_split: missing operand");
    if (!gemm_split_supported(N, K, cpn::EPI_LN_GELU_BF16))
        return fail("cpn_gemm_ln_gelu_split: N=%d must be a multiple of 128 and K=%d of 8", N, K);
    return gemm_split(plan, cpn::EPI_LN_GELU_BF16, A_hi, A_lo, Bw_hi, Bw_lo, M, N, K, bias, nullptr, stats, nparts, colsum,
                      eps, out_hi, out_lo, stream);
}

}  // extern "C"
