// C ABI of libcpnb200.so (see include/cpn_b200.h). Host side only: argument checks, TMA descriptor encoding,
// kernel selection and launch, and the plan recorder. No torch types, no allocation of user-visible memory.
#include "api_common.h"

#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <functional>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "convnext_kernels.cuh"
#include "dwconv_stats.cuh"
#include "gemm_tcgen05.cuh"
#include "gemm_tma_epi.cuh"
#include "eval_kernels.cuh"
#include "head_kernels.cuh"
#include "layout_kernels.cuh"
#include "mlp_fused.cuh"

namespace cpn_api {

thread_local std::string g_err;

int fail(const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return -1;
}

// Algorithmic work of the op being submitted (for cpn_plan_run_timed's per-op report); set by each entry point.
struct OpInfo {
    const char* name = "op";
    double flops = 0.0;
    double bytes = 0.0;
    std::string detail;  // shape of this launch, for per-op reports
};
thread_local OpInfo g_info;
void set_info(const char* name, double flops, double bytes, const char* fmt, ...) {
    g_info.name = name;
    g_info.flops = flops;
    g_info.bytes = bytes;
    g_info.detail.clear();
    if (fmt != nullptr) {
        char buf[160];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, sizeof(buf), fmt, ap);
        va_end(ap);
        g_info.detail = buf;
    }
}

// ------------------------------------------------------------------------------------------------ TMA descriptors
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

// bf16 row-major [rows, K] matrix, box = 64 (K) x box_rows, 128-byte swizzle; out-of-bounds reads give zeros.
int make_tmap(CUtensorMap* m, const void* ptr, long long rows, int K, int box_rows) {
    EncodeTiledFn enc = get_encode_fn();
    if (enc == nullptr) return fail("cuTensorMapEncodeTiled not available from the driver");
    if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0) return fail("TMA operand not 16-byte aligned");
    if (K % 8 != 0) return fail("TMA operand row pitch must be a multiple of 16 bytes (K %% 8 == 0), K=%d", K);
    cuuint64_t dims[2] = {static_cast<cuuint64_t>(K), static_cast<cuuint64_t>(rows)};
    cuuint64_t strides[1] = {static_cast<cuuint64_t>(K) * 2};
    cuuint32_t box[2] = {static_cast<cuuint32_t>(cpn::GEMM_BK), static_cast<cuuint32_t>(box_rows)};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled failed with CUresult %d (rows=%lld K=%d box_rows=%d)",
                                       static_cast<int>(r), rows, K, box_rows);
    return 0;
}

int make_tmap_tiled(CUtensorMap* m, CUtensorMapDataType dtype, int rank, const void* ptr, const cuuint64_t* dims,
                    const cuuint64_t* strides_bytes, const cuuint32_t* box, CUtensorMapSwizzle swizzle) {
    EncodeTiledFn enc = get_encode_fn();
    if (enc == nullptr) return fail("cuTensorMapEncodeTiled not available from the driver");
    if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0) return fail("TMA operand not 16-byte aligned");
    cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    CUresult r = enc(m, dtype, static_cast<cuuint32_t>(rank), const_cast<void*>(ptr), dims, strides_bytes, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        return fail("cuTensorMapEncodeTiled (rank %d) failed with CUresult %d (dims %llu %llu %llu %llu, box %u %u %u %u)",
                    rank, static_cast<int>(r), (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
                    (unsigned long long)(rank > 2 ? dims[2] : 0), (unsigned long long)(rank > 3 ? dims[3] : 0), box[0],
                    rank > 1 ? box[1] : 0, rank > 2 ? box[2] : 0, rank > 3 ? box[3] : 0);
    }
    return 0;
}

// Row-major [rows, cols] output (bf16 or fp32) for the TMA epilogue: box = 128 bytes x 32 rows, 128-byte swizzle.
int make_tmap_out(CUtensorMap* m, const void* ptr, long long rows, int cols, bool bf16) {
    EncodeTiledFn enc = get_encode_fn();
    if (enc == nullptr) return fail("cuTensorMapEncodeTiled not available from the driver");
    if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0) return fail("TMA output not 16-byte aligned");
    const int esz = bf16 ? 2 : 4;
    cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
    cuuint64_t strides[1] = {static_cast<cuuint64_t>(cols) * esz};
    cuuint32_t box[2] = {static_cast<cuuint32_t>(128 / esz), 32};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(m, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                     const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return fail("cuTensorMapEncodeTiled (output) failed with CUresult %d (rows=%lld cols=%d)", static_cast<int>(r),
                    rows, cols);
    return 0;
}

// Per-device caches. The SM count and the opt-in to more than 48 KB of dynamic shared memory belong to a device, and one
// process may drive several of them (nn.DataParallel runs one thread per device: reference main.py:118).
constexpr int MAX_DEVICES = 64;
int cur_device() {
    int dev = 0;
    cudaGetDevice(&dev);
    return (dev >= 0 && dev < MAX_DEVICES) ? dev : 0;
}

int num_sms() {
    static std::atomic<int> cache[MAX_DEVICES];
    const int dev = cur_device();
    int n = cache[dev].load(std::memory_order_relaxed);
    if (n == 0) {
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
        cache[dev].store(n, std::memory_order_relaxed);
    }
    return n;
}

// Launch with programmatic dependent launch (and an optional CTA-pair cluster): the kernel may become resident and run
// its on-chip prologue while the previous kernel of the stream drains. ONLY for kernels that execute ptx::pdl_wait()
// before their first global access. Off unless CPN_PDL=1: measured on B200 (profiles/r01_pdl_ab.md) the 48-launch step
// is 4.27 ms with it and 4.24 ms without -- the launch gaps it hides are smaller than what early-resident CTAs cost.
bool pdl_enabled() {
    static const bool on = [] { const char* e = getenv("CPN_PDL"); return e != nullptr && atoi(e) != 0; }();
    return on;
}
}  // namespace cpn_api
using namespace cpn_api;

namespace {
// ------------------------------------------------------------------------------------------------ GEMM dispatch
struct GemmLaunch {
    CUtensorMap ta, tb, ta_lo, tb_lo;
    cpn::GemmParams p;
    int grid;
};

template <int BN, int ACT, bool SPLIT>
int launch_gemm_inst(const GemmLaunch& g, cudaStream_t st) {
    using S = cpn::GemmSmem<BN, SPLIT>;
    auto kern = cpn::gemm_bf16_tcgen05_kernel<BN, ACT, SPLIT>;
    CPN_CUDA(ensure_smem(kern, S::TOTAL));
    kern<<<g.grid, cpn::GEMM_THREADS, S::TOTAL, st>>>(g.ta, g.tb, g.ta_lo, g.tb_lo, g.p);
    CPN_CUDA(cudaGetLastError());
    return 0;
}

template <int BN, bool SPLIT>
int launch_gemm_act(const GemmLaunch& g, int act, cudaStream_t st) {
    switch (act) {
        case CPN_ACT_NONE: return launch_gemm_inst<BN, cpn::ACT_NONE, SPLIT>(g, st);
        case CPN_ACT_GELU: return launch_gemm_inst<BN, cpn::ACT_GELU, SPLIT>(g, st);
        case CPN_ACT_RELU: return launch_gemm_inst<BN, cpn::ACT_RELU, SPLIT>(g, st);
    }
    return fail("unknown activation %d", act);
}

template <bool SPLIT>
int launch_gemm_bn(const GemmLaunch& g, int bn, int act, cudaStream_t st) {
    switch (bn) {
        case 32: return launch_gemm_act<32, SPLIT>(g, act, st);
        case 64: return launch_gemm_act<64, SPLIT>(g, act, st);
        case 96: return launch_gemm_act<96, SPLIT>(g, act, st);
        case 128: return launch_gemm_act<128, SPLIT>(g, act, st);
        case 192: return launch_gemm_act<192, SPLIT>(g, act, st);
        case 256: return launch_gemm_act<256, SPLIT>(g, act, st);
    }
    return fail("no GEMM instantiation for BN=%d", bn);
}

// Largest tile width that divides N; fall back to narrower tiles when the grid would not fill the SMs.
int pick_bn(int M, int N, bool split) {
    static const int cands[] = {256, 192, 128, 96, 64, 32};
    const int m_blocks = (M + cpn::GEMM_BM - 1) / cpn::GEMM_BM;
    int best = 0;
    for (int bn : cands) {
        if (N % bn != 0) continue;
        if (split && bn > 128) continue;  // hi+lo planes double the stage size
        if (best == 0) best = bn;
        if (m_blocks * (N / bn) >= num_sms()) return bn;  // first (largest) width with at least one full wave
    }
    // no width fills the machine: take the narrowest divisor >= 64 for more CTAs
    for (int i = 5; i >= 0; --i) {
        const int bn = cands[i];
        if (N % bn == 0 && bn >= 64 && !(split && bn > 128)) return bn;
    }
    return best;
}

struct Gemm2Launch {
    CUtensorMap ta, tb, tout;
    cpn::Gemm2Params p;
    int grid;
    int cg;  // 1: one CTA per 128-row tile; 2: CTA pairs (cta_group::2) on 256-row tiles
};

template <int BN, int EPI, int CG, int EW = cpn::GEMM_EPI_WARPS>
int launch_gemm2_inst(const Gemm2Launch& g, cudaStream_t st) {
    using S = cpn::Gemm2Cfg<BN, EPI, CG, EW>;
    auto kern = cpn::gemm_tma_epi_kernel<BN, EPI, CG, EW>;
    CPN_CUDA(ensure_smem(kern, S::TOTAL));
    CPN_CUDA(launch_pdl(kern, dim3(g.grid), dim3(S::THREADS), S::TOTAL, st, CG, g.ta, g.tb, g.tout, g.p));
    CPN_CUDA(cudaGetLastError());
    return 0;
}

template <int EPI>
int launch_gemm2_bn(const Gemm2Launch& g, int bn, cudaStream_t st) {
    if (g.cg == 2) {
        switch (bn) {
            case 256: return launch_gemm2_inst<256, EPI, 2>(g, st);
            case 192: return launch_gemm2_inst<192, EPI, 2>(g, st);
            case 128: return launch_gemm2_inst<128, EPI, 2>(g, st);
        }
        return fail("no CTA-pair GEMM instantiation for BN=%d", bn);
    }
    if (cpn::epi_out_bf16(EPI)) {
        // bf16-output (GELU) epilogues: CPN_GEMM_EW=16 selects 16 epilogue warps. Measured: no gain at K=384 (0.0736 vs
        // 0.0716 ms) and a loss at K=768 (0.235 vs 0.193 ms), so 8 stays the default (profiles/r01_gemm_pairs.md).
        static const int ew = [] { const char* e = getenv("CPN_GEMM_EW"); return e == nullptr ? 8 : atoi(e); }();
        constexpr int EPB = cpn::epi_out_bf16(EPI) ? EPI : cpn::EPI_GELU_BF16;
        if (ew == 16) {
            switch (bn) {
                case 256: return launch_gemm2_inst<256, EPB, 1, 16>(g, st);
                case 192: return launch_gemm2_inst<192, EPB, 1, 16>(g, st);
                case 128: return launch_gemm2_inst<128, EPB, 1, 16>(g, st);
                case 64: return launch_gemm2_inst<64, EPB, 1, 16>(g, st);
            }
        }
    }
    switch (bn) {
        case 256: return launch_gemm2_inst<256, EPI, 1>(g, st);
        case 192: return launch_gemm2_inst<192, EPI, 1>(g, st);
        case 128: return launch_gemm2_inst<128, EPI, 1>(g, st);
        case 64: return launch_gemm2_inst<64, EPI, 1>(g, st);
    }
    if (!cpn::epi_out_bf16(EPI)) {
        if (bn == 96) return launch_gemm2_inst<96, cpn::epi_out_bf16(EPI) ? cpn::EPI_BIAS_F32 : EPI, 1>(g, st);
        if (bn == 32) return launch_gemm2_inst<32, cpn::epi_out_bf16(EPI) ? cpn::EPI_BIAS_F32 : EPI, 1>(g, st);
    }
    return fail("no TMA-epilogue GEMM instantiation for BN=%d", bn);
}

int launch_gemm2(const Gemm2Launch& g, int bn, int epi, cudaStream_t st) {
    switch (epi) {
        case cpn::EPI_GELU_BF16: return launch_gemm2_bn<cpn::EPI_GELU_BF16>(g, bn, st);
        case cpn::EPI_SCALE_RES_F32: return launch_gemm2_bn<cpn::EPI_SCALE_RES_F32>(g, bn, st);
        case cpn::EPI_LN_GELU_BF16: return launch_gemm2_bn<cpn::EPI_LN_GELU_BF16>(g, bn, st);
        default: return launch_gemm2_bn<cpn::EPI_BIAS_F32>(g, bn, st);
    }
}

// Tile width for the TMA-epilogue kernel: a multiple of the 128-byte output chunk that divides N.
int pick_bn2(int M, int N, int epi) {
    static const int cands[] = {256, 192, 128, 96, 64, 32};
    const int chunk = cpn::epi_out_bf16(epi) ? 64 : 32;
    if (const char* e = getenv("CPN_GEMM_BN")) {  // tuning knob (tools/run_op.py)
        const int bn = atoi(e);
        if (bn > 0 && N % bn == 0 && bn % chunk == 0 && (bn == 256 || bn == 192 || bn == 128 || bn == 64)) return bn;
    }
    const int m_blocks = (M + cpn::GEMM_BM - 1) / cpn::GEMM_BM;
    int best = 0;
    for (int bn : cands) {
        if (N % bn != 0 || bn % chunk != 0) continue;
        if (best == 0) best = bn;
        if (m_blocks * (N / bn) >= num_sms()) return bn;
    }
    for (int i = 5; i >= 0; --i) {
        const int bn = cands[i];
        if (N % bn == 0 && bn % chunk == 0 && bn >= 64) return bn;
    }
    return best;
}

// Tile geometry of a TMA-epilogue GEMM: CTA pairs (cta_group::2, 256-row tiles, each CTA stages half of the B tile)
// when the tile is wide enough to split and there is at least one pair tile per SM pair; CPN_GEMM_2SM=0 disables.
int gemm2_geometry(Gemm2Launch& g, const void* Bw, int M, int N, int K, int bn, int epi) {
    static const int allow = [] { const char* e = getenv("CPN_GEMM_2SM"); return e == nullptr ? 1 : atoi(e); }();
    const int pairs = num_sms() / 2;
    const long long pair_tiles = static_cast<long long>((M + 2 * cpn::GEMM_BM - 1) / (2 * cpn::GEMM_BM)) * (N / bn);
    // Measured (profiles/r01_gemm_pairs.md): pairs help the K-heavy fp32-output GEMMs (fc2: +4 % at K=1536, +14 % at
    // K=3072) and cost a few percent on the epilogue-heavy bf16-output ones (fc1), so only the former use them.
    // Since the pair's epilogue arrives no longer carry a GPU-scope fence (ptx::mbar_arrive_leader) pairs also gain
    // 2 % on fc1 at K = 768 (0.1970 vs 0.2009 ms); at K = 384 the difference is within noise.
    const bool wanted = !cpn::epi_out_bf16(epi) || K >= 768 || allow >= 2;  // CPN_GEMM_2SM=2: pairs everywhere (tests)
    g.cg = (allow != 0 && wanted && (bn == 256 || bn == 192 || bn == 128) && (pair_tiles >= pairs || allow == 2)) ? 2 : 1;
    if (make_tmap(&g.tb, Bw, N, K, bn / g.cg) != 0) return -1;
    g.p.num_m_blocks = (M + g.cg * cpn::GEMM_BM - 1) / (g.cg * cpn::GEMM_BM);
    g.p.num_n_blocks = N / bn;
    g.p.num_k_blocks = (K + cpn::GEMM_BK - 1) / cpn::GEMM_BK;
    static const int order = [] { const char* e = getenv("CPN_GEMM_ORDER"); return e == nullptr ? 0 : atoi(e); }();
    g.p.m_fastest = order;
#ifdef CPN_DEV_KNOBS
    static const int dbg = [] { const char* e = getenv("CPN_GEMM_DBG"); return e == nullptr ? 0 : atoi(e); }();
    g.p.dbg = dbg;
#else
    g.p.dbg = 0;
#endif
    const int tiles = g.p.num_m_blocks * g.p.num_n_blocks;
    const int slots = num_sms() / g.cg;
    g.grid = g.cg * (tiles < slots ? tiles : slots);
    return 0;
}

}  // namespace

// ------------------------------------------------------------------------------------------------ plan
struct cpn_plan {
    std::vector<Launch> ops;
    std::vector<OpInfo> info;
    cudaGraphExec_t exec = nullptr;  // the same launches as one CUDA graph (cpn_plan_capture); replayed by cpn_plan_run
    size_t captured_ops = 0;
    ~cpn_plan() {
        if (exec != nullptr) cudaGraphExecDestroy(exec);
    }
};

namespace cpn_api {
int submit(cpn_plan* plan, void* stream, Launch fn) {
    if (plan != nullptr) {
        plan->ops.push_back(std::move(fn));
        plan->info.push_back(g_info);
        return 0;
    }
    return fn(static_cast<cudaStream_t>(stream));
}
}  // namespace cpn_api

namespace {
// dwconv7_stats_kernel: 192 registers per thread (98 for the taps) -> 2 warps per scheduler, 4 CTAs of 2 warps per SM.
// The TW = 3 / 168-register variants of round 1 brought nothing (profiles/r01_ncu_full_v10.md) and are gone.
template <int C>
int launch_dws(const float* x, const float* w49, const float* bias, __nv_bfloat16* out, __nv_bfloat16* out_lo, float* stats,
               int H, int W, int nstrips, int rows, dim3 grid, cudaStream_t st) {
    if (out_lo != nullptr)
        CPN_CUDA(launch_pdl(cpn::dwconv7_stats_kernel<C, 200, cpn::DWS_TW, true>, grid, dim3(cpn::DWS_WARPS * 32), 0, st, 1, x,
                            w49, bias, out, out_lo, stats, H, W, nstrips, rows));
    else
        CPN_CUDA(launch_pdl(cpn::dwconv7_stats_kernel<C, 200, cpn::DWS_TW, false>, grid, dim3(cpn::DWS_WARPS * 32), 0, st, 1, x,
                            w49, bias, out, out_lo, stats, H, W, nstrips, rows));
    CPN_CUDA(cudaGetLastError());
    return 0;
}
}  // namespace

namespace {
struct MlpLaunch {
    CUtensorMap tx, tw1, tw2;
    cpn::MlpParams p;
    int grid;
};

template <int C, int CG>
int launch_mlp_inst(const MlpLaunch& g, cudaStream_t st) {
    using S = cpn::MlpCfg<C, CG>;
    auto kern = cpn::mlp_fused_kernel<C, CG>;
    CPN_CUDA(ensure_smem(kern, S::TOTAL));
    CPN_CUDA(launch_pdl(kern, dim3(g.grid), dim3(S::THREADS), S::TOTAL, st, CG, g.tx, g.tw1, g.tw2, g.p));
    CPN_CUDA(cudaGetLastError());
    return 0;
}

int mlp_cg(int C) {
    static const int env = [] { const char* e = getenv("CPN_MLP_CG"); return e == nullptr ? 0 : atoi(e); }();
    if (C == 384) return 2;  // single CTAs have no room for the weight ring next to the 96 KB X tile
    return env == 1 ? 1 : 2;
}
}  // namespace

extern "C" {

int cpn_version(void) { return 100; }

const char* cpn_last_error(void) { return g_err.c_str(); }

int cpn_check_device(void) {
    int dev = 0;
    CPN_CUDA(cudaGetDevice(&dev));
    int major = 0, minor = 0;
    CPN_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    CPN_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
    if (major != 10) return fail("libcpnb200 is built for sm_100a only; device %d is sm_%d%d", dev, major, minor);
    if (get_encode_fn() == nullptr) return fail("driver does not export cuTensorMapEncodeTiled");
    return 0;
}

int cpn_plan_create(cpn_plan** out) {
    if (out == nullptr) return fail("cpn_plan_create: null output");
    *out = new cpn_plan();
    return 0;
}

int cpn_plan_capture(cpn_plan* plan, void* stream) {
    if (plan == nullptr) return fail("cpn_plan_capture: null plan");
    if (plan->ops.empty()) return fail("cpn_plan_capture: empty plan");
    (void)stream;  // the capture only records: it runs on a private stream (torch's default stream is the legacy stream,
                   // which cannot be captured) and the graph is later launched into whatever stream cpn_plan_run is given
    if (plan->exec != nullptr) {
        cudaGraphExecDestroy(plan->exec);
        plan->exec = nullptr;
    }
    cudaStream_t st = nullptr;
    CPN_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    // ThreadLocal: other host threads (a second device's DataParallel thread, the allocator) are not affected
    cudaError_t e = cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
    if (e != cudaSuccess) {
        cudaStreamDestroy(st);
        return fail("cudaStreamBeginCapture: %s", cudaGetErrorString(e));
    }
    int rc = 0;
    for (auto& op : plan->ops) {
        rc = op(st);
        if (rc != 0) break;
    }
    cudaGraph_t graph = nullptr;
    e = cudaStreamEndCapture(st, &graph);  // always: leaves the stream out of capture mode, also after a failed op
    cudaStreamDestroy(st);
    if (rc != 0) {
        if (graph != nullptr) cudaGraphDestroy(graph);
        cudaGetLastError();  // a failed capture leaves a sticky-looking error behind; it is reported through rc
        return rc;
    }
    if (e != cudaSuccess || graph == nullptr) {
        cudaGetLastError();
        return fail("cudaStreamEndCapture: %s", cudaGetErrorString(e));
    }
    const cudaError_t ei = cudaGraphInstantiate(&plan->exec, graph, 0);
    cudaGraphDestroy(graph);
    if (ei != cudaSuccess) {
        plan->exec = nullptr;
        cudaGetLastError();
        return fail("cudaGraphInstantiate: %s", cudaGetErrorString(ei));
    }
    plan->captured_ops = plan->ops.size();
    return 0;
}

int cpn_plan_has_graph(const cpn_plan* plan) {
    return plan != nullptr && plan->exec != nullptr && plan->captured_ops == plan->ops.size() ? 1 : 0;
}

int cpn_plan_run(cpn_plan* plan, void* stream) {
    if (plan == nullptr) return fail("cpn_plan_run: null plan");
    if (plan->exec != nullptr && plan->captured_ops == plan->ops.size()) {
        CPN_CUDA(cudaGraphLaunch(plan->exec, static_cast<cudaStream_t>(stream)));
        return 0;
    }
    for (auto& op : plan->ops) {
        const int rc = op(static_cast<cudaStream_t>(stream));
        if (rc != 0) return rc;
    }
    return 0;
}

int cpn_plan_run_timed(cpn_plan* plan, void* stream, float* ms_out, int n) {
    if (plan == nullptr || ms_out == nullptr) return fail("cpn_plan_run_timed: null argument");
    const int nops = static_cast<int>(plan->ops.size());
    if (n < nops) return fail("cpn_plan_run_timed: output holds %d entries, plan has %d ops", n, nops);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    std::vector<cudaEvent_t> ev(nops + 1);
    for (auto& e : ev) CPN_CUDA(cudaEventCreate(&e));
    CPN_CUDA(cudaEventRecord(ev[0], st));
    int rc = 0;
    for (int i = 0; i < nops && rc == 0; ++i) {
        rc = plan->ops[i](st);
        if (rc == 0) CPN_CUDA(cudaEventRecord(ev[i + 1], st));
    }
    if (rc == 0) {
        CPN_CUDA(cudaEventSynchronize(ev[nops]));
        for (int i = 0; i < nops; ++i) CPN_CUDA(cudaEventElapsedTime(&ms_out[i], ev[i], ev[i + 1]));
    }
    for (auto& e : ev) cudaEventDestroy(e);
    return rc;
}

int cpn_plan_op_info(const cpn_plan* plan, int i, const char** name, double* flops, double* bytes) {
    if (plan == nullptr || i < 0 || i >= static_cast<int>(plan->info.size())) return fail("cpn_plan_op_info: bad index");
    if (name != nullptr) *name = plan->info[i].name;
    if (flops != nullptr) *flops = plan->info[i].flops;
    if (bytes != nullptr) *bytes = plan->info[i].bytes;
    return 0;
}

const char* cpn_plan_op_detail(const cpn_plan* plan, int i) {
    if (plan == nullptr || i < 0 || i >= static_cast<int>(plan->info.size())) return "";
    return plan->info[i].detail.c_str();
}

int cpn_plan_num_ops(const cpn_plan* plan) { return plan == nullptr ? -1 : static_cast<int>(plan->ops.size()); }

int cpn_plan_destroy(cpn_plan* plan) {
    delete plan;
    return 0;
}

// ------------------------------------------------------------------------------------------------ ConvNeXt ops
int cpn_stem_ln(cpn_plan* plan, const float* x, const float* w, const float* bias, const float* ln_w,
                const float* ln_b, float* out, int B, int H, int W, float eps, void* stream) {
    if (B <= 0 || H < 4 || W < 4) return fail("cpn_stem_ln: bad shape B=%d H=%d W=%d", B, H, W);
    const int Ho = (H - 4) / 4 + 1, Wo = (W - 4) / 4 + 1;
    const long long groups = static_cast<long long>(B) * Ho * ((Wo + cpn::STEM_TOK - 1) / cpn::STEM_TOK);
    long long blocks = (groups + cpn::STEM_WARPS - 1) / cpn::STEM_WARPS;
    const long long cap = static_cast<long long>(num_sms()) * 2;  // weights live in registers: keep CTAs persistent
    if (blocks > cap) blocks = cap;
    const int grid = static_cast<int>(blocks);
    set_info("stem_ln", 2.0 * 48 * 96 * B * Ho * Wo, 4.0 * B * (3.0 * H * W + 96.0 * Ho * Wo), "B=%d H=%d W=%d", B, H, W);
    const size_t smem2 = (12 * 96 * 4 + 2 * 12 * 4 * static_cast<size_t>(Wo)) * sizeof(float);
    if (W % 4 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0 && smem2 <= 160 * 1024 && getenv("CPN_STEM_V1") == nullptr) {
        long long rows = static_cast<long long>(B) * Ho;
        const long long cap2 = 2LL * num_sms();  // 124 registers x 256 threads: two resident CTAs per SM
        const int grid2 = static_cast<int>(rows < cap2 ? rows : cap2);
        return submit(plan, stream, [=](cudaStream_t st) -> int {
            CPN_CUDA(ensure_smem(cpn::stem_ln_rows_kernel, smem2));
            CPN_CUDA(launch_pdl(cpn::stem_ln_rows_kernel, grid2, dim3(cpn::STEM2_THREADS), smem2, st, 1, x, w, bias, ln_w, ln_b,
                                out, B, H, W, Ho, Wo, eps));
            CPN_CUDA(cudaGetLastError());
            return 0;
        });
    }
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        cpn::stem_ln_kernel<<<grid, cpn::STEM_WARPS * 32, 0, st>>>(x, w, bias, ln_w, ln_b, out, B, H, W, Ho, Wo, eps);
        CPN_CUDA(cudaGetLastError());
        return 0;
    });
}

int cpn_dwconv7_ln(cpn_plan* plan, const float* x, const float* w49, const float* bias, const float* ln_w,
                   const float* ln_b, void* out_hi, void* out_lo, int B, int H, int W, int C, float eps,
                   void* stream) {
    if (C % 2 != 0 || C > 2 * 32 * cpn::DW_MAX_WARPS) return fail("cpn_dwconv7_ln: unsupported C=%d", C);
    if (B <= 0 || B > 65535) return fail("cpn_dwconv7_ln: unsupported B=%d", B);
    const int threads = ((C / 2 + 31) / 32) * 32;
    const int tw = (threads > 192) ? 2 : 4;
    // rows per CTA: whole column strips when the batch already fills the machine, shorter segments otherwise
    const int strips = (W + tw - 1) / tw;
    int rows = H;
    while (rows > 8 && static_cast<long long>(B) * strips * ((H + rows - 1) / rows) < 2LL * num_sms()) rows = (rows + 1) / 2;
    const dim3 grid(strips, (H + rows - 1) / rows, B);
    auto* hi = static_cast<__nv_bfloat16*>(out_hi);
    auto* lo = static_cast<__nv_bfloat16*>(out_lo);
    set_info("dwconv7_ln", 2.0 * 49 * B * H * W * C, 1.0 * B * H * W * C * (4 + 2 + (out_lo != nullptr ? 2 : 0)),
             "B=%d H=%d W=%d C=%d", B, H, W, C);
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        if (tw == 2)
            cpn::dwconv7_ln_kernel<2><<<grid, threads, 0, st>>>(x, w49, bias, ln_w, ln_b, hi, lo, H, W, C, rows, eps);
        else
            cpn::dwconv7_ln_kernel<4><<<grid, threads, 0, st>>>(x, w49, bias, ln_w, ln_b, hi, lo, H, W, C, rows, eps);
        CPN_CUDA(cudaGetLastError());
        return 0;
    });
}

}  // extern "C"

// First-generation kernel (taps in registers, 8 warps per SM); cpn_dwconv7_stats (api_dwconv.cu) selects it with
// CPN_DW_V1=1 for A/B measurements.
int cpn_api::dwconv7_stats_v1(cpn_plan* plan, const float* x, const float* w49, const float* bias, void* out_bf16,
                              void* out_lo, float* stats, int B, int H, int W, int C, void* stream) {
    if (C != 96 && C != 192 && C != 384 && C != 768)
        return fail("cpn_dwconv7_stats: C=%d is not a ConvNeXt-tiny width (96, 192, 384, 768)", C);
    if (B <= 0 || B > 65535 || H <= 0 || W <= 0) return fail("cpn_dwconv7_stats: bad shape B=%d H=%d W=%d", B, H, W);
    constexpr int tw = cpn::DWS_TW;
    const int nstrips = (W + tw - 1) / tw;
    const int warps = (nstrips * (C / 32) + 1) / 2;  // a warp = two (strip, 32-channel group) units
    const int gx = (warps + cpn::DWS_WARPS - 1) / cpn::DWS_WARPS;
    // row segments: every segment pays 6 halo rows of loads AND of FMAs (the body is branch-free), so split H only
    // when the grid would not give each SM its resident CTAs a few times over
    int nseg = 1;
    while (static_cast<long long>(gx) * nseg * B < 5LL * num_sms() && (H + nseg) / (nseg + 1) >= 12) ++nseg;
    if (const char* e = getenv("CPN_DWS_NSEG")) nseg = atoi(e) > 0 ? atoi(e) : nseg;  // tuning knob (tools/run_op.py)
    const int rows = (H + nseg - 1) / nseg;
    nseg = (H + rows - 1) / rows;
    const dim3 grid(gx, nseg, B);
    auto* out = static_cast<__nv_bfloat16*>(out_bf16);
    auto* lo = static_cast<__nv_bfloat16*>(out_lo);
    const double tok = static_cast<double>(B) * H * W;
    set_info(lo != nullptr ? "dwconv7_split" : "dwconv7_stats", 2.0 * 49 * tok * C,
             tok * C * (4 + 2 + (lo != nullptr ? 2 : 0)) + tok * (C / 32) * 8, "B=%d H=%d W=%d C=%d rows/seg=%d ctas=%d", B, H,
             W, C, rows, gx * nseg * B);
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        switch (C) {
            case 96: return launch_dws<96>(x, w49, bias, out, lo, stats, H, W, nstrips, rows, grid, st);
            case 192: return launch_dws<192>(x, w49, bias, out, lo, stats, H, W, nstrips, rows, grid, st);
            case 384: return launch_dws<384>(x, w49, bias, out, lo, stats, H, W, nstrips, rows, grid, st);
            default: return launch_dws<768>(x, w49, bias, out, lo, stats, H, W, nstrips, rows, grid, st);
        }
    });
}

extern "C" {

int cpn_gemm_ln_gelu(cpn_plan* plan, const void* A, const void* Bw, int M, int N, int K, const float* stats,
                     int nparts, const float* colsum, const float* bias, float eps, void* out_bf16, void* stream) {
    if (M <= 0 || N <= 0 || K <= 0) return fail("cpn_gemm_ln_gelu: bad shape M=%d N=%d K=%d", M, N, K);
    if (K % 8 != 0) return fail("cpn_gemm_ln_gelu: K must be a multiple of 8, K=%d", K);
    if (nparts > cpn::LN_MAX_PARTS) return fail("cpn_gemm_ln_gelu: nparts=%d exceeds %d", nparts, cpn::LN_MAX_PARTS);
    if (nparts <= 0 || stats == nullptr || colsum == nullptr || bias == nullptr || out_bf16 == nullptr)
        return fail("cpn_gemm_ln_gelu: missing operand");
    const int epi = cpn::EPI_LN_GELU_BF16;
    const int bn2 = pick_bn2(M, N, epi);
    if (bn2 == 0 || bn2 == 96 || bn2 == 32) return fail("cpn_gemm_ln_gelu: no tile width for N=%d", N);
    Gemm2Launch g;
    memset(&g, 0, sizeof(g));
    if (make_tmap(&g.ta, A, M, K, cpn::GEMM_BM) != 0) return -1;
    if (make_tmap_out(&g.tout, out_bf16, M, N, true) != 0) return -1;
    g.p.M = M; g.p.N = N; g.p.K = K;
    if (gemm2_geometry(g, Bw, M, N, K, bn2, epi) != 0) return -1;
    g.p.bias = bias; g.p.scale = nullptr;
    g.p.stats = stats; g.p.colsum = colsum; g.p.nparts = nparts; g.p.inv_k = 1.0f / static_cast<float>(K); g.p.eps = eps;
    const double mn = static_cast<double>(M) * N;
    set_info("gemm_bf16", 2.0 * mn * K, 2.0 * (static_cast<double>(M) * K + static_cast<double>(N) * K) + 2.0 * mn,
             "M=%d N=%d K=%d BN=%d cg=%d epi=ln_gelu_bf16", M, N, K, bn2, g.cg);
    return submit(plan, stream, [=](cudaStream_t st) -> int { return launch_gemm2(g, bn2, epi, st); });
}


int cpn_mlp_fused(cpn_plan* plan, const void* y_bf16, const void* w1g, const void* w2, int M, int C, const float* stats,
                  int nparts, const float* colsum, const float* bias1, const float* bias2, const float* scale, float eps,
                  float* x, void* stream) {
    if (M <= 0) return fail("cpn_mlp_fused: bad M=%d", M);
    if (C != 96 && C != 192 && C != 384) return fail("cpn_mlp_fused: C must be 96, 192 or 384 (C=%d)", C);
    if (nparts != C / 32) return fail("cpn_mlp_fused: nparts=%d, expected C/32=%d", nparts, C / 32);
    if (y_bf16 == nullptr || w1g == nullptr || w2 == nullptr || stats == nullptr || colsum == nullptr ||
        bias1 == nullptr || bias2 == nullptr || scale == nullptr || x == nullptr)
        return fail("cpn_mlp_fused: missing operand");
    if ((reinterpret_cast<uintptr_t>(x) & 15) != 0) return fail("cpn_mlp_fused: x not 16-byte aligned");
    const int cg = mlp_cg(C);
    const int n2 = C > 256 ? C / 2 : C;
    MlpLaunch g;
    memset(&g, 0, sizeof(g));
    if (make_tmap(&g.tx, y_bf16, M, C, cpn::GEMM_BM) != 0) return -1;
    if (make_tmap(&g.tw1, w1g, 4LL * C, C, 64 / cg) != 0) return -1;
    if (make_tmap(&g.tw2, w2, C, 4 * C, n2 / cg) != 0) return -1;
    g.p.M = M;
    g.p.num_tiles = (M + cg * cpn::GEMM_BM - 1) / (cg * cpn::GEMM_BM);
    g.p.stats = stats; g.p.nparts = nparts; g.p.inv_c = 1.0f / static_cast<float>(C); g.p.eps = eps;
    g.p.colsum = colsum; g.p.bias1 = bias1; g.p.bias2 = bias2; g.p.scale = scale; g.p.x = x;
    // Groups that run one tile less than the others start half a tile late (estimated from the chunk count): the
    // final epilogues of all CTAs otherwise hit HBM/L2 in the same few microseconds (profiles/r01_mlp_fused.md).
    static const int stagger_pct = [] { const char* e = getenv("CPN_MLP_STAGGER"); return e == nullptr ? 50 : atoi(e); }();
    const int tile_cycles = C == 384 ? 75000 : C == 192 ? 32000 : 15500;  // measured, profiles/r01_mlp_fused.md
    g.p.stagger = tile_cycles / 100 * stagger_pct;  // negative percentage: delays spread over 0.5 .. 1.44 of it
#ifdef CPN_DEV_KNOBS
    static const int mlp_dbg = [] { const char* e = getenv("CPN_MLP_DBG"); return e == nullptr ? 0 : atoi(e); }();
    g.p.dbg = mlp_dbg;
#else
    g.p.dbg = 0;
#endif
    const int slots = num_sms() / cg;
    g.grid = cg * (g.p.num_tiles < slots ? g.p.num_tiles : slots);
    const double mc = static_cast<double>(M) * C;
    set_info("mlp_fused", 16.0 * mc * C, 2.0 * mc + 8.0 * mc + 16.0 * C * C, "M=%d C=%d cg=%d", M, C, cg);
#ifdef CPN_DEV_KNOBS
    static long long* ts_dev = nullptr;  // dev only: CPN_MLP_TS=1 prints a clock64 trace of CTA 0's first tile
    static const bool want_ts = getenv("CPN_MLP_TS") != nullptr;
    if (want_ts && plan == nullptr) {
        if (ts_dev == nullptr) CPN_CUDA(cudaMalloc(&ts_dev, 512 * sizeof(long long)));
        CPN_CUDA(cudaMemset(ts_dev, 0, 512 * sizeof(long long)));
        g.p.ts = ts_dev;
    }
#endif
    auto run = [=](cudaStream_t st) -> int {
        if (cg == 2) {
            switch (C) {
                case 96: return launch_mlp_inst<96, 2>(g, st);
                case 192: return launch_mlp_inst<192, 2>(g, st);
                default: return launch_mlp_inst<384, 2>(g, st);
            }
        }
        return C == 96 ? launch_mlp_inst<96, 1>(g, st) : launch_mlp_inst<192, 1>(g, st);
    };
#ifdef CPN_DEV_KNOBS
    if (g.p.ts != nullptr) {
        if (run(static_cast<cudaStream_t>(stream)) != 0) return -1;
        CPN_CUDA(cudaDeviceSynchronize());
        static long long h[512];
        CPN_CUDA(cudaMemcpy(h, ts_dev, sizeof(h), cudaMemcpyDeviceToHost));
        long long t0 = h[0];
        for (int j = 0; j < 512; ++j) if (h[j] != 0 && h[j] < t0) t0 = h[j];
        fprintf(stderr, "# epilogue warp 4 (chunk: start, a1_full, tmem_ld, a1_empty arrive, h_empty, math+sts, fence, h_full arrive)\n");
        for (int j = 0; j < C / 16 && j < 32; ++j) {
            fprintf(stderr, "E %2d:", j);
            for (int k = 0; k < 8; ++k) fprintf(stderr, " %7lld", h[j * 8 + k] - t0);
            fprintf(stderr, "\n");
        }
        fprintf(stderr, "# tiles of CTA 0 (chunk loop start, end, acc2 ready, final epilogue end)\n");
        for (int t = 0; t < 8; ++t)
            if (h[480 + 4 * t] != 0)
                fprintf(stderr, "T %d: %7lld %7lld %7lld %7lld\n", t, h[480 + 4 * t] - h[480], h[481 + 4 * t] - h[480],
                        h[482 + 4 * t] - h[480], h[483 + 4 * t] - h[480]);
        fprintf(stderr, "# MMA thread (step: G1 start, w_full, a1_empty | G2 start, w_full, h_full)\n");
        for (int j = 0; j <= C / 16 && j < 32; ++j) {
            fprintf(stderr, "M %2d:", j);
            for (int k = 0; k < 6; ++k) fprintf(stderr, " %7lld", h[256 + j * 8 + k] ? h[256 + j * 8 + k] - t0 : 0);
            fprintf(stderr, "\n");
        }
        return 0;
    }
#endif
    return submit(plan, stream, run);
}

int cpn_ln_im2col2(cpn_plan* plan, const float* x, const float* ln_w, const float* ln_b, void* out_hi, void* out_lo,
                   int B, int H, int W, int C, int stride, float eps, void* stream) {
    if (C % 32 != 0 || C > 32 * cpn::LN_MAX_PER_LANE) return fail("cpn_ln_im2col2: unsupported C=%d", C);
    if (stride != 1 && stride != 2) return fail("cpn_ln_im2col2: stride must be 1 or 2");
    if (H < 2 || W < 2) return fail("cpn_ln_im2col2: map too small");
    const int Ho = (H - 2) / stride + 1, Wo = (W - 2) / stride + 1;
    const long long ntok = static_cast<long long>(B) * H * W;
    const int grid = static_cast<int>((ntok + 7) / 8);
    auto* hi = static_cast<__nv_bfloat16*>(out_hi);
    auto* lo = static_cast<__nv_bfloat16*>(out_lo);
    set_info("ln_im2col2", 8.0 * ntok * C, 4.0 * ntok * C + 1.0 * B * Ho * Wo * 4 * C * (out_lo != nullptr ? 4 : 2),
             "B=%d H=%d W=%d C=%d s=%d", B, H, W, C, stride);
    if (C % 4 == 0 && C <= 768 && ntok < (1LL << 31) - 65536 && (lo == nullptr || C == 96 || C == 192 || C == 384)) {
        // vectorised persistent kernels, two tokens in flight per warp
        const int nv = (C / 4 + 31) / 32;
        long long want = (ntok + 15) / 16;
        const long long cap = 8LL * num_sms();
        const int pgrid = static_cast<int>(want < cap ? want : cap);
        static const bool v3 = getenv("CPN_LNIM_V2") == nullptr;
        if ((v3 || lo != nullptr) && (C == 96 || C == 192 || C == 384)) {
            const int tpw = 32 / (C / 12);
            long long want3 = (ntok + 16 * tpw - 1) / (16 * tpw);  // 8 warps x 2 passes x tpw tokens per CTA
            const int pgrid3 = static_cast<int>(want3 < cap ? want3 : cap);
            return submit(plan, stream, [=](cudaStream_t st) -> int {
#define CPN_LNIM3(LPT)                                                                                                     \
    if (lo != nullptr)                                                                                                     \
        CPN_CUDA(launch_pdl(cpn::ln_im2col3_bf16_kernel<LPT, true>, pgrid3, dim3(256), 0, st, 1, x, ln_w, ln_b, hi, lo, B, H, W, \
                            stride, Ho, Wo, eps));                                                                         \
    else                                                                                                                   \
        CPN_CUDA(launch_pdl(cpn::ln_im2col3_bf16_kernel<LPT, false>, pgrid3, dim3(256), 0, st, 1, x, ln_w, ln_b, hi, lo, B, H, W, \
                            stride, Ho, Wo, eps));
                if (C == 96) { CPN_LNIM3(8) }
                else if (C == 192) { CPN_LNIM3(16) }
                else { CPN_LNIM3(32) }
#undef CPN_LNIM3
                CPN_CUDA(cudaGetLastError());
                return 0;
            });
        }
        return submit(plan, stream, [=](cudaStream_t st) -> int {
            switch (nv) {
                case 1: cpn::ln_im2col2_bf16_kernel<1><<<pgrid, 256, 0, st>>>(x, ln_w, ln_b, hi, B, H, W, C, stride, Ho, Wo, eps); break;
                case 2: cpn::ln_im2col2_bf16_kernel<2><<<pgrid, 256, 0, st>>>(x, ln_w, ln_b, hi, B, H, W, C, stride, Ho, Wo, eps); break;
                case 3: cpn::ln_im2col2_bf16_kernel<3><<<pgrid, 256, 0, st>>>(x, ln_w, ln_b, hi, B, H, W, C, stride, Ho, Wo, eps); break;
                case 4: cpn::ln_im2col2_bf16_kernel<4><<<pgrid, 256, 0, st>>>(x, ln_w, ln_b, hi, B, H, W, C, stride, Ho, Wo, eps); break;
                default: cpn::ln_im2col2_bf16_kernel<6><<<pgrid, 256, 0, st>>>(x, ln_w, ln_b, hi, B, H, W, C, stride, Ho, Wo, eps); break;
            }
            CPN_CUDA(cudaGetLastError());
            return 0;
        });
    }
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        cpn::ln_im2col2_kernel<<<grid, 256, 0, st>>>(x, ln_w, ln_b, hi, lo, B, H, W, C, stride, Ho, Wo, eps);
        CPN_CUDA(cudaGetLastError());
        return 0;
    });
}

int cpn_gemm_bf16(cpn_plan* plan, const void* A, const void* A_lo, const void* Bw, const void* Bw_lo, int M, int N,
                  int K, const float* bias, const float* scale, const float* res_f32, const void* res_bf16,
                  float* out_f32, void* out_bf16, void* out_bf16_lo, int act, int relu_after_res, void* stream) {
    if (M <= 0 || N <= 0 || K <= 0) return fail("cpn_gemm_bf16: bad shape M=%d N=%d K=%d", M, N, K);
    if (N % 32 != 0) return fail("cpn_gemm_bf16: N must be a multiple of 32, N=%d", N);
    if (K % 8 != 0) return fail("cpn_gemm_bf16: K must be a multiple of 8, K=%d", K);
    if ((A_lo == nullptr) != (Bw_lo == nullptr)) return fail("cpn_gemm_bf16: A_lo and Bw_lo must both be given");
    if (out_f32 == nullptr && out_bf16 == nullptr) return fail("cpn_gemm_bf16: no output");
    if (out_bf16_lo != nullptr && out_bf16 == nullptr) return fail("cpn_gemm_bf16: lo plane needs the hi plane");
    const bool split = A_lo != nullptr;
    if (!split && bias != nullptr && out_bf16_lo == nullptr && res_bf16 == nullptr && !relu_after_res &&
        getenv("CPN_GEMM_LEGACY") == nullptr) {
        int epi = -1;
        if (act == CPN_ACT_GELU && out_bf16 != nullptr && out_f32 == nullptr && scale == nullptr && res_f32 == nullptr)
            epi = cpn::EPI_GELU_BF16;
        else if (act == CPN_ACT_NONE && out_f32 != nullptr && out_bf16 == nullptr && scale != nullptr &&
                 res_f32 == out_f32)
            epi = cpn::EPI_SCALE_RES_F32;
        else if (act == CPN_ACT_NONE && out_f32 != nullptr && out_bf16 == nullptr && scale == nullptr &&
                 res_f32 == nullptr)
            epi = cpn::EPI_BIAS_F32;
        const int bn2 = epi >= 0 ? pick_bn2(M, N, epi) : 0;
        if (bn2 != 0) {
            Gemm2Launch g;
            memset(&g, 0, sizeof(g));
            if (make_tmap(&g.ta, A, M, K, cpn::GEMM_BM) != 0) return -1;
            const bool obf = epi == cpn::EPI_GELU_BF16;
            if (make_tmap_out(&g.tout, obf ? out_bf16 : static_cast<void*>(out_f32), M, N, obf) != 0) return -1;
            g.p.M = M; g.p.N = N; g.p.K = K;
            if (gemm2_geometry(g, Bw, M, N, K, bn2, epi) != 0) return -1;
            g.p.bias = bias; g.p.scale = scale;
            const double mn = static_cast<double>(M) * N;
            const double bytes = 2.0 * (static_cast<double>(M) * K + static_cast<double>(N) * K) +
                                 (obf ? 2.0 : (epi == cpn::EPI_SCALE_RES_F32 ? 8.0 : 4.0)) * mn;
            set_info("gemm_bf16", 2.0 * mn * K, bytes, "M=%d N=%d K=%d BN=%d cg=%d epi=%s", M, N, K, bn2, g.cg,
                     obf ? "gelu_bf16" : (epi == cpn::EPI_SCALE_RES_F32 ? "scale_res_f32" : "bias_f32"));
            return submit(plan, stream, [=](cudaStream_t st) -> int { return launch_gemm2(g, bn2, epi, st); });
        }
    }
    if (split && bias != nullptr && out_bf16 == nullptr && out_f32 != nullptr && res_bf16 == nullptr && !relu_after_res &&
        act == CPN_ACT_NONE && getenv("CPN_GEMM_LEGACY") == nullptr) {
        // fp32 mode, fp32 output: three-pass kernel with the TMA epilogue (api_gemm_split.cu)
        int epi = -1;
        if (scale != nullptr && res_f32 == out_f32) epi = cpn::EPI_SCALE_RES_F32;
        else if (scale == nullptr && res_f32 == nullptr) epi = cpn::EPI_BIAS_F32;
        if (epi >= 0 && gemm_split_supported(N, K, epi))
            return gemm_split(plan, epi, A, A_lo, Bw, Bw_lo, M, N, K, bias, scale, nullptr, 0, nullptr, 0.f, out_f32, nullptr,
                              stream);
    }
    const int bn = pick_bn(M, N, split);
    if (bn == 0) return fail("cpn_gemm_bf16: no tile width divides N=%d", N);
    GemmLaunch g;
    memset(&g, 0, sizeof(g));
    if (make_tmap(&g.ta, A, M, K, cpn::GEMM_BM) != 0) return -1;
    if (make_tmap(&g.tb, Bw, N, K, bn) != 0) return -1;
    if (split) {
        if (make_tmap(&g.ta_lo, A_lo, M, K, cpn::GEMM_BM) != 0) return -1;
        if (make_tmap(&g.tb_lo, Bw_lo, N, K, bn) != 0) return -1;
    } else {
        g.ta_lo = g.ta;
        g.tb_lo = g.tb;
    }
    g.p.M = M; g.p.N = N; g.p.K = K;
    g.p.num_m_blocks = (M + cpn::GEMM_BM - 1) / cpn::GEMM_BM;
    g.p.num_n_blocks = N / bn;
    g.p.num_k_blocks = (K + cpn::GEMM_BK - 1) / cpn::GEMM_BK;
    g.p.bias = bias; g.p.scale = scale; g.p.res_f32 = res_f32;
    g.p.res_bf16 = static_cast<const __nv_bfloat16*>(res_bf16);
    g.p.out_f32 = out_f32;
    g.p.out_bf16 = static_cast<__nv_bfloat16*>(out_bf16);
    g.p.out_bf16_lo = static_cast<__nv_bfloat16*>(out_bf16_lo);
    g.p.relu_after_res = relu_after_res;
    const int tiles = g.p.num_m_blocks * g.p.num_n_blocks;
    g.grid = tiles < num_sms() ? tiles : num_sms();
    {
        double bytes = 2.0 * (static_cast<double>(M) * K + static_cast<double>(N) * K) * (split ? 2 : 1);
        const double mn = static_cast<double>(M) * N;
        if (res_f32 != nullptr) bytes += 4.0 * mn;
        if (res_bf16 != nullptr) bytes += 2.0 * mn;
        if (out_f32 != nullptr) bytes += 4.0 * mn;
        if (out_bf16 != nullptr) bytes += 2.0 * mn;
        if (out_bf16_lo != nullptr) bytes += 2.0 * mn;
        set_info(split ? "gemm_bf16x3" : "gemm_bf16", 2.0 * mn * K * (split ? 3 : 1), bytes, "M=%d N=%d K=%d BN=%d act=%d res=%d",
                 M, N, K, bn, act, res_f32 != nullptr || res_bf16 != nullptr);
    }
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        return split ? launch_gemm_bn<true>(g, bn, act, st) : launch_gemm_bn<false>(g, bn, act, st);
    });
}

// ------------------------------------------------------------------------------------------------ head ops
int cpn_addon_conv1x1(cpn_plan* plan, const float* feat, const float* w, const float* bias, float* z, long long T,
                      int C, int P, void* stream) {
    const size_t smem = static_cast<size_t>(P) * C * sizeof(float);
    if (smem > 200 * 1024) return fail("cpn_addon_conv1x1: P*C=%d*%d too large for the small-P kernel", P, C);
    long long blocks = (T + 7) / 8;
    if (blocks > 4LL * num_sms()) blocks = 4LL * num_sms();
    const int grid = static_cast<int>(blocks);
    set_info("addon_conv1x1", 2.0 * T * C * P, 4.0 * T * (C + P));
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        CPN_CUDA(ensure_smem(cpn::addon_conv1x1_kernel, smem));
        cpn::addon_conv1x1_kernel<<<grid, 256, smem, st>>>(feat, w, bias, z, T, C, P);
        CPN_CUDA(cudaGetLastError());
        return 0;
    });
}

int cpn_addon_softmax_pool(cpn_plan* plan, const float* feat, const float* w, const float* bias, const float* gumbel,
                           float tau, int hard, int mode, float* map, float* partial, int* partial_arg, int B, int HW, int C,
                           int P, int chunks, void* stream) {
    if (P != 16 && P != 32) return fail("cpn_addon_softmax_pool: P must be 16 or 32 (P=%d)", P);
    if (C <= 0 || C % 64 != 0) return fail("cpn_addon_softmax_pool: C must be a multiple of 64 (C=%d)", C);
    if (chunks <= 0 || B <= 0 || B > 65535 || HW <= 0) return fail("cpn_addon_softmax_pool: bad chunks / B / HW");
    if (feat == nullptr || w == nullptr || partial == nullptr) return fail("cpn_addon_softmax_pool: missing operand");
    const size_t smem = (static_cast<size_t>(C) * P + cpn::ADDON_WARPS * (32 * 65 + 2 * P)) * sizeof(float);
    if (smem > 200 * 1024) return fail("cpn_addon_softmax_pool: C=%d P=%d needs %zu bytes of shared memory", C, P, smem);
    const dim3 grid(chunks, B);
    int* parg = (partial_arg != nullptr && mode == CPN_POOL_MAX) ? partial_arg : nullptr;
    set_info("addon_head", 2.0 * B * HW * C * P, 4.0 * B * HW * (C + P * ((gumbel != nullptr) + (map != nullptr))),
             "B=%d HW=%d C=%d P=%d chunks=%d", B, HW, C, P, chunks);
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        if (P == 16) {
            CPN_CUDA(ensure_smem(cpn::addon_head_kernel<16>, smem));
            CPN_CUDA(launch_pdl(cpn::addon_head_kernel<16>, grid, dim3(32 * cpn::ADDON_WARPS), smem, st, 1, feat, w, bias, gumbel, tau, hard, mode,
                                map, partial, parg, HW, C, chunks));
        } else {
            CPN_CUDA(ensure_smem(cpn::addon_head_kernel<32>, smem));
            CPN_CUDA(launch_pdl(cpn::addon_head_kernel<32>, grid, dim3(32 * cpn::ADDON_WARPS), smem, st, 1, feat, w, bias, gumbel, tau, hard, mode,
                                map, partial, parg, HW, C, chunks));
        }
        CPN_CUDA(cudaGetLastError());
        return 0;
    });
}

int cpn_softmax_pool_arg(cpn_plan* plan, const float* z, const float* gumbel, float tau, int hard, int mode, float* map,
                         float* partial, int* partial_arg, int B, int HW, int P, int chunks, void* stream) {
    if (chunks <= 0 || B <= 0 || B > 65535) return fail("cpn_softmax_pool: bad chunks/B");
    const bool arg = partial_arg != nullptr && mode == CPN_POOL_MAX;
    const size_t smem = 8 * static_cast<size_t>(P) * sizeof(float) * (arg ? 2 : 1);
    if (smem > 200 * 1024) return fail("cpn_softmax_pool: P=%d too large", P);
    const dim3 grid(chunks, B);
    set_info("softmax_pool", 4.0 * B * HW * P, 4.0 * B * HW * P * (1 + (gumbel != nullptr) + (map != nullptr)));
    if (gumbel == nullptr && !hard && map == nullptr && P % 128 == 0 && (P <= 1024 || P == 2048)) {
        // plain softmax, map not requested: register-resident kernel
        const int nv = P / 128;
        return submit(plan, stream, [=](cudaStream_t st) -> int {
#define CPN_SMP_CASE(NVV)                                                                                              \
    case NVV:                                                                                                          \
        if (arg) {                                                                                                     \
            CPN_CUDA(ensure_smem(cpn::softmax_pool_reg_kernel<NVV, true>, smem));   \
            CPN_CUDA(launch_pdl(cpn::softmax_pool_reg_kernel<NVV, true>, grid, dim3(256), smem, st, 1, z, mode, partial,  \
                                partial_arg, HW, chunks));                                                            \
        } else {                                                                                                       \
            CPN_CUDA(launch_pdl(cpn::softmax_pool_reg_kernel<NVV, false>, grid, dim3(256), smem, st, 1, z, mode, partial, \
                                static_cast<int*>(nullptr), HW, chunks));                                              \
        }                                                                                                              \
        break;
            switch (nv) {
                CPN_SMP_CASE(1) CPN_SMP_CASE(2) CPN_SMP_CASE(3) CPN_SMP_CASE(4)
                CPN_SMP_CASE(5) CPN_SMP_CASE(6) CPN_SMP_CASE(7) CPN_SMP_CASE(8)
                case 16:  // P = 2048 (ResNet-50): one token in flight per warp
                    if (arg) {
                        CPN_CUDA(ensure_smem(cpn::softmax_pool_reg_kernel<16, true, 1>, smem));
                        CPN_CUDA(launch_pdl(cpn::softmax_pool_reg_kernel<16, true, 1>, grid, dim3(256), smem, st, 1, z, mode,
                                            partial, partial_arg, HW, chunks));
                    } else {
                        CPN_CUDA(ensure_smem(cpn::softmax_pool_reg_kernel<16, false, 1>, smem));
                        CPN_CUDA(launch_pdl(cpn::softmax_pool_reg_kernel<16, false, 1>, grid, dim3(256), smem, st, 1, z, mode,
                                            partial, static_cast<int*>(nullptr), HW, chunks));
                    }
                    break;
            }
#undef CPN_SMP_CASE
            CPN_CUDA(cudaGetLastError());
            return 0;
        });
    }
    if (gumbel == nullptr && !hard && map != nullptr && !arg && P % 128 == 0 && P <= 1024) {
        // plain softmax with the map written: staged through shared memory so that the map rows go out coalesced
        const int nv = P / 128;
        const size_t msmem = (32 + 8) * static_cast<size_t>(P) * sizeof(float);
        return submit(plan, stream, [=](cudaStream_t st) -> int {
#define CPN_SMM_CASE(NVV)                                                                                            \
    case NVV:                                                                                                        \
        CPN_CUDA(ensure_smem(cpn::softmax_pool_map_kernel<NVV>, msmem));                                             \
        CPN_CUDA(launch_pdl(cpn::softmax_pool_map_kernel<NVV>, grid, dim3(256), msmem, st, 1, z, mode, map, partial, HW,  \
                            chunks));                                                                               \
        break;
            switch (nv) {
                CPN_SMM_CASE(1) CPN_SMM_CASE(2) CPN_SMM_CASE(3) CPN_SMM_CASE(4)
                CPN_SMM_CASE(5) CPN_SMM_CASE(6) CPN_SMM_CASE(7) CPN_SMM_CASE(8)
            }
#undef CPN_SMM_CASE
            CPN_CUDA(cudaGetLastError());
            return 0;
        });
    }
    int* parg = arg ? partial_arg : nullptr;
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        CPN_CUDA(ensure_smem(cpn::softmax_pool_kernel, smem));
        cpn::softmax_pool_kernel<<<grid, 256, smem, st>>>(z, gumbel, tau, hard, mode, map, partial, parg, HW, P, chunks);
        CPN_CUDA(cudaGetLastError());
        return 0;
    });
}

int cpn_softmax_pool(cpn_plan* plan, const float* z, const float* gumbel, float tau, int hard, int mode, float* map,
                     float* partial, int B, int HW, int P, int chunks, void* stream) {
    return cpn_softmax_pool_arg(plan, z, gumbel, tau, hard, mode, map, partial, nullptr, B, HW, P, chunks, stream);
}

int cpn_pool_argmax_reduce(cpn_plan* plan, const float* partial, const int* partial_arg, int chunks, int* arg_out, int B,
                           int P, void* stream) {
    if (partial == nullptr || partial_arg == nullptr || arg_out == nullptr || chunks <= 0 || B <= 0 || B > 65535)
        return fail("cpn_pool_argmax_reduce: bad argument");
    const dim3 grid((P + 127) / 128, B);
    set_info("pool_argmax", 0.0, 8.0 * B * chunks * P + 4.0 * B * P);
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        cpn::pool_argmax_reduce_kernel<<<grid, 128, 0, st>>>(partial, partial_arg, chunks, arg_out, P);
        CPN_CUDA(cudaGetLastError());
        return 0;
    });
}

int cpn_topk_update(cpn_plan* plan, const float* pooled, const int* argmax, int B, int P, int k, long long image_base,
                    float* top_val, long long* top_img, int* top_tok, int* top_cnt, void* stream) {
    if (pooled == nullptr || argmax == nullptr || top_val == nullptr || top_img == nullptr || top_tok == nullptr ||
        top_cnt == nullptr)
        return fail("cpn_topk_update: missing operand");
    if (B <= 0 || P <= 0 || k <= 0) return fail("cpn_topk_update: bad shape B=%d P=%d k=%d", B, P, k);
    set_info("topk_update", 0.0, 8.0 * B * P + 16.0 * k * P);
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        cpn::topk_update_kernel<<<(P + 127) / 128, 128, 0, st>>>(pooled, argmax, B, P, k, image_base, top_val, top_img,
                                                                 top_tok, top_cnt);
        CPN_CUDA(cudaGetLastError());
        return 0;
    });
}

int cpn_eval_batch(cpn_plan* plan, const float* pooled, const float* out, const float* w, const long long* ys,
                   float mult, float thr, int B, int P, int K, int* pred, float* pred_prob, float* max_score,
                   int* any_size, int* pred_size, int* anz, int* top1, int* top5, unsigned long long* cm,
                   unsigned char* relevant, int* protos_per_class, void* stream) {
    if (pooled == nullptr || out == nullptr || w == nullptr || pred == nullptr || pred_prob == nullptr ||
        max_score == nullptr || any_size == nullptr || pred_size == nullptr || anz == nullptr || top1 == nullptr ||
        top5 == nullptr || relevant == nullptr || protos_per_class == nullptr)
        return fail("cpn_eval_batch: missing operand");
    if (B <= 0 || P <= 0 || K <= 0) return fail("cpn_eval_batch: bad shape B=%d P=%d K=%d", B, P, K);
    if (P > 200 * 1024) return fail("cpn_eval_batch: P=%d too large", P);
    cpn::EvalParams ep;
    ep.pooled = pooled; ep.out = out; ep.w = w; ep.ys = ys; ep.mult = mult; ep.thr = thr;
    ep.B = B; ep.P = P; ep.K = K; ep.pred = pred; ep.pred_prob = pred_prob; ep.max_score = max_score;
    ep.any_size = any_size; ep.pred_size = pred_size; ep.anz = anz; ep.top1 = top1; ep.top5 = top5; ep.cm = cm;
    ep.relevant = relevant;
    set_info("eval_batch", 0.0, 4.0 * B * K * P + 4.0 * B * (P + K) + 2.0 * K * P);
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        CPN_CUDA(cudaMemsetAsync(relevant, 0, static_cast<size_t>(K) * P, st));  // per-batch statistic
        CPN_CUDA(ensure_smem(cpn::eval_batch_kernel, P));
        cpn::eval_batch_kernel<<<B, cpn::EVAL_THREADS, static_cast<size_t>(P), st>>>(ep);
        CPN_CUDA(cudaGetLastError());
        cpn::eval_class_count_kernel<<<K, 128, 0, st>>>(relevant, protos_per_class, P);
        CPN_CUDA(cudaGetLastError());
        return 0;
    });
}

int cpn_head_tail(cpn_plan* plan, const float* partial, int chunks, int mode, int inference, int use_ste,
                  int max_count, int encoding, const float* enc_w0, const float* enc_w1, const float* enc_w2,
                  const float* cls_w, const float* cls_b, float* pooled_out, float* logits_out, int B, int P, int D,
                  int K, void* stream) {
    cpn::TailParams tp;
    tp.partial = partial; tp.chunks = chunks; tp.mode = mode; tp.inference = inference; tp.use_ste = use_ste;
    tp.max_count = max_count; tp.encoding = encoding; tp.enc_w0 = enc_w0; tp.enc_w1 = enc_w1; tp.enc_w2 = enc_w2;
    tp.cls_w = cls_w; tp.cls_b = cls_b; tp.pooled_out = pooled_out; tp.logits_out = logits_out;
    tp.P = P; tp.D = D; tp.K = K;
    if (mode == CPN_POOL_SUM) {
        if (max_count <= 0) return fail("cpn_head_tail: max_count must be positive");
        if ((encoding == CPN_ENC_ONEHOT || encoding == CPN_ENC_LINEAR) && D % P != 0)
            return fail("cpn_head_tail: D=%d not a multiple of P=%d", D, P);
        if ((encoding == CPN_ENC_LINEAR || encoding == CPN_ENC_LINEAR_FULL || encoding == CPN_ENC_BILINEAR) &&
            enc_w0 == nullptr)
            return fail("cpn_head_tail: encoding %d needs weights", encoding);
        if (encoding == CPN_ENC_BILINEAR && (enc_w1 == nullptr || enc_w2 == nullptr))
            return fail("cpn_head_tail: bilinear needs W and V");
        if ((encoding == CPN_ENC_IDENTITY || encoding == CPN_ENC_NONE) && D != P)
            return fail("cpn_head_tail: identity encoding needs D == P");
    } else if (D != P) {
        return fail("cpn_head_tail: max pooling needs D == P");
    }
    const size_t smem = (static_cast<size_t>(P) + 2 * static_cast<size_t>(D)) * sizeof(float);
    if (smem > 200 * 1024) return fail("cpn_head_tail: P=%d D=%d too large", P, D);
    // class slices per image: enough CTAs to fill the GPU when the batch alone does not (64 images on 148 SMs)
    int ks = (4 * num_sms() + B - 1) / B;
    ks = ks < 1 ? 1 : (ks > 8 ? 8 : ks);
    if (ks > (K + 7) / 8) ks = (K + 7) / 8;  // at least one class per warp
    if (D > 4096) ks = 1;                    // dense count encodings are not worth repeating
    set_info("head_tail", 2.0 * B * D * K, 4.0 * B * (1.0 * chunks * P + 1.0 * D * K));
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        CPN_CUDA(ensure_smem(cpn::head_tail_kernel, smem));
        CPN_CUDA(launch_pdl(cpn::head_tail_kernel, dim3(B, ks), dim3(256), smem, st, 1, tp));
        CPN_CUDA(cudaGetLastError());
        return 0;
    });
}

// ------------------------------------------------------------------------------------------------ layout / ResNet helpers
int cpn_nhwc_to_nchw(cpn_plan* plan, const float* x, float* out, int B, int HW, int C, void* stream) {
    const dim3 grid((HW + 31) / 32, (C + 31) / 32, B);
    set_info("nhwc_to_nchw", 0.0, 8.0 * B * HW * C);
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        cpn::transpose_last2_kernel<<<grid, 256, 0, st>>>(x, out, HW, C);
        CPN_CUDA(cudaGetLastError());
        return 0;
    });
}

int cpn_nchw_to_nhwc(cpn_plan* plan, const float* x, float* out, int B, int HW, int C, void* stream) {
    const dim3 grid((C + 31) / 32, (HW + 31) / 32, B);
    set_info("nchw_to_nhwc", 0.0, 8.0 * B * HW * C);
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        cpn::transpose_last2_kernel<<<grid, 256, 0, st>>>(x, out, C, HW);
        CPN_CUDA(cudaGetLastError());
        return 0;
    });
}

static int gather_grid(long long total) {
    long long blocks = (total + 255) / 256;
    const long long cap = 16LL * num_sms();
    return static_cast<int>(blocks > cap ? cap : (blocks < 1 ? 1 : blocks));
}

int cpn_im2col_bf16(cpn_plan* plan, const void* x, void* out, int B, int H, int W, int C, int k, int stride,
                    int pad, int Kpad, void* stream) {
    if (C % 8 != 0 || Kpad % 8 != 0 || Kpad < k * k * C) return fail("cpn_im2col_bf16: bad C=%d Kpad=%d", C, Kpad);
    const int Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
    const int grid = gather_grid(static_cast<long long>(B) * Ho * Wo * (Kpad / 8));
    set_info("im2col_bf16", 0.0, 4.0 * B * Ho * Wo * Kpad);
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        cpn::im2col_bf16_kernel<<<grid, 256, 0, st>>>(static_cast<const uint4*>(x), static_cast<uint4*>(out), B, H, W, C,
                                                      k, stride, pad, Ho, Wo, Kpad);
        CPN_CUDA(cudaGetLastError());
        return 0;
    });
}

int cpn_im2col_nchw_f32(cpn_plan* plan, const float* x, void* out, void* out_lo, int B, int C, int H, int W, int k,
                        int stride, int pad, int Kpad, void* stream) {
    if (Kpad < k * k * C || Kpad % 8 != 0) return fail("cpn_im2col_nchw_f32: Kpad must be a multiple of 8 and >= k*k*C");
    if ((reinterpret_cast<uintptr_t>(out) & 15) != 0 || (reinterpret_cast<uintptr_t>(out_lo) & 15) != 0)
        return fail("cpn_im2col_nchw_f32: outputs must be 16-byte aligned");
    const int Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
    const int grid = gather_grid(static_cast<long long>(B) * Ho * Wo * (Kpad / 8));
    const size_t tab_bytes = 2 * static_cast<size_t>(Kpad) * sizeof(int);
    set_info("im2col_nchw_f32", 0.0, 4.0 * B * C * H * W + 2.0 * B * Ho * Wo * Kpad * (out_lo != nullptr ? 2 : 1));
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        cpn::im2col_nchw_f32_kernel<<<grid, 256, tab_bytes, st>>>(x, static_cast<__nv_bfloat16*>(out),
                                                          static_cast<__nv_bfloat16*>(out_lo), B, C, H, W, k, stride,
                                                          pad, Ho, Wo, Kpad);
        CPN_CUDA(cudaGetLastError());
        return 0;
    });
}

int cpn_maxpool3s2(cpn_plan* plan, const float* x, float* out_f32, void* out_hi, void* out_lo, int B, int H, int W,
                   int C, void* stream) {
    const int Ho = (H + 2 - 3) / 2 + 1, Wo = (W + 2 - 3) / 2 + 1;
    const int grid = gather_grid(static_cast<long long>(B) * Ho * Wo * C);
    set_info("maxpool3s2", 0.0, 4.0 * B * H * W * C + 1.0 * B * Ho * Wo * C * ((out_f32 != nullptr) * 4 + (out_hi != nullptr) * 2 + (out_lo != nullptr) * 2));
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        cpn::maxpool3s2_kernel<<<grid, 256, 0, st>>>(x, out_f32, static_cast<__nv_bfloat16*>(out_hi),
                                                     static_cast<__nv_bfloat16*>(out_lo), B, H, W, C, Ho, Wo);
        CPN_CUDA(cudaGetLastError());
        return 0;
    });
}

int cpn_maxpool3s2_bf16(cpn_plan* plan, const void* x, void* out, int B, int H, int W, int C, void* stream) {
    if (x == nullptr || out == nullptr) return fail("cpn_maxpool3s2_bf16: missing operand");
    if (C % 8 != 0 || ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15) != 0)
        return fail("cpn_maxpool3s2_bf16: C must be a multiple of 8 and the maps 16-byte aligned");
    const int Ho = (H + 2 - 3) / 2 + 1, Wo = (W + 2 - 3) / 2 + 1;
    const int grid = gather_grid(static_cast<long long>(B) * Ho * Wo * (C / 8));
    set_info("maxpool3s2", 0.0, 2.0 * B * H * W * C + 2.0 * B * Ho * Wo * C);
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        cpn::maxpool3s2_bf16_kernel<<<grid, 256, 0, st>>>(static_cast<const uint4*>(x), static_cast<uint4*>(out), B, H, W, C,
                                                          Ho, Wo);
        CPN_CUDA(cudaGetLastError());
        return 0;
    });
}

int cpn_subsample2_bf16(cpn_plan* plan, const void* x, void* out, int B, int H, int W, int C, void* stream) {
    if (C % 8 != 0) return fail("cpn_subsample2_bf16: C must be a multiple of 8");
    const int Ho = (H + 1) / 2, Wo = (W + 1) / 2;
    const int grid = gather_grid(static_cast<long long>(B) * Ho * Wo * (C / 8));
    set_info("subsample2_bf16", 0.0, 4.0 * B * Ho * Wo * C);
    return submit(plan, stream, [=](cudaStream_t st) -> int {
        cpn::subsample2_bf16_kernel<<<grid, 256, 0, st>>>(static_cast<const uint4*>(x), static_cast<uint4*>(out), B, H,
                                                          W, C, Ho, Wo);
        CPN_CUDA(cudaGetLastError());
        return 0;
    });
}

}  // extern "C"
