// Host-side helpers shared by the translation units of libcpnb200.so (api.cu defines them): error reporting, the plan
// recorder, TMA descriptor encoding, per-device caches and launch wrappers. Nothing here crosses the C ABI.
#pragma once
#include "../../include/cpn_b200.h"

#include <cuda.h>
#include <cuda_runtime.h>

#include <cstring>
#include <functional>
#include <map>
#include <mutex>
#include <utility>

namespace cpn_api {

int fail(const char* fmt, ...);

#define CPN_CUDA(expr)                                                                                              \
    do {                                                                                                            \
        cudaError_t _e = (expr);                                                                                    \
        if (_e != cudaSuccess)                                                                                      \
            return ::cpn_api::fail("%s: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__);            \
    } while (0)

using Launch = std::function<int(cudaStream_t)>;

// Algorithmic work of the op being submitted (cpn_plan_run_timed's per-op report); every entry point sets it first.
void set_info(const char* name, double flops, double bytes, const char* fmt = nullptr, ...);
// Launch now (plan == nullptr) or record into the plan.
int submit(cpn_plan* plan, void* stream, Launch fn);

// bf16 row-major [rows, K] matrix, box = 64 (K) x box_rows, 128-byte swizzle; out-of-bounds reads give zeros.
int make_tmap(CUtensorMap* m, const void* ptr, long long rows, int K, int box_rows);
// Row-major [rows, cols] output (bf16 or fp32) for the TMA epilogues: box = 128 bytes x 32 rows, 128-byte swizzle.
int make_tmap_out(CUtensorMap* m, const void* ptr, long long rows, int cols, bool bf16);
// Generic tiled descriptor (rank <= 5): dims / strides in elements-of-dtype and bytes as cuTensorMapEncodeTiled wants.
int make_tmap_tiled(CUtensorMap* m, CUtensorMapDataType dtype, int rank, const void* ptr, const cuuint64_t* dims,
                    const cuuint64_t* strides_bytes, const cuuint32_t* box, CUtensorMapSwizzle swizzle);

// dwconv7_stats_kernel (api.cu): the taps-in-registers depthwise conv, ConvNeXt-tiny widths; out_lo != nullptr = fp32 mode.
int dwconv7_stats_v1(cpn_plan* plan, const float* x, const float* w49, const float* bias, void* out_bf16, void* out_lo,
                     float* stats, int B, int H, int W, int C, void* stream);

// Three-pass bf16 hi/lo GEMM with the TMA epilogue (api_gemm_split.cu): epi is a cpn::GemmEpi code (LN+GELU with hi/lo
// bf16 output planes, scale+residual fp32, bias fp32). cpn_gemm_bf16 routes its split cases here.
bool gemm_split_supported(int N, int K, int epi);
int gemm_split(cpn_plan* plan, int epi, const void* A, const void* A_lo, const void* Bw, const void* Bw_lo, int M, int N,
               int K, const float* bias, const float* scale, const float* stats, int nparts, const float* colsum,
               float eps, void* out, void* out_lo, void* stream);

int cur_device();
int num_sms();
bool pdl_enabled();

// cudaFuncAttributeMaxDynamicSharedMemorySize, once per (kernel, device) and again only if a launch needs more. The
// attribute belongs to a device, and one process may drive several (nn.DataParallel threads: reference main.py:118).
template <class Kern>
cudaError_t ensure_smem(Kern kern, size_t bytes) {
    if (bytes <= 48 * 1024) return cudaSuccess;
    static std::mutex mu;
    static std::map<std::pair<const void*, int>, size_t> done;
    const std::pair<const void*, int> key(reinterpret_cast<const void*>(kern), cur_device());
    std::lock_guard<std::mutex> lock(mu);
    auto it = done.find(key);
    if (it != done.end() && it->second >= bytes) return cudaSuccess;
    const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bytes));
    if (e == cudaSuccess) done[key] = bytes;
    return e;
}

// Launch with an optional thread-block cluster and (CPN_PDL=1) programmatic dependent launch: the kernel may become
// resident and run its on-chip prologue while the previous kernel of the stream drains. ONLY for kernels that execute
// ptx::pdl_wait() before their first global access. Measured on B200 (profiles/r01_pdl_ab.md): the 48-launch step is
// 4.27 ms with it and 4.24 ms without, so it is off by default.
template <class... KArgs, class... Args>
cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, int cluster,
                       Args... args) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    unsigned n = 0;
    if (cluster > 1) {
        attr[n].id = cudaLaunchAttributeClusterDimension;
        attr[n].val.clusterDim.x = cluster;
        attr[n].val.clusterDim.y = 1;
        attr[n].val.clusterDim.z = 1;
        ++n;
    }
    if (pdl_enabled()) {
        attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[n].val.programmaticStreamSerializationAllowed = 1;
        ++n;
    }
    cfg.attrs = attr;
    cfg.numAttrs = n;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

}  // namespace cpn_api
