// cpn_softmax_pool_static: softmax over prototypes + spatial max / sum per image for callers that may not allocate
// (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). No counterpart in the
// reference. The arithmetic is cpn_softmax_pool's (pipnet/pipnet.py:96,
// 103-105 softmax + adaptive max pool; pipnet/count_pipnet.py:88 spatial sum) followed by the fold over chunks that
// cpn_head_tail does.
//
// cpn_softmax_pool makes the caller bring `partial [B, chunks, P]`; here that scratch is the file-scope device array
// below: CPN_SOFTMAX_POOL_STATIC_MAX_BYTES (4 MiB) in every CUDA context that loads the library, held until exit.
//   * In HOST code `g_pool_scratch` is only a stand-in for the symbol, not a device address. The address is looked up
//     with cudaGetSymbolAddress once per device, with that device current, and passed to the kernels as a parameter.
//   * All calls share the array and it keeps the previous call's contents. Nothing has to start at zero: the pool kernel
//     overwrites every (image, chunk, prototype) entry the fold kernel reads, and both use the same `chunks`.
//   * Calls that overlap on the GPU (different streams, or several host threads on one device) RACE on the array. Calls
//     on one stream are ordered by the stream and are safe.
//   * The entry point calls no cudaMalloc / cudaFree / allocating library routine and records nothing in a plan; it is
//     two plain kernel launches with at most 48 KB of dynamic shared memory (no function attribute needed).
#include "api_common.h"

#include <cmath>

using namespace cpn_api;

namespace {

constexpr size_t SCRATCH_FLOATS = CPN_SOFTMAX_POOL_STATIC_MAX_BYTES / sizeof(float);
constexpr int MAX_DEV = 64;
constexpr int MAX_CHUNKS = 64;
constexpr int TOKENS_PER_CHUNK = 32;

static __device__ __align__(256) float g_pool_scratch[SCRATCH_FLOATS];

// Device address of g_pool_scratch on the current device (one lookup per device).
int scratch_address(float** out) {
    static std::mutex mu;
    static float* addr[MAX_DEV];
    int dev = -1;
    CPN_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEV) return fail("cpn_softmax_pool_static: device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (addr[dev] == nullptr) {
        float* p = nullptr;
        CPN_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_pool_scratch));
        if (p == nullptr) return fail("cpn_softmax_pool_static: cudaGetSymbolAddress returned a null address");
        addr[dev] = p;
    }
    *out = addr[dev];
    return 0;
}

__device__ __forceinline__ float warp_reduce_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_reduce_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// grid = (chunks, B): a CTA owns a contiguous range of one image's tokens, one warp per token at a time.
// z [B, HW, P] fp32; partial [B, chunks, P] (the static scratch).
__global__ void softmax_pool_static_kernel(const float* __restrict__ z, int mode, float* __restrict__ partial, int HW,
                                           int P, int chunks) {
    extern __shared__ float s_acc[];  // [warps][P]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
    const int b = blockIdx.y, chunk = blockIdx.x;
    const int per_chunk = (HW + chunks - 1) / chunks;
    const int t0 = chunk * per_chunk, t1 = min(HW, t0 + per_chunk);
    float* mine = s_acc + warp * P;
    for (int p = lane; p < P; p += 32) mine[p] = 0.f;  // probabilities are >= 0: neutral for max and for sum
    __syncwarp();
    for (int t = t0 + warp; t < t1; t += wpb) {
        const float* zr = z + (static_cast<size_t>(b) * HW + t) * P;
        float m = -INFINITY;
        for (int p = lane; p < P; p += 32) m = fmaxf(m, __ldg(zr + p));
        m = warp_reduce_max(m);
        float s = 0.f;
        for (int p = lane; p < P; p += 32) s += expf(__ldg(zr + p) - m);
        s = warp_reduce_sum(s);
        const float inv_s = 1.0f / s;
        for (int p = lane; p < P; p += 32) {
            const float y = expf(__ldg(zr + p) - m) * inv_s;
            mine[p] = mode == CPN_POOL_MAX ? fmaxf(mine[p], y) : mine[p] + y;
        }
    }
    __syncthreads();
    float* out = partial + (static_cast<size_t>(b) * chunks + chunk) * P;
    for (int p = threadIdx.x; p < P; p += blockDim.x) {
        float a = s_acc[p];
        for (int w = 1; w < wpb; ++w) a = mode == CPN_POOL_MAX ? fmaxf(a, s_acc[w * P + p]) : a + s_acc[w * P + p];
        out[p] = a;
    }
}

// grid = (ceil(P / 128), B): pooled[b, p] = max / sum over the chunks, in chunk order (deterministic).
__global__ void pool_fold_static_kernel(const float* __restrict__ partial, int mode, float* __restrict__ pooled, int P,
                                        int chunks) {
    const int b = blockIdx.y, p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const float* src = partial + static_cast<size_t>(b) * chunks * P + p;
    float a = src[0];
    for (int c = 1; c < chunks; ++c) {
        const float o = src[static_cast<size_t>(c) * P];
        a = mode == CPN_POOL_MAX ? fmaxf(a, o) : a + o;
    }
    pooled[static_cast<size_t>(b) * P + p] = a;
}

}  // namespace

extern "C" int cpn_softmax_pool_static(const float* z, int mode, float* pooled, int B, int HW, int P, void* stream) {
    if (z == nullptr || pooled == nullptr) return fail("cpn_softmax_pool_static: missing operand");
    if (mode != CPN_POOL_MAX && mode != CPN_POOL_SUM) return fail("cpn_softmax_pool_static: mode %d", mode);
    if (B <= 0 || B > 65535 || HW <= 0 || P <= 0) return fail("cpn_softmax_pool_static: bad B=%d HW=%d P=%d", B, HW, P);
    if (P > 3072) return fail("cpn_softmax_pool_static: P=%d exceeds the 3072 prototypes that fit 48 KB of shared memory", P);
    const size_t per_chunk_floats = static_cast<size_t>(B) * P;
    if (per_chunk_floats > SCRATCH_FLOATS)
        return fail("cpn_softmax_pool_static: B * P = %zu floats exceeds the static scratch of %zu floats (%u bytes); use "
                    "cpn_softmax_pool with caller-owned scratch", per_chunk_floats, SCRATCH_FLOATS,
                    static_cast<unsigned>(CPN_SOFTMAX_POOL_STATIC_MAX_BYTES));
    int chunks = (HW + TOKENS_PER_CHUNK - 1) / TOKENS_PER_CHUNK;
    if (chunks > MAX_CHUNKS) chunks = MAX_CHUNKS;
    const size_t fit = SCRATCH_FLOATS / per_chunk_floats;
    if (static_cast<size_t>(chunks) > fit) chunks = static_cast<int>(fit);
    float* scratch = nullptr;
    if (int rc = scratch_address(&scratch)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int threads = P <= 1536 ? 256 : 128;  // [warps][P] floats <= 48 KB
    const size_t smem = static_cast<size_t>(threads / 32) * P * sizeof(float);
    softmax_pool_static_kernel<<<dim3(chunks, B), threads, smem, st>>>(z, mode, scratch, HW, P, chunks);
    CPN_CUDA(cudaGetLastError());
    pool_fold_static_kernel<<<dim3((P + 127) / 128, B), 128, 0, st>>>(scratch, mode, pooled, P, chunks);
    CPN_CUDA(cudaGetLastError());
    return 0;
}
