// CUDA-core kernels of the ConvNeXt feature path (channels-last, fp32 residual stream):
//   stem_ln_kernel       conv 4x4 stride 4 (3 -> 96) + bias + per-token LayerNorm        NCHW fp32 -> NHWC fp32
//   dwconv7_ln_kernel    depthwise 7x7 (pad 3) + bias + per-token LayerNorm               NHWC fp32 -> token-major bf16
//   ln_im2col2_kernel    per-token LayerNorm + 2x2 patch gather (stride 1 or 2)           NHWC fp32 -> im2col bf16
// Reference arithmetic: torchvision/models/convnext.py:31-36 (LayerNorm2d), :51-57 (CNBlock head), :120-131 (stem),
// :146-150 (downsample); stride surgery features/convnext_features.py:5-15.
#pragma once
#include "device_utils.cuh"

namespace cpn {

// ------------------------------------------------------------------------------------------------ stem
// One warp handles 8 horizontally adjacent output tokens; lane l owns output channels l, l+32, l+64.
// Weights [96][48] (k = c*16 + kh*4 + kw, the torch conv weight layout) live in registers for the whole kernel.
constexpr int STEM_COUT = 96;
constexpr int STEM_K = 48;
constexpr int STEM_TOK = 8;
constexpr int STEM_WARPS = 8;

__global__ void __launch_bounds__(STEM_WARPS * 32)
stem_ln_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
               const float* __restrict__ ln_w, const float* __restrict__ ln_b, float* __restrict__ out, int B, int H,
               int W, int Ho, int Wo, float eps) {
    __shared__ __align__(16) float s_in[STEM_WARPS][STEM_K][STEM_TOK];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float wr[STEM_K][3];
#pragma unroll
    for (int k = 0; k < STEM_K; ++k)
#pragma unroll
        for (int i = 0; i < 3; ++i) wr[k][i] = __ldg(w + (lane + 32 * i) * STEM_K + k);
    float br[3], gw[3], gb[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        br[i] = __ldg(bias + lane + 32 * i);
        gw[i] = __ldg(ln_w + lane + 32 * i);
        gb[i] = __ldg(ln_b + lane + 32 * i);
    }
    const int groups_per_row = (Wo + STEM_TOK - 1) / STEM_TOK;
    const long long num_groups = static_cast<long long>(B) * Ho * groups_per_row;
    for (long long g = static_cast<long long>(blockIdx.x) * STEM_WARPS + warp; g < num_groups;
         g += static_cast<long long>(gridDim.x) * STEM_WARPS) {
        const int gx = static_cast<int>(g % groups_per_row);
        const long long t = g / groups_per_row;
        const int ho = static_cast<int>(t % Ho);
        const int b = static_cast<int>(t / Ho);
        const int wo0 = gx * STEM_TOK;
        // gather the 3 x 4 input row segments (32 contiguous floats each) into s_in[k][token]
        __syncwarp();
#pragma unroll
        for (int seg = 0; seg < 12; ++seg) {
            const int c = seg >> 2, kh = seg & 3;
            const int col = 4 * wo0 + lane;
            float v = 0.0f;
            if (col < W) v = __ldg(x + ((static_cast<size_t>(b) * 3 + c) * H + (4 * ho + kh)) * W + col);
            s_in[warp][c * 16 + kh * 4 + (lane & 3)][lane >> 2] = v;
        }
        __syncwarp();
        float acc[STEM_TOK][3];
#pragma unroll
        for (int tkn = 0; tkn < STEM_TOK; ++tkn)
#pragma unroll
            for (int i = 0; i < 3; ++i) acc[tkn][i] = br[i];
#pragma unroll
        for (int k = 0; k < STEM_K; ++k) {
            const float4 a0 = *reinterpret_cast<const float4*>(&s_in[warp][k][0]);
            const float4 a1 = *reinterpret_cast<const float4*>(&s_in[warp][k][4]);
            const float in8[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
            for (int tkn = 0; tkn < STEM_TOK; ++tkn)
#pragma unroll
                for (int i = 0; i < 3; ++i) acc[tkn][i] = fmaf(in8[tkn], wr[k][i], acc[tkn][i]);
        }
#pragma unroll
        for (int tkn = 0; tkn < STEM_TOK; ++tkn) {
            const float mean = warp_sum(acc[tkn][0] + acc[tkn][1] + acc[tkn][2]) * (1.0f / STEM_COUT);
            const float d0 = acc[tkn][0] - mean, d1 = acc[tkn][1] - mean, d2 = acc[tkn][2] - mean;
            const float var = warp_sum(d0 * d0 + d1 * d1 + d2 * d2) * (1.0f / STEM_COUT);
            const float rstd = rsqrtf(var + eps);
            if (wo0 + tkn < Wo) {
                float* o = out + ((static_cast<size_t>(b) * Ho + ho) * Wo + wo0 + tkn) * STEM_COUT;
                o[lane] = d0 * rstd * gw[0] + gb[0];
                o[lane + 32] = d1 * rstd * gw[1] + gb[1];
                o[lane + 64] = d2 * rstd * gw[2] + gb[2];
            }
        }
    }
}

// Second stem kernel (used when W % 4 == 0): a persistent CTA walks output rows (b, ho). The 12 input row segments
// of a row (3 channels x 4 kernel rows, 4*Wo floats each) arrive by cp.async one row ahead (double buffer), the
// weights sit in shared memory as [c*4+kh][n][kw] so that a lane's four kw taps are one 16-byte load, and a thread
// computes 4 tokens x 6 channels (channels cg, cg+16, ..., cg+80): per (c,kh) step 4 broadcast input loads + 6 weight
// loads feed 96 FMAs. LayerNorm (two passes, as the reference) reduces over the 16 lanes that share a token.
constexpr int STEM2_THREADS = 256;

__device__ __forceinline__ float half_warp_sum(float v) {
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__global__ void __launch_bounds__(STEM2_THREADS)
stem_ln_rows_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                    const float* __restrict__ ln_w, const float* __restrict__ ln_b, float* __restrict__ out, int B,
                    int H, int W, int Ho, int Wo, float eps) {
    extern __shared__ __align__(16) float stem_smem[];
    float* ws = stem_smem;                      // [12][96][4]
    const int rowf = 12 * 4 * Wo;               // floats of one staged output row
    float* inb = stem_smem + 12 * 96 * 4;       // [2][12][4 * Wo]
    const int tid = threadIdx.x;
    for (int idx = tid; idx < STEM_COUT * STEM_K; idx += STEM2_THREADS) {
        const int n = idx / STEM_K, k = idx - n * STEM_K;
        ws[((k >> 2) * STEM_COUT + n) * 4 + (k & 3)] = __ldg(w + idx);
    }
    const int cg = tid & 15, tq = tid >> 4;
    float br[6], gw[6], gb[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        br[i] = __ldg(bias + i * 16 + cg);
        gw[i] = __ldg(ln_w + i * 16 + cg);
        gb[i] = __ldg(ln_b + i * 16 + cg);
    }
    ptx::pdl_wait();  // weights above are constants of the plan; the activations come from the previous kernel
    ptx::pdl_launch_dependents();
    const int nrows = B * Ho;
    auto issue = [&](int r, int buf) {
        const int b = r / Ho, ho = r - b * Ho;
        const uint32_t dst0 = ptx::smem_u32(inb + buf * rowf);
        for (int j = tid; j < 12 * Wo; j += STEM2_THREADS) {
            const int seg = j / Wo, q = j - seg * Wo;
            const int c = seg >> 2, kh = seg & 3;
            const float* src = x + ((static_cast<size_t>(b) * 3 + c) * H + (4 * ho + kh)) * W + 4 * q;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst0 + (seg * 4 * Wo + 4 * q) * 4), "l"(src)
                         : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    int r = blockIdx.x;
    if (r < nrows) issue(r, 0);
    int buf = 0;
    for (; r < nrows; r += gridDim.x, buf ^= 1) {
        if (r + static_cast<int>(gridDim.x) < nrows) {
            issue(r + gridDim.x, buf ^ 1);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncthreads();  // this row's inputs (and, the first time, the weights) are visible to every thread
        const float* in = inb + buf * rowf;
        float* orow = out + static_cast<size_t>(r) * Wo * STEM_COUT;
        for (int t0 = 0; t0 < Wo; t0 += 64) {
            int tok[4];
#pragma unroll
            for (int tt = 0; tt < 4; ++tt) tok[tt] = min(t0 + tq * 4 + tt, Wo - 1);
            float acc[4][6];
#pragma unroll
            for (int tt = 0; tt < 4; ++tt)
#pragma unroll
                for (int i = 0; i < 6; ++i) acc[tt][i] = br[i];
#pragma unroll 2
            for (int ckh = 0; ckh < 12; ++ckh) {
                float4 a[4], wv[6];
#pragma unroll
                for (int tt = 0; tt < 4; ++tt) a[tt] = *reinterpret_cast<const float4*>(in + ckh * 4 * Wo + 4 * tok[tt]);
#pragma unroll
                for (int i = 0; i < 6; ++i) wv[i] = *reinterpret_cast<const float4*>(ws + (ckh * STEM_COUT + i * 16 + cg) * 4);
#pragma unroll
                for (int tt = 0; tt < 4; ++tt)
#pragma unroll
                    for (int i = 0; i < 6; ++i) {
                        acc[tt][i] = fmaf(a[tt].x, wv[i].x, acc[tt][i]);
                        acc[tt][i] = fmaf(a[tt].y, wv[i].y, acc[tt][i]);
                        acc[tt][i] = fmaf(a[tt].z, wv[i].z, acc[tt][i]);
                        acc[tt][i] = fmaf(a[tt].w, wv[i].w, acc[tt][i]);
                    }
            }
#pragma unroll
            for (int tt = 0; tt < 4; ++tt) {
                float s = 0.f;
#pragma unroll
                for (int i = 0; i < 6; ++i) s += acc[tt][i];
                const float mean = half_warp_sum(s) * (1.0f / STEM_COUT);
                float ss = 0.f;
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    acc[tt][i] -= mean;
                    ss = fmaf(acc[tt][i], acc[tt][i], ss);
                }
                const float rstd = rsqrtf(half_warp_sum(ss) * (1.0f / STEM_COUT) + eps);
                if (t0 + tq * 4 + tt < Wo) {
                    float* o = orow + static_cast<size_t>(tok[tt]) * STEM_COUT + cg;
#pragma unroll
                    for (int i = 0; i < 6; ++i) o[i * 16] = acc[tt][i] * rstd * gw[i] + gb[i];
                }
            }
        }
        __syncthreads();  // everyone is done with this buffer before the next iteration's prefetch overwrites it
    }
}

// ------------------------------------------------------------------------------------------------ dwconv7 + LN
// CTA = (column strip of TW tokens, row segment, image); thread = one channel pair; it marches down the input rows
// keeping the 7 in-flight output rows in registers, so every loaded input feeds 7*TW packed FMAs and the 49 weights
// of its two channels stay in registers. A finished output row is normalised across the CTA (channels = threads)
// and stored as bf16 (optionally with the bf16 residual "lo" plane for fp32 mode).
constexpr int DW_MAX_WARPS = 12;

// TW = 4 runs with <= 192 threads (C <= 384), TW = 2 with up to 384 threads (C = 768): the bound sets the register cap.
template <int TW>
__global__ void __launch_bounds__(TW == 2 ? DW_MAX_WARPS * 32 : 192)
dwconv7_ln_kernel(const float* __restrict__ x, const float* __restrict__ w49, const float* __restrict__ bias,
                  const float* __restrict__ ln_w, const float* __restrict__ ln_b, __nv_bfloat16* __restrict__ out_hi,
                  __nv_bfloat16* __restrict__ out_lo, int H, int W, int C, int rows_per_cta, float eps) {
    __shared__ __align__(16) float red[2][DW_MAX_WARPS][2 * TW];
    const int cp = threadIdx.x;            // channel pair
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nwarps = blockDim.x >> 5;
    const bool active = 2 * cp < C;
    const int c0 = active ? 2 * cp : 0;
    const int w0 = blockIdx.x * TW;
    const int h0 = blockIdx.y * rows_per_cta;
    const int nout = min(rows_per_cta, H - h0);
    const int b = blockIdx.z;
    const float inv_c = 1.0f / static_cast<float>(C);

    float2 wr[49];
#pragma unroll
    for (int k = 0; k < 49; ++k)
        wr[k] = active ? __ldg(reinterpret_cast<const float2*>(w49 + static_cast<size_t>(k) * C + c0))
                       : make_float2(0.f, 0.f);
    const float2 bs = active ? __ldg(reinterpret_cast<const float2*>(bias + c0)) : make_float2(0.f, 0.f);
    const float2 gw = active ? __ldg(reinterpret_cast<const float2*>(ln_w + c0)) : make_float2(0.f, 0.f);
    const float2 gb = active ? __ldg(reinterpret_cast<const float2*>(ln_b + c0)) : make_float2(0.f, 0.f);

    float2 acc[7][TW];
    const float* xb = x + static_cast<size_t>(b) * H * W * C + c0;
    const int nrows_in = nout + 6;
    int emitted = 0;

    for (int rb = 0; rb < nrows_in; rb += 7) {
#pragma unroll
        for (int j = 0; j < 7; ++j) {
            const int r = rb + j;
            if (r >= nrows_in) break;
            const int hi = h0 - 3 + r;
            float2 in[TW + 6];
            const bool row_ok = active && hi >= 0 && hi < H;
#pragma unroll
            for (int i = 0; i < TW + 6; ++i) {
                const int col = w0 - 3 + i;
                in[i] = (row_ok && col >= 0 && col < W)
                            ? __ldg(reinterpret_cast<const float2*>(xb + (static_cast<size_t>(hi) * W + col) * C))
                            : make_float2(0.f, 0.f);
            }
            // input row r feeds output row (r - kh) with filter row kh; kh == 0 starts that output row.
#pragma unroll
            for (int kh = 0; kh < 7; ++kh) {
                const int slot = (j - kh + 7) % 7;
#pragma unroll
                for (int t = 0; t < TW; ++t) {
                    float2 a = (kh == 0) ? bs : acc[slot][t];
#pragma unroll
                    for (int kw = 0; kw < 7; ++kw) a = ffma2(in[t + kw], wr[kh * 7 + kw], a);
                    acc[slot][t] = a;
                }
            }
            if (r >= 6) {
                // output row (r - 6) is complete; it sits in slot (j + 1) % 7
                const int slot = (j + 1) % 7;
                const int ho = h0 + r - 6;
                const int buf = emitted & 1;
                ++emitted;
                float part[2 * TW];
#pragma unroll
                for (int t = 0; t < TW; ++t) {
                    const float2 v = acc[slot][t];
                    part[2 * t] = warp_sum(v.x + v.y);
                    part[2 * t + 1] = warp_sum(v.x * v.x + v.y * v.y);
                }
                if (lane == 0) {
#pragma unroll
                    for (int i = 0; i < 2 * TW; ++i) red[buf][warp][i] = part[i];
                }
                __syncthreads();
                float tot[2 * TW];
#pragma unroll
                for (int i = 0; i < 2 * TW; ++i) tot[i] = 0.f;
                for (int wv = 0; wv < nwarps; ++wv) {
#pragma unroll
                    for (int i = 0; i < 2 * TW; ++i) tot[i] += red[buf][wv][i];
                }
#pragma unroll
                for (int t = 0; t < TW; ++t) {
                    const int wo = w0 + t;
                    if (active && wo < W) {
                        const float mean = tot[2 * t] * inv_c;
                        const float var = fmaxf(tot[2 * t + 1] * inv_c - mean * mean, 0.f);
                        const float rstd = rsqrtf(var + eps);
                        const float2 v = acc[slot][t];
                        const float y0 = (v.x - mean) * rstd * gw.x + gb.x;
                        const float y1 = (v.y - mean) * rstd * gw.y + gb.y;
                        const size_t o = ((static_cast<size_t>(b) * H + ho) * W + wo) * C + c0;
                        *reinterpret_cast<uint32_t*>(out_hi + o) = pack_bf16x2(y0, y1);
                        if (out_lo != nullptr)
                            *reinterpret_cast<uint32_t*>(out_lo + o) = pack_bf16x2(y0 - bf16_round(y0), y1 - bf16_round(y1));
                    }
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------ LN + 2x2 im2col
// One warp per input token: LayerNorm over C, then scatter the normalised bf16 row into the (up to four) im2col
// rows of the 2x2 / stride-s convolution that read it:  out[(b,ho,wo), (kh*2+kw)*C + c] = LN(x[b, ho*s+kh, wo*s+kw, c]).
constexpr int LN_MAX_PER_LANE = 24;  // C <= 768

__global__ void __launch_bounds__(256)
ln_im2col2_kernel(const float* __restrict__ x, const float* __restrict__ ln_w, const float* __restrict__ ln_b,
                  __nv_bfloat16* __restrict__ out_hi, __nv_bfloat16* __restrict__ out_lo, int B, int H, int W, int C,
                  int stride, int Ho, int Wo, float eps) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long token = static_cast<long long>(blockIdx.x) * (blockDim.x >> 5) + warp;
    const long long ntok = static_cast<long long>(B) * H * W;
    if (token >= ntok) return;
    const int w = static_cast<int>(token % W);
    const int h = static_cast<int>((token / W) % H);
    const int b = static_cast<int>(token / (static_cast<long long>(W) * H));
    const int per_lane = C >> 5;
    const float* xp = x + token * C;
    float v[LN_MAX_PER_LANE];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < LN_MAX_PER_LANE; ++i) {
        if (i < per_lane) {
            v[i] = __ldg(xp + lane + 32 * i);
            s += v[i];
        }
    }
    const float mean = warp_sum(s) / static_cast<float>(C);
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < LN_MAX_PER_LANE; ++i) {
        if (i < per_lane) {
            const float d = v[i] - mean;
            ss += d * d;
        }
    }
    const float rstd = rsqrtf(warp_sum(ss) / static_cast<float>(C) + eps);
#pragma unroll
    for (int i = 0; i < LN_MAX_PER_LANE; ++i) {
        if (i < per_lane) {
            const int c = lane + 32 * i;
            v[i] = (v[i] - mean) * rstd * __ldg(ln_w + c) + __ldg(ln_b + c);
        }
    }
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
        const int kh = kk >> 1, kw = kk & 1;
        const int hh = h - kh, ww = w - kw;
        if (hh < 0 || ww < 0 || (hh % stride) != 0 || (ww % stride) != 0) continue;
        const int ho = hh / stride, wo = ww / stride;
        if (ho >= Ho || wo >= Wo) continue;
        const size_t o = ((static_cast<size_t>(b) * Ho + ho) * Wo + wo) * (4 * static_cast<size_t>(C)) +
                         static_cast<size_t>(kk) * C;
#pragma unroll
        for (int i = 0; i < LN_MAX_PER_LANE; ++i) {
            if (i < per_lane) {
                const __nv_bfloat16 hv = __float2bfloat16_rn(v[i]);
                out_hi[o + lane + 32 * i] = hv;
                if (out_lo != nullptr) out_lo[o + lane + 32 * i] = __float2bfloat16_rn(v[i] - __bfloat162float(hv));
            }
        }
    }
}

// bf16-only variant (no lo plane), NV = ceil(C / 128) float4 per lane: a lane owns 4 consecutive channels per chunk,
// so a token row is read with 16-byte loads (512 B per warp and instruction) and each of its im2col copies is written
// with 8-byte stores; two tokens are in flight per warp and warps stride over the token list (persistent grid).
template <int NV>
__global__ void __launch_bounds__(256)
ln_im2col2_bf16_kernel(const float* __restrict__ x, const float* __restrict__ ln_w, const float* __restrict__ ln_b,
                       __nv_bfloat16* __restrict__ out, int B, int H, int W, int C, int stride, int Ho, int Wo,
                       float eps) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ntok = B * H * W;  // < 2^31 (checked by the host): 32-bit index arithmetic, no 64-bit divisions
    const int nwarps = gridDim.x * (blockDim.x >> 5);
    const int nchunk = C >> 2;  // float4 chunks per token
    float4 gw[NV], gb[NV];
    bool act[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        act[i] = lane + 32 * i < nchunk;
        gw[i] = act[i] ? __ldg(reinterpret_cast<const float4*>(ln_w) + lane + 32 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
        gb[i] = act[i] ? __ldg(reinterpret_cast<const float4*>(ln_b) + lane + 32 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    const float inv_c = 1.0f / static_cast<float>(C);
    for (int t = (blockIdx.x * (blockDim.x >> 5) + warp) * 2; t < ntok; t += 2 * nwarps) {
        const int ntk = t + 1 < ntok ? 2 : 1;
        float4 v[2][NV];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const float4* xr = reinterpret_cast<const float4*>(x + static_cast<size_t>(t + (u < ntk ? u : 0)) * C);
#pragma unroll
            for (int i = 0; i < NV; ++i) v[u][i] = act[i] ? __ldg(xr + lane + 32 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            if (u >= ntk) break;
            float s = 0.f;
#pragma unroll
            for (int i = 0; i < NV; ++i) s += (v[u][i].x + v[u][i].y) + (v[u][i].z + v[u][i].w);
            const float mean = warp_sum(s) * inv_c;
            float ss = 0.f;
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                if (act[i]) {
                    v[u][i].x -= mean; v[u][i].y -= mean; v[u][i].z -= mean; v[u][i].w -= mean;
                    ss += (v[u][i].x * v[u][i].x + v[u][i].y * v[u][i].y) + (v[u][i].z * v[u][i].z + v[u][i].w * v[u][i].w);
                }
            }
            const float rstd = rsqrtf(warp_sum(ss) * inv_c + eps);
            uint2 pk[NV];
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                pk[i].x = pack_bf16x2(fmaf(v[u][i].x * rstd, gw[i].x, gb[i].x), fmaf(v[u][i].y * rstd, gw[i].y, gb[i].y));
                pk[i].y = pack_bf16x2(fmaf(v[u][i].z * rstd, gw[i].z, gb[i].z), fmaf(v[u][i].w * rstd, gw[i].w, gb[i].w));
            }
            const int tok = t + u;
            const int row = tok / W;
            const int w = tok - row * W;
            const int b = row / H;
            const int h = row - b * H;
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                const int kh = kk >> 1, kw = kk & 1;
                const int hh = h - kh, ww = w - kw;
                if (hh < 0 || ww < 0) continue;
                int ho = hh, wo = ww;
                if (stride == 2) {
                    if ((hh | ww) & 1) continue;
                    ho = hh >> 1;
                    wo = ww >> 1;
                }
                if (ho >= Ho || wo >= Wo) continue;
                uint2* o = reinterpret_cast<uint2*>(out + ((static_cast<size_t>(b) * Ho + ho) * Wo + wo) * (4 * static_cast<size_t>(C)) +
                                                    static_cast<size_t>(kk) * C);
#pragma unroll
                for (int i = 0; i < NV; ++i)
                    if (act[i]) o[lane + 32 * i] = pk[i];
            }
        }
    }
}

// C = 12 * LPT channels (96 / 192 / 384 = the three ConvNeXt-tiny downsample inputs): LPT lanes share a token, three
// float4 per lane, so a warp normalises 32 / LPT tokens per pass with log2(LPT) shuffle levels and every lane busy
// (the generic kernel above leaves 8 of 32 lanes idle at C = 96 and spends most of its time on per-token index math).
// Two passes are in flight per warp; warps stride over the token list (persistent grid).
// SPLIT (fp32 mode): the normalised value is written as hi + lo bf16 planes.
template <int LPT, bool SPLIT = false>
__global__ void __launch_bounds__(256)
ln_im2col3_bf16_kernel(const float* __restrict__ x, const float* __restrict__ ln_w, const float* __restrict__ ln_b,
                       __nv_bfloat16* __restrict__ out, __nv_bfloat16* __restrict__ out_lo, int B, int H, int W, int stride,
                       int Ho, int Wo, float eps) {
    constexpr int TPW = 32 / LPT;  // tokens per warp and pass
    constexpr int C = 12 * LPT;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int sub = lane / LPT, l = lane % LPT;
    const int ntok = B * H * W;
    const int nwarps = gridDim.x * (blockDim.x >> 5);
    float4 gw[3], gb[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        gw[i] = __ldg(reinterpret_cast<const float4*>(ln_w) + l + LPT * i);
        gb[i] = __ldg(reinterpret_cast<const float4*>(ln_b) + l + LPT * i);
    }
    ptx::pdl_wait();  // weights above are constants of the plan; the activations come from the previous kernel
    ptx::pdl_launch_dependents();
    const float inv_c = 1.0f / static_cast<float>(C);
    for (int g = (blockIdx.x * (blockDim.x >> 5) + warp) * 2; g * TPW < ntok; g += 2 * nwarps) {
        float4 v[2][3];
        int tok[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            tok[u] = (g + u) * TPW + sub;
            const float4* xr = reinterpret_cast<const float4*>(x + static_cast<size_t>(min(tok[u], ntok - 1)) * C);
#pragma unroll
            for (int i = 0; i < 3; ++i) v[u][i] = __ldg(xr + l + LPT * i);
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            float s = 0.f;
#pragma unroll
            for (int i = 0; i < 3; ++i) s += (v[u][i].x + v[u][i].y) + (v[u][i].z + v[u][i].w);
#pragma unroll
            for (int o = LPT / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            const float mean = s * inv_c;
            float ss = 0.f;
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                v[u][i].x -= mean; v[u][i].y -= mean; v[u][i].z -= mean; v[u][i].w -= mean;
                ss += (v[u][i].x * v[u][i].x + v[u][i].y * v[u][i].y) + (v[u][i].z * v[u][i].z + v[u][i].w * v[u][i].w);
            }
#pragma unroll
            for (int o = LPT / 2; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
            const float rstd = rsqrtf(ss * inv_c + eps);
            if (tok[u] >= ntok) continue;
            uint2 pk[3], pl[3];
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                const float y0 = fmaf(v[u][i].x * rstd, gw[i].x, gb[i].x), y1 = fmaf(v[u][i].y * rstd, gw[i].y, gb[i].y);
                const float y2 = fmaf(v[u][i].z * rstd, gw[i].z, gb[i].z), y3 = fmaf(v[u][i].w * rstd, gw[i].w, gb[i].w);
                pk[i].x = pack_bf16x2(y0, y1);
                pk[i].y = pack_bf16x2(y2, y3);
                if (SPLIT) {
                    pl[i].x = pack_bf16x2(y0 - __uint_as_float(pk[i].x << 16), y1 - __uint_as_float(pk[i].x & 0xffff0000u));
                    pl[i].y = pack_bf16x2(y2 - __uint_as_float(pk[i].y << 16), y3 - __uint_as_float(pk[i].y & 0xffff0000u));
                }
            }
            const int row = tok[u] / W;
            const int w = tok[u] - row * W;
            const int b = row / H;
            const int h = row - b * H;
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                const int kh = kk >> 1, kw = kk & 1;
                const int hh = h - kh, ww = w - kw;
                if (hh < 0 || ww < 0) continue;
                int ho = hh, wo = ww;
                if (stride == 2) {
                    if ((hh | ww) & 1) continue;
                    ho = hh >> 1;
                    wo = ww >> 1;
                }
                if (ho >= Ho || wo >= Wo) continue;
                const size_t off = ((static_cast<size_t>(b) * Ho + ho) * Wo + wo) * (4 * static_cast<size_t>(C)) +
                                   static_cast<size_t>(kk) * C;
                uint2* o = reinterpret_cast<uint2*>(out + off);
#pragma unroll
                for (int i = 0; i < 3; ++i) o[l + LPT * i] = pk[i];
                if (SPLIT) {
                    uint2* ol = reinterpret_cast<uint2*>(out_lo + off);
#pragma unroll
                    for (int i = 0; i < 3; ++i) ol[l + LPT * i] = pl[i];
                }
            }
        }
    }
}

}  // namespace cpn
