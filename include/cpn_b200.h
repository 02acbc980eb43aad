/* cpn_b200.h -- C ABI of libcpnb200.so: the B200 (sm_100a) forward hot path of PIP-Net / Count-PIP-Net.
 *
 * The reference (TarasKutsyk/Count_PIPNet) has no FFI: its boundary for this path is the Python nn.Module API
 * (pipnet/pipnet.py:31-41 PIPNet.forward, pipnet/count_pipnet.py:70-110 CountPIPNet.forward). This library is what a
 * drop-in module binds underneath that API (count_pipnet_b200/_lib.py is the ctypes binding; INTEGRATION.md shows
 * the reference-side stub). Each entry point names the reference code whose arithmetic it replaces.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer into caller-owned memory (torch storage); the library allocates nothing;
 *   - `stream` is a cudaStream_t passed as void*; calls enqueue work and return without synchronising;
 *   - return value 0 = ok, negative = error (cpn_last_error() gives a thread-local message);
 *   - `plan` may be NULL (launch now) or a plan handle (record the launch, replayed by cpn_plan_run);
 *   - activations are channels-last: a feature map is [B, H, W, C] = [tokens, C] with C contiguous;
 *   - "bf16" buffers hold __nv_bfloat16; a non-NULL `*_lo` pointer asks for the residual plane of the bf16 hi/lo
 *     split (x = hi + lo) used by fp32 mode (three-pass bf16 GEMM).
 */
#ifndef CPN_B200_H
#define CPN_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct cpn_plan cpn_plan;

/* activation codes for cpn_gemm_bf16 */
#define CPN_ACT_NONE 0
#define CPN_ACT_GELU 1 /* exact erf GELU, torchvision/models/convnext.py:56 */
#define CPN_ACT_RELU 2

/* pooling modes for cpn_softmax_pool / cpn_head_tail */
#define CPN_POOL_MAX 0 /* PIP-Net: AdaptiveMaxPool2d((1,1)), pipnet/pipnet.py:105-108 */
#define CPN_POOL_SUM 1 /* Count-PIP-Net: proto_features.sum(dim=(2,3)), pipnet/count_pipnet.py:88 */

/* count encodings for cpn_head_tail (pipnet/count_pipnet.py:392-417) */
#define CPN_ENC_NONE 0
#define CPN_ENC_ONEHOT 1      /* count_pipnet_utils.py:141-185 */
#define CPN_ENC_LINEAR 2      /* count_pipnet_utils.py:498-519 */
#define CPN_ENC_LINEAR_FULL 3 /* count_pipnet_utils.py:429-439 */
#define CPN_ENC_BILINEAR 4    /* count_pipnet_utils.py:371-385 */
#define CPN_ENC_IDENTITY 5    /* count_pipnet_utils.py:461-462 */

/* ---- library ---------------------------------------------------------------------------------------------- */
int cpn_version(void);
const char* cpn_last_error(void);
/* 0 if the current device is sm_100 (B200) and the kernels can run on it. */
int cpn_check_device(void);

/* ---- plans: a recorded sequence of launches with pre-encoded TMA descriptors ------------------------------- */
int cpn_plan_create(cpn_plan** out);
int cpn_plan_run(cpn_plan* plan, void* stream);
/* Capture the recorded launches into one CUDA graph on `stream` (cudaStreamBeginCapture .. cudaGraphInstantiate); from
 * then on cpn_plan_run replays the graph with one cudaGraphLaunch. For small batches (the reference's visualisation and
 * attribution callers run B = 1 forwards: util/vis_pipnet.py:208-213, util/interpret_idg.py:168-172) the step is
 * launch-bound otherwise. Call after the plan has run once eagerly (lazy kernel loading is not capturable). */
int cpn_plan_capture(cpn_plan* plan, void* stream);
int cpn_plan_has_graph(const cpn_plan* plan);
int cpn_plan_num_ops(const cpn_plan* plan);
/* Replay with a CUDA event between consecutive ops; ms_out[i] = device time of op i (synchronises at the end). */
int cpn_plan_run_timed(cpn_plan* plan, void* stream, float* ms_out, int n);
/* Kernel family name and algorithmic FLOPs / HBM bytes of op i (what the roofline fractions are computed from). */
int cpn_plan_op_info(const cpn_plan* plan, int i, const char** name, double* flops, double* bytes);
/* Shape string of op i ("M=.. N=.. K=.."), for per-launch reports; "" if none. */
const char* cpn_plan_op_detail(const cpn_plan* plan, int i);
int cpn_plan_destroy(cpn_plan* plan);

/* ---- ConvNeXt feature path --------------------------------------------------------------------------------- */
/* Stem: Conv2d(3,96,k=4,s=4)+bias then LayerNorm2d(96) -- torchvision/models/convnext.py:120-131.
 * x [B,3,H,W] fp32 NCHW; w [96,48] (= conv weight [96,3,4,4]); out [B,H/4,W/4,96] fp32 channels-last. */
int cpn_stem_ln(cpn_plan* plan, const float* x, const float* w, const float* bias, const float* ln_w,
                const float* ln_b, float* out, int B, int H, int W, float eps, void* stream);

/* Depthwise 7x7 (pad 3) + bias, then LayerNorm(C) -- CNBlock.block[0..2], torchvision/models/convnext.py:52-54.
 * x [B,H,W,C] fp32; w49 [49,C] fp32 (conv weight [C,1,7,7] transposed); out_hi/out_lo [B*H*W, C] bf16. */
int cpn_dwconv7_ln(cpn_plan* plan, const float* x, const float* w49, const float* bias, const float* ln_w,
                   const float* ln_b, void* out_hi, void* out_lo, int B, int H, int W, int C, float eps,
                   void* stream);

/* bf16-mode split of the same CNBlock head (torchvision/models/convnext.py:52-55): the depthwise conv writes the
 * UN-normalised result and per-token partial LayerNorm sums, and the LayerNorm is folded into the fc1 GEMM.
 * cpn_dwconv7_stats: x [B,H,W,C] fp32, w49 [49,C], bias [C]  ->  out [B*H*W, C] bf16 = round(conv + bias),
 *   stats [B*H*W, C/32, 2] fp32 = (sum, sum of squares) of the rounded values over each 32-channel group.
 *   C must be a multiple of 32.
 * cpn_dwconv7_stats_split (fp32 mode): the same with the value kept to 16 mantissa bits as two planes,
 *   out_hi = bf16(v), out_lo = bf16(v - out_hi); stats are those of hi + lo (what the three-pass GEMM sees).
 * cpn_gemm_ln_gelu: out[M,N] bf16 = gelu(rstd_m * (A[M,K] * Bw[N,K]^T - mean_m * colsum[n]) + bias[n]) with
 *   Bw = bf16(W1 * ln_weight), colsum[n] = sum_k Bw[n,k], bias = b1 + W1 * ln_bias, and mean_m / rstd_m rebuilt from
 *   stats [M, nparts, 2] (K = C channels per token) -- equals gelu(fc1(LayerNorm(A))) of convnext.py:53-56. */
int cpn_dwconv7_stats(cpn_plan* plan, const float* x, const float* w49, const float* bias, void* out_bf16,
                      float* stats, int B, int H, int W, int C, void* stream);
int cpn_dwconv7_stats_split(cpn_plan* plan, const float* x, const float* w49, const float* bias, void* out_hi,
                            void* out_lo, float* stats, int B, int H, int W, int C, void* stream);
int cpn_gemm_ln_gelu(cpn_plan* plan, const void* A, const void* Bw, int M, int N, int K, const float* stats,
                     int nparts, const float* colsum, const float* bias, float eps, void* out_bf16, void* stream);
/* fp32 mode of cpn_gemm_ln_gelu: every operand is a (hi, lo) pair of bf16 planes, three MMA passes per k-step
 * (hi*hi + hi*lo + lo*hi), the GELU is the exact erf form and the result is written as (out_hi, out_lo) planes.
 * colsum[n] = sum_k (Bw_hi + Bw_lo)[n,k]; stats from cpn_dwconv7_stats_split. N must be a multiple of 128. */
int cpn_gemm_ln_gelu_split(cpn_plan* plan, const void* A_hi, const void* A_lo, const void* Bw_hi, const void* Bw_lo,
                           int M, int N, int K, const float* stats, int nparts, const float* colsum, const float* bias,
                           float eps, void* out_hi, void* out_lo, void* stream);

/* Whole CNBlock MLP in one kernel (bf16 mode; torchvision convnext.py:53-57, 64-66):
 *   x[M,C] += scale * (gelu(rstd_m * (y * w1g^T - mean_m * colsum) + bias1) * w2^T + bias2)
 * y [M,C] bf16 = un-normalised depthwise-conv output with its per-token partial sums stats [M, C/32, 2] (both from
 * cpn_dwconv7_stats), w1g [4C,C] bf16 = W1 * gamma, colsum [4C] its row sums, bias1 [4C] = b1 + W1 * beta,
 * w2 [C,4C] bf16, bias2 / scale [C]; x is the fp32 residual stream, updated in place. The 4C-wide hidden activations
 * stay in shared / tensor memory. C in {96, 192, 384}. */
int cpn_mlp_fused(cpn_plan* plan, const void* y_bf16, const void* w1g, const void* w2, int M, int C, const float* stats,
                  int nparts, const float* colsum, const float* bias1, const float* bias2, const float* scale, float eps,
                  float* x, void* stream);

/* LayerNorm2d(C) then the im2col rows of Conv2d(C,C',k=2,stride) -- torchvision/models/convnext.py:146-150 with
 * the stride surgery of features/convnext_features.py:5-15.
 * x [B,H,W,C] fp32; out [B*Ho*Wo, 4*C] bf16 with column (kh*2+kw)*C + c, Ho = (H-2)/stride+1. */
int cpn_ln_im2col2(cpn_plan* plan, const float* x, const float* ln_w, const float* ln_b, void* out_hi, void* out_lo,
                   int B, int H, int W, int C, int stride, float eps, void* stream);

/* tcgen05/TMA GEMM with fused epilogue:  v = A[M,K] * Bw[N,K]^T + bias;  v = act(v);  v *= scale;  v += residual;
 * (relu_after_res: the ReLU is applied after the residual add instead);  outputs fp32 and/or bf16 (hi, lo).
 * Replaces nn.Linear+GELU / nn.Linear*layer_scale+residual (torchvision/models/convnext.py:55-57,64-66), the 2x2
 * downsample convs (:146-150) and conv+BatchNorm+ReLU of features/resnet_features.py:97-117.
 * A [M,K] bf16 and Bw [N,K] bf16 row-major (K contiguous), K % 8 == 0, N % 32 == 0.
 * A_lo / Bw_lo non-NULL selects fp32 mode: acc = A*Bw + A*Bw_lo + A_lo*Bw (three bf16 passes). */
int cpn_gemm_bf16(cpn_plan* plan, const void* A, const void* A_lo, const void* Bw, const void* Bw_lo, int M, int N,
                  int K, const float* bias, const float* scale, const float* res_f32, const void* res_bf16,
                  float* out_f32, void* out_bf16, void* out_bf16_lo, int act, int relu_after_res, void* stream);

/* ---- prototype head ---------------------------------------------------------------------------------------- */
/* Add-on Conv2d(C,P,1)+bias on channels-last features -- pipnet/pipnet.py:102, pipnet/count_pipnet.py:378-379.
 * feat [T,C] fp32; w [P,C]; bias [P] or NULL; z [T,P] fp32. */
int cpn_addon_conv1x1(cpn_plan* plan, const float* feat, const float* w, const float* bias, float* z, long long T,
                      int C, int P, void* stream);
/* The same add-on conv fused with what follows it (pipnet/pipnet.py:102-108; count_pipnet.py:378-381, 88): 1x1 conv C -> P,
 * softmax over P (or hard Gumbel-softmax with the given noise), spatial max / sum per chunk of tokens -- one kernel, the
 * patch logits never reach HBM. Arguments as cpn_addon_conv1x1 + cpn_softmax_pool_arg. P = 16 or 32, C % 64 == 0. */
int cpn_addon_softmax_pool(cpn_plan* plan, const float* feat, const float* w, const float* bias, const float* gumbel,
                           float tau, int hard, int mode, float* map, float* partial, int* partial_arg, int B, int HW, int C,
                           int P, int chunks, void* stream);

/* Per-patch softmax over P (nn.Softmax(dim=1), pipnet/pipnet.py:96,103) or hard Gumbel-softmax with caller-supplied
 * noise (count_pipnet_utils.py:33-38), then spatial max / sum per image as `chunks` partial results.
 * z [B,HW,P] fp32; gumbel [B,P,HW] fp32 or NULL; map [B,P,HW] fp32 (NCHW proto_features) or NULL;
 * partial [B,chunks,P] fp32. */
int cpn_softmax_pool(cpn_plan* plan, const float* z, const float* gumbel, float tau, int hard, int mode, float* map,
                     float* partial, int B, int HW, int P, int chunks, void* stream);

/* cpn_softmax_pool with, for max pooling, the patch token (h * W + w) of every chunk maximum: partial_arg
 * [B,chunks,P] int32 (NULL = cpn_softmax_pool). cpn_pool_argmax_reduce folds the chunks into arg_out [B,P]: the
 * location of each prototype's spatial maximum per image, lowest token on ties like torch.max -- what the reference's
 * visualisation computes from the full map (util/vis_pipnet.py:209-274, 1095-1147) and what the multi-GPU
 * per-prototype max gather of SURVEY section 8(e) exchanges. */
int cpn_softmax_pool_arg(cpn_plan* plan, const float* z, const float* gumbel, float tau, int hard, int mode,
                         float* map, float* partial, int* partial_arg, int B, int HW, int P, int chunks,
                         void* stream);
int cpn_pool_argmax_reduce(cpn_plan* plan, const float* partial, const int* partial_arg, int chunks, int* arg_out,
                           int B, int P, void* stream);

/* Allocation-free variant of cpn_softmax_pool + the chunk fold of cpn_head_tail. It has NO counterpart in the reference;
 * it was added for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it). Plain softmax over P per patch, then spatial max / sum:
 * z [B,HW,P] fp32 -> pooled [B,P] fp32. The per-chunk partial results live in a static device array of
 * CPN_SOFTMAX_POOL_STATIC_MAX_BYTES inside the library (held by every GPU that loads it, until exit) instead of a
 * caller buffer: B * P * 4 bytes beyond that limit, or P > 3072, is rejected with an error. The call runs no cudaMalloc,
 * cudaFree or allocating library routine and takes no plan. All calls share the array (it keeps the last call's
 * contents; every entry that is read is overwritten first, so nothing needs clearing): calls on one stream are safe,
 * calls that OVERLAP on the GPU (other streams / threads on the same device) race. */
#define CPN_SOFTMAX_POOL_STATIC_MAX_BYTES (4u << 20)
int cpn_softmax_pool_static(const float* z, int mode, float* pooled, int B, int HW, int P, void* stream);

/* Running top-k images per prototype over a stream of batches (the search loop of util/vis_pipnet.py:209-271, which
 * runs one image per forward on the host): merges pooled [B,P] / argmax [B,P] (cpn_pool_argmax_reduce) of images
 * image_base .. image_base + B - 1 into top_val / top_img / top_tok [k,P] and top_cnt [P] (zero-initialised by the
 * caller before the first batch). An image enters a list shorter than k, or when its score is strictly greater than
 * the last entry's; lists stay sorted by descending score, equal scores in arrival order -- entry for entry what the
 * reference's loop keeps. */
int cpn_topk_update(cpn_plan* plan, const float* pooled, const int* argmax, int B, int P, int k, long long image_base,
                    float* top_val, long long* top_img, int* top_tok, int* top_cnt, void* stream);

/* Per-batch evaluation statistics of the reference's test loop (replaces pipnet/test.py:82-137 and
 * compute_local_explanation_sizes, pipnet/test.py:266-318, which build a [K,B,P] tensor and walk the batch on the host).
 * pooled [B,P], out [B,K]: the forward's 2nd and 3rd outputs; w [K,P]: the weights the reference multiplies pooled with
 * (_classification.weight as stored for PIP-Net; get_prototype_importance_per_class stacked for Count-PIP-Net);
 * ys [B] int64 labels (NULL: no labels -- top1/top5 are 0 and cm is untouched); mult = normalization_multiplier;
 * thr = 1e-3 in the reference. Per image: pred (first arg-max class), pred_prob (amax softmax(log1p(out**mult))),
 * max_score (0 = abstained), any_size / pred_size (local explanation sizes), anz (|pooled| > thr), top1 / top5
 * (label rank, ties towards the lower class). cm [K,K] uint64 ACCUMULATES cm[y][pred] += 1 (NULL to skip).
 * relevant [K,P] bytes is scratch (zeroed by the call); protos_per_class [K] = prototypes with a non-zero thresholded
 * contribution to the class in this batch (pipnet/test.py:108-113). Integer results are bit-exact. */
int cpn_eval_batch(cpn_plan* plan, const float* pooled, const float* out, const float* w, const long long* ys,
                   float mult, float thr, int B, int P, int K, int* pred, float* pred_prob, float* max_score,
                   int* any_size, int* pred_size, int* anz, int* top1, int* top5, unsigned long long* cm,
                   unsigned char* relevant, int* protos_per_class, void* stream);

/* Chunk reduction, inference threshold (pipnet/pipnet.py:36) or round+clamp (pipnet/count_pipnet.py:90-97),
 * count encoding, and the non-negative classifier F.linear(x, relu(W), b) (pipnet/pipnet.py:70-71).
 * pooled_out [B,P] is the module's second return value; logits_out [B,K] the third. */
int cpn_head_tail(cpn_plan* plan, const float* partial, int chunks, int mode, int inference, int use_ste,
                  int max_count, int encoding, const float* enc_w0, const float* enc_w1, const float* enc_w2,
                  const float* cls_w, const float* cls_b, float* pooled_out, float* logits_out, int B, int P, int D,
                  int K, void* stream);

/* ---- layout / ResNet helpers ------------------------------------------------------------------------------- */
/* [B,H,W,C] fp32 channels-last -> [B,C,H,W] fp32 (what reference callers of net._net(x) expect). */
int cpn_nhwc_to_nchw(cpn_plan* plan, const float* x, float* out, int B, int HW, int C, void* stream);
/* [B,C,H,W] fp32 -> [B,H,W,C] fp32 (entry of net._add_on(features) when called on its own). */
int cpn_nchw_to_nhwc(cpn_plan* plan, const float* x, float* out, int B, int HW, int C, void* stream);

/* ResNet conv + folded BatchNorm (+ ReLU / + identity) as ONE tcgen05 GEMM without an im2col matrix
 * (features/resnet_features.py:97-117 Bottleneck.forward, :55-74 BasicBlock.forward; eval-mode BN is an affine map).
 * x_bf16 [B,H,W,Cin] channels-last bf16; w_bf16 [Cout, k*k*Cin] bf16 with K ordered (kh, kw, cin) and the BN scale folded
 * in; bias [Cout] = the folded BN shift. ksize = 1 (plain GEMM) or 3 (stride 1, pad 1: the nine taps are read from the
 * activation with 4-D TMA boxes, out-of-map rows / columns arrive as zeros = the padding). Output combinations:
 *   out_bf16 only, relu = 1                       out_bf16 = relu(conv + bias)                  conv1, conv2
 *   out_bf16 only, relu = 0                       out_bf16 = conv + bias                        projection shortcut (bf16 stream)
 *   res_f32 + out_f32 + out_bf16, relu = 1        out_f32 = relu(res_f32 + conv + bias), out_bf16 = bf16(out_f32)   block end
 *   out_f32 only, relu = 0                        out_f32 = conv + bias                         projection shortcut
 * Cout % 64 == 0; ksize 3 needs Cin % 64 == 0 and cpn_conv3x3_tile_rows(H, W) > 0 (image rows per 128-row M tile; 0 =
 * this map size needs the im2col path: cpn_im2col_bf16 + cpn_gemm_bf16). */
int cpn_conv3x3_tile_rows(int H, int W);
int cpn_conv_bn_act(cpn_plan* plan, const void* x_bf16, const void* w_bf16, const float* bias, const float* res_f32,
                    float* out_f32, void* out_bf16, int B, int H, int W, int Cin, int Cout, int ksize, int relu,
                    void* stream);
/* The same conv + folded BN + identity + ReLU on a bf16 residual stream (bf16 mode of the ResNet program: a block end then
 * moves 2 + 2 bytes per element instead of 4 + 4 + 2; these 1x1 convs are HBM-bound). res_bf16 [B*H*W, Cout] bf16 is the
 * identity (features/resnet_features.py:110-115 `out += identity; out = self.relu(out)`); exactly one output:
 *   out_bf16   = bf16(relu(res_bf16 + conv + bias))      every block but the last
 *   out_f32    = relu(res_bf16 + conv + bias)            the last block (fp32 features for the head)
 * cpn_conv_bn_act with out_bf16 only and relu = 0 gives the projection shortcut as bf16. */
int cpn_conv_bn_act_resb(cpn_plan* plan, const void* x_bf16, const void* w_bf16, const float* bias, const void* res_bf16,
                         float* out_f32, void* out_bf16, int B, int H, int W, int Cin, int Cout, int ksize, void* stream);

/* im2col rows of a kxk convolution (pad, stride) over a channels-last bf16 map, K padded to Kpad with zeros:
 * out[(b,ho,wo), (kh*k+kw)*C + c] -- feeds cpn_gemm_bf16 for features/resnet_features.py convs. */
int cpn_im2col_bf16(cpn_plan* plan, const void* x, void* out, int B, int H, int W, int C, int k, int stride,
                    int pad, int Kpad, void* stream);
/* Same gather from the fp32 NCHW network input (ResNet conv1 7x7 s2, features/resnet_features.py:139). */
int cpn_im2col_nchw_f32(cpn_plan* plan, const float* x, void* out, void* out_lo, int B, int C, int H, int W, int k,
                        int stride, int pad, int Kpad, void* stream);
/* MaxPool2d(3, stride 2, pad 1) on a channels-last fp32 map (features/resnet_features.py:142); any of the outputs
 * (fp32, bf16 hi, bf16 lo) may be NULL. */
int cpn_maxpool3s2(cpn_plan* plan, const float* x, float* out_f32, void* out_hi, void* out_lo, int B, int H, int W,
                   int C, void* stream);
/* The same pooling on a channels-last bf16 map (the ResNet program's bf16 residual stream: rounding is monotonic, so the
 * result equals bf16(maxpool(fp32 map))). C % 8 == 0. */
int cpn_maxpool3s2_bf16(cpn_plan* plan, const void* x, void* out, int B, int H, int W, int C, void* stream);
/* Strided row subsample of a channels-last bf16 map (1x1 stride-2 downsample conv input, resnet_features.py:178). */
int cpn_subsample2_bf16(cpn_plan* plan, const void* x, void* out, int B, int H, int W, int C, void* stream);

/* ---- input pipeline -------------------------------------------------------------------- */
/* The reference's evaluation transform `Compose([Resize((Ho, Wo)), ToTensor(), Normalize(mean, std)])` (util/data.py:381-385;
 * same steps at :265-275, :296-307, :465-480, :522-531) for a batch of decoded RGB images of one size, bit-exact with
 * Pillow's BILINEAR resize (Resample.c: widened triangle filter, 22-bit fixed-point coefficients, horizontal then vertical
 * pass, each rounded to uint8) and torchvision's fp32 ToTensor / Normalize.
 * src [B, Hs, Ws, 3] uint8 (device) -> out [B, 3, Ho, Wo] fp32 NCHW (device); mean3 / std3: three HOST floats each, read
 * during the call; ws: device workspace of cpn_preprocess_ws_ints(Hs, Ws, Ho, Wo) ints (coefficient tables, rewritten by
 * every call). Source sides up to 16384, down-scaling up to 64x. */
long long cpn_preprocess_ws_ints(int Hs, int Ws, int Ho, int Wo);
int cpn_preprocess_u8(cpn_plan* plan, const unsigned char* src, float* out, int* ws, int B, int Hs, int Ws, int Ho, int Wo,
                      const float* mean3, const float* std3, void* stream);

/* ---- training side of the prototype head ----------------------------------------------- */
/* These four take the map as the reference's modules return it: [N, P, H*W] fp32, NCHW. No plans: a training step is
 * not replayed. All reductions run in a fixed order (two runs give the same bits).
 *
 * calculate_loss + align_loss (pipnet/train.py:165-243, 253-265) and every gradient of the loss in one call:
 *   align = -mean_n log(sum_c pf1[n,c] pf2[n,c] + 1e-12)         (train.py:173, both directions have this value)
 *   tanh  = -(mean_p log(tanh(sum_b coeff*pooled1[b,p]) + eps) + the same for pooled2) / 2          (train.py:198-201)
 *   class = NLLLoss(mean)(log_softmax(use_log1p ? log1p(out^m) : out), cat(ys, ys))               (train.py:216-222)
 *   loss  = finetune ? class_w*class : align_w*align + tanh_w*tanh (+ class_w*class unless pretrain) (train.py:212-229)
 * proto_features [2B, P, HW], pooled [2B, P], out [2B, K] (view 1 = rows [0, B), view 2 = rows [B, 2B): the halves of
 * train.py:167-168's chunk(2)), ys [B] int64 labels in [0, K), norm_mult = the classifier's normalization_multiplier (one
 * device float). out / ys / norm_mult are ignored when pretrain != 0. Outputs: grad_pf / grad_pooled / grad_out = d loss
 * / d input (same shapes; any may be NULL; grad_pf and grad_pooled must be NULL when finetune != 0, the loss does not
 * depend on them); result [8] = {align, tanh, class, loss, number of rows whose arg-max(out) is the label
 * (train.py:233-235), d loss / d norm_mult, 0, 0}; workspace: cpn_pipnet_loss_ws_floats(B, P, HW) floats. */
long long cpn_pipnet_loss_ws_floats(int B, int P, int HW);
int cpn_pipnet_loss(const float* proto_features, const float* pooled, const float* out, const long long* ys,
                    const float* norm_mult, int B, int P, int HW, int K, float align_w, float tanh_w, float class_w,
                    float tanh_coeff, float eps_tanh, int pretrain, int finetune, int use_log1p, float* grad_pf,
                    float* grad_pooled, float* grad_out, float* result, float* workspace, void* stream);
/* buf[0..n) *= *scale unless *scale == 1 (scale: one device float). The backward of cpn_pipnet_loss's autograd node:
 * loss.backward() (train.py:103) arrives with 1, and then no pass over the gradients is spent. */
int cpn_scale_unless_one(float* buf, long long n, const float* scale, void* stream);
/* Softmax(dim=1) over the prototypes of an NCHW map and the spatial pooling of the autograd (training) forward:
 * pipnet/pipnet.py:32-33 with :96 and :105-108 (mode CPN_POOL_MAX; argmax [N, P] int32 = flattened position of the
 * maximum, the first one on ties as AdaptiveMaxPool2d's backward picks it), pipnet/count_pipnet.py:356-360 and :88
 * (mode CPN_POOL_SUM; argmax unused, may be NULL). z -> proto_features [N, P, HW], pooled [N, P];
 * workspace: cpn_softmax_pool_nchw_ws_floats(N, P, HW) floats. */
long long cpn_softmax_pool_nchw_ws_floats(int N, int P, int HW);
int cpn_softmax_pool_nchw(const float* z, int mode, float* proto_features, float* pooled, int* argmax, int N, int P,
                          int HW, float* workspace, void* stream);
/* Its backward (what autograd does for those modules, as one kernel): grad_z = pf * (g - sum_c g*pf) with
 * g = grad_pf (NULL = zeros) + grad_pooled (NULL = zeros) routed to the arg-max position (max) or to every position (sum). */
int cpn_softmax_pool_nchw_bwd(const float* proto_features, const float* grad_pf, const float* grad_pooled,
                              const int* argmax, int mode, float* grad_z, int N, int P, int HW, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CPN_B200_H */
