/* libb200infer — C ABI of the B200-native batched-forward engine.
 *
 * Drop-in scope: the path the reference scheduler (293-project/src/scheduler.py) runs per batch inside a model actor:
 *   RequestQueue.get_batch (:258-322)  ->  GPUWorker.process_batch (:435-475):  torch.stack -> .cuda() -> model(inputs) -> .cpu()
 * plus the client-side preprocessing the reference leaves to callers (profiling/run_profiler.py:63-70).
 *
 * Conventions: every entry returns 0 on success or a negative B200_ERR_* code; b200_last_error() then describes the
 * failure (thread-local).  Nothing throws across the boundary.  All pointers named `*_dev`, and all in/out tensor
 * pointers of the operator-level entries, are device pointers on the current CUDA device; `stream` is a cudaStream_t
 * passed as void* (NULL = legacy default stream).  Activations are NHWC bf16; `*_pix_stride` is the distance in
 * elements between consecutive pixels (= channel count of the underlying buffer, 0 = dense), which lets a call read
 * or write a channel slice of a wider buffer (concat / chunk without a copy).
 */
#ifndef B200INFER_H_
#define B200INFER_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_ABI_VERSION 1

#if defined(__GNUC__)
#define B200_API __attribute__((visibility("default")))
#else
#define B200_API
#endif

#define B200_OK 0
#define B200_ERR_INVALID (-1)
#define B200_ERR_CUDA (-2)
#define B200_ERR_BLOB (-3)
#define B200_ERR_NO_DEVICE (-4)

/* activation codes for the fused epilogues */
#define B200_ACT_NONE 0
#define B200_ACT_RELU 1
#define B200_ACT_SILU 2
#define B200_ACT_SIGMOID 3

B200_API const char* b200_last_error(void);
B200_API int b200_abi_version(void);

/* ---- operator level (one entry per kernel; what the parity tests bind) ------------------------------------- */

/* Conv2d (k in {1,3}, stride in {1,2}) + folded BatchNorm bias (+ residual add) + activation, tcgen05 implicit GEMM.
 * Replaces torch.nn.Conv2d + BatchNorm2d + ReLU/SiLU (+ `out += identity`) as reached from scheduler.py:450 through
 * torchvision (models/resnet.py:143-161 Bottleneck.forward; models/efficientnet.py:105-231; ops/misc.py Conv2dNormActivation).
 * weight: bf16 [cout][k*k][cin] (tap-major, then input channel); bias: fp32 [cout]. cin, cout multiples of 8. */
B200_API int b200_conv2d_nhwc(const void* in, int n, int h, int w, int cin, int in_pix_stride, const void* weight,
                     const float* bias, const void* residual, int res_pix_stride, void* out, int cout,
                     int out_pix_stride, int ksize, int stride, int pad, int act, void* stream);

/* Two 1x1 convolutions that are summed, as ONE GEMM with concatenated K: out = act(W[:, :cin] . in + W[:, cin:] .
 * in2[::stride2, ::stride2] + bias). Replaces the last conv of a ResNet bottleneck together with its `downsample` branch
 * (tv resnet.py:155-161: `out = bn3(conv3(out)); identity = downsample(x); out += identity; relu`): the identity tensor
 * is neither written nor read back. weight: bf16 [cout][cin + cin2], bias fp32 [cout] (= both folded BN biases added);
 * cin a multiple of 64; in2 is [n][h2][w2][cin2] with (h2 - 1) / stride2 + 1 == h. */
B200_API int b200_conv1x1_dual_nhwc(const void* in, int n, int h, int w, int cin, int in_pix_stride, const void* in2, int h2,
                                    int w2, int cin2, int in2_pix_stride, int stride2, const void* weight,
                                    const float* bias, void* out, int cout, int out_pix_stride, int act, void* stream);

/* 1x1 convolution whose input is first multiplied by a per-(image, channel) factor: out = act(W . (in * in_scale[n][c]) +
 * bias (+ residual)). Replaces `scale * input` of SqueezeExcitation.forward (tv ops/misc.py:256-258) together with the
 * MBConv project conv that consumes it (tv efficientnet.py:157-163): the scaled activation is never written. The product
 * is rounded to bf16 in shared memory, exactly where b200_channel_scale_nhwc would have stored it. in_scale fp32 [n][cin]. */
B200_API int b200_conv1x1_scaled_nhwc(const void* in, int n, int h, int w, int cin, int in_pix_stride, const float* in_scale,
                                      const void* weight, const float* bias, const void* residual, int res_pix_stride,
                                      void* out, int cout, int out_pix_stride, int act, void* stream);

/* The tail of a ShuffleNetV2 unit in one kernel: out = channel_shuffle(cat(other, relu(W . in + bias)), groups = 2)
 * (tv models/shufflenetv2.py:24-38 channel_shuffle, :96-103 InvertedResidual.forward), written by the epilogue of the 1x1
 * conv that ends branch2; replaces b200_conv2d_nhwc followed by b200_shuffle_concat_nhwc, and `other`'s values are copied
 * bit for bit. other: c channels in logical order, padded to cp; out: the halves layout of b200_shuffle_concat_nhwc with
 * half width cp (logical l at l for l < c, at cp + l - c above; pads zero). cp = 64 or a multiple of 128, c even.
 * weight: bf16 [cp][cin] in "slots" -- output channel i < c/2 in row i, i >= c/2 in row cp/2 + (i - c/2), other rows zero --
 * and bias fp32 [cp] likewise: slot s lands on physical channel 2 s + 1, other's value of the same slot on 2 s. */
B200_API int b200_conv1x1_shuffle_nhwc(const void* in, int n, int h, int w, int cin, int in_pix_stride, const void* other,
                                       int other_pix_stride, int c, int cp, const void* weight, const float* bias,
                                       void* out, int out_pix_stride, void* stream);

/* Same contract as b200_conv2d_nhwc for layers with few input channels (cin <= 56, cout <= 256): GEMM-K packs the filter
 * taps (k = tap * cin + c), the im2col tile is gathered into shared memory by cp.async producers. Serves the
 * FusedMBConv 3x3 convolutions of EfficientNetV2 (tv efficientnet.py:172-231) and YOLOv8n's early layers.
 * weight: bf16 [b200_gather_conv_npad(cout)][b200_gather_conv_kpad(cin, k)], zero padded. */
B200_API int b200_gather_conv_npad(int cout);
B200_API int b200_gather_conv_kpad(int cin, int ksize);
B200_API int b200_gather_conv2d_nhwc(const void* in, int n, int h, int w, int cin, int in_pix_stride, const void* weight,
                                     const float* bias, const void* residual, int res_pix_stride, void* out, int cout,
                                     int out_pix_stride, int ksize, int stride, int pad, int act, int res_after_act,
                                     void* stream);

/* Stem convolution with 3 input channels (input NHWC with C padded to 4) + folded BN + activation, direct FP32-pipe kernel.
 * Replaces tv resnet.py:268-270 (conv1 7x7/2 + bn1 + relu), shufflenetv2.py:112-116, efficientnet.py:283-287 (3x3/2 stems).
 * weight: fp32 [k*k][3][cout_pad], cout_pad = cout rounded up to the channel tile b200_stem_cout_tile(cout). */
B200_API int b200_stem_cout_tile(int cout);
B200_API int b200_stem_conv_nhwc(const void* in, int n, int h, int w, const float* weight, const float* bias, void* out,
                                 int cout, int out_pix_stride, int ksize, int stride, int pad, int act, void* stream);

/* Tensor-core variant of the stem (7x7/2 pad 3 or 3x3/2 pad 1): the producer warps build the im2col tile in shared
 * memory with 16-byte cp.async copies, tcgen05.mma consumes it. weight: bf16 [b200_stem_tc_nout(cout)][b200_stem_tc_kpad(k)],
 * GEMM-K index = r * (k == 7 ? 32 : 16) + (s + 1) * 4 + c (kernel row r, kernel column s, channel c), zero elsewhere. */
B200_API int b200_stem_tc_kpad(int ksize);
B200_API int b200_stem_tc_nout(int cout);
B200_API int b200_stem_conv_tc_nhwc(const void* in, int n, int h, int w, const void* weight, const float* bias, void* out,
                                    int cout, int out_pix_stride, int ksize, int stride, int pad, int act, void* stream);

/* ResNet stem (7x7/2 pad 3, 3 -> 64, + folded BN + ReLU) with no im2col copy: one TMA box brings the 7 input rows of an
 * output row and non-swizzled tcgen05 descriptors (LBO 16 B, SBO 128 B) read the overlapping 8-pixel windows in place.
 * Replaces tv resnet.py:268-270. weight: bf16 image [7][4][64][8] (b200_stem_row_weight_bytes() bytes): element e of K chunk
 * k of kernel row r = weight[cout][c = e % 4][r][s = 2k + e / 4 - 1], zero where c == 3 or s outside [0, 6]. */
B200_API int b200_stem_row_weight_bytes(void);
B200_API int b200_stem_row_conv_nhwc(const void* in, int n, int h, int w, const void* weight, const float* bias, void* out,
                                     int out_pix_stride, void* stream);

/* Same stem fused with MaxPool2d(3,2,1) (tv resnet.py:268-271): the 112x112x64 activation stays in shared memory, `out` is
 * the pooled [n][h/4][w/4][64] tensor. Stem width (w/2) must be <= 128; weight layout as b200_stem_row_conv_nhwc. */
B200_API int b200_stem_pool_nhwc(const void* in, int n, int h, int w, const void* weight, const float* bias, void* out,
                                 int out_pix_stride, void* stream);

/* The same two kernels for every stride-2 stem the served models have: ksize 7 (pad 3) or 3 (pad 1), cout a multiple of 8
 * up to 64, act none / relu / silu (tv shufflenetv2.py:112-118 conv1 + maxpool, efficientnet.py:283-287, YOLOv8 layer 0).
 * The aligned window of output pixel ow starts at input pixel 2*ow - pad - 1 and spans kc 16-byte chunks (kc = 4 for 7x7,
 * 2 for 3x3). weight: bf16 image [ksize][kc][cout <= 32 ? 32 : 64][8] (b200_stem_row_weight_bytes_k(ksize, cout) bytes):
 * element e of chunk k of kernel row r = weight[cout][c = e % 4][r][s = 2k + e / 4 - 1], zero where c == 3, s outside the
 * kernel, or the row is past cout; bias fp32 padded to the same width. The pooled variant needs w/2 <= 128. */
B200_API int b200_stem_row_weight_bytes_k(int ksize, int cout);
B200_API int b200_stem_row_conv_k_nhwc(const void* in, int n, int h, int w, const void* weight, const float* bias, void* out,
                                       int cout, int out_pix_stride, int ksize, int act, void* stream);
B200_API int b200_stem_pool_k_nhwc(const void* in, int n, int h, int w, const void* weight, const float* bias, void* out,
                                   int cout, int out_pix_stride, int ksize, int act, void* stream);

/* Depthwise 3x3 (pad 1) + folded BN + activation. Replaces tv shufflenetv2.py:62-76 `depthwise_conv` and the depthwise
 * stage of MBConv (tv efficientnet.py:140-151). weight: fp32 [9][c]; bias fp32 [c]. */
B200_API int b200_dwconv3x3_nhwc(const void* in, int n, int h, int w, int c, int in_pix_stride, const float* weight,
                                 const float* bias, void* out, int out_pix_stride, int stride, int act, void* stream);

/* The same depthwise convolution with the squeeze of the SqueezeExcitation block that follows it in MBConv fused in
 * (tv efficientnet.py:153-155 -> tv ops/misc.py:249-250, AdaptiveAvgPool2d(1)): besides `out`, writes fp32
 * strip_sums[n][rows][c], channel sums of the stored outputs, rows = b200_dwconv3x3_squeeze_rows(c, ho, wo, stride):
 * 1 when the layer runs on the TMA-tile kernel (stride 1, one row band per image: the block adds its strips itself), else
 * one row per strip of the strip kernel (b200_dwconv3x3_strips); 0: this output geometry has neither and the call is
 * rejected -- widths that are a multiple of 7 have one. b200_se_gate / b200_se_hidden add the rows in index order, so
 * the result is deterministic. */
B200_API int b200_dwconv3x3_squeeze_rows(int c, int ho, int wo, int stride);
B200_API int b200_dwconv3x3_strips(int ho, int wo, int stride);
B200_API int b200_dwconv3x3_squeeze_nhwc(const void* in, int n, int h, int w, int c, int in_pix_stride,
                                         const float* weight, const float* bias, void* out, int out_pix_stride,
                                         int stride, int act, float* strip_sums, void* stream);

/* Squeeze-excitation gate, one launch: gate[n][c] = sigmoid(fc2(silu(fc1(sum(strip_sums[n]) * inv_hw)))). Replaces
 * SqueezeExcitation._scale (tv ops/misc.py:249-254). w1 fp32 [squeeze][c] (fc1 weight), b1 [squeeze], w2t fp32
 * [squeeze][c] (fc2 weight TRANSPOSED), b2 [c]; gate fp32 [n][c] feeds b200_channel_scale_nhwc. */
B200_API int b200_se_gate(const float* strip_sums, int n, int strips, int c, float inv_hw, const float* w1,
                          const float* b1, const float* w2t, const float* b2, int squeeze, float* gate, void* stream);

/* The same block in two launches, so that no thread block streams a whole fc matrix (the form the engine uses):
 * b200_se_hidden: hidden[n][j] = silu(b1[j] + inv_hw * sum_c rows(strip_sums[n])[c] * w1[j][c]), fp32 [n][squeeze];
 * b200_se_scale_nhwc: out = in * sigmoid(b2[c] + sum_j hidden[n][j] * w2t[j][c]) -- fc2, scale_activation and the
 * `scale * input` of SqueezeExcitation.forward (tv ops/misc.py:252-258) in one pass over the activation. squeeze <= 256. */
B200_API int b200_se_hidden(const float* strip_sums, int n, int rows, int c, float inv_hw, const float* w1, const float* b1,
                            int squeeze, float* hidden, void* stream);
B200_API int b200_se_scale_nhwc(const void* in, int n, int h, int w, int c, int in_pix_stride, const float* hidden,
                                const float* w2t, const float* b2, int squeeze, void* out, int out_pix_stride,
                                void* stream);

/* Max pooling (k, stride, pad). Replaces nn.MaxPool2d(3,2,1) (tv resnet.py:271, shufflenetv2.py:118) and SPPF's 5x5/1/2. */
B200_API int b200_maxpool_nhwc(const void* in, int n, int h, int w, int c, int in_pix_stride, void* out,
                               int out_pix_stride, int ksize, int stride, int pad, void* stream);

/* Global average pool to fp32 [n][c]. Replaces nn.AdaptiveAvgPool2d(1) (tv resnet.py:278) / x.mean([2,3]) (shufflenetv2.py:161). */
B200_API int b200_global_avgpool_nhwc(const void* in, int n, int h, int w, int c, int in_pix_stride, float* out,
                                      void* stream);

/* out[n][j] = act(x_scale * sum_k x[n][k] w[j][k] + b[j]), all fp32. Replaces nn.Linear (tv resnet.py:280) and the two
 * 1x1 convs of SqueezeExcitation (tv ops/misc.py). */
B200_API int b200_fc(const float* x, const float* w, const float* b, float* out, int n, int k, int j, float x_scale,
                     int act, void* stream);

/* y = x * gate[n][c] (SE excitation; tv ops/misc.py SqueezeExcitation.forward `scale * input`). */
B200_API int b200_channel_scale_nhwc(const void* in, int n, int h, int w, int c, int in_pix_stride, const float* gate,
                                     void* out, int out_pix_stride, void* stream);

/* Channel-slice copy / nearest 2x upsample / ShuffleNetV2 concat + channel_shuffle(groups=2)
 * (tv shufflenetv2.py:28-40,98-104: logical out[2i] = a[i], out[2i+1] = b[i], i < c). a and b hold c channels padded to
 * cp = ceil8(c); out has 2*cp channels stored as two halves: logical l at physical (l < c ? l : cp + l - c), pads zero,
 * so the next unit's x.chunk(2) is two 16-byte aligned slices. */
B200_API int b200_copy_channels_nhwc(const void* in, int n, int h, int w, int c, int in_pix_stride, void* out,
                                     int out_pix_stride, void* stream);
B200_API int b200_upsample2x_nhwc(const void* in, int n, int h, int w, int c, int in_pix_stride, void* out,
                                  int out_pix_stride, void* stream);
B200_API int b200_shuffle_concat_nhwc(const void* a, int a_pix_stride, const void* b, int b_pix_stride, int n, int h,
                                      int w, int c, void* out, int out_pix_stride, void* stream);

/* ---- engine level (what B200Model / B200GPUWorker.process_batch bind) ------------------------------------------ */

typedef struct b200_engine b200_engine;

/* Builds an engine from a plan+weights blob (produced by rdb_b200.lowering from a torchvision-style module; BatchNorm
 * folded, bf16 conv weights). Replaces GPUWorker model load `model_registry[name].cpu().to(device).eval()`
 * (scheduler.py:409-417, 508-515). model_kind may be NULL/"" (accept whatever the blob holds). Allocates all device
 * memory for `max_batch` up front. Not thread-safe: one engine per actor process, like the reference's models dict. */
B200_API int b200_engine_create(const char* model_kind, const void* blob, size_t nbytes, int max_batch, int device,
                                b200_engine** out);
/* Same, with options.  B200_ENGINE_KEEP_BUFFERS: every activation buffer keeps memory of its own, so b200_engine_buffer /
 * b200_engine_read_buffer see each layer's output after a forward (the per-layer parity tests).  Default (flags = 0):
 * activation buffers whose lifetimes in the plan do not overlap share addresses -- the engine then holds about what
 * torch holds for the same forward (`peak_memory_mb` of the reference's profiles, profiling/ModelProfiler.py:85-112,
 * which 293-project/src/nexus.py:222-226 sums to decide what fits on a GPU); only the model input, the fp32 vectors
 * and the output stay readable after a forward. */
#define B200_ENGINE_KEEP_BUFFERS 1u
B200_API int b200_engine_create_ex(const char* model_kind, const void* blob, size_t nbytes, int max_batch, int device,
                                   unsigned flags, b200_engine** out);
/* Prepares `n` batch sizes ahead of traffic: TMA descriptors, one eager pass and the CUDA-graph capture each (the
 * reference's actor loads its models before serving, scheduler.py:409-417; its first batches pay cuDNN autotuning
 * instead).  A worker that pads batches to a fixed set of bucket sizes calls this once at load; afterwards b200_infer of
 * those sizes only replays graphs.  Synchronises `stream`. */
B200_API int b200_engine_warmup(b200_engine* e, const int* batches, int n, void* stream);
/* Device memory the engine holds right now (weights, activation arena, tables, staging) and the activation part of it. */
B200_API int b200_engine_memory(b200_engine* e, size_t* device_bytes, size_t* activation_bytes);
/* What b200_engine_create_ex(blob, max_batch, flags) would hold, without touching a device: the figure a batch-size
 * profile reports as peak memory (ModelProfiler.py:355-357 `peak_memory_mb`). */
B200_API int b200_plan_memory(const void* blob, size_t nbytes, int max_batch, unsigned flags, size_t* device_bytes);
B200_API void b200_engine_destroy(b200_engine* e);
B200_API int b200_engine_info(b200_engine* e, int* in_h, int* in_w, int* out_dim, int* max_batch, int* n_ops);
B200_API int b200_engine_set_graphs(b200_engine* e, int enabled); /* CUDA-graph replay per batch size (default on) */
B200_API long long b200_engine_launch_count(b200_engine* e);      /* kernels launched so far by this engine */
/* Device pointer of an activation buffer (-1 = model input NHWC4 bf16, -2 = model output fp32), for tests/zero-copy. */
B200_API int b200_engine_buffer(b200_engine* e, int buf, void** dev_ptr, size_t* bytes_per_image);

/* Per-op device time of one eager pass (CUDA events on `stream` around every op; synchronises). ms_per_op / flops_per_op
 * have n_ops entries (b200_engine_info); flops are 2*M*N*K of the tensor-core convolutions, 0 for other ops. */
B200_API int b200_engine_profile(b200_engine* e, int batch, float* ms_per_op, double* flops_per_op, int n_ops,
                                 void* stream);

/* Synchronising debug read of an activation buffer (first `batch` images) into host memory; used by per-layer parity tests. */
B200_API int b200_engine_read_buffer(b200_engine* e, int buf, int batch, void* host_dst);

/* Batch collate fused with preprocessing: n uint8 HWC images (packed in one buffer at `offsets`, sizes hw[n][2]) ->
 * padded NHWC(C=4) bf16 batch of `padded_batch` images, normalised with the blob's mean/std; slots >= n are zero.
 * Replaces torch.stack(batch.inputs) (scheduler.py:443) + the client-side ToTensor/Normalize/Resize
 * (profiling/run_profiler.py:63-70). mode: 0 identity (sizes must equal the model input), 1 pad top-left, 2 bilinear
 * resize (no antialias), 3 letterbox, 4 = transforms.Resize((H, W)): bilinear WITH antialias stretched to the model
 * input, 5 = transforms.Resize(S) + CenterCrop(S) for a square S x S model input (short side to S with antialias,
 * long side int(S * long / short), centre crop) -- both with torch's antialias weights, one rounding at the end: within
 * 1 in the last uint8 bit of torchvision.transforms.v2.functional.resize(uint8, antialias=True) before normalisation.
 * A request with hw[i][1] < 0 is a single-channel (grayscale) image of width -hw[i][1], h * w bytes, expanded L -> RGB
 * like PIL's convert("RGB") (31 of the 2000 images of sample-models/yolov8/workload are grayscale). src_on_device=0: `packed` is host memory (pinned for an async copy) and is staged with one H2D
 * copy; 1: it is already a device pointer. out = NULL writes straight into the engine's input buffer. */
B200_API int b200_collate(b200_engine* e, const void* packed, const uint64_t* offsets, const int32_t* hw, int n,
                          int padded_batch, int mode, int src_on_device, size_t packed_bytes, void* out_nhwc_bf16,
                          void* stream);
/* (No counterpart in the reference.)  b200_collate for callers that may not allocate device memory while serving: an
 * actor placed on a GPU by the profiled peak memory of its sessions has no headroom for, and cannot afford the
 * device-wide synchronisation of, a cudaMalloc inside its duty cycle.  The caller passes the packed uint8 images (already
 * on the device) and the output; the per-image table and the normalisation table -- the only other memory the kernel
 * touches -- live in a fixed array inside the library image: B200_NOALLOC_MAX_IMAGES * 16 B + 1.5 KB = 17.5 KB held by every
 * GPU that loads the library, until exit.  n > B200_NOALLOC_MAX_IMAGES is refused with B200_ERR_INVALID.  Never calls
 * cudaMalloc/cudaFree or any allocating routine.  The first call on each device also resolves the array's device
 * address, and with lazy module loading the CUDA runtime loads the kernel code onto the device then: make one warm-up
 * call per device during actor start-up (before any stream capture); from the second call on, device memory in use does
 * not change across a call.  A stream that is being captured is refused with B200_ERR_INVALID (the table uploads read
 * host stack memory).  packed_dev that is not 16-byte aligned is served by the byte-wise kernel.  All calls share the one array, which keeps the last
 * call's tables: two calls whose GPU work overlaps (different streams, or threads) race on it -- keep them on one
 * stream, or order them with events.  mean3/std3: per-channel normalisation ((u8/255 - mean) / std); other arguments as
 * b200_collate. */
#define B200_NOALLOC_MAX_IMAGES 1024
B200_API int b200_collate_u8_noalloc(const void* packed_dev, const uint64_t* offsets, const int32_t* hw, int n,
                                     int padded_batch, int out_h, int out_w, int mode, int fill, const float* mean3,
                                     const float* std3, size_t packed_bytes, void* out_nhwc4_bf16_dev, void* stream);

/* Same for reference-native requests: n already-normalised fp32 CHW tensors stacked as [n][3][H][W]
 * (293-project/src/test_scheduler.py:89 submits torch.randn(3,224,224)). */
B200_API int b200_collate_f32(b200_engine* e, const float* stacked_chw, int n, int padded_batch, int src_on_device,
                              void* out_nhwc_bf16, void* stream);

/* YOLOv8 post-processing on the device: class-aware greedy NMS over the engine's [batch][84][A] head output (head = NULL
 * uses the engine's output buffer; otherwise a device pointer of that layout). boxes: device fp32 [batch][max_det][6]
 * = x1,y1,x2,y2,score,class (input pixels); counts: device int32 [batch]. max_det <= 300. Restates ultralytics
 * utils/ops.py non_max_suppression (conf 0.25 / IoU 0.45 / max_det 300 are its defaults); no counterpart runs in the
 * reference (YOLOv8 is workload-only there). */
B200_API int b200_decode_yolo(b200_engine* e, const float* head, int batch, float conf, float iou, int max_det,
                              float* boxes, int32_t* counts, void* stream);

/* Forward pass of `batch` images. in = NULL or the engine input buffer: run in place; otherwise it is copied in first.
 * out (host or device, fp32 [batch][out_dim]) receives the result with an async copy on `stream`; NULL skips the copy.
 * Replaces `outputs = model(inputs)` and `outputs.cpu()` (scheduler.py:449-450, 468). No host synchronisation inside. */
B200_API int b200_infer(b200_engine* e, const void* in_nhwc_bf16, int batch, float* out, void* stream);

/* ---- request ring (what B200Scheduler.submit_request / SlabRequestQueue.get_batch bind) ------------------------------
 * The per-model FIFO of the reference (293-project/src/scheduler.py:190-372: RequestQueue over ray.util.queue.Queue, a
 * queue actor reached by RPC, python/ray/util/queue.py:20-305) as a lock-free multi-producer / multi-consumer ring in
 * caller-provided memory (a shared-memory segment mapped by the submitting process and by every model-actor process).
 * A request is its raw bytes (uint8 HWC image) in a fixed-size slot plus a 128-byte descriptor; a consumer owns the
 * slots it popped until it releases them, so queued or in-flight requests are never overwritten.  Host code only. */
#define B200_RING_ID_BYTES 100
/* Bytes of memory a ring of `nslots` slots of `slot_bytes` needs (header + descriptors + payload area). */
B200_API size_t b200_ring_bytes(uint32_t nslots, uint32_t slot_bytes);
/* RequestQueue.__init__ (scheduler.py:193-222): max_queued is its max_size (2000 at scheduler.py:632); nslots >=
 * max_queued leaves room for requests that were popped but whose batch is still being copied to a GPU.
 * `mem` 128-byte aligned, zero or garbage; call once, before any other process uses the ring. */
B200_API int b200_ring_init(void* mem, size_t bytes, uint32_t nslots, uint32_t slot_bytes, uint32_t max_queued);
B200_API int b200_ring_layout(void* mem, uint32_t* nslots, uint32_t* slot_bytes, uint32_t* max_queued, uint64_t* payload_off,
                              uint64_t* total_bytes);
/* RequestQueue.add_request (scheduler.py:238-254): 1 = queued, 0 = queue full (dropped and counted), < 0 = misuse.
 * `arrival` is the caller's time.time() (scheduler.py:246). */
B200_API int b200_ring_push(void* mem, const char* request_id, const void* payload, uint32_t nbytes, int h, int w, int c,
                            double arrival);
/* The same for n equal-size requests `stride` bytes apart (an ingress thread that received a burst); ids: n entries of
 * B200_RING_ID_BYTES, or NULL for "<id_prefix><first_index + i>".  Returns how many were queued. */
B200_API int b200_ring_push_many(void* mem, int n, const char* ids, const char* id_prefix, long long first_index,
                                 const void* payloads, size_t stride, uint32_t nbytes, int h, int w, int c, double arrival);
/* One round of RequestQueue.get_batch (scheduler.py:275-289): claims up to max_items requests in FIFO order, drops the
 * ones for which  arrival + slo_s < now + expected_latency_s  (scheduler.py:281; counted in *discarded) and returns
 * the others: slot numbers (payload of slot s at payload_off + s * slot_bytes), sequence numbers (for release), arrival
 * times, [h, w] pairs (w negated for a single-channel request, the convention b200_collate takes), byte counts, ids
 * (B200_RING_ID_BYTES each).  Any output except slots / seqs may be NULL. */
B200_API int b200_ring_pop(void* mem, int max_items, double now, double expected_latency_s, double slo_s, uint32_t* slots,
                           uint64_t* seqs, double* arrivals, int32_t* hw, uint32_t* nbytes, char* ids, int* discarded);
/* Gives popped slots back to the producers (after the batch's host-to-device copy has completed). */
B200_API int b200_ring_release(void* mem, const uint64_t* seqs, int n);
B200_API long long b200_ring_qsize(void* mem);
/* RequestQueue.record_batch_completion (scheduler.py:324-341): returns the number of SLO violations recorded. */
B200_API int b200_ring_record(void* mem, double latency_ms, int n_requests, double slo_ms);
/* RequestQueue.get_stats (scheduler.py:343-372) over the same window of latencies the reference's list would hold:
 * out[0..10] = total_requests, dropped_requests, slo_violations, queue_size, last_latency, avg_latency, p95_latency,
 * p99_latency, discarded_stale, completed_requests, completed_batches, and with n_out >= 12 out[11] = the largest batch
 * latency recorded since init (the window above holds only the last 900-1000 requests).  Visible to every process that maps the ring
 * (the reference's actors update a pickled copy of the metrics that the driver never sees). */
B200_API int b200_ring_stats(void* mem, double* out, int n_out);
/* The latency window those percentiles are taken over (the reference's metrics['latencies'] list, scheduler.py:333-336),
 * oldest first, at most `cap` values; returns the count.  For merging the windows of a queue sharded over several rings. */
B200_API int b200_ring_latencies(void* mem, double* out, int cap);

#ifdef __cplusplus
}
#endif
#endif /* B200INFER_H_ */
