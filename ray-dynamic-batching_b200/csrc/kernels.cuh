// Host-side launchers of the bandwidth-bound kernels (collate, small-Cin stem conv, depthwise conv, pooling, FC, SE,
// upsample/concat, YOLO head decode).  All activations NHWC bf16 described by b200::View.
#pragma once
#include "common.h"

namespace b200 {

// ---- collate (collate.cu) ----
struct CollateItem {
  uint64_t offset;  // byte offset of the image inside the staged uint8 buffer (multiple of 16 for the fast path)
  int32_t h, w;     // source height / width (HWC, 3 channels)
};
enum : int { COLLATE_IDENTITY = 0, COLLATE_PAD = 1, COLLATE_RESIZE = 2, COLLATE_LETTERBOX = 3, COLLATE_RESIZE_AA = 4,
              COLLATE_RESIZE_SHORT_AA = 5 };
void build_norm_lut(const float mean[3], const float stdv[3], uint16_t lut[3 * 256]);
int collate_u8_launch(const uint8_t* src_base, const CollateItem* items_dev, const CollateItem* items_host, int n,
                      int padded_batch, int out_h, int out_w, int mode, int fill, const uint16_t* lut_dev, void* out,
                      cudaStream_t stream);
int collate_f32chw_launch(const float* src, int n, int padded_batch, int h, int w, void* out, cudaStream_t stream);

// ---- small-Cin direct convolution for the network stems (stem.cu): in C==4 (3 real), fp32 weights [k*k][4][cout] ----
int stem_cout_tile(int cout);
int stem_conv_launch(const View& in, const View& out, const float* w, const float* bias, int ksize, int stride, int pad,
                     int act, cudaStream_t stream);

// tensor-core variant (stem_tc.cu): bf16 weights [stem_tc_nout(cout)][stem_tc_kpad(k)], K index = r*ROW + (s+1)*4 + c
int stem_tc_kpad(int ksize);
int stem_tc_nout(int cout);
int stem_tc_launch(const View& in, const View& out, const void* w, const float* bias, int ksize, int stride, int pad,
                   int act, cudaStream_t stream);

// Toeplitz-descriptor ResNet stem (stem_row.cu): 7x7/2 pad 3, cout 64, ReLU; weights = bf16 smem image (stem_row_weight_bytes())
int stem_row_weight_bytes();
int stem_row_launch(const View& in, const View& out, const void* w, const float* bias, cudaStream_t stream);
// fused conv1 + bn1 + relu + maxpool(3,2,1): `out` is the pooled tensor
int stem_pool_launch(const View& in, const View& out, const void* w, const float* bias, cudaStream_t stream);
// the same two kernels for any served stride-2 stem: ksize 7 (pad 3) or 3 (pad 1), cout <= 64 (multiple of 8),
// act none / relu / silu; weights = bf16 image [ksize][ksize == 7 ? 4 : 2][cout <= 32 ? 32 : 64][8]
int stem_row_weight_bytes_k(int ksize, int cout);
int stem_row_launch_k(const View& in, const View& out, const void* w, const float* bias, int ksize, int act,
                      cudaStream_t stream);
int stem_pool_launch_k(const View& in, const View& out, const void* w, const float* bias, int ksize, int act,
                       cudaStream_t stream);

// ---- packed-K gather convolution for small Cin (gather_conv.cu): bf16 weights [npad][kpad], K index = tap*cin_p + c ----
int gather_conv_npad(int cout);
int gather_conv_kpad(int cin_p, int ksize);
int gather_conv_launch(const View& in, const View& out, const View& res, const void* w, const float* bias, int ksize,
                       int stride, int pad, int act, bool res_after_act, cudaStream_t stream);

// ---- depthwise 3x3 (dwconv.cu): fp32 weights [9][C], bias [C]; optional fused squeeze: per-strip fp32 channel sums of
// the stored outputs, [N][dwconv3x3_strips()][C] (0 strips = this geometry has no strip kernel) ----
int dwconv3x3_strips(int Ho, int Wo, int stride);              // strip kernel only
int dwconv3x3_strips_c(int C, int Ho, int Wo, int stride);     // rows the fused squeeze writes for this layer (tile kernel: 1)
int dwconv3x3_launch(const View& in, const View& out, const float* w, const float* bias, int stride, int act,
                     float* se_partial /* or null */, cudaStream_t stream);

// ---- pooling (pool.cu) ----
int maxpool_launch(const View& in, const View& out, int ksize, int stride, int pad, cudaStream_t stream);
int global_avgpool_launch(const View& in, float* out /* [N][C] fp32 */, cudaStream_t stream);

// ---- heads (head.cu) ----
// out[n][j] = act(sum_k x[n][k] * w[j][k] + b[j]);  x fp32 [N][K] (optionally pre-scaled), w fp32 [J][K]
int fc_launch(const float* x, const float* w, const float* b, float* out, int n, int k, int j, float x_scale, int act,
              cudaStream_t stream);
// gate[n][c] = sigmoid(fc2(silu(fc1(mean)))) from the depthwise kernel's strip sums; w1 [S][C], w2t [S][C] (fc2 transposed)
int se_gate_launch(const float* partial, int n, int strips, int c, float inv_hw, const float* w1, const float* b1,
                   const float* w2t, const float* b2, int s, float* gate, cudaStream_t stream);
// the same block in two launches that never pull a whole weight matrix through one SM:
// hid[n][j] = silu(fc1(mean)) fp32 [N][S], then y = x * sigmoid(fc2(hid)) with the gate computed per 64-channel slice
int se_hidden_launch(const float* sums, int n, int rows, int c, float inv_hw, const float* w1, const float* b1, int s,
                     float* hid, cudaStream_t stream);
int se_scale_launch(const View& in, const View& out, const float* hid, const float* w2t, const float* b2, int s,
                    cudaStream_t stream);
// y = x * gate[n][c]  (SE excitation), gate fp32 already passed through sigmoid
int channel_scale_launch(const View& in, const View& out, const float* gate, cudaStream_t stream);

// ---- data movement (move.cu) ----
int copy_channels_launch(const View& in, const View& out, cudaStream_t stream);             // slice -> slice copy
int upsample2x_launch(const View& in, const View& out, cudaStream_t stream);                // nearest, into a slice
// logical out[2i] = a[i], out[2i+1] = b[i] (i < c); logical channel l is stored at physical (l < c ? l : cp + l - c)
int shuffle_concat_launch(const View& a, const View& b, const View& out, int c, int cp, cudaStream_t stream);

// ---- YOLOv8 head (yolo.cu) ----
// hb/hc: hidden activations of the box / class branch (64 / 80 channels); wb [64][64], wc [80][80] fp32 final 1x1 convs;
// out fp32 [N][84][a_total], this level's anchors at [a_off, a_off + H*W)
int yolo_decode_launch(const View& hb, const View& hc, const float* wb, const float* bb, const float* wc, const float* bc,
                       float* out, int a_off, int a_total, float stride, cudaStream_t stream);
int yolo_nms_launch(const float* pred, int batch, int nc, int a_total, float conf, float iou, int max_det, float* boxes,
                    int* counts, cudaStream_t stream);

}  // namespace b200
