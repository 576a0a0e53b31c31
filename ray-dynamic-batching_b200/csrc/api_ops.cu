// C-ABI, operator level: one entry per kernel so each can be parity-tested on its own (tests/ call these via ctypes).
// The engine-level entries (b200_engine_*) live in engine.cu and compose the same launchers.
#include "../../include/b200infer.h"
#include "common.h"
#include "conv_gemm.cuh"
#include "kernels.cuh"

using namespace b200;

extern "C" {

const char* b200_last_error(void) { return last_error_cstr(); }

int b200_abi_version(void) { return B200_ABI_VERSION; }

int b200_conv2d_nhwc(const void* in, int n, int h, int w, int cin, int in_pix_stride, const void* weight,
                     const float* bias, const void* residual, int res_pix_stride, void* out, int cout,
                     int out_pix_stride, int ksize, int stride, int pad, int act, void* stream) {
  if (n <= 0 || h <= 0 || w <= 0) return fail(B200_ERR_INVALID, "b200_conv2d_nhwc: empty input");
  const int ho = (h + 2 * pad - ksize) / stride + 1, wo = (w + 2 * pad - ksize) / stride + 1;
  ConvSpec s;
  s.in = make_view(const_cast<void*>(in), n, h, w, cin, in_pix_stride);
  s.out = make_view(out, n, ho, wo, cout, out_pix_stride);
  if (residual) s.res = make_view(const_cast<void*>(residual), n, ho, wo, cout, res_pix_stride);
  s.w = weight;
  s.bias = bias;
  s.ksize = ksize;
  s.stride = stride;
  s.pad = pad;
  s.act = act;
  ConvLaunch l;
  B200_TRY(conv_prepare(s, &l));
  return conv_launch(l, static_cast<cudaStream_t>(stream));
}

int b200_conv1x1_dual_nhwc(const void* in, int n, int h, int w, int cin, int in_pix_stride, const void* in2, int h2, int w2,
                           int cin2, int in2_pix_stride, int stride2, const void* weight, const float* bias, void* out,
                           int cout, int out_pix_stride, int act, void* stream) {
  if (n <= 0 || h <= 0 || w <= 0 || !in2) return fail(B200_ERR_INVALID, "b200_conv1x1_dual_nhwc: empty input");
  ConvSpec s;
  s.in = make_view(const_cast<void*>(in), n, h, w, cin, in_pix_stride);
  s.in2 = make_view(const_cast<void*>(in2), n, h2, w2, cin2, in2_pix_stride);
  s.stride2 = stride2;
  s.out = make_view(out, n, h, w, cout, out_pix_stride);
  s.w = weight;
  s.bias = bias;
  s.ksize = 1;
  s.stride = 1;
  s.pad = 0;
  s.act = act;
  ConvLaunch l;
  B200_TRY(conv_prepare(s, &l));
  return conv_launch(l, static_cast<cudaStream_t>(stream));
}

int b200_conv1x1_scaled_nhwc(const void* in, int n, int h, int w, int cin, int in_pix_stride, const float* in_scale,
                             const void* weight, const float* bias, const void* residual, int res_pix_stride, void* out,
                             int cout, int out_pix_stride, int act, void* stream) {
  if (n <= 0 || h <= 0 || w <= 0 || !in_scale) return fail(B200_ERR_INVALID, "b200_conv1x1_scaled_nhwc: empty input");
  ConvSpec s;
  s.in = make_view(const_cast<void*>(in), n, h, w, cin, in_pix_stride);
  s.a_scale = in_scale;
  s.out = make_view(out, n, h, w, cout, out_pix_stride);
  if (residual) s.res = make_view(const_cast<void*>(residual), n, h, w, cout, res_pix_stride);
  s.w = weight;
  s.bias = bias;
  s.ksize = 1;
  s.stride = 1;
  s.pad = 0;
  s.act = act;
  ConvLaunch l;
  B200_TRY(conv_prepare(s, &l));
  return conv_launch(l, static_cast<cudaStream_t>(stream));
}

int b200_conv1x1_shuffle_nhwc(const void* in, int n, int h, int w, int cin, int in_pix_stride, const void* other,
                              int other_pix_stride, int c, int cp, const void* weight, const float* bias, void* out,
                              int out_pix_stride, void* stream) {
  if (n <= 0 || h <= 0 || w <= 0 || !other || c <= 0) return fail(B200_ERR_INVALID, "b200_conv1x1_shuffle_nhwc: empty input");
  ConvSpec s;
  s.in = make_view(const_cast<void*>(in), n, h, w, cin, in_pix_stride);
  s.res = make_view(const_cast<void*>(other), n, h, w, cp, other_pix_stride);
  s.out = make_view(out, n, h, w, 2 * cp, out_pix_stride);
  s.interleave_c = c;
  s.w = weight;
  s.bias = bias;
  s.ksize = 1;
  s.stride = 1;
  s.pad = 0;
  s.act = 1;      // b200::ACT_RELU
  ConvLaunch l;
  B200_TRY(conv_prepare(s, &l));
  return conv_launch(l, static_cast<cudaStream_t>(stream));
}

int b200_gather_conv_npad(int cout) { return gather_conv_npad(cout); }
int b200_gather_conv_kpad(int cin, int ksize) { return gather_conv_kpad(cin, ksize); }

int b200_gather_conv2d_nhwc(const void* in, int n, int h, int w, int cin, int in_pix_stride, const void* weight,
                            const float* bias, const void* residual, int res_pix_stride, void* out, int cout,
                            int out_pix_stride, int ksize, int stride, int pad, int act, int res_after_act, void* stream) {
  const int ho = (h + 2 * pad - ksize) / stride + 1, wo = (w + 2 * pad - ksize) / stride + 1;
  View res;
  if (residual) res = make_view(const_cast<void*>(residual), n, ho, wo, cout, res_pix_stride);
  return gather_conv_launch(make_view(const_cast<void*>(in), n, h, w, cin, in_pix_stride),
                            make_view(out, n, ho, wo, cout, out_pix_stride), res, weight, bias, ksize, stride, pad, act,
                            res_after_act != 0, static_cast<cudaStream_t>(stream));
}

int b200_stem_cout_tile(int cout) { return stem_cout_tile(cout); }

int b200_stem_conv_nhwc(const void* in, int n, int h, int w, const float* weight, const float* bias, void* out, int cout,
                        int out_pix_stride, int ksize, int stride, int pad, int act, void* stream) {
  const int ho = (h + 2 * pad - ksize) / stride + 1, wo = (w + 2 * pad - ksize) / stride + 1;
  return stem_conv_launch(make_view(const_cast<void*>(in), n, h, w, 4, 4), make_view(out, n, ho, wo, cout, out_pix_stride),
                          weight, bias, ksize, stride, pad, act, static_cast<cudaStream_t>(stream));
}

int b200_stem_row_weight_bytes(void) { return stem_row_weight_bytes(); }

int b200_stem_row_conv_nhwc(const void* in, int n, int h, int w, const void* weight, const float* bias, void* out,
                            int out_pix_stride, void* stream) {
  const int ho = (h + 6 - 7) / 2 + 1, wo = (w + 6 - 7) / 2 + 1;
  return stem_row_launch(make_view(const_cast<void*>(in), n, h, w, 4, 4), make_view(out, n, ho, wo, 64, out_pix_stride),
                         weight, bias, static_cast<cudaStream_t>(stream));
}

int b200_stem_pool_nhwc(const void* in, int n, int h, int w, const void* weight, const float* bias, void* out,
                        int out_pix_stride, void* stream) {
  const int ho = ((h + 6 - 7) / 2 + 1) / 2, wo = ((w + 6 - 7) / 2 + 1) / 2;
  return stem_pool_launch(make_view(const_cast<void*>(in), n, h, w, 4, 4), make_view(out, n, ho, wo, 64, out_pix_stride),
                          weight, bias, static_cast<cudaStream_t>(stream));
}

int b200_stem_row_weight_bytes_k(int ksize, int cout) { return stem_row_weight_bytes_k(ksize, cout); }

int b200_stem_row_conv_k_nhwc(const void* in, int n, int h, int w, const void* weight, const float* bias, void* out,
                              int cout, int out_pix_stride, int ksize, int act, void* stream) {
  if (ksize != 7 && ksize != 3) return fail(B200_ERR_INVALID, "stem_row: kernel must be 7x7 or 3x3");
  const int pad = ksize / 2;
  const int ho = (h + 2 * pad - ksize) / 2 + 1, wo = (w + 2 * pad - ksize) / 2 + 1;
  return stem_row_launch_k(make_view(const_cast<void*>(in), n, h, w, 4, 4), make_view(out, n, ho, wo, cout, out_pix_stride),
                           weight, bias, ksize, act, static_cast<cudaStream_t>(stream));
}

int b200_stem_pool_k_nhwc(const void* in, int n, int h, int w, const void* weight, const float* bias, void* out,
                          int cout, int out_pix_stride, int ksize, int act, void* stream) {
  if (ksize != 7 && ksize != 3) return fail(B200_ERR_INVALID, "stem_pool: kernel must be 7x7 or 3x3");
  const int pad = ksize / 2;
  const int ho = ((h + 2 * pad - ksize) / 2 + 1) / 2, wo = ((w + 2 * pad - ksize) / 2 + 1) / 2;
  return stem_pool_launch_k(make_view(const_cast<void*>(in), n, h, w, 4, 4), make_view(out, n, ho, wo, cout, out_pix_stride),
                            weight, bias, ksize, act, static_cast<cudaStream_t>(stream));
}

int b200_stem_tc_kpad(int ksize) { return stem_tc_kpad(ksize); }
int b200_stem_tc_nout(int cout) { return stem_tc_nout(cout); }

int b200_stem_conv_tc_nhwc(const void* in, int n, int h, int w, const void* weight, const float* bias, void* out, int cout,
                           int out_pix_stride, int ksize, int stride, int pad, int act, void* stream) {
  const int ho = (h + 2 * pad - ksize) / stride + 1, wo = (w + 2 * pad - ksize) / stride + 1;
  return stem_tc_launch(make_view(const_cast<void*>(in), n, h, w, 4, 4), make_view(out, n, ho, wo, cout, out_pix_stride),
                        weight, bias, ksize, stride, pad, act, static_cast<cudaStream_t>(stream));
}

int b200_dwconv3x3_nhwc(const void* in, int n, int h, int w, int c, int in_pix_stride, const float* weight,
                        const float* bias, void* out, int out_pix_stride, int stride, int act, void* stream) {
  const int ho = (h - 1) / stride + 1, wo = (w - 1) / stride + 1;
  return dwconv3x3_launch(make_view(const_cast<void*>(in), n, h, w, c, in_pix_stride),
                          make_view(out, n, ho, wo, c, out_pix_stride), weight, bias, stride, act, nullptr,
                          static_cast<cudaStream_t>(stream));
}

int b200_dwconv3x3_strips(int ho, int wo, int stride) { return dwconv3x3_strips(ho, wo, stride); }
int b200_dwconv3x3_squeeze_rows(int c, int ho, int wo, int stride) { return dwconv3x3_strips_c(c, ho, wo, stride); }

int b200_dwconv3x3_squeeze_nhwc(const void* in, int n, int h, int w, int c, int in_pix_stride, const float* weight,
                                const float* bias, void* out, int out_pix_stride, int stride, int act,
                                float* strip_sums, void* stream) {
  if (!strip_sums) return fail(B200_ERR_INVALID, "dwconv3x3_squeeze: strip_sums is null");
  if (stride != 1 && stride != 2) return fail(B200_ERR_INVALID, "dwconv3x3_squeeze: stride must be 1 or 2");
  const int ho = (h - 1) / stride + 1, wo = (w - 1) / stride + 1;
  return dwconv3x3_launch(make_view(const_cast<void*>(in), n, h, w, c, in_pix_stride),
                          make_view(out, n, ho, wo, c, out_pix_stride), weight, bias, stride, act, strip_sums,
                          static_cast<cudaStream_t>(stream));
}

int b200_se_gate(const float* strip_sums, int n, int strips, int c, float inv_hw, const float* w1, const float* b1,
                 const float* w2t, const float* b2, int squeeze, float* gate, void* stream) {
  return se_gate_launch(strip_sums, n, strips, c, inv_hw, w1, b1, w2t, b2, squeeze, gate,
                        static_cast<cudaStream_t>(stream));
}

int b200_se_hidden(const float* strip_sums, int n, int rows, int c, float inv_hw, const float* w1, const float* b1,
                   int squeeze, float* hidden, void* stream) {
  return se_hidden_launch(strip_sums, n, rows, c, inv_hw, w1, b1, squeeze, hidden, static_cast<cudaStream_t>(stream));
}

int b200_se_scale_nhwc(const void* in, int n, int h, int w, int c, int in_pix_stride, const float* hidden,
                       const float* w2t, const float* b2, int squeeze, void* out, int out_pix_stride, void* stream) {
  return se_scale_launch(make_view(const_cast<void*>(in), n, h, w, c, in_pix_stride),
                         make_view(out, n, h, w, c, out_pix_stride), hidden, w2t, b2, squeeze,
                         static_cast<cudaStream_t>(stream));
}

int b200_maxpool_nhwc(const void* in, int n, int h, int w, int c, int in_pix_stride, void* out, int out_pix_stride,
                      int ksize, int stride, int pad, void* stream) {
  const int ho = (h + 2 * pad - ksize) / stride + 1, wo = (w + 2 * pad - ksize) / stride + 1;
  return maxpool_launch(make_view(const_cast<void*>(in), n, h, w, c, in_pix_stride),
                        make_view(out, n, ho, wo, c, out_pix_stride), ksize, stride, pad,
                        static_cast<cudaStream_t>(stream));
}

int b200_global_avgpool_nhwc(const void* in, int n, int h, int w, int c, int in_pix_stride, float* out, void* stream) {
  return global_avgpool_launch(make_view(const_cast<void*>(in), n, h, w, c, in_pix_stride), out,
                               static_cast<cudaStream_t>(stream));
}

int b200_fc(const float* x, const float* w, const float* b, float* out, int n, int k, int j, float x_scale, int act,
            void* stream) {
  return fc_launch(x, w, b, out, n, k, j, x_scale, act, static_cast<cudaStream_t>(stream));
}

int b200_channel_scale_nhwc(const void* in, int n, int h, int w, int c, int in_pix_stride, const float* gate, void* out,
                            int out_pix_stride, void* stream) {
  return channel_scale_launch(make_view(const_cast<void*>(in), n, h, w, c, in_pix_stride),
                              make_view(out, n, h, w, c, out_pix_stride), gate, static_cast<cudaStream_t>(stream));
}

int b200_copy_channels_nhwc(const void* in, int n, int h, int w, int c, int in_pix_stride, void* out, int out_pix_stride,
                            void* stream) {
  return copy_channels_launch(make_view(const_cast<void*>(in), n, h, w, c, in_pix_stride),
                              make_view(out, n, h, w, c, out_pix_stride), static_cast<cudaStream_t>(stream));
}

int b200_upsample2x_nhwc(const void* in, int n, int h, int w, int c, int in_pix_stride, void* out, int out_pix_stride,
                         void* stream) {
  return upsample2x_launch(make_view(const_cast<void*>(in), n, h, w, c, in_pix_stride),
                           make_view(out, n, 2 * h, 2 * w, c, out_pix_stride), static_cast<cudaStream_t>(stream));
}

int b200_shuffle_concat_nhwc(const void* a, int a_pix_stride, const void* b, int b_pix_stride, int n, int h, int w,
                             int c, void* out, int out_pix_stride, void* stream) {
  const int cp = (c + 7) / 8 * 8;
  return shuffle_concat_launch(make_view(const_cast<void*>(a), n, h, w, cp, a_pix_stride),
                               make_view(const_cast<void*>(b), n, h, w, cp, b_pix_stride),
                               make_view(out, n, h, w, 2 * cp, out_pix_stride), c, cp, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
