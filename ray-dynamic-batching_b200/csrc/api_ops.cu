// C-ABI, operator level: one entry per kernel so each can be parity-tested on its own (tests/ call these via ctypes).
// The engine-level entries (b200_engine_*) live in engine.cu and compose the same launchers.
#include "../../include/b200infer.h"
#include "common.h"
#include "conv_gemm.cuh"

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

}  // extern "C"
