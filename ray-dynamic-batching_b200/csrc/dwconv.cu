// Depthwise 3x3 convolution (stride 1 / 2, pad 1) + folded BatchNorm bias + optional activation on NHWC bf16.
// Bandwidth-bound (9 MAC per output element): one thread owns 8 channels (16 B) of one output pixel, so a warp reads
// 512 contiguous bytes per tap; the 3x3 halo re-reads hit L1/L2.  fp32 weights [9][C] and fp32 accumulation.
//
// Replaces the grouped nn.Conv2d(groups=C)+BatchNorm2d of tv shufflenetv2.py:62-76,86-96 (`depthwise_conv`) and the
// depthwise stage of MBConv (tv efficientnet.py:140-151).
#include "kernels.cuh"
#include "ptx.cuh"

namespace b200 {

__global__ void __launch_bounds__(256)
dwconv3x3_kernel(const uint16_t* __restrict__ in, int H, int W, int in_ps, uint16_t* __restrict__ out, int Ho, int Wo,
                 int out_ps, int C8, int stride, const float* __restrict__ w, const float* __restrict__ bias, int act,
                 long long total) {
  const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int c8 = static_cast<int>(idx % C8);
  long long pix = idx / C8;
  const int ow = static_cast<int>(pix % Wo);
  pix /= Wo;
  const int oh = static_cast<int>(pix % Ho);
  const int n = static_cast<int>(pix / Ho);
  const int C = C8 * 8;
  float acc[8];
  {
    const float4 b0 = *reinterpret_cast<const float4*>(bias + c8 * 8);
    const float4 b1 = *reinterpret_cast<const float4*>(bias + c8 * 8 + 4);
    acc[0] = b0.x; acc[1] = b0.y; acc[2] = b0.z; acc[3] = b0.w;
    acc[4] = b1.x; acc[5] = b1.y; acc[6] = b1.z; acc[7] = b1.w;
  }
  const int y0 = oh * stride - 1, x0 = ow * stride - 1;
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const int y = y0 + r;
    if (y < 0 || y >= H) continue;
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      const int x = x0 + q;
      if (x < 0 || x >= W) continue;
      const uint4 v = *reinterpret_cast<const uint4*>(in + ((static_cast<size_t>(n) * H + y) * W + x) * in_ps + c8 * 8);
      const float4 w0 = *reinterpret_cast<const float4*>(w + (r * 3 + q) * C + c8 * 8);
      const float4 w1 = *reinterpret_cast<const float4*>(w + (r * 3 + q) * C + c8 * 8 + 4);
      acc[0] = fmaf(bf16_lo(v.x), w0.x, acc[0]); acc[1] = fmaf(bf16_hi(v.x), w0.y, acc[1]);
      acc[2] = fmaf(bf16_lo(v.y), w0.z, acc[2]); acc[3] = fmaf(bf16_hi(v.y), w0.w, acc[3]);
      acc[4] = fmaf(bf16_lo(v.z), w1.x, acc[4]); acc[5] = fmaf(bf16_hi(v.z), w1.y, acc[5]);
      acc[6] = fmaf(bf16_lo(v.w), w1.z, acc[6]); acc[7] = fmaf(bf16_hi(v.w), w1.w, acc[7]);
    }
  }
  uint4 o;
  o.x = pack_bf16x2(apply_act(acc[0], act), apply_act(acc[1], act));
  o.y = pack_bf16x2(apply_act(acc[2], act), apply_act(acc[3], act));
  o.z = pack_bf16x2(apply_act(acc[4], act), apply_act(acc[5], act));
  o.w = pack_bf16x2(apply_act(acc[6], act), apply_act(acc[7], act));
  *reinterpret_cast<uint4*>(out + ((static_cast<size_t>(n) * Ho + oh) * Wo + ow) * out_ps + c8 * 8) = o;
}

int dwconv3x3_launch(const View& in, const View& out, const float* w, const float* bias, int stride, int act,
                     float* se_sum, cudaStream_t stream) {
  if (se_sum) return fail(B200_ERR_INVALID, "dwconv: fused SE squeeze not implemented");
  if (in.C != out.C || in.C % 8 || in.pix_stride % 8 || out.pix_stride % 8)
    return fail(B200_ERR_INVALID, "dwconv: channels must match and be multiples of 8");
  if ((in.H + 2 - 3) / stride + 1 != out.H || (in.W + 2 - 3) / stride + 1 != out.W || in.N != out.N)
    return fail(B200_ERR_INVALID, "dwconv: output geometry mismatch");
  const long long total = static_cast<long long>(out.N) * out.H * out.W * (out.C / 8);
  dwconv3x3_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, stream>>>(
      static_cast<const uint16_t*>(in.base), in.H, in.W, in.pix_stride, static_cast<uint16_t*>(out.base), out.H, out.W,
      out.pix_stride, out.C / 8, stride, w, bias, act, total);
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}

}  // namespace b200
