// Network stems: Conv2d with 3 input channels (+ folded BatchNorm + ReLU/SiLU) as a direct convolution on the FP32 pipes.
// With K = k*k*3 (147 or 27) the GEMM is too thin for a tcgen05 tile pipeline fed by TMA boxes, and the layer is 3 % of
// ResNet-50's FLOPs; inputs are staged once per 16x16 output tile in shared memory (fp32), weights are broadcast from
// shared memory, every thread keeps 2 pixels x CT output channels in registers.
//
// Replaces torchvision's stem as reached from scheduler.py:450:  resnet.py:268-270 (conv1/bn1/relu),
// shufflenetv2.py:112-116 (conv1), efficientnet.py:283-287 (features[0]); YOLOv8 model.0.
#include "kernels.cuh"
#include "ptx.cuh"

namespace b200 {

static constexpr int ST_TW = 16, ST_TH = 16;  // output tile; thread t -> pixels (t/16, t%16) and (t/16 + 8, t%16)

template <int K, int S, int CT>
__global__ void __launch_bounds__(256)
stem_conv_kernel(const uint16_t* __restrict__ in, int H, int W, int in_ps, uint16_t* __restrict__ out, int Ho, int Wo,
                 int out_ps, int cout, const float* __restrict__ w, const float* __restrict__ bias, int pad, int act) {
  constexpr int IW = (ST_TW - 1) * S + K, IH = (ST_TH - 1) * S + K;
  extern __shared__ __align__(16) uint8_t smem[];
  float4* s_in = reinterpret_cast<float4*>(smem);                    // [IH][IW]
  float* s_w = reinterpret_cast<float*>(smem + IH * IW * 16);        // [K*K*3][cout_pad]
  const int groups = blockDim.x / 128;
  const int cout_pad = groups * CT;
  const int n = blockIdx.z;
  const int oh0 = blockIdx.y * ST_TH, ow0 = blockIdx.x * ST_TW;
  const int ih0 = oh0 * S - pad, iw0 = ow0 * S - pad;

  for (int i = threadIdx.x; i < K * K * 3 * cout_pad; i += blockDim.x) s_w[i] = w[i];
  const uint2* src = reinterpret_cast<const uint2*>(in);  // one pixel (4 x bf16) = 8 bytes when in_ps == 4
  for (int i = threadIdx.x; i < IH * IW; i += blockDim.x) {
    const int y = i / IW, x = i - y * IW;
    const int iy = ih0 + y, ix = iw0 + x;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (iy >= 0 && iy < H && ix >= 0 && ix < W) {
      const uint2 p = src[(static_cast<size_t>(n) * H + iy) * W + ix];
      v = make_float4(bf16_lo(p.x), bf16_hi(p.x), bf16_lo(p.y), 0.f);
    }
    s_in[i] = v;
  }
  __syncthreads();

  const int t = threadIdx.x & 127, cg = threadIdx.x >> 7;
  const int tx = t & 15, ty = t >> 4;
  float acc0[CT], acc1[CT];
#pragma unroll
  for (int j = 0; j < CT; ++j) {
    const int c = cg * CT + j;
    const float b = c < cout ? bias[c] : 0.f;
    acc0[j] = b; acc1[j] = b;
  }
  const float4* p0 = s_in + (ty * S) * IW + tx * S;
  const float4* p1 = s_in + ((ty + 8) * S) * IW + tx * S;
  const float* wp = s_w + cg * CT;
#pragma unroll 1
  for (int r = 0; r < K; ++r) {
#pragma unroll
    for (int q = 0; q < K; ++q) {
      const float4 a0 = p0[r * IW + q];
      const float4 a1 = p1[r * IW + q];
      const float* wt = wp + ((r * K + q) * 3) * cout_pad;
#pragma unroll
      for (int j4 = 0; j4 < CT / 4; ++j4) {
        const float4 w0 = *reinterpret_cast<const float4*>(wt + j4 * 4);
        const float4 w1 = *reinterpret_cast<const float4*>(wt + cout_pad + j4 * 4);
        const float4 w2 = *reinterpret_cast<const float4*>(wt + 2 * cout_pad + j4 * 4);
        float* c0 = acc0 + j4 * 4;
        float* c1 = acc1 + j4 * 4;
        c0[0] = fmaf(a0.x, w0.x, c0[0]); c0[1] = fmaf(a0.x, w0.y, c0[1]); c0[2] = fmaf(a0.x, w0.z, c0[2]); c0[3] = fmaf(a0.x, w0.w, c0[3]);
        c1[0] = fmaf(a1.x, w0.x, c1[0]); c1[1] = fmaf(a1.x, w0.y, c1[1]); c1[2] = fmaf(a1.x, w0.z, c1[2]); c1[3] = fmaf(a1.x, w0.w, c1[3]);
        c0[0] = fmaf(a0.y, w1.x, c0[0]); c0[1] = fmaf(a0.y, w1.y, c0[1]); c0[2] = fmaf(a0.y, w1.z, c0[2]); c0[3] = fmaf(a0.y, w1.w, c0[3]);
        c1[0] = fmaf(a1.y, w1.x, c1[0]); c1[1] = fmaf(a1.y, w1.y, c1[1]); c1[2] = fmaf(a1.y, w1.z, c1[2]); c1[3] = fmaf(a1.y, w1.w, c1[3]);
        c0[0] = fmaf(a0.z, w2.x, c0[0]); c0[1] = fmaf(a0.z, w2.y, c0[1]); c0[2] = fmaf(a0.z, w2.z, c0[2]); c0[3] = fmaf(a0.z, w2.w, c0[3]);
        c1[0] = fmaf(a1.z, w2.x, c1[0]); c1[1] = fmaf(a1.z, w2.y, c1[1]); c1[2] = fmaf(a1.z, w2.z, c1[2]); c1[3] = fmaf(a1.z, w2.w, c1[3]);
      }
    }
  }
  const int ow = ow0 + tx;
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    const int oh = oh0 + ty + half * 8;
    if (oh >= Ho || ow >= Wo) continue;
    const float* a = half ? acc1 : acc0;
    uint16_t* dst = out + ((static_cast<size_t>(n) * Ho + oh) * Wo + ow) * out_ps + cg * CT;
#pragma unroll
    for (int j8 = 0; j8 < CT / 8; ++j8) {
      if (cg * CT + j8 * 8 >= cout) break;
      uint4 o;
      o.x = pack_bf16x2(apply_act(a[j8 * 8 + 0], act), apply_act(a[j8 * 8 + 1], act));
      o.y = pack_bf16x2(apply_act(a[j8 * 8 + 2], act), apply_act(a[j8 * 8 + 3], act));
      o.z = pack_bf16x2(apply_act(a[j8 * 8 + 4], act), apply_act(a[j8 * 8 + 5], act));
      o.w = pack_bf16x2(apply_act(a[j8 * 8 + 6], act), apply_act(a[j8 * 8 + 7], act));
      *reinterpret_cast<uint4*>(dst + j8 * 8) = o;
    }
  }
}

template <int K, int S, int CT>
static int stem_launch_t(const View& in, const View& out, const float* w, const float* bias, int pad, int act,
                         cudaStream_t stream) {
  constexpr int IW = (ST_TW - 1) * S + K, IH = (ST_TH - 1) * S + K;
  const int groups = (out.C + CT - 1) / CT;
  const size_t smem = size_t(IH) * IW * 16 + size_t(K) * K * 3 * groups * CT * 4;
  static bool configured[64] = {};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(stem_conv_kernel<K, S, CT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    configured[dev & 63] = true;
  }
  dim3 grid((out.W + ST_TW - 1) / ST_TW, (out.H + ST_TH - 1) / ST_TH, out.N);
  stem_conv_kernel<K, S, CT><<<grid, 128 * groups, smem, stream>>>(
      static_cast<const uint16_t*>(in.base), in.H, in.W, in.pix_stride, static_cast<uint16_t*>(out.base), out.H, out.W,
      out.pix_stride, out.C, w, bias, pad, act);
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}

// weights: fp32 [k*k][3][cout_pad] with cout_pad = ceil(cout / CT) * CT (CT = 32 for cout >= 32 and cout % 32 == 0,
// 24 for cout == 24, else 16); see stem_cout_pad().
int stem_cout_tile(int cout) { return (cout % 32 == 0) ? 32 : (cout % 24 == 0 ? 24 : 16); }

int stem_conv_launch(const View& in, const View& out, const float* w, const float* bias, int ksize, int stride, int pad,
                     int act, cudaStream_t stream) {
  if (in.C != 4 || in.pix_stride != 4) return fail(B200_ERR_INVALID, "stem: input must be NHWC with C padded to 4");
  if (out.C % 8 || out.pix_stride % 8) return fail(B200_ERR_INVALID, "stem: cout must be a multiple of 8");
  const int ct = stem_cout_tile(out.C);
  const int groups = (out.C + ct - 1) / ct;
  if (groups > 2) return fail(B200_ERR_INVALID, "stem: cout too large for the direct kernel");
  if ((in.H + 2 * pad - ksize) / stride + 1 != out.H || (in.W + 2 * pad - ksize) / stride + 1 != out.W)
    return fail(B200_ERR_INVALID, "stem: output geometry mismatch");
  if (ksize == 7 && stride == 2 && ct == 32) return stem_launch_t<7, 2, 32>(in, out, w, bias, pad, act, stream);
  if (ksize == 3 && stride == 2 && ct == 32) return stem_launch_t<3, 2, 32>(in, out, w, bias, pad, act, stream);
  if (ksize == 3 && stride == 2 && ct == 24) return stem_launch_t<3, 2, 24>(in, out, w, bias, pad, act, stream);
  if (ksize == 3 && stride == 2 && ct == 16) return stem_launch_t<3, 2, 16>(in, out, w, bias, pad, act, stream);
  return fail(B200_ERR_INVALID, "stem: unsupported (ksize, stride, cout) combination");
}

}  // namespace b200
