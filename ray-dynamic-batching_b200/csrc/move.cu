// Pure data-movement kernels on NHWC bf16 channel slices: slice copy, nearest 2x upsample into a concat slice, and the
// ShuffleNetV2 concat + channel_shuffle(groups=2) interleave.  16 bytes per thread, coalesced.
//
// Replaces torch.cat / x.chunk(2) / channel_shuffle (tv shufflenetv2.py:28-40,98-104) and nn.Upsample(2,'nearest') +
// Concat of the YOLOv8 head.
#include "kernels.cuh"
#include "ptx.cuh"

namespace b200 {

__global__ void __launch_bounds__(256)
copy_channels_kernel(const uint16_t* __restrict__ in, int in_ps, uint16_t* __restrict__ out, int out_ps, int C8,
                     long long total) {
  pdl_wait();   // launched with programmatic stream serialization: inputs are valid from here on (no early trigger:
                // a multi-wave kernel must not share its SMs with the waiting blocks of its successor)
  const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int c8 = static_cast<int>(idx % C8);
  const long long pix = idx / C8;
  *reinterpret_cast<uint4*>(out + pix * out_ps + c8 * 8) = *reinterpret_cast<const uint4*>(in + pix * in_ps + c8 * 8);
}

int copy_channels_launch(const View& in, const View& out, cudaStream_t stream) {
  if (in.C != out.C || in.C % 8 || in.N != out.N || in.H != out.H || in.W != out.W)
    return fail(B200_ERR_INVALID, "copy_channels: shape mismatch");
  const long long total = static_cast<long long>(in.N) * in.H * in.W * (in.C / 8);
  B200_CUDA_OK(launch_pdl(copy_channels_kernel, dim3(static_cast<unsigned>((total + 255) / 256)), dim3(256), size_t(0), stream, static_cast<const uint16_t*>(in.base), in.pix_stride, static_cast<uint16_t*>(out.base), out.pix_stride, in.C / 8, total));
  return B200_OK;
}

__global__ void __launch_bounds__(256)
upsample2x_kernel(const uint16_t* __restrict__ in, int H, int W, int in_ps, uint16_t* __restrict__ out, int out_ps,
                  int C8, long long total) {
  pdl_wait();   // launched with programmatic stream serialization: inputs are valid from here on (no early trigger:
                // a multi-wave kernel must not share its SMs with the waiting blocks of its successor)
  const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int c8 = static_cast<int>(idx % C8);
  long long pix = idx / C8;
  const int ow = static_cast<int>(pix % (2 * W));
  pix /= 2 * W;
  const int oh = static_cast<int>(pix % (2 * H));
  const int n = static_cast<int>(pix / (2 * H));
  const uint4 v =
      *reinterpret_cast<const uint4*>(in + ((static_cast<size_t>(n) * H + (oh >> 1)) * W + (ow >> 1)) * in_ps + c8 * 8);
  *reinterpret_cast<uint4*>(out + ((static_cast<size_t>(n) * 2 * H + oh) * 2 * W + ow) * out_ps + c8 * 8) = v;
}

int upsample2x_launch(const View& in, const View& out, cudaStream_t stream) {
  if (in.C != out.C || in.C % 8 || in.N != out.N || 2 * in.H != out.H || 2 * in.W != out.W)
    return fail(B200_ERR_INVALID, "upsample2x: shape mismatch");
  const long long total = static_cast<long long>(out.N) * out.H * out.W * (out.C / 8);
  B200_CUDA_OK(launch_pdl(upsample2x_kernel, dim3(static_cast<unsigned>((total + 255) / 256)), dim3(256), size_t(0), stream, static_cast<const uint16_t*>(in.base), in.H, in.W, in.pix_stride, static_cast<uint16_t*>(out.base), out.pix_stride, in.C / 8, total));
  return B200_OK;
}

// ShuffleNetV2 concat + channel_shuffle(groups=2): logical out[2i] = a[i], out[2i+1] = b[i] (i < c).  The output keeps
// the two halves of the *shuffled* tensor in separately padded slices (logical l -> physical l < c ? l : cp + l - c), so the
// next unit's chunk(2) needs no copy.  A thread assembles 8 output channels (16 B store) from 2-byte gathers that hit L1.
__global__ void __launch_bounds__(256)
shuffle_concat_kernel(const uint16_t* __restrict__ a, int a_ps, const uint16_t* __restrict__ b, int b_ps,
                      uint16_t* __restrict__ out, int out_ps, int c, int cp, long long total) {
  pdl_wait();   // launched with programmatic stream serialization: inputs are valid from here on (no early trigger:
                // a multi-wave kernel must not share its SMs with the waiting blocks of its successor)
  const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int chunks = (2 * cp) >> 3;
  const int ch = static_cast<int>(idx % chunks);
  const long long pix = idx / chunks;
  const uint16_t* pa = a + pix * a_ps;
  const uint16_t* pb = b + pix * b_ps;
  uint16_t v[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int p = ch * 8 + e;
    const int half = p >= cp ? 1 : 0;
    const int q = p - half * cp;
    uint16_t x = 0;
    if (q < c) {
      const int l = half * c + q;
      x = (l & 1) ? pb[l >> 1] : pa[l >> 1];
    }
    v[e] = x;
  }
  uint4 o;
  o.x = uint32_t(v[0]) | (uint32_t(v[1]) << 16); o.y = uint32_t(v[2]) | (uint32_t(v[3]) << 16);
  o.z = uint32_t(v[4]) | (uint32_t(v[5]) << 16); o.w = uint32_t(v[6]) | (uint32_t(v[7]) << 16);
  *reinterpret_cast<uint4*>(out + pix * out_ps + ch * 8) = o;
}

int shuffle_concat_launch(const View& a, const View& b, const View& out, int c, int cp, cudaStream_t stream) {
  if (cp % 8 || c > cp || c <= 0 || a.C < c || b.C < c || out.C != 2 * cp || a.N != out.N || a.H != out.H ||
      a.W != out.W || b.N != a.N || b.H != a.H || b.W != a.W || out.pix_stride % 8)
    return fail(B200_ERR_INVALID, "shuffle_concat: shape mismatch");
  const long long total = static_cast<long long>(a.N) * a.H * a.W * (2 * cp / 8);
  B200_CUDA_OK(launch_pdl(shuffle_concat_kernel, dim3(static_cast<unsigned>((total + 255) / 256)), dim3(256), size_t(0), stream, static_cast<const uint16_t*>(a.base), a.pix_stride, static_cast<const uint16_t*>(b.base), b.pix_stride, static_cast<uint16_t*>(out.base), out.pix_stride, c, cp, total));
  return B200_OK;
}

}  // namespace b200
