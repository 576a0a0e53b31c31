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
  copy_channels_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, stream>>>(
      static_cast<const uint16_t*>(in.base), in.pix_stride, static_cast<uint16_t*>(out.base), out.pix_stride, in.C / 8,
      total);
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}

__global__ void __launch_bounds__(256)
upsample2x_kernel(const uint16_t* __restrict__ in, int H, int W, int in_ps, uint16_t* __restrict__ out, int out_ps,
                  int C8, long long total) {
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
  upsample2x_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, stream>>>(
      static_cast<const uint16_t*>(in.base), in.H, in.W, in.pix_stride, static_cast<uint16_t*>(out.base), out.pix_stride,
      in.C / 8, total);
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}

// out[..., 2i] = a[..., i]; out[..., 2i+1] = b[..., i]   == channel_shuffle(cat(a, b), groups=2).
// A thread reads 16 B (8 channels) from each of a and b and writes 32 contiguous bytes.
__global__ void __launch_bounds__(256)
shuffle_concat_kernel(const uint16_t* __restrict__ a, int a_ps, const uint16_t* __restrict__ b, int b_ps,
                      uint16_t* __restrict__ out, int out_ps, int C8, long long total) {
  const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int c8 = static_cast<int>(idx % C8);
  const long long pix = idx / C8;
  const uint4 va = *reinterpret_cast<const uint4*>(a + pix * a_ps + c8 * 8);
  const uint4 vb = *reinterpret_cast<const uint4*>(b + pix * b_ps + c8 * 8);
  uint4 o0, o1;
  o0.x = __byte_perm(va.x, vb.x, 0x5410); o0.y = __byte_perm(va.x, vb.x, 0x7632);
  o0.z = __byte_perm(va.y, vb.y, 0x5410); o0.w = __byte_perm(va.y, vb.y, 0x7632);
  o1.x = __byte_perm(va.z, vb.z, 0x5410); o1.y = __byte_perm(va.z, vb.z, 0x7632);
  o1.z = __byte_perm(va.w, vb.w, 0x5410); o1.w = __byte_perm(va.w, vb.w, 0x7632);
  uint4* dst = reinterpret_cast<uint4*>(out + pix * out_ps + c8 * 16);
  dst[0] = o0;
  dst[1] = o1;
}

int shuffle_concat_launch(const View& a, const View& b, const View& out, cudaStream_t stream) {
  if (a.C != b.C || out.C != 2 * a.C || a.C % 8 || a.N != out.N || a.H != out.H || a.W != out.W || b.N != a.N ||
      b.H != a.H || b.W != a.W)
    return fail(B200_ERR_INVALID, "shuffle_concat: shape mismatch");
  const long long total = static_cast<long long>(a.N) * a.H * a.W * (a.C / 8);
  shuffle_concat_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, stream>>>(
      static_cast<const uint16_t*>(a.base), a.pix_stride, static_cast<const uint16_t*>(b.base), b.pix_stride,
      static_cast<uint16_t*>(out.base), out.pix_stride, a.C / 8, total);
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}

}  // namespace b200
