// Pooling on NHWC bf16: max-pool (k x k, stride s, pad p; padding never wins the max) and global average pool.
// Bandwidth-bound; one thread owns 8 channels (16 bytes) of one output pixel so a warp reads/writes 512 contiguous bytes.
//
// Replaces nn.MaxPool2d(3,2,1) (tv resnet.py:271, shufflenetv2.py:118), the 5x5/1/2 max-pools of YOLOv8's SPPF, and
// nn.AdaptiveAvgPool2d(1) / x.mean([2,3]) (resnet.py:278, shufflenetv2.py:161, efficientnet.py:303).
#include "kernels.cuh"
#include "ptx.cuh"

namespace b200 {

__device__ __forceinline__ uint32_t bf16x2_max(uint32_t a, uint32_t b) {
  __nv_bfloat162 x = *reinterpret_cast<__nv_bfloat162*>(&a), y = *reinterpret_cast<__nv_bfloat162*>(&b);
  __nv_bfloat162 m = __hmax2(x, y);
  return *reinterpret_cast<uint32_t*>(&m);
}

__global__ void __launch_bounds__(256)
maxpool_kernel(const uint16_t* __restrict__ in, int H, int W, int in_ps, uint16_t* __restrict__ out, FastDiv Hod,
               FastDiv Wod, int out_ps, FastDiv C8, int k, int s, int p, unsigned total) {
  pdl_wait();   // launched with programmatic stream serialization: inputs are valid from here on (no early trigger:
                // a multi-wave kernel must not share its SMs with the waiting blocks of its successor)
  const unsigned idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int Ho = static_cast<int>(Hod.d), Wo = static_cast<int>(Wod.d);
  const unsigned pix = fast_div(idx, C8);
  const int c8 = static_cast<int>(idx - pix * C8.d);
  const unsigned row = fast_div(pix, Wod);
  const int ow = static_cast<int>(pix - row * Wod.d);
  const int n = static_cast<int>(fast_div(row, Hod));
  const int oh = static_cast<int>(row - n * Hod.d);
  const uint32_t NEG = 0xFF80FF80u;  // bf16 -inf pair
  uint4 m = make_uint4(NEG, NEG, NEG, NEG);
  const int y0 = oh * s - p, x0 = ow * s - p;
  for (int r = 0; r < k; ++r) {
    const int y = y0 + r;
    if (y < 0 || y >= H) continue;
    for (int q = 0; q < k; ++q) {
      const int x = x0 + q;
      if (x < 0 || x >= W) continue;
      const uint4 v = *reinterpret_cast<const uint4*>(in + ((static_cast<size_t>(n) * H + y) * W + x) * in_ps + c8 * 8);
      m.x = bf16x2_max(m.x, v.x); m.y = bf16x2_max(m.y, v.y);
      m.z = bf16x2_max(m.z, v.z); m.w = bf16x2_max(m.w, v.w);
    }
  }
  *reinterpret_cast<uint4*>(out + ((static_cast<size_t>(n) * Ho + oh) * Wo + ow) * out_ps + c8 * 8) = m;
}

int maxpool_launch(const View& in, const View& out, int ksize, int stride, int pad, cudaStream_t stream) {
  if (in.C != out.C || in.C % 8 || in.pix_stride % 8 || out.pix_stride % 8)
    return fail(B200_ERR_INVALID, "maxpool: channels must match and be multiples of 8");
  if ((in.H + 2 * pad - ksize) / stride + 1 != out.H || (in.W + 2 * pad - ksize) / stride + 1 != out.W || in.N != out.N)
    return fail(B200_ERR_INVALID, "maxpool: output geometry mismatch");
  const long long total = static_cast<long long>(out.N) * out.H * out.W * (out.C / 8);
  if (total >= (1ll << 31)) return fail(B200_ERR_INVALID, "maxpool: tensor too large");
  const unsigned blocks = static_cast<unsigned>((total + 255) / 256);
  B200_CUDA_OK(launch_pdl(maxpool_kernel, dim3(blocks), dim3(256), size_t(0), stream, static_cast<const uint16_t*>(in.base), in.H, in.W, in.pix_stride, static_cast<uint16_t*>(out.base), make_fast_div(out.H), make_fast_div(out.W), out.pix_stride, make_fast_div(out.C / 8), ksize, stride, pad, static_cast<unsigned>(total)));
  return B200_OK;
}

// Global average pool: block = (image, up to 1024 threads); thread owns a 16-byte channel chunk and strides over pixels in
// `gy` slices that are combined through shared memory.  fp32 accumulation, sum * (1/HW) like aten's mean.
__global__ void __launch_bounds__(1024)
global_avgpool_kernel(const uint16_t* __restrict__ in, int HW, int in_ps, int C8, float* __restrict__ out) {
  pdl_wait();   // launched with programmatic stream serialization: inputs are valid from here on (no early trigger:
                // a multi-wave kernel must not share its SMs with the waiting blocks of its successor)
  extern __shared__ float s_part[];  // [slices][C8*8]
  const int n = blockIdx.x;
  const int slices = blockDim.x / C8 > 0 ? blockDim.x / C8 : 1;
  const int c8 = threadIdx.x % C8, sl = threadIdx.x / C8;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (sl < slices) {
    const uint16_t* base = in + static_cast<size_t>(n) * HW * in_ps + c8 * 8;
#pragma unroll 4
    for (int p = sl; p < HW; p += slices) {
      const uint4 v = *reinterpret_cast<const uint4*>(base + static_cast<size_t>(p) * in_ps);
      acc[0] += bf16_lo(v.x); acc[1] += bf16_hi(v.x); acc[2] += bf16_lo(v.y); acc[3] += bf16_hi(v.y);
      acc[4] += bf16_lo(v.z); acc[5] += bf16_hi(v.z); acc[6] += bf16_lo(v.w); acc[7] += bf16_hi(v.w);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) s_part[(sl * C8 + c8) * 8 + i] = acc[i];
  }
  __syncthreads();
  const float inv = 1.0f / static_cast<float>(HW);
  for (int c = threadIdx.x; c < C8 * 8; c += blockDim.x) {
    float s = 0.f;
    for (int j = 0; j < slices; ++j) s += s_part[j * C8 * 8 + c];
    out[static_cast<size_t>(n) * C8 * 8 + c] = s * inv;
  }
}

int global_avgpool_launch(const View& in, float* out, cudaStream_t stream) {
  if (in.C % 8 || in.pix_stride % 8) return fail(B200_ERR_INVALID, "avgpool: channels must be a multiple of 8");
  const int C8 = in.C / 8;
  // when C8 > 256 a block walks several channel chunks: launch ceil(C8/256) column blocks via grid.y
  if (C8 <= 256) {
    // as many pixel slices as fit 1024 threads: a 7 x 7 x 2048 image is 49 loads per thread with 256 threads and the
    // 256 blocks of a batch were 1.7 waves of 8 warps; with 4 slices all blocks are resident at once
    const int slices = 1024 / C8 > 0 ? 1024 / C8 : 1;
    const int threads = slices * C8;
    const size_t smem = size_t(slices) * C8 * 8 * sizeof(float);
    B200_CUDA_OK(launch_pdl(global_avgpool_kernel, dim3(in.N), dim3(threads), size_t(smem), stream, static_cast<const uint16_t*>(in.base), in.H * in.W, in.pix_stride, C8, out));
  } else {
    return fail(B200_ERR_INVALID, "avgpool: more than 2048 channels unsupported");
  }
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}

}  // namespace b200
