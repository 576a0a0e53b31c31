// Pure data-movement kernels on NHWC bf16 channel slices: slice copy, nearest 2x upsample into a concat slice, and the
// ShuffleNetV2 concat + channel_shuffle(groups=2) interleave.  16 bytes per thread, coalesced.
//
// Replaces torch.cat / x.chunk(2) / channel_shuffle (tv shufflenetv2.py:28-40,98-104) and nn.Upsample(2,'nearest') +
// Concat of the YOLOv8 head.
#include "kernels.cuh"
#include "ptx.cuh"

namespace b200 {

__global__ void __launch_bounds__(256)
copy_channels_kernel(const uint16_t* __restrict__ in, int in_ps, uint16_t* __restrict__ out, int out_ps, FastDiv C8,
                     unsigned total) {
  pdl_wait();   // launched with programmatic stream serialization: inputs are valid from here on (no early trigger:
                // a multi-wave kernel must not share its SMs with the waiting blocks of its successor)
  const unsigned idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const unsigned pix = fast_div(idx, C8);
  const int c8 = static_cast<int>(idx - pix * C8.d);
  *reinterpret_cast<uint4*>(out + static_cast<size_t>(pix) * out_ps + c8 * 8) =
      *reinterpret_cast<const uint4*>(in + static_cast<size_t>(pix) * in_ps + c8 * 8);
}

int copy_channels_launch(const View& in, const View& out, cudaStream_t stream) {
  if (in.C != out.C || in.C % 8 || in.N != out.N || in.H != out.H || in.W != out.W)
    return fail(B200_ERR_INVALID, "copy_channels: shape mismatch");
  const long long total = static_cast<long long>(in.N) * in.H * in.W * (in.C / 8);
  if (total >= (1ll << 31)) return fail(B200_ERR_INVALID, "copy_channels: tensor too large");
  B200_CUDA_OK(launch_pdl(copy_channels_kernel, dim3(static_cast<unsigned>((total + 255) / 256)), dim3(256), size_t(0), stream, static_cast<const uint16_t*>(in.base), in.pix_stride, static_cast<uint16_t*>(out.base), out.pix_stride, make_fast_div(in.C / 8), static_cast<unsigned>(total)));
  return B200_OK;
}

__global__ void __launch_bounds__(256)
upsample2x_kernel(const uint16_t* __restrict__ in, int H, int W, int in_ps, uint16_t* __restrict__ out, int out_ps,
                  FastDiv C8, FastDiv W2, FastDiv H2, unsigned total) {
  pdl_wait();   // launched with programmatic stream serialization: inputs are valid from here on (no early trigger:
                // a multi-wave kernel must not share its SMs with the waiting blocks of its successor)
  const unsigned idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const unsigned pix = fast_div(idx, C8);
  const int c8 = static_cast<int>(idx - pix * C8.d);
  const unsigned row = fast_div(pix, W2);
  const int ow = static_cast<int>(pix - row * W2.d);
  const int n = static_cast<int>(fast_div(row, H2));
  const int oh = static_cast<int>(row - n * H2.d);
  const uint4 v =
      *reinterpret_cast<const uint4*>(in + ((static_cast<size_t>(n) * H + (oh >> 1)) * W + (ow >> 1)) * in_ps + c8 * 8);
  *reinterpret_cast<uint4*>(out + ((static_cast<size_t>(n) * 2 * H + oh) * 2 * W + ow) * out_ps + c8 * 8) = v;
}

int upsample2x_launch(const View& in, const View& out, cudaStream_t stream) {
  if (in.C != out.C || in.C % 8 || in.N != out.N || 2 * in.H != out.H || 2 * in.W != out.W)
    return fail(B200_ERR_INVALID, "upsample2x: shape mismatch");
  const long long total = static_cast<long long>(out.N) * out.H * out.W * (out.C / 8);
  if (total >= (1ll << 31)) return fail(B200_ERR_INVALID, "upsample2x: tensor too large");
  B200_CUDA_OK(launch_pdl(upsample2x_kernel, dim3(static_cast<unsigned>((total + 255) / 256)), dim3(256), size_t(0), stream, static_cast<const uint16_t*>(in.base), in.H, in.W, in.pix_stride, static_cast<uint16_t*>(out.base), out.pix_stride, make_fast_div(in.C / 8), make_fast_div(2 * in.W), make_fast_div(2 * in.H), static_cast<unsigned>(total)));
  return B200_OK;
}

// ShuffleNetV2 concat + channel_shuffle(groups=2): logical out[2i] = a[i], out[2i+1] = b[i] (i < c).  The output keeps
// the two halves of the *shuffled* tensor in separately padded slices (logical l -> physical l < c ? l : cp + l - c), so the
// next unit's chunk(2) needs no copy.  A thread owns ONE 16-byte output chunk position and walks pixels; the source runs
// depend only on the chunk, so they are resolved once.
__global__ void __launch_bounds__(256)
shuffle_concat_kernel(const uint16_t* __restrict__ a, int a_ps, const uint16_t* __restrict__ b, int b_ps,
                      uint16_t* __restrict__ out, int out_ps, int c, int cp, long long npix, int chunks, int pix_slots) {
  pdl_wait();   // launched with programmatic stream serialization: inputs are valid from here on (no early trigger:
                // a multi-wave kernel must not share its SMs with the waiting blocks of its successor)
  const int ppb = blockDim.x / chunks;                       // pixels a block handles per step
  const int ch = threadIdx.x % chunks;
  const int lp = threadIdx.x / chunks;
  if (lp >= ppb) return;
  // source window of this chunk: with l0 = first logical channel (always even: c and cp are), it needs a[i0 .. i0+3] and
  // b[i0 .. i0+3], i0 = l0 / 2 -- two 8-byte runs that are only 2-byte aligned in the second half (i0 = 4k - (cp - c) / 2).
  // Each run is read as two ALIGNED 8-byte loads and funnel-shifted, so a warp's loads cover 256 contiguous bytes each.
  const int p0 = ch * 8;
  const int half = p0 >= cp ? 1 : 0;
  const int q0 = p0 - half * cp;
  const int l0 = half * c + q0;
  const int i0 = l0 >> 1;
  const int ia = i0 & ~3, sh = (i0 & 3) * 16;
  const int nvalid = max(0, min(8, c - q0));                 // elements of this chunk that are real channels
  const uint64_t keep_lo = nvalid >= 4 ? ~0ull : (nvalid == 0 ? 0ull : (~0ull >> (64 - 16 * nvalid)));      // chunk elements 0-3
  const uint64_t keep_hi = nvalid >= 8 ? ~0ull : (nvalid <= 4 ? 0ull : (~0ull >> (64 - 16 * (nvalid - 4))));  // elements 4-7
  for (long long pix = static_cast<long long>(blockIdx.x) * ppb + lp; pix < npix; pix += static_cast<long long>(pix_slots)) {
    const uint16_t* pa = a + pix * a_ps + ia;
    const uint16_t* pb = b + pix * b_ps + ia;
    uint64_t av = *reinterpret_cast<const uint64_t*>(pa), bv = *reinterpret_cast<const uint64_t*>(pb);
    if (sh) {
      const uint64_t a2 = *reinterpret_cast<const uint64_t*>(pa + 4), b2 = *reinterpret_cast<const uint64_t*>(pb + 4);
      av = (av >> sh) | (a2 << (64 - sh));
      bv = (bv >> sh) | (b2 << (64 - sh));
    }
    // av = a[i0..i0+3], bv = b[i0..i0+3] (16 bits each): interleave to a0 b0 a1 b1 | a2 b2 a3 b3
    const uint32_t a01 = static_cast<uint32_t>(av), a23 = static_cast<uint32_t>(av >> 32);
    const uint32_t b01 = static_cast<uint32_t>(bv), b23 = static_cast<uint32_t>(bv >> 32);
    uint64_t lo = static_cast<uint64_t>(__byte_perm(a01, b01, 0x5410)) | (static_cast<uint64_t>(__byte_perm(a01, b01, 0x7632)) << 32);
    uint64_t hi = static_cast<uint64_t>(__byte_perm(a23, b23, 0x5410)) | (static_cast<uint64_t>(__byte_perm(a23, b23, 0x7632)) << 32);
    lo &= keep_lo;
    hi &= keep_hi;
    uint4 o;
    o.x = static_cast<uint32_t>(lo); o.y = static_cast<uint32_t>(lo >> 32);
    o.z = static_cast<uint32_t>(hi); o.w = static_cast<uint32_t>(hi >> 32);
    *reinterpret_cast<uint4*>(out + pix * out_ps + ch * 8) = o;
  }
}

int shuffle_concat_launch(const View& a, const View& b, const View& out, int c, int cp, cudaStream_t stream) {
  if (cp % 8 || c > cp || c <= 0 || a.C < c || b.C < c || out.C != 2 * cp || a.N != out.N || a.H != out.H ||
      a.W != out.W || b.N != a.N || b.H != a.H || b.W != a.W || out.pix_stride % 8 || a.pix_stride % 4 || b.pix_stride % 4 ||
      (c & 1) || (reinterpret_cast<uintptr_t>(a.base) & 7) || (reinterpret_cast<uintptr_t>(b.base) & 7))
    return fail(B200_ERR_INVALID, "shuffle_concat: shape mismatch (c must be even, inputs 8-byte aligned)");
  const int chunks = (2 * cp) >> 3;
  if (chunks > 256) return fail(B200_ERR_INVALID, "shuffle_concat: more than 2048 output channels unsupported");
  const long long npix = static_cast<long long>(a.N) * a.H * a.W;
  const int ppb = 256 / chunks;
  const int threads = ppb * chunks;
  long long blocks = (npix + ppb - 1) / ppb;
  const long long cap = static_cast<long long>(device_sm_count()) * 16;    // grid-stride beyond ~16 blocks per SM
  if (blocks > cap) blocks = cap;
  const int pix_slots = static_cast<int>(blocks) * ppb;
  B200_CUDA_OK(launch_pdl(shuffle_concat_kernel, dim3(static_cast<unsigned>(blocks)), dim3(threads), size_t(0), stream,
                          static_cast<const uint16_t*>(a.base), a.pix_stride, static_cast<const uint16_t*>(b.base),
                          b.pix_stride, static_cast<uint16_t*>(out.base), out.pix_stride, c, cp, npix, chunks, pix_slots));
  return B200_OK;
}

}  // namespace b200
