// K1: batch collate fused with preprocessing.
//
// Replaces, in one pass over HBM, the reference's client-side `transforms.ToTensor()` + `Normalize(mean, std)`
// (profiling/run_profiler.py:63-70; 293-project/src/milind-code/scheduler.py:24-27) and the host gather
// `torch.stack(batch.inputs)` of GPUWorker.process_batch (293-project/src/scheduler.py:443).
//
// Output layout: NHWC bf16 with C padded 3 -> 4 (8 bytes per pixel, channel 3 = 0), batch padded to `padded_batch`
// images (slots >= n are zero).  Values are bit-exact w.r.t. torchvision's fp32 arithmetic
// ((u8 / 255) - mean) / std followed by round-to-nearest-even to bf16: they come from a 3x256 table built on the host
// with exactly those IEEE fp32 operations.
#include "kernels.cuh"
#include "ptx.cuh"

namespace b200 {

// ----------------------------------------------------------------------------------------------------------------
// Fast path: source image has exactly the output size -> the image is one flat stream of pixels.
// A block stages 12 KB of uint8 (4096 pixels) with coalesced 16-byte loads, then every thread emits 16-byte stores
// (2 pixels) so that a warp writes 512 contiguous bytes per instruction.
// ----------------------------------------------------------------------------------------------------------------
static constexpr int CT_THREADS = 256;
static constexpr int CT_PIX = 4096;  // pixels per block

__global__ void __launch_bounds__(CT_THREADS)
collate_u8_flat_kernel(const uint8_t* __restrict__ src_base, const CollateItem* __restrict__ items, int n,
                       int pixels_per_image, const uint16_t* __restrict__ lut, uint16_t* __restrict__ out) {
  __shared__ __align__(16) uint8_t s_in[CT_PIX * 3];
  __shared__ uint16_t s_lut[3 * 256];
  const int slot = blockIdx.y;
  const int p0 = blockIdx.x * CT_PIX;
  const int npix = min(CT_PIX, pixels_per_image - p0);
  uint4* dst = reinterpret_cast<uint4*>(out + (static_cast<size_t>(slot) * pixels_per_image + p0) * 4);
  const int nvec_out = (npix + 1) / 2;  // 16-byte output chunks (2 pixels each); pixels_per_image is even-checked on host

  if (slot >= n) {  // batch padding: zero image
    for (int q = threadIdx.x; q < nvec_out; q += CT_THREADS) dst[q] = make_uint4(0, 0, 0, 0);
    return;
  }
  for (int i = threadIdx.x; i < 3 * 256; i += CT_THREADS) s_lut[i] = lut[i];
  const uint8_t* src = src_base + items[slot].offset + static_cast<size_t>(p0) * 3;
  const int nbytes = npix * 3;
  const int nvec_in = nbytes >> 4;
  const uint4* src4 = reinterpret_cast<const uint4*>(src);  // offset % 16 == 0 and p0*3 % 16 == 0 (checked on host)
  for (int i = threadIdx.x; i < nvec_in; i += CT_THREADS) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(src4 + i));
    reinterpret_cast<uint4*>(s_in)[i] = v;
  }
  for (int i = (nvec_in << 4) + threadIdx.x; i < nbytes; i += CT_THREADS) s_in[i] = src[i];
  __syncthreads();
  for (int q = threadIdx.x; q < nvec_out; q += CT_THREADS) {
    const uint8_t* px = s_in + q * 6;
    uint4 o;
    o.x = uint32_t(s_lut[px[0]]) | (uint32_t(s_lut[256 + px[1]]) << 16);
    o.y = uint32_t(s_lut[512 + px[2]]);
    if (2 * q + 1 < npix) {
      o.z = uint32_t(s_lut[px[3]]) | (uint32_t(s_lut[256 + px[4]]) << 16);
      o.w = uint32_t(s_lut[512 + px[5]]);
    } else {
      o.z = 0; o.w = 0;
    }
    dst[q] = o;
  }
}

// ----------------------------------------------------------------------------------------------------------------
// General path: per-image source size; nearest / bilinear resize or letterbox with constant fill.
// One thread per output pixel (8-byte store).  Index arithmetic is integer-exact and mirrored by oracle/preprocess.py.
// ----------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
collate_u8_general_kernel(const uint8_t* __restrict__ src_base, const CollateItem* __restrict__ items, int n, int out_h,
                          int out_w, int mode, int fill, const uint16_t* __restrict__ lut,
                          uint16_t* __restrict__ out) {
  __shared__ uint16_t s_lut[3 * 256];
  for (int i = threadIdx.x; i < 3 * 256; i += blockDim.x) s_lut[i] = lut[i];
  __syncthreads();
  const int slot = blockIdx.y;
  const int pix = blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= out_h * out_w) return;
  uint2* dst = reinterpret_cast<uint2*>(out) + static_cast<size_t>(slot) * out_h * out_w + pix;
  if (slot >= n) {
    *dst = make_uint2(0, 0);
    return;
  }
  CollateItem it = items[slot];
  const int nch = it.w < 0 ? 1 : 3;          // w < 0: a single-channel (grayscale) request of width -w, expanded L -> RGB
  it.w = abs(it.w);
  const uint8_t* src = src_base + it.offset;
  const int oy = pix / out_w, ox = pix - oy * out_w;
  int r, g, b;
  // placement rectangle of the (possibly resized) image inside the output canvas
  int rh, rw, top, left;
  if (mode == COLLATE_PAD) {          // no resize: top-left aligned, rest filled
    rh = min(it.h, out_h); rw = min(it.w, out_w); top = 0; left = 0;
  } else if (mode == COLLATE_LETTERBOX) {  // aspect-preserving resize, centred, rest filled
    // scale = min(out_h/h, out_w/w) expressed in integers: the limiting side fills the canvas
    if (static_cast<long long>(out_h) * it.w <= static_cast<long long>(out_w) * it.h) {
      rh = out_h; rw = max(1, static_cast<int>((static_cast<long long>(it.w) * out_h + it.h / 2) / it.h));
    } else {
      rw = out_w; rh = max(1, static_cast<int>((static_cast<long long>(it.h) * out_w + it.w / 2) / it.w));
    }
    rw = min(rw, out_w); rh = min(rh, out_h);
    top = (out_h - rh) / 2; left = (out_w - rw) / 2;
  } else {                            // COLLATE_RESIZE: stretch to the full canvas
    rh = out_h; rw = out_w; top = 0; left = 0;
  }
  const int y = oy - top, x = ox - left;
  if (y < 0 || y >= rh || x < 0 || x >= rw) {
    r = g = b = fill;
  } else if (mode == COLLATE_PAD || (rh == it.h && rw == it.w)) {
    const uint8_t* p = src + (static_cast<size_t>(y) * it.w + x) * nch;
    r = p[0]; g = p[nch == 3 ? 1 : 0]; b = p[nch == 3 ? 2 : 0];
  } else {
    // bilinear, half-pixel centres, fixed point with 10 fractional bits; result rounded to uint8 (like a uint8 resize)
    const long long sx_num = static_cast<long long>(2 * x + 1) * it.w - rw;   // = (2*srcx) * rw   (srcx = (x+.5)*w/rw-.5)
    const long long sy_num = static_cast<long long>(2 * y + 1) * it.h - rh;
    const long long denx = 2LL * rw, deny = 2LL * rh;
    long long fx = (sx_num << 10) / denx, fy = (sy_num << 10) / deny;         // floor for >= 0; clamp below handles < 0
    if (sx_num < 0) fx = 0;
    if (sy_num < 0) fy = 0;
    int x0 = static_cast<int>(fx >> 10), y0 = static_cast<int>(fy >> 10);
    int wx = static_cast<int>(fx & 1023), wy = static_cast<int>(fy & 1023);
    if (x0 >= it.w - 1) { x0 = it.w - 1; wx = 0; }
    if (y0 >= it.h - 1) { y0 = it.h - 1; wy = 0; }
    const int x1 = min(x0 + 1, it.w - 1), y1 = min(y0 + 1, it.h - 1);
    const uint8_t* p00 = src + (static_cast<size_t>(y0) * it.w + x0) * nch;
    const uint8_t* p01 = src + (static_cast<size_t>(y0) * it.w + x1) * nch;
    const uint8_t* p10 = src + (static_cast<size_t>(y1) * it.w + x0) * nch;
    const uint8_t* p11 = src + (static_cast<size_t>(y1) * it.w + x1) * nch;
    int c[3];
#pragma unroll
    for (int kk = 0; kk < 3; ++kk) {
      const int k = nch == 3 ? kk : 0;
      const int top_v = p00[k] * (1024 - wx) + p01[k] * wx;
      const int bot_v = p10[k] * (1024 - wx) + p11[k] * wx;
      c[kk] = static_cast<int>((static_cast<long long>(top_v) * (1024 - wy) + static_cast<long long>(bot_v) * wy +
                                (1 << 19)) >> 20);
    }
    r = c[0]; g = c[1]; b = c[2];
  }
  uint2 o;
  o.x = uint32_t(s_lut[r]) | (uint32_t(s_lut[256 + g]) << 16);
  o.y = uint32_t(s_lut[512 + b]);
  *dst = o;
}

// ----------------------------------------------------------------------------------------------------------------
// The reference's own resize: transforms.Resize(224) / Resize((224, 224)) (profiling/run_profiler.py:63-70,
// 293-project/src/milind-code/scheduler.py:24-27) = bilinear WITH antialias.  The arithmetic is torch's
// (_compute_indices_min_size_weights_aa, ATen/native/cpu/UpSampleKernel.cpp): per axis, scale = in / out, support =
// max(scale, 1), centre = scale * (i + 0.5), taps [int(centre - support + 0.5), int(centre + support + 0.5)) clipped to the
// image, triangle weights 1 - |(j - centre + 0.5) / max(scale, 1)| normalised to 1.  One thread per output pixel walks
// its (taps_y x taps_x) window once -- a single rounding at the end, where torch's uint8 path rounds between its two
// passes: within 1 in the last bit of torchvision (oracle/preprocess.py resize_aa_u8, pinned to torchvision's output).
// COLLATE_RESIZE_AA stretches to the canvas; COLLATE_RESIZE_SHORT_AA resizes the short side to the (square) canvas size
// and centre-crops the long one, torchvision's Resize(s) + CenterCrop(s) with its integer size and offset rules.
// ----------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
collate_u8_aa_kernel(const uint8_t* __restrict__ src_base, const CollateItem* __restrict__ items, int n, int out_h,
                     int out_w, int mode, const uint16_t* __restrict__ lut, uint16_t* __restrict__ out) {
  __shared__ uint16_t s_lut[3 * 256];
  for (int i = threadIdx.x; i < 3 * 256; i += blockDim.x) s_lut[i] = lut[i];
  __syncthreads();
  const int slot = blockIdx.y;
  const int pix = blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= out_h * out_w) return;
  uint2* dst = reinterpret_cast<uint2*>(out) + static_cast<size_t>(slot) * out_h * out_w + pix;
  if (slot >= n) {
    *dst = make_uint2(0, 0);
    return;
  }
  CollateItem it = items[slot];
  const int nch = it.w < 0 ? 1 : 3;
  it.w = abs(it.w);
  const uint8_t* src = src_base + it.offset;
  const int oy = pix / out_w, ox = pix - oy * out_w;
  int nh = out_h, nw = out_w, top = 0, left = 0;
  if (mode == COLLATE_RESIZE_SHORT_AA) {
    // torchvision _compute_resized_output_size: short side -> s, long side -> int(s * long / short)
    const int sz = min(out_h, out_w);
    if (it.w <= it.h) { nw = sz; nh = static_cast<int>(static_cast<long long>(sz) * it.h / it.w); }
    else              { nh = sz; nw = static_cast<int>(static_cast<long long>(sz) * it.w / it.h); }
    // CenterCrop: int(round((N - out) / 2.0)), Python's round (ties to even)
    const int dy = nh - out_h, dx = nw - out_w;
    top = (dy & 1) ? ((dy >> 1) + ((dy >> 1) & 1)) : (dy >> 1);
    left = (dx & 1) ? ((dx >> 1) + ((dx >> 1) & 1)) : (dx >> 1);
  }
  const int ry = oy + top, rx = ox + left;              // position in the resized image
  const float sy = static_cast<float>(it.h) / nh, sx = static_cast<float>(it.w) / nw;
  const float sup_y = fmaxf(sy, 1.f), sup_x = fmaxf(sx, 1.f);
  const float inv_y = 1.f / sup_y, inv_x = 1.f / sup_x;
  const float cy = sy * (ry + 0.5f), cx = sx * (rx + 0.5f);
  const int y0 = max(static_cast<int>(cy - sup_y + 0.5f), 0), y1 = min(static_cast<int>(cy + sup_y + 0.5f), it.h);
  const int x0 = max(static_cast<int>(cx - sup_x + 0.5f), 0), x1 = min(static_cast<int>(cx + sup_x + 0.5f), it.w);
  float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, wsum_y = 0.f, wsum_x = 0.f;
  for (int x = x0; x < x1; ++x) wsum_x += fmaxf(1.f - fabsf((x - cx + 0.5f) * inv_x), 0.f);
  for (int y = y0; y < y1; ++y) {
    const float wy = fmaxf(1.f - fabsf((y - cy + 0.5f) * inv_y), 0.f);
    wsum_y += wy;
    const uint8_t* row = src + (static_cast<size_t>(y) * it.w + x0) * nch;
    float r0 = 0.f, r1 = 0.f, r2 = 0.f;
    if (nch == 3) {
      for (int x = x0; x < x1; ++x, row += 3) {
        const float wx = fmaxf(1.f - fabsf((x - cx + 0.5f) * inv_x), 0.f);
        r0 += wx * row[0]; r1 += wx * row[1]; r2 += wx * row[2];
      }
    } else {
      for (int x = x0; x < x1; ++x, ++row) r0 += fmaxf(1.f - fabsf((x - cx + 0.5f) * inv_x), 0.f) * row[0];
      r1 = r2 = r0;
    }
    acc0 += wy * r0; acc1 += wy * r1; acc2 += wy * r2;
  }
  const float norm = 1.f / (wsum_x * wsum_y);
  const int r = min(255, max(0, static_cast<int>(floorf(acc0 * norm + 0.5f))));
  const int g = min(255, max(0, static_cast<int>(floorf(acc1 * norm + 0.5f))));
  const int b = min(255, max(0, static_cast<int>(floorf(acc2 * norm + 0.5f))));
  uint2 o;
  o.x = uint32_t(s_lut[r]) | (uint32_t(s_lut[256 + g]) << 16);
  o.y = uint32_t(s_lut[512 + b]);
  *dst = o;
}

// ----------------------------------------------------------------------------------------------------------------
// Reference-native requests: already-normalised fp32 CHW tensors (293-project/src/test_scheduler.py:89 submits
// torch.randn(3,224,224)).  Gathers them into the padded NHWC4 bf16 batch; 2 pixels (16 B) per thread.
// ----------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
collate_f32chw_kernel(const float* __restrict__ src, int n, int hw, uint16_t* __restrict__ out) {
  const int slot = blockIdx.y;
  const int q = blockIdx.x * blockDim.x + threadIdx.x;  // pixel pair
  if (2 * q >= hw) return;
  uint4* dst = reinterpret_cast<uint4*>(out + (static_cast<size_t>(slot) * hw + 2 * q) * 4);
  if (slot >= n) {
    *dst = make_uint4(0, 0, 0, 0);
    return;
  }
  const float* s = src + static_cast<size_t>(slot) * 3 * hw + 2 * q;
  const float2 r = *reinterpret_cast<const float2*>(s);
  const float2 g = *reinterpret_cast<const float2*>(s + hw);
  const float2 b = *reinterpret_cast<const float2*>(s + 2 * hw);
  uint4 o;
  o.x = pack_bf16x2(r.x, g.x);
  o.y = pack_bf16x2(b.x, 0.f);
  o.z = pack_bf16x2(r.y, g.y);
  o.w = pack_bf16x2(b.y, 0.f);
  *dst = o;
}

// ------------------------------------------------------------------------------------------------ host side

void build_norm_lut(const float mean[3], const float stdv[3], uint16_t lut[3 * 256]) {
  for (int c = 0; c < 3; ++c)
    for (int u = 0; u < 256; ++u) {
      volatile float x = static_cast<float>(u) / 255.0f;  // ToTensor: .div(255)       (IEEE fp32)
      volatile float y = x - mean[c];                     // Normalize: sub_(mean)
      volatile float z = y / stdv[c];                     //            .div_(std)
      const float zf = z;
      uint32_t bits;
      memcpy(&bits, &zf, 4);
      const uint32_t rounded = bits + 0x7FFFu + ((bits >> 16) & 1u);  // fp32 -> bf16 round-to-nearest-even
      lut[c * 256 + u] = static_cast<uint16_t>(rounded >> 16);
    }
}

int collate_u8_launch(const uint8_t* src_base, const CollateItem* items_dev, const CollateItem* items_host, int n,
                      int padded_batch, int out_h, int out_w, int mode, int fill, const uint16_t* lut_dev, void* out,
                      cudaStream_t stream) {
  if (n < 0 || padded_batch < n || padded_batch <= 0) return fail(B200_ERR_INVALID, "collate: bad batch counts");
  if (out_h <= 0 || out_w <= 0) return fail(B200_ERR_INVALID, "collate: bad output size");
  // the flat kernel reads 16-byte vectors: the base pointer and every offset must be 16-byte aligned, otherwise the
  // general kernel (byte loads) takes the batch
  bool flat = (out_h * out_w) % 2 == 0 && (reinterpret_cast<uintptr_t>(src_base) % 16) == 0;
  for (int i = 0; i < n && flat; ++i)
    flat = items_host[i].h == out_h && items_host[i].w == out_w && (items_host[i].offset % 16) == 0;
  if (flat) {
    dim3 grid((out_h * out_w + CT_PIX - 1) / CT_PIX, padded_batch);
    collate_u8_flat_kernel<<<grid, CT_THREADS, 0, stream>>>(src_base, items_dev, n, out_h * out_w, lut_dev,
                                                            static_cast<uint16_t*>(out));
  } else {
    if (mode == COLLATE_RESIZE_AA || mode == COLLATE_RESIZE_SHORT_AA) {
      if (mode == COLLATE_RESIZE_SHORT_AA && out_h != out_w)
        return fail(B200_ERR_INVALID, "collate: short-side resize + centre crop needs a square model input");
      dim3 grid((out_h * out_w + 255) / 256, padded_batch);
      collate_u8_aa_kernel<<<grid, 256, 0, stream>>>(src_base, items_dev, n, out_h, out_w, mode, lut_dev,
                                                     static_cast<uint16_t*>(out));
      B200_CUDA_OK(cudaGetLastError());
      return B200_OK;
    }
    if (mode != COLLATE_IDENTITY && mode != COLLATE_PAD && mode != COLLATE_RESIZE && mode != COLLATE_LETTERBOX)
      return fail(B200_ERR_INVALID, "collate: unknown preprocess mode");
    if (mode == COLLATE_IDENTITY) {   // same-size images that missed the vector path only for alignment: plain copy
      for (int i = 0; i < n; ++i)
        if (items_host[i].h != out_h || items_host[i].w != out_w)
          return fail(B200_ERR_INVALID, "collate: image size differs from the model input and no resize mode given");
      mode = COLLATE_PAD;
    }
    dim3 grid((out_h * out_w + 255) / 256, padded_batch);
    collate_u8_general_kernel<<<grid, 256, 0, stream>>>(src_base, items_dev, n, out_h, out_w, mode, fill, lut_dev,
                                                        static_cast<uint16_t*>(out));
  }
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}

int collate_f32chw_launch(const float* src, int n, int padded_batch, int h, int w, void* out, cudaStream_t stream) {
  if (n < 0 || padded_batch < n || padded_batch <= 0) return fail(B200_ERR_INVALID, "collate: bad batch counts");
  if ((h * w) % 2) return fail(B200_ERR_INVALID, "collate_f32: odd pixel count");
  dim3 grid((h * w / 2 + 255) / 256, padded_batch);
  collate_f32chw_kernel<<<grid, 256, 0, stream>>>(src, n, h * w, static_cast<uint16_t*>(out));
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}

}  // namespace b200
