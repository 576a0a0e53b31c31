// Depthwise 3x3 convolution (stride 1 / 2, pad 1) + folded BatchNorm bias + optional activation on NHWC bf16, with an
// optional fused squeeze (per-strip channel sums for the squeeze-excitation block that follows in MBConv).
//
// Bandwidth-bound (9 MAC per output element).  Strip kernel: one thread owns 4 channels (8 B) of a strip of DW_TW
// output columns x R output rows (R = 2 at stride 1, 1 at stride 2), keeps its 36 weights in registers and slides a
// 3-column register window along the strip, so every input element is loaded once per strip (+ the 2-column halo)
// instead of once per tap: L1 wavefronts per output drop ~10x against the per-pixel kernel below, which is L1-bound.
// A warp covers 128 contiguous channels of one pixel per load (256 B).  fp32 weights [9][C], fp32 accumulation, taps
// accumulated in (row, column) order on top of the bias -- the same order as the per-pixel kernel, so both agree bit
// for bit.  Widths that are not a multiple of DW_TW fall back to the per-pixel kernel.
//
// Replaces the grouped nn.Conv2d(groups=C)+BatchNorm2d of tv shufflenetv2.py:62-76,86-96 (`depthwise_conv`) and the
// depthwise stage of MBConv (tv efficientnet.py:140-151); the fused sums replace the AdaptiveAvgPool2d(1) that opens
// SqueezeExcitation (tv ops/misc.py:244-254).
#include "kernels.cuh"
#include "ptx.cuh"

namespace b200 {

int dwconv3x3_tile_try(const View& in, const View& out, const float* w, const float* bias, int stride, int act,
                       float* sums, cudaStream_t stream);       // dwconv_tile.cu
int dwconv3x3_tile_squeeze_rows(int C, int H, int W, int stride);
int dwconv3x3_strips(int Ho, int Wo, int stride);

constexpr int DW_STRIP_THREADS = 128;
constexpr int DW_TW = 7;   // strip width: every spatial size of the served classifiers (224 / 2^k) is a multiple of 7

__device__ __forceinline__ uint2 dw_load(const uint16_t* __restrict__ p, bool ok) {
  uint2 r = make_uint2(0u, 0u);
  if (ok) r = *reinterpret_cast<const uint2*>(p);
  return r;
}
__device__ __forceinline__ void dw_unpack(const uint2 r, float (&v)[4]) {
  v[0] = bf16_lo(r.x); v[1] = bf16_hi(r.x); v[2] = bf16_lo(r.y); v[3] = bf16_hi(r.y);
}

// se_partial: null, or fp32 [N][strips_h * strips_w][C]: sums of the STORED (bf16-rounded, activated) outputs of each
// strip; the gate kernel adds the strips of an image in index order, so the squeeze is deterministic.
// All loads of a strip (ROWS x NCOL 8-byte words, still packed) are issued before the first FMA: the kernel is
// latency-bound otherwise (4 loads in flight per thread gave ~2.4 TB/s).
// SQ (fused squeeze, an image has at most 32 strips): a block is (slots = strips rounded up to a power of two) x
// (128 / slots channel quads) of ONE image, so the strips of a channel meet in shared memory and the block writes the
// image's channel sums [N][C] directly, added in strip order (deterministic).  Otherwise threads are laid out channel
// quad fastest over (image, strip) and, if se_partial is given, every strip writes its own partial row.
template <int STRIDE, int ACT, bool SQ>
__global__ void __launch_bounds__(DW_STRIP_THREADS, 3)
dwconv3x3_strip_kernel(const uint16_t* __restrict__ in, int H, int W, int in_ps, uint16_t* __restrict__ out, int Ho,
                       int Wo, int out_ps, FastDiv C4d, FastDiv strips_hd, FastDiv strips_wd, const float* __restrict__ w,
                       const float* __restrict__ bias, float* __restrict__ se_partial, unsigned total, int slots) {
  pdl_wait();   // launched with programmatic stream serialization: inputs are valid from here on (no early trigger:
                // a multi-wave kernel must not share its SMs with the waiting blocks of its successor)
  constexpr int R = STRIDE == 1 ? 2 : 1;                        // output rows per thread
  constexpr int ROWS = STRIDE == 1 ? 4 : 3;                     // input rows per thread
  constexpr int NCOL = STRIDE == 1 ? DW_TW + 2 : 2 * DW_TW + 1;   // input columns per thread, first one = halo
  __shared__ float4 red[SQ ? DW_STRIP_THREADS : 1];
  const int C4 = static_cast<int>(C4d.d), strips_h = static_cast<int>(strips_hd.d), strips_w = static_cast<int>(strips_wd.d);
  int c4, sx, sy, n;
  bool active;
  if (SQ) {
    const int qn = DW_STRIP_THREADS / slots;                    // channel quads per block
    const int qblocks = (C4 + qn - 1) / qn;
    const int slot = threadIdx.x / qn;
    c4 = (blockIdx.x % qblocks) * qn + threadIdx.x % qn;
    n = blockIdx.x / qblocks;
    sy = slot / strips_w;
    sx = slot - sy * strips_w;
    active = slot < strips_h * strips_w && c4 < C4;
    if (!active) { c4 = 0; sx = 0; sy = 0; }                    // idle lanes still reach the barrier below
  } else {
    const unsigned idx = blockIdx.x * blockDim.x + threadIdx.x;     // 32-bit, multiply-high division (see fast_div)
    if (idx >= total) return;
    const unsigned t = fast_div(idx, C4d);
    c4 = static_cast<int>(idx - t * C4d.d);
    const unsigned t2 = fast_div(t, strips_wd);
    sx = static_cast<int>(t - t2 * strips_wd.d);
    n = static_cast<int>(fast_div(t2, strips_hd));
    sy = static_cast<int>(t2 - n * strips_hd.d);
    active = true;
  }
  const int C = C4 * 4;
  const int oy0 = sy * R, ox0 = sx * DW_TW;
  const int iy0 = oy0 * STRIDE - 1, ix0 = ox0 * STRIDE - 1;

  uint2 raw[ROWS][NCOL];
#pragma unroll
  for (int i = 0; i < ROWS; ++i) {
    const int y = iy0 + i;
    const bool rowok = y >= 0 && y < H;
    const uint16_t* rowp = in + ((static_cast<size_t>(n) * H + (rowok ? y : 0)) * W) * in_ps + c4 * 4;
#pragma unroll
    for (int k = 0; k < NCOL; ++k) {
      const int x = ix0 + k;
      raw[i][k] = dw_load(rowp + static_cast<ptrdiff_t>(x) * in_ps, active && rowok && x >= 0 && x < W);
    }
  }
  float wt[9][4], b[4];
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    const float4 v = *reinterpret_cast<const float4*>(w + k * C + c4 * 4);
    wt[k][0] = v.x; wt[k][1] = v.y; wt[k][2] = v.z; wt[k][3] = v.w;
  }
  {
    const float4 v = *reinterpret_cast<const float4*>(bias + c4 * 4);
    b[0] = v.x; b[1] = v.y; b[2] = v.z; b[3] = v.w;
  }
  float win[ROWS][3][4];
  float sum[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int i = 0; i < ROWS; ++i) {
    if (STRIDE == 1) {
      dw_unpack(raw[i][0], win[i][1]);
      dw_unpack(raw[i][1], win[i][2]);
    } else {
      dw_unpack(raw[i][0], win[i][2]);
    }
  }
#pragma unroll
  for (int j = 0; j < DW_TW; ++j) {
    const int ox = ox0 + j;
#pragma unroll
    for (int i = 0; i < ROWS; ++i) {
      if (STRIDE == 1) {
#pragma unroll
        for (int c = 0; c < 4; ++c) { win[i][0][c] = win[i][1][c]; win[i][1][c] = win[i][2][c]; }
        dw_unpack(raw[i][j + 2], win[i][2]);
      } else {
#pragma unroll
        for (int c = 0; c < 4; ++c) win[i][0][c] = win[i][2][c];
        dw_unpack(raw[i][2 * j + 1], win[i][1]);
        dw_unpack(raw[i][2 * j + 2], win[i][2]);
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if (oy0 + r >= Ho || !active) continue;
      float acc[4] = {b[0], b[1], b[2], b[3]};
#pragma unroll
      for (int dr = 0; dr < 3; ++dr) {
        // a tap outside the image is skipped by the per-pixel kernel; here it multiplies an exact zero: same result
        // unless a weight is inf/nan, which a folded BatchNorm never produces
#pragma unroll
        for (int q = 0; q < 3; ++q) {
#pragma unroll
          for (int c = 0; c < 4; ++c) acc[c] = fmaf(win[r + dr][q][c], wt[dr * 3 + q][c], acc[c]);
        }
      }
      uint2 o;
      o.x = pack_bf16x2(apply_act(acc[0], ACT), apply_act(acc[1], ACT));
      o.y = pack_bf16x2(apply_act(acc[2], ACT), apply_act(acc[3], ACT));
      *reinterpret_cast<uint2*>(out + ((static_cast<size_t>(n) * Ho + oy0 + r) * Wo + ox) * out_ps + c4 * 4) = o;
      sum[0] += bf16_lo(o.x); sum[1] += bf16_hi(o.x); sum[2] += bf16_lo(o.y); sum[3] += bf16_hi(o.y);
    }
  }
  if (SQ) {
    const int qn = DW_STRIP_THREADS / slots;
    red[threadIdx.x] = active ? make_float4(sum[0], sum[1], sum[2], sum[3]) : make_float4(0.f, 0.f, 0.f, 0.f);
    __syncthreads();
    if (threadIdx.x < qn && active) {
      float4 s4 = red[threadIdx.x];
      const int strips = strips_h * strips_w;
      for (int j = 1; j < strips; ++j) {            // fixed order
        const float4 v = red[j * qn + threadIdx.x];
        s4.x += v.x; s4.y += v.y; s4.z += v.z; s4.w += v.w;
      }
      *reinterpret_cast<float4*>(se_partial + static_cast<size_t>(n) * C + c4 * 4) = s4;
    }
  } else if (se_partial) {
    const size_t strip = (static_cast<size_t>(n) * strips_h + sy) * strips_w + sx;
    *reinterpret_cast<float4*>(se_partial + strip * C + c4 * 4) = make_float4(sum[0], sum[1], sum[2], sum[3]);
  }
}

// Per-pixel fallback: one thread owns 8 channels (16 B) of one output pixel.
__global__ void __launch_bounds__(256)
dwconv3x3_kernel(const uint16_t* __restrict__ in, int H, int W, int in_ps, uint16_t* __restrict__ out, int Ho, int Wo,
                 int out_ps, int C8, int stride, const float* __restrict__ w, const float* __restrict__ bias, int act,
                 long long total) {
  pdl_wait();   // launched with programmatic stream serialization: inputs are valid from here on (no early trigger:
                // a multi-wave kernel must not share its SMs with the waiting blocks of its successor)
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

static int dw_strip_count(int Ho, int Wo, int stride) {
  if (Wo % DW_TW) return 0;
  const int r = stride == 1 ? 2 : 1;
  return ((Ho + r - 1) / r) * (Wo / DW_TW);
}

// rows of channel sums the fused squeeze writes per image: one per strip (or 1 with B200_DW_SQ=1 when the strips of an
// image fit one block); 0 = this geometry has no strip kernel
static bool dw_block_reduce() {      // B200_DW_SQ=1: reduce an image's strips inside the kernel (one row of sums per image)
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200_DW_SQ");
    v = (e && e[0] == '1') ? 1 : 0;    // measured on EfficientNetV2-S @128: the strip-major layout it needs costs more
  }                                    // in the depthwise kernel (+1.5 us a layer) than the gate saves
  return v == 1;
}

int dwconv3x3_strips_c(int C, int Ho, int Wo, int stride) {
  if (dwconv3x3_tile_squeeze_rows(C, Ho, Wo, stride)) return 1;      // the tile kernel reduces the strips itself
  return dwconv3x3_strips(Ho, Wo, stride);
}

int dwconv3x3_strips(int Ho, int Wo, int stride) {
  const int s = dw_strip_count(Ho, Wo, stride);
  return (dw_block_reduce() && s <= 32) ? (s > 0 ? 1 : 0) : s;
}

int dwconv3x3_launch(const View& in, const View& out, const float* w, const float* bias, int stride, int act,
                     float* se_partial, cudaStream_t stream) {
  if (in.C != out.C || in.C % 8 || in.pix_stride % 8 || out.pix_stride % 8)
    return fail(B200_ERR_INVALID, "dwconv: channels must match and be multiples of 8");
  if (stride != 1 && stride != 2) return fail(B200_ERR_INVALID, "dwconv: stride must be 1 or 2");
  if ((in.H + 2 - 3) / stride + 1 != out.H || (in.W + 2 - 3) / stride + 1 != out.W || in.N != out.N)
    return fail(B200_ERR_INVALID, "dwconv: output geometry mismatch");
  static int use_strips = -1;
  if (use_strips < 0) {
    const char* e = getenv("B200_DW_STRIPS");
    use_strips = (e && e[0] == '0') ? 0 : 1;
  }
  {
    const int rc = dwconv3x3_tile_try(in, out, w, bias, stride, act, se_partial, stream);
    if (rc <= 0) return rc;        // launched (or failed); 1 = not a layer for the tile kernel
  }
  const int strips = dw_strip_count(out.H, out.W, stride);
  if (strips > 0 && (use_strips || se_partial)) {
    const int r = stride == 1 ? 2 : 1;
    const int strips_h = (out.H + r - 1) / r, strips_w = out.W / DW_TW;
    const bool sq = se_partial != nullptr && strips <= 32 && dw_block_reduce();
    int slots = 1;
    while (slots < strips) slots <<= 1;
    const int qn = DW_STRIP_THREADS / slots;
    const long long total = static_cast<long long>(out.N) * strips * (out.C / 4);
    if (total >= (1ll << 31)) return fail(B200_ERR_INVALID, "dwconv: tensor too large");
    const unsigned grid = sq ? static_cast<unsigned>(out.N) * static_cast<unsigned>((out.C / 4 + qn - 1) / qn)
                             : static_cast<unsigned>((total + DW_STRIP_THREADS - 1) / DW_STRIP_THREADS);
    const uint16_t* ip = static_cast<const uint16_t*>(in.base);
    uint16_t* op = static_cast<uint16_t*>(out.base);
#define B200_DW_LAUNCH2(S, A, Q)                                                                                       \
  B200_CUDA_OK(launch_pdl(dwconv3x3_strip_kernel<S, A, Q>, dim3(grid), dim3(DW_STRIP_THREADS), size_t(0), stream, ip, in.H,  \
                          in.W, in.pix_stride, op, out.H, out.W, out.pix_stride, make_fast_div(out.C / 4),                   \
                          make_fast_div(strips_h), make_fast_div(strips_w), w, bias, se_partial, static_cast<unsigned>(total), \
                          slots))
#define B200_DW_LAUNCH(S, A) do { if (sq) B200_DW_LAUNCH2(S, A, true); else B200_DW_LAUNCH2(S, A, false); } while (0)
    if (act != ACT_NONE && act != ACT_RELU && act != ACT_SILU)
      return fail(B200_ERR_INVALID, "dwconv: activation must be none, relu or silu");
    if (stride == 1) {
      if (act == ACT_NONE) B200_DW_LAUNCH(1, ACT_NONE);
      else if (act == ACT_RELU) B200_DW_LAUNCH(1, ACT_RELU);
      else B200_DW_LAUNCH(1, ACT_SILU);
    } else {
      if (act == ACT_NONE) B200_DW_LAUNCH(2, ACT_NONE);
      else if (act == ACT_RELU) B200_DW_LAUNCH(2, ACT_RELU);
      else B200_DW_LAUNCH(2, ACT_SILU);
    }
#undef B200_DW_LAUNCH2
#undef B200_DW_LAUNCH
    B200_CUDA_OK(cudaGetLastError());
    return B200_OK;
  }
  if (se_partial) return fail(B200_ERR_INVALID, "dwconv: fused squeeze needs an output width that is a multiple of 7");
  const long long total = static_cast<long long>(out.N) * out.H * out.W * (out.C / 8);
  B200_CUDA_OK(launch_pdl(dwconv3x3_kernel, dim3(static_cast<unsigned>((total + 255) / 256)), dim3(256), size_t(0), stream, static_cast<const uint16_t*>(in.base), in.H, in.W, in.pix_stride, static_cast<uint16_t*>(out.base), out.H, out.W, out.pix_stride, out.C / 8, stride, w, bias, act, total));
  return B200_OK;
}

}  // namespace b200
