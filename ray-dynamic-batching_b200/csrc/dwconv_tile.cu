// Depthwise 3x3 / stride 1 or 2 / pad 1 from TMA-staged tiles (+ folded BatchNorm bias, activation, optional fused squeeze).
//
// The strip kernel of dwconv.cu reads every input element 2.6 times (each thread loads its own 4 x 9 halo window) and is
// bound by what one SM can pull from L2.  Here a persistent block takes work items (image, row band, channel group);
// ONE TMA box per item brings the (rows + 2) x (W + 2) x CG input tile into shared memory -- coordinates outside the image
// are zero-filled, which is the padding -- so global reads are 1.0x (+ one halo row per band), and the 256 compute
// threads read their windows from shared memory.  A 3-4 deep ring of tiles keeps the loads ahead of the math.
//   warps 0-7 : compute.  thread = (channel quad, strip slot); a strip is 2 output rows x 7 columns, windows slide in
//               registers exactly as in the strip kernel (same FMA order: results are bit-identical)
//   warp 8    : TMA producer (whole warp in the loop, one elected lane issues)
// Stride 2: the box is (2 * rows + 1) x (2 * W + 1) input pixels, a strip is 1 output row x 7 columns and the window
// slides by two columns (six shared-memory loads per output instead of four per two).
// With one band per image the strips of an (image, channel group) all live in one block: their channel sums are added in
// strip order in shared memory and written as ONE row sums[n][c] (the SE squeeze, deterministic).
//
// Replaces the stride-1 depthwise convs of tv shufflenetv2.py:86-96 and efficientnet.py:140-151 (+ the AdaptiveAvgPool2d(1)
// of ops/misc.py:249-250).
#include "kernels.cuh"
#include "ptx.cuh"

namespace b200 {

static constexpr int DT_COMPUTE = 256;
static constexpr int DT_THREADS = DT_COMPUTE + 32;
static constexpr int DT_TW = 7;
static constexpr int DT_MAX_STAGES = 4;
static constexpr int DT_SMEM_MAX = 216 * 1024;   // dynamic shared memory the kernel may ask for (8 KB static on top)

struct DwTileParams {
  const float* w;        // [9][C]
  const float* bias;     // [C]
  uint16_t* out;
  float* sums;           // [N][C] or null
  int out_ps;
  int C, H, W, N;
  int cg;                // channels per work item (32, 64, 128 or 256) = TMA box inner dimension
  int stride;
  int groups;            // ceil(C / cg)
  int band_h, bands;     // output rows per item, items per image column of groups
  int stages;
  int tile_pitch;        // tile_bytes rounded up to 128 (TMA destinations are 128-byte aligned)
  int tile_bytes;        // (band_h + 2) * (W + 2) * cg * 2
  int total_items;       // N * bands * groups
  unsigned mul_strips_w; // fast_div multiplier of W / 7
};

template <int ACT, int CG, bool SQ, int STRIDE>
__global__ void __launch_bounds__(DT_THREADS, 1)
dwconv3x3_tile_kernel(const __grid_constant__ CUtensorMap in_map, const __grid_constant__ DwTileParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[DT_MAX_STAGES], bar_empty[DT_MAX_STAGES];
  __shared__ __align__(16) float red[2][DT_COMPUTE * 4];
  const int warp = threadIdx.x >> 5;
  const uint32_t ring = (smem_u32(smem_raw) + 127u) & ~127u;
  // A block takes a CONTIGUOUS run of items, ordered (channel group, image, band): the group -- and with it the 9 + 1
  // weight vectors every thread keeps in registers -- changes once or twice per block instead of with every item.
  const int per = (p.total_items + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const int item0 = blockIdx.x * per;
  const int my_items = max(0, min(per, p.total_items - item0));
  const int per_group = p.N * p.bands;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&in_map);
    for (int s = 0; s < DT_MAX_STAGES; ++s) {
      mbar_init(smem_u32(&bar_full[s]), 1);
      mbar_init(smem_u32(&bar_empty[s]), DT_COMPUTE / 32);     // one arrival per compute warp
    }
    fence_barrier_init();
  }
  __syncthreads();
  pdl_launch_dependents();
  pdl_wait();

  if (warp == DT_COMPUTE / 32) {
    // =========================== TMA producer ===========================
    int g = item0 / per_group, r = item0 - g * per_group;
    int n = r / p.bands, band = r - n * p.bands;       // running (group, image, band) and ring (stage, phase) counters:
    int s = 0;                                         // no division per item in a warp that runs alone
    uint32_t ph = 0;
    for (int it = 0; it < my_items; ++it) {
      mbar_wait(smem_u32(&bar_empty[s]), ph ^ 1);
      if (elect_one()) {
        const uint32_t full = smem_u32(&bar_full[s]);
        mbar_expect_tx(full, p.tile_bytes);
        // box [CG, W + 2, band_h + 2, 1] at (g * CG, -1, band * band_h - 1, n): outside the image -> zeros
        // (stride 2: [CG, 2 W + 1, 2 band_h + 1, 1] at input row 2 * band * band_h - 1)
        tma_load_4d(&in_map, full, ring + s * p.tile_pitch, g * CG, -1, STRIDE * band * p.band_h - 1, n);
      }
      if (++s == p.stages) { s = 0; ph ^= 1; }
      if (++band == p.bands) { band = 0; ++n; }
      if (++r == per_group) { r = 0; n = 0; ++g; }
    }
    __syncwarp();
    return;
  }

  // =========================== compute ===========================
  // The inner loop is instruction-issue bound (ncu: 111 instructions per 4-channel output before this layout, 18 of
  // them the FFMA2s), so everything that is not math is hoisted: CG is a template parameter (shared-memory offsets of
  // the window loads are immediates off four row registers), stores walk two row pointers, validity is two
  // per-strip predicates instead of a branch per output, SiLU and the squeeze sums run on packed f32x2 lanes.
  constexpr int QUADS = CG >> 2;                // channel quads per item
  constexpr int NSLOTS = DT_COMPUTE / QUADS;    // strips worked on concurrently
  const int quad = threadIdx.x % QUADS;
  const int slot = threadIdx.x / QUADS;
  const int strips_w = p.W / DT_TW;
  const uint32_t row_pitch = static_cast<uint32_t>(STRIDE == 1 ? p.W + 2 : 2 * p.W + 1) * CG * 2;   // bytes per staged input row
  const size_t out_row = static_cast<size_t>(p.W) * p.out_ps;              // elements per output row
  float2 wt[9][2], b[2];
  int g = item0 / per_group, r = item0 - g * per_group;
  int n = r / p.bands, band = r - n * p.bands;
  int s = 0;
  uint32_t ph = 0;
  int g_loaded = -1;
  const uint32_t mul_strips_w = p.mul_strips_w;
  for (int it = 0; it < my_items; ++it) {
    const int c0 = g * CG + quad * 4;
    const bool c_ok = c0 < p.C;                 // the last group of a ragged channel count
    if (g != g_loaded) {
      g_loaded = g;
      const int cc = c_ok ? c0 : 0;
#pragma unroll
      for (int k = 0; k < 9; ++k) {
        const float4 v = *reinterpret_cast<const float4*>(p.w + k * p.C + cc);
        wt[k][0] = make_float2(v.x, v.y); wt[k][1] = make_float2(v.z, v.w);
      }
      const float4 v = *reinterpret_cast<const float4*>(p.bias + cc);
      b[0] = make_float2(v.x, v.y); b[1] = make_float2(v.z, v.w);
    }
    const int rows = min(p.band_h, p.H - band * p.band_h);      // output rows of this item
    const int strips = (STRIDE == 1 ? (rows + 1) >> 1 : rows) * strips_w;
    mbar_wait(smem_u32(&bar_full[s]), ph);
    const uint32_t tile = ring + s * p.tile_pitch + quad * 8;
    uint16_t* const out_item = p.out + (static_cast<size_t>(n) * p.H + band * p.band_h) * out_row + c0;
    float2 sum0 = make_float2(0.f, 0.f), sum1 = make_float2(0.f, 0.f);
    for (int st = slot; st < strips; st += NSLOTS) {
      const int sy = fast_div(st, strips_w, mul_strips_w), sx = st - sy * strips_w;
#define DT_LDS(i, j, v)                                                                                              \
  {                                                                                                                  \
    uint32_t lo_, hi_;                                                                                               \
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(lo_), "=r"(hi_) : "r"(rowa[i] + (j) * (CG * 2)));        \
    v[0] = make_float2(bf16_lo(lo_), bf16_hi(lo_));                                                                  \
    v[1] = make_float2(bf16_lo(hi_), bf16_hi(hi_));                                                                  \
  }
      if constexpr (STRIDE == 1) {
        const int oy0 = sy * 2, ox0 = sx * DT_TW;                 // within the band
        const bool ok[2] = {c_ok, c_ok && oy0 + 1 < rows};        // an odd row count leaves the strip's second row outside
        // staged tile: row i <-> input row (band0 - 1 + i), column j <-> input column (j - 1)
        uint32_t rowa[4];
        rowa[0] = tile + oy0 * row_pitch + ox0 * (CG * 2);
#pragma unroll
        for (int i = 1; i < 4; ++i) rowa[i] = rowa[i - 1] + row_pitch;
        uint16_t* outp[2];
        outp[0] = out_item + oy0 * out_row + static_cast<size_t>(ox0) * p.out_ps;
        outp[1] = outp[0] + out_row;
        float2 win[4][3][2];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          DT_LDS(i, 0, win[i][1]);
          DT_LDS(i, 1, win[i][2]);
        }
#pragma unroll
        for (int j = 0; j < DT_TW; ++j) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            win[i][0][0] = win[i][1][0]; win[i][0][1] = win[i][1][1];
            win[i][1][0] = win[i][2][0]; win[i][1][1] = win[i][2][1];
            DT_LDS(i, j + 2, win[i][2]);
          }
#pragma unroll
          for (int rr = 0; rr < 2; ++rr) {
            float2 acc0 = b[0], acc1 = b[1];
#pragma unroll
            for (int dr = 0; dr < 3; ++dr)
#pragma unroll
              for (int q = 0; q < 3; ++q) {      // same (row, column) order on top of the bias as the strip kernel: bit-identical
                acc0 = ffma2(win[rr + dr][q][0], wt[dr * 3 + q][0], acc0);
                acc1 = ffma2(win[rr + dr][q][1], wt[dr * 3 + q][1], acc1);
              }
            acc0 = apply_act2<ACT>(acc0);
            acc1 = apply_act2<ACT>(acc1);
            uint2 o;
            o.x = pack_bf16x2(acc0.x, acc0.y);
            o.y = pack_bf16x2(acc1.x, acc1.y);
            if (ok[rr]) {
              *reinterpret_cast<uint2*>(outp[rr]) = o;
              if (SQ) {     // sums of the ROUNDED outputs, as the squeeze of the stored tensor would see them
                sum0 = fadd2(sum0, make_float2(bf16_lo(o.x), bf16_hi(o.x)));
                sum1 = fadd2(sum1, make_float2(bf16_lo(o.y), bf16_hi(o.y)));
              }
            }
            outp[rr] += p.out_ps;
          }
        }
      } else {
        // staged tile: row i <-> input row (2 * band0 - 1 + i), column j <-> input column (j - 1); output (oy0, ox0 + j)
        // reads rows 2 oy0 + dr and columns 2 (ox0 + j) + q of it
        const int oy0 = sy, ox0 = sx * DT_TW;
        uint32_t rowa[3];
        rowa[0] = tile + (2 * oy0) * row_pitch + (2 * ox0) * (CG * 2);
        rowa[1] = rowa[0] + row_pitch;
        rowa[2] = rowa[1] + row_pitch;
        uint16_t* outp = out_item + oy0 * out_row + static_cast<size_t>(ox0) * p.out_ps;
        float2 win[3][3][2];
#pragma unroll
        for (int i = 0; i < 3; ++i) DT_LDS(i, 0, win[i][2]);
#pragma unroll
        for (int j = 0; j < DT_TW; ++j) {
#pragma unroll
          for (int i = 0; i < 3; ++i) {
            win[i][0][0] = win[i][2][0]; win[i][0][1] = win[i][2][1];
            DT_LDS(i, 2 * j + 1, win[i][1]);
            DT_LDS(i, 2 * j + 2, win[i][2]);
          }
          float2 acc0 = b[0], acc1 = b[1];
#pragma unroll
          for (int dr = 0; dr < 3; ++dr)
#pragma unroll
            for (int q = 0; q < 3; ++q) {      // same order as the strip kernel: bit-identical
              acc0 = ffma2(win[dr][q][0], wt[dr * 3 + q][0], acc0);
              acc1 = ffma2(win[dr][q][1], wt[dr * 3 + q][1], acc1);
            }
          acc0 = apply_act2<ACT>(acc0);
          acc1 = apply_act2<ACT>(acc1);
          uint2 o;
          o.x = pack_bf16x2(acc0.x, acc0.y);
          o.y = pack_bf16x2(acc1.x, acc1.y);
          if (c_ok) {
            *reinterpret_cast<uint2*>(outp) = o;
            if (SQ) {
              sum0 = fadd2(sum0, make_float2(bf16_lo(o.x), bf16_hi(o.x)));
              sum1 = fadd2(sum1, make_float2(bf16_lo(o.y), bf16_hi(o.y)));
            }
          }
          outp += p.out_ps;
        }
      }
#undef DT_LDS
    }
    __syncwarp();
    if ((threadIdx.x & 31) == 0) mbar_arrive(smem_u32(&bar_empty[s]));      // this warp is done reading the tile
    if (SQ) {
      // one band per image (checked by the host): every strip of (n, channel quad) is in this block.  red[] is double
      // buffered by item parity: whoever writes red[it & 1] has passed the barrier of item it - 1, which the reducing
      // threads reach only after they have read red[it & 1] for item it - 2.
      float* const rd = red[it & 1];
      *reinterpret_cast<float4*>(&rd[threadIdx.x * 4]) = make_float4(sum0.x, sum0.y, sum1.x, sum1.y);
      named_bar_sync(1, DT_COMPUTE);
      if (slot == 0 && c_ok) {
        float4 t = *reinterpret_cast<float4*>(&rd[quad * 4]);
#pragma unroll
        for (int k = 1; k < NSLOTS; ++k) {                                   // slot order = strip order: deterministic
          const float4 v = *reinterpret_cast<float4*>(&rd[(k * QUADS + quad) * 4]);
          t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w;
        }
        *reinterpret_cast<float4*>(p.sums + static_cast<size_t>(n) * p.C + c0) = t;
      }
    }
    if (++s == p.stages) { s = 0; ph ^= 1; }
    if (++band == p.bands) { band = 0; ++n; }
    if (++r == per_group) { r = 0; n = 0; ++g; }
  }
}

// ---------------------------------------------------------------------------------------------------------------- host
static bool dw_tile_s2_enabled() {      // B200_DW_TILE_S2=0: stride-2 layers stay on the strip kernel
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200_DW_TILE_S2");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1;
}

static bool dw_tile_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200_DW_TILE");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1;
}

// Geometry the tile kernel takes: OUTPUT width a multiple of 7 and at most 56, at least two tiles in the ring.
// Stride 1: bands of 14 rows, channel group by how many strips a band has.  Stride 2 (a tile is four times the output):
// the (channel group, band height) pair with the best product of slot use, halo overhead and channel padding.
// one_band: the caller wants the fused squeeze, which needs a whole image per item.
static bool dw_tile_plan(int C, int H, int W, int stride, bool one_band, DwTileParams* p) {
  if (W % DT_TW || W > 56 || H < 1 || C % 8) return false;
  p->stride = stride;
  if (stride == 1) {
    p->band_h = H <= 14 ? H : 14;
    p->bands = (H + p->band_h - 1) / p->band_h;
    const int strips = ((p->band_h + 1) / 2) * (W / DT_TW);
    int cg = strips >= 12 ? 64 : (strips >= 6 ? 128 : 256);
    const int c64 = (C + 63) / 64 * 64;
    if (cg > c64) cg = c64 >= 256 ? 256 : (c64 >= 128 ? 128 : 64);
    p->cg = cg;
    p->groups = (C + cg - 1) / cg;
    p->tile_bytes = (p->band_h + 2) * (W + 2) * cg * 2;
    const int row_bytes = (W + 2) * cg * 2;      // slack after the ring: odd row counts read one row past a tile
    p->stages = (DT_SMEM_MAX - row_bytes - 256) / p->tile_bytes;
    if (p->stages > DT_MAX_STAGES) p->stages = DT_MAX_STAGES;
    return p->stages >= 2;
  }
  double best = 0.0;
  static const int cgs[4] = {32, 64, 128, 256};
  for (int ci = 0; ci < 4; ++ci) {
    const int cg = cgs[ci];
    if (cg == 32 && C > 32) continue;
    if (cg > 64 && cg > (C + 63) / 64 * 64) continue;
    const int groups = (C + cg - 1) / cg, nslots = DT_COMPUTE / (cg / 4);
    for (int bh = H; bh >= (one_band ? H : 1); --bh) {
      const int tile_bytes = (2 * bh + 1) * (2 * W + 1) * cg * 2;
      int stages = (DT_SMEM_MAX - 256) / ((tile_bytes + 127) & ~127);
      if (stages < 2) continue;
      if (stages > DT_MAX_STAGES) stages = DT_MAX_STAGES;
      const int strips = bh * (W / DT_TW);
      const int rounds = (strips + nslots - 1) / nslots;
      const double score = double(strips) / (rounds * nslots) * (2.0 * bh) / (2 * bh + 1) * double(C) / (groups * cg) *
                           (stages >= 3 ? 1.0 : 0.85) * ((H + bh - 1) / bh * bh == H ? 1.0 : 0.9);
      if (score > best + 1e-9) {
        best = score;
        p->cg = cg; p->groups = groups; p->band_h = bh; p->bands = (H + bh - 1) / bh;
        p->tile_bytes = tile_bytes; p->stages = stages;
      }
    }
  }
  return best > 0.0;
}

// 1 when the tile kernel would take this layer and writes ONE row of channel sums per image, else 0
int dwconv3x3_tile_squeeze_rows(int C, int H, int W, int stride) {
  DwTileParams p;
  if (!dw_tile_enabled() || (stride == 2 && !dw_tile_s2_enabled()) || !dw_tile_plan(C, H, W, stride, true, &p)) return 0;
  return p.bands == 1 ? 1 : 0;
}

template <int ACT, int CG, bool SQ, int STRIDE>
static int dw_tile_launch_t(const CUtensorMap& map, const DwTileParams& p, size_t smem, cudaStream_t stream) {
  static bool configured[64] = {};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(dwconv3x3_tile_kernel<ACT, CG, SQ, STRIDE>, cudaFuncAttributeMaxDynamicSharedMemorySize, DT_SMEM_MAX));
    configured[dev & 63] = true;
  }
  const int ctas = p.total_items < device_sm_count() ? p.total_items : device_sm_count();
  B200_CUDA_OK(launch_pdl(dwconv3x3_tile_kernel<ACT, CG, SQ, STRIDE>, dim3(ctas), dim3(DT_THREADS), smem, stream, map, p));
  return B200_OK;
}

template <int ACT, int STRIDE>
static int dw_tile_launch_a(const CUtensorMap& map, const DwTileParams& p, size_t smem, cudaStream_t stream) {
  const bool sq = p.sums != nullptr;
  switch (p.cg) {
    case 32: return sq ? dw_tile_launch_t<ACT, 32, true, STRIDE>(map, p, smem, stream) : dw_tile_launch_t<ACT, 32, false, STRIDE>(map, p, smem, stream);
    case 64: return sq ? dw_tile_launch_t<ACT, 64, true, STRIDE>(map, p, smem, stream) : dw_tile_launch_t<ACT, 64, false, STRIDE>(map, p, smem, stream);
    case 128: return sq ? dw_tile_launch_t<ACT, 128, true, STRIDE>(map, p, smem, stream) : dw_tile_launch_t<ACT, 128, false, STRIDE>(map, p, smem, stream);
    default: return sq ? dw_tile_launch_t<ACT, 256, true, STRIDE>(map, p, smem, stream) : dw_tile_launch_t<ACT, 256, false, STRIDE>(map, p, smem, stream);
  }
}

template <int STRIDE>
static int dw_tile_launch_s(int act, const CUtensorMap& map, const DwTileParams& p, size_t smem, cudaStream_t stream) {
  if (act == ACT_NONE) return dw_tile_launch_a<ACT_NONE, STRIDE>(map, p, smem, stream);
  if (act == ACT_RELU) return dw_tile_launch_a<ACT_RELU, STRIDE>(map, p, smem, stream);
  return dw_tile_launch_a<ACT_SILU, STRIDE>(map, p, smem, stream);
}

// returns B200_OK after launching, or 1 if this layer is not for the tile kernel (the caller falls back to the strip kernel)
int dwconv3x3_tile_try(const View& in, const View& out, const float* w, const float* bias, int stride, int act,
                       float* sums, cudaStream_t stream) {
  DwTileParams p;
  if (!dw_tile_enabled() || (stride == 2 && !dw_tile_s2_enabled()) || !dw_tile_plan(out.C, out.H, out.W, stride, sums != nullptr, &p)) return 1;
  if (sums && p.bands != 1) return 1;
  if (act != ACT_NONE && act != ACT_RELU && act != ACT_SILU) return 1;
  if (in.pix_stride % 8 || (reinterpret_cast<uintptr_t>(in.base) & 15)) return 1;
  p.w = w; p.bias = bias;
  p.out = static_cast<uint16_t*>(out.base);
  p.sums = sums;
  p.out_ps = out.pix_stride;
  p.C = out.C; p.H = out.H; p.W = out.W; p.N = out.N;
  p.total_items = out.N * p.bands * p.groups;
  p.mul_strips_w = fast_div_mul(out.W / DT_TW);
  CUtensorMap map;
  {
    const uint64_t ips = uint64_t(in.pix_stride) * 2;
    const uint64_t dims[4] = {uint64_t(in.C), uint64_t(in.W), uint64_t(in.H), uint64_t(in.N)};
    const uint64_t str[3] = {ips, ips * in.W, ips * in.W * in.H};
    const uint32_t box[4] = {uint32_t(p.cg), uint32_t(stride == 1 ? out.W + 2 : 2 * out.W + 1),
                             uint32_t(stride == 1 ? p.band_h + 2 : 2 * p.band_h + 1), 1u};
    B200_TRY(encode_tmap(&map, in.base, 4, dims, str, box, SWZ_NONE));
  }
  p.tile_pitch = (p.tile_bytes + 127) & ~127;
  // stride 1: + one row, odd row counts read one row past the tile
  const size_t smem = size_t(p.stages) * p.tile_pitch + (stride == 1 ? size_t(in.W + 2) * p.cg * 2 : 0) + 256;
  return stride == 1 ? dw_tile_launch_s<1>(act, map, p, smem, stream) : dw_tile_launch_s<2>(act, map, p, smem, stream);
}

}  // namespace b200
