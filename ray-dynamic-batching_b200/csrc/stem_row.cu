// ResNet stem (7x7 / stride 2 / pad 3, 3 -> 64, + folded BN + ReLU) without any im2col copy: the tensor core reads the
// sliding windows straight out of the staged input rows.
//
// With the input stored NHWC with C padded to 4, the 8-pixel aligned window of output pixel ow in kernel row r is the 64
// contiguous bytes starting at pixel 2*ow - 4 of input row 2*oh + r - 3, i.e. at byte ow * 16 of the row: consecutive output
// pixels' windows are the same bytes shifted by 16.  That Toeplitz structure is exactly what a *non-swizzled* K-major
// shared-memory descriptor expresses when its strides are chosen as LBO (next 16-byte K chunk) = 16 B and SBO (next group
// of 8 rows) = 128 B:  address(row m, chunk k) = base + m*16 + k*16.  So one TMA box per tile brings the 7 input rows of
// an output row (zero-filled outside the image = the padding), and 7 x 2 tcgen05.mma (M=128, N=64, K=16) consume them in
// place; GEMM-K is (kernel row, window pixel, channel) = 7 x 32 with zero weights for window pixel 0 and channel 3.
//
//   warp 0 : TMA producer (4-stage ring of 7-row boxes)        warp 1 : TMEM alloc + single-thread MMA issue
//   warps 4-7 : epilogue (TMEM -> +bias -> ReLU -> bf16 -> swizzled smem -> TMA store clipped at the row end)
//
// Replaces tv resnet.py:268-270 (conv1 + bn1 + relu) as reached from scheduler.py:450.
#include <stdlib.h>

#include "kernels.cuh"
#include "ptx.cuh"

namespace b200 {

static constexpr int SR_THREADS = 256;
static constexpr int SR_WGS = 4;          // stem_row_kernel: epilogue warpgroups = accumulators in TMEM
static constexpr int SR2_THREADS = 128 + SR_WGS * 128;
// An input row is staged as 18 groups of 16 pixels (128 B each): TMA moves whole 128-byte lines (a box with a 16-byte inner
// dimension costs 8x more TMA work).  The box starts 16 pixels left of the tile's first *output* pixel pair, i.e. 12 pixels
// (96 B) before the first window, so descriptors start at +96.
static constexpr int SR_GROUPS = 18;
static constexpr int SR_ROW_BYTES = SR_GROUPS * 128; // 2304
static constexpr int SR_OUT_BYTES = 128 * 128;       // one staged output tile (128 pixels x 64 channels bf16)

// Geometry of a stride-2 stem with a KS x KS kernel and pad KS / 2 (7x7 pad 3: ResNet; 3x3 pad 1: ShuffleNetV2,
// EfficientNetV2, YOLOv8), NOUT = accumulator width (cout rounded up to 32 or 64).  The aligned window of output pixel ow
// starts at input pixel 2*ow - pad - 1 and is KC 16-byte chunks (2 pixels each) long; GEMM-K per kernel row is KC * 8.
template <int KS, int NOUT_>
struct StemCfg {
  static constexpr int KR = KS;                       // kernel rows = input rows per output row
  // input boxes in flight.  A 3x3 stem's box is 6.9 KB: four of them per SM are 4 MB over the GPU, and with ~1.5 us of loaded
  // latency that caps the input stream at ~2.7 TB/s (ncu: the epilogue groups wait for accumulators, the tensor pipe is 6 %
  // busy) -> ten.  A 7x7 stem's box is 16 KB and the kernel is paced by its 14 MMAs per row -> four.
  static constexpr int STAGES = KS == 3 ? 10 : 4;
  // fused max-pool: stem rows kept in shared memory.  Pooled row ph needs rows 2ph-1 .. 2ph+1 and frees two of them only
  // when it is done; with four slots the drain groups could run one row ahead of the pooling group and waited for a
  // free slot half of the time (ncu, 3x3 stem).  Eight (six next to the 7x7 stem's larger input ring) decouple them.
  static constexpr int POOL_RING = KS == 3 ? 8 : 6;
  // pooling threads of the fused kernel: two warpgroups for the 3x3 stems (measured 197.7 -> 174.8 us, ShuffleNetV2 @512);
  // the 7x7 / 64-wide stem is paced by its 14 MMAs per row and ran 3 % slower with the second group (188 -> 194 us)
  static constexpr int POOL_THREADS = KS == 3 ? 256 : 128;
  static constexpr int PAD = KS / 2;
  static constexpr int KC = (KS + 2) / 2 & ~1;        // 7 -> 4 chunks (8 pixels), 3 -> 2 chunks (4 pixels)
  static constexpr int MMAS = KC / 2;                 // K = 16 (two chunks) per MMA
  static constexpr int NOUT = NOUT_;
  static constexpr int LEAD_BYTES = (15 - PAD) * 8;   // the box starts 16 pixels left of the tile's first output pixel pair
  static constexpr int STAGE_BYTES = KR * SR_ROW_BYTES;
  static constexpr int W_BYTES = KR * KC * NOUT * 16; // weights: [kernel row][k chunk][cout][8 bf16]
  static constexpr int TMEM_COLS = 2 * NOUT < 32 ? 32 : 2 * NOUT;
  // fused kernels: four accumulators, so the MMAs of row it + 2 do not wait for the drain of row it
  // (measured on the ring kernel: 174.8 -> 172.8 us for the 3x3 stem, 188.4 -> 185.8 us for the 7x7 one)
  static constexpr int POOL_ACCS = 4;
  static constexpr int POOL_TMEM_COLS = POOL_ACCS * NOUT;
  static_assert(KS == 7 || KS == 3, "stem kernel size");
  static_assert(NOUT == 32 || NOUT == 64, "stem accumulator width");
};

// Non-swizzled K-major descriptor: start | LBO (stride between the two 16-byte K chunks of one MMA) | SBO (stride between
// groups of 8 rows) | version 1 | layout type 0.
__device__ __forceinline__ uint64_t umma_desc_noswz(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  const uint32_t lo = ((addr >> 4) & 0x3FFFu) | ((lbo_bytes >> 4) << 16);
  const uint32_t hi = (sbo_bytes >> 4) | (1u << 14);
  return (static_cast<uint64_t>(hi) << 32) | lo;
}

struct StemRowParams {
  const uint16_t* w;     // bf16, smem image of the weights (SR_W_BYTES)
  const float* bias;     // [64]
  int Ho, Wo, segs;      // segs = ceil(Wo / 128)
  int total_tiles;       // N * Ho * segs
  unsigned mul_segs, mul_Ho;   // fast_div multipliers
};

template <int KS, int NOUT, int ACT>
__global__ void __launch_bounds__(SR2_THREADS, 1)
stem_row_kernel(const __grid_constant__ CUtensorMap in_map, const __grid_constant__ CUtensorMap out_map,
                const __grid_constant__ StemRowParams p) {
  using G = StemCfg<KS, NOUT>;
  constexpr int SR_STAGE_BYTES = G::STAGE_BYTES, SR_W_BYTES = G::W_BYTES, SR_LEAD_BYTES = G::LEAD_BYTES;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[G::STAGES], bar_empty[G::STAGES], bar_t_full[SR_WGS], bar_t_empty[SR_WGS];
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float bias_s[64];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t ring = base;
  const uint32_t wsm = ring + G::STAGES * SR_STAGE_BYTES + 256;   // +256: rows m >= 112 of the last kernel row read past it
  const uint32_t stage_out = (wsm + SR_W_BYTES + 1023u) & ~1023u;
  const int per_cta = (p.total_tiles + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const int tile0 = static_cast<int>(blockIdx.x) * per_cta;
  const int my_tiles = max(0, min(per_cta, p.total_tiles - tile0));

  // weights -> smem (already in the descriptor layout), bias
  for (int i = threadIdx.x; i < SR_W_BYTES / 16; i += SR2_THREADS) {
    const uint4 v = reinterpret_cast<const uint4*>(p.w)[i];
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(wsm + i * 16), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
  }
  if (threadIdx.x < NOUT) bias_s[threadIdx.x] = p.bias[threadIdx.x];
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&in_map);
    tma_prefetch_desc(&out_map);
    for (int s = 0; s < G::STAGES; ++s) {
      mbar_init(smem_u32(&bar_full[s]), 1);
      mbar_init(smem_u32(&bar_empty[s]), 1);
    }
    for (int b = 0; b < SR_WGS; ++b) {
      mbar_init(smem_u32(&bar_t_full[b]), 1);
      mbar_init(smem_u32(&bar_t_empty[b]), 128);
    }
    fence_barrier_init();
  } else if (warp == 1) {
    tmem_alloc<SR_WGS * NOUT>(smem_u32(&tmem_base_s));
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  pdl_launch_dependents();
  pdl_wait();      // everything above overlapped the previous kernel's tail; activations are read from here on

  auto tile_coords = [&](int it, int& n_oh, int& oh, int& ow0) {
    const int t = tile0 + it;
    n_oh = fast_div(t, p.segs, p.mul_segs);               // = n * Ho + oh
    const int seg = t - n_oh * p.segs;
    oh = n_oh - static_cast<int>(fast_div(n_oh, p.Ho, p.mul_Ho)) * p.Ho;
    ow0 = seg * 128;
  };

  if (warp == 0) {
    // whole warp in the loop, one elected lane issues (see elect_one in ptx.cuh)
    for (int it = 0; it < my_tiles; ++it) {
      int n_oh, oh, ow0;
      tile_coords(it, n_oh, oh, ow0);
      const int s = it % G::STAGES;
      mbar_wait(smem_u32(&bar_empty[s]), ((it / G::STAGES) & 1) ^ 1);
      const uint32_t full = smem_u32(&bar_full[s]);
      if (elect_one()) {
        mbar_expect_tx(full, G::KR * SR_ROW_BYTES);
        // box [64 (16 pixels), 18 groups, 7 rows, 1 image] at (0, ow0/8 - 1, 2*oh - 3, n); outside the image -> zeros
        tma_load_4d(&in_map, full, ring + s * SR_STAGE_BYTES, 0, ow0 / 8 - 1, 2 * oh - G::PAD, fast_div(n_oh, p.Ho, p.mul_Ho));
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    {
      constexpr uint32_t idesc = umma_idesc_bf16_f32(128, NOUT);
      for (int it = 0; it < my_tiles; ++it) {
        const int b = it % SR_WGS, s = it % G::STAGES;
        mbar_wait(smem_u32(&bar_t_empty[b]), ((it / SR_WGS) & 1) ^ 1);
        mbar_wait(smem_u32(&bar_full[s]), (it / G::STAGES) & 1);
        tc_fence_after();
        const uint32_t a0 = ring + s * SR_STAGE_BYTES;
        if (elect_one()) {
#pragma unroll
          for (int r = 0; r < G::KR; ++r) {
#pragma unroll
            for (int kk = 0; kk < G::MMAS; ++kk) {
              // A: row m at +m*16, K chunk at +16  (overlapping windows);  B: [k chunk][cout][8]: chunk stride NOUT*16, 8-row stride 128
              const uint64_t adesc = umma_desc_noswz(a0 + r * SR_ROW_BYTES + SR_LEAD_BYTES + kk * 32, 16, 128);
              const uint64_t bdesc = umma_desc_noswz(wsm + (r * G::KC + kk * 2) * (NOUT * 16), NOUT * 16, 128);
              umma_bf16(tmem_base + b * NOUT, adesc, bdesc, idesc, (r | kk) != 0 ? 1u : 0u);
            }
          }
          umma_commit(smem_u32(&bar_empty[s]));
          umma_commit(smem_u32(&bar_t_full[b]));
        }
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    // SR_WGS epilogue warpgroups: group g drains accumulator g of every SR_WGS-th tile, so the per-tile latency chains
    // (TMEM load -> convert -> smem -> store, one warp per scheduler) of SR_WGS tiles overlap
    const int wg = (warp - 4) >> 2;              // 0 .. SR_WGS - 1
    const int epi_tid = (threadIdx.x - 128) & 127;
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int sw = row & 7;
    for (int it = wg; it < my_tiles; it += SR_WGS) {
      const int b = wg;
      int n_oh, oh, ow0;
      tile_coords(it, n_oh, oh, ow0);
      if (epi_tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // this group's previous store has read staging[b]
      named_bar_sync(1 + 2 * wg, 128);
      mbar_wait(smem_u32(&bar_t_full[b]), (it / SR_WGS) & 1);
      tc_fence_after();
      const uint32_t dst_row = stage_out + b * SR_OUT_BYTES + row * 128;
      const uint32_t acc = tmem_base + b * NOUT + (static_cast<uint32_t>(quarter * 32) << 16);
#pragma unroll
      for (int cc = 0; cc < NOUT / 32; ++cc) {
        uint32_t v[32];
        tmem_ld_32x32(acc + cc * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int chunk = cc * 4 + g;
          float f[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) f[i] = apply_act(__uint_as_float(v[g * 8 + i]) + bias_s[chunk * 8 + i], ACT);
          const uint32_t o0 = pack_bf16x2(f[0], f[1]), o1 = pack_bf16x2(f[2], f[3]);
          const uint32_t o2 = pack_bf16x2(f[4], f[5]), o3 = pack_bf16x2(f[6], f[7]);
          asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(dst_row + ((chunk ^ sw) << 4)), "r"(o0), "r"(o1),
                       "r"(o2), "r"(o3)
                       : "memory");
        }
      }
      tc_fence_before();
      mbar_arrive(smem_u32(&bar_t_empty[b]));
      fence_proxy_async_smem();
      named_bar_sync(2 + 2 * wg, 128);
      if (epi_tid == 0) {
        // 3-D store [64 ch, 128 pixels, 1 row] at (0, ow0, n*Ho + oh): pixels past Wo are clipped
        asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4}], [%1];"
                     ::"l"(reinterpret_cast<uint64_t>(&out_map)), "r"(stage_out + b * SR_OUT_BYTES), "r"(0), "r"(ow0), "r"(n_oh)
                     : "memory");
        tma_store_commit();
      }
    }
    if (epi_tid == 0) tma_store_wait_all();
  }

  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<SR_WGS * NOUT>(tmem_base);
  }
}

// ----------------------------------------------------------------------------------------------------------------
// Fused stem + MaxPool2d(3, 2, 1): the 112x112x64 stem activation (1.6 MB per image) never goes to HBM.
// A work unit is (image, band of SP_BAND pooled rows); the CTA walks the stem rows of the band in order, the epilogue
// keeps the last three rows (bf16, swizzled) in shared memory and emits pooled row ph as soon as stem row 2*ph+1 exists.
// One stem row per band is recomputed (its top halo).  Requires Wo <= 128 (one MMA tile per stem row) and even Ho, Wo.
// Replaces tv resnet.py:268-271 (conv1, bn1, relu, maxpool).
// ----------------------------------------------------------------------------------------------------------------
static constexpr int SP_BAND = 7;

struct StemPoolParams {
  const uint16_t* w;
  const float* bias;
  uint16_t* out;        // pooled NHWC [N][Ho/2][Wo/2][out_ps]
  int out_ps;
  int cout;             // stored channels (multiple of 8, <= NOUT)
  int Ho, Wo;           // stem output size
  int bands;            // bands per image = ceil((Ho/2) / SP_BAND)
  int total_units;      // N * bands
  unsigned mul_ch;      // fast_div multiplier of cout / 8
};

__device__ __forceinline__ uint32_t bf16x2_max_u32(uint32_t a, uint32_t b) {
  __nv_bfloat162 x = *reinterpret_cast<__nv_bfloat162*>(&a), y = *reinterpret_cast<__nv_bfloat162*>(&b);
  __nv_bfloat162 m = __hmax2(x, y);
  return *reinterpret_cast<uint32_t*>(&m);
}

// warps 0: TMA, 1: MMA, 4-7 / 8-11: drain of accumulator 0 / 1, 12-19: pooling.  Two pooling warpgroups: an ncu capture of the
// 3x3 / 32-wide variant (ShuffleNetV2 @512) had the drain warps waiting for a free row slot 57 % of the time and the single
// pooling warpgroup busy or waiting for the last row of its window the rest -- the pooled row (224 to 448 items of nine
// dependent shared-memory loads and a max tree each) was the serial stage of the kernel; with 256 threads it is one
// item per thread for 32 channels and under two for 64.
static constexpr int SP_THREADS = 384 + 256;     // upper bound; a variant runs 384 + StemCfg::POOL_THREADS

template <int KS, int NOUT, int ACT>
__global__ void __launch_bounds__(SP_THREADS, 1)
stem_pool_kernel(const __grid_constant__ CUtensorMap in_map, const __grid_constant__ StemPoolParams p) {
  using G = StemCfg<KS, NOUT>;
  constexpr int SR_STAGE_BYTES = G::STAGE_BYTES, SR_W_BYTES = G::W_BYTES, SR_LEAD_BYTES = G::LEAD_BYTES;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[G::STAGES], bar_empty[G::STAGES], bar_t_full[G::POOL_ACCS], bar_t_empty[G::POOL_ACCS];
  __shared__ __align__(8) uint64_t bar_row_ready[G::POOL_RING], bar_row_free[G::POOL_RING];
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float bias_s[64];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t ring = base;
  const uint32_t wsm = ring + G::STAGES * SR_STAGE_BYTES + 256;
  const uint32_t rows = (wsm + SR_W_BYTES + 1023u) & ~1023u;     // G::POOL_RING stem rows [128 px][128 B], chunk-swizzled
  const int per_cta = (p.total_units + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const int unit0 = static_cast<int>(blockIdx.x) * per_cta;
  const int my_units = max(0, min(per_cta, p.total_units - unit0));
  const int Hp = p.Ho >> 1, Wp = p.Wo >> 1;

  for (int i = threadIdx.x; i < SR_W_BYTES / 16; i += static_cast<int>(blockDim.x)) {
    const uint4 v = reinterpret_cast<const uint4*>(p.w)[i];
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(wsm + i * 16), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
  }
  if (threadIdx.x < NOUT) bias_s[threadIdx.x] = p.bias[threadIdx.x];
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&in_map);
    for (int s = 0; s < G::STAGES; ++s) {
      mbar_init(smem_u32(&bar_full[s]), 1);
      mbar_init(smem_u32(&bar_empty[s]), 1);
    }
    for (int b = 0; b < G::POOL_ACCS; ++b) {
      mbar_init(smem_u32(&bar_t_full[b]), 1);
      mbar_init(smem_u32(&bar_t_empty[b]), 128);
    }
    for (int r = 0; r < G::POOL_RING; ++r) {
      mbar_init(smem_u32(&bar_row_ready[r]), 128);
      mbar_init(smem_u32(&bar_row_free[r]), G::POOL_THREADS);
    }
    fence_barrier_init();
  } else if (warp == 1) {
    tmem_alloc<G::POOL_TMEM_COLS>(smem_u32(&tmem_base_s));
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  pdl_launch_dependents();
  pdl_wait();      // everything above overlapped the previous kernel's tail; activations are read from here on

  // stem rows of unit u: y in [y_lo, y_hi]
  auto unit_rows = [&](int u, int& n, int& ph0, int& ph1, int& y_lo, int& y_hi) {
    n = u / p.bands;
    ph0 = (u - n * p.bands) * SP_BAND;
    ph1 = min(ph0 + SP_BAND, Hp);                 // pooled rows [ph0, ph1)
    y_lo = max(0, 2 * ph0 - 1);
    y_hi = 2 * (ph1 - 1) + 1;
  };

  if (warp == 0 || warp == 3) {
    // two producer warps and two MMA-issuing warps, rows dealt alternately: a single warp's ~110 dependent instructions per
    // row (coordinates, descriptors, barriers) are ~760 cycles -- that loop, not the MMAs or the epilogue, paced the kernel
    const int mine = warp == 0 ? 0 : 1;
    {
      int it = 0;
      for (int u = unit0; u < unit0 + my_units; ++u) {
        int n, ph0, ph1, y_lo, y_hi;
        unit_rows(u, n, ph0, ph1, y_lo, y_hi);
        for (int y = y_lo; y <= y_hi; ++y, ++it) {
          if ((it & 1) != mine) continue;
          const int s = it % G::STAGES;
          mbar_wait(smem_u32(&bar_empty[s]), ((it / G::STAGES) & 1) ^ 1);
          const uint32_t full = smem_u32(&bar_full[s]);
          if (elect_one()) {
            mbar_expect_tx(full, G::KR * SR_ROW_BYTES);
            tma_load_4d(&in_map, full, ring + s * SR_STAGE_BYTES, 0, -1, 2 * y - G::PAD, n);
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1 || warp == 2) {
    const int mine = warp - 1;
    {
      constexpr uint32_t idesc = umma_idesc_bf16_f32(128, NOUT);
      int it = 0;
      for (int u = unit0; u < unit0 + my_units; ++u) {
        int n, ph0, ph1, y_lo, y_hi;
        unit_rows(u, n, ph0, ph1, y_lo, y_hi);
        for (int y = y_lo; y <= y_hi; ++y, ++it) {
          if ((it & 1) != mine) continue;
          const int b = it & (G::POOL_ACCS - 1), s = it % G::STAGES;
          mbar_wait(smem_u32(&bar_t_empty[b]), ((it / G::POOL_ACCS) & 1) ^ 1);
          mbar_wait(smem_u32(&bar_full[s]), (it / G::STAGES) & 1);
          tc_fence_after();
          const uint32_t a0 = ring + s * SR_STAGE_BYTES;
          if (elect_one()) {
#pragma unroll
            for (int r = 0; r < G::KR; ++r) {
#pragma unroll
              for (int kk = 0; kk < G::MMAS; ++kk) {
                const uint64_t adesc = umma_desc_noswz(a0 + r * SR_ROW_BYTES + SR_LEAD_BYTES + kk * 32, 16, 128);
                const uint64_t bdesc = umma_desc_noswz(wsm + (r * G::KC + kk * 2) * (NOUT * 16), NOUT * 16, 128);
                umma_bf16(tmem_base + b * NOUT, adesc, bdesc, idesc, (r | kk) != 0 ? 1u : 0u);
              }
            }
            umma_commit(smem_u32(&bar_empty[s]));
            umma_commit(smem_u32(&bar_t_full[b]));
          }
        }
      }
    }
    __syncwarp();
  } else if (warp >= 4 && warp < 12) {
    // =========================== accumulator drain: TMEM -> +bias -> ReLU -> bf16 -> row ring ===========================
    // Two warpgroups, one per accumulator: group g drains every second stem row, so the drain (~1000 cycles per row:
    // two dependent TMEM loads, bias, ReLU, pack, 8 stores) keeps up with the ~450-cycle MMA of a row.
    const int dwg = (warp - 4) >> 2;
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;     // stem pixel (ow) owned by this thread
    const int sw = row & 7;
    int it = 0;                               // stem rows processed by this CTA (ring slot = it % G::POOL_RING)
    for (int u = unit0; u < unit0 + my_units; ++u) {
      int n, ph0, ph1, y_lo, y_hi;
      unit_rows(u, n, ph0, ph1, y_lo, y_hi);
      for (int y = y_lo; y <= y_hi; ++y, ++it) {
        const int b = it & (G::POOL_ACCS - 1), slot = it % G::POOL_RING;
        if ((b & 1) != dwg) continue;
        mbar_wait(smem_u32(&bar_row_free[slot]), ((it / G::POOL_RING) & 1) ^ 1);   // pooling is done with the row that lived here
        mbar_wait(smem_u32(&bar_t_full[b]), (it / G::POOL_ACCS) & 1);
        tc_fence_after();
        const uint32_t dst_row = rows + slot * SR_OUT_BYTES + row * 128;
        const uint32_t acc = tmem_base + b * NOUT + (static_cast<uint32_t>(quarter * 32) << 16);
#pragma unroll
        for (int cc = 0; cc < NOUT / 32; ++cc) {
          uint32_t v[32];
          tmem_ld_32x32(acc + cc * 32, v);
          tmem_ld_wait();
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const int chunk = cc * 4 + g;
            const float4 b0 = *reinterpret_cast<const float4*>(&bias_s[chunk * 8]);
            const float4 b1 = *reinterpret_cast<const float4*>(&bias_s[chunk * 8 + 4]);
            float f[8];
            f[0] = apply_act(__uint_as_float(v[g * 8 + 0]) + b0.x, ACT); f[1] = apply_act(__uint_as_float(v[g * 8 + 1]) + b0.y, ACT);
            f[2] = apply_act(__uint_as_float(v[g * 8 + 2]) + b0.z, ACT); f[3] = apply_act(__uint_as_float(v[g * 8 + 3]) + b0.w, ACT);
            f[4] = apply_act(__uint_as_float(v[g * 8 + 4]) + b1.x, ACT); f[5] = apply_act(__uint_as_float(v[g * 8 + 5]) + b1.y, ACT);
            f[6] = apply_act(__uint_as_float(v[g * 8 + 6]) + b1.z, ACT); f[7] = apply_act(__uint_as_float(v[g * 8 + 7]) + b1.w, ACT);
            const uint32_t o0 = pack_bf16x2(f[0], f[1]), o1 = pack_bf16x2(f[2], f[3]);
            const uint32_t o2 = pack_bf16x2(f[4], f[5]), o3 = pack_bf16x2(f[6], f[7]);
            asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(dst_row + ((chunk ^ sw) << 4)), "r"(o0), "r"(o1),
                         "r"(o2), "r"(o3)
                         : "memory");
          }
        }
        tc_fence_before();
        mbar_arrive(smem_u32(&bar_t_empty[b]));
        mbar_arrive(smem_u32(&bar_row_ready[slot]));
      }
    }
  } else if (warp >= 12) {
    // =========================== pooling warps: 3x3 / stride 2 max over the row ring, coalesced stores ===========================
    const int ptid = threadIdx.x - 384;       // 0..G::POOL_THREADS-1
    int it = 0;
    for (int u = unit0; u < unit0 + my_units; ++u) {
      int n, ph0, ph1, y_lo, y_hi;
      unit_rows(u, n, ph0, ph1, y_lo, y_hi);
      for (int y = y_lo; y <= y_hi; ++y, ++it) {
        const int slot = it % G::POOL_RING;
        mbar_wait(smem_u32(&bar_row_ready[slot]), (it / G::POOL_RING) & 1);
        const bool pool = (y & 1) && (y >> 1) >= ph0;     // y = 2*ph + 1 completes pooled row ph (rows y-2, y-1, y)
        if (pool) {
          const int ph = y >> 1;
          uint16_t* orow = p.out + (static_cast<size_t>(n) * Hp + ph) * Wp * p.out_ps;
          const int CH = p.cout >> 3;               // 16-byte chunks per pooled pixel
          for (int item = ptid; item < Wp * CH; item += G::POOL_THREADS) {
            const int pw = fast_div(item, CH, p.mul_ch), chunk = item - pw * CH;
            // Branch-free window: a tap outside the image is replaced by its in-window neighbour (max is idempotent), so
            // the nine 16-byte loads are independent and issue back to back; with a branch per tap this loop was
            // latency-bound and the pooling warps were the kernel's bottleneck.  Only the left column (pw = 0) and the top
            // row of the image (y = 1) ever clamp: Wo and Ho are even.
            uint4 v[9];
#pragma unroll
            for (int dy = 0; dy < 3; ++dy) {
              const int back = (y - 2 + dy < y_lo) ? 1 : 2 - dy;      // rows y-2, y-1, y live in slots it-2, it-1, it
              const uint32_t rbase = rows + ((it - back) % G::POOL_RING) * SR_OUT_BYTES;
#pragma unroll
              for (int dx = 0; dx < 3; ++dx) {
                const int xx = max(2 * pw - 1 + dx, 0);
                asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                             : "=r"(v[dy * 3 + dx].x), "=r"(v[dy * 3 + dx].y), "=r"(v[dy * 3 + dx].z), "=r"(v[dy * 3 + dx].w)
                             : "r"(rbase + xx * 128 + ((chunk ^ (xx & 7)) << 4)));
              }
            }
            uint4 m = v[0];
#pragma unroll
            for (int t = 1; t < 9; ++t) {
              m.x = bf16x2_max_u32(m.x, v[t].x); m.y = bf16x2_max_u32(m.y, v[t].y);
              m.z = bf16x2_max_u32(m.z, v[t].z); m.w = bf16x2_max_u32(m.w, v[t].w);
            }
            *reinterpret_cast<uint4*>(orow + static_cast<size_t>(pw) * p.out_ps + chunk * 8) = m;
          }
          // rows y-2 (top, second and last use) and y-1 (middle, only use) are no longer needed
          if (y - 2 >= y_lo) mbar_arrive(smem_u32(&bar_row_free[(it - 2) % G::POOL_RING]));
          mbar_arrive(smem_u32(&bar_row_free[(it - 1) % G::POOL_RING]));
        }
        if (y == y_hi) mbar_arrive(smem_u32(&bar_row_free[slot]));   // last row of the band: nobody below will need it
      }
    }
  }

  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<G::POOL_TMEM_COLS>(tmem_base);
  }
}

// ----------------------------------------------------------------------------------------------------------------
// The same fused stem + MaxPool2d(3, 2, 1) with the pooling done in REGISTERS by the warps that drain the accumulators.
// An ncu capture of the kernel above (3x3 / 32-wide, ShuffleNetV2 @512) after its waits had been tuned away showed the
// two stages simply executing: 1,920 warp instructions per stem row -- ~870 to drain a row into the shared-memory ring
// (TMEM load, bias, ReLU, pack, swizzled stores, two barriers) and ~940 to pool it back out (nine 16-byte loads and a max
// tree per item, index arithmetic, three barriers) -- 64 % of the issue slots of the SM, for a kernel whose MMAs need 240
// cycles per row.  Here a thread owns stem pixel ow (its TMEM lane) for ALL rows and HALF = NOUT / 2 channels (warpgroup
// g drains channel half g): the vertical 3-max is a running bf16x2 max in registers, the horizontal one takes the two
// neighbour pixels from the neighbouring lanes by shuffle (the one neighbour that sits in another warp comes through a
// small shared-memory mailbox and one named barrier per pooled row), even pixels store their pooled pixel.  No row ring,
// no pooling warps: ~800 warp instructions per row.  Bit-identical with the unfused pair (max is exact in bf16).
// ----------------------------------------------------------------------------------------------------------------
static constexpr int SPR_THREADS = 384;   // warps 0: TMA, 1: MMA, 4-7: channel half 0, 8-11: channel half 1

template <int KS, int NOUT, int ACT>
__global__ void __launch_bounds__(SPR_THREADS, 1)
stem_pool_reg_kernel(const __grid_constant__ CUtensorMap in_map, const __grid_constant__ StemPoolParams p) {
  using G = StemCfg<KS, NOUT>;
  constexpr int SR_STAGE_BYTES = G::STAGE_BYTES, SR_W_BYTES = G::W_BYTES, SR_LEAD_BYTES = G::LEAD_BYTES;
  constexpr int HALF = NOUT / 2;            // channels per drain thread
  constexpr int HP = HALF / 2;              // packed bf16x2 registers per thread
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[G::STAGES], bar_empty[G::STAGES], bar_t_full[G::POOL_ACCS], bar_t_empty[G::POOL_ACCS];
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float bias_s[64];
  __shared__ __align__(16) uint32_t xch[2][2][4][HP];   // [mailbox copy][channel half][warp quarter]: lane 31's running max

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t ring = base;
  const uint32_t wsm = ring + G::STAGES * SR_STAGE_BYTES + 256;
  const int per_cta = (p.total_units + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const int unit0 = static_cast<int>(blockIdx.x) * per_cta;
  const int my_units = max(0, min(per_cta, p.total_units - unit0));
  const int Hp = p.Ho >> 1, Wp = p.Wo >> 1;

  for (int i = threadIdx.x; i < SR_W_BYTES / 16; i += SPR_THREADS) {
    const uint4 v = reinterpret_cast<const uint4*>(p.w)[i];
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(wsm + i * 16), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
  }
  if (threadIdx.x < NOUT) bias_s[threadIdx.x] = p.bias[threadIdx.x];
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&in_map);
    for (int s = 0; s < G::STAGES; ++s) {
      mbar_init(smem_u32(&bar_full[s]), 1);
      mbar_init(smem_u32(&bar_empty[s]), 1);
    }
    for (int b = 0; b < G::POOL_ACCS; ++b) {
      mbar_init(smem_u32(&bar_t_full[b]), 1);
      mbar_init(smem_u32(&bar_t_empty[b]), 256);      // both channel halves have read the accumulator
    }
    fence_barrier_init();
  } else if (warp == 1) {
    tmem_alloc<G::POOL_TMEM_COLS>(smem_u32(&tmem_base_s));
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  pdl_launch_dependents();
  pdl_wait();

  auto unit_rows = [&](int u, int& n, int& ph0, int& ph1, int& y_lo, int& y_hi) {
    n = u / p.bands;
    ph0 = (u - n * p.bands) * SP_BAND;
    ph1 = min(ph0 + SP_BAND, Hp);                 // pooled rows [ph0, ph1)
    y_lo = max(0, 2 * ph0 - 1);
    y_hi = 2 * (ph1 - 1) + 1;
  };

  if (warp == 0 || warp == 3) {
    // two producer warps and two MMA-issuing warps, rows dealt alternately: a single warp's ~110 dependent instructions per
    // row (coordinates, descriptors, barriers) are ~760 cycles -- that loop, not the MMAs or the epilogue, paced the kernel
    const int mine = warp == 0 ? 0 : 1;
    int it = 0;
    for (int u = unit0; u < unit0 + my_units; ++u) {
      int n, ph0, ph1, y_lo, y_hi;
      unit_rows(u, n, ph0, ph1, y_lo, y_hi);
      for (int y = y_lo; y <= y_hi; ++y, ++it) {
        if ((it & 1) != mine) continue;
        const int s = it % G::STAGES;
        mbar_wait(smem_u32(&bar_empty[s]), ((it / G::STAGES) & 1) ^ 1);
        const uint32_t full = smem_u32(&bar_full[s]);
        if (elect_one()) {
          mbar_expect_tx(full, G::KR * SR_ROW_BYTES);
          tma_load_4d(&in_map, full, ring + s * SR_STAGE_BYTES, 0, -1, 2 * y - G::PAD, n);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1 || warp == 2) {
    const int mine = warp - 1;
    constexpr uint32_t idesc = umma_idesc_bf16_f32(128, NOUT);
    int it = 0;
    for (int u = unit0; u < unit0 + my_units; ++u) {
      int n, ph0, ph1, y_lo, y_hi;
      unit_rows(u, n, ph0, ph1, y_lo, y_hi);
      for (int y = y_lo; y <= y_hi; ++y, ++it) {
        if ((it & 1) != mine) continue;
          const int b = it & (G::POOL_ACCS - 1), s = it % G::STAGES;
        mbar_wait(smem_u32(&bar_t_empty[b]), ((it / G::POOL_ACCS) & 1) ^ 1);
        mbar_wait(smem_u32(&bar_full[s]), (it / G::STAGES) & 1);
        tc_fence_after();
        const uint32_t a0 = ring + s * SR_STAGE_BYTES;
        if (elect_one()) {
#pragma unroll
          for (int r = 0; r < G::KR; ++r) {
#pragma unroll
            for (int kk = 0; kk < G::MMAS; ++kk) {
              const uint64_t adesc = umma_desc_noswz(a0 + r * SR_ROW_BYTES + SR_LEAD_BYTES + kk * 32, 16, 128);
              const uint64_t bdesc = umma_desc_noswz(wsm + (r * G::KC + kk * 2) * (NOUT * 16), NOUT * 16, 128);
              umma_bf16(tmem_base + b * NOUT, adesc, bdesc, idesc, (r | kk) != 0 ? 1u : 0u);
            }
          }
          umma_commit(smem_u32(&bar_empty[s]));
          umma_commit(smem_u32(&bar_t_full[b]));
        }
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    const int g = (warp - 4) >> 2;            // channel half
    const int quarter = warp & 3;
    const int ow = quarter * 32 + lane;       // stem pixel owned by this thread
    const bool emits = (ow & 1) == 0 && ow < p.Wo;
    const int nchunk = max(0, min(HALF / 8, (p.cout - g * HALF) / 8));   // 16-byte chunks of this half that are stored
    const uint32_t lane_base = tmem_base + g * HALF + (static_cast<uint32_t>(quarter * 32) << 16);
    // The rows of this CTA as one sequence, so that the TMEM load of row it + 1 is issued before the arithmetic of row it:
    // every drain warp sees every row, and with the wait for the MMA, the load and the pooling in one dependent chain a
    // row took ~760 cycles whatever the instruction count (first version of this kernel: no faster than the row ring).
    struct RowIt { int u, n, ph0, y_lo, y_hi, y; bool valid; };
    auto open_unit = [&](RowIt& r) {
      r.valid = r.u < unit0 + my_units;
      if (r.valid) {
        int ph1;
        unit_rows(r.u, r.n, r.ph0, ph1, r.y_lo, r.y_hi);
        r.y = r.y_lo;
      }
    };
    auto advance = [&](RowIt& r) {
      if (r.y < r.y_hi) { ++r.y; return; }
      ++r.u;
      open_unit(r);
    };
    auto issue_load = [&](int it, uint32_t (&v)[HALF]) {
      const int b = it & (G::POOL_ACCS - 1);
      mbar_wait(smem_u32(&bar_t_full[b]), (it / G::POOL_ACCS) & 1);
      tc_fence_after();
      if constexpr (HALF == 32) tmem_ld_32x32(lane_base + b * NOUT, v); else tmem_ld_32x16(lane_base + b * NOUT, v);
    };
    uint32_t vm[HP];
    int emitted = 0;                          // alternates the mailbox: a writer is never one barrier ahead of a reader of the same copy
    // one row: `v` holds its accumulator (load already issued), `vnext` receives the next row's
    auto row = [&](int it, RowIt& cur, uint32_t (&v)[HALF], uint32_t (&vnext)[HALF]) {
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(smem_u32(&bar_t_empty[it & (G::POOL_ACCS - 1)]));   // the accumulator is free as soon as it is in registers
      const RowIt me = cur;
      advance(cur);
      if (cur.valid) issue_load(it + 1, vnext);
      uint32_t pk[HP];
#pragma unroll
      for (int j = 0; j < HP; ++j)
        pk[j] = pack_bf16x2(apply_act(__uint_as_float(v[2 * j]) + bias_s[g * HALF + 2 * j], ACT),
                            apply_act(__uint_as_float(v[2 * j + 1]) + bias_s[g * HALF + 2 * j + 1], ACT));
      if (me.y == me.y_lo) {
#pragma unroll
        for (int j = 0; j < HP; ++j) vm[j] = pk[j];
      } else {
#pragma unroll
        for (int j = 0; j < HP; ++j) vm[j] = bf16x2_max_u32(vm[j], pk[j]);
      }
      if ((me.y & 1) && (me.y >> 1) >= me.ph0) {       // row 2*ph + 1 completes pooled row ph (rows y-2, y-1, y are in vm)
        const int ph = me.y >> 1;
        const int par = emitted++ & 1;
        uint32_t* box = xch[par][g][quarter];
        if (lane == 31) {
#pragma unroll
          for (int j = 0; j < HP; ++j) box[j] = vm[j];
        }
        named_bar_sync(1 + g, 128);
        uint32_t m[HP];
#pragma unroll
        for (int j = 0; j < HP; ++j) {
          uint32_t left = __shfl_up_sync(0xffffffffu, vm[j], 1);
          const uint32_t right = __shfl_down_sync(0xffffffffu, vm[j], 1);
          // pixel ow - 1 of lane 0 lives in the previous warp (mailbox); pixel -1 is padding: max is idempotent, use own
          if (lane == 0) left = quarter > 0 ? xch[par][g][quarter - 1][j] : vm[j];
          m[j] = bf16x2_max_u32(bf16x2_max_u32(left, vm[j]), right);
        }
        if (emits) {
          uint16_t* dst = p.out + ((static_cast<size_t>(me.n) * Hp + ph) * Wp + (ow >> 1)) * p.out_ps + g * HALF;
#pragma unroll
          for (int c = 0; c < HALF / 8; ++c)
            if (c < nchunk) *reinterpret_cast<uint4*>(dst + c * 8) = make_uint4(m[4 * c], m[4 * c + 1], m[4 * c + 2], m[4 * c + 3]);
        }
#pragma unroll
        for (int j = 0; j < HP; ++j) vm[j] = pk[j];   // row y is also the top row of the next pooled row's window
      }
    };
    RowIt cur;
    cur.u = unit0;
    open_unit(cur);
    uint32_t va[HALF], vb[HALF];
    if (cur.valid) issue_load(0, va);
    for (int it = 0; cur.valid; it += 2) {
      row(it, cur, va, vb);
      if (!cur.valid) break;
      row(it + 1, cur, vb, va);
    }
  }

  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<G::POOL_TMEM_COLS>(tmem_base);
  }
}

// ---------------------------------------------------------------------------------------------------------------- host
static int stem_nout(int cout) { return cout <= 32 ? 32 : 64; }

int stem_row_weight_bytes_k(int ksize, int cout) {
  if ((ksize != 7 && ksize != 3) || cout <= 0 || cout > 64) return 0;
  const int kc = ksize == 7 ? 4 : 2;
  return ksize * kc * stem_nout(cout) * 16;
}
int stem_row_weight_bytes() { return stem_row_weight_bytes_k(7, 64); }

static int stem_check(const char* who, const View& in, const View& out, int ksize, int act, int Ho, int Wo) {
  if (in.C != 4 || in.pix_stride != 4 || in.W % 16)
    return fail(B200_ERR_INVALID, std::string(who) + ": input must be NHWC4 with width a multiple of 16");
  if (ksize != 7 && ksize != 3) return fail(B200_ERR_INVALID, std::string(who) + ": kernel must be 7x7 or 3x3 (stride 2, pad k/2)");
  if (out.C % 8 || out.C > 64 || out.pix_stride % 8) return fail(B200_ERR_INVALID, std::string(who) + ": cout must be a multiple of 8, at most 64");
  if (act != ACT_NONE && act != ACT_RELU && act != ACT_SILU) return fail(B200_ERR_INVALID, std::string(who) + ": activation must be none, relu or silu");
  if ((in.H + 2 * (ksize / 2) - ksize) / 2 + 1 != Ho || (in.W + 2 * (ksize / 2) - ksize) / 2 + 1 != Wo)
    return fail(B200_ERR_INVALID, std::string(who) + ": output geometry mismatch");
  return B200_OK;
}

static int stem_in_map(CUtensorMap* m, const View& in, int krows) {
  const uint64_t dims[4] = {64, uint64_t(in.W / 16), uint64_t(in.H), uint64_t(in.N)};
  const uint64_t str[3] = {128, uint64_t(in.W) * 8, uint64_t(in.W) * 8 * in.H};
  const uint32_t box[4] = {64, uint32_t(SR_GROUPS), uint32_t(krows), 1};
  return encode_tmap(m, in.base, 4, dims, str, box, SWZ_NONE);
}

template <int KS, int NOUT, int ACT>
static int stem_pool_reg_launch_t(const CUtensorMap& in_map, const StemPoolParams& p, cudaStream_t stream) {
  using G = StemCfg<KS, NOUT>;
  const size_t smem = size_t(G::STAGES) * G::STAGE_BYTES + 256 + G::W_BYTES + 2048;
  static bool configured[64] = {};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(stem_pool_reg_kernel<KS, NOUT, ACT>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    configured[dev & 63] = true;
  }
  const int ctas = p.total_units < device_sm_count() ? p.total_units : device_sm_count();
  B200_CUDA_OK(launch_pdl_if(p.total_units >= pdl_min_waves() * device_sm_count(), stem_pool_reg_kernel<KS, NOUT, ACT>, dim3(ctas), dim3(SPR_THREADS), smem, stream, in_map, p));
  return B200_OK;
}

template <int KS, int NOUT, int ACT>
static int stem_pool_launch_t(const CUtensorMap& in_map, const StemPoolParams& p, cudaStream_t stream) {
  using G = StemCfg<KS, NOUT>;
  const size_t smem = size_t(G::STAGES) * G::STAGE_BYTES + 256 + G::W_BYTES + 1024 + G::POOL_RING * SR_OUT_BYTES + 1024;
  static bool configured[64] = {};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(stem_pool_kernel<KS, NOUT, ACT>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    configured[dev & 63] = true;
  }
  const int ctas = p.total_units < device_sm_count() ? p.total_units : device_sm_count();
  B200_CUDA_OK(launch_pdl_if(p.total_units >= pdl_min_waves() * device_sm_count(), stem_pool_kernel<KS, NOUT, ACT>, dim3(ctas), dim3(384 + G::POOL_THREADS), smem, stream, in_map, p));
  return B200_OK;
}

template <int KS, int NOUT, int ACT>
static int stem_row_launch_t(const CUtensorMap& in_map, const CUtensorMap& out_map, const StemRowParams& p, cudaStream_t stream) {
  using G = StemCfg<KS, NOUT>;
  const size_t smem = size_t(G::STAGES) * G::STAGE_BYTES + 256 + G::W_BYTES + 1024 + SR_WGS * SR_OUT_BYTES + 1024;
  static bool configured[64] = {};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(stem_row_kernel<KS, NOUT, ACT>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    configured[dev & 63] = true;
  }
  const int ctas = p.total_tiles < device_sm_count() ? p.total_tiles : device_sm_count();
  B200_CUDA_OK(launch_pdl_if(p.total_tiles >= pdl_min_waves() * device_sm_count(), stem_row_kernel<KS, NOUT, ACT>, dim3(ctas), dim3(SR2_THREADS), smem, stream, in_map, out_map, p));
  return B200_OK;
}

// the served stems: 7x7 -> 64 ReLU (ResNet), 3x3 -> <= 32 with ReLU (ShuffleNetV2) or SiLU (EfficientNetV2, YOLOv8)
#define B200_STEM_DISPATCH(FN, ...)                                                                       \
  do {                                                                                                    \
    if (ksize == 7 && nout == 64 && act == ACT_RELU) return FN<7, 64, ACT_RELU>(__VA_ARGS__);             \
    if (ksize == 3 && nout == 32 && act == ACT_RELU) return FN<3, 32, ACT_RELU>(__VA_ARGS__);             \
    if (ksize == 3 && nout == 32 && act == ACT_SILU) return FN<3, 32, ACT_SILU>(__VA_ARGS__);             \
    if (ksize == 3 && nout == 64 && act == ACT_SILU) return FN<3, 64, ACT_SILU>(__VA_ARGS__);             \
    if (ksize == 3 && nout == 64 && act == ACT_RELU) return FN<3, 64, ACT_RELU>(__VA_ARGS__);             \
    return fail(B200_ERR_INVALID, "stem: this (kernel, width, activation) combination is not instantiated"); \
  } while (0)

// Fused stem conv + bn + act + maxpool(3,2,1).  `out` is the pooled tensor [N][Ho/2][Wo/2][cout].
int stem_pool_launch_k(const View& in, const View& out, const void* w, const float* bias, int ksize, int act,
                       cudaStream_t stream) {
  const int Ho = (in.H + 2 * (ksize / 2) - ksize) / 2 + 1, Wo = (in.W + 2 * (ksize / 2) - ksize) / 2 + 1;
  B200_TRY(stem_check("stem_pool", in, out, ksize, act, Ho, Wo));
  if (Wo > 128 || (Ho & 1) || (Wo & 1)) return fail(B200_ERR_INVALID, "stem_pool: needs stem width <= 128 and even stem size");
  if (out.H != Ho / 2 || out.W != Wo / 2 || out.N != in.N) return fail(B200_ERR_INVALID, "stem_pool: output geometry mismatch");
  CUtensorMap in_map;
  B200_TRY(stem_in_map(&in_map, in, ksize));
  StemPoolParams p;
  p.w = static_cast<const uint16_t*>(w);
  p.bias = bias;
  p.out = static_cast<uint16_t*>(out.base);
  p.out_ps = out.pix_stride;
  p.cout = out.C;
  p.Ho = Ho; p.Wo = Wo;
  p.bands = (Ho / 2 + SP_BAND - 1) / SP_BAND;
  p.total_units = in.N * p.bands;
  p.mul_ch = fast_div_mul(p.cout >> 3);
  const int nout = stem_nout(out.C);
  static int reg_pool = -1;
  if (reg_pool < 0) {
    // B200_STEM_POOL_REG=1: the register-pooling kernel.  Measured on the same box (ShuffleNetV2 @512 / ResNet-50 @256):
    // 1.511 / 3.133 ms against 1.489 / 3.096 ms for the row ring -- it executes fewer instructions (91 M vs 118 M) but every
    // drain warp walks every row, ~190 dependent instructions per row per warp, and a row takes as long as that chain;
    // in the ring kernel two drain groups and the pooling groups work on different rows at once.  Kept as an opt-in.
    const char* e = getenv("B200_STEM_POOL_REG");
    reg_pool = (e && e[0] == '1') ? 1 : 0;
  }
  if (reg_pool && out.pix_stride % 8 == 0) B200_STEM_DISPATCH(stem_pool_reg_launch_t, in_map, p, stream);
  B200_STEM_DISPATCH(stem_pool_launch_t, in_map, p, stream);
}

int stem_pool_launch(const View& in, const View& out, const void* w, const float* bias, cudaStream_t stream) {
  if (out.C != 64) return fail(B200_ERR_INVALID, "stem_pool: cout must be 64");
  return stem_pool_launch_k(in, out, w, bias, 7, ACT_RELU, stream);
}

// weights: bf16 smem image [kernel row][k chunk][NOUT][8]: element e of chunk k = window pixel 2k + e/4 (= kernel
// column + 1), channel e%4; zero for window pixel 0, channel 3 and rows >= cout.  bias fp32 [NOUT].
int stem_row_launch_k(const View& in, const View& out, const void* w, const float* bias, int ksize, int act,
                      cudaStream_t stream) {
  B200_TRY(stem_check("stem_row", in, out, ksize, act, out.H, out.W));
  if (in.N != out.N) return fail(B200_ERR_INVALID, "stem_row: output geometry mismatch");
  CUtensorMap in_map, out_map;
  B200_TRY(stem_in_map(&in_map, in, ksize));
  {
    const uint64_t ops = uint64_t(out.pix_stride) * 2;
    const uint64_t dims[3] = {uint64_t(out.C), uint64_t(out.W), uint64_t(out.N) * out.H};   // channels past cout are clipped
    const uint64_t str[2] = {ops, ops * out.W};
    const uint32_t box[3] = {64, 128, 1};
    B200_TRY(encode_tmap(&out_map, out.base, 3, dims, str, box, SWZ_128B));
  }
  StemRowParams p;
  p.w = static_cast<const uint16_t*>(w);
  p.bias = bias;
  p.Ho = out.H; p.Wo = out.W;
  p.segs = (out.W + 127) / 128;
  p.total_tiles = out.N * out.H * p.segs;
  p.mul_segs = fast_div_mul(p.segs);
  p.mul_Ho = fast_div_mul(p.Ho);
  const int nout = stem_nout(out.C);
  B200_STEM_DISPATCH(stem_row_launch_t, in_map, out_map, p, stream);
}

int stem_row_launch(const View& in, const View& out, const void* w, const float* bias, cudaStream_t stream) {
  if (out.C != 64) return fail(B200_ERR_INVALID, "stem_row: cout must be 64");
  return stem_row_launch_k(in, out, w, bias, 7, ACT_RELU, stream);
}

}  // namespace b200
