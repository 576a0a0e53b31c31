// Implicit-GEMM convolution for sm_100a: D[pixels, Cout] = act(sum_taps A_tap[pixels, Cin] * W_tap[Cout, Cin]^T + bias (+res)).
//
// Replaces (as part of the model forward) what the reference reaches through `model(inputs)`
// (293-project/src/scheduler.py:449-450 -> torchvision Conv2d+BatchNorm2d+ReLU/SiLU, e.g. tv models/resnet.py:143-161).
//
// Structure (one 128-pixel x BLOCK_N-channel tile per CTA, 2 CTAs resident per SM):
//   warp 0  : TMA producer. Per k-block one 4-D box of the NHWC input (64 channels x TW x TH x TN pixels, shifted by the
//             filter tap; out-of-image coordinates are zero-filled by TMA = the conv padding) and one box of weights.
//   warp 1  : allocates TMEM, single thread issues tcgen05.mma (M=128, N=BLOCK_N, K=16) into a fp32 TMEM accumulator.
//   warps 2-5: epilogue. tcgen05.ld -> +bias (+residual) -> ReLU/SiLU -> bf16 -> swizzled smem -> TMA store (NHWC).
// Stride-2 convolutions read four parity views of the input (strided tensor maps), so every tap is a plain box load.
#include "conv_gemm.cuh"
#include "ptx.cuh"

#include <stdlib.h>

namespace b200 {

static constexpr int BLOCK_M = 128;
static constexpr int BLOCK_K = 64;

struct ConvParams;
// k-block kb of a tile -> (tap / input index t, channel offset c0 in that input, weight coordinates (cw, tw)).
// A K-concatenated 1x1 conv (kb_split > 0) reads input 0 for the first kb_split blocks and input 1 after them, against
// one weight row [Ca + Cb]; a plain conv walks its taps.
__device__ __forceinline__ void kb_source(const ConvParams& p, int kb, int& t, int& c0, int& cw, int& tw);
static constexpr int A_STAGE_BYTES = BLOCK_M * BLOCK_K * 2;  // 16 KB
static constexpr int CONV_THREADS = 192;

template <int BLOCK_N>
struct ConvCfg {
  static constexpr int STAGES = BLOCK_N == 64 ? 4 : (BLOCK_N == 128 ? 3 : 4);
  static constexpr int B_STAGE_BYTES = BLOCK_N * BLOCK_K * 2;
  static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024;  // +1024: manual 1 KB alignment of the ring
  static_assert(BLOCK_N * BLOCK_M * 2 <= STAGES * STAGE_BYTES, "output staging must fit in the (drained) ring");
};

template <int BLOCK_N>
__global__ void __launch_bounds__(CONV_THREADS)
conv_gemm_kernel(const __grid_constant__ ConvMaps maps, const __grid_constant__ ConvParams p) {
  using Cfg = ConvCfg<BLOCK_N>;
  constexpr int STAGES = Cfg::STAGES;

  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[STAGES];
  __shared__ __align__(8) uint64_t bar_empty[STAGES];
  __shared__ __align__(8) uint64_t bar_tmem_full;
  __shared__ __align__(8) uint64_t bar_res_full;
  __shared__ uint32_t tmem_base_s;
  __shared__ float bias_s[BLOCK_N];

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t ring = (smem_u32(smem_raw) + 1023u) & ~1023u;

  // ---- tile coordinates ----
  const int n_tile = blockIdx.x % p.n_tiles;
  int m_tile = blockIdx.x / p.n_tiles;
  const int t_w = m_tile % p.tiles_w;
  m_tile /= p.tiles_w;
  const int t_h = m_tile % p.tiles_h;
  const int t_n = m_tile / p.tiles_h;
  const int w0 = t_w * p.tw, h0 = t_h * p.th, nb0 = t_n * p.tn;
  const int n0 = n_tile * BLOCK_N;
  const int num_kb = p.num_kb;

  // ---- one-time setup ----
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.a[0]);
    tma_prefetch_desc(&maps.b);
    tma_prefetch_desc(&maps.d);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(smem_u32(&bar_full[s]), 1);
      mbar_init(smem_u32(&bar_empty[s]), 1);
    }
    mbar_init(smem_u32(&bar_tmem_full), 1);
    mbar_init(smem_u32(&bar_res_full), 1);
    fence_barrier_init();
  } else if (warp == 1) {
    tmem_alloc<BLOCK_N>(smem_u32(&tmem_base_s));
  } else if (warp >= 2) {
    for (int i = threadIdx.x - 64; i < BLOCK_N; i += 128) {
      const int c = n0 + i;
      bias_s[i] = (c < p.cout) ? p.bias[c] : 0.f;
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  pdl_launch_dependents();
  pdl_wait();      // everything above overlapped the previous kernel's tail; activations are read from here on

  if (warp == 0) {
    // =========================== TMA producer ===========================
    if (lane == 0) {
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % STAGES;
        const uint32_t ph = (kb / STAGES) & 1;
        mbar_wait(smem_u32(&bar_empty[s]), ph ^ 1);
        const uint32_t full = smem_u32(&bar_full[s]);
        mbar_expect_tx(full, Cfg::STAGE_BYTES);
        int t, c0, cw, tw_;
        kb_source(p, kb, t, c0, cw, tw_);
        const uint32_t a_dst = ring + s * Cfg::STAGE_BYTES;
        tma_load_4d(&maps.a[p.tap_map[t]], full, a_dst, c0, w0 + p.tap_dx[t], h0 + p.tap_dy[t], nb0);
        tma_load_3d(&maps.b, full, a_dst + A_STAGE_BYTES, cw, tw_, n0);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16_f32(BLOCK_M, BLOCK_N);
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % STAGES;
        const uint32_t ph = (kb / STAGES) & 1;
        mbar_wait(smem_u32(&bar_full[s]), ph);
        tc_fence_after();
        const uint32_t a_src = ring + s * Cfg::STAGE_BYTES;
        const uint64_t adesc = umma_desc_sw128(a_src);
        const uint64_t bdesc = umma_desc_sw128(a_src + A_STAGE_BYTES);
#pragma unroll
        for (int k = 0; k < BLOCK_K / 16; ++k) {
          // advancing 16 bf16 (= 32 B) inside the 128 B swizzle row: +2 in the (addr >> 4) field
          umma_bf16(tmem_base, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
        }
        umma_commit(smem_u32(&bar_empty[s]));  // frees the smem slot once these MMAs have read it
      }
      umma_commit(smem_u32(&bar_tmem_full));  // accumulator complete
    }
    __syncwarp();
  } else {
    // =========================== epilogue ===========================
    const int epi_tid = threadIdx.x - 64;     // 0..127
    const int quarter = warp & 3;             // TMEM lane quarter this warp may access
    const int row = quarter * 32 + lane;      // tile row (pixel) owned by this thread
    const uint32_t stage_d = ring;            // output staging aliases the drained ring
    mbar_wait(smem_u32(&bar_tmem_full), 0);
    tc_fence_after();
    if (p.has_res) {
      if (epi_tid == 0) {
        const uint32_t rb = smem_u32(&bar_res_full);
        mbar_expect_tx(rb, BLOCK_N * BLOCK_M * 2);
#pragma unroll
        for (int j = 0; j < BLOCK_N / 64; ++j)
          tma_load_4d(&maps.r, rb, stage_d + j * (BLOCK_M * 128), n0 + j * 64, w0, h0, nb0);
      }
      mbar_wait(smem_u32(&bar_res_full), 0);
    }
    const uint32_t row_off = row * 128;
    const int sw = row & 7;
#pragma unroll 1
    for (int cc = 0; cc < BLOCK_N / 32; ++cc) {
      uint32_t v[32];
      tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + cc * 32, v);
      tmem_ld_wait();
      const uint32_t sub = stage_d + (cc >> 1) * (BLOCK_M * 128) + row_off;
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        const int chunk = (cc & 1) * 4 + g;  // 16-byte chunk inside the 128 B row
        const uint32_t addr = sub + ((chunk ^ sw) << 4);
        float f[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = __uint_as_float(v[g * 8 + i]) + bias_s[cc * 32 + g * 8 + i];
        float rr[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (p.has_res) {
          uint32_t r0, r1, r2, r3;
          asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
          rr[0] = bf16_lo(r0); rr[1] = bf16_hi(r0); rr[2] = bf16_lo(r1); rr[3] = bf16_hi(r1);
          rr[4] = bf16_lo(r2); rr[5] = bf16_hi(r2); rr[6] = bf16_lo(r3); rr[7] = bf16_hi(r3);
        }
        if (p.has_res == 2) {
#pragma unroll
          for (int i = 0; i < 8; ++i) f[i] = apply_act(f[i], p.act) + rr[i];
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) f[i] = apply_act(f[i] + rr[i], p.act);
        }
        const uint32_t o0 = pack_bf16x2(f[0], f[1]), o1 = pack_bf16x2(f[2], f[3]);
        const uint32_t o2 = pack_bf16x2(f[4], f[5]), o3 = pack_bf16x2(f[6], f[7]);
        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(o0), "r"(o1), "r"(o2), "r"(o3)
                     : "memory");
      }
    }
    tc_fence_before();
    fence_proxy_async_smem();  // generic-proxy smem writes -> visible to the TMA store
    named_bar_sync(1, 128);
    if (epi_tid == 0) {
#pragma unroll
      for (int j = 0; j < BLOCK_N / 64; ++j) {
        if (n0 + j * 64 < p.cout) tma_store_4d(&maps.d, stage_d + j * (BLOCK_M * 128), n0 + j * 64, w0, h0, nb0);
      }
      tma_store_commit();
      tma_store_wait_all();
    }
  }

  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<BLOCK_N>(tmem_base);
  }
}

// ================================================================================================================
// Persistent variant: one CTA per SM loops over tiles; loads, MMAs, the epilogue and the stores of consecutive tiles
// overlap (smem ring continues across tiles, two TMEM accumulators, three output-staging buffers).
//   warp 0 : TMA producer (A/B ring)              warp 1 : MMA issuer (TMEM accumulator it & 1)
//   warp 2 : "DMA" thread: prefetches the residual tile into staging[it % 3] and issues the TMA store of tile it-1
//   warps 4-7, 8-11 : two epilogue warpgroups (TMEM -> +bias/+res/act -> bf16 -> staging, in place over the residual);
//                     group g drains accumulator g of every second tile, so thin-K layers, whose tile time IS the
//                     epilogue (~0.6 us with SiLU), overlap two of them
// ================================================================================================================
static constexpr int PCONV_THREADS = 384;
static constexpr int PCONV_THREADS_MAX = 512;   // + warps 12-15: A-operand scaler, launched only for layers with ConvSpec::a_scale
static constexpr int HALO_WGS = 4;          // epilogue warpgroups = accumulators in TMEM = tiles in their epilogue at once
static constexpr int HALO_THREADS = 128 + HALO_WGS * 128;   // warps 0-2: TMA / MMA / DMA, then one warpgroup per accumulator

// Without a residual the staging ring only has to cover the store of the previous tile (2 buffers); the shared memory that
// frees goes into a deeper A/B ring, which is what the HBM-bound layers need (bytes in flight per SM ~ latency x bandwidth).
// DEEP: long-K layers with few tiles per SM (EfficientNetV2 project convs: K = 960 / 1536, 2-3 tiles per SM) are paced by
// the ring depth (tile period = (TMA latency + MMA time) / stages), not by the store path: two staging buffers, five stages.
template <int BLOCK_N, int RES, int DEEP = 0>
struct PConvCfg {
  // DEEP == 2: thin-K layers (one or two k-blocks per tile: 1x1 convs with Cin <= 128 and no residual) are paced by the
  // store path -- measured on a 24 -> 64 1x1 conv: the DMA thread sat in cp.async.bulk.wait_group.read 80 % of the time and
  // the epilogue waited for a staging buffer -- so the shared memory goes into staging buffers instead of ring stages.
  // RES == 3 (ShuffleNetV2 unit tail, BLOCK_N = 64 only): OUT_BUFS counts the ring of the other branch's tiles (X1_BYTES
  // each, loaded OUT_BUFS tiles ahead); the interleaved output tiles of twice the width have two buffers of their own,
  // one per epilogue warpgroup.  These layers have one to four k-blocks, three stages do.
  static constexpr int STAGES = RES == 3 ? 3
                              : DEEP == 3 ? (BLOCK_N == 64 ? 6 : 4)
                              : DEEP == 2 ? (BLOCK_N == 64 ? 4 : 3)
                              : DEEP == 1 ? (BLOCK_N == 64 ? 8 : 5)
                                          : (RES ? (BLOCK_N == 64 ? 4 : 3) : (BLOCK_N == 64 ? 7 : 4));   // A/B ring
  static constexpr int OUT_BUFS = RES == 3 ? 4
                                : DEEP == 3 ? (BLOCK_N == 64 ? 5 : 3)
                                : DEEP == 2 ? (BLOCK_N == 64 ? 8 : 4)
                                : DEEP == 1 ? 2 : (RES ? (BLOCK_N == 64 ? 6 : 4) : 3);         // output / residual staging ring
  static constexpr int LOOKAHEAD = OUT_BUFS - 2;            // tiles the residual prefetch runs ahead of the epilogue
  static constexpr int B_STAGE_BYTES = BLOCK_N * BLOCK_K * 2;
  static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  // the other branch's tile: 64 channels (128-byte swizzled rows) + the next 8 (16-byte rows): its channels for a tile's
  // slots start at a channel that is not a multiple of 8 in the second half-group, and TMA coordinates must be
  static constexpr int X1_BYTES = RES == 3 ? BLOCK_M * (BLOCK_N + 8) * 2 : 0;
  static constexpr int IL_OUT_BYTES = BLOCK_M * BLOCK_N * 4;
  static constexpr int OUT_BYTES = RES == 3 ? X1_BYTES : BLOCK_M * BLOCK_N * 2;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + OUT_BUFS * OUT_BYTES + (RES == 3 ? 2 * IL_OUT_BYTES : 0) + 1024;
  static_assert(SMEM_BYTES <= 230 * 1024, "smem budget");
};

__device__ __forceinline__ void kb_source(const ConvParams& p, int kb, int& t, int& c0, int& cw, int& tw) {
  if (p.kb_split > 0) {
    t = kb >= p.kb_split ? 1 : 0;
    c0 = (kb - (t ? p.kb_split : 0)) * BLOCK_K;
    cw = kb * BLOCK_K;
    tw = 0;
  } else {
    t = fast_div(kb, p.kb_per_tap, p.mul_kb_per_tap);
    c0 = (kb - t * p.kb_per_tap) * BLOCK_K;
    cw = c0;
    tw = t;
  }
}

__device__ __forceinline__ void tile_coords(const ConvParams& p, int tile, int& n_tile, int& w0, int& h0, int& nb0) {
  const int m = fast_div(tile, p.n_tiles, p.mul_n_tiles);
  n_tile = tile - m * p.n_tiles;
  const int mw = fast_div(m, p.tiles_w, p.mul_tiles_w);
  const int t_w = m - mw * p.tiles_w;
  const int t_n = fast_div(mw, p.tiles_h, p.mul_tiles_h);
  const int t_h = mw - t_n * p.tiles_h;
  w0 = t_w * p.tw; h0 = t_h * p.th; nb0 = t_n * p.tn;
}

template <int ACT>
__device__ __forceinline__ float act_t(float x) {
  if (ACT == ACT_RELU) return fmaxf(x, 0.f);
  if (ACT == ACT_SILU) return silu_f(x);
  if (ACT == ACT_SIGMOID) return sigmoid_f(x);
  return x;
}

// RES: 0 no residual, 1 act(acc + bias + res), 2 act(acc + bias) + res.  ACT: b200::Act.
template <int BLOCK_N, int RES, int ACT, int DEEP = 0>
__global__ void __launch_bounds__(PCONV_THREADS_MAX, 1)
conv_gemm_persistent_kernel(const __grid_constant__ ConvMaps maps, const __grid_constant__ ConvParams p,
                            const int total_tiles) {
  using Cfg = PConvCfg<BLOCK_N, RES, DEEP>;
  constexpr int STAGES = Cfg::STAGES;
  constexpr int NB = Cfg::OUT_BUFS;
  constexpr int LA = Cfg::LOOKAHEAD;

  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[STAGES];
  __shared__ __align__(8) uint64_t bar_empty[STAGES];
  __shared__ __align__(8) uint64_t bar_scaled[STAGES];   // A tile multiplied by the per-image channel factors (p.a_scale)
  __shared__ __align__(8) uint64_t bar_tmem_full[2];
  __shared__ __align__(8) uint64_t bar_tmem_empty[2];
  __shared__ __align__(8) uint64_t bar_stage_ready[NB];  // staging buffer free (and residual landed, if any)
  __shared__ __align__(8) uint64_t bar_out_ready[NB];    // epilogue finished writing the staging buffer
  __shared__ __align__(8) uint64_t bar_out_free[2];      // RES == 3: the store of the output buffer's previous tile has read it
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float bias_s[2][BLOCK_N];   // one per epilogue warpgroup
  __shared__ __align__(16) uint32_t keep_s[2][BLOCK_N / 2];   // RES == 3: 0xFFFF per valid slot of the other branch (pairs of slots)

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t ring = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t staging = ring + STAGES * Cfg::STAGE_BYTES;
  const int num_kb = p.num_kb;
  const int my_tiles = (total_tiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.a[0]);
    tma_prefetch_desc(&maps.b);
    tma_prefetch_desc(&maps.d);
    if (RES) tma_prefetch_desc(&maps.r);
    if (RES == 3) tma_prefetch_desc(&maps.a[3]);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(smem_u32(&bar_full[s]), 1);
      mbar_init(smem_u32(&bar_empty[s]), 1);
      mbar_init(smem_u32(&bar_scaled[s]), 128);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(smem_u32(&bar_tmem_full[b]), 1);
      mbar_init(smem_u32(&bar_tmem_empty[b]), 128);
    }
    for (int b = 0; b < NB; ++b) {
      mbar_init(smem_u32(&bar_stage_ready[b]), 1);
      mbar_init(smem_u32(&bar_out_ready[b]), 128);
    }
    mbar_init(smem_u32(&bar_out_free[0]), 1);
    mbar_init(smem_u32(&bar_out_free[1]), 1);
    fence_barrier_init();
  } else if (warp == 1) {
    tmem_alloc<2 * BLOCK_N>(smem_u32(&tmem_base_s));
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  pdl_launch_dependents();
  pdl_wait();      // everything above overlapped the previous kernel's tail; activations are read from here on

  if (warp == 0) {
    // =========================== TMA producer ===========================
    // Every single-thread role below runs its loops with the WHOLE warp (uniform control flow) and issues under
    // elect_one(): see ptx.cuh.
    {
      uint32_t g = 0;  // k-block counter across all tiles of this CTA
      for (int it = 0; it < my_tiles; ++it) {
        const int tile0_ = blockIdx.x + it * gridDim.x;
        const int tile = (p.reverse & 1) ? total_tiles - 1 - tile0_ : tile0_;
        int n_tile, w0, h0, nb0;
        tile_coords(p, tile, n_tile, w0, h0, nb0);
        const int n0 = n_tile * BLOCK_N;
        for (int kb = 0; kb < num_kb; ++kb, ++g) {
          const int s = g % STAGES;
          const uint32_t ph = (g / STAGES) & 1;
          mbar_wait(smem_u32(&bar_empty[s]), ph ^ 1);
          const uint32_t full = smem_u32(&bar_full[s]);
          int t, c0, cw, tw_;
          kb_source(p, kb, t, c0, cw, tw_);
          const uint32_t a_dst = ring + s * Cfg::STAGE_BYTES;
          if (elect_one()) {
            mbar_expect_tx(full, Cfg::STAGE_BYTES);
            tma_load_4d(&maps.a[p.tap_map[t]], full, a_dst, c0, w0 + p.tap_dx[t], h0 + p.tap_dy[t], nb0);
            tma_load_3d(&maps.b, full, a_dst + A_STAGE_BYTES, cw, tw_, n0);
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    {
      constexpr uint32_t idesc = umma_idesc_bf16_f32(BLOCK_M, BLOCK_N);
      uint32_t g = 0;
      for (int it = 0; it < my_tiles; ++it) {
        const int b = it & 1;
        mbar_wait(smem_u32(&bar_tmem_empty[b]), ((it >> 1) & 1) ^ 1);  // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t acc = tmem_base + b * BLOCK_N;
        for (int kb = 0; kb < num_kb; ++kb, ++g) {
          const int s = g % STAGES;
          const uint32_t ph = (g / STAGES) & 1;
          mbar_wait(smem_u32(p.a_scale ? &bar_scaled[s] : &bar_full[s]), ph);
          tc_fence_after();
          const uint32_t a_src = ring + s * Cfg::STAGE_BYTES;
          const uint64_t adesc = umma_desc_sw128(a_src);
          const uint64_t bdesc = umma_desc_sw128(a_src + A_STAGE_BYTES);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BLOCK_K / 16; ++k)
              umma_bf16(acc, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
            umma_commit(smem_u32(&bar_empty[s]));
            if (kb == num_kb - 1) umma_commit(smem_u32(&bar_tmem_full[b]));
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 2) {
    // =========================== DMA: residual prefetch (LA tiles ahead) + output stores ===========================
    {
      // whole warp in the loops, single-thread work under elect_one() (always the same lane, so the bulk-group
      // bookkeeping of the TMA stores stays with one thread)
      auto acquire_staging = [&](int j) {  // make staging[j % NB] available to the epilogue of tile j
        const int tile0_ = blockIdx.x + j * gridDim.x;
        const int tile = (p.reverse & 1) ? total_tiles - 1 - tile0_ : tile0_;
        const uint32_t rdy = smem_u32(&bar_stage_ready[j % NB]);
        if (RES == 3) {
          // the other branch's channels for this tile's 64 slots: slot s holds channel s below il_group and channel
          // il_half + (s - il_group) above it.  TMA wants the channel coordinate on a 16-byte boundary, il_half is not:
          // the boxes start at the multiple of 8 below the first channel (64 + 8 channels), the epilogue shifts.
          int n_tile, w0, h0, nb0;
          tile_coords(p, tile, n_tile, w0, h0, nb0);
          if (elect_one()) {
            mbar_expect_tx(rdy, Cfg::X1_BYTES);
            const uint32_t dst = staging + (j % NB) * Cfg::OUT_BYTES;
            const int slot = n_tile * BLOCK_N;
            const int ch8 = (slot < p.il_group ? slot : p.il_half + (slot - p.il_group)) & ~7;
            tma_load_4d(&maps.r, rdy, dst, ch8, w0, h0, nb0);
            tma_load_4d(&maps.a[3], rdy, dst + BLOCK_M * BLOCK_N * 2, ch8 + BLOCK_N, w0, h0, nb0);
          }
        } else if (RES) {
          int n_tile, w0, h0, nb0;
          tile_coords(p, tile, n_tile, w0, h0, nb0);
          if (elect_one()) {
            mbar_expect_tx(rdy, Cfg::OUT_BYTES);
#pragma unroll
            for (int q = 0; q < BLOCK_N / 64; ++q)
              tma_load_4d(&maps.r, rdy, staging + (j % NB) * Cfg::OUT_BYTES + q * (BLOCK_M * 128), n_tile * BLOCK_N + q * 64,
                          w0, h0, nb0);
          }
        } else {
          if (elect_one()) mbar_arrive(rdy);
        }
      };
      if (RES == 3) {
        // The other branch's tiles run NB tiles ahead in a ring of their own: slot it % NB is free again once the epilogue
        // of tile it is through (bar_out_ready).  Output buffer it & 1 goes back to its warpgroup when the store has read it.
        for (int j = 0; j < NB && j < my_tiles; ++j) acquire_staging(j);
        for (int it = 0; it < my_tiles; ++it) {
          const int tile0_ = blockIdx.x + it * gridDim.x;
          const int tile = (p.reverse & 1) ? total_tiles - 1 - tile0_ : tile0_;
          int n_tile, w0, h0, nb0;
          tile_coords(p, tile, n_tile, w0, h0, nb0);
          const int n0 = n_tile * BLOCK_N;
          const int ob = it & 1;
          mbar_wait(smem_u32(&bar_out_ready[ob]), (it >> 1) & 1);
          if (elect_one()) {      // slots [n0, n0 + 64) -> physical channels [2 n0, 2 n0 + 128) of the unit's output
#pragma unroll
            for (int q = 0; q < BLOCK_N / 32; ++q)
              if (n0 + q * 32 < p.cout)
                tma_store_4d(&maps.d, staging + NB * Cfg::X1_BYTES + ob * Cfg::IL_OUT_BYTES + q * (BLOCK_M * 128),
                             2 * n0 + q * 64, w0, h0, nb0);
            tma_store_commit();
          }
          if (it + NB < my_tiles) acquire_staging(it + NB);
          if (elect_one()) {
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            mbar_arrive(smem_u32(&bar_out_free[ob]));
          }
        }
      } else {
      for (int j = 0; j < LA && j < my_tiles; ++j) acquire_staging(j);
      for (int it = 0; it < my_tiles; ++it) {
        if (LA == 0) {
          // no prefetch distance: staging[it % NB] was last read by the store of tile it - NB; stores issued so far: <= it - 1
          if (elect_one()) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(NB - 1) : "memory");
          acquire_staging(it);
        }
        // store of tile it, once its epilogue is done
        {
          const int tile0_ = blockIdx.x + it * gridDim.x;
          const int tile = (p.reverse & 1) ? total_tiles - 1 - tile0_ : tile0_;
          int n_tile, w0, h0, nb0;
          tile_coords(p, tile, n_tile, w0, h0, nb0);
          const int n0 = n_tile * BLOCK_N;
          const int ob = it % NB;
          mbar_wait(smem_u32(&bar_out_ready[ob]), (it / NB) & 1);
          if (elect_one()) {
#pragma unroll
            for (int q = 0; q < BLOCK_N / 64; ++q)
              if (n0 + q * 64 < p.cout)
                tma_store_4d(&maps.d, staging + ob * Cfg::OUT_BYTES + q * (BLOCK_M * 128), n0 + q * 64, w0, h0, nb0);
            tma_store_commit();
          }
        }
        // staging for tile it + LA: its buffer was last read by the store of tile it + LA - NB (= it - 2, or it - 1 with
        // the three big buffers of RES == 3); stores issued so far are tiles <= it, so allowing the NB - LA most recent
        // groups to be pending is enough.
        const int j = it + LA;
        if (LA > 0 && j < my_tiles) {
          if (elect_one()) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(NB - LA) : "memory");
          acquire_staging(j);
        }
      }
      }
      if (elect_one()) tma_store_wait_all();
    }
    __syncwarp();
  } else if (warp >= 12) {
    // =========================== A-operand scaler (launched only when p.a_scale is set) ===========================
    // Measured on B200 (EfficientNetV2-S project convs, K = 960 / 1536): this fusion LOSES -- 73 us against 22 us (conv) +
    // 18 us (separate scale pass).  The read-modify-write of the tile competes with the tensor core's operand reads for the
    // same shared-memory bandwidth and adds a hop to every ring slot, so it stays an opt-in (B200_FOLD_SE_SCALE=1).
    // SE excitation folded into the project conv: row r of the A tile (one pixel, 64 channels of this k-block, 128-byte
    // swizzled) is multiplied by a_scale[image of the pixel][channels] between the TMA load and the MMA, rounding to bf16
    // exactly where the separate scale kernel stored its result.
    const int r = threadIdx.x - 384;                       // 0..127: tile row
    const int sw = r & 7;
    // row of a_scale of this thread's pixel in tile `it` (null past the tensor): one division per TILE, not per k-block
    auto gate_row = [&](int it) -> const float* {
      if (it >= my_tiles) return nullptr;
      const int tile0_ = blockIdx.x + it * gridDim.x;
      const int tile = (p.reverse & 1) ? total_tiles - 1 - tile0_ : tile0_;
      int n_tile, w0, h0, nb0;
      tile_coords(p, tile, n_tile, w0, h0, nb0);
      const unsigned m = static_cast<unsigned>(w0 + r);    // flat pixel index (1x1 / stride 1 layers are tiled flat)
      if (static_cast<long long>(m) >= p.a_pixels) return nullptr;
      return p.a_scale + static_cast<size_t>(m / static_cast<unsigned>(p.a_scale_hw)) * p.a_scale_c;
    };
    auto gate_load = [&](const float* grow, int kb, float4 (&gv)[16]) {
#pragma unroll
      for (int c = 0; c < 16; ++c)
        gv[c] = (grow && kb * BLOCK_K + c * 4 < p.a_scale_c) ? *reinterpret_cast<const float4*>(grow + kb * BLOCK_K + c * 4)
                                                              : make_float4(0.f, 0.f, 0.f, 0.f);
    };
    float4 cur[16], nxt[16];
    const float* grow = gate_row(0);
    gate_load(grow, 0, cur);
    uint32_t g = 0;
    for (int it = 0; it < my_tiles; ++it) {
      const float* grow_next = nullptr;
      for (int kb = 0; kb < num_kb; ++kb, ++g) {
        const int s = g % STAGES;
        if (kb + 1 == num_kb) {
          grow_next = gate_row(it + 1);
          gate_load(grow_next, 0, nxt);
        } else {
          gate_load(grow, kb + 1, nxt);
        }
        mbar_wait(smem_u32(&bar_full[s]), (g / STAGES) & 1);
        const uint32_t row = ring + s * Cfg::STAGE_BYTES + r * 128;
        uint32_t v[8][4];
#pragma unroll
        for (int c = 0; c < 8; ++c)
          asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v[c][0]), "=r"(v[c][1]), "=r"(v[c][2]), "=r"(v[c][3])
                       : "r"(row + ((c ^ sw) << 4)));
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const float4 g0 = cur[2 * c], g1 = cur[2 * c + 1];
          const uint32_t o0 = pack_bf16x2(bf16_lo(v[c][0]) * g0.x, bf16_hi(v[c][0]) * g0.y);
          const uint32_t o1 = pack_bf16x2(bf16_lo(v[c][1]) * g0.z, bf16_hi(v[c][1]) * g0.w);
          const uint32_t o2 = pack_bf16x2(bf16_lo(v[c][2]) * g1.x, bf16_hi(v[c][2]) * g1.y);
          const uint32_t o3 = pack_bf16x2(bf16_lo(v[c][3]) * g1.z, bf16_hi(v[c][3]) * g1.w);
          asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(row + ((c ^ sw) << 4)), "r"(o0), "r"(o1), "r"(o2), "r"(o3)
                       : "memory");
        }
        fence_proxy_async_smem();                          // generic-proxy writes -> visible to the tensor core's reads
        mbar_arrive(smem_u32(&bar_scaled[s]));
#pragma unroll
        for (int c = 0; c < 16; ++c) cur[c] = nxt[c];
      }
      grow = grow_next;
    }
  } else if (warp >= 4) {
    // =========================== epilogue ===========================
    const int wg = (warp - 4) >> 2;                  // epilogue warpgroup 0 / 1 <-> accumulator 0 / 1
    const int epi_tid = (threadIdx.x - 128) & 127;   // 0..127 within the group
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t row_off = row * 128;
    const int sw = row & 7;
    for (int it = wg; it < my_tiles; it += 2) {
      const int tile0_ = blockIdx.x + it * gridDim.x;
      const int tile = (p.reverse & 1) ? total_tiles - 1 - tile0_ : tile0_;
      const int b = it & 1;
      const int bb = wg;
      const int sb = it % NB;                       // staging buffer (RES == 3: slot of the other branch's tile)
      const int ob = RES == 3 ? (it & 1) : sb;      // buffer the output is staged in
      const int n0 = (tile - static_cast<int>(fast_div(tile, p.n_tiles, p.mul_n_tiles)) * p.n_tiles) * BLOCK_N;
      const int x1_ch8 = RES == 3 ? ((n0 < p.il_group ? n0 : p.il_half + (n0 - p.il_group)) & ~7) : 0;   // as in the DMA warp
      named_bar_sync(1 + wg, 128);               // every thread of the group is done reading the previous tile's bias
      for (int i = epi_tid; i < BLOCK_N; i += 128) {
        const int c = n0 + i;
        bias_s[bb][i] = (c < p.cout) ? p.bias[c] : 0.f;
      }
      if (RES == 3 && epi_tid < BLOCK_N / 2) {      // slots past the valid channels of a half-group hold the NEXT channels
        uint32_t m = 0;                             // of the other branch (the boxes are contiguous): masked to zero
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int slot = n0 + 2 * epi_tid + e;
          const int in_group = slot < p.il_group ? slot : slot - p.il_group;
          if (in_group < p.il_half) m |= 0xFFFFu << (16 * e);
        }
        keep_s[bb][epi_tid] = m;
      }
      named_bar_sync(1 + wg, 128);
      mbar_wait(smem_u32(&bar_tmem_full[b]), (it >> 1) & 1);
      tc_fence_after();
      mbar_wait(smem_u32(&bar_stage_ready[sb]), (it / NB) & 1);
      if (RES == 3) mbar_wait(smem_u32(&bar_out_free[ob]), ((it >> 1) & 1) ^ 1);
      const uint32_t stage_d = staging + sb * Cfg::OUT_BYTES;
      const uint32_t out_d = staging + NB * Cfg::OUT_BYTES + ob * Cfg::IL_OUT_BYTES;      // RES == 3 only
      const uint32_t acc = tmem_base + b * BLOCK_N + (static_cast<uint32_t>(quarter * 32) << 16);
#pragma unroll 1
      for (int cc = 0; cc < BLOCK_N / 32; ++cc) {
        uint32_t v[32];
        tmem_ld_32x32(acc + cc * 32, v);
        tmem_ld_wait();
        const uint32_t sub = stage_d + (cc >> 1) * (BLOCK_M * 128) + row_off;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int chunk = (cc & 1) * 4 + g;
          const uint32_t addr = sub + ((chunk ^ sw) << 4);
          const float4 b0 = *reinterpret_cast<const float4*>(&bias_s[bb][cc * 32 + g * 8]);
          const float4 b1 = *reinterpret_cast<const float4*>(&bias_s[bb][cc * 32 + g * 8 + 4]);
          float f[8];
          f[0] = __uint_as_float(v[g * 8 + 0]) + b0.x; f[1] = __uint_as_float(v[g * 8 + 1]) + b0.y;
          f[2] = __uint_as_float(v[g * 8 + 2]) + b0.z; f[3] = __uint_as_float(v[g * 8 + 3]) + b0.w;
          f[4] = __uint_as_float(v[g * 8 + 4]) + b1.x; f[5] = __uint_as_float(v[g * 8 + 5]) + b1.y;
          f[6] = __uint_as_float(v[g * 8 + 6]) + b1.z; f[7] = __uint_as_float(v[g * 8 + 7]) + b1.w;
          if (RES == 3) {
            // cat + channel_shuffle(groups = 2): slots 8 k8 .. 8 k8 + 7 of this row and the other branch's values of the
            // same slots become 16 consecutive output channels (other branch on the even ones, bits copied as they are).
            // The other branch's tile starts at channel x1_ch8 (a multiple of 8): 64 channels in 128-byte swizzled rows.
            const int k8 = cc * 4 + g;
            const int slot0 = n0 + k8 * 8;
            const int off = (slot0 < p.il_group ? slot0 : p.il_half + (slot0 - p.il_group)) - x1_ch8;   // first channel, from the box start
            const int q = off >> 3, dl = off & 7;
            uint32_t r0, r1, r2, r3;
            if (dl == 0) {
              asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                           : "r"(stage_d + row_off + ((q ^ sw) << 4)));
            } else {
              // 16 bytes that start 2 dl bytes into chunk q: five words from chunks q and q + 1 (chunk 8 = the 8-channel
              // box behind the 64-channel one), then a 16-bit funnel shift when dl is odd
              uint32_t wv[5];
#pragma unroll
              for (int t = 0; t < 5; ++t) {
                const int wi = (dl >> 1) + t;
                const int ch = q + (wi >> 2);
                const uint32_t a = ch < 8 ? stage_d + row_off + ((ch ^ sw) << 4) + ((wi & 3) << 2)
                                          : stage_d + BLOCK_M * BLOCK_N * 2 + row * 16 + ((wi & 3) << 2);
                asm volatile("ld.shared.u32 %0, [%1];" : "=r"(wv[t]) : "r"(a));
              }
              const uint32_t sh = (dl & 1) * 16;
              r0 = __funnelshift_r(wv[0], wv[1], sh); r1 = __funnelshift_r(wv[1], wv[2], sh);
              r2 = __funnelshift_r(wv[2], wv[3], sh); r3 = __funnelshift_r(wv[3], wv[4], sh);
            }
            const uint4 kp = *reinterpret_cast<const uint4*>(&keep_s[bb][k8 * 4]);
            r0 &= kp.x; r1 &= kp.y; r2 &= kp.z; r3 &= kp.w;
#pragma unroll
            for (int i = 0; i < 8; ++i) f[i] = act_t<ACT>(f[i]);
            const uint32_t q0 = pack_bf16x2(f[0], f[1]), q1 = pack_bf16x2(f[2], f[3]);
            const uint32_t q2 = pack_bf16x2(f[4], f[5]), q3 = pack_bf16x2(f[6], f[7]);
            const uint32_t osub = out_d + cc * (BLOCK_M * 128) + row_off;
            asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(osub + (((2 * g) ^ sw) << 4)),
                         "r"(__byte_perm(r0, q0, 0x5410)), "r"(__byte_perm(r0, q0, 0x7632)),
                         "r"(__byte_perm(r1, q1, 0x5410)), "r"(__byte_perm(r1, q1, 0x7632)) : "memory");
            asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(osub + (((2 * g + 1) ^ sw) << 4)),
                         "r"(__byte_perm(r2, q2, 0x5410)), "r"(__byte_perm(r2, q2, 0x7632)),
                         "r"(__byte_perm(r3, q3, 0x5410)), "r"(__byte_perm(r3, q3, 0x7632)) : "memory");
            continue;
          }
          if (RES) {
            uint32_t r0, r1, r2, r3;
            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
            const float rr[8] = {bf16_lo(r0), bf16_hi(r0), bf16_lo(r1), bf16_hi(r1),
                                 bf16_lo(r2), bf16_hi(r2), bf16_lo(r3), bf16_hi(r3)};
#pragma unroll
            for (int i = 0; i < 8; ++i) f[i] = (RES == 2) ? act_t<ACT>(f[i]) + rr[i] : act_t<ACT>(f[i] + rr[i]);
          } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) f[i] = act_t<ACT>(f[i]);
          }
          const uint32_t o0 = pack_bf16x2(f[0], f[1]), o1 = pack_bf16x2(f[2], f[3]);
          const uint32_t o2 = pack_bf16x2(f[4], f[5]), o3 = pack_bf16x2(f[6], f[7]);
          asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(o0), "r"(o1), "r"(o2), "r"(o3)
                       : "memory");
        }
      }
      tc_fence_before();
      mbar_arrive(smem_u32(&bar_tmem_empty[b]));   // accumulator may be overwritten by tile it+2
      fence_proxy_async_smem();                    // staging writes -> visible to the TMA store
      mbar_arrive(smem_u32(&bar_out_ready[ob]));
    }
  }

  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<2 * BLOCK_N>(tmem_base);
  }
}

// ================================================================================================================
// 3x3 / stride 1 / pad 1 convolution with few input channels (Cin <= 64) from a halo tile.
// The generic kernel loads the 128-pixel input tile nine times (once per tap) and the weights once per tile, which makes a
// narrow layer shared-memory-bandwidth bound (~31 % of the tensor peak at 64 channels, far less below).  Here a tile is
// 16 rows x 8 pixels of one image; ONE TMA box brings its 18-row x 10-pixel halo and the nine taps are nine shifted *views*
// of it: descriptor start = halo + (dy*10 + dx) * PB, SBO = 10 * PB (one halo row per 8-pixel group), base_offset 0 (the
// hardware swizzles on absolute address bits, so shifted starts stay consistent with what TMA wrote).
// PB = bytes per staged pixel = the swizzle span: 128 B (33..64 channels), 64 B (17..32) or 32 B (<= 16); channels between
// Cin and PB/2 are out of bounds of the tensor map and arrive as zeros, so HBM keeps the compact layout.  GEMM-K per tap is
// PB/2.  The 9 x N x PB/2 weights of the CTA's output-channel tile (N = 32 or 64) stay resident in shared memory; with
// Cout > 64 the grid is a multiple of the channel tiles so that a CTA always works on the same one.
// Measured floor (ablation on B200, 256 x 56 x 56 x 64 -> 64): with loads, stores and the epilogue all disabled the 36
// M128 x N64 x K16 MMAs of a tile still take ~1.7 us (~90 cycles each, aligned or shifted views alike): operands stream
// from shared memory at ~64 B/clk/SM (A 4 KB + B 2 KB per MMA), so 64-wide layers are operand-bandwidth bound at ~45 %
// of the tensor peak whatever the pipelining; the same law gives 128 cycles for N = 128 and peak only for CTA pairs.
// ================================================================================================================
static constexpr int HALO_ROWS = 18, HALO_PITCH_PX = 10;   // exactly the 18 x 10 pixels a 16 x 8 tile needs
static constexpr int HALO_MAX_STAGES = 4;

struct HaloCfg {
  static constexpr int MAX_OUT_BUFS = 8;    // staging buffers (4, 6 or 8: what fits); residual prefetch runs half of them ahead
  static constexpr int OUT_BYTES = BLOCK_M * 64 * 2;
  static constexpr int halo_bytes(int pb) { return HALO_ROWS * HALO_PITCH_PX * pb; }               // what TMA delivers
  static constexpr int halo_stride(int pb) { return (halo_bytes(pb) + 1023) / 1024 * 1024; }       // stage pitch
  static constexpr int w_bytes(int pb, int n) { return 9 * n * pb; }
  static constexpr int smem_bytes(int pb, int n, int stages, int out_bufs) {
    return w_bytes(pb, n) + stages * halo_stride(pb) + out_bufs * OUT_BYTES + 1024;
  }
};

__device__ __forceinline__ uint64_t make_u64(uint32_t lo, uint32_t hi) {
  uint64_t d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(hi));
  return d;
}

// K-major swizzled descriptor (span = pb bytes) with explicit 8-row-group stride; start need not be atom aligned
__device__ __forceinline__ uint64_t umma_desc_swz(uint32_t smem_addr, uint32_t sbo_bytes, int pb) {
  const uint32_t layout = pb == 128 ? 2u : (pb == 64 ? 4u : 6u);
  const uint32_t lo = ((smem_addr >> 4) & 0x3FFFu) | (1u << 16);
  const uint32_t hi = (sbo_bytes >> 4) | (1u << 14) | (layout << 29);
  return (static_cast<uint64_t>(hi) << 32) | lo;
}

template <int PB, int RES, int ACT>
__global__ void __launch_bounds__(HALO_THREADS, 1)
conv3x3_halo_kernel(const __grid_constant__ ConvMaps maps, const __grid_constant__ ConvParams p, const int total_tiles) {
  using Cfg = HaloCfg;
  constexpr int BLOCK_N = 64;          // accumulator stride in TMEM and bias buffer size; the tile's width is p.block_n
  constexpr int NACC = HALO_WGS;
  const int NB = p.halo_out_bufs, LA = NB >> 1;
  const int N = p.block_n, HS = p.halo_stages;
  constexpr int halo_bytes = Cfg::halo_bytes(PB), halo_stride = Cfg::halo_stride(PB);
  const int tap_bytes = N * PB;

  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_halo_full[HALO_MAX_STAGES];
  __shared__ __align__(8) uint64_t bar_halo_empty[HALO_MAX_STAGES];
  __shared__ __align__(8) uint64_t bar_w_full;
  __shared__ __align__(8) uint64_t bar_tmem_full[NACC];
  __shared__ __align__(8) uint64_t bar_tmem_empty[NACC];
  __shared__ __align__(8) uint64_t bar_stage_ready[Cfg::MAX_OUT_BUFS];
  __shared__ __align__(8) uint64_t bar_out_ready[Cfg::MAX_OUT_BUFS];
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float bias_s[NACC][BLOCK_N];

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t wres = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t halo = wres + 9 * tap_bytes;
  const int planes = p.halo_s2 ? 4 : 1;                    // stride 2: parity planes (py, px) of the input, plane = py * 2 + px
  const int stage_bytes = planes * halo_stride;
  const uint32_t staging = halo + HS * stage_bytes;
  const int my_tiles = (total_tiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.a[0]);
    tma_prefetch_desc(&maps.b);
    tma_prefetch_desc(&maps.d);
    if (RES) tma_prefetch_desc(&maps.r);
    for (int b = 0; b < HALO_MAX_STAGES; ++b) {
      mbar_init(smem_u32(&bar_halo_full[b]), 1);
      mbar_init(smem_u32(&bar_halo_empty[b]), 1);
    }
    for (int b = 0; b < NACC; ++b) {
      mbar_init(smem_u32(&bar_tmem_full[b]), 1);
      mbar_init(smem_u32(&bar_tmem_empty[b]), 128);
    }
    mbar_init(smem_u32(&bar_w_full), 1);
    for (int b = 0; b < NB; ++b) {
      mbar_init(smem_u32(&bar_stage_ready[b]), 1);
      mbar_init(smem_u32(&bar_out_ready[b]), 128);
    }
    fence_barrier_init();
  } else if (warp == 1) {
    tmem_alloc<NACC * BLOCK_N>(smem_u32(&tmem_base_s));
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  pdl_launch_dependents();
  pdl_wait();      // everything above overlapped the previous kernel's tail; activations are read from here on

  if (warp == 0) {
    // =========================== TMA producer: weights once, one halo box per tile ===========================
    if (my_tiles > 0) {
      {
        // the grid is a multiple of n_tiles, so every tile of this CTA has the channel tile of its first one
        const int first = (p.reverse & 1) ? total_tiles - 1 - static_cast<int>(blockIdx.x) : static_cast<int>(blockIdx.x);
        const int n0 = (first % p.n_tiles) * N;
        if (elect_one()) {
          mbar_expect_tx(smem_u32(&bar_w_full), 9 * tap_bytes);
          for (int t = 0; t < 9; ++t) tma_load_3d(&maps.b, smem_u32(&bar_w_full), wres + t * tap_bytes, 0, t, n0);
        }
      }
      // ring positions are running counters (stage, phase): HS and NB are run-time values, and a division per tile in a
      // warp that runs alone costs more than the TMA it is about to issue
      int hb = 0;
      uint32_t hph = 0;
      for (int it = 0; it < my_tiles; ++it) {
        const int tile0_ = blockIdx.x + it * gridDim.x;
        const int tile = (p.reverse & 1) ? total_tiles - 1 - tile0_ : tile0_;
        int n_tile, w0, h0, nb0;
        tile_coords(p, tile, n_tile, w0, h0, nb0);
        mbar_wait(smem_u32(&bar_halo_empty[hb]), hph ^ 1);
        const uint32_t full = smem_u32(&bar_halo_full[hb]);
        if (elect_one()) {
          mbar_expect_tx(full, planes * halo_bytes);
          for (int pl = 0; pl < planes; ++pl)     // box [PB/2, 10, 18, 1] of the input (stride 1) or of each parity plane
            tma_load_4d(&maps.a[pl], full, halo + hb * stage_bytes + pl * halo_stride, 0, w0 - 1, h0 - 1, nb0);
        }
        if (++hb == HS) { hb = 0; hph ^= 1; }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // =========================== MMA issuer: nine shifted views of the halo ===========================
    // The whole warp walks the tile loop (uniform control flow); one elected lane issues the MMAs.  This one thread
    // issues every MMA of the CTA, so the issue sequence is kept to ~3 uniform-datapath instructions per MMA: the
    // descriptors' high words are loop constants, the low words (start address >> 4) are a per-tile base plus
    // compile-time tap / k offsets.  base_offset stays 0: measured on B200, the swizzle is applied to the absolute
    // shared-memory address bits, so a start shifted by dx pixels sees exactly the pattern TMA wrote.
    if (my_tiles > 0) {
      const uint32_t idesc = umma_idesc_bf16_f32(BLOCK_M, N);
      constexpr int KSTEPS = PB / 32;     // K = 16 (32 bytes) per MMA
      const uint64_t adesc0 = umma_desc_swz(halo, HALO_PITCH_PX * PB, PB);
      const uint64_t bdesc0 = umma_desc_swz(wres, 8 * PB, PB);
      const uint32_t a_hi = static_cast<uint32_t>(adesc0 >> 32), b_hi = static_cast<uint32_t>(bdesc0 >> 32);
      const uint32_t a_lo0 = static_cast<uint32_t>(adesc0), b_lo0 = static_cast<uint32_t>(bdesc0);
      const uint32_t wtap = static_cast<uint32_t>(tap_bytes) >> 4;
      mbar_wait(smem_u32(&bar_w_full), 0);
      int hb = 0;
      uint32_t hph = 0;
      for (int it = 0; it < my_tiles; ++it) {
        const int b = it % NACC;
        mbar_wait(smem_u32(&bar_tmem_empty[b]), ((it / NACC) & 1) ^ 1);
        mbar_wait(smem_u32(&bar_halo_full[hb]), hph);
        tc_fence_after();
        const uint32_t acc = tmem_base + b * BLOCK_N;
        const uint32_t a_lo = a_lo0 + hb * (stage_bytes >> 4);   // no carry out of the 14-bit address field: smem < 256 KB
        const bool s2 = p.halo_s2 != 0;
        const int ksteps = p.halo_ksteps;
        if (elect_one()) {
#pragma unroll
          for (int t = 0; t < 9; ++t) {
            // stride 1: view shifted by (dy, dx) pixels.  stride 2: tap (dy, dx) reads in(2y + dy - 1, 2x + dx - 1) = parity
            // plane (dy != 1, dx != 1) at (y - (dy == 0), x - (dx == 0)): with the box origin at (y0 - 1, x0 - 1) that is a
            // shift of (dy != 0, dx != 0) pixels inside the plane's 18 x 10 box.  Both offsets are compile-time constants.
            const int dy = t / 3, dx = t % 3;
            const uint32_t off1 = static_cast<uint32_t>((dy * HALO_PITCH_PX + dx) * PB) >> 4;
            const uint32_t off2 = static_cast<uint32_t>(((dy != 1 ? 2 : 0) + (dx != 1 ? 1 : 0)) * halo_stride +
                                                        ((dy != 0 ? 1 : 0) * HALO_PITCH_PX + (dx != 0 ? 1 : 0)) * PB) >> 4;
            const uint32_t a_t = a_lo + (s2 ? off2 : off1);
            const uint32_t b_t = b_lo0 + t * wtap;
#pragma unroll
            for (int k = 0; k < KSTEPS; ++k)
              if (k < ksteps)     // channels from 16 * ksteps on are zero padding of the PB-byte pixel slot
                umma_bf16(acc, make_u64(a_t + 2 * k, a_hi), make_u64(b_t + 2 * k, b_hi), idesc, (t | k) != 0 ? 1u : 0u);
          }
          umma_commit(smem_u32(&bar_halo_empty[hb]));
          umma_commit(smem_u32(&bar_tmem_full[b]));
        }
        __syncwarp();
        if (++hb == HS) { hb = 0; hph ^= 1; }
      }
    }
  } else if (warp == 2) {
    // =========================== DMA: residual prefetch (LA tiles ahead) + output stores ===========================
    {
      // whole warp in the loops, single-thread work under elect_one() (always the same lane, so the bulk-group
      // bookkeeping of the TMA stores stays with one thread)
      auto acquire_staging = [&](int j, int jb) {  // make staging[jb = j % NB] available to the epilogue of tile j
        const int tile0_ = blockIdx.x + j * gridDim.x;
        const int tile = (p.reverse & 1) ? total_tiles - 1 - tile0_ : tile0_;
        const uint32_t rdy = smem_u32(&bar_stage_ready[jb]);
        if (RES) {
          int n_tile, w0, h0, nb0;
          tile_coords(p, tile, n_tile, w0, h0, nb0);
          if (elect_one()) {
            mbar_expect_tx(rdy, Cfg::OUT_BYTES);
#pragma unroll
            for (int q = 0; q < BLOCK_N / 64; ++q)
              tma_load_4d(&maps.r, rdy, staging + jb * Cfg::OUT_BYTES + q * (BLOCK_M * 128), n_tile * N + q * 64,
                          w0, h0, nb0);
          }
        } else {
          if (elect_one()) mbar_arrive(rdy);
        }
      };
      for (int j = 0; j < LA && j < my_tiles; ++j) acquire_staging(j, j);
      int ob = 0;
      uint32_t oph = 0;
      for (int it = 0; it < my_tiles; ++it) {
        // store of tile it, once its epilogue is done
        {
          const int tile0_ = blockIdx.x + it * gridDim.x;
          const int tile = (p.reverse & 1) ? total_tiles - 1 - tile0_ : tile0_;
          int n_tile, w0, h0, nb0;
          tile_coords(p, tile, n_tile, w0, h0, nb0);
          const int n0 = n_tile * N;
          mbar_wait(smem_u32(&bar_out_ready[ob]), oph);
          if (elect_one()) {
#pragma unroll
            for (int q = 0; q < BLOCK_N / 64; ++q)
              if (n0 + q * 64 < p.cout)
                tma_store_4d(&maps.d, staging + ob * Cfg::OUT_BYTES + q * (BLOCK_M * 128), n0 + q * 64, w0, h0, nb0);
            tma_store_commit();
          }
        }
        // staging for tile it + LA: its buffer was last read by the store of tile it + LA - NB (= it - NB / 2);
        // stores issued so far are tiles <= it, so allowing the NB / 2 most recent groups to be pending is enough.
        const int j = it + LA;
        if (j < my_tiles) {
          if (elect_one()) {
            if (LA == 2) asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory");
            else if (LA == 3) asm volatile("cp.async.bulk.wait_group.read 3;" ::: "memory");
            else asm volatile("cp.async.bulk.wait_group.read 4;" ::: "memory");
          }
          acquire_staging(j, ob + LA >= NB ? ob + LA - NB : ob + LA);
        }
        if (++ob == NB) { ob = 0; oph ^= 1; }
      }
      if (elect_one()) tma_store_wait_all();
    }
    __syncwarp();
  } else if (warp >= 4) {
    // =========================== epilogue ===========================
    // HALO_WGS warpgroups, one per accumulator: group g drains tiles g, g + HALO_WGS, ...  A group is one warp per
    // scheduler walking a dependent chain (TMEM load -> bias/residual/activation -> pack -> staging): ~1500-2700 cycles
    // of mostly latency per tile, longer than a thin layer's MMAs.  Such layers are paced by how many of these chains run
    // side by side, hence four groups and four accumulators.
    const int wg = (warp - 4) >> 2;
    const int epi_tid = (threadIdx.x - 128) & 127;  // 0..127 within the group
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t row_off = row * 128;
    const int sw = row & 7;
    {
      // the channel tile is the same for every tile of this CTA: the bias is staged once per group
      const int first = (p.reverse & 1) ? total_tiles - 1 - static_cast<int>(blockIdx.x) : static_cast<int>(blockIdx.x);
      const int n0 = (first % p.n_tiles) * N;
      for (int i = epi_tid; i < N; i += 128) {
        const int c = n0 + i;
        bias_s[wg][i] = (c < p.cout) ? p.bias[c] : 0.f;
      }
      named_bar_sync(1 + wg, 128);
    }
    int ob = wg;                  // = it % NB (NB >= NACC), advanced by NACC per tile of this group
    uint32_t oph = 0;             // = (it / NB) & 1
    for (int it = wg; it < my_tiles; it += NACC) {
      const int b = wg;
      mbar_wait(smem_u32(&bar_tmem_full[b]), (it / NACC) & 1);
      tc_fence_after();
      mbar_wait(smem_u32(&bar_stage_ready[ob]), oph);
      const uint32_t stage_d = staging + ob * Cfg::OUT_BYTES;
      const uint32_t acc = tmem_base + b * BLOCK_N + (static_cast<uint32_t>(quarter * 32) << 16);
#pragma unroll 1
      for (int cc = 0; cc < N / 32; ++cc) {
        uint32_t v[32];
        tmem_ld_32x32(acc + cc * 32, v);
        tmem_ld_wait();
        const uint32_t sub = stage_d + (cc >> 1) * (BLOCK_M * 128) + row_off;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int chunk = (cc & 1) * 4 + g;
          const uint32_t addr = sub + ((chunk ^ sw) << 4);
          const float4 b0 = *reinterpret_cast<const float4*>(&bias_s[b][cc * 32 + g * 8]);
          const float4 b1 = *reinterpret_cast<const float4*>(&bias_s[b][cc * 32 + g * 8 + 4]);
          float f[8];
          f[0] = __uint_as_float(v[g * 8 + 0]) + b0.x; f[1] = __uint_as_float(v[g * 8 + 1]) + b0.y;
          f[2] = __uint_as_float(v[g * 8 + 2]) + b0.z; f[3] = __uint_as_float(v[g * 8 + 3]) + b0.w;
          f[4] = __uint_as_float(v[g * 8 + 4]) + b1.x; f[5] = __uint_as_float(v[g * 8 + 5]) + b1.y;
          f[6] = __uint_as_float(v[g * 8 + 6]) + b1.z; f[7] = __uint_as_float(v[g * 8 + 7]) + b1.w;
          if (RES) {
            uint32_t r0, r1, r2, r3;
            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
            const float rr[8] = {bf16_lo(r0), bf16_hi(r0), bf16_lo(r1), bf16_hi(r1),
                                 bf16_lo(r2), bf16_hi(r2), bf16_lo(r3), bf16_hi(r3)};
#pragma unroll
            for (int i = 0; i < 8; ++i) f[i] = (RES == 2) ? act_t<ACT>(f[i]) + rr[i] : act_t<ACT>(f[i] + rr[i]);
          } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) f[i] = act_t<ACT>(f[i]);
          }
          const uint32_t o0 = pack_bf16x2(f[0], f[1]), o1 = pack_bf16x2(f[2], f[3]);
          const uint32_t o2 = pack_bf16x2(f[4], f[5]), o3 = pack_bf16x2(f[6], f[7]);
          asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(o0), "r"(o1), "r"(o2), "r"(o3)
                       : "memory");
        }
      }
      tc_fence_before();
      mbar_arrive(smem_u32(&bar_tmem_empty[b]));   // accumulator may be overwritten by tile it+2
      fence_proxy_async_smem();                    // staging writes -> visible to the TMA store
      mbar_arrive(smem_u32(&bar_out_ready[ob]));
      ob += NACC;
      if (ob >= NB) { ob -= NB; oph ^= 1; }
    }
  }

  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<NACC * BLOCK_N>(tmem_base);
  }
}

// ================================================================================================================
// CTA-pair variant (tcgen05.mma.cta_group::2): two CTAs on the two SMs of a TPC compute one 256-pixel x BLOCK_N tile.
// Each CTA loads its own 128 pixels of A and *half* of the weight tile (BLOCK_N/2 rows); the leader issues M=256 MMAs
// that read both halves, so per SM the shared-memory operand traffic per FLOP halves -- single-CTA tiles are bound by
// shared-memory bandwidth (write by TMA + read by the tensor core) at ~50 % of the tensor peak for N=128.
// Accumulator rows 0-127 live in the leader's TMEM, 128-255 in the peer's; epilogue, residual prefetch and stores are
// per CTA exactly as in the single-CTA persistent kernel.  Barrier plumbing: both CTAs' TMA loads credit the leader's
// `full` barrier; tcgen05.commit multicasts `empty` / `tmem_full` arrivals to both CTAs; both epilogues release the
// accumulator on the leader's `tmem_empty` barrier (count 256).
// ================================================================================================================
static constexpr int P2CONV_THREADS = 384;   // warps 0-2: TMA / MMA / DMA, 4-7 and 8-11: one epilogue warpgroup per 128 columns

template <int BLOCK_N, int RES>
struct P2ConvCfg {
  static constexpr int STAGES = RES ? (BLOCK_N == 256 ? 3 : 4) : (BLOCK_N == 256 ? 5 : 6);
  static constexpr int SUB_N = 128;                  // the epilogue / residual / store pipeline works on 128-column units
  static constexpr int SUBS = BLOCK_N / SUB_N;
  static constexpr int OUT_BUFS = RES ? 4 : 2;
  static constexpr int LOOKAHEAD = OUT_BUFS - 2;     // units the residual prefetch runs ahead
  static constexpr int B_STAGE_BYTES = (BLOCK_N / 2) * BLOCK_K * 2;  // this CTA's half of the weight tile
  static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  static constexpr int OUT_BYTES = BLOCK_M * SUB_N * 2;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + OUT_BUFS * OUT_BYTES + 1024;
  static_assert(SMEM_BYTES <= 230 * 1024, "smem budget");
  static_assert(2 * BLOCK_N <= 512, "TMEM budget");
};

template <int BLOCK_N, int RES, int ACT>
__global__ void __launch_bounds__(P2CONV_THREADS, 1)
conv_gemm_2cta_kernel(const __grid_constant__ ConvMaps maps, const __grid_constant__ ConvParams p,
                      const int total_pair_tiles, const int m_tiles) {
  using Cfg = P2ConvCfg<BLOCK_N, RES>;
  constexpr int STAGES = Cfg::STAGES;
  constexpr int NB = Cfg::OUT_BUFS;
  constexpr int LA = Cfg::LOOKAHEAD;

  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[STAGES];
  __shared__ __align__(8) uint64_t bar_empty[STAGES];
  __shared__ __align__(8) uint64_t bar_tmem_full[2];
  __shared__ __align__(8) uint64_t bar_tmem_empty[2];
  __shared__ __align__(8) uint64_t bar_stage_ready[NB];
  __shared__ __align__(8) uint64_t bar_out_ready[NB];
  __shared__ uint32_t tmem_base_s;
  constexpr int NBIAS = BLOCK_N == 256 ? 1 : 2;   // N=256 is 128 bytes short of smem with two bias buffers
  __shared__ __align__(16) float bias_s[NBIAS][BLOCK_N];

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();        // 0 = leader
  const int pair = blockIdx.x >> 1;
  const int num_pairs = gridDim.x >> 1;
  const uint32_t ring = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t staging = ring + STAGES * Cfg::STAGE_BYTES;
  const int num_kb = p.num_kb;
  const int my_tiles = (total_pair_tiles - pair + num_pairs - 1) / num_pairs;

  // pair tile -> this CTA's 128-pixel m-tile and the shared n-tile
  auto my_coords = [&](int it, int& n0, int& w0, int& h0, int& nb0) {
    const int pt0 = pair + it * num_pairs;
    const int pt = (p.reverse & 1) ? total_pair_tiles - 1 - pt0 : pt0;
    const int pm = fast_div(pt, p.n_tiles, p.mul_n_tiles);
    const int n_tile = pt - pm * p.n_tiles;
    int m = pm * 2 + static_cast<int>(rank);
    if (m >= m_tiles) m = m_tiles;            // phantom tile of an odd tail: coordinates past the tensor (TMA clips)
    const int mw = fast_div(m, p.tiles_w, p.mul_tiles_w);
    const int t_w = m - mw * p.tiles_w;
    const int t_n = fast_div(mw, p.tiles_h, p.mul_tiles_h);
    const int t_h = mw - t_n * p.tiles_h;
    w0 = t_w * p.tw; h0 = t_h * p.th; nb0 = t_n * p.tn;
    n0 = n_tile * BLOCK_N;
  };

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.a[0]);
    tma_prefetch_desc(&maps.b);
    tma_prefetch_desc(&maps.d);
    if (RES) tma_prefetch_desc(&maps.r);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(smem_u32(&bar_full[s]), 1);
      mbar_init(smem_u32(&bar_empty[s]), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(smem_u32(&bar_tmem_full[b]), 1);
      mbar_init(smem_u32(&bar_tmem_empty[b]), 2 * 128 * Cfg::SUBS);   // every epilogue thread of both CTAs (leader's copy is used)
    }
    for (int b = 0; b < NB; ++b) {
      mbar_init(smem_u32(&bar_stage_ready[b]), 1);
      mbar_init(smem_u32(&bar_out_ready[b]), 128);
    }
    fence_barrier_init();
  } else if (warp == 1) {
    tmem_alloc_2cta<2 * BLOCK_N>(smem_u32(&tmem_base_s));
  }
  tc_fence_before();
  cluster_sync_all();      // peer barriers are initialised and TMEM is allocated in both CTAs
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  pdl_launch_dependents();
  pdl_wait();      // everything above overlapped the previous kernel's tail; activations are read from here on

  if (warp == 0) {
    // =========================== TMA producer (both CTAs) ===========================
    {
      uint32_t g = 0;
      for (int it = 0; it < my_tiles; ++it) {
        int n0, w0, h0, nb0;
        my_coords(it, n0, w0, h0, nb0);
        for (int kb = 0; kb < num_kb; ++kb, ++g) {
          const int s = g % STAGES;
          const uint32_t ph = (g / STAGES) & 1;
          mbar_wait(smem_u32(&bar_empty[s]), ph ^ 1);
          const uint32_t full = smem_u32(&bar_full[s]);
          int t, c0, cw, tw_;
          kb_source(p, kb, t, c0, cw, tw_);
          const uint32_t a_dst = ring + s * Cfg::STAGE_BYTES;
          if (elect_one()) {
            if (rank == 0) mbar_expect_tx(full, 2 * Cfg::STAGE_BYTES);   // both CTAs' bytes land on the leader's barrier
            tma_load_4d_2cta(&maps.a[p.tap_map[t]], full, a_dst, c0, w0 + p.tap_dx[t], h0 + p.tap_dy[t], nb0);
            tma_load_3d_2cta(&maps.b, full, a_dst + A_STAGE_BYTES, cw, tw_, n0 + static_cast<int>(rank) * (BLOCK_N / 2));
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // =========================== MMA issuer (leader CTA only) ===========================
    if (rank == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16_f32(2 * BLOCK_M, BLOCK_N);
      uint32_t g = 0;
      for (int it = 0; it < my_tiles; ++it) {
        const int b = it & 1;
        mbar_wait_cluster(smem_u32(&bar_tmem_empty[b]), ((it >> 1) & 1) ^ 1);  // both epilogues drained this accumulator
        tc_fence_after();
        const uint32_t acc = tmem_base + b * BLOCK_N;
        for (int kb = 0; kb < num_kb; ++kb, ++g) {
          const int s = g % STAGES;
          const uint32_t ph = (g / STAGES) & 1;
          mbar_wait_cluster(smem_u32(&bar_full[s]), ph);
          tc_fence_after();
          const uint32_t a_src = ring + s * Cfg::STAGE_BYTES;
          const uint64_t adesc = umma_desc_sw128(a_src);
          const uint64_t bdesc = umma_desc_sw128(a_src + A_STAGE_BYTES);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BLOCK_K / 16; ++k)
              umma_bf16_2cta(acc, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
            umma_commit_2cta(smem_u32(&bar_empty[s]));       // frees the slot in both CTAs
            if (kb == num_kb - 1) umma_commit_2cta(smem_u32(&bar_tmem_full[b]));     // accumulator ready in both CTAs
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 2) {
    // =========================== DMA: residual prefetch + output stores (per CTA, 128-column units) ===========================
    {
      constexpr int SUBS = Cfg::SUBS, SUB_N = Cfg::SUB_N;
      const int total_units = my_tiles * SUBS;
      auto unit_coords = [&](int u, int& n0, int& w0, int& h0, int& nb0) {
        my_coords(u / SUBS, n0, w0, h0, nb0);
        n0 += (u % SUBS) * SUB_N;
      };
      auto acquire_staging = [&](int u) {
        const uint32_t rdy = smem_u32(&bar_stage_ready[u % NB]);
        if (RES) {
          int n0, w0, h0, nb0;
          unit_coords(u, n0, w0, h0, nb0);
          if (elect_one()) {
            mbar_expect_tx(rdy, Cfg::OUT_BYTES);
#pragma unroll
            for (int q = 0; q < SUB_N / 64; ++q)
              tma_load_4d(&maps.r, rdy, staging + (u % NB) * Cfg::OUT_BYTES + q * (BLOCK_M * 128), n0 + q * 64, w0, h0, nb0);
          }
        } else {
          if (elect_one()) mbar_arrive(rdy);
        }
      };
      for (int u = 0; u < LA && u < total_units; ++u) acquire_staging(u);
      for (int u = 0; u < total_units; ++u) {
        if (LA == 0) {
          if (elect_one()) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(NB - 1) : "memory");
          acquire_staging(u);
        }
        {
          int n0, w0, h0, nb0;
          unit_coords(u, n0, w0, h0, nb0);
          const int ob = u % NB;
          mbar_wait(smem_u32(&bar_out_ready[ob]), (u / NB) & 1);
          if (elect_one()) {
#pragma unroll
            for (int q = 0; q < SUB_N / 64; ++q)
              if (n0 + q * 64 < p.cout)
                tma_store_4d(&maps.d, staging + ob * Cfg::OUT_BYTES + q * (BLOCK_M * 128), n0 + q * 64, w0, h0, nb0);
            tma_store_commit();
          }
        }
        const int j = u + LA;   // its buffer was last read by the store of unit j - NB = u - 2; stores issued so far: <= u
        if (LA > 0 && j < total_units) {
          if (elect_one()) asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory");
          acquire_staging(j);
        }
      }
      if (elect_one()) tma_store_wait_all();
    }
    __syncwarp();
  } else if (warp >= 4) {
    // =========================== epilogue (per CTA, its own 128 rows) ===========================
    // One warpgroup per 128-column unit: with N = 256 the two halves of a tile are drained concurrently (warps 4-7 and
    // 8-11 both cover TMEM lane quarters 0-3), which keeps short-K layers MMA-bound instead of epilogue-bound.
    constexpr int SUBS = Cfg::SUBS, SUB_N = Cfg::SUB_N;
    const int h = (warp - 4) >> 2;
    const int epi_tid = (threadIdx.x - 128) & 127;
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t row_off = row * 128;
    const int sw = row & 7;
    for (int it = 0; it < my_tiles && h < SUBS; ++it) {
      const int b = it & 1;
      const int pt0_ = pair + it * num_pairs;
      const int ptn_ = (p.reverse & 1) ? total_pair_tiles - 1 - pt0_ : pt0_;
      const int n0 = (ptn_ - static_cast<int>(fast_div(ptn_, p.n_tiles, p.mul_n_tiles)) * p.n_tiles) * BLOCK_N;
      for (int i = epi_tid; i < SUB_N; i += 128) {
        const int c = n0 + h * SUB_N + i;
        bias_s[b % NBIAS][h * SUB_N + i] = (c < p.cout) ? p.bias[c] : 0.f;
      }
      named_bar_sync(1 + h, 128);
      mbar_wait_cluster(smem_u32(&bar_tmem_full[b]), (it >> 1) & 1);
      tc_fence_after();
      const uint32_t acc = tmem_base + b * BLOCK_N + (static_cast<uint32_t>(quarter * 32) << 16);
      {
        const int u = it * SUBS + h;
        const int ob = u % NB;
        mbar_wait(smem_u32(&bar_stage_ready[ob]), (u / NB) & 1);
        const uint32_t stage_d = staging + ob * Cfg::OUT_BYTES;
#pragma unroll 1
        for (int cc = 0; cc < SUB_N / 32; ++cc) {
          const int col = h * SUB_N + cc * 32;
          uint32_t v[32];
          tmem_ld_32x32(acc + col, v);
          tmem_ld_wait();
          const uint32_t sub = stage_d + (cc >> 1) * (BLOCK_M * 128) + row_off;
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const int chunk = (cc & 1) * 4 + g;
            const uint32_t addr = sub + ((chunk ^ sw) << 4);
            const float4 b0 = *reinterpret_cast<const float4*>(&bias_s[b % NBIAS][col + g * 8]);
            const float4 b1 = *reinterpret_cast<const float4*>(&bias_s[b % NBIAS][col + g * 8 + 4]);
            float f[8];
            f[0] = __uint_as_float(v[g * 8 + 0]) + b0.x; f[1] = __uint_as_float(v[g * 8 + 1]) + b0.y;
            f[2] = __uint_as_float(v[g * 8 + 2]) + b0.z; f[3] = __uint_as_float(v[g * 8 + 3]) + b0.w;
            f[4] = __uint_as_float(v[g * 8 + 4]) + b1.x; f[5] = __uint_as_float(v[g * 8 + 5]) + b1.y;
            f[6] = __uint_as_float(v[g * 8 + 6]) + b1.z; f[7] = __uint_as_float(v[g * 8 + 7]) + b1.w;
            if (RES) {
              uint32_t r0, r1, r2, r3;
              asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
              const float rr[8] = {bf16_lo(r0), bf16_hi(r0), bf16_lo(r1), bf16_hi(r1),
                                   bf16_lo(r2), bf16_hi(r2), bf16_lo(r3), bf16_hi(r3)};
#pragma unroll
              for (int i = 0; i < 8; ++i) f[i] = (RES == 2) ? act_t<ACT>(f[i]) + rr[i] : act_t<ACT>(f[i] + rr[i]);
            } else {
#pragma unroll
              for (int i = 0; i < 8; ++i) f[i] = act_t<ACT>(f[i]);
            }
            const uint32_t o0 = pack_bf16x2(f[0], f[1]), o1 = pack_bf16x2(f[2], f[3]);
            const uint32_t o2 = pack_bf16x2(f[4], f[5]), o3 = pack_bf16x2(f[6], f[7]);
            asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(o0), "r"(o1), "r"(o2), "r"(o3)
                         : "memory");
          }
        }
        fence_proxy_async_smem();
        mbar_arrive(smem_u32(&bar_out_ready[ob]));        // this 128-column unit may be stored
      }
      tc_fence_before();
      mbar_arrive_leader(smem_u32(&bar_tmem_empty[b]));   // release the accumulator to the leader's MMA issuer
      if (NBIAS == 1) named_bar_sync(1 + h, 128);         // this group is done reading its half of the single bias buffer
    }
  }

  tc_fence_before();
  cluster_sync_all();      // neither CTA may free TMEM or exit while the peer still uses the pair's resources
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_2cta<2 * BLOCK_N>(tmem_base);
  }
}

// ------------------------------------------------------------------------------------------------ host side

static int use_2cta();

static void pick_tile(int W, int H, int N, int* tw, int* th, int* tn) {
  // Rectangular 128-pixel tile (powers of two) with the least padding waste; ties -> wider in W, then H.
  double best = 1e30;
  int bw = 128, bh = 1, bn = 1;
  for (int w = 128; w >= 1; w >>= 1) {
    for (int h = 128 / w; h >= 1; h >>= 1) {
      const int n = 128 / (w * h);
      const double cover = double((W + w - 1) / w * w) * double((H + h - 1) / h * h) * double((N + n - 1) / n * n);
      if (cover < best - 0.5) {
        best = cover;
        bw = w; bh = h; bn = n;
      }
    }
  }
  *tw = bw; *th = bh; *tn = bn;
}

static int encode_act_view(CUtensorMap* m, const void* base, int C, int W, int H, int N, uint64_t sW, uint64_t sH,
                           uint64_t sN, int bw, int bh, int bn) {
  const uint64_t dims[4] = {uint64_t(C), uint64_t(W), uint64_t(H), uint64_t(N)};
  const uint64_t str[3] = {sW, sH, sN};
  const uint32_t box[4] = {64u, uint32_t(bw), uint32_t(bh), uint32_t(bn)};
  return encode_tmap(m, base, 4, dims, str, box, SWZ_128B);
}

int conv_prepare(const ConvSpec& s, ConvLaunch* L) {
  const View& in = s.in;
  const View& out = s.out;
  if (!in.base || !out.base || !s.w || !s.bias) return fail(B200_ERR_INVALID, "conv: null pointer");
  if (s.ksize != 1 && s.ksize != 3) return fail(B200_ERR_INVALID, "conv: ksize must be 1 or 3");
  if (s.stride != 1 && s.stride != 2) return fail(B200_ERR_INVALID, "conv: stride must be 1 or 2");
  if (in.C % 8 || out.C % 8 || in.pix_stride % 8 || out.pix_stride % 8)
    return fail(B200_ERR_INVALID, "conv: channel counts must be multiples of 8");
  const int Ho = (in.H + 2 * s.pad - s.ksize) / s.stride + 1;
  const int Wo = (in.W + 2 * s.pad - s.ksize) / s.stride + 1;
  if (Ho != out.H || Wo != out.W || in.N != out.N) return fail(B200_ERR_INVALID, "conv: output geometry mismatch");
  const bool has_res = s.res.base != nullptr;
  const bool dual = s.in2.base != nullptr;
  const bool il = s.interleave_c > 0;      // output = channel_shuffle(cat(res, conv), 2): see ConvSpec::interleave_c
  const int cout = il ? out.C / 2 : out.C;   // GEMM-N (weight rows): the conv's own channels ("slots" when interleaving)
  if (il) {
    if (s.ksize != 1 || s.stride != 1 || s.pad != 0 || dual || s.a_scale || !has_res || s.res_after_act)
      return fail(B200_ERR_INVALID, "conv: interleaving needs a plain 1x1 / stride 1 conv and the other branch as `res`");
    if (out.C % 128 || (cout != 64 && cout % 128) || (s.interleave_c & 1) || s.interleave_c > cout || s.act != ACT_RELU)
      return fail(B200_ERR_INVALID, "conv: interleaving needs 64 or a multiple of 128 slots per branch, an even channel count and ReLU");
  }
  if (s.a_scale && (s.ksize != 1 || s.stride != 1 || s.pad != 0 || dual))
    return fail(B200_ERR_INVALID, "conv: input scaling needs a plain 1x1 / stride 1 conv");
  if (dual) {
    const View& b2 = s.in2;
    if (s.ksize != 1 || s.stride != 1 || s.pad != 0) return fail(B200_ERR_INVALID, "conv: a second input needs a 1x1 / stride 1 conv");
    if (s.stride2 != 1 && s.stride2 != 2) return fail(B200_ERR_INVALID, "conv: second-input stride must be 1 or 2");
    if (in.C % BLOCK_K || b2.C % 8 || b2.pix_stride % 8)
      return fail(B200_ERR_INVALID, "conv: first input of a K-concatenated conv needs a multiple of 64 channels");
    if (b2.N != out.N || (b2.H - 1) / s.stride2 + 1 != out.H || (b2.W - 1) / s.stride2 + 1 != out.W)
      return fail(B200_ERR_INVALID, "conv: second input geometry mismatch");
    if (has_res) return fail(B200_ERR_INVALID, "conv: a K-concatenated conv takes no residual (the second input replaces it)");
  }
  if (has_res && (s.res.H != out.H || s.res.W != out.W || s.res.C != cout || s.res.N != out.N))
    return fail(B200_ERR_INVALID, "conv: residual geometry mismatch");

  ConvParams& p = L->p;
  memset(&p, 0, sizeof(p));
  memset(&L->maps, 0, sizeof(L->maps));
  const int taps = s.ksize * s.ksize;
  p.num_taps = taps;
  p.kb_per_tap = (in.C + BLOCK_K - 1) / BLOCK_K;
  p.num_kb = taps * p.kb_per_tap;
  p.kb_split = 0;
  if (dual) {
    p.kb_split = in.C / BLOCK_K;
    p.num_kb = p.kb_split + (s.in2.C + BLOCK_K - 1) / BLOCK_K;
  }
  p.cout = cout;
  p.act = s.act;
  p.has_res = il ? 3 : (has_res ? (s.res_after_act ? 2 : 1) : 0);
  p.il_group = cout / 2;
  p.il_half = s.interleave_c / 2;

  p.bias = s.bias;
  p.reverse = s.reverse_tiles ? 1 : 0;
  p.a_scale = s.a_scale;
  p.a_scale_hw = in.H * in.W;
  p.a_scale_c = in.C;
  p.a_pixels = static_cast<long long>(in.N) * in.H * in.W;
  // Tile choice.  Layers with enough arithmetic per byte (K >= 256, Cout >= 128) are bound by shared-memory operand
  // bandwidth in single-CTA mode -> CTA pairs (M = 256), N = 256 when Cout allows.  Thin layers stay single-CTA.
  const int k_total = in.C * taps + (dual ? s.in2.C : 0);
  static int res_mode = -1;   // residual layers: 0 single CTA, 1 CTA pair N=128 (deep residual prefetch), 2 CTA pair N=256
  if (res_mode < 0) {
    const char* e = getenv("B200_2CTA_RES_MODE");
    res_mode = e ? atoi(e) : 2;
  }
  static int res_min_k = -1;       // residual layers go to CTA pairs from this K on
  if (res_min_k < 0) {
    const char* e = getenv("B200_2CTA_RES_MIN_K");
    res_min_k = e ? atoi(e) : 512;
  }
  static int pair128_min_k = -1;   // Cout in [128, 256): CTA pairs with N = 128 from this K on (0 = never)
  if (pair128_min_k < 0) {
    const char* e = getenv("B200_PAIR128_MIN_K");
    pair128_min_k = e ? atoi(e) : 0;
  }
  bool two_cta = use_2cta() && ((cout >= 256 && k_total >= 256) ||
                                (cout >= 128 && pair128_min_k > 0 && k_total >= pair128_min_k));
  int block_n = two_cta ? (cout >= 256 ? 256 : 128) : (cout <= 64 ? 64 : 128);
  if (s.a_scale) two_cta = false, block_n = cout <= 64 ? 64 : 128;      // the scaler role lives in the single-CTA kernel
  if (il) two_cta = false, block_n = 64;      // the interleaving epilogue lives in the single-CTA kernel, 64 slots per tile
  if (two_cta && has_res) {
    // residual layers: CTA pairs pay off only with enough K per tile to hide the residual prefetch (layer4 conv3: K = 512)
    if (res_mode == 0 || (res_mode == 2 && k_total < res_min_k)) { two_cta = false; block_n = 128; }
    else if (res_mode == 1) block_n = 128;
  }
  static int use_halo = -1;
  if (use_halo < 0) {
    const char* e = getenv("B200_CONV_HALO");
    use_halo = (e && e[0] == '0') ? 0 : 1;
  }
  // halo kernel: 3x3/s1/p1, Cin <= 64, 16x8-pixel tiles with little padding waste; Cout <= 64, or wider while the layer
  // is too thin in K for the generic kernel (Cin < 64: each of its k-blocks would be mostly padding)
  const long long halo_cover = (long long)((out.W + 7) / 8 * 8) * ((out.H + 15) / 16 * 16);   // tiles are laid over the OUTPUT
  static int use_halo_s2 = -1;
  if (use_halo_s2 < 0) {
    const char* e = getenv("B200_CONV_HALO_S2");
    use_halo_s2 = (e && e[0] == '0') ? 0 : 1;
  }
  // stride 2 (Cin <= 32 only: a stage holds four parity planes): YOLOv8n layers 1 and 3, EfficientNetV2 features.2.0
  const bool halo_s2 = use_halo_s2 && s.stride == 2 && in.C <= 32 && in.W >= 2 && in.H >= 2;
  const bool halo = use_halo && !dual && s.ksize == 3 && (s.stride == 1 || halo_s2) && s.pad == 1 && in.C <= 64 &&
                    s.act != ACT_SIGMOID && (cout <= 64 || (in.C < 64 && cout <= 256)) &&
                    halo_cover * 100 <= (long long)out.W * out.H * 135;
  if (halo) {
    two_cta = false;
    block_n = cout <= 32 ? 32 : 64;
    p.pb = in.C <= 16 ? 32 : (in.C <= 32 ? 64 : 128);
    p.block_n = block_n;
    p.halo_s2 = s.stride == 2 ? 1 : 0;
    p.halo_ksteps = (in.C + 15) / 16;
    p.halo_stages = p.halo_s2 ? (p.pb == 64 ? 2 : 4) : (p.pb == 128 ? 3 : 4);
    p.halo_out_bufs = 8;       // as many staging buffers as fit next to the weights and the halo ring (216 KB in all)
    while (p.halo_out_bufs > 4 &&
           HaloCfg::smem_bytes(p.pb, block_n, p.halo_stages * (p.halo_s2 ? 4 : 1), p.halo_out_bufs) > 216 * 1024)
      p.halo_out_bufs -= 2;
    L->halo_smem = HaloCfg::smem_bytes(p.pb, block_n, p.halo_stages * (p.halo_s2 ? 4 : 1), p.halo_out_bufs);
  }
  L->halo = halo;
  L->block_n = block_n;
  L->two_cta = two_cta;
  p.n_tiles = (cout + block_n - 1) / block_n;

  const uint64_t ips = uint64_t(in.pix_stride) * 2, ops = uint64_t(out.pix_stride) * 2;
  const bool flat = (s.ksize == 1 && s.stride == 1 && s.pad == 0) && !(dual && s.stride2 != 1);
  int oW = out.W, oH = out.H, oN = out.N;  // logical output pixel grid used for tiling
  if (flat) {
    oW = out.N * out.H * out.W; oH = 1; oN = 1;
    p.tw = 128; p.th = 1; p.tn = 1;
    B200_TRY(encode_act_view(&L->maps.a[0], in.base, in.C, oW, 1, 1, ips, ips * oW, ips * oW, 128, 1, 1));
    p.tap_map[0] = 0; p.tap_dx[0] = 0; p.tap_dy[0] = 0;
    if (dual) {
      const uint64_t ips2 = uint64_t(s.in2.pix_stride) * 2;
      B200_TRY(encode_act_view(&L->maps.a[1], s.in2.base, s.in2.C, oW, 1, 1, ips2, ips2 * oW, ips2 * oW, 128, 1, 1));
      p.tap_map[1] = 1; p.tap_dx[1] = 0; p.tap_dy[1] = 0;
    }
  } else if (halo) {
    p.tw = 8; p.th = 16; p.tn = 1;
    const uint64_t rowb = ips * in.W, imgb = rowb * in.H;
    const uint64_t dims[4] = {uint64_t(in.C), uint64_t(in.W), uint64_t(in.H), uint64_t(in.N)};
    const uint64_t str[3] = {ips, rowb, imgb};
    const uint32_t box[4] = {uint32_t(p.pb / 2), uint32_t(HALO_PITCH_PX), uint32_t(HALO_ROWS), 1u};
    const Swizzle swz = p.pb == 128 ? SWZ_128B : (p.pb == 64 ? SWZ_64B : SWZ_32B);
    if (!p.halo_s2) {
      B200_TRY(encode_tmap(&L->maps.a[0], in.base, 4, dims, str, box, swz));
    } else {
      for (int py = 0; py < 2; ++py)
        for (int px = 0; px < 2; ++px) {      // parity plane (py, px)(i, j) = in(2i + py, 2j + px): a view with doubled strides
          const uint64_t pd[4] = {uint64_t(in.C), uint64_t((in.W - px + 1) / 2), uint64_t((in.H - py + 1) / 2), uint64_t(in.N)};
          const uint64_t ps[3] = {2 * ips, 2 * rowb, imgb};
          const char* base = static_cast<const char*>(in.base) + (uint64_t(py) * in.W + px) * ips;
          B200_TRY(encode_tmap(&L->maps.a[py * 2 + px], base, 4, pd, ps, box, swz));
        }
    }
  } else {
    pick_tile(oW, oH, oN, &p.tw, &p.th, &p.tn);
    const uint64_t rowb = ips * in.W, imgb = rowb * in.H;
    if (s.stride == 1) {
      B200_TRY(encode_act_view(&L->maps.a[0], in.base, in.C, in.W, in.H, in.N, ips, rowb, imgb, p.tw, p.th, p.tn));
    } else {
      for (int py = 0; py < 2; ++py)
        for (int px = 0; px < 2; ++px) {
          const int Wp = (in.W - px + 1) / 2, Hp = (in.H - py + 1) / 2;
          if (Wp <= 0 || Hp <= 0) continue;
          const char* base = static_cast<const char*>(in.base) + (uint64_t(py) * in.W + px) * ips;
          B200_TRY(encode_act_view(&L->maps.a[py * 2 + px], base, in.C, Wp, Hp, in.N, 2 * ips, 2 * rowb, imgb, p.tw,
                                   p.th, p.tn));
        }
    }
    if (dual) {
      // second input sampled at (stride2 * y, stride2 * x): a view with multiplied strides, tiled like the output
      const View& b2 = s.in2;
      const uint64_t ips2 = uint64_t(b2.pix_stride) * 2, rowb2 = ips2 * b2.W, imgb2 = rowb2 * b2.H;
      B200_TRY(encode_act_view(&L->maps.a[1], b2.base, b2.C, out.W, out.H, out.N, s.stride2 * ips2, s.stride2 * rowb2, imgb2,
                               p.tw, p.th, p.tn));
      p.tap_map[1] = 1; p.tap_dx[1] = 0; p.tap_dy[1] = 0;
    }
    for (int r = 0; r < s.ksize; ++r)
      for (int q = 0; q < s.ksize; ++q) {
        const int t = r * s.ksize + q;
        const int oy = r - s.pad, ox = q - s.pad;
        if (s.stride == 1) {
          p.tap_map[t] = 0; p.tap_dy[t] = (signed char)oy; p.tap_dx[t] = (signed char)ox;
        } else {
          const int py = ((oy % 2) + 2) % 2, px = ((ox % 2) + 2) % 2;
          p.tap_map[t] = (signed char)(py * 2 + px);
          p.tap_dy[t] = (signed char)((oy - py) / 2);
          p.tap_dx[t] = (signed char)((ox - px) / 2);
        }
      }
  }
  p.tiles_w = (oW + p.tw - 1) / p.tw;
  p.tiles_h = (oH + p.th - 1) / p.th;
  p.tiles_n = (oN + p.tn - 1) / p.tn;
  p.mul_n_tiles = fast_div_mul(p.n_tiles);
  p.mul_tiles_w = fast_div_mul(p.tiles_w);
  p.mul_tiles_h = fast_div_mul(p.tiles_h);
  p.mul_kb_per_tap = fast_div_mul(p.kb_per_tap);

  // output (+ residual) maps share the tile geometry
  {
    const uint64_t rowb = ops * oW, imgb = rowb * oH;
    B200_TRY(encode_act_view(&L->maps.d, out.base, out.C, oW, oH, oN, ops, rowb, imgb, p.tw, p.th, p.tn));
    if (il) {      // the other branch: a 64-channel box and the 8 channels behind it (see the kernel's DMA warp)
      const uint64_t rps = uint64_t(s.res.pix_stride) * 2;
      const uint64_t dims[4] = {uint64_t(cout), uint64_t(oW), 1u, 1u};
      const uint64_t str[3] = {rps, rps * oW, rps * oW};
      const uint32_t box[4] = {64u, 128u, 1u, 1u}, box8[4] = {8u, 128u, 1u, 1u};
      B200_TRY(encode_tmap(&L->maps.r, s.res.base, 4, dims, str, box, SWZ_128B));
      B200_TRY(encode_tmap(&L->maps.a[3], s.res.base, 4, dims, str, box8, SWZ_NONE));
    } else if (has_res) {
      const uint64_t rps = uint64_t(s.res.pix_stride) * 2;
      B200_TRY(encode_act_view(&L->maps.r, s.res.base, cout, oW, oH, oN, rps, rps * oW, rps * oW * oH, p.tw, p.th,
                               p.tn));
    }
  }
  // weights: [Cin, taps, Cout]
  {
    const uint64_t kc_min = uint64_t(in.C) + (dual ? uint64_t(s.in2.C) : 0);   // K-concatenated: one weight row [Ca + Cb]
    if ((s.w_cin > 0 && uint64_t(s.w_cin) < kc_min) || (s.w_rows > 0 && s.w_rows < cout) || s.w_cin % 8)
      return fail(B200_ERR_INVALID, "conv: padded weight extent smaller than the layer");
    const uint64_t kc = s.w_cin > 0 ? uint64_t(s.w_cin) : kc_min;
    const uint64_t dims[3] = {kc, uint64_t(taps), uint64_t(s.w_rows > 0 ? s.w_rows : cout)};
    const uint64_t str[2] = {kc * 2, kc * 2 * taps};
    const uint32_t box[3] = {halo ? uint32_t(p.pb / 2) : 64u, 1u, uint32_t(two_cta ? block_n / 2 : block_n)};  // a CTA pair loads half each
    B200_TRY(encode_tmap(&L->maps.b, s.w, 3, dims, str, box,
                         !halo || p.pb == 128 ? SWZ_128B : (p.pb == 64 ? SWZ_64B : SWZ_32B)));
  }
  const long long m_tiles = (long long)p.tiles_w * p.tiles_h * p.tiles_n;
  if (m_tiles * p.n_tiles > 0x7fffffffLL) return fail(B200_ERR_INVALID, "conv: grid too large");
  L->grid = unsigned(m_tiles * p.n_tiles);
  L->flops = 2.0 * double(out.N) * out.H * out.W * (il ? s.interleave_c : cout) * double(k_total);
  return B200_OK;
}

static int use_2cta() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200_CONV_2CTA");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v;
}

static bool use_persistent() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200_CONV_PERSISTENT");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1;
}

template <int BLOCK_N, int RES, int ACT, int DEEP = 0>
static int launch_persistent_t(const ConvLaunch& l, cudaStream_t stream) {
  static bool configured[64] = {};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  using Cfg = PConvCfg<BLOCK_N, RES, DEEP>;
  if (!configured[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(conv_gemm_persistent_kernel<BLOCK_N, RES, ACT, DEEP>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    configured[dev & 63] = true;
  }
  const int total = static_cast<int>(l.grid);
  const int ctas = total < device_sm_count() ? total : device_sm_count();
  B200_CUDA_OK(launch_pdl_if(total >= pdl_min_waves() * device_sm_count(), conv_gemm_persistent_kernel<BLOCK_N, RES, ACT, DEEP>,
                             dim3(ctas), dim3(l.p.a_scale ? PCONV_THREADS_MAX : PCONV_THREADS), Cfg::SMEM_BYTES, stream, l.maps,
                             l.p, total));
  return B200_OK;
}

static bool deep_ring(const ConvLaunch& l) {     // B200_CONV_DEEP=0 turns the deep-ring variant off
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200_CONV_DEEP");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1 && l.p.num_kb >= 8 && static_cast<long long>(l.grid) <= 8LL * device_sm_count();
}

// Thin-K layers: 2 = four ring stages + eight staging buffers (the store path was the limit when this was tuned), 3 = six
// stages + five buffers (ncu on ShuffleNetV2's 24 -> 58 conv at 56 x 56: producer 66 % of its time waiting for a free
// stage while the MMA warp waits for a full one -- load latency, not stores).  B200_CONV_THIN_MODE picks.
static int thin_mode() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200_CONV_THIN_MODE");
    v = e ? atoi(e) : 2;
  }
  return v;
}

static bool deep_staging(const ConvLaunch& l) {     // B200_CONV_THIN=0 turns the deep-staging variant off
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200_CONV_THIN");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1 && l.p.num_kb <= 2 && static_cast<long long>(l.grid) >= 8LL * device_sm_count();
}

template <int BLOCK_N, int RES>
static int launch_persistent_act(const ConvLaunch& l, cudaStream_t stream) {
  switch (l.p.act) {
    case ACT_NONE:
      if (BLOCK_N == 128 && RES != 2 && deep_ring(l)) return launch_persistent_t<BLOCK_N, RES, ACT_NONE, (BLOCK_N == 128 && RES != 2) ? 1 : 0>(l, stream);
      if (RES == 0 && deep_staging(l) && thin_mode() == 3) return launch_persistent_t<BLOCK_N, RES, ACT_NONE, RES == 0 ? 3 : 0>(l, stream);
      if (RES == 0 && deep_staging(l)) return launch_persistent_t<BLOCK_N, RES, ACT_NONE, RES == 0 ? 2 : 0>(l, stream);
      return launch_persistent_t<BLOCK_N, RES, ACT_NONE>(l, stream);
    case ACT_RELU:
      if (RES == 0 && deep_staging(l) && thin_mode() == 3) return launch_persistent_t<BLOCK_N, RES, ACT_RELU, RES == 0 ? 3 : 0>(l, stream);
      if (RES == 0 && deep_staging(l)) return launch_persistent_t<BLOCK_N, RES, ACT_RELU, RES == 0 ? 2 : 0>(l, stream);
      return launch_persistent_t<BLOCK_N, RES, ACT_RELU>(l, stream);
    case ACT_SILU:
      if (RES == 0 && deep_staging(l) && thin_mode() == 3) return launch_persistent_t<BLOCK_N, RES, ACT_SILU, RES == 0 ? 3 : 0>(l, stream);
      if (RES == 0 && deep_staging(l)) return launch_persistent_t<BLOCK_N, RES, ACT_SILU, RES == 0 ? 2 : 0>(l, stream);
      return launch_persistent_t<BLOCK_N, RES, ACT_SILU>(l, stream);
    case ACT_SIGMOID: return launch_persistent_t<BLOCK_N, RES, ACT_SIGMOID>(l, stream);
  }
  return fail(B200_ERR_INVALID, "conv: unknown activation");
}

template <int BLOCK_N>
static int launch_persistent_res(const ConvLaunch& l, cudaStream_t stream) {
  if (l.p.has_res == 3) {      // ShuffleNetV2 unit tail: conv_prepare only lets 64-slot tiles with ReLU through
    if constexpr (BLOCK_N == 64) return launch_persistent_t<64, 3, ACT_RELU>(l, stream);
    return fail(B200_ERR_INVALID, "conv: interleaving runs on 64-slot tiles");
  }
  if (l.p.has_res == 0) return launch_persistent_act<BLOCK_N, 0>(l, stream);
  if (l.p.has_res == 1) return launch_persistent_act<BLOCK_N, 1>(l, stream);
  return launch_persistent_act<BLOCK_N, 2>(l, stream);
}

template <int BLOCK_N, int RES, int ACT>
static int launch_2cta_t(const ConvLaunch& l, cudaStream_t stream) {
  static bool configured[64] = {};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  auto kern = conv_gemm_2cta_kernel<BLOCK_N, RES, ACT>;
  if (!configured[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, P2ConvCfg<BLOCK_N, RES>::SMEM_BYTES));
    configured[dev & 63] = true;
  }
  const int m_tiles = l.p.tiles_w * l.p.tiles_h * l.p.tiles_n;
  const int pair_tiles = ((m_tiles + 1) / 2) * l.p.n_tiles;
  const int max_pairs = device_sm_count() / 2;
  const int pairs = pair_tiles < max_pairs ? pair_tiles : max_pairs;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * pairs);
  cfg.blockDim = dim3(P2CONV_THREADS);
  cfg.dynamicSmemBytes = P2ConvCfg<BLOCK_N, RES>::SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = (pdl_enabled() && pair_tiles >= pdl_min_waves() * max_pairs) ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  B200_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, l.maps, l.p, pair_tiles, m_tiles));
  return B200_OK;
}

template <int BLOCK_N, int RES>
static int launch_2cta_act(const ConvLaunch& l, cudaStream_t stream) {
  switch (l.p.act) {
    case ACT_NONE: return launch_2cta_t<BLOCK_N, RES, ACT_NONE>(l, stream);
    case ACT_RELU: return launch_2cta_t<BLOCK_N, RES, ACT_RELU>(l, stream);
    case ACT_SILU: return launch_2cta_t<BLOCK_N, RES, ACT_SILU>(l, stream);
    case ACT_SIGMOID: return launch_2cta_t<BLOCK_N, RES, ACT_SIGMOID>(l, stream);
  }
  return fail(B200_ERR_INVALID, "conv: unknown activation");
}

template <int BLOCK_N>
static int launch_2cta_res(const ConvLaunch& l, cudaStream_t stream) {
  if (l.p.has_res == 0) return launch_2cta_act<BLOCK_N, 0>(l, stream);
  if (l.p.has_res == 1) return launch_2cta_act<BLOCK_N, 1>(l, stream);
  return launch_2cta_act<BLOCK_N, 2>(l, stream);
}

template <int PB, int RES, int ACT>
static int launch_halo_t(const ConvLaunch& l, cudaStream_t stream) {
  static bool configured[64] = {};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(conv3x3_halo_kernel<PB, RES, ACT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      216 * 1024));      // >= every (swizzle span, stride, stages) configuration
    configured[dev & 63] = true;
  }
  const int total = static_cast<int>(l.grid);
  int ctas = total < device_sm_count() ? total : device_sm_count();
  ctas -= ctas % l.p.n_tiles;    // a CTA keeps ONE channel tile's weights resident: tile stride must preserve tile % n_tiles
  B200_CUDA_OK(launch_pdl_if(total >= pdl_min_waves() * device_sm_count(), conv3x3_halo_kernel<PB, RES, ACT>, dim3(ctas), dim3(HALO_THREADS), size_t(l.halo_smem), stream,
                          l.maps, l.p, total));
  return B200_OK;
}

template <int PB, int RES>
static int launch_halo_act(const ConvLaunch& l, cudaStream_t stream) {
  switch (l.p.act) {
    case ACT_NONE: return launch_halo_t<PB, RES, ACT_NONE>(l, stream);
    case ACT_RELU: return launch_halo_t<PB, RES, ACT_RELU>(l, stream);
    case ACT_SILU: return launch_halo_t<PB, RES, ACT_SILU>(l, stream);
  }
  return fail(B200_ERR_INVALID, "conv: activation not supported by the halo kernel");
}

template <int PB>
static int launch_halo_res(const ConvLaunch& l, cudaStream_t stream) {
  if (l.p.has_res == 0) return launch_halo_act<PB, 0>(l, stream);
  if (l.p.has_res == 1) return launch_halo_act<PB, 1>(l, stream);
  return launch_halo_act<PB, 2>(l, stream);
}

template <int BLOCK_N>
static int launch_t(const ConvLaunch& l, cudaStream_t stream) {
  static bool configured[64] = {};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(conv_gemm_kernel<BLOCK_N>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      ConvCfg<BLOCK_N>::SMEM_BYTES));
    configured[dev & 63] = true;
  }
  conv_gemm_kernel<BLOCK_N><<<l.grid, CONV_THREADS, ConvCfg<BLOCK_N>::SMEM_BYTES, stream>>>(l.maps, l.p);
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}

int conv_launch(const ConvLaunch& l, cudaStream_t stream) {
  if (l.p.has_res == 3) return launch_persistent_res<64>(l, stream);      // the interleaving epilogue exists in this kernel only
  if (l.halo) {
    if (l.p.pb == 128) return launch_halo_res<128>(l, stream);
    if (l.p.pb == 64) return launch_halo_res<64>(l, stream);
    return launch_halo_res<32>(l, stream);
  }
  if (l.two_cta) {
    if (l.block_n == 256) return launch_2cta_res<256>(l, stream);
    return launch_2cta_res<128>(l, stream);
  }
  if (use_persistent()) {
    if (l.block_n == 64) return launch_persistent_res<64>(l, stream);
    return launch_persistent_res<128>(l, stream);
  }
  if (l.block_n == 64) return launch_t<64>(l, stream);
  return launch_t<128>(l, stream);
}

}  // namespace b200
