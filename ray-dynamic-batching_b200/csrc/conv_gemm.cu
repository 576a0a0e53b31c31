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

namespace b200 {

static constexpr int BLOCK_M = 128;
static constexpr int BLOCK_K = 64;
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
  const int num_kb = p.num_taps * p.kb_per_tap;

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

  if (warp == 0) {
    // =========================== TMA producer ===========================
    if (lane == 0) {
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % STAGES;
        const uint32_t ph = (kb / STAGES) & 1;
        mbar_wait(smem_u32(&bar_empty[s]), ph ^ 1);
        const uint32_t full = smem_u32(&bar_full[s]);
        mbar_expect_tx(full, Cfg::STAGE_BYTES);
        const int t = kb / p.kb_per_tap;
        const int c0 = (kb - t * p.kb_per_tap) * BLOCK_K;
        const uint32_t a_dst = ring + s * Cfg::STAGE_BYTES;
        tma_load_4d(&maps.a[p.tap_map[t]], full, a_dst, c0, w0 + p.tap_dx[t], h0 + p.tap_dy[t], nb0);
        tma_load_3d(&maps.b, full, a_dst + A_STAGE_BYTES, c0, t, n0);
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

// ------------------------------------------------------------------------------------------------ host side

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
  if (has_res && (s.res.H != out.H || s.res.W != out.W || s.res.C != out.C || s.res.N != out.N))
    return fail(B200_ERR_INVALID, "conv: residual geometry mismatch");

  ConvParams& p = L->p;
  memset(&p, 0, sizeof(p));
  memset(&L->maps, 0, sizeof(L->maps));
  const int taps = s.ksize * s.ksize;
  p.num_taps = taps;
  p.kb_per_tap = (in.C + BLOCK_K - 1) / BLOCK_K;
  p.cout = out.C;
  p.act = s.act;
  p.has_res = has_res ? (s.res_after_act ? 2 : 1) : 0;
  p.bias = s.bias;
  const int block_n = out.C <= 64 ? 64 : 128;
  L->block_n = block_n;
  p.n_tiles = (out.C + block_n - 1) / block_n;

  const uint64_t ips = uint64_t(in.pix_stride) * 2, ops = uint64_t(out.pix_stride) * 2;
  const bool flat = (s.ksize == 1 && s.stride == 1 && s.pad == 0);
  int oW = out.W, oH = out.H, oN = out.N;  // logical output pixel grid used for tiling
  if (flat) {
    oW = out.N * out.H * out.W; oH = 1; oN = 1;
    p.tw = 128; p.th = 1; p.tn = 1;
    B200_TRY(encode_act_view(&L->maps.a[0], in.base, in.C, oW, 1, 1, ips, ips * oW, ips * oW, 128, 1, 1));
    p.tap_map[0] = 0; p.tap_dx[0] = 0; p.tap_dy[0] = 0;
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

  // output (+ residual) maps share the tile geometry
  {
    const uint64_t rowb = ops * oW, imgb = rowb * oH;
    B200_TRY(encode_act_view(&L->maps.d, out.base, out.C, oW, oH, oN, ops, rowb, imgb, p.tw, p.th, p.tn));
    if (has_res) {
      const uint64_t rps = uint64_t(s.res.pix_stride) * 2;
      B200_TRY(encode_act_view(&L->maps.r, s.res.base, out.C, oW, oH, oN, rps, rps * oW, rps * oW * oH, p.tw, p.th,
                               p.tn));
    }
  }
  // weights: [Cin, taps, Cout]
  {
    const uint64_t dims[3] = {uint64_t(in.C), uint64_t(taps), uint64_t(out.C)};
    const uint64_t str[2] = {uint64_t(in.C) * 2, uint64_t(in.C) * 2 * taps};
    const uint32_t box[3] = {64u, 1u, uint32_t(block_n)};
    B200_TRY(encode_tmap(&L->maps.b, s.w, 3, dims, str, box, SWZ_128B));
  }
  const long long m_tiles = (long long)p.tiles_w * p.tiles_h * p.tiles_n;
  if (m_tiles * p.n_tiles > 0x7fffffffLL) return fail(B200_ERR_INVALID, "conv: grid too large");
  L->grid = unsigned(m_tiles * p.n_tiles);
  L->flops = 2.0 * double(out.N) * out.H * out.W * out.C * double(in.C) * taps;
  return B200_OK;
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
  if (l.block_n == 64) return launch_t<64>(l, stream);
  return launch_t<128>(l, stream);
}

}  // namespace b200
