// Packed-K implicit-GEMM convolution for layers with few input channels (Cin <= 56): the EfficientNetV2 FusedMBConv
// 3x3 convolutions (24 / 48 channels at 112^2 / 56^2) and YOLOv8n's early layers (16 / 32 channels at 320^2 / 160^2).
//
// The TMA kernel spends one 64-wide k-block per filter tap, so with Cin = 24 only 37 % of every MMA is useful (and with
// Cout = 24 only 37 % of every tile).  Here GEMM-K is the *packed* concatenation of the taps, k = tap * Cin + c: 9 x 24 =
// 216 -> 4 k-blocks instead of 9.  A TMA box cannot write a fraction of a swizzled 128-byte row, so the A tile is
// gathered by the producer warps: thread m owns output pixel m of the tile and copies, per tap, Cin / 8 16-byte pieces
// (zero-filled outside the image) with cp.async into the K-major, 128-byte-swizzled layout tcgen05.mma reads.  Weights
// ([N_pad][K_pad] bf16, same packed K order) stream by TMA, one k-block per ring stage.
//
//   warps 0-3 : gather producers (cp.async, 2 k-blocks in flight)      warp 8 : TMEM alloc, weight TMA, MMA issue
//   warps 4-7 : epilogue (TMEM -> +bias (+residual) -> act -> bf16 -> 16-byte global stores, one pixel row per thread)
//
// Replaces the same torchvision Conv2d+BN+SiLU(+residual) call sites as conv_gemm.cu (tv efficientnet.py:172-231).
#include "kernels.cuh"
#include "ptx.cuh"

namespace b200 {

static constexpr int GC_THREADS = 288;
static constexpr int GC_STAGES = 4;
static constexpr int GC_LAG = 2;  // cp.async groups (k-blocks) a producer keeps in flight

struct GatherParams {
  const uint16_t* in;
  uint16_t* out;
  const uint16_t* res;
  const float* bias;
  int H, W, in_ps, Ho, Wo, out_ps, res_ps;
  int cpt;        // 16-byte chunks per tap = Cin_p / 8
  int ksize, stride, pad;
  int n_chunks;   // ksize*ksize*cpt
  int kb;         // k-blocks = ceil(n_chunks / 8)
  int cout;       // real (8-padded) output channels to store
  int act, res_mode;
  long long total_pix;
  int total_tiles;
  FastDiv Wod, Hod;     // multiply-high division by Wo and Ho (128 gather threads decompose a pixel index per tile)
};

template <int NPAD>
__global__ void __launch_bounds__(GC_THREADS, 1)
gather_conv_kernel(const __grid_constant__ CUtensorMap wmap, const __grid_constant__ GatherParams p) {
  constexpr int STAGE_BYTES = 128 * 128 + NPAD * 128;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[GC_STAGES], bar_empty[GC_STAGES], bar_t_full[2], bar_t_empty[2];
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float bias_s[NPAD];
  __shared__ int tap_tab[64];   // per 16-byte chunk of a GEMM-K row: (kernel row << 16) | (kernel col << 8) | channel chunk; -1 = padding

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t ring = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int per_cta = (p.total_tiles + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const int tile0 = static_cast<int>(blockIdx.x) * per_cta;
  const int my_tiles = max(0, min(per_cta, p.total_tiles - tile0));

  for (int i = threadIdx.x; i < NPAD; i += GC_THREADS) bias_s[i] = i < p.cout ? p.bias[i] : 0.f;
  for (int ci = threadIdx.x; ci < 64; ci += GC_THREADS) {
    int e = -1;
    if (ci < p.n_chunks) {
      const int tap = ci / p.cpt, j = ci - tap * p.cpt;
      const int r = tap / p.ksize, q = tap - r * p.ksize;
      e = (r << 16) | (q << 8) | j;
    }
    tap_tab[ci] = e;
  }
  if (warp == 8) {
    if (lane == 0) {
      tma_prefetch_desc(&wmap);
      for (int s = 0; s < GC_STAGES; ++s) {
        mbar_init(smem_u32(&bar_full[s]), 128 + 1);   // 128 gather threads + the weight TMA's expect_tx arrival
        mbar_init(smem_u32(&bar_empty[s]), 1);
      }
      for (int b = 0; b < 2; ++b) {
        mbar_init(smem_u32(&bar_t_full[b]), 1);
        mbar_init(smem_u32(&bar_t_empty[b]), 128);
      }
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc<2 * NPAD>(smem_u32(&tmem_base_s));
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  pdl_launch_dependents();
  pdl_wait();      // everything above overlapped the previous kernel's tail; activations are read from here on

  if (warp < 4) {
    // =========================== gather producers ===========================
    const int m = threadIdx.x;
    const int sw = m & 7;
    uint32_t g = 0;          // k-blocks issued
    uint32_t signalled = 0;  // k-blocks whose completion has been published
    for (int it = 0; it < my_tiles; ++it) {
      const long long pix = static_cast<long long>(tile0 + it) * 128 + m;
      const bool valid = pix < p.total_pix;
      const long long pp = valid ? pix : 0;
      const unsigned t1 = fast_div(static_cast<unsigned>(pp), p.Wod);        // total_pix < 2^31 (checked by the host)
      const int ow = static_cast<int>(static_cast<unsigned>(pp) - t1 * p.Wod.d);
      const int n = static_cast<int>(fast_div(t1, p.Hod));
      const int oh = static_cast<int>(t1 - n * p.Hod.d);
      const int y0 = oh * p.stride - p.pad, x0 = ow * p.stride - p.pad;
      const uint16_t* img = p.in + static_cast<size_t>(n) * p.H * p.W * p.in_ps;
      for (int kb = 0; kb < p.kb; ++kb, ++g) {
        const int s = g % GC_STAGES;
        mbar_wait(smem_u32(&bar_empty[s]), ((g / GC_STAGES) & 1) ^ 1);
        const uint32_t row = ring + s * STAGE_BYTES + m * 128;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const int e = tap_tab[kb * 8 + c];
          const int r = e >> 16, q = (e >> 8) & 0xff, j = e & 0xff;
          const int y = y0 + r, x = x0 + q;
          const bool ok = valid && e >= 0 && y >= 0 && y < p.H && x >= 0 && x < p.W;
          const uint16_t* src = ok ? img + (static_cast<size_t>(y) * p.W + x) * p.in_ps + j * 8 : p.in;
          asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" ::"r"(row + ((c ^ sw) << 4)), "l"(src),
                       "r"(ok ? 16u : 0u)
                       : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        if (g + 1 - signalled > GC_LAG) {   // keep GC_LAG groups in flight, publish the oldest
          asm volatile("cp.async.wait_group %0;" ::"n"(GC_LAG) : "memory");
          fence_proxy_async_smem();
          mbar_arrive(smem_u32(&bar_full[signalled % GC_STAGES]));
          ++signalled;
        }
      }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    fence_proxy_async_smem();
    for (; signalled < g; ++signalled) mbar_arrive(smem_u32(&bar_full[signalled % GC_STAGES]));
  } else if (warp < 8) {
    // =========================== epilogue ===========================
    const int quarter = warp & 3;
    const int m = quarter * 32 + lane;
    for (int it = 0; it < my_tiles; ++it) {
      const int b = it & 1;
      mbar_wait(smem_u32(&bar_t_full[b]), (it >> 1) & 1);
      tc_fence_after();
      const long long pix = static_cast<long long>(tile0 + it) * 128 + m;
      const bool valid = pix < p.total_pix;
      uint16_t* dst = p.out + pix * p.out_ps;
      const uint16_t* rsrc = p.res ? p.res + pix * p.res_ps : nullptr;
      const uint32_t acc = tmem_base + b * NPAD + (static_cast<uint32_t>(quarter * 32) << 16);
#pragma unroll 1
      for (int cc = 0; cc < NPAD / 32; ++cc) {
        uint32_t v[32];
        tmem_ld_32x32(acc + cc * 32, v);
        tmem_ld_wait();
        if (valid) {
#pragma unroll
          for (int gq = 0; gq < 4; ++gq) {
            const int c0 = cc * 32 + gq * 8;
            if (c0 < p.cout) {
              float f[8];
#pragma unroll
              for (int i = 0; i < 8; ++i) f[i] = __uint_as_float(v[gq * 8 + i]) + bias_s[c0 + i];
              float rr[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
              if (rsrc) {
                const uint4 rv = *reinterpret_cast<const uint4*>(rsrc + c0);
                rr[0] = bf16_lo(rv.x); rr[1] = bf16_hi(rv.x); rr[2] = bf16_lo(rv.y); rr[3] = bf16_hi(rv.y);
                rr[4] = bf16_lo(rv.z); rr[5] = bf16_hi(rv.z); rr[6] = bf16_lo(rv.w); rr[7] = bf16_hi(rv.w);
              }
#pragma unroll
              for (int i = 0; i < 8; ++i)
                f[i] = (p.res_mode == 2) ? apply_act(f[i], p.act) + rr[i] : apply_act(f[i] + rr[i], p.act);
              uint4 o;
              o.x = pack_bf16x2(f[0], f[1]); o.y = pack_bf16x2(f[2], f[3]);
              o.z = pack_bf16x2(f[4], f[5]); o.w = pack_bf16x2(f[6], f[7]);
              *reinterpret_cast<uint4*>(dst + c0) = o;
            }
          }
        }
      }
      tc_fence_before();
      mbar_arrive(smem_u32(&bar_t_empty[b]));
    }
  } else {
    // =========================== weight TMA + MMA issuer ===========================
    // whole warp in the loops, one elected lane issues (see elect_one in ptx.cuh)
    {
      constexpr uint32_t idesc = umma_idesc_bf16_f32(128, NPAD);
      const uint32_t total_kb = static_cast<uint32_t>(my_tiles) * p.kb;
      uint32_t wl = 0;  // weight k-blocks requested
      auto request_weights = [&]() {
        const int s = wl % GC_STAGES;
        mbar_wait(smem_u32(&bar_empty[s]), ((wl / GC_STAGES) & 1) ^ 1);
        const uint32_t full = smem_u32(&bar_full[s]);
        if (elect_one()) {
          mbar_expect_tx(full, NPAD * 128);
          tma_load_2d(&wmap, full, ring + s * STAGE_BYTES + 128 * 128, static_cast<int>(wl % p.kb) * 64, 0);
        }
        ++wl;
      };
      for (int i = 0; i < GC_STAGES - 1 && wl < total_kb; ++i) request_weights();
      uint32_t g = 0;
      for (int it = 0; it < my_tiles; ++it) {
        const int b = it & 1;
        mbar_wait(smem_u32(&bar_t_empty[b]), ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        for (int kb = 0; kb < p.kb; ++kb, ++g) {
          const int s = g % GC_STAGES;
          mbar_wait(smem_u32(&bar_full[s]), (g / GC_STAGES) & 1);
          tc_fence_after();
          const uint32_t a_src = ring + s * STAGE_BYTES;
          const uint64_t adesc = umma_desc_sw128(a_src);
          const uint64_t bdesc = umma_desc_sw128(a_src + 128 * 128);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_bf16(tmem_base + b * NPAD, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
            umma_commit(smem_u32(&bar_empty[s]));
            if (kb == p.kb - 1) umma_commit(smem_u32(&bar_t_full[b]));
          }
          if (wl < total_kb) request_weights();   // refill the ring (waits for the slot freed GC_STAGES-1 blocks ago)
        }
      }
    }
    __syncwarp();
  }

  __syncthreads();
  if (warp == 8) {
    tc_fence_after();
    tmem_dealloc<2 * NPAD>(tmem_base);
  }
}

int gather_conv_npad(int cout) { return cout <= 32 ? 32 : (cout <= 64 ? 64 : (cout <= 128 ? 128 : 256)); }
int gather_conv_kpad(int cin_p, int ksize) { return (ksize * ksize * (cin_p / 8) + 7) / 8 * 64; }

template <int NPAD>
static int gather_launch_t(const CUtensorMap& wmap, const GatherParams& p, cudaStream_t stream) {
  const size_t smem = size_t(GC_STAGES) * (128 * 128 + NPAD * 128) + 1024;
  static bool configured[64] = {};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(gather_conv_kernel<NPAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    configured[dev & 63] = true;
  }
  const int ctas = p.total_tiles < device_sm_count() ? p.total_tiles : device_sm_count();
  B200_CUDA_OK(launch_pdl(gather_conv_kernel<NPAD>, dim3(ctas), dim3(GC_THREADS), size_t(smem), stream, wmap, p));
  return B200_OK;
}

// weights: bf16 [gather_conv_npad(cout)][gather_conv_kpad(cin_p, k)], GEMM-K index = tap * cin_p + c; bias fp32 [cout].
int gather_conv_launch(const View& in, const View& out, const View& res, const void* w, const float* bias, int ksize,
                       int stride, int pad, int act, bool res_after_act, cudaStream_t stream) {
  if (in.C % 8 || ksize * ksize * (in.C / 8) > 64 || in.pix_stride % 8)
    return fail(B200_ERR_INVALID, "gather_conv: cin must be a multiple of 8 with k*k*cin <= 512");
  if (out.C % 8 || out.C > 256 || out.pix_stride % 8) return fail(B200_ERR_INVALID, "gather_conv: cout must be a multiple of 8, <= 256");
  if (ksize != 1 && ksize != 3) return fail(B200_ERR_INVALID, "gather_conv: ksize must be 1 or 3");
  if ((in.H + 2 * pad - ksize) / stride + 1 != out.H || (in.W + 2 * pad - ksize) / stride + 1 != out.W || in.N != out.N)
    return fail(B200_ERR_INVALID, "gather_conv: output geometry mismatch");
  if (res.base && (res.C != out.C || res.H != out.H || res.W != out.W || res.N != out.N || res.pix_stride % 8))
    return fail(B200_ERR_INVALID, "gather_conv: residual geometry mismatch");
  const int npad = gather_conv_npad(out.C);
  const int kpad = gather_conv_kpad(in.C, ksize);
  CUtensorMap wmap;
  const uint64_t dims[2] = {uint64_t(kpad), uint64_t(npad)};
  const uint64_t str[1] = {uint64_t(kpad) * 2};
  const uint32_t box[2] = {64u, uint32_t(npad)};
  B200_TRY(encode_tmap(&wmap, w, 2, dims, str, box, SWZ_128B));
  GatherParams p;
  p.in = static_cast<const uint16_t*>(in.base);
  p.out = static_cast<uint16_t*>(out.base);
  p.res = static_cast<const uint16_t*>(res.base);
  p.bias = bias;
  p.H = in.H; p.W = in.W; p.in_ps = in.pix_stride; p.Ho = out.H; p.Wo = out.W; p.out_ps = out.pix_stride;
  p.res_ps = res.base ? res.pix_stride : 0;
  p.cpt = in.C / 8;
  p.ksize = ksize; p.stride = stride; p.pad = pad;
  p.n_chunks = ksize * ksize * p.cpt;
  p.kb = (p.n_chunks + 7) / 8;
  p.cout = out.C;
  p.act = act;
  p.res_mode = res.base ? (res_after_act ? 2 : 1) : 0;
  p.total_pix = static_cast<long long>(out.N) * out.H * out.W;
  if (p.total_pix >= (1ll << 31) - 128) return fail(B200_ERR_INVALID, "gather_conv: tensor too large");
  p.total_tiles = static_cast<int>((p.total_pix + 127) / 128);
  p.Wod = make_fast_div(out.W);
  p.Hod = make_fast_div(out.H);
  if (npad == 32) return gather_launch_t<32>(wmap, p, stream);
  if (npad == 64) return gather_launch_t<64>(wmap, p, stream);
  if (npad == 128) return gather_launch_t<128>(wmap, p, stream);
  return gather_launch_t<256>(wmap, p, stream);
}

}  // namespace b200
