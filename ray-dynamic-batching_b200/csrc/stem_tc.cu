// Tensor-core network stem: stride-2 KxK convolution (K = 7, pad 3 or K = 3, pad 1) with 3 input channels + folded
// BatchNorm + ReLU/SiLU, as an implicit GEMM whose A tile is built in shared memory by the producer warps.
//
// Why not TMA: with C padded to 4 a kernel row of one output pixel is K*4 contiguous bf16 (56 or 24 bytes) -- far below a
// TMA box row -- but an *aligned* window of 8 (K=7) or 4 (K=3) input pixels starting at the even x = 2*ow - pad - 1 is
// 64 / 32 contiguous, 16-byte aligned bytes.  So GEMM-K is laid out as (kernel row, window pixel, channel): 32 (16)
// elements per kernel row, window pixel 0 and channel 3 carry zero weights.  Each producer thread owns one output pixel
// (one row of the 128 x K_pad A tile) and fills it with 16-byte cp.async copies (zero-filled outside the image) straight
// into the 128-byte-swizzled K-major layout tcgen05.mma reads; consecutive threads read consecutive 16-byte pieces of
// the input row, i.e. fully coalesced.  Weights ([Cout][K_pad] bf16, same K layout) are TMA-loaded once per CTA.
//
//   warps 0-3 : im2col producers (tile it+1 is in flight while tile it is consumed; two A buffers)
//   warps 4-7 : epilogue  (TMEM -> +bias -> act -> bf16 -> 16-byte global stores, one pixel row per thread)
//   warp  8   : TMEM allocation, weight load, single-thread tcgen05.mma issue (two TMEM accumulators)
//
// Replaces tv resnet.py:268-270 (conv1 7x7/2 + bn1 + relu), shufflenetv2.py:112-116, efficientnet.py:283-287 and YOLOv8
// layer 0 as reached from scheduler.py:450.
#include "kernels.cuh"
#include "ptx.cuh"

namespace b200 {

static constexpr int ST_THREADS = 288;

template <int K>
struct StemGeo {
  static constexpr int WP = (K == 7) ? 8 : 4;               // window pixels per kernel row
  static constexpr int ROW_ELEMS = WP * 4;                  // GEMM-K elements per kernel row
  static constexpr int ROWS_PER_KB = 64 / ROW_ELEMS;        // kernel rows per 64-wide k-block (2 or 4)
  static constexpr int KB = (K + ROWS_PER_KB - 1) / ROWS_PER_KB;  // k-blocks (4 or 1)
  static constexpr int A_BYTES = KB * 128 * 128;            // one A buffer
  static constexpr int PAIRS = WP / 2;                      // 16-byte copies per kernel row
};

__device__ __forceinline__ void cp_async16_zfill(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}

template <int K, int NOUT>  // NOUT: UMMA N (16, 32 or 64) >= cout
__global__ void __launch_bounds__(ST_THREADS, 1)
stem_tc_kernel(const __grid_constant__ CUtensorMap wmap, const uint16_t* __restrict__ in, int H, int W,
               uint16_t* __restrict__ out, int Ho, int Wo, int out_ps, int cout, const float* __restrict__ bias, int pad,
               int act, long long total_pix, int total_tiles) {
  using G = StemGeo<K>;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_a_full[2], bar_a_empty[2], bar_t_full[2], bar_t_empty[2], bar_w;
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float bias_s[NOUT];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_buf[2] = {base, base + G::A_BYTES};
  const uint32_t w_buf = base + 2 * G::A_BYTES;
  // contiguous tile range per CTA: consecutive tiles are vertically adjacent output rows, whose 7-row input windows
  // overlap -> the re-reads hit this SM's L1 (cp.async.ca) instead of L2
  const int per_cta = (total_tiles + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const int tile0 = static_cast<int>(blockIdx.x) * per_cta;
  const int my_tiles = max(0, min(per_cta, total_tiles - tile0));

  // zero both A buffers once: padded K positions (zero kernel row, tail) are never written again
  for (uint32_t i = threadIdx.x; i < 2 * G::A_BYTES / 16; i += ST_THREADS)
    asm volatile("st.shared.v4.u32 [%0], {%1, %1, %1, %1};" ::"r"(base + i * 16), "r"(0u) : "memory");
  if (threadIdx.x < NOUT) bias_s[threadIdx.x] = threadIdx.x < cout ? bias[threadIdx.x] : 0.f;
  if (warp == 8) {
    if (lane == 0) {
      tma_prefetch_desc(&wmap);
      for (int b = 0; b < 2; ++b) {
        mbar_init(smem_u32(&bar_a_full[b]), 128);
        mbar_init(smem_u32(&bar_a_empty[b]), 1);
        mbar_init(smem_u32(&bar_t_full[b]), 1);
        mbar_init(smem_u32(&bar_t_empty[b]), 128);
      }
      mbar_init(smem_u32(&bar_w), 1);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc<(2 * NOUT < 32 ? 32 : 2 * NOUT)>(smem_u32(&tmem_base_s));
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  pdl_launch_dependents();
  pdl_wait();      // everything above overlapped the previous kernel's tail; activations are read from here on

  if (warp < 4) {
    // =========================== im2col producers ===========================
    const int m = threadIdx.x;  // A-tile row
    const int sw = m & 7;
    auto issue = [&](int it) {
      const long long pix = static_cast<long long>(tile0 + it) * 128 + m;
      const bool valid = pix < total_pix;
      const long long pp = valid ? pix : 0;
      const int ow = static_cast<int>(pp % Wo);
      const long long t = pp / Wo;
      const int oh = static_cast<int>(t % Ho);
      const int n = static_cast<int>(t / Ho);
      const int y0 = 2 * oh - pad, x0 = 2 * ow - pad - 1;  // x0 is even
      const uint32_t row = a_buf[it & 1] + m * 128;
#pragma unroll
      for (int r = 0; r < K; ++r) {
        const int y = y0 + r;
        const bool yok = valid && y >= 0 && y < H;
        const uint16_t* src_row = in + (static_cast<size_t>(n) * H + (yok ? y : 0)) * W * 4;
        const uint32_t kb_off = (r / G::ROWS_PER_KB) * (128 * 128);
        const int c_base = (r % G::ROWS_PER_KB) * G::PAIRS;
#pragma unroll
        for (int j = 0; j < G::PAIRS; ++j) {
          const int x = x0 + 2 * j;
          const bool ok = yok && x >= 0 && x < W;  // W even, x even: the pair is entirely inside or outside
          cp_async16_zfill(row + kb_off + (((c_base + j) ^ sw) << 4), src_row + (ok ? x : 0) * 4, ok ? 16u : 0u);
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    if (my_tiles > 0) issue(0);
    for (int it = 0; it < my_tiles; ++it) {
      if (it + 1 < my_tiles) {
        const int nb = (it + 1) & 1;
        mbar_wait(smem_u32(&bar_a_empty[nb]), (((it + 1) >> 1) & 1) ^ 1);  // MMAs of tile it-1 have read this buffer
        issue(it + 1);
        asm volatile("cp.async.wait_group 1;" ::: "memory");
      } else {
        asm volatile("cp.async.wait_group 0;" ::: "memory");
      }
      fence_proxy_async_smem();  // cp.async (generic proxy) writes -> visible to tcgen05.mma (async proxy)
      mbar_arrive(smem_u32(&bar_a_full[it & 1]));
    }
  } else if (warp < 8) {
    // =========================== epilogue ===========================
    const int quarter = warp & 3;
    const int m = quarter * 32 + lane;
    for (int it = 0; it < my_tiles; ++it) {
      const int b = it & 1;
      mbar_wait(smem_u32(&bar_t_full[b]), (it >> 1) & 1);
      tc_fence_after();
      const long long pix = static_cast<long long>(tile0 + it) * 128 + m;
      uint16_t* dst = out + pix * out_ps;
      const uint32_t acc = tmem_base + b * NOUT + (static_cast<uint32_t>(quarter * 32) << 16);
      constexpr int CH = NOUT < 32 ? NOUT : 32;
#pragma unroll
      for (int cc = 0; cc < NOUT / CH; ++cc) {
        uint32_t v[32];
        if (CH == 32) {
          tmem_ld_32x32(acc + cc * 32, v);
        } else {
          asm volatile(
              "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
              : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
                "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
              : "r"(acc)
              : "memory");
        }
        tmem_ld_wait();
        if (pix < total_pix) {
#pragma unroll
          for (int g = 0; g < CH / 8; ++g) {
            const int c0 = cc * CH + g * 8;
            if (c0 < cout) {
              float f[8];
#pragma unroll
              for (int i = 0; i < 8; ++i) f[i] = apply_act(__uint_as_float(v[g * 8 + i]) + bias_s[c0 + i], act);
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
    // =========================== weights + MMA issuer ===========================
    // whole warp in the loop, one elected lane issues (see elect_one in ptx.cuh)
    if (elect_one()) {
      mbar_expect_tx(smem_u32(&bar_w), G::KB * NOUT * 128);
      for (int kb = 0; kb < G::KB; ++kb) tma_load_2d(&wmap, smem_u32(&bar_w), w_buf + kb * (NOUT * 128), kb * 64, 0);
    }
    mbar_wait(smem_u32(&bar_w), 0);
    constexpr uint32_t idesc = umma_idesc_bf16_f32(128, NOUT);
    for (int it = 0; it < my_tiles; ++it) {
      const int b = it & 1;
      mbar_wait(smem_u32(&bar_t_empty[b]), ((it >> 1) & 1) ^ 1);
      mbar_wait(smem_u32(&bar_a_full[b]), (it >> 1) & 1);
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int kb = 0; kb < G::KB; ++kb) {
          const uint64_t adesc = umma_desc_sw128(a_buf[b] + kb * (128 * 128));
          const uint64_t bdesc = umma_desc_sw128(w_buf + kb * (NOUT * 128));
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16(tmem_base + b * NOUT, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
        }
        umma_commit(smem_u32(&bar_a_empty[b]));
        umma_commit(smem_u32(&bar_t_full[b]));
      }
    }
    __syncwarp();
  }

  __syncthreads();
  if (warp == 8) {
    tc_fence_after();
    tmem_dealloc<(2 * NOUT < 32 ? 32 : 2 * NOUT)>(tmem_base);
  }
}

int stem_tc_kpad(int ksize) { return ksize == 7 ? StemGeo<7>::KB * 64 : StemGeo<3>::KB * 64; }
int stem_tc_nout(int cout) { return cout <= 16 ? 16 : (cout <= 32 ? 32 : 64); }

template <int K, int NOUT>
static int stem_tc_launch_t(const View& in, const View& out, const void* w, const float* bias, int pad, int act,
                            cudaStream_t stream) {
  using G = StemGeo<K>;
  CUtensorMap wmap;
  const uint64_t dims[2] = {uint64_t(G::KB * 64), uint64_t(NOUT)};
  const uint64_t str[1] = {uint64_t(G::KB * 64) * 2};
  const uint32_t box[2] = {64u, uint32_t(NOUT)};
  B200_TRY(encode_tmap(&wmap, w, 2, dims, str, box, SWZ_128B));
  const size_t smem = 2 * G::A_BYTES + G::KB * NOUT * 128 + 1024;
  static bool configured[64] = {};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(stem_tc_kernel<K, NOUT>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    configured[dev & 63] = true;
  }
  const long long total_pix = static_cast<long long>(out.N) * out.H * out.W;
  const int tiles = static_cast<int>((total_pix + 127) / 128);
  const int ctas = tiles < device_sm_count() ? tiles : device_sm_count();
  stem_tc_kernel<K, NOUT><<<ctas, ST_THREADS, smem, stream>>>(wmap, static_cast<const uint16_t*>(in.base), in.H, in.W,
                                                             static_cast<uint16_t*>(out.base), out.H, out.W, out.pix_stride,
                                                             out.C, bias, pad, act, total_pix, tiles);
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}

// weights: bf16 [stem_tc_nout(cout)][stem_tc_kpad(k)] with GEMM-K index = r * ROW_ELEMS + s' * 4 + c
// (r kernel row, s' = s + 1 window pixel, c channel); bias fp32 [cout].
int stem_tc_launch(const View& in, const View& out, const void* w, const float* bias, int ksize, int stride, int pad,
                   int act, cudaStream_t stream) {
  if (in.C != 4 || in.pix_stride != 4) return fail(B200_ERR_INVALID, "stem: input must be NHWC with C padded to 4");
  if (stride != 2 || !((ksize == 7 && pad == 3) || (ksize == 3 && pad == 1)))
    return fail(B200_ERR_INVALID, "stem_tc: supports 7x7/2 pad 3 and 3x3/2 pad 1");
  if (in.W % 2) return fail(B200_ERR_INVALID, "stem_tc: input width must be even");
  if (out.C % 8 || out.pix_stride % 8 || out.C > 64) return fail(B200_ERR_INVALID, "stem_tc: cout must be a multiple of 8, <= 64");
  if ((in.H + 2 * pad - ksize) / 2 + 1 != out.H || (in.W + 2 * pad - ksize) / 2 + 1 != out.W || in.N != out.N)
    return fail(B200_ERR_INVALID, "stem: output geometry mismatch");
  const int nout = stem_tc_nout(out.C);
  if (ksize == 7) {
    if (nout == 64) return stem_tc_launch_t<7, 64>(in, out, w, bias, pad, act, stream);
    if (nout == 32) return stem_tc_launch_t<7, 32>(in, out, w, bias, pad, act, stream);
    return stem_tc_launch_t<7, 16>(in, out, w, bias, pad, act, stream);
  }
  if (nout == 64) return stem_tc_launch_t<3, 64>(in, out, w, bias, pad, act, stream);
  if (nout == 32) return stem_tc_launch_t<3, 32>(in, out, w, bias, pad, act, stream);
  return stem_tc_launch_t<3, 16>(in, out, w, bias, pad, act, stream);
}

}  // namespace b200
