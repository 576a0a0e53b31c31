// YOLOv8 Detect head: fused final 1x1 convolutions (fp32) + DFL + dist2bbox + stride scaling + class sigmoid, and NMS.
//
// The reference names YOLOv8n only through its image workload (sample-models/yolov8/workload); the arithmetic restated
// here is ultralytics' Detect.forward / DFL / dist2bbox / non_max_suppression (see ray-dynamic-batching_b200/yolov8n_arch.py).
//
// decode: one thread per anchor.  It reads the two hidden vectors of its pixel (64 box-branch + 80 class-branch bf16
// channels, outputs of the tensor-core 3x3 convs), applies the last 1x1 convs in fp32 with weights broadcast from shared
// memory (so the DFL logits never see bf16 rounding), turns 4 x 16 bins into distances by a softmax expectation, builds
// the xywh box around (grid + 0.5) * stride and writes [84][A] channel-major, coalesced over anchors.
#include "kernels.cuh"
#include "ptx.cuh"

namespace b200 {

static constexpr int YD_THREADS = 128;
static constexpr int REG_MAX = 16;

template <int CB, int NC>  // CB: box-branch hidden channels (= 4*REG_MAX outputs), NC: classes (= class-branch hidden channels)
__global__ void __launch_bounds__(YD_THREADS)
yolo_decode_kernel(const uint16_t* __restrict__ hb, int hb_ps, const uint16_t* __restrict__ hc, int hc_ps, int H, int W,
                   int N, const float* __restrict__ wb, const float* __restrict__ bb, const float* __restrict__ wc,
                   const float* __restrict__ bc, float* __restrict__ out, int a_off, int a_total, float stride) {
  pdl_wait();   // launched with programmatic stream serialization: inputs are valid from here on (no early trigger:
                // a multi-wave kernel must not share its SMs with the waiting blocks of its successor)
  extern __shared__ __align__(16) float s_w[];  // wb [64][CB] | wc [NC][NC] | bb [64] | bc [NC]
  float* s_wb = s_w;
  float* s_wc = s_wb + 4 * REG_MAX * CB;
  float* s_bb = s_wc + NC * NC;
  float* s_bc = s_bb + 4 * REG_MAX;
  // weights -> shared memory once per block (cp.async: every 16-byte copy in flight at once); the block then walks
  // anchor tiles with a grid stride, so the 42 KB are fetched ~600 times per launch instead of once per 128 anchors
  for (int i = threadIdx.x; i < (4 * REG_MAX * CB) / 4; i += YD_THREADS)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(s_wb + i * 4)), "l"(wb + i * 4) : "memory");
  for (int i = threadIdx.x; i < (NC * NC) / 4; i += YD_THREADS)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(s_wc + i * 4)), "l"(wc + i * 4) : "memory");
  for (int i = threadIdx.x; i < 4 * REG_MAX; i += YD_THREADS) s_bb[i] = bb[i];
  for (int i = threadIdx.x; i < NC; i += YD_THREADS) s_bc[i] = bc[i];
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  const long long total_anchors = static_cast<long long>(N) * H * W;
  for (long long idx = static_cast<long long>(blockIdx.x) * YD_THREADS + threadIdx.x; idx < total_anchors;
       idx += static_cast<long long>(gridDim.x) * YD_THREADS) {
  const int a = static_cast<int>(idx % (H * W));
  const int n = static_cast<int>(idx / (H * W));
  const int y = a / W, x = a - y * W;
  float* o = out + (static_cast<size_t>(n) * (4 + NC)) * a_total + a_off + a;

  {  // ---- box branch ----
    float h[CB];
    const uint4* src = reinterpret_cast<const uint4*>(hb + idx * hb_ps);
#pragma unroll
    for (int i = 0; i < CB / 8; ++i) {
      const uint4 v = src[i];
      h[i * 8 + 0] = bf16_lo(v.x); h[i * 8 + 1] = bf16_hi(v.x); h[i * 8 + 2] = bf16_lo(v.y); h[i * 8 + 3] = bf16_hi(v.y);
      h[i * 8 + 4] = bf16_lo(v.z); h[i * 8 + 5] = bf16_hi(v.z); h[i * 8 + 6] = bf16_lo(v.w); h[i * 8 + 7] = bf16_hi(v.w);
    }
    float dist[4];
#pragma unroll 1
    for (int s = 0; s < 4; ++s) {
      float lg[REG_MAX];
#pragma unroll
      for (int j = 0; j < REG_MAX; ++j) {
        const float4* wr = reinterpret_cast<const float4*>(s_wb + (s * REG_MAX + j) * CB);
        float acc = s_bb[s * REG_MAX + j];
#pragma unroll
        for (int k4 = 0; k4 < CB / 4; ++k4) {
          const float4 w4 = wr[k4];
          acc = fmaf(h[k4 * 4 + 0], w4.x, acc); acc = fmaf(h[k4 * 4 + 1], w4.y, acc);
          acc = fmaf(h[k4 * 4 + 2], w4.z, acc); acc = fmaf(h[k4 * 4 + 3], w4.w, acc);
        }
        lg[j] = acc;
      }
      float m = lg[0];
#pragma unroll
      for (int j = 1; j < REG_MAX; ++j) m = fmaxf(m, lg[j]);
      float den = 0.f, num = 0.f;
#pragma unroll
      for (int j = 0; j < REG_MAX; ++j) {
        const float e = expf(lg[j] - m);
        den += e;
        num += e * static_cast<float>(j);
      }
      dist[s] = num / den;
    }
    const float ax = static_cast<float>(x) + 0.5f, ay = static_cast<float>(y) + 0.5f;
    const float x1 = ax - dist[0], y1 = ay - dist[1], x2 = ax + dist[2], y2 = ay + dist[3];
    o[0] = (x1 + x2) * 0.5f * stride;
    o[static_cast<size_t>(a_total)] = (y1 + y2) * 0.5f * stride;
    o[static_cast<size_t>(2) * a_total] = (x2 - x1) * stride;
    o[static_cast<size_t>(3) * a_total] = (y2 - y1) * stride;
  }
  {  // ---- class branch ----
    float h[NC];
    const uint4* src = reinterpret_cast<const uint4*>(hc + idx * hc_ps);
#pragma unroll
    for (int i = 0; i < NC / 8; ++i) {
      const uint4 v = src[i];
      h[i * 8 + 0] = bf16_lo(v.x); h[i * 8 + 1] = bf16_hi(v.x); h[i * 8 + 2] = bf16_lo(v.y); h[i * 8 + 3] = bf16_hi(v.y);
      h[i * 8 + 4] = bf16_lo(v.z); h[i * 8 + 5] = bf16_hi(v.z); h[i * 8 + 6] = bf16_lo(v.w); h[i * 8 + 7] = bf16_hi(v.w);
    }
#pragma unroll 1
    for (int j = 0; j < NC; ++j) {
      const float4* wr = reinterpret_cast<const float4*>(s_wc + j * NC);
      float acc = s_bc[j];
#pragma unroll
      for (int k4 = 0; k4 < NC / 4; ++k4) {
        const float4 w4 = wr[k4];
        acc = fmaf(h[k4 * 4 + 0], w4.x, acc); acc = fmaf(h[k4 * 4 + 1], w4.y, acc);
        acc = fmaf(h[k4 * 4 + 2], w4.z, acc); acc = fmaf(h[k4 * 4 + 3], w4.w, acc);
      }
      o[static_cast<size_t>(4 + j) * a_total] = 1.f / (1.f + expf(-acc));
    }
  }
  }
}

// ----------------------------------------------------------------------------------------------------------------
// The same decode with the two final 1x1 convolutions on the tensor cores (warp-level mma.sync m16n8k16, bf16 x bf16 ->
// fp32).  The SIMT version above spends 10.5 k FMAs per anchor with one broadcast shared-memory load per 4 FMAs and runs
// at ~30 % of the FP32 pipe (185 us for the 80 x 80 level at batch 32).  Precision is kept by splitting every fp32
// weight into two bf16 terms, w = hi + lo (|w - hi - lo| <= 2^-17 |w|): the hidden activations are bf16 already, so
// x . hi + x . lo accumulated in fp32 differs from the fp32 product only in the last bits, and the DFL logits still
// never see a bf16 rounding.  A warp owns 32 anchors (two m16 tiles); fragments follow the PTX m16n8k16 layout:
//   A (row = anchor): a0 (g, 2t..), a1 (g+8, 2t..), a2 (g, 2t+8..), a3 (g+8, 2t+8..)   g = lane / 4, t = lane % 4
//   B (col = output): b0 (k = 2t.., n = g), b1 (k = 2t+8.., n = g)       C: c0,c1 (g, 2t, 2t+1), c2,c3 (g+8, ...)
// A side's 16 DFL bins of one anchor therefore sit in the four lanes of a quad: softmax by two xor-shuffles.
// ----------------------------------------------------------------------------------------------------------------
static constexpr int YM_THREADS = 128;

template <int CB, int NC>
__global__ void __launch_bounds__(YM_THREADS)
yolo_decode_mma_kernel(const uint16_t* __restrict__ hb, int hb_ps, const uint16_t* __restrict__ hc, int hc_ps, int H, int W,
                       int N, const float* __restrict__ wb, const float* __restrict__ bb, const float* __restrict__ wc,
                       const float* __restrict__ bc, float* __restrict__ out, int a_off, int a_total, float stride) {
  constexpr int NB = 4 * REG_MAX;            // box-branch outputs (64)
  constexpr int KPB = CB + 8, KPC = NC + 8;  // padded row lengths: a quad's 32-bit loads of 8 rows hit 32 distinct banks
  static_assert(CB % 16 == 0 && NC % 16 == 0 && NC % 8 == 0, "K must be a multiple of the MMA K");
  pdl_wait();
  extern __shared__ __align__(16) uint16_t s_y[];   // wb_hi [NB][KPB] | wb_lo | wc_hi [NC][KPC] | wc_lo | bb [NB] f32 | bc [NC] f32
  uint16_t* wbh = s_y;
  uint16_t* wbl = wbh + NB * KPB;
  uint16_t* wch = wbl + NB * KPB;
  uint16_t* wcl = wch + NC * KPC;
  float* s_bb = reinterpret_cast<float*>(wcl + NC * KPC);
  float* s_bc = s_bb + NB;
  auto split = [](float w, uint16_t& hi, uint16_t& lo) {
    const __nv_bfloat16 h = __float2bfloat16_rn(w);
    const __nv_bfloat16 l = __float2bfloat16_rn(w - __bfloat162float(h));
    hi = *reinterpret_cast<const uint16_t*>(&h);
    lo = *reinterpret_cast<const uint16_t*>(&l);
  };
  for (int i = threadIdx.x; i < NB * CB; i += YM_THREADS) split(wb[i], wbh[(i / CB) * KPB + i % CB], wbl[(i / CB) * KPB + i % CB]);
  for (int i = threadIdx.x; i < NC * NC; i += YM_THREADS) split(wc[i], wch[(i / NC) * KPC + i % NC], wcl[(i / NC) * KPC + i % NC]);
  for (int i = threadIdx.x; i < NB; i += YM_THREADS) s_bb[i] = bb[i];
  for (int i = threadIdx.x; i < NC; i += YM_THREADS) s_bc[i] = bc[i];
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const long long total_anchors = static_cast<long long>(N) * H * W;
  const int HW = H * W;
  for (long long base = (static_cast<long long>(blockIdx.x) * (YM_THREADS / 32) + warp) * 32; base < total_anchors;
       base += static_cast<long long>(gridDim.x) * (YM_THREADS / 32) * 32) {
    // the four anchor rows this lane touches: tile mt, half h -> anchor base + mt * 16 + h * 8 + g (clamped for loads)
    long long arow[2][2];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int h = 0; h < 2; ++h) arow[mt][h] = base + mt * 16 + h * 8 + g;
    auto clampa = [&](long long a) { return a < total_anchors ? a : total_anchors - 1; };

    // ---------------- box branch: 64 hidden -> 4 sides x 16 bins, DFL, dist2bbox ----------------
    {
      uint32_t afr[2][CB / 16][4];
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        const uint16_t* r0 = hb + clampa(arow[mt][0]) * hb_ps;
        const uint16_t* r1 = hb + clampa(arow[mt][1]) * hb_ps;
#pragma unroll
        for (int ks = 0; ks < CB / 16; ++ks) {
          afr[mt][ks][0] = *reinterpret_cast<const uint32_t*>(r0 + ks * 16 + 2 * t);
          afr[mt][ks][1] = *reinterpret_cast<const uint32_t*>(r1 + ks * 16 + 2 * t);
          afr[mt][ks][2] = *reinterpret_cast<const uint32_t*>(r0 + ks * 16 + 2 * t + 8);
          afr[mt][ks][3] = *reinterpret_cast<const uint32_t*>(r1 + ks * 16 + 2 * t + 8);
        }
      }
      float dist[2][2][4];      // [tile][half][side]
#pragma unroll 1
      for (int s = 0; s < 4; ++s) {
        float acc[2][2][4];     // [tile][n-tile of the side][c0..c3]
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
          const float b0 = s_bb[s * REG_MAX + nt * 8 + 2 * t], b1 = s_bb[s * REG_MAX + nt * 8 + 2 * t + 1];
#pragma unroll
          for (int mt = 0; mt < 2; ++mt) { acc[mt][nt][0] = b0; acc[mt][nt][1] = b1; acc[mt][nt][2] = b0; acc[mt][nt][3] = b1; }
        }
#pragma unroll
        for (int ks = 0; ks < CB / 16; ++ks)
#pragma unroll
          for (int nt = 0; nt < 2; ++nt) {
            const int off = (s * REG_MAX + nt * 8 + g) * KPB + ks * 16 + 2 * t;
            const uint32_t h0 = *reinterpret_cast<const uint32_t*>(wbh + off), h1 = *reinterpret_cast<const uint32_t*>(wbh + off + 8);
            const uint32_t l0 = *reinterpret_cast<const uint32_t*>(wbl + off), l1 = *reinterpret_cast<const uint32_t*>(wbl + off + 8);
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) {
              mma_bf16_16816(acc[mt][nt], afr[mt][ks], h0, h1);
              mma_bf16_16816(acc[mt][nt], afr[mt][ks], l0, l1);
            }
          }
        // softmax expectation over the side's 16 bins: this lane holds bins 2t, 2t+1 (n-tile 0) and 8+2t, 9+2t (n-tile 1)
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const float v0 = acc[mt][0][2 * h], v1 = acc[mt][0][2 * h + 1], v2 = acc[mt][1][2 * h], v3 = acc[mt][1][2 * h + 1];
            float m = fmaxf(fmaxf(v0, v1), fmaxf(v2, v3));
            m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
            m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 2));
            const float e0 = expf(v0 - m), e1 = expf(v1 - m), e2 = expf(v2 - m), e3 = expf(v3 - m);
            float den = e0 + e1 + e2 + e3;
            float num = e0 * (2 * t) + e1 * (2 * t + 1) + e2 * (8 + 2 * t) + e3 * (9 + 2 * t);
            den += __shfl_xor_sync(0xffffffffu, den, 1); num += __shfl_xor_sync(0xffffffffu, num, 1);
            den += __shfl_xor_sync(0xffffffffu, den, 2); num += __shfl_xor_sync(0xffffffffu, num, 2);
            dist[mt][h][s] = num / den;
          }
      }
      // the four lanes of a quad hold the same four distances: lane t writes output component t (cx, cy, w, h)
#pragma unroll
      for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const long long idx = arow[mt][h];
          if (idx >= total_anchors) continue;
          const int a = static_cast<int>(idx % HW), n = static_cast<int>(idx / HW);
          const int y = a / W, x = a - y * W;
          const float ax = static_cast<float>(x) + 0.5f, ay = static_cast<float>(y) + 0.5f;
          const float x1 = ax - dist[mt][h][0], y1 = ay - dist[mt][h][1], x2 = ax + dist[mt][h][2], y2 = ay + dist[mt][h][3];
          const float comp = t == 0 ? (x1 + x2) * 0.5f * stride : t == 1 ? (y1 + y2) * 0.5f * stride
                           : t == 2 ? (x2 - x1) * stride : (y2 - y1) * stride;
          out[(static_cast<size_t>(n) * (4 + NC) + t) * a_total + a_off + a] = comp;
        }
    }
    // ---------------- class branch: 80 hidden -> 80 classes, sigmoid ----------------
    {
      uint32_t afr[2][NC / 16][4];
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        const uint16_t* r0 = hc + clampa(arow[mt][0]) * hc_ps;
        const uint16_t* r1 = hc + clampa(arow[mt][1]) * hc_ps;
#pragma unroll
        for (int ks = 0; ks < NC / 16; ++ks) {
          afr[mt][ks][0] = *reinterpret_cast<const uint32_t*>(r0 + ks * 16 + 2 * t);
          afr[mt][ks][1] = *reinterpret_cast<const uint32_t*>(r1 + ks * 16 + 2 * t);
          afr[mt][ks][2] = *reinterpret_cast<const uint32_t*>(r0 + ks * 16 + 2 * t + 8);
          afr[mt][ks][3] = *reinterpret_cast<const uint32_t*>(r1 + ks * 16 + 2 * t + 8);
        }
      }
      size_t obase[2][2];
      bool live[2][2];
#pragma unroll
      for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const long long idx = clampa(arow[mt][h]);
          live[mt][h] = arow[mt][h] < total_anchors;
          obase[mt][h] = static_cast<size_t>(idx / HW) * (4 + NC) * a_total + a_off + static_cast<size_t>(idx % HW);
        }
#pragma unroll 1
      for (int nt = 0; nt < NC / 8; ++nt) {
        const float b0 = s_bc[nt * 8 + 2 * t], b1 = s_bc[nt * 8 + 2 * t + 1];
        float acc[2][4];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) { acc[mt][0] = b0; acc[mt][1] = b1; acc[mt][2] = b0; acc[mt][3] = b1; }
#pragma unroll
        for (int ks = 0; ks < NC / 16; ++ks) {
          const int off = (nt * 8 + g) * KPC + ks * 16 + 2 * t;
          const uint32_t h0 = *reinterpret_cast<const uint32_t*>(wch + off), h1 = *reinterpret_cast<const uint32_t*>(wch + off + 8);
          const uint32_t l0 = *reinterpret_cast<const uint32_t*>(wcl + off), l1 = *reinterpret_cast<const uint32_t*>(wcl + off + 8);
#pragma unroll
          for (int mt = 0; mt < 2; ++mt) {
            mma_bf16_16816(acc[mt], afr[mt][ks], h0, h1);
            mma_bf16_16816(acc[mt], afr[mt][ks], l0, l1);
          }
        }
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            if (!live[mt][h]) continue;
            float* o = out + obase[mt][h] + static_cast<size_t>(4 + nt * 8 + 2 * t) * a_total;
            o[0] = 1.f / (1.f + expf(-acc[mt][2 * h]));
            o[a_total] = 1.f / (1.f + expf(-acc[mt][2 * h + 1]));
          }
      }
    }
  }
}

int yolo_decode_launch(const View& hb, const View& hc, const float* wb, const float* bb, const float* wc, const float* bc,
                       float* out, int a_off, int a_total, float stride, cudaStream_t stream) {
  if (hb.C != 64 || hc.C != 80) return fail(B200_ERR_INVALID, "yolo_decode: expects 64 box-hidden and 80 class-hidden channels");
  if (hb.N != hc.N || hb.H != hc.H || hb.W != hc.W) return fail(B200_ERR_INVALID, "yolo_decode: branch shapes differ");
  if (a_off < 0 || a_off + hb.H * hb.W > a_total) return fail(B200_ERR_INVALID, "yolo_decode: anchor range out of bounds");
  const size_t smem = (size_t(64) * 64 + 80 * 80 + 64 + 80) * sizeof(float);
  static bool configured[64] = {};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(yolo_decode_kernel<64, 80>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    configured[dev & 63] = true;
  }
  const long long total = static_cast<long long>(hb.N) * hb.H * hb.W;
  static int use_mma = -1;
  if (use_mma < 0) {
    const char* e = getenv("B200_YOLO_MMA");
    use_mma = (e && e[0] == '0') ? 0 : 1;      // B200_YOLO_MMA=0: the SIMT version above
  }
  if (use_mma && hb.pix_stride % 2 == 0 && hc.pix_stride % 2 == 0) {
    const size_t sm = (size_t(2) * 64 * (64 + 8) + size_t(2) * 80 * (80 + 8)) * 2 + (64 + 80) * sizeof(float);
    static bool conf_mma[64] = {};
    if (!conf_mma[dev & 63]) {
      B200_CUDA_OK(cudaFuncSetAttribute(yolo_decode_mma_kernel<64, 80>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
      conf_mma[dev & 63] = true;
    }
    long long blk = (total + YM_THREADS - 1) / YM_THREADS;
    if (blk > 4LL * device_sm_count()) blk = 4LL * device_sm_count();
    B200_CUDA_OK(launch_pdl(yolo_decode_mma_kernel<64, 80>, dim3(static_cast<unsigned>(blk)), dim3(YM_THREADS), sm, stream,
                            static_cast<const uint16_t*>(hb.base), hb.pix_stride, static_cast<const uint16_t*>(hc.base),
                            hc.pix_stride, hb.H, hb.W, hb.N, wb, bb, wc, bc, out, a_off, a_total, stride));
    return B200_OK;
  }
  long long blocks = (total + YD_THREADS - 1) / YD_THREADS;
  if (blocks > 4LL * device_sm_count()) blocks = 4LL * device_sm_count();     // 4 blocks of 42 KB of weights per SM, grid-stride over anchors
  B200_CUDA_OK(launch_pdl(yolo_decode_kernel<64, 80>, dim3(static_cast<unsigned>(blocks)), dim3(YD_THREADS), size_t(smem), stream, static_cast<const uint16_t*>(hb.base), hb.pix_stride, static_cast<const uint16_t*>(hc.base), hc.pix_stride, hb.H, hb.W, hb.N, wb, bb, wc, bc, out, a_off, a_total, stride));
  return B200_OK;
}

// ----------------------------------------------------------------------------------------------------------------
// NMS: one block per image.  (1) best class per anchor, keep score > conf; (2) bitonic sort of the candidates by
// (score desc, anchor asc) in shared memory; (3) greedy suppression, class-aware via the max_wh class offset, each
// candidate checked against the kept list in parallel.  Output [max_det][6] = x1,y1,x2,y2,score,class + count.
// ----------------------------------------------------------------------------------------------------------------
static constexpr int NMS_THREADS = 1024;
static constexpr int NMS_CAP = 16384;  // >= anchors (8400 at 640x640), power of two for the bitonic network

__global__ void __launch_bounds__(NMS_THREADS)
yolo_nms_kernel(const float* __restrict__ pred, int nc, int a_total, float conf, float iou_thr, int max_det,
                float max_wh, float* __restrict__ boxes, int* __restrict__ counts) {
  extern __shared__ __align__(16) unsigned long long s_key[];  // [NMS_CAP] (score bits << 32) | ~anchor, then [NMS_CAP] class bytes
  unsigned char* s_cls = reinterpret_cast<unsigned char*>(s_key + NMS_CAP);   // best class of anchor a (candidates only)
  __shared__ int s_n;
  __shared__ float s_kept[300 * 4];  // class-shifted x1,y1,x2,y2 of the boxes kept so far
  const int n = blockIdx.x;
  const float* p = pred + static_cast<size_t>(n) * (4 + nc) * a_total;
  if (threadIdx.x == 0) s_n = 0;
  for (int i = threadIdx.x; i < NMS_CAP; i += NMS_THREADS) s_key[i] = 0ull;
  __syncthreads();
  for (int a = threadIdx.x; a < a_total; a += NMS_THREADS) {
    float best = -1.f;
    int cls = 0;
    for (int c = 0; c < nc; ++c) {       // first class attaining the best score (torch.max semantics)
      const float v = p[static_cast<size_t>(4 + c) * a_total + a];
      if (v > best) { best = v; cls = c; }
    }
    if (best > conf) {
      const int slot = atomicAdd(&s_n, 1);
      // scores are in (0,1): their float bits order like unsigned ints; ~anchor makes the lower anchor win ties
      s_key[slot] = (static_cast<unsigned long long>(__float_as_uint(best)) << 32) | (0xFFFFFFFFu - static_cast<unsigned>(a));
      s_cls[a] = static_cast<unsigned char>(cls);
    }
  }
  __syncthreads();
  const int ncand = s_n;
  int len = 1;
  while (len < ncand) len <<= 1;
  for (int k = 2; k <= len; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < len; i += NMS_THREADS) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const unsigned long long x = s_key[i], y = s_key[ixj];
          const bool desc = (i & k) == 0;  // descending overall
          if (desc ? (x < y) : (x > y)) { s_key[i] = y; s_key[ixj] = x; }
        }
      }
      __syncthreads();
    }
  // Greedy suppression by ONE warp, no block-wide barriers: each lane fetches the box of one of the next 32 candidates (the
  // global loads of a whole group are in flight together), then the group is walked in score order, every candidate
  // broadcast by shuffles and tested by the 32 lanes against the kept list in shared memory.
  if (threadIdx.x >= 32) return;
  const int lane = threadIdx.x;
  float* ob = boxes + static_cast<size_t>(n) * max_det * 6;
  const int cap = max_det < 300 ? max_det : 300;
  int nk = 0;
  for (int base = 0; base < ncand && nk < cap; base += 32) {
    float x1 = 0.f, y1 = 0.f, x2 = 0.f, y2 = 0.f, score = 0.f, off = 0.f;
    int cls = 0;
    if (base + lane < ncand) {
      const unsigned long long key = s_key[base + lane];
      const int a = static_cast<int>(0xFFFFFFFFu - static_cast<unsigned>(key & 0xFFFFFFFFull));
      score = __uint_as_float(static_cast<unsigned>(key >> 32));
      cls = s_cls[a];
      const float cx = p[a], cy = p[static_cast<size_t>(a_total) + a], w = p[static_cast<size_t>(2) * a_total + a],
                  h = p[static_cast<size_t>(3) * a_total + a];
      x1 = cx - w / 2; y1 = cy - h / 2; x2 = cx + w / 2; y2 = cy + h / 2;
      off = static_cast<float>(cls) * max_wh;
    }
    const int in_group = min(32, ncand - base);
    for (int j = 0; j < in_group && nk < cap; ++j) {
      const float bx1 = __shfl_sync(0xffffffffu, x1, j), by1 = __shfl_sync(0xffffffffu, y1, j);
      const float bx2 = __shfl_sync(0xffffffffu, x2, j), by2 = __shfl_sync(0xffffffffu, y2, j);
      const float bo = __shfl_sync(0xffffffffu, off, j);
      const float sx1 = bx1 + bo, sy1 = by1 + bo, sx2 = bx2 + bo, sy2 = by2 + bo;
      bool sup = false;
      for (int k = lane; k < nk; k += 32) {
        const float kx1 = s_kept[k * 4], ky1 = s_kept[k * 4 + 1], kx2 = s_kept[k * 4 + 2], ky2 = s_kept[k * 4 + 3];
        const float iw = fmaxf(fminf(sx2, kx2) - fmaxf(sx1, kx1), 0.f), ih = fmaxf(fminf(sy2, ky2) - fmaxf(sy1, ky1), 0.f);
        const float inter = iw * ih;
        const float iou = inter / ((sx2 - sx1) * (sy2 - sy1) + (kx2 - kx1) * (ky2 - ky1) - inter + 1e-7f);
        sup = sup || (iou > iou_thr);
      }
      if (!__any_sync(0xffffffffu, sup)) {
        if (lane == j) {
          s_kept[nk * 4] = sx1; s_kept[nk * 4 + 1] = sy1; s_kept[nk * 4 + 2] = sx2; s_kept[nk * 4 + 3] = sy2;
          ob[nk * 6] = x1; ob[nk * 6 + 1] = y1; ob[nk * 6 + 2] = x2; ob[nk * 6 + 3] = y2; ob[nk * 6 + 4] = score;
          ob[nk * 6 + 5] = static_cast<float>(cls);
        }
        ++nk;
        __syncwarp();
      }
    }
  }
  if (lane == 0) counts[n] = nk;
}

int yolo_nms_launch(const float* pred, int batch, int nc, int a_total, float conf, float iou, int max_det, float* boxes,
                    int* counts, cudaStream_t stream) {
  if (a_total > NMS_CAP) return fail(B200_ERR_INVALID, "yolo_nms: too many anchors");
  if (max_det <= 0 || max_det > 300) return fail(B200_ERR_INVALID, "yolo_nms: max_det must be in [1, 300]");
  static bool configured[64] = {};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(yolo_nms_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, NMS_CAP * 9));
    configured[dev & 63] = true;
  }
  yolo_nms_kernel<<<batch, NMS_THREADS, NMS_CAP * 9, stream>>>(pred, nc, a_total, conf, iou, max_det, 7680.f, boxes, counts);
  B200_CUDA_OK(cudaGetLastError());
  return B200_OK;
}

}  // namespace b200
