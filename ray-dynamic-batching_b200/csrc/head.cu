// Fully-connected heads (classifier FC, Squeeze-Excitation FCs) and the SE channel scaling.
//
// fc: out[n][j] = act(x_scale * sum_k x[n][k] * w[j][k] + b[j]) in fp32 (the classifier is kept in fp32: logits carry the
// parity bound).  A block stages 8 batch rows of x in shared memory; each warp produces 4 outputs for those 8 rows,
// lanes split K with float4 loads (coalesced over a weight row) and combine with warp shuffles.
// Replaces nn.Linear (tv resnet.py:280 `fc`, shufflenetv2.py:162, efficientnet.py:306 classifier) and the two 1x1
// convolutions of SqueezeExcitation (tv ops/misc.py `SqueezeExcitation._scale`).
#include "kernels.cuh"
#include "ptx.cuh"

namespace b200 {

static constexpr int FC_NB = 8;   // batch rows per block
static constexpr int FC_JW = 4;   // outputs per warp
static constexpr int FC_WARPS = 8;

__global__ void __launch_bounds__(FC_WARPS * 32)
fc_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b, float* __restrict__ out,
          int N, int K, int J, float x_scale, int act) {
  pdl_wait();   // launched with programmatic stream serialization: inputs are valid from here on (no early trigger:
                // a multi-wave kernel must not share its SMs with the waiting blocks of its successor)
  extern __shared__ __align__(16) float s_x[];  // [FC_NB][K]
  const int n0 = blockIdx.y * FC_NB;
  const int K4 = K >> 2;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int r = warp; r < FC_NB; r += FC_WARPS) {          // one warp per staged row: no index division
    const bool ok = n0 + r < N;
    const float4* src = reinterpret_cast<const float4*>(x + static_cast<size_t>(ok ? n0 + r : 0) * K);
    for (int c = lane; c < K4; c += 32)
      reinterpret_cast<float4*>(s_x)[r * K4 + c] = ok ? src[c] : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  __syncthreads();
  const int j0 = (blockIdx.x * FC_WARPS + warp) * FC_JW;
  if (j0 >= J) return;
  // packed f32x2 accumulators (even / odd k of every float4): the kernel is issue-bound (ncu: 30 M warp instructions,
  // 16 M of them FFMA), FFMA2 halves the math instructions; the two lanes are added before the warp reduction
  float2 acc2[FC_JW][FC_NB];
#pragma unroll
  for (int a = 0; a < FC_JW; ++a)
#pragma unroll
    for (int r = 0; r < FC_NB; ++r) acc2[a][r] = make_float2(0.f, 0.f);
  const float4* wr[FC_JW];
#pragma unroll
  for (int a = 0; a < FC_JW; ++a) wr[a] = reinterpret_cast<const float4*>(w + static_cast<size_t>(min(j0 + a, J - 1)) * K);
  for (int c = lane; c < K4; c += 32) {
    float4 wv[FC_JW];
#pragma unroll
    for (int a = 0; a < FC_JW; ++a) wv[a] = wr[a][c];
#pragma unroll
    for (int r = 0; r < FC_NB; ++r) {
      const float4 xv = reinterpret_cast<const float4*>(s_x)[r * K4 + c];
#pragma unroll
      for (int a = 0; a < FC_JW; ++a) {
        acc2[a][r] = ffma2(make_float2(xv.x, xv.y), make_float2(wv[a].x, wv[a].y), acc2[a][r]);
        acc2[a][r] = ffma2(make_float2(xv.z, xv.w), make_float2(wv[a].z, wv[a].w), acc2[a][r]);
      }
    }
  }
  float acc[FC_JW][FC_NB];
#pragma unroll
  for (int a = 0; a < FC_JW; ++a)
#pragma unroll
    for (int r = 0; r < FC_NB; ++r) acc[a][r] = acc2[a][r].x + acc2[a][r].y;
#pragma unroll
  for (int a = 0; a < FC_JW; ++a)
#pragma unroll
    for (int r = 0; r < FC_NB; ++r) {
      float v = acc[a][r];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      acc[a][r] = v;
    }
  // lane (a * FC_NB + r) writes output (j0 + a, n0 + r)
  const int a = lane / FC_NB, r = lane % FC_NB;
  float v = 0.f;
#pragma unroll
  for (int aa = 0; aa < FC_JW; ++aa)
#pragma unroll
    for (int rr = 0; rr < FC_NB; ++rr)
      if (aa == a && rr == r) v = acc[aa][rr];
  if (j0 + a < J && n0 + r < N) out[static_cast<size_t>(n0 + r) * J + j0 + a] = apply_act(v * x_scale + b[j0 + a], act);
}

int fc_launch(const float* x, const float* w, const float* b, float* out, int n, int k, int j, float x_scale, int act,
              cudaStream_t stream) {
  if (k % 4) return fail(B200_ERR_INVALID, "fc: K must be a multiple of 4");
  if (k > 6144) return fail(B200_ERR_INVALID, "fc: K too large (max 6144)");
  const size_t smem = size_t(FC_NB) * k * sizeof(float);
  static bool configured[64] = {};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(fc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FC_NB * 6144 * 4));
    configured[dev & 63] = true;
  }
  dim3 grid((j + FC_WARPS * FC_JW - 1) / (FC_WARPS * FC_JW), (n + FC_NB - 1) / FC_NB);
  B200_CUDA_OK(launch_pdl(fc_kernel, dim3(grid), dim3(FC_WARPS * 32), size_t(smem), stream, x, w, b, out, n, k, j, x_scale, act));
  return B200_OK;
}

// Squeeze-excitation gate in one launch: mean over the strips the depthwise kernel summed -> fc1 + SiLU -> fc2 +
// sigmoid.  One 1024-thread block per image; the two matrices are tiny (C x C/4 .. C/24) and every block reads them
// from L2, so the kernel is three dependent rounds of L2 latency: all loads are 16-byte and issued in unrolled batches.
// w1 fp32 [S][C] (row per squeeze unit), w2t fp32 [S][C] (fc2 TRANSPOSED so a warp reads consecutive channels).
// Replaces SqueezeExcitation._scale (tv ops/misc.py:249-254): avgpool, fc1, activation, fc2, scale_activation.
constexpr int SE_THREADS = 1024;
__global__ void __launch_bounds__(SE_THREADS)
se_gate_kernel(const float* __restrict__ partial, int P, int C, float inv_hw, const float* __restrict__ w1,
               const float* __restrict__ b1, const float* __restrict__ w2t, const float* __restrict__ b2, int S, int JS,
               float* __restrict__ gate) {
  pdl_wait();   // launched with programmatic stream serialization: inputs are valid from here on (no early trigger:
                // a multi-wave kernel must not share its SMs with the waiting blocks of its successor)
  extern __shared__ __align__(16) float se_smem[];   // mean[C], hid[S], part[JS][C]
  float* mean = se_smem;
  float* hid = se_smem + C;
  float* part = hid + S;
  const int n = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Q = C >> 2;
  for (int q = threadIdx.x; q < Q; q += SE_THREADS) {
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 8
    for (int p = 0; p < P; ++p) {      // fixed order: deterministic
      const float4 v = *reinterpret_cast<const float4*>(partial + (static_cast<size_t>(n) * P + p) * C + q * 4);
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    *reinterpret_cast<float4*>(mean + q * 4) = make_float4(s.x * inv_hw, s.y * inv_hw, s.z * inv_hw, s.w * inv_hw);
  }
  __syncthreads();
  for (int j = warp; j < S; j += SE_THREADS / 32) {
    float acc = 0.f;
    const float* wr = w1 + static_cast<size_t>(j) * C;
#pragma unroll 4
    for (int q = lane; q < Q; q += 32) {
      const float4 wv = *reinterpret_cast<const float4*>(wr + q * 4);
      const float4 mv = *reinterpret_cast<const float4*>(mean + q * 4);
      acc = fmaf(mv.x, wv.x, acc); acc = fmaf(mv.y, wv.y, acc); acc = fmaf(mv.z, wv.z, acc); acc = fmaf(mv.w, wv.w, acc);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) {
      const float z = acc + b1[j];
      hid[j] = z / (1.f + expf(-z));
    }
  }
  __syncthreads();
  // fc2: thread = (channel quad, j-slice js); slices are combined through shared memory in slice order
  for (int item = threadIdx.x; item < Q * JS; item += SE_THREADS) {
    const int q = item % Q, js = item / Q;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 8
    for (int j = js; j < S; j += JS) {
      const float4 wv = *reinterpret_cast<const float4*>(w2t + static_cast<size_t>(j) * C + q * 4);
      const float h = hid[j];
      acc.x = fmaf(h, wv.x, acc.x); acc.y = fmaf(h, wv.y, acc.y); acc.z = fmaf(h, wv.z, acc.z); acc.w = fmaf(h, wv.w, acc.w);
    }
    *reinterpret_cast<float4*>(part + static_cast<size_t>(js) * C + q * 4) = acc;
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += SE_THREADS) {
    float acc = b2[c];
    for (int js = 0; js < JS; ++js) acc += part[static_cast<size_t>(js) * C + c];
    gate[static_cast<size_t>(n) * C + c] = 1.f / (1.f + expf(-acc));
  }
}

// The same gate spread over clusters of 8 CTAs.  The one-block-per-image kernel above drags both weight matrices
// (2 * S * C * 4 bytes, 786 KB at C = 1536) through every SM behind chains of dependent L2 round trips: 8-19 us a layer,
// 30 layers.  Here CTA r of a cluster owns the channel slice [r * KS, (r + 1) * KS) of BOTH matrices, staged in shared
// memory once, and the cluster walks over groups of 8 images:
//   1. stage its slice of the group's sums (cp.async)
//   2. partial fc1 over its slice: part[g][j]                                   (a split-K of fc1 over the cluster)
//   3. cluster barrier; CTA r adds the 8 partials of image r in rank order through distributed shared memory
//      (deterministic), applies bias + SiLU and writes the hidden units into every CTA's hid[][]; cluster barrier
//   4. fc2 + sigmoid for its channel slice of the 8 images
// Each CTA moves 2 * S * KS * 4 bytes of weights (98 KB at C = 1536) once, and up to 18 clusters share the images.
constexpr int SEC_R = 8, SEC_G = 8, SEC_THREADS = 512, SEC_SMAX = 64, SEC_PAD = 4;
__global__ void __cluster_dims__(SEC_R, 1, 1) __launch_bounds__(SEC_THREADS)
se_gate_cluster_kernel(const float* __restrict__ partial, int N, int P, int C, float inv_hw, const float* __restrict__ w1,
                       const float* __restrict__ b1, const float* __restrict__ w2t, const float* __restrict__ b2, int S,
                       int KS, float* __restrict__ gate) {
  pdl_wait();
  extern __shared__ __align__(16) float sec_smem[];   // xs[8][KS + 4], w1s[S][KS + 4], w2s[S][KS], part[8][64], hid[8][64]
  const int ld = KS + SEC_PAD;
  float* xs = sec_smem;
  float* w1s = xs + SEC_G * ld;
  float* w2s = w1s + S * ld;
  float* part = w2s + S * KS;
  float* hid = part + SEC_G * SEC_SMAX;
  const uint32_t rank = cluster_ctarank();
  const int cluster = blockIdx.x / SEC_R, clusters = gridDim.x / SEC_R;
  const int k0 = rank * KS;
  const int kw = max(0, min(KS, C - k0));             // channels of this slice (multiple of 4; 0 for a surplus CTA)
  const int Q = KS >> 2, qw = kw >> 2;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  auto copy16 = [](float* dst, const float* src, bool ok) {   // !ok: no bytes are read, the 16 bytes are zero-filled
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(ok ? 16 : 0) : "memory");
  };
  for (int r = warp; r < 2 * S; r += SEC_THREADS / 32) {       // the two weight slices: once per CTA
    const bool first = r < S;
    const float* src = (first ? w1 + static_cast<size_t>(r) * C : w2t + static_cast<size_t>(r - S) * C) + k0;
    float* dst = first ? w1s + r * ld : w2s + (r - S) * KS;
    for (int q = lane; q < Q; q += 32) copy16(dst + q * 4, q < qw ? src + q * 4 : w1, q < qw);
  }
  const int groups = (N + SEC_G - 1) / SEC_G;
  for (int grp = cluster; grp < groups; grp += clusters) {
    const int g0 = grp * SEC_G;
    // ---- 1. this group's sums (row 0 of every image; further rows are added below)
    if (warp < SEC_G) {
      const bool img = g0 + warp < N;
      const float* src = partial + static_cast<size_t>(img ? g0 + warp : 0) * P * C + k0;
      for (int q = lane; q < Q; q += 32) copy16(xs + warp * ld + q * 4, q < qw ? src + q * 4 : partial, img && q < qw);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    if (P > 1) {                                       // the depthwise kernel left one row of sums per strip: add them in order
      if (warp < SEC_G && g0 + warp < N) {
        for (int q = lane; q < qw; q += 32) {
          float4 v = *reinterpret_cast<float4*>(xs + warp * ld + q * 4);
          for (int p = 1; p < P; ++p) {
            const float4 u = *reinterpret_cast<const float4*>(partial + (static_cast<size_t>(g0 + warp) * P + p) * C + k0 + q * 4);
            v.x += u.x; v.y += u.y; v.z += u.z; v.w += u.w;
          }
          *reinterpret_cast<float4*>(xs + warp * ld + q * 4) = v;
        }
      }
      __syncthreads();
    }
    // ---- 2. partial fc1: thread = (image g, unit j); a warp shares g (its x loads are broadcasts), rows of w1s are
    //         KS + 4 floats apart, so the 8 lanes of a 16-byte phase hit 8 different bank quads
    {
      const int g = threadIdx.x >> 6, jj = threadIdx.x & 63;
      float acc = 0.f;
      if (jj < S) {
        const float* xr = xs + g * ld;
        const float* wr = w1s + jj * ld;
#pragma unroll 4
        for (int q = 0; q < qw; ++q) {
          const float4 a = *reinterpret_cast<const float4*>(xr + q * 4);
          const float4 u = *reinterpret_cast<const float4*>(wr + q * 4);
          acc = fmaf(a.x, u.x, acc); acc = fmaf(a.y, u.y, acc); acc = fmaf(a.z, u.z, acc); acc = fmaf(a.w, u.w, acc);
        }
      }
      part[g * SEC_SMAX + jj] = acc;
    }
    cluster_sync_all();
    // ---- 3. CTA r: the 8 partials of image r (rank order), bias + SiLU, broadcast to every CTA's hid[][]
    if (threadIdx.x < SEC_SMAX && threadIdx.x < S) {
      const int jj = threadIdx.x;
      const uint32_t pa = smem_u32(part + rank * SEC_SMAX + jj);
      float v = 0.f;
#pragma unroll
      for (uint32_t r = 0; r < SEC_R; ++r) v += ld_cluster_f32(cluster_map(pa, r));
      const float z = v * inv_hw + b1[jj];
      const float h = z / (1.f + expf(-z));
      const uint32_t ha = smem_u32(hid + rank * SEC_SMAX + jj);
#pragma unroll
      for (uint32_t r = 0; r < SEC_R; ++r) st_cluster_f32(cluster_map(ha, r), h);
    }
    cluster_sync_all();
    // ---- 4. fc2 + sigmoid: thread = (channel quad, image), hidden units broadcast from shared memory
    for (int item = threadIdx.x; item < qw * SEC_G; item += SEC_THREADS) {
      const int g = item / qw, q = item - g * qw;
      if (g0 + g >= N) continue;
      float4 a = *reinterpret_cast<const float4*>(b2 + k0 + q * 4);
      const float* hr = hid + g * SEC_SMAX;
#pragma unroll 8
      for (int jj = 0; jj < S; ++jj) {
        const float4 wv = *reinterpret_cast<const float4*>(w2s + jj * KS + q * 4);
        const float h0 = hr[jj];
        a.x = fmaf(h0, wv.x, a.x); a.y = fmaf(h0, wv.y, a.y); a.z = fmaf(h0, wv.z, a.z); a.w = fmaf(h0, wv.w, a.w);
      }
      float4 o;
      o.x = 1.f / (1.f + expf(-a.x)); o.y = 1.f / (1.f + expf(-a.y)); o.z = 1.f / (1.f + expf(-a.z)); o.w = 1.f / (1.f + expf(-a.w));
      *reinterpret_cast<float4*>(gate + static_cast<size_t>(g0 + g) * C + k0 + q * 4) = o;
    }
    // the next group's part[] / hid[] writes are ordered behind this group's reads by the two cluster barriers: a CTA
    // writes part[] only after its own step 4, and hid[] only after every CTA has passed the first barrier of the next group
    __syncthreads();                                   // xs is restaged next
  }
}

static bool se_cluster_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200_SE_CLUSTER");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1;
}

int se_gate_launch(const float* partial, int n, int strips, int c, float inv_hw, const float* w1, const float* b1,
                   const float* w2t, const float* b2, int s, float* gate, cudaStream_t stream) {
  if (n <= 0 || strips <= 0 || c <= 0 || s <= 0) return fail(B200_ERR_INVALID, "se_gate: empty shape");
  if (c % 4) return fail(B200_ERR_INVALID, "se_gate: channels must be a multiple of 4");
  if (se_cluster_enabled() && s <= SEC_SMAX) {
    const int ks = (c + 4 * SEC_R - 1) / (4 * SEC_R) * 4;
    const size_t smem = sizeof(float) * (static_cast<size_t>(SEC_G + s) * (ks + SEC_PAD) + static_cast<size_t>(s) * ks + 2 * SEC_G * SEC_SMAX);
    if (smem <= 200 * 1024) {
      static bool configured[64] = {};
      int dev = 0;
      B200_CUDA_OK(cudaGetDevice(&dev));
      if (!configured[dev & 63]) {
        B200_CUDA_OK(cudaFuncSetAttribute(se_gate_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        configured[dev & 63] = true;
      }
      const int groups = (n + SEC_G - 1) / SEC_G, max_clusters = device_sm_count() / SEC_R;
      const int clusters = groups < max_clusters ? groups : max_clusters;
      B200_CUDA_OK(launch_pdl(se_gate_cluster_kernel, dim3(clusters * SEC_R), dim3(SEC_THREADS), size_t(smem), stream, partial, n, strips, c, inv_hw, w1, b1, w2t, b2, s, ks, gate));
      return B200_OK;
    }
  }
  int js = SE_THREADS / (c / 4);
  js = js < 1 ? 1 : (js > 8 ? 8 : js);
  if (js > s) js = s;
  const size_t smem = sizeof(float) * (static_cast<size_t>(c) * (1 + js) + s);
  if (smem > 48 * 1024) return fail(B200_ERR_INVALID, "se_gate: more than ~6k channels unsupported");
  B200_CUDA_OK(launch_pdl(se_gate_kernel, dim3(n), dim3(SE_THREADS), size_t(smem), stream, partial, strips, c, inv_hw, w1, b1, w2t, b2, s, js, gate));
  return B200_OK;
}

// Squeeze-excitation split so that no block drags a whole weight matrix through its SM (the one-launch gate reads
// 2 * S * C * 4 bytes per image: up to 786 KB per block at ~100 GB/s per SM):
//   se_hidden : block = 16 images x 16 squeeze units; hid[n][j] = silu(b1[j] + inv_hw * sum_c sums[n][c] * w1[j][c])
//   se_scale  : block = one image x 64 channels; gate = sigmoid(b2 + hid[n] . w2t[:, c]) computed on the fly, then
//               y = x * gate over the image's pixels (replaces channel_scale + the fc2 half of the gate)
constexpr int SEH_G = 16, SEH_J = 16, SEH_THREADS = 512, SEH_PAD = 4;
// block = 16 images x 16 squeeze units: their 32 rows (C floats each) are staged in shared memory in one burst of
// independent 16-byte copies, then 8 k-slices x 64 register-tiled (2 x 2) threads reduce them from shared memory.
__global__ void __launch_bounds__(SEH_THREADS)
se_hidden_kernel(const float* __restrict__ sums, int N, int P, int C, float inv_hw, const float* __restrict__ w1,
                 const float* __restrict__ b1, int S, float* __restrict__ hid) {
  pdl_wait();   // launched with programmatic stream serialization: inputs are valid from here on (no early trigger:
                // a multi-wave kernel must not share its SMs with the waiting blocks of its successor)
  extern __shared__ __align__(16) float seh_smem[];   // x[16][C + PAD], w[16][C + PAD], part[8 * 64 * 4]
  const int ld = C + SEH_PAD;
  float* xs = seh_smem;
  float* ws = seh_smem + SEH_G * ld;
  float* part = ws + SEH_J * ld;
  const int g0 = blockIdx.x * SEH_G, j0 = blockIdx.y * SEH_J;
  const int Q = C >> 2;
  // Both tiles go global -> shared with cp.async (zero-filled past the last image / unit): every 16-byte copy is in
  // flight at once, where a load + store loop waits one L2 round trip per iteration.
  for (int i = threadIdx.x; i < (SEH_G + SEH_J) * Q; i += SEH_THREADS) {
    const int r = i / Q, q = i - r * Q;
    const bool is_w = r >= SEH_G;
    const int row = is_w ? j0 + r - SEH_G : g0 + r;
    const bool ok = is_w ? row < S : row < N;
    const float* src = is_w ? w1 + static_cast<size_t>(ok ? row : 0) * C + q * 4
                            : sums + static_cast<size_t>(ok ? row : 0) * P * C + q * 4;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(seh_smem + r * ld + q * 4)), "l"(src),
                 "r"(ok ? 16 : 0)
                 : "memory");
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  if (P > 1) {                                           // more than 32 strips per image: add the remaining rows in order
    for (int i = threadIdx.x; i < SEH_G * Q; i += SEH_THREADS) {
      const int r = i / Q, q = i - r * Q;
      if (g0 + r >= N) continue;
      float4 v = *reinterpret_cast<float4*>(xs + r * ld + q * 4);
      for (int p = 1; p < P; ++p) {
        const float4 u = *reinterpret_cast<const float4*>(sums + (static_cast<size_t>(g0 + r) * P + p) * C + q * 4);
        v.x += u.x; v.y += u.y; v.z += u.z; v.w += u.w;
      }
      *reinterpret_cast<float4*>(xs + r * ld + q * 4) = v;
    }
    __syncthreads();
  }
  // thread = (k-slice 0..7, 2 x 2 micro-tile): per 16-byte step 4 shared-memory loads feed 16 FMAs, and a warp's loads
  // touch 4 image rows / 8 unit rows at a 16-byte bank offset from each other (ld = C + 4): one wavefront each
  const int ks = threadIdx.x >> 6, tile = threadIdx.x & 63;
  const int ta = tile >> 3, tj = tile & 7;
  const float* x0 = xs + (2 * ta) * ld;
  const float* x1 = x0 + ld;
  const float* w0 = ws + (2 * tj) * ld;
  const float* w1r = w0 + ld;
  const int qs = (Q + 7) >> 3;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};      // (image 2ta, unit 2tj), (2ta, 2tj+1), (2ta+1, 2tj), (2ta+1, 2tj+1)
  for (int q = ks * qs; q < min(Q, (ks + 1) * qs); ++q) {
    const float4 a0 = *reinterpret_cast<const float4*>(x0 + q * 4);
    const float4 a1 = *reinterpret_cast<const float4*>(x1 + q * 4);
    const float4 u0 = *reinterpret_cast<const float4*>(w0 + q * 4);
    const float4 u1 = *reinterpret_cast<const float4*>(w1r + q * 4);
    acc[0] = fmaf(a0.x, u0.x, acc[0]); acc[0] = fmaf(a0.y, u0.y, acc[0]); acc[0] = fmaf(a0.z, u0.z, acc[0]); acc[0] = fmaf(a0.w, u0.w, acc[0]);
    acc[1] = fmaf(a0.x, u1.x, acc[1]); acc[1] = fmaf(a0.y, u1.y, acc[1]); acc[1] = fmaf(a0.z, u1.z, acc[1]); acc[1] = fmaf(a0.w, u1.w, acc[1]);
    acc[2] = fmaf(a1.x, u0.x, acc[2]); acc[2] = fmaf(a1.y, u0.y, acc[2]); acc[2] = fmaf(a1.z, u0.z, acc[2]); acc[2] = fmaf(a1.w, u0.w, acc[2]);
    acc[3] = fmaf(a1.x, u1.x, acc[3]); acc[3] = fmaf(a1.y, u1.y, acc[3]); acc[3] = fmaf(a1.z, u1.z, acc[3]); acc[3] = fmaf(a1.w, u1.w, acc[3]);
  }
  *reinterpret_cast<float4*>(part + (ks * 64 + tile) * 4) = make_float4(acc[0], acc[1], acc[2], acc[3]);
  __syncthreads();
  if (threadIdx.x < 256) {
    const int tl = threadIdx.x >> 2, e = threadIdx.x & 3;       // micro-tile, element of it
    const int ia = 2 * (tl >> 3) + (e >> 1), ij = 2 * (tl & 7) + (e & 1);
    float v = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) v += part[(k * 64 + tl) * 4 + e];          // fixed order
    if (g0 + ia < N && j0 + ij < S) {
      const float z = v * inv_hw + b1[j0 + ij];
      hid[static_cast<size_t>(g0 + ia) * S + j0 + ij] = z / (1.f + expf(-z));
    }
  }
}

int se_hidden_launch(const float* sums, int n, int rows, int c, float inv_hw, const float* w1, const float* b1, int s,
                     float* hid, cudaStream_t stream) {
  if (n <= 0 || rows <= 0 || c <= 0 || s <= 0 || c % 4) return fail(B200_ERR_INVALID, "se_hidden: bad shape");
  const size_t smem = sizeof(float) * (static_cast<size_t>(SEH_G + SEH_J) * (c + SEH_PAD) + 8 * 64 * 4);
  if (smem > 220 * 1024) return fail(B200_ERR_INVALID, "se_hidden: more than ~1700 channels unsupported");
  static bool configured[64] = {};
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    B200_CUDA_OK(cudaFuncSetAttribute(se_hidden_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    configured[dev & 63] = true;
  }
  const dim3 grid((n + SEH_G - 1) / SEH_G, (s + SEH_J - 1) / SEH_J);
  B200_CUDA_OK(launch_pdl(se_hidden_kernel, dim3(grid), dim3(SEH_THREADS), size_t(smem), stream, sums, n, rows, c, inv_hw, w1, b1, s, hid));
  return B200_OK;
}

constexpr int SES_C = 64, SES_THREADS = 256;
__global__ void __launch_bounds__(SES_THREADS)
se_scale_kernel(const uint16_t* __restrict__ in, int in_ps, uint16_t* __restrict__ out, int out_ps, int HW, int C,
                const float* __restrict__ hid, const float* __restrict__ w2t, const float* __restrict__ b2, int S) {
  pdl_wait();   // launched with programmatic stream serialization: inputs are valid from here on (no early trigger:
                // a multi-wave kernel must not share its SMs with the waiting blocks of its successor)
  __shared__ __align__(16) float hid_s[256];
  __shared__ __align__(16) float w_s[256 / 4][SES_C];   // up to 64 squeeze units at a time
  __shared__ float part[4][SES_C];
  __shared__ __align__(16) float gate_s[SES_C];
  const int n = blockIdx.y;
  const int c0 = blockIdx.x * SES_C;
  const int cw = min(SES_C, C - c0);                    // channels of this slice (multiple of 8)
  const int c = threadIdx.x & (SES_C - 1), jg = threadIdx.x >> 6;        // 64 channels x 4 slices of the squeeze units
  float acc = 0.f;
  for (int s0 = 0; s0 < S; s0 += 64) {                  // one pass for every served model (S <= 64)
    const int sn = min(64, S - s0);
    __syncthreads();
    // hidden units and the [sn][cw] slice of fc2^T go to shared memory with cp.async: all copies in flight at once
    for (int i = threadIdx.x; i < sn * (cw >> 2); i += SES_THREADS) {
      const int j = i / (cw >> 2), q = i - j * (cw >> 2);
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(&w_s[j][q * 4])),
                   "l"(w2t + static_cast<size_t>(s0 + j) * C + c0 + q * 4)
                   : "memory");
    }
    for (int j = threadIdx.x; j < sn; j += SES_THREADS)
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(&hid_s[j])),
                   "l"(hid + static_cast<size_t>(n) * S + s0 + j)
                   : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    if (c < cw)
      for (int j = jg; j < sn; j += 4) acc = fmaf(hid_s[j], w_s[j][c], acc);
  }
  part[jg][c] = acc;
  __syncthreads();
  if (threadIdx.x < SES_C) {
    float z = (c < cw) ? b2[c0 + c] : 0.f;
#pragma unroll
    for (int g = 0; g < 4; ++g) z += part[g][c];                          // fixed order
    gate_s[c] = 1.f / (1.f + expf(-z));
  }
  __syncthreads();
  const int chunks = min(SES_C, C - c0) >> 3;                             // 16-byte chunks of this slice
  const uint16_t* ib = in + static_cast<size_t>(n) * HW * in_ps + c0;
  uint16_t* ob = out + static_cast<size_t>(n) * HW * out_ps + c0;
  for (int item = threadIdx.x; item < HW * chunks; item += SES_THREADS) {
    const int px = item / chunks, ch = item - px * chunks;
    const uint4 v = *reinterpret_cast<const uint4*>(ib + static_cast<size_t>(px) * in_ps + ch * 8);
    const float4 g0 = *reinterpret_cast<const float4*>(&gate_s[ch * 8]);
    const float4 g1 = *reinterpret_cast<const float4*>(&gate_s[ch * 8 + 4]);
    uint4 o;
    o.x = pack_bf16x2(bf16_lo(v.x) * g0.x, bf16_hi(v.x) * g0.y);
    o.y = pack_bf16x2(bf16_lo(v.y) * g0.z, bf16_hi(v.y) * g0.w);
    o.z = pack_bf16x2(bf16_lo(v.z) * g1.x, bf16_hi(v.z) * g1.y);
    o.w = pack_bf16x2(bf16_lo(v.w) * g1.z, bf16_hi(v.w) * g1.w);
    *reinterpret_cast<uint4*>(ob + static_cast<size_t>(px) * out_ps + ch * 8) = o;
  }
}

int se_scale_launch(const View& in, const View& out, const float* hid, const float* w2t, const float* b2, int s,
                    cudaStream_t stream) {
  if (in.C != out.C || in.C % 8 || in.N != out.N || in.H != out.H || in.W != out.W || in.pix_stride % 8 || out.pix_stride % 8)
    return fail(B200_ERR_INVALID, "se_scale: shape mismatch");
  if (s <= 0 || s > 256) return fail(B200_ERR_INVALID, "se_scale: squeeze width must be in [1, 256]");
  const dim3 grid((in.C + SES_C - 1) / SES_C, in.N);
  B200_CUDA_OK(launch_pdl(se_scale_kernel, dim3(grid), dim3(SES_THREADS), size_t(0), stream, static_cast<const uint16_t*>(in.base), in.pix_stride, static_cast<uint16_t*>(out.base), out.pix_stride, in.H * in.W, in.C, hid, w2t, b2, s));
  return B200_OK;
}

// y[n,h,w,c] = x[n,h,w,c] * gate[n][c]
__global__ void __launch_bounds__(256)
channel_scale_kernel(const uint16_t* __restrict__ in, int in_ps, uint16_t* __restrict__ out, int out_ps, FastDiv HW,
                     FastDiv C8d, const float* __restrict__ gate, unsigned total) {
  pdl_wait();   // launched with programmatic stream serialization: inputs are valid from here on (no early trigger:
                // a multi-wave kernel must not share its SMs with the waiting blocks of its successor)
  const unsigned idx = blockIdx.x * blockDim.x + threadIdx.x;     // 32-bit index math with multiply-high division: the
  if (idx >= total) return;                                       // 64-bit % and / cost more than the 16 bytes a thread moves
  const int C8 = static_cast<int>(C8d.d);
  const unsigned pix = fast_div(idx, C8d);
  const int c8 = static_cast<int>(idx - pix * C8d.d);
  const int n = static_cast<int>(fast_div(pix, HW));
  const uint4 v = *reinterpret_cast<const uint4*>(in + static_cast<size_t>(pix) * in_ps + c8 * 8);
  const float4 g0 = *reinterpret_cast<const float4*>(gate + static_cast<size_t>(n) * C8 * 8 + c8 * 8);
  const float4 g1 = *reinterpret_cast<const float4*>(gate + static_cast<size_t>(n) * C8 * 8 + c8 * 8 + 4);
  uint4 o;
  o.x = pack_bf16x2(bf16_lo(v.x) * g0.x, bf16_hi(v.x) * g0.y);
  o.y = pack_bf16x2(bf16_lo(v.y) * g0.z, bf16_hi(v.y) * g0.w);
  o.z = pack_bf16x2(bf16_lo(v.z) * g1.x, bf16_hi(v.z) * g1.y);
  o.w = pack_bf16x2(bf16_lo(v.w) * g1.z, bf16_hi(v.w) * g1.w);
  *reinterpret_cast<uint4*>(out + static_cast<size_t>(pix) * out_ps + c8 * 8) = o;
}

int channel_scale_launch(const View& in, const View& out, const float* gate, cudaStream_t stream) {
  if (in.C != out.C || in.C % 8 || in.N != out.N || in.H != out.H || in.W != out.W)
    return fail(B200_ERR_INVALID, "channel_scale: shape mismatch");
  const long long total = static_cast<long long>(in.N) * in.H * in.W * (in.C / 8);
  if (total >= (1ll << 31)) return fail(B200_ERR_INVALID, "channel_scale: tensor too large");
  B200_CUDA_OK(launch_pdl(channel_scale_kernel, dim3(static_cast<unsigned>((total + 255) / 256)), dim3(256), size_t(0), stream, static_cast<const uint16_t*>(in.base), in.pix_stride, static_cast<uint16_t*>(out.base), out.pix_stride, make_fast_div(in.H * in.W), make_fast_div(in.C / 8), gate, static_cast<unsigned>(total)));
  return B200_OK;
}

}  // namespace b200
