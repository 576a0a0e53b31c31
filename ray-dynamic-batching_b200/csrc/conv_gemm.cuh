// Host interface of the tcgen05 implicit-GEMM convolution (1x1 / 3x3, stride 1 / 2) on NHWC bf16.
#pragma once
#include "common.h"

namespace b200 {

struct ConvSpec {
  View in, out;
  View res;                  // optional residual (base == nullptr -> none); same geometry as `out`
  const float* a_scale = nullptr;   // optional per-(image, input channel) fp32 factor [N][Cin] applied to the input (SE
                                    //   excitation folded into the project conv); 1x1 / stride 1 only
  View in2;                  // optional second input of a K-concatenated 1x1 conv (base == nullptr -> none): the output is
  int stride2 = 1;           //   act(W[:, :Ca] . in + W[:, Ca:] . in2[::stride2, ::stride2] + bias), w = [Cout][Ca + Cb]
  const void* w = nullptr;   // bf16 [Cout][k*k][Cin], BatchNorm already folded in
  int w_cin = 0, w_rows = 0;  // physical extent of `w`: [w_rows >= Cout][k*k][w_cin >= Cin (+ Cin2)], zero padded; 0 = exact.
                              //   Measured on B200: a weight box that is partly out of bounds of its tensor map (zero-filled
                              //   by TMA) loads about twice as slowly, and a thin layer loads one per 128-pixel tile
                              //   (24 -> 58 1x1 at 56 x 56 x 512: 99 -> 55 us with the weights padded to the 64 x 64 box).
                              //   Input boxes clipped the same way, clipped stores and residual boxes cost nothing measurable.
  const float* bias = nullptr;  // fp32 [Cout]
  int ksize = 1, stride = 1, pad = 0;
  int act = 0;               // b200::Act
  bool res_after_act = false;
  bool reverse_tiles = false;  // false: act(acc + bias + res) (ResNet); true: act(acc + bias) + res (FusedMBConv e=1)
  // ShuffleNetV2 unit tail (tv models/shufflenetv2.py:24-38,96-103: cat + channel_shuffle(groups=2)) folded into the 1x1
  // conv that ends branch2: interleave_c = c > 0 logical channels per branch.  `res` is then the OTHER branch (x1, c
  // channels in logical order, res.C = cp physical) and `out` the unit's whole output in halves layout [cp | cp]
  // (out.C = 2 cp): logical channel l = 2 i + e (e = 0: x1[i], e = 1: this conv's channel i) lives at l for l < c and at
  // cp + (l - c) above.  The weights hold cp "slots": channel i < c/2 in slot i, channel i >= c/2 in slot cp/2 + i - c/2,
  // the others zero, so that slot s goes to physical channel 2 s + 1 and the x1 value beside it to 2 s.
  int interleave_c = 0;
};

struct alignas(64) ConvMaps {
  CUtensorMap a[4];  // input views (index = parity (py*2+px) for stride 2, else 0; a[3] of an interleaving conv: see r)
  CUtensorMap b;     // weights [Cin, taps, Cout]
  CUtensorMap d;     // output
  CUtensorMap r;     // residual (aliases d's geometry)
};

struct ConvParams {
  int tiles_w, tiles_h, tiles_n;  // M-tile grid (pixels)
  int tw, th, tn;                 // pixels per tile along W, H, N  (tw*th*tn == 128)
  int n_tiles;                    // tiles along Cout
  unsigned mul_n_tiles, mul_tiles_w, mul_tiles_h, mul_kb_per_tap;   // fast_div multipliers of the four divisors above
  int num_taps, kb_per_tap;
  int num_kb;                     // k-blocks per tile (num_taps * kb_per_tap, or both inputs' blocks of a K-concatenated conv)
  int kb_split;                   // K-concatenated conv: k-blocks [0, kb_split) come from input 0, the rest from input 1; 0 = off
  int cout, act, has_res;  // has_res: 0 none, 1 add before the activation, 2 add after it, 3 interleave with `res` (cout = slots)
  int il_group, il_half;   // interleave: slots per half (cp / 2), valid channels per half-group (c / 2)
  int pb;                  // halo kernel: bytes per pixel of the staged halo (32 / 64 / 128 = swizzle span), K per tap = pb / 2
  int block_n;             // halo kernel: accumulator width (32 or 64)
  int halo_stages;         // halo kernel: halo boxes in flight (2 or 4)
  int halo_out_bufs;       // halo kernel: output staging buffers (4, 6 or 8)
  int halo_ksteps;         // halo kernel: K16 steps per tap = ceil(Cin / 16) (<= PB / 32: a 48-channel layer skips the all-zero fourth)
  int halo_s2;             // halo kernel: stride 2 -- a stage holds the four parity planes of the input (18 x 10 each)
  int reverse;             // walk the tiles last-to-first (alternated layer by layer so a layer starts on what is still in L2)
  signed char tap_map[9], tap_dx[9], tap_dy[9];
  const float* bias;
  const float* a_scale;    // null, or [N][a_scale_c] fp32: the A tile is multiplied by it in shared memory before the MMA
  int a_scale_hw, a_scale_c;   // pixels per image, channels per image row of a_scale
  long long a_pixels;      // N * H * W (rows past it are TMA zero fill)
};

struct ConvLaunch {
  ConvMaps maps;
  ConvParams p;
  unsigned grid = 0;
  int block_n = 0;
  bool two_cta = false;
  bool halo = false;     // 3x3 halo-tile kernel (Cin <= 64, resident weights)
  int halo_smem = 0;
  double flops = 0;  // 2*M*N*K (useful, un-padded)
};

int conv_prepare(const ConvSpec& spec, ConvLaunch* out);
int conv_launch(const ConvLaunch& l, cudaStream_t stream);

}  // namespace b200
