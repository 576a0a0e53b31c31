// Host-side shared declarations for libb200infer: error reporting, NHWC tensor views, TMA descriptor encoding.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <string.h>

#include "../../include/b200infer.h"

namespace b200 {

// ---- error plumbing: every C-ABI entry returns 0 or a negative code and records a message ----
void set_error(const std::string& msg);
int fail(int code, const std::string& msg);
const char* last_error_cstr();

#define B200_CUDA_OK(expr)                                                                             \
  do {                                                                                                 \
    cudaError_t _e = (expr);                                                                           \
    if (_e != cudaSuccess)                                                                             \
      return ::b200::fail(B200_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
  } while (0)

#define B200_TRY(expr)       \
  do {                       \
    int _rc = (expr);        \
    if (_rc != 0) return _rc; \
  } while (0)

// ---- NHWC bf16 view: `base` already points at the first channel of the slice; consecutive pixels are
// `pix_stride` elements apart (= total channels of the underlying buffer), rows/images are dense. ----
struct View {
  void* base = nullptr;
  int N = 0, H = 0, W = 0, C = 0;
  int pix_stride = 0;
};
inline View make_view(void* base, int N, int H, int W, int C, int pix_stride = 0, int c_off = 0) {
  View v;
  v.base = static_cast<char*>(base) + static_cast<size_t>(c_off) * 2;
  v.N = N; v.H = H; v.W = W; v.C = C;
  v.pix_stride = pix_stride ? pix_stride : C;
  return v;
}

// ---- TMA descriptors (driver entry point fetched at run time; no link-time libcuda dependency) ----
enum Swizzle : int { SWZ_NONE = 0, SWZ_32B = 1, SWZ_64B = 2, SWZ_128B = 3 };
// rank-4 tiled map over bf16: dims[0] is innermost (contiguous); strides_bytes[i] is the stride of dims[i+1].
int encode_tmap(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                const uint32_t* box, Swizzle swz);

int device_sm_count();

// Launch with programmatic stream serialization (B200_PDL=0 turns it off): the kernel MUST call pdl_wait() (ptx.cuh) before
// reading anything an earlier kernel of the stream wrote.
// Measured: the programmatic edge pays off between large persistent kernels (ResNet-50 @256: -2 %) and costs ~2.5 us per
// edge in graphs of small kernels (YOLOv8n @32: +10 %), so callers pass `big` = "this launch has several waves of tiles".
bool pdl_enabled();
int pdl_min_waves();   // B200_PDL_MIN_WAVES (default 3): tiles per SM from which a persistent kernel is launched with the programmatic edge
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_if(bool big, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                 Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = (big && pdl_enabled()) ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  return launch_pdl_if(false, kern, grid, block, smem, stream, static_cast<Args&&>(args)...);
}

// multiplier for fast_div (ptx.cuh): floor(2^32 / d) + 1; unused for d < 2
struct FastDiv { unsigned d, mul; };   // divisor and its multiplier, passed to kernels by value
inline unsigned fast_div_mul(int d) { return d < 2 ? 0u : static_cast<unsigned>((1ull << 32) / static_cast<unsigned>(d)) + 1u; }

inline FastDiv make_fast_div(long long d) { return FastDiv{static_cast<unsigned>(d), fast_div_mul(static_cast<int>(d))}; }

}  // namespace b200
