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

}  // namespace b200
