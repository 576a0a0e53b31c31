#include "common.h"

#include <cudaTypedefs.h>
#include <mutex>
#include <stdlib.h>

namespace b200 {

static thread_local std::string g_last_error;

void set_error(const std::string& msg) { g_last_error = msg; }
int fail(int code, const std::string& msg) {
  g_last_error = msg;
  return code;
}
const char* last_error_cstr() { return g_last_error.c_str(); }

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

int encode_tmap(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                const uint32_t* box, Swizzle swz) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return fail(B200_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable (driver too old?)");
  cuuint64_t gdim[5];
  cuuint64_t gstr[4];
  cuuint32_t bx[5];
  cuuint32_t es[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
    if (i + 1 < rank) gstr[i] = strides_bytes[i];
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15) != 0) return fail(B200_ERR_INVALID, "tensor map base not 16B aligned");
  for (int i = 0; i + 1 < rank; ++i)
    if (gstr[i] % 16 != 0) return fail(B200_ERR_INVALID, "tensor map stride not a multiple of 16 bytes");
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, static_cast<cuuint32_t>(rank), const_cast<void*>(base), gdim,
                  gstr, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  swz == SWZ_128B ? CU_TENSOR_MAP_SWIZZLE_128B
                  : swz == SWZ_64B ? CU_TENSOR_MAP_SWIZZLE_64B
                  : swz == SWZ_32B ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    std::string m = "cuTensorMapEncodeTiled failed rc=" + std::to_string(static_cast<int>(r)) + " rank=" +
                    std::to_string(rank) + " dims=";
    for (int i = 0; i < rank; ++i) m += std::to_string(dims[i]) + (i + 1 < rank ? "," : "");
    m += " box=";
    for (int i = 0; i < rank; ++i) m += std::to_string(box[i]) + (i + 1 < rank ? "," : "");
    m += " strides=";
    for (int i = 0; i + 1 < rank; ++i) m += std::to_string(strides_bytes[i]) + (i + 2 < rank ? "," : "");
    return fail(B200_ERR_CUDA, m);
  }
  return B200_OK;
}

bool pdl_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200_PDL");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1;
}

int pdl_min_waves() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200_PDL_MIN_WAVES");
    v = e ? atoi(e) : 3;
  }
  return v;
}

int device_sm_count() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;
  }
  return sms;
}

}  // namespace b200
