// b200_collate_u8_noalloc: the collate + preprocess kernel for callers that must not allocate device memory.
//
// Who that is: a model actor that the scheduler has packed onto a GPU by the *profiled* peak memory of its sessions
// (the bin-packing policy budgets each GPU from the profile's peak_memory_mb column).  Such an actor has handed all the
// memory it is entitled to to its engines up front; a cudaMalloc in the middle of a duty cycle can fail on a fully packed
// GPU, and even when it succeeds it synchronises the whole device and stalls the other sessions sharing it.  The regular
// b200_collate grows an engine-owned staging buffer on demand (cudaMalloc/cudaFree inside the call); this variant keeps
// everything the kernel needs beyond the caller's input and output -- the per-image descriptor table and the 3x256
// normalisation table -- in one fixed array that lives inside the loaded library image.
//
// "Allocates nothing" holds from the second call on a device onwards: the first call resolves the array's address, and
// with lazy module loading the CUDA runtime itself loads the library image and this kernel's code onto the device at
// that moment (driver-internal device memory).  An actor therefore makes one warm-up call per device while it is
// still initialising, before it reports its memory as settled.
//
// Cost: every GPU that loads libb200infer.so holds B200_NOALLOC_MAX_IMAGES * 16 B + 1.5 KB (= 17.5 KB) until the process
// exits, whether or not this entry point is ever called.
#include "kernels.cuh"

namespace b200 {

struct NoAllocScratch {
  CollateItem items[B200_NOALLOC_MAX_IMAGES];
  uint16_t lut[3 * 256];
};

// File-scope device array: on the host this name is only a handle for cudaGetSymbolAddress, never a device pointer.
static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(NoAllocScratch)];

static constexpr int MAX_DEVICES = 64;
static void* g_scratch_addr[MAX_DEVICES] = {};  // device address of g_noalloc_scratch on each GPU, filled on first use

// Once per device, with that device current: resolve the array's address in that device's copy of the library image.
static int noalloc_setup(int dev, NoAllocScratch** out) {
  if (dev < 0 || dev >= MAX_DEVICES) return fail(B200_ERR_INVALID, "b200_collate_u8_noalloc: device index out of range");
  if (!g_scratch_addr[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p)
      return fail(B200_ERR_CUDA, std::string("b200_collate_u8_noalloc: cudaGetSymbolAddress failed: ") +
                                     cudaGetErrorString(e));
    g_scratch_addr[dev] = p;
  }
  *out = static_cast<NoAllocScratch*>(g_scratch_addr[dev]);
  return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_collate_u8_noalloc(const void* packed_dev, const uint64_t* offsets, const int32_t* hw, int n,
                                       int padded_batch, int out_h, int out_w, int mode, int fill, const float* mean3,
                                       const float* std3, size_t packed_bytes, void* out_nhwc4_bf16_dev, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (n > B200_NOALLOC_MAX_IMAGES)
    return fail(B200_ERR_INVALID, "b200_collate_u8_noalloc: " + std::to_string(n) + " images exceed the fixed scratch (max " +
                                      std::to_string(B200_NOALLOC_MAX_IMAGES) + "); split the batch or use b200_collate");
  if (n < 0 || padded_batch < n || padded_batch <= 0)
    return fail(B200_ERR_INVALID, "b200_collate_u8_noalloc: need 0 <= n <= padded_batch");
  if (!out_nhwc4_bf16_dev || !mean3 || !std3 || (n > 0 && (!packed_dev || !offsets || !hw)))
    return fail(B200_ERR_INVALID, "b200_collate_u8_noalloc: null argument");
  if (fill < 0 || fill > 255) return fail(B200_ERR_INVALID, "b200_collate_u8_noalloc: fill must be a byte value");
  // The two table uploads below read host stack memory that is gone when this function returns; recorded into a graph
  // they would be replayed from dead memory, so a stream that is being captured is refused.
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  B200_CUDA_OK(cudaStreamIsCapturing(stream, &cap));
  if (cap != cudaStreamCaptureStatusNone)
    return fail(B200_ERR_INVALID, "b200_collate_u8_noalloc: the stream is being captured; call it outside stream capture");
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  NoAllocScratch* scratch = nullptr;
  B200_TRY(noalloc_setup(dev, &scratch));

  // Host copies of the two tables (stack; pageable sources are staged by the copy before it returns).
  CollateItem items[B200_NOALLOC_MAX_IMAGES];
  for (int i = 0; i < n; ++i) {
    items[i].offset = offsets[i];
    items[i].h = hw[2 * i];
    items[i].w = hw[2 * i + 1];
    if (items[i].h <= 0 || items[i].w == 0) return fail(B200_ERR_INVALID, "b200_collate_u8_noalloc: empty image");
    const uint64_t wpx = items[i].w < 0 ? uint64_t(-int64_t(items[i].w)) : uint64_t(items[i].w);   // w < 0: grayscale
    if (offsets[i] + uint64_t(items[i].h) * wpx * (items[i].w < 0 ? 1 : 3) > packed_bytes)
      return fail(B200_ERR_INVALID, "b200_collate_u8_noalloc: image " + std::to_string(i) + " exceeds the packed buffer");
    if (mode == COLLATE_IDENTITY && (items[i].h != out_h || items[i].w != out_w))
      return fail(B200_ERR_INVALID, "b200_collate_u8_noalloc: image " + std::to_string(i) + " is not out_h x out_w");
  }
  uint16_t lut[3 * 256];
  build_norm_lut(mean3, std3, lut);
  // Both tables are rewritten in full on every call, so nothing in the shared array has to start out as zero.
  if (n > 0)
    B200_CUDA_OK(cudaMemcpyAsync(scratch->items, items, sizeof(CollateItem) * n, cudaMemcpyHostToDevice, stream));
  B200_CUDA_OK(cudaMemcpyAsync(scratch->lut, lut, sizeof(lut), cudaMemcpyHostToDevice, stream));
  return collate_u8_launch(static_cast<const uint8_t*>(packed_dev), scratch->items, items, n, padded_batch, out_h, out_w,
                           mode, fill, scratch->lut, out_nhwc4_bf16_dev, stream);
}
