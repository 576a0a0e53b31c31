/* libb200infer — C ABI of the B200-native batched-forward engine.
 *
 * Drop-in scope: the path the reference scheduler (293-project/src/scheduler.py) runs per batch inside a model actor:
 *   RequestQueue.get_batch (:258-322)  ->  GPUWorker.process_batch (:435-475):  torch.stack -> .cuda() -> model(inputs) -> .cpu()
 * plus the client-side preprocessing the reference leaves to callers (profiling/run_profiler.py:63-70).
 *
 * Conventions: every entry returns 0 on success or a negative B200_ERR_* code; b200_last_error() then describes the
 * failure (thread-local).  Nothing throws across the boundary.  All pointers named `*_dev`, and all in/out tensor
 * pointers of the operator-level entries, are device pointers on the current CUDA device; `stream` is a cudaStream_t
 * passed as void* (NULL = legacy default stream).  Activations are NHWC bf16; `*_pix_stride` is the distance in
 * elements between consecutive pixels (= channel count of the underlying buffer, 0 = dense), which lets a call read
 * or write a channel slice of a wider buffer (concat / chunk without a copy).
 */
#ifndef B200INFER_H_
#define B200INFER_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_ABI_VERSION 1

#if defined(__GNUC__)
#define B200_API __attribute__((visibility("default")))
#else
#define B200_API
#endif

#define B200_OK 0
#define B200_ERR_INVALID (-1)
#define B200_ERR_CUDA (-2)
#define B200_ERR_BLOB (-3)
#define B200_ERR_NO_DEVICE (-4)

/* activation codes for the fused epilogues */
#define B200_ACT_NONE 0
#define B200_ACT_RELU 1
#define B200_ACT_SILU 2
#define B200_ACT_SIGMOID 3

B200_API const char* b200_last_error(void);
B200_API int b200_abi_version(void);

/* ---- operator level (one entry per kernel; what the parity tests bind) ------------------------------------- */

/* Conv2d (k in {1,3}, stride in {1,2}) + folded BatchNorm bias (+ residual add) + activation, tcgen05 implicit GEMM.
 * Replaces torch.nn.Conv2d + BatchNorm2d + ReLU/SiLU (+ `out += identity`) as reached from scheduler.py:450 through
 * torchvision (models/resnet.py:143-161 Bottleneck.forward; models/efficientnet.py:105-231; ops/misc.py Conv2dNormActivation).
 * weight: bf16 [cout][k*k][cin] (tap-major, then input channel); bias: fp32 [cout]. cin, cout multiples of 8. */
B200_API int b200_conv2d_nhwc(const void* in, int n, int h, int w, int cin, int in_pix_stride, const void* weight,
                     const float* bias, const void* residual, int res_pix_stride, void* out, int cout,
                     int out_pix_stride, int ksize, int stride, int pad, int act, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200INFER_H_ */
