// Plan-driven forward engine: the Python side lowers a torchvision-style module tree into a flat list of fused ops
// (BatchNorm folded, activations/residuals fused) plus a weights section; this executor owns device memory for weights
// and activations, prepares TMA descriptors per batch size and replays the op list as one CUDA graph per batch size.
#pragma once
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "common.h"
#include "conv_gemm.cuh"
#include "kernels.cuh"

namespace b200 {

#pragma pack(push, 1)
struct BlobHeader {
  char magic[8];  // "B200PLAN"
  uint32_t version;
  uint32_t n_bufs;
  uint32_t n_ops;
  uint32_t reserved;
  char model_kind[32];
  int32_t in_h, in_w;
  int32_t in_buf, out_buf;
  int32_t out_dim;   // fp32 elements per image in out_buf
  int32_t pre_fill;  // letterbox / pad fill value (uint8)
  float mean[3], stdv[3];
  uint64_t weights_offset, weights_bytes;
};
struct BufRec {
  int32_t h, w, c, dtype;  // dtype 0 = bf16, 1 = fp32
};
struct OpRec {
  int32_t type;
  int32_t p[23];
  int64_t w_off, b_off, w2_off, b2_off;
};
#pragma pack(pop)
static_assert(sizeof(OpRec) == 128, "OpRec layout");

enum OpType : int {
  OP_CONV = 1, OP_STEM = 2, OP_MAXPOOL = 3, OP_AVGPOOL = 4, OP_FC = 5, OP_DWCONV = 6, OP_SCALE = 7, OP_COPY = 8,
  OP_UPSAMPLE = 9, OP_SHUFFLE = 10, OP_YOLO_DECODE = 11, OP_GATHER_CONV = 12, OP_SE_GATE = 13,
  OP_SE_HIDDEN = 14, OP_SE_SCALE = 15
};

struct PreparedOp {
  OpRec rec;
  ConvLaunch conv;  // OP_CONV only
};

struct BatchPlan {
  int batch = 0;
  std::vector<PreparedOp> ops;
  cudaGraphExec_t graph = nullptr;
  bool warmed = false;
};

struct Engine {
  std::string kind;
  BlobHeader hdr;
  std::vector<BufRec> bufs;
  std::vector<OpRec> ops;
  int max_batch = 0;
  int device = 0;
  bool use_graphs = true;
  char* weights_dev = nullptr;
  std::vector<char*> buf_dev;        // per buffer: its address inside `arena_dev` (bf16) or `vec_dev` (fp32)
  char* arena_dev = nullptr;         // bf16 activations, buffers with disjoint lifetimes share addresses (plan_arena)
  char* vec_dev = nullptr;           // fp32 vector buffers (pooled features, SE sums / gates, logits): never shared
  size_t arena_bytes = 0, vec_bytes = 0;
  size_t device_bytes = 0;           // everything this engine holds on the device right now
  bool keep_buffers = false;         // B200_ENGINE_KEEP_BUFFERS: every buffer keeps its own memory (debug reads)
  uint16_t* lut_dev = nullptr;
  // collate staging
  uint8_t* stage_dev = nullptr;
  size_t stage_bytes = 0;
  CollateItem* items_dev = nullptr;
  CollateItem* items_host = nullptr;  // pinned ring of ITEM_SLOTS x max_batch records (async H2D source)
  static constexpr int ITEM_SLOTS = 8;
  cudaEvent_t item_events[ITEM_SLOTS] = {};
  unsigned item_slot = 0;
  float* stage_f32_dev = nullptr;
  std::unordered_map<int, std::unique_ptr<BatchPlan>> plans;
  cudaStream_t capture_stream = nullptr;
  long long launches = 0;  // kernels launched so far (graph nodes counted per replay)

  ~Engine();
  int init(const char* kind, const void* blob, size_t nbytes, int max_batch, int device, unsigned flags = 0);
  int parse(const char* kind, const void* blob, size_t nbytes, int max_batch);   // header + tables, validated; no CUDA
  // offsets of every buffer inside the two device allocations for `max_batch`; returns their sizes
  void plan_arena(bool reuse, std::vector<size_t>* offsets, size_t* arena, size_t* vec) const;
  size_t fixed_bytes() const;        // weights + tables + collate item ring: what init() allocates besides the buffers
  int warm(int batch, cudaStream_t stream);   // eager pass + graph capture for one batch size
  View view(int buf, int coff, int c, int batch) const;
  int plan_for(int batch, BatchPlan** out);
  int run_ops(BatchPlan& plan, cudaStream_t stream, int first, int last);
  int run(int batch, cudaStream_t stream);
  size_t buf_bytes_per_image(int b) const {
    return size_t(bufs[b].h) * bufs[b].w * bufs[b].c * (bufs[b].dtype ? 4 : 2);
  }
};

}  // namespace b200
