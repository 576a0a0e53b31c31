#include "engine.h"

#include <algorithm>
#include <stdlib.h>

using namespace b200;

namespace b200 {

Engine::~Engine() {
  cudaSetDevice(device);
  for (auto& kv : plans)
    if (kv.second->graph) cudaGraphExecDestroy(kv.second->graph);
  cudaFree(arena_dev);
  cudaFree(vec_dev);
  cudaFree(weights_dev);
  cudaFree(lut_dev);
  cudaFree(stage_dev);
  cudaFree(items_dev);
  cudaFree(stage_f32_dev);
  if (items_host) cudaFreeHost(items_host);
  for (auto& ev : item_events)
    if (ev) cudaEventDestroy(ev);
  if (capture_stream) cudaStreamDestroy(capture_stream);
}

int Engine::parse(const char* kind_in, const void* blob, size_t nbytes, int max_batch_in) {
  if (!blob || nbytes < sizeof(BlobHeader)) return fail(B200_ERR_BLOB, "engine: blob too small");
  memcpy(&hdr, blob, sizeof(hdr));
  if (memcmp(hdr.magic, "B200PLAN", 8) != 0 || hdr.version != 1) return fail(B200_ERR_BLOB, "engine: bad magic/version");
  hdr.model_kind[31] = 0;
  kind = hdr.model_kind;
  if (kind_in && kind_in[0] && kind != kind_in)
    return fail(B200_ERR_BLOB, "engine: blob is for model '" + kind + "', not '" + kind_in + "'");
  const size_t tables = sizeof(BlobHeader) + size_t(hdr.n_bufs) * sizeof(BufRec) + size_t(hdr.n_ops) * sizeof(OpRec);
  if (tables > nbytes || hdr.weights_offset < tables || hdr.weights_offset + hdr.weights_bytes > nbytes)
    return fail(B200_ERR_BLOB, "engine: truncated blob");
  if (max_batch_in <= 0) return fail(B200_ERR_INVALID, "engine: max_batch must be positive");
  max_batch = max_batch_in;

  const char* base = static_cast<const char*>(blob);
  bufs.resize(hdr.n_bufs);
  memcpy(bufs.data(), base + sizeof(BlobHeader), size_t(hdr.n_bufs) * sizeof(BufRec));
  ops.resize(hdr.n_ops);
  memcpy(ops.data(), base + sizeof(BlobHeader) + size_t(hdr.n_bufs) * sizeof(BufRec), size_t(hdr.n_ops) * sizeof(OpRec));
  if (hdr.in_buf < 0 || hdr.in_buf >= int(hdr.n_bufs) || hdr.out_buf < 0 || hdr.out_buf >= int(hdr.n_bufs))
    return fail(B200_ERR_BLOB, "engine: bad in/out buffer ids");
  for (const OpRec& o : ops) {
    const int ids[3] = {o.p[0], o.p[3], o.p[6]};
    for (int k = 0; k < 3; ++k)
      if (ids[k] >= int(hdr.n_bufs) || (k < 2 && ids[k] < 0)) return fail(B200_ERR_BLOB, "engine: op references unknown buffer");
    if ((o.type == OP_DWCONV || o.type == OP_SCALE || o.type == OP_SE_SCALE || o.type == OP_SHUFFLE ||
         o.type == OP_YOLO_DECODE) && (o.p[12] < 0 || o.p[12] >= int(hdr.n_bufs)))
      return fail(B200_ERR_BLOB, "engine: op references unknown side buffer");
    if (o.type == OP_DWCONV && o.p[12] > 0) {   // fused squeeze: the side buffer must hold every strip's channel sums
      const BufRec& sb = bufs[o.p[12]];
      const BufRec& ob = bufs[o.p[3]];
      const int strips = dwconv3x3_strips_c(o.p[5], ob.h, ob.w, o.p[9]);
      if (strips <= 0 || sb.dtype != 1 || int64_t(sb.h) * sb.w * sb.c < int64_t(strips) * o.p[5])
        return fail(B200_ERR_BLOB, "engine: depthwise squeeze buffer too small for this geometry");
    }
    if (o.type == OP_SE_HIDDEN) {
      const BufRec& sb = bufs[o.p[0]];
      const BufRec& hb = bufs[o.p[3]];
      if (o.p[1] <= 0 || o.p[2] <= 0 || o.p[8] <= 0 || sb.dtype != 1 || hb.dtype != 1 ||
          int64_t(sb.h) * sb.w * sb.c < int64_t(o.p[1]) * o.p[2] || int64_t(hb.h) * hb.w * hb.c < o.p[8])
        return fail(B200_ERR_BLOB, "engine: squeeze-excitation hidden buffers do not match its shape");
    }
    if (o.type == OP_SE_SCALE) {
      const BufRec& hb = bufs[o.p[12]];
      if (o.p[8] <= 0 || o.p[8] > 256 || hb.dtype != 1 || int64_t(hb.h) * hb.w * hb.c < o.p[8])
        return fail(B200_ERR_BLOB, "engine: squeeze-excitation scale does not match its hidden buffer");
    }
    if (o.type == OP_SE_GATE) {
      const BufRec& sb = bufs[o.p[0]];
      const BufRec& gb = bufs[o.p[3]];
      if (o.p[1] <= 0 || o.p[2] <= 0 || o.p[8] <= 0 || sb.dtype != 1 || gb.dtype != 1 ||
          int64_t(sb.h) * sb.w * sb.c < int64_t(o.p[1]) * o.p[2] || int64_t(gb.h) * gb.w * gb.c < o.p[2])
        return fail(B200_ERR_BLOB, "engine: squeeze-excitation gate buffers do not match its shape");
    }
    if (o.type == OP_CONV && o.p[17] != 0) {
      if (o.p[17] < 0 || o.p[17] > int(hdr.n_bufs) || bufs[o.p[17] - 1].dtype != 1 ||
          int64_t(bufs[o.p[17] - 1].h) * bufs[o.p[17] - 1].w * bufs[o.p[17] - 1].c < o.p[2])
        return fail(B200_ERR_BLOB, "engine: conv input-scale buffer does not match its input channels");
    }
    if (o.type == OP_CONV && o.p[18] != 0 && (o.p[18] < 0 || o.p[6] < 0 || o.p[18] > o.p[5] / 2))
      return fail(B200_ERR_BLOB, "engine: interleaving conv needs the other branch and at most half the output channels");
    if (o.type == OP_CONV && (o.p[13] < 0 || o.p[13] > int(hdr.n_bufs)))
      return fail(B200_ERR_BLOB, "engine: conv references an unknown second input");
    const int64_t offs[4] = {o.w_off, o.b_off, o.w2_off, o.b2_off};
    for (int64_t off : offs)
      if (off >= int64_t(hdr.weights_bytes)) return fail(B200_ERR_BLOB, "engine: weight offset out of range");
  }
  return B200_OK;
}

// Buffers an op reads or writes (mirrors run_ops): ids[] gets up to 6 entries, returns how many.
static int op_buffers(const OpRec& o, int ids[6]) {
  int n = 0;
  const int32_t* p = o.p;
  ids[n++] = p[0];
  ids[n++] = p[3];
  switch (o.type) {
    case OP_CONV:
      if (p[6] >= 0) ids[n++] = p[6];
      if (p[13] > 0) ids[n++] = p[13] - 1;
      if (p[17] > 0) ids[n++] = p[17] - 1;
      break;
    case OP_GATHER_CONV:
      if (p[6] >= 0) ids[n++] = p[6];
      break;
    case OP_DWCONV:
      if (p[12] > 0) ids[n++] = p[12];
      break;
    case OP_SCALE: case OP_SE_SCALE: case OP_SHUFFLE: case OP_YOLO_DECODE:
      ids[n++] = p[12];
      break;
    default:
      break;
  }
  return n;
}

// Activation memory by liveness.  A bf16 buffer is live from the first op that touches it to the last one; two buffers
// whose live ranges do not intersect may share addresses (the ops run in plan order on one stream, and a kernel launched
// with the programmatic edge still waits for its predecessor to finish before it touches global memory).  Placement:
// largest buffer first, each at the lowest offset that is free of every already-placed buffer alive at the same time.
// Not shared, on purpose: the model input (a caller may read it back, or collate the next batch, at any time) and the
// fp32 vector buffers (tiny; and stale fp32 bits read as bf16 padding channels could be NaN patterns, which zero weights
// do not cancel -- stale bf16 activations are always finite).
void Engine::plan_arena(bool reuse, std::vector<size_t>* offsets, size_t* arena, size_t* vec) const {
  const int nb = static_cast<int>(bufs.size());
  std::vector<int> first(nb, 1 << 30), last(nb, -1);
  for (int i = 0; i < static_cast<int>(ops.size()); ++i) {
    int ids[6];
    const int n = op_buffers(ops[i], ids);
    for (int k = 0; k < n; ++k) {
      first[ids[k]] = std::min(first[ids[k]], i);
      last[ids[k]] = std::max(last[ids[k]], i);
    }
  }
  auto bytes_of = [&](int b) { return (buf_bytes_per_image(b) * size_t(max_batch) + 256 + 255) / 256 * 256; };
  offsets->assign(nb, 0);
  size_t vtop = 0, atop = 0;
  std::vector<int> order;
  for (int b = 0; b < nb; ++b) {
    if (bufs[b].dtype == 1) {
      (*offsets)[b] = vtop;
      vtop += bytes_of(b);
    } else {
      order.push_back(b);
    }
  }
  std::sort(order.begin(), order.end(), [&](int x, int y) {
    const size_t bx = bytes_of(x), by = bytes_of(y);
    return bx != by ? bx > by : x < y;
  });
  std::vector<int> placed;
  for (int b : order) {
    const bool pinned = !reuse || b == hdr.in_buf || last[b] < 0;
    size_t off = 0;
    if (pinned) {
      first[b] = -1;            // alive for the whole plan
      last[b] = 1 << 30;
    }
    // candidate offsets: 0 and the end of every conflicting buffer; take the lowest that fits
    std::vector<std::pair<size_t, size_t>> busy;   // [begin, end) of placed buffers alive together with b
    for (int q : placed)
      if (!(last[q] < first[b] || last[b] < first[q])) busy.emplace_back((*offsets)[q], (*offsets)[q] + bytes_of(q));
    std::sort(busy.begin(), busy.end());
    for (const auto& iv : busy) {
      if (off + bytes_of(b) <= iv.first) break;
      off = std::max(off, iv.second);
    }
    (*offsets)[b] = off;
    atop = std::max(atop, off + bytes_of(b));
    placed.push_back(b);
  }
  *arena = std::max<size_t>(atop, 256);
  *vec = std::max<size_t>(vtop, 256);
}

size_t Engine::fixed_bytes() const {
  return std::max<uint64_t>(hdr.weights_bytes, 256) + sizeof(uint16_t) * 3 * 256 + sizeof(CollateItem) * size_t(max_batch);
}

int Engine::init(const char* kind_in, const void* blob, size_t nbytes, int max_batch_in, int device_in, unsigned flags) {
  B200_TRY(parse(kind_in, blob, nbytes, max_batch_in));
  device = device_in;
  keep_buffers = (flags & B200_ENGINE_KEEP_BUFFERS) != 0;
  if (const char* e = getenv("B200_ARENA_REUSE"))
    if (e[0] == '0') keep_buffers = true;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(B200_ERR_NO_DEVICE, "engine: no CUDA device");
  B200_CUDA_OK(cudaSetDevice(device));
  cudaDeviceProp prop;
  B200_CUDA_OK(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) return fail(B200_ERR_NO_DEVICE, std::string("engine: needs an sm_100 device, found ") + prop.name);

  const char* base = static_cast<const char*>(blob);
  B200_CUDA_OK(cudaMalloc(&weights_dev, std::max<uint64_t>(hdr.weights_bytes, 256)));
  B200_CUDA_OK(cudaMemcpy(weights_dev, base + hdr.weights_offset, hdr.weights_bytes, cudaMemcpyHostToDevice));
  std::vector<size_t> offsets;
  plan_arena(!keep_buffers, &offsets, &arena_bytes, &vec_bytes);
  B200_CUDA_OK(cudaMalloc(&arena_dev, arena_bytes));
  B200_CUDA_OK(cudaMemset(arena_dev, 0, arena_bytes));
  B200_CUDA_OK(cudaMalloc(&vec_dev, vec_bytes));
  B200_CUDA_OK(cudaMemset(vec_dev, 0, vec_bytes));
  buf_dev.assign(hdr.n_bufs, nullptr);
  for (uint32_t b = 0; b < hdr.n_bufs; ++b) buf_dev[b] = (bufs[b].dtype == 1 ? vec_dev : arena_dev) + offsets[b];
  uint16_t lut[3 * 256];
  build_norm_lut(hdr.mean, hdr.stdv, lut);
  B200_CUDA_OK(cudaMalloc(&lut_dev, sizeof(lut)));
  B200_CUDA_OK(cudaMemcpy(lut_dev, lut, sizeof(lut), cudaMemcpyHostToDevice));
  B200_CUDA_OK(cudaMalloc(&items_dev, sizeof(CollateItem) * max_batch));
  B200_CUDA_OK(cudaMallocHost(&items_host, sizeof(CollateItem) * max_batch * ITEM_SLOTS));
  for (auto& ev : item_events) B200_CUDA_OK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  B200_CUDA_OK(cudaStreamCreateWithFlags(&capture_stream, cudaStreamNonBlocking));
  device_bytes = fixed_bytes() + arena_bytes + vec_bytes;
  return B200_OK;
}

View Engine::view(int buf, int coff, int c, int batch) const {
  const BufRec& b = bufs[buf];
  return make_view(buf_dev[buf], batch, b.h, b.w, c, b.c, coff);
}

int Engine::plan_for(int batch, BatchPlan** out) {
  auto it = plans.find(batch);
  if (it != plans.end()) {
    *out = it->second.get();
    return B200_OK;
  }
  std::unique_ptr<BatchPlan> plan(new BatchPlan);
  plan->batch = batch;
  plan->ops.resize(ops.size());
  static int alt = -1;
  if (alt < 0) {
    const char* e = getenv("B200_ALTERNATE_TILES");
    alt = (e && e[0] == '0') ? 0 : 1;
  }
  int conv_index = 0;
  for (size_t i = 0; i < ops.size(); ++i) {
    PreparedOp& po = plan->ops[i];
    po.rec = ops[i];
    const int32_t* p = po.rec.p;
    if (po.rec.type == OP_CONV) {
      ConvSpec s;
      s.in = view(p[0], p[1], p[2], batch);
      s.out = view(p[3], p[4], p[5], batch);
      s.interleave_c = p[18];      // ShuffleNetV2 unit tail: p[5] = 2 cp output channels, residual slot = the other branch (cp)
      if (p[6] >= 0) s.res = view(p[6], p[7], p[18] > 0 ? p[5] / 2 : p[5], batch);
      s.w = weights_dev + po.rec.w_off;
      s.bias = reinterpret_cast<const float*>(weights_dev + po.rec.b_off);
      s.ksize = p[8]; s.stride = p[9]; s.pad = p[10]; s.act = p[11];
      s.w_cin = p[19]; s.w_rows = p[20];      // weights stored padded to the TMA box sizes (0 = exact)
      s.res_after_act = p[12] != 0;
      if (p[17] > 0) s.a_scale = reinterpret_cast<const float*>(buf_dev[p[17] - 1]);   // SE gate [N][Cin] fp32, buffer id + 1
      if (p[13] > 0) {      // K-concatenated second input: buffer id + 1, channel offset, channels, stride
        s.in2 = view(p[13] - 1, p[14], p[15], batch);
        s.stride2 = p[16];
      }
      s.reverse_tiles = alt && (conv_index++ & 1);   // consecutive conv layers walk their tiles in opposite directions
      int rc = conv_prepare(s, &po.conv);
      if (rc != 0) return fail(rc, "op " + std::to_string(i) + ": " + last_error_cstr());
    }
  }
  *out = plan.get();
  plans[batch] = std::move(plan);
  return B200_OK;
}

int Engine::run_ops(BatchPlan& plan, cudaStream_t stream, int first, int last) {
  const int batch = plan.batch;
  for (int i = first; i <= last; ++i) {
    PreparedOp& po = plan.ops[i];
    const int32_t* p = po.rec.p;
    const char* W = weights_dev;
    int rc = B200_OK;
    switch (po.rec.type) {
      case OP_CONV:
        rc = conv_launch(po.conv, stream);
        break;
      case OP_STEM:
        if (p[12] == 3) {
          rc = stem_pool_launch_k(view(p[0], p[1], p[2], batch), view(p[3], p[4], p[5], batch), W + po.rec.w_off,
                                  reinterpret_cast<const float*>(W + po.rec.b_off), p[8], p[11], stream);
          break;
        }
        if (p[12] == 2) {
          rc = stem_row_launch_k(view(p[0], p[1], p[2], batch), view(p[3], p[4], p[5], batch), W + po.rec.w_off,
                                 reinterpret_cast<const float*>(W + po.rec.b_off), p[8], p[11], stream);
          break;
        }
        if (p[12] == 1) {
          rc = stem_tc_launch(view(p[0], p[1], p[2], batch), view(p[3], p[4], p[5], batch), W + po.rec.w_off,
                              reinterpret_cast<const float*>(W + po.rec.b_off), p[8], p[9], p[10], p[11], stream);
          break;
        }
        rc = stem_conv_launch(view(p[0], p[1], p[2], batch), view(p[3], p[4], p[5], batch),
                              reinterpret_cast<const float*>(W + po.rec.w_off),
                              reinterpret_cast<const float*>(W + po.rec.b_off), p[8], p[9], p[10], p[11], stream);
        break;
      case OP_MAXPOOL:
        rc = maxpool_launch(view(p[0], p[1], p[2], batch), view(p[3], p[4], p[5], batch), p[8], p[9], p[10], stream);
        break;
      case OP_AVGPOOL:
        rc = global_avgpool_launch(view(p[0], p[1], p[2], batch), reinterpret_cast<float*>(buf_dev[p[3]]), stream);
        break;
      case OP_FC: {
        float scale;
        memcpy(&scale, &p[12], 4);
        rc = fc_launch(reinterpret_cast<const float*>(buf_dev[p[0]]), reinterpret_cast<const float*>(W + po.rec.w_off),
                       reinterpret_cast<const float*>(W + po.rec.b_off), reinterpret_cast<float*>(buf_dev[p[3]]), batch,
                       p[2], p[5], scale, p[11], stream);
        break;
      }
      case OP_DWCONV:
        rc = dwconv3x3_launch(view(p[0], p[1], p[2], batch), view(p[3], p[4], p[5], batch),
                              reinterpret_cast<const float*>(W + po.rec.w_off),
                              reinterpret_cast<const float*>(W + po.rec.b_off), p[9], p[11],
                              p[12] > 0 ? reinterpret_cast<float*>(buf_dev[p[12]]) : nullptr, stream);
        break;
      case OP_SE_GATE: {
        float inv_hw;
        memcpy(&inv_hw, &p[12], 4);
        rc = se_gate_launch(reinterpret_cast<const float*>(buf_dev[p[0]]), batch, p[1], p[2], inv_hw,
                            reinterpret_cast<const float*>(W + po.rec.w_off),
                            reinterpret_cast<const float*>(W + po.rec.b_off),
                            reinterpret_cast<const float*>(W + po.rec.w2_off),
                            reinterpret_cast<const float*>(W + po.rec.b2_off), p[8],
                            reinterpret_cast<float*>(buf_dev[p[3]]), stream);
        break;
      }
      case OP_SE_HIDDEN: {
        float inv_hw;
        memcpy(&inv_hw, &p[12], 4);
        rc = se_hidden_launch(reinterpret_cast<const float*>(buf_dev[p[0]]), batch, p[1], p[2], inv_hw,
                              reinterpret_cast<const float*>(W + po.rec.w_off),
                              reinterpret_cast<const float*>(W + po.rec.b_off), p[8],
                              reinterpret_cast<float*>(buf_dev[p[3]]), stream);
        break;
      }
      case OP_SE_SCALE:
        rc = se_scale_launch(view(p[0], p[1], p[2], batch), view(p[3], p[4], p[5], batch),
                             reinterpret_cast<const float*>(buf_dev[p[12]]),
                             reinterpret_cast<const float*>(W + po.rec.w_off),
                             reinterpret_cast<const float*>(W + po.rec.b_off), p[8], stream);
        break;
      case OP_SCALE:
        rc = channel_scale_launch(view(p[0], p[1], p[2], batch), view(p[3], p[4], p[5], batch),
                                  reinterpret_cast<const float*>(buf_dev[p[12]]), stream);
        break;
      case OP_COPY:
        rc = copy_channels_launch(view(p[0], p[1], p[2], batch), view(p[3], p[4], p[5], batch), stream);
        break;
      case OP_UPSAMPLE:
        rc = upsample2x_launch(view(p[0], p[1], p[2], batch), view(p[3], p[4], p[5], batch), stream);
        break;
      case OP_SHUFFLE:
        rc = shuffle_concat_launch(view(p[0], p[1], p[2], batch), view(p[12], p[13], p[2], batch),
                                   view(p[3], p[4], p[5], batch), p[14], p[15], stream);
        break;
      case OP_GATHER_CONV: {
        View res;
        if (p[6] >= 0) res = view(p[6], p[7], p[5], batch);
        rc = gather_conv_launch(view(p[0], p[1], p[2], batch), view(p[3], p[4], p[5], batch), res, W + po.rec.w_off,
                                reinterpret_cast<const float*>(W + po.rec.b_off), p[8], p[9], p[10], p[11], p[12] != 0,
                                stream);
        break;
      }
      case OP_YOLO_DECODE:
        rc = yolo_decode_launch(view(p[0], p[1], p[2], batch), view(p[12], p[13], p[14], batch),
                                reinterpret_cast<const float*>(W + po.rec.w_off),
                                reinterpret_cast<const float*>(W + po.rec.b_off),
                                reinterpret_cast<const float*>(W + po.rec.w2_off),
                                reinterpret_cast<const float*>(W + po.rec.b2_off), reinterpret_cast<float*>(buf_dev[p[3]]),
                                p[15], p[17], static_cast<float>(p[16]), stream);
        break;
      default:
        rc = fail(B200_ERR_BLOB, "engine: unknown op type " + std::to_string(po.rec.type));
    }
    if (rc != 0) return fail(rc, "op " + std::to_string(i) + " (type " + std::to_string(po.rec.type) + "): " + last_error_cstr());
  }
  return B200_OK;
}

static int capture_plan(Engine& e, BatchPlan& plan) {
  const int n = static_cast<int>(plan.ops.size());
  cudaGraph_t g = nullptr;
  B200_CUDA_OK(cudaStreamBeginCapture(e.capture_stream, cudaStreamCaptureModeThreadLocal));
  int rc = e.run_ops(plan, e.capture_stream, 0, n - 1);
  cudaError_t err = cudaStreamEndCapture(e.capture_stream, &g);
  if (rc != 0) {
    if (g) cudaGraphDestroy(g);
    return rc;
  }
  if (err != cudaSuccess) return fail(B200_ERR_CUDA, std::string("graph capture: ") + cudaGetErrorString(err));
  err = cudaGraphInstantiate(&plan.graph, g, 0);
  cudaGraphDestroy(g);
  if (err != cudaSuccess) return fail(B200_ERR_CUDA, std::string("graph instantiate: ") + cudaGetErrorString(err));
  return B200_OK;
}

int Engine::run(int batch, cudaStream_t stream) {
  if (batch <= 0 || batch > max_batch)
    return fail(B200_ERR_INVALID, "engine: batch " + std::to_string(batch) + " outside [1, " + std::to_string(max_batch) + "]");
  BatchPlan* plan = nullptr;
  B200_TRY(plan_for(batch, &plan));
  const int n = static_cast<int>(plan->ops.size());
  if (!use_graphs || !plan->warmed) {
    // first run of a batch size is eager: it sets kernel attributes and surfaces shape errors before any capture
    B200_TRY(run_ops(*plan, stream, 0, n - 1));
    plan->warmed = true;
    launches += n;
    return B200_OK;
  }
  if (!plan->graph) B200_TRY(capture_plan(*this, *plan));
  B200_CUDA_OK(cudaGraphLaunch(plan->graph, stream));
  launches += n;
  return B200_OK;
}

// Everything the first two calls of a batch size would do (descriptors, eager pass, graph capture), done up front on
// whatever the buffers hold, so that the first served batch of that size already replays a graph.
int Engine::warm(int batch, cudaStream_t stream) {
  if (batch <= 0 || batch > max_batch) return fail(B200_ERR_INVALID, "engine: warm-up batch outside [1, max_batch]");
  BatchPlan* plan = nullptr;
  B200_TRY(plan_for(batch, &plan));
  const int n = static_cast<int>(plan->ops.size());
  if (!plan->warmed) {
    B200_TRY(run_ops(*plan, stream, 0, n - 1));
    B200_CUDA_OK(cudaStreamSynchronize(stream));
    plan->warmed = true;
    launches += n;
  }
  if (use_graphs && !plan->graph) B200_TRY(capture_plan(*this, *plan));
  return B200_OK;
}

}  // namespace b200

// ================================================================================================ C ABI, engine level
struct b200_engine {
  Engine e;
};

extern "C" {

int b200_engine_create(const char* model_kind, const void* blob, size_t nbytes, int max_batch, int device,
                       b200_engine** out) {
  return b200_engine_create_ex(model_kind, blob, nbytes, max_batch, device, 0u, out);
}

int b200_engine_create_ex(const char* model_kind, const void* blob, size_t nbytes, int max_batch, int device,
                          unsigned flags, b200_engine** out) {
  if (!out) return fail(B200_ERR_INVALID, "b200_engine_create: out is null");
  *out = nullptr;
  b200_engine* h = new (std::nothrow) b200_engine;
  if (!h) return fail(B200_ERR_INVALID, "b200_engine_create: out of host memory");
  int rc = h->e.init(model_kind, blob, nbytes, max_batch, device, flags);
  if (rc != 0) {
    std::string msg = last_error_cstr();
    delete h;
    return fail(rc, msg);
  }
  *out = h;
  return B200_OK;
}

void b200_engine_destroy(b200_engine* h) { delete h; }

int b200_engine_warmup(b200_engine* h, const int* batches, int n, void* stream_) {
  if (!h || (n > 0 && !batches)) return fail(B200_ERR_INVALID, "b200_engine_warmup: null argument");
  B200_CUDA_OK(cudaSetDevice(h->e.device));
  for (int i = 0; i < n; ++i) B200_TRY(h->e.warm(batches[i], static_cast<cudaStream_t>(stream_)));
  B200_CUDA_OK(cudaStreamSynchronize(static_cast<cudaStream_t>(stream_)));
  return B200_OK;
}

int b200_engine_memory(b200_engine* h, size_t* device_bytes, size_t* activation_bytes) {
  if (!h) return fail(B200_ERR_INVALID, "null engine");
  if (device_bytes) *device_bytes = h->e.device_bytes;
  if (activation_bytes) *activation_bytes = h->e.arena_bytes + h->e.vec_bytes;
  return B200_OK;
}

int b200_plan_memory(const void* blob, size_t nbytes, int max_batch, unsigned flags, size_t* device_bytes) {
  if (!device_bytes) return fail(B200_ERR_INVALID, "b200_plan_memory: null output");
  Engine e;
  B200_TRY(e.parse(nullptr, blob, nbytes, max_batch));
  std::vector<size_t> offsets;
  size_t arena = 0, vec = 0;
  e.plan_arena((flags & B200_ENGINE_KEEP_BUFFERS) == 0, &offsets, &arena, &vec);
  *device_bytes = e.fixed_bytes() + arena + vec;
  return B200_OK;
}

int b200_engine_info(b200_engine* h, int* in_h, int* in_w, int* out_dim, int* max_batch, int* n_ops) {
  if (!h) return fail(B200_ERR_INVALID, "null engine");
  if (in_h) *in_h = h->e.hdr.in_h;
  if (in_w) *in_w = h->e.hdr.in_w;
  if (out_dim) *out_dim = h->e.hdr.out_dim;
  if (max_batch) *max_batch = h->e.max_batch;
  if (n_ops) *n_ops = static_cast<int>(h->e.ops.size());
  return B200_OK;
}

int b200_engine_set_graphs(b200_engine* h, int enabled) {
  if (!h) return fail(B200_ERR_INVALID, "null engine");
  h->e.use_graphs = enabled != 0;
  return B200_OK;
}

long long b200_engine_launch_count(b200_engine* h) { return h ? h->e.launches : 0; }

int b200_engine_buffer(b200_engine* h, int buf, void** dev_ptr, size_t* bytes_per_image) {
  if (!h || buf < -2 || buf >= static_cast<int>(h->e.bufs.size())) return fail(B200_ERR_INVALID, "bad buffer id");
  if (buf == -1) buf = h->e.hdr.in_buf;
  if (buf == -2) buf = h->e.hdr.out_buf;
  if (dev_ptr) *dev_ptr = h->e.buf_dev[buf];
  if (bytes_per_image) *bytes_per_image = h->e.buf_bytes_per_image(buf);
  return B200_OK;
}

int b200_engine_profile(b200_engine* h, int batch, float* ms_per_op, double* flops_per_op, int n_ops, void* stream_) {
  if (!h) return fail(B200_ERR_INVALID, "null engine");
  Engine& e = h->e;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (n_ops != static_cast<int>(e.ops.size())) return fail(B200_ERR_INVALID, "b200_engine_profile: n_ops mismatch");
  if (batch <= 0 || batch > e.max_batch) return fail(B200_ERR_INVALID, "b200_engine_profile: bad batch");
  BatchPlan* plan = nullptr;
  B200_TRY(e.plan_for(batch, &plan));
  std::vector<cudaEvent_t> ev(n_ops + 1);
  for (auto& x : ev) B200_CUDA_OK(cudaEventCreate(&x));
  // The launches and the event records between them are captured into one graph, so the gaps between events are kernel
  // time and not the host's launch rate (eager launches put a ~10-15 us floor under every small op).
  static int graph_profile = -1;
  if (graph_profile < 0) {
    const char* g = getenv("B200_PROFILE_GRAPH");
    graph_profile = (g && g[0] == '0') ? 0 : 1;
  }
  if (graph_profile) {
    cudaStream_t cs = e.capture_stream;
    B200_CUDA_OK(cudaStreamSynchronize(stream));
    B200_CUDA_OK(cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal));
    int rc = B200_OK;
    cudaError_t ce = cudaEventRecordWithFlags(ev[0], cs, cudaEventRecordExternal);   // a real event-record node
    for (int i = 0; i < n_ops && rc == B200_OK && ce == cudaSuccess; ++i) {
      rc = e.run_ops(*plan, cs, i, i);
      if (rc == B200_OK) ce = cudaEventRecordWithFlags(ev[i + 1], cs, cudaEventRecordExternal);
    }
    cudaGraph_t graph = nullptr;
    const cudaError_t ee = cudaStreamEndCapture(cs, &graph);
    if (rc != B200_OK) { if (graph) cudaGraphDestroy(graph); return rc; }
    B200_CUDA_OK(ce);
    B200_CUDA_OK(ee);
    cudaGraphExec_t exec = nullptr;
    B200_CUDA_OK(cudaGraphInstantiate(&exec, graph, 0));
    B200_CUDA_OK(cudaGraphLaunch(exec, cs));     // warm-up replay
    B200_CUDA_OK(cudaGraphLaunch(exec, cs));
    B200_CUDA_OK(cudaStreamSynchronize(cs));
    cudaGraphExecDestroy(exec);
    cudaGraphDestroy(graph);
    e.launches += 2 * n_ops;
  } else {
  B200_CUDA_OK(cudaEventRecord(ev[0], stream));
  for (int i = 0; i < n_ops; ++i) {
    int rc = e.run_ops(*plan, stream, i, i);
    if (rc != 0) return rc;
    B200_CUDA_OK(cudaEventRecord(ev[i + 1], stream));
  }
  B200_CUDA_OK(cudaStreamSynchronize(stream));
  e.launches += n_ops;
  }
  for (int i = 0; i < n_ops; ++i) {
    B200_CUDA_OK(cudaEventElapsedTime(&ms_per_op[i], ev[i], ev[i + 1]));
    if (flops_per_op) flops_per_op[i] = plan->ops[i].rec.type == OP_CONV ? plan->ops[i].conv.flops : 0.0;
  }
  for (auto& x : ev) cudaEventDestroy(x);
  return B200_OK;
}

int b200_engine_read_buffer(b200_engine* h, int buf, int batch, void* host_dst) {
  if (!h || buf < -2 || buf >= static_cast<int>(h->e.bufs.size())) return fail(B200_ERR_INVALID, "bad buffer id");
  if (batch <= 0 || batch > h->e.max_batch || !host_dst) return fail(B200_ERR_INVALID, "bad batch / destination");
  if (buf == -1) buf = h->e.hdr.in_buf;
  if (buf == -2) buf = h->e.hdr.out_buf;
  B200_CUDA_OK(cudaDeviceSynchronize());
  B200_CUDA_OK(cudaMemcpy(host_dst, h->e.buf_dev[buf], h->e.buf_bytes_per_image(buf) * batch, cudaMemcpyDeviceToHost));
  return B200_OK;
}

int b200_collate(b200_engine* h, const void* packed, const uint64_t* offsets, const int32_t* hw, int n, int padded_batch,
                 int mode, int src_on_device, size_t packed_bytes, void* out_nhwc_bf16, void* stream_) {
  if (!h) return fail(B200_ERR_INVALID, "null engine");
  Engine& e = h->e;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (n < 0 || padded_batch < n || padded_batch > e.max_batch || padded_batch <= 0)
    return fail(B200_ERR_INVALID, "b200_collate: need 0 <= n <= padded_batch <= max_batch");
  if (n > 0 && (!packed || !offsets || !hw)) return fail(B200_ERR_INVALID, "b200_collate: null input");
  // the item table is the source of an async copy: rotate through pinned slots, waiting only if a slot is still in flight
  const unsigned slot = e.item_slot++ % Engine::ITEM_SLOTS;
  B200_CUDA_OK(cudaEventSynchronize(e.item_events[slot]));
  CollateItem* items_host = e.items_host + size_t(slot) * e.max_batch;
  for (int i = 0; i < n; ++i) {
    items_host[i].offset = offsets[i];
    items_host[i].h = hw[2 * i];
    items_host[i].w = hw[2 * i + 1];
    if (hw[2 * i] <= 0 || hw[2 * i + 1] == 0) return fail(B200_ERR_INVALID, "b200_collate: empty image");
    const uint64_t wpx = hw[2 * i + 1] < 0 ? uint64_t(-int64_t(hw[2 * i + 1])) : uint64_t(hw[2 * i + 1]);   // w < 0: grayscale
    if (offsets[i] + uint64_t(hw[2 * i]) * wpx * (hw[2 * i + 1] < 0 ? 1 : 3) > packed_bytes)
      return fail(B200_ERR_INVALID, "b200_collate: image " + std::to_string(i) + " exceeds the packed buffer");
    if (mode == COLLATE_IDENTITY && (hw[2 * i] != e.hdr.in_h || hw[2 * i + 1] != e.hdr.in_w))
      return fail(B200_ERR_INVALID, "b200_collate: image " + std::to_string(i) + " is not the model input size");
  }
  const uint8_t* src = static_cast<const uint8_t*>(packed);
  if (!src_on_device && n > 0) {
    if (packed_bytes > e.stage_bytes) {
      if (e.stage_dev) B200_CUDA_OK(cudaFree(e.stage_dev));
      e.device_bytes -= e.stage_bytes;
      e.stage_dev = nullptr;
      e.stage_bytes = 0;
      const size_t want = packed_bytes + packed_bytes / 4 + 4096;
      B200_CUDA_OK(cudaMalloc(&e.stage_dev, want));
      e.device_bytes += want;
      e.stage_bytes = want;
    }
    B200_CUDA_OK(cudaMemcpyAsync(e.stage_dev, packed, packed_bytes, cudaMemcpyHostToDevice, stream));
    src = e.stage_dev;
  }
  if (n > 0) {
    B200_CUDA_OK(cudaMemcpyAsync(e.items_dev, items_host, sizeof(CollateItem) * n, cudaMemcpyHostToDevice, stream));
    B200_CUDA_OK(cudaEventRecord(e.item_events[slot], stream));
  }
  void* out = out_nhwc_bf16 ? out_nhwc_bf16 : e.buf_dev[e.hdr.in_buf];
  B200_TRY(collate_u8_launch(src, e.items_dev, items_host, n, padded_batch, e.hdr.in_h, e.hdr.in_w, mode,
                             e.hdr.pre_fill, e.lut_dev, out, stream));
  e.launches += 1;
  return B200_OK;
}

int b200_collate_f32(b200_engine* h, const float* stacked_chw, int n, int padded_batch, int src_on_device,
                     void* out_nhwc_bf16, void* stream_) {
  if (!h) return fail(B200_ERR_INVALID, "null engine");
  Engine& e = h->e;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (n < 0 || padded_batch < n || padded_batch > e.max_batch || padded_batch <= 0)
    return fail(B200_ERR_INVALID, "b200_collate_f32: need 0 <= n <= padded_batch <= max_batch");
  const size_t per = size_t(3) * e.hdr.in_h * e.hdr.in_w;
  const float* src = stacked_chw;
  if (!src_on_device && n > 0) {
    if (!e.stage_f32_dev) {
      B200_CUDA_OK(cudaMalloc(&e.stage_f32_dev, per * sizeof(float) * e.max_batch));
      e.device_bytes += per * sizeof(float) * e.max_batch;
    }
    B200_CUDA_OK(cudaMemcpyAsync(e.stage_f32_dev, stacked_chw, per * sizeof(float) * n, cudaMemcpyHostToDevice, stream));
    src = e.stage_f32_dev;
  }
  void* out = out_nhwc_bf16 ? out_nhwc_bf16 : e.buf_dev[e.hdr.in_buf];
  B200_TRY(collate_f32chw_launch(src, n, padded_batch, e.hdr.in_h, e.hdr.in_w, out, stream));
  e.launches += 1;
  return B200_OK;
}

int b200_decode_yolo(b200_engine* h, const float* head, int batch, float conf, float iou, int max_det, float* boxes,
                     int32_t* counts, void* stream_) {
  if (!h) return fail(B200_ERR_INVALID, "null engine");
  Engine& e = h->e;
  if (batch <= 0 || batch > e.max_batch || !boxes || !counts) return fail(B200_ERR_INVALID, "b200_decode_yolo: bad arguments");
  if (e.hdr.out_dim % 84 != 0) return fail(B200_ERR_INVALID, "b200_decode_yolo: engine output is not a [84][A] detection head");
  const float* pred = head ? head : reinterpret_cast<const float*>(e.buf_dev[e.hdr.out_buf]);
  B200_TRY(yolo_nms_launch(pred, batch, 80, e.hdr.out_dim / 84, conf, iou, max_det, boxes, counts,
                           static_cast<cudaStream_t>(stream_)));
  e.launches += 1;
  return B200_OK;
}

int b200_infer(b200_engine* h, const void* in_nhwc_bf16, int batch, float* out, void* stream_) {
  if (!h) return fail(B200_ERR_INVALID, "null engine");
  Engine& e = h->e;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (batch <= 0 || batch > e.max_batch) return fail(B200_ERR_INVALID, "b200_infer: batch outside [1, max_batch]");
  char* in_buf = e.buf_dev[e.hdr.in_buf];
  if (in_nhwc_bf16 && in_nhwc_bf16 != in_buf)
    B200_CUDA_OK(cudaMemcpyAsync(in_buf, in_nhwc_bf16, e.buf_bytes_per_image(e.hdr.in_buf) * batch,
                                 cudaMemcpyDeviceToDevice, stream));
  B200_TRY(e.run(batch, stream));
  if (out)
    B200_CUDA_OK(cudaMemcpyAsync(out, e.buf_dev[e.hdr.out_buf], size_t(e.hdr.out_dim) * sizeof(float) * batch,
                                 cudaMemcpyDefault, stream));
  return B200_OK;
}

}  // extern "C"
