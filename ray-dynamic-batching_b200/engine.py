"""B200Model: the callable that takes the place of a torchvision nn.Module in the reference's `model_registry` /
`GPUWorker.models[name]` (293-project/src/scheduler.py:40-44, 441-450).

    model = B200Model("resnet50", module)         # lowering + blob + b200_engine_create
    logits = model(inputs)                        # inputs: [B,3,H,W] fp32 (cuda or cpu), like the reference call site
    logits = model.infer_u8(list_of_uint8_HWC)    # engine-native path: fused collate + preprocess on the GPU

Everything numeric happens in libb200infer.so; torch is used for device memory, streams and pinned host buffers only.
"""
import ctypes

import numpy as np
import torch

from . import lib as _lib
from . import lowering as _lowering

COLLATE_IDENTITY, COLLATE_PAD, COLLATE_RESIZE, COLLATE_LETTERBOX, COLLATE_RESIZE_AA, COLLATE_RESIZE_SHORT_AA = 0, 1, 2, 3, 4, 5


def default_buckets(max_batch):
    """1, 2, 4, 8, then steps of 8 to 64, of 32 to 256, of 64 beyond: padding waste stays under ~25 % from 8 up, under
    ~12 % from 64 up, and a full-size engine (512) prepares 26 graphs."""
    b = [1, 2, 4, 8]
    b += list(range(16, 65, 8)) + list(range(96, 257, 32)) + list(range(320, 1025, 64))
    return [x for x in b if x <= max_batch] + ([max_batch] if max_batch not in b else [])


def plan_memory_bytes(blob, max_batch, keep_buffers=False):
    """Device bytes an engine created from `blob` at `max_batch` holds (no device needed)."""
    out = ctypes.c_size_t()
    _lib.check(_lib.load().b200_plan_memory(blob, len(blob), int(max_batch), _lib.ENGINE_KEEP_BUFFERS if keep_buffers else 0,
                                            ctypes.byref(out)), "b200_plan_memory")
    return out.value


class B200Model:
    is_b200_engine = True

    def __init__(self, kind, module=None, blob=None, max_batch=256, device=0, use_graphs=True, in_hw=None,
                 keep_buffers=False, buckets=None):
        if not torch.cuda.is_available():
            raise _lib.B200Error("B200Model needs a CUDA device (sm_100); there is no CPU fallback")
        self.lib = _lib.load()
        self.kind = kind
        self.device = torch.device("cuda", device)
        if blob is None:
            if module is None:
                raise ValueError("need a module to lower or a prebuilt blob")
            pb = _lowering.lower(kind, module, **({} if in_hw is None else {'in_hw': tuple(in_hw)}))
            blob = pb.to_blob()
            self.out_logical = pb.out_logical
            self.op_names = list(pb.names)
        else:
            self.out_logical = None
            self.op_names = None
        self.blob = blob
        self.max_batch = max_batch
        self._h = ctypes.c_void_p()
        torch.cuda.set_device(self.device)
        flags = _lib.ENGINE_KEEP_BUFFERS if keep_buffers else 0
        _lib.check(self.lib.b200_engine_create_ex(kind.encode(), blob, len(blob), max_batch, device, flags,
                                                  ctypes.byref(self._h)), "b200_engine_create")
        ih, iw, od, mb, nops = (ctypes.c_int() for _ in range(5))
        _lib.check(self.lib.b200_engine_info(self._h, ih, iw, od, mb, nops))
        self.in_h, self.in_w, self.out_dim, self.n_ops = ih.value, iw.value, od.value, nops.value
        if self.out_logical is None:
            self.out_logical = self.out_dim
        if not use_graphs:
            _lib.check(self.lib.b200_engine_set_graphs(self._h, 0))
        self._pinned = None
        self._slots = None
        self._slot_i = 0
        # detection models: {"conf", "iou", "max_det"} makes the asynchronous path run the class-aware NMS on the device
        # and hand back detections [n, max_det, 6]; None hands back the raw [84][A] head.  Set before the first async call.
        self.nms = None
        self.timing = {"batches": 0, "h2d_ms": 0.0, "gpu_ms": 0.0}
        # batch-size buckets: a served batch of n requests is padded (zero images) to the next bucket, so the engine only
        # ever runs the sizes it captured graphs for at load.  None = every size runs as is (one graph per distinct size).
        self.buckets = None
        if buckets is not None:
            self.set_buckets(default_buckets(max_batch) if buckets == "auto" else buckets)

    # ---- lifecycle (mirrors .eval()/.to()/.cpu() so the reference's load/unload code keeps working) ----
    def eval(self):
        return self

    def to(self, *_a, **_k):
        return self

    def cpu(self):
        return self

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self.lib.b200_engine_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- buckets / memory ----
    def set_buckets(self, buckets, warm=True):
        b = sorted({int(x) for x in buckets if 0 < int(x) <= self.max_batch} | {self.max_batch})
        self.buckets = b
        if warm:
            self.warmup(b)

    def bucket_for(self, n):
        """Smallest prepared batch size >= n (n itself when the engine has no buckets)."""
        if not self.buckets:
            return n
        import bisect
        i = bisect.bisect_left(self.buckets, n)
        if i == len(self.buckets):
            raise _lib.B200Error("batch %d exceeds engine max_batch %d" % (n, self.max_batch))
        return self.buckets[i]

    def warmup(self, batches):
        arr = (ctypes.c_int * len(batches))(*[int(x) for x in batches])
        _lib.check(self.lib.b200_engine_warmup(self._h, arr, len(batches), self._stream()), "b200_engine_warmup")

    def memory_bytes(self):
        """(device bytes the engine holds now, the activation part of it)."""
        tot, act = ctypes.c_size_t(), ctypes.c_size_t()
        _lib.check(self.lib.b200_engine_memory(self._h, ctypes.byref(tot), ctypes.byref(act)))
        return tot.value, act.value

    # ---- helpers ----
    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def launch_count(self):
        return int(self.lib.b200_engine_launch_count(self._h))

    def buffer(self, buf_id, batch):
        """Debug copy of an engine activation buffer: uint8 host tensor [batch, bytes_per_image] (-1 input, -2 output)."""
        p, nb = ctypes.c_void_p(), ctypes.c_size_t()
        _lib.check(self.lib.b200_engine_buffer(self._h, buf_id, ctypes.byref(p), ctypes.byref(nb)))
        host = torch.empty(batch * nb.value, dtype=torch.uint8)
        _lib.check(self.lib.b200_engine_read_buffer(self._h, buf_id, batch, _lib.ptr(host)), "b200_engine_read_buffer")
        return host.view(batch, nb.value)

    def profile_ops(self, batch):
        """One eager pass with CUDA events around every op: returns (ms_per_op, flops_per_op) as numpy arrays."""
        ms = np.zeros(self.n_ops, dtype=np.float32)
        fl = np.zeros(self.n_ops, dtype=np.float64)
        _lib.check(self.lib.b200_engine_profile(self._h, batch, ms.ctypes.data_as(ctypes.c_void_p),
                                                fl.ctypes.data_as(ctypes.c_void_p), self.n_ops, self._stream()),
                   "b200_engine_profile")
        return ms, fl

    # ---- reference call site: outputs = model(inputs) ----
    def __call__(self, inputs):
        return self.forward(inputs)

    def forward(self, inputs):
        if inputs.dim() != 4 or inputs.shape[1] != 3 or inputs.shape[2] != self.in_h or inputs.shape[3] != self.in_w:
            raise _lib.B200Error("expected [B,3,%d,%d] input, got %s" % (self.in_h, self.in_w, tuple(inputs.shape)))
        b = inputs.shape[0]
        if b > self.max_batch:
            raise _lib.B200Error("batch %d exceeds engine max_batch %d" % (b, self.max_batch))
        x = inputs.contiguous().float()
        on_dev = 1 if x.is_cuda else 0
        s = self._stream()
        _lib.check(self.lib.b200_collate_f32(self._h, _lib.ptr(x), b, b, on_dev, None, s), "b200_collate_f32")
        out = torch.empty(b, self.out_dim, device=self.device, dtype=torch.float32)
        _lib.check(self.lib.b200_infer(self._h, None, b, _lib.ptr(out), s), "b200_infer")
        return out[:, :self.out_logical]

    # ---- engine-native path ----
    def pack_u8(self, images):
        """Packs uint8 HWC images into one pinned buffer (16-byte aligned slots). Returns (packed, offsets, hw)."""
        n = len(images)
        offsets = np.zeros(n, dtype=np.uint64)
        hw = np.zeros((n, 2), dtype=np.int32)
        total = 0
        for i, im in enumerate(images):
            offsets[i] = total
            gray = im.dim() == 2 or im.shape[2] == 1            # single-channel request: width passed negative (L -> RGB)
            hw[i] = (im.shape[0], -im.shape[1] if gray else im.shape[1])
            total += (im.numel() + 15) // 16 * 16
        if self._pinned is None or self._pinned.numel() < total:
            self._pinned = torch.empty(max(total, 1), dtype=torch.uint8, pin_memory=True)
        for i, im in enumerate(images):
            o = int(offsets[i])
            self._pinned[o:o + im.numel()].copy_(im.reshape(-1))
        return self._pinned, offsets, hw, total

    def collate_u8(self, packed, offsets, hw, n, padded_batch=None, mode=COLLATE_IDENTITY, total_bytes=None):
        padded_batch = padded_batch or n
        on_dev = 1 if packed.is_cuda else 0
        total = int(total_bytes if total_bytes is not None else packed.numel())
        _lib.check(self.lib.b200_collate(self._h, _lib.ptr(packed), offsets.ctypes.data_as(ctypes.c_void_p),
                                         hw.ctypes.data_as(ctypes.c_void_p), n, padded_batch, mode, on_dev, total, None,
                                         self._stream()), "b200_collate")

    def infer_u8(self, images, padded_batch=None, mode=COLLATE_IDENTITY, out=None):
        """images: list of uint8 HWC tensors (host). Returns fp32 logits [n, out] on the device (padding rows dropped)."""
        n = len(images)
        packed, offsets, hw, total = self.pack_u8(images)
        pb = padded_batch or n
        self.collate_u8(packed, offsets, hw, n, pb, mode, total)
        if out is None:
            out = torch.empty(pb, self.out_dim, device=self.device, dtype=torch.float32)
        _lib.check(self.lib.b200_infer(self._h, None, pb, _lib.ptr(out), self._stream()), "b200_infer")
        return out[:n, :self.out_logical]

    # ---- pipelined host path: the H2D copy of batch i+1 overlaps the forward of batch i ----
    N_SLOTS = 3

    class _Slot:
        def __init__(self, model):
            self.staging = None
            self.nms = model.nms
            if self.nms:          # detections leave the GPU, not the [84][A] head (2.8 MB per 640 x 640 image)
                md = int(self.nms["max_det"])
                self.boxes_dev = torch.zeros(model.max_batch, md, 6, device=model.device, dtype=torch.float32)
                self.counts_dev = torch.zeros(model.max_batch, device=model.device, dtype=torch.int32)
                self.out_host = torch.empty(model.max_batch, md, 6, dtype=torch.float32, pin_memory=True)
                self.counts_host = torch.empty(model.max_batch, dtype=torch.int32, pin_memory=True)
            else:
                self.out_host = torch.empty(model.max_batch, model.out_dim, dtype=torch.float32, pin_memory=True)
            self.copy_start = torch.cuda.Event(enable_timing=True)
            self.copied = torch.cuda.Event(enable_timing=True)
            self.done = torch.cuda.Event(enable_timing=True)
            self.n = 0
            self.busy = False
            self.out_logical = model.out_logical
            self.model = model

        def result(self):
            """Blocks until this batch is finished; returns the fp32 outputs [n, out] -- a VIEW of this slot's pinned
            buffer, valid until the slot is handed out again (N_SLOTS - 1 submissions later): copy it to keep it.
            With NMS on (model.nms): [n, max_det, 6] = x1, y1, x2, y2, score, class; rows past an image's count are zero."""
            self.done.synchronize()
            self.busy = False
            t = self.model.timing                   # device-side split of this batch: upload, then wait + collate + forward + D2H
            t["batches"] += 1
            t["h2d_ms"] += self.copy_start.elapsed_time(self.copied)
            t["gpu_ms"] += self.copied.elapsed_time(self.done)
            if self.nms:
                return self.out_host[:self.n]
            return self.out_host[:self.n, :self.out_logical]

    def prepare_async(self, total_bytes=None):
        """Allocates everything the asynchronous path needs (pinned output slots, device staging for a full batch, the copy
        stream) now, so that the first served batch does not pay for it: an actor calls this while it loads."""
        total = int(total_bytes or self.max_batch * self.in_h * self.in_w * 3)
        for _ in range(self.N_SLOTS):
            self._next_slot(total)
        self._slot_i = 0

    def _next_slot(self, total):
        if self._slots is None:
            self._slots = [B200Model._Slot(self) for _ in range(self.N_SLOTS)]
            self._copy_stream = torch.cuda.Stream(self.device)
        slot = self._slots[self._slot_i % self.N_SLOTS]
        self._slot_i += 1
        if slot.busy:
            raise _lib.B200Error("more than %d batches in flight: collect .result() of earlier batches first"
                                 % (self.N_SLOTS - 1))
        slot.done.synchronize()                      # the previous user of this slot has left the GPU
        if slot.staging is None or slot.staging.numel() < total:
            slot.staging = torch.empty(total + total // 4 + 4096, dtype=torch.uint8, device=self.device)
        return slot

    def _launch_slot(self, slot, offsets, hw, n, padded_batch, mode, total):
        compute = torch.cuda.current_stream(self.device)
        compute.wait_event(slot.copied)
        pb = padded_batch or n
        self.collate_u8(slot.staging, offsets, hw, n, pb, mode, total)
        if slot.nms:
            _lib.check(self.lib.b200_infer(self._h, None, pb, None, self._stream()), "b200_infer")
            slot.boxes_dev[:pb].zero_()
            _lib.check(self.lib.b200_decode_yolo(self._h, None, pb, float(slot.nms["conf"]), float(slot.nms["iou"]),
                                                 int(slot.nms["max_det"]), _lib.ptr(slot.boxes_dev),
                                                 _lib.ptr(slot.counts_dev), self._stream()), "b200_decode_yolo")
            slot.out_host[:pb].copy_(slot.boxes_dev[:pb], non_blocking=True)
            slot.counts_host[:pb].copy_(slot.counts_dev[:pb], non_blocking=True)
        else:
            _lib.check(self.lib.b200_infer(self._h, None, pb, _lib.ptr(slot.out_host), self._stream()), "b200_infer")
        slot.n = n
        slot.busy = True
        slot.done.record(compute)
        return slot

    def infer_packed_async(self, packed_host, offsets, hw, n, padded_batch=None, mode=COLLATE_IDENTITY, total_bytes=None):
        """Asynchronous engine-native call: `packed_host` (pinned uint8) -> H2D on a side stream -> collate -> forward ->
        D2H into a pinned slot.  Returns a handle whose .result() waits for this batch only.  N_SLOTS - 1 batches may be
        in flight: the copy of one overlaps the compute of the other."""
        total = int(total_bytes if total_bytes is not None else packed_host.numel())
        slot = self._next_slot(total)
        with torch.cuda.stream(self._copy_stream):
            slot.copy_start.record(self._copy_stream)
            slot.staging[:total].copy_(packed_host[:total], non_blocking=True)
            slot.copied.record(self._copy_stream)
        return self._launch_slot(slot, offsets, hw, n, padded_batch, mode, total)

    def infer_runs_async(self, slab, runs, offsets, hw, n, padded_batch=None, mode=COLLATE_IDENTITY, total_bytes=None):
        """Same for a batch that sits in a page-locked request slab as a few contiguous runs of slots:
        runs = [(slab byte offset, byte count, staging byte offset)], one DMA each; `offsets` are the images' positions
        in the staging buffer."""
        total = int(total_bytes)
        slot = self._next_slot(total)
        with torch.cuda.stream(self._copy_stream):
            slot.copy_start.record(self._copy_stream)
            for src, nbytes, dst in runs:
                slot.staging[dst:dst + nbytes].copy_(slab[src:src + nbytes], non_blocking=True)
            slot.copied.record(self._copy_stream)
        return self._launch_slot(slot, offsets, hw, n, padded_batch, mode, total)

    def decode_yolo(self, batch, conf=0.25, iou=0.45, max_det=300, head=None):
        """Class-aware NMS on the device over the [batch, 84, A] head output of the last forward (or `head`).
        Returns (boxes [batch, max_det, 6] = x1,y1,x2,y2,score,class ; counts [batch]) as device tensors."""
        boxes = torch.zeros(batch, max_det, 6, device=self.device, dtype=torch.float32)
        counts = torch.zeros(batch, device=self.device, dtype=torch.int32)
        _lib.check(self.lib.b200_decode_yolo(self._h, _lib.ptr(head), batch, conf, iou, max_det, _lib.ptr(boxes),
                                             _lib.ptr(counts), self._stream()), "b200_decode_yolo")
        return boxes, counts

    def infer_resident(self, batch, out):
        """Forward over whatever the engine input buffer holds (inputs already resident in HBM)."""
        _lib.check(self.lib.b200_infer(self._h, None, batch, _lib.ptr(out), self._stream()), "b200_infer")
        return out
