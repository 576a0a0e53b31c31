"""ZMQ ingress: the reference's draft wire format (293-project/src/milind-code/request_simulator.py:33-41, receiver
milind-code/scheduler.py:74-100) carrying the image itself instead of a path on the receiver's disk.

The draft sends one JSON object per request over PUSH/PULL on tcp://*:5555,
    {"timestamp", "model_name", "request_id", "SLO", "image_path"},
and its receiver ends up submitting `torch.rand(3, 224, 224)` (the image load is commented out, milind-code/scheduler.py:91-92).
Here a request is a two-frame message:
    frame 0  the same JSON header, `image_path` optional, plus "shape": [H, W, C] (C = 3, or 1 for grayscale) and "dtype": "uint8"
    frame 1  the raw pixels, H * W * C bytes, row-major HWC
and the receiver copies frame 1 once, straight into the model's request slot in shared memory
(B200Scheduler.submit_request -> b200_ring_push); nothing is decoded, converted to fp32 or pickled on the way.  A
single-frame message in the draft's exact format is still accepted when it names a readable `image_path` and PIL is there to
decode it (JPEG decode is the step before the hot path).

    ing = ZmqIngress(scheduler, "tcp://127.0.0.1:5555"); ing.start(); ...; ing.stop()
    sender = RequestSender("tcp://127.0.0.1:5555"); sender.send("resnet", "resnet17", image_u8_hwc, slo_ms=2000)
"""
import json
import logging
import threading
import time

import numpy as np


def _zmq():
    import zmq            # pyzmq is the reference project's only requirement (293-project/requirements.txt)
    return zmq


def encode_request(model_name, request_id, image, slo_ms=0.0, timestamp=None):
    """[header frame, pixel frame] for a uint8 HWC (or HW) numpy array / torch tensor."""
    arr = image.numpy() if hasattr(image, "numpy") else np.asarray(image)
    if arr.dtype != np.uint8 or arr.ndim not in (2, 3):
        raise TypeError("requests carry uint8 HWC (or HW) images, got %s %s" % (arr.dtype, arr.shape))
    if arr.ndim == 2:
        arr = arr[:, :, None]
    arr = np.ascontiguousarray(arr)
    header = {"timestamp": time.time() if timestamp is None else timestamp, "model_name": model_name,
              "request_id": request_id, "SLO": slo_ms, "shape": list(arr.shape), "dtype": "uint8"}
    return [json.dumps(header).encode(), memoryview(arr).cast("B")]


def decode_request(frames):
    """(header dict, uint8 HWC array viewing frame 1) -- or, for the draft's single JSON frame, the decoded file."""
    header = json.loads(bytes(frames[0]))
    if len(frames) >= 2:
        h, w, c = header["shape"]
        buf = frames[1]
        arr = np.frombuffer(buf, dtype=np.uint8)
        if arr.size != h * w * c:
            raise ValueError("request %r: %d payload bytes for shape %s" % (header.get("request_id"), arr.size, header["shape"]))
        return header, arr.reshape(h, w, c)
    path = header.get("image_path")
    if not path:
        raise ValueError("request %r carries neither pixels nor an image_path" % header.get("request_id"))
    from PIL import Image
    img = Image.open(path)
    arr = np.asarray(img if img.mode in ("RGB", "L") else img.convert("RGB"), dtype=np.uint8)
    return header, (arr[:, :, None] if arr.ndim == 2 else arr)


class RequestSender:
    """Client side (RequestSimulator's socket, request_simulator.py:12-16): PUSH, connects to the ingress."""

    def __init__(self, endpoint="tcp://127.0.0.1:5555", hwm=4096):
        zmq = _zmq()
        self.context = zmq.Context.instance()
        self.socket = self.context.socket(zmq.PUSH)
        self.socket.setsockopt(zmq.SNDHWM, hwm)
        self.socket.setsockopt(zmq.LINGER, 2000)
        self.socket.connect(endpoint)

    def send(self, model_name, request_id, image, slo_ms=0.0):
        self.socket.send_multipart(encode_request(model_name, request_id, image, slo_ms), copy=False)

    def close(self):
        self.socket.close()


class ZmqIngress:
    """Receiver (RequestHandle.listen_for_requests / process_request, milind-code/scheduler.py:74-100): PULL, binds, and
    hands every request to `scheduler.submit_request(model_name, request_id, image)`."""

    def __init__(self, scheduler, endpoint="tcp://127.0.0.1:5555", hwm=4096):
        zmq = _zmq()
        self.scheduler = scheduler
        self.endpoint = endpoint
        self.context = zmq.Context.instance()
        self.socket = self.context.socket(zmq.PULL)
        self.socket.setsockopt(zmq.RCVHWM, hwm)
        self.socket.bind(endpoint)
        self.endpoint = self.socket.getsockopt(zmq.LAST_ENDPOINT).decode()     # the actual port for "tcp://127.0.0.1:*"
        self._logger = logging.getLogger("ZmqIngress")
        self.thread = None
        self.running = False
        self.received = self.accepted = self.rejected = self.malformed = 0

    def process_request(self, frames) -> bool:
        """milind-code/scheduler.py:82-100: False when the model has no queue, the queue is full, or the message is bad."""
        self.received += 1
        try:
            header, image = decode_request(frames)
            ok = self.scheduler.submit_request(header["model_name"], str(header["request_id"]), image)
        except Exception as e:  # noqa: BLE001 - the reference logs and returns False
            self._logger.error("Error adding request: %s", e)
            self.malformed += 1
            return False
        if ok:
            self.accepted += 1
        else:
            self.rejected += 1
        return ok

    def listen_for_requests(self):
        zmq = _zmq()
        poller = zmq.Poller()
        poller.register(self.socket, zmq.POLLIN)
        while self.running:
            if not poller.poll(50):
                continue
            while True:
                try:
                    frames = self.socket.recv_multipart(zmq.NOBLOCK, copy=False)
                except zmq.Again:
                    break
                self.process_request([f.buffer for f in frames])

    def start(self):
        if self.thread is None:
            self.running = True
            self.thread = threading.Thread(target=self.listen_for_requests, daemon=True)
            self.thread.start()

    def stop(self):
        self.running = False
        if self.thread is not None:
            self.thread.join(timeout=2.0)
            self.thread = None
        self.socket.close(0)

    def stats(self):
        return {"received": self.received, "accepted": self.accepted, "rejected": self.rejected, "malformed": self.malformed}
