"""B200-native batched-forward engine behind the ray-dynamic-batching scheduler interface.

Import as ``rdb_b200`` (see ``rdb_b200.py`` at the repo root: the directory name carries a hyphen).

  rdb_b200.lib            ctypes binding of libb200infer.so (include/b200infer.h)
  rdb_b200.engine         B200Model: the callable that replaces a torchvision module in the reference's model registry
  rdb_b200.request_queue  B200RequestQueue / BatchRequest (reference RequestQueue surface)
  rdb_b200.worker         B200GPUWorker (reference GPUWorker surface)
  rdb_b200.lowering       nn.Module -> plan + weights blob
  rdb_b200.synth          seeded synthetic weights / requests
"""
from . import lib  # noqa: F401

__all__ = ["lib"]
