"""swiftsnails_b200 — a Blackwell-native parameter server with the capabilities of SwiftSnails.

Layers:
  * ``csrc/core``     C++17 host core: buffers, config, hash fragments, ZeroMQ transfer,
                      CPU sparse table, master / server / worker roles (control plane + CPU path)
  * ``csrc/kernels``  hand-written sm_100a CUDA kernels behind the C ABI of ``include/ssb200.h``
  * ``swiftsnails_b200.ops``       Python wrappers over the kernels
  * ``swiftsnails_b200.parallel``  hash fragments, peer (NVLink) memory, NCCL baseline, launcher
  * ``swiftsnails_b200.models``    word2vec SGNS, sparse logistic regression, embedding bench
  * ``swiftsnails_b200.utils``     config parser, timers, text dump reader/writer
"""
__version__ = "0.1.0"

from . import _lib  # noqa: F401
