"""channel_gpu_b200 — B200-native KMM channel-flow RK3 time-stepper.

Python is only the launcher: every computation happens in
lib/libchannel_b200.so (hand-written CUDA for sm_100a behind the C ABI of
include/channel_b200.h).  Importing this package without that library raises:
there is no CPU or PyTorch fallback.
"""
from .api import Channel, ChannelError, lib_path, load_library, partition, exchange_layout, peer_put_layout, FIELDS  # noqa: F401
