"""wsinsight_b200: B200-native (sm_100a) implementation of WSInsight's patch-classification hot path.

Host code is Python/PyTorch (device memory, streams, torch.distributed); all arithmetic of the
path runs in ``libwsi_b200.so`` (C ABI in ``include/wsi_b200.h``).  There is no CPU fallback.
"""

__version__ = "0.1.0"
