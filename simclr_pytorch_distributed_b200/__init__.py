"""B200-native (sm_100a) implementation of the two data-parallel hot paths of
Dyfine/SimCLR_pytorch_distributed, behind the reference's own Python API:

  SupConLoss      /root/reference/losses.py:7-93
  SyncBatchNorm   the torch.nn.SyncBatchNorm layers of --syncBN (/root/reference/main_supcon.py:222-224)
  ProjectionTail  F.normalize(head[2](h)) + view stacking (/root/reference/networks/resnet_big.py:168-181)

Host code is Python/PyTorch (device memory, streams, torch.distributed); the arithmetic is in
hand-written CUDA behind the C ABI of include/simclr_b200.h (csrc/libsimclr_b200.so).
"""
from .losses import SupConLoss  # noqa: F401
from .sync_batchnorm import SyncBatchNorm  # noqa: F401
from .head import ProjectionTail  # noqa: F401

__all__ = ['SupConLoss', 'SyncBatchNorm', 'ProjectionTail']
