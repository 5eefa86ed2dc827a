import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from simclr_pytorch_distributed_b200 import SyncBatchNorm
from simclr_pytorch_distributed_b200.sync_batchnorm import _Kernels as K, _layout
for shape, cl, dt in [((256,256,8,8), True, torch.float32), ((256,256,8,8), True, torch.bfloat16), ((32,64,32,32), True, torch.float32), ((16,2048,4,4), True, torch.float32), ((8,24,5,7), True, torch.float32), ((64,128), False, torch.float32), ((16,64,16,16), False, torch.float32)]:
    bn = SyncBatchNorm(shape[1]).cuda()
    x = torch.randn(shape, device='cuda').to(dt)
    if cl: x = x.contiguous(memory_format=torch.channels_last)
    xx, geom = _layout(x)
    bn(x); bn(x)
    torch.cuda.synchronize()
    print(shape, cl, dt, 'nbt', int(bn.num_batches_tracked), 'fused fwd', K.fused_supported(xx, geom, 0, 0))
