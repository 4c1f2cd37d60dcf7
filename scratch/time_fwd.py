import torch, sys
sys.path.insert(0, '.')
from pytorch_mnist_ddp_b200.ops import native_ops as ops
from pytorch_mnist_ddp_b200.models.net import Net, ParamLayout
torch.manual_seed(1)
net = Net().cuda()
w = {k: ParamLayout.to_internal(k, v.cpu()).cuda().contiguous() for k, v in net.state_dict().items()}
a1 = torch.rand(200, 26, 26, 32, device='cuda')
ctrl = ops.make_ctrl(train=True, step=3)
for tc in (2, 3):
    for _ in range(5): ops.conv2_fwd(a1, w['conv2.weight'], w['conv2.bias'], ctrl, tc=tc)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50): ops.conv2_fwd(a1, w['conv2.weight'], w['conv2.bias'], ctrl, tc=tc)
    e1.record(); torch.cuda.synchronize()
    print('fwd variant', tc, 'us/call', e0.elapsed_time(e1) / 50 * 1000)
