# Does adadelta (SIMT, no smem) co-reside with the persistent conv kernels when launched while they run?
import sys, torch
sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
from pytorch_mnist_ddp_b200 import _native as N
from pytorch_mnist_ddp_b200.ops import native_ops as ops
L = N.lib()
B = 3000
dev = "cuda"
dy = torch.randn(B, 24, 24, 64, device=dev)
w = torch.randn(9 * 64 * 32, device=dev)
a1 = torch.randn(B, 26, 26, 32, device=dev)
da1 = torch.empty_like(a1)
part = torch.empty(148 * (9 * 64 * 32 + 64), device=dev)
n = 1181184
p, g, v, u = (torch.randn(n, device=dev).abs() for _ in range(4))
ctrl = ops.make_ctrl()
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
_ptr = lambda t: t.data_ptr()

def big(kind):
    if kind == "dgrad":
        N.check(L.b200_conv2_dgrad_tc3(_ptr(dy), _ptr(w), _ptr(a1), _ptr(da1), B, s1.cuda_stream))
    elif kind == "wgrad":
        N.check(L.b200_conv2_wgrad_tc2(_ptr(dy), _ptr(a1), _ptr(part), None, B, s1.cuda_stream))

def small():
    N.check(L.b200_adadelta(_ptr(p), _ptr(g), _ptr(v), _ptr(u), n, _ptr(ctrl), 0.9, 1e-6, -1, None, s2.cuda_stream))

for kind in ("none", "dgrad", "wgrad"):
    for rep in range(4):
        torch.cuda.synchronize()
        eb0, eb1, e0, e1 = (torch.cuda.Event(enable_timing=True) for _ in range(4))
        with torch.cuda.stream(s1):
            eb0.record(s1)
            big(kind)
            eb1.record(s1)
        with torch.cuda.stream(s2):
            torch.cuda._sleep(100000)  # ~50 us: the big kernel is running by now
            e0.record(s2)
            small()
            e1.record(s2)
        torch.cuda.synchronize()
        if rep:
            print(f"{kind:6s} big kernel {eb0.elapsed_time(eb1)*1e3:8.1f} us | adadelta {e0.elapsed_time(e1)*1e3:8.1f} us, "
                  f"started {eb0.elapsed_time(e0)*1e3:7.1f} us after the big kernel's start")
