"""Probe: cuDNN time of the ResNet stem conv (7x7/2, 3 input channels) with the input channels zero-padded to 4 / 8."""
import torch
import torch.nn.functional as F
dev = "cuda:0"
torch.backends.cudnn.benchmark = False
x3 = torch.randn(16, 3, 448, 800, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
w3 = torch.randn(64, 3, 7, 7, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
b = torch.randn(64, device=dev, dtype=torch.bfloat16)
def bench(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
ref = torch.cudnn_convolution_relu(x3, w3, b, (2, 2), (3, 3), (1, 1), 1)
for c in (3, 4, 8):
    x = F.pad(x3, (0, 0, 0, 0, 0, c - 3)).contiguous(memory_format=torch.channels_last)
    w = F.pad(w3, (0, 0, 0, 0, 0, c - 3)).contiguous(memory_format=torch.channels_last)
    for name, fn in (("cudnn_convolution_relu", lambda: torch.cudnn_convolution_relu(x, w, b, (2, 2), (3, 3), (1, 1), 1)),
                     ("conv2d+relu_", lambda: F.conv2d(x, w, b, (2, 2), (3, 3)).relu_())):
        try:
            y = fn()
            err = float((y.float() - ref.float()).abs().max() / ref.float().abs().max())
            print("C=%d %-24s %8.1f us  rel err %.3g" % (c, name, bench(fn), err))
        except Exception as e:
            print("C=%d %s failed: %s" % (c, name, str(e)[:100]))
    print("  pad cost %.1f us" % bench(lambda: F.pad(x3, (0, 0, 0, 0, 0, c - 3)).contiguous(memory_format=torch.channels_last)))
torch.backends.cudnn.benchmark = True
for c in (3, 8):
    x = F.pad(x3, (0, 0, 0, 0, 0, c - 3)).contiguous(memory_format=torch.channels_last)
    w = F.pad(w3, (0, 0, 0, 0, 0, c - 3)).contiguous(memory_format=torch.channels_last)
    print("benchmark=True C=%d conv2d+relu_ %8.1f us" % (c, bench(lambda: F.conv2d(x, w, b, (2, 2), (3, 3)).relu_())))
mp = torch.nn.MaxPool2d(3, 2, 1)
print("maxpool %.1f us" % bench(lambda: mp(ref)))
