"""HBM write / copy bandwidth of torch's own fill_ and copy_ on a 734 MB tensor (the stem output size): the ceiling for a
kernel that only writes."""
import torch
n = 64 * 224 * 400 * 64
x = torch.empty(n, dtype=torch.bfloat16, device="cuda")
y = torch.empty(n, dtype=torch.bfloat16, device="cuda")


def t(fn, reps=20):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps


us = t(lambda: x.fill_(1.0))
print("fill_ %d MB: %.1f us -> %.2f TB/s written" % (n * 2 // 10**6, us, n * 2 / us / 1e6))
us = t(lambda: y.copy_(x))
print("copy_ %d MB: %.1f us -> %.2f TB/s read+written" % (n * 2 // 10**6, us, 2 * n * 2 / us / 1e6))
us = t(lambda: torch.cuda.memset if False else x.zero_())
print("zero_ %d MB: %.1f us -> %.2f TB/s written" % (n * 2 // 10**6, us, n * 2 / us / 1e6))
