"""Is an in-place elementwise pass slower than an out-of-place one on this GPU?  (rotate_kernel is in place.)"""
import torch
n = 2891 * 1021 * 65
a = torch.randn(n, device="cuda"); b = torch.empty_like(a); c = torch.randn(n, device="cuda"); d = torch.empty_like(a)
def t(fn, reps=10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for _ in range(reps):
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1))
    return best
gb = n * 4 * 2 / 1e6
print("out of place  a*2 -> b      : %.4f ms  %.0f GB/s" % ((ms := t(lambda: torch.mul(a, 2.0, out=b))), gb / ms))
print("in place      a*2 -> a      : %.4f ms  %.0f GB/s" % ((ms := t(lambda: a.mul_(1.0001))), gb / ms))
print("two arrays out of place     : %.4f ms  %.0f GB/s" % ((ms := t(lambda: (torch.mul(a, 2.0, out=b), torch.mul(c, 2.0, out=d)))), 2 * gb / ms))
print("two arrays in place         : %.4f ms  %.0f GB/s" % ((ms := t(lambda: (a.mul_(1.0001), c.mul_(1.0001)))), 2 * gb / ms))
print("copy a -> b                 : %.4f ms  %.0f GB/s" % ((ms := t(lambda: b.copy_(a))), gb / ms))
