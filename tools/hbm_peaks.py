"""Write-only / read-only / copy HBM bandwidth of this GPU (the train-step kernels are write- or read-bound on their
records, MEASURED_PEAKS.json only has the copy figure)."""
import torch
dev = torch.device("cuda")
n = 1 << 30   # 4 GiB of fp32
a = torch.empty(n, dtype=torch.float32, device=dev)
b = torch.empty(n, dtype=torch.float32, device=dev)


def t(fn, reps=5):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e-3)
    return best

print("write-only (fill_)   %.1f GB/s" % (4 * n / t(lambda: a.fill_(1.0)) / 1e9))
print("write-only (memset)  %.1f GB/s" % (4 * n / t(lambda: a.zero_()) / 1e9))
print("read-only  (sum)     %.1f GB/s" % (4 * n / t(lambda: a.sum()) / 1e9))
print("copy (read + write)  %.1f GB/s" % (8 * n / t(lambda: b.copy_(a)) / 1e9))
