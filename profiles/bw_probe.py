"""Read-only / write-only / copy DRAM bandwidth probe (torch library kernels) to
put the per-kernel roofline fractions in context: a write-dominated kernel
cannot reach the copy figure of MEASURED_PEAKS.json."""
import torch
n = 1 << 30   # 4 GiB fp32
x = torch.empty(n, dtype=torch.float32, device="cuda")
y = torch.empty(n, dtype=torch.float32, device="cuda")
def t(f, reps=10):
    f(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(reps): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e-3
tw = t(lambda: x.zero_())
tr = t(lambda: x.sum())
tc = t(lambda: y.copy_(x))
ta = t(lambda: torch.add(x, 1.0, out=y))
print(f"write-only {4*n/tw/1e9:.0f} GB/s  read-only {4*n/tr/1e9:.0f} GB/s  copy {8*n/tc/1e9:.0f} GB/s  add(out) {8*n/ta/1e9:.0f} GB/s")
