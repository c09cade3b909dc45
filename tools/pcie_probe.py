"""D2H bandwidth of every visible GPU into pinned host memory, one at a time and all together (GPU box)."""
import threading, time, torch
n = torch.cuda.device_count()
N = 1 << 30
src = [torch.empty(N, dtype=torch.uint8, device=f"cuda:{i}") for i in range(n)]
dst = [torch.empty(N, dtype=torch.uint8).pin_memory() for i in range(n)]
def run(i, reps=4):
    torch.cuda.set_device(i)
    s = torch.cuda.Stream(i)
    with torch.cuda.stream(s):
        dst[i].copy_(src[i], non_blocking=True); s.synchronize()
        t = time.perf_counter()
        for _ in range(reps): dst[i].copy_(src[i], non_blocking=True)
        s.synchronize()
    return reps * N / (time.perf_counter() - t) / 1e9
for i in range(n): print(f"gpu{i} alone: {run(i):.1f} GB/s", flush=True)
res = [0] * n
th = [threading.Thread(target=lambda i=i: res.__setitem__(i, run(i, 8))) for i in range(n)]
[t.start() for t in th]; [t.join() for t in th]
print("together:", [round(x, 1) for x in res])
import subprocess
print(subprocess.run("nvidia-smi --query-gpu=index,pcie.link.gen.current,pcie.link.width.current,pcie.link.gen.max,pcie.link.width.max --format=csv", shell=True, capture_output=True, text=True).stdout)
