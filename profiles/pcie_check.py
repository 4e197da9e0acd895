"""Device-to-host bandwidth into pinned memory (what bounds the e2e number) and the PCIe link state."""
import subprocess, time, torch
n_el = 308 * 1024 * 1024 // 4
x = torch.empty(n_el, device="cuda")
h = torch.empty(n_el, dtype=x.dtype, pin_memory=True)
for n in (1, 2, 4):
    step = n_el // n
    parts = [(x[i * step:(i + 1) * step], h[i * step:(i + 1) * step]) for i in range(n)]
    streams = [torch.cuda.Stream() for _ in range(n)]
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(5):
        t = time.perf_counter()
        for (src, dst), s in zip(parts, streams):
            with torch.cuda.stream(s):
                dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t)
    print("D2H 308 MiB in %d concurrent copies: %.2f ms = %.1f GB/s" % (n, best * 1e3, step * n * 4 / best / 1e9))
print(subprocess.run(["nvidia-smi", "--query-gpu=pcie.link.gen.current,pcie.link.gen.max,pcie.link.width.current", "--format=csv"],
                     capture_output=True, text=True).stdout)
