"""PCIe H2D probe: pinned copy bandwidth with 1..4 concurrent streams (is one copy engine the limit?)."""
import torch, time
dev = torch.device("cuda:0")
for mb in (10, 20, 256):
    n = mb << 20
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    d = torch.empty(n, dtype=torch.uint8, device=dev)
    for ns in (1, 2, 4):
        streams = [torch.cuda.Stream() for _ in range(ns)]
        chunk = n // ns
        def go():
            for i, s in enumerate(streams):
                with torch.cuda.stream(s):
                    d[i * chunk:(i + 1) * chunk].copy_(h[i * chunk:(i + 1) * chunk], non_blocking=True)
        for _ in range(3):
            go()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        R = 20
        for _ in range(R):
            go()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / R
        print(f"{mb} MiB, {ns} streams: {dt*1e3:.3f} ms  {n/dt/1e9:.1f} GB/s", flush=True)
import subprocess
print(subprocess.run("nvidia-smi --query-gpu=pcie.link.gen.current,pcie.link.width.current,pcie.link.gen.max --format=csv", shell=True, capture_output=True, text=True).stdout)
