#!/usr/bin/env python3
"""Probe the peer path between GPU 0 and GPU 1: topology, copy bandwidth and small-copy latency."""
import subprocess, time, torch
print(subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True).stdout[:1500])
print("can access peer:", torch.cuda.can_device_access_peer(0, 1))
for mb in (1, 16, 256):
    a = torch.empty(mb << 20, dtype=torch.uint8, device="cuda:0"); b = torch.empty(mb << 20, dtype=torch.uint8, device="cuda:1")
    for _ in range(3): b.copy_(a)
    torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.device(0):
        s.record()
        for _ in range(20): b.copy_(a)
        e.record(); torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    ms = s.elapsed_time(e) / 20
    print(f"{mb:4d} MiB copy 0->1: {ms*1e3:8.1f} us  {mb/1024/(ms/1e3):7.1f} GiB/s")
