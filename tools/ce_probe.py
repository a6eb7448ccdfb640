"""Does a scalar read-back wait behind a bulk device-to-host copy on another stream?"""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kmls_b200
from kmls_b200 import _lib
dev = torch.device("cuda", 0)
big = torch.zeros(256 << 20, dtype=torch.int32, device=dev)      # 1 GiB
hbig = torch.empty(256 << 20, dtype=torch.int32).pin_memory()
small = torch.arange(4, dtype=torch.int64, device=dev)
hsmall = torch.zeros(4, dtype=torch.int64).pin_memory()
side = torch.cuda.Stream(device=dev)
torch.cuda.synchronize()
def bulk(chunks):
    with torch.cuda.stream(side):
        n = big.numel() // chunks
        for c in range(chunks):
            hbig[c * n:(c + 1) * n].copy_(big[c * n:(c + 1) * n], non_blocking=True)
for chunks in (1, 256):
    for mode in ("item", "kernel"):
        torch.cuda.synchronize()
        t0 = time.perf_counter(); bulk(chunks); t1 = time.perf_counter()
        if mode == "item":
            v = int(small[3].item())
        else:
            _lib.call("kmls_words_to_host", small, 8, hsmall, torch.cuda.current_stream().cuda_stream)
            torch.cuda.current_stream().synchronize()
            v = int(hsmall[3])
        t2 = time.perf_counter()
        torch.cuda.synchronize(); t3 = time.perf_counter()
        print(f"chunks={chunks:4d} {mode:6s}: enqueue {1e3*(t1-t0):6.2f} ms, scalar read {1e3*(t2-t1):6.2f} ms, bulk done after {1e3*(t3-t0):6.2f} ms, v={v}")
