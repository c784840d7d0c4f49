"""Fill (almost) all free GPU memory with a byte pattern and release it: the next process then sees
that pattern in memory it never wrote (debugging aid for reads of uninitialised device memory)."""
import sys, torch
mode = sys.argv[1] if len(sys.argv) > 1 else "random"
free, total = torch.cuda.mem_get_info()
n = int(free * 0.9)
x = torch.empty(n, dtype=torch.uint8, device="cuda")
if mode == "random":
    step = 1 << 30
    for o in range(0, n, step):
        x[o:o + step].random_(0, 256)
else:
    x.fill_(int(mode))
torch.cuda.synchronize()
print("scribbled", n, "bytes with", mode)
