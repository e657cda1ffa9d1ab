"""Developer tooling: pinned H2D / D2H bandwidth of this box, alone and both directions at once."""
import time, torch
n = 256 << 20
h1 = torch.empty(n, dtype=torch.uint8).pin_memory(); h2 = torch.empty(n, dtype=torch.uint8).pin_memory()
d1 = torch.empty(n, dtype=torch.uint8, device="cuda"); d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(h2d, d2h, reps=5):
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(reps):
        if h2d:
            with torch.cuda.stream(s1): d1.copy_(h1, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t) / reps
    return n / dt / 1e9
run(True, True)
print("H2D alone %.1f GB/s" % run(True, False)); print("D2H alone %.1f GB/s" % run(False, True))
print("both at once: %.1f GB/s per direction" % run(True, True))
# many 4 MiB copies instead of one big one
def chunks(h2d, d2h, reps=3, sz=4 << 20):
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(reps):
        for o in range(0, n, sz):
            if h2d:
                with torch.cuda.stream(s1): d1[o:o+sz].copy_(h1[o:o+sz], non_blocking=True)
            if d2h:
                with torch.cuda.stream(s2): h2[o:o+sz].copy_(d2[o:o+sz], non_blocking=True)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t) / reps
    return n / dt / 1e9
print("4 MiB pieces, both: %.1f GB/s per direction" % chunks(True, True))
print("1.6 MiB pieces D2H: %.1f GB/s" % chunks(False, True, sz=1600000))
# the e2e step's pattern: 4 MiB pieces in both directions, spread over several streams each way
def multi(nstreams, reps=3, sz=4 << 20, sz2=None):
    ss = [torch.cuda.Stream() for _ in range(2 * nstreams)]
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(reps):
        for k, o in enumerate(range(0, n, sz)):
            with torch.cuda.stream(ss[k % nstreams]): d1[o:o+sz].copy_(h1[o:o+sz], non_blocking=True)
            with torch.cuda.stream(ss[nstreams + k % nstreams]): h2[o:o+sz].copy_(d2[o:o+sz], non_blocking=True)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t) / reps
    return n / dt / 1e9
for ns in (1, 2, 6, 12):
    print("4 MiB pieces, both directions, %d streams each way: %.1f GB/s per direction" % (ns, multi(ns)))
print("32 MiB pieces, both directions, 6 streams each way: %.1f GB/s per direction" % multi(6, sz=32 << 20))
