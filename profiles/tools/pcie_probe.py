"""Why does the chunked bidirectional pipeline see ~38 GB/s per direction when two whole-vector
copies in opposite directions see ~50?  Pure-torch copy experiments, no SpMV."""
import sys
import time

import torch

N = 256 ** 3
torch.cuda.set_device(0)
dev = torch.device("cuda:0")
xh = torch.randn(N, dtype=torch.float64).pin_memory()
yh = torch.empty(N, dtype=torch.float64).pin_memory()
xd = torch.empty(N, dtype=torch.float64, device=dev)
yd = torch.randn(N, dtype=torch.float64, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def wall(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(n):
        fn()
        torch.cuda.synchronize()
    return (time.perf_counter() - t) / n * 1e3


def chunked(C, shift, deps, sync_each=True):
    """H2D chunk k on s1, D2H chunk k-shift on s2 (deps: D2H k waits for H2D k via an event)."""
    b = [N * k // C for k in range(C + 1)]

    def fn():
        evs = []
        for k in range(C + shift):
            if k < C:
                with torch.cuda.stream(s1):
                    xd[b[k]:b[k + 1]].copy_(xh[b[k]:b[k + 1]], non_blocking=True)
                    if deps:
                        e = torch.cuda.Event()
                        e.record(s1)
                        evs.append(e)
            j = k - shift
            if j >= 0:
                with torch.cuda.stream(s2):
                    if deps:
                        s2.wait_event(evs[j])
                    yh[b[j]:b[j + 1]].copy_(yd[b[j]:b[j + 1]], non_blocking=True)
    return fn


mb = N * 8 / 1e6
for C in (1, 8, 32):
    for shift in (0, 1):
        for deps in (False, True):
            ms = wall(chunked(C, shift, deps))
            ideal = 2.707 * (1 + shift / C)
            print(f"chunks {C:3d} shift {shift} deps {int(deps)}: {ms:7.3f} ms   ({2 * mb / ms:6.1f} GB/s both ways; "
                  f"whole-vector concurrency would give {ideal:.2f} ms)")


def hopped(C, stream3, timing):
    """mult_host's event chain without the kernels: H2D k on s1 -> event -> stream3 waits, records ->
    s2 waits -> D2H k."""
    b = [N * k // C for k in range(C + 1)]

    def fn():
        for k in range(C):
            with torch.cuda.stream(s1):
                xd[b[k]:b[k + 1]].copy_(xh[b[k]:b[k + 1]], non_blocking=True)
                e = torch.cuda.Event(enable_timing=timing)
                e.record(s1)
            stream3.wait_event(e)
            e2 = torch.cuda.Event(enable_timing=timing)
            e2.record(stream3)
            with torch.cuda.stream(s2):
                s2.wait_event(e2)
                yh[b[k]:b[k + 1]].copy_(yd[b[k]:b[k + 1]], non_blocking=True)
    return fn


s3 = torch.cuda.Stream()
for name, st in (("side stream", s3), ("default stream", torch.cuda.default_stream())):
    for timing in (False, True):
        print(f"chunks 8, hop through {name}, timing events {int(timing)}: {wall(hopped(8, st, timing)):7.3f} ms")
# separate device buffers per direction vs one buffer both ways
xd2 = torch.empty(N, dtype=torch.float64, device=dev)
