#!/usr/bin/env python3
"""Host->device link probe: pinned 64 MB copies on 1, 2 and 4 streams, page-locked (default) and write-combined host memory."""
import ctypes as C
import torch

rt = C.CDLL("libcudart.so.12")
N = 64 << 20
dev = [torch.empty(N, dtype=torch.uint8, device="cuda") for _ in range(4)]


def alloc(flags):
    p = C.c_void_p()
    assert rt.cudaHostAlloc(C.byref(p), C.c_size_t(N), C.c_uint(flags)) == 0
    C.memset(p, 1, N)
    return p


def run(hosts, nstreams, d2h=False):
    streams = [torch.cuda.Stream() for _ in range(nstreams)]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for s in streams:
        s.wait_event(e0)
    for rep in range(4):
        for i, s in enumerate(streams):
            a, b = (hosts[i], C.c_void_p(dev[i].data_ptr())) if d2h else (C.c_void_p(dev[i].data_ptr()), hosts[i])
            assert rt.cudaMemcpyAsync(a, b, C.c_size_t(N), C.c_int(4), C.c_void_p(s.cuda_stream)) == 0
    for s in streams:
        torch.cuda.current_stream().wait_stream(s)
    e1.record()
    torch.cuda.synchronize()
    return 4 * nstreams * N / (e0.elapsed_time(e1) / 1e3) / 1e9


for name, flags in (("pinned", 0), ("portable", 1), ("write-combined", 4)):
    hosts = [alloc(flags) for _ in range(4)]
    for ns in (1, 2, 4):
        run(hosts, ns)
        print("%-15s h2d %d stream(s): %6.2f GB/s   d2h: %6.2f GB/s" % (name, ns, run(hosts, ns), run(hosts, ns, True)), flush=True)
    for h in hosts:
        rt.cudaFreeHost(h)
