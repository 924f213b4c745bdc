"""What the host-buffer pipeline's copy pattern costs on PCIe, without any kernel: 512 MiB up + 320 MiB down (one 16 Mi-ray batch),
(a) as two big copies at once, (b) as 16 x 32 MiB uploads and 16 x 5 x 4 MiB downloads on two streams, the download of chunk k
queued behind the upload of chunk k (event), (c) like (b) with an event record + wait between every pair of copies as
igb200_trace_rays does, (d) like (b) with an HBM-bound kernel (0.2 ms) between upload and download of a chunk, (e) with a kernel that only sleeps.  One GPU, pinned memory."""
import time

import torch

up, down, chunks = 512 << 20, 320 << 20, 16
h_in, h_out = torch.empty(up, dtype=torch.uint8).pin_memory(), torch.empty(down, dtype=torch.uint8).pin_memory()
d_in, d_out = torch.empty(up, dtype=torch.uint8, device="cuda"), torch.empty(down, dtype=torch.uint8, device="cuda")
scratch = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
s_up, s_k, s_down = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.Stream()
cu, cd = up // chunks, down // chunks // 5


def big():
    with torch.cuda.stream(s_up):
        d_in.copy_(h_in, non_blocking=True)
    with torch.cuda.stream(s_down):
        h_out.copy_(d_out, non_blocking=True)


def chunked(kernel=False, timing=False, sleep=False, align=False, one=False):
    pend = None
    for c in range(chunks):
        with torch.cuda.stream(s_up):
            d_in[c * cu:(c + 1) * cu].copy_(h_in[c * cu:(c + 1) * cu], non_blocking=True)
            e = torch.cuda.Event(enable_timing=timing)
            e.record(s_up)
        last = e
        if kernel or sleep:
            with torch.cuda.stream(s_k):
                s_k.wait_event(e)
                if sleep:
                    torch.cuda._sleep(500_000)  # ~0.26 ms at 1.9 GHz, one thread, no memory traffic
                else:
                    scratch.add_(1)  # 256 MiB read + write: about 0.1 ms of HBM traffic per call
                    scratch.add_(1)
                last = torch.cuda.Event(enable_timing=timing)
                last.record(s_k)
        def download(cc, ev):
            with torch.cuda.stream(s_down):
                for x in ev:
                    s_down.wait_event(x)
                if one:
                    o = 5 * cc * cd
                    h_out[o:o + 5 * cd].copy_(d_out[o:o + 5 * cd], non_blocking=True)
                else:
                    for k in range(5):
                        o = (5 * cc + k) * cd
                        h_out[o:o + cd].copy_(d_out[o:o + cd], non_blocking=True)

        if align:  # the download of chunk c-1 starts when the upload of chunk c ends (and its kernel is done)
            if pend is not None:
                download(c - 1, [pend, e])
            pend = last
            if c == chunks - 1:
                download(c, [last])
        else:
            download(c, [last])


def timed(fn, reps=10):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
        torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3


for rep in range(2):
    print(f"(f) (e) + download(c-1) starts at the end of upload(c): {timed(lambda: chunked(sleep=True, align=True)):7.3f} ms")
    print(f"(g) (e) with one 20 MiB download per chunk             : {timed(lambda: chunked(sleep=True, one=True)):7.3f} ms")
    print(f"(h) (b) with one 20 MiB download per chunk             : {timed(lambda: chunked(one=True)):7.3f} ms")
    print(f"(a) two big copies at once            : {timed(big):7.3f} ms")
    print(f"(b) 16 chunks, download behind upload  : {timed(chunked):7.3f} ms")
    print(f"(c) same with timing events            : {timed(lambda: chunked(timing=True)):7.3f} ms")
    print(f"(d) same with an HBM-bound kernel/chunk: {timed(lambda: chunked(kernel=True)):7.3f} ms")
    print(f"(e) same with a 0.26 ms sleep kernel   : {timed(lambda: chunked(sleep=True)):7.3f} ms")
