"""Projection head's last layer + normalise: the fused kernel against torch's Linear followed by ntx_l2norm.

    python tools/project_times.py [n] [hdim] [d]
"""
import os
import statistics
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from simclr_b200 import _cabi  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
hdim = int(sys.argv[2]) if len(sys.argv) > 2 else 512
d = int(sys.argv[3]) if len(sys.argv) > 3 else 128
dev = torch.device("cuda:0")
be = _cabi.Backend()
g = torch.Generator().manual_seed(7)
h0 = torch.randn(n, hdim, generator=g).bfloat16().to(dev)
h1 = torch.randn(n, hdim, generator=g).bfloat16().to(dev)
w = (torch.randn(d, hdim, generator=g) / hdim ** 0.5).bfloat16().to(dev)
ld = be.row_pitch(d, torch.bfloat16)
rows = torch.empty(2 * n, ld, dtype=torch.bfloat16, device=dev)
inv = torch.empty(2 * n, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def fused():
    be.project_l2norm(h0, h1, w, rows, inv)


def unfused():
    z0 = torch.nn.functional.linear(h0, w)
    z1 = torch.nn.functional.linear(h1, w)
    be.l2norm(z0, z1, rows, inv)


for name, fn in (("fused ntx_project_l2norm", fused), ("2 x F.linear + ntx_l2norm", unfused)):
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        for _ in range(3):
            fn()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side, capture_error_mode="thread_local"):
            fn()
        ts = []
        for it in range(25):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            graph.replay()                    # host work (tensor maps, launches) stays out of the timed region
            e1.record()
            torch.cuda.synchronize()
            if it >= 5:
                ts.append(e0.elapsed_time(e1) * 1e3)
    bytes_min = 2 * n * hdim * 2 + d * hdim * 2 + 2 * n * ld * 2
    print(f"{name:28s} {statistics.median(ts):7.1f} us   (compulsory traffic {bytes_min / 1e6:.1f} MB = "
          f"{bytes_min / statistics.median(ts) / 1e3:.0f} GB/s), CUDA graph replay, L2 flushed")
