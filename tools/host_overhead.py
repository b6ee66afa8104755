"""Where does the host time of one eager NT_Xent step go?  (cProfile, small problem so the GPU is idle.)"""
import cProfile
import os
import pstats
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from simclr_b200 import NT_Xent  # noqa: E402

dev = torch.device("cuda:0")
n, d = 256, 128
a = torch.randn(n, d, device=dev, requires_grad=True)
b = torch.randn(n, d, device=dev, requires_grad=True)
crit = NT_Xent(0.5)


def step():
    a.grad = b.grad = None
    crit(a, b).backward()


for _ in range(20):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(500):
    step()
torch.cuda.synchronize()
print(f"eager step: {(time.perf_counter() - t0) / 500 * 1e6:.1f} us wall")
pr = cProfile.Profile()
pr.enable()
for _ in range(500):
    step()
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("cumulative").print_stats(28)
