#!/usr/bin/env python
"""Where the PyTorch baseline's step goes: torch.profiler over a few eager steps (1 GPU), top CUDA kernels by
total time.  `python bench/prof_baseline.py > gpurun_out/baseline_prof.txt`"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from baseline.wide_deep_torch import TorchWideDeep, synthetic_pool  # noqa: E402


def main():
    args = argparse.Namespace(emb_rows=10_000_000, batch=8192, pool=4)
    dev = torch.device("cuda", 0)
    m = TorchWideDeep(args.emb_rows, args.batch, 0, 1, dev)
    pool = [tuple(t.to(dev) for t in b) for b in synthetic_pool(args, 0, pin=False)]
    for i in range(3):
        m.step(*pool[i % 4])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(10):
        m.step(*pool[i % 4])
    e1.record()
    torch.cuda.synchronize()
    print("eager: %.3f ms/step" % (e0.elapsed_time(e1) / 10))
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA, torch.profiler.ProfilerActivity.CPU]) as prof:
        for i in range(5):
            m.step(*pool[i % 4])
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=35, max_name_column_width=70))


if __name__ == "__main__":
    main()
