#!/usr/bin/env python3
"""Resident-set bandwidth probe (SURVEY 8d asks for the L2 figure next to the HBM one): device-to-device copies of
buffers that fit the 126 MB L2 against one that does not.  torch is only the launcher of the copy kernel here.
    python scripts/l2_bandwidth.py > profiles/rNN_l2_bandwidth.json"""
import json

import torch


def copy_gbs(nbytes, iters=200):
    a = torch.empty(nbytes // 4, dtype=torch.float32, device="cuda")
    b = torch.empty_like(a)
    a.normal_()
    for _ in range(10):
        b.copy_(a)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        b.copy_(a)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    return 2.0 * nbytes * iters / (ms * 1e-3) / 1e9      # read + write


def main():
    out = {"device": torch.cuda.get_device_name(0), "l2_bytes": torch.cuda.get_device_properties(0).L2_cache_size, "copy_gbs": {}}
    for mb in (4, 8, 16, 32, 48, 512, 2048):
        out["copy_gbs"][f"{mb} MB per buffer"] = round(copy_gbs(mb << 20, iters=200 if mb <= 48 else 20), 1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
