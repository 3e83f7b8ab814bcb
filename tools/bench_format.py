"""Throughput of the cube-text formatter on device-resident arrays: values/s and HBM GB/s
(algorithmic bytes = 8 B read + field bytes written per value) against the measured copy bandwidth.  One JSON line per
(fmt, order).  Never touches oracle/: the CPU baseline (C oracle, glibc "%.*E" + OpenMP on all host cores, bounded
sample) is supplied by bench.py (`python bench.py --aux format`)."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from critic2_b200 import gpu  # noqa: E402


def main(argv=None, cpu=None):
    """cpu: optional object with text(sample, fmt) -> cpu_baseline dict (bench.py supplies it)."""
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=400)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--cpu-n", type=int, default=128)
    a = ap.parse_args(argv)

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    HBM = float(peaks.get("hbm_gbs", 6540.0)) if isinstance(peaks, dict) else 6540.0   # MEASURED_PEAKS.json, else the recipe fallback

    gpu.init_devices([0])
    dev = torch.device("cuda", 0)
    n = [a.n] * 3
    rng = np.random.default_rng(0)
    host = rng.normal(0, 1, a.n ** 3) * 10.0 ** rng.integers(-8, 3, a.n ** 3)
    g = torch.from_numpy(host).to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream().cuda_stream

    sample = host[: a.cpu_n ** 3].reshape(a.cpu_n, a.cpu_n, a.cpu_n)
    cpu_res = {fmt: (cpu.text(sample, fmt) if cpu else None) for fmt in (0, 1, 2)}

    for fmt in (0, 1, 2):
        nb = gpu.cube_text_size(n, fmt)
        text = torch.empty(nb, dtype=torch.uint8, device=dev)
        for order in (0, 1):
            ms = []
            for i in range(a.warmup + a.steps):
                flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                gpu.cube_text_dev(g, n, text, fmt, order, stream=stream)
                e1.record()
                torch.cuda.synchronize()
                if i >= a.warmup:
                    ms.append(e0.elapsed_time(e1))
            t = float(np.mean(ms))
            nv = a.n ** 3
            byts = 8.0 * nv + nb
            print(json.dumps({"kernel": "format_records_kernel", "fmt": fmt, "order": order, "grid": f"{a.n}^3", "values": nv,
                              "text_bytes": nb, "ms": t, "values_per_s": nv / t * 1e3,
                              "roofline": {"bound": "hbm", "achieved": byts / t / 1e6, "peak": HBM, "unit": "GB/s",
                                           "frac": byts / t / 1e6 / HBM, "algorithmic_bytes_per_value": byts / nv},
                              "cpu_baseline": cpu_res[fmt]}), flush=True)
        del text


if __name__ == "__main__":
    main()
