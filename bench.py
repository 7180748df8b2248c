#!/usr/bin/env python
"""bench.py -- ID-estimates/s of the manifold_dimension hot path on BASELINE.json configs[1]
(50-sphere in R^200, models/fcn.py score net H=2048 L=5 random-init, K=1501 perturbations per anchor,
512 anchors per GPU = 4096 anchors over 8 x B200; weak scaling).

    python bench.py --gpus N --steps K --warmup W            # our arm (torchrun launches it for N > 1)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU algorithm on the host cores

A step = one pass of the whole path (perturb -> score net -> centre/Gram -> spectrum -> ID) over this rank's
anchors.  One JSON line is printed by rank 0.  See DESIGN.md "Measurement" for what each key means.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

D, H, L, B_LOADER, MANIFOLD = 200, 2048, 5, 500, 50
K = 3 * B_LOADER + 1  # reference K schedule for Euclidean anchors (dim_reduction.py:166-171): 1501
SIGMA_MIN, SIGMA_MAX, N_SCALES, T_EPS = 1e-2, 4.0, 1000, 1e-5
FLOPS_PER_EVAL = 2.0 * ((D + 1) * H + L * H * H + H * D)  # 43.6 MFLOP


def make_config():
    from iddiff_b200.config import ConfigDict
    return ConfigDict(model=dict(name="fcn", state_size=D, hidden_layers=L, hidden_nodes=H, dropout=0.0,
                                 sigma_min=SIGMA_MIN, sigma_max=SIGMA_MAX, num_scales=N_SCALES,
                                 checkpoint_path=None),
                      training=dict(sde="vesde", batch_size=B_LOADER), data=dict(use_data_mean=False))


def make_model(seed=0):
    """Random-init FCN (nn.Linear default init, torch.manual_seed(seed)): no checkpoint exists offline."""
    from iddiff_b200 import models
    torch.manual_seed(seed)
    return models.create_model(make_config())


def make_anchors(n, seed=1):
    """n points of the 50-sphere in R^200 (KSphereDataset generator, random isometry embedding)."""
    from iddiff_b200 import data
    torch.manual_seed(seed)
    return data.ksphere_points(n, 1, D, MANIFOLD, 0.0, "random_isometry").float().contiguous()


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return p, "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        self.th.join(timeout=2)
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for j, n in enumerate(names) if any(len(r) > 2 + j and r[2 + j] == "Active" for r in self.rows)]
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}
