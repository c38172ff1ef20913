#!/usr/bin/env python
"""bench.py — the planning hot path on N B200s of one node.

A "step" is one plan: noise -> fused generator + T-step rollout + terminal cost (K1) ->
segmented argmin/top-k (K3) -> best action (-> one min-allreduce when sharded).
Metric: MPC candidate trajectory-steps/s = G*K*T / t.  Workload at N=1: BASELINE config 2
(G=1, K=65,536, T=8, E=128/C=256, H=1024, nd=2).  N>1: weak scaling, every rank scores
K=65,536 candidates of the same goal (K_total = N*65,536), one allreduce-min per plan.
The ndiv loss (BASELINE config 3: k=4096, fwd+bwd) is timed in the same run and reported
under "ndiv".

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

F_ALG = 666_112          # algorithmic FLOP per candidate-step (goal half of fc1 folded)
CFG = dict(G=1, K=65536, T=8, E=128, H=1024, nd=2, topk=16)
NDIV = dict(N=1, k=4096, Dx=32, Dz=16)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], bf16=d["bf16_tflops"], bf16_sus=d["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sus=1400.0, src="fallback")


class ClockSampler(threading.Thread):
    """samples SM clock + throttle reasons through NVML while the timed region runs"""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.samples, self.reasons, self.max_mhz, self._stop_evt, self.ok = [], set(), None, threading.Event(), False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            pass

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
                 0x80: "hw_power_brake", 0x2: "applications_clocks_setting", 0x100: "display_clock_setting",
                 0x10: "sync_boost"}
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, n in names.items():
                    if r & bit:
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.01)

    def finish(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join()
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def l2_flusher(dev):
    buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # 256 MB > 126 MB L2
    return lambda: buf.fill_(1)


def ev():
    return torch.cuda.Event(enable_timing=True)


# --------------------------------------------------------------------------------------------
# CPU baseline / reference arm: the oracle port (torch-CPU restatement of the reference loop)
# --------------------------------------------------------------------------------------------
def cpu_port_plan_rate(K_sample, reps, threads):
    from ndivplanning_b200 import synthetic
    from oracle import ndiv_oracle as O
    torch.set_num_threads(threads)
    c = CFG
    gen_p, fwd_p = synthetic.make_models(nd=c["nd"], E=c["E"], H=c["H"])
    gen, fwd = O.GeneratorMLP(c["nd"]).eval(), O.LatentForward(c["E"], c["H"]).eval()
    gen.load_state_dict(gen_p.state_dict()); fwd.load_state_dict(fwd_p.state_dict())
    e0, goal, noise = synthetic.make_plan_inputs(1, K_sample, c["T"], c["nd"], c["E"], seed=1)
    times = []
    for r in range(reps + 1):
        t0 = time.perf_counter()
        O.mpc_plan(gen, fwd, e0[0], goal[0], noise[:, 0], topk=c["topk"])   # incl. the python argmin scan
        times.append(time.perf_counter() - t0)
    return times[1:], K_sample * c["T"]


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    Ks = 8192
    # warmup W (bounded) + K timed steps; each step = one plan over a Ks-candidate sample
    cpu_port_plan_rate(Ks, max(0, min(args.warmup, 2) - 1), threads)
    times, units = cpu_port_plan_rate(Ks, args.steps, threads)
    t = sum(times) / len(times)
    v = units / t
    out = {"impl": "reference", "metric": "mpc_candidate_steps_per_sec", "value": v, "unit": "candidate-steps/s",
           "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3,
           "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": dict(workload="mpc_cfg2_K65536_T8_E128_H1024_nd2", **CFG),
           "cpu_baseline": {"value": v, "unit": "candidate-steps/s", "cores": threads, "kind": "port",
                            "sample": f"{Ks} of {CFG['K']} candidates per step x T={CFG['T']}, oracle port of "
                                      f"mpc_eval.py:124-165 incl. python argmin, torch {torch.__version__} CPU fp32"},
           "e2e": {"value": v, "unit": "candidate-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out))
