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


# --------------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------------
def run_gpu(args):
    import torch.distributed as dist
    from ndivplanning_b200 import FusedPlanner, diversity, launch_count, synthetic
    from ndivplanning_b200.sharded import ShardedPlanner

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world} (launch with torch.distributed.run)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    c = CFG
    G, K, T, E, nd, topk = c["G"], c["K"], c["T"], c["E"], c["nd"], c["topk"]
    K_total = K * world
    gen, fwd = synthetic.make_models(nd=nd, E=E, H=c["H"])
    planner = FusedPlanner(gen, fwd, horizon=T, device=dev)
    sharded = ShardedPlanner(planner) if world > 1 else None
    e0_h, goal_h, noise_h = synthetic.make_plan_inputs(G, K, T, nd, E, seed=1 + rank)
    e0_h, goal_h, _ = synthetic.make_plan_inputs(G, 1, 1, nd, E, seed=1)       # same goal on every rank
    e0_h, goal_h, noise_h = e0_h.pin_memory(), goal_h.pin_memory(), noise_h.pin_memory()
    e0, goal, noise = e0_h.to(dev), goal_h.to(dev), noise_h.to(dev)
    flush = l2_flusher(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def plan_resident():
        if sharded:
            return sharded.plan(e0, goal, noise, K_total, topk=topk)
        return planner.plan(e0, goal, noise, topk=topk)

    res_h = {"idx": torch.empty(G, dtype=torch.int32).pin_memory(), "act": torch.empty(G, 4).pin_memory(),
             "key": torch.empty(G, dtype=torch.int64).pin_memory()}

    def plan_e2e():
        # host buffers in, host result out: H2D of this step's inputs + D2H of the plan inside the timed region
        e0_d = e0_h.to(dev, non_blocking=True); goal_d = goal_h.to(dev, non_blocking=True)
        nz_d = noise_h.to(dev, non_blocking=True)
        p = sharded.plan(e0_d, goal_d, nz_d, K_total, topk=topk) if sharded else planner.plan(e0_d, goal_d, nz_d, topk=topk)
        res_h["idx"].copy_(p.best_idx, non_blocking=True)
        res_h["act"].copy_(p.best_action, non_blocking=True)
        res_h["key"].copy_(p.packed_min, non_blocking=True)
        return p

    def timed(fn, steps, warmup, split=False):
        for _ in range(warmup):
            fn()
        barrier()
        tot, k1 = 0.0, 0.0
        pairs = []
        for _ in range(steps):
            flush()                                   # L2 flush between timed iterations (untimed)
            a, b = ev(), ev()
            a.record(); fn(); b.record()
            pairs.append((a, b))
        barrier()
        tot = sum(a.elapsed_time(b) for a, b in pairs) / steps
        return tot

    def timed_k1(steps):
        pairs = []
        for _ in range(steps):
            flush()
            a, b = ev(), ev()
            a.record(); planner.rollout_score(e0, goal, noise); b.record()
            pairs.append((a, b))
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in pairs) / steps

    sampler = ClockSampler(local)
    plan0 = plan_resident()                            # sanity: a plan exists and is finite
    assert torch.isfinite(plan0.cost).all()
    n_before = None
    for _ in range(args.warmup):
        plan_resident()
    barrier()
    sampler.start()
    n_before = launch_count()
    ms = timed(plan_resident, args.steps, 0)
    n_launch = launch_count() - n_before
    k1_ms = timed_k1(args.steps)
    ms_e2e = timed(plan_e2e, args.steps, args.warmup)
    # ---- ndiv (BASELINE config 3) fwd+bwd, rows sharded when world > 1 is not timed here (N=1 figure per rank)
    x_h, z_h = synthetic.make_ndiv_inputs(NDIV["N"], NDIV["k"], NDIV["Dx"], NDIV["Dz"])
    x, z = x_h.to(dev), z_h.to(dev)

    def ndiv_step():
        xg = x.detach().requires_grad_(True)
        loss = diversity.compute_pairwise_divergence(xg, z)
        loss.backward()
        return loss

    ndiv_ms = timed(ndiv_step, args.steps, args.warmup)
    clocks = sampler.finish()

    def maxr(v):
        if world == 1:
            return v
        t = torch.tensor([v], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    ms, k1_ms, ms_e2e, ndiv_ms = maxr(ms), maxr(k1_ms), maxr(ms_e2e), maxr(ndiv_ms)
    if rank == 0:
        pk = peaks()
        units = G * K_total * T
        value = units / (ms * 1e-3)
        k1_tflops = G * K * T * F_ALG / (k1_ms * 1e-3) / 1e12
        pairs = NDIV["N"] * NDIV["k"] ** 2
        out = {
            "metric": "mpc_candidate_steps_per_sec", "value": value, "unit": "candidate-steps/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(workload="mpc_cfg2_K65536_T8_E128_H1024_nd2", per_gpu_K=K, K_total=K_total,
                           l2_flush_between_steps=True, precision=planner.precision, weights="default-init sep30",
                           **{k: v for k, v in c.items() if k != "K"}),
            "e2e": {"value": units / (ms_e2e * 1e-3), "unit": "candidate-steps/s", "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": int(noise_h.numel() * 4 + e0_h.numel() * 4 + goal_h.numel() * 4),
                    "d2h_bytes_per_step": int(G * (4 + 16 + 8))},
            "gpu_launches": int(n_launch),
            "clocks": clocks,
            "roofline": {"kernel": "rollout_f32_kernel", "bound": "tensor", "achieved": k1_tflops, "peak": pk["bf16"],
                         "unit": "TFLOP/s", "frac": k1_tflops / pk["bf16"], "traffic": None,
                         "peak_source": pk["src"] + " bf16 burst (kernel timed alone)",
                         "kernel_ms": k1_ms, "flop_per_candidate_step": F_ALG,
                         "note": "fp32 parity mode issues FFMA on CUDA cores (no tensor pipe); fraction is vs the "
                                 "tensor roofline the path is ultimately bound by"},
            "ndiv": {"metric": "ndiv_pairs_per_sec_fwd_bwd", "value": pairs / (ndiv_ms * 1e-3), "unit": "ordered pairs/s",
                     "ms_per_step": ndiv_ms, "config": NDIV},
        }
        if world == 1:
            t0 = time.perf_counter()
            threads = os.cpu_count() or 1
            times, cu = cpu_port_plan_rate(8192, 3, threads)
            out["cpu_baseline"] = {"value": cu / min(times), "unit": "candidate-steps/s", "cores": threads, "kind": "port",
                                   "sample": f"8192 of {K} candidates x T={T}, best of 3 after 1 warm-up, oracle port of "
                                             f"mpc_eval.py:124-165 incl. python argmin scan; {time.perf_counter() - t0:.1f}s"}
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    a = ap.parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_gpu(a)
