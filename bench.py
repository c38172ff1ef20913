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
import gc
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


_REAL_STDOUT = None


def quiet_stdout():
    """libraries (NCCL's version banner) print to fd 1; the contract is ONE JSON line on stdout.  Everything
    written to fd 1 from now on goes to stderr; emit() writes the JSON line to the real stdout."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(obj):
    line = (json.dumps(obj) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(line.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, line)


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
    Ks = CFG["K"]                       # the whole cfg2 candidate set per step (~0.6 s of CPU work per step)
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
                            "sample": f"all {Ks} candidates per step x T={CFG['T']} (no sub-sampling), oracle port of "
                                      f"mpc_eval.py:124-165 incl. python argmin, torch {torch.__version__} CPU fp32"},
           "e2e": {"value": v, "unit": "candidate-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(out)


# --------------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------------
K1_KERNEL = {"fp32": "rollout_f32_kernel", "f16": "rollout_tc_kernel<1>", "f16x2": "rollout_tc_kernel<2>",
             "f16x3": "rollout_tc_kernel<3>", "f16+refine": "rollout_tc_kernel<1>"}
MMA_PASSES = {"fp32": 0, "f16": 1, "f16x2": 2, "f16x3": 3, "f16+refine": 1}
NCU_TRAFFIC = {"rollout_tc_kernel<1>": 4.98e6, "rollout_f32_kernel": 5.69e6}   # bytes per launch, ncu (profiles/)
POOL = 40   # distinct noise sets rotated through the timed region: 40 x 4.2 MB = 168 MB > 126 MB of L2


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
    prec = args.precision
    gen, fwd = synthetic.make_models(nd=nd, E=E, H=c["H"])
    planner = FusedPlanner(gen, fwd, horizon=T, device=dev, precision=prec)
    sharded = ShardedPlanner(planner) if world > 1 else None
    # steady-state loop object: pre-allocated plan slots, tails (refine / exchange / merge) on side streams
    ring = sharded.ring(G, K_total, topk) if sharded else planner.ring(G, K, topk)
    e0_h, goal_h, _ = synthetic.make_plan_inputs(G, 1, 1, nd, E, seed=1)       # same goal on every rank
    g_ = torch.Generator().manual_seed(100 + rank)
    pool_h = [torch.rand(T, G, K, nd, generator=g_).pin_memory() for _ in range(POOL)]   # U[0,1) like the sampler
    e0_h, goal_h = e0_h.pin_memory(), goal_h.pin_memory()
    e0, goal = e0_h.to(dev), goal_h.to(dev)
    pool = [n.to(dev) for n in pool_h]
    flush = l2_flusher(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def plan_dev(nz):
        return ring.plan(e0, goal, nz)

    def run_resident(steps, first):
        """`steps` plans back to back on device-resident inputs; one event pair around all of them"""
        a, b = ev(), ev()
        a.record()
        plans = []
        for i in range(steps):
            plans.append(plan_dev(pool[(first + i) % POOL]))
            if len(plans) > 2:                      # plan i-2 is consumed (read on this stream) before plan i+1 starts
                plans.pop(0).wait()
        for p_ in plans:
            p_.wait()
        b.record()
        return a, b, plans

    res_h = [{"idx": torch.empty(G, dtype=torch.int32).pin_memory(), "act": torch.empty(G, 4).pin_memory(),
              "key": torch.empty(G, dtype=torch.int64).pin_memory(), "done": torch.cuda.Event()} for _ in range(4)]

    copy_stream = torch.cuda.Stream(device=dev)
    NBUF = 4                                           # input staging buffers = ring depth (a slot's tail frees both)
    in_d = [{"nz": torch.empty(T, G, K, nd, device=dev), "e0": torch.empty(G, E, device=dev),
             "goal": torch.empty(G, E, device=dev), "up": torch.cuda.Event(), "free": torch.cuda.Event()}
            for _ in range(NBUF)]

    out_stream = torch.cuda.Stream(device=dev)

    def read_back(q, r):
        """D2H of a finished plan on its own stream: the rollout stream never waits for a tail"""
        out_stream.wait_event(q.ready)
        with torch.cuda.stream(out_stream):
            r["idx"].copy_(q.best_idx, non_blocking=True); r["act"].copy_(q.best_action, non_blocking=True)
            r["key"].copy_(q.packed_min, non_blocking=True)
            r["done"].record(out_stream)

    e2e_host = {}

    def run_e2e(steps, first):
        """host buffers in, host results out, every step: H2D of the step's noise/e0/goal from pinned memory
        (copy stream, NBUF staging buffers so it overlaps the previous plans' rollouts), plan, D2H of the plan"""
        a, b = ev(), ev()
        main = torch.cuda.current_stream(dev)
        t_host = time.perf_counter()
        a.record()
        copy_stream.wait_stream(main)
        prev = None
        for i in range(steps):
            d = in_d[i % NBUF]
            with torch.cuda.stream(copy_stream):
                if i >= NBUF:
                    copy_stream.wait_event(d["free"])          # the plan that read this buffer has finished
                d["nz"].copy_(pool_h[(first + i) % POOL], non_blocking=True)
                d["e0"].copy_(e0_h, non_blocking=True); d["goal"].copy_(goal_h, non_blocking=True)
                d["up"].record(copy_stream)
            main.wait_event(d["up"])
            if i >= 4:
                main.wait_event(res_h[i & 3]["done"])          # plan i-4 (same ring slot) has been read back
            p_ = ring.plan(d["e0"], d["goal"], d["nz"])
            d["free"] = p_.ready
            if prev is not None:                       # results of the previous plan leave while this one runs
                read_back(*prev)
            prev = (p_, res_h[i & 3])
        read_back(*prev)
        main.wait_stream(out_stream)
        b.record()
        e2e_host["ms"] = (time.perf_counter() - t_host) * 1e3 / steps   # host time to ISSUE a step (no sync inside)
        return a, b

    def timed_k1(pl, steps):
        pairs = []
        for i in range(steps):
            flush()
            torch.cuda._sleep(1_000_000)               # ~0.5 ms: the host enqueues the launch before `a` is reached
            a, b = ev(), ev()
            a.record(); pl.rollout_score(e0, goal, pool[i % POOL]); b.record()
            pairs.append((a, b))
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in pairs) / steps

    # ---- parity of the benchmarked configuration against the fp32 CUDA path (outside the timed region)
    parity = None
    if world == 1:
        exact = FusedPlanner(gen, fwd, horizon=T, device=dev, precision="fp32").plan(e0, goal, pool[0], topk=topk)
        got = planner.plan(e0, goal, pool[0], topk=topk)
        torch.cuda.synchronize()
        sel = sorted(set(got.topk_idx[0].tolist() + [int(got.best_idx[0])]))
        rel = ((got.cost - exact.cost).abs() / exact.cost)
        parity = {"vs": "fp32 CUDA-core path (itself parity-tested against the oracle)",
                  "argmin_equal": bool(int(got.best_idx[0]) == int(exact.best_idx[0])),
                  "topk_set_equal": bool(set(got.topk_idx[0].tolist()) == set(exact.topk_idx[0].tolist())),
                  "verified": got.verified.tolist(),
                  "cost_rel_err_selected_max": float(rel[0, sel].max()), "cost_rel_err_all_max": float(rel.max())}
        del exact
    # setup (untimed, not counted as warm-up): first use of every stream / NCCL channel / host buffer of both loops
    run_resident(8, 3)
    run_e2e(4, 3)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    run_resident(args.warmup, POOL - args.warmup)      # W warm-up steps: the same pipelined loop as the timed region
    gc.collect(); gc.disable()                         # no collector pauses inside the timed regions; collected BEFORE
    barrier()                                          # the barrier so the ranks enter the timed region together
    n_before = launch_count()
    a, b, plans = run_resident(args.steps, 0)
    barrier()
    ms = a.elapsed_time(b) / args.steps
    n_launch = launch_count() - n_before
    assert all(torch.isfinite(p_.cost).all() for p_ in plans[-1:])
    del plans
    k1_ms = timed_k1(planner, min(args.steps, 10))
    run_e2e(min(args.warmup, 3), 0)
    barrier()
    a, b = run_e2e(args.steps, 7)
    barrier()
    ms_e2e = a.elapsed_time(b) / args.steps
    # ---- ndiv (BASELINE config 3) fwd+bwd (N=1 figure per rank)
    x_h, z_h = synthetic.make_ndiv_inputs(NDIV["N"], NDIV["k"], NDIV["Dx"], NDIV["Dz"])
    x, z = x_h.to(dev), z_h.to(dev)

    def ndiv_step():
        xg = x.detach().requires_grad_(True)
        loss = diversity.compute_pairwise_divergence(xg, z)
        loss.backward()
        return loss

    for _ in range(args.warmup):
        ndiv_step()
    pairs_ = []
    for _ in range(args.steps):
        flush()
        torch.cuda._sleep(1_000_000)                   # the step's launches are queued before `a` is reached
        a, b = ev(), ev()
        a.record(); ndiv_step(); b.record()
        pairs_.append((a, b))
    torch.cuda.synchronize()
    ndiv_ms = sum(a.elapsed_time(b) for a, b in pairs_) / args.steps
    clocks = sampler.finish()
    # ---- GAN training step around ndiv (context, rank 0 at N=1): CUDA-graph replays of one
    # ndp_gan_train_step at the reference's batch (F=56 codes x S=6 samples) and at F=4096 x S=16
    gan_steps = {}
    if world == 1:
        from ndivplanning_b200 import GanTrainer
        for Fg, Sg in ((56, 6), (4096, 16)):
            ggen, gdis = synthetic.make_gan_models(2, seed=0)
            tr = GanTrainer(ggen, gdis, num_sample=Sg, noise_dim=2, max_codes=Fg, device=dev)
            gb = [t.to(dev) for t in synthetic.make_gan_batch(Fg, Sg, 2, seed=5)]
            for _ in range(2):
                tr.step(*gb)
            l0 = launch_count()
            tr.step(*gb)
            per_step = launch_count() - l0
            cg = tr.capture(*gb)
            for _ in range(3):
                cg.replay()
            torch.cuda.synchronize()
            n_ = 20
            a, b = ev(), ev()
            a.record()
            for _ in range(n_):
                cg.replay()
            b.record(); torch.cuda.synchronize()
            g_ms = a.elapsed_time(b) / n_
            gan_steps[f"F{Fg}_S{Sg}"] = {"ms_per_step": g_ms, "samples_per_s": Fg * Sg / (g_ms * 1e-3),
                                         "kernel_launches_per_step": per_step,
                                         "losses_D_G_div": [float(v) for v in tr.losses.cpu()]}
            del cg, tr
    # ---- image encoder in front of the path (first slice; context, rank 0 at N=1): CUDA-graph replays
    # of ndp_encoder_forward at a plan's two frames, a trajectory's 8 and a GAN batch's 64
    enc_runs = {}
    if world == 1:
        from ndivplanning_b200.encoder import FusedEncoder
        fe = FusedEncoder(synthetic.make_image_encoder(seed=7), max_images=64, device=dev)
        for Be in (2, 8, 64):
            fr = synthetic.make_frames(Be, seed=3).to(dev)
            l0 = launch_count()
            fe(fr)
            per_call = launch_count() - l0
            cg, _codes = fe.capture(fr)
            for _ in range(3):
                cg.replay()
            torch.cuda.synchronize()
            n_ = 20
            a, b = ev(), ev()
            a.record()
            for _ in range(n_):
                cg.replay()
            b.record(); torch.cuda.synchronize()
            e_ms = a.elapsed_time(b) / n_
            enc_runs[f"B{Be}"] = {"ms": e_ms, "images_per_s": Be / (e_ms * 1e-3),
                                  "fp32_tflops": Be * 0.6224e9 / (e_ms * 1e-3) / 1e12, "kernel_launches": per_call}
            del cg
        del fe
    # ---- other precision modes of the same kernel family, for context (short runs)
    modes = {}
    if world == 1:
        for m in ("f16", "f16x3", "fp32"):
            pl = FusedPlanner(gen, fwd, horizon=T, device=dev, precision=m)
            n_ = 3 if m == "fp32" else 10
            for i in range(2):
                pl.plan(e0, goal, pool[i], topk=topk)
            torch.cuda.synchronize()
            a, b = ev(), ev()
            a.record()
            for i in range(n_):
                pl.plan(e0, goal, pool[(3 + i) % POOL], topk=topk)
            b.record(); torch.cuda.synchronize()
            pms = a.elapsed_time(b) / n_
            kms = timed_k1(pl, n_)
            modes[m] = {"plan_ms": pms, "candidate_steps_per_s": G * K * T / (pms * 1e-3), "k1_ms": kms,
                        "k1_algorithmic_tflops": G * K * T * F_ALG / (kms * 1e-3) / 1e12}
            del pl

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
        passes = MMA_PASSES[prec]
        pairs = NDIV["N"] * NDIV["k"] ** 2
        out = {
            "metric": "mpc_candidate_steps_per_sec", "value": value, "unit": "candidate-steps/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if prec == "fp32" else "f16 operands (split hi/lo where stated), f32 accumulate/state",
            "data": "synthetic",
            "config": dict(workload="mpc_cfg2_K65536_T8_E128_H1024_nd2", per_gpu_K=K, K_total=K_total,
                           l2="inputs rotate through a %d MB pool (> 126 MB L2); weights stay L2-resident by design"
                              % (POOL * T * G * K * nd * 4 >> 20),
                           precision=prec, weights="default-init sep30",
                           step="one plan: rollout of all K + top-%d/argmin%s%s" % (
                               topk, " + split-operand refine of the 128 best" if prec == "f16+refine" else "",
                               "; tails (refine/select/gather%s) of consecutive plans run on 2 side streams under the next "
                               "bulk rollouts (PlanRing, 4 slots)" % (" + record all-gather + merge" if world > 1 else "")),
                           **{k: v for k, v in c.items() if k != "K"}),
            "e2e": {"value": units / (ms_e2e * 1e-3), "unit": "candidate-steps/s", "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": int(pool_h[0].numel() * 4 + e0_h.numel() * 4 + goal_h.numel() * 4),
                    "d2h_bytes_per_step": int(G * (4 + 16 + 8)),
                    "host_issue_ms_per_step": e2e_host.get("ms")},
            "gpu_launches": int(n_launch),
            "clocks": clocks,
            "roofline": {"kernel": K1_KERNEL[prec], "bound": "tensor", "achieved": k1_tflops, "peak": pk["bf16"],
                         "unit": "TFLOP/s", "frac": k1_tflops / pk["bf16"],
                         "traffic": NCU_TRAFFIC.get(K1_KERNEL[prec]),
                         "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed "
                                           "ncu --set full captures (profiles/r01_summary.md); algorithmic HBM bytes per "
                                           "launch: 4.19 MB noise + 0.7 MB fp16 weights in, 1.3 MB cost/a0 out (stays in L2)",
                         "peak_source": pk["src"] + " bf16 burst (kernel timed alone, L2 flushed before each launch)",
                         "kernel_ms": k1_ms, "flop_per_candidate_step": F_ALG, "mma_passes": passes,
                         "issued_mma_frac": (k1_tflops * max(passes, 1) / pk["bf16"]) if passes else None,
                         "note": "achieved = algorithmic FLOP (666,112 per candidate-step) / CUDA-event time of the bulk "
                                 "rollout kernel; fp16 and bf16 share the tcgen05 kind::f16 rate"},
            "ndiv": {"metric": "ndiv_pairs_per_sec_fwd_bwd", "value": pairs / (ndiv_ms * 1e-3), "unit": "ordered pairs/s",
                     "ms_per_step": ndiv_ms, "config": NDIV,
                     "roofline": {"bound": "MUFU (3 special-function ops per ordered pair: 2 sqrt + 1 rcp; SURVEY 8d)",
                                  "floor_ms": pairs * 3 / (148 * 16 * 1.965e9) * 1e3,
                                  "frac": pairs * 3 / (148 * 16 * 1.965e9) * 1e3 / ndiv_ms,
                                  "note": "Gram-form distances on tcgen05 (fp16 split operands), fused hinge/gradient "
                                          "epilogue, exact direct-difference kernels gated behind it; floor = pairs x 3 / "
                                          "(148 SMs x 16 MUFU/clk x 1.965 GHz); ms_per_step includes the host autograd "
                                          "wrapper's 11 launches"}},
            "parity": parity,
            "modes": modes,
            "gan_train_step": {"what": "one iteration of train_gan.py:159-203 on encoded inputs (generator + discriminator "
                                       "forward/backward, BCE, ndiv, two Adam updates), CUDA-graph replay",
                               "shapes": gan_steps},
            "image_encoder": {"what": "models/image_autoencoder.py Encoder forward (eval), [B,3,128,128] -> [B,128], exact "
                                      "fp32 implicit-GEMM kernels, CUDA-graph replay; 0.622 GFLOP per image",
                              "batches": enc_runs},
        }
        if world == 1:
            t0 = time.perf_counter()
            threads = os.cpu_count() or 1
            times, cu = cpu_port_plan_rate(K, 10, threads)
            out["cpu_baseline"] = {"value": cu / min(times), "unit": "candidate-steps/s", "cores": threads, "kind": "port",
                                   "sample": f"all {K} candidates x T={T}, best of 10 plans after 1 warm-up, oracle port of "
                                             f"mpc_eval.py:124-165 incl. python argmin scan; {time.perf_counter() - t0:.1f}s"}
        emit(out)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--precision", default="f16+refine", choices=["fp32", "f16", "f16x2", "f16x3", "f16+refine"])
    a = ap.parse_args()
    quiet_stdout()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_gpu(a)
