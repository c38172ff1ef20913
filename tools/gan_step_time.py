"""Times the device-resident GAN training step (ndp_gan_train_step) eagerly and as a CUDA-graph replay, and the
oracle restatement (torch CPU, the reference's ops) beside it.  python tools/gan_step_time.py [F S [F S ...]]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ndivplanning_b200 import GanTrainer, launch_count, synthetic  # noqa: E402

# per generated row: generator fwd 2*83,200 + bwd 2x; discriminator (real rows amortised): fwd x2 (+1/S), bwd dW+dX
G_MAC = 258 * 128 + 128 * 64 + 64 * 128 + 128 * 256 + 256 * 4
D_MAC = 260 * 64 + 64 * 128 + 128 * 256 + 256


def flop_per_step(F, S):
    B = F * S
    g = 2 * G_MAC * B * 3                                       # fwd + dX + dW
    d = 2 * D_MAC * (B + F) * 3 + 2 * D_MAC * B * 2             # D step (fwd,dX,dW on B+F rows) + G step (fwd,dX)
    return g + d


def main():
    dev = torch.device("cuda", 0)
    shapes = [(56, 6), (4096, 16)]
    if len(sys.argv) > 2:
        a = [int(v) for v in sys.argv[1:]]
        shapes = list(zip(a[0::2], a[1::2]))
    for F, S in shapes:
        gen, dis = synthetic.make_gan_models(2, seed=0)
        tr = GanTrainer(gen, dis, num_sample=S, noise_dim=2, max_codes=F, device=dev)
        batch = [t.to(dev) for t in synthetic.make_gan_batch(F, S, 2, seed=5)]
        for _ in range(3):
            tr.step(*batch)
        torch.cuda.synchronize()
        n = 50
        l0 = launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            tr.step(*batch)
        e1.record(); torch.cuda.synchronize()
        eager = e0.elapsed_time(e1) / n
        launches = (launch_count() - l0) / n
        g = tr.capture(*batch)
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(n):
            g.replay()
        e1.record(); torch.cuda.synchronize()
        graph = e0.elapsed_time(e1) / n
        fl = flop_per_step(F, S)
        print(f"F={F} S={S} rows={F*S}: eager {eager*1e3:.1f} us, graph {graph*1e3:.1f} us, {launches:.0f} launches/step, "
              f"{F*S/graph*1e3:.3e} samples/s, {fl/graph/1e9:.2f} TFLOP/s fp32", flush=True)
        if os.environ.get("GAN_CPU", "1") == "1":
            from oracle import ndiv_oracle as O
            og, od = O.GeneratorMLP(2), O.ActionDiscriminator()
            G_opt, D_opt = O.gan_optimizers(og, od)
            c, a, z = synthetic.make_gan_batch(F, S, 2, seed=5)
            O.gan_train_step(og, od, G_opt, D_opt, c, a, z)
            reps = 5 if F * S > 10000 else 20
            t0 = time.perf_counter()
            for _ in range(reps):
                O.gan_train_step(og, od, G_opt, D_opt, c, a, z)
            cpu = (time.perf_counter() - t0) / reps
            print(f"   reference ops on {torch.get_num_threads()} host threads: {cpu*1e3:.2f} ms/step "
                  f"({F*S/cpu:.3e} samples/s) -> x{cpu*1e3/graph:.0f}", flush=True)


if __name__ == "__main__":
    main()
