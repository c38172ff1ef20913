"""step-by-step comparison of ndp_gan_train_step with the oracle restatement (fp32 and fp64) on the box's CPU"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ndivplanning_b200 import GanTrainer, synthetic
from ndivplanning_b200.gan_train import D_LAYERS, G_LAYERS
from oracle import ndiv_oracle as O

Fb, S, nd, nsteps = 56, 6, 2, 3
dev = torch.device("cuda", 0)
gen, dis = synthetic.make_gan_models(nd, seed=0)
tr = GanTrainer(gen, dis, num_sample=S, noise_dim=nd, max_codes=Fb, device=dev)
refs = {}
for dt in (torch.float32, torch.float64):
    og, od = O.GeneratorMLP(nd).to(dt), O.ActionDiscriminator().to(dt)
    og.load_state_dict({k: v.to(dt) for k, v in gen.state_dict().items()})
    od.load_state_dict({k: v.to(dt) for k, v in dis.state_dict().items()})
    refs[dt] = (og, od) + O.gan_optimizers(og, od)

def flat(m, grad=False):
    return torch.cat([(p.grad if grad else p.detach()).reshape(-1).double() for p in m.parameters()]).numpy()
def flat_sd(sd, layers):
    return torch.cat([sd[f"{n}.{k}"].reshape(-1) for n in layers for k in ("weight", "bias")]).double().cpu().numpy()
def stat(name, mine, r32, r64):
    em, er = np.abs(mine - r64), np.abs(r32 - r64)
    print(f"  {name:12s} mine-vs-64 max {em.max():.3e} med {np.median(em):.3e} rms {np.sqrt((em**2).mean()):.3e} | "
          f"ref32-vs-64 max {er.max():.3e} med {np.median(er):.3e} rms {np.sqrt((er**2).mean()):.3e}")

for it in range(nsteps):
    c, a, z = synthetic.make_gan_batch(Fb, S, nd, seed=31 + 100 * it)
    l = tr.step(c.to(dev), a.to(dev), z.to(dev)).cpu().numpy()
    gg, dg = [t.double().cpu().numpy() for t in tr.grads()]
    out = {}
    for dt in refs:
        og, od, Go, Do = refs[dt]
        dl, gl, pl, ah = O.gan_train_step(og, od, Go, Do, c.to(dt), a.to(dt), z.to(dt))
        out[dt] = (np.array([float(dl), float(gl), float(pl)]), flat(og, True), flat(og), flat(od), ah.double().numpy())
    print(f"step {it}: losses mine {l} ref32 {out[torch.float32][0]} ref64 {out[torch.float64][0]}")
    stat("a_hat", tr.last_actions(Fb).double().cpu().numpy(), out[torch.float32][4], out[torch.float64][4])
    stat("G grad", gg, out[torch.float32][1], out[torch.float64][1])
    gsd, dsd = tr.state_dicts()
    stat("G params", flat_sd(gsd, G_LAYERS), out[torch.float32][2], out[torch.float64][2])
    stat("D params", flat_sd(dsd, D_LAYERS), out[torch.float32][3], out[torch.float64][3])
    # per-layer generator gradient error
    off = 0
    for n in G_LAYERS:
        for k in ("weight", "bias"):
            sz = gsd[f"{n}.{k}"].numel()
            em = np.abs(gg[off:off+sz] - out[torch.float64][1][off:off+sz]); er = np.abs(out[torch.float32][1][off:off+sz] - out[torch.float64][1][off:off+sz])
            sc = np.abs(out[torch.float64][1][off:off+sz]).max()
            print(f"     grad {n}.{k:6s} scale {sc:.3e} mine max {em.max():.3e} med {np.median(em):.3e} | ref32 max {er.max():.3e} med {np.median(er):.3e}")
            off += sz
