import sys, numpy as np, torch
sys.path.insert(0, '/root/repo')
from ndivplanning_b200 import GanTrainer, synthetic
cuda = torch.device('cuda', 0)
Fb, S, nd = 3000, 8, 2
gen, dis = synthetic.make_gan_models(nd, seed=2)
batch = [t.to(cuda) for t in synthetic.make_gan_batch(Fb, S, nd, seed=11)]
res = []
for rep in range(4):
    tr = GanTrainer(gen, dis, num_sample=S, noise_dim=nd, max_codes=Fb, pairwise_div_factor=0.0, device=cuda)
    l = tr.step(*batch).cpu().numpy()
    gg, dg = [t.cpu().numpy() for t in tr.grads()]
    res.append((l, gg, dg))
for r in res[1:]:
    print('losses equal', np.array_equal(r[0], res[0][0]), 'G grad equal', np.array_equal(r[1], res[0][1]), 'max diff', np.abs(r[1]-res[0][1]).max(), 'D grad equal', np.array_equal(r[2], res[0][2]), np.abs(r[2]-res[0][2]).max())
