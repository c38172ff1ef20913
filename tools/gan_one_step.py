"""two eager GAN steps at F S (for an ncu launch list): python tools/gan_one_step.py F S"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ndivplanning_b200 import GanTrainer, synthetic
F, S = int(sys.argv[1]), int(sys.argv[2])
dev = torch.device("cuda", 0)
gen, dis = synthetic.make_gan_models(2, seed=0)
tr = GanTrainer(gen, dis, num_sample=S, noise_dim=2, max_codes=F, device=dev)
b = [t.to(dev) for t in synthetic.make_gan_batch(F, S, 2, seed=5)]
for _ in range(2):
    tr.step(*b)
torch.cuda.synchronize()
print("ok", tr.losses.cpu().numpy())
