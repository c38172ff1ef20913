"""One warm + a few timed encoder forwards at a fixed batch, for ncu launch lists.  Usage: encoder_once.py [B]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ndivplanning_b200 import synthetic  # noqa: E402
from ndivplanning_b200.encoder import FusedEncoder  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
fe = FusedEncoder(synthetic.make_image_encoder(seed=7), max_images=B, device="cuda:0")
frames = synthetic.make_frames(B, seed=3).to("cuda:0")
for _ in range(3):
    y = fe(frames)
torch.cuda.synchronize()
print("ok", float(y.abs().max()))
