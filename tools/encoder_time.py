"""Time ndp_encoder_forward (csrc/k5_encoder.cu) with CUDA events; the torch/cuDNN eval-mode forward of the same
module is printed beside it as a library yardstick (not a product path).  Usage: python tools/encoder_time.py"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ndivplanning_b200 import synthetic  # noqa: E402
from ndivplanning_b200.encoder import FusedEncoder  # noqa: E402

FLOPS_PER_IMAGE = 2 * sum(oh * oh * co * 9 * ci for oh, ci, co in
                          ((64, 3, 64), (32, 64, 128), (16, 128, 256), (8, 256, 512), (4, 512, 1024))) + 2 * 128 * 16384


def time_ms(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    dev = torch.device("cuda:0")
    enc = synthetic.make_image_encoder(seed=7)
    fe = FusedEncoder(enc, max_images=64, device=dev)
    lib = enc.to(dev)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    for B in (1, 2, 8, 16, 64):
        frames = synthetic.make_frames(B, seed=3).to(dev)
        t = time_ms(lambda: fe(frames))
        with torch.no_grad():
            tl = time_ms(lambda: lib(frames))
            err = float((fe(frames) - lib(frames)).abs().max())
        print(json.dumps({"B": B, "ms": round(t, 4), "images_per_s": round(B / t * 1e3, 1),
                          "fp32_tflops": round(B * FLOPS_PER_IMAGE / t / 1e9, 2), "cudnn_fp32_ms": round(tl, 4),
                          "max_abs_diff_vs_cudnn": err}))


if __name__ == "__main__":
    main()
