"""Dev check of the tcgen05 encoder layers (csrc/k5_encoder_tc.cu): codes of B = 16 / 27 / 64 frames against the fp64
oracle and against the CUDA-core kernels (child process with NDP_ENC_NO_TC=1), plus per-kernel device times
(torch.profiler) of one B = 64 forward."""
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from ndivplanning_b200 import synthetic  # noqa: E402
from ndivplanning_b200.encoder import FusedEncoder  # noqa: E402

BS = (16, 27, 64)


def run(path):
    enc = synthetic.make_image_encoder(seed=7)
    fe = FusedEncoder(enc, max_images=64, device="cuda:0")
    out = {}
    for B in BS:
        frames = synthetic.make_frames(B, seed=31)
        out[B] = fe(frames.cuda()).view(B, 128).cpu()
    torch.save(out, path)
    return enc, fe


if __name__ == "__main__":
    if len(sys.argv) > 1:
        run(sys.argv[1])
        sys.exit(0)
    from oracle import ndiv_oracle as O
    from torch.profiler import ProfilerActivity, profile
    with tempfile.TemporaryDirectory() as d:
        a, b = os.path.join(d, "tc.pt"), os.path.join(d, "cc.pt")
        enc, fe = run(a)
        subprocess.check_call([sys.executable, os.path.abspath(__file__), b], env=dict(os.environ, NDP_ENC_NO_TC="1"))
        tc, cc = torch.load(a), torch.load(b)
        ok = True
        for B in BS:
            frames = synthetic.make_frames(B, seed=31)
            with torch.no_grad():
                want = O.encoder_forward(enc.state_dict(), frames.double()).view(B, 128)
            e_tc = float((tc[B].double() - want).abs().max())
            e_cc = float((cc[B].double() - want).abs().max())
            print(f"B={B}: |tc - fp64| {e_tc:.3e}   |cuda-core fp32 - fp64| {e_cc:.3e}   |tc - cuda-core| "
                  f"{float((tc[B] - cc[B]).abs().max()):.3e}   code magnitude {float(want.abs().max()):.3f}", flush=True)
            ok &= e_tc < 8e-6
        frames = synthetic.make_frames(64, seed=3).cuda()
        for _ in range(3):
            fe(frames)
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for _ in range(5):
                fe(frames)
            torch.cuda.synchronize()
        tot = 0.0
        for e in sorted(prof.key_averages(), key=lambda e: -e.device_time_total):
            if e.device_time_total > 0:
                print(f"  {e.device_time_total / 5:8.1f} us/forward x{e.count // 5:3d}  {e.key[:100]}")
                tot += e.device_time_total / 5
        print(f"  sum {tot:.1f} us per B=64 forward")
        print("encoder_tc_check", "ok" if ok else "MISMATCH")
        sys.exit(0 if ok else 1)
