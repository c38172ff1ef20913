"""Device-resident image encoder — SURVEY.md 8f row N3 (first slice): the reference's
``models.image_autoencoder.Encoder.forward`` (models/image_autoencoder.py:35-49) in eval mode, as called at
mpc_eval.py:135-136 (``image_encoder(state_cur)``) and train_gan.py:152-153.

``FusedEncoder(module)`` takes a reference ``Encoder`` (or the product's ``ImageEncoder`` container), folds the
eval-mode BatchNorm of conv1..conv3 into the convolution weights, permutes them to the kernel's
[Cout][KH][KW][Cin] order and runs every frame batch through ``ndp_encoder_forward`` (csrc/k5_encoder.cu).
Calling it returns ``[B,128,1,1]`` like the reference.  Training-mode BatchNorm (batch statistics) is outside this
slice: a module in training mode is refused.  No compute happens in Python and there is no CPU path.
"""
import ctypes as C

import torch

from . import _lib

CONVS = ("conv1", "conv2", "conv3", "conv4", "conv5", "conv6")
BN_AFTER = ("conv1", "conv2", "conv3")        # image_autoencoder.py:37-41; conv4_bn/conv5_bn are never applied
IMG = 128


def fold_state_dict(sd, eps: float = 1e-5):
    """[(W [Cout,KH,KW,Cin], b [Cout])] per layer, fp32, BatchNorm (running statistics) folded in.  The fold is
    done in fp64 and rounded once."""
    out = []
    for name in CONVS:
        W = sd[f"{name}.weight"].detach().double()
        b = sd[f"{name}.bias"].detach().double()
        if name in BN_AFTER:
            g = sd[f"{name}_bn.weight"].detach().double()
            beta = sd[f"{name}_bn.bias"].detach().double()
            mean = sd[f"{name}_bn.running_mean"].detach().double()
            var = sd[f"{name}_bn.running_var"].detach().double()
            scale = g / torch.sqrt(var + eps)
            W = W * scale.view(-1, 1, 1, 1)
            b = (b - mean) * scale + beta
        out.append((W.permute(0, 2, 3, 1).contiguous().float(), b.float().contiguous()))
    return out


class FusedEncoder:
    def __init__(self, module, *, max_images: int = 64, device=None):
        if not torch.cuda.is_available():
            raise _lib.NdpError("FusedEncoder needs an sm_100a GPU; there is no CPU fallback")
        self.device = torch.device(device if device is not None else torch.cuda.current_device())
        if self.device.type != "cuda":
            raise _lib.NdpError("FusedEncoder needs an sm_100a GPU; there is no CPU fallback")
        if module.training:
            raise ValueError("FusedEncoder runs eval-mode BatchNorm (running statistics): call module.eval() first")
        self.module, self.max_images = module, int(max_images)
        self._h = C.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().ndp_encoder_create(C.byref(self._h), self.max_images, IMG))
        self.load(module.state_dict(), eps=getattr(getattr(module, "conv1_bn", None), "eps", 1e-5))

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            try:
                _lib.lib().ndp_encoder_destroy(h)
            except Exception:       # interpreter shutdown
                pass

    def load(self, state_dict, eps: float = 1e-5) -> None:
        folded = [(W.to(self.device), b.to(self.device)) for W, b in fold_state_dict(state_dict, eps)]
        Wp = (C.c_void_p * 6)(*[W.data_ptr() for W, _ in folded])
        bp = (C.c_void_p * 6)(*[b.data_ptr() for _, b in folded])
        with torch.cuda.device(self.device):
            st = torch.cuda.current_stream(self.device)
            _lib.check(_lib.lib().ndp_encoder_load(self._h, Wp, bp, C.c_void_p(st.cuda_stream)))
            st.synchronize()        # the staged copies may be freed after this

    def __call__(self, images: torch.Tensor) -> torch.Tensor:
        if images.device != self.device or images.dtype != torch.float32:
            raise _lib.NdpError("FusedEncoder: images must be fp32 on %s (no CPU path)" % self.device)
        if images.dim() != 4 or tuple(images.shape[1:]) != (3, IMG, IMG):
            raise ValueError("FusedEncoder: images must be [B,3,%d,%d], got %s" % (IMG, IMG, tuple(images.shape)))
        B = images.shape[0]
        if B == 0:
            return torch.empty(0, 128, 1, 1, dtype=torch.float32, device=self.device)
        images = images.contiguous()
        codes = torch.empty(B, 128, dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            st = torch.cuda.current_stream(self.device)
            for lo in range(0, B, self.max_images):
                n = min(self.max_images, B - lo)
                _lib.check(_lib.lib().ndp_encoder_forward(self._h, C.c_void_p(images[lo:lo + n].data_ptr()), n,
                                                          C.c_void_p(codes[lo:lo + n].data_ptr()),
                                                          C.c_void_p(st.cuda_stream)))
        return codes.view(B, 128, 1, 1)

    def capture(self, images: torch.Tensor):
        """(graph, codes): one forward over the static `images` tensor as a CUDA graph; refill `images` in place and
        replay() — `codes` [B,128,1,1] is overwritten by every replay.  B must be <= max_images."""
        if images.shape[0] > self.max_images:
            raise ValueError("capture: B=%d > max_images=%d" % (images.shape[0], self.max_images))
        self(images)                                    # warm: one-time function attributes are set outside capture
        g = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            with torch.cuda.graph(g, stream=side):
                codes = self(images)
        torch.cuda.current_stream(self.device).wait_stream(side)
        return g, codes
