"""Host-side module definitions the planning path consumes.

These are *containers for weights* with the reference's call signatures; the
B200 path never runs their ``forward`` — `FusedPlanner` reads ``state_dict()``
and runs the sm_100a kernels.  ``forward`` is the reference-style definition so
the modules stay usable in the reference's own scripts.

* ``GeneratorMLP`` has the layout of ``models.gan.Decoder`` (reference
  models/gan.py:61-86): fc1..fc5, (256+nd)->128->64->128->256->4.  A reference
  ``Decoder`` instance can be passed anywhere a ``GeneratorMLP`` is accepted.
* ``LatentForward`` is the embedding-space forward model with the call contract of
  ``ForwardAutoencoder.forward`` (reference models/forward_encoder.py:105-114):
  ``cat([code, actions], 1)`` (state first), eval mode returns ``state + residual``.
  The reference has no latent dynamics module; body per SURVEY.md §8a-F.
* ``ActionDiscriminator`` has the layout of ``models.gan.Discriminator`` (reference
  models/gan.py:89-110): fc1..fc4, (256+4)->64->128->256->1, leaky ReLU, input
  ``cat([action, state_code], 1)``.  Weight container for ``GanTrainer``.
* ``ImageEncoder`` has the layout of ``models.image_autoencoder.Encoder`` (reference
  models/image_autoencoder.py:14-49): conv1..conv5 3x3 stride 2 (3->64->128->256->512->1024),
  BatchNorm applied after conv1..conv3 only (conv4_bn / conv5_bn exist as parameters and are never
  applied, as in the reference), conv6 4x4 valid 1024->128.  Weight container for ``FusedEncoder``.
"""
import torch
from torch import nn
from torch.nn import functional as F


class GeneratorMLP(nn.Module):
    def __init__(self, noise_dim: int, cond_dim: int = 256):
        super().__init__()
        self.fc1 = nn.Linear(cond_dim + noise_dim, 128)
        self.fc2 = nn.Linear(128, 64)
        self.fc3 = nn.Linear(64, 128)
        self.fc4 = nn.Linear(128, 256)
        self.fc5 = nn.Linear(256, 4)

    def forward(self, z):
        h = F.relu(self.fc1(z))
        h = F.relu(self.fc2(h))
        h = F.relu(self.fc3(h))
        h = F.relu(self.fc4(h))
        return self.fc5(h)


class LatentForward(nn.Module):
    def __init__(self, E: int = 128, H: int = 1024, A: int = 4):
        super().__init__()
        self.fc1 = nn.Linear(E + A, H)
        self.fc2 = nn.Linear(H, E)

    def forward(self, state_cur, actions):
        x = torch.cat([state_cur.flatten(1), actions], dim=1)
        r = self.fc2(F.relu(self.fc1(x))).view_as(state_cur)
        return r if self.training else state_cur + r


class ActionDiscriminator(nn.Module):
    def __init__(self, cond_dim: int = 256, A: int = 4):
        super().__init__()
        self.fc1 = nn.Linear(cond_dim + A, 64)
        self.fc2 = nn.Linear(64, 128)
        self.fc3 = nn.Linear(128, 256)
        self.fc4 = nn.Linear(256, 1)

    def forward(self, action, state_code):
        h = F.leaky_relu(self.fc1(torch.cat([action, state_code], dim=1)))
        h = F.leaky_relu(self.fc2(h))
        h = F.leaky_relu(self.fc3(h))
        return self.fc4(h)


class ImageEncoder(nn.Module):
    WIDTHS = (3, 64, 128, 256, 512, 1024)

    def __init__(self):
        super().__init__()
        for i in range(5):          # conv_i then conv_i_bn: the reference's registration (and RNG draw) order
            setattr(self, f"conv{i + 1}", nn.Conv2d(self.WIDTHS[i], self.WIDTHS[i + 1], 3, 2, 1))
            setattr(self, f"conv{i + 1}_bn", nn.BatchNorm2d(self.WIDTHS[i + 1]))
        self.conv6 = nn.Conv2d(1024, 128, 4, 1, 0)

    def forward(self, x):
        for i in (1, 2, 3):
            x = F.relu(getattr(self, f"conv{i}_bn")(getattr(self, f"conv{i}")(x)))
        x = F.relu(self.conv5(F.relu(self.conv4(x))))
        return self.conv6(x)
