"""ndivplanning_b200 — B200-native (sm_100a) planning hot path of goodmattg/ndivplanning.

    from ndivplanning_b200 import FusedPlanner, diversity, fetch_push_control_evaluation

Everything computes through libndivplan.so (hand-written CUDA behind a C ABI,
include/ndivplan.h); there is no CPU fallback.
"""
from . import diversity  # noqa: F401
from ._lib import LIB_PATH, NdpError, launch_count  # noqa: F401
from .encoder import FusedEncoder  # noqa: F401
from .gan_train import GanTrainer, train_gan_epoch  # noqa: F401
from .models import ActionDiscriminator, GeneratorMLP, ImageEncoder, LatentForward  # noqa: F401
from .mpc_eval import fetch_push_control_evaluation, sample_noise, sample_noise_device  # noqa: F401
from .planner import FusedPlanner, Plan, PlanRing, select_plan, unpack_index  # noqa: F401
from .trajectory_loader import PushDataset  # noqa: F401

__all__ = ["diversity", "FusedPlanner", "Plan", "PlanRing", "select_plan", "unpack_index", "GeneratorMLP", "LatentForward", "ActionDiscriminator", "GanTrainer", "train_gan_epoch", "FusedEncoder", "ImageEncoder",
           "fetch_push_control_evaluation", "sample_noise", "sample_noise_device", "NdpError", "launch_count", "LIB_PATH", "PushDataset"]
