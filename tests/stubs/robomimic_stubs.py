"""Stand-in modules for robomimic's NON-numerical dependencies that are absent from this image: with
them ``import robomimic.models.policy_nets`` works, so the real ICRT network can be built around the tokenizer seam
(test infrastructure only; nothing here computes anything)."""
import sys
import types


class _Any:
    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Any()

    def __getattr__(self, n):
        return _Any()


def _getattr(name):
    if name.startswith("__"):
        raise AttributeError(name)
    return _Any


STUBBED = ["clip", "h5py", "imageio", "tensorboardX", "matplotlib", "matplotlib.pyplot", "matplotlib.cm",
           "diffusers", "diffusers.optimization", "diffusers.training_utils", "diffusers.schedulers",
           "diffusers.schedulers.scheduling_ddpm", "diffusers.schedulers.scheduling_ddim",
           "tianshou", "tianshou.env", "egl_probe", "mujoco", "robosuite"]


def install(reference_root="/root/reference"):
    """Put the stubs and the reference tree on the import path (idempotent)."""
    if "termcolor" not in sys.modules:
        m = types.ModuleType("termcolor")
        m.colored = lambda s, *a, **k: s
        m.cprint = lambda *a, **k: None
        sys.modules["termcolor"] = m
    for n in STUBBED:
        if n not in sys.modules:
            m = types.ModuleType(n)
            m.__getattr__ = _getattr
            m.__path__ = []
            sys.modules[n] = m
    if reference_root not in sys.path:
        sys.path.insert(0, reference_root)
