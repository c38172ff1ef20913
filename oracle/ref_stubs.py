"""Import shims that let the UNMODIFIED reference be imported from
/root/reference in the build container.

The reference's hot-path modules pull in dead imports (matplotlib, imageio,
spectral_normalization, h5py, visdom, plotly, imp, dotmap).  We register empty
stand-ins so `import diversity, mpc_eval; from models.gan import Decoder` works.
Used only by oracle/make_golden.py and tests that run when /root/reference exists.
"""
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("NDIV_REFERENCE_ROOT", "/root/reference")


class DotMap(dict):
    """Attribute-access dict (stand-in for dotmap.DotMap; mpc_eval.py:24)."""

    def __init__(self, *a, **k):
        super().__init__(*a, **k)
        for key, v in list(self.items()):
            if isinstance(v, dict) and not isinstance(v, DotMap):
                self[key] = DotMap(v)

    def __getattr__(self, name):
        try:
            return self[name]
        except KeyError as e:  # pragma: no cover
            raise AttributeError(name) from e

    __setattr__ = dict.__setitem__


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "diversity.py"))


def install_stubs() -> None:
    def mod(name, **attrs):
        if name in sys.modules:
            return sys.modules[name]
        m = types.ModuleType(name)
        for k, v in attrs.items():
            setattr(m, k, v)
        sys.modules[name] = m
        return m

    class _Anything:
        def __init__(self, *a, **k):
            pass

        def __getattr__(self, name):
            return lambda *a, **k: None

    try:
        import matplotlib  # noqa: F401
        import matplotlib.pyplot  # noqa: F401
    except Exception:
        mpl = mod("matplotlib", use=lambda *a, **k: None)
        plt = mod("matplotlib.pyplot", switch_backend=lambda *a, **k: None)
        mpl.pyplot = plt
    mod("imageio")
    mod("spectral_normalization", SpectralNorm=lambda m: m)
    mod("h5py", File=_Anything)
    mod("visdom", Visdom=_Anything)
    pl = mod("plotly")
    pl.tools = mod("plotly.tools", mpl_to_plotly=lambda *a, **k: None)
    mod("imp")
    mod("dotmap", DotMap=DotMap)
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)


def import_reference():
    """Returns (diversity, mpc_eval, gan.Decoder) from the unmodified reference."""
    if not reference_available():
        raise RuntimeError(f"reference tree not found at {REFERENCE_ROOT}")
    install_stubs()
    import diversity  # type: ignore
    import mpc_eval  # type: ignore
    from models.gan import Decoder  # type: ignore

    return diversity, mpc_eval, Decoder
