"""Import the UNMODIFIED reference modules (models, kan, utils, mdct, window) from /root/reference.

TEST INFRASTRUCTURE ONLY.  Works only where /root/reference is mounted (the build container), never on
the GPU box.  Used by tests/golden/make_golden.py to manufacture golden vectors, and by the optional
`-m "not gpu"` cross-checks that skip when the mount is absent.

The reference imports plotting / IO packages that are not installed here:
matplotlib (window.py:8, utils.py:9,13), librosa (utils.py:19), torchsummary (utils.py:23),
rff (models.py:12).  None of them is touched on the sine / KAN / MSE path, so empty stub modules
are injected into sys.modules before the import.
"""
import os
import sys
import types

REF_ROOT = os.environ.get("INR_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "models.py"))


def _stub(name, **attrs):
    if name in sys.modules:
        return sys.modules[name]
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def load():
    """Returns a namespace with the reference modules: .models .kan .utils .mdct .window"""
    if not available():
        raise RuntimeError("reference tree not mounted at %s" % REF_ROOT)
    mpl = _stub("matplotlib")
    plt = _stub("matplotlib.pyplot")
    mpl.pyplot = plt
    _stub("librosa")
    _stub("torchsummary", summary=lambda *a, **k: None)
    rff = _stub("rff")
    rff.layers = _stub("rff.layers")
    try:
        import torchaudio  # noqa: F401
    except Exception:  # pragma: no cover
        _stub("torchaudio")
    saved = list(sys.path)
    # the reference's flat module names collide with our drop-in shims: import under a clean path
    shadow = {k: sys.modules.pop(k) for k in ("models", "kan", "utils", "mdct", "window") if k in sys.modules}
    sys.path.insert(0, REF_ROOT)
    try:
        import importlib
        ns = types.SimpleNamespace()
        for name in ("window", "mdct", "utils", "models", "kan"):
            mod = importlib.import_module(name)
            assert os.path.dirname(os.path.abspath(mod.__file__)) == os.path.abspath(REF_ROOT), mod.__file__
            setattr(ns, name, mod)
    finally:
        sys.path[:] = saved
        for name in ("window", "mdct", "utils", "models", "kan"):
            sys.modules.pop(name, None)
        sys.modules.update(shadow)
    return ns
