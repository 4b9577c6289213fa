"""Import the UNMODIFIED reference (``/root/reference/pinot``) in the build container.  ORACLE tooling.

The reference needs ``dgl``, ``rdkit`` and ``matplotlib``, none of which is installed.  ``dgl`` is
satisfied by the CPU restatement ``oracle.dgl_ref``; the other two are empty stubs (they are only
touched by featurisation / plotting code that the hot path never reaches).  Recipe verified in
SURVEY.md appendix B.  /root/reference does not exist on the GPU box: only
``tests/golden/make_golden.py`` and ``-m "not gpu"`` tests (which skip when it is absent) use this.
"""
import importlib.machinery
import os
import sys
import types
import warnings

REFERENCE_ROOT = os.environ.get("PINOT_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "pinot"))


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__spec__ = importlib.machinery.ModuleSpec(name, None)
    m.__path__ = []
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def load():
    """Returns the reference ``pinot`` package (imported once)."""
    if "pinot" in sys.modules and getattr(sys.modules["pinot"], "__file__", "").startswith(REFERENCE_ROOT):
        return sys.modules["pinot"]
    if not available():
        raise RuntimeError("reference tree not found at %s" % REFERENCE_ROOT)
    from . import dgl_ref
    dgl_ref.install("dgl")
    if "rdkit" not in sys.modules:
        _stub("rdkit", Chem=object())
        _stub("rdkit.Chem")
    if "matplotlib" not in sys.modules:
        mpl = _stub("matplotlib")
        mpl.pyplot = _stub("matplotlib.pyplot")
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        import pinot  # noqa: E402
    return pinot
