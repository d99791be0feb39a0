"""Import the reference's own modules from /root/reference with stand-ins for its missing dependencies.

TEST INFRASTRUCTURE ONLY, and only usable in the build container: /root/reference does not exist on
the GPU box, so nothing marked ``gpu`` may call this.  Used by ``make_goldens.py`` and by the CPU
tests that compare the restatements with the reference while it is available.
"""
from __future__ import annotations

import importlib
import os
import sys

REFERENCE_DIR = os.environ.get("QSB200_REFERENCE_DIR", "/root/reference")
_HERE = os.path.dirname(os.path.abspath(__file__))


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_DIR, "dipolar_ensemble_with_rare.py"))


def _prepend(path: str) -> None:
    if path not in sys.path:
        sys.path.insert(0, path)


def load_reference_model(backend: str = "exact"):
    """The reference ``dipolar_ensemble_with_rare`` module, executed unmodified on top of fake_qutip."""
    if not reference_available():
        raise FileNotFoundError(REFERENCE_DIR)
    _prepend(os.path.dirname(_HERE))                 # so that `import oracle` works inside the fake
    _prepend(os.path.join(_HERE, "fake_qutip"))
    import qutip  # the fake

    qutip.SESOLVE_BACKEND = backend
    name = "dipolar_ensemble_with_rare"
    mod = sys.modules.get(name)
    if mod is not None and not getattr(mod, "__file__", "").startswith(REFERENCE_DIR):
        del sys.modules[name]                        # a drop-in with the same name was imported earlier
        mod = None
    if mod is None:
        _prepend(REFERENCE_DIR)
        mod = importlib.import_module(name)
        sys.path.remove(REFERENCE_DIR)
    return mod


def load_reference_sweep():
    """The reference ``sweep_sea_detuning`` module on top of inert matplotlib stubs (plots become no-ops)."""
    load_reference_model()
    _prepend(os.path.join(_HERE, "fake_mpl"))
    name = "sweep_sea_detuning"
    if name in sys.modules:
        return sys.modules[name]
    _prepend(REFERENCE_DIR)
    try:
        return importlib.import_module(name)
    finally:
        sys.path.remove(REFERENCE_DIR)
