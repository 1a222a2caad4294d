"""Import harness for the UNMODIFIED reference (TEST INFRASTRUCTURE ONLY).

Puts inert `matplotlib` / `matlab.engine` stand-ins (oracle/stubs) and the
read-only reference checkout on `sys.path`, so `import modulation`, `import
mp_model`, ... resolve to the reference's own modules.  Used only

  * by `tests/golden/make_golden.py` to generate the committed golden vectors, and
  * by `tests/test_oracle_vs_reference.py` (skipped when the checkout is absent,
    e.g. on the GPU box) to pin `oracle/link_oracle.py` against the reference.

Nothing in the product package imports this file.
"""
import os
import sys

REFERENCE_DIR = os.environ.get("MMIMO_REFERENCE_DIR", "/root/reference")
_STUBS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "stubs")

_REF_MODULES = ("modulation", "transceiver", "distortion", "antenna_array", "channel",
                "noise", "corrector", "mp_model", "utilities", "speedup", "plot_settings")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_DIR, "mp_model.py"))


def import_reference():
    """Return a dict name -> reference module.  Raises if the checkout is absent."""
    if not reference_available():
        raise RuntimeError("reference checkout not found at %s" % REFERENCE_DIR)
    import numpy as np
    sys.dont_write_bytecode = True  # the checkout is read-only
    if not hasattr(np, "complex"):
        np.complex = complex  # channel.py:328 uses the alias removed in numpy 1.24
    for mod in list(sys.modules):
        if mod in _REF_MODULES or mod == "matplotlib" or mod.startswith("matplotlib."):
            # a product drop-in module of the same name may be loaded; the reference
            # must win inside this harness
            del sys.modules[mod]
    saved = list(sys.path)
    sys.path.insert(0, REFERENCE_DIR)
    sys.path.insert(0, _STUBS)
    try:
        import importlib
        mods = {name: importlib.import_module(name) for name in _REF_MODULES}
    finally:
        sys.path[:] = saved
    for name, m in mods.items():
        if name in ("speedup",):
            continue
        assert os.path.dirname(os.path.abspath(m.__file__)) == os.path.abspath(REFERENCE_DIR), (name, m.__file__)
    # detach the names again so that later `import modulation` can resolve to something else
    for name in _REF_MODULES:
        sys.modules.pop(name, None)
    return mods
