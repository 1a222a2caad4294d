"""Drop-in modules with the reference's own module and class names.

Put this directory at the front of `sys.path` (or call `install()`), and the reference's unchanged
`main_*` scripts resolve `import modulation`, `import antenna_array`, `import mp_model`, ... to these
thin wrappers over the sm_100a kernels of libmmimo_b200.so.  Numeric work is CUDA;
there is no CPU fallback - without the library or a GPU every numeric method raises."""
import os
import sys

DROPIN_DIR = os.path.dirname(os.path.abspath(__file__))
MODULES = ("utilities", "speedup", "plot_settings", "modulation", "distortion", "transceiver", "antenna_array",
           "channel", "noise", "corrector", "mp_model")


def install():
    """Make `import modulation` etc. resolve to the drop-in modules in this interpreter."""
    root = os.path.dirname(os.path.dirname(DROPIN_DIR))
    for p in (root, DROPIN_DIR):
        if p in sys.path:
            sys.path.remove(p)
    sys.path.insert(0, root)
    sys.path.insert(0, DROPIN_DIR)
    for name in MODULES:
        m = sys.modules.get(name)
        if m is not None and os.path.dirname(os.path.abspath(getattr(m, "__file__", ""))) != DROPIN_DIR:
            del sys.modules[name]
    return DROPIN_DIR
