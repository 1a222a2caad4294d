"""Dump the public API surface of the UNMODIFIED reference modules on the hot path (class names, method names and
parameter lists incl. defaults, module-level functions) to tests/golden/api_manifest.json.  Run in the build
container only (needs /root/reference).  tests/test_dropin_api.py checks the drop-in modules against it."""
import inspect
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.abspath(os.path.join(HERE, "..", "..")))

from oracle.refharness import import_reference  # noqa: E402

MODULES = ["modulation", "transceiver", "distortion", "antenna_array", "channel", "noise", "corrector", "mp_model",
           "utilities"]


def sig(fn):
    out = []
    for p in inspect.signature(fn).parameters.values():
        d = None if p.default is inspect.Parameter.empty else repr(p.default)
        out.append([p.name, d])
    return out


def main():
    ref = import_reference()
    man = {}
    for name in MODULES:
        mod = ref[name]
        entry = {"functions": {}, "classes": {}}
        for n, obj in vars(mod).items():
            if getattr(obj, "__module__", None) != name or n.startswith("__"):
                continue
            if inspect.isclass(obj):
                methods = {}
                for mn, m in vars(obj).items():
                    if inspect.isfunction(m) and (not mn.startswith("_") or mn in ("__init__", "__str__")):
                        methods[mn] = sig(m)
                props = sorted(mn for mn, m in vars(obj).items() if isinstance(m, property))
                entry["classes"][n] = {"bases": [b.__name__ for b in obj.__bases__], "methods": methods, "properties": props}
            elif inspect.isfunction(obj) or hasattr(obj, "py_func"):
                f = getattr(obj, "py_func", obj)
                entry["functions"][n] = sig(f)
        man[name] = entry
    with open(os.path.join(HERE, "api_manifest.json"), "w") as fh:
        json.dump(man, fh, indent=1, sort_keys=True)
    print({k: (len(v["classes"]), len(v["functions"])) for k, v in man.items()})


if __name__ == "__main__":
    main()
