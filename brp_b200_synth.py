"""The numpy-only helpers of the package (`synth`: synthetic work-units and banks, `sharding`: bank split and hit-list
exchange), loaded by path WITHOUT importing the package itself, i.e. without loading libbrp_b200.so.  Used where the CUDA
library must stay out of the process: the reference arm of bench.py and the golden-vector generators."""
import importlib.util as _ilu
import pathlib as _pl
import sys as _sys

_pkg = _pl.Path(__file__).resolve().parent / "boinc-app-eah-brp_b200"


def _load(name):
    key = f"brp_b200_standalone_{name}"
    if key in _sys.modules:
        return _sys.modules[key]
    spec = _ilu.spec_from_file_location(key, _pkg / f"{name}.py")
    mod = _ilu.module_from_spec(spec)
    _sys.modules[key] = mod
    spec.loader.exec_module(mod)
    return mod


synth = _load("synth")
sharding = _load("sharding")
