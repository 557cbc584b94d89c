"""Import alias.  The package directory is named after the reference repository
(``boinc-app-eah-brp_b200/``), which is not a valid Python identifier, so ``import brp_b200``
loads it from that directory under this shorter name."""
import importlib.util as _ilu
import pathlib as _pl
import sys as _sys

_pkg = _pl.Path(__file__).resolve().parent / "boinc-app-eah-brp_b200"
_spec = _ilu.spec_from_file_location("brp_b200", _pkg / "__init__.py", submodule_search_locations=[str(_pkg)])
_mod = _ilu.module_from_spec(_spec)
_sys.modules["brp_b200"] = _mod
_spec.loader.exec_module(_mod)
