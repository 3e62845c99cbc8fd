"""Import shim: the package directory is named after the reference repository
(`fast-genomic-data-processing_b200/`), which is not a valid Python identifier, so this module loads it
under the name `fgdp_b200`."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "fast-genomic-data-processing_b200")
_spec = importlib.util.spec_from_file_location("fgdp_b200", os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["fgdp_b200"] = _mod
_spec.loader.exec_module(_mod)
