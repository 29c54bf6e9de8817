"""Import shim: the package directory is named ``dcegm.jl_b200`` (after the reference repo
DCEGM.jl), which is not a valid Python identifier.  ``import dcegm_jl_b200`` loads that
directory as a regular package under this name."""
import importlib.util
import os
import sys

_here = os.path.dirname(os.path.abspath(__file__))
_pkg_dir = os.path.join(_here, "dcegm.jl_b200")
_spec = importlib.util.spec_from_file_location(
    __name__, os.path.join(_pkg_dir, "__init__.py"), submodule_search_locations=[_pkg_dir]
)
_mod = importlib.util.module_from_spec(_spec)
sys.modules[__name__] = _mod
_spec.loader.exec_module(_mod)
