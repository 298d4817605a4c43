"""Import shim: exposes the package in ``self-supervise-sfm_b200/`` (a directory name Python cannot import) as
``sailrecon_b200``.  ``import sailrecon_b200`` from the repository root (or with it on sys.path)."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "self-supervise-sfm_b200")
_spec = importlib.util.spec_from_file_location(__name__, os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules[__name__] = _mod
_spec.loader.exec_module(_mod)
