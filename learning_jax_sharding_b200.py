"""Import shim: the product package lives in the directory ``learning-jax-sharding_b200/`` (a name
Python cannot import directly); importing this module loads that package under the importable
name ``learning_jax_sharding_b200``."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "learning-jax-sharding_b200")
_spec = importlib.util.spec_from_file_location(__name__, os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules[__name__] = _mod
_spec.loader.exec_module(_mod)
