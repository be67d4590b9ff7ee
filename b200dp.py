"""Import alias: ``import b200dp`` loads the package that lives in the directory
``neural-networks-parallel-training-with-mpi_b200/`` (a name Python cannot import directly
because of the hyphens).  After this module runs, ``sys.modules['b200dp']`` is the real
package, so ``b200dp.models``, ``b200dp.ops`` ... resolve as ordinary sub-packages.
"""
import importlib.util
import os
import sys

PACKAGE_DIRNAME = "neural-networks-parallel-training-with-mpi_b200"
_pkg_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), PACKAGE_DIRNAME)
_spec = importlib.util.spec_from_file_location(
    "b200dp", os.path.join(_pkg_dir, "__init__.py"), submodule_search_locations=[_pkg_dir]
)
_mod = importlib.util.module_from_spec(_spec)
sys.modules["b200dp"] = _mod
_spec.loader.exec_module(_mod)
