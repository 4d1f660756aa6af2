"""Import shim: ``import distributed_pipeline_b200`` loads the package that
lives in the (hyphenated, hence not directly importable) directory
``distributed-pipeline_b200/`` next to this file."""
import importlib.util
import os
import sys

_pkg_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "distributed-pipeline_b200")
_spec = importlib.util.spec_from_file_location(
    __name__, os.path.join(_pkg_dir, "__init__.py"), submodule_search_locations=[_pkg_dir])
_module = importlib.util.module_from_spec(_spec)
sys.modules[__name__] = _module
_spec.loader.exec_module(_module)
