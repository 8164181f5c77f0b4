"""Import alias: ``import dmlp_b200`` loads the package that lives in
``distributed-machine-learning-pipeline_b200/`` (a directory name Python cannot import directly).

The alias replaces itself in ``sys.modules`` with the real package, so
``dmlp_b200.models``, ``dmlp_b200.ops`` ... resolve to the hyphenated directory's sub-packages.
"""
import importlib.util as _ilu
import os as _os
import sys as _sys

_real_dir = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                          "distributed-machine-learning-pipeline_b200")
_spec = _ilu.spec_from_file_location("dmlp_b200", _os.path.join(_real_dir, "__init__.py"),
                                     submodule_search_locations=[_real_dir])
_mod = _ilu.module_from_spec(_spec)
_sys.modules["dmlp_b200"] = _mod
_spec.loader.exec_module(_mod)
