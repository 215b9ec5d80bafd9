"""Import alias: the package directory is named `hipcomp-core_b200` (not a valid
identifier), so `import hipcomp_core_b200` loads it through importlib."""
import importlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
_pkg = importlib.import_module("hipcomp-core_b200")
sys.modules[__name__] = _pkg
