"""Importable alias: the package directory is named after the reference repo and so contains hyphens
(`distributed-lms-raft-llm_b200/`); `import lms_b200` gives the same module object."""
import importlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
_pkg = importlib.import_module("distributed-lms-raft-llm_b200")
sys.modules[__name__] = _pkg
