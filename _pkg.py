"""Loads the package directory `ytk-learn_b200/` (hyphenated, so not importable by name)
as module `ytk_learn_b200`."""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
NAME = "ytk_learn_b200"


def load():
    if NAME in sys.modules:
        return sys.modules[NAME]
    path = os.path.join(ROOT, "ytk-learn_b200", "__init__.py")
    spec = importlib.util.spec_from_file_location(NAME, path, submodule_search_locations=[os.path.dirname(path)])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[NAME] = mod
    spec.loader.exec_module(mod)
    return mod
