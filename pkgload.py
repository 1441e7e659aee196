"""Load the hyphenated package directory ``monodepth-estimation_b200/`` under the
importable name ``monodepth_estimation_b200``.

The directory name is fixed by the project layout and is not a valid Python
identifier, so bench.py, __graft_entry__.py and tests/conftest.py call
``load()`` once and then ``import monodepth_estimation_b200`` works everywhere.
"""
import importlib.util
import os
import sys

PKG_NAME = "monodepth_estimation_b200"
PKG_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "monodepth-estimation_b200")


def load():
    if PKG_NAME in sys.modules:
        return sys.modules[PKG_NAME]
    spec = importlib.util.spec_from_file_location(
        PKG_NAME, os.path.join(PKG_DIR, "__init__.py"), submodule_search_locations=[PKG_DIR]
    )
    mod = importlib.util.module_from_spec(spec)
    sys.modules[PKG_NAME] = mod
    spec.loader.exec_module(mod)
    return mod
