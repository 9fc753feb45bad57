"""Loader for the package directory ``erlangnetwork-gnsslib-sdr_b200`` (the
hyphens make it un-importable by name); ``import sdrb200_pkg as sdr`` gives the
package module."""
import importlib.util
import os
import sys

_NAME = "erlangnetwork_gnsslib_sdr_b200"
_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "erlangnetwork-gnsslib-sdr_b200")


def load():
    if _NAME in sys.modules:
        return sys.modules[_NAME]
    spec = importlib.util.spec_from_file_location(
        _NAME, os.path.join(_DIR, "__init__.py"), submodule_search_locations=[_DIR])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[_NAME] = mod
    try:
        spec.loader.exec_module(mod)
    except BaseException:
        del sys.modules[_NAME]
        raise
    return mod
