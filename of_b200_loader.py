"""Imports the package directory ``optical-flow_b200/`` (hyphenated, so not importable by
name) as module ``optical_flow_b200``.  Usage: ``from of_b200_loader import ofb``."""
import importlib.util
import os
import sys

_ROOT = os.path.dirname(os.path.abspath(__file__))
_PKG = os.path.join(_ROOT, "optical-flow_b200")
_NAME = "optical_flow_b200"

if _NAME in sys.modules:
    ofb = sys.modules[_NAME]
else:
    _spec = importlib.util.spec_from_file_location(
        _NAME, os.path.join(_PKG, "__init__.py"), submodule_search_locations=[_PKG])
    ofb = importlib.util.module_from_spec(_spec)
    sys.modules[_NAME] = ofb
    _spec.loader.exec_module(ofb)
