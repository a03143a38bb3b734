"""Importable alias of the package directory (whose mandated name contains hyphens):
`import b200cnn` gives `distributed-machine-learning-using-cnn-cifar-10-dataset-_b200`."""
import importlib
import os
import sys

_ROOT = os.path.dirname(os.path.abspath(__file__))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
_pkg = importlib.import_module("distributed-machine-learning-using-cnn-cifar-10-dataset-_b200")
sys.modules[__name__] = _pkg
