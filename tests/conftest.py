"""pytest configuration: registers the ``gpu`` marker and makes the repo root importable.

``python -m pytest tests -m "not gpu"`` runs everywhere (CPU, gloo, sim backend);
``python -m pytest tests -m gpu`` needs a B200 and the in-tree native library.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (B200, sm_100a) and the native library")
    config.addinivalue_line("markers", "multigpu: test needs at least two CUDA devices")
