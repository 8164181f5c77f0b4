"""`pip install .` for dmlp_b200 (optional: the repo runs in place with `python build_native.py`).

The import package `dmlp_b200` lives in `distributed-machine-learning-pipeline_b200/`; the native library is built by
build_native.py (nvcc, sm_100a; no GPU needed to build) into that package's `_native/` and installed as package data.
`DMLP_SKIP_NATIVE_BUILD=1` installs the Python side only (services on the `sim` backend, proto, client).
"""
import os
import subprocess
import sys

from setuptools import setup
from setuptools.command.build_py import build_py

ROOT = os.path.dirname(os.path.abspath(__file__))
SRC = "distributed-machine-learning-pipeline_b200"


def _packages():
    out = []
    for d, _, files in os.walk(os.path.join(ROOT, SRC)):
        if "__init__.py" in files:
            rel = os.path.relpath(d, os.path.join(ROOT, SRC))
            out.append("dmlp_b200" if rel == "." else "dmlp_b200." + rel.replace(os.sep, "."))
    return sorted(out)


class BuildWithNative(build_py):
    def run(self):
        if os.environ.get("DMLP_SKIP_NATIVE_BUILD", "0") in ("", "0"):
            subprocess.check_call([sys.executable, os.path.join(ROOT, "build_native.py")], cwd=ROOT)
        super().run()


setup(
    name="dmlp_b200",
    version="0.1.0",
    description="B200-native (sm_100a) collective runtime, gRPC GPU services and data-parallel MLP trainer",
    packages=_packages(),
    package_dir={"dmlp_b200": SRC},
    package_data={"dmlp_b200": ["_native/*.so"]},
    python_requires=">=3.10",
    install_requires=["numpy", "grpcio", "protobuf"],
    extras_require={"gpu": ["torch"]},
    cmdclass={"build_py": BuildWithNative},
    entry_points={"console_scripts": [
        "dmlp-train=dmlp_b200.train:main",
        "dmlp-serve-devices=dmlp_b200.serve_devices:main",
        "dmlp-serve-coordinator=dmlp_b200.serve_coordinator:main",
        "dmlp-grpc-client=dmlp_b200.grpc_client:main",
    ]},
)
