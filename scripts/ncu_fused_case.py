"""Smallest program that runs the cluster-fused step kernel (8 steps per launch) a few times: the ncu target."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dmlp_b200.parallel.trainer import Trainer, TrainerConfig  # noqa: E402
from dmlp_b200.utils.data import synthetic_mnist  # noqa: E402

t = Trainer(TrainerConfig(dims=(784, 128, 64, 10), batch=64, lr=0.01), device=0)
x, y = synthetic_mnist(64 * 64, seed=0)
st = t.fit(x, y, 64)
print("ok", st["loss"], t.launches_last_run)
t.close()
