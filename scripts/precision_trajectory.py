"""tf32 tensor-core training against the fp64 oracle over a long run (the decision record for a 3xTF32 mode).

The reference computes in fp32 scalars (DSML/client/client.go:112-202); this framework's small-model kernels feed
tf32 operands (10-bit mantissa, fp32 accumulate and storage) to tcgen05. SURVEY section 5.6 lists a `3xtf32` precision
mode as the fp32-parity option. This script measures what such a mode would buy: the same 10 epochs (156 batches of 64,
SGD lr 0.01, the reference's schedule) run (a) by the native trainer and (b) by the NumPy fp64 restatement of the
reference math, from the same initial weights on the same batches. Reported per epoch: loss / accuracy of both and the
largest absolute weight difference; at the end the accuracy of both models on a held-out set.

    python scripts/precision_trajectory.py [--out profiles/r2_precision_trajectory.json]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=10)
    ap.add_argument("--samples", type=int, default=10000)
    ap.add_argument("--lr", type=float, default=0.01)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    from dmlp_b200.models import oracle
    from dmlp_b200.parallel.trainer import Trainer, TrainerConfig
    from dmlp_b200.utils.data import synthetic_mnist
    dims, B = (784, 128, 64, 10), 64
    x, y = synthetic_mnist(a.samples, seed=21)
    xt, yt = synthetic_mnist(5000, seed=22)
    spe = a.samples // B
    t = Trainer(TrainerConfig(dims=dims, batch=B, lr=a.lr, seed=3), device=0)
    ref = [(w.astype(np.float64), b.astype(np.float64)) for w, b in t.weights()]
    rows = []
    for ep in range(1, a.epochs + 1):
        st = t.fit(x, y, spe)
        losses, hits = [], 0
        for s in range(spe):
            b0 = ((ep - 1) * spe + s) % spe
            xb, yb = x[b0 * B:(b0 + 1) * B].astype(np.float64), y[b0 * B:(b0 + 1) * B]
            ref, l, c = oracle.dp_train_step(ref, [(xb, yb)], a.lr)
            losses.append(l)
            hits += c
        wdiff = max(max(np.abs(w - rw).max(), np.abs(b - rb).max()) for (w, b), (rw, rb) in zip(t.weights(), ref))
        wmax = max(np.abs(rw).max() for rw, _ in ref)
        rows.append({"epoch": ep, "tf32_loss": round(st["loss"], 6), "fp64_loss": round(float(np.mean(losses)), 6),
                     "tf32_acc": round(st["accuracy"], 5), "fp64_acc": round(hits / (spe * B), 5),
                     "max_abs_weight_diff": float(wdiff), "max_abs_weight": float(wmax)})
        print(json.dumps(rows[-1]), flush=True)
    ev = t.evaluate(xt, yt)
    out = {"config": "MLP 784-128-64-10, batch 64, SGD lr %g, %d epochs x %d batches, synthetic MNIST-shaped data, 1 x B200" %
                     (a.lr, a.epochs, spe), "epochs": rows, "tf32_heldout_accuracy": ev["accuracy"]}
    try:
        l, c, _ = oracle.loss_and_grads(ref, xt[:5000 // B * B].astype(np.float64), yt[:5000 // B * B])
        out["fp64_heldout_accuracy"] = c / (5000 // B * B)
    except Exception as e:  # noqa: BLE001
        out["fp64_heldout_accuracy"] = str(e)
    print(json.dumps({k: v for k, v in out.items() if k != "epochs"}))
    if a.out:
        with open(a.out, "w") as f:
            json.dump(out, f, indent=1)
    t.close()


if __name__ == "__main__":
    main()
