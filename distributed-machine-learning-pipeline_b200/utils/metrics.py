"""Logging and metrics. Human-readable lines keep the reference's wording
(``DSML/client/client.go:557,566,578,652,655,501``) so outputs are comparable; in addition every
record can be appended to a JSONL file (samples/s, device step time, loss, accuracy), which the
reference has no equivalent of."""
from __future__ import annotations

import json
import logging
import time
from typing import Optional


def setup_logging(level: int = logging.INFO) -> None:
    logging.basicConfig(level=level, format="%(asctime)s %(message)s", datefmt="%Y/%m/%d %H:%M:%S")


class MetricsWriter:
    def __init__(self, path: Optional[str] = None):
        self.f = open(path, "a") if path else None
        self.t0 = time.time()

    def write(self, kind: str, **fields) -> None:
        if self.f is None:
            return
        rec = {"t": round(time.time() - self.t0, 4), "kind": kind}
        rec.update({k: (float(v) if hasattr(v, "__float__") and not isinstance(v, (int, bool, str)) else v)
                    for k, v in fields.items() if k != "per_step_loss"})
        self.f.write(json.dumps(rec) + "\n")
        self.f.flush()

    def close(self) -> None:
        if self.f:
            self.f.close()
            self.f = None
