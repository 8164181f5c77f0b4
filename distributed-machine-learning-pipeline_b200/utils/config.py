"""One configuration object + flags/env for everything the reference hard-codes.

The reference has no flag, env var or config file: hyper-parameters, addresses, ports, dataset
paths, health interval and timeouts are compile-time constants (``DSML/client/client.go:22-33,518,
527,545-548``, ``cmd/*/main.go``, ``gpu_coordinator_server.go:57,103,136,144,498``).
``Config`` carries the same defaults; every field can be overridden by ``DMLP_<FIELD>`` in the
environment or ``--<field>`` on the command line (``add_arguments`` / ``from_args``).
"""
from __future__ import annotations

import argparse
import dataclasses
import os
from dataclasses import dataclass
from typing import Tuple


@dataclass
class Config:
    # model / optimisation (client.go:22-33)
    dims: Tuple[int, ...] = (784, 128, 64, 10)
    batch: int = 64
    epochs: int = 10
    lr: float = 0.01
    schedule: str = "constant"          # constant | step | cosine | warmup-cosine
    precision: str = "tf32"             # tf32 | bf16
    seed: int = 0
    # data (client.go:545-548)
    data_dir: str = "/root/reference/DSML/data"
    synthetic: bool = True
    train_samples: int = 60000
    test_samples: int = 10000
    # collectives
    allreduce: str = "auto"             # auto | oneshot | twoshot | ring
    overlap: bool = False
    watchdog_ms: int = 10000
    allreduce_table: str = ""           # rows of the auto dispatcher: "world:oneshot_max:nvls_min,..." (K/M/G, never)
    # services (cmd/*/main.go, gpu_coordinator_server.go:57,103)
    backend: str = "cuda"               # cuda | sim
    coordinator_addr: str = "127.0.0.1:50051"
    device_ports: Tuple[int, ...] = (5003, 5004, 5005)
    device_ids: Tuple[int, ...] = (1, 2, 3)
    mem_size: int = 0x3000
    health_interval: float = 5.0
    probe_timeout: float = 2.0
    # fault injection: "" | "dead-rank:<r>@step:<s>"
    fault: str = ""
    metrics_jsonl: str = ""
    weights_out: str = ""
    weights_in: str = ""

    @staticmethod
    def _convert(field: dataclasses.Field, text: str):
        if field.type in ("bool", bool):
            return text.lower() in ("1", "true", "yes", "on")
        if field.type in ("int", int):
            return int(text, 0)
        if field.type in ("float", float):
            return float(text)
        if "Tuple" in str(field.type):
            return tuple(int(t, 0) for t in text.replace("-", ",").split(",") if t)
        return text

    @classmethod
    def from_env(cls, **overrides) -> "Config":
        cfg = cls(**overrides)
        for f in dataclasses.fields(cls):
            v = os.environ.get("DMLP_" + f.name.upper())
            if v is not None:
                setattr(cfg, f.name, cls._convert(f, v))
        return cfg

    @classmethod
    def add_arguments(cls, ap: argparse.ArgumentParser) -> None:
        for f in dataclasses.fields(cls):
            ap.add_argument("--" + f.name.replace("_", "-"), default=None, help=f"default {f.default!r}")

    @classmethod
    def from_args(cls, ns: argparse.Namespace) -> "Config":
        cfg = cls.from_env()
        for f in dataclasses.fields(cls):
            v = getattr(ns, f.name, None)
            if v is not None:
                setattr(cfg, f.name, cls._convert(f, v))
        return cfg
