"""Config / flag system (the reference has none: SURVEY §5.6) and metrics writer."""
import argparse
import json
import os

from dmlp_b200.utils.config import Config
from dmlp_b200.utils.metrics import MetricsWriter


def test_defaults_match_reference_constants():
    c = Config()
    assert c.batch == 64 and c.lr == 0.01 and c.epochs == 10            # client.go:23,27,28
    assert c.device_ports == (5003, 5004, 5005) and c.device_ids == (1, 2, 3) and c.mem_size == 0x3000
    assert c.coordinator_addr.endswith(":50051") and c.health_interval == 5.0 and c.probe_timeout == 2.0


def test_env_and_flags_override(monkeypatch):
    monkeypatch.setenv("DMLP_BATCH", "128")
    monkeypatch.setenv("DMLP_DIMS", "784,128,10")
    monkeypatch.setenv("DMLP_OVERLAP", "true")
    ap = argparse.ArgumentParser()
    Config.add_arguments(ap)
    c = Config.from_args(ap.parse_args(["--lr", "0.1", "--mem-size", "0x2000"]))
    assert c.batch == 128 and c.dims == (784, 128, 10) and c.overlap is True
    assert c.lr == 0.1 and c.mem_size == 0x2000


def test_metrics_jsonl(tmp_path):
    p = os.path.join(tmp_path, "m.jsonl")
    w = MetricsWriter(p)
    w.write("epoch", epoch=1, loss=0.5, accuracy=0.9)
    w.close()
    rec = json.loads(open(p).read().strip())
    assert rec["kind"] == "epoch" and rec["loss"] == 0.5
