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


def test_bench_reference_arm_reports_unavailable():
    """`bench.py --impl reference`: the reference is a Go module that cannot be built here; one JSON line, exit 0."""
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--gpus", "1", "--steps", "5",
                        "--warmup", "3"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and "unavailable" in line and len(line["unavailable"]) > 10


def test_graft_entry_build_is_importable_and_links():
    """__graft_entry__.build() is the driver's "does it build" check: here we only require that the module imports
    and that the in-tree shared library (when already built) exposes the entry points the bindings declare."""
    import importlib
    import os
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    ge = importlib.import_module("__graft_entry__")
    assert callable(ge.build) and callable(ge.smoke)
    from dmlp_b200.runtime import native
    if not native.available():
        import pytest
        pytest.skip("native library not built yet")
    lib = native.lib()
    for sym in ("dmlp_gemm", "dmlp_allreduce", "dmlp_allreduce_nvls", "dmlp_mlp_fused_steps", "dmlp_trainer_run",
                "dmlp_allreduce_sum_f32_noalloc", "dmlp_comm_is_emulated", "dmlp_gemm_prof_read",
                "dmlp_allreduce_noalloc_status", "dmlp_mlp_fused_steps_gx", "dmlp_mlp_fused_set_fault", "dmlp_input_exchange",
                "dmlp_trainer_set_fault", "dmlp_trainer_state", "dmlp_allreduce_pick_algo_ex", "dmlp_trainer_pipelined_layers", "dmlp_allreduce_set_table",
                "dmlp_wgrad_rs_prof_read"):
        assert hasattr(lib, sym), sym


def test_allreduce_dispatcher_table():
    """`algo=auto` (csrc/collectives.cu, table from the sweeps under profiles/): one-shot for latency-bound sizes, two-shot
    above, NVLS only when the caller says the buffers are multicast-bound and only from the size at which it wins; the ring
    (the RPC's namesake, gpu_coordinator_server.go:272-376) is the comparison mode and never picked."""
    from dmlp_b200.runtime import native
    if not native.available():
        import pytest
        pytest.skip("native library not built yet")
    pick = native.lib().dmlp_allreduce_pick_algo_ex
    one, two, ring, nvls = (native.ALGO_BY_NAME[k] for k in ("oneshot", "twoshot", "ring", "nvls"))
    assert pick(1, 1 << 20, 1) == one
    for world in (2, 3, 4, 8):
        assert pick(world, 1024, 1) == one
        assert pick(world, 437544, 0) == one          # the small model's gradient
        assert pick(world, 1 << 30, 0) == two
        picks = {pick(world, 1 << s, n) for s in range(10, 31) for n in (0, 1)}
        assert ring not in picks
    assert pick(2, 4 << 20, 0) == one and pick(2, 16 << 20, 0) == two
    assert pick(8, 512 << 10, 0) == one and pick(8, 1 << 20, 0) == two
    assert pick(8, 16 << 20, 1) == two and pick(8, 64 << 20, 1) == nvls and pick(8, 64 << 20, 0) == two
    assert pick(4, 64 << 20, 1) == nvls and pick(2, 1 << 30, 1) == two  # (no multicast win on two ranks)


def test_allreduce_dispatcher_table_can_be_overridden():
    """Thresholds are data, not code: rows can be replaced per process (`set_allreduce_table`, Config.allreduce_table,
    DMLP_ALLREDUCE_TABLE) and the built-in table comes back with an empty spec."""
    import pytest
    from dmlp_b200.runtime import native
    if not native.available():
        pytest.skip("native library not built yet")
    from dmlp_b200.runtime.comm import pick_algo, set_allreduce_table
    one, two, nvls = (native.ALGO_BY_NAME[k] for k in ("oneshot", "twoshot", "nvls"))
    try:
        set_allreduce_table("8:256K:32M,2:1M:never,5:64K:1G")
        assert pick_algo(8, 256 << 10) == one and pick_algo(8, (256 << 10) + 4) == two
        assert pick_algo(8, 32 << 20, True) == nvls and pick_algo(8, (32 << 20) - 4, True) == two
        assert pick_algo(2, 1 << 20) == one and pick_algo(2, 2 << 20) == two and pick_algo(2, 1 << 30, True) == two
        assert pick_algo(5, 64 << 10) == one and pick_algo(5, 128 << 10) == two  # a new row between 4 and 6
        assert pick_algo(4, 1 << 20) == one  # untouched rows keep the built-in values
        assert pick_algo(7, 512 << 10) == one and pick_algo(7, 1 << 20) == two  # 7 ranks read the row of 6
        for bad in ("8:256K", "x:1:2", "8:1Q:2", "1:1K:1K", "9:1K:1K", "8:1K:1K,,"):
            with pytest.raises(native.NativeError):
                set_allreduce_table(bad)
        assert pick_algo(8, 256 << 10) == one and pick_algo(8, (256 << 10) + 4) == two  # a bad spec changes nothing
    finally:
        set_allreduce_table(None)
    assert pick_algo(8, 512 << 10) == one and pick_algo(8, 1 << 20) == two and pick_algo(8, 64 << 20, True) == nvls


def test_tsan_run_of_the_host_runtime():
    """ThreadSanitizer over the native runtime's shared host state (dispatcher table, per-thread error slot): opt-in
    (DMLP_RUN_TSAN=1, ~2 min of nvcc), the recorded run is profiles/r2_tsan_host.txt."""
    import os
    import shutil
    import subprocess
    import pytest
    if os.environ.get("DMLP_RUN_TSAN", "0") in ("", "0"):
        pytest.skip("set DMLP_RUN_TSAN=1 to rebuild the host runtime with -fsanitize=thread and run the driver")
    if not (shutil.which("nvcc") or os.path.exists("/usr/local/cuda/bin/nvcc")):
        pytest.skip("no nvcc")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run(["bash", os.path.join(root, "scripts", "tsan_host_runtime.sh")], capture_output=True, text=True, timeout=1200)
    out = r.stdout + r.stderr
    assert r.returncode == 0 and "0 inconsistencies, table restored: yes" in out, out[-3000:]
    assert "WARNING: ThreadSanitizer" not in out, out[-3000:]


def test_pip_install_into_a_target_directory(tmp_path):
    """setup.py: the hyphenated source directory installs as the import package `dmlp_b200` (sub-packages, console
    scripts, the in-tree native library as package data). Python side only here (DMLP_SKIP_NATIVE_BUILD=1)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    target = os.path.join(root, "build", "pipcheck")
    env = dict(os.environ, DMLP_SKIP_NATIVE_BUILD="1")
    r = subprocess.run([sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--no-deps", "--upgrade",
                        "--target", target, root], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    code = ("import dmlp_b200, dmlp_b200.services.coordinator, dmlp_b200.proto.gpu_sim, dmlp_b200.parallel.schedule;"
            "print(dmlp_b200.__file__)")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=str(tmp_path),
                       env=dict(os.environ, PYTHONPATH=target), timeout=300)
    assert r.returncode == 0 and os.path.join("pipcheck", "dmlp_b200") in r.stdout, r.stdout + r.stderr
    assert os.path.exists(os.path.join(target, "bin", "dmlp-train"))
