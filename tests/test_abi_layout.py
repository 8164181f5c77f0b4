"""The ctypes mirrors of the C ABI structs must match include/dmlp_b200.h byte for byte (CPU only).

A tiny C program that includes the public header is compiled with the host compiler and prints
sizeof / offsetof of the LAST field for every descriptor struct; the same numbers are computed
from the ctypes classes. A field appended on one side only (the usual way these drift apart)
changes at least one of them.
"""
import ctypes as C
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# (C type, last field in C, ctypes class path, last field in ctypes)
STRUCTS = [
    ("dmlp_gemm_desc_t", "ce_correct_dev", "dmlp_b200.runtime.native:GemmDesc", "ce_correct"),
    ("dmlp_device_info_t", "l2_bytes", "dmlp_b200.runtime.native:DeviceInfo", "l2_bytes"),
    ("dmlp_allreduce_desc_t", "prof_dev", "dmlp_b200.runtime.native:AllReduceDesc", "prof"),
    ("dmlp_nvls_allreduce_desc_t", "oneshot", "dmlp_b200.runtime.nvls:NvlsAllReduceDesc", "oneshot"),
    ("dmlp_mlp_fused_desc_t", "gx_enable", "dmlp_b200.runtime.native:MlpFusedDesc", "gx_enable"),
    ("dmlp_trainer_config_t", "fused", "dmlp_b200.parallel.trainer:_NativeCfg", "fused"),
]


def _c_layout(tmp_path):
    cc = shutil.which("gcc") or shutil.which("cc")
    if cc is None:
        pytest.skip("no host C compiler")
    lines = ["#include <stdio.h>", "#include <stddef.h>", '#include "dmlp_b200.h"', "int main(void) {"]
    for ctype, last, _, _ in STRUCTS:
        lines.append(f'  printf("{ctype} %zu %zu\\n", sizeof({ctype}), offsetof({ctype}, {last}));')
    lines += ["  return 0;", "}"]
    src = tmp_path / "abi.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "abi"
    subprocess.run([cc, "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src)], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout
    return {l.split()[0]: (int(l.split()[1]), int(l.split()[2])) for l in out.splitlines()}


def _load(path):
    mod, name = path.split(":")
    import importlib
    return getattr(importlib.import_module(mod), name)


def test_ctypes_mirrors_match_the_header(tmp_path):
    layout = _c_layout(tmp_path)
    for ctype, _, pypath, pylast in STRUCTS:
        cls = _load(pypath)
        size, off = layout[ctype]
        assert C.sizeof(cls) == size, f"{pypath}: sizeof {C.sizeof(cls)} != {ctype} {size}"
        assert getattr(cls, pylast).offset == off, f"{pypath}.{pylast}: offset {getattr(cls, pylast).offset} != {off}"
