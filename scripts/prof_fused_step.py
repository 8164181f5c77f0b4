"""Phase timing of the cluster-fused step kernel from in-kernel globaltimer stamps (1 GPU)."""
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dmlp_b200.models.layout import ParamLayout  # noqa: E402
from dmlp_b200.runtime import native  # noqa: E402


Desc = native.MlpFusedDesc


def multi_step_profile(lib, dims, lay, B, slots=8):
    """Timeline of the LAST step of an 8-step launch (in-place SGD with lr = 0), relative to that step's start."""
    x = torch.rand(slots * B, 784, device="cuda")
    y = torch.randint(0, 10, (slots * B,), device="cuda", dtype=torch.int32)
    params = torch.from_numpy(lay.init_reference(0)).cuda()
    grads = torch.zeros_like(params)
    metrics = torch.zeros(2 * slots, device="cuda")
    lr = torch.zeros(1, device="cuda")
    prof = torch.zeros(8, 16, dtype=torch.int64, device="cuda")
    d = Desc()
    d.num_layers = 3
    for i, v in enumerate(dims):
        d.dims[i] = v
    d.batch = B
    d.x, d.labels, d.params, d.grads = x.data_ptr(), y.data_ptr(), params.data_ptr(), grads.data_ptr()
    for l, s in enumerate(lay.layers):
        d.w_off[l], d.b_off[l] = s.w_off, s.b_off
    d.metrics, d.prof = metrics.data_ptr(), prof.data_ptr()
    d.sgd_params, d.lr, d.x_slots = params.data_ptr(), lr.data_ptr(), slots
    plan = C.c_void_p()
    native.check(lib.dmlp_mlp_fused_plan_create(C.byref(d), C.byref(plan)))
    lib.dmlp_mlp_fused_steps.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    acc = np.zeros((8, 16))
    n = 0
    for it in range(30):
        native.check(lib.dmlp_mlp_fused_steps(plan, 0, slots, None))
        torch.cuda.synchronize()
        if it >= 10:
            p = prof.cpu().numpy().astype(np.float64)
            acc += p - p[:, 13:14].min()
            n += 1
    acc /= n
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        native.check(lib.dmlp_mlp_fused_steps(plan, 0, slots, None))
    e1.record()
    torch.cuda.synchronize()
    names = ["-", "-", "F1 mma+scatter", "RS1 arrived", "H1+F2+scatter", "RS2 arrived", "H2+F3+allreduce",
             "logits arrived", "CE+dgL+gather dZ2", "dgrad2 done", "gather dZ1", "wgrads drained + wg1 done", "stores",
             "step start", "end-of-step barrier passed"]
    mean = acc.mean(axis=0)
    order = [13, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 14]
    out = {"multi_step_launch_us_per_step": round(e0.elapsed_time(e1) * 1e3 / 50 / slots, 2),
           "last_step_phase_end_us": {names[i]: round(mean[i] / 1e3, 2) for i in order},
           "last_step_phase_us": {names[order[k]]: round((mean[order[k]] - mean[order[k - 1]]) / 1e3, 2) for k in range(1, len(order))}}
    print(json.dumps(out, indent=1))


def main():
    lib = native.lib()
    dims = (784, 128, 64, 10)
    lay = ParamLayout(dims)
    B = 64
    x = torch.rand(B, 784, device="cuda")
    y = torch.randint(0, 10, (B,), device="cuda", dtype=torch.int32)
    params = torch.from_numpy(lay.init_reference(0)).cuda()
    grads = torch.zeros_like(params)
    metrics = torch.zeros(4, device="cuda")
    prof = torch.zeros(8, 16, dtype=torch.int64, device="cuda")
    d = Desc()
    d.num_layers = 3
    for i, v in enumerate(dims):
        d.dims[i] = v
    d.batch = B
    d.x, d.labels, d.params, d.grads = x.data_ptr(), y.data_ptr(), params.data_ptr(), grads.data_ptr()
    for l, s in enumerate(lay.layers):
        d.w_off[l], d.b_off[l] = s.w_off, s.b_off
    d.metrics, d.prof = metrics.data_ptr(), prof.data_ptr()
    plan = C.c_void_p()
    lib.dmlp_mlp_fused_plan_create.argtypes = [C.POINTER(Desc), C.POINTER(C.c_void_p)]
    lib.dmlp_mlp_fused_step.argtypes = [C.c_void_p, C.c_void_p]
    native.check(lib.dmlp_mlp_fused_plan_create(C.byref(d), C.byref(plan)))
    st = torch.cuda.Stream()
    acc = np.zeros((8, 16))
    n = 0
    with torch.cuda.stream(st):
        for it in range(30):
            native.check(lib.dmlp_mlp_fused_step(plan, st.cuda_stream))
            torch.cuda.synchronize()
            if it >= 10:
                p = prof.cpu().numpy().astype(np.float64)
                acc += p - p[:, :1].min()
                n += 1
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            for _ in range(50):
                native.check(lib.dmlp_mlp_fused_step(plan, st.cuda_stream))
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        g.replay()
        e1.record(st)
        torch.cuda.synchronize()
    acc /= n
    multi_step_profile(lib, dims, lay, B)
    names = ["entry", "cluster live", "F1 mma+scatter", "RS1 arrived", "H1+F2+scatter", "RS2 arrived", "H2+F3+allreduce",
             "logits arrived", "CE+dgL+gather dZ2", "dgrad2 done", "gather dZ1", "wgrads drained + wg1 done", "stores"]
    mean = acc.mean(axis=0)
    out = {"us_per_launch_graph": round(e0.elapsed_time(e1) * 1e3 / 50, 2),
           "phase_end_us_mean_over_ctas": {names[i]: round(mean[i] / 1e3, 2) for i in range(13)},
           "phase_us": {names[i]: round((mean[i] - mean[i - 1]) / 1e3, 2) for i in range(1, 13)}}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
