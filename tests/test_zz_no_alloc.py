"""The allocation-free all-reduce entry point, called through the C ABI (GPU only).

``dmlp_allreduce_sum_f32_noalloc`` is for callers that may not allocate GPU memory when they call
into the library — a process whose GPU memory is already fully owned by its own pool, or a loop
that cannot accept the device-wide stall of cudaMalloc/cudaFree. It keeps all of its working
memory in a fixed array inside the shared library. This test drives it with ctypes exactly as a C
program would, twice in a row on fresh data (the array is shared by every call and keeps the
previous call's contents, so the second call proves the control words are re-initialised), and
compares each result with the communicator-based ``dmlp_allreduce`` on the same inputs AND with a plain PyTorch
fp64 sum of them; after each call the watchdog record kept inside the static array is read back through
``dmlp_allreduce_noalloc_status`` (a kernel that gave up must not look like a success).

The file name sorts last so a ``pytest -x`` run reaches it only after every other test has run.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_FLOATS_PER_RANK = 65536  # documented limit: count <= world * 65536


def _c_abi():
    from dmlp_b200.runtime import native
    lib = C.CDLL(native.so_path())
    fn = lib.dmlp_allreduce_sum_f32_noalloc
    fn.argtypes = [C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_size_t, C.c_void_p]
    fn.restype = C.c_int
    lib.dmlp_last_error.restype = C.c_char_p
    lib.dmlp_allreduce_noalloc_status.argtypes = [C.POINTER(C.c_uint32)]
    lib.dmlp_allreduce_noalloc_status.restype = C.c_int
    return lib, fn


def _status(lib):
    st = (C.c_uint32 * 4)()
    assert lib.dmlp_allreduce_noalloc_status(st) == 0, lib.dmlp_last_error().decode()
    return list(st)


def _ptr_table(tensors):
    return (C.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])


@pytest.mark.parametrize("world,count", [(2, 109386), (3, 101770), (8, 4099), (1, 1000)])
def test_noalloc_matches_communicator_allreduce_twice(world, count):
    from dmlp_b200.runtime import native
    from dmlp_b200.runtime.comm import Communicator
    lib, fn = _c_abi()
    comm = Communicator.local([0] * world, staging_bytes=1 << 20)
    stream = torch.cuda.current_stream().cuda_stream
    for call in range(2):  # two calls in a row share the static array
        torch.manual_seed(100 * world + call)
        ins = [torch.randn(count, device="cuda") for _ in range(world)]
        got = [torch.full((count,), float("nan"), device="cuda") for _ in range(world)]
        want = [torch.full((count,), float("nan"), device="cuda") for _ in range(world)]
        rc = fn(world, _ptr_table(ins), _ptr_table(got), count, stream)
        assert rc == 0, lib.dmlp_last_error().decode()
        comm.allreduce([t.data_ptr() for t in ins], [t.data_ptr() for t in want], count,
                       dtype=native.F32, op=native.SUM, algo=native.ALGO_RING, stream=stream)
        torch.cuda.synchronize()
        comm.check()
        assert _status(lib)[0] == 0, f"the allocation-free all-reduce reported {_status(lib)}"
        scale = max(1.0, want[0].abs().max().item())
        ref = torch.stack([t.double() for t in ins]).sum(0)  # plain PyTorch reference of the same op
        for g, w in zip(got, want):
            assert (g.double() - w.double()).abs().max().item() < 1e-5 * scale
            assert (g.double() - ref).abs().max().item() < 1e-5 * scale
        for g in got[1:]:
            assert torch.equal(g, got[0]), "every logical rank must receive the same bits"
    comm.destroy()


def test_noalloc_in_place_and_limits():
    lib, fn = _c_abi()
    world = 2
    bufs = [torch.full((1024,), float(r + 1), device="cuda") for r in range(world)]
    assert fn(world, _ptr_table(bufs), _ptr_table(bufs), 1024, None) == 0  # out aliases in
    torch.cuda.synchronize()
    assert _status(lib) == [0, 0, 0, 0]
    assert all(torch.equal(b, torch.full_like(b, 3.0)) for b in bufs)
    # one element past the documented capacity is refused with a clear message, nothing is launched
    too_many = world * MAX_FLOATS_PER_RANK + 1
    big = [torch.empty(too_many, device="cuda") for _ in range(world)]
    rc = fn(world, _ptr_table(big), _ptr_table(big), too_many, None)
    assert rc == 1 and b"exceeds the static scratch capacity" in lib.dmlp_last_error()
    assert fn(world, _ptr_table(big), _ptr_table(big), too_many - 1, None) == 0  # exactly at the limit
    torch.cuda.synchronize()
    assert fn(9, _ptr_table(big), _ptr_table(big), 16, None) == 1  # more ranks than the array has slices for
