"""The runtime-built gpu_sim contract matches the reference schema (DSML/proto/gpu_sim.proto)."""
import re

import pytest

from dmlp_b200.proto import gpu_sim as pb


def _parse_reference_proto():
    """Parses message/field/enum/rpc declarations out of the reference .proto if it is mounted."""
    path = "/root/reference/DSML/proto/gpu_sim.proto"
    try:
        text = open(path).read()
    except OSError:
        pytest.skip("reference proto not mounted")
    text = re.sub(r"//.*", "", text)
    msgs = {}
    for m in re.finditer(r"message\s+(\w+)\s*\{((?:[^{}]|\{[^{}]*\})*)\}", text):
        body = m.group(2)
        fields = {}
        for f in re.finditer(r"(?:repeated\s+)?(map<[^>]+>|[\w.]+)\s+(\w+)\s*=\s*(\d+)\s*;", body):
            fields[f.group(2)] = int(f.group(3))
        msgs[m.group(1)] = fields
    rpcs = {}
    for s in re.finditer(r"service\s+(\w+)\s*\{((?:[^{}]|\{[^{}]*\})*)\}", text):
        rpcs[s.group(1)] = re.findall(r"rpc\s+(\w+)\s*\(\s*(stream\s+)?(\w+)\s*\)\s*returns\s*\(\s*(\w+)\s*\)", s.group(2))
    enums = {e.group(1): dict((n, int(v)) for n, v in re.findall(r"(\w+)\s*=\s*(\d+)\s*;", e.group(2)))
             for e in re.finditer(r"enum\s+(\w+)\s*\{([^}]*)\}", text)}
    return msgs, rpcs, enums


def test_messages_fields_and_numbers_match_reference():
    msgs, _, _ = _parse_reference_proto()
    ours = {name: {f[0]: f[2] for f in fields} for name, fields in pb.MESSAGES}
    for name, fields in msgs.items():
        assert name in ours, f"missing message {name}"
        assert ours[name] == fields, f"{name}: {ours[name]} != {fields}"


def test_services_and_full_method_names():
    _, rpcs, _ = _parse_reference_proto()
    for svc, methods in rpcs.items():
        mine = {m[0]: m for m in pb.SERVICES[svc]}
        for name, stream, req, resp in methods:
            assert name in mine
            assert mine[name][1] == req and mine[name][2] == resp and mine[name][3] == bool(stream)
    assert pb.full_method("GPUDevice", "StreamSend") == "/gpu_sim.GPUDevice/StreamSend"
    assert pb.full_method("GPUCoordinator", "AllReduceRing") == "/gpu_sim.GPUCoordinator/AllReduceRing"
    assert len([m for m in pb.SERVICES["GPUDevice"] if ("GPUDevice", m[0]) not in pb.EXTENSION_METHODS]) == 6
    assert len(pb.SERVICES["GPUCoordinator"]) == 9


def test_enums_match_reference():
    _, _, enums = _parse_reference_proto()
    assert dict(pb.ENUMS["Status"]) == enums["Status"] == {"IN_PROGRESS": 0, "SUCCESS": 1, "FAILED": 2}
    assert dict(pb.ENUMS["ReduceOp"]) == enums["ReduceOp"] == {"SUM": 0, "PROD": 1, "MIN": 2, "MAX": 3}


def test_wire_roundtrip_oneof_and_map():
    req = pb.MemcpyRequest(hostToDevice=pb.MemcpyHostToDeviceRequest(
        hostSrcData=b"Hello GPU!", dstDeviceId=pb.DeviceId(value=7), dstMemAddr=pb.MemAddr(value=0x1000)))
    back = pb.MemcpyRequest.FromString(req.SerializeToString())
    assert back.WhichOneof("either") == "hostToDevice" and back.hostToDevice.hostSrcData == b"Hello GPU!"
    ar = pb.AllReduceRingRequest(commId=3, count=16, op=pb.MAX)
    ar.memAddrs[1].value = 0x1800
    back = pb.AllReduceRingRequest.FromString(ar.SerializeToString())
    assert back.memAddrs[1].value == 0x1800 and back.op == pb.MAX
    # golden bytes: field 1 varint 3, field 2 varint 16 — what any protoc-generated peer emits
    assert pb.AllReduceRingRequest(commId=3, count=16).SerializeToString() == b"\x08\x03\x10\x10"


def test_get_comm_status_accepts_stale_and_current_request_shapes():
    # the reference's generated stubs predate field 2; both decode
    assert pb.GetCommStatusRequest.FromString(b"\x08\x05").commId == 5
    cur = pb.GetCommStatusRequest(commId=5, device_addresses=["a:1"])
    assert pb.GetCommStatusRequest.FromString(cur.SerializeToString()).device_addresses == ["a:1"]
