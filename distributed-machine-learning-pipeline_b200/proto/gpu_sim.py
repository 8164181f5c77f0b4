"""The ``gpu_sim`` wire contract, built at import time without protoc.

Neither ``protoc`` nor ``grpc_tools`` exists on the target image, so the file descriptor is
assembled with ``descriptor_pb2`` and message classes come from ``message_factory``. Package,
service, method, message and field names and numbers are those of the reference contract
(``DSML/proto/gpu_sim.proto:1-244``), so a client generated from the reference ``.proto``
interoperates byte-for-byte. Two additions are tolerated on the wire: the ``RunForward`` /
``RunBackward`` RPCs that only exist in the reference's stale generated stubs
(``DSML/proto/gpu_sim.pb.go:1225-1425``), and ``GetCommStatusRequest`` with or without field 2.

A human-readable rendering of the same schema is in ``gpu_sim.proto`` beside this file; a unit
test keeps the two in sync.
"""
from __future__ import annotations

from typing import Dict, List, Tuple

from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

PACKAGE = "gpu_sim"
FD = descriptor_pb2.FieldDescriptorProto

_SCALARS = {
    "uint64": FD.TYPE_UINT64, "uint32": FD.TYPE_UINT32, "int64": FD.TYPE_INT64,
    "bool": FD.TYPE_BOOL, "bytes": FD.TYPE_BYTES, "string": FD.TYPE_STRING,
}

# (name, [(field, type, number, label?)]); type is a scalar keyword, a message name, or "enum:Name".
# label: "" singular, "repeated", "oneof:<group>", "map:<keytype>:<valuetype>".
MESSAGES: List[Tuple[str, List[Tuple[str, str, int, str]]]] = [
    ("DeviceId", [("value", "uint64", 1, "")]),
    ("Rank", [("value", "uint32", 1, "")]),
    ("MemAddr", [("value", "uint64", 1, "")]),
    ("StreamId", [("value", "uint64", 1, "")]),
    ("DeviceMetadata", [("deviceId", "DeviceId", 1, ""), ("minMemAddr", "MemAddr", 2, ""),
                        ("maxMemAddr", "MemAddr", 3, "")]),
    ("GetDeviceMetadataRequest", []),
    ("GetDeviceMetadataResponse", [("metadata", "DeviceMetadata", 1, "")]),
    ("BeginSendRequest", [("sendBuffAddr", "MemAddr", 1, ""), ("numBytes", "uint64", 2, ""),
                          ("dstRank", "Rank", 3, "")]),
    ("BeginSendResponse", [("initiated", "bool", 1, ""), ("streamId", "StreamId", 2, "")]),
    ("BeginReceiveRequest", [("streamId", "StreamId", 1, ""), ("recvBuffAddr", "MemAddr", 2, ""),
                             ("numBytes", "uint64", 3, ""), ("srcRank", "Rank", 4, "")]),
    ("BeginReceiveResponse", [("initiated", "bool", 1, "")]),
    ("DataChunk", [("data", "bytes", 1, ""), ("streamId", "uint64", 2, "")]),
    ("StreamSendResponse", [("success", "bool", 1, "")]),
    ("GetStreamStatusRequest", [("streamId", "StreamId", 1, "")]),
    ("GetStreamStatusResponse", [("status", "enum:Status", 1, "")]),
    ("CommInitRequest", [("numDevices", "uint32", 1, ""), ("device_addresses", "string", 2, "repeated")]),
    ("CommInitResponse", [("success", "bool", 1, ""), ("commId", "uint64", 2, ""),
                          ("devices", "DeviceMetadata", 3, "repeated")]),
    ("GetCommStatusRequest", [("commId", "uint64", 1, ""), ("device_addresses", "string", 2, "repeated")]),
    ("GetCommStatusResponse", [("status", "enum:Status", 1, "")]),
    ("GroupStartRequest", [("commId", "uint64", 1, "")]),
    ("GroupStartResponse", [("success", "bool", 1, "")]),
    ("GroupEndRequest", [("commId", "uint64", 1, "")]),
    ("GroupEndResponse", [("success", "bool", 1, "")]),
    ("AllReduceRingRequest", [("commId", "uint64", 1, ""), ("count", "uint64", 2, ""),
                              ("op", "enum:ReduceOp", 3, ""), ("memAddrs", "MemAddr", 4, "map:uint32")]),
    ("AllReduceRingResponse", [("success", "bool", 1, "")]),
    ("MemcpyHostToDeviceRequest", [("hostSrcData", "bytes", 1, ""), ("dstDeviceId", "DeviceId", 2, ""),
                                   ("dstMemAddr", "MemAddr", 3, "")]),
    ("MemcpyHostToDeviceResponse", [("success", "bool", 1, "")]),
    ("MemcpyDeviceToHostRequest", [("srcDeviceId", "DeviceId", 1, ""), ("srcMemAddr", "MemAddr", 2, ""),
                                   ("numBytes", "uint64", 3, "")]),
    ("MemcpyDeviceToHostResponse", [("dstData", "bytes", 1, "")]),
    ("MemcpyRequest", [("hostToDevice", "MemcpyHostToDeviceRequest", 1, "oneof:either"),
                       ("deviceToHost", "MemcpyDeviceToHostRequest", 2, "oneof:either")]),
    ("MemcpyResponse", [("hostToDevice", "MemcpyHostToDeviceResponse", 1, "oneof:either"),
                        ("deviceToHost", "MemcpyDeviceToHostResponse", 2, "oneof:either")]),
    ("CommDestroyRequest", [("commId", "uint64", 1, "")]),
    ("CommDestroyResponse", [("success", "bool", 1, "")]),
    ("CommFinalizeRequest", [("commId", "uint64", 1, "")]),
    ("CommFinalizeResponse", [("success", "bool", 1, "")]),
    ("NaiveAllReduceRequest", [("commId", "uint64", 1, ""), ("dataSize", "uint64", 2, ""),
                               ("latencyMs", "uint32", 3, "")]),
    ("NaiveAllReduceResponse", [("success", "bool", 1, ""), ("totalTimeMs", "int64", 2, ""),
                                ("totalDataTransferred", "int64", 3, "")]),
    # Only in the reference's stale generated stubs; kept so such clients still connect.
    ("RunForwardRequest", [("deviceId", "uint64", 1, ""), ("inputAddr", "uint64", 2, ""),
                           ("outputAddr", "uint64", 3, "")]),
    ("RunForwardResponse", [("success", "bool", 1, "")]),
    ("RunBackwardRequest", [("deviceId", "uint64", 1, ""), ("gradientAddr", "uint64", 2, "")]),
    ("RunBackwardResponse", [("success", "bool", 1, "")]),
]

ENUMS: Dict[str, List[Tuple[str, int]]] = {
    "Status": [("IN_PROGRESS", 0), ("SUCCESS", 1), ("FAILED", 2)],
    "ReduceOp": [("SUM", 0), ("PROD", 1), ("MIN", 2), ("MAX", 3)],
}

# service -> [(method, request, response, client_streaming)]
SERVICES: Dict[str, List[Tuple[str, str, str, bool]]] = {
    "GPUDevice": [
        ("GetDeviceMetadata", "GetDeviceMetadataRequest", "GetDeviceMetadataResponse", False),
        ("BeginSend", "BeginSendRequest", "BeginSendResponse", False),
        ("BeginReceive", "BeginReceiveRequest", "BeginReceiveResponse", False),
        ("StreamSend", "DataChunk", "StreamSendResponse", True),
        ("GetStreamStatus", "GetStreamStatusRequest", "GetStreamStatusResponse", False),
        ("Memcpy", "MemcpyRequest", "MemcpyResponse", False),
        ("RunForward", "RunForwardRequest", "RunForwardResponse", False),
        ("RunBackward", "RunBackwardRequest", "RunBackwardResponse", False),
    ],
    "GPUCoordinator": [
        ("CommInit", "CommInitRequest", "CommInitResponse", False),
        ("GetCommStatus", "GetCommStatusRequest", "GetCommStatusResponse", False),
        ("CommDestroy", "CommDestroyRequest", "CommDestroyResponse", False),
        ("CommFinalize", "CommFinalizeRequest", "CommFinalizeResponse", False),
        ("GroupStart", "GroupStartRequest", "GroupStartResponse", False),
        ("GroupEnd", "GroupEndRequest", "GroupEndResponse", False),
        ("AllReduceRing", "AllReduceRingRequest", "AllReduceRingResponse", False),
        ("NaiveAllReduce", "NaiveAllReduceRequest", "NaiveAllReduceResponse", False),
        ("Memcpy", "MemcpyRequest", "MemcpyResponse", False),
    ],
}
# Methods that are part of the authored .proto (the two extra device RPCs are wire-tolerated only).
EXTENSION_METHODS = {("GPUDevice", "RunForward"), ("GPUDevice", "RunBackward")}


def _camel_entry(field: str) -> str:
    return field[0].upper() + field[1:] + "Entry"


def build_file_descriptor() -> descriptor_pb2.FileDescriptorProto:
    f = descriptor_pb2.FileDescriptorProto()
    f.name = "dmlp_b200/gpu_sim.proto"
    f.package = PACKAGE
    f.syntax = "proto3"
    f.options.go_package = "/proto"
    for ename, values in ENUMS.items():
        e = f.enum_type.add()
        e.name = ename
        for vname, num in values:
            v = e.value.add()
            v.name, v.number = vname, num
    for mname, fields in MESSAGES:
        m = f.message_type.add()
        m.name = mname
        oneofs: Dict[str, int] = {}
        for fname, ftype, num, label in fields:
            fd = m.field.add()
            fd.name, fd.number = fname, num
            fd.json_name = fname
            fd.label = FD.LABEL_OPTIONAL
            if label == "repeated":
                fd.label = FD.LABEL_REPEATED
            elif label.startswith("oneof:"):
                group = label.split(":", 1)[1]
                if group not in oneofs:
                    oneofs[group] = len(m.oneof_decl)
                    m.oneof_decl.add().name = group
                fd.oneof_index = oneofs[group]
            if label.startswith("map:"):
                key_t = label.split(":")[1]
                entry = m.nested_type.add()
                entry.name = _camel_entry(fname)
                entry.options.map_entry = True
                kf = entry.field.add()
                kf.name, kf.number, kf.label, kf.type, kf.json_name = "key", 1, FD.LABEL_OPTIONAL, _SCALARS[key_t], "key"
                vf = entry.field.add()
                vf.name, vf.number, vf.label, vf.json_name = "value", 2, FD.LABEL_OPTIONAL, "value"
                vf.type, vf.type_name = FD.TYPE_MESSAGE, f".{PACKAGE}.{ftype}"
                fd.label = FD.LABEL_REPEATED
                fd.type, fd.type_name = FD.TYPE_MESSAGE, f".{PACKAGE}.{mname}.{entry.name}"
            elif ftype in _SCALARS:
                fd.type = _SCALARS[ftype]
            elif ftype.startswith("enum:"):
                fd.type, fd.type_name = FD.TYPE_ENUM, f".{PACKAGE}.{ftype[5:]}"
            else:
                fd.type, fd.type_name = FD.TYPE_MESSAGE, f".{PACKAGE}.{ftype}"
    for sname, methods in SERVICES.items():
        s = f.service.add()
        s.name = sname
        for mname, req, resp, cstream in methods:
            md = s.method.add()
            md.name = mname
            md.input_type, md.output_type = f".{PACKAGE}.{req}", f".{PACKAGE}.{resp}"
            md.client_streaming = cstream
    return f


_pool = descriptor_pool.DescriptorPool()
FILE_DESCRIPTOR_PROTO = build_file_descriptor()
_file_desc = _pool.Add(FILE_DESCRIPTOR_PROTO) or _pool.FindFileByName(FILE_DESCRIPTOR_PROTO.name)


def _cls(name: str):
    return message_factory.GetMessageClass(_pool.FindMessageTypeByName(f"{PACKAGE}.{name}"))


# Message classes are exported as module attributes: gpu_sim.MemAddr, gpu_sim.CommInitRequest, ...
_g = globals()
for _name, _ in MESSAGES:
    _g[_name] = _cls(_name)

Status = _pool.FindEnumTypeByName(f"{PACKAGE}.Status")
ReduceOp = _pool.FindEnumTypeByName(f"{PACKAGE}.ReduceOp")
IN_PROGRESS, SUCCESS, FAILED = 0, 1, 2
SUM, PROD, MIN, MAX = 0, 1, 2, 3


def status_name(v: int) -> str:
    return Status.values_by_number[v].name


def full_method(service: str, method: str) -> str:
    """'/gpu_sim.GPUDevice/Memcpy' — same full names as the reference's generated stubs
    (``DSML/proto/gpu_sim_grpc.pb.go:21-30,393-403``)."""
    return f"/{PACKAGE}.{service}/{method}"


def message_class(name: str):
    return _g[name]
