"""kubelet device-plugin v1beta1 endpoint over a unix socket (Contract A on the wire).

Serves the same five RPCs the reference registers (server.go:171-218, 221-533; wire format
vendor/k8s.io/kubelet/pkg/apis/deviceplugin/v1beta1/api.proto:23-76) on top of `plugin.py`, whose
response building happens behind the C-ABI.  protobuf is encoded/decoded by the ~60-line proto3
wire codec below (the messages only use strings, bools, varints, nested and repeated fields), so
no generated code or protoc is needed.  Errors surface the way grpc-go surfaces a plain Go error:
status UNKNOWN with the error text as message.
"""
from __future__ import annotations

import os
from concurrent import futures
from typing import Dict, Iterable, List, Tuple

import grpc

from . import plugin as P

# ----------------------------------------------------------------------------- proto3 wire codec


def _varint(n: int) -> bytes:
    n &= (1 << 64) - 1
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        out.append(b | (0x80 if n else 0))
        if not n:
            return bytes(out)


def _f_bytes(field: int, b: bytes) -> bytes:
    return _varint((field << 3) | 2) + _varint(len(b)) + b


def _f_str(field: int, s: str) -> bytes:
    return _f_bytes(field, s.encode()) if s else b""


def _f_varint(field: int, v: int) -> bytes:
    return _varint((field << 3) | 0) + _varint(v) if v else b""


def parse(buf: bytes) -> List[Tuple[int, int, object]]:
    """-> [(field, wire_type, value)]; value is int for varint, bytes for length-delimited."""
    out, i = [], 0
    while i < len(buf):
        key = shift = 0
        while True:
            b = buf[i]; i += 1
            key |= (b & 0x7F) << shift; shift += 7
            if not b & 0x80:
                break
        field, wt = key >> 3, key & 7
        if wt == 0:
            v = shift = 0
            while True:
                b = buf[i]; i += 1
                v |= (b & 0x7F) << shift; shift += 7
                if not b & 0x80:
                    break
            out.append((field, wt, v))
        elif wt == 2:
            n = shift = 0
            while True:
                b = buf[i]; i += 1
                n |= (b & 0x7F) << shift; shift += 7
                if not b & 0x80:
                    break
            out.append((field, wt, bytes(buf[i:i + n]))); i += n
        elif wt == 1:
            out.append((field, wt, bytes(buf[i:i + 8]))); i += 8
        elif wt == 5:
            out.append((field, wt, bytes(buf[i:i + 4]))); i += 4
        else:
            raise ValueError(f"unsupported wire type {wt}")
    return out


# ----------------------------------------------------------------------------- messages (api.proto)
def enc_options(o: dict) -> bytes:  # DevicePluginOptions, api.proto:27-32
    return _f_varint(1, int(bool(o.get("PreStartRequired")))) + _f_varint(2, int(bool(o.get("GetPreferredAllocationAvailable"))))


def enc_register(r: dict) -> bytes:  # RegisterRequest, api.proto:34-44
    return _f_str(1, r["Version"]) + _f_str(2, r["Endpoint"]) + _f_str(3, r["ResourceName"]) + _f_bytes(4, enc_options(r["Options"]))


def dec_register(b: bytes) -> dict:
    r = {"Version": "", "Endpoint": "", "ResourceName": "", "Options": {"PreStartRequired": False, "GetPreferredAllocationAvailable": False}}
    for f, _, v in parse(b):
        if f == 1: r["Version"] = v.decode()
        elif f == 2: r["Endpoint"] = v.decode()
        elif f == 3: r["ResourceName"] = v.decode()
        elif f == 4:
            for g, _, w in parse(v):
                if g == 1: r["Options"]["PreStartRequired"] = bool(w)
                elif g == 2: r["Options"]["GetPreferredAllocationAvailable"] = bool(w)
    return r


def enc_device(d: dict) -> bytes:  # Device, api.proto:101-110
    out = _f_str(1, d["ID"]) + _f_str(2, d["Health"])
    if d.get("Topology"):
        nodes = b"".join(_f_bytes(1, _f_varint(1, n["ID"])) for n in d["Topology"]["Nodes"])
        out += _f_bytes(3, nodes)
    return out


def enc_list_and_watch(devs: Iterable[dict]) -> bytes:
    return b"".join(_f_bytes(1, enc_device(d)) for d in devs)


def dec_list_and_watch(b: bytes) -> List[dict]:
    out = []
    for f, _, v in parse(b):
        if f != 1:
            continue
        d = {"ID": "", "Health": "", "Topology": None}
        for g, _, w in parse(v):
            if g == 1: d["ID"] = w.decode()
            elif g == 2: d["Health"] = w.decode()
            elif g == 3:
                nodes = []
                for h, _, x in parse(w):
                    if h == 1:
                        nid = 0
                        for k, _, y in parse(x):
                            if k == 1: nid = y
                        nodes.append({"ID": nid})
                d["Topology"] = {"Nodes": nodes}
        out.append(d)
    return out


def enc_allocate_request(containers: Iterable[Iterable[str]]) -> bytes:
    return b"".join(_f_bytes(1, b"".join(_f_bytes(1, i.encode()) for i in ids)) for ids in containers)


def dec_allocate_request(b: bytes) -> List[List[str]]:
    return [[w.decode() for g, _, w in parse(v) if g == 1] for f, _, v in parse(b) if f == 1]


def _enc_map(field: int, m: Dict[str, str]) -> bytes:
    return b"".join(_f_bytes(field, _f_str(1, k) + _f_str(2, v)) for k, v in m.items())


def enc_allocate_response(resps: Iterable[P.ContainerAllocateResponse]) -> bytes:
    out = b""
    for r in resps:
        body = _enc_map(1, r.Envs)
        for m in r.Mounts:
            body += _f_bytes(2, _f_str(1, m.ContainerPath) + _f_str(2, m.HostPath) + _f_varint(3, int(m.ReadOnly)))
        for d in r.Devices:
            body += _f_bytes(3, _f_str(1, d.ContainerPath) + _f_str(2, d.HostPath) + _f_str(3, d.Permissions))
        body += _enc_map(4, r.Annotations)
        out += _f_bytes(1, body)
    return out


def enc_allocate_container(r: P.ContainerAllocateResponse) -> bytes:
    """One ContainerAllocateResponse (what the kubelet stores as AllocResp in its checkpoint)."""
    return parse(enc_allocate_response([r]))[0][2]


def dec_allocate_container(v: bytes) -> dict:
    return dec_allocate_response(_f_bytes(1, v))[0]


def dec_allocate_response(b: bytes) -> List[dict]:
    out = []
    for f, _, v in parse(b):
        if f != 1:
            continue
        r = {"Envs": {}, "Mounts": [], "Devices": [], "Annotations": {}}
        for g, _, w in parse(v):
            if g in (1, 4):
                kv = {h: x.decode() for h, _, x in parse(w)}
                r["Envs" if g == 1 else "Annotations"][kv.get(1, "")] = kv.get(2, "")
            elif g == 2:
                m = {"ContainerPath": "", "HostPath": "", "ReadOnly": False}
                for h, _, x in parse(w):
                    if h == 1: m["ContainerPath"] = x.decode()
                    elif h == 2: m["HostPath"] = x.decode()
                    elif h == 3: m["ReadOnly"] = bool(x)
                r["Mounts"].append(m)
            elif g == 3:
                d = {h: x.decode() for h, _, x in parse(w)}
                r["Devices"].append({"ContainerPath": d.get(1, ""), "HostPath": d.get(2, ""), "Permissions": d.get(3, "")})
        out.append(r)
    return out


def enc_preferred_request(reqs: Iterable[dict]) -> bytes:
    out = b""
    for r in reqs:
        body = b"".join(_f_bytes(1, i.encode()) for i in r.get("AvailableDeviceIDs", []))
        body += b"".join(_f_bytes(2, i.encode()) for i in r.get("MustIncludeDeviceIDs", []))
        body += _f_varint(3, int(r.get("AllocationSize", 0)))
        out += _f_bytes(1, body)
    return out


def dec_preferred_request(b: bytes) -> List[dict]:
    out = []
    for f, _, v in parse(b):
        if f != 1:
            continue
        r = {"AvailableDeviceIDs": [], "MustIncludeDeviceIDs": [], "AllocationSize": 0}
        for g, _, w in parse(v):
            if g == 1: r["AvailableDeviceIDs"].append(w.decode())
            elif g == 2: r["MustIncludeDeviceIDs"].append(w.decode())
            elif g == 3: r["AllocationSize"] = w
        out.append(r)
    return out


def enc_preferred_response(resps: Iterable[List[str]]) -> bytes:
    return b"".join(_f_bytes(1, b"".join(_f_bytes(1, i.encode()) for i in ids)) for ids in resps)


def dec_preferred_response(b: bytes) -> List[List[str]]:
    return [[w.decode() for g, _, w in parse(v) if g == 1] for f, _, v in parse(b) if f == 1]


# ----------------------------------------------------------------------------- server
_ident = lambda b: b  # noqa: E731  (messages travel as raw bytes; we do the (de)serialisation)


class _Handler(grpc.GenericRpcHandler):
    def __init__(self, plugin: P.NvidiaDevicePlugin):
        self.p = plugin

    def service(self, details):
        m = details.method
        if m == "/v1beta1.DevicePlugin/GetDevicePluginOptions":
            return grpc.unary_unary_rpc_method_handler(lambda req, ctx: enc_options(self.p.GetDevicePluginOptions()), _ident, _ident)
        if m == "/v1beta1.DevicePlugin/ListAndWatch":
            def law(req, ctx):
                for devs in self.p.ListAndWatch():
                    yield enc_list_and_watch(devs)
            return grpc.unary_stream_rpc_method_handler(law, _ident, _ident)
        if m == "/v1beta1.DevicePlugin/Allocate":
            def alloc(req, ctx):
                try:
                    return enc_allocate_response(self.p.Allocate(dec_allocate_request(req)))
                except P.PluginError as e:
                    ctx.abort(grpc.StatusCode.UNKNOWN, str(e))
            return grpc.unary_unary_rpc_method_handler(alloc, _ident, _ident)
        if m == "/v1beta1.DevicePlugin/GetPreferredAllocation":
            def pref(req, ctx):
                try:
                    r = self.p.GetPreferredAllocation(dec_preferred_request(req))
                except P.PluginError as e:
                    ctx.abort(grpc.StatusCode.UNKNOWN, str(e))
                return enc_preferred_response(r or [])
            return grpc.unary_unary_rpc_method_handler(pref, _ident, _ident)
        if m == "/v1beta1.DevicePlugin/PreStartContainer":
            return grpc.unary_unary_rpc_method_handler(lambda req, ctx: b"", _ident, _ident)  # server.go:536-538
        return None


def Serve(plugin: P.NvidiaDevicePlugin, socket_path: str) -> grpc.Server:
    """server.go:171-218 (without the crash-restart loop): listen on the plugin's unix socket."""
    if os.path.exists(socket_path):
        os.unlink(socket_path)
    srv = grpc.server(futures.ThreadPoolExecutor(max_workers=4))
    srv.add_generic_rpc_handlers((_Handler(plugin),))
    srv.add_insecure_port("unix://" + socket_path)
    srv.start()
    return srv


def Register(plugin: P.NvidiaDevicePlugin, kubelet_socket: str, timeout: float = 5.0) -> None:
    """server.go:221-243: tell the kubelet our endpoint and resource name."""
    with grpc.insecure_channel("unix://" + kubelet_socket) as ch:
        grpc.channel_ready_future(ch).result(timeout=timeout)
        call = ch.unary_unary("/v1beta1.Registration/Register", request_serializer=_ident, response_deserializer=_ident)
        call(enc_register(plugin.Register()), timeout=timeout)
