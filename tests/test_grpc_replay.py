"""Config 1 of BASELINE.json: Register / ListAndWatch / Allocate replayed over real gRPC on unix
sockets against a kubelet stub, 1 mocked GPU split 10 ways (SURVEY.md §8 c.1 known answers)."""
import json
import os
import re
import tempfile
from concurrent import futures

import grpc
import pytest

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "contract_a.json")))
UUID = GOLD["mock"]["uuid"]
ident = lambda b: b  # noqa: E731


class KubeletStub(grpc.GenericRpcHandler):
    """The kubelet side: a Registration service that records what the plugin announces."""

    def __init__(self):
        self.registered = []

    def service(self, details):
        if details.method == "/v1beta1.Registration/Register":
            def reg(req, ctx):
                self.registered.append(req)
                return b""
            return grpc.unary_unary_rpc_method_handler(reg, ident, ident)
        return None


@pytest.fixture()
def rig(vgpu):
    from vgpu_b200 import grpc_server as G, plugin as P
    td = tempfile.mkdtemp(prefix="dp")
    kubelet_sock, plugin_sock = os.path.join(td, "kubelet.sock"), os.path.join(td, "nvidia-gpu.sock")
    stub = KubeletStub()
    ksrv = grpc.server(futures.ThreadPoolExecutor(max_workers=2))
    ksrv.add_generic_rpc_handlers((stub,))
    ksrv.add_insecure_port("unix://" + kubelet_sock)
    ksrv.start()
    flags = P.Flags(deviceSplitCount=10, pcibusfile=GOLD["mock"]["pcibusfile"])
    plugin = P.NewNvidiaDevicePlugin(P.ResourceManager([P.Device(UUID, Index="0", Paths=["/dev/nvidia0"], memory_mib=GOLD["mock"]["total_mib"], numa=0)]), flags)
    plugin.initialize()
    psrv = G.Serve(plugin, plugin_sock)
    G.Register(plugin, kubelet_sock)
    ch = grpc.insecure_channel("unix://" + plugin_sock)
    yield G, stub, ch, plugin
    ch.close()
    psrv.stop(0)
    ksrv.stop(0)


def test_register_announces_endpoint_resource_and_options(rig):
    G, stub, ch, plugin = rig
    assert len(stub.registered) == 1
    assert G.dec_register(stub.registered[0]) == {"Version": "v1beta1", "Endpoint": "nvidia-gpu.sock", "ResourceName": "nvidia.com/gpu",
                                                  "Options": {"PreStartRequired": False, "GetPreferredAllocationAvailable": True}}
    opts = ch.unary_unary("/v1beta1.DevicePlugin/GetDevicePluginOptions", request_serializer=ident, response_deserializer=ident)(b"")
    assert opts == G.enc_options({"GetPreferredAllocationAvailable": True})


def test_list_and_watch_streams_ten_healthy_vgpus_then_health_updates(rig):
    G, stub, ch, plugin = rig
    plugin.markUnhealthy(plugin.cachedDevices[0])
    stream = ch.unary_stream("/v1beta1.DevicePlugin/ListAndWatch", request_serializer=ident, response_deserializer=ident)(b"")
    first, second = (G.dec_list_and_watch(m) for m in list(stream)[:2])
    assert [d["ID"] for d in first] == [f"{UUID}-{i}" for i in range(10)]
    assert all(d["Health"] == "Healthy" and d["Topology"] == {"Nodes": [{"ID": 0}]} for d in first)
    assert all(d["Health"] == "Unhealthy" for d in second)


def test_allocate_every_vgpu_and_a_pair_over_the_wire(rig):
    G, stub, ch, plugin = rig
    alloc = ch.unary_unary("/v1beta1.DevicePlugin/Allocate", request_serializer=ident, response_deserializer=ident)
    want_mounts = [{"ContainerPath": c, "HostPath": h, "ReadOnly": r} for c, h, r in GOLD["mounts"]]
    caches = set()
    for i in range(10):
        (r,) = G.dec_allocate_response(alloc(G.enc_allocate_request([[f"{UUID}-{i}"]])))
        cache = r["Envs"].pop("CUDA_DEVICE_MEMORY_SHARED_CACHE")
        assert re.match(r"^/tmp/[0-9a-f-]{36}\.cache$", cache)
        caches.add(cache)
        assert r["Envs"] == GOLD["cases"][0]["envs"]
        assert r["Mounts"] == want_mounts and r["Devices"] == [] and r["Annotations"] == {}
    assert len(caches) == 10
    # one request, two containers; second container holds two vGPUs of the same physical GPU
    r1, r2 = G.dec_allocate_response(alloc(G.enc_allocate_request([[f"{UUID}-0"], [f"{UUID}-3", f"{UUID}-7"]])))
    r2["Envs"].pop("CUDA_DEVICE_MEMORY_SHARED_CACHE")
    assert r2["Envs"] == GOLD["cases"][1]["envs"]


def test_allocate_unknown_device_is_a_grpc_error_with_the_go_text(rig):
    G, stub, ch, plugin = rig
    alloc = ch.unary_unary("/v1beta1.DevicePlugin/Allocate", request_serializer=ident, response_deserializer=ident)
    with pytest.raises(grpc.RpcError) as e:
        alloc(G.enc_allocate_request([[GOLD["unknown"]["id"]]]))
    assert e.value.code() == grpc.StatusCode.UNKNOWN and e.value.details() == GOLD["unknown"]["error"]


def test_preferred_allocation_over_the_wire(rig):
    G, stub, ch, plugin = rig
    pref = ch.unary_unary("/v1beta1.DevicePlugin/GetPreferredAllocation", request_serializer=ident, response_deserializer=ident)
    avail = [f"{UUID}-{i}" for i in (5, 1, 8)]
    out = G.dec_preferred_response(pref(G.enc_preferred_request([{"AvailableDeviceIDs": avail, "AllocationSize": 1},
                                                                   {"AvailableDeviceIDs": avail, "AllocationSize": 2}])))
    assert out == [[avail[0]], []]


def test_wire_codec_matches_google_protobuf(vgpu):
    """Cross-check the hand-written codec against google.protobuf on a descriptor built from api.proto's fields."""
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory
    from vgpu_b200 import grpc_server as G, plugin as P
    fd = descriptor_pb2.FileDescriptorProto(name="t_api.proto", package="t_v1beta1", syntax="proto3")
    m = fd.message_type.add(name="Mount")
    for n, num, t in (("container_path", 1, 9), ("host_path", 2, 9), ("read_only", 3, 8)):
        m.field.add(name=n, number=num, type=t, label=1)
    c = fd.message_type.add(name="ContainerAllocateResponse")
    e = c.nested_type.add(name="EnvsEntry")
    e.options.map_entry = True
    e.field.add(name="key", number=1, type=9, label=1)
    e.field.add(name="value", number=2, type=9, label=1)
    c.field.add(name="envs", number=1, type=11, label=3, type_name=".t_v1beta1.ContainerAllocateResponse.EnvsEntry")
    c.field.add(name="mounts", number=2, type=11, label=3, type_name=".t_v1beta1.Mount")
    a = fd.message_type.add(name="AllocateResponse")
    a.field.add(name="container_responses", number=1, type=11, label=3, type_name=".t_v1beta1.ContainerAllocateResponse")
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    Msg = message_factory.GetMessageClass(pool.FindMessageTypeByName("t_v1beta1.AllocateResponse"))
    resp = P.ContainerAllocateResponse(Envs={"A": "1", "NVIDIA_DEVICE_MAP": "0:GPU-x 1:GPU-x"},
                                       Mounts=[P.Mount("/etc/ld.so.preload", "/usr/local/vgpu/ld.so.preload", True), P.Mount("/a", "/b", False)])
    pb = Msg()
    pb.ParseFromString(G.enc_allocate_response([resp]))
    cr = pb.container_responses[0]
    assert dict(cr.envs) == resp.Envs
    assert [(x.container_path, x.host_path, x.read_only) for x in cr.mounts] == [("/etc/ld.so.preload", "/usr/local/vgpu/ld.so.preload", True), ("/a", "/b", False)]
    assert G.dec_allocate_response(pb.SerializeToString())[0]["Envs"] == resp.Envs
