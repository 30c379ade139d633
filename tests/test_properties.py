"""Property tests (hypothesis) of the host-side pieces: protobuf wire codec round trips, the limit
parser restatement against strtoul semantics, and Contract A invariants over random flag values."""
import re

from hypothesis import given, settings, strategies as st

from oracle import plugin_model as M
from oracle.quota_model import limit_from_env

ident = st.text(alphabet=st.characters(min_codepoint=33, max_codepoint=126), min_size=0, max_size=40)


@settings(max_examples=200, deadline=None)
@given(envs=st.dictionaries(ident.filter(lambda s: s != ""), ident, max_size=8),
       mounts=st.lists(st.tuples(ident, ident, st.booleans()), max_size=6),
       ann=st.dictionaries(ident.filter(lambda s: s != ""), ident, max_size=3))
def test_allocate_response_wire_round_trip(vgpu, envs, mounts, ann):
    from vgpu_b200 import grpc_server as G, plugin as P
    r = P.ContainerAllocateResponse(Envs=envs, Mounts=[P.Mount(*m) for m in mounts], Annotations=ann)
    (d,) = G.dec_allocate_response(G.enc_allocate_response([r]))
    assert d["Envs"] == envs and d["Annotations"] == ann
    assert [(m["ContainerPath"], m["HostPath"], m["ReadOnly"]) for m in d["Mounts"]] == [tuple(m) for m in mounts]
    assert G.dec_allocate_container(G.enc_allocate_container(r))["Envs"] == envs


@settings(max_examples=200, deadline=None)
@given(reqs=st.lists(st.fixed_dictionaries({"AvailableDeviceIDs": st.lists(ident.filter(bool), max_size=5),
                                            "MustIncludeDeviceIDs": st.lists(ident.filter(bool), max_size=3),
                                            "AllocationSize": st.integers(0, 1 << 20)}), max_size=4))
def test_preferred_request_wire_round_trip(vgpu, reqs):
    from vgpu_b200 import grpc_server as G
    assert G.dec_preferred_request(G.enc_preferred_request(reqs)) == reqs


@settings(max_examples=300, deadline=None)
@given(n=st.integers(0, (1 << 64) - 1), suffix=st.sampled_from(["", "k", "K", "m", "M", "g", "G"]), base=st.sampled_from([10, 16, 8]))
def test_limit_parser_restatement(n, suffix, base):
    text = {10: str(n), 16: hex(n), 8: ("0" + oct(n)[2:]) if n else "0"}[base] + suffix
    scale = {"": 1, "k": 1 << 10, "m": 1 << 20, "g": 1 << 30}[suffix.lower()]
    want = n * scale if n * scale < (1 << 64) else 0   # overflow of the multiplication -> "no limit"
    assert limit_from_env(text) == want


@settings(max_examples=200, deadline=None)
@given(mib=st.integers(1, 1 << 22), split=st.integers(1, 64), ms=st.floats(0.01, 16.0), cs=st.floats(0.01, 16.0), k=st.integers(1, 6))
def test_contract_a_invariants(vgpu, mib, split, ms, cs, k):
    """Every vGPU of a GPU gets the same quota; the quotas never exceed scaling x physical; the SM
    limit never exceeds 100*coresScaling/split; ids are unique and ordered; C-ABI == restatement."""
    P = vgpu.plugin
    uuid = "GPU-aaaaaaaa-0000-0000-0000-000000000001"
    p = P.NewNvidiaDevicePlugin(P.ResourceManager([P.Device(uuid, memory_mib=mib)]),
                                P.Flags(deviceSplitCount=split, deviceMemoryScaling=ms, deviceCoresScaling=cs))
    p.initialize()
    ids = [v.ID for v in p.vDevices]
    assert ids == [f"{uuid}-{i}" for i in range(split)]
    mems = {v.memory for v in p.vDevices}
    assert len(mems) == 1 and mems.pop() * split <= mib * ms + split
    take = ids[: min(k, split)]
    (r,) = p.Allocate([take])
    assert int(r.Envs["CUDA_DEVICE_SM_LIMIT"]) == M.sm_limit(split, cs) <= 100 * cs / split
    assert ("CUDA_OVERSUBSCRIBE" in r.Envs) == (ms > 1)
    assert r.Envs["NVIDIA_DEVICE_MAP"].split(" ") == [f"{i}:{uuid}" for i in range(len(take))]
    assert re.match(r"^/tmp/[0-9a-f-]{36}\.cache$", r.Envs["CUDA_DEVICE_MEMORY_SHARED_CACHE"])
    want = M.device2vdevice([{"ID": uuid, "memory_mib": mib}], split, ms)
    assert [v.memory for v in p.vDevices] == [w["memory"] for w in want]
