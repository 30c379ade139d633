import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, vgpu_b200
from vgpu_b200 import capi, datapath as dp
dev = torch.device("cuda", 0)
for pct, n in ((25, 4), (10, 10), (33, 3), (6, 15)):
    for flags in (0, 1):
        parts = (capi.SmPartition * n)()
        rc = capi.lib.vgpu_sm_partition_create_ex(0, pct, n, flags, parts)
        if rc != 0:
            print(pct, n, flags, "ERR", capi.last_error()); continue
        sms = [p.sm_count for p in parts]
        # a cuBLAS bf16 GEMM (2-CTA clusters on Blackwell) inside partition 0, and memtest rates
        st = torch.cuda.ExternalStream(parts[0].stream, device=dev)
        x = torch.randn(4096, 4096, device=dev, dtype=torch.bfloat16)
        torch.cuda.synchronize()
        ok = "?"
        try:
            with torch.cuda.stream(st):
                y = x @ x
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(st)
                for _ in range(10): y = x @ x
                e1.record(st)
            torch.cuda.synchronize()
            tf = 10 * 2 * 4096**3 / (e0.elapsed_time(e1) * 1e-3) / 1e12
            ref = x[:32].float() @ x.float()
            ok = f"gemm ok={bool(torch.allclose(y[:32].float(), ref, rtol=2e-2, atol=2.0))} {tf:.0f} TF/s"
        except Exception as ex:
            ok = "gemm FAILED " + str(ex)[:80]
        buf = torch.empty(4 << 30, dtype=torch.uint8, device=dev)
        res = torch.zeros(3, dtype=torch.int64, device=dev)
        rates = []
        for pat in (0, 1):
            dp.memtest_fill(buf, pat, 1, stream=parts[0].stream)
            e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            e[0].record(st); dp.memtest_fill(buf, pat, 1, stream=parts[0].stream); e[1].record(st)
            dp.memtest_verify(buf, pat, 1, result=res, stream=parts[0].stream); e[2].record(st)
            torch.cuda.synchronize()
            rates.append(f"{(4<<30)/(e[0].elapsed_time(e[1])*1e6):.0f}/{(4<<30)/(e[1].elapsed_time(e[2])*1e6):.0f}")
        print(f"pct={pct} n={n} fine={flags}: sms={sms} sum={sum(sms)} | {ok} | memtest fill/verify GB/s walk {rates[0]} rand {rates[1]}")
        del buf
        capi.lib.vgpu_sm_partition_destroy(parts, n)
