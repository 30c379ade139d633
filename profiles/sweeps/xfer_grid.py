import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1:
    import torch, vgpu_b200
    from vgpu_b200 import capi, datapath as dp
    dev = torch.device("cuda", 0)
    n = 4096; cb = capi.CHUNK_BYTES
    a = torch.empty(n * cb, dtype=torch.uint8, device=dev); b = torch.empty_like(a)
    dp.memtest_fill(a, 1, 1)
    out = []
    for nn in (64, 512, 4096):
        for v in (1,):
            for _ in range(2): dp.spill(a, b, nn, variant=v)
            best = 1e9
            for _ in range(5):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); dp.spill(a, b, nn, variant=v); e1.record(); torch.cuda.synchronize()
                best = min(best, e0.elapsed_time(e1))
            out.append(f"n={nn} v={v}: {nn*cb/(best*1e6):.0f}")
    print(sys.argv[1], " | ".join(out))
else:
    for u in ("2", "4", "8"):
        for cap in ("8", "16", "24", "32", "48", "64", "128", "100000"):
            env = dict(os.environ, VGPU_XFER_CTAS_PER_SM=cap, VGPU_XFER_U=u)
            subprocess.run([sys.executable, __file__, f"U={u} cap={cap}"], env=env)
