import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1:
    import torch, vgpu_b200
    from vgpu_b200 import capi, datapath as dp
    dev = torch.device("cuda", 0)
    n = 2048; cb = capi.CHUNK_BYTES
    a = torch.empty(n * cb, dtype=torch.uint8, device=dev); b = torch.empty_like(a)
    host = torch.empty(512 * cb, dtype=torch.uint8).pin_memory()
    dp.memtest_fill(a, 1, 1)
    ref = dp.checksum(a, n)
    out = []
    for name, dst, tier, nn in (("local", b, 0, n), ("host", host, 2, 512)):
        for _ in range(2): ck = dp.spill(a, dst, nn, variant=2, tier=tier)
        best = 1e9
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); ck = dp.spill(a, dst, nn, variant=2, tier=tier); e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        ok = torch.equal(ck, ref[:nn]) and (torch.equal(b[:nn*cb], a[:nn*cb]) if tier == 0 else torch.equal(host[:nn*cb], a[:nn*cb].cpu()))
        out.append(f"{name} {nn*cb/(best*1e6):.0f} GB/s ok={ok}")
    print(sys.argv[1], " | ".join(out))
else:
    subprocess.run([sys.executable, __file__, "v1 default:"], env=dict(os.environ))
    names = {1: "16K x2 4cw", 2: "16K x2 2cw", 3: "32K x2 2cw", 4: "32K x2 4cw", 5: "32K x2 8cw", 6: "64K x2 4cw", 7: "48K x2 4cw"}
    for c in range(1, 8):
        for m in (4, 16):
            subprocess.run([sys.executable, __file__, f"ws cfg{c} {names[c]} grid x{m}:"], env=dict(os.environ, VGPU_TMA_WS=str(c), VGPU_TMA_GRID_MULT=str(m)))
