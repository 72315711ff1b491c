"""GPU timing experiment (not a test): flash forward with parts switched off (SPB_FA_ABLATE), C3 shape.
bits: 1 no exp (MUFU), 2 no row max, 4 no P store, 8 no PV MMA, 16 no QK MMA, 32 softmax = protocol only, 64 no S load."""
import math, os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1 and sys.argv[1] == "x":
    import torch
    from solvingpapers_b200 import _lib as L, ops
    B, T, Hq, Hkv, hd = 4, 2048, 32, 8, 128
    g = torch.Generator(device="cuda").manual_seed(1)
    mk = lambda h: torch.randn(B, T, h, hd, generator=g, device="cuda").bfloat16()
    q, k, v = mk(Hq), mk(Hkv), mk(Hkv)
    causal = os.environ.get("FA_CAUSAL", "1") == "1"
    fn = lambda: ops.flash_fwd(q, k, v, causal, 1 / math.sqrt(hd), impl=L.IMPL_TCGEN05)
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): fn()
    e1.record(); torch.cuda.synchronize()
    print(f"causal={int(causal)} ablate={os.environ.get('SPB_FA_ABLATE','0'):>3s}: {e0.elapsed_time(e1)/10:.4f} ms")
else:
    masks = [int(a) for a in sys.argv[1:]] or (0, 1, 2, 4, 3, 7, 64, 71, 32, 8, 16, 24, 56, 39)
    for c in ("1", "0"):
        for m in masks:
            subprocess.run([sys.executable, __file__, "x"], env=dict(os.environ, SPB_FA_ABLATE=str(m), FA_CAUSAL=c))
