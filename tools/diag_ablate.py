"""GPU timing experiment (not a test): flash backward with parts switched off (SPB_FB_ABLATE), C3 shape.
The switches exist only in a library built with `tools/build_variant.sh ablate -DSPB_FB_ABLATE`; point SPB_LIB_PATH at it."""
import math, os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1:
    import torch
    from solvingpapers_b200 import _lib as L, ops
    B, T, Hq, Hkv, hd = 4, 2048, 32, 8, 128
    g = torch.Generator(device="cuda").manual_seed(1)
    mk = lambda h: torch.randn(B, T, h, hd, generator=g, device="cuda").bfloat16()
    q, k, v, do = mk(Hq), mk(Hkv), mk(Hkv), mk(Hq)
    o, lse = ops.flash_fwd(q, k, v, True, 1 / math.sqrt(hd), impl=L.IMPL_TCGEN05)
    fn = lambda: ops.flash_bwd(q, k, v, o, do, lse, True, 1 / math.sqrt(hd), impl=L.IMPL_TCGEN05)
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): fn()
    e1.record(); torch.cuda.synchronize()
    print(f"ablate={os.environ.get('SPB_FB_ABLATE','0'):>3s}: {e0.elapsed_time(e1)/10:.3f} ms")
else:
    for m in (0, 1, 2, 3, 4, 8, 12, 15, 16, 31, 32, 47, 63):
        subprocess.run([sys.executable, __file__, "x"], env=dict(os.environ, SPB_FB_ABLATE=str(m)))
