"""GPU diagnostic (not a test): tcgen05 flash attention forward/backward vs a torch fp32 restatement on the GPU's own
inputs (checker runs on the GPU in fp32 to keep big shapes fast); error localisation by 128-row tile and head."""
import math
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from solvingpapers_b200 import _lib as L
from solvingpapers_b200 import ops


def ref_attn(q, k, v, causal, scale):
    B, T, Hq, hd = q.shape
    n_rep = Hq // k.shape[2]
    qf, kf, vf = (t.float() for t in (q, k, v))
    kf = kf.repeat_interleave(n_rep, dim=2)
    vf = vf.repeat_interleave(n_rep, dim=2)
    s = torch.einsum("bthd,bshd->bhts", qf, kf) * scale
    if causal:
        s = s + torch.triu(torch.full((T, T), float("-inf"), device=q.device), diagonal=1)
    lse = torch.logsumexp(s, dim=-1)
    o = torch.einsum("bhts,bshd->bthd", torch.softmax(s, dim=-1), vf)
    return o, lse


def rel(a, b):
    return ((a.float() - b.float()).abs().max() / b.float().abs().max().clamp_min(1e-30)).item()


def case(B, T, Hq, Hkv, hd, causal, bwd=False, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    mk = lambda h: torch.randn(B, T, h, hd, generator=g, device="cuda").bfloat16()
    q, k, v, do = mk(Hq), mk(Hkv), mk(Hkv), mk(Hq)
    scale = 1.0 / math.sqrt(hd)
    tag = f"B{B} T{T} Hq{Hq} Hkv{Hkv} hd{hd} causal{int(causal)}"
    try:
        o, lse = ops.flash_fwd(q, k, v, causal, scale, impl=L.IMPL_TCGEN05)
        torch.cuda.synchronize()
    except Exception as ex:
        print(f"{tag}: FWD EXCEPTION {ex}")
        return False
    qr, kr, vr = (t.float().requires_grad_(True) for t in (q, k, v))
    oref, lseref = ref_attn(qr, kr, vr, causal, scale)
    eo, el = rel(o, oref), rel(lse, lseref)
    ok = eo < 1e-2 and el < 1e-3 and bool(torch.isfinite(o.float()).all())
    print(f"{tag}: fwd o {eo:.3e} lse {el:.3e} {'OK' if ok else 'FAIL'}")
    if not ok:
        err = (o.float() - oref).abs().amax(dim=-1)  # [B,T,Hq]
        nt = (T + 127) // 128
        for hh in range(min(Hq, 2)):
            row = []
            for t in range(nt):
                row.append(f"{err[0, t * 128:(t + 1) * 128, hh].max().item():.2e}")
            print(f"   head {hh} per-q-tile max abs err: {' '.join(row)}")
        print("   o[0,0,0,:6]  ", o[0, 0, 0, :6].float().tolist(), "\n   ref          ", oref[0, 0, 0, :6].tolist())
        print("   o[0,T-1,0,:6]", o[0, T - 1, 0, :6].float().tolist(), "\n   ref          ", oref[0, T - 1, 0, :6].tolist())
        print("   lse[0,0,:4]", lse[0, 0, :4].tolist(), " ref", lseref[0, 0, :4].tolist())
        return False
    if bwd:
        try:
            dq, dk, dv = ops.flash_bwd(q, k, v, o, do, lse, causal, scale, impl=L.IMPL_TCGEN05)
            torch.cuda.synchronize()
        except Exception as ex:
            print(f"{tag}: BWD EXCEPTION {ex}")
            return False
        oref.backward(do.float())
        e = {"dq": rel(dq, qr.grad), "dk": rel(dk, kr.grad), "dv": rel(dv, vr.grad)}
        okb = all(x < 2e-2 for x in e.values())
        print(f"{tag}: bwd " + " ".join(f"{k_} {x:.3e}" for k_, x in e.items()) + (" OK" if okb else " FAIL"))
        if not okb:
            for name, got, want in (("dq", dq, qr.grad), ("dk", dk, kr.grad), ("dv", dv, vr.grad)):
                err = (got.float() - want).abs().amax(dim=-1)
                nt = (T + 127) // 128
                print(f"   {name} head0 per-tile max abs err:", " ".join(f"{err[0, t * 128:(t + 1) * 128, 0].max().item():.2e}" for t in range(nt)),
                      f"| max|ref| {want.abs().max().item():.2e}")
                print(f"   {name}[0,0,0,:4]", got[0, 0, 0, :4].float().tolist(), "ref", want[0, 0, 0, :4].tolist())
        return okb
    return True


def timeit(B, T, Hq, Hkv, hd, causal, bwd):
    g = torch.Generator(device="cuda").manual_seed(1)
    mk = lambda h: torch.randn(B, T, h, hd, generator=g, device="cuda").bfloat16()
    q, k, v, do = mk(Hq), mk(Hkv), mk(Hkv), mk(Hq)
    scale = 1.0 / math.sqrt(hd)
    o, lse = ops.flash_fwd(q, k, v, causal, scale, impl=L.IMPL_TCGEN05)
    m = (T + 1) / (2 * T) if causal else 1.0
    fl = 4.0 * B * Hq * T * T * hd * m
    for name, fn, f in (("fwd", lambda: ops.flash_fwd(q, k, v, causal, scale, impl=L.IMPL_TCGEN05), fl),
                        ("bwd", lambda: ops.flash_bwd(q, k, v, o, do, lse, causal, scale, impl=L.IMPL_TCGEN05), 2.5 * fl)):
        if name == "bwd" and not bwd:
            continue
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print(f"time {name} B{B} T{T} Hq{Hq} Hkv{Hkv} hd{hd} causal{int(causal)}: {ms:.3f} ms  {f / ms / 1e9:.1f} TFLOP/s")


if __name__ == "__main__":
    bwd = "--bwd" in sys.argv
    print(torch.cuda.get_device_name(0), L.load().spb_version(), "bwd" if bwd else "fwd only")
    ok = True
    ok &= case(1, 128, 1, 1, 128, False, bwd)
    ok &= case(1, 128, 1, 1, 128, True, bwd)
    ok &= case(1, 256, 1, 1, 128, False, bwd)
    ok &= case(1, 512, 2, 1, 128, True, bwd)
    ok &= case(2, 1024, 4, 2, 128, True, bwd)
    ok &= case(1, 128, 1, 1, 64, False, bwd)
    ok &= case(2, 384, 4, 2, 64, True, bwd)
    ok &= case(2, 197, 3, 3, 64, False, bwd)
    ok &= case(1, 200, 2, 2, 128, True, bwd)
    ok &= case(1, 2048, 8, 2, 128, True, bwd)
    ok &= case(1, 128, 1, 1, 256, False, bwd)
    ok &= case(2, 384, 2, 1, 256, True, bwd)
    ok &= case(1, 1024, 4, 1, 256, True, bwd)
    print("ALL OK" if ok else "SOME FAILED")
    if ok:
        timeit(4, 2048, 32, 8, 128, True, bwd)
        timeit(4, 2048, 32, 8, 128, False, bwd)
        timeit(16, 1024, 12, 12, 64, False, bwd)
        timeit(64, 1024, 1, 1, 256, True, bwd)
        timeit(4, 4096, 8, 1, 256, True, bwd)
