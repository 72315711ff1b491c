"""GPU diagnostic (not a test): tcgen05 GEMM, every operand-major combination, with an error map that localises
descriptor / swizzle mistakes (per 8-row x 64-column block)."""
import math
import sys
import os

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from solvingpapers_b200 import _lib as L
from solvingpapers_b200 import ops


def case(M, N, K, tA, tB, out_dtype=torch.bfloat16):
    g = torch.Generator().manual_seed(M * 7 + N * 3 + K + 2 * tA + tB)
    A = torch.randn((K, M) if tA else (M, K), generator=g).bfloat16()
    B = (torch.randn((N, K) if tB else (K, N), generator=g) / math.sqrt(K)).bfloat16()
    ref = (A.double().t() if tA else A.double()) @ (B.double().t() if tB else B.double())
    try:
        d = ops.gemm(A.cuda(), B.cuda(), transA=tA, transB=tB, out_dtype=out_dtype, impl=L.IMPL_TCGEN05)
        torch.cuda.synchronize()
    except Exception as ex:
        print(f"M{M} N{N} K{K} tA{int(tA)} tB{int(tB)}: EXCEPTION {ex}")
        return False
    err = (d.double().cpu() - ref).abs()
    rel = err.max().item() / ref.abs().max().item()
    ok = rel < 5e-3
    print(f"M{M} N{N} K{K} tA{int(tA)} tB{int(tB)} out={str(out_dtype)[6:]}: max_rel_err {rel:.3e} {'OK' if ok else 'FAIL'}")
    if not ok:
        # error map: rows grouped by 8, columns by 64
        em = err[: (M // 8) * 8, : (N // 64) * 64].reshape(M // 8, 8, N // 64, 64).amax(dim=(1, 3)) if M >= 8 and N >= 64 else err
        bad = (em > 5e-3 * ref.abs().max().item())
        print("  bad 8x64 blocks: %d of %d; first rows:" % (int(bad.sum()), bad.numel()))
        for r in range(min(bad.shape[0], 4)):
            print("   ", "".join("X" if b else "." for b in bad[r].tolist()))
        print("  got[0,:8] ", [round(v, 3) for v in d[0, :8].float().cpu().tolist()])
        print("  ref[0,:8] ", [round(v, 3) for v in ref[0, :8].tolist()])
        print("  got[1,:4] ", [round(v, 3) for v in d[1, :4].float().cpu().tolist()], " ref[1,:4]", [round(v, 3) for v in ref[1, :4].tolist()])
        print("  got[8,:4] ", [round(v, 3) for v in d[8, :4].float().cpu().tolist()], " ref[8,:4]", [round(v, 3) for v in ref[8, :4].tolist()])
        print("  got[0,64:68]", [round(v, 3) for v in d[0, 64:68].float().cpu().tolist()], " ref", [round(v, 3) for v in ref[0, 64:68].tolist()])
    return ok


if __name__ == "__main__":
    print(torch.cuda.get_device_name(0), L.load().spb_version())
    for (tA, tB) in ((False, True), (False, False), (True, True), (True, False)):
        for (M, N, K) in ((128, 128, 16), (128, 128, 64), (128, 256, 64), (128, 256, 256), (256, 512, 512)):
            case(M, N, K, tA, tB)
    # big enough for the CTA-pair kernel (>= 74 tiles of 256x256), all majors, ragged edges, fp32 out
    for (tA, tB) in ((False, True), (False, False), (True, True), (True, False)):
        case(2560, 2048, 512, tA, tB)
    case(2504, 1992, 520, False, True)
    case(2504, 1992, 520, True, False, torch.float32)
    case(256, 768, 16384, True, False, torch.float32)
    case(4096, 6144, 4096, False, False)
