"""Tolerance metric shared by every parity test (oracle = test infrastructure).

BASELINE.json asks for "max relative error 1e-4 in fp32 mode and 2e-2 in bf16
mode".  A per-element relative error is unbounded near zero, so - as SURVEY.md
section 8(c) prescribes - the metric is  max|a-b| / max|b|  per tensor, b being the
oracle.  Every number quoted in tests, DESIGN.md and profiles/ uses this.
"""
import torch

TOL_F32 = 1e-4
TOL_BF16 = 2e-2


def max_rel_err(a, b) -> float:
    a = torch.as_tensor(a).detach().double().cpu()
    b = torch.as_tensor(b).detach().double().cpu()
    assert a.shape == b.shape, (a.shape, b.shape)
    if b.numel() == 0:
        return 0.0
    denom = b.abs().max().item()
    if denom == 0.0:
        denom = 1.0
    return (a - b).abs().max().item() / denom


def tol_for(dtype) -> float:
    return TOL_BF16 if dtype in (torch.bfloat16, "bf16", 1) else TOL_F32
