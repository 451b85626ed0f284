"""Shared comparison helpers for the parity suite (rules from SURVEY.md §8c)."""
import os
from types import SimpleNamespace

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')

# Tolerances (stated once, used everywhere):
OPERATOR_ROW_REL = 1e-5     # |ours - ref64| <= 1e-5 * max_row |ref64|  (north_star: operator entries rel 1e-5, row-scaled)
LOGIT_REL = 1e-4            # north_star: network logits rel 1e-4 in fp32
GRAD_REL = 2e-3             # parameter gradients through 4+ BN layers, relative to the tensor's max


def load(name):
    z = np.load(os.path.join(GOLDEN, name + '.npz'))
    return SimpleNamespace(**{k.replace('.', '__'): z[k] for k in z.files}), z


def t(a, dtype=None):
    out = torch.from_numpy(np.asarray(a))
    return out if dtype is None else out.to(dtype)


def row_scaled_err(ours, ref, rows, num_rows):
    """max over entries of |ours-ref| / max_{entries in the same row}|ref|."""
    ours, ref = ours.double(), ref.double()
    row_max = torch.zeros(num_rows, dtype=torch.float64).scatter_reduce(0, rows, ref.abs(), 'amax')
    return ((ours - ref).abs() / row_max[rows].clamp(min=1e-300)).max().item()


def rel_err(ours, ref):
    """max |ours-ref| / max |ref| (tensor-scaled)."""
    ours, ref = ours.double(), ref.double()
    return ((ours - ref).abs().max() / ref.abs().max().clamp(min=1e-300)).item()


def elem_err(ours, ref):
    """Per-element worst case max_i |ours_i - ref_i| / |ref_i| (entries below 1e-6 of the tensor's scale are measured against
    that floor).  Reported beside the tensor-scaled `rel_err`, which is the reading the tolerances are stated in: a logit that
    is small next to the largest one may be off by more than 1e-4 of ITSELF while the tensor-scaled error is within bounds."""
    ours, ref = ours.double(), ref.double()
    floor = 1e-6 * ref.abs().max().clamp(min=1e-300)
    return ((ours - ref).abs() / ref.abs().clamp(min=floor)).max().item()


def check_frames(normal, xb, yb, tol=1e-5):
    """orthonormal + right-handed (reference tests: test_grad_div_mls.py:12-24)."""
    for a in (normal, xb, yb):
        assert (a.norm(dim=1) - 1).abs().max() < tol
    assert (normal * xb).sum(1).abs().max() < tol
    assert (normal * yb).sum(1).abs().max() < tol
    assert (xb * yb).sum(1).abs().max() < tol
    assert (torch.linalg.cross(xb, yb, dim=-1) - normal).abs().max() < 10 * tol


def report(name, value):
    """Append a measured error to gpurun_out/parity_report.txt (margins, not just pass/fail)."""
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'gpurun_out')
    try:
        os.makedirs(out, exist_ok=True)
        with open(os.path.join(out, 'parity_report.txt'), 'a') as f:
            f.write(f'{name}\t{value:.3e}\n')
    except OSError:
        pass
    return value
