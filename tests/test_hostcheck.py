"""CPU-only: the per-point math the CUDA kernels share (csrc/geom_core.cuh), compiled for the
host into libdc_hostcheck.so, checked against the reference's fp64 outputs (tests/golden)."""
import ctypes
import os

import numpy as np
import pytest
import torch

from helpers import load, t, row_scaled_err, check_frames, OPERATOR_ROW_REL
from oracle import geometry as og

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'deltaconv_b200', 'libdc_hostcheck.so')
pytestmark = pytest.mark.skipif(not os.path.exists(LIB), reason='host check library not built')


def _fp(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _build(g, normal, xb, yb, idx):
    hc = ctypes.CDLL(LIB)
    pos = np.ascontiguousarray(g.pos, dtype=np.float32)
    N, k = idx.shape
    ptr = og.cloud_ptr(t(g.batch), N).numpy().astype(np.int32)
    G = np.zeros((N, k, 2), np.float32)
    D = np.zeros((N, k, 2), np.float32)
    coords = np.zeros((N, k, 2), np.float32)
    weights = np.zeros((N, k), np.float32)
    mapping = np.zeros((N, k, 2, 2), np.float32)
    lam_shape = float(g.lam_shape)
    hc.hc_build_grad_div(_fp(pos), _fp(normal), _fp(xb), _fp(yb), _fp(idx), _fp(ptr), ctypes.c_int(len(ptr) - 1),
                         ctypes.c_int(N), ctypes.c_int(k), ctypes.c_float(float(g.width)), ctypes.c_float(float(g.lam)),
                         ctypes.c_float(max(lam_shape, 0.0)), ctypes.c_int(int(lam_shape >= 0)),
                         ctypes.c_int(int(g.normalized)), _fp(G), _fp(D), _fp(coords), _fp(weights), _fp(mapping))
    return G, D, coords, weights, mapping


@pytest.mark.parametrize('name', ['geom_normals_k10', 'geom_ragged_k20', 'geom_estimated_k16', 'geom_unnormalized_shape'])
def test_fit_core_matches_reference(name):
    g, _ = load(name)
    N, k = g.pos.shape[0], int(g.k)
    idx = og.ell_from_edge_index(t(g.ref64_edge_index), N).numpy().astype(np.int32)
    normal, xb, yb = (np.ascontiguousarray(getattr(g, f'ref64_{n}'), dtype=np.float32) for n in ('normal', 'xb', 'yb'))
    G, D, coords, weights, mapping = _build(g, normal, xb, yb, idx)
    (gr, gc, gv), (dr, dc, dv) = og.ell_to_coo(t(idx).long(), t(G), t(D))
    assert row_scaled_err(gv, t(g.ref64_grad_val), gr, 2 * N) < OPERATOR_ROW_REL
    assert row_scaled_err(dv, t(g.ref64_div_val), dr, N) < OPERATOR_ROW_REL
    E = N * k
    assert np.abs(coords.reshape(E, 2) - g.ref64_coords).max() < 1e-6
    assert np.abs(weights.reshape(E) - g.ref64_weights).max() < 1e-6
    assert np.abs(mapping.reshape(E, 2, 2) - g.ref64_mapping).max() < 2e-5


def test_sentinel_padding_matches_oracle():
    g, _ = load('geom_normals_k10')
    N, k = g.pos.shape[0], int(g.k)
    idx = og.ell_from_edge_index(t(g.ref64_edge_index), N).clone()
    idx[::7, -1] = N                        # pad the last slot of every 7th point
    idx[3, 4:] = N
    normal, xb, yb = (np.ascontiguousarray(getattr(g, f'ref64_{n}'), dtype=np.float32) for n in ('normal', 'xb', 'yb'))
    G, D, *_ = _build(g, normal, xb, yb, idx.numpy().astype(np.int32))
    Go, Do = og.fit_operators(t(g.pos, torch.float64), t(normal, torch.float64), t(xb, torch.float64),
                              t(yb, torch.float64), idx, t(g.batch), float(g.width), float(g.lam))
    scale = Go.abs().amax(dim=(1, 2), keepdim=True).clamp(min=1e-30)
    assert ((t(G).double() - Go).abs() / scale).max() < OPERATOR_ROW_REL
    scale = Do.abs().amax(dim=(1, 2), keepdim=True).clamp(min=1e-30)
    assert ((t(D).double() - Do).abs() / scale).max() < OPERATOR_ROW_REL
    assert (t(G)[idx == N] == 0).all() and (t(D)[idx == N] == 0).all()


def test_frame_estimation_core():
    g, _ = load('geom_estimated_k16')
    hc = ctypes.CDLL(LIB)
    pos = np.ascontiguousarray(g.pos, dtype=np.float32)
    N = pos.shape[0]
    idx10 = og.ell_from_edge_index(t(g.ref32_edge_index10), N).numpy().astype(np.int32)
    n, xb, yb = (np.zeros((N, 3), np.float32) for _ in range(3))
    hc.hc_estimate_basis(_fp(pos), _fp(idx10), ctypes.c_int(N), ctypes.c_int(10), _fp(pos), _fp(n), _fp(xb), _fp(yb))
    check_frames(t(n), t(xb), t(yb), 1e-5)
    # gauge-invariant agreement with the reference's SVD
    assert ((t(n).double() * t(g.ref64_normal)).sum(1) - 1).abs().max() < 1e-5
    assert ((t(xb).double() * t(g.ref64_xb)).sum(1).abs() - 1).abs().max() < 1e-4


def test_tangent_basis_core():
    g, _ = load('geom_normals_k10')
    hc = ctypes.CDLL(LIB)
    normal = np.ascontiguousarray(g.input_normal, dtype=np.float32)
    N = normal.shape[0]
    xb, yb = np.zeros((N, 3), np.float32), np.zeros((N, 3), np.float32)
    hc.hc_tangent_basis(_fp(normal), ctypes.c_int(N), _fp(xb), _fp(yb))
    assert np.abs(xb - g.ref64_xb).max() < 1e-6 and np.abs(yb - g.ref64_yb).max() < 1e-6


def test_knn_distance_formula_is_the_pinned_one():
    rng = np.random.default_rng(0)
    hc = ctypes.CDLL(LIB)
    q = rng.standard_normal(3).astype(np.float32)
    p = rng.standard_normal((4096, 3)).astype(np.float32)
    out = np.zeros(4096, np.float32)
    hc.hc_knn_dist2(_fp(q), _fp(p), ctypes.c_int(4096), _fp(out))
    d = p - q
    want = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]
    assert np.array_equal(out, want)
