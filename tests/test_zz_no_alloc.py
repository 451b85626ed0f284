"""GPU test of dc_build_grad_div_static, the operator build for callers that cannot allocate
device memory (hosts with a fixed memory partition, loops that cannot afford cudaMalloc): its
scratch is a fixed array inside the library instead of a caller-provided workspace.

Called through the C ABI, twice in a row on different batches (the array is shared between calls
and keeps the previous call's contents — results must not depend on them), and compared with
dc_build_grad_div within the operator tolerance of the parity suite.  Collected last on purpose."""
import pytest
import torch

from helpers import OPERATOR_ROW_REL

pytestmark = pytest.mark.gpu
DEV = 'cuda'


def _inputs(sizes, k, cfg):
    import deltaconv_b200 as D
    data = D.utils.synthetic_batch(len(sizes), 0, config=cfg, sizes=sizes).to(DEV)
    graph = D.ops.knn(data.pos, k, data.batch, num_clouds=len(sizes), want_edge_index=False)
    xb, yb = D.ops.tangent_basis(data.norm)
    return data.pos, data.norm, xb, yb, graph


def _build(static, pos, normal, xb, yb, graph, normalized, lam, lam_shape):
    from deltaconv_b200 import _lib
    from deltaconv_b200._lib import call, ptr, query
    N, k, B = graph.N, graph.k, graph.B
    G = torch.full((N, k, 2), float('nan'), device=DEV)
    D_ = torch.full((N, k, 2), float('nan'), device=DEV)
    head = (ptr(pos), ptr(normal), ptr(xb), ptr(yb), ptr(graph.idx))
    tail = (ptr(graph.cloud_ptr), B, N, k, 1.0, lam, lam_shape if lam_shape is not None else 0.0, int(lam_shape is not None),
            int(normalized), ptr(G), ptr(D_))
    common = head + tail
    if static:
        torch.cuda.synchronize()
        # count allocation EVENTS (monotonic), not bytes in use: a garbage-collection pass may free unrelated
        # tensors of earlier tests while we are in here
        before = torch.cuda.memory_stats().get('allocation.all.allocated', 0)
        call('dc_build_grad_div_static', *common)
        assert torch.cuda.memory_stats().get('allocation.all.allocated', 0) == before     # nothing was allocated on our behalf
    else:
        nbytes = query('dc_build_ws_bytes', N, B)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=DEV)
        call('dc_build_grad_div', *head, None, *tail, None, None, None, None, ptr(ws), nbytes)       # no kNN distances, no deferred scale
    return G, D_


def _close(a, b):
    scale = b.abs().amax(dim=(1, 2), keepdim=True).clamp(min=1e-30)
    return not a.isnan().any() and ((a - b).abs() / scale).max().item() <= OPERATOR_ROW_REL


def test_static_scratch_build_matches_workspace_build_on_two_calls_in_a_row():
    first = _inputs([1024] * 8, 20, 31)
    second = _inputs([300, 77, 500], 12, 32)
    for inputs, normalized, lam, lam_shape in ((first, True, 1e-3, None), (second, False, 1e-2, 1e-3), (first, True, 1e-3, None)):
        G0, D0 = _build(False, *inputs, normalized, lam, lam_shape)
        G1, D1 = _build(True, *inputs, normalized, lam, lam_shape)
        assert _close(G1, G0) and _close(D1, D0)


def test_static_scratch_build_matches_the_fp64_oracle():
    """The same entry point against the CPU oracle directly (not only against the workspace build)."""
    from oracle import geometry as og
    pos, normal, xb, yb, graph = _inputs([256, 300, 77], 12, 34)
    G, D_ = _build(True, pos, normal, xb, yb, graph, True, 1e-3, None)
    batch = torch.repeat_interleave(torch.arange(3), torch.tensor([256, 300, 77]))
    Go, Do = og.fit_operators(pos.double().cpu(), normal.double().cpu(), xb.double().cpu(), yb.double().cpu(),
                              graph.idx.long().cpu(), batch, 1.0, 1e-3)
    assert _close(G.double().cpu(), Go) and _close(D_.double().cpu(), Do)


def test_static_scratch_build_rejects_batches_beyond_its_limit():
    from deltaconv_b200 import _lib
    pos, normal, xb, yb, graph = _inputs([64, 64], 8, 33)
    max_points = _lib.query('dc_build_static_max_points')
    max_clouds = _lib.query('dc_build_static_max_clouds')
    assert max_points == 262144 and max_clouds == 4096
    out = torch.empty(128, 8, 2, device=DEV)
    args = (_lib.ptr(pos), _lib.ptr(normal), _lib.ptr(xb), _lib.ptr(yb), _lib.ptr(graph.idx), _lib.ptr(graph.cloud_ptr))
    tail = (8, 1.0, 1e-3, 0.0, 0, 1, _lib.ptr(out), _lib.ptr(out))
    with pytest.raises(_lib.DeltaConvLibraryError, match='exceed the built-in scratch'):
        _lib.call('dc_build_grad_div_static', *args, 2, max_points + 1, *tail)
    with pytest.raises(_lib.DeltaConvLibraryError, match='exceed the built-in scratch'):
        _lib.call('dc_build_grad_div_static', *args, max_clouds + 1, 128, *tail)
