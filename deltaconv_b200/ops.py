"""Host side of the CUDA operators: graph/operator containers and `torch.autograd.Function`
wrappers around the C ABI (include/deltaconv_b200.h).  PyTorch supplies device memory, the
current stream and the autograd tape; all arithmetic happens in the library.

Every function here requires contiguous fp32 CUDA tensors and raises otherwise — there is no
CPU or eager-PyTorch fallback.
"""
import os

import torch

from . import _lib
from ._lib import call, ptr, query, rows_ld

_F32, _I32, _I64 = torch.float32, torch.int32, torch.int64

_workspaces = {}
_CHECK_LAYOUT = bool(int(os.environ.get('DC_CHECK_LAYOUT', '0')))


def workspace(nbytes, device, slot=0):
    """Cached scratch buffer (one per device and slot, grows on demand).  All kernels that use slot 0 are enqueued on the
    current stream, so successive users are ordered; slot 1 belongs to the side stream of `forked_backward`."""
    key = (device.type, device.index, slot)
    buf = _workspaces.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(int(nbytes), 1 << 20), dtype=torch.uint8, device=device)
        _workspaces[key] = buf
    return buf


def live_workspaces():
    """The scratch buffers currently installed (one per device).  A captured CUDA graph has their addresses baked in,
    so whoever owns the graph keeps this list: a buffer that `workspace()` later replaces by a larger one then stays
    allocated instead of being handed to other tensors while replays still write into it."""
    return list(_workspaces.values())


def _f32(t, name='tensor'):
    if not (torch.is_tensor(t) and t.is_cuda and t.dtype == _F32):
        raise _lib.DeltaConvLibraryError(
            f"{name}: expected an fp32 CUDA tensor, got {getattr(t, 'dtype', type(t))} on "
            f"{getattr(t, 'device', '?')} (deltaconv_b200 has no CPU path)")
    return t


def _mat(t, name='tensor'):
    """2-D fp32 CUDA tensor with unit column stride -> (tensor, ld)."""
    _f32(t, name)
    if t.dim() != 2:
        raise _lib.DeltaConvLibraryError(f"{name}: expected a 2-D tensor, got shape {tuple(t.shape)}")
    return rows_ld(t)


# ------------------------------------------------------------------------------------------------
# clouds and graphs
# ------------------------------------------------------------------------------------------------

def cloud_layout(batch, num_points, device, num_clouds=None, max_points=None):
    """(cloud_ptr int32 [B+1], B, size of the largest cloud) for a sorted PyG batch vector; cached on the
    tensor object.  `batch=None` means one cloud.  Without the hints the two integers are read back from
    the device (one sync, as the reference's scatter calls do).

    Contract of the cache and the hints: the entry is keyed on the tensor's in-place version counter, so writing a new
    layout into the same `batch` tensor through torch (copy_, index assignment, ...) invalidates it.  Under CUDA-graph replay
    nothing is recomputed at all: every replayed batch must have the captured layout (GraphedTrainStep / GraphedInference).
    `num_clouds` / `max_points` are trusted when given: `num_clouds` must equal batch[-1] + 1 and `max_points` must bound the
    largest cloud (a smaller `num_clouds` silently drops clouds).  Set DC_CHECK_LAYOUT=1 to verify `num_clouds` (one sync)."""
    if batch is None:
        return torch.tensor([0, num_points], dtype=_I32, device=device), 1, num_points
    cached = getattr(batch, '_dc_cloud', None)
    if cached is not None and cached[3] == num_points and cached[4] == batch._version:
        return cached[0], cached[1], cached[2]
    if not batch.is_cuda or batch.dtype != _I64:
        raise _lib.DeltaConvLibraryError("batch must be an int64 CUDA tensor")
    batch = batch.contiguous()
    B = int(num_clouds) if num_clouds is not None else (int(batch[-1].item()) + 1 if num_points else 1)
    if num_clouds is not None and num_points and _CHECK_LAYOUT and int(batch[-1].item()) + 1 != B:
        raise _lib.DeltaConvLibraryError(f"cloud_layout: num_clouds={B} but batch[-1] + 1 = {int(batch[-1].item()) + 1}")
    cp = torch.empty(B + 1, dtype=_I32, device=device)
    call('dc_cloud_ptr', ptr(batch), num_points, B, ptr(cp))
    maxp = int(max_points) if max_points is not None else (int((cp[1:] - cp[:-1]).max().item()) if num_points else 0)
    try:
        batch._dc_cloud = (cp, B, maxp, num_points, batch._version)
    except Exception:
        pass
    return cp, B, maxp


def cloud_ptr_of(batch, num_points, device, num_clouds=None):
    cp, B, _ = cloud_layout(batch, num_points, device, num_clouds)
    return cp, B


class EllGraph:
    """Fixed-k neighbourhood graph in ELL form plus, lazily, its incoming-edge CSR."""

    def __init__(self, idx, cloud_ptr=None, num_clouds=None, max_cloud=0):
        self.idx = idx
        self.N, self.k = idx.shape
        self.cloud_ptr, self.B, self.max_cloud = cloud_ptr, num_clouds, int(max_cloud or 0)
        self._transpose = None
        self._edge_index = None
        self.d2 = self.d2_of = None

    def clouds(self):
        """(cloud_ptr pointer or None, B, largest cloud) as the SpMM entry points take them."""
        if self.cloud_ptr is None:
            return None, 0, 0
        return self.cloud_ptr.data_ptr(), self.B, self.max_cloud

    def transpose(self):
        if self._transpose is None:
            dev = self.idx.device
            in_ptr = torch.empty(self.N + 1, dtype=_I32, device=dev)
            in_edge = torch.empty(max(self.N * self.k, 1), dtype=_I32, device=dev)
            nbytes = query('dc_ell_transpose_ws_bytes', self.N)
            ws = workspace(nbytes, dev)
            call('dc_ell_transpose', ptr(self.idx), self.N, self.k, ptr(in_ptr), ptr(in_edge), ptr(ws), nbytes)
            self._transpose = (in_ptr, in_edge)
        return self._transpose

    def edge_index(self):
        if self._edge_index is None:
            ei = torch.empty(2, self.N * self.k, dtype=_I64, device=self.idx.device)
            call('dc_edge_index_from_ell', ptr(self.idx), self.N, self.k, ptr(ei))
            ei._dc_graph = self
            self._edge_index = ei
        return self._edge_index


KNN_GRID_FROM = 1536      # points in the largest cloud from which the grid search replaces the brute-force scan (measured crossover)
KNN_GRID_FROM_K20 = 1000  # ... for k >= 20: with ~6 points per cell the grid already wins at 1024-point clouds (130 vs 145 us per 32 clouds at k = 20, 173 vs 265 at k = 40)
KNN_METHOD = 'auto'       # 'auto' | 'brute' | 'grid' (tests and benchmarks force one)


def knn(pos, k, batch=None, num_clouds=None, want_edge_index=True, max_points=None, want_d2=True):
    """ELL kNN graph of `pos` [N,3] per cloud (self included, ascending (d2, index)).  The squared distances ride along as
    `graph.d2` [N,k] (+inf in padded slots): the operator build takes the mean edge lengths from them."""
    pos = _f32(pos, 'pos').contiguous()
    if pos.dim() != 2 or pos.size(1) != 3:
        raise _lib.DeltaConvLibraryError(f"knn: expected pos of shape [N,3], got {tuple(pos.shape)}")
    N = pos.size(0)
    cp, B, maxp = cloud_layout(batch, N, pos.device, num_clouds, max_points)
    idx = torch.empty(N, k, dtype=_I32, device=pos.device)
    ei = torch.empty(2, N * k, dtype=_I64, device=pos.device) if want_edge_index else None
    d2 = torch.empty(N, k, dtype=_F32, device=pos.device) if want_d2 else None
    method = KNN_METHOD if KNN_METHOD != 'auto' else ('grid' if maxp > (KNN_GRID_FROM_K20 if k >= 20 else KNN_GRID_FROM) else 'brute')
    if method == 'grid':                 # large clouds: uniform grid + ring search, same exact result (dc_knn_grid)
        nb = query('dc_knn_grid_ws_bytes', N, B)
        ws = workspace(nb, pos.device)
        call('dc_knn_grid', ptr(pos), ptr(cp), B, N, k, ptr(idx), ptr(ei), ptr(d2), ptr(ws), nb)
    else:
        call('dc_knn', ptr(pos), ptr(cp), B, N, k, ptr(idx), ptr(ei), ptr(d2))
    graph = EllGraph(idx, cp, B, maxp)
    graph.d2, graph.d2_of = d2, pos              # distances belong to these positions (build_operators checks the tensor)
    if ei is not None:
        ei._dc_graph = graph
        graph._edge_index = ei
    return graph


def spatial_order(pos, batch=None, num_clouds=None):
    """Permutation that orders the points of every cloud along a Morton curve (clouds stay where they are, so a sorted
    `batch` vector is unchanged by it).  For clouds of 10^5 points and more the operators built on `pos[perm]` gather from
    nearby rows: 2x on the SpMMs of a 2^20-point cloud.  The sort itself is torch's (a pre-transform, not on the hot path)."""
    pos = _f32(pos, 'pos').contiguous()
    N = pos.size(0)
    cp, B, _ = cloud_layout(batch, N, pos.device, num_clouds)
    keys = torch.empty(N, dtype=_I64, device=pos.device)
    ws = workspace(B * 24, pos.device)
    call('dc_morton_keys', ptr(pos), ptr(cp), B, N, ptr(keys), ptr(ws), B * 24)
    return torch.argsort(keys, stable=True)


def geodesic_fps(pos, n_samples, batch=None, num_clouds=None, start=None, max_points=None, return_dist=False, ordered=True):
    """Geodesic farthest-point sampling of every cloud at once (dc_geodesic_fps; reference sampling.cpp:21-81 does one cloud
    per call on one CPU thread).  `pos` [N,3] CUDA, fp32 or fp64 (the reference computes in fp64 whatever it is given);
    `start` [B] = first sample of every cloud as an index inside the cloud, drawn uniformly when None like the reference.
    `ordered`: visit the points of every cloud along a Morton curve (the neighbour search then skips most of the cloud per
    query; the result is the same either way).
    Returns [B, n_samples] int64 indices inside each cloud (and the distance of every point to its nearest sample)."""
    if not (torch.is_tensor(pos) and pos.is_cuda and pos.dtype in (_F32, torch.float64)):
        raise _lib.DeltaConvLibraryError("geodesic_fps: expected an fp32 or fp64 CUDA tensor (deltaconv_b200 has no CPU path)")
    if pos.dim() != 2 or pos.size(1) != 3:
        raise _lib.DeltaConvLibraryError(f"geodesic_fps: expected pos of shape [N,3], got {tuple(pos.shape)}")
    N = pos.size(0)
    cp, B, maxp = cloud_layout(batch, N, pos.device, num_clouds, max_points)
    pos64 = pos.to(torch.float64).contiguous()
    if start is None:
        sizes = (cp[1:] - cp[:-1]).to(torch.float64)
        start = (torch.rand(B, device=pos.device, dtype=torch.float64) * sizes).floor().clamp_(min=0).to(_I32)
        start = torch.minimum(start, (sizes - 1).clamp_(min=0).to(_I32))
    else:
        start = torch.as_tensor(start, device=pos.device).to(_I32).reshape(B).contiguous()
    samples = torch.empty(B, n_samples, dtype=_I32, device=pos.device)
    dist = torch.empty(N, dtype=torch.float64, device=pos.device) if return_dist else None
    nb = query('dc_geodesic_fps_ws_bytes', N)
    ws = workspace(nb, pos.device)
    order = None
    if ordered and maxp > 2 * 64:
        # visiting order: Morton curve inside every cloud.  Any order gives the same samples, so the sort key may be coarse:
        # with up to 1023 clouds, (cloud, 21 Morton bits) fits 32 bits and torch's radix sort does half the passes on half
        # the bytes of the 64-bit key (cloud << 30 | 30 bits)
        keys = torch.empty(N, dtype=_I64, device=pos.device)
        kws = workspace(B * 24, pos.device)
        call('dc_morton_keys', ptr(pos.float().contiguous()), ptr(cp), B, N, ptr(keys), ptr(kws), B * 24)
        if B <= 1023:
            keys = (((keys >> 30) << 21) | ((keys & ((1 << 30) - 1)) >> 9)).to(_I32)
        order = torch.argsort(keys, stable=True).to(_I32)
    call('dc_geodesic_fps', ptr(pos64), ptr(cp), B, N, min(maxp, N), int(n_samples), ptr(order), ptr(start), ptr(samples), ptr(dist),
         ptr(ws), nb)
    samples = samples.long()
    return (samples, dist) if return_dist else samples


def graph_of(edge_index, num_points, strict=True):
    """EllGraph behind a [2,E] int64 edge_index: the one `knn_graph` attached, else a conversion
    under the reference's own assumption (rows grouped by centre, same k everywhere —
    grad_div_mls.py:24-25,224).  A graph that does not have that form raises, or returns None with strict=False
    (DeltaConv then takes its generic path)."""
    if isinstance(edge_index, EllGraph):
        return edge_index
    g = getattr(edge_index, '_dc_graph', None)
    if g is not None and g.N == num_points:
        return g
    if isinstance(edge_index, (tuple, list)):
        edge_index = torch.stack(list(edge_index))
    if not (edge_index.is_cuda and edge_index.dtype == _I64 and edge_index.dim() == 2 and edge_index.size(0) == 2):
        raise _lib.DeltaConvLibraryError("edge_index must be a [2,E] int64 CUDA tensor")
    E = edge_index.size(1)
    if num_points == 0 or E % num_points:
        if not strict:
            return None
        raise _lib.DeltaConvLibraryError(f"edge_index with {E} edges is not a fixed-k graph on {num_points} points")
    k = E // num_points
    ei = edge_index.contiguous()
    # row 0 must hold the centres, grouped: point i owns edges [i*k, (i+1)*k).  (A `source_to_target` graph, or any other
    # layout, would otherwise be read as a degenerate ELL without complaint.)
    centres = torch.arange(num_points, device=ei.device).repeat_interleave(k)
    if not torch.equal(ei[0], centres):
        if not strict:
            return None
        raise _lib.DeltaConvLibraryError("edge_index is not grouped by centre with the same k for every point "
                                         "(expected flow='target_to_source' kNN layout, grad_div_mls.py:24-25)")
    idx = torch.empty(num_points, k, dtype=_I32, device=ei.device)
    call('dc_ell_from_edge_index', ptr(ei), num_points, k, ptr(idx))
    g = EllGraph(idx)
    try:
        edge_index._dc_graph = g
    except Exception:
        pass
    return g


# ------------------------------------------------------------------------------------------------
# frames and operator build
# ------------------------------------------------------------------------------------------------

def tangent_basis(normal):
    normal = _f32(normal, 'normal').contiguous()
    xb, yb = torch.empty_like(normal), torch.empty_like(normal)
    call('dc_tangent_basis', ptr(normal), normal.size(0), ptr(xb), ptr(yb))
    return xb, yb


def estimate_frames(pos, graph, orientation=None):
    pos = _f32(pos, 'pos').contiguous()
    if orientation is not None:
        orientation = _f32(orientation, 'orientation').contiguous()
    n, xb, yb = (torch.empty_like(pos) for _ in range(3))
    call('dc_estimate_basis', ptr(pos), ptr(graph.idx), graph.N, graph.k, ptr(orientation), ptr(n), ptr(xb), ptr(yb))
    return n, xb, yb


def build_operators(pos, normal, xb, yb, graph, cloud_ptr, B, kernel_width=1.0, lam=1e-3, normalized=True,
                    lam_shape=None, debug=False, deferred=None):
    """G, D [N,k,2] of dc_build_grad_div (normalised entries).  With `deferred` given (True / False) the return is
    (G, D, point_scale): deferred=True leaves the entries un-normalised and returns the factor [N] the operator applications
    multiply their rows with; deferred=False normalises in place and returns None as the factor."""
    pos, normal, xb, yb = (_f32(t, n).contiguous() for t, n in ((pos, 'pos'), (normal, 'normal'), (xb, 'x_basis'), (yb, 'y_basis')))
    N, k = graph.N, graph.k
    dev = pos.device
    G = torch.empty(N, k, 2, dtype=_F32, device=dev)
    D = torch.empty(N, k, 2, dtype=_F32, device=dev)
    dbg = [None, None, None]
    if debug:
        dbg = [torch.empty(N, k, 2, dtype=_F32, device=dev), torch.empty(N, k, dtype=_F32, device=dev),
               torch.empty(N, k, 2, 2, dtype=_F32, device=dev)]
    mode = (2 if deferred else 1) if normalized else 0
    scale = torch.empty(N, dtype=_F32, device=dev) if mode == 2 else None
    # the kNN's squared distances, if the graph was built on these very positions
    d2 = graph.d2 if (graph.d2 is not None and graph.d2_of is not None and graph.d2_of.data_ptr() == pos.data_ptr()
                      and graph.d2_of._version == pos._version) else None
    nbytes = query('dc_build_ws_bytes', N, B)
    ws = workspace(nbytes, dev)
    call('dc_build_grad_div', ptr(pos), ptr(normal), ptr(xb), ptr(yb), ptr(graph.idx), ptr(d2), ptr(cloud_ptr), B, N, k,
         float(kernel_width), float(lam), float(lam_shape if lam_shape is not None else 0.0),
         int(lam_shape is not None), mode, ptr(G), ptr(D), ptr(scale), ptr(dbg[0]), ptr(dbg[1]), ptr(dbg[2]),
         ptr(ws), nbytes)
    if deferred is not None:
        return (G, D, scale, dbg) if debug else (G, D, scale)
    return (G, D, dbg) if debug else (G, D)


# ------------------------------------------------------------------------------------------------
# operator application
# ------------------------------------------------------------------------------------------------

class _GradApply(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, op):
        x, ldx = _mat(x, 'x')
        g = op.graph
        if x.size(0) != g.N:
            raise _lib.DeltaConvLibraryError(f"grad @ x: x has {x.size(0)} rows, operator has {g.N} columns")
        out = torch.empty(2 * g.N, x.size(1), dtype=_F32, device=x.device)
        call('dc_grad_fwd', ptr(g.idx), ptr(op.G_raw), g.N, g.k, x.size(1), ptr(x), ldx, ptr(out), out.size(1), *op.tail())
        ctx.op = op
        return out

    @staticmethod
    def backward(ctx, dv):
        op, g = ctx.op, ctx.op.graph
        dv, ld = _mat(dv, 'grad_output')
        in_ptr, in_edge = g.transpose()
        dx = torch.empty(g.N, dv.size(1), dtype=_F32, device=dv.device)
        call('dc_grad_bwd', ptr(in_ptr), ptr(in_edge), ptr(op.G_raw), g.N, g.k, dv.size(1), ptr(dv), ld, ptr(dx), dx.size(1), *op.tail())
        return dx, None


class _DivApply(torch.autograd.Function):
    @staticmethod
    def forward(ctx, v, op):
        v, ldv = _mat(v, 'v')
        g = op.graph
        if v.size(0) != 2 * g.N:
            raise _lib.DeltaConvLibraryError(f"div @ v: v has {v.size(0)} rows, operator has {2 * g.N} columns")
        C = v.size(1)
        out = torch.empty(g.N, C, dtype=_F32, device=v.device)
        call('dc_divcurlnorm_fwd', ptr(g.idx), ptr(op.D_raw), g.N, g.k, C, ptr(v), ldv, ptr(out), C, None, 0, None, 0, None, 0, None, 0, *op.tail())
        ctx.op = op
        return out

    @staticmethod
    def backward(ctx, dd):
        op, g = ctx.op, ctx.op.graph
        dd, ld = _mat(dd, 'grad_output')
        in_ptr, in_edge = g.transpose()
        C = dd.size(1)
        dv = torch.empty(2 * g.N, C, dtype=_F32, device=dd.device)
        call('dc_divcurlnorm_bwd', ptr(in_ptr), ptr(in_edge), ptr(op.D_raw), g.N, g.k, C, ptr(dd), ld, None, 0, None, 0,
             None, 0, ptr(dv), C, *op.tail())
        return dv, None


class EllOperator:
    """What `build_grad_div` returns in place of a torch_sparse.SparseTensor: supports `.size(i)`,
    `.sizes()` and `@ dense` (differentiable w.r.t. the dense operand), like the reference's
    operators are used (deltanet_base.py:78, deltaconv.py:57,66, operators.py:23-46).

    Storage: the kernels read `G_raw` / `D_raw` [N,k,2] and, when `scale` [N] is given, multiply their result rows by it (the
    per-cloud normalisation of grad_div_mls.py:258-260, deferred: dc_build_grad_div normalized = 2).  `G` / `D` are the
    reference's entries (normalised), materialised on first access for inspection, `coo()` and `to_dense()`."""

    def __init__(self, kind, graph, G, D, scale=None):
        self.kind, self.graph, self.G_raw, self.D_raw, self.scale = kind, graph, G, D, scale
        self._entries = None

    def _normalised(self):
        if self.scale is None:
            return self.G_raw, self.D_raw
        if self._entries is None:
            f = self.scale[:, None, None]
            self._entries = (self.G_raw * f, self.D_raw * f)
        return self._entries

    @property
    def G(self):
        return self._normalised()[0]

    @property
    def D(self):
        return self._normalised()[1]

    def tail(self):
        """(row_scale pointer, cloud_ptr pointer, B, largest cloud): the trailing arguments of the operator applications."""
        return (ptr(self.scale),) + self.graph.clouds()

    def size(self, dim):
        N = self.graph.N
        return (2 * N, N)[dim] if self.kind == 'grad' else (N, 2 * N)[dim]

    def sizes(self):
        return [self.size(0), self.size(1)]

    def __matmul__(self, dense):
        return _GradApply.apply(dense, self) if self.kind == 'grad' else _DivApply.apply(dense, self)

    matmul = __matmul__

    def coo(self):
        """(row, col, value) sorted by (row, col) — the reference's SparseTensor storage order
        (grad_div_mls.py:253-255,263,273-275).  For parity tests and interoperability."""
        g = self.graph
        N, k = g.N, g.k
        idx = g.idx.long()
        i = torch.arange(N, device=idx.device)[:, None].expand(N, k)
        valid = (idx != N).reshape(-1).repeat_interleave(2)
        if self.kind == 'grad':
            row = torch.stack([2 * i, 2 * i + 1], -1).reshape(-1)
            col = torch.stack([idx, idx], -1).reshape(-1)
            val, ncol = self.G.reshape(-1), N
        else:
            row = torch.stack([i, i], -1).reshape(-1)
            col = torch.stack([2 * idx, 2 * idx + 1], -1).reshape(-1)
            val, ncol = self.D.reshape(-1), 2 * N
        row, col, val = row[valid], col[valid], val[valid]
        perm = torch.argsort(row * ncol + col, stable=True)
        return row[perm], col[perm], val[perm]

    def to_dense(self):
        row, col, val = self.coo()
        out = torch.zeros(self.size(0), self.size(1), dtype=_F32, device=val.device)
        out.index_put_((row, col), val, accumulate=True)
        return out

    def __repr__(self):
        return f'EllOperator({self.kind}, N={self.graph.N}, k={self.graph.k})'


class _ScalarCat(torch.autograd.Function):
    """[x | div v | curl v | |v|] (deltaconv.py:57) in one pass over v (x is copied into its column block by
    the same kernel); the second output is the [div|curl] column block, a view, for the Hodge Laplacian of the
    vector stream."""

    @staticmethod
    def forward(ctx, x, v, op):
        x, ldx = _mat(x, 'x')
        v, ldv = _mat(v, 'v')
        g = op.graph
        C = x.size(1)
        if v.size(0) != 2 * g.N or v.size(1) != C or x.size(0) != g.N:
            raise _lib.DeltaConvLibraryError(f"DeltaConv: x {tuple(x.shape)} / v {tuple(v.shape)} do not match N={g.N}")
        cat = torch.empty(g.N, 4 * C, dtype=_F32, device=x.device)
        call('dc_divcurlnorm_fwd', ptr(g.idx), ptr(op.D_raw), g.N, g.k, C, ptr(v), ldv,
             cat.data_ptr() + 4 * C, 4 * C, cat.data_ptr() + 8 * C, 4 * C, cat.data_ptr() + 12 * C, 4 * C,
             ptr(x), ldx, cat.data_ptr(), 4 * C, *op.tail())
        dc = cat[:, C:3 * C]
        ctx.op, ctx.C = op, C
        ctx.save_for_backward(v)
        return cat, dc

    @staticmethod
    def backward(ctx, dcat, ddc):
        op, g, C = ctx.op, ctx.op.graph, ctx.C
        (v,) = ctx.saved_tensors
        v, ldv = _mat(v)
        dcat, ld = _mat(dcat.contiguous(), 'grad_output')
        if ddc is not None:
            ddc, ld_ddc = _mat(ddc, 'grad_output')
            t = torch.empty(g.N, 2 * C, dtype=_F32, device=dcat.device)
            call('dc_add_rows', dcat.data_ptr() + 4 * C, ld, ptr(ddc), ld_ddc, g.N, 2 * C, ptr(t), 2 * C)
            d_div, d_curl, ldt = t.data_ptr(), t.data_ptr() + 4 * C, 2 * C
        else:
            d_div, d_curl, ldt = dcat.data_ptr() + 4 * C, dcat.data_ptr() + 8 * C, ld
        in_ptr, in_edge = g.transpose()
        dv = torch.empty(2 * g.N, C, dtype=_F32, device=dcat.device)
        call('dc_divcurlnorm_bwd', ptr(in_ptr), ptr(in_edge), ptr(op.D_raw), g.N, g.k, C, d_div, ldt, d_curl, ldt,
             dcat.data_ptr() + 12 * C, ld, ptr(v), ldv, ptr(dv), C, *op.tail())
        return dcat[:, :C], dv, None


_colmaps = {}


def _vector_cat_layout(Cin, Cout, pad, device):
    """(column of the hodge block, column of the grad block, total width, column map or None)."""
    if not pad or (Cin % 4 == 0 and Cout % 4 == 0):
        return Cin, 2 * Cin, 2 * Cin + Cout, None
    Cp = (Cin + 3) // 4 * 4
    W = 2 * Cp + (Cout + 3) // 4 * 4
    key = (Cin, Cout, device.index)
    if key not in _colmaps:
        _colmaps[key] = torch.cat([torch.arange(Cin), Cp + torch.arange(Cin), 2 * Cp + torch.arange(Cout)]).to(device=device, dtype=_I32)
    return Cp, 2 * Cp, W, _colmaps[key]


class _VectorCat(torch.autograd.Function):
    """[v | hodge_laplacian(v) | grad x'] (deltaconv.py:66, operators.py:35-46) from the div/curl
    already computed for the scalar stream.  pad=True: every block starts at a multiple of 4 floats (the first layer's widths
    3, 3, 64 would otherwise put the 64-wide block at column 6 and send the kernels that read or write it down their scalar
    path); the columns in between are zero and meet zero weight columns in the GEMM."""

    @staticmethod
    def forward(ctx, v, dc, x_new, op, pad=False):
        v, ldv = _mat(v, 'v')
        dc, ld_dc = _mat(dc, 'div_curl')
        x_new, ldx = _mat(x_new, 'x')
        g = op.graph
        Cin, Cout = v.size(1), x_new.size(1)
        c_h, c_g, W, col_map = _vector_cat_layout(Cin, Cout, pad, v.device)
        if col_map is None:
            cat = torch.empty(2 * g.N, W, dtype=_F32, device=v.device)
        else:
            cat = torch.zeros(2 * g.N, W, dtype=_F32, device=v.device)
        call('dc_hodge_fwd', ptr(g.idx), ptr(op.G_raw), g.N, g.k, Cin, ptr(dc), ld_dc, dc.data_ptr() + 4 * Cin, ld_dc,
             cat.data_ptr() + 4 * c_h, W, ptr(v), ldv, cat.data_ptr(), W, *op.tail())
        call('dc_grad_fwd', ptr(g.idx), ptr(op.G_raw), g.N, g.k, Cout, ptr(x_new), ldx, cat.data_ptr() + 4 * c_g, W, *op.tail())
        ctx.op, ctx.Cin, ctx.Cout, ctx.c_h, ctx.c_g = op, Cin, Cout, c_h, c_g
        if col_map is not None:
            cat._dc_colmap = col_map
        return cat

    @staticmethod
    def backward(ctx, dcat):
        op, g, Cin, Cout = ctx.op, ctx.op.graph, ctx.Cin, ctx.Cout
        dcat, ld = _mat(dcat.contiguous(), 'grad_output')
        in_ptr, in_edge = g.transpose()
        ddc = torch.empty(g.N, 2 * Cin, dtype=_F32, device=dcat.device)
        call('dc_hodge_bwd', ptr(in_ptr), ptr(in_edge), ptr(op.G_raw), g.N, g.k, Cin, dcat.data_ptr() + 4 * ctx.c_h, ld,
             ptr(ddc), 2 * Cin, ddc.data_ptr() + 4 * Cin, 2 * Cin, *op.tail())
        dx = torch.empty(g.N, Cout, dtype=_F32, device=dcat.device)
        call('dc_grad_bwd', ptr(in_ptr), ptr(in_edge), ptr(op.G_raw), g.N, g.k, Cout, dcat.data_ptr() + 4 * ctx.c_g, ld, ptr(dx), Cout, *op.tail())
        return dcat[:, :Cin], ddc, dx, None, None


def scalar_cat(x, v, div_op):
    return _ScalarCat.apply(x, v, div_op)


def vector_cat(v, div_curl, x_new, grad_op, pad=False):
    """[v | hodge_laplacian(v) | grad x'] as one [2N, 2 Cin + Cout] buffer; pad=True starts every block at a multiple of 4
    floats, with zero columns in between (the result carries the column map ops.vector_linear_with_rotation needs)."""
    return _VectorCat.apply(v, div_curl, x_new, grad_op, pad)


def curl_of(v, div_op):
    """curl(v, div) = -div @ J(v)  (operators.py:23-27) without materialising J(v)."""
    return _CurlApply.apply(v, div_op)


class _CurlApply(torch.autograd.Function):
    @staticmethod
    def forward(ctx, v, op):
        v, ldv = _mat(v, 'v')
        g = op.graph
        C = v.size(1)
        out = torch.empty(g.N, C, dtype=_F32, device=v.device)
        call('dc_divcurlnorm_fwd', ptr(g.idx), ptr(op.D_raw), g.N, g.k, C, ptr(v), ldv, None, 0, ptr(out), C, None, 0, None, 0, None, 0, *op.tail())
        ctx.op = op
        return out

    @staticmethod
    def backward(ctx, dcurl):
        op, g = ctx.op, ctx.op.graph
        dcurl, ld = _mat(dcurl, 'grad_output')
        in_ptr, in_edge = g.transpose()
        C = dcurl.size(1)
        dv = torch.empty(2 * g.N, C, dtype=_F32, device=dcurl.device)
        call('dc_divcurlnorm_bwd', ptr(in_ptr), ptr(in_edge), ptr(op.D_raw), g.N, g.k, C, None, 0, ptr(dcurl), ld, None, 0,
             None, 0, ptr(dv), C, *op.tail())
        return dv, None


# ------------------------------------------------------------------------------------------------
# column blocks: producers write straight into their slice of a concatenation buffer
# ------------------------------------------------------------------------------------------------

class ColumnBuffer:
    """One [rows, width] fp32 buffer that several operators fill block of columns by block (through their output leading
    dimension), so that `torch.cat(parts, dim=1)` (deltanet_classification.py:44, deltanet_segmentation.py:57,66) never
    copies: `join_columns(parts)` hands the whole buffer to the consumer and routes the gradient back block by block.
    `block(col, width)` is an alias of the block's memory that is NOT a view in autograd's eyes (no base tensor), so an
    operator can return it as a fresh output."""

    def __init__(self, rows, width, device):
        self.buf = torch.empty(rows, width, dtype=_F32, device=device)

    def block(self, col, width):
        b = self.buf
        if col < 0 or col + width > b.size(1):
            raise _lib.DeltaConvLibraryError(f"ColumnBuffer: block [{col}, {col + width}) outside a buffer of width {b.size(1)}")
        alias = torch.empty(0, dtype=_F32, device=b.device).set_(b.untyped_storage(), b.storage_offset() + col,
                                                                 (b.size(0), width), (b.stride(0), 1))
        alias._dc_colbuf = self
        return alias


def _out_block(slot, rows, width, device):
    """Output tensor of an operator: a fresh [rows, width] tensor, or the block `slot = (ColumnBuffer, first column)`."""
    if slot is None:
        return torch.empty(rows, width, dtype=_F32, device=device)
    cb, col = slot
    if cb.buf.size(0) != rows:
        raise _lib.DeltaConvLibraryError(f"ColumnBuffer has {cb.buf.size(0)} rows, the operator produces {rows}")
    return cb.block(col, width)


class _JoinColumns(torch.autograd.Function):
    @staticmethod
    def forward(ctx, cb, first_col, *parts):
        ctx.widths = [p.size(1) for p in parts]
        b = cb.buf
        return torch.empty(0, dtype=_F32, device=b.device).set_(b.untyped_storage(), b.storage_offset() + first_col,
                                                                 (b.size(0), sum(ctx.widths)), (b.stride(0), 1))

    @staticmethod
    def backward(ctx, dbuf):
        grads, col = [], 0
        for w in ctx.widths:
            grads.append(dbuf[:, col:col + w])
            col += w
        return (None, None) + tuple(grads)


def join_columns(parts):
    """torch.cat(parts, dim=1); free when the parts are adjacent blocks, in order, of one ColumnBuffer: the result is then
    the buffer's column range itself (row pitch = the buffer's width; every kernel takes a leading dimension) and the gradient
    is handed back block by block as views."""
    parts = list(parts)
    cb = getattr(parts[0], '_dc_colbuf', None)
    if cb is not None and all(getattr(p, '_dc_colbuf', None) is cb for p in parts):
        b = cb.buf
        first = (parts[0].data_ptr() - b.data_ptr()) // 4
        col, ok = first, 0 <= first < b.size(1)
        for p in parts:
            ok = ok and p.dim() == 2 and p.data_ptr() == b.data_ptr() + 4 * col and p.size(0) == b.size(0) and \
                (p.stride(0) == b.stride(0) or p.size(0) <= 1) and (p.stride(1) == 1 or p.size(1) <= 1)
            col += p.size(1)
        if ok and col <= b.size(1):
            return _JoinColumns.apply(cb, first, *parts)
    return torch.cat(parts, dim=1)


# ------------------------------------------------------------------------------------------------
# Linear
# ------------------------------------------------------------------------------------------------

def _gemm_ws(M, N, K, device, slot=0):
    nbytes = query('dc_gemm_ws_bytes', M, N, K)
    return workspace(nbytes, device, slot), nbytes


# The input gradient and the weight gradient of a Linear are independent products.  While a step is being CAPTURED into a
# CUDA graph the weight gradient is issued on a side stream (a parallel branch of the graph): the persistent GEMM kernels end
# in a partially filled wave on the narrow layers (32768 x 64 is 1.7 waves of tiles), and the other product's CTAs fill the
# SMs that fall idle.  Eager launches stay on one stream (the extra event calls would cost the host more than they gain).
FORK_BACKWARD_GEMMS = not os.environ.get('DC_NO_FORK')
_side_streams = {}


def _linear_backward_pair(dy, ld, weight, ldw, x, ldx, M, K, Nout, want_dx, want_dw):
    """(dX, dW) of Y = X W^T from dY; the two products run concurrently under graph capture."""
    dev = dy.device
    dx = torch.empty(M, K, dtype=_F32, device=dev) if want_dx else None
    dw = torch.empty(Nout, K, dtype=_F32, device=dev) if want_dw else None
    forkable = FORK_BACKWARD_GEMMS and want_dx and want_dw and M >= 4096
    fork = forkable and torch.cuda.is_current_stream_capturing()
    if forkable and not fork:
        _gemm_ws(Nout, K, M, dev, slot=1)           # the side stream's scratch exists before any capture (warm-up steps are eager)
    if want_dw:
        ws_w, nb_w = _gemm_ws(Nout, K, M, dev, slot=1 if fork else 0)
    if want_dx:
        ws_x, nb_x = _gemm_ws(M, K, Nout, dev)
    if fork:
        cur = torch.cuda.current_stream()
        side = _side_streams.get(dev.index)
        if side is None:
            side = _side_streams[dev.index] = torch.cuda.Stream(device=dev)
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            call('dc_linear_bwd_weight', ptr(dy), ld, ptr(x), ldx, M, K, Nout, ptr(dw), K, ptr(ws_w), nb_w)
        call('dc_linear_bwd_input', ptr(dy), ld, ptr(weight), ldw, M, K, Nout, ptr(dx), K, ptr(ws_x), nb_x)
        cur.wait_stream(side)
        return dx, dw
    if want_dx:
        call('dc_linear_bwd_input', ptr(dy), ld, ptr(weight), ldw, M, K, Nout, ptr(dx), K, ptr(ws_x), nb_x)
    if want_dw:
        call('dc_linear_bwd_weight', ptr(dy), ld, ptr(x), ldx, M, K, Nout, ptr(dw), K, ptr(ws_w), nb_w)
    return dx, dw


class _Linear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias):
        x, ldx = _mat(x, 'input')
        weight, ldw = _mat(weight, 'weight')
        if x.size(1) != weight.size(1):
            raise _lib.DeltaConvLibraryError(f"Linear: input has {x.size(1)} features, weight expects {weight.size(1)}")
        M, K, Nout = x.size(0), x.size(1), weight.size(0)
        y = torch.empty(M, Nout, dtype=_F32, device=x.device)
        if M:
            ws, nb = _gemm_ws(M, Nout, K, x.device)
            call('dc_linear_fwd', ptr(x), ldx, ptr(weight), ldw, ptr(_f32(bias, 'bias').contiguous()) if bias is not None else None,
                 M, K, Nout, ptr(y), Nout, ptr(ws), nb)
        ctx.save_for_backward(x, weight)
        ctx.has_bias = bias is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        x, ldx = _mat(x)
        weight, ldw = _mat(weight)
        dy, ld = _mat(dy, 'grad_output')
        M, K, Nout = x.size(0), x.size(1), weight.size(0)
        db = None
        dx, dw = _linear_backward_pair(dy, ld, weight, ldw, x, ldx, M, K, Nout, ctx.needs_input_grad[0], ctx.needs_input_grad[1])
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = torch.empty(Nout, dtype=_F32, device=dy.device)
            nb = query('dc_colreduce_ws_bytes', Nout)
            ws = workspace(nb, dy.device)
            call('dc_col_sum', ptr(dy), ld, M, Nout, ptr(db), ptr(ws), nb)
        return dx, dw, db


def linear(x, weight, bias=None):
    return _Linear.apply(x, weight, bias)


class _ExpandVecWeight(torch.autograd.Function):
    """[W1 | W2] -> [[W1, -W2], [W2, W1]] — folds I_J (operators.py:19-21) into the weights.  `Kp` >= K is the width of the
    (padded) operand it multiplies, `col_map` the operand column of every weight column."""

    @staticmethod
    def forward(ctx, weight, Kp, col_map):
        weight = _f32(weight, 'weight').contiguous()
        Cout, K2 = weight.shape
        out = torch.empty(2 * Cout, 2 * Kp, dtype=_F32, device=weight.device)
        call('dc_expand_vec_weight', ptr(weight), Cout, K2 // 2, Kp, ptr(col_map), ptr(out))
        ctx.K, ctx.col_map = K2 // 2, col_map
        return out

    @staticmethod
    def backward(ctx, dwe):
        dwe = dwe.contiguous()
        Cout, Kp, K = dwe.size(0) // 2, dwe.size(1) // 2, ctx.K
        dw = torch.empty(Cout, 2 * K, dtype=_F32, device=dwe.device)
        call('dc_fold_vec_weight_grad', ptr(dwe), Cout, K, Kp, ptr(ctx.col_map), ptr(dw))
        return dw, None, None


def vector_linear_with_rotation(v_cat, weight):
    """Linear(I_J(v_cat)) (deltaconv.py:68) as one GEMM on the [N, 2K] view of the interleaved field with the expanded weight;
    v_cat [2N,Kp], weight [Cout,2K] -> [2N,Cout].  A `v_cat` from ops.vector_cat(pad=True) carries its column map
    (`_dc_colmap`): its blocks start at multiples of 4 floats, the columns in between are zero."""
    rows, Kp = v_cat.shape
    K = weight.size(1) // 2
    col_map = getattr(v_cat, '_dc_colmap', None)
    if weight.size(1) != 2 * K or (Kp != K and (col_map is None or col_map.numel() != K)):
        raise _lib.DeltaConvLibraryError(f"vector Linear: weight expects {weight.size(1)} inputs, I_J(v) has {2 * Kp}")
    we = _ExpandVecWeight.apply(weight, Kp, col_map)
    out = _Linear.apply(v_cat.contiguous().view(rows // 2, 2 * Kp), we, None)
    return out.view(rows, weight.size(0))


# ------------------------------------------------------------------------------------------------
# BatchNorm + activation, VectorNonLin
# ------------------------------------------------------------------------------------------------

def _running_update(bn):
    """(update, momentum, counter pointer) for a training-mode statistics pass.  With a fixed momentum the kernel
    increments `num_batches_tracked` itself; the cumulative average (momentum=None) needs the count on the host."""
    update = bn.training and bn.track_running_stats and bn.running_mean is not None
    if not update:
        return False, 0.0, None
    if bn.momentum is None:
        bn.num_batches_tracked += 1
        return True, 1.0 / float(bn.num_batches_tracked), None
    nbt = bn.num_batches_tracked
    counted = nbt is not None and nbt.is_cuda and nbt.dtype == torch.int64
    if nbt is not None and not counted:
        bn.num_batches_tracked += 1
    return True, float(bn.momentum), (nbt.data_ptr() if counted else None)


def _running_invstd(bn):
    """1 / sqrt(running_var + eps) of a BatchNorm in evaluation mode, computed once per state of `running_var` (its in-place
    version counter) instead of with two small launches per layer and forward.  A captured inference graph holds the tensor
    that was current at capture: re-capture after loading new statistics."""
    rv = bn.running_var
    cached = getattr(bn, '_dc_invstd', None)
    if cached is None or cached[0] != rv._version or cached[1] != bn.eps or cached[2].device != rv.device or cached[3] != rv.data_ptr():
        cached = (rv._version, bn.eps, torch.rsqrt(rv + bn.eps), rv.data_ptr())
        bn._dc_invstd = cached
    return cached[2]


def _bn_moments(bn, x, ldx, M, C, mode):
    """(mean, invstd) to normalise with: batch statistics in training mode (running stats are
    updated in place like torch.nn.BatchNorm1d), running statistics in eval mode."""
    dev = x.device
    use_batch = bn.training or not bn.track_running_stats or bn.running_mean is None
    if use_batch:
        mean = torch.empty(C, dtype=_F32, device=dev)
        invstd = torch.empty(C, dtype=_F32, device=dev)
        update, momentum, nbt = _running_update(bn)
        nb = query('dc_colreduce_ws_bytes', C)
        ws = workspace(nb, dev)
        call('dc_bn_stats', x.data_ptr(), ldx, M, C, mode, float(bn.eps), float(momentum), ptr(mean), ptr(invstd),
             ptr(bn.running_mean) if update else None, ptr(bn.running_var) if update else None, nbt, ptr(ws), nb)
        return mean, invstd, True
    return bn.running_mean.contiguous(), _running_invstd(bn), False


class _BNAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, y, gamma, beta, addend, bn, slope, slot=None):
        y, ldy = _mat(y, 'input')
        M, C = y.shape
        if M == 0:
            raise _lib.DeltaConvLibraryError("BatchNorm1d: empty batch")
        mean, invstd, batch_stats = _bn_moments(bn, y, ldy, M, C, 0)
        out = _out_block(slot, M, C, y.device)
        add_p, ld_add = (None, 0)
        if addend is not None:
            addend, ld_add = _mat(addend, 'addend')
            add_p = ptr(addend)
        call('dc_bn_act_fwd', ptr(y), ldy, M, C, ptr(mean), ptr(invstd), ptr(gamma), ptr(beta), float(slope),
             add_p, ld_add, ptr(out), out.stride(0) if M > 1 else C)
        ctx.save_for_backward(y, mean, invstd, gamma, beta)
        ctx.slope, ctx.batch_stats, ctx.has_add = slope, batch_stats, addend is not None
        return out

    @staticmethod
    def backward(ctx, dout):
        y, mean, invstd, gamma, beta = ctx.saved_tensors
        y, ldy = _mat(y)
        dout, ldd = _mat(dout, 'grad_output')
        M, C = y.shape
        dgamma = torch.empty(C, dtype=_F32, device=y.device)
        dbeta = torch.empty(C, dtype=_F32, device=y.device)
        dy = torch.empty(M, C, dtype=_F32, device=y.device)
        nb = query('dc_colreduce_ws_bytes', C)
        ws = workspace(nb, y.device)
        call('dc_bn_act_bwd', ptr(y), ldy, ptr(dout), ldd, M, C, ptr(mean), ptr(invstd), ptr(gamma), ptr(beta),
             float(ctx.slope), int(ctx.batch_stats), ptr(dgamma), ptr(dbeta), ptr(dy), C, ptr(ws), nb)
        return (dy, dgamma if gamma is not None else None, dbeta if beta is not None else None,
                dout if ctx.has_add else None, None, None, None)


def bn_act(y, bn, slope=1.0, addend=None, slot=None):
    """act(BatchNorm1d(y)) + addend with act = LeakyReLU(slope) (slope 1 = identity); `slot` = (ColumnBuffer, column)
    makes the result land in that block."""
    return _BNAct.apply(y, bn.weight, bn.bias, addend, bn, float(slope), slot)


class _LinearBNAct(torch.autograd.Function):
    """Linear (no bias) -> BatchNorm1d (batch statistics) -> LeakyReLU(slope) (+ addend) as ONE autograd node
    (mlp.py:9) — one node instead of two keeps the pre-activation the only saved intermediate."""

    @staticmethod
    def forward(ctx, x, weight, gamma, beta, addend, bn, slope, slot=None):
        x, ldx = _mat(x, 'input')
        weight, ldw = _mat(weight, 'weight')
        if x.size(1) != weight.size(1):
            raise _lib.DeltaConvLibraryError(f"Linear: input has {x.size(1)} features, weight expects {weight.size(1)}")
        M, K, C = x.size(0), x.size(1), weight.size(0)
        if M == 0:
            raise _lib.DeltaConvLibraryError("BatchNorm1d: empty batch")
        dev = x.device
        y = torch.empty(M, C, dtype=_F32, device=dev)
        mean = torch.empty(C, dtype=_F32, device=dev)
        invstd = torch.empty(C, dtype=_F32, device=dev)
        update, momentum, nbt = _running_update(bn)
        ws, nb = _gemm_ws(M, C, K, dev)
        call('dc_linear_fwd', ptr(x), ldx, ptr(weight), ldw, None, M, K, C, ptr(y), C, ptr(ws), nb)
        nb = query('dc_colreduce_ws_bytes', C)
        ws = workspace(nb, dev)
        call('dc_bn_stats', ptr(y), C, M, C, 0, float(bn.eps), float(momentum), ptr(mean), ptr(invstd),
             ptr(bn.running_mean) if update else None, ptr(bn.running_var) if update else None, nbt, ptr(ws), nb)
        out = _out_block(slot, M, C, dev)
        add_p, ld_add = None, 0
        if addend is not None:
            addend, ld_add = _mat(addend, 'addend')
            add_p = ptr(addend)
        call('dc_bn_act_fwd', ptr(y), C, M, C, ptr(mean), ptr(invstd), ptr(gamma), ptr(beta), float(slope), add_p, ld_add,
             ptr(out), out.stride(0) if M > 1 else C)
        ctx.save_for_backward(x, weight, y, mean, invstd, gamma, beta)
        ctx.slope, ctx.has_add = slope, addend is not None
        return out

    @staticmethod
    def backward(ctx, dout):
        x, weight, y, mean, invstd, gamma, beta = ctx.saved_tensors
        x, ldx = _mat(x)
        weight, ldw = _mat(weight)
        dout, ldd = _mat(dout, 'grad_output')
        M, K, C = x.size(0), x.size(1), weight.size(0)
        dev = x.device
        dgamma = torch.empty(C, dtype=_F32, device=dev)
        dbeta = torch.empty(C, dtype=_F32, device=dev)
        dy = torch.empty(M, C, dtype=_F32, device=dev)
        nb = query('dc_colreduce_ws_bytes', C)
        ws = workspace(nb, dev)
        call('dc_bn_act_bwd', ptr(y), C, ptr(dout), ldd, M, C, ptr(mean), ptr(invstd), ptr(gamma), ptr(beta), float(ctx.slope), 1,
             ptr(dgamma), ptr(dbeta), ptr(dy), C, ptr(ws), nb)
        dx, dw = _linear_backward_pair(dy, C, weight, ldw, x, ldx, M, K, C, ctx.needs_input_grad[0], ctx.needs_input_grad[1])
        return (dx, dw, dgamma if gamma is not None else None, dbeta if beta is not None else None,
                dout if ctx.has_add else None, None, None, None)


def linear_bn_act(x, weight, bn, slope=1.0, addend=None, slot=None):
    """act(BatchNorm1d(x @ weight.T)) + addend with batch statistics (training mode); `slot` = (ColumnBuffer, column) makes
    the result land in that block."""
    return _LinearBNAct.apply(x, weight, bn.weight, bn.bias, addend, bn, float(slope), slot)


class _VecNonLin(torch.autograd.Function):
    @staticmethod
    def forward(ctx, v, gamma, beta, bias, bn):
        v, ldv = _mat(v, 'input')
        rows, C = v.shape
        if rows % 2:
            raise _lib.DeltaConvLibraryError("VectorNonLin: vector fields have an even number of rows")
        M = rows // 2
        out = torch.empty(rows, C, dtype=_F32, device=v.device)
        if bn is not None:
            mean, invstd, batch_stats = _bn_moments(bn, v, ldv, M, C, 1)
            call('dc_vecnonlin_fwd', ptr(v), ldv, M, C, ptr(mean), ptr(invstd), ptr(gamma), ptr(beta), None, ptr(out), C)
        else:
            mean = invstd = None
            batch_stats = False
            call('dc_vecnonlin_fwd', ptr(v), ldv, M, C, None, None, None, None, ptr(bias), ptr(out), C)
        ctx.save_for_backward(v, mean, invstd, gamma, beta, bias)
        ctx.batch_stats, ctx.with_bn = batch_stats, bn is not None
        return out

    @staticmethod
    def backward(ctx, dout):
        v, mean, invstd, gamma, beta, bias = ctx.saved_tensors
        v, ldv = _mat(v)
        dout, ldd = _mat(dout, 'grad_output')
        rows, C = v.shape
        M = rows // 2
        dv = torch.empty(rows, C, dtype=_F32, device=v.device)
        nb = query('dc_colreduce_ws_bytes', C)
        ws = workspace(nb, v.device)
        if ctx.with_bn:
            dgamma = torch.empty(C, dtype=_F32, device=v.device)
            dbeta = torch.empty(C, dtype=_F32, device=v.device)
            call('dc_vecnonlin_bwd', ptr(v), ldv, ptr(dout), ldd, M, C, ptr(mean), ptr(invstd), ptr(gamma), ptr(beta), None,
                 int(ctx.batch_stats), ptr(dgamma), ptr(dbeta), None, ptr(dv), C, ptr(ws), nb)
            return dv, dgamma if gamma is not None else None, dbeta if beta is not None else None, None, None
        rows_grad = torch.empty(M, C, dtype=_F32, device=v.device)
        call('dc_vecnonlin_bwd', ptr(v), ldv, ptr(dout), ldd, M, C, None, None, None, None, ptr(bias), 0,
             None, None, ptr(rows_grad), ptr(dv), C, ptr(ws), nb)
        dbias = torch.empty(C, dtype=_F32, device=v.device)
        call('dc_col_sum', ptr(rows_grad), C, M, C, ptr(dbias), ptr(ws), nb)
        return dv, None, None, dbias, None


def vec_nonlin(v, bias, bn):
    if bn is None:
        _f32(bias, 'VectorNonLin.bias')              # a parameter left on the CPU would hand the kernel a host address
    if bn is not None:
        return _VecNonLin.apply(v, bn.weight, bn.bias, None, bn)
    return _VecNonLin.apply(v, None, None, bias, None)


# ------------------------------------------------------------------------------------------------
# neighbourhood max, pools
# ------------------------------------------------------------------------------------------------

class _GatherMax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, y, graph):
        y, ldy = _mat(y, 'input')
        N, k, C = graph.N, graph.k, y.size(1)
        out = torch.empty(N, C, dtype=_F32, device=y.device)
        arg = torch.empty(N, C, dtype=torch.uint8, device=y.device)
        call('dc_gathermax_fwd', ptr(graph.idx), N, k, C, ptr(y), ldy, ptr(out), C, ptr(arg))
        ctx.graph = graph
        ctx.save_for_backward(arg)
        return out

    @staticmethod
    def backward(ctx, dout):
        g = ctx.graph
        (arg,) = ctx.saved_tensors
        dout, ldd = _mat(dout, 'grad_output')
        in_ptr, in_edge = g.transpose()
        C = dout.size(1)
        dy = torch.empty(g.N, C, dtype=_F32, device=dout.device)
        call('dc_gathermax_bwd', ptr(in_ptr), ptr(in_edge), g.N, g.k, C, ptr(arg), ptr(dout), ldd, ptr(dy), C)
        return dy, None


def gather_max(y, graph):
    return _GatherMax.apply(y, graph)


class _LinearBNActGatherMax(torch.autograd.Function):
    """max over the neighbours of act(BatchNorm1d(x @ W.T)) (deltaconv.py:54 with the last block of s_mlp_max) as ONE node: the
    GEMM writes the pre-activation, the statistics pass reads it, and the gather-max applies BatchNorm + LeakyReLU to the
    per-channel extreme it selects (dc_gathermax_bn_fwd) — the [N,C] activation is neither written nor read.  Backward: the
    plain gather-max backward gives the gradient of the activation, then the BatchNorm and Linear backward of _LinearBNAct."""

    @staticmethod
    def forward(ctx, x, weight, gamma, beta, bn, slope, graph):
        x, ldx = _mat(x, 'input')
        weight, ldw = _mat(weight, 'weight')
        if x.size(1) != weight.size(1):
            raise _lib.DeltaConvLibraryError(f"Linear: input has {x.size(1)} features, weight expects {weight.size(1)}")
        M, K, C = x.size(0), x.size(1), weight.size(0)
        if M != graph.N:
            raise _lib.DeltaConvLibraryError(f"gather-max: {M} feature rows for a graph on {graph.N} points")
        dev = x.device
        y = torch.empty(M, C, dtype=_F32, device=dev)
        ws, nb = _gemm_ws(M, C, K, dev)
        call('dc_linear_fwd', ptr(x), ldx, ptr(weight), ldw, None, M, K, C, ptr(y), C, ptr(ws), nb)
        mean, invstd, batch_stats = _bn_moments(bn, y, C, M, C, 0)
        out = torch.empty(M, C, dtype=_F32, device=dev)
        arg = torch.empty(M, C, dtype=torch.uint8, device=dev)
        call('dc_gathermax_bn_fwd', ptr(graph.idx), M, graph.k, C, ptr(y), C, ptr(mean), ptr(invstd), ptr(gamma), ptr(beta),
             float(slope), ptr(out), C, ptr(arg))
        ctx.save_for_backward(x, weight, y, mean, invstd, gamma, beta, arg)
        ctx.slope, ctx.batch_stats, ctx.graph = slope, batch_stats, graph
        return out

    @staticmethod
    def backward(ctx, dout):
        x, weight, y, mean, invstd, gamma, beta, arg = ctx.saved_tensors
        g = ctx.graph
        x, ldx = _mat(x)
        weight, ldw = _mat(weight)
        dout, ldd = _mat(dout, 'grad_output')
        M, K, C = x.size(0), x.size(1), weight.size(0)
        dev = x.device
        in_ptr, in_edge = g.transpose()
        da = torch.empty(M, C, dtype=_F32, device=dev)
        call('dc_gathermax_bwd', ptr(in_ptr), ptr(in_edge), M, g.k, C, ptr(arg), ptr(dout), ldd, ptr(da), C)
        dgamma = torch.empty(C, dtype=_F32, device=dev)
        dbeta = torch.empty(C, dtype=_F32, device=dev)
        dy = torch.empty(M, C, dtype=_F32, device=dev)
        nb = query('dc_colreduce_ws_bytes', C)
        ws = workspace(nb, dev)
        call('dc_bn_act_bwd', ptr(y), C, ptr(da), C, M, C, ptr(mean), ptr(invstd), ptr(gamma), ptr(beta), float(ctx.slope),
             int(ctx.batch_stats), ptr(dgamma), ptr(dbeta), ptr(dy), C, ptr(ws), nb)
        dx, dw = _linear_backward_pair(dy, C, weight, ldw, x, ldx, M, K, C, ctx.needs_input_grad[0], ctx.needs_input_grad[1])
        return (dx, dw, dgamma if gamma is not None else None, dbeta if beta is not None else None, None, None, None)


def gathermax_fusable(lin, slope, x):
    """The last block of s_mlp_max can run as the fused node: no Linear bias, a strictly increasing LeakyReLU-family activation
    (slope > 0; ReLU's flat part would make the arg ambiguous but not the value — kept out for exact parity of the routing),
    a channel count the float4 kernels take."""
    return lin.bias is None and slope is not None and slope > 0 and lin.out_features % 4 == 0 and x.dim() == 2


def linear_bn_act_gathermax(x, weight, bn, slope, graph):
    return _LinearBNActGatherMax.apply(x, weight, bn.weight, bn.bias, bn, float(slope), graph)


class _EdgeDiff(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, graph):
        x, ldx = _mat(x, 'x')
        N, k, C = graph.N, graph.k, x.size(1)
        out = torch.empty(N * k, C, dtype=_F32, device=x.device)
        call('dc_edge_diff_fwd', ptr(graph.idx), N, k, C, ptr(x), ldx, ptr(out))
        ctx.graph = graph
        return out

    @staticmethod
    def backward(ctx, de):
        g = ctx.graph
        de = de.contiguous()
        in_ptr, in_edge = g.transpose()
        C = de.size(1)
        dx = torch.empty(g.N, C, dtype=_F32, device=de.device)
        call('dc_edge_diff_bwd', ptr(in_ptr), ptr(in_edge), g.N, g.k, C, ptr(de), ptr(dx), C)
        return dx, None


class _SegMax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, t, graph):
        t = _f32(t, 'input').contiguous()
        N, k, C = graph.N, graph.k, t.size(1)
        out = torch.empty(N, C, dtype=_F32, device=t.device)
        arg = torch.empty(N, C, dtype=torch.uint8, device=t.device)
        call('dc_segmax_fwd', ptr(t), N, k, C, ptr(graph.idx), ptr(out), C, ptr(arg))
        ctx.graph = graph
        ctx.save_for_backward(arg)
        return out

    @staticmethod
    def backward(ctx, dout):
        g = ctx.graph
        (arg,) = ctx.saved_tensors
        dout, ldd = _mat(dout, 'grad_output')
        C = dout.size(1)
        din = torch.empty(g.N * g.k, C, dtype=_F32, device=dout.device)
        call('dc_segmax_bwd', ptr(dout), ldd, g.N, g.k, C, ptr(arg), ptr(din))
        return din, None


def edge_diff(x, graph):
    return _EdgeDiff.apply(x, graph)


def segment_max(t, graph):
    return _SegMax.apply(t, graph)


class _CentralizedMax(torch.autograd.Function):
    """max over each point's k edges of act(BatchNorm(Linear(x_j - x_i))) — the centralized scalar aggregation of
    the first DeltaConv layer (deltaconv.py:50-52) — without materialising any per-edge tensor."""

    @staticmethod
    def forward(ctx, x, weight, gamma, beta, bn, graph, slope):
        x, ldx = _mat(x, 'x')
        weight = _f32(weight, 'weight').contiguous()
        N, k, Cin, Cout = graph.N, graph.k, x.size(1), weight.size(0)
        dev = x.device
        nbytes = query('dc_central_ws_bytes', N, k, Cin, Cout)
        ws = workspace(nbytes, dev)
        use_batch = bn.training or not bn.track_running_stats or bn.running_mean is None
        moments = None
        if use_batch:
            moments = torch.empty(20, dtype=torch.float64, device=dev)
            mean = torch.empty(Cout, dtype=_F32, device=dev)
            invstd = torch.empty(Cout, dtype=_F32, device=dev)
            update, momentum, nbt = _running_update(bn)
            call('dc_central_stats', ptr(graph.idx), N, k, Cin, Cout, ptr(x), ldx, ptr(weight), float(bn.eps), float(momentum),
                 ptr(moments), ptr(mean), ptr(invstd), ptr(bn.running_mean) if update else None,
                 ptr(bn.running_var) if update else None, nbt, ptr(ws), nbytes)
        else:
            mean = bn.running_mean.contiguous()
            invstd = _running_invstd(bn)
        out = torch.empty(N, Cout, dtype=_F32, device=dev)
        arg = torch.empty(N, Cout, dtype=torch.uint8, device=dev)
        call('dc_central_fwd', ptr(graph.idx), N, k, Cin, Cout, ptr(x), ldx, ptr(weight), ptr(gamma), ptr(beta), ptr(mean), ptr(invstd),
             float(slope), ptr(out), Cout, ptr(arg))
        ctx.graph, ctx.slope, ctx.use_batch = graph, slope, use_batch
        ctx.save_for_backward(x, weight, gamma, beta, mean, invstd, moments, arg)
        return out

    @staticmethod
    def backward(ctx, dout):
        x, weight, gamma, beta, mean, invstd, moments, arg = ctx.saved_tensors
        g = ctx.graph
        x, ldx = _mat(x)
        dout, ldd = _mat(dout, 'grad_output')
        N, k, Cin, Cout = g.N, g.k, x.size(1), weight.size(0)
        dev = x.device
        nbytes = query('dc_central_ws_bytes', N, k, Cin, Cout)
        ws = workspace(nbytes, dev)
        dW = torch.empty(Cout, Cin, dtype=_F32, device=dev)
        dgamma = torch.empty(Cout, dtype=_F32, device=dev)
        dbeta = torch.empty(Cout, dtype=_F32, device=dev)
        dx, in_ptr, in_edge = None, None, None
        if ctx.needs_input_grad[0]:
            dx = torch.empty(N, Cin, dtype=_F32, device=dev)
            in_ptr, in_edge = g.transpose()
        call('dc_central_bwd', ptr(g.idx), ptr(in_ptr), ptr(in_edge), N, k, Cin, Cout, ptr(x), ldx, ptr(weight), ptr(gamma), ptr(beta),
             ptr(mean), ptr(invstd), ptr(moments), float(ctx.slope), int(ctx.use_batch), ptr(arg), ptr(dout), ldd,
             ptr(dW), ptr(dgamma), ptr(dbeta), ptr(dx), Cin, ptr(ws), nbytes)
        return (dx, dW, dgamma if gamma is not None else None, dbeta if beta is not None else None, None, None, None)


def centralized_max_supported(in_channels, k):
    return in_channels <= query('dc_central_max_cin') and k <= 64


def centralized_max(x, weight, bn, graph, slope):
    return _CentralizedMax.apply(x, weight, bn.weight, bn.bias, bn, graph, float(slope))


class _CentralEdgeAct(torch.autograd.Function):
    """First block of a centralized edge MLP of depth >= 2 (deltaconv.py:50-52 with mlp_depth = 2, the ShapeNet configuration):
    h [N*k, Cout] = act(BatchNorm(Linear(x_j - x_i))) written ONCE from the coordinates.  Statistics over the E edge rows from
    the moments of the offsets (dc_central_stats); backward from the dense gradient of h in one pass over it
    (dc_central_edge_bwd).  No gradient w.r.t. x: callers that need it take the unfused sequence."""

    @staticmethod
    def forward(ctx, x, weight, gamma, beta, bn, graph, slope):
        x, ldx = _mat(x, 'x')
        weight = _f32(weight, 'weight').contiguous()
        N, k, Cin, Cout = graph.N, graph.k, x.size(1), weight.size(0)
        dev = x.device
        use_batch = bn.training or not bn.track_running_stats or bn.running_mean is None
        moments = None
        if use_batch:
            nbytes = query('dc_central_ws_bytes', N, k, Cin, Cout)
            ws = workspace(nbytes, dev)
            moments = torch.empty(20, dtype=torch.float64, device=dev)
            mean = torch.empty(Cout, dtype=_F32, device=dev)
            invstd = torch.empty(Cout, dtype=_F32, device=dev)
            update, momentum, nbt = _running_update(bn)
            call('dc_central_stats', ptr(graph.idx), N, k, Cin, Cout, ptr(x), ldx, ptr(weight), float(bn.eps), float(momentum),
                 ptr(moments), ptr(mean), ptr(invstd), ptr(bn.running_mean) if update else None,
                 ptr(bn.running_var) if update else None, nbt, ptr(ws), nbytes)
        else:
            mean = bn.running_mean.contiguous()
            invstd = _running_invstd(bn)
        out = torch.empty(N * k, Cout, dtype=_F32, device=dev)
        call('dc_central_edge_fwd', ptr(graph.idx), N, k, Cin, Cout, ptr(x), ldx, ptr(weight), ptr(gamma), ptr(beta), ptr(mean), ptr(invstd),
             float(slope), ptr(out))
        ctx.graph, ctx.slope, ctx.use_batch = graph, slope, use_batch
        ctx.save_for_backward(x, weight, gamma, beta, mean, invstd, moments)
        return out

    @staticmethod
    def backward(ctx, dh):
        if ctx.needs_input_grad[0]:
            raise _lib.DeltaConvLibraryError("central_edge_act: no gradient w.r.t. the point features on this path")
        x, weight, gamma, beta, mean, invstd, moments = ctx.saved_tensors
        g = ctx.graph
        x, ldx = _mat(x)
        dh, ldd = _mat(dh, 'grad_output')
        N, k, Cin, Cout = g.N, g.k, x.size(1), weight.size(0)
        dev = x.device
        nbytes = query('dc_central_ws_bytes', N, k, Cin, Cout)
        ws = workspace(nbytes, dev)
        dW = torch.empty(Cout, Cin, dtype=_F32, device=dev)
        dgamma = torch.empty(Cout, dtype=_F32, device=dev)
        dbeta = torch.empty(Cout, dtype=_F32, device=dev)
        call('dc_central_edge_bwd', ptr(g.idx), N, k, Cin, Cout, ptr(x), ldx, ptr(weight), ptr(gamma), ptr(beta), ptr(mean), ptr(invstd),
             ptr(moments), float(ctx.slope), int(ctx.use_batch), ptr(dh), ldd, ptr(dW), ptr(dgamma), ptr(dbeta), ptr(ws), nbytes)
        return (None, dW, dgamma if gamma is not None else None, dbeta if beta is not None else None, None, None, None)


def central_edge_act(x, weight, bn, graph, slope):
    return _CentralEdgeAct.apply(x, weight, bn.weight, bn.bias, bn, graph, float(slope))


class _LinearBNActSegMax(torch.autograd.Function):
    """max over each point's k edge rows of act(BatchNorm1d(h @ W.T)) — the LAST block of a centralized edge MLP and the
    scatter-max of deltaconv.py:52 as one node: the GEMM writes the pre-activation [N*k, C], the statistics pass reads it, the
    max applies BatchNorm + LeakyReLU to the extreme it selects (dc_segmax_bn_fwd).  Backward: the gradient of the activation
    is non-zero at one edge per (point, channel), so the BatchNorm sums run over the points and the [N*k, C] gradient of the
    pre-activation is written directly (dc_segmax_bn_bwd) — the dense gradient of the activation is never formed."""

    @staticmethod
    def forward(ctx, h, weight, gamma, beta, bn, slope, graph):
        h, ldh = _mat(h, 'input')
        weight, ldw = _mat(weight, 'weight')
        if h.size(1) != weight.size(1):
            raise _lib.DeltaConvLibraryError(f"Linear: input has {h.size(1)} features, weight expects {weight.size(1)}")
        M, K, C = h.size(0), h.size(1), weight.size(0)
        N, k = graph.N, graph.k
        if M != N * k:
            raise _lib.DeltaConvLibraryError(f"segment-max: {M} edge rows for a graph with {N} x {k} slots")
        dev = h.device
        y = torch.empty(M, C, dtype=_F32, device=dev)
        ws, nb = _gemm_ws(M, C, K, dev)
        call('dc_linear_fwd', ptr(h), ldh, ptr(weight), ldw, None, M, K, C, ptr(y), C, ptr(ws), nb)
        mean, invstd, batch_stats = _bn_moments(bn, y, C, M, C, 0)
        out = torch.empty(N, C, dtype=_F32, device=dev)
        arg = torch.empty(N, C, dtype=torch.uint8, device=dev)
        call('dc_segmax_bn_fwd', ptr(y), N, k, C, ptr(graph.idx), ptr(mean), ptr(invstd), ptr(gamma), ptr(beta), float(slope),
             ptr(out), C, ptr(arg))
        ctx.save_for_backward(h, weight, y, mean, invstd, gamma, beta, arg)
        ctx.slope, ctx.batch_stats, ctx.graph = slope, batch_stats, graph
        return out

    @staticmethod
    def backward(ctx, dout):
        h, weight, y, mean, invstd, gamma, beta, arg = ctx.saved_tensors
        g = ctx.graph
        h, ldh = _mat(h)
        weight, ldw = _mat(weight)
        dout, ldd = _mat(dout, 'grad_output')
        M, K, C = h.size(0), h.size(1), weight.size(0)
        dev = h.device
        dgamma = torch.empty(C, dtype=_F32, device=dev)
        dbeta = torch.empty(C, dtype=_F32, device=dev)
        dy = torch.empty(M, C, dtype=_F32, device=dev)
        nb = query('dc_colreduce_ws_bytes', C)
        ws = workspace(nb, dev)
        call('dc_segmax_bn_bwd', ptr(y), g.N, g.k, C, ptr(arg), ptr(dout), ldd, ptr(mean), ptr(invstd), ptr(gamma), ptr(beta),
             float(ctx.slope), int(ctx.batch_stats), ptr(dgamma), ptr(dbeta), ptr(dy), ptr(ws), nb)
        dh, dw = _linear_backward_pair(dy, C, weight, ldw, h, ldh, M, K, C, ctx.needs_input_grad[0], ctx.needs_input_grad[1])
        return (dh, dw, dgamma if gamma is not None else None, dbeta if beta is not None else None, None, None, None)


def linear_bn_act_segmax(h, weight, bn, slope, graph):
    return _LinearBNActSegMax.apply(h, weight, bn.weight, bn.bias, bn, float(slope), graph)


def segmax_fusable(lin, slope):
    """The last block of a centralized edge MLP can run as the fused node (same conditions as gathermax_fusable)."""
    return lin.bias is None and slope is not None and slope > 0 and lin.out_features % 4 == 0


class _Pool(torch.autograd.Function):
    """mode 'max' | 'mean' | 'maxmean' ([max | mean] side by side, deltanet_classification.py:46-49)."""

    @staticmethod
    def forward(ctx, x, cloud_ptr, B, mode):
        x, ldx = _mat(x, 'x')
        N, C = x.shape
        want_max, want_mean = mode in ('max', 'maxmean'), mode in ('mean', 'maxmean')
        width = C * (2 if mode == 'maxmean' else 1)
        out = torch.empty(B, width, dtype=_F32, device=x.device)
        arg = torch.empty(B, C, dtype=_I32, device=x.device) if want_max else None
        mean_p = (out.data_ptr() + (4 * C if mode == 'maxmean' else 0)) if want_mean else None
        call('dc_pool_fwd', ptr(x), ldx, ptr(cloud_ptr), B, C, ptr(out) if want_max else None, width, mean_p, width, ptr(arg))
        ctx.mode, ctx.N, ctx.C, ctx.B = mode, N, C, B
        ctx.save_for_backward(cloud_ptr, arg)
        return out

    @staticmethod
    def backward(ctx, dout):
        cloud_ptr, arg = ctx.saved_tensors
        dout, ld = _mat(dout, 'grad_output')
        N, C, B, mode = ctx.N, ctx.C, ctx.B, ctx.mode
        dx = torch.empty(N, C, dtype=_F32, device=dout.device)
        dmax = dout.data_ptr() if mode in ('max', 'maxmean') else None
        dmean = (dout.data_ptr() + (4 * C if mode == 'maxmean' else 0)) if mode in ('mean', 'maxmean') else None
        call('dc_pool_bwd', ptr(cloud_ptr), B, N, C, ptr(arg), dmax, ld, dmean, ld, ptr(dx), C)
        return dx, None, None, None


def pool(x, batch, mode, size=None):
    cp, B = cloud_ptr_of(batch, x.size(0), x.device, size)
    return _Pool.apply(x, cp, B, mode)


class _LinearBNActPool(torch.autograd.Function):
    """pool(act(BatchNorm1d(x @ W.T))) with batch statistics — the embedding layer and the global pools that consume it
    (deltanet_classification.py:45-49, deltanet_segmentation.py:58-61) as one node: the [N,C] activation and its
    gradient are never written (dc_bn_act_pool_fwd / dc_bn_act_pool_bwd)."""

    @staticmethod
    def forward(ctx, x, weight, gamma, beta, bn, slope, cloud_ptr, B, mode):
        x, ldx = _mat(x, 'input')
        weight, ldw = _mat(weight, 'weight')
        if x.size(1) != weight.size(1):
            raise _lib.DeltaConvLibraryError(f"Linear: input has {x.size(1)} features, weight expects {weight.size(1)}")
        M, K, C = x.size(0), x.size(1), weight.size(0)
        if M == 0:
            raise _lib.DeltaConvLibraryError("BatchNorm1d: empty batch")
        dev = x.device
        y = torch.empty(M, C, dtype=_F32, device=dev)
        mean = torch.empty(C, dtype=_F32, device=dev)
        invstd = torch.empty(C, dtype=_F32, device=dev)
        update, momentum, nbt = _running_update(bn)
        ws, nb = _gemm_ws(M, C, K, dev)
        call('dc_linear_fwd', ptr(x), ldx, ptr(weight), ldw, None, M, K, C, ptr(y), C, ptr(ws), nb)
        nb = query('dc_colreduce_ws_bytes', C)
        ws = workspace(nb, dev)
        call('dc_bn_stats', ptr(y), C, M, C, 0, float(bn.eps), float(momentum), ptr(mean), ptr(invstd),
             ptr(bn.running_mean) if update else None, ptr(bn.running_var) if update else None, nbt, ptr(ws), nb)
        want_max, want_mean = mode in ('max', 'maxmean'), mode in ('mean', 'maxmean')
        width = C * (2 if mode == 'maxmean' else 1)
        out = torch.empty(B, width, dtype=_F32, device=dev)
        arg = torch.empty(B, C, dtype=_I32, device=dev) if want_max else None
        mean_p = (out.data_ptr() + (4 * C if mode == 'maxmean' else 0)) if want_mean else None
        call('dc_bn_act_pool_fwd', ptr(y), C, ptr(cloud_ptr), B, C, ptr(mean), ptr(invstd), ptr(gamma), ptr(beta), float(slope),
             ptr(out) if want_max else None, width, mean_p, width, ptr(arg))
        ctx.save_for_backward(x, weight, y, mean, invstd, gamma, beta, cloud_ptr, arg)
        ctx.slope, ctx.mode, ctx.B = slope, mode, B
        return out

    @staticmethod
    def backward(ctx, dout):
        x, weight, y, mean, invstd, gamma, beta, cloud_ptr, arg = ctx.saved_tensors
        x, ldx = _mat(x)
        weight, ldw = _mat(weight)
        dout, ld = _mat(dout, 'grad_output')
        M, K, C = x.size(0), x.size(1), weight.size(0)
        dev, mode = x.device, ctx.mode
        dgamma = torch.empty(C, dtype=_F32, device=dev)
        dbeta = torch.empty(C, dtype=_F32, device=dev)
        dy = torch.empty(M, C, dtype=_F32, device=dev)
        dmax = dout.data_ptr() if mode in ('max', 'maxmean') else None
        dmean = (dout.data_ptr() + (4 * C if mode == 'maxmean' else 0)) if mode in ('mean', 'maxmean') else None
        nb = query('dc_colreduce_ws_bytes', C)
        ws = workspace(nb, dev)
        call('dc_bn_act_pool_bwd', ptr(y), C, M, C, ptr(cloud_ptr), ctx.B, ptr(arg), dmax, ld, dmean, ld,
             ptr(mean), ptr(invstd), ptr(gamma), ptr(beta), float(ctx.slope), ptr(dgamma), ptr(dbeta), ptr(dy), C, ptr(ws), nb)
        dx, dw = _linear_backward_pair(dy, C, weight, ldw, x, ldx, M, K, C, ctx.needs_input_grad[0], ctx.needs_input_grad[1])
        return (dx, dw, dgamma if gamma is not None else None, dbeta if beta is not None else None) + (None,) * 5


def linear_bn_act_pool(x, weight, bn, slope, batch, mode, size=None):
    """pool(act(BatchNorm1d(x @ weight.T)), batch) with batch statistics; mode 'max' | 'mean' | 'maxmean'.
    Needs out_features % 4 == 0 (see `pool_fusable`)."""
    cp, B = cloud_ptr_of(batch, x.size(0), x.device, size)
    return _LinearBNActPool.apply(x, weight, bn.weight, bn.bias, bn, float(slope), cp, B, mode)


def linear_bn_act_pool_inference(x, weight, bn, slope, batch, mode, size=None):
    """The same fused Linear -> BatchNorm1d -> LeakyReLU -> pool with the layer's RUNNING statistics (evaluation mode),
    for forward passes that record no autograd tape: the [N,C] activation is never written."""
    cp, B = cloud_ptr_of(batch, x.size(0), x.device, size)
    x, ldx = _mat(x, 'input')
    weight, ldw = _mat(weight, 'weight')
    M, K, C = x.size(0), x.size(1), weight.size(0)
    dev = x.device
    y = torch.empty(M, C, dtype=_F32, device=dev)
    ws, nb = _gemm_ws(M, C, K, dev)
    call('dc_linear_fwd', ptr(x), ldx, ptr(weight), ldw, None, M, K, C, ptr(y), C, ptr(ws), nb)
    mean = bn.running_mean.contiguous()
    invstd = _running_invstd(bn)
    want_max, want_mean = mode in ('max', 'maxmean'), mode in ('mean', 'maxmean')
    width = C * (2 if mode == 'maxmean' else 1)
    out = torch.empty(B, width, dtype=_F32, device=dev)
    arg = torch.empty(B, C, dtype=_I32, device=dev) if want_max else None
    mean_p = (out.data_ptr() + (4 * C if mode == 'maxmean' else 0)) if want_mean else None
    call('dc_bn_act_pool_fwd', ptr(y), C, ptr(cp), B, C, ptr(mean), ptr(invstd), ptr(bn.weight), ptr(bn.bias), float(slope),
         ptr(out) if want_max else None, width, mean_p, width, ptr(arg))
    return out


def pool_fusable_inference(bn, out_features, bias, slope, x):
    running = not (bn.training or not bn.track_running_stats or bn.running_mean is None)
    no_tape = not (torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in bn.parameters())))
    return running and no_tape and bias is None and slope is not None and out_features % 4 == 0


def pool_fusable(bn, out_features, bias, slope):
    """True when Linear -> BatchNorm1d -> LeakyReLU -> pool can run as the fused node: training-mode statistics, no
    Linear bias, a LeakyReLU-family activation and a channel count the float4 kernels take."""
    batch_stats = bn.training or not bn.track_running_stats or bn.running_mean is None
    return batch_stats and bias is None and slope is not None and out_features % 4 == 0


class _BcastRows(torch.autograd.Function):
    @staticmethod
    def forward(ctx, src, cloud_ptr, B, N, slot=None):
        src, lds = _mat(src, 'src')
        C = src.size(1)
        out = _out_block(slot, N, C, src.device)
        call('dc_bcast_rows', ptr(src), lds, ptr(cloud_ptr), B, N, C, ptr(out), out.stride(0) if N > 1 else C)
        ctx.B, ctx.N = B, N
        ctx.save_for_backward(cloud_ptr)
        return out

    @staticmethod
    def backward(ctx, dout):
        (cloud_ptr,) = ctx.saved_tensors
        dout, ld = _mat(dout, 'grad_output')
        C = dout.size(1)
        # per-cloud column sums = mean-pool backward's transpose: reuse the pool kernel (mean * count)
        dsrc = torch.empty(ctx.B, C, dtype=_F32, device=dout.device)
        call('dc_pool_fwd', ptr(dout), ld, ptr(cloud_ptr), ctx.B, C, None, 0, ptr(dsrc), C, None)
        counts = (cloud_ptr[1:] - cloud_ptr[:-1]).to(_F32)
        return dsrc * counts[:, None], None, None, None, None


class _LinearBcastBNAct(torch.autograd.Function):
    """act(BatchNorm1d([g[batch] | x] @ W.T)) — the first block of the segmentation head (deltanet_segmentation.py:61-66: the
    pooled embedding and the category features are broadcast to every point and concatenated with the per-point features)
    WITHOUT the broadcast: the product splits into x @ W_x.T over the N points plus (g @ W_g.T)[batch] over the B clouds, so
    the [N, Kg] broadcast block is never written, never multiplied and never receives a gradient of its own (Kg = 1088 of the
    1536 input columns in the ShapeNet configuration).  Weight layout as in the reference: columns [0,Kg) multiply g."""

    @staticmethod
    def forward(ctx, x, g, weight, gamma, beta, bn, slope, cloud_ptr, B):
        x, ldx = _mat(x, 'input')
        g, ldg = _mat(g, 'per-cloud input')
        _f32(weight, 'weight')
        N, Kx, Kg, C = x.size(0), x.size(1), g.size(1), weight.size(0)
        if weight.dim() != 2 or weight.size(1) != Kg + Kx:
            raise _lib.DeltaConvLibraryError(f"Linear: input has {Kg} + {Kx} features, weight expects {weight.size(-1)}")
        if g.size(0) != B:
            raise _lib.DeltaConvLibraryError(f"broadcast: {g.size(0)} per-cloud rows for {B} clouds")
        weight = weight.contiguous()
        Wg, ldw = _mat(weight[:, :Kg])
        Wx, _ = _mat(weight[:, Kg:])
        dev = x.device
        yg = torch.empty(B, C, dtype=_F32, device=dev)
        ws, nb = _gemm_ws(B, C, Kg, dev)
        call('dc_linear_fwd', ptr(g), ldg, ptr(Wg), ldw, None, B, Kg, C, ptr(yg), C, ptr(ws), nb)
        y = torch.empty(N, C, dtype=_F32, device=dev)
        ws, nb = _gemm_ws(N, C, Kx, dev)
        call('dc_linear_fwd', ptr(x), ldx, ptr(Wx), ldw, None, N, Kx, C, ptr(y), C, ptr(ws), nb)
        call('dc_bcast_rows_add', ptr(yg), C, ptr(cloud_ptr), B, N, C, ptr(y), C)
        mean, invstd, batch_stats = _bn_moments(bn, y, C, N, C, 0)
        out = torch.empty(N, C, dtype=_F32, device=dev)
        call('dc_bn_act_fwd', ptr(y), C, N, C, ptr(mean), ptr(invstd), ptr(gamma), ptr(beta), float(slope), None, 0, ptr(out), C)
        ctx.save_for_backward(x, g, weight, y, mean, invstd, gamma, beta, cloud_ptr)
        ctx.slope, ctx.batch_stats, ctx.B = slope, batch_stats, B
        return out

    @staticmethod
    def backward(ctx, dout):
        x, g, weight, y, mean, invstd, gamma, beta, cloud_ptr = ctx.saved_tensors
        x, ldx = _mat(x)
        g, ldg = _mat(g)
        dout, ldd = _mat(dout, 'grad_output')
        N, Kx, Kg, C, B = x.size(0), x.size(1), g.size(1), weight.size(0), ctx.B
        Wg, ldw = _mat(weight[:, :Kg])
        Wx, _ = _mat(weight[:, Kg:])
        dev = x.device
        dgamma = torch.empty(C, dtype=_F32, device=dev)
        dbeta = torch.empty(C, dtype=_F32, device=dev)
        dy = torch.empty(N, C, dtype=_F32, device=dev)
        nb = query('dc_colreduce_ws_bytes', C)
        ws = workspace(nb, dev)
        call('dc_bn_act_bwd', ptr(y), C, ptr(dout), ldd, N, C, ptr(mean), ptr(invstd), ptr(gamma), ptr(beta), float(ctx.slope),
             int(ctx.batch_stats), ptr(dgamma), ptr(dbeta), ptr(dy), C, ptr(ws), nb)
        # gradient of the per-cloud rows: column sums of dy over each cloud (the mean pool's kernel, times the cloud sizes)
        dyg = torch.empty(B, C, dtype=_F32, device=dev)
        call('dc_pool_fwd', ptr(dy), C, ptr(cloud_ptr), B, C, None, 0, ptr(dyg), C, None)
        dyg = dyg * (cloud_ptr[1:] - cloud_ptr[:-1]).to(_F32)[:, None]
        want_x, want_g, want_w = ctx.needs_input_grad[0], ctx.needs_input_grad[1], ctx.needs_input_grad[2]
        dx = dg = dw = None
        if want_w:
            dw = torch.empty(C, Kg + Kx, dtype=_F32, device=dev)
            ws_w, nb_w = _gemm_ws(C, Kx, N, dev)
            call('dc_linear_bwd_weight', ptr(dy), C, ptr(x), ldx, N, Kx, C, ptr(dw[:, Kg:]), Kg + Kx, ptr(ws_w), nb_w)
            ws_w, nb_w = _gemm_ws(C, Kg, B, dev)
            call('dc_linear_bwd_weight', ptr(dyg), C, ptr(g), ldg, B, Kg, C, ptr(dw), Kg + Kx, ptr(ws_w), nb_w)
        if want_x:
            dx = torch.empty(N, Kx, dtype=_F32, device=dev)
            ws_x, nb_x = _gemm_ws(N, Kx, C, dev)
            call('dc_linear_bwd_input', ptr(dy), C, ptr(Wx), ldw, N, Kx, C, ptr(dx), Kx, ptr(ws_x), nb_x)
        if want_g:
            dg = torch.empty(B, Kg, dtype=_F32, device=dev)
            ws_x, nb_x = _gemm_ws(B, Kg, C, dev)
            call('dc_linear_bwd_input', ptr(dyg), C, ptr(Wg), ldw, B, Kg, C, ptr(dg), Kg, ptr(ws_x), nb_x)
        return (dx, dg, dw, dgamma if gamma is not None else None, dbeta if beta is not None else None, None, None, None, None)


def linear_bcast_bn_act(x, g, weight, bn, slope, batch, num_clouds=None):
    """act(BatchNorm1d(cat([g[batch], x], 1) @ weight.T)) without forming g[batch]: x [N,Kx] per point, g [B,Kg] per cloud."""
    cp, B = cloud_ptr_of(batch, x.size(0), x.device, num_clouds)
    return _LinearBcastBNAct.apply(x, g, weight, bn.weight, bn.bias, bn, float(slope), cp, B)


def broadcast_rows(src, batch, num_points, slot=None):
    cp, B = cloud_ptr_of(batch, num_points, src.device, src.size(0))
    return _BcastRows.apply(src, cp, B, num_points, slot)
