"""Augmentation / normalisation transforms with the reference's names and arguments (deltaconv/transforms/*.py), written
so that they also run on a whole device-resident `Batch`: with a `batch` vector every cloud draws its own random
parameters and all clouds are transformed in one vectorised pass on the GPU; without one they act on a
single cloud exactly like the reference's per-sample CPU transforms.  Plain torch — they are not on the hot path.

`GeodesicFPS` is the exception: it runs on the CUDA library (dc_geodesic_fps), on one cloud or on a whole batch.
`SamplePoints` and `NormalizeArea` are the reference's two mesh pre-transforms (`data.face` [3,F]); they act on one mesh.
"""
import math
import numbers

import torch

__all__ = ['NormalizeScale', 'NormalizeAxes', 'RandomScale', 'RandomTranslateGlobal', 'RandomRotate', 'RandomNormals', 'GeodesicFPS',
           'SamplePoints', 'NormalizeArea', 'MortonOrder']


def _clouds(data):
    """(cloud index per point [N] int64, number of clouds) — one cloud when the object carries no batch vector."""
    pos = data.pos
    batch = getattr(data, 'batch', None)
    if batch is None:
        return torch.zeros(pos.size(0), dtype=torch.long, device=pos.device), 1
    hint = getattr(data, '_num_clouds', None)
    return batch, int(hint) if hint is not None else (int(batch.max()) + 1 if batch.numel() else 1)


def _per_cloud_extreme(values, batch, B, largest):
    """max (or min) over the points of every cloud -> [B, C]"""
    fill = float('-inf') if largest else float('inf')
    out = torch.full((B, values.size(1)), fill, dtype=values.dtype, device=values.device)
    index = batch[:, None].expand_as(values)
    return out.scatter_reduce(0, index, values, 'amax' if largest else 'amin', include_self=True)


def _has_normals(data):
    return getattr(data, 'norm', None) is not None


def _per_axis(value, dim):
    if isinstance(value, numbers.Number):
        return [float(value)] * dim
    assert len(value) == dim
    return [float(v) for v in value]


class NormalizeScale(object):
    """Centre every cloud on the middle of its bounding box and scale it into the unit ball (normalize_scale.py:14-22)."""

    def __init__(self, norm_ord=2, scaling_factor=None):
        self.norm_ord = norm_ord
        self.scaling_factor = scaling_factor

    def __call__(self, data):
        batch, B = _clouds(data)
        pos = data.pos
        centre = (_per_cloud_extreme(pos, batch, B, True) + _per_cloud_extreme(pos, batch, B, False)) / 2
        pos = pos - centre[batch]
        if self.scaling_factor is None:
            radius = torch.linalg.norm(pos, ord=self.norm_ord, dim=1, keepdim=True)
            scale = (1 / _per_cloud_extreme(radius, batch, B, True)) * 0.999999
            pos = pos * scale[batch]
        else:
            pos = pos * ((1 / self.scaling_factor) * 0.999999)
        data.pos = pos
        return data

    def __repr__(self):
        return '{}()'.format(self.__class__.__name__)


class NormalizeAxes(object):
    """Permute the axes of every cloud from smallest to widest standard deviation, then scale so that the last axis
    reaches 1/2 (normalize_axes.py:17-26)."""

    def __init__(self, max_points=-1):
        self.max_points = max_points

    def __call__(self, data):
        batch, B = _clouds(data)
        pos = data.pos
        count = torch.zeros(B, 1, dtype=pos.dtype, device=pos.device).index_add_(0, batch, torch.ones_like(pos[:, :1]))
        mean = torch.zeros(B, pos.size(1), dtype=pos.dtype, device=pos.device).index_add_(0, batch, pos) / count.clamp(min=1)
        var = torch.zeros_like(mean).index_add_(0, batch, (pos - mean[batch]) ** 2) / (count - 1).clamp(min=1)     # unbiased, as torch.std
        order = torch.sort(var, dim=1)[1]                                  # sorting by variance = sorting by std
        pos = torch.gather(pos, 1, order[batch])
        top = _per_cloud_extreme(pos, batch, B, True)[:, -1:]
        data.pos = pos * (1 / (2 * top))[batch]
        return data

    def __repr__(self):
        return '{}()'.format(self.__class__.__name__)


class RandomScale(object):
    """Scale every cloud by its own random factor per axis drawn from `scales = (a, b)`; normals are transformed with the
    inverse scale and renormalised (random_scale.py:28-34)."""

    def __init__(self, scales):
        assert isinstance(scales, (tuple, list)) and len(scales) == 2
        self.scales = scales

    def __call__(self, data):
        batch, B = _clouds(data)
        scale = data.pos.new_empty(B, data.pos.size(1)).uniform_(*self.scales)
        data.pos = data.pos * scale[batch]
        if _has_normals(data):
            n = data.norm * (1 / scale)[batch]
            data.norm = n / torch.linalg.norm(n, dim=1, keepdim=True)
        return data

    def __repr__(self):
        return '{}({})'.format(self.__class__.__name__, self.scales)


class RandomTranslateGlobal(object):
    """Translate every cloud as a whole by its own random offset in (-translate, +translate) per axis
    (random_translate_global.py:23-34)."""

    def __init__(self, translate):
        self.translate = translate

    def __call__(self, data):
        batch, B = _clouds(data)
        limit = data.pos.new_tensor([abs(t) for t in _per_axis(self.translate, data.pos.size(1))])
        offset = (data.pos.new_empty(B, data.pos.size(1)).uniform_(-1, 1)) * limit
        data.pos = data.pos + offset[batch]
        return data

    def __repr__(self):
        return '{}({})'.format(self.__class__.__name__, self.translate)


class RandomRotate(object):
    """Rotate every cloud (positions and normals) about `axis` by its own random angle from `degrees`
    (random_rotate.py:20-45; same matrices, applied as `pos @ matrix`)."""

    def __init__(self, degrees, axis=0):
        if isinstance(degrees, numbers.Number):
            degrees = (-abs(degrees), abs(degrees))
        assert isinstance(degrees, (tuple, list)) and len(degrees) == 2
        self.degrees = degrees
        self.axis = axis

    def __call__(self, data):
        batch, B = _clouds(data)
        pos = data.pos
        angle = pos.new_empty(B).uniform_(*self.degrees) * (math.pi / 180.0)
        sin, cos = torch.sin(angle), torch.cos(angle)
        one, zero = torch.ones_like(sin), torch.zeros_like(sin)
        if pos.size(-1) == 2:
            rows = [[cos, sin], [-sin, cos]]
        elif self.axis == 0:
            rows = [[one, zero, zero], [zero, cos, sin], [zero, -sin, cos]]
        elif self.axis == 1:
            rows = [[cos, zero, -sin], [zero, one, zero], [sin, zero, cos]]
        else:
            rows = [[cos, sin, zero], [-sin, cos, zero], [zero, zero, one]]
        matrix = torch.stack([torch.stack(r, dim=-1) for r in rows], dim=-2)          # [B, d, d]
        per_point = matrix[batch]
        data.pos = torch.bmm(pos.unsqueeze(1), per_point).squeeze(1)
        if _has_normals(data):
            data.norm = torch.bmm(data.norm.unsqueeze(1), per_point.to(data.norm.dtype)).squeeze(1)
        return data

    def __repr__(self):
        return '{}({}, axis={})'.format(self.__class__.__name__, self.degrees, self.axis)


class RandomNormals(object):
    """Jitter every normal by a random offset in (-translate, +translate) per axis and renormalise
    (random_normals.py:23-35)."""

    def __init__(self, translate):
        self.translate = translate

    def __call__(self, data):
        n = data.norm
        limit = n.new_tensor([abs(t) for t in _per_axis(self.translate, n.size(1))])
        n = n + n.new_empty(n.shape).uniform_(-1, 1) * limit
        data.norm = n / torch.linalg.norm(n, dim=-1, keepdim=True).clamp(1e-5)
        return data

    def __repr__(self):
        return '{}({})'.format(self.__class__.__name__, self.translate)


class GeodesicFPS(object):
    """Keep `n_samples` points of every cloud, chosen by geodesic farthest-point sampling (geodesic_fps.py:14-43; the
    sampler is the CUDA one, `ops.geodesic_fps`).  On a single cloud the fields are handled exactly like the reference's:
    `sample_idx` is stored, `pos`, `norm` (or `normal`, moved to `norm`), `x` and a per-point `y` are indexed, a cloud
    smaller than `n_samples` is tiled.  With a `batch` vector every cloud is sampled in the same launch and `batch` is
    rebuilt; `sample_idx` then indexes the rows of the incoming batch."""

    def __init__(self, n_samples=None, store_original=False):
        self.n_samples = n_samples
        self.store_original = store_original

    def __call__(self, data):
        from . import ops
        pos = data.pos
        if self.n_samples is None:
            self.n_samples = pos.size(0)
        n = self.n_samples
        batch = getattr(data, 'batch', None)
        device_pos = pos if pos.is_cuda else pos.cuda()
        if batch is None:
            P = pos.size(0)
            idx = ops.geodesic_fps(device_pos, n)[0]
            if P < n:
                idx = idx[:P].repeat(math.ceil(n / P))
            idx = idx[:n]
        else:
            device_batch = batch if batch.is_cuda else batch.cuda()
            cp, B, _ = ops.cloud_layout(device_batch, pos.size(0), device_pos.device, getattr(data, '_num_clouds', None),
                                        getattr(data, '_max_points', None))
            local = ops.geodesic_fps(device_pos, n, batch=device_batch, num_clouds=B)           # [B, n]
            sizes = (cp[1:] - cp[:-1]).long()
            # a cloud smaller than n: its first P samples, tiled (geodesic_fps.py:24-25)
            slot = torch.arange(n, device=local.device)[None, :] % sizes.clamp(min=1)[:, None]
            idx = (torch.gather(local, 1, slot) + cp[:-1].long()[:, None]).reshape(-1)
            data.batch = torch.arange(B, device=batch.device).repeat_interleave(n)
            if hasattr(data, '_max_points'):
                data._num_clouds, data._max_points = B, n
        idx = idx.to(pos.device)
        assert idx.max() <= pos.size(0)
        assert idx.min() >= 0
        data.sample_idx = idx
        if self.store_original:
            data.pos_original = data.pos
            data.y_original = data.y
        data.pos = pos[idx]
        if getattr(data, 'norm', None) is not None:
            data.norm = data.norm[idx]
        if getattr(data, 'normal', None) is not None:
            data.norm = data.normal[idx]
        if getattr(data, 'x', None) is not None:
            data.x = data.x[idx]
        y = getattr(data, 'y', None)
        if y is not None and type(y) is not int and y.size(0) > 1 and y.size(0) == pos.size(0):
            data.y = y[idx]
        return data

    def __repr__(self):
        return '{}()'.format(self.__class__.__name__)


class SamplePoints(object):
    """Draw `num` points on the faces of a mesh, faces chosen in proportion to their area, uniformly inside each face
    (sample_points.py:23-60).  The random draws come in the reference's order — one `multinomial` over the faces, then one
    `rand(num, 2)` — so a seeded run gives the reference's points.  `include_normals` stores the face normal of every
    sample in `data.norm`; `include_labels` takes the per-vertex label of the sampled face's first corner."""

    def __init__(self, num, remove_faces=True, include_normals=False, include_labels=False):
        self.num = num
        self.remove_faces = remove_faces
        self.include_normals = include_normals
        self.include_labels = include_labels

    def __call__(self, data):
        vertices, face = data.pos, data.face
        assert vertices.size(1) == 3 and face.size(0) == 3
        scale = vertices.max()                                    # the reference samples on the mesh scaled by its largest coordinate
        vertices = vertices / scale
        corner = vertices[face[0]]
        doubled_area = torch.linalg.cross(vertices[face[1]] - corner, vertices[face[2]] - corner, dim=1).norm(p=2, dim=1)
        weight = doubled_area.abs() / 2
        chosen = face[:, torch.multinomial(weight / weight.sum(), self.num, replacement=True)]
        st = torch.rand(self.num, 2, device=vertices.device)
        outside = st.sum(dim=-1) > 1                              # fold the unit square onto the triangle
        st[outside] = 1 - st[outside]
        origin = vertices[chosen[0]]
        edge_a, edge_b = vertices[chosen[1]] - origin, vertices[chosen[2]] - origin
        if self.include_normals:
            data.norm = torch.nn.functional.normalize(torch.linalg.cross(edge_a, edge_b, dim=1), p=2)
        data.pos = (origin + st[:, :1] * edge_a + st[:, 1:] * edge_b) * scale
        if self.include_labels:
            data.y = data.y[chosen[0]]
        if self.remove_faces:
            data.face = None
        return data

    def __repr__(self):
        return '{}({})'.format(self.__class__.__name__, self.num)


class NormalizeArea(object):
    """Centre a mesh on the middle of its bounding box and scale it by 1/sqrt(area) (normalize_area.py:12-20).  Kept
    result-compatible with the reference, which indexes `data.face[:, 0..2]` on the [3,F] face tensor: the "area" it sums is
    that of the three triangles spanned by the corner triples of faces 0, 1 and 2, not the surface area of the mesh."""

    def __call__(self, data):
        pos = data.pos
        pos = pos - (pos.max(dim=0)[0] + pos.min(dim=0)[0]) / 2
        first, second, third = pos[data.face[:, 0]], pos[data.face[:, 1]], pos[data.face[:, 2]]
        summed = torch.linalg.norm(torch.linalg.cross(second - first, third - first, dim=-1), dim=-1).sum() / 2
        data.pos = pos * (1 / torch.sqrt(summed))
        return data

    def __repr__(self):
        return '{}()'.format(self.__class__.__name__)


class MortonOrder(object):
    """Pre-transform without a counterpart in the reference: reorder the points of every cloud along a Morton (Z-order) curve
    (`ops.spatial_order`).  Clouds stay where they are and `batch` is unchanged; `pos`, `norm` / `normal`, `x` and per-point
    `y` are permuted together, `morton_perm` keeps the permutation.  Consecutive rows then are spatial neighbours, so the
    kNN graph's gathers (operator build, SpMMs, neighbourhood max) hit the SM's L1 instead of crossing to L2: 2x on the SpMMs
    of a 2^20-point cloud, 1.6x on the operator build.  Results are those of the permuted cloud: a network output per cloud is
    unchanged up to fp32 summation order, per-point outputs come back in the new order (exact distance ties in the kNN
    resolve by index, so a reordering can change which of two equidistant neighbours is kept)."""

    def __call__(self, data):
        from . import ops
        pos = data.pos
        device_pos = pos if pos.is_cuda else pos.cuda()
        batch = getattr(data, 'batch', None)
        device_batch = None if batch is None else (batch if batch.is_cuda else batch.cuda())
        perm = ops.spatial_order(device_pos.float(), device_batch, getattr(data, '_num_clouds', None)).to(pos.device)
        n = pos.size(0)
        for name in ('pos', 'norm', 'normal', 'x', 'y'):
            t = getattr(data, name, None)
            if torch.is_tensor(t) and t.dim() >= 1 and t.size(0) == n:          # per-point fields only (a per-cloud `y` has B rows)
                setattr(data, name, t[perm].contiguous())
        data.morton_perm = perm
        return data

    def __repr__(self):
        return f'{self.__class__.__name__}()'

