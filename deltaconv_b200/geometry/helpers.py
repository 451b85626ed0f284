"""The reference's intermediate geometry helpers (deltaconv/geometry/grad_div_mls.py:72-194) with their names and
signatures, as device-agnostic torch expressions.  `build_grad_div` does NOT call them — the CUDA build kernel fuses
all four steps per point (csrc/build.cu) — they exist so that code written against the reference's public API (its
tests, notebooks) keeps working, on CPU or GPU tensors alike.  Edge tensors are [E] / [E,*] in the row-grouped
order of `edge_index` (E = N*k)."""
import torch

__all__ = ['coords_projected', 'gaussian_weights', 'weighted_least_squares', 'fit_vector_mapping']

_EPS = 1e-5


def _k_of(edge_index, k):
    return int((edge_index[0] == 0).sum()) if k is None else int(k)


def coords_projected(pos, normal, x_basis, y_basis, edge_index, k=None):
    """Tangent-plane coordinates (u, v) of every neighbour in its centre's frame -> [E,2] (grad_div_mls.py:72-97)."""
    row, col = edge_index
    offset = pos[col] - pos[row]
    n = normal[row]
    in_plane = offset - n * (offset * n).sum(-1, keepdim=True)
    return torch.stack([(in_plane * x_basis[row]).sum(-1), (in_plane * y_basis[row]).sum(-1)], dim=1)


def gaussian_weights(dist, k, batch=None, kernel_width=1):
    """Gaussian edge weights, kernel size relative to the mean edge length of each cloud, normalised per
    neighbourhood -> [E] (grad_div_mls.py:100-116)."""
    k = int(k)
    d = dist.view(-1, k)
    N = d.size(0)
    if batch is None:
        batch = torch.zeros(N, dtype=torch.long, device=dist.device)
    B = int(batch.max()) + 1 if N else 0
    per_point = d.mean(dim=1)
    total = torch.zeros(B, dtype=d.dtype, device=d.device).index_add(0, batch, per_point)
    count = torch.zeros(B, dtype=d.dtype, device=d.device).index_add(0, batch, torch.ones_like(per_point)).clamp(min=1)
    scale = (kernel_width * total / count)[batch].unsqueeze(1)
    w = torch.exp(-(d * d) / (scale * scale))
    return (w / w.sum(dim=1, keepdim=True).clamp(_EPS)).reshape(-1)


def weighted_least_squares(coords, weights, k, regularizer, shape_regularizer=None):
    """Rows of the regularised pseudo-inverse of the quadric fit [1,u,v,u^2,uv,v^2] per edge -> [E,6]; with a second
    regulariser also the variant used for the surface fit (grad_div_mls.py:119-152)."""
    k = int(k)
    out_dtype = coords.dtype
    # The normal equations are formed and inverted in fp64 whatever the input precision: with regularizer -> 0 (as the
    # reference's own tests call this) the 6x6 systems are badly conditioned, and a batched fp32 inverse on the GPU does not
    # reach the 1e-6 the reference's CPU LAPACK path delivers.
    u, v = coords[:, 0].double(), coords[:, 1].double()
    basis = torch.stack([torch.ones_like(u), u, v, u * u, u * v, v * v], dim=1).view(-1, k, 6)
    wb = basis * weights.double().view(-1, k, 1)
    gram = torch.einsum('nki,nkj->nij', wb, basis)
    # the reference adds `regularizer * torch.eye(6)` built in the default dtype: lambda is rounded to fp32
    eye = torch.eye(6, device=coords.device, dtype=torch.float32)

    def solve(reg):
        inv = torch.linalg.inv(gram + (reg * eye).to(gram.dtype))
        return torch.einsum('nij,nkj->nki', inv, wb).reshape(-1, 6).to(out_dtype)
    if shape_regularizer is None:
        return solve(regularizer)
    return solve(regularizer), solve(shape_regularizer)


def fit_vector_mapping(pos, normal, x_basis, y_basis, edge_index, wls, coords):
    """2x2 map from each neighbour's tangent frame to the centre's frame pushed along the fitted surface -> [E,2,2]
    (grad_div_mls.py:155-194, eq. 15 of the DeltaConv supplement)."""
    row, col = edge_index
    N = pos.size(0)
    n = normal[row]
    height = (n * (pos[col] - pos[row])).sum(-1, keepdim=True)
    coeff = torch.zeros(N, 6, dtype=wls.dtype, device=wls.device).index_add(0, row, wls * height)[row]
    u, v = coords[:, 0], coords[:, 1]
    hx = coeff[:, 1] + 2 * coeff[:, 3] * u + coeff[:, 4] * v
    hy = coeff[:, 2] + coeff[:, 4] * u + 2 * coeff[:, 5] * v
    gam_x = x_basis[row] + n * hx.unsqueeze(-1)
    gam_y = y_basis[row] + n * hy.unsqueeze(-1)
    det = 1 + hx * hx + hy * hy
    inv_metric = torch.stack([1 + hy * hy, -hx * hy, -hx * hy, 1 + hx * hx], dim=-1).view(-1, 2, 2) / det.view(-1, 1, 1)
    xj, yj = x_basis[col], y_basis[col]
    change = torch.stack([(gam_x * xj).sum(-1), (gam_x * yj).sum(-1), (gam_y * xj).sum(-1), (gam_y * yj).sum(-1)], dim=-1).view(-1, 2, 2)
    return torch.bmm(inv_metric, change)
