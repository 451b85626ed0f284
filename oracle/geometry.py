"""Oracle (CPU torch) for the geometric half of the hot path, in fixed-k ELL form.

TEST INFRASTRUCTURE — see oracle/__init__.py.  Everything is dtype-generic (run it in fp64 to
get the reference value the fp32 kernels are compared against).

Layout used throughout (and by the CUDA kernels):
  idx [N,k]   int64  neighbour j of centre i in slot e (slot order = ascending distance)
  G   [N,k,2]        grad entries:  grad[2i+a, idx[i,e]] = G[i,e,a]
  D   [N,k,2]        div  entries:  div[i, 2*idx[i,e]+a] = D[i,e,a]
  vector fields are [2N,C] with rows (2i, 2i+1) = (x, y) components (operators.py:7,12-16).
"""
import torch

EPS = 1e-5  # grad_div_mls.py:7


# --------------------------------------------------------------------------------------
# graph
# --------------------------------------------------------------------------------------

def cloud_ptr(batch, num_points):
    """[B+1] offsets of the (contiguous, sorted) clouds named by `batch` (None -> one cloud)."""
    if batch is None:
        return torch.tensor([0, num_points], dtype=torch.long)
    counts = torch.bincount(batch.cpu())
    return torch.cat([counts.new_zeros(1), counts.cumsum(0)])


def knn_ell(pos, k, batch=None):
    """k nearest neighbours per cloud, self included, as idx [N,k].

    Replaces `knn_graph(pos, k, batch, loop=True, flow='target_to_source')`
    (deltanet_base.py:52,63; third-party torch_cluster, absent from the tree).  Order pinned by
    this repo: ascending (d², index) with
    d² = (dx*dx + dy*dy) + dz*dz, every operation rounded in the input dtype.
    """
    ptr = cloud_ptr(batch, pos.size(0)).tolist()
    out = torch.empty(pos.size(0), k, dtype=torch.long)
    for s, e in zip(ptr[:-1], ptr[1:]):
        if e == s:
            continue
        if e - s < k:
            raise ValueError(f"cloud with {e - s} points has fewer than k={k} points")
        p = pos[s:e]
        dx = p[None, :, 0] - p[:, None, 0]
        dy = p[None, :, 1] - p[:, None, 1]
        dz = p[None, :, 2] - p[:, None, 2]
        d2 = (dx * dx + dy * dy) + dz * dz
        out[s:e] = torch.argsort(d2, dim=1, stable=True)[:, :k] + s
    return out


def edge_index_from_ell(idx):
    """[2,E] int64 (row 0 = centre, row 1 = neighbour) as torch_cluster returns it."""
    N, k = idx.shape
    return torch.stack([torch.arange(N).repeat_interleave(k), idx.reshape(-1)])


def ell_from_edge_index(edge_index, num_points):
    """Inverse of the above under the reference's own assumption (grad_div_mls.py:24-25,224):
    rows grouped by centre, the same k for every centre."""
    k = edge_index.size(1) // num_points
    return edge_index[1].reshape(num_points, k)


# --------------------------------------------------------------------------------------
# tangent frames
# --------------------------------------------------------------------------------------

def _unit(a, floor):
    return a / a.norm(dim=-1, keepdim=True).clamp(floor)


def tangent_basis_from_normal(normal):
    """build_tangent_basis, grad_div_mls.py:50-69."""
    e1 = normal.new_tensor([1.0, 0.0, 0.0]).expand_as(normal)
    e2 = normal.new_tensor([0.0, 1.0, 0.0]).expand_as(normal)
    probe = torch.where((normal[:, :1]).abs() > 0.9, e2, e1)       # |n . e1| = |n_x|
    xb = _unit(torch.linalg.cross(probe, normal, dim=-1), EPS)
    yb = _unit(torch.linalg.cross(normal, xb, dim=-1), EPS)
    return xb, yb


def estimate_basis(pos, idx, orientation=None):
    """estimate_basis, grad_div_mls.py:10-47: left singular vectors of the 3 x k matrix of
    neighbour offsets; normal = smallest, x = largest.  The sign of x is LAPACK's (gauge)."""
    offs = (pos[idx] - pos[:, None, :]).transpose(1, 2)           # [N,3,k]
    U = torch.linalg.svd(offs).U
    normal = _unit(U[:, :, 2], EPS)
    if orientation is not None:
        flip = (normal * orientation).sum(-1, keepdim=True) < 0
        normal = torch.where(flip, -normal, normal)
    xb = _unit(U[:, :, 0], EPS)
    yb = _unit(torch.linalg.cross(normal, xb, dim=-1), EPS)
    return normal, xb, yb


# --------------------------------------------------------------------------------------
# WLS operator fit
# --------------------------------------------------------------------------------------

def fit_operators(pos, normal, xb, yb, idx, batch=None, kernel_width=1.0, lam=1e-3,
                  normalized=True, lam_shape=None, return_parts=False):
    """build_grad_div (grad_div_mls.py:197-277) and its helpers coords_projected (:72-97),
    gaussian_weights (:100-116), weighted_least_squares (:119-152), fit_vector_mapping
    (:155-194), emitting ELL values instead of two COO SparseTensors.

    A slot whose index equals N is the reference's padding sentinel (:226-227): it behaves
    like neighbour 0 with weight forced to zero after normalisation, and its entries are 0.
    """
    N, k = idx.shape
    dt = pos.dtype
    valid = idx != N
    j = torch.where(valid, idx, torch.zeros_like(idx))
    if batch is None:
        batch = torch.zeros(N, dtype=torch.long)

    # A.1 tangent-plane coordinates
    d = pos[j] - pos[:, None, :]                                   # [N,k,3]
    height = (d * normal[:, None, :]).sum(-1)                      # n_i . d
    d_tan = d - normal[:, None, :] * height[..., None]
    u = (d_tan * xb[:, None, :]).sum(-1)
    v = (d_tan * yb[:, None, :]).sum(-1)

    # A.2 gaussian weights, kernel size relative to the per-cloud mean edge length
    r = d.norm(dim=-1)
    per_point = r.mean(dim=1)
    B = int(batch.max()) + 1 if N else 0
    cloud_sum = torch.zeros(B, dtype=dt).index_add(0, batch, per_point)
    cloud_cnt = torch.zeros(B, dtype=dt).index_add(0, batch, torch.ones_like(per_point)).clamp(min=1)
    scale = (kernel_width * (cloud_sum / cloud_cnt))[batch][:, None]
    w = torch.exp(-(r * r) / (scale * scale))
    w = w / w.sum(dim=1, keepdim=True).clamp(EPS)
    w = torch.where(valid, w, torch.zeros_like(w))

    # A.3 quadric basis and regularised normal equations
    basis = torch.stack([torch.ones_like(u), u, v, u * u, u * v, v * v], dim=-1)   # [N,k,6]
    wb = basis * w[..., None]
    normal_mat = torch.einsum('nki,nkj->nij', wb, basis)
    eye = torch.eye(6)          # fp32 on purpose: the reference builds `regularizer * torch.eye(6, 6)`
                                # in the default dtype (:140), so lambda is rounded to fp32 even in an fp64 run

    def solve(reg):
        Z = torch.linalg.inv(normal_mat + (reg * eye).to(dt))
        return torch.einsum('nij,nkj->nki', Z, wb)                 # [N,k,6]

    wls = solve(lam)
    wls_surface = wls if lam_shape is None else solve(lam_shape)

    # A.4-5 gradient entries, per-cloud infinity-norm normalisation
    g = wls[..., 1:3] * valid[..., None]
    if normalized:
        row_abs = g.abs().sum(dim=1)                               # [N,2]
        s = row_abs.norm(dim=1)
        S = torch.zeros(B, dtype=dt).scatter_reduce(0, batch, s, reduce='amax', include_self=True)
        Sp = S[batch][:, None, None]
        g = torch.where(Sp > 1e-5, g / Sp, g)

    # A.6 surface fit and frame mapping
    coeff = (wls_surface * height[..., None]).sum(dim=1)           # [N,6]
    c = coeff[:, None, :]
    hx = c[..., 1] + 2 * c[..., 3] * u + c[..., 4] * v
    hy = c[..., 2] + c[..., 4] * u + 2 * c[..., 5] * v
    gam_x = xb[:, None, :] + normal[:, None, :] * hx[..., None]
    gam_y = yb[:, None, :] + normal[:, None, :] * hy[..., None]
    det = 1 + hx * hx + hy * hy
    Eb, Fb, Gb = 1 + hx * hx, hx * hy, 1 + hy * hy
    xj, yj = xb[j], yb[j]
    t00, t01 = (gam_x * xj).sum(-1), (gam_x * yj).sum(-1)
    t10, t11 = (gam_y * xj).sum(-1), (gam_y * yj).sum(-1)
    m00 = (Gb * t00 - Fb * t10) / det
    m01 = (Gb * t01 - Fb * t11) / det
    m10 = (-Fb * t00 + Eb * t10) / det
    m11 = (-Fb * t01 + Eb * t11) / det

    # A.7 divergence entries
    D = torch.stack([g[..., 0] * m00 + g[..., 1] * m10,
                     g[..., 0] * m01 + g[..., 1] * m11], dim=-1) * valid[..., None]
    if return_parts:
        parts = dict(coords=torch.stack([u, v], -1), weights=w, wls=wls,
                     mapping=torch.stack([m00, m01, m10, m11], -1).reshape(N, k, 2, 2))
        return g, D, parts
    return g, D


# --------------------------------------------------------------------------------------
# operator application and algebra (operators.py:4-46, SparseTensor @ dense)
# --------------------------------------------------------------------------------------

def apply_grad(idx, G, x):
    """(grad @ x): [N,C] -> [2N,C]."""
    N, k = idx.shape
    j = torch.where(idx != N, idx, torch.zeros_like(idx))
    out = torch.einsum('nka,nkc->nac', G, x[j])
    return out.reshape(2 * N, -1)


def apply_div(idx, D, v):
    """(div @ v): [2N,C] -> [N,C]."""
    N, k = idx.shape
    j = torch.where(idx != N, idx, torch.zeros_like(idx))
    vv = v.reshape(N, 2, -1)[j]                                    # [N,k,2,C]
    return torch.einsum('nka,nkac->nc', D, vv)


def vec_norm(v):
    return v.reshape(-1, 2, v.size(1)).norm(dim=1)


def rot90(v):
    vv = v.reshape(-1, 2, v.size(1))
    return torch.stack([-vv[:, 1], vv[:, 0]], dim=1).reshape(v.shape)


def with_rot90(v):
    return torch.cat([v, rot90(v)], dim=1)


def curl(idx, D, v):
    return -apply_div(idx, D, rot90(v))


def laplacian(idx, G, D, x):
    return -apply_div(idx, D, apply_grad(idx, G, x))


def hodge_laplacian(idx, G, D, v):
    return -(apply_grad(idx, G, apply_div(idx, D, v)) + rot90(apply_grad(idx, G, curl(idx, D, v))))


def ell_to_coo(idx, G, D):
    """COO triplets of both operators sorted by (row, col), for entry-wise comparison with the
    reference's SparseTensors (grad_div_mls.py:253-255,263,273-275)."""
    N, k = idx.shape
    i = torch.arange(N)[:, None].expand(N, k)
    g_row = torch.stack([2 * i, 2 * i + 1], -1).reshape(-1)
    g_col = torch.stack([idx, idx], -1).reshape(-1)
    g_val = G.reshape(-1)
    d_row = torch.stack([i, i], -1).reshape(-1)
    d_col = torch.stack([2 * idx, 2 * idx + 1], -1).reshape(-1)
    d_val = D.reshape(-1)

    def srt(r, c, val, ncol):
        p = torch.argsort(r * ncol + c, stable=True)
        return r[p], c[p], val[p]
    return srt(g_row, g_col, g_val, N), srt(d_row, d_col, d_val, 2 * N)
