"""Stand-in for torch_cluster.knn_graph with the tie-break this repo pins: neighbours by
(squared L2 in fp32 arithmetic of the input dtype, index) ascending, self included when
loop=True, clouds never mix.  Test infrastructure only."""
import torch


def _knn_idx(x, k, batch, loop):
    N = x.size(0)
    if batch is None:
        batch = torch.zeros(N, dtype=torch.long, device=x.device)
    out = torch.empty(N, k, dtype=torch.long, device=x.device)
    start = 0
    counts = torch.bincount(batch).tolist()
    for cnt in counts:
        if cnt == 0:
            continue
        p = x[start:start + cnt]
        diff = p[None, :, :] - p[:, None, :]
        sq = diff * diff
        d2 = sq[..., 0]
        for a in range(1, x.size(1)):
            d2 = d2 + sq[..., a]
        if not loop:
            d2 = d2 + torch.diag(torch.full((cnt,), float('inf'), dtype=d2.dtype))
        order = torch.argsort(d2, dim=1, stable=True)[:, :k]
        out[start:start + cnt] = order + start
        start += cnt
    return out


def knn_graph(x, k, batch=None, loop=False, flow='source_to_target', **_):
    idx = _knn_idx(x, k, batch, loop)
    N = x.size(0)
    query = torch.arange(N, device=x.device).repeat_interleave(k)
    nbr = idx.reshape(-1)
    if flow == 'target_to_source':
        return torch.stack([query, nbr], dim=0)
    return torch.stack([nbr, query], dim=0)
