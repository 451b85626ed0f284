"""GPU parity of geodesic farthest-point sampling (dc_geodesic_fps): the sample SEQUENCE and the fp64
distance vector must equal the reference's bit for bit when both start from the same point — checked against the
reference's committed outputs (tests/golden/geodesic_fps.npz), against the C restatement of oracle/ on ragged batches,
and against the compiled reference itself where oracle/_ref/libref_fps.so travelled to this box."""
import warnings

import numpy as np
import pytest
import torch

from helpers import load
from oracle import fps as ofps

pytestmark = pytest.mark.gpu
DEV = 'cuda'


def dc():
    import deltaconv_b200
    return deltaconv_b200


def _shape(n, rng, squash=(1.0, 0.7, 0.4)):
    u = rng.standard_normal((n, 3))
    u /= np.linalg.norm(u, axis=1, keepdims=True)
    r = 1 + 0.15 * np.sin(3 * np.arctan2(u[:, 1], u[:, 0])) * u[:, 2]
    return (u * r[:, None] * np.array(squash)).astype(np.float32)


def _batch(clouds):
    pos = torch.from_numpy(np.concatenate(clouds)).to(DEV)
    batch = torch.arange(len(clouds)).repeat_interleave(torch.tensor([len(c) for c in clouds])).to(DEV)
    return pos, batch


def test_golden_sequences_of_the_reference_bit_exact():
    _, z = load('geodesic_fps')
    for c in range(len(z['kinds'])):
        pos, want = z[f'pos_{c}'], z[f'samples_{c}']
        got = dc().ops.geodesic_fps(torch.from_numpy(pos).to(DEV), len(want), start=[int(want[0])])
        assert got.dtype == torch.int64 and tuple(got.shape) == (1, len(want))
        assert np.array_equal(got[0].cpu().numpy(), want), f'case {c}'


@pytest.mark.parametrize('sizes,n_samples', [([11, 50, 300, 2048, 5000], 64), ([700, 40, 1200], 512), ([1024] * 40, 256),
                                             ([8192, 8192, 3000], 1024)])
def test_ragged_batch_matches_the_restatement_samples_and_distances(sizes, n_samples):
    rng = np.random.default_rng(sum(sizes) + n_samples)
    clouds = [_shape(n, rng) for n in sizes]
    start = [int(rng.integers(0, n)) for n in sizes]
    pos, batch = _batch(clouds)
    got, dist = dc().ops.geodesic_fps(pos, n_samples, batch=batch, start=start, return_dist=True)
    got, dist = got.cpu().numpy(), dist.cpu().numpy()
    off = 0
    for b, x in enumerate(clouds):
        want, want_dist = ofps.restated_fps(x, n_samples, start[b], want_dist=True)
        assert np.array_equal(got[b], want), f'cloud {b}'
        assert np.array_equal(dist[off:off + len(x)], want_dist), f'cloud {b}: distances differ'     # bit-exact fp64
        off += len(x)


def test_cloud_too_large_for_shared_memory_takes_the_workspace_path():
    rng = np.random.default_rng(5)
    # above the 18,432-point limit of the shared-memory path, far below it, and just below it (one CTA per SM), one launch
    clouds = [_shape(20000, rng), _shape(500, rng), _shape(17000, rng)]
    pos, batch = _batch(clouds)
    got, dist = dc().ops.geodesic_fps(pos, 200, batch=batch, start=[123, 7, 9999], return_dist=True)
    for b, x in enumerate(clouds):
        want, want_dist = ofps.restated_fps(x, 200, [123, 7, 9999][b], want_dist=True)
        assert np.array_equal(got[b].cpu().numpy(), want)
    # the same large cloud alone with a bound that is too low: it still takes the workspace path and stays correct
    one = torch.zeros(20000, dtype=torch.int64, device=DEV)
    alone = dc().ops.geodesic_fps(pos[:20000].clone(), 200, batch=one, num_clouds=1, start=[123], max_points=1000)
    assert torch.equal(alone[0], got[0])
    assert np.array_equal(dist[:20000].cpu().numpy(), ofps.restated_fps(clouds[0], 200, 123, want_dist=True)[1])


def test_disconnected_graph_and_fp64_input():
    rng = np.random.default_rng(11)
    x = rng.standard_normal((400, 3)) * 0.1
    x[150:] += np.array([9.0, 0.0, 0.0])                            # two components: infinite distances until both are hit
    got = dc().ops.geodesic_fps(torch.from_numpy(x).to(DEV), 50, start=[3])       # genuinely double coordinates
    assert np.array_equal(got[0].cpu().numpy(), ofps.restated_fps(x, 50, 3))


def test_visiting_order_changes_nothing_and_exact_ties_go_to_the_lower_index():
    rng = np.random.default_rng(21)
    # a lattice (many exactly equal distances) with every point present twice (zero distances), shuffled
    g = np.stack(np.meshgrid(np.arange(12), np.arange(12), np.arange(3), indexing='ij'), -1).reshape(-1, 3).astype(np.float32) * 0.25
    x = np.concatenate([g, g])[rng.permutation(2 * len(g))]
    clouds = [x, _shape(3000, rng), x[:500]]
    start = [5, 17, 0]
    pos, batch = _batch(clouds)
    a, da = dc().ops.geodesic_fps(pos, 300, batch=batch, start=start, return_dist=True, ordered=True)
    b, db = dc().ops.geodesic_fps(pos, 300, batch=batch, start=start, return_dist=True, ordered=False)
    assert torch.equal(a, b) and torch.equal(da, db)
    off = 0
    for i, c in enumerate(clouds):
        want, want_dist = ofps.restated_fps(c, 300, start[i], want_dist=True)
        assert np.array_equal(a[i].cpu().numpy(), want), f'cloud {i}'
        assert np.array_equal(da[off:off + len(c)].cpu().numpy(), want_dist), f'cloud {i}'
        off += len(c)


@pytest.mark.parametrize('offset,scale', [(0.0, 1.0), (3.0e4, 1.0), (-7.0e6, 1.0), (0.0, 1e-11), (0.0, 3e19), (1e-3, 1e-7)])
def test_fp32_screening_never_costs_a_neighbour(offset, scale):
    """The neighbour search screens candidates in fp32 with a margin derived from the largest coordinate.  Far from the
    origin that margin is wide (coordinates lose most of their fp32 digits), at extreme scales the screening is off:
    either way graph, sample sequence and fp64 distances must stay those of the all-fp64 restatement."""
    rng = np.random.default_rng(17)
    clouds = [(_shape(n, rng).astype(np.float64) + offset) * scale for n in (1500, 700)]      # genuinely double coordinates
    pos = torch.from_numpy(np.concatenate(clouds)).to(DEV)
    batch = torch.arange(2, device=DEV).repeat_interleave(torch.tensor([1500, 700], device=DEV))
    got, dist = dc().ops.geodesic_fps(pos, 200, batch=batch, start=[11, 5], return_dist=True)
    off = 0
    for b, x in enumerate(clouds):
        want, want_dist = ofps.restated_fps(x, 200, [11, 5][b], want_dist=True)
        assert np.array_equal(got[b].cpu().numpy(), want), f'cloud {b}'
        assert np.array_equal(dist[off:off + len(x)].cpu().numpy(), want_dist), f'cloud {b}'
        off += len(x)


@pytest.mark.skipif(not ofps.have_reference(), reason='compiled reference (oracle/_ref) not on this box')
def test_against_the_compiled_reference_itself():
    rng = np.random.default_rng(99)
    for n, m in ((1000, 1000), (4096, 1024)):
        x = _shape(n, rng)
        ref = ofps.reference_fps(x, m)
        got = dc().ops.geodesic_fps(torch.from_numpy(x).to(DEV), m, start=[int(ref[0])])
        assert np.array_equal(got[0].cpu().numpy(), ref)


def test_reference_shaped_numpy_face():
    g = dc().geometry.geodesic_fps
    rng = np.random.default_rng(2)
    x = _shape(300, rng)
    idx = g(x, 100)
    assert isinstance(idx, np.ndarray) and idx.shape == (100,) and len(set(idx.tolist())) == 100
    assert np.array_equal(g(x, 100, start=int(idx[0])), ofps.restated_fps(x, 100, idx[0]))
    with pytest.raises(ValueError):
        g(torch.from_numpy(x), 10)
    with pytest.raises(ValueError):
        g(x[:, :2], 10)
    with pytest.raises(RuntimeError):
        g(x[:8], 4)
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter('always')
        idx = g(x[:30], 45, start=0)
    assert any('larger than number of points' in str(m.message) for m in w)
    assert (idx[30:] == 0).all() and sorted(idx[:30].tolist()) == list(range(30))


def test_transform_on_one_cloud_and_on_a_batch():
    D = dc()
    rng = np.random.default_rng(4)
    # one cloud, as the reference's pre_transform sees it (CPU tensors in, CPU tensors out)
    x = torch.from_numpy(_shape(500, rng))
    data = D.utils.Data(pos=x.clone(), norm=torch.nn.functional.normalize(x, dim=1), y=torch.tensor([3]))
    out = D.transforms.GeodesicFPS(128)(data)
    assert out.pos.shape == (128, 3) and out.norm.shape == (128, 3) and out.y.tolist() == [3]
    assert torch.equal(out.pos, x[out.sample_idx]) and out.sample_idx.dtype == torch.int64
    assert np.array_equal(out.sample_idx.numpy(), ofps.restated_fps(x.numpy(), 128, int(out.sample_idx[0])))
    # fewer points than samples: tiled like the reference (geodesic_fps.py:24-27)
    small = D.utils.Data(pos=x[:40].clone())
    out = D.transforms.GeodesicFPS(100)(small)
    assert out.pos.shape == (100, 3) and torch.equal(out.sample_idx[:40], out.sample_idx[40:80])
    # a device batch: every cloud sampled in one launch, batch vector rebuilt
    clouds = [D.utils.Data(pos=torch.from_numpy(_shape(n, rng)), y=torch.tensor([i])) for i, n in enumerate((300, 64, 900))]
    big = D.utils.Batch.from_data_list(clouds).to(DEV)
    before = big.pos.clone()
    out = D.transforms.GeodesicFPS(128)(big)
    assert out.pos.shape == (3 * 128, 3) and torch.equal(out.batch.cpu(), torch.arange(3).repeat_interleave(128))
    assert torch.equal(out.pos, before[out.sample_idx]) and out.y.tolist() == [0, 1, 2]
    off = 0
    for i, c in enumerate(clouds):
        local = (out.sample_idx[i * 128:(i + 1) * 128] - off).cpu().numpy()
        n = c.pos.size(0)
        want = ofps.restated_fps(c.pos.numpy(), 128, int(local[0]))
        want = np.tile(want[:n], -(-128 // n))[:128] if n < 128 else want
        assert np.array_equal(local, want)
        off += n


def test_full_size_batch_properties():
    """BASELINE's pre-transform size (8 x 1024 points -> 1024 samples, train_modelnet.py:32-33) for a batch that fills the
    GPU: the oracle checks one cloud of it, the rest goes through size-independent properties — samples distinct and in
    range, identical clouds with identical starts give identical sequences wherever they sit in the batch."""
    rng = np.random.default_rng(8)
    B, P, n = 148, 8192, 1024
    clouds = [_shape(P, rng) for _ in range(4)]
    pos = torch.from_numpy(np.concatenate([clouds[b % 4] for b in range(B)])).to(DEV)
    batch = torch.arange(B, device=DEV).repeat_interleave(P)
    start = [int(s) for s in rng.integers(0, P, B)]
    start[4] = start[0]                                                # same cloud, same start -> same sequence
    got = dc().ops.geodesic_fps(pos, n, batch=batch, start=start, num_clouds=B, max_points=P)
    assert tuple(got.shape) == (B, n) and int(got.min()) >= 0 and int(got.max()) < P
    assert all(len(set(row.tolist())) == n for row in got[:16].cpu())
    assert torch.equal(got[0], got[4])
    assert np.array_equal(got[1].cpu().numpy(), ofps.restated_fps(clouds[1], n, start[1]))
