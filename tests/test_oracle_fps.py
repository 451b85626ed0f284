"""CPU: the restated geodesic farthest-point sampler (oracle/geodesic_fps.c) against
  * the committed outputs of the reference's own sampler (tests/golden/geodesic_fps.npz, oracle/gen_golden_fps.py),
  * the compiled reference itself where oracle/_ref/libref_fps.so exists (fresh seeded clouds, every call),
and the claim the CUDA kernel rests on: label-correcting relaxation rounds in any order end in exactly the distance
vector Dijkstra returns, bit for bit (fps.cu header)."""
import numpy as np
import pytest

from helpers import load
from oracle import fps

pytestmark = pytest.mark.skipif(not fps.have_restated(), reason='oracle/liboracle_fps.so not built (make -C oracle)')


def _golden_cases():
    g, z = load('geodesic_fps')
    n = len(z['kinds'])
    return [(z[f'pos_{c}'], z[f'samples_{c}'], z[f'knn_{c}']) for c in range(n)]


def test_restatement_reproduces_the_reference_sequences_of_the_golden_file():
    for pos, samples, knn in _golden_cases():
        mine = fps.restated_fps(pos, len(samples), samples[0])
        assert np.array_equal(mine, samples)                       # bit-exact index sequence
        if knn.shape[0]:
            nbr, _ = fps.restated_knn(pos)
            assert np.array_equal(np.sort(nbr, axis=1), np.sort(knn, axis=1))    # same neighbour sets as the kd-tree


def test_more_samples_than_points_repeats_index_zero_like_the_reference():
    pos, samples, _ = _golden_cases()[-1]                          # 40 points, 64 samples
    assert len(set(samples[:40].tolist())) == 40 and (samples[40:] == 0).all()
    assert np.array_equal(fps.restated_fps(pos, 64, samples[0]), samples)


@pytest.mark.skipif(not fps.have_reference(), reason='compiled reference (oracle/_ref) not present')
def test_restatement_against_the_compiled_reference_on_fresh_clouds():
    rng = np.random.default_rng(7)
    for n, m in ((64, 64), (500, 333), (1500, 1024), (3000, 100)):
        x = (rng.standard_normal((n, 3)) * np.array([1.0, 0.6, 0.3])).astype(np.float32)
        ref = fps.reference_fps(x, m)
        assert np.array_equal(fps.restated_fps(x, m, ref[0]), ref)
        # one Dijkstra pass of the reference into a vector that already holds distances
        mine_samples, dist = fps.restated_fps(x, 5, ref[0], want_dist=True)
        d = np.full(n, np.inf)
        for s in mine_samples:
            d = fps.reference_dijkstra(x, s, d)
        assert np.array_equal(d, dist)                             # bit-exact doubles


def _relaxation_rounds(nbr, length, source, dist, rng):
    """The CUDA kernel's scheme on the host: a work queue per round, relaxations inside a round in random order."""
    d = dist.copy()
    d[source] = 0.0
    queue = [source]
    while queue:
        items = [(u, j) for u in queue for j in range(nbr.shape[1]) if nbr[u, j] >= 0]
        rng.shuffle(items)
        nxt = set()
        for u, j in items:
            cand = d[u] + length[u, j]
            if cand < d[nbr[u, j]]:
                d[nbr[u, j]] = cand
                nxt.add(int(nbr[u, j]))
        queue = list(nxt)
    return d


def test_relaxation_rounds_in_any_order_end_in_dijkstras_vector_bit_for_bit():
    rng = np.random.default_rng(3)
    x = rng.uniform(-1, 1, (600, 3)).astype(np.float32)
    nbr, length = fps.restated_knn(x)
    samples, final = fps.restated_fps(x, 6, 17, want_dist=True)
    d = np.full(600, np.inf)
    for s in samples:
        d = _relaxation_rounds(nbr, length, int(s), d, rng)
    assert np.array_equal(d, final)


def test_fp32_screening_margin_holds_on_random_and_adversarial_pairs():
    """The neighbour search of fps.cu drops a candidate without an fp64 evaluation when its fp32 squared distance exceeds
    R^2, R = (sqrt(T)(1 + 1e-15) + 2 sqrt(3) u S) / (1 - 3u), u = 2^-24, S = largest |coordinate|, T = current 10th d2.
    The claim behind it: then the fp64 d2 exceeds T.  Checked here in numpy (fp32 arithmetic emulated with float32 ops) on
    pairs placed right at the threshold, where a wrong margin would show."""
    rng = np.random.default_rng(1)
    u = 2.0 ** -24
    n = 200000
    for S_scale, offset in ((1.0, 0.0), (1.0, 50.0), (1.0, 4.0e4), (1e-6, 0.0), (1e9, 3e9)):
        q = (rng.uniform(-1, 1, (n, 3)) + offset) * S_scale
        direction = rng.standard_normal((n, 3))
        direction /= np.linalg.norm(direction, axis=1, keepdims=True)
        radius = rng.uniform(1e-3, 0.3, (n, 1)) * S_scale
        p = q + direction * radius                                   # fp64 points at distance ~radius
        S = max(np.abs(q).max(), np.abs(p).max())
        d = q - p
        d64 = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]
        # thresholds just around the true distance: T in [d64 (1 - 1e-5), d64 (1 + 1e-5)]
        T = d64 * (1 + rng.uniform(-1e-5, 1e-5, n))
        R = (np.sqrt(T) * (1 + 1e-15) + 2 * np.sqrt(3.0) * u * S * (1 + 1e-6)) / (1 - 3 * u)
        thr = np.nextafter((R * R * (1 + 1e-12)).astype(np.float32), np.float32(np.inf))       # rounded up, as __double2float_ru
        q32, p32 = q.astype(np.float32), p.astype(np.float32)
        e = q32 - p32
        d32 = (e[:, 0] * e[:, 0] + e[:, 1] * e[:, 1]) + e[:, 2] * e[:, 2]
        dropped = d32 > thr
        assert not (dropped & (d64 <= T)).any(), (S_scale, offset)
        # and the screening is not vacuous: near the origin a candidate twice as far as the 10th neighbour is always dropped
        if offset == 0.0:
            R_far = (np.sqrt(d64 / 4) * (1 + 1e-15) + 2 * np.sqrt(3.0) * u * S * (1 + 1e-6)) / (1 - 3 * u)
            assert (d32 > (R_far * R_far * (1 + 1e-12)).astype(np.float32)).mean() > 0.999
    # a wrong (too small) margin is caught by this construction: without the 2 sqrt(3) u S term pairs at offset 4e4 fail
    q = (rng.uniform(-1, 1, (n, 3)) + 4.0e4)
    direction = rng.standard_normal((n, 3)); direction /= np.linalg.norm(direction, axis=1, keepdims=True)
    p = q + direction * rng.uniform(1e-3, 0.3, (n, 1))
    d = q - p
    d64 = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]
    T = d64 * (1 + rng.uniform(0, 1e-5, n))
    e = q.astype(np.float32) - p.astype(np.float32)
    d32 = (e[:, 0] * e[:, 0] + e[:, 1] * e[:, 1]) + e[:, 2] * e[:, 2]
    naive = (T * (1 + 1e-12)).astype(np.float32)
    assert ((d32 > naive) & (d64 <= T)).any()
