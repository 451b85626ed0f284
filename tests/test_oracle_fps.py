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
