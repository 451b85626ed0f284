"""CPU oracle for the DeltaConv hot path — TEST INFRASTRUCTURE, not product code.

This package restates, in plain CPU torch and in this repo's fixed-k ELL formulation, the
algorithm of the reference's hot path (rubenwiersma/deltaconv @186fec3; citations below are
relative to /root/reference):

  oracle.geometry   kNN graph, tangent frames, WLS grad/div operator fit, operator algebra
                    (deltaconv/geometry/grad_div_mls.py:10-277, operators.py:4-46,
                     third-party semantics per SURVEY.md App. B)
  oracle.network    DeltaConv, MLP, VectorMLP, VectorNonLin, DeltaNetBase and the two heads
                    (deltaconv/nn/*.py, deltaconv/models/*.py) with the reference's state-dict keys

Pinning.  The reference stores no golden vectors (its tests are property-based),
so the oracle is pinned against *outputs of the reference itself*: `oracle/gen_golden.py`
imports the unmodified reference from /root/reference (through `oracle/ref_standins`, pure-torch
stand-ins for the four third-party packages absent from this image) and writes
`tests/golden/*.npz`; `tests/test_oracle_vs_golden.py` checks every oracle function against
those fixtures.  Parity status: **pinned for everything the reference itself computes**;
kNN tie order and arg-max tie routing are *unpinned by the reference* (third-party behaviour,
no source in tree) and pinned by this repo's own rule ((d², index)-lexicographic; lowest slot).

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference`
legs may import this package.  The product (`deltaconv_b200`) never does.
"""
