"""Stub for the reference's pybind11 geodesic-FPS binding (not on the hot path; unbuildable here:
needs cmake + geometry-central + Eigen).  Only lets `import deltaconv` succeed."""


def geodesicFPS(*_args, **_kwargs):
    raise NotImplementedError("deltaconv_bindings.geodesicFPS is outside the hot path (SURVEY.md §2 #13)")
