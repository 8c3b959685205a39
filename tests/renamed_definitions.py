"""A GTScript stencil written for the library-matching test (build container only: needs gt4py).  It is the
4-point Laplacian smoothing the library implements as `hdiff` without limiter, under other names."""
import numpy as np

from gt4py.cartesian.gtscript import PARALLEL, Field, computation, interval  # noqa: F401


def smoothing(p: Field[np.float64], q: Field[np.float64], r: Field[np.float64]):
    with computation(PARALLEL), interval(...):
        curv = 4.0 * p[0, 0, 0] - (p[1, 0, 0] + p[-1, 0, 0] + p[0, 1, 0] + p[0, -1, 0])
        east = curv[1, 0, 0] - curv[0, 0, 0]
        north = curv[0, 1, 0] - curv[0, 0, 0]
        r = p[0, 0, 0] - q[0, 0, 0] * (east[0, 0, 0] - east[-1, 0, 0] + north[0, 0, 0] - north[0, -1, 0])
