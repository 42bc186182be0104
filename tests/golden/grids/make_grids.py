"""Writes the small regular_grid resources the golden media of tests/support.py point at, in the reference's own on-disk form
(math/GPFunctions.hpp:346-384: a grid JSON {type, bounds_min, bounds_max, res, path, interpolate} next to a raw file of doubles,
element (x res + y) res + z, Vec3d elements interleaved).  Committed together with its outputs; run once:  python tests/golden/grids/make_grids.py"""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def smooth(rng, res, comps, lo, hi):
    """A few random low-frequency cosines per component, mapped into [lo, hi]."""
    ax = (np.arange(res) + 0.5) / res
    x, y, z = np.meshgrid(ax, ax, ax, indexing="ij")
    out = np.zeros((res, res, res, comps))
    for c in range(comps):
        v = np.zeros_like(x)
        for _ in range(4):
            k = rng.uniform(0.5, 2.5, 3); ph = rng.uniform(0, 2 * np.pi)
            v += rng.uniform(0.3, 1.0) * np.cos(2 * np.pi * (k[0] * x + k[1] * y + k[2] * z) + ph)
        v = (v - v.min()) / (v.max() - v.min())
        out[..., c] = lo + (hi - lo) * v
    return out


def write(name, values, lo, hi, interp):
    res = values.shape[0]
    raw = name + "-values.bin"
    np.ascontiguousarray(values, dtype="<f8").tofile(os.path.join(HERE, raw))
    json.dump({"type": "regular_grid", "bounds_min": [lo] * 3, "bounds_max": [hi] * 3, "res": res, "path": raw, "interpolate": interp},
              open(os.path.join(HERE, name + ".json"), "w"), indent=1)


def main():
    rng = np.random.default_rng(20261105)
    write("var8_linear", smooth(rng, 8, 1, 0.5, 2.0)[..., 0], -1.6, 1.6, "linear")
    write("var5_quadratic", smooth(rng, 5, 1, 0.6, 1.8)[..., 0], -1.5, 1.5, "quadratic")
    write("ls6_quadratic", smooth(rng, 6, 3, 0.5, 3.0), -1.5, 1.5, "quadratic")   # Vec3d: the largest component is the lateral scale
    write("ls7_linear", smooth(rng, 7, 3, 0.45, 2.4), -1.7, 1.7, "linear")
    write("aniso4_point", smooth(rng, 4, 1, 0.1, 0.9)[..., 0], -1.5, 1.5, "point")
    # a tabulated signed distance (ProceduralMean over a regular_grid "f"): a torus of radii (0.8, 0.3) in the xz plane, sampled at cell centres
    res = 12; ax = -1.6 + 3.2 * (np.arange(res) + 0.5) / res
    x, y, z = np.meshgrid(ax, ax, ax, indexing="ij")
    write("torus12_sdf_linear", np.sqrt((np.sqrt(x * x + z * z) - 0.8) ** 2 + y * y) - 0.3, -1.6, 1.6, "linear")


if __name__ == "__main__":
    main()
