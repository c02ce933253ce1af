"""Regenerates the committed known-answer fixtures from the reference's demo data.

Run in the build container (needs /root/reference):  python tests/golden/make_golden.py
 - agaricus_kat.npz / dermatology_kat.npz: binned train matrix (A1 applied by the
   oracle), labels, weights, raw test values, bin centres -- inputs of the two README
   known-answer tests, so they can run where /root/reference is absent (the GPU box).
 - bins_golden.npz: value -> bin vectors for A1 edge cases, from the oracle.
The expected outputs themselves are the reference's printed numbers, kept in tests/kat.py.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import kat  # noqa: E402
import oracle_binding as ob  # noqa: E402


def main():
    np.savez_compressed(os.path.join(HERE, "agaricus_kat.npz"), **kat.make_kat1_inputs())
    np.savez_compressed(os.path.join(HERE, "dermatology_kat.npz"), **kat.make_kat2_inputs())
    rng = np.random.default_rng(20261106)
    cuts = np.sort(rng.normal(size=37).astype(np.float32))
    cuts[5] = cuts[6]  # duplicate centre
    mids = ((cuts[1:] + cuts[:-1]).astype(np.float32).astype(np.float64) * 0.5).astype(np.float32)
    vals = np.concatenate([cuts, mids, np.nextafter(mids, np.float32(9)), np.nextafter(mids, np.float32(-9)),
                           rng.normal(size=200).astype(np.float32) * 2,
                           np.array([-np.inf, np.inf, np.nan, 0.0, -0.0, 1e-38, -1e38], np.float32)])
    start = np.array([0, len(cuts)], np.int32)
    bins = ob.find_bins(vals.reshape(-1, 1), cuts, start).reshape(-1)
    np.savez_compressed(os.path.join(HERE, "bins_golden.npz"), cuts=cuts, vals=vals, bins=bins)
    # the first rows of the reference's demo data, verbatim: input of the row-format parser test
    import ytk_data as yd
    for name, rows in (("agaricus.train.ytklearn", 25), ("dermatology.train.ytklearn", 25)):
        with open(os.path.join(yd.REF, "demo/data/ytklearn", name)) as src, open(os.path.join(HERE, name + ".head"), "w") as dst:
            for _ in range(rows):
                dst.write(src.readline())


if __name__ == "__main__":
    main()
