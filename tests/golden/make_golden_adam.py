"""Generates tests/golden/ouradam.npz by running the UNMODIFIED reference optimizer (/root/reference/scene/OurAdam.py,
imported by path) on CPU over the seeded scenarios of oracle/ouradam.py. Run in the build container:

    python tests/golden/make_golden_adam.py

The .npz holds only OUTPUTS (param, exp_avg, exp_avg_sq per group after 4 steps); the inputs are regenerated from the seed."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ouradam  # noqa: E402

CASES = [(257, 1e-15, 11), (257, 1e-8, 12)]   # (rows, eps, seed): eps of gaussian_model.py:937 and of train_post.py:816


def main():
    ref = ouradam.load_reference()
    assert ref is not None, "needs /root/reference"
    out = {}
    for rows, eps, seed in CASES:
        res = ouradam.run_scenario(ref.Adam, rows, eps, seed)
        for k, (p, m, v) in res.items():
            tag = "r%d_s%d_%s" % (rows, seed, k)
            out[tag + "_p"], out[tag + "_m"], out[tag + "_v"] = p.numpy(), m.numpy(), v.numpy()
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ouradam.npz"), **out)
    print("wrote ouradam.npz with", len(out), "arrays")


if __name__ == "__main__":
    main()
