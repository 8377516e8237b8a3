"""Known-answer container for tests/test_container.py::test_container_known_answer.

The container format has no counterpart in the reference (it never serialises its streams), so this vector is made by
the CPU restatement oracle/container_oracle.py itself: it pins the byte layout against drift, nothing more.
    python tests/golden/make_container_kat.py
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import container_oracle as cor  # noqa: E402


def main():
    rng = np.random.default_rng(20261104)
    J, nvox, D = 6, 700, 3
    keys = np.sort(rng.choice(1 << (3 * J), size=nvox, replace=False)).astype(np.int64)
    q = np.zeros((D, nvox), dtype=np.int32)
    q[0] = rng.integers(-25, 26, nvox)
    q[1] = rng.integers(-2, 3, nvox) * (rng.random(nvox) < 0.15)
    q[2, [3, 300, 650]] = [70000, -1, 9]
    vmin = np.array([-1.25, 0.5, 3.0], dtype=np.float32)
    args = dict(J=J, n_in=1234, step=0.125, vmin=vmin, voxel_size=0.03125, width=2.0)
    data = cor.pack_scene(J, 1234, 0.125, tuple(vmin), 0.03125, 2.0, keys, q, True)
    hb = int.from_bytes(data[8:12], "little")
    np.savez_compressed(os.path.join(HERE, "container_kat.npz"), keys=keys, q=q, sha256=hashlib.sha256(data).hexdigest(),
                        head=np.frombuffer(data[:hb], dtype=np.uint8), **{k: np.asarray(v) for k, v in args.items()})
    print(len(data), hashlib.sha256(data).hexdigest())


if __name__ == "__main__":
    main()
