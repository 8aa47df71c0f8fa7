import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_vectors.npz")
CASES = [("cornell", 64, 48, False, 10), ("gallery", 80, 60, True, 6), ("atrium:8", 64, 40, False, 5)]


def load():
    return np.load(GOLDEN)


def key(name):
    return name.replace(":", "_")
