import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")

STOF_CKPTS = ["different-armadillo", "clean-serenity", "graceful-snow", "breezy-violet"]
ESPCN_CKPTS = ["vital-puddle", "wobbly-sponge"]
EDSR_CKPTS = ["proud-cherry", "snowy-dragon"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_npz(name):
    with np.load(os.path.join(GOLDEN, name), allow_pickle=False) as z:
        return {k: z[k] for k in z.files}


@pytest.fixture(scope="session")
def golden():
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = load_npz(name)
        return cache[name]
    return get


def th_key(th):
    return "none" if th is None else ("%g" % th).replace(".", "p").replace("-", "m")


def th_list(g):
    th = float(g["th"])
    ths = [None, 0.5]
    if not np.isnan(th) and th != 0.5:
        ths.append(th)
    return ths
