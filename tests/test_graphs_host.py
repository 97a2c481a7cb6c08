"""Host-side logic of bsarec_b200/graphs.py (no GPU): the seed arithmetic of graph replays is 64-bit wrap-around
addition — the same sum the kernels form on the device (common.cuh dropout_key) and oracle/rng_ref.py accepts."""
import numpy as np
import pytest
import torch

from bsarec_b200 import graphs
from oracle import rng_ref


def test_replay_seed_is_wraparound_addition():
    base = (1 << 62) + 12345
    for r in (0, 1, 2, 7, 1000):
        s = graphs.replay_seed(base, r)
        assert 0 <= s < 1 << 64
        with np.errstate(over="ignore"):
            want = np.uint64(base) + np.uint64(graphs.SEED_STRIDE) * np.uint64(r)
        assert s == int(want)
    # the int64 tensor the graph adds on the device carries the same bits
    assert graphs._signed64(graphs.SEED_STRIDE) & ((1 << 64) - 1) == graphs.SEED_STRIDE
    assert graphs.SEED_STRIDE % 2 == 1 and graphs.SEED_STRIDE != int(rng_ref.GOLDEN)


def test_replay_streams_do_not_collide_with_site_keys():
    """(replay, site) -> key must be injective over the ranges in use: 64 replays x 41 sites of one captured step."""
    keys = {int(rng_ref.site_key(graphs.replay_seed(99, r), site)) for r in range(64) for site in range(41)}
    assert len(keys) == 64 * 41


def test_graphed_step_refuses_other_optimizers():
    lin = torch.nn.Linear(2, 2)
    with pytest.raises(TypeError):
        graphs.GraphedTrainStep(lin, torch.optim.Adam(lin.parameters()))
