"""The package's input generators (multidisttorch_b200/synth.py) must be bit-identical to the ones the golden
fixtures were generated with (oracle/vae_numpy.py), so that bench.py's golden replay feeds the same inputs."""
import numpy as np

from multidisttorch_b200 import synth
from oracle import vae_numpy as O
from tests import golden_cases as G


def test_generators_match_the_oracle_bit_for_bit():
    assert synth.SEGMENTS == O.SEGMENTS and synth.PARAM_SEED == G.PARAM_SEED and synth.DATA_SEED == G.DATA_SEED
    for H in (400, 4096):
        a, b = synth.init_params(77, H=H), O.init_params(77, O.Dims(H=H))
        for k in O.SEGMENTS:
            assert np.array_equal(a[k], b[k]), (H, k)
    for seed, B in ((5, 1), (6, 48), (7, 128)):
        xa, ea = synth.make_batch(seed, B)
        xb, eb = O.make_batch(seed, B)
        assert np.array_equal(xa, xb) and np.array_equal(ea, eb)
    assert synth.batch_seed(3, 2) == G.batch_seed(3, 2)
