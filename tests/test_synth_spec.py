"""The numpy statement of the injected drift equals the plain one (CPU)."""
import numpy as np

from nexus_configuration_controller_b200._ffi import SynthCfg
from tests import synth_spec


def test_numpy_spec_equals_python_spec():
    c = SynthCfg()
    c.groups, c.n_secrets, c.n_configmaps = 700, 2, 1
    c.tmpl_len_lo = c.tmpl_len_hi = 512
    c.dep_len_lo, c.dep_len_hi = 1, 100
    c.seed, c.drift_seed = 5, 6
    c.p_flip_ppm, c.p_missing_ppm, c.p_len_ppm, c.p_owner_ppm, c.p_rogue_ppm = 50000, 30000, 20000, 40000, 25000
    a = synth_spec.expected_codes(c, 5)
    b = synth_spec.expected_codes_np(c, 5)
    np.testing.assert_array_equal(a, b)
    assert set(np.unique(a).tolist()) == {0, 1, 2, 4, 6, 8}
