"""CPU: pin the oracle to the reference's own Maya-hosted solver tests, restated as array
scenes in tests/kat_scenes.py:
test1.py, test3.py, test_min_max_values.py (box constraints), test_marker_weight.py (weight
normalisation by the per-frame maximum), test_marker_enabled.py."""
import pytest

from kat_scenes import KATS
from oraclelib import oracle_mmsolver


@pytest.mark.parametrize("name", list(KATS))
def test_oracle_reproduces_reference_kat(name):
    sb, flags, expected = KATS[name]()
    r, values = oracle_mmsolver(sb, **flags)
    assert r.success == 1                      # `self.assertEqual(result[0], 'success=1')`
    for plug, (want, eps) in expected.items():
        assert abs(values[plug] - want) < eps, (name, plug, values[plug], want)
