"""SeriesStore time handling: device time is int32 ms from the store's origin; a Unix-epoch `now`
must be converted (set_wall_time) or refused, never wrapped (the reference's Prometheus uses the
wall clock: mlflow_operator.py:302-303 queries "now")."""
import pytest
import torch

import canary_b200  # noqa: F401
from canary_b200 import operator_api as api, params as P


def _store(epoch_ms):
    z = torch.zeros(1)
    return api.SeriesStore(z, z, z, torch.zeros((1, 2), dtype=torch.int32), P.make_params(), epoch_ms=epoch_ms)


def test_wall_time_is_taken_relative_to_the_origin():
    st = _store(1_700_000_000_000)
    st.set_wall_time(1_700_000_600_000)
    assert st.now_ms == 600_000
    assert st._device_now(60) == 600_000
    st.advance(10)
    assert st._device_now(60) == 610_000


@pytest.mark.parametrize("now", [1_700_000_600_000, -(1 << 31), (1 << 30)])
def test_times_outside_the_device_range_are_refused(now):
    st = _store(1_700_000_000_000)
    st.set_time(now)
    with pytest.raises(ValueError, match="outside"):
        st._device_now(60)
