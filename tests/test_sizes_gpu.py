"""World counts that do not line up with anything — one world, fewer worlds than a warp, one more
than a warp / a CTA / a step-graph slice boundary — against the reference CPU backend over an
episode and its reset, ragged teams (0-3 hiders, 1-3 seekers) with grab / lock actions: CTA tails,
the one-slice and two-slice step graphs and the solver's world-order buffers at their edges."""
import pytest

import refsim

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(not refsim.available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("worlds,seed", [(1, 3), (3, 4), (31, 6), (33, 7), (513, 8), (1025, 9), (2049, 10)])
def test_odd_world_counts(worlds, seed):
    from full_parity import run, grablock_tape
    s = run(worlds, 250, seed, [0, 1, 100, 239, 240, 249], hs=(0, 3), sk=(1, 3), tape=grablock_tape)
    assert not s["errors"], s["errors"][:10]
    assert s["overflow"] == 0
