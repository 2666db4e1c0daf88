"""BASELINE.json configs[1] itself, diffed against the reference CPU backend at full size: 16000
worlds, seed 0, 2 + 2 agents, randint(-5, 5) x/y actions, 245 steps (through the step-239 reset
wave), every world and every export compared at steps {0, 1, 96, 200, 239, 240, 245} — plus the
candidate pairs (order-exact), the contact rows of the last substep and the body every lidar ray
hit, which no export shows.  A second run covers seed 5 x 1024 worlds with
ragged 0-3 teams and grab / lock actions.  Tolerances: tests/full_parity.py."""
import pytest

import refsim

pytestmark = pytest.mark.gpu


def _report(s):
    for c in s["checked_steps"]:
        st = c["stats"]
        print(f"  step {c['step']}: max_pose {st.get('max_pose', 0):.3g} max_vel {st.get('max_vel', 0):.3g} "
              f"max_export {st.get('max_export_diff', 0):.3g} mean_cand {st.get('mean_cand', 0):.2f} "
              f"max_cand {st.get('max_cand', 0)} max_contacts {st.get('max_contacts', 0)} "
              f"lidar hits {st.get('lidar_hits', 0)}/{st.get('lidar_rays', 0)}")
    print(f"  reference {s['seconds_reference']:.1f} s, CUDA {s['seconds_cuda']:.1f} s, max {s['max']}")


@pytest.mark.skipif(not refsim.available(), reason="oracle/_ref not built")
def test_headline_config_full_size():
    from full_parity import run
    s = run(16000, 245, 0, [0, 1, 96, 200, 239, 240, 245])
    _report(s)
    assert not s["errors"], s["errors"][:10]
    assert s["overflow"] == 0


@pytest.mark.skipif(not refsim.available(), reason="oracle/_ref not built")
def test_ragged_teams_grab_lock_1024():
    from full_parity import run, grablock_tape
    s = run(1024, 245, 5, [0, 1, 50, 120, 239, 240, 245], hs=(0, 3), sk=(0, 3), tape=grablock_tape)
    _report(s)
    assert not s["errors"], s["errors"][:10]
    assert s["overflow"] == 0
