"""Loading and replaying the committed golden fixtures (tests/golden/*.npz, produced from
the reference CPU backend by tests/golden/make_golden.py)."""
import glob
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
EXPORT_KEYS = ["prep_counter", "action", "self_obs", "self_type", "self_mask", "agent_obs", "box_obs",
               "ramp_obs", "agent_vis", "box_vis", "ramp_vis", "lidar", "seed", "reward", "done",
               "global_pos", "episode_result"]


def case_names():
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))


class Golden:
    def __init__(self, name):
        d = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
        self.d = d
        (self.N, self.seed, self.flags, self.min_h, self.max_h, self.min_s, self.max_s,
         self.steps) = (int(x) for x in d["config"])
        self.A = self.max_h + self.max_s
        self.snaps = [int(x) for x in d["snaps"]]
        self.tape = d["tape"].astype(np.int32)
        # optional: level ids passed to triggerReset(world, level) before step 1
        self.levels = [int(x) for x in d["levels"]] if "levels" in d.files else None

    def worlds(self, step):
        return self.d[f"s{step}_worlds"]

    def exports(self, step):
        return {k: self.d[f"s{step}_ex_{k}"] for k in EXPORT_KEYS}

    def kwargs(self):
        return dict(sim_flags=self.flags, min_hiders=self.min_h, max_hiders=self.max_h,
                    min_seekers=self.min_s, max_seekers=self.max_s)
