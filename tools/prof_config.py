"""Per-kernel CUDA-event table (hs_step_profiled) for a configuration:
    python tools/prof_config.py train     # jax_train.py flags, 3+3 agents, seed 5
    python tools/prof_config.py bench     # benchmark.py configuration
"""
import collections, sys, os
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "marl-hideandseek_b200"))
import gpu_hideseek as gh
N = 16000
which = sys.argv[1] if len(sys.argv) > 1 else "bench"
kw = dict(rand_seed=5, sim_flags=1 | 4 | 8, min_hiders=3, max_hiders=3, min_seekers=3, max_seekers=3) if which == "train" \
    else dict(rand_seed=0, sim_flags=0, min_hiders=2, max_hiders=2, min_seekers=2, max_seekers=2)
sim = gh.HideAndSeekSimulator(num_worlds=N, **kw)
sim.init()
A = sim.agents_per_world
act = sim.action_tensor().to_torch()
dev = act.device
acc, cnt = collections.defaultdict(float), collections.Counter()
for i in range(140):
    if which == "train":
        act[:, 0:3].copy_(torch.randint(0, 5, (N * A, 3), dtype=torch.int32, device=dev))
        act[:, 3:5].copy_(torch.randint(0, 2, (N * A, 2), dtype=torch.int32, device=dev))
    else:
        act[:, 0:2].copy_(torch.randint(-5, 5, (N * A, 2), dtype=torch.int32, device=dev))
    if i < 120:
        sim.step()
    else:
        for n, ms in sim.step_profiled():
            acc[n] += ms; cnt[n] += 1
print(which, "sum ms/step", sum(acc.values()) / 20, sim.world_stats())
for n in sorted(acc, key=lambda n: -acc[n]):
    print(f"{n:20s} x{cnt[n]/20:4.1f} avg {acc[n]/cnt[n]:.4f} ms  total/step {acc[n]/20:.4f}")
