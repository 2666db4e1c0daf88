import os, sys, collections
import torch
sys.path.insert(0, "marl-hideandseek_b200")
import gpu_hideseek as gh
N=16000
sim = gh.HideAndSeekSimulator(num_worlds=N, rand_seed=0, min_hiders=2, max_hiders=2, min_seekers=2, max_seekers=2)
sim.init()
act = sim.action_tensor().to_torch(); rst = sim.reset_tensor().to_torch()
dev = act.device
acc = collections.defaultdict(float); cnt = collections.Counter()
for i in range(60):
    act[:, 0:2].copy_(torch.randint(-5, 5, (N*4, 2), dtype=torch.int32, device=dev))
    rst.copy_((torch.rand(N, 1, device=dev) < 0.25).int())
    if i < 40:
        sim.step()
    else:
        for n, ms in sim.step_profiled():
            acc[n] += ms; cnt[n] += 1
tot = sum(acc.values())
print("sum ms/step", tot/20)
for n in sorted(acc, key=lambda n: -acc[n]):
    print(f"{n:20s} x{cnt[n]/20:4.1f} avg {acc[n]/cnt[n]:.4f} ms  total/step {acc[n]/20:.4f}")
