"""Small end-to-end run for compute-sanitizer: init, steps with grab/lock actions and random
resets, debug levels, checkpoints, the JAX buffer entry points."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "marl-hideandseek_b200"))
import gpu_hideseek as gh

N = 48
sim = gh.HideAndSeekSimulator(num_worlds=N, rand_seed=3, min_hiders=1, max_hiders=3, min_seekers=1, max_seekers=3)
sim.init()
act = sim.action_tensor().to_torch()
rst = sim.reset_tensor().to_torch()
gen = torch.Generator(device="cpu").manual_seed(1)
for step in range(int(os.environ.get("STEPS", "24"))):
    a = torch.zeros(act.shape, dtype=torch.int32)
    a[:, 0:3] = torch.randint(0, 11, (act.shape[0], 3), generator=gen, dtype=torch.int32)
    a[:, 3] = (torch.rand(act.shape[0], generator=gen) < 0.3).int()
    a[:, 4] = (torch.rand(act.shape[0], generator=gen) < 0.2).int()
    act.copy_(a.cuda())
    rst.copy_((torch.rand(N, 1, generator=gen) < 0.1).int().cuda())
    if step == 3:
        for w, lvl in enumerate(range(2, 9)):
            sim.trigger_reset(w, lvl)
    sim.step()
for w in range(N):
    sim.save_checkpoint(w)
sim.step()
sim.ckpt_ctrl_tensor().to_torch().fill_(1)
sim.load_checkpoints()
sim.step()
torch.cuda.synchronize()
print("sanitize run ok, overflow", sim.overflow_count())
