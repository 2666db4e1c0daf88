"""Consecutive device-timed 240-step windows, optionally after the host-buffer sections bench.py
runs before its full-episode window (diagnostic for the bench sequence)."""
import sys, time, torch
sys.path.insert(0, "marl-hideandseek_b200")
import gpu_hideseek as gh
N, A = 16000, 4
sim = gh.HideAndSeekSimulator(num_worlds=N, rand_seed=0)
sim.init()
act = sim.action_tensor().to_torch()
stream = torch.cuda.current_stream().cuda_stream
def one():
    act[:, 0:2] = torch.randint(-5, 5, (N * A, 2), dtype=torch.int32, device="cuda")
    sim.step_async(stream)
def window(tag, n=240, extra=None):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        one()
        if extra: extra(i)
    e1.record(); torch.cuda.synchronize()
    print(f"{tag}: {n} steps {e0.elapsed_time(e1)/n:.4f} ms/step", flush=True)
for _ in range(5): one()
window("plain 0"); window("plain 1")
R = N * A
h_act = torch.zeros((R, 5), dtype=torch.int32).pin_memory(); h_act[:, 2] = 5
h_rew = torch.zeros((R,), dtype=torch.float32).pin_memory(); h_done = torch.zeros((R,), dtype=torch.int32).pin_memory()
tape = torch.randint(-5, 5, (8, R, 2), dtype=torch.int32)
for i in range(128):
    h_act[:, 0:2] = tape[i % 8]
    sim.step_host(h_act.data_ptr(), h_rew.data_ptr(), h_done.data_ptr(), stream); torch.cuda.synchronize()
window("after step_host x128")
names = ("prep_counter", "self_data", "self_type", "self_mask", "lidar", "agent_data", "box_data", "ramp_data",
         "visible_agents_mask", "visible_boxes_mask", "visible_ramps_mask", "reward", "done")
outs = [getattr(sim, n + "_tensor")().to_torch() for n in names]
h_outs = [torch.empty(t.shape, dtype=t.dtype).pin_memory() for t in outs]
for i in range(26):
    act.copy_(h_act, non_blocking=True); sim.step_async(stream)
    for s, d in zip(outs, h_outs): d.copy_(s, non_blocking=True)
    torch.cuda.synchronize()
window("after obs-to-host x26")
reward = sim.reward_tensor().to_torch()
acc = torch.zeros((), device="cuda", dtype=torch.float64)
def ex(i):
    global acc
    if i in (120, 239): acc += reward.abs().sum()
window("with reward checksum", extra=ex)
window("plain again")
