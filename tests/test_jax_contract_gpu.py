"""The JAX / XLA entry points (Manager::gpuJAXInit / gpuJAXStep / gpuJAXSaveCheckpoints /
gpuJAXLoadCheckpoints, src/mgr.cpp:365-437) driven the way XLA drives them — a flat list of
device buffers on a non-default stream — with PyTorch standing in for XLA (jax is not
installed).  The results must equal the ordinary Manager API on a
second simulator fed the same inputs."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

N = 64
KW = dict(num_worlds=N, rand_seed=5, sim_flags=1 | 4 | 8, min_hiders=3, max_hiders=3, min_seekers=3,
          max_seekers=3)        # scripts/jax_train.py:69-81
OBS = ["prep_counter", "self_data", "self_type", "self_mask", "self_lidar", "agent_data", "box_data",
       "ramp_data", "vis_agents_mask", "vis_boxes_mask", "vis_ramps_mask"]


def _like(t):
    return torch.empty_like(t.to_torch())


def _xla_call(sim, which, stream, bufs):
    """Invoke the custom-call entry point exactly as XLA would: (stream, buffers, opaque, len)."""
    cc = sim.xla_custom_calls()
    fn = C.CFUNCTYPE(None, C.c_void_p, C.POINTER(C.c_void_p), C.c_char_p, C.c_size_t)(cc["fns"][which])
    ptrs = [0, 0] + [b.data_ptr() for b in bufs]          # CPU-backend slot, ordering token
    arr = (C.c_void_p * len(ptrs))(*ptrs)
    fn(C.c_void_p(stream.cuda_stream), arr, cc["opaque"], len(cc["opaque"]))


def test_train_interface_names():
    import gpu_hideseek as gh
    sim = gh.HideAndSeekSimulator(**KW)
    ti = sim.train_interface()
    assert list(ti["step_inputs"]["actions"]) == ["actions"]
    assert list(ti["step_inputs"]["pbt"]) == ["policy_assignments"]
    assert list(ti["step_outputs"]["observations"]) == OBS
    assert list(ti["step_outputs"]["pbt"]) == ["episode_results"]
    assert ti["step_outputs"]["rewards"].shape == (N * 6, 1)
    assert ti["checkpointing"]["checkpoint_data"].shape[0] == N
    # the lowering's operand / result table: from the live TrainInterface == from the configuration
    from gpu_hideseek import jax_register
    assert sim.jax_interface_table() == jax_register.expected_interface_table(N, 6, sim.ckpt_tensor().shape[1])
    import importlib.util
    if importlib.util.find_spec("jax") is None:
        with pytest.raises(RuntimeError, match="jax is not installed"):
            sim.jax(True)       # the registration needs jax itself; everything it registers is tested below
    with pytest.raises(RuntimeError, match="no CPU backend"):
        sim.jax(False)


def test_xla_buffers_equal_manager_api():
    import gpu_hideseek as gh
    ref = gh.HideAndSeekSimulator(**KW)
    ref.init()
    sim = gh.HideAndSeekSimulator(**KW)
    ti = sim.train_interface()
    rti = ref.train_interface()
    stream = torch.cuda.Stream()
    obs = [_like(t) for t in ti["step_outputs"]["observations"].values()]
    robs = [t.to_torch() for t in rti["step_outputs"]["observations"].values()]

    _xla_call(sim, "init", stream, obs)                   # waits for the stream itself
    for name, a, b in zip(OBS, obs, robs):
        assert torch.equal(a, b), ("init", name)

    act = _like(ti["step_inputs"]["actions"]["actions"])
    rst = _like(ti["step_inputs"]["resets"])
    pol = torch.zeros_like(ti["step_inputs"]["pbt"]["policy_assignments"].to_torch())
    rew, done = _like(ti["step_outputs"]["rewards"]), _like(ti["step_outputs"]["dones"])
    epr = _like(ti["step_outputs"]["pbt"]["episode_results"])
    ract, rrst = rti["step_inputs"]["actions"]["actions"].to_torch(), rti["step_inputs"]["resets"].to_torch()
    gen = torch.Generator(device="cpu").manual_seed(3)
    for step in range(12):
        a = torch.zeros(act.shape, dtype=torch.int32)
        a[:, 0:3] = torch.randint(0, 5, (act.shape[0], 3), generator=gen, dtype=torch.int32)
        a[:, 3] = (torch.rand(act.shape[0], generator=gen) < 0.3).int()
        r = (torch.rand(N, 1, generator=gen) < 0.1).int()
        ract.copy_(a.cuda()); rrst.copy_(r.cuda())
        ref.step()
        with torch.cuda.stream(stream):
            act.copy_(a.cuda(), non_blocking=True); rst.copy_(r.cuda(), non_blocking=True)
        _xla_call(sim, "step", stream, [act, rst, pol] + obs + [rew, done, epr])
        stream.synchronize()                              # the step call itself does not wait
        for name, x, y in zip(OBS, obs, robs):
            assert torch.equal(x, y), (step, name)
        assert torch.equal(rew, rti["step_outputs"]["rewards"].to_torch()), step
        assert torch.equal(done, rti["step_outputs"]["dones"].to_torch()), step
        assert torch.equal(epr, rti["step_outputs"]["pbt"]["episode_results"].to_torch()), step

    # checkpoints: save every world through the XLA entry, compare with Manager::saveCheckpoint
    ctrl = torch.ones_like(sim.ckpt_ctrl_tensor().to_torch())
    blob = _like(ti["checkpointing"]["checkpoint_data"])
    _xla_call(sim, "save_ckpts", stream, [ctrl, blob])
    stream.synchronize()
    for w in range(N):
        ref.save_checkpoint(w)
    assert torch.equal(blob, ref.ckpt_tensor().to_torch())

    # ... step on, then restore: the observations returned by the restore call are those of the
    # restored state, identical on both simulators
    _xla_call(sim, "step", stream, [act, rst, pol] + obs + [rew, done, epr])
    ref.step()
    _xla_call(sim, "restore_ckpts", stream, [ctrl, blob] + obs)
    stream.synchronize()
    ref.ckpt_ctrl_tensor().to_torch().fill_(1)
    ref.load_checkpoints()
    for name, x, y in zip(OBS, obs, robs):
        assert torch.equal(x, y), ("restore", name)
    assert sim.overflow_count() == 0
