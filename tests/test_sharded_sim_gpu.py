"""gpu_hideseek.ShardedSimulator: one object, several shards, one step().  With
both shards on device 0 (what a 1-GPU box can run; on a multi-GPU box pass different ids) the
concatenated exports must equal one simulator over all the worlds, bit for bit: worlds are
independent and the RNG is keyed by global world index."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

NAMES = ("self_data", "box_data", "lidar", "reward", "done", "seed", "visible_boxes_mask", "global_positions")


def test_sharded_equals_single():
    import gpu_hideseek as gh
    N, A, steps = 1000, 4, 25
    ids = [0, 1 % torch.cuda.device_count(), 0]                  # uneven 3-way split: 334 / 333 / 333
    one = gh.HideAndSeekSimulator(num_worlds=N, rand_seed=9)
    many = gh.ShardedSimulator(ids, N, rand_seed=9)
    assert [n for _, n in many.worlds] == [334, 333, 333] and many.worlds[1][0] == 334
    one.init(); many.init()
    a1 = one.action_tensor().to_torch()
    am = [t.to_torch() for t in many.action_tensor()]
    gen = torch.Generator(device="cpu").manual_seed(2)
    for step in range(steps):
        a = torch.zeros((N * A, 5), dtype=torch.int32)
        a[:, 0:3] = torch.randint(0, 11, (N * A, 3), generator=gen, dtype=torch.int32)
        a[:, 3] = (torch.rand(N * A, generator=gen) < 0.2).int()
        a1.copy_(a.to(a1.device))
        for t, (b, n) in zip(am, many.worlds):
            t.copy_(a[b * A:(b + n) * A].to(t.device))
        if step == 10:
            one.trigger_reset(700, 1); many.trigger_reset(700, 1)       # lands in the third shard
        one.step(); many.step()
    for n in NAMES:
        want = getattr(one, n + "_tensor")().to_torch().cpu().numpy()
        got = np.concatenate([t.to_torch().cpu().numpy() for t in getattr(many, n + "_tensor")()], axis=0)
        assert got.shape == want.shape, n
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), n
    assert many.overflow_count() == 0
