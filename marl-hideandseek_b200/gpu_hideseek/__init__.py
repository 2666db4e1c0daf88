"""gpu_hideseek — drop-in Python module for the B200 batch simulator.

Mirrors the reference's nanobind module of the same name (src/bindings.cpp:20-121
and M/src/python/bindings.cpp:320-378): `HideAndSeekSimulator(...)` with the
same keyword arguments, `init()`, `step()`, the `*_tensor()` getters returning
`madrona.Tensor` objects with `.to_torch()`, `SimFlags`, and
`madrona.ExecMode`.  Underneath it binds the C ABI of
`lib/libhideseek_b200.so` (include/hideseek/c_api.h) with ctypes; tensors are
zero-copy views of the library's device buffers (CUDA array interface).

There is no CPU execution mode and no fallback: constructing a simulator
without the CUDA library or without a GPU raises.
"""
import ctypes as C
import enum
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(os.path.dirname(_HERE), "lib", "libhideseek_b200.so")

_lib = None

EXPORT_NAMES = [
    "reset", "prep_counter", "action", "self_data", "self_type", "self_mask",
    "agent_data", "box_data", "ramp_data", "visible_agents_mask", "visible_boxes_mask",
    "visible_ramps_mask", "lidar", "seed", "reward", "done", "global_positions",
    "policy_assignments", "episode_result", "ckpt_ctrl", "ckpt",
]
EXPORT_ID = {n: i for i, n in enumerate(EXPORT_NAMES)}

C_API_SYMBOLS = [
    "hs_create", "hs_destroy", "hs_init", "hs_step", "hs_sync", "hs_export",
    "hs_trigger_reset", "hs_set_action", "hs_step_host", "hs_dump_world",
    "hs_world_dump_bytes", "hs_last_step_launches", "hs_overflow_count", "hs_last_error",
    "hs_step_profiled", "hs_profiled_kernel_name", "hs_world_stats",
]


class hs_config(C.Structure):
    _fields_ = [("gpu_id", C.c_int32), ("num_worlds", C.c_int32), ("world_base", C.c_int32),
                ("sim_flags", C.c_uint32), ("rand_seed", C.c_uint32),
                ("min_hiders", C.c_int32), ("max_hiders", C.c_int32),
                ("min_seekers", C.c_int32), ("max_seekers", C.c_int32),
                ("num_pbt_policies", C.c_int32)]


def load_library():
    """dlopen the CUDA library and declare the C-ABI prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise RuntimeError(
            f"{_LIB_PATH} is missing: build it with `make -C marl-hideandseek_b200` "
            "(there is no CPU fallback)")
    l = C.CDLL(_LIB_PATH)
    l.hs_create.argtypes = [C.POINTER(hs_config), C.POINTER(C.c_void_p)]
    l.hs_destroy.argtypes = [C.c_void_p]
    l.hs_destroy.restype = None
    l.hs_init.argtypes = [C.c_void_p, C.c_void_p]
    l.hs_step.argtypes = [C.c_void_p, C.c_void_p]
    l.hs_sync.argtypes = [C.c_void_p, C.c_void_p]
    l.hs_export.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int),
                            C.POINTER(C.c_int), C.POINTER(C.c_int64)]
    l.hs_trigger_reset.argtypes = [C.c_void_p, C.c_int, C.c_int]
    l.hs_set_action.argtypes = [C.c_void_p] + [C.c_int] * 6
    l.hs_step_host.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    l.hs_dump_world.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    l.hs_last_step_launches.argtypes = [C.c_void_p]
    l.hs_overflow_count.argtypes = [C.c_void_p, C.POINTER(C.c_uint32)]
    l.hs_last_error.restype = C.c_char_p
    l.hs_step_profiled.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_float), C.c_int, C.POINTER(C.c_int)]
    l.hs_profiled_kernel_name.argtypes = [C.c_int]
    l.hs_profiled_kernel_name.restype = C.c_char_p
    _lib = l
    return l


def _check(rc):
    if rc != 0:
        raise RuntimeError("hideseek: " + load_library().hs_last_error().decode())


class SimFlags(enum.IntFlag):
    """src/sim_flags.hpp:7-13"""
    Default = 0
    UseFixedWorld = 1 << 0
    IgnoreEpisodeLength = 1 << 1
    RandomFlipTeams = 1 << 2
    ZeroAgentVelocity = 1 << 3


class _Madrona:
    """Subset of the `madrona` helper submodule the scripts use."""

    class ExecMode(enum.IntEnum):
        CPU = 0
        CUDA = 1

    class Tensor:
        """Non-owning device tensor (M/include/madrona/py/utils.hpp:73-141)."""
        _TYPESTR = {0: "|u1", 1: "<i4", 2: "<f4"}

        def __init__(self, ptr, dtype, shape, gpu_id, owner):
            self._ptr, self._dtype, self._shape, self._gpu = ptr, dtype, tuple(shape), gpu_id
            self._owner = owner     # keeps the simulator (and its buffers) alive

        @property
        def __cuda_array_interface__(self):
            return {"shape": self._shape, "typestr": self._TYPESTR[self._dtype],
                    "data": (self._ptr, False), "version": 3, "strides": None}

        def devicePtr(self):
            return self._ptr

        def gpuID(self):
            return self._gpu

        def to_torch(self):
            import torch
            t = torch.as_tensor(self, device=f"cuda:{self._gpu}")
            t._hs_owner = self._owner
            return t

        def to_jax(self):
            raise RuntimeError("jax is not installed in this environment")


madrona = _Madrona


class HideAndSeekSimulator:
    """Same constructor keywords as the reference binding (src/bindings.cpp:32-73)."""

    def __init__(self, exec_mode=_Madrona.ExecMode.CUDA, gpu_id=0, num_worlds=1, sim_flags=0,
                 rand_seed=0, min_hiders=2, max_hiders=2, min_seekers=2, max_seekers=2,
                 num_pbt_policies=0, enable_batch_renderer=False, batch_render_width=64,
                 batch_render_height=64, world_base=0):
        if int(exec_mode) != int(_Madrona.ExecMode.CUDA):
            raise RuntimeError("gpu_hideseek (B200): only ExecMode.CUDA exists; there is no CPU backend")
        if enable_batch_renderer:
            raise RuntimeError("batch renderer (Vulkan) is unavailable on this box")
        self._lib = load_library()
        self._cfg = hs_config(int(gpu_id), int(num_worlds), int(world_base), int(sim_flags),
                              int(rand_seed), int(min_hiders), int(max_hiders),
                              int(min_seekers), int(max_seekers), int(num_pbt_policies))
        h = C.c_void_p()
        _check(self._lib.hs_create(C.byref(self._cfg), C.byref(h)))
        self._h = h
        self.num_worlds = int(num_worlds)
        self.agents_per_world = int(max_hiders) + int(max_seekers)
        self.gpu_id = int(gpu_id)

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            self._lib.hs_destroy(h)
            self._h = None

    # Manager::init / step are synchronous on the legacy default stream, like the
    # reference's cuGraphLaunch + cudaStreamSynchronize.
    def init(self):
        _check(self._lib.hs_init(self._h, None))
        _check(self._lib.hs_sync(self._h, None))

    def step(self):
        _check(self._lib.hs_step(self._h, None))
        _check(self._lib.hs_sync(self._h, None))

    def step_async(self, stream=0):
        _check(self._lib.hs_step(self._h, C.c_void_p(stream)))

    def sync(self, stream=0):
        _check(self._lib.hs_sync(self._h, C.c_void_p(stream)))

    def step_host(self, actions_ptr, reward_ptr, done_ptr, stream=0):
        _check(self._lib.hs_step_host(self._h, C.c_void_p(stream), C.c_void_p(actions_ptr),
                                      C.c_void_p(reward_ptr), C.c_void_p(done_ptr)))

    def _tensor(self, name):
        ptr, dt, nd = C.c_void_p(), C.c_int(), C.c_int()
        dims = (C.c_int64 * 4)()
        _check(self._lib.hs_export(self._h, EXPORT_ID[name], C.byref(ptr), C.byref(dt),
                                   C.byref(nd), dims))
        return _Madrona.Tensor(ptr.value, dt.value, [dims[i] for i in range(nd.value)],
                               self.gpu_id, self)

    def trigger_reset(self, world, level=1):
        _check(self._lib.hs_trigger_reset(self._h, world, level))

    def set_action(self, agent_row, x, y, r, g=0, l=0):
        _check(self._lib.hs_set_action(self._h, agent_row, x, y, r, int(g), int(l)))

    def dump_world(self, world):
        from .dump import world_dt
        out = np.zeros(1, dtype=world_dt)
        assert self._lib.hs_world_dump_bytes() == world_dt.itemsize
        _check(self._lib.hs_dump_world(self._h, world, out.ctypes.data, world_dt.itemsize))
        return out[0]

    def step_profiled(self, stream=0):
        """One un-graphed step with a CUDA event after every kernel; returns [(name, ms)]."""
        ms = (C.c_float * 128)()
        n = C.c_int()
        _check(self._lib.hs_step_profiled(self._h, C.c_void_p(stream), ms, 128, C.byref(n)))
        return [(self._lib.hs_profiled_kernel_name(i).decode(), ms[i]) for i in range(n.value)]

    def world_stats(self):
        """Mean live bodies, agents, candidates, contacts per world (measurement only)."""
        out = (C.c_int64 * 4)()
        self._lib.hs_world_stats.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
        _check(self._lib.hs_world_stats(self._h, out))
        n = float(self.num_worlds)
        return dict(bodies=out[0] / n, agents=out[1] / n, candidates=out[2] / n, contacts=out[3] / n)

    def last_step_launches(self):
        return self._lib.hs_last_step_launches(self._h)

    def overflow_count(self):
        v = C.c_uint32()
        _check(self._lib.hs_overflow_count(self._h, C.byref(v)))
        return v.value

    def depth_tensor(self):
        raise RuntimeError("depth_tensor: batch renderer unavailable on this box")

    def rgb_tensor(self):
        raise RuntimeError("rgb_tensor: batch renderer unavailable on this box")

    def jax(self, xla_gpu):
        raise RuntimeError("jax interface: jax is not installed in this environment")


def _add_getter(name):
    def getter(self):
        return self._tensor(name)
    getter.__name__ = name + "_tensor"
    setattr(HideAndSeekSimulator, name + "_tensor", getter)


for _n in EXPORT_NAMES:
    _add_getter(_n)
