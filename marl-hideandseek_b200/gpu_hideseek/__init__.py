"""gpu_hideseek — drop-in Python module for the B200 hide-and-seek batch simulator.

Mirrors the reference's nanobind module of the same name (src/bindings.cpp:20-121 and the
`madrona` submodule of M/src/python/bindings.cpp:320-378): `HideAndSeekSimulator(...)` with
the same keyword arguments, `init()`, `step()`, the `*_tensor()` getters returning
`madrona.Tensor` objects with `.to_torch()` (zero-copy DLPack), `SimFlags`, and
`madrona.ExecMode`.

Layers (DESIGN.md "Boundary"):
    gpu_hideseek._native            pybind11 module (host/bindings.cpp)
      -> GPUHideSeek::Manager       C++ host class, API of src/mgr.hpp (host/mgr.cpp)
        -> C ABI                    include/hideseek/c_api.h (lib/libhideseek_b200.so)
          -> sm_100a CUDA kernels   csrc/*.cuh

There is no CPU execution mode and no fallback: importing this package without the built
libraries, or constructing a simulator without a GPU, raises.
"""
import ctypes as C
import enum
import os

import numpy as np

# the Manager aborts the process on failure like the reference (FATAL); tests and
# interactive users get exceptions instead unless they opt back in with HIDESEEK_THROW=0
os.environ.setdefault("HIDESEEK_THROW", "1")

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_DIR = os.path.join(os.path.dirname(_HERE), "lib")
_LIB_PATH = os.path.join(_LIB_DIR, "libhideseek_b200.so")

try:
    from . import _native
except ImportError as e:      # pragma: no cover - build problem, fail loudly
    raise ImportError(
        f"gpu_hideseek: the native module is not built ({e}); run "
        "`make -C marl-hideandseek_b200` (there is no Python/CPU fallback)") from e

madrona = _native.madrona


class SimFlags(enum.IntFlag):
    """src/sim_flags.hpp:7-13; arithmetic like the reference's nb::is_arithmetic() enum."""
    Default = 0
    UseFixedWorld = 1 << 0
    IgnoreEpisodeLength = 1 << 1
    RandomFlipTeams = 1 << 2
    ZeroAgentVelocity = 1 << 3


EXPORT_NAMES = [
    "reset", "prep_counter", "action", "self_data", "self_type", "self_mask",
    "agent_data", "box_data", "ramp_data", "visible_agents_mask", "visible_boxes_mask",
    "visible_ramps_mask", "lidar", "seed", "reward", "done", "global_positions",
    "policy_assignments", "episode_result", "ckpt_ctrl", "ckpt",
]

_lib = None


def c_api_symbols():
    """Function names declared in include/hideseek/c_api.h (parsed from the header)."""
    import re
    hdr = os.path.join(os.path.dirname(os.path.dirname(_HERE)), "include", "hideseek", "c_api.h")
    src = re.sub(r"/\*.*?\*/", "", open(hdr).read(), flags=re.S)
    return sorted(set(re.findall(r"\b(hs_[a-z0-9_]+)\s*\(", src)))


def load_library():
    """ctypes handle on the C-ABI library (diagnostic / measurement entry points)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise RuntimeError(f"{_LIB_PATH} is missing: build it with `make -C marl-hideandseek_b200` "
                           "(there is no CPU fallback)")
    l = C.CDLL(_LIB_PATH)
    l.hs_dump_world.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    l.hs_last_step_launches.argtypes = [C.c_void_p]
    l.hs_overflow_count.argtypes = [C.c_void_p, C.POINTER(C.c_uint32)]
    l.hs_last_error.restype = C.c_char_p
    l.hs_step_profiled.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_float), C.c_int, C.POINTER(C.c_int)]
    l.hs_profiled_kernel_name.argtypes = [C.c_int]
    l.hs_profiled_kernel_name.restype = C.c_char_p
    l.hs_world_stats.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
    l.hs_build_view_mesh_table.argtypes = [C.c_char_p, C.c_void_p, C.c_int]
    l.hs_build_view_tex_table.argtypes = [C.c_char_p, C.c_void_p, C.c_int]
    _lib = l
    return l


def _check(rc):
    if rc != 0:
        raise RuntimeError("hideseek: " + load_library().hs_last_error().decode())


def view_mesh_table(data_dir=None):
    """The render-mesh table of the agent views (csrc/hs_views.cuh) as a numpy record: built-in
    meshes, or the OBJ files of `data_dir` (the reference's data/ directory).  Host only."""
    from .dump import view_mesh_dt
    lib = load_library()
    out = np.zeros(1, dtype=view_mesh_dt)
    assert lib.hs_view_mesh_table_bytes() == view_mesh_dt.itemsize
    _check(lib.hs_build_view_mesh_table((data_dir or "").encode(), out.ctypes.data, view_mesh_dt.itemsize))
    return out[0]


def view_tex_table(data_dir=None):
    """Per-triangle texture coordinates and material -> texture bindings of the agent views
    (hs::ViewTexTable) for `data_dir`.  Host only."""
    from .dump import view_tex_dt
    lib = load_library()
    out = np.zeros(1, dtype=view_tex_dt)
    assert lib.hs_view_tex_table_bytes() == view_tex_dt.itemsize
    _check(lib.hs_build_view_tex_table((data_dir or "").encode(), out.ctypes.data, view_tex_dt.itemsize))
    return out[0]


class HideAndSeekSimulator(_native.HideAndSeekSimulator):
    """The reference's constructor keywords (src/bindings.cpp:32-73), all optional here.

    On top of the reference API this subclass adds the diagnostic helpers the parity tests and
    bench.py use (dump_world, step_profiled, world_stats ...); they call the C ABI directly on
    the Manager's `hs_sim*`.
    """

    def __init__(self, exec_mode=madrona.ExecMode.CUDA, gpu_id=0, num_worlds=1, sim_flags=0,
                 rand_seed=0, min_hiders=2, max_hiders=2, min_seekers=2, max_seekers=2,
                 num_pbt_policies=0, enable_batch_renderer=False, batch_render_width=64,
                 batch_render_height=64, world_base=0, render_data_dir=None):
        super().__init__(exec_mode, int(gpu_id), int(num_worlds), int(sim_flags), int(rand_seed),
                         int(min_hiders), int(max_hiders), int(min_seekers), int(max_seekers),
                         int(num_pbt_policies), bool(enable_batch_renderer),
                         int(batch_render_width), int(batch_render_height), int(world_base),
                         str(render_data_dir or ""))
        self._clib = load_library()
        self._h = C.c_void_p(self.handle())

    def dump_world(self, world):
        from .dump import world_dt
        out = np.zeros(1, dtype=world_dt)
        assert self._clib.hs_world_dump_bytes() == world_dt.itemsize
        _check(self._clib.hs_dump_world(self._h, world, out.ctypes.data, world_dt.itemsize))
        return out[0]

    def step_profiled(self, stream=0):
        """One un-graphed step with a CUDA event after every kernel; returns [(name, ms)]."""
        ms = (C.c_float * 128)()
        n = C.c_int()
        _check(self._clib.hs_step_profiled(self._h, C.c_void_p(stream), ms, 128, C.byref(n)))
        return [(self._clib.hs_profiled_kernel_name(i).decode(), ms[i]) for i in range(n.value)]

    def world_stats(self):
        """Mean live bodies, agents, candidates, contacts per world (measurement only)."""
        out = (C.c_int64 * 4)()
        _check(self._clib.hs_world_stats(self._h, out))
        n = float(self.num_worlds)
        return dict(bodies=out[0] / n, agents=out[1] / n, candidates=out[2] / n, contacts=out[3] / n)

    def xla_custom_calls(self):
        """What an XLA GPU custom-call registration needs (M/src/python/jax_register.py): the
        addresses of the four entry points `void fn(cudaStream_t, void **buffers, const char
        *opaque, size_t opaque_len)` in libhideseek_b200.so and the `opaque` bytes (this
        simulator's pointer).  Buffers 0 and 1 of a call are the CPU slot and JAX's token."""
        import struct
        fns = {}
        for key, sym in (("init", "hs_xla_gpu_init"), ("step", "hs_xla_gpu_step"),
                         ("save_ckpts", "hs_xla_gpu_save_checkpoints"),
                         ("restore_ckpts", "hs_xla_gpu_load_checkpoints")):
            fns[key] = C.cast(getattr(self._clib, sym), C.c_void_p).value
        return dict(fns=fns, opaque=struct.pack("P", self.handle()))

    def jax(self, xla_gpu=True):
        """src/bindings.cpp:118 — {'init', 'step', 'save_ckpts', 'restore_ckpts'} as jitted JAX
        functions over XLA custom calls into the C ABI (gpu_hideseek/jax_register.py).  Raises
        RuntimeError when jax is not installed."""
        from . import jax_register
        return jax_register.build(self, xla_gpu)

    def jax_interface_table(self):
        """Operand / result (name, shape, dtype) lists of the four XLA custom calls; needs no jax."""
        from . import jax_register
        return jax_register.interface_table(self)

    def last_step_launches(self):
        return self._clib.hs_last_step_launches(self._h)

    def overflow_count(self):
        v = C.c_uint32()
        _check(self._clib.hs_overflow_count(self._h, C.byref(v)))
        return v.value


def shard_worlds(total_worlds, num_shards):
    """World-index partition used for multi-GPU runs: shard g owns the
    contiguous range [base, base + count); worlds never interact, so there is no exchange.
    Returns [(world_base, num_worlds)] per shard."""
    if num_shards <= 0 or total_worlds < 0:
        raise ValueError("shard_worlds: bad arguments")
    q, r = divmod(total_worlds, num_shards)
    out, base = [], 0
    for g in range(num_shards):
        n = q + (1 if g < r else 0)
        out.append((base, n))
        base += n
    return out


class ShardedSimulator:
    """One simulator object over several GPUs of one process: `num_worlds` worlds
    are cut into contiguous shards (shard_worlds), one HideAndSeekSimulator per entry of `gpu_ids`
    with `world_base` set so that the episode RNG stays keyed by GLOBAL world index — results are
    those of one big simulator.  Worlds never interact, so there is no collective: `step()` launches
    every shard's step graph on its device's default stream from this thread, then waits for all
    of them.  A `*_tensor()` getter returns the LIST of per-shard tensors (a contiguous cross-GPU
    tensor does not exist); `worlds` gives each shard's [base, base + count).  The reference has
    no multi-GPU mode (`gpuID` is a single device)."""

    def __init__(self, gpu_ids, num_worlds, **kwargs):
        if not gpu_ids:
            raise ValueError("ShardedSimulator: gpu_ids is empty")
        for k in ("gpu_id", "world_base"):
            if k in kwargs:
                raise ValueError(f"ShardedSimulator: {k} is set per shard")
        self.gpu_ids = [int(g) for g in gpu_ids]
        self.worlds = shard_worlds(int(num_worlds), len(self.gpu_ids))
        self.num_worlds = int(num_worlds)
        self.shards = [HideAndSeekSimulator(gpu_id=g, num_worlds=n, world_base=b, **kwargs)
                       for g, (b, n) in zip(self.gpu_ids, self.worlds) if n > 0]

    def init(self):
        for s in self.shards:
            s.init()

    def step_async(self):
        for s in self.shards:
            s.step_async(0)

    def sync(self):
        for s in self.shards:
            s.sync(0)

    def step(self):
        self.step_async()
        self.sync()

    def overflow_count(self):
        return sum(s.overflow_count() for s in self.shards)

    def __getattr__(self, name):
        # tensor getters and per-world pokes of the reference API, fanned out
        if name.endswith("_tensor"):
            return lambda: [getattr(s, name)() for s in self.shards]
        raise AttributeError(name)

    def _locate(self, world):
        for s, (b, n) in zip(self.shards, [w for w in self.worlds if w[1] > 0]):
            if b <= world < b + n:
                return s, world - b
        raise IndexError(f"world {world} out of range")

    def trigger_reset(self, world, level=1):
        s, w = self._locate(world)
        s.trigger_reset(w, level)

    def set_action(self, agent_row, x, y, r, g=False, l=False):
        A = self.shards[0].agents_per_world
        s, w = self._locate(agent_row // A)
        s.set_action(w * A + agent_row % A, x, y, r, g, l)
