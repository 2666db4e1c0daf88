"""`sim.jax(xla_gpu)` — the simulator as four jitted JAX functions.

What the reference builds in C++ + embedded Python (JAXInterface::buildEntry / setup,
M/include/madrona/py/bindings.inl:18-100, M/src/python/bindings.cpp:227-285, and the script
M/src/python/jax_register.py:13-207 it executes), restated for this package:

  * the four XLA GPU custom-call targets are the C-ABI functions hs_xla_gpu_init / _step /
    _save_checkpoints / _load_checkpoints of libhideseek_b200.so
    (`void fn(cudaStream_t, void **buffers, const char *opaque, size_t)`, the signature of
    JAXInterface::gpuEntryFn, bindings.inl:107-119); `opaque` carries the hs_sim pointer, buffer 0
    is the CPU-backend slot and buffer 1 JAX's ordering token — both skipped by the C side;
  * every call threads a zero-size float32 "token" (first operand, last result) so XLA keeps the
    calls ordered and never elides a step whose outputs go unused (jax_register.py:38-58);
  * operand / result shapes and dtypes come from the Manager's TrainInterface
    (src/mgr.cpp:1338-1375): `interface_table` below lists them without needing jax.

    fns = sim.jax(True)
    out = fns['init']()                       # {'state', 'obs': {...11 tensors...}}
    out = fns['step']({'state': out['state'], 'actions': {'actions': a}, 'resets': r,
                       'sim_ctrl': None, 'pbt': {'policy_assignments': p}})
    ck = fns['save_ckpts']({'state': s, 'should_save': m})          # {'state', 'ckpts'}
    ob = fns['restore_ckpts']({'state': s, 'should_restore': m, 'ckpt_data': c})

`sim_ctrl` (an empty tensor in the reference, src/mgr.cpp:1333-1336) is accepted and ignored: the
C entry point, like the reference's gpuStreamStep (src/mgr.cpp:379-398), consumes exactly
actions, resets and the policy assignments.

There is no CPU backend in this package: `xla_gpu=False` raises.
"""
import numpy as np


def interface_table(sim):
    """Operand / result lists of the four custom calls as (name, shape, dtype-name) tuples, token
    excluded — the contract between the Python lowering and the C entry points:

      init:          ()                                          -> obs x 11
      step:          actions, resets, policy_assignments         -> obs x 11, rewards, dones, episode_results
      save_ckpts:    should_save                                 -> checkpoint data
      restore_ckpts: should_restore, checkpoint data             -> obs x 11
    """
    ti = sim.train_interface()

    def sd(name, t):
        return (name, tuple(int(d) for d in t.shape), str(t.dtype))

    obs = [sd(k, t) for k, t in ti["step_outputs"]["observations"].items()]
    step_in = [sd(k, t) for k, t in ti["step_inputs"]["actions"].items()]
    step_in.append(sd("resets", ti["step_inputs"]["resets"]))
    step_in += [sd(k, t) for k, t in ti["step_inputs"]["pbt"].items()]
    step_out = obs + [sd("rewards", ti["step_outputs"]["rewards"]), sd("dones", ti["step_outputs"]["dones"])]
    step_out += [sd(k, t) for k, t in ti["step_outputs"]["pbt"].items()]
    ckpt = sd("ckpt_data", ti["checkpointing"]["checkpoint_data"])
    ctrl = sd("ckpt_ctrl", sim.ckpt_ctrl_tensor())
    return {
        "init": {"operands": [], "results": obs},
        "step": {"operands": step_in, "results": step_out},
        "save_ckpts": {"operands": [ctrl], "results": [ckpt]},
        "restore_ckpts": {"operands": [ctrl, ckpt], "results": obs},
    }


def expected_interface_table(num_worlds, agents_per_world, checkpoint_bytes=1392):
    """The same table from the configuration alone (shapes of src/mgr.cpp:1062-1331 / SURVEY.md
    §8a): what `interface_table(sim)` must return for a simulator of `num_worlds` worlds with
    `agents_per_world` agent slots.  Lets the contract be checked where no GPU (and no jax) exists."""
    N, R = int(num_worlds), int(num_worlds) * int(agents_per_world)
    obs = [("prep_counter", (R, 1), "int32"), ("self_data", (R, 13), "float32"), ("self_type", (R, 1), "int32"),
           ("self_mask", (R, 1), "float32"), ("self_lidar", (R, 30), "float32"), ("agent_data", (R, 5, 14), "float32"),
           ("box_data", (R, 9, 17), "float32"), ("ramp_data", (R, 2, 14), "float32"),
           ("vis_agents_mask", (R, 5, 1), "float32"), ("vis_boxes_mask", (R, 9, 1), "float32"),
           ("vis_ramps_mask", (R, 2, 1), "float32")]
    step_in = [("actions", (R, 5), "int32"), ("resets", (N, 1), "int32"), ("policy_assignments", (R, 1), "int32")]
    step_out = obs + [("rewards", (R, 1), "float32"), ("dones", (R, 1), "int32"), ("episode_results", (N, 2), "float32")]
    ctrl, ckpt = ("ckpt_ctrl", (N, 1), "int32"), ("ckpt_data", (N, checkpoint_bytes), "uint8")
    return {
        "init": {"operands": [], "results": obs},
        "step": {"operands": step_in, "results": step_out},
        "save_ckpts": {"operands": [ctrl], "results": [ckpt]},
        "restore_ckpts": {"operands": [ctrl, ckpt], "results": obs},
    }


def _capsule(address):
    """PyCapsule named like XLA expects (nb::capsule(fn, "xla._CUSTOM_CALL_TARGET"))."""
    import ctypes
    new = ctypes.pythonapi.PyCapsule_New
    new.restype = ctypes.py_object
    new.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_void_p]
    return new(address, b"xla._CUSTOM_CALL_TARGET", None)


def build(sim, xla_gpu=True):
    """Registers the custom calls for `sim` and returns {'init', 'step', 'save_ckpts',
    'restore_ckpts'} (jitted), like JAXInterface::setup.  Raises RuntimeError when jax is missing."""
    if not xla_gpu:
        raise RuntimeError("gpu_hideseek.jax: this package has no CPU backend; call sim.jax(True)")
    try:
        import jax
        from jax import numpy as jnp
        from jax.interpreters import mlir
        from jax.interpreters.mlir import ir
    except ImportError as e:
        raise RuntimeError(f"gpu_hideseek.jax: jax is not installed in this environment ({e}); the "
                           "custom-call entry points are available through sim.xla_custom_calls()") from e
    from functools import partial

    try:                                        # jax >= 0.4.16
        from jax.extend.core import Primitive
    except ImportError:                         # pragma: no cover - older jax
        from jax.core import Primitive
    from jax.core import ShapedArray
    try:
        from jax.interpreters import xla
        apply_primitive = xla.apply_primitive
    except (ImportError, AttributeError):       # pragma: no cover - newer jax
        from jax._src import dispatch
        apply_primitive = dispatch.apply_primitive
    dtype_to_ir_type = mlir.dtype_to_ir_type

    def custom_call(name, *, operands, operand_layouts, result_types, result_layouts, backend_config):
        # the original "status-less" custom-call ABI (api_version 1), side-effecting
        if hasattr(mlir, "custom_call"):
            return mlir.custom_call(name, result_types=result_types, operands=operands,
                                    backend_config=backend_config, has_side_effect=True, api_version=1,
                                    operand_layouts=operand_layouts, result_layouts=result_layouts).results
        from jaxlib.hlo_helpers import custom_call as legacy      # pragma: no cover - older jaxlib
        return legacy(name, backend_config=backend_config, operands=operands, operand_layouts=operand_layouts,
                      result_types=result_types, result_layouts=result_layouts, has_side_effect=True).results

    def register(name, address):
        cap = _capsule(address)
        try:                                    # jax >= 0.4.31
            import jax.ffi as ffi
            ffi.register_ffi_target(name, cap, platform="CUDA", api_version=0)
        except (ImportError, AttributeError):   # pragma: no cover - older jax
            try:
                from jax.extend import ffi
                ffi.register_ffi_target(name, cap, platform="CUDA", api_version=0)
            except (ImportError, AttributeError):
                from jax.lib import xla_client
                xla_client.register_custom_call_target(name, cap, platform="gpu")

    cc = sim.xla_custom_calls()
    table = interface_table(sim)
    prefix = f"{type(sim).__name__}_{id(sim)}"
    opaque = cc["opaque"]
    sim_ptr = np.uint64(sim.handle())

    def row_major(shape):
        return tuple(range(len(shape) - 1, -1, -1))

    def token_type():
        return ir.RankedTensorType.get((0,), dtype_to_ir_type(np.dtype("float32")))

    def abstract(entries):
        return tuple(ShapedArray(shape, np.dtype(dt)) for _, shape, dt in entries)

    def make_primitive(key):
        name = f"{prefix}_{key}"
        register(name, cc["fns"][key])
        results = table[key]["results"]

        def lowering(ctx, token, *flat_inputs):
            in_types = [ir.RankedTensorType(i.type) for i in flat_inputs]
            in_layouts = [row_major(t.shape) for t in in_types]
            res_types = [ir.RankedTensorType.get(shape, dtype_to_ir_type(np.dtype(dt))) for _, shape, dt in results]
            res_layouts = [row_major(shape) for _, shape, _ in results]
            out = custom_call(
                name, backend_config=opaque,
                # operand 0: the CPU-backend slot of the reference's contract, operand 1: the token
                operands=[mlir.ir_constant(sim_ptr), token, *flat_inputs],
                operand_layouts=[(), (0,), *in_layouts],
                result_types=[*res_types, token_type()],
                result_layouts=[*res_layouts, (0,)])
            *out, tok = out
            return [tok, *out]

        prim = Primitive(name)
        prim.multiple_results = True
        prim.def_impl(partial(apply_primitive, prim))
        prim.def_abstract_eval(lambda *a: (ShapedArray((0,), np.dtype("float32")), *abstract(results)))
        mlir.register_lowering(prim, lowering, platform="cuda")
        return prim

    prims = {k: make_primitive(k) for k in ("init", "step", "save_ckpts", "restore_ckpts")}
    obs_names = [n for n, _, _ in table["init"]["results"]]
    action_names = list(sim.train_interface()["step_inputs"]["actions"])
    pbt_in = list(sim.train_interface()["step_inputs"]["pbt"])
    pbt_out = list(sim.train_interface()["step_outputs"]["pbt"])

    def init_func():
        state, *flat = prims["init"].bind(jnp.empty((0,), jnp.float32))
        return {"state": state, "obs": dict(zip(obs_names, flat))}

    def step_func(step_inputs):
        flat_in = [step_inputs["state"]]
        flat_in += [step_inputs["actions"][k] for k in action_names]
        flat_in.append(step_inputs["resets"])
        flat_in += [step_inputs["pbt"][k] for k in pbt_in]          # 'sim_ctrl' is empty: not passed on
        state, *flat = prims["step"].bind(*flat_in)
        n = len(obs_names)
        return {"state": state, "obs": dict(zip(obs_names, flat[:n])), "rewards": flat[n], "dones": flat[n + 1],
                "pbt": dict(zip(pbt_out, flat[n + 2:]))}

    def save_ckpts_func(save_inputs):
        state, data = prims["save_ckpts"].bind(save_inputs["state"], save_inputs["should_save"])
        return {"state": state, "ckpts": data}

    def restore_ckpts_func(restore_inputs):
        state, *flat = prims["restore_ckpts"].bind(restore_inputs["state"], restore_inputs["should_restore"],
                                                   restore_inputs["ckpt_data"])
        return {"state": state, "obs": dict(zip(obs_names, flat))}

    return {"init": jax.jit(init_func), "step": jax.jit(step_func),
            "save_ckpts": jax.jit(save_ckpts_func), "restore_ckpts": jax.jit(restore_ckpts_func)}
