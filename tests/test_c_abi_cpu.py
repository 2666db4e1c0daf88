"""The drop-in boundary on a machine without a GPU: the C-ABI library loads and exports every
function include/hideseek/c_api.h declares, the Manager / Python module expose the reference's
API surface (src/mgr.hpp:34-91, src/bindings.cpp:20-121), and construction fails loudly —
never falling back to a CPU path — when there is no CUDA device."""
import ctypes as C
import inspect
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "marl-hideandseek_b200")

# method names of the reference's Python class (src/bindings.cpp:75-98)
REFERENCE_METHODS = [
    "init", "step", "reset_tensor", "done_tensor", "prep_counter_tensor", "action_tensor",
    "reward_tensor", "self_data_tensor", "self_type_tensor", "self_mask_tensor", "agent_data_tensor",
    "box_data_tensor", "ramp_data_tensor", "visible_agents_mask_tensor", "visible_boxes_mask_tensor",
    "visible_ramps_mask_tensor", "global_positions_tensor", "depth_tensor", "rgb_tensor",
    "lidar_tensor", "seed_tensor", "ckpt_ctrl_tensor", "ckpt_tensor", "jax",
]
# constructor keywords (src/bindings.cpp:61-73)
REFERENCE_KWARGS = ["exec_mode", "gpu_id", "num_worlds", "sim_flags", "rand_seed", "min_hiders",
                    "max_hiders", "min_seekers", "max_seekers", "num_pbt_policies",
                    "enable_batch_renderer", "batch_render_width", "batch_render_height"]
# C++ Manager methods (src/mgr.hpp:34-91)
REFERENCE_MANAGER = [
    "init", "step", "saveCheckpoint", "loadCheckpoint", "loadCheckpoints", "resetTensor", "doneTensor",
    "prepCounterTensor", "actionTensor", "rewardTensor", "selfDataTensor", "selfTypeTensor",
    "selfMaskTensor", "agentDataTensor", "boxDataTensor", "rampDataTensor", "visibleAgentsMaskTensor",
    "visibleBoxesMaskTensor", "visibleRampsMaskTensor", "globalPositionsTensor", "lidarTensor",
    "seedTensor", "checkpointControlTensor", "checkpointTensor", "depthTensor", "rgbTensor",
    "triggerReset", "setAction", "policyAssignmentsTensor", "episodeResultTensor",
]


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_library_exports_every_declared_symbol():
    import gpu_hideseek as gh
    declared = gh.c_api_symbols()
    assert len(declared) >= 20
    lib = C.CDLL(os.path.join(PKG, "lib", "libhideseek_b200.so"))
    missing = [s for s in declared if not hasattr(lib, s)]
    assert not missing, missing


def test_header_is_plain_c():
    # the boundary must be consumable from C (cgo / JNI / ctypes style FFI): compile it as C
    src = '#include <hideseek/c_api.h>\nint main(void){ hs_config c; (void)c; return HS_NUM_EXPORTS == 21 ? 0 : 1; }\n'
    out = os.path.join(ROOT, "tests", "_build", "hdr_check")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-x", "c", "-", "-I" + os.path.join(ROOT, "include"),
                    "-o", out], input=src.encode(), check=True)
    assert subprocess.run([out]).returncode == 0


def test_manager_library_exports_reference_methods():
    out = subprocess.run(["nm", "-DC", os.path.join(PKG, "lib", "libhideseek_mgr.so")],
                         capture_output=True, text=True, check=True).stdout
    for m in REFERENCE_MANAGER:
        assert re.search(rf"GPUHideSeek::Manager::{m}\(", out), m
    assert "GPUHideSeek::Manager::Manager(GPUHideSeek::Manager::Config const&)" in out


def test_python_api_surface():
    import gpu_hideseek as gh
    for m in REFERENCE_METHODS:
        assert callable(getattr(gh.HideAndSeekSimulator, m)), m
    params = list(inspect.signature(gh.HideAndSeekSimulator.__init__).parameters)[1:]
    assert params[: len(REFERENCE_KWARGS)] == REFERENCE_KWARGS
    assert int(gh.SimFlags.UseFixedWorld | gh.SimFlags.RandomFlipTeams | gh.SimFlags.ZeroAgentVelocity) == 13
    assert hasattr(gh.madrona, "ExecMode") and hasattr(gh.madrona.ExecMode, "CUDA")
    assert hasattr(gh.madrona.Tensor, "to_torch") and hasattr(gh.madrona.Tensor, "to_jax")


def test_no_cpu_backend():
    import gpu_hideseek as gh
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        gh.HideAndSeekSimulator(exec_mode=gh.madrona.ExecMode.CPU, num_worlds=2)


def test_view_entry_points_reject_bad_arguments():
    # host-side argument checks of the agent-view entry points (no device needed)
    import gpu_hideseek as gh
    lib = gh.load_library()
    assert lib.hs_enable_views(None, 64, 64, None) != 0
    assert lib.hs_render_views(None, None) != 0
    with pytest.raises(RuntimeError, match="cannot open"):
        gh.view_mesh_table("/nonexistent-data-dir")
    t = gh.view_mesh_table()
    assert t["num_tris"] > 0 and abs(float(t["fov_scale"]) - 0.8390996) < 1e-6


@pytest.mark.skipif(_has_gpu(), reason="checks the no-GPU failure path")
def test_fails_loudly_without_gpu():
    import gpu_hideseek as gh
    with pytest.raises(RuntimeError, match="no CUDA device"):
        gh.HideAndSeekSimulator(num_worlds=2)


def test_product_does_not_reference_oracle():
    # the product tree must not import / link / load anything under oracle/
    bad = []
    for dirpath, _, files in os.walk(PKG):
        for f in files:
            if f.endswith((".cu", ".cuh", ".cpp", ".hpp", ".h", ".py")) or f == "Makefile":
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                if re.search(r"oracle/|libhs_ref|refsim", text):
                    bad.append(os.path.join(dirpath, f))
    assert not bad, bad
    needed = subprocess.run(["ldd", os.path.join(PKG, "lib", "libhideseek_b200.so")],
                            capture_output=True, text=True).stdout
    assert "hs_ref" not in needed
