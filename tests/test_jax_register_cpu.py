"""sim.jax(): the registration module must import and describe its contract without jax or a GPU;
with jax absent `build` fails with a RuntimeError that names jax (not an ImportError from deep
inside), and the operand / result table follows SURVEY.md §8a's shapes."""
import importlib.util

import numpy as np
import pytest

HAVE_JAX = importlib.util.find_spec("jax") is not None


def test_module_imports_without_jax():
    from gpu_hideseek import jax_register
    assert callable(jax_register.build) and callable(jax_register.interface_table)


@pytest.mark.skipif(HAVE_JAX, reason="jax is installed here")
def test_build_names_the_missing_dependency():
    from gpu_hideseek import jax_register
    with pytest.raises(RuntimeError, match="jax is not installed"):
        jax_register.build(object(), True)
    with pytest.raises(RuntimeError, match="no CPU backend"):
        jax_register.build(object(), False)


def test_expected_table_matches_survey_byte_counts():
    from gpu_hideseek import jax_register
    N, A = 16000, 4
    t = jax_register.expected_interface_table(N, A)

    def nbytes(entries):
        return sum(int(np.prod(s)) * np.dtype(d).itemsize for _, s, d in entries)

    # the 11 observation tensors + reward + done are 1260 B per agent row
    per_row = (nbytes(t["step"]["results"]) - N * 2 * 4) / (N * A)
    assert per_row == 1260
    assert nbytes(t["step"]["operands"]) == N * A * 5 * 4 + N * 4 + N * A * 4
    assert [n for n, _, _ in t["init"]["results"]] == [n for n, _, _ in t["restore_ckpts"]["results"]]
    assert t["save_ckpts"]["results"][0][1] == (N, 1392)
    # capsules carry the name XLA looks for
    import ctypes
    cap = jax_register._capsule(0x1000)
    ctypes.pythonapi.PyCapsule_GetName.restype = ctypes.c_char_p
    ctypes.pythonapi.PyCapsule_GetName.argtypes = [ctypes.py_object]
    assert ctypes.pythonapi.PyCapsule_GetName(cap) == b"xla._CUSTOM_CALL_TARGET"
