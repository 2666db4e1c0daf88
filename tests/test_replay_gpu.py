"""bin/replay — the windowless version of the reference viewer's replay loop
(src/viewer.cpp:12-27, 186-216): a log recorded step by step (all worlds' Checkpoint records)
must replay to the same observations, and --dump-views must write the agent views."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REPLAY = os.path.join(ROOT, "marl-hideandseek_b200", "bin", "replay")


def _lines(out):
    return [l for l in out.splitlines() if l.startswith("step ")]


def test_record_then_replay(tmp_path):
    log = str(tmp_path / "run.log")
    N, steps = 6, 14
    rec = subprocess.run([REPLAY, str(N), "--record", log, "--steps", str(steps)], capture_output=True, text=True, check=True)
    assert os.path.getsize(log) == steps * N * 1392
    views = tmp_path / "views"
    views.mkdir()
    rep = subprocess.run([REPLAY, str(N), "--replay", log, "--print-obs", "--dump-views", str(views)],
                         capture_output=True, text=True, check=True)
    a, b = _lines(rec.stdout), _lines(rep.stdout)
    assert len(a) == steps and len(b) == steps
    for la, lb in zip(a, b):
        fa, fb = la.split(), lb.split()
        assert fa[:4] == fb[:4], (la, lb)                       # step index and prep counter
        assert abs(float(fa[5]) - float(fb[5])) <= 1e-3 * max(1.0, abs(float(fa[5]))), (la, lb)
    assert "Self Data" in rep.stdout and "Reward" in rep.stdout
    files = sorted(os.listdir(views))
    assert len(files) == steps * 6 * 2
    pgm = open(views / "step0003_agent0.pgm", "rb").read()
    assert pgm.startswith(b"P5\n64 64\n255\n") and len(pgm) == len(b"P5\n64 64\n255\n") + 64 * 64
    assert max(pgm[-64 * 64:]) > 0                              # the view sees something
    ppm = open(views / "step0003_agent0.ppm", "rb").read()
    assert ppm.startswith(b"P6\n64 64\n255\n") and len(ppm) == len(b"P6\n64 64\n255\n") + 64 * 64 * 3


def test_usage_errors():
    assert subprocess.run([REPLAY, "4"], capture_output=True).returncode != 0
    assert subprocess.run([REPLAY, "4", "--record", "/tmp/a", "--replay", "/tmp/b"], capture_output=True).returncode != 0
