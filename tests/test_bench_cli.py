"""bench.py's command line on a box without a GPU: the driver's flags parse, the training-step
record is on unless --no-train-step, the GPU arm refuses to run without a CUDA device (no CPU
fallback behind the headline number), and rank != 0 of the reference arm leaves without work."""
import importlib
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench():
    sys.path.insert(0, ROOT)
    try:
        return importlib.import_module("bench")
    finally:
        sys.path.remove(ROOT)


def _parse(bench, argv, monkeypatch):
    monkeypatch.setattr(sys, "argv", ["bench.py"] + argv)
    return bench.parse()


def test_driver_flags_parse_and_defaults(monkeypatch):
    bench = _bench()
    a = _parse(bench, [], monkeypatch)
    assert (a.gpus, a.impl) == (1, "ours") and a.warmup >= 3 and not a.no_train_step
    a = _parse(bench, ["--gpus", "8", "--steps", "20", "--warmup", "3"], monkeypatch)
    assert (a.gpus, a.steps, a.warmup) == (8, 20, 3)
    a = _parse(bench, ["--impl", "reference", "--gpus", "2", "--steps", "2", "--warmup", "1"], monkeypatch)
    assert a.impl == "reference"
    # command lines of the round's earlier profiles still parse; the leg can be switched off
    assert not _parse(bench, ["--train-step"], monkeypatch).no_train_step
    assert _parse(bench, ["--no-train-step"], monkeypatch).no_train_step
    assert 60 <= bench.TRAIN_STEP_LIMIT_S <= 600


@pytest.mark.skipif(torch.cuda.is_available(), reason="needs a box without a GPU")
def test_gpu_arm_refuses_to_run_without_a_gpu():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "1"],
                       capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode != 0
    assert "no CPU fallback" in (r.stderr + r.stdout)
    assert "{" not in r.stdout          # no bench line


def test_reference_arm_other_ranks_exit_without_work():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                       capture_output=True, text=True, timeout=300, cwd=ROOT, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""
