"""tools/train_step.py (BASELINE.json configs[4] harness) stays runnable: a 1+1-layer model, two
steps, eager and as one CUDA graph -- the op and the fused front-end must be capturable (they
enqueue on the capturing stream, allocate through torch only and never synchronise)."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("extra", [[], ["--cuda-graph"], ["--torch-frontend"]], ids=["eager", "graph", "torch_frontend"])
def test_train_step_runs(extra):
    cmd = [sys.executable, os.path.join(ROOT, "tools", "train_step.py"), "--layers", "1", "--batch", "2",
           "--steps", "2", "--warmup", "1"] + extra
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert res.returncode == 0, res.stderr[-2000:]
    line = [l for l in res.stdout.splitlines() if l.startswith("{")][-1]
    rec = json.loads(line)
    assert rec["n_gpus"] == 1 and rec["value"] > 0 and rec["loss"] == rec["loss"]      # finite, not NaN
    # 1 encoder + 1 decoder layer.  Default: the encoder layer runs the fused module op (query-order
    # pre-pass + forward, backward reusing the plan: 3 launches), the decoder layer the front-end kernel +
    # op, both ways (4).  --torch-frontend: 2 op forwards + 2 backwards + the encoder's pre-pass.
    assert rec["op_launches_per_step"] == (5 if "--torch-frontend" in extra else 7)
