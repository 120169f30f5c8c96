"""GPU: the reference's own trainer and densification strategies on top of the b200 backend.

Runs `tools/run_simple_trainer.py` — the UNMODIFIED `examples/simple_trainer.py` and `gsplat/strategy/*` from
`baseline/_ref` (tools/install_reference.sh; git-ignored, travels to the GPU box) — for a short schedule in which
duplicate / split / remove / reset_opa (DefaultStrategy) resp. relocate / sample_add / noise injection (MCMCStrategy, on
gsb_relocation and gsb_selective_adam) all execute, and checks that the scene is being learnt. Skipped when the
reference install is not present (a fresh clone without the build container).
"""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HAVE_REF = os.path.isdir(os.path.join(ROOT, "baseline", "_ref", "examples"))


@pytest.mark.skipif(not HAVE_REF, reason="baseline/_ref (the reference install) is not present")
@pytest.mark.parametrize("extra,min_psnr", [([], 17.0), (["--strategy", "mcmc", "--packed", "--visible-adam"], 22.0)])
def test_reference_trainer_runs_on_the_b200_backend(tmp_path, extra, min_psnr):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "run_simple_trainer.py"), "--steps", "120",
                        "--result-dir", str(tmp_path)] + extra, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("TRAINER ")][-1]
    out = json.loads(line[len("TRAINER "):])
    assert out["b200_kernel_launches"] > 1000
    assert out["gaussians_end"] != out["gaussians_start"], "the densification strategy never changed the scene"
    assert out["val_psnr"] > min_psnr, out
