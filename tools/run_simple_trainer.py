"""Runs the REFERENCE's trainer — `examples/simple_trainer.py`, unmodified, from baseline/_ref — on top of the b200
backend (SURVEY 8f-1 / VERDICT r1 "next" 9), through INTEGRATION.md option B: this repo's `gsplat` package provides
`rasterization()`, `gsplat.distributed`, `gsplat.relocation`, `SelectiveAdam`; everything else the trainer imports from
`gsplat` (`strategy/`, `compression/`, `utils.py`) is the reference's own code via GSPLAT_REFERENCE_PATH.

    python tools/run_simple_trainer.py [--steps 300] [--strategy default|mcmc] [--packed] [--visible-adam]

Offline stand-ins (tools/trainer_stubs/): the COLMAP data loader becomes a synthetic multi-view capture rendered from
the reference's garden fixture; imageio / viser / nerfview / torchmetrics / fused_ssim / matplotlib are minimal
replacements (SSIM is computed in torch, LPIPS reports NaN: no pretrained weights offline); trajectory videos are
skipped. The loss (L1 + SSIM), the optimisers, the SH-degree schedule, `DefaultStrategy` / `MCMCStrategy` with their
`strategy/ops.py` (duplicate / split / remove / reset_opa / relocate / sample_add) and the evaluation loop all run as
written by the reference. Prints the trainer's log and a final `TRAINER {...}` JSON line.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--strategy", default="default", choices=["default", "mcmc"])
    ap.add_argument("--packed", action="store_true")
    ap.add_argument("--visible-adam", action="store_true")
    ap.add_argument("--result-dir", default="/tmp/gsb_simple_trainer")
    args = ap.parse_args()
    if not os.path.isdir(os.path.join(REF, "examples")):
        raise SystemExit("baseline/_ref/examples is missing: run tools/install_reference.sh in the build container")
    os.environ["GSPLAT_BACKEND"] = "b200"
    os.environ["GSPLAT_REFERENCE_PATH"] = REF
    sys.path[:0] = [os.path.join(ROOT, "tools", "trainer_stubs"), os.path.join(ROOT, "gsplat-triton_b200"),
                    os.path.join(REF, "examples")]
    import torch

    import simple_trainer as ST  # the reference's file
    from gsplat.b200_impl._lib import lib

    assert os.path.realpath(ST.__file__).startswith(os.path.realpath(REF)), ST.__file__
    assert "gsplat-triton_b200" in ST.rasterization.__module__ or "gsplat.rendering" == ST.rasterization.__module__
    import gsplat.rendering as R

    assert R.BACKEND == "b200" and os.path.realpath(R.__file__).startswith(os.path.realpath(ROOT))
    import gsplat.strategy as S

    assert os.path.realpath(S.__file__).startswith(os.path.realpath(REF)), "strategies must be the reference's"

    steps = args.steps
    if args.strategy == "default":
        # the reference's schedule (refine 500..15000 every 100, reset every 3000) compressed onto this short run so
        # that duplicate / split / remove and reset_opa all execute
        strategy = ST.DefaultStrategy(verbose=True, refine_start_iter=steps // 6, refine_stop_iter=steps,
                                      refine_every=max(steps // 12, 10), reset_every=steps // 2, pause_refine_after_reset=0)
    else:
        strategy = ST.MCMCStrategy(verbose=True, cap_max=60_000, refine_start_iter=steps // 6, refine_stop_iter=steps,
                                   refine_every=max(steps // 12, 10))
    cfg = ST.Config(disable_viewer=True, data_dir="synthetic", data_factor=1, result_dir=args.result_dir,
                    max_steps=steps, eval_steps=[steps], save_steps=[steps], ply_steps=[], save_ply=False,
                    sh_degree=3, sh_degree_interval=max(steps // 4, 1), packed=args.packed,
                    visible_adam=args.visible_adam, strategy=strategy, tb_every=50,
                    **({"opacity_reg": 0.01, "scale_reg": 0.01, "init_opa": 0.5, "init_scale": 0.1}
                       if args.strategy == "mcmc" else {}))
    ST.Runner.render_traj = lambda self, step: None  # needs a video writer
    lib.gsb_reset_launch_count()
    t0 = time.time()
    runner = ST.Runner(0, 0, 1, cfg)
    n0 = len(runner.splats["means"])
    runner.train()
    torch.cuda.synchronize()
    wall = time.time() - t0
    stats = json.load(open(os.path.join(args.result_dir, "stats", f"val_step{steps - 1:04d}.json")))
    print("TRAINER " + json.dumps({
        "trainer": "reference examples/simple_trainer.py (unmodified) on GSPLAT_BACKEND=b200", "strategy": args.strategy,
        "packed": args.packed, "visible_adam": args.visible_adam, "steps": steps,
        "gaussians_start": n0, "gaussians_end": len(runner.splats["means"]),
        "val_psnr": stats.get("psnr"), "val_ssim": stats.get("ssim"), "wall_s": round(wall, 1),
        "b200_kernel_launches": int(lib.gsb_launch_count())}), flush=True)


if __name__ == "__main__":
    main()
