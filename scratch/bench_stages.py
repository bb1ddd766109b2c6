"""Experiment harness: run bench.py (kernels only) on every .so in scratch/variants and print the stage rates."""
import glob, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in sorted(glob.glob(os.path.join(ROOT, "scratch", "variants", "*.so"))):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--lib", p, "--no-cpu-baseline", "--no-e2e", "--steps", "10", "--batch", "64"],
                       capture_output=True, text=True)
    try:
        d = json.loads(r.stdout.strip().splitlines()[-1])
    except Exception:
        print("FAILED", p, r.stderr[-500:]); continue
    o = d["other_stages"]
    print(os.path.basename(p), "apply", round(d["stage_mpix_s"]["apply"] / 1e3, 1), "gen", round(d["stage_mpix_s"]["generate"] / 1e3, 1),
          " ".join(f"{k.replace('apply_','').replace('api0_fused_tonemap_generate','fused')[:14]}={v['mpix_s']/1e3:.0f}" for k, v in o.items()), flush=True)
