#!/usr/bin/env python
"""Run bench.py once per BASELINE.json config (workload) on this box and collect the JSON lines.

  python tools/run_configs.py [--out gpurun_out/configs.jsonl] [--gpus N] [workload ...]

Each line is exactly what `python bench.py --workload W` printed; the curated copy lives under profiles/.
"""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "configs.jsonl"))
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--extra", default="")
    ap.add_argument("workloads", nargs="*", default=["bunny", "atrium", "bathroom", "micro", "4k"])
    args = ap.parse_args()
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    with open(args.out, "a") as out:
        for w in args.workloads:
            if args.gpus > 1:
                cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
                       "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.join(ROOT, "bench.py"), "--gpus", str(args.gpus)]
            else:
                cmd = [sys.executable, os.path.join(ROOT, "bench.py")]
            cmd += ["--workload", w] + args.extra.split()
            r = subprocess.run(cmd, capture_output=True, text=True, cwd=ROOT)
            lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
            if r.returncode != 0 or not lines:
                sys.stderr.write(f"[{w}] failed rc={r.returncode}\n{r.stderr[-2000:]}\n")
                out.write(json.dumps({"workload": w, "error": r.stderr[-500:]}) + "\n")
                continue
            rec = json.loads(lines[-1]); rec["workload_key"] = w
            out.write(json.dumps(rec) + "\n"); out.flush()
            print(f"[{w}] {rec['value']:.1f} {rec['unit']}  ({rec['ms_per_step']:.3f} ms/step)")


if __name__ == "__main__":
    main()
