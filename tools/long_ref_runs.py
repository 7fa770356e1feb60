"""Long CPU runs of the UNMODIFIED reference at headline scale.

  python tools/long_ref_runs.py [out.jsonl]

Runs baseline/_ref/a.out (main.cpp as shipped, -Ofast, in-tree MPI shim) with 8 ranks on the host:
  1. n = 32760, m = 126  (the reference's 3x3-blocked fast path), residual included
  2. n = 32768, m = 128  (config 2 exactly; scalar path), stopped after the glob_time line
Each result is appended as one JSON line with the program's own stdout.
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "baseline", "ref_cpu"))
import run_reference as rr  # noqa: E402

out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "ref_long_runs.jsonl")
err = rr.ensure_built()
if err:
    sys.exit(err)
done = set()
if os.path.exists(out):
    for ln in open(out):
        try:
            r = json.loads(ln)
            if r.get("glob_time"):
                done.add((r["n"], r["m"]))
        except ValueError:
            pass
for n, m, want_res, limit in ((32760, 126, True, 3 * 3600), (32768, 128, False, 6 * 3600)):
    if (n, m) in done:      # finished in an earlier session
        continue
    t0 = time.time()
    r = rr.run_once(n, m, 8, timeout_s=limit, want_residual=want_res)
    r.update(n=n, m=m, ranks=8, started=t0, host_cpus=os.cpu_count())
    if r["glob_time"]:
        r["gflops_2n3"] = 2.0 * n ** 3 / r["glob_time"] / 1e9
    with open(out, "a") as f:
        f.write(json.dumps(r) + "\n")
