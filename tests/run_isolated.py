"""Runs each collected test in its own process (a faulting kernel poisons the CUDA context of
its process only) and writes one line per test plus the failing output to gpurun_out/.

    python tests/run_isolated.py [-k expr] [--timeout 120] [pytest paths...]
"""
import argparse
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("-k", default=None)
    ap.add_argument("--timeout", type=int, default=180)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "isolated.log"))
    ap.add_argument("--per-case", action="store_true")
    ap.add_argument("paths", nargs="*", default=["tests"])
    a = ap.parse_args()
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    cmd = [sys.executable, "-m", "pytest", "--collect-only", "-q", "-m", "gpu", *a.paths]
    if a.k:
        cmd += ["-k", a.k]
    ids = [l.strip() for l in subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True).stdout.splitlines() if "::" in l]
    if not a.per_case:  # one process per test FUNCTION (all its parametrisations together)
        ids = list(dict.fromkeys(i.split("[")[0] for i in ids))
    npass = 0
    with open(a.out, "w") as log:
        for tid in ids:
            t0 = time.time()
            try:
                r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-m", "gpu", tid, "--tb=short", "--no-header", "-p", "no:cacheprovider"],
                                   cwd=ROOT, capture_output=True, text=True, timeout=a.timeout)
                ok, txt = r.returncode == 0, r.stdout[-6000:] + r.stderr[-1500:]
            except subprocess.TimeoutExpired as e:
                ok, txt = False, f"TIMEOUT after {a.timeout}s\n{(e.stdout or b'')[-2000:]}"
            npass += ok
            summary = [l for l in txt.splitlines() if " passed" in l or " failed" in l or " skipped" in l or " error" in l]
            line = f"{'PASS' if ok else 'FAIL'} {time.time() - t0:6.1f}s {tid}  [{summary[-1].strip() if summary else '?'}]"
            print(line, flush=True)
            log.write(line + "\n")
            if not ok:
                log.write(txt + "\n" + "-" * 100 + "\n")
            log.flush()
    print(f"{npass}/{len(ids)} passed")
    return 0 if npass == len(ids) else 1


if __name__ == "__main__":
    sys.exit(main())
