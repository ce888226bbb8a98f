"""Run the reference's experiment drivers (code/setups/*.py), UNMODIFIED, against the drop-in package.

Each script runs in a scratch working directory as `python <script>` with the package directory first on the module
path, so `from network import *` / `from soup import *` / `import keras.backend` resolve to
the B200 engine.  The scripts come from /root/reference when it exists (this container) and otherwise from the copies
`oracle/stage_ref.py` staged under oracle/_ref/setups/ (GPU box); either way the file is executed as is.

    python tools/run_ref_setups.py [--only name.py ...] [--out gpurun_out/ref_setups.json] [--timeout 900]

Prints one JSON object per script: wall seconds, return code, the experiment's log.txt, the dill files it wrote.
Used by tests/test_gpu_unmodified_setups.py and by bench.py's `facade` leg.
"""
import argparse
import glob
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "self-replicating-neural-networks_b200")
REF_SETUPS = "/root/reference/code/setups"
STAGED = os.path.join(ROOT, "oracle", "_ref", "setups")


def setups_dir():
    for d in (REF_SETUPS, STAGED):
        if glob.glob(os.path.join(d, "*.py")):
            return d
    return None


def run_script(script, workdir=None, seed=1234, timeout=900, env_extra=None):
    """Execute one unmodified setup script; returns a dict (rc, seconds, log, files, workdir, tail)."""
    workdir = workdir or tempfile.mkdtemp(prefix="srnn_setup_")
    env = dict(os.environ)
    env["PYTHONPATH"] = PKG + os.pathsep + env.get("PYTHONPATH", "")
    env.setdefault("SRNN_SEED", str(seed))
    env["PYTHONHASHSEED"] = "0"
    if env_extra:
        env.update(env_extra)
    t0 = time.time()
    try:
        proc = subprocess.run([sys.executable, script], cwd=workdir, env=env, stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, timeout=timeout)
        rc, out = proc.returncode, proc.stdout.decode(errors="replace")
    except subprocess.TimeoutExpired as exc:
        rc, out = -9, (exc.stdout or b"").decode(errors="replace") + "\n[timeout]"
    seconds = time.time() - t0
    exps = sorted(glob.glob(os.path.join(workdir, "experiments", "exp-*")))
    logs, files = {}, {}
    for e in exps:
        name = os.path.basename(e)
        lp = os.path.join(e, "log.txt")
        logs[name] = open(lp).read() if os.path.exists(lp) else None
        files[name] = sorted(os.path.basename(p) for p in glob.glob(os.path.join(e, "*")))
    return {"script": os.path.basename(script), "rc": rc, "seconds": round(seconds, 3), "logs": logs, "files": files,
            "workdir": workdir, "tail": out[-1500:]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", nargs="*", default=None)
    ap.add_argument("--out", default=None)
    ap.add_argument("--timeout", type=int, default=900)
    args = ap.parse_args()
    d = setups_dir()
    if d is None:
        print(json.dumps({"unavailable": "no reference setups: run oracle/stage_ref.py where /root/reference exists"}))
        return 0
    results = []
    for script in sorted(glob.glob(os.path.join(d, "*.py"))):
        if args.only and os.path.basename(script) not in args.only:
            continue
        res = run_script(script, timeout=args.timeout)
        results.append(res)
        print(json.dumps(res), flush=True)
    if args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        with open(args.out, "w") as fh:
            json.dump(results, fh, indent=1)
    return 0 if all(r["rc"] == 0 for r in results) else 1


if __name__ == "__main__":
    sys.exit(main())
