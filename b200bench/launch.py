"""Single-host launcher: one harness process per GPU (or per CPU rank with ``--device cpu``).

Replaces the reference's pod-per-GPU Kubernetes launch (``scripts/launch_multi.sh:57-82`` +
``k8s/job-{master,workers}.template.yaml``) for the 8xB200 single-box target, and adds what the
reference leaves to the cluster: a per-run timeout, rank-failure propagation (the
whole process group is killed as soon as one rank dies) and a non-zero exit code on failure.

    python -m b200bench.launch --nproc 8 --timeout 900 -- --strategy ddp --tier A --seq-len 2048 \
        --steps 100 --per-device-batch 1 --grad-accum 4 --results-dir results/ddp_ws8
"""
from __future__ import annotations

import argparse
import os
import signal
import socket
import subprocess
import sys
import time
from typing import List


def free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def launch(nproc: int, harness_args: List[str], timeout: float = 900.0, master_port: int = 0,
           log_dir: str = "", env_extra: dict = None) -> int:
    port = master_port or free_port()
    procs: List[subprocess.Popen] = []
    logs = []
    for r in range(nproc):
        cmd = [sys.executable, "-u", "-m", "b200bench.cli", "--world-size", str(nproc), "--rank", str(r),
               "--local-rank", str(r), "--master-addr", "127.0.0.1", "--master-port", str(port)] + harness_args
        env = dict(os.environ)
        env.update(env_extra or {})
        env.setdefault("PYTHONFAULTHANDLER", "1")
        out = None
        if r > 0 and log_dir:
            os.makedirs(log_dir, exist_ok=True)
            out = open(os.path.join(log_dir, f"rank{r}.log"), "w")
            logs.append(out)
        elif r > 0:
            out = subprocess.DEVNULL
        procs.append(subprocess.Popen(cmd, env=env, stdout=out, stderr=subprocess.STDOUT if out else None,
                                      start_new_session=True))
    deadline = time.time() + timeout
    rc = 0
    try:
        while True:
            states = [p.poll() for p in procs]
            if any(s not in (None, 0) for s in states):
                bad = [i for i, s in enumerate(states) if s not in (None, 0)]
                print(f"[launch] rank(s) {bad} failed with {[states[i] for i in bad]}; killing the group",
                      file=sys.stderr)
                rc = next(s for s in states if s not in (None, 0))
                break
            if all(s == 0 for s in states):
                break
            if time.time() > deadline:
                print(f"[launch] timeout after {timeout:.0f}s; killing the group", file=sys.stderr)
                rc = 124
                break
            time.sleep(0.2)
    finally:
        for p in procs:
            if p.poll() is None:
                try:
                    os.killpg(p.pid, signal.SIGTERM)
                except ProcessLookupError:
                    pass
        t_end = time.time() + 5
        for p in procs:
            try:
                p.wait(timeout=max(0.1, t_end - time.time()))
            except subprocess.TimeoutExpired:
                try:
                    os.killpg(p.pid, signal.SIGKILL)
                except ProcessLookupError:
                    pass
        for f in logs:
            f.close()
    return rc


def main() -> None:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--nproc", type=int, required=True)
    ap.add_argument("--timeout", type=float, default=900.0)
    ap.add_argument("--master-port", type=int, default=0)
    ap.add_argument("--log-dir", type=str, default="")
    ap.add_argument("rest", nargs=argparse.REMAINDER)
    a = ap.parse_args()
    rest = a.rest[1:] if a.rest and a.rest[0] == "--" else a.rest
    sys.exit(launch(a.nproc, rest, a.timeout, a.master_port, a.log_dir))


if __name__ == "__main__":
    main()
