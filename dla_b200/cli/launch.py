"""One-process-per-GPU launcher (stands in for ``accelerate launch --config_file config/accelerate_config.yaml``,
``R/scripts/launch_sft.sh:7-8``).

    python -m dla_b200.cli.launch [--nproc N] [--topology config/accelerate_config.yaml] -m dla_b200.cli.train_dpo --config …

Topology comes from ``--nproc``, else ``num_processes`` of the accelerate-style YAML
(``R/config/accelerate_config.yaml:12``), else the visible GPU count; ``use_cpu: true`` (or no GPU) launches
gloo ranks on the CPU.  Rendezvous is static on 127.0.0.1:``main_process_port`` (29500).  All ranks are watched:
the first one that dies takes the others down at once (no waiting for a collective to time out) and its exit code
propagates; ``--max-restarts N`` (or ``DLA_MAX_RESTARTS``) relaunches the job up to N times with ``--resume latest``
(resume-from-latest as crash recovery; the reference has neither).
"""
from __future__ import annotations

import argparse
import os
import signal
import subprocess
import sys
from typing import List

import yaml


def main() -> None:
    ap = argparse.ArgumentParser(add_help=False)
    ap.add_argument("--nproc", type=int, default=None)
    ap.add_argument("--topology", default="config/accelerate_config.yaml")
    ap.add_argument("--port", type=int, default=None)
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--max-restarts", type=int, default=int(os.environ.get("DLA_MAX_RESTARTS", 0)),
                    help="relaunch all ranks with --resume latest when a rank exits non-zero (crash recovery)")
    ap.add_argument("-m", dest="module", required=True)
    args, rest = ap.parse_known_args()
    topo = {}
    if os.path.exists(args.topology):
        with open(args.topology) as fh:
            topo = yaml.safe_load(fh) or {}
    import torch

    use_cpu = args.cpu or bool(topo.get("use_cpu")) or not torch.cuda.is_available()
    n = args.nproc or int(os.environ.get("DLA_NPROC", 0)) or int(topo.get("num_processes", 0) or 0)
    if not use_cpu:
        n = min(n, torch.cuda.device_count()) if n else torch.cuda.device_count()
    n = max(n, 1)
    port = args.port or int(topo.get("main_process_port", 29500))
    cmd = [sys.executable, "-m", args.module, *rest]
    if n == 1 and args.max_restarts <= 0:
        os.execvp(cmd[0], cmd)
    rc = 0
    for attempt in range(args.max_restarts + 1):
        if attempt > 0:
            # crash recovery: every rank is relaunched from the last complete checkpoint; a fresh port
            # avoids the previous rendezvous socket
            if "--resume" not in cmd:
                cmd = cmd + ["--resume", "latest"]
            print(f"[launch] attempt {attempt + 1}/{args.max_restarts + 1}: restarting {n} rank(s) with --resume latest", flush=True)
        rc = _run_once(cmd, n, port + attempt, use_cpu)
        if rc == 0 or rc == 130:
            break
    sys.exit(rc)


def _run_once(cmd: List[str], n: int, port: int, use_cpu: bool) -> int:
    """Start the ranks and watch ALL of them: the first non-zero exit takes the others down (SIGTERM, SIGKILL after a
    grace period) instead of leaving them blocked in a collective until a timeout."""
    import time

    procs: List[subprocess.Popen] = []
    for r in range(n):
        env = dict(os.environ, RANK=str(r), LOCAL_RANK=str(r), WORLD_SIZE=str(n), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        if use_cpu:
            env["CUDA_VISIBLE_DEVICES"] = ""
        procs.append(subprocess.Popen(cmd, env=env))
    rc = 0
    try:
        while any(p.poll() is None for p in procs):
            for p in procs:
                code = p.poll()
                if code is not None and code != 0 and rc == 0:
                    rc = code
                    for q in procs:
                        if q.poll() is None:
                            q.send_signal(signal.SIGTERM)
                    deadline = time.monotonic() + 15.0
                    while any(q.poll() is None for q in procs) and time.monotonic() < deadline:
                        time.sleep(0.1)
                    for q in procs:
                        if q.poll() is None:
                            q.kill()
            time.sleep(0.05)
        for p in procs:
            if p.returncode != 0 and rc == 0:
                rc = p.returncode
    except KeyboardInterrupt:
        for q in procs:
            if q.poll() is None:
                q.send_signal(signal.SIGTERM)
        rc = 130
    return rc


if __name__ == "__main__":
    main()
