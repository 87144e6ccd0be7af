"""One-process-per-GPU launcher (stands in for ``accelerate launch --config_file config/accelerate_config.yaml``,
``R/scripts/launch_sft.sh:7-8``).

    python -m dla_b200.cli.launch [--nproc N] [--topology config/accelerate_config.yaml] -m dla_b200.cli.train_dpo --config …

Topology comes from ``--nproc``, else ``num_processes`` of the accelerate-style YAML
(``R/config/accelerate_config.yaml:12``), else the visible GPU count; ``use_cpu: true`` (or no GPU) launches
gloo ranks on the CPU.  Rendezvous is static on 127.0.0.1:``main_process_port`` (29500).  A rank that dies
takes the job down (the children are one process group); exit codes propagate.
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
    if n == 1:
        os.execvp(cmd[0], cmd)
    procs: List[subprocess.Popen] = []
    for r in range(n):
        env = dict(os.environ, RANK=str(r), LOCAL_RANK=str(r), WORLD_SIZE=str(n), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        if use_cpu:
            env["CUDA_VISIBLE_DEVICES"] = ""
        procs.append(subprocess.Popen(cmd, env=env))
    rc = 0
    try:
        for p in procs:
            code = p.wait()
            if code != 0 and rc == 0:
                rc = code
                for q in procs:
                    if q.poll() is None:
                        q.send_signal(signal.SIGTERM)
    except KeyboardInterrupt:
        for q in procs:
            q.send_signal(signal.SIGTERM)
        rc = 130
    sys.exit(rc)


if __name__ == "__main__":
    main()
