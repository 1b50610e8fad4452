"""``python -m mpit_b200.launch.run -np N script.py [args...]`` -- the ``mpirun -np N th script.lua``
of this framework (/root/reference/README.md:30,54,60): starts N local processes, gives each its
rank/size/rendezvous through the environment, forwards output, and tears everything down if one rank
fails (MPI's default error handler aborts the job; so does this)."""
from __future__ import annotations

import argparse
import os
import signal
import socket
import subprocess
import sys
import time


def free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def launch(nproc: int, argv: list, env_extra: dict | None = None, timeout: float | None = None,
           module: bool = False) -> int:
    port = free_port()
    procs = []
    for r in range(nproc):
        env = dict(os.environ)
        env.update({"MPIT_RANK": str(r), "MPIT_WORLD_SIZE": str(nproc), "MPIT_ADDR": "127.0.0.1",
                    "MPIT_PORT": str(port), "OMP_NUM_THREADS": env.get("OMP_NUM_THREADS", "1")})
        if env_extra:
            env.update(env_extra)
        cmd = [sys.executable] + (["-m"] if module else []) + list(argv)
        procs.append(subprocess.Popen(cmd, env=env, start_new_session=True))
    t0 = time.time()
    rc = 0
    live = set(range(nproc))
    try:
        while live:
            for r in list(live):
                code = procs[r].poll()
                if code is None:
                    continue
                live.discard(r)
                if code != 0 and rc == 0:
                    rc = code
                    sys.stderr.write(f"[mpit run] rank {r} exited with {code}; aborting the job\n")
                    for q in live:
                        _kill(procs[q])
            if timeout is not None and time.time() - t0 > timeout:
                sys.stderr.write("[mpit run] timeout; aborting the job\n")
                for q in live:
                    _kill(procs[q])
                rc = rc or 124
                break
            time.sleep(0.02)
    finally:
        for p in procs:
            if p.poll() is None:
                _kill(p)
    return rc


def _kill(p: subprocess.Popen) -> None:
    try:
        os.killpg(p.pid, signal.SIGKILL)   # exactly the process group we started
    except ProcessLookupError:
        pass


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="mpit_b200.launch.run")
    ap.add_argument("-np", "-n", dest="np", type=int, required=True)
    ap.add_argument("-m", dest="module", action="store_true", help="run a module instead of a script")
    ap.add_argument("--timeout", type=float, default=None)
    ap.add_argument("cmd", nargs=argparse.REMAINDER)
    a = ap.parse_args(argv)
    return launch(a.np, a.cmd, timeout=a.timeout, module=a.module)


if __name__ == "__main__":
    sys.exit(main())
