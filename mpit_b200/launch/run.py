"""``python -m mpit_b200.launch.run -np N script.py [args...]`` -- the ``mpirun -np N th script.lua``
of this framework (/root/reference/README.md:30,54,60): starts N processes, gives each its
rank/size/rendezvous through the environment, forwards output, and tears everything down if one rank
fails (MPI's default error handler aborts the job; so does this).

``--hostfile FILE`` spreads the ranks over several machines the way ``mpirun --hostfile`` does with the
reference's ``host:slots`` lines (/root/reference/BiCNN/hostfiles:1-6; ``host slots=N`` is read too): ranks fill
the hosts slot by slot, entries naming this machine are started directly, the others through a remote shell
(``--rsh``, default ``ssh -o BatchMode=yes``) that must find the same checkout at the same path.  Rank 0's host
is the rendezvous address unless ``--master-addr`` says otherwise."""
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


def parse_hostfile(text: str) -> list:
    """-> [(host, slots), ...] from ``host:slots`` / ``host slots=N`` / bare ``host`` lines; ``#`` starts a comment."""
    hosts = []
    for line in text.splitlines():
        line = line.split("#", 1)[0].strip()
        if not line:
            continue
        name, slots = line, 1
        if "slots=" in line:
            name, rest = line.split(None, 1)
            slots = int(rest.split("slots=", 1)[1].split()[0])
        elif ":" in line:
            name, n = line.rsplit(":", 1)
            slots = int(n)
        if slots < 1:
            raise ValueError(f"hostfile: {line!r} offers no slot")
        hosts.append((name.strip(), slots))
    if not hosts:
        raise ValueError("hostfile lists no host")
    return hosts


def place_ranks(nproc: int, hosts: list) -> list:
    """Host of every rank, filling each host's slots before moving on; wraps around when the job is larger than the
    total number of slots (oversubscription, as ``mpirun`` allows on request)."""
    order = [h for h, n in hosts for _ in range(n)]
    return [order[r % len(order)] for r in range(nproc)]


def _is_local(host: str) -> bool:
    return host in ("localhost", "127.0.0.1", "::1", socket.gethostname(), socket.getfqdn())


_FORWARD = ("MPIT_", "MASTER_", "OMP_NUM_THREADS", "MEGS", "CUDA_VISIBLE_DEVICES", "NCCL_", "PYTHONPATH")


def _remote_command(rsh: str, host: str, env: dict, cmd: list) -> list:
    import shlex
    keep = {k: v for k, v in env.items() if k.startswith(_FORWARD)}
    line = "cd " + shlex.quote(os.getcwd()) + " && exec env " + \
        " ".join(f"{k}={shlex.quote(v)}" for k, v in sorted(keep.items())) + " " + " ".join(shlex.quote(c) for c in cmd)
    return shlex.split(rsh) + [host, line]


def launch(nproc: int, argv: list, env_extra: dict | None = None, timeout: float | None = None,
           module: bool = False, hosts: list | None = None, rsh: str = "ssh -o BatchMode=yes",
           master_addr: str | None = None) -> int:
    port = free_port()
    placement = place_ranks(nproc, hosts) if hosts else ["localhost"] * nproc
    if master_addr is None:
        master_addr = "127.0.0.1" if _is_local(placement[0]) and all(_is_local(h) for h in placement) else \
            (socket.gethostname() if _is_local(placement[0]) else placement[0])
    procs = []
    for r in range(nproc):
        env = dict(os.environ)
        env.update({"MPIT_RANK": str(r), "MPIT_WORLD_SIZE": str(nproc), "MPIT_ADDR": master_addr,
                    "MPIT_PORT": str(port), "OMP_NUM_THREADS": env.get("OMP_NUM_THREADS", "1")})
        if env_extra:
            env.update(env_extra)
        cmd = [sys.executable] + (["-m"] if module else []) + list(argv)
        if not _is_local(placement[r]):
            cmd = _remote_command(rsh, placement[r], env, cmd)
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
    ap.add_argument("--hostfile", "-hostfile", "--machinefile", dest="hostfile", default=None,
                    help="lines of host:slots (BiCNN/hostfiles) or host slots=N")
    ap.add_argument("--rsh", default="ssh -o BatchMode=yes", help="remote shell used for hosts other than this one")
    ap.add_argument("--master-addr", default=None)
    ap.add_argument("-x", dest="env", action="append", default=[], metavar="NAME[=VALUE]",
                    help="export an environment variable to every rank (mpirun -x); NAME alone forwards its current value")
    ap.add_argument("--oversubscribe", action="store_true", help="accepted for mpirun compatibility (always allowed)")
    ap.add_argument("cmd", nargs=argparse.REMAINDER)
    a = ap.parse_args(argv)
    hosts = parse_hostfile(open(a.hostfile).read()) if a.hostfile else None
    extra = {}
    for item in a.env:
        name, eq, value = item.partition("=")
        extra[name] = value if eq else os.environ.get(name, "")
    global _FORWARD
    _FORWARD = _FORWARD + tuple(extra)               # also across the remote shell
    return launch(a.np, a.cmd, env_extra=extra or None, timeout=a.timeout, module=a.module, hosts=hosts, rsh=a.rsh,
                  master_addr=a.master_addr)


if __name__ == "__main__":
    sys.exit(main())
