"""Inter-communicators, ports and the name service ("Process Creation and Management" plus the inter-communicator
part of "Groups, Contexts, Communicators").

/root/reference/mpifuncs.c wraps ``Intercomm_create / Intercomm_merge / Comm_remote_size / Comm_remote_group``,
``Open_port / Close_port / Publish_name / Lookup_name / Unpublish_name``, ``Comm_accept / Comm_connect / Comm_join /
Comm_disconnect / Comm_get_parent`` ("Process mgmt 10/12"; ``Comm_spawn[_multiple]`` are the two
it leaves out, /root/reference/readspec.py:73-84).  No script in the tree calls any of them.

The native message layer fixes its world at ``Init`` (a full TCP mesh between the ranks of one launch), so everything
here connects GROUPS OF THE SAME JOB -- e.g. the server ranks and the client ranks of a parameter-server run -- which is
also the only use MPI programs without a spawn-capable launcher can make of these calls.  An inter-communicator is two
groups plus a private "bridge" communicator over their union, created without a collective through
``mpit_msg_comm_from_ranks``; point-to-point calls given an :class:`Intercomm` address ranks of the REMOTE group, as
MPI specifies.  ``Comm_spawn`` raises ``ERR_SPAWN``: start every process with ``python -m mpit_b200.launch.run`` (or
torchrun) and split roles by rank, as the reference's launchers do (/root/reference/asyncsgd/mlaunch.lua:40-46).
"""
from __future__ import annotations

import ctypes
import os
import struct
from typing import Dict, List, Optional, Tuple

import torch

from .._native import host

ERR_COMM, ERR_ARG, ERR_NAME, ERR_PORT, ERR_SERVICE, ERR_SPAWN = 5, 13, 33, 38, 41, 42
MAX_PORT_NAME = 256
_TAG_PORT, _TAG_JOIN = 32300, 32290


def _m():
    from .. import msg
    return msg


def _err(code, what):
    return _m().MPIError(code, what)


class Intercomm:
    """Two disjoint groups of one job.  ``bridge`` spans both, low group first; ``offset`` maps a remote rank to
    its rank in the bridge."""

    def __init__(self, local_comm: int, local_ranks: List[int], remote_ranks: List[int], bridge: int, low: bool):
        self.local_comm, self.bridge, self.low = local_comm, bridge, low
        self.local_ranks, self.remote_ranks = list(local_ranks), list(remote_ranks)
        self.offset = len(local_ranks) if low else 0
        self.name = ""

    def __repr__(self):
        return f"Intercomm(local={self.local_ranks}, remote={self.remote_ranks})"


_seq = [0]


def _world_ranks(comm: int) -> List[int]:
    return _m().Comm_group(comm).ranks


def _build(local_comm: int, local_leader: int, peer_comm: int, remote_leader: Optional[int], tag: int) -> Intercomm:
    """Leaders swap (sequence number, member list) over ``peer_comm``, tell their groups, and everybody derives the
    same bridge communicator.  ``remote_leader`` None = learn it from whoever contacts us (the accepting side)."""
    m = _m()
    me_local = m.Comm_rank(local_comm)
    mine = _world_ranks(local_comm)                        # collective over local_comm
    # the local group agrees on a sequence number that no member has used before
    top = torch.tensor([_seq[0]], dtype=torch.int64)
    m.Allreduce(top.clone(), top, 1, m.LONG, m.MAX, local_comm)
    header = torch.zeros(2, dtype=torch.int64)             # (remote sequence number, remote group size)
    if me_local == local_leader:
        out = torch.tensor([int(top[0]), len(mine)] + mine, dtype=torch.int64)
        if remote_leader is None:
            st = m.Status()
            m.Probe(m.ANY_SOURCE, tag, peer_comm, st)
            remote_leader = st.source
        m.Send(out, out.numel(), m.LONG, remote_leader, tag, peer_comm)
        st = m.Status()
        m.Probe(remote_leader, tag, peer_comm, st)
        n = m.Get_count(st, m.LONG)
        got = torch.empty(n, dtype=torch.int64)
        m.Recv(got, n, m.LONG, remote_leader, tag, peer_comm)
        header[:] = got[:2]
        theirs_t = got[2:].clone()
    m.Bcast(header, 2, m.LONG, local_leader, local_comm)
    if me_local != local_leader:
        theirs_t = torch.empty(int(header[1]), dtype=torch.int64)
    m.Bcast(theirs_t, theirs_t.numel(), m.LONG, local_leader, local_comm)
    theirs = theirs_t.tolist()
    if set(theirs) & set(mine):
        raise _err(ERR_COMM, "inter-communicator: the two groups overlap")
    seq = max(int(top[0]), int(header[0])) + 1
    _seq[0] = seq
    low = min(mine) < min(theirs)
    union = (mine + theirs) if low else (theirs + mine)
    seed = ((min(mine[0], theirs[0]) * 1009 + max(mine[0], theirs[0])) * 4099 + tag * 31 + seq) & 0x7FFFFFFF
    arr = (ctypes.c_int * len(union))(*union)
    bridge = ctypes.c_int(-1)
    m._chk(host().mpit_msg_comm_from_ranks(arr, len(union), seed, ctypes.byref(bridge)), "Intercomm_create")
    return Intercomm(m.Comm_dup(local_comm), mine, theirs, bridge.value, low)


def Intercomm_create(local_comm: int, local_leader: int, peer_comm: int, remote_leader: int, tag: int) -> Intercomm:
    """Collective over both local communicators.  ``peer_comm`` / ``remote_leader`` matter at the two leaders only."""
    return _build(local_comm, local_leader, peer_comm, remote_leader, tag)


def Intercomm_merge(intercomm: Intercomm, high: bool) -> int:
    """An intra-communicator over both groups; the group that passes ``high=False`` is ordered first (both passing the
    same value: the group holding the lowest world rank first)."""
    m = _m()
    # learn the other side's flag: leaders swap it over the bridge, then tell their group
    flag = torch.tensor([int(bool(high))], dtype=torch.int64)
    other = torch.zeros(1, dtype=torch.int64)
    if m.Comm_rank(intercomm.local_comm) == 0:
        peer = intercomm.offset                                   # remote rank 0 in the bridge
        m.Sendrecv(flag, 1, m.LONG, peer, 32291, other, 1, m.LONG, peer, 32291, intercomm.bridge)
    m.Bcast(other, 1, m.LONG, 0, intercomm.local_comm)
    mine_first = intercomm.low if int(other[0]) == int(flag[0]) else not high
    nloc, nrem = len(intercomm.local_ranks), len(intercomm.remote_ranks)
    me = m.Comm_rank(intercomm.local_comm)
    key = me if mine_first else nrem + me
    return m.Comm_split(intercomm.bridge, 0, key)


def Comm_remote_size(intercomm: Intercomm) -> int:
    return len(_inter(intercomm).remote_ranks)


def Comm_remote_group(intercomm: Intercomm):
    return _m().Group(_inter(intercomm).remote_ranks)


def _inter(c) -> Intercomm:
    if not isinstance(c, Intercomm):
        raise _err(ERR_COMM, "not an inter-communicator")
    return c


def Comm_disconnect(comm) -> int:
    """Waits for nothing more than ``Comm_free`` does here: sends are handed to the progress thread when they return.
    Returns COMM_NULL."""
    m = _m()
    if isinstance(comm, Intercomm):
        m.Comm_free(comm.local_comm)
        m.Comm_free(comm.bridge)
    else:
        m.Comm_free(comm)
    return m.COMM_NULL


def Comm_get_parent() -> int:
    """COMM_NULL: no process of this framework is created by ``Comm_spawn``."""
    return _m().COMM_NULL


def Comm_spawn(command, argv=None, maxprocs: int = 1, info=None, root: int = 0, comm: int = 0):
    raise _err(ERR_SPAWN, "Comm_spawn: the process set is fixed at launch (python -m mpit_b200.launch.run -np N); "
                          "split roles by rank instead")


def Comm_spawn_multiple(count, commands, argvs=None, maxprocs=None, infos=None, root: int = 0, comm: int = 0):
    raise _err(ERR_SPAWN, "Comm_spawn_multiple: the process set is fixed at launch")


# ---- ports and the name service ------------------------------------------------------------------------------------
_open_ports: Dict[str, int] = {}
_port_seq = [0]


def Open_port(info=None) -> str:
    """A port name other groups of this job can ``Comm_connect`` to: ``mpit-port:<world rank>:<number>``."""
    m = _m()
    _port_seq[0] += 1
    name = f"mpit-port:{m.Comm_rank(m.COMM_WORLD)}:{_port_seq[0]}"
    _open_ports[name] = _port_seq[0]
    return name


def Close_port(port_name: str) -> None:
    if _open_ports.pop(port_name, None) is None:
        raise _err(ERR_PORT, f"Close_port({port_name!r})")


def _parse_port(port_name: str) -> Tuple[int, int]:
    try:
        kind, rank, num = port_name.split(":")
        if kind != "mpit-port":
            raise ValueError
        return int(rank), int(num)
    except ValueError:
        raise _err(ERR_PORT, f"malformed port name {port_name!r}") from None


def _names_dir() -> str:
    d = os.environ.get("MPIT_NAMESERVICE_DIR")
    if not d:
        job = os.environ.get("MPIT_PORT", os.environ.get("MASTER_PORT", "0"))
        d = os.path.join(os.environ.get("TMPDIR", "/tmp"), f"mpit-names-{os.getuid()}-{job}")
    os.makedirs(d, exist_ok=True)
    return d


def _name_file(service_name: str) -> str:
    return os.path.join(_names_dir(), service_name.encode().hex())


def Publish_name(service_name: str, info, port_name: str) -> None:
    """The name service is a directory shared by the job's processes (``MPIT_NAMESERVICE_DIR``, default under /tmp,
    keyed by the job's rendezvous port)."""
    path = _name_file(service_name)
    try:
        fd = os.open(path, os.O_WRONLY | os.O_CREAT | os.O_EXCL, 0o600)
    except FileExistsError:
        raise _err(ERR_SERVICE, f"Publish_name: {service_name!r} is already published") from None
    with os.fdopen(fd, "w") as f:
        f.write(port_name)


def Unpublish_name(service_name: str, info=None, port_name: Optional[str] = None) -> None:
    try:
        os.unlink(_name_file(service_name))
    except FileNotFoundError:
        raise _err(ERR_SERVICE, f"Unpublish_name: {service_name!r} is not published") from None


def Lookup_name(service_name: str, info=None) -> str:
    try:
        with open(_name_file(service_name)) as f:
            return f.read()
    except FileNotFoundError:
        raise _err(ERR_NAME, f"Lookup_name: {service_name!r} is not published") from None


def Comm_accept(port_name: str, info, root: int, comm: int) -> Intercomm:
    """Collective over ``comm``; ``root`` must be the process that opened the port."""
    m = _m()
    owner, num = _parse_port(port_name)
    if m.Comm_rank(comm) == root and (port_name not in _open_ports or owner != m.Comm_rank(m.COMM_WORLD)):
        raise _err(ERR_PORT, "Comm_accept: the port is not open on the root")
    return _build(comm, root, m.COMM_WORLD, None, _TAG_PORT + num)


def Comm_connect(port_name: str, info, root: int, comm: int) -> Intercomm:
    """Collective over ``comm``; blocks until the owner of the port calls ``Comm_accept``."""
    m = _m()
    owner, num = _parse_port(port_name)
    return _build(comm, root, m.COMM_WORLD, owner, _TAG_PORT + num)


def Comm_join(fd: int) -> Intercomm:
    """Two processes of the job that share a connected socket become an inter-communicator of one process each."""
    m = _m()
    me = m.Comm_rank(m.COMM_WORLD)
    os.write(fd, struct.pack("<i", me))
    buf = b""
    while len(buf) < 4:
        chunk = os.read(fd, 4 - len(buf))
        if not chunk:
            raise _err(ERR_ARG, "Comm_join: the socket closed during the handshake")
        buf += chunk
    peer = struct.unpack("<i", buf)[0]
    return _build(m.COMM_SELF, 0, m.COMM_WORLD, peer, _TAG_JOIN)


# ---- routing of point-to-point calls that were handed an inter-communicator ----------------------------------------
def route(comm, peer: int) -> Tuple[int, int, int]:
    """-> (native communicator, native peer rank, offset to subtract from a status' source)."""
    if isinstance(comm, Intercomm):
        return comm.bridge, (peer + comm.offset if peer >= 0 else peer), comm.offset
    return comm, peer, 0


__all__ = [_n for _n in dir() if _n[0].isupper() and _n not in ("Any", "Callable", "Dict", "List", "Optional", "Sequence",
                                                                 "Tuple")]
