"""Process topologies: cartesian grids and graphs attached to a communicator.

/root/reference/mpifuncs.c wraps the sixteen MPI-1 topology calls ("Topology 16/16":
``Cart_create/coords/rank/shift/sub/get/map``, ``Cartdim_get``, ``Dims_create``, ``Graph_create/get/map/neighbors/
neighbors_count``, ``Graphdims_get``, ``Topo_test``); nothing in the tree calls them.  On one NVSwitch box every GPU
is one hop from every other, so rank reordering has nothing to optimise: ``reorder`` is accepted and ranks keep
their order.  A topology is a record keyed by the communicator handle; the communicators themselves come from the
native ``Comm_split``.
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

GRAPH, CART, DIST_GRAPH = 1, 2, 3
ERR_TOPOLOGY, ERR_DIMS, ERR_ARG, ERR_RANK = 11, 12, 13, 6

_TOPO: Dict[int, dict] = {}


def _m():
    from .. import msg
    return msg


def _err(code, what):
    return _m().MPIError(code, what)


def _cart(comm: int) -> dict:
    t = _TOPO.get(comm)
    if not t or t["kind"] != CART:
        raise _err(ERR_TOPOLOGY, "communicator has no cartesian topology")
    return t


def _graph(comm: int) -> dict:
    t = _TOPO.get(comm)
    if not t or t["kind"] != GRAPH:
        raise _err(ERR_TOPOLOGY, "communicator has no graph topology")
    return t


def Dims_create(nnodes: int, dims: Sequence[int]) -> List[int]:
    """Balanced factorisation of ``nnodes`` over the entries of ``dims`` that are 0 (others are kept);
    the free dimensions come out in non-increasing order."""
    dims = list(dims)
    fixed = 1
    for d in dims:
        if d < 0:
            raise _err(ERR_DIMS, "Dims_create: negative dimension")
        if d:
            fixed *= d
    free = [i for i, d in enumerate(dims) if d == 0]
    if fixed == 0 or nnodes % fixed:
        raise _err(ERR_DIMS, "Dims_create: nnodes is not a multiple of the fixed dimensions")
    rest = nnodes // fixed
    if not free:
        if rest != 1:
            raise _err(ERR_DIMS, "Dims_create: fixed dimensions do not multiply to nnodes")
        return dims
    # prime factors, largest first, each given to the currently smallest dimension
    factors, p = [], 2
    while p * p <= rest:
        while rest % p == 0:
            factors.append(p)
            rest //= p
        p += 1
    if rest > 1:
        factors.append(rest)
    out = [1] * len(free)
    for f in sorted(factors, reverse=True):
        out[out.index(min(out))] *= f
    out.sort(reverse=True)
    for i, v in zip(free, out):
        dims[i] = v
    return dims


def Cart_create(comm_old: int, dims: Sequence[int], periods: Sequence[bool], reorder: bool = False) -> int:
    """Collective.  Ranks beyond prod(dims) get COMM_NULL.  Row-major rank order."""
    m = _m()
    n = 1
    for d in dims:
        n *= d
    if n > m.Comm_size(comm_old):
        raise _err(ERR_DIMS, "Cart_create: the grid is larger than the communicator")
    me = m.Comm_rank(comm_old)
    new = m.Comm_split(comm_old, 0 if me < n else m.UNDEFINED, me)
    if new != m.COMM_NULL:
        _TOPO[new] = {"kind": CART, "dims": list(dims), "periods": [bool(p) for p in periods]}
    return new


def Cartdim_get(comm: int) -> int:
    return len(_cart(comm)["dims"])


def Cart_get(comm: int) -> Tuple[List[int], List[bool], List[int]]:
    """-> (dims, periods, coords of the calling process)."""
    t = _cart(comm)
    return list(t["dims"]), list(t["periods"]), Cart_coords(comm, _m().Comm_rank(comm))


def Cart_coords(comm: int, rank: int) -> List[int]:
    dims = _cart(comm)["dims"]
    coords = []
    for d in reversed(dims):
        coords.insert(0, rank % d)
        rank //= d
    return coords


def Cart_rank(comm: int, coords: Sequence[int]) -> int:
    t = _cart(comm)
    r = 0
    for c, d, per in zip(coords, t["dims"], t["periods"]):
        if per:
            c %= d
        elif not 0 <= c < d:
            raise _err(ERR_RANK, "Cart_rank: coordinate out of range in a non-periodic dimension")
        r = r * d + c
    return r


def Cart_shift(comm: int, direction: int, disp: int) -> Tuple[int, int]:
    """-> (rank_source, rank_dest) for a shift by ``disp`` along ``direction``; PROC_NULL past a non-periodic edge."""
    m = _m()
    t = _cart(comm)
    if not 0 <= direction < len(t["dims"]):
        raise _err(ERR_DIMS, "Cart_shift: no such dimension")
    me = Cart_coords(comm, m.Comm_rank(comm))

    def at(offset):
        c = list(me)
        c[direction] += offset
        d = t["dims"][direction]
        if t["periods"][direction]:
            c[direction] %= d
        elif not 0 <= c[direction] < d:
            return m.PROC_NULL
        return Cart_rank(comm, c)

    return at(-disp), at(disp)


def Cart_sub(comm: int, remain_dims: Sequence[bool]) -> int:
    """Split the grid into sub-grids that keep the dimensions flagged in ``remain_dims``."""
    m = _m()
    t = _cart(comm)
    me = Cart_coords(comm, m.Comm_rank(comm))
    color = key = 0
    for c, d, keep in zip(me, t["dims"], remain_dims):
        if keep:
            key = key * d + c
        else:
            color = color * d + c
    new = m.Comm_split(comm, color, key)
    _TOPO[new] = {"kind": CART, "dims": [d for d, k in zip(t["dims"], remain_dims) if k],
                  "periods": [p for p, k in zip(t["periods"], remain_dims) if k]}
    return new


def Cart_map(comm: int, dims: Sequence[int], periods: Sequence[bool]) -> int:
    """The rank this process would get in ``Cart_create(comm, dims, periods, reorder=True)`` (UNDEFINED if none)."""
    m = _m()
    n = 1
    for d in dims:
        n *= d
    me = m.Comm_rank(comm)
    return me if me < n else m.UNDEFINED


def Graph_create(comm_old: int, index: Sequence[int], edges: Sequence[int], reorder: bool = False) -> int:
    """``index[i]`` = total number of neighbours of nodes 0..i; ``edges`` = the concatenated adjacency lists."""
    m = _m()
    nnodes = len(index)
    if nnodes > m.Comm_size(comm_old):
        raise _err(ERR_ARG, "Graph_create: more nodes than processes")
    if nnodes and index[-1] != len(edges) or any(not 0 <= e < nnodes for e in edges):
        raise _err(ERR_ARG, "Graph_create: malformed index / edges")
    me = m.Comm_rank(comm_old)
    new = m.Comm_split(comm_old, 0 if me < nnodes else m.UNDEFINED, me)
    if new != m.COMM_NULL:
        _TOPO[new] = {"kind": GRAPH, "index": list(index), "edges": list(edges)}
    return new


def Graphdims_get(comm: int) -> Tuple[int, int]:
    t = _graph(comm)
    return len(t["index"]), len(t["edges"])


def Graph_get(comm: int) -> Tuple[List[int], List[int]]:
    t = _graph(comm)
    return list(t["index"]), list(t["edges"])


def Graph_neighbors_count(comm: int, rank: int) -> int:
    idx = _graph(comm)["index"]
    return idx[rank] - (idx[rank - 1] if rank else 0)


def Graph_neighbors(comm: int, rank: int) -> List[int]:
    t = _graph(comm)
    lo = t["index"][rank - 1] if rank else 0
    return t["edges"][lo:t["index"][rank]]


def Graph_map(comm: int, index: Sequence[int], edges: Sequence[int]) -> int:
    m = _m()
    me = m.Comm_rank(comm)
    return me if me < len(index) else m.UNDEFINED


def Topo_test(comm: int) -> int:
    """GRAPH, CART or UNDEFINED."""
    t = _TOPO.get(comm)
    return t["kind"] if t else _m().UNDEFINED


def forget(comm: int) -> None:
    _TOPO.pop(comm, None)


__all__ = [_n for _n in dir() if _n[0].isupper() and _n not in ("Any", "Callable", "Dict", "List", "Optional", "Sequence",
                                                                 "Tuple")]
