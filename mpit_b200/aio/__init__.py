"""Coroutine-driven non-blocking IO: ``aio_send`` / ``aio_recv`` / ``co_execute`` / ``co_ping`` /
``co_wait`` and the FIFO ``Queue`` (reference: /root/reference/init.lua:41-192, queue.lua:3-47).

Lua coroutines become Python generators.  A coroutine body is a generator function that starts with
``yield signal_INIT`` (the reference bodies start with ``coroutine.yield(mpiT.signal_INIT)``,
/root/reference/asyncsgd/pserver.lua:39) and ends with ``yield signal_DONE``; inside it, IO is
``yield from aio_send(...)`` / ``yield from aio_recv(...)``.  The scheduler is the reference's
single-threaded round robin: ``co_ping`` pops one coroutine, resumes it once and pushes it back unless
it signalled DONE.

Differences that are deliberate:
* ``aio_recv`` cancels for real.  The reference breaks out of its probe loop when ``state.io`` goes
  false but then still posts the Irecv and never reaches its Cancel branch, leaking a receive
  (/root/reference/init.lua:81-101); here a receive is only posted once a matching message has been
  probed, and if ``state.io`` drops while it is in flight it is cancelled, waited and checked.
* Progress does not depend on pinging: the native layer moves bytes in a background thread, the
  coroutines only observe completion.
"""
from __future__ import annotations

import time
from typing import Callable, Generator, Iterable, Optional

signal_INIT, signal_EXEC, signal_OK, signal_ERR, signal_DONE = 1, 2, 3, 4, 5


class Queue:
    """FIFO with the reference's interface (queue.lua:3-47): push / pop / empty / len / clear."""

    def __init__(self):
        self.first = 0
        self.last = -1
        self.q = {}

    def push(self, value) -> None:
        self.last += 1
        self.q[self.last] = value

    def empty(self) -> bool:
        return self.first > self.last

    def pop(self):
        if self.empty():
            return None
        v = self.q.pop(self.first)
        if self.first == self.last:
            self.first, self.last = 0, -1
        else:
            self.first += 1
        return v

    def len(self) -> int:
        return self.last - self.first + 1

    __len__ = len

    def clear(self) -> None:
        self.q.clear()
        self.first, self.last = 0, -1


class IOState(dict):
    """The ``state`` table shared by a server/client and its coroutines: ``state.io`` gates IO,
    ``state.on`` gates service loops (/root/reference/asyncsgd/pserver.lua:26-29)."""

    def __init__(self, io: bool = True, on: bool = True, **kw):
        super().__init__(io=io, on=on, **kw)

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v


def aio_send(buf, count: int, dtype: int, dest: int, tag: int, comm: int, state: Optional[IOState] = None,
             cb: Optional[Callable] = None) -> Generator[int, None, None]:
    """Isend, then yield EXEC until it completes; if ``state.io`` goes false first: Cancel + Wait +
    Test_cancelled and stop when the cancel took (/root/reference/init.lua:41-67)."""
    from .. import msg as mpiT
    state = state if state is not None else IOState()
    req, status = mpiT.Request(1), mpiT.Status(1)
    mpiT.Isend(buf, count, dtype, dest, tag, comm, req)
    finished = False
    while not finished:
        if state["io"]:
            finished = mpiT.Test(req, status)
        else:
            mpiT.Cancel(req)
            mpiT.Wait(req, status)
            if mpiT.Test_cancelled(status):
                break
            finished = True   # too late to cancel: it went out
        if not finished:
            yield signal_EXEC
    if cb:
        cb(state)
    yield signal_OK


def aio_recv(buf, count: int, dtype: int, source: int, tag: int, comm: int, state: Optional[IOState] = None,
             cb: Optional[Callable] = None, status=None) -> Generator[int, None, None]:
    """Iprobe until a matching message exists (so the buffer is never posted early -- no unexpected-
    message buffering of GB-sized parameters, /root/reference/init.lua:77-84), then Irecv + Test
    (:88-103).  The generator's return value (``ok = yield from aio_recv(...)``) tells whether data
    actually landed in ``buf``; it is False when ``state.io`` dropped first."""
    from .. import msg as mpiT
    state = state if state is not None else IOState()
    req = mpiT.Request(1)
    status = status if status is not None else mpiT.Status(1)
    found = False
    while not found:
        if not state["io"]:
            if cb:
                cb(state)
            yield signal_OK
            return False
        found = mpiT.Iprobe(source, tag, comm, status)
        yield signal_EXEC
    mpiT.Irecv(buf, count, dtype, source, tag, comm, req)
    finished = False
    while not finished:
        if state["io"]:
            finished = mpiT.Test(req, status)
            if not finished:
                yield signal_EXEC
        else:
            mpiT.Cancel(req)
            mpiT.Wait(req, status)
            finished = not mpiT.Test_cancelled(status)
            break
    if cb:
        cb(state)
    yield signal_OK
    return finished


def co_execute(fn: Callable[..., Generator], args: Iterable = ()):
    """Create the coroutine and run it to its first yield (/root/reference/init.lua:139-150)."""
    co = fn(*args)
    try:
        first = next(co)
    except StopIteration as e:
        raise AssertionError("a coroutine body must yield at least once (signal_INIT)") from e
    assert first is not None
    return co


def co_ping(coq: Queue) -> bool:
    """Resume one queued coroutine; False when the queue is empty (/root/reference/init.lua:154-181).
    A coroutine that yields DONE is resumed once more so that it can return, then dropped; any
    exception inside a coroutine propagates (the reference asserts)."""
    if coq.empty():
        return False
    co = coq.pop()
    try:
        sig = next(co)
    except StopIteration:
        return True   # finished without announcing DONE: drop it
    if sig == signal_DONE:
        try:
            next(co)
        except StopIteration:
            pass
    else:
        coq.push(co)
    return True


def co_wait(coq: Queue, usec: int = 0) -> None:
    """Ping until every coroutine in the queue has signalled DONE; ``usec`` sleeps between pings
    (default busy poll, /root/reference/init.lua:185-192, README.md:65)."""
    while co_ping(coq):
        if usec > 0:
            time.sleep(usec * 1e-6)
