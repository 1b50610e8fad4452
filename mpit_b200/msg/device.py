"""Device backend of the message layer: ``mpiT.Send / Recv / Isend / Irecv / Iprobe / Test / Wait / Cancel`` on CUDA
tensors move the payload GPU to GPU over NVLink instead of bouncing through pinned host memory and a socket.

Reference behaviour: any Torch Storage -- a ``CudaStorage`` too -- is handed to MPI by pointer
(``/root/reference/lua-mpi.h:70-78``), and ``aio_send`` / ``aio_recv`` poll ``Iprobe`` / ``Test`` from coroutines
(``/root/reference/init.lua:41-108``) with a working ``Cancel`` on shutdown (``:52-59``).

How a message travels (``csrc/kernels/mailbox.cu``):

* every ordered pair (src, dst) owns a ring of ``K`` descriptors ``{seq, tag, bytes, kind, state, done}`` in a POSIX
  shared-memory segment that all processes of the box map and register with CUDA;
* *eager* (<= ``slot_bytes``): the sender's kernel stores the payload into slot ``seq % K`` of the receiver's
  symmetric memory, then its GPU publishes ``seq`` (``st.release.sys``) in stream order;
* *rendezvous* (larger): nothing is copied on the send side -- the descriptor carries a CUDA-IPC handle of the
  sender's allocation and the receiver's kernel pulls the bytes straight into the user buffer; the send completes when
  the receiver has acknowledged (``done``), which is MPI's synchronous-mode completion for large messages;
* zero-byte messages are a descriptor write by the CPU, no GPU work at all;
* matching is done by the receiver's CPU on the shared words: oldest message first per source (non-overtaking),
  ``ANY_SOURCE`` / ``ANY_TAG``, messages skipped by a tag-selective receive stay queued; ``Iprobe`` / ``Test`` are
  memory reads, not kernel launches; ``Cancel`` of a send is a compare-and-swap against the claiming receive;
* flow control: slot ``seq % K`` is reused once the receiver has set ``done`` for message ``seq - K``; sends that find
  their slot busy wait in a per-destination queue that every call into the transport progresses.

Sender and receiver have to agree on the buffer kind: a CUDA tensor travels on this transport, a host buffer on the
host transport (zero-byte control messages of the parameter-server protocol use ``None`` buffers and stay on the
host layer on both sides).  Only ``COMM_WORLD`` is served here; other communicators fall back to host staging.
"""
from __future__ import annotations

import ctypes
import os
import time
from collections import OrderedDict, deque
from typing import Deque, Dict, List, Optional

import torch

from .._native import kernels, NativeError

ANY = -1
_LIVE = -2          # marker stored in Request._arr for a live device request (REQUEST_NULL is the host layer's)


class DevRequest:
    __slots__ = ("kind", "peer", "tag", "nbytes", "tensor", "ptr", "done", "cancelled", "error", "src", "rtag",
                 "count_bytes", "event", "desc", "seq", "posted", "wait_ack", "keep")

    def __init__(self, kind: str, peer: int, tag: int, nbytes: int, tensor):
        self.kind, self.peer, self.tag, self.nbytes, self.tensor = kind, peer, tag, nbytes, tensor
        self.ptr = tensor.data_ptr() if tensor is not None else 0
        self.done = self.cancelled = False
        self.error = 0
        self.src, self.rtag, self.count_bytes = peer, tag, 0
        self.event: Optional[torch.cuda.Event] = None
        self.desc = 0           # host address of the descriptor this request is bound to
        self.seq = 0
        self.posted = False     # send: descriptor filled and published
        self.wait_ack = False   # send, rendezvous: complete when the receiver has set `done`
        self.keep = None


class DeviceTransport:
    def __init__(self, heap, rank: int, size: int, job: str, K: int = 8, slot_bytes: int = 64 << 10):
        self.k = kernels()
        self.heap, self.rank, self.size, self.K, self.slot_bytes = heap, rank, size, K, slot_bytes
        self.dev = heap.device
        self.dsz = int(self.k.mpit_mbx_desc_bytes())
        # payload slots in MY symmetric memory: [src][K][slot_bytes]
        self.slots = heap.alloc(size * K * slot_bytes // 4)
        # descriptor rings of the whole job in one shared segment: [dst][src][K]
        self.shm_name = f"/mpit_mbx_{job}"
        self.shm_bytes = size * size * K * self.dsz
        from ..parallel.symm import _barrier
        hp, dp = ctypes.c_void_p(0), ctypes.c_void_p(0)
        if rank == 0:
            self._map(1, hp, dp)
        _barrier(heap.group)
        if rank != 0:
            self._map(0, hp, dp)
        _barrier(heap.group)
        self.host_base, self.dev_base = hp.value, dp.value
        self.next_seq = [1] * size                       # per destination: next message number
        self.sendq: List[Deque[DevRequest]] = [deque() for _ in range(size)]
        self.inflight_send: List[DevRequest] = []
        self.base = (ctypes.c_uint * size)(*([1] * size))            # per source: lowest unconsumed message
        self.mask = (ctypes.c_ulonglong * size)(*([0] * size))      # consumed out of order
        self.recvq: List[DevRequest] = []
        self.inflight_recv: List[DevRequest] = []
        self._rr = 0
        self._ipc_out: Dict[int, tuple] = {}             # allocation base -> (handle bytes)
        self._ipc_in: "OrderedDict[bytes, int]" = OrderedDict()      # handle -> mapped base address
        self.stream = torch.cuda.Stream(device=self.dev)
        self.stats = {"eager": 0, "rendezvous": 0, "zero": 0, "probes": 0}

    def _map(self, create, hp, dp):
        rc = self.k.mpit_shm_map(self.shm_name.encode(), self.shm_bytes, create, ctypes.byref(hp), ctypes.byref(dp))
        if rc != 0:
            raise NativeError(f"cannot map the descriptor segment {self.shm_name}: code {rc}")

    def close(self):
        if getattr(self, "host_base", None):
            self.k.mpit_shm_unmap(self.shm_name.encode(), ctypes.c_void_p(self.host_base), self.shm_bytes,
                                  int(self.rank == 0))
            self.host_base = None

    # -- addressing -----------------------------------------------------------------------------------------------
    def _desc(self, dst: int, src: int, seq: int, dev: bool = False) -> int:
        off = ((dst * self.size + src) * self.K + seq % self.K) * self.dsz
        return (self.dev_base if dev else self.host_base) + off

    def _slot_addr(self, dst: int, src: int, seq: int) -> int:
        return self.slots.ptrs[dst] + (src * self.K + seq % self.K) * self.slot_bytes

    def _word(self, desc: int, which: int) -> int:
        return self.k.mpit_mbx_word(ctypes.c_void_p(desc), which) & 0xFFFFFFFF

    # -- send side --------------------------------------------------------------------------------------------------
    def isend(self, tensor, nbytes: int, dest: int, tag: int) -> DevRequest:
        if tag < 0:
            raise ValueError("a send needs a non-negative tag")
        req = DevRequest("send", dest, tag, nbytes, tensor)
        if nbytes > 0:
            # the payload is ready once the work already queued on the caller's stream has run
            req.event = torch.cuda.Event()
            req.event.record(torch.cuda.current_stream(self.dev))
        self.sendq[dest].append(req)
        self.progress()
        return req

    def _post(self, req: DevRequest) -> bool:
        dest = req.peer
        seq = self.next_seq[dest]
        desc = self._desc(dest, self.rank, seq)
        if seq > self.K and self._word(desc, 2) != seq - self.K:
            return False                                  # the slot's previous message has not been consumed yet
        k, st = self.k, ctypes.c_void_p(self.stream.cuda_stream)
        req.desc, req.seq = desc, seq
        if req.nbytes == 0:
            k.mpit_mbx_fill(ctypes.c_void_p(desc), req.tag, 0, 0, 0, None)
            k.mpit_mbx_publish_host(ctypes.c_void_p(desc), seq)
            self.stats["zero"] += 1
        else:
            self.stream.wait_event(req.event)
            if req.nbytes <= self.slot_bytes:
                k.mpit_mbx_fill(ctypes.c_void_p(desc), req.tag, req.nbytes, 0, 0, None)
                rc = k.mpit_mbx_copy(ctypes.c_void_p(self._slot_addr(dest, self.rank, seq)), ctypes.c_void_p(req.ptr),
                                     req.nbytes, st)
                self.stats["eager"] += 1
            else:
                handle, off = self._export(req.ptr)
                k.mpit_mbx_fill(ctypes.c_void_p(desc), req.tag, req.nbytes, 1, off, handle)
                req.wait_ack = True
                rc = 0
                self.stats["rendezvous"] += 1
            if rc == 0:
                rc = k.mpit_mbx_publish(ctypes.c_void_p(self._desc(dest, self.rank, seq, dev=True)), seq, st)
            if rc != 0:
                raise NativeError(f"device transport: posting a message failed with code {rc}")
            req.event = torch.cuda.Event()
            req.event.record(self.stream)
        self.next_seq[dest] = seq + 1
        req.posted = True
        self.inflight_send.append(req)
        return True

    def _export(self, ptr: int):
        handle = (ctypes.c_char * 64)()
        off = ctypes.c_ulonglong(0)
        rc = self.k.mpit_ipc_export_range(ptr, handle, ctypes.byref(off))
        if rc != 0:
            raise NativeError(f"device transport: cannot export the send buffer through CUDA IPC (code {rc}); "
                              f"buffers larger than {self.slot_bytes} bytes must come from cudaMalloc memory")
        return handle, off.value

    # -- receive side -------------------------------------------------------------------------------------------------
    def irecv(self, tensor, nbytes: int, source: int, tag: int) -> DevRequest:
        req = DevRequest("recv", source, tag, nbytes, tensor)
        self.recvq.append(req)
        self.progress()
        return req

    def _scan(self, source: int, tag: int):
        out5 = (ctypes.c_int * 5)()
        nb = ctypes.c_longlong(0)
        ring = self.host_base + self.rank * self.size * self.K * self.dsz
        found = self.k.mpit_mbx_scan(ctypes.c_void_p(ring), self.size, self.K, self.base, self.mask, source, tag,
                                     self._rr, out5, ctypes.byref(nb))
        if not found:
            return None
        return out5[0], out5[1] & 0xFFFFFFFF, out5[2], out5[3], out5[4], nb.value

    def _consume(self, src: int, seq: int) -> None:
        j = seq - self.base[src]
        self.mask[src] |= (1 << j)
        while self.mask[src] & 1:                         # advance past everything consumed in order
            self.mask[src] >>= 1
            self.base[src] += 1

    def _match_one(self, req: DevRequest) -> bool:
        """Try to bind the receive to a message; True when it was bound (or failed for good)."""
        while True:
            hit = self._scan(req.peer, req.tag)
            if hit is None:
                return False
            src, seq, mtag, kind, state, nbytes = hit
            desc = self._desc(self.rank, src, seq)
            if state == 3 or self.k.mpit_mbx_cas_state(ctypes.c_void_p(desc), 1, 2) != 1:
                # withdrawn by its sender (MPI_Cancel): retire the slot and look again
                self._consume(src, seq)
                self.k.mpit_mbx_set_done(ctypes.c_void_p(desc), seq)
                continue
            self._consume(src, seq)
            self._rr = (src + 1) % self.size
            req.src, req.rtag, req.count_bytes, req.desc, req.seq = src, mtag, nbytes, desc, seq
            if nbytes > req.nbytes:
                req.error = 15                            # MPI_ERR_TRUNCATE
                req.count_bytes = req.nbytes
                self.k.mpit_mbx_set_done(ctypes.c_void_p(desc), seq)
                req.done = True
                return True
            if nbytes == 0:
                self.k.mpit_mbx_set_done(ctypes.c_void_p(desc), seq)
                req.done = True
                return True
            if kind == 0:
                src_addr = self._slot_addr(self.rank, src, seq)
            else:
                src_addr = self._import(desc, src)
            st = ctypes.c_void_p(self.stream.cuda_stream)
            rc = self.k.mpit_mbx_copy(ctypes.c_void_p(req.ptr), ctypes.c_void_p(src_addr), nbytes, st)
            if rc != 0:
                raise NativeError(f"device transport: receive copy failed with code {rc}")
            req.event = torch.cuda.Event()
            req.event.record(self.stream)
            self.inflight_recv.append(req)
            return True

    def _import(self, desc: int, src: int) -> int:
        handle = (ctypes.c_char * 64)()
        off = ctypes.c_ulonglong(0)
        self.k.mpit_mbx_desc_info(ctypes.c_void_p(desc), ctypes.byref(off), handle)
        if src == self.rank:
            raise NativeError("device transport: a rendezvous message to oneself cannot be mapped; use a smaller message")
        key = bytes(handle)
        base = self._ipc_in.get(key)
        if base is None:
            p = ctypes.c_ulonglong(0)
            rc = self.k.mpit_symm_ipc_import(handle, ctypes.byref(p))
            if rc != 0:
                raise NativeError(f"device transport: cudaIpcOpenMemHandle failed with code {rc}")
            base = p.value
            self._ipc_in[key] = base
            if len(self._ipc_in) > 64:                    # bound the number of foreign mappings kept open
                _, old = self._ipc_in.popitem(last=False)
                torch.cuda.synchronize(self.dev)
                self.k.mpit_symm_ipc_close(old)
        else:
            self._ipc_in.move_to_end(key)
        return base + off.value

    # -- progress -----------------------------------------------------------------------------------------------------
    def progress(self) -> None:
        for q in self.sendq:
            while q and not q[0].cancelled and self._post(q[0]):
                q.popleft()
            while q and q[0].cancelled:
                q.popleft()
        if self.inflight_send:
            still = []
            for r in self.inflight_send:
                if r.done:
                    continue
                if r.wait_ack or r.nbytes == 0:
                    # rendezvous: the receiver has pulled the data; zero-byte: fire and forget
                    if r.nbytes == 0 or self._word(r.desc, 2) == r.seq or self._word(r.desc, 1) == 3:
                        r.done = True
                        continue
                elif r.event.query():
                    r.done = True
                    continue
                still.append(r)
            self.inflight_send = still
        if self.recvq:
            left = []
            for r in self.recvq:                          # posted order: earlier receives match first
                if r.cancelled:
                    continue
                # a message that arrives between two scans of this pass must go to the EARLIER receive it matches:
                # a receive is not tried while an earlier unmatched one could take the same messages
                blocked = any((e.peer < 0 or r.peer < 0 or e.peer == r.peer) and (e.tag < 0 or r.tag < 0 or e.tag == r.tag)
                              for e in left)
                if blocked or not self._match_one(r):
                    left.append(r)
            self.recvq = left
        if self.inflight_recv:
            still = []
            for r in self.inflight_recv:
                if r.event.query():
                    self.k.mpit_mbx_set_done(ctypes.c_void_p(r.desc), r.seq)     # slot / sender buffer reusable
                    r.done = True
                else:
                    still.append(r)
            self.inflight_recv = still

    def test(self, req: DevRequest) -> bool:
        if not req.done and not req.cancelled:
            self.progress()
        return req.done or req.cancelled

    def wait(self, req: DevRequest, timeout_s: float = 120.0) -> None:
        t0 = time.time()
        spins = 0
        while not self.test(req):
            spins += 1
            if spins > 200:
                time.sleep(0)
            if time.time() - t0 > timeout_s:
                raise TimeoutError("device transport: a request did not complete (no matching message or receive)")

    def cancel(self, req: DevRequest) -> None:
        """MPI_Cancel: a receive that has not matched is withdrawn; a send is withdrawn if no receive has claimed it
        (compare-and-swap on the descriptor's state), otherwise it completes normally."""
        if req.done or req.cancelled:
            return
        if req.kind == "recv":
            if req in self.recvq:
                self.recvq.remove(req)
                req.cancelled = True
            return
        if not req.posted:
            req.cancelled = True
            return
        if self.k.mpit_mbx_cas_state(ctypes.c_void_p(req.desc), 1, 3) == 1:
            req.cancelled = True

    def iprobe(self, source: int, tag: int):
        """(flag, source, tag, bytes) of the oldest matching message that no receive has claimed."""
        self.stats["probes"] += 1
        self.progress()
        while True:
            hit = self._scan(source, tag)
            if hit is None:
                return False, 0, 0, 0
            src, seq, mtag, kind, state, nbytes = hit
            if state == 3:
                self._consume(src, seq)
                self.k.mpit_mbx_set_done(ctypes.c_void_p(self._desc(self.rank, src, seq)), seq)
                continue
            return True, src, mtag, nbytes


_transport: Optional[DeviceTransport] = None


def enable(heap=None, K: int = 8, slot_bytes: int = 64 << 10) -> DeviceTransport:
    """Collective over the job: build the device transport (symmetric payload slots + shared descriptor segment).
    Call after ``mpiT.Init`` (or after ``torch.distributed.init_process_group``) on every rank."""
    global _transport
    if _transport is not None:
        return _transport
    from ..parallel.symm import SymmetricHeap, _dist_info
    if heap is None:
        heap = SymmetricHeap(torch.device("cuda", torch.cuda.current_device()))
    rank, size = _dist_info(heap.group)
    job = os.environ.get("MPIT_PORT", os.environ.get("MASTER_PORT", "0")) + "_" + str(os.getuid())
    _transport = DeviceTransport(heap, rank, size, job, K, slot_bytes)
    return _transport


def disable() -> None:
    global _transport
    if _transport is not None:
        torch.cuda.synchronize(_transport.dev)
        _transport.close()
        _transport = None


def transport() -> Optional[DeviceTransport]:
    return _transport


# ---- collectives on CUDA tensors ---------------------------------------------------------------------------------
_coll = {"obj": None, "cap": 0}


def _collectives(count: int):
    """Device collectives sized for `count` floats; (re)allocated collectively -- every rank calls a collective with
    the same count, so every rank takes the same branch."""
    tr = _transport
    if tr is None:
        return None
    if _coll["obj"] is None or _coll["cap"] < count:
        from ..parallel.collectives import DeviceCollectives
        cap = max(1 << 16, 1 << (count - 1).bit_length())
        _coll["obj"], _coll["cap"] = DeviceCollectives(tr.heap, cap), cap
    return _coll["obj"]


def device_allreduce(sbuf, rbuf, count: int, comm) -> bool:
    """MPI_Allreduce(SUM, FLOAT) of CUDA tensors over COMM_WORLD on the NVLink kernels (test/testreduceall.lua:28)."""
    from . import COMM_WORLD
    if _transport is None or comm != COMM_WORLD:
        return False
    c = _collectives(count)
    out = c.allreduce(sbuf.view(-1)[:count])
    rbuf.view(-1)[:count].copy_(out)
    return True


def device_bcast(buf, count: int, root: int, comm) -> bool:
    from . import COMM_WORLD
    if _transport is None or comm != COMM_WORLD:
        return False
    c = _collectives(count)
    if _transport.rank == root:
        c.out.local[:count].copy_(buf.view(-1)[:count])
    c.bcast(count, root)
    if _transport.rank != root:
        buf.view(-1)[:count].copy_(c.out.local[:count])
    return True
