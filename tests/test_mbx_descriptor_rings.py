"""The CPU half of the device transport of the message layer (csrc/kernels/mailbox.cu, mpit_b200/msg/device.py): the
descriptor rings are plain words in (shared) host memory, and matching, claiming and cancelling are CPU atomics on
them -- so MPI's matching rules can be checked here without a GPU: oldest first per source, non-overtaking (a message
that has not arrived hides everything behind it), tag-selective receives leaving other messages queued, ANY_SOURCE /
ANY_TAG, ring wrap-around, and the claim-versus-cancel race that exactly one side may win
(reference semantics: MPI_Recv / MPI_Iprobe / MPI_Cancel as wrapped in /root/reference/mpifuncs.c and driven by
/root/reference/init.lua:41-108)."""
import ctypes
import threading

import numpy as np
import pytest

from mpit_b200._native import kernels

K, NSRC = 4, 3


class Ring:
    """Descriptors of ONE receiver: [nsrc][K] of 128 bytes each, plus the receiver's bookkeeping (device.py)."""

    def __init__(self):
        self.k = kernels()
        self.dsz = int(self.k.mpit_mbx_desc_bytes())
        assert self.dsz == 128
        self.mem = np.zeros(NSRC * K * self.dsz, dtype=np.uint8)
        self.base = (ctypes.c_uint * NSRC)(*([1] * NSRC))          # message numbers are 1-based
        self.mask = (ctypes.c_ulonglong * NSRC)()
        self.sent = [0] * NSRC

    def desc(self, src, seq):
        return ctypes.c_void_p(self.mem.ctypes.data + (src * K + seq % K) * self.dsz)

    def send(self, src, tag, nbytes=4, publish=True):
        self.sent[src] += 1
        seq = self.sent[src]
        self.k.mpit_mbx_fill(self.desc(src, seq), tag, nbytes, 0, 0, None)
        if publish:
            self.k.mpit_mbx_publish_host(self.desc(src, seq), seq)
        return seq

    def scan(self, src=-1, tag=-1, start=0):
        out5 = (ctypes.c_int * 5)()
        nb = ctypes.c_longlong(0)
        hit = self.k.mpit_mbx_scan(ctypes.c_void_p(self.mem.ctypes.data), NSRC, K, self.base, self.mask, src, tag, start,
                                   out5, ctypes.byref(nb))
        return (out5[0], out5[1], out5[2], out5[4], nb.value) if hit else None      # src, seq, tag, state, bytes

    def consume(self, src, seq):
        j = seq - self.base[src]
        self.mask[src] |= 1 << j
        while self.mask[src] & 1:
            self.mask[src] >>= 1
            self.base[src] += 1
        self.k.mpit_mbx_set_done(self.desc(src, seq), seq)


def test_oldest_first_and_tag_selective_matching():
    r = Ring()
    r.send(1, tag=7)
    r.send(1, tag=9, nbytes=12)
    r.send(1, tag=7)
    assert r.scan(src=1, tag=7)[:3] == (1, 1, 7)                  # the older of the two tag-7 messages
    assert r.scan(src=1, tag=9) == (1, 2, 9, 1, 12)               # a tag-selective receive passes over message 1 ...
    r.consume(1, 2)
    assert r.scan(src=1, tag=-1)[:3] == (1, 1, 7)                 # ... which stays queued, still first
    r.consume(1, 1)
    assert r.scan(src=1, tag=9) is None                           # consumed messages are not matched again
    assert r.scan(src=1, tag=7)[:3] == (1, 3, 7)
    assert r.base[1] == 3 and r.mask[1] == 0                      # base advanced over both consumed messages
    assert r.scan(src=0) is None and r.scan(src=2) is None


def test_a_message_that_has_not_arrived_hides_what_is_behind_it():
    r = Ring()
    first = r.send(0, tag=1, publish=False)                       # descriptor filled, sequence word not yet visible
    r.send(0, tag=2)
    assert r.scan(src=0, tag=2) is None                           # non-overtaking: message 2 may not pass message 1
    assert r.scan(src=0) is None
    r.k.mpit_mbx_publish_host(r.desc(0, first), first)
    assert r.scan(src=0, tag=2)[:3] == (0, 2, 2)
    assert r.scan(src=0)[:3] == (0, 1, 1)


def test_any_source_round_robin_and_any_tag():
    r = Ring()
    r.send(0, tag=5)
    r.send(2, tag=6)
    assert r.scan(src=-1, tag=-1, start=0)[:3] == (0, 1, 5)
    assert r.scan(src=-1, tag=-1, start=1)[:3] == (2, 1, 6)       # the scan starts after the source served last
    assert r.scan(src=-1, tag=6, start=0)[:3] == (2, 1, 6)
    assert r.scan(src=-1, tag=8) is None


def test_ring_wraps_and_a_slot_is_reused_only_after_done():
    r = Ring()
    for i in range(3 * K):                                        # three times around the ring, one message in flight
        seq = r.send(1, tag=i)
        d = r.desc(1, seq)
        assert r.k.mpit_mbx_word(d, 0) == seq and r.k.mpit_mbx_word(d, 2) != seq       # posted, not yet done
        hit = r.scan(src=1, tag=i)
        assert hit[:3] == (1, seq, i)
        assert r.k.mpit_mbx_cas_state(d, 1, 2) == 1               # claimed
        r.consume(1, seq)
        assert r.k.mpit_mbx_word(d, 2) == seq                     # the sender may reuse slot seq % K from here on
    assert r.base[1] == 3 * K + 1
    # a full ring: K messages queued, every one found in order
    seqs = [r.send(2, tag=100 + j) for j in range(K)]
    for j, s in enumerate(seqs):
        assert r.scan(src=2)[:3] == (2, s, 100 + j)
        r.consume(2, s)


def test_cancelled_messages_are_reported_for_clean_up_and_skipped():
    r = Ring()
    s1 = r.send(0, tag=3)
    s2 = r.send(0, tag=3)
    assert r.k.mpit_mbx_cas_state(r.desc(0, s1), 1, 3) == 1       # MPI_Cancel of the first send wins
    hit = r.scan(src=0, tag=3)
    assert hit[1] == s1 and hit[3] == 3                           # reported with state 3 so the receiver retires the slot
    r.consume(0, s1)
    assert r.scan(src=0, tag=3)[:3] == (0, s2, 3)
    assert r.k.mpit_mbx_cas_state(r.desc(0, s2), 1, 2) == 1
    assert r.k.mpit_mbx_cas_state(r.desc(0, s2), 1, 3) == 2       # too late to cancel: the receive owns it


@pytest.mark.parametrize("rounds", [300])
def test_claim_and_cancel_race_has_exactly_one_winner(rounds):
    r = Ring()
    wins = {"recv": 0, "cancel": 0}
    for i in range(rounds):
        seq = r.send(1, tag=1)
        d = r.desc(1, seq)
        res = {}
        go = threading.Barrier(2)

        def claim():
            go.wait()
            res["recv"] = r.k.mpit_mbx_cas_state(d, 1, 2)

        def cancel():
            go.wait()
            res["cancel"] = r.k.mpit_mbx_cas_state(d, 1, 3)

        ts = [threading.Thread(target=claim), threading.Thread(target=cancel)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        won = [name for name, old in res.items() if old == 1]
        assert len(won) == 1, res                                 # the CAS returns the old state: 1 = this side won
        wins[won[0]] += 1
        assert r.k.mpit_mbx_word(d, 1) == (2 if won[0] == "recv" else 3)
        r.consume(1, seq)
    assert wins["recv"] + wins["cancel"] == rounds
