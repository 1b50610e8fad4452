"""Host message layer, coroutine IO and parameter-server protocol: N local processes (the reference's
way of testing, ``mpirun -np N`` on one host, /root/reference/README.md:30,54), with assertions."""
import os
import sys

import pytest

from mpit_b200.launch.run import launch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _run(case, nproc, env=None, timeout=120):
    rc = launch(nproc, ["-c", f"import sys; sys.path.insert(0, {ROOT!r}); sys.argv=['mp', {case!r}]; "
                              f"import runpy; runpy.run_path({os.path.join(HERE, 'mp_cases.py')!r}, run_name='__main__')"],
                env_extra=env, timeout=timeout)
    assert rc == 0, f"case {case} failed with exit code {rc}"


@pytest.mark.parametrize("nproc", [2, 5])
def test_ring_send_recv(nproc):
    _run("ring", nproc)


def test_requests_of_a_dead_peer_complete_with_an_error():
    _run("peer_death", 3, timeout=60)


def test_symmetric_heap_rendezvous_over_host_layer():
    _run("symm_rendezvous", 3)


@pytest.mark.parametrize("nproc", [2, 4])
def test_allreduce(nproc):
    _run("allreduce", nproc)


def test_iallreduce_test_wait_state_machine():
    _run("iallreduce", 3)


def test_tag_matching_probe_truncation():
    _run("matching", 2)


def test_cancel_semantics():
    _run("cancel", 2)


def test_communicators_split_dup():
    _run("comm", 4)


def test_collectives():
    _run("collectives", 4)


def test_ps_protocol_partition_seed_stop():
    _run("ps_protocol", 4)


def test_ps_protocol_uneven():
    _run("ps_protocol", 6)


def test_ps_eamsgd_converges():
    _run("ps_eamsgd", 4)


@pytest.mark.parametrize("su", ["1", "3"])
def test_ps_downpour_converges(su):
    _run("ps_downpour", 4, env={"SU": su})


def test_ps_server_side_adam_matches_oracle():
    _run("ps_server_adam", 2)


def test_ps_single_mode_writer_reader():
    _run("ps_single_mode", 3)


def test_ptest_bandwidth():
    _run("ptest", 4)


@pytest.mark.parametrize("case,nproc", [("x_vector_collectives", 4), ("x_requests", 3), ("x_groups_ops", 4),
                                        ("x_window", 3)])
def test_wider_mpi_surface(case, nproc):
    """Send modes, persistent requests, vector collectives, groups / user ops, fence windows (msg/extras.py)."""
    _run(case, nproc)


@pytest.mark.parametrize("case,nproc", [("c_derived_messages", 4), ("c_topology", 6), ("c_file_io", 4),
                                        ("c_pscw_and_attrs", 4), ("c_intercomm", 4),
                                        ("c_maxloc_and_wide_types", 4), ("c_sharded_checkpoint", 3),
                                        ("c_ring_allreduce", 5), ("c_ring_allreduce", 3),
                                        ("c_big_collectives", 4)])
def test_mpi_chapters_multi_process(case, nproc):
    """Derived datatypes on the wire, topologies, MPI-IO with views and shared pointers, post/start/complete/wait,
    attributes across Comm_dup, inter-communicators / ports / Comm_join (msg/datatypes.py, topo.py, fileio.py, rma.py,
    attrs.py, dynproc.py)."""
    _run(case, nproc)


@pytest.mark.parametrize("cma", ["1", "0"])
def test_large_message_rendezvous(cma):
    """Single-copy rendezvous for large Isends on one host (csrc/msg/hostcomm.cpp) and the socket path it falls back
    to: same observable behaviour except that the rendezvous buffers nothing at the receiver."""
    _run("c_rendezvous", 3, env={"MPIT_CMA": cma})
