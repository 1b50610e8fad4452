"""End-to-end CPU launchers over the host message layer (BASELINE.json config 1 and the reference's
mlaunch process layout)."""
import os
import sys

from mpit_b200.launch.run import launch


def test_claunch_mlp_easgd_one_server_two_clients(tmp_path):
    rc = launch(3, ["mpit_b200.launch.claunch", "--epochs", "1", "--trsize", "256", "--tesize", "128", "--mb", "64",
                    "--hidden", "64", "--save", str(tmp_path)], module=True, timeout=300)
    assert rc == 0
    logs = sorted(os.listdir(tmp_path))
    assert "train1.log" in logs and "test2.log" in logs
    rows = open(tmp_path / "train1.log").read().strip().split("\n")
    assert rows[0].startswith("% train accuracy") and len(rows) == 2


def test_mlaunch_cpu_parity_roles_mnist(tmp_path):
    rc = launch(4, ["mpit_b200.launch.mlaunch", "--device", "cpu", "--model", "mnist", "--trsize", "256", "--tesize",
                    "128", "--mb", "64", "--save", str(tmp_path)], module=True, timeout=300)
    assert rc == 0
    assert os.path.exists(tmp_path / "train1.log") and os.path.exists(tmp_path / "train3.log")


def test_failing_rank_aborts_job():
    rc = launch(2, ["-c", "import os,sys,time; sys.exit(3) if os.environ['MPIT_RANK']=='1' else time.sleep(30)"],
                timeout=20)
    assert rc == 3


def test_ptest_cpu_bandwidth():
    assert launch(4, ["mpit_b200.launch.ptest", "--device", "cpu", "--ssize", "65536", "--rounds", "3"], module=True,
                  timeout=120) == 0


def test_testreduceall_cpu():
    assert launch(4, ["mpit_b200.launch.testreduceall", "--ssize", "65536", "--megs", "0.25"], module=True,
                  timeout=120) == 0


def test_ptest2_stragglers():
    assert launch(4, ["mpit_b200.launch.ptest2", "--ssize", "4096", "--rounds", "2", "--skew-size", "16"], module=True,
                  timeout=120) == 0


def test_ring_smoke_script():
    """/root/reference/test.lua: one float around the ring."""
    assert launch(3, ["mpit_b200.launch.test"], module=True, timeout=60) == 0
