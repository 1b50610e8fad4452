"""The BiCNN helper functions (utils.lua restated) and hostfile placement / remote start of the job launcher."""
import io
import os
import stat
import sys

import torch

from mpit_b200.launch import run
from mpit_b200.utils import helpers as H

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_tensor_descriptions_and_sizes():
    assert H.tsize([3, 28, 28]) == "3x28x28" and H.tsize(None) == "(no tensor)"
    t = torch.tensor([[1, 2], [3, 6]], dtype=torch.int64)
    assert H.tdims(t) == "2x2"
    assert H.tinfo(t) == "(2x2 min: 1.000000 mean: 3.000000 max: 6.000000 type: torch.int64)"
    assert H.tinfo(None) == "(no tensor)" and H.tinfo(torch.zeros(0)) == "(empty tensor)" and H.tinfo(5) == "5"
    out = io.StringIO()
    assert H.conv_size(28, 28, 5, 5, 1, 1, 3, 64, out) == (24, 24)
    assert H.pool_size(24, 24, 2, 2, 2, 2, out) == (12, 12)
    assert H.pad_size(12, 12, 2, 1, out) == (16, 14)
    H.pool_size(5, 5, 2, 2, 2, 2, out)
    assert "convolution\t3x28x28\t* 5x5 (stride 1x1)\t-> 64x24x24" in out.getvalue()
    assert "warning: expected integers in 2.5x2.5" in out.getvalue()
    buf = io.StringIO()
    H.tflat(torch.tensor([1.0, 2.5]), buf)
    assert buf.getvalue() == "1.0 2.5 \n"
    assert len(H.tstamp(0)) == 15 and H.tstamp(0)[8] == "."
    assert H.lnarrow(list("abcdef"), 2, 2) == ["b", "c", "d"]          # o .. o+s inclusive, 1-based


def test_dict_helpers_and_clones():
    acc = {"feval": 1.0, "note": "x"}
    H.table_acc(acc, {"feval": 2.5, "sync": 1.0, "note": "y"})
    assert acc == {"feval": 3.5, "sync": 1.0, "note": "y"}
    assert H.sorted_keys({"b": 1, "a": 2}) == ["a", "b"] and H.count_list({"a": 1, "b": 2}) == 2
    assert H.table_string({"sync": 0.5, "feval": 2.0, "sub": {"x": 1.0}}, 1000.0, " %s %.0fms") == \
        " feval 2000ms sub { x 1000ms } sync 500ms"
    src = {"w": torch.ones(2, 2, dtype=torch.float64), "n": 3, "inner": {"idx": torch.tensor([1, 2])}}
    c = H.clone_table(src, torch.float32)
    assert c["w"].dtype == torch.float32 and c["inner"]["idx"].dtype == torch.int64 and c["n"] == 3
    c["w"].zero_()
    assert float(src["w"].sum()) == 4.0
    net = torch.nn.Sequential(torch.nn.Linear(4, 3), torch.nn.Tanh())
    net[0].scratch = torch.zeros(5)
    twin = H.clone_module(net, torch.float64, exclude=("scratch", "bias"))
    assert twin[0].weight.dtype == torch.float64 and twin[0].bias is None and not hasattr(twin[0], "scratch")
    assert torch.equal(twin[0].weight.float(), net[0].weight) and net[0].bias is not None

    class Conf:
        def __init__(self):
            self.rank, self.buf = 3, torch.arange(3)

    o = Conf()
    d = H.clone_object(o)
    d.buf += 1
    assert isinstance(d, Conf) and d.rank == 3 and o.buf.tolist() == [0, 1, 2]


def test_cap_weight_norms():
    conv = torch.ones(4, 3, 2, 2)          # per-feature norm sqrt(12)
    conv[0] *= 0.1
    lin = torch.tensor([[3.0, 4.0], [0.3, 0.4]])
    bias = torch.tensor([-5.0, 0.2, 5.0])
    H.cap_weight_norms([conv, lin, bias], cap_conv=1.0, cap_linear=1.0, cap_bias=1.0)
    norms = conv.flatten(1).norm(dim=1)
    assert torch.allclose(norms[1:], torch.ones(3)) and abs(float(norms[0]) - 0.1 * 12 ** 0.5) < 1e-6
    assert torch.allclose(lin, torch.tensor([[0.6, 0.8], [0.3, 0.4]])) and torch.equal(bias, torch.tensor([-1.0, 0.2, 1.0]))
    assert H.tdinfo is H.cap_weight_norms


def test_hostfile_parsing_and_placement():
    hosts = run.parse_hostfile("bluejgpu1:16\nbluejgpu2:16   # second node\n\nnode3 slots=2 max_slots=4\nsolo\n")
    assert hosts == [("bluejgpu1", 16), ("bluejgpu2", 16), ("node3", 2), ("solo", 1)]
    where = run.place_ranks(5, [("a", 2), ("b", 2)])
    assert where == ["a", "a", "b", "b", "a"]                           # by slot, then wrap (oversubscribed)


def test_ring_over_a_hostfile_with_a_remote_shell(tmp_path):
    """Two ranks 'here', two on a host reached through a stand-in remote shell that runs the command locally:
    exercises the environment forwarding and the quoting of the remote command line."""
    fake = tmp_path / "fake_rsh"
    fake.write_text("#!/bin/sh\n# usage: fake_rsh HOST COMMAND\necho \"$1\" >> " + str(tmp_path / "hosts_seen") +
                    "\nshift\nexec sh -c \"$1\"\n")
    fake.chmod(fake.stat().st_mode | stat.S_IEXEC)
    hf = tmp_path / "hostfile"
    hf.write_text("localhost:2\nnode-b:2\n")
    code = (f"import sys; sys.path.insert(0, {ROOT!r}); sys.argv=['mp', 'ring']; import runpy; "
            f"runpy.run_path({os.path.join(ROOT, 'tests', 'mp_cases.py')!r}, run_name='__main__')")
    cwd = os.getcwd()
    os.chdir(ROOT)
    try:
        rc = run.launch(4, ["-c", code], hosts=run.parse_hostfile(hf.read_text()), rsh=str(fake),
                        master_addr="127.0.0.1", timeout=120)
    finally:
        os.chdir(cwd)
    assert rc == 0
    assert (tmp_path / "hosts_seen").read_text().split() == ["node-b", "node-b"]


def test_launcher_exports_environment_like_mpirun_x(tmp_path, monkeypatch):
    script = tmp_path / "envcheck.py"
    script.write_text("import os, sys\n"
                      "ok = os.environ.get('MEGS') == '3' and os.environ.get('FOO_BAR') == 'baz'\n"
                      "sys.exit(0 if ok else 7)\n")
    monkeypatch.setenv("FOO_BAR", "baz")
    assert run.main(["-np", "2", "-x", "MEGS=3", "-x", "FOO_BAR", "--oversubscribe", str(script)]) == 0
    monkeypatch.delenv("FOO_BAR")
    assert run.main(["-np", "1", "-x", "MEGS=3", str(script)]) == 7
