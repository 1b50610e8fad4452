"""Which ingredient keeps the native server thread's drain from running next to a spinning push kernel?
One process, one GPU; prints the longest flow-control wait for each variant."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch

from mpit_b200.parallel.ps_device import AsyncDevicePS
from mpit_b200.parallel.symm import LocalPeerHeap, SymmetricHeap
from mpit_b200.parallel.center import ShardedCenter


def run(name, make, matmuls, sync, n=400003, rounds=6, warm=False):
    ps = make(n)
    ps.start_server(interval_us=30)
    g = torch.ones(n, device="cuda")
    x = torch.empty(n, device="cuda")
    a = torch.randn(2048, 2048, device="cuda")
    if warm:                       # every kernel of the exchange has run once before anything spins
        ps.push(gflat=g); torch.cuda.synchronize(); time.sleep(0.05)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for t in range(rounds):
        for _ in range(matmuls):
            a = torch.tanh(a @ a * 1e-3)
        ps.push(gflat=g)
        ps.pull(x)
    if sync == "device":
        torch.cuda.synchronize()
    else:
        torch.cuda.current_stream().synchronize()
    dt = time.perf_counter() - t0
    time.sleep(0.1)
    drains = ps.stop_server()
    print(f"{name:45s} {dt * 1e3:9.1f} ms  max wait {ps.max_wait_us:10.1f} us  drains {drains}  applied {ps.applied}", flush=True)


def local(n):
    return AsyncDevicePS.local_group(LocalPeerHeap(1, devices=[0]), n, "add", {})[0]


def symm(mode):
    def make(n):
        heap = SymmetricHeap(torch.device("cuda", 0), mode=mode)
        return AsyncDevicePS(heap, n, "add", {}, ShardedCenter.create(heap, n))
    return make


if __name__ == "__main__":
    torch.cuda.set_device(0)
    run("local heap, no matmul, stream sync", local, 0, "stream")
    run("local heap, no matmul, stream sync (2nd)", local, 0, "stream")
    run("local heap, matmul, stream sync", local, 4, "stream")
    run("local heap, matmul, device sync", local, 4, "device")
    run("ipc heap, matmul, device sync", symm("ipc"), 4, "device")
    run("vmm heap, matmul, device sync", symm("vmm"), 4, "device")
    run("vmm heap, no matmul, stream sync", symm("vmm"), 0, "stream")
    run("vmm heap, matmul, device sync, warmed", symm("vmm"), 4, "device", warm=True)
