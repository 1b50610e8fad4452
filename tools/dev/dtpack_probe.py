"""Device pack / unpack throughput of mpit_dt_move_device for a few derived datatypes (CUDA events, warm),
and the target of the ncu capture in profiles/ (``ncu -k regex:dt_move ... python tools/dev/dtpack_probe.py``)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from mpit_b200 import msg as mpiT                                   # noqa: E402
from mpit_b200.msg import datatypes as dt                           # noqa: E402


def main():
    dev = torch.device("cuda:0")
    cases = {
        # name: (datatype, count, floats in the strided buffer)
        "rows_1KiB_of_4KiB": (mpiT.Type_commit(mpiT.Type_vector(65536, 256, 1024, mpiT.FLOAT)), 1, 65536 * 1024),
        "tile_512x512_of_4096x4096": (mpiT.Type_commit(mpiT.Type_create_subarray(
            [4096, 4096], [512, 512], [1024, 2048], mpiT.ORDER_C, mpiT.FLOAT)), 1, 4096 * 4096),
        "column_pairs_8B": (mpiT.Type_commit(mpiT.Type_create_resized(
            mpiT.Type_vector(4096, 2, 4096, mpiT.FLOAT), 0, 8)), 256, 4096 * 4096),
    }
    out = {}
    for name, (t, count, nfloats) in cases.items():
        src = torch.randn(nfloats, device=dev)
        T = dt.get(t)
        nbytes = count * T.size
        stream = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        flat = src.view(torch.uint8)
        for _ in range(3):
            dt._device_move(T, flat, stream, nbytes, True)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 20
        e0.record()
        for _ in range(reps):
            dt._device_move(T, flat, stream, nbytes, True)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / reps
        # reference point: torch's own strided copy of the same region where it can express it
        out[name] = {"packed_MiB": round(nbytes / 2 ** 20, 2), "us": round(us, 2),
                     "GBps_read_plus_write": round(2 * nbytes / us / 1e3, 1), "unit_bytes": T._dev_tables[dev][2]}
    # torch baseline for the tile case: slicing + contiguous()
    a = torch.randn(4096, 4096, device=dev)
    for _ in range(3):
        a[1024:1536, 2048:2560].contiguous()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        a[1024:1536, 2048:2560].contiguous()
    e1.record()
    torch.cuda.synchronize()
    out["torch_tile_contiguous_us"] = round(e0.elapsed_time(e1) * 1e3 / 20, 2)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
