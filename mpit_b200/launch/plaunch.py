"""plaunch: BiCNN answer selection with the parameter server and the parallel / single optimizers
(reference /root/reference/BiCNN/plaunch.lua: 58 ``torch.CmdLine`` options :7-70, roles ``ps`` / ``pt`` /
``pe`` from ``masterFreq`` / ``testerfirst`` / ``testerlast`` :124-163, tester table :166-177, per-rank seeds
:114-115).

    python -m mpit_b200.launch.run -np 3 -m mpit_b200.launch.plaunch -optimization adam -testerlast -epoch 1

Ranks above ``-maxrank`` exit immediately (the reference makes them spin forever, plaunch.lua:90-96).
Without corpus files a synthetic corpus of the same structure is used.
"""
from __future__ import annotations

import argparse
import os
import random
import sys

import torch

OPTIONS = [   # (name, default, help) -- plaunch.lua:10-69, same names, same defaults
    ("threads", 1, "number of threads"),
    ("optimization", "downpour", "sgd | downpour | eamsgd | adam (adamax) | adamsingle (adamaxsingle) | rmsprop | "
                                 "rmspropsingle | adagrad | adagradsingle | adadelta | adadeltasingle"),
    ("learningRate", 1e-2, "learning rate at t=0"), ("batchSize", 1, "mini-batch size (1 = pure stochastic)"),
    ("lrAdagrad", 1e-3, "learning rate for adagrad"), ("lrDecayAdagrad", 1e-6, "learning rate decay for adagrad"),
    ("epsilonAdagrad", 1e-10, "epsilon for adagrad"), ("modeAdagrad", "global", "mode for distributed adagrad"),
    ("rhoAdadelta", 0.9, "rho for adadelta"), ("lrAdadelta", 1.0, "lr for adadelta"),
    ("epsilonAdadelta", 1e-6, "epsilon for adadelta"), ("modeAdadelta", "global", "mode for distributed adadelta"),
    ("lrAdam", 1e-3, "learning rate for adam"), ("beta1Adam", 0.9, "beta1 for adam"), ("beta2Adam", 0.999, "beta2 for adam"),
    ("epsilonAdam", 1e-8, "epsilon for adam"), ("stepDivAdam", 72, "step divide for adam"),
    ("modeAdam", "global", "mode for adam"), ("gradClip", 0.5, "boundary for gradient clip"),
    ("weightDecay", 1e-6, "weight decay"), ("decayRMSProp", 0.95, "decay for rmsprop"),
    ("lrRMSProp", 1e-4, "learning rate for rmsprop"), ("momentumRMSProp", 0.9, "momentum for rmsprop"),
    ("epsilonRMSProp", 1e-4, "epsilon for rmsprop"), ("modeRMSProp", "global", "mode for rmsprop"),
    ("momentum", 0.0, "momentum (SGD only)"), ("commperiod", 1, "sync updates"), ("movingrate", 0.05, "moving rate"),
    ("type", "float", "float | cuda"),
    ("trainFile", "release/Question_answerlabels.train.ibmprep.ibminput", "input training file"),
    ("validFile", "release/Question_answerlabels.dev.ibmprep.ibminput", "input validation file"),
    ("testFile1", "release/Question_answerlabels.test1.ibmprep.ibminput", "input test file 1"),
    ("testFile2", "release/Question_answerlabels.test2.ibmprep.ibminput", "input test file 2"),
    ("label2answFile", "release/label2answers.ibmprep", "preprocessed label2answers file"),
    ("embeddingFile", "release/filtered.wordvec.vecs100.txt", "input embedding file"),
    ("embeddingDim", 100, "input word embedding dimension"), ("contConvWidth", 2, "continuous convolution filter width"),
    ("wordHiddenDim", 200, "first hidden layer output dimension"), ("numFilters", 3000, "CNN filters amount"),
    ("epoch", 50, "maximum epoch"), ("L1reg", 0.0, "L1 regularization coefficient"),
    ("L2reg", 1e-4, "L2 regularization coefficient"), ("margin", 0.02, "margin for hinge loss"),
    ("maxnegsample", 100, "maximum sampling times for negative examples"),
    ("validMode", "additionalTester", "none | lastClient | additionalTester"),
    ("validSleepTime", 1.0, "validation sleep time in seconds, only for additionalTester"),
    ("servRecvgrad", True, "server recvgrad"), ("servSendparam", True, "server send param to workers"),
    ("mmode", 1, "1|2"), ("outputprefix", "none", "output file prefix"), ("prevtime", 0.0, "time start point"),
    ("loadmodel", "none", "load model file name"), ("preloadBinary", False, "load data from binary files"),
    ("testerfirst", False, "rank 0 is the tester"), ("testerlast", False, "last rank is the tester"),
    ("masterFreq", 2, "this parameter control the ratio of master and client"), ("maxrank", 120, "max rank used"),
    ("singlemode", False, "single mode, server recv params instead of grads"),
    # additions of this framework (bounded runs, synthetic corpus)
    ("maxEvals", 3, "tester iterations before it stops (the reference never stops)"),
    ("synthTrain", 64, "synthetic corpus: training pairs"), ("synthAnswers", 60, "synthetic corpus: answer space"),
    ("synthVocab", 500, "synthetic corpus: vocabulary"),
    ("psServer", "thread", "device path: thread (native server thread per shard owner) | inline (drain once per step)"),
]


def parse(argv=None) -> dict:
    ap = argparse.ArgumentParser(prog="plaunch", description="Options:")
    for name, default, help_ in OPTIONS:
        if isinstance(default, bool):
            ap.add_argument(f"-{name}", action="store_true", default=default, help=help_)
        else:
            ap.add_argument(f"-{name}", type=type(default), default=default, help=help_)
    return vars(ap.parse_args(argv))


def assign_roles(rank: int, size: int, opt: dict):
    """plaunch.lua:124-177 (ranks start at 0): returns (role, sranks, cranks, tranks)."""
    sranks, cranks, tranks, role = [], [], set(), None
    mf = opt["masterFreq"]
    if opt["testerfirst"]:
        cranks.append(0)
        if rank == 0:
            role = "pe"
        for i in range(1, size):
            if i % mf != 0:
                cranks.append(i)
                role = "pt" if rank == i else role
            else:
                sranks.append(i)
                role = "ps" if rank == i else role
    elif opt["testerlast"]:
        for i in range(0, size - 1):
            if (i + 1) % mf != 0:
                cranks.append(i)
                role = "pt" if rank == i else role
            else:
                sranks.append(i)
                role = "ps" if rank == i else role
        cranks.append(size - 1)
        if rank == size - 1:
            role = "pe"
    else:   # no dedicated tester: alternate trainers and servers
        for i in range(size):
            if (i + 1) % mf != 0:
                cranks.append(i)
                role = "pt" if rank == i else role
            else:
                sranks.append(i)
                role = "ps" if rank == i else role
    if opt["validMode"] == "lastClient":
        tranks.add(cranks[-1])
    elif opt["validMode"] == "additionalTester":
        if opt["testerfirst"]:
            tranks.add(0)
        elif opt["testerlast"]:
            tranks.add(size - 1)
        else:
            raise ValueError("Incorrect configuration: additionalTester needs -testerfirst or -testerlast")
    return role, sranks, cranks, tranks


def _corpus(opt: dict, W):
    if opt["preloadBinary"] or os.path.exists(opt["trainFile"]):
        return W.parse_corpus(opt)
    return W.synthetic_corpus(opt["synthVocab"], opt["synthAnswers"], opt["synthTrain"], 20, 10,
                              opt["embeddingDim"], opt["contConvWidth"], seed=0)


def _gpu_main(opt: dict) -> int:
    """``-type cuda`` under torchrun: one process per B200, no server processes -- every GPU trains and owns one shard
    of the parameters together with the server-side rule state; gradients, parameters and acknowledgements move over
    NVLink from inside the kernels (parallel/ps_device.py).  ``-testerlast`` / ``-testerfirst`` turn that rank into
    the shadow tester of bicnn.lua:576-597 (it only pulls, evaluates and saves, and still serves its shard).

        python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 -m mpit_b200.launch.plaunch \
            -type cuda -optimization adam -batchSize 4 -epoch 1"""
    import torch.distributed as dist
    from ..parallel.symm import SymmetricHeap
    from ..workloads import bicnn as W
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    barrier = (lambda: (torch.cuda.synchronize(dev), dist.barrier(), torch.cuda.synchronize(dev))) if world > 1 else None
    if rank == min(1, world - 1):
        print(opt)
    torch.manual_seed(rank)
    random.seed(rank)
    tester = -1
    if opt["validMode"] == "additionalTester" and world > 1:
        tester = 0 if opt["testerfirst"] else (world - 1 if opt["testerlast"] else -1)
    also_test = opt["validMode"] == "lastClient" and rank == world - 1
    server = "thread" if (tester >= 0 or opt["psServer"] == "thread") else "inline"
    heap = SymmetricHeap(dev)
    tr = W.BiCNNDeviceTrainer(_corpus(opt, W), opt, dev, rank, heap, server=server)
    if barrier:
        barrier()                      # the first trainer's parameters are in the shards (pclient.lua:131-133)
    print(f"Client {rank} ready to run {'testing' if rank == tester else 'training'} (gpu {local}, server {server})")
    out = W.run_device_trainer(tr, rank == tester, also_test, barrier)
    if rank != tester:
        print(f"[rank {rank}] {out['steps']} steps in {out['seconds']:.2f} s, "
              f"{out['steps'] * opt['batchSize'] / max(out['seconds'], 1e-9):.1f} questions/s, "
              f"loss {out['loss'][-1] if out['loss'] else float('nan'):.5f}")
    heap.check_timeout()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main(argv=None) -> int:
    opt = parse(argv)
    if opt["type"] == "cuda" and "LOCAL_RANK" in os.environ and "MPIT_RANK" not in os.environ:
        return _gpu_main(opt)
    from .. import msg as mpiT, ps
    from ..workloads import bicnn as W
    torch.set_num_threads(opt["threads"])
    mpiT.Init()
    rank, world = mpiT.get_rank(), mpiT.get_size()
    size = min(world, opt["maxrank"] + 1)
    if rank > opt["maxrank"]:
        print(f"rank {rank} do nothing ")
        mpiT.Finalize()
        return 0
    if rank == 1:
        print(opt)
    torch.manual_seed(rank)
    random.seed(rank)
    role, sranks, cranks, tranks = assign_roles(rank, size, opt)
    conf = {"lr": opt["learningRate"], "rank": rank, "size": size, "world": mpiT.COMM_WORLD, "sranks": sranks,
            "cranks": cranks, "tranks": tranks, "oncuda": opt["type"] == "cuda", "opt": dict(opt)}
    if opt["singlemode"]:
        trainers = [c for c in cranks if c not in tranks]
        conf["opt"]["single_writer"] = trainers[0]
        conf["opt"]["single_readers"] = sorted(tranks)
    if role == "ps":
        ps.pServer(conf).start()
    else:
        corpus = _corpus(opt, W)
        dev = "cuda" if opt["type"] == "cuda" and torch.cuda.is_available() else "cpu"
        if dev == "cuda":     # B200 path: tcgen05 encoder + fused GESD/margin kernels
            torch.cuda.set_device(rank % torch.cuda.device_count())
            tr = W.BiCNNKernelTrainer(corpus, opt, f"cuda:{rank % torch.cuda.device_count()}", rank)
        else:
            tr = W.BiCNNTrainer(corpus, opt, dev, rank)
        pc = ps.pClient(conf)
        is_tester = opt["validMode"] == "additionalTester" and rank in tranks
        also_test = opt["validMode"] == "lastClient" and rank in tranks
        print(f"Client {rank} ready to run {'testing' if is_tester else 'training'}")
        W.run_trainer(tr, pc, is_tester, also_test)
    mpiT.Finalize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
