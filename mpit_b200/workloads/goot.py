"""The reference's generic train/test driver (/root/reference/asyncsgd/goot.lua) for torch modules on
the host path: flat parameters, optimizer selection, epoch loop with shuffling and random crops,
confusion matrix, ``train.log``/``test.log``, periodic Torch7 checkpoints, timing buckets.

The mini-batch is evaluated as one batched forward/backward (the reference loops over samples,
goot.lua:198-210; the result -- mean loss, mean gradient -- is the same).
"""
from __future__ import annotations

import os
import time
from typing import Callable, Optional

import torch

from .. import t7
from ..models.mnist import flat_params
from ..optim import reference as R
from ..utils import Logger, ConfusionMatrix, Timers, progress

DEFAULTS = dict(name="sgd", full=False, lr=1e-2, mom=0.99, mb=128, mva=0.0, su=1, maxepoch=1, saveep=10,
                save="results", seed=1, trsize=2000, tesize=1000, crop=0, l2wd=0.0, lrd=0.0, lrp=0.0,
                mommax_quirk=False, quiet=False)


def optimizer_config(opt: dict, pclient, rank: int):
    """goot.lua:108-135.  ``mommax_quirk=True`` reproduces the reference's 'sgd' branch, which passes
    ``mommax`` instead of ``mom`` and therefore runs plain SGD (goot.lua:112-115)."""
    name = opt["name"]
    if name in ("sgd", "msgd"):
        cfg = {"lr": opt["lr"], "mommax": opt["mom"]} if opt["mommax_quirk"] else {"lr": opt["lr"], "mom": opt["mom"]}
        cfg.update(l2wd=opt["l2wd"], lrd=opt["lrd"], lrp=opt["lrp"])
        return R.msgd, cfg
    if name == "downpour":
        return R.downpour, {"lr": opt["lr"], "pclient": pclient, "su": opt["su"], "l2wd": opt["l2wd"], "lrd": opt["lrd"]}
    if name in ("eamsgd", "easgd"):
        return R.eamsgd, {"lr": opt["lr"], "pclient": pclient, "su": opt["su"], "mva": opt["mva"],
                          "mom": 0.0 if name == "easgd" else opt["mom"], "rank": rank, "l2wd": opt["l2wd"],
                          "lrd": opt["lrd"], "lrp": opt["lrp"]}
    if name in R.OPTIMIZERS:
        cfg = dict(opt.get("rule_hyper", {}))
        cfg.update(pclient=pclient, su=opt["su"])
        return R.OPTIMIZERS[name], cfg
    raise ValueError(f"unknown optimizer {name}")


def _crop_batch(x: torch.Tensor, crop: int, gen: Optional[torch.Generator], center: bool) -> torch.Tensor:
    if not crop or crop == x.shape[-1]:
        return x
    B, _, H, W = x.shape
    if center:   # goot.lua:283-285
        y0 = x0 = (H - crop + 1) // 2
        return x[:, :, y0:y0 + crop, x0:x0 + crop]
    offs = torch.randint(0, H - crop + 1, (B, 2), generator=gen)   # goot.lua:177
    return torch.stack([x[i, :, offs[i, 0]:offs[i, 0] + crop, offs[i, 1]:offs[i, 1] + crop] for i in range(B)])


def run(opt_in: dict, model: torch.nn.Module, train_xy, test_xy, pclient=None, rank: int = -1,
        save_fn: Optional[Callable] = None) -> dict:
    opt = dict(DEFAULTS)
    opt.update(opt_in)
    torch.manual_seed(opt["seed"])
    gen = torch.Generator().manual_seed(opt["seed"])
    theta, grad = flat_params(model)
    tm = Timers()
    classes = [str(i) for i in range(10)]
    confusion = ConfusionMatrix(classes)
    os.makedirs(opt["save"], exist_ok=True)
    train_log = Logger(os.path.join(opt["save"], f"train{'' if rank < 0 else rank}.log"))
    test_log = Logger(os.path.join(opt["save"], f"test{'' if rank < 0 else rank}.log"))
    opti, cfg = optimizer_config(opt, pclient, rank)
    if pclient is not None:
        pclient.start(theta, grad)
    (trx, try_), (tex, tey) = train_xy, test_xy
    trsize, tesize = min(opt["trsize"], trx.shape[0]), min(opt["tesize"], tex.shape[0])
    mb = opt["mb"]
    t_start = time.perf_counter()
    hist = {"train_loss": [], "train_acc": [], "test_acc": [], "test_loss": []}
    it = 0
    avg_err = 0.0

    def test():
        model.eval()
        confusion.zero()
        tot = 0.0
        with torch.no_grad():
            for t in range(0, tesize, mb):
                xb = _crop_batch(tex[t:t + mb], opt["crop"], None, True)
                out = model(xb)
                tot += float(torch.nn.functional.nll_loss(out, tey[t:t + mb].long(), reduction="sum"))
                confusion.batch_add(out, tey[t:t + mb])
        acc, loss = confusion.totalValid, tot / tesize
        test_log.add({"% test accuracy": f"{acc:0.4f}",
                      "% wall-clock elapsed time(min)": f"{(time.perf_counter() - t_start) / 60:0.2f}",
                      "% test error": f"{loss:0.2f}"})
        confusion.zero()
        return acc, loss

    for epoch in range(1, opt["maxepoch"] + 1):
        model.train()
        shuffle = torch.randperm(trsize, generator=gen)
        train_err, nb = 0.0, 0
        for t in range(0, trsize, mb):
            progress(t + mb, trsize, 0 if opt["quiet"] else 50 * mb)
            idx = shuffle[t:t + mb]
            xb = _crop_batch(trx[idx], opt["crop"], gen, False)
            yb = try_[idx].long()

            def feval(x):
                with tm("feval"):
                    if x is not theta:
                        theta.copy_(x)
                    grad.zero_()
                    out = model(xb)
                    loss = torch.nn.functional.nll_loss(out, yb)   # mean over the mini-batch
                    loss.backward()
                    confusion.batch_add(out.detach(), yb)
                    return float(loss.detach()), grad

            _, fx = opti(feval, theta, cfg)
            train_err += fx[0]
            avg_err += fx[0]
            nb += 1
            it += 1
        train_acc = confusion.totalValid
        train_log.add({"% train accuracy": f"{train_acc:0.4f}",
                       "% wall-clock elapsed time(min)": f"{(time.perf_counter() - t_start) / 60:0.2f}",
                       "% train loss": f"{train_err / max(nb, 1):0.4f}"})
        hist["train_loss"].append(train_err / max(nb, 1))
        hist["train_acc"].append(train_acc)
        if epoch % opt["saveep"] == 0:   # goot.lua:246-254
            fn = os.path.join(opt["save"], f"model-epoch-{epoch}.net")
            (save_fn or (lambda path, th: t7.save(path, th.clone())))(fn, theta)
        confusion.zero()
        acc, loss = test()
        hist["test_acc"].append(acc)
        hist["test_loss"].append(loss)
    if pclient is not None:
        pclient.stop()
    total = time.perf_counter() - t_start
    tm["sync"] = cfg.get("dusync", 0.0)
    if not opt["quiet"]:   # goot.lua:330-337
        print(rank, "total training time is", total)
        print(rank, "total function eval time is", tm.get("feval", 0.0))
        print(f"\n===> time to learn 1 sample = {total / max(trsize * opt['maxepoch'], 1) * 1000:.4f}ms")
        print(rank, "total sync time is", tm["sync"])
    train_log.close()
    test_log.close()
    hist.update(total=total, feval=tm.get("feval", 0.0), sync=tm["sync"], theta=theta, avg_err=avg_err / max(it, 1))
    return hist
