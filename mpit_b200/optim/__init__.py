"""Distributed optimizers.

``reference``  torch-tensor restatements in the reference's ``optim.<name>(opfunc, w, config)``
               convention (CPU workers + oracles): msgd, downpour, eamsgd/easgd, and the
               parallel/single rmsprop, adam, adamax, adagrad, adadelta.
``device``     the same algorithms as fused sm_100a kernel schedules over an NVLink-sharded centre.
"""
from .reference import (OPTIMIZERS, msgd, downpour, eamsgd, easgd, rmsprop, adam, adamax, adagrad, adadelta,
                        rmspropsingle, adamsingle, adamaxsingle, adagradsingle, adadeltasingle, msgdsingle,
                        apply_rule, rule_state)
