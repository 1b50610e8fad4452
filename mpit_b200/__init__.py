"""mpit_b200 -- asynchronous / elastic-averaging distributed SGD for one 8xB200 NVSwitch box.

A from-scratch, Blackwell-first framework with the capabilities of fanshiqing/mpiT:

* ``mpit_b200.msg``      tagged message layer (Init/Send/Recv/Isend/Irecv/Iprobe/Test/Wait/
                         Cancel/Bcast/Allreduce/...) over a C++ host transport
* ``mpit_b200.aio``      coroutine-driven non-blocking IO (aio_send/aio_recv/co_execute/co_ping/
                         co_wait, Queue)
* ``mpit_b200.ps``       parameter server / client protocol (host path) and the NVLink-sharded
                         centre variable (device path)
* ``mpit_b200.optim``    msgd, downpour, eamsgd/easgd, and parallel/single
                         rmsprop/adam/adamax/adagrad/adadelta
* ``mpit_b200.models``   the CIFAR 7-layer CNN, MNIST-shaped MLP/CNN, BiCNN
* ``mpit_b200.ops``      Python front-ends of the hand-written sm_100a kernels
* ``mpit_b200.t7``       Torch7 binary (de)serialiser for checkpoint compatibility
* ``mpit_b200.launch``   claunch / glaunch / mlaunch / plaunch / ptest / testreduceall equivalents
"""
__version__ = "0.1.0"
