"""Multi-GPU plumbing: symmetric heap, sharded centre variable, device collectives."""
from .symm import SymmetricHeap, SymmBuffer, LocalPeerHeap
from .center import ShardedCenter, partition
from .collectives import DeviceCollectives, DeviceRequest
from .window import DeviceWindow
