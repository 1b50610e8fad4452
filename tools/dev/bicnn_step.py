"""One BiCNN engine step at the reference's dimensions, for ncu launch lists: python tools/dev/bicnn_step.py"""
import sys
sys.path.insert(0, ".")
import torch
from mpit_b200.models.bicnn import BiCNN, pad_sentences
from mpit_b200.models.bicnn_engine import BiCnnEngine
dev = torch.device("cuda:0")
V, E_, H_, F_, S = 22354, 100, 200, 3000, 256
torch.manual_seed(1)
bm = BiCNN(V, E_, H_, F_, 2)
sents = [torch.randint(0, V, (int(k),)) for k in torch.randint(6, 31, (S,))]
tokens, lens = pad_sentences(sents)
q = torch.arange(0, 64).repeat_interleave(2)
qi = torch.stack([q, q], 1).reshape(-1)[:256].int().to(dev)
ai = torch.stack([64 + q, 128 + torch.arange(128)], 1).reshape(-1)[:256].int().to(dev)
theta = torch.cat([p_.detach().reshape(-1) for p_ in bm.param_list()]).to(dev)
be = BiCnnEngine(V, S, 32, E_, H_, F_, 2, margin=0.5)
be.load_flat(theta)
be.set_sentences(tokens.to(dev).int(), lens.to(dev).int())
g = torch.zeros_like(theta)
def step():
    be.encode(); be.loss_backward(qi, ai, g); be.flat_grad(g)
for _ in range(3): step()
torch.cuda.synchronize()
torch.cuda.profiler.start()
step()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
be.check()
print("ok")
