import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from mpit_b200.models.bicnn import BiCNN, pad_sentences
from mpit_b200.models.bicnn_engine import BiCnnEngine
E, H, F = 100, 200, 3000
torch.manual_seed(0)
V, S = 300, 12
model = BiCNN(V, E, H, F, 2)
sents = [torch.randint(0, V, (int(n),)) for n in torch.randint(4, 30, (S,))]
tokens, lens = pad_sentences(sents)
print("lens", lens.tolist())
theta = torch.cat([p.detach().reshape(-1) for p in model.param_list()])
eng = BiCnnEngine(V, S, 32, E, H, F, 2)
eng.load_flat(theta.cuda())
eng.set_sentences(tokens.cuda().int(), lens.cuda().int())
y = eng.encode()
torch.cuda.synchronize()
print("err", int(eng.err.item()))
tk = torch.zeros(S, 32, dtype=torch.long); tk[:, :tokens.shape[1]] = tokens
emb_ref = model.lookup[tk]
h_ref = torch.tanh(emb_ref @ model.lin_w.t() + model.lin_b).detach()
emb = eng.emb.view(S, 32, eng.Ep)[:, :, :E].cpu()
h = eng.h[:S * 32 * eng.Hp].view(S, 32, eng.Hp)[:, :, :H].cpu()
rel = lambda a, b: float((a - b).norm() / (b.norm() + 1e-12))
print("emb rel", rel(emb, emb_ref.detach()))
for s in range(S):
    print(s, "h rel", round(rel(h[s], h_ref[s]), 5), end=" | ")
print()
frames = torch.cat([h_ref[:, :-1], h_ref[:, 1:]], 2)
c = frames @ model.conv_w.detach().t() + model.conv_b.detach()
for s in range(S):
    T = int(lens[s]) - 1
    r_ref = torch.relu(c[s, :T].max(0).values)
    r = eng.r.view(S, eng.Fp)[s, :F].cpu()
    a_ref = c[s, :T].argmax(0)
    a = eng.arg.view(S, eng.Fp)[s, :F].cpu().long()
    print(s, "T", T, "r rel", round(rel(r, r_ref), 5), "zeros", int((r == 0).sum()), "ref zeros", int((r_ref == 0).sum()),
          "arg match", float((a == a_ref)[r_ref > 0].float().mean()), "argmax>", int(a.max()))
