"""Diagnostic (GPU box): error budget of the CUDA path vs the fp64 oracle on awkward inputs.  Test infrastructure — it
lives under tests/ because it imports the oracle; run as `python tests/tools/diag_parity.py`."""
import sys, os
_TESTS = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(_TESTS))
sys.path.insert(0, _TESTS)
import torch
import torch.nn.functional as F
import deltaconv_b200 as D
from oracle import geometry as og, network as on
from helpers import load, t, rel_err

DEV = 'cuda'
torch.manual_seed(1)
N = 1000
x = torch.rand(N, 3)
xg = x.to(DEV)
ei = D.geometry.knn_graph(xg, 20, loop=True, flow='target_to_source')
normal, xb, yb = D.geometry.estimate_basis(xg, ei)
idx = og.knn_ell(x, 20)
print('knn equal', torch.equal(ei._dc_graph.idx.cpu().long(), idx))
for lam in (1e-3, 1e-8):
    grad, div = D.geometry.build_grad_div(xg, normal, xb, yb, ei, regularizer=lam)
    G64, D64 = og.fit_operators(x.double(), normal.cpu().double(), xb.cpu().double(), yb.cpu().double(), idx, None, 1.0, lam)
    G32, D32 = og.fit_operators(x, normal.cpu(), xb.cpu(), yb.cpu(), idx, None, 1.0, lam)
    for nm, ours, o32, ref in (('G', grad.G, G32, G64), ('D', div.D, D32, D64)):
        sc = ref.abs().amax(dim=(1, 2), keepdim=True)
        print(f'lam={lam} {nm}: cuda row-scaled err {((ours.cpu().double()-ref).abs()/sc).max():.3e}   oracle-fp32 err {((o32.double()-ref).abs()/sc).max():.3e}')

# stage-wise DeltaConv comparison
grad, div = D.geometry.build_grad_div(xg, normal, xb, yb, ei, regularizer=1e-3)
G64, D64 = grad.G.cpu().double(), div.D.cpu().double()
conv = D.nn.DeltaConv(3, 1, depth=1, centralized=False).to(DEV)
oc = on.DeltaConv(3, 1, depth=1, centralized=False).double()
oc.load_state_dict({k: v.double() if v.is_floating_point() else v for k, v in conv.state_dict().items()})
target = torch.rand(N, 1)
v = grad @ xg
out, vout = conv(xg, v, grad, div, ei)
loss = F.l1_loss(out, target.to(DEV)); loss.backward()
xo = x.double()
vo = og.apply_grad(idx, G64, xo)
oo, ovo = oc(xo, vo, idx, G64, D64)
lo = F.l1_loss(oo, target.double()); lo.backward()
print('v0', rel_err(v.cpu(), vo), 'x_out', rel_err(out.detach().cpu(), oo.detach()), 'v_out', rel_err(vout.detach().cpu(), ovo.detach()))
for (n1, p1), (n2, p2) in zip(conv.named_parameters(), oc.named_parameters()):
    if p2.grad is not None:
        print(n1, 'grad abs err', (p1.grad.cpu().double() - p2.grad).abs().max().item(), 'scale', p2.grad.abs().max().item())

# model without normals
g, z = load('model_cls_nonormal')
model = D.utils.fill_deterministic(D.models.DeltaNetClassification(3, 5, conv_channels=[8, 16], num_neighbors=int(g.k)), 2)
for m in model.modules():
    if isinstance(m, torch.nn.Dropout): m.p = 0.0
model = model.to(DEV)
data = D.utils.Data(pos=t(g.pos), batch=t(g.batch), y=t(g.y)); data.norm = None
data = data.to(DEV)
model.eval()
with torch.no_grad():
    print('nonormal eval logits err', rel_err(model(data).cpu(), t(g.ref64_logits_eval)))
model.train()
lg = model(data)
print('nonormal train logits err', rel_err(lg.detach().cpu(), t(g.ref64_logits_train)), ' (reference fp32 vs fp64:', rel_err(t(g.ref32_logits_train), t(g.ref64_logits_train)), ')')
