"""Per-tensor gradient distance of the CUDA path from the fp64 oracle at a BASELINE config shape, next to the fp32 oracle's own
distance (the floor).  Diagnostic for tests/test_gpu_configs.py; run on the GPU box:  python tests/tools/diag_config_grads.py seg"""
import copy
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'tests')]
import torch                       # noqa: E402
import torch.nn.functional as F    # noqa: E402
import deltaconv_b200 as D         # noqa: E402
from oracle import network as on   # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else 'seg'
if which == 'seg':
    data = D.utils.synthetic_batch(2, 2048, config=3, num_classes=50, per_point_labels=True, categories=16)
    kw = dict(conv_channels=[64, 128, 256], mlp_depth=2, embedding_size=1024, categorical_vector=True, num_neighbors=30)
    build, loss, seed = (lambda m: m.DeltaNetSegmentation(3, 50, **kw)), F.cross_entropy, 7
else:
    data = D.utils.synthetic_batch(32, 1024, config=2, num_classes=40)
    build, loss, seed = (lambda m: m.DeltaNetClassification(3, 40, num_neighbors=20)), on.smoothed_cross_entropy, 1


def nodrop(m):
    for x in m.modules():
        if isinstance(x, torch.nn.Dropout):
            x.p = 0.0
    return m


model = nodrop(D.utils.fill_deterministic(build(D.models), seed))
o64 = nodrop(build(on)).double()
o64.load_state_dict({k: (v.double() if v.is_floating_point() else v) for k, v in model.state_dict().items()})
o32 = copy.deepcopy(o64).float()
model = model.cuda().train()
o64.train()
o32.train()
dev = data.to('cuda')
loss(model(dev), dev.y).backward()
loss(o64(data.double()), data.y).backward()
loss(o32(data), data.y).backward()
g64, g32 = dict(o64.named_parameters()), dict(o32.named_parameters())
scale = max(float(p.grad.abs().max()) for p in g64.values() if p.grad is not None)
rows = []
for n, p in model.named_parameters():
    if g64[n].grad is None:
        continue
    r = g64[n].grad
    den = max(float(r.abs().max()), 1e-3 * scale)
    rows.append((float((p.grad.cpu().double() - r).abs().max()) / den, float((g32[n].grad.double() - r).abs().max()) / den,
                 float((p.grad.cpu().double() - g32[n].grad.double()).abs().max()) / den, float(r.abs().max()), n))
rows.sort(reverse=True)
print('ours-vs-64   oracle32-vs-64   ours-vs-oracle32   max|ref|   tensor')
for r in rows:
    print(f'{r[0]:.3e}   {r[1]:.3e}   {r[2]:.3e}   {r[3]:.3e}   {r[4]}')
