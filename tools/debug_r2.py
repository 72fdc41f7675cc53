import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'pytorch_geometric-1_b200'))
import torch, torch.nn.functional as F, copy
from oracle import ref_cpu as O
from torch_geometric.datasets import synthetic
from torch_geometric.nn import GCNConv, EdgeConv
from torch_geometric.transforms import NormalizeFeatures
from torch_geometric import ops
dev = torch.device('cuda:0')

def err(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp(min=1e-30))

def sync(c, r):
    r.load_state_dict({k: v.detach().cpu().float() for k, v in c.state_dict().items()})

print('== cfg1 noise')
data = NormalizeFeatures()(synthetic.cora_like())
for noise in (False, True):
    ei = synthetic.add_noise_edges(data.edge_index, data.x.size(0), 0.01, 0.01, 11) if noise else data.edge_index
    torch.manual_seed(0)
    c1, c2 = GCNConv(1433, 16, cached=True).to(dev), GCNConv(16, 7, cached=True).to(dev)
    r1, r2 = O.GCNConv(1433, 16, cached=True), O.GCNConv(16, 7, cached=True)
    sync(c1, r1); sync(c2, r2)
    eid = ei.to(dev)
    h1 = c1(data.x.to(dev), eid); h1.retain_grad()
    a1 = F.relu(h1); h2 = c2(a1, eid); h2.retain_grad()
    out = F.log_softmax(h2, dim=1)
    g1 = r1(data.x, ei); g1.retain_grad()
    b1 = F.relu(g1); g2 = r2(b1, ei); g2.retain_grad()
    oref = F.log_softmax(g2, dim=1)
    F.nll_loss(out[data.train_mask.to(dev)], data.y[data.train_mask].to(dev)).backward()
    F.nll_loss(oref[data.train_mask], data.y[data.train_mask]).backward()
    print('noise', noise, 'out', err(out, oref), 'dh2', err(h2.grad, g2.grad), 'dh1', err(h1.grad, g1.grad),
          'dW2', err(c2.weight.grad, r2.weight.grad), 'dW1', err(c1.weight.grad, r1.weight.grad),
          'db1', err(c1.bias.grad, r1.bias.grad))
    # dW1 recomputed from the GPU's own dh1 chain in fp64
    ei2, norm = c1.cached_result
    n = data.x.size(0)
    A = torch.zeros(n, n, dtype=torch.float64, device=dev)
    A.index_put_((ei2[1], ei2[0]), norm.double(), accumulate=True)
    gxw = A.t() @ h1.grad.double()
    dW1 = data.x.to(dev).double().t() @ gxw
    print('   dW1 gpu vs fp64 from gpu dh1:', err(c1.weight.grad, dW1), ' oracle vs same:', err(r1.weight.grad, dW1))

print('== cfg4')
d4 = synthetic.knn_clouds(32, 1024, 20, seed=3)
torch.manual_seed(1)
mlp = torch.nn.Sequential(torch.nn.Linear(6, 64), torch.nn.ReLU(), torch.nn.Linear(64, 64))
ref = O.EdgeConv(copy.deepcopy(mlp)); conv = EdgeConv(mlp).to(dev)
ei = d4.edge_index
ra = ref(d4.x, ei); rb = conv(d4.x.to(dev), ei.to(dev))
print('edgeconv out', err(rb, ra))
xi = ops.index_select0(d4.x.to(dev), ei[1].to(dev)); print('gather x_i', err(xi, d4.x[ei[1]]))
xj = ops.index_select0(d4.x.to(dev), ei[0].to(dev)); print('gather x_j', err(xj, d4.x[ei[0]]))
msg_g = conv.nn(torch.cat([xi, xj - xi], dim=1)); msg_c = ref.nn(torch.cat([d4.x[ei[1]], d4.x[ei[0]] - d4.x[ei[1]]], dim=1))
print('msg', err(msg_g, msg_c))
from torch_geometric.utils import scatter_
print('scatter_max of cpu msgs', err(scatter_('max', msg_c.detach().to(dev), ei[1].to(dev), 0, 32768), O.scatter_('max', msg_c.detach(), ei[1], 0, 32768)))
diff = (rb.detach().cpu() - ra.detach()).abs()
bad = (diff > 1e-3).nonzero()
print('bad elems', bad.shape, bad[:10].tolist(), 'max ref', float(ra.abs().max()))

print('== cfg5 small')
n, pairs = 2449029 // 16, 31000000 // 16
data = synthetic.products_like(seed=4, num_nodes=n, num_pairs=pairs, num_features=100)
ei = data.edge_index
torch.manual_seed(0)
convs = [GCNConv(100, 256, cached=True), GCNConv(256, 256, cached=True), GCNConv(256, 47, cached=True)]
refs = [O.GCNConv(100, 256, cached=True), O.GCNConv(256, 256, cached=True), O.GCNConv(256, 47, cached=True)]
for c, r in zip(convs, refs):
    c.to(dev); sync(c, r)
def fwd(layers, x, e):
    hs = []; h = x
    for k, c in enumerate(layers):
        h = c(h, e); h.retain_grad(); hs.append(h)
        if k < 2: h = F.relu(h)
    return F.log_softmax(h, dim=1), hs
out, hs = fwd(convs, data.x.to(dev), ei.to(dev))
oref, hr = fwd(refs, data.x, ei)
F.nll_loss(out, data.y.to(dev)).backward(); F.nll_loss(oref, data.y).backward()
print('out', err(out, oref))
for k in range(3):
    print('layer', k, 'h', err(hs[k], hr[k]), 'dh', err(hs[k].grad, hr[k].grad), 'dW', err(convs[k].weight.grad, refs[k].weight.grad), 'db', err(convs[k].bias.grad, refs[k].bias.grad))
# fp64 dW0 from gpu's own dh0
ei2, norm = convs[0].cached_result
ax = torch.zeros(n, 100, dtype=torch.float64, device=dev).index_add_(0, ei2[1], norm.double().view(-1,1) * data.x.to(dev).double()[ei2[0]])
dW0 = ax.t() @ hs[0].grad.double()
print('dW0 gpu vs fp64(from gpu dh0)', err(convs[0].weight.grad, dW0), 'oracle vs same', err(refs[0].weight.grad, dW0))
axo = torch.zeros(n, 100, dtype=torch.float64).index_add_(0, ei2[1].cpu(), norm.double().cpu().view(-1,1) * data.x.double()[ei2[0].cpu()])
print('oracle dW0 vs fp64(from oracle dh0)', err(refs[0].weight.grad, axo.t() @ hr[0].grad.double()))
