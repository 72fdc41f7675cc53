"""CPU ORACLE — TEST INFRASTRUCTURE ONLY.  **Parity unpinned.**

Restatement (plain torch on CPU, on top of ``oracle.ref_cpu.MessagePassing``) of
the operators SURVEY.md §8(f) ranks next to the hot path: the ``aggr='add'|'mean'``
convolutions whose message is ``w_e * x_j`` or a per-target softmax (ChebConv,
GraphConv, SGConv, APPNP, SAGEConv, GINConv, AGNNConv — selectable models of the
fork, ConvexPruning.py:227-275, README.md:136-146), global pooling
(ConvexPruning.py:22,301-338; examples/enzymes_topk_pool.py:7-8,38-50),
``Batch.from_data_list`` (ConvexPruning.py:459-460,530) and ``knn_graph``
(README.md:149-150).

Provenance as for ``ref_cpu``: the code of these operators lives in the pinned,
un-vendored ``torch-geometric==1.4.3`` (requirement.txt:7); what is written here
is its published algorithm [UPSTREAM-RECALL], each class written the way that
package writes it — an explicit ``message()``/``update()`` over the generic
gather -> message -> scatter_ propagate — so the CUDA fast paths are compared
against the unfused formulation.  The reference holds no tests or golden vectors
for them.  Only ``tests/`` may import this module.
"""
import torch
import torch.nn.functional as F
from torch.nn import Parameter

from . import ref_cpu as O


class ChebConv(O.MessagePassing):
    """T_0 = x, T_1 = L~ x, T_k = 2 L~ T_{k-1} - T_{k-2}; L~ = -D^-1/2 A D^-1/2 (lambda_max = 2,
    self-loops removed, degree by source row)."""

    def __init__(self, in_channels, out_channels, K, bias=True):
        super().__init__(aggr='add')
        self.weight = Parameter(torch.zeros(K, in_channels, out_channels))
        self.bias = Parameter(torch.zeros(out_channels)) if bias else None

    @staticmethod
    def norm(edge_index, num_nodes, edge_weight, dtype=None):
        edge_index, edge_weight = O.remove_self_loops(edge_index, edge_weight)
        if edge_weight is None:
            edge_weight = torch.ones((edge_index.size(1),), dtype=dtype)
        row, col = edge_index
        deg = O.scatter_add(edge_weight, row, dim=0, dim_size=num_nodes)
        dis = deg.pow(-0.5)
        dis[dis == float('inf')] = 0
        return edge_index, -dis[row] * edge_weight * dis[col]

    def forward(self, x, edge_index, edge_weight=None):
        edge_index, norm = self.norm(edge_index, x.size(0), edge_weight, x.dtype)
        Tx_0 = x
        out = torch.matmul(Tx_0, self.weight[0])
        if self.weight.size(0) > 1:
            Tx_1 = self.propagate(edge_index, x=x, norm=norm)
            out = out + torch.matmul(Tx_1, self.weight[1])
        for k in range(2, self.weight.size(0)):
            Tx_2 = 2 * self.propagate(edge_index, x=Tx_1, norm=norm) - Tx_0
            out = out + torch.matmul(Tx_2, self.weight[k])
            Tx_0, Tx_1 = Tx_1, Tx_2
        return out + self.bias if self.bias is not None else out

    def message(self, x_j, norm):
        return norm.view(-1, 1) * x_j


class GraphConv(O.MessagePassing):
    """x_i' = Theta_1 x_i + aggr_j e_ji Theta_2 x_j."""

    def __init__(self, in_channels, out_channels, aggr='add', bias=True):
        super().__init__(aggr=aggr)
        self.weight = Parameter(torch.zeros(in_channels, out_channels))
        self.lin = torch.nn.Linear(in_channels, out_channels, bias=bias)

    def forward(self, x, edge_index, edge_weight=None, size=None):
        h = torch.matmul(x, self.weight)
        return self.propagate(edge_index, size=size, x=x, h=h, edge_weight=edge_weight)

    def message(self, h_j, edge_weight):
        return h_j if edge_weight is None else edge_weight.view(-1, 1) * h_j

    def update(self, aggr_out, x):
        return aggr_out + self.lin(x)


class SGConv(O.MessagePassing):
    """X' = (D^-1/2 (A+I) D^-1/2)^K X Theta."""

    def __init__(self, in_channels, out_channels, K=1, bias=True):
        super().__init__(aggr='add')
        self.K = K
        self.lin = torch.nn.Linear(in_channels, out_channels, bias=bias)

    def forward(self, x, edge_index, edge_weight=None):
        edge_index, norm = O.gcn_norm(edge_index, x.size(0), edge_weight, dtype=x.dtype)
        for _ in range(self.K):
            x = self.propagate(edge_index, x=x, norm=norm)
        return self.lin(x)

    def message(self, x_j, norm):
        return norm.view(-1, 1) * x_j


class APPNP(O.MessagePassing):
    """X^(k) = (1 - alpha) A^ X^(k-1) + alpha X^(0)."""

    def __init__(self, K, alpha):
        super().__init__(aggr='add')
        self.K, self.alpha = K, alpha

    def forward(self, x, edge_index, edge_weight=None):
        edge_index, norm = O.gcn_norm(edge_index, x.size(0), edge_weight, dtype=x.dtype)
        hidden = x
        for _ in range(self.K):
            x = self.propagate(edge_index, x=x, norm=norm)
            x = x * (1 - self.alpha) + self.alpha * hidden
        return x

    def message(self, x_j, norm):
        return norm.view(-1, 1) * x_j


class SAGEConv(O.MessagePassing):
    """mean over the neighbourhood (self-loop added), then one weight matrix."""

    def __init__(self, in_channels, out_channels, normalize=False, bias=True):
        super().__init__(aggr='mean')
        self.normalize = normalize
        self.weight = Parameter(torch.zeros(in_channels, out_channels))
        self.bias = Parameter(torch.zeros(out_channels)) if bias else None

    def forward(self, x, edge_index):
        edge_index, _ = O.remove_self_loops(edge_index)
        edge_index, _ = O.add_self_loops(edge_index, num_nodes=x.size(0))
        return self.propagate(edge_index, x=x)

    def update(self, aggr_out):
        aggr_out = torch.matmul(aggr_out, self.weight)
        if self.bias is not None:
            aggr_out = aggr_out + self.bias
        if self.normalize:
            aggr_out = F.normalize(aggr_out, p=2, dim=-1)
        return aggr_out


class GINConv(O.MessagePassing):
    """x_i' = nn((1 + eps) x_i + sum_j x_j), self-loops removed."""

    def __init__(self, nn, eps=0.0):
        super().__init__(aggr='add')
        self.nn, self.eps = nn, eps

    def forward(self, x, edge_index):
        edge_index, _ = O.remove_self_loops(edge_index)
        return self.nn((1 + self.eps) * x + self.propagate(edge_index, x=x))


class AGNNConv(O.MessagePassing):
    """alpha_ij = softmax_i(beta * cos(x_i, x_j)); x_i' = sum_j alpha_ij x_j (self-loops added)."""

    def __init__(self, beta=1.0):
        super().__init__(aggr='add')
        self.beta = Parameter(torch.tensor([beta]))

    def forward(self, x, edge_index):
        edge_index, _ = O.remove_self_loops(edge_index)
        edge_index, _ = O.add_self_loops(edge_index, num_nodes=x.size(0))
        x_norm = F.normalize(x, p=2, dim=-1)
        return self.propagate(edge_index, x=x, x_norm=x_norm, num_nodes=x.size(0))

    def message(self, edge_index_i, x_j, x_norm_i, x_norm_j, num_nodes):
        alpha = self.beta * (x_norm_i * x_norm_j).sum(dim=-1)
        alpha = O.softmax(alpha, edge_index_i, num_nodes)
        return x_j * alpha.view(-1, 1)


# ---- global pooling: scatter_ keyed on the batch vector ----------------------------
def global_add_pool(x, batch, size=None):
    size = int(batch.max()) + 1 if size is None else size
    return O.scatter_('add', x, batch, dim=0, dim_size=size)


def global_mean_pool(x, batch, size=None):
    size = int(batch.max()) + 1 if size is None else size
    return O.scatter_('mean', x, batch, dim=0, dim_size=size)


def global_max_pool(x, batch, size=None):
    size = int(batch.max()) + 1 if size is None else size
    return O.scatter_('max', x, batch, dim=0, dim_size=size)


# ---- Batch.from_data_list: block-diagonal stacking, pure Python loops ----------------
def batch_from_list(graphs):
    """graphs: list of dicts {x [n,F], edge_index [2,e], y [...]}.  Returns dict with the
    concatenated x / y, edge_index shifted by the cumulative node count, and ``batch``."""
    xs, eis, ys, batch = [], [], [], []
    cum = 0
    for gi, g in enumerate(graphs):
        n = g['x'].size(0)
        xs.append(g['x'])
        eis.append(g['edge_index'] + cum)
        if 'y' in g:
            ys.append(g['y'])
        batch += [gi] * n
        cum += n
    out = {'x': torch.cat(xs, 0), 'edge_index': torch.cat(eis, 1),
           'batch': torch.tensor(batch, dtype=torch.long)}
    if ys:
        out['y'] = torch.cat(ys, 0)
    return out


# ---- knn_graph: exact k nearest neighbours inside each example, loops in Python -------
def knn_graph(x, k, batch=None, loop=False):
    """edge_index [2, n*k]: row 0 = neighbour (source), row 1 = centre (target), the
    neighbours of a centre ordered by distance (ties by index)."""
    n = x.size(0)
    batch = torch.zeros(n, dtype=torch.long) if batch is None else batch
    src, dst = [], []
    xd = x.double()
    for i in range(n):
        cand = [j for j in range(n) if batch[j] == batch[i] and (loop or j != i)]
        d = [float(((xd[i] - xd[j]) ** 2).sum()) for j in cand]
        order = sorted(range(len(cand)), key=lambda t: (d[t], cand[t]))[:k]
        src += [cand[t] for t in order]
        dst += [i] * len(order)
    return torch.tensor([src, dst], dtype=torch.long)


# ---- TopKPooling (ConvexPruning.py:301-338): per-graph top-k by projection score --------
def topk_pooling(x, edge_index, batch, weight, ratio):
    """score = tanh(x.w / ||w||); per graph keep the ceil(ratio * n) best nodes (descending
    score, graphs in order); x' = x[perm] * score[perm]; edges of the induced subgraph,
    renumbered.  Plain Python loops.  Returns (x', edge_index', batch', perm, score[perm])."""
    score = torch.tanh((x * weight.view(1, -1)).sum(-1) / weight.norm(p=2))
    perm = []
    for g in range(int(batch.max()) + 1):
        ids = [i for i in range(x.size(0)) if int(batch[i]) == g]
        k = -(-int(round(ratio * len(ids) * 1e6)) // 10 ** 6)          # ceil, float-safe
        ids.sort(key=lambda i: (-float(score[i].detach()), i))
        perm += ids[:k]
    perm = torch.tensor(perm, dtype=torch.long)
    pos = {int(v): i for i, v in enumerate(perm.tolist())}
    keep = [(pos[int(a)], pos[int(b)]) for a, b in edge_index.t().tolist()
            if int(a) in pos and int(b) in pos]
    ei = torch.tensor(keep, dtype=torch.long).t().reshape(2, -1)
    return x[perm] * score[perm].view(-1, 1), ei, batch[perm], perm, score[perm]


# ---- fps / radius / radius_graph (torch_cluster 1.5.x semantics, [UPSTREAM-RECALL]; ref
# sites examples/pointnet++.py:39-41,46-48): plain Python loops ---------------------------
def fps(x, batch=None, ratio=0.5, start=None):
    """Farthest point sampling per example: ceil(ratio * n_b) indices in selection order;
    the first sample is ``start[b]`` (local index, default 0), every next one the point with
    the largest squared distance to the selected set, ties to the lower index."""
    n = x.size(0)
    batch = torch.zeros(n, dtype=torch.long) if batch is None else batch
    xd = x.double()
    out = []
    for b in range(int(batch.max()) + 1 if n else 0):
        ids = [i for i in range(n) if int(batch[i]) == b]
        if not ids:
            continue
        m = -(-int(round(ratio * len(ids) * 1e6)) // 10 ** 6)
        cur = ids[0 if start is None else int(start[b])]
        dmin = {i: float('inf') for i in ids}
        for _ in range(m):
            out.append(cur)
            for i in ids:
                dmin[i] = min(dmin[i], float(((xd[i] - xd[cur]) ** 2).sum()))
            cur = max(ids, key=lambda i: (dmin[i], -i))
    return torch.tensor(out, dtype=torch.long)


def radius(x, y, r, batch_x=None, batch_y=None, max_num_neighbors=32):
    """[2, E]: row 0 = y index, row 1 = x index; per y the first max_num_neighbors x of the
    same example with |x - y| <= r, in ascending x index."""
    batch_x = torch.zeros(x.size(0), dtype=torch.long) if batch_x is None else batch_x
    batch_y = torch.zeros(y.size(0), dtype=torch.long) if batch_y is None else batch_y
    xd, yd = x.double(), y.double()
    rows, cols = [], []
    for q in range(y.size(0)):
        found = 0
        for j in range(x.size(0)):
            if int(batch_x[j]) != int(batch_y[q]) or found >= max_num_neighbors:
                continue
            if float(((xd[j] - yd[q]) ** 2).sum()) <= r * r:
                rows.append(q)
                cols.append(j)
                found += 1
    return torch.tensor([rows, cols], dtype=torch.long).reshape(2, -1)


def radius_graph(x, r, batch=None, loop=False, max_num_neighbors=32, flow='source_to_target'):
    row, col = radius(x, x, r, batch, batch, max_num_neighbors if loop else max_num_neighbors + 1)
    row, col = (col, row) if flow == 'source_to_target' else (row, col)
    if not loop:
        mask = row != col
        row, col = row[mask], col[mask]
    return torch.stack([row, col], dim=0)
