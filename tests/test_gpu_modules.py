"""Stand-alone forwards of the sub-modules (reference models/mpn.py:32-54 MetaLayer, :59-69 EdgeModel, :71-101 NodeModel) on
caller-supplied tensors, against the same formulas in plain torch on the CPU."""
import copy
import importlib

import pytest
import torch

from oracle import mpn_oracle

pytestmark = pytest.mark.gpu
cfgmod = importlib.import_module("gnn-cca_b200.config")


def _ref_step(model, x, ei, e, agg):
    row, col = ei[0], ei[1]
    em, nm = model.MPNet.edge_model.edge_mlp, model.MPNet.node_model.node_mlp
    We, be = em.linear(0).weight.cpu(), em.linear(0).bias.cpu()
    Wn, bn = nm.linear(0).weight.cpu(), nm.linear(0).bias.cpu()
    e_new = torch.relu(torch.nn.functional.linear(torch.cat([x[row], x[col], e], dim=1), We, be))          # mpn.py:48,68-69
    m = torch.relu(torch.nn.functional.linear(torch.cat([x[row], e_new], dim=1), Wn, bn))                  # mpn.py:97
    fn = {"sum": mpn_oracle.scatter_add, "mean": mpn_oracle.scatter_mean, "max": lambda *a: mpn_oracle.scatter_max(*a)[0]}[agg]
    return fn(m, row, 0, x.size(0)), e_new, m


@pytest.mark.parametrize("reattach", [False, True])
@pytest.mark.parametrize("agg", ["sum", "mean", "max"])
def test_meta_layer_edge_model_node_model_forward(agg, reattach, cuda_device):
    import gnncca
    params = cfgmod.graph_net_params()
    params["node_agg_fn"] = agg
    params["reattach_initial_nodes"] = params["reattach_initial_edges"] = reattach
    torch.manual_seed(3)
    model = gnncca.MOTMPNet(copy.deepcopy(params), None, "bdnet_market").to(cuda_device).eval()
    g = torch.Generator().manual_seed(5)
    N, E = 300, 7000
    x = torch.randn(N, 64 if reattach else 32, generator=g)
    e = torch.randn(E, 12 if reattach else 6, generator=g)
    ei = torch.randint(0, N, (2, E), generator=g)
    ei[0][ei[0] == 17] = 18                                    # node 17 has no out-edge: aggregates to 0
    with torch.no_grad():
        h_ref, e_ref, m_ref = _ref_step(model, x, ei, e, agg)
        xd, ed, eid = x.to(cuda_device), e.to(cuda_device), ei.to(cuda_device)
        h, e_new = model.MPNet(xd, eid, ed)                    # MetaLayer.forward, caller's (unsorted) edge order
        assert h.shape == (N, 32) and e_new.shape == (E, 6)
        tol = lambda got, ref: float(((got.cpu() - ref).abs() / (1 + ref.abs() + 0.01 * ref.abs().max())).max())
        assert tol(e_new, e_ref) <= 1e-4 and tol(h, h_ref) <= 1e-4
        assert float(h[17].abs().max()) == 0.0
        e2 = model.MPNet.edge_model(xd[eid[0]], xd[eid[1]], ed)          # EdgeModel.forward
        assert tol(e2, e_ref) <= 1e-4
        h2 = model.MPNet.node_model(xd, eid, e2)                         # NodeModel.forward
        assert tol(h2, h_ref) <= 1e-4
    with pytest.raises(NotImplementedError):                             # autograd: only through MOTMPNet.forward
        model.MPNet(xd, eid, ed)
