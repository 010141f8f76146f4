"""Training path (autograd) of MOTMPNet.forward -- to be filled in."""


def forward_training(model, x, edge_index, edge_attr, plan=None):
    raise NotImplementedError("gnn-cca_b200: training (autograd) path not built yet")
