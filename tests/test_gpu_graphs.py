"""CUDA-graph replay of the training step (gnncca.GraphedTrainStep) against the eager step: same loss, same gradients,
same BatchNorm buffers, on a stream of batches whose shape repeats while their structure and the weights change."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu


def _batch(dev, split, seed, frames=6, total=12):
    """Frames of `total` detections over 2 cameras, `split` of them in camera 0: the same (N, E) for split and
    total - split, different edge_index."""
    import gnncca
    from importlib import import_module
    syn = import_module("gnn-cca_b200.synthetic")
    det = syn.make_detections(2, total // 2, frames, seed=seed, device=dev, max_dist=44.23)
    cam = torch.tensor(([0] * split + [1] * (total - split)) * frames, dtype=torch.int32, device=dev)
    ids = torch.tensor((list(range(split)) + list(range(total - split))) * frames, dtype=torch.int64, device=dev)
    det["id_cam"], det["ids"] = cam, ids
    return gnncca.build_graph_batch(**det)


@pytest.mark.parametrize("variant", ["plain", "bncls", "reattach_max"])
def test_graph_replay_equals_the_eager_step(variant, cuda_device):
    import gnncca
    from importlib import import_module
    params = import_module("gnn-cca_b200.config").graph_net_params()
    if variant == "bncls":
        params["classifier_feats_dict"]["use_batchnorm"] = True
    if variant == "reattach_max":
        params["reattach_initial_nodes"] = params["reattach_initial_edges"] = True
        params["node_agg_fn"] = "max"
    torch.manual_seed(5)
    eager = gnncca.MOTMPNet(copy.deepcopy(params), None, "resnet50", math="fp32").to(cuda_device).train()
    graphed = copy.deepcopy(eager)
    opt_e = torch.optim.SGD(eager.parameters(), lr=1e-2, momentum=0.9)
    opt_g = torch.optim.SGD(graphed.parameters(), lr=1e-2, momentum=0.9)
    step = gnncca.GraphedTrainStep(graphed, pos_weight=2.0)
    seq = [_batch(cuda_device, 5, 1), _batch(cuda_device, 7, 2), _batch(cuda_device, 5, 3), _batch(cuda_device, 4, 4, total=10),
           _batch(cuda_device, 7, 5), _batch(cuda_device, 6, 6, total=10)]
    assert seq[0].edge_index.shape == seq[1].edge_index.shape and not torch.equal(seq[0].edge_index, seq[1].edge_index)
    for i, b in enumerate(seq):
        opt_e.zero_grad(set_to_none=True)
        st = gnncca.loss_stats(eager(b)["classified_edges"], b.edge_labels, pos_weight=2.0)
        st.loss.backward()
        loss_g = step(b)
        assert torch.equal(loss_g, st.loss.detach()), (variant, i)
        assert step.last_stats.counts() == st.counts()
        for (n, pe), (_, pg) in zip(eager.named_parameters(), graphed.named_parameters()):
            assert torch.equal(pe.grad, pg.grad), (variant, i, n)
        for (n, be), (_, bg) in zip(eager.named_buffers(), graphed.named_buffers()):
            assert torch.equal(be, bg), (variant, i, n)
        opt_e.step(); opt_g.step()
    assert len(step.graphs) == 2 and step.replays == len(seq) and step.eager == 0
    # a batch that did not come from build_graph_batch runs eagerly, with the same result
    b = seq[0]
    plain = gnncca.Batch(x=b.x, edge_index=b.edge_index, edge_attr=b.edge_attr, edge_labels=b.edge_labels)
    opt_e.zero_grad(set_to_none=True)
    st = gnncca.loss_stats(eager(b)["classified_edges"], b.edge_labels, pos_weight=2.0)
    st.loss.backward()
    assert torch.equal(step(plain), st.loss.detach()) and step.eager == 1
