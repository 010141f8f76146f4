"""Host-side mirror of the reference module API (models/mlp.py, models/mpn.py) -- no GPU needed."""
import copy
import importlib
import json
import os

import pytest
import torch

from conftest import golden_cases, load_golden

gnncca = importlib.import_module("gnn-cca_b200")
cfgmod = importlib.import_module("gnn-cca_b200.config")
engine = importlib.import_module("gnn-cca_b200.engine")


@pytest.mark.parametrize("case", golden_cases("mpn_"))
def test_state_dict_keys_and_shapes_equal_the_reference(case):
    """utils.load_pretrained_weights matches on key AND size and then loads strictly (utils.py:482-493): the drop-in
    must expose exactly the reference's parameter tree (golden files hold the reference's own state_dict)."""
    z = load_golden("mpn_" + case)
    params = json.loads(str(z["config"]))
    model = gnncca.MOTMPNet(copy.deepcopy(params), None, str(z["arch"]))
    ref = {k[3:]: v for k, v in z.items() if k.startswith("sd/")}
    sd = model.state_dict()
    assert list(sd.keys()) == list(ref.keys())
    for k in ref:
        assert tuple(sd[k].shape) == tuple(ref[k].shape), k
    model.load_state_dict({k: torch.from_numpy(v) for k, v in ref.items()}, strict=True)


def test_constructor_mutates_the_config_like_the_reference():
    params = cfgmod.graph_net_params()
    gnncca.MOTMPNet(params, None, "resnet50")
    assert params["encoder_feats_dict"]["edges"]["node_in_dim"] == 2048      # mpn.py:167-170 merges the node dict in place
    with pytest.raises(AssertionError):
        bad = cfgmod.graph_net_params(); bad["node_agg_fn"] = "median"
        gnncca.MOTMPNet(bad, None, "resnet50")
    with pytest.raises(AssertionError):
        gnncca.MLP(4, 6)                                                        # mlp.py:8: fc_dims must be a list / tuple


def test_seeded_construction_draws_the_same_weights_as_the_reference():
    from oracle import ref_shim
    if not ref_shim.reference_available():
        pytest.skip("/root/reference not present")
    mpn = ref_shim.import_reference("models.mpn")
    for bn in (False, True):
        params = cfgmod.graph_net_params(classifier_batchnorm=bn)
        torch.manual_seed(7)
        ref = mpn.MOTMPNet(copy.deepcopy(params), None, "resnet50")
        after_ref = torch.rand(1)
        torch.manual_seed(7)
        ours = gnncca.MOTMPNet(copy.deepcopy(params), None, "resnet50")
        after_ours = torch.rand(1)
        for (ka, va), (kb, vb) in zip(ref.state_dict().items(), ours.state_dict().items()):
            assert ka == kb and torch.equal(va, vb)
        assert torch.equal(after_ref, after_ours)                               # the RNG stream continues identically


def test_mlp_slot_numbering():
    mlp = importlib.import_module("gnn-cca_b200.mlp")
    specs = mlp.plan_layers(6, [4, 1], 0, True)        # Linear BN ReLU Dropout Linear -> fc_layers.0, .1, .4
    assert [s.linear for s in specs] == [0, 4] and specs[0].bn == 1 and specs[1].bn is None and not specs[1].relu
    specs = mlp.plan_layers(2048, [128, 32], None, False)
    assert [s.linear for s in specs] == [0, 2]


def test_cpu_tensors_are_rejected_not_computed():
    model = gnncca.MOTMPNet(cfgmod.graph_net_params(), None, "bdnet_market").eval()

    class Dd:
        x, edge_index, edge_attr = torch.zeros(4, 512), torch.tensor([[0, 1], [1, 0]]), torch.zeros(2, 4)
    with pytest.raises(RuntimeError):
        with torch.no_grad():
            model(Dd())
    with pytest.raises(RuntimeError):
        gnncca.build_graph_batch(torch.tensor([0, 2]), torch.tensor([0, 1]), torch.tensor([0, 0]), torch.zeros(2, dtype=torch.float64),
                                 torch.zeros(2, dtype=torch.float64), torch.ones(1, dtype=torch.float64), torch.zeros(2, 8), torch.zeros(2, 16))


def test_frame_capacity_classes():
    model = gnncca.MOTMPNet(cfgmod.graph_net_params(), None, "bdnet_market")
    assert engine.frame_capacity(model, 24576, 442368, 1024, (24, 432)) == (24, 448)     # exact maxima, rounded to 8 / 32
    assert engine.frame_capacity(model, 24576, 442368, 1024) == (32, 1024)                # 1.25 x average -> class
    assert engine.frame_capacity(model, 2048, 917504, 4, (512, 229376)) is None           # cfg 4 frames: per-step kernels
    assert engine.frame_capacity(model, 10, 10, 0) is None


def test_checkpoints_travel_both_ways_through_the_reference_functions(tmp_path):
    """the checkpoint dict of main_training.py:422-433 written from the drop-in is read by the reference's own
    load_pretrained_weights (utils.py:458-507) into the reference's MOTMPNet with every layer matched, and a reference checkpoint
    (also with DataParallel's 'module.' prefix) loads into the drop-in.  Needs /root/reference (build container)."""
    import collections
    import os.path as osp
    import shutil
    import warnings
    import yaml
    from oracle import ref_shim
    if not ref_shim.reference_available():
        pytest.skip("/root/reference not present")
    ns = {"torch": torch, "osp": osp, "warnings": warnings, "OrderedDict": collections.OrderedDict, "shutil": shutil, "yaml": yaml, "os": os}
    ref_shim.reference_def("libs/utils.py", "load_checkpoint", ns)
    load_pretrained = ref_shim.reference_def("libs/utils.py", "load_pretrained_weights", ns)
    save_checkpoint = ref_shim.reference_def("libs/utils.py", "save_checkpoint", ns)
    mpn = ref_shim.import_reference("models.mpn")
    params = cfgmod.graph_net_params(classifier_batchnorm=True)
    torch.manual_seed(11)
    ours = gnncca.MOTMPNet(copy.deepcopy(params), None, "resnet50")
    torch.manual_seed(12)
    ref = mpn.MOTMPNet(copy.deepcopy(params), None, "resnet50")
    os.makedirs(tmp_path / "files")
    opt = torch.optim.SGD(ours.parameters(), lr=0.1)
    save_checkpoint({"epoch": 3, "model_state_dict": ours.state_dict(), "optimizer_state_dict": opt.state_dict(), "prec": 90.0, "prec1": [91.0],
                     "prec0": [89.0], "best_loss": 0.5, "model_parameters": 0.268, "CONFIG": {"DATASET_TRAIN": {"NAME": ["EPFL-Terrace"]}}},
                    False, str(tmp_path), "ours")                                                      # utils.py:406-407
    load_pretrained(ref, str(tmp_path / "files" / "ours_latest.pth.tar"))
    for (ka, va), (kb, vb) in zip(ours.state_dict().items(), ref.state_dict().items()):
        assert ka == kb and torch.equal(va, vb)                                                         # nothing discarded
    torch.manual_seed(13)
    ref2 = mpn.MOTMPNet(copy.deepcopy(params), None, "resnet50")
    torch.save({"model_state_dict": {"module." + k: v for k, v in ref2.state_dict().items()}}, tmp_path / "ref.pth.tar")
    load_pretrained(ours, str(tmp_path / "ref.pth.tar"))
    for (ka, va), (kb, vb) in zip(ref2.state_dict().items(), ours.state_dict().items()):
        assert ka == kb and torch.equal(va, vb)
