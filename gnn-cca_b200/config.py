"""The reference's shipped MPN hyper-parameters as data (facts from its YAML files, not code).

``config/config_training.yaml:94-181`` (no BatchNorm) and ``config/config_inference.yaml:76-163``
(classifier with BatchNorm).  Only the architectures' encoder input width differs (2048 for resnet50 /
resnext101_cars, 512 for the others).  Returned dicts are fresh copies: ``MOTMPNet.__init__`` mutates them.
"""
from __future__ import annotations

import copy

_NODE_IN = {"resnet50": 2048, "bdnet_market": 512, "bdnet_cuhk": 512, "osnet_ms_c_d": 512, "osnet_market": 512,
            "resnext101_cars": 2048}

CONV_TO_M = {"EPFL-Laboratory": 44.23, "EPFL-Terrace": 50.83, "EPFL-Basketball": 153.54}   # config_training.yaml:40-47


def graph_net_params(classifier_batchnorm: bool = False) -> dict:
    nodes = {a: dict(node_in_dim=d, node_fc_dims=[128], node_out_dim=32, dropout_p=0, use_batchnorm=False)
             for a, d in _NODE_IN.items()}
    return copy.deepcopy(dict(
        node_agg_fn="sum", num_enc_steps=4, num_class_steps=3, reattach_initial_nodes=False, reattach_initial_edges=False,
        encoder_feats_dict=dict(edges=dict(edge_in_dim=4, edge_fc_dims=[], edge_out_dim=6), nodes=nodes),
        edge_model_feats_dict=dict(fc_dims=[6], dropout_p=0, use_batchnorm=False),
        node_model_feats_dict=dict(fc_dims=[32], dropout_p=0, use_batchnorm=False),
        classifier_feats_dict=dict(edge_in_dim=6, edge_fc_dims=[4], edge_out_dim=1, dropout_p=0,
                                   use_batchnorm=bool(classifier_batchnorm))))
