"""resnet_18() with the reference's signature (model/resnet.py:152-351), executed by libxvector.

Modified ResNet-18 on [batch, length, dim] log-mel features: conv0_1 3x3 stem, four stages of conv_block /
identity_block (3x3, SAME, no bias; 64/128/256/512 channels; frequency stride 2 and optional time stride from the
second stage on), conv5 (1 x F/8) VALID, dense1, dense2, pooling, tdnn6, tdnn7.  Only `pooling`, `tdnn6_*` and
`tdnn7_*` are published as endpoints, exactly as in the reference (model/resnet.py:303-349).

The reference asserts a 40-dim input (model/resnet.py:189-190); this build accepts any multiple of 8 because
BASELINE.json config 4 also names 64-dim fbank (the assertion is kept when `strict_dim=True`)."""
from .tdnn import _inject_defaults, _run_network


def endpoint_names(params):
    names = ["pooling", "tdnn6_dense", "tdnn6_bn", "tdnn6_relu", "tdnn7_dense"]
    if not params.last_layer_no_bn:
        names.append("tdnn7_bn")
    if not params.last_layer_linear:
        names.append("tdnn7_relu")
    return names


def resnet_18(features, params, is_training=None, reuse_variables=None, aux_features=None, strict_dim=False):
    """Build (= run) a ResNet-18.  Returns (output of the last layer, endpoints) like the reference."""
    if is_training:
        raise NotImplementedError("resnet_18(is_training=True): the training graph is outside this build's scope")
    if strict_dim:
        assert features.shape[-1] == 40                      # model/resnet.py:190
    if "resnet_blocks" not in params.dict:                   # model/resnet.py:203-204
        params.dict["resnet_blocks"] = [2, 2, 2, 2]
    _inject_defaults(params, pooling_layer=False)
    return _run_network(features, params, "resnet_18", endpoint_names(params), "tdnn7")
