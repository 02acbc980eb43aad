"""CPU: the host side of the drop-in - module tree / state-dict names, factory behaviour, composed (autograd) path
against the oracle, dispatch rules, and that the C-ABI library loads and exports everything include/cpn_b200.h declares."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

import cases
import oracle
from count_pipnet_b200 import _lib, engine, synth
from count_pipnet_b200.features import convnext_features as cf
from count_pipnet_b200.pipnet import count_pipnet as my_count
from count_pipnet_b200.pipnet import pipnet as my_pipnet

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_symbols_exported():
    header = open(os.path.join(ROOT, "include", "cpn_b200.h")).read()
    declared = set(re.findall(r"\b(cpn_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 20
    lib = ctypes.CDLL(_lib.LIB_PATH)
    missing = [s for s in sorted(declared) if not hasattr(lib, s)]
    assert not missing, missing
    assert set(_lib.SIGNATURES) | set(_lib.OTHER_SIGNATURES) == declared


def test_library_loads_and_reports_errors_without_gpu():
    lib = _lib.load()
    assert lib.cpn_version() >= 100
    h = ctypes.c_void_p()
    assert lib.cpn_plan_create(ctypes.byref(h)) == 0 and lib.cpn_plan_num_ops(h) == 0
    assert lib.cpn_plan_destroy(h) == 0
    # argument validation happens before any CUDA call
    rc = lib.cpn_gemm_bf16(None, None, None, None, None, 128, 100, 64, None, None, None, None, None, None, None, 0, 0,
                           None)
    assert rc != 0 and b"multiple of 32" in lib.cpn_last_error()


def test_state_dict_names_convnext():
    model, p = my_pipnet.get_pipnet(200, synth.make_args())
    keys = list(model.state_dict().keys())
    assert p == 768
    for k in ["_net.features.0.0.weight", "_net.features.0.1.bias", "_net.features.1.0.layer_scale",
              "_net.features.1.2.block.0.weight", "_net.features.5.8.block.5.bias", "_net.features.6.1.weight",
              "_net.features.7.2.block.3.weight", "_classification.weight",
              "_classification.normalization_multiplier", "_multiplier"]:
        assert k in keys, k
    assert model.state_dict()["_net.features.7.2.block.3.weight"].shape == (3072, 768)
    # stride surgery of the "26" variant: 96->192 keeps stride 2, the two later downsample convs run at stride 1
    strides = [m.stride for m in model._net.modules() if isinstance(m, torch.nn.Conv2d) and m.kernel_size == (2, 2)]
    assert strides == [(2, 2), (1, 1), (1, 1)]
    m13 = cf.convnext_tiny_13_features()
    strides = [m.stride for m in m13.modules() if isinstance(m, torch.nn.Conv2d) and m.kernel_size == (2, 2)]
    assert strides == [(2, 2), (2, 2), (1, 1)]


@pytest.mark.parametrize("num_stages,channels,grid192", [(1, 96, 48), (2, 192, 24), (3, 192, 24), (4, 384, 23),
                                                          (5, 384, 23), (6, 768, 22), (7, 768, 22)])
def test_mid_layer_shapes(num_stages, channels, grid192):
    """Per-stage shape table of the reference (tests/test-convnext-backbone.py)."""
    shape = cf.get_feature_dimensions(use_mid_layers=True, num_stages=num_stages, input_size=192)
    assert tuple(shape) == (1, channels, grid192, grid192)


def test_count_network_factory_and_errors():
    args = synth.make_args(num_features=16, use_mid_layers=True, num_stages=3, intermediate_layer="bilinear",
                           activation="gumbel_softmax")
    model, p = my_count.get_count_network(9, args, max_count=3, use_ste=True)
    assert p == 16 and model._classification.weight.shape == (9, 48) and hasattr(model, "_max_count")
    assert "_intermediate.embed.weight" in model.state_dict()
    with pytest.raises(ValueError):
        my_count.get_count_network(9, synth.make_args(net="resnet50"))
    with pytest.raises(ValueError):
        my_count.get_count_network(9, synth.make_args(intermediate_layer="nope"))
    with pytest.raises(Exception):
        my_pipnet.get_pipnet(9, synth.make_args(net="vgg"))
    model.update_temperature(0.3)
    assert model._add_on[-1].tau == 0.3
    onehot, _ = my_count.get_count_network(9, synth.make_args(num_features=16, use_mid_layers=True, num_stages=3))
    assert onehot.get_prototype_importance_per_class(0).shape == (9,)


@pytest.mark.parametrize("name", ["count_linear_softmax", "count_identity_softmax_noste",
                                  "count_linear_full_softmax_train_outputs", "pipnet_convnext13_addon32_bias",
                                  "pipnet_resnet18_64"])
def test_composed_path_equals_oracle(name):
    """With autograd enabled the modules run composed PyTorch ops; that path must equal the reference arithmetic."""
    case = cases.CASES[name]
    model = cases.build(case, my_pipnet, my_count)
    x = cases.inputs(case)
    sd = {k: v.detach() for k, v in model.state_dict().items()}
    with torch.enable_grad():
        proto, pooled, out = model(x, inference=case["inference"])
    assert out.requires_grad
    o_proto, o_pooled, o_out = oracle.forward(sd, cases.oracle_cfg(case), x, inference=case["inference"])
    assert torch.allclose(proto, o_proto, atol=1e-6) and torch.allclose(pooled, o_pooled, atol=1e-5)
    assert torch.allclose(out, o_out, rtol=1e-5, atol=1e-5)


def test_inference_on_cpu_raises_no_fallback():
    model, _ = my_pipnet.get_pipnet(5, synth.make_args(use_mid_layers=True, num_stages=1))
    model.eval()
    with torch.no_grad(), pytest.raises(_lib.CpnError):
        model(torch.zeros(1, 3, 32, 32), inference=True)
    model.train()
    assert engine.use_fused(model, torch.zeros(1, 3, 32, 32)) is False


def test_synth_weights_are_deterministic():
    model, _ = my_pipnet.get_pipnet(7, synth.make_args(use_mid_layers=True, num_stages=2))
    a = synth.synth_state_dict(model.state_dict(), seed=3)
    b = synth.synth_state_dict(model.state_dict(), seed=3)
    assert all(torch.equal(a[k], b[k]) for k in a)
    w = a["_classification.weight"]
    assert ((w > 0).sum(1) == 8).all() and float(a["_multiplier"]) == 2.0


# ---------------------------------------------------------------------------------------------- checkpoint compatibility
def _state_key_table():
    import json
    return json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "state_dict_keys.json")))


@pytest.mark.parametrize("name", list(cases.CASES))
def test_state_dict_matches_reference_models(name):
    """Ordered keys, shapes and dtypes equal those of the REAL reference model of the same configuration
    (tests/golden/state_dict_keys.json, written by make_golden_state_keys.py from /root/reference): a reference
    checkpoint loads with strict=True (util/checkpoint_manager.py:66)."""
    ref = _state_key_table()[name]
    mine = [[k, list(v.shape), str(v.dtype)] for k, v in cases.build(cases.CASES[name], my_pipnet, my_count).state_dict().items()]
    assert [r[0] for r in mine] == [r[0] for r in ref]
    assert mine == ref


def test_load_reference_checkpoint_prefixes(tmp_path):
    from count_pipnet_b200 import checkpoint as ck
    case = cases.CASES["count_linear_softmax"]
    src = cases.build(case, my_pipnet, my_count)
    with torch.no_grad():
        for p in src.parameters():
            p.add_(0.25)
    wrapped_state = {"module." + k: v.clone() for k, v in src.state_dict().items()}  # as nn.DataParallel saves it
    path = tmp_path / "net_trained"
    torch.save({"model_state_dict": wrapped_state, "epoch": 3, "accuracy": 0.5}, path)
    # bare model <- DataParallel checkpoint file
    dst = cases.build(case, my_pipnet, my_count)
    res = ck.load_reference_checkpoint(dst, str(path))
    assert not res.missing_keys and not res.unexpected_keys
    for (k, a), (_, b) in zip(src.state_dict().items(), dst.state_dict().items()):
        assert torch.equal(a, b), k
    # wrapped model <- bare state dict
    dst2 = torch.nn.DataParallel(cases.build(case, my_pipnet, my_count))
    ck.load_reference_checkpoint(dst2, src.state_dict())
    assert torch.equal(dst2.module._classification.weight, src._classification.weight)
    # mismatches raise as in the reference (strict=True)
    broken = dict(wrapped_state)
    broken.pop("module._classification.weight")
    with pytest.raises(RuntimeError):
        ck.load_reference_checkpoint(cases.build(case, my_pipnet, my_count), {"model_state_dict": broken})
    with pytest.raises(ValueError):
        ck.load_reference_checkpoint(dst, {"epoch": 1})


def test_bench_configs_match_parity_cases():
    """bench.py's workloads (count_pipnet_b200/configs.py) are the BASELINE.json configs the parity cases pin."""
    import cases
    from count_pipnet_b200 import configs
    pairs = {"cfg1": "cfg1_count_shapes_onehot_gumbel", "cfg2": "cfg2_pipnet_convnext26_224",
             "cfg3": "cfg3_pipnet_resnet50_224", "cfg4": "cfg4_count_bilinear_gumbel",
             "cfg5": "cfg5_pipnet_convnext26_448_map"}
    for name, case_name in pairs.items():
        cfg, case = configs.CONFIGS[name], cases.CASES[case_name]
        for key in ("model", "classes", "size", "args"):
            assert cfg[key] == case[key], (name, key)
        assert cfg.get("max_count", 3) == case.get("max_count", 3)
        assert configs.oracle_cfg(name) == cases.oracle_cfg(case)
    assert configs.batch_per_gpu("cfg4", 1) == 64 and configs.batch_per_gpu("cfg4", 4) == 128
    assert configs.batch_per_gpu("cfg3", 8) == 32 and configs.batch_per_gpu("cfg5", 8) == 16


def test_multi_device_dataparallel_replica_is_rejected():
    """A DataParallel replica (torch marks it `_is_replica`) cannot keep packed weights or plans: clear error."""
    import torch
    from count_pipnet_b200 import _lib, engine
    model, _ = my_pipnet.get_pipnet(10, synth.make_args(use_mid_layers=True, num_stages=2))
    model._is_replica = True
    with pytest.raises(_lib.CpnError, match="DataParallel"):
        engine.state_for(model, torch.device("cuda", 0))
