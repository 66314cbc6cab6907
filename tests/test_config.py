"""The mini-Hydra loader (neural_cg/utils/config.py) on the reference's YAML files kept verbatim under
learningsparsepreconditioner4gpu_b200/config/, and the Hydra-style infer entry point.  Expected values: SURVEY.md Appendix A
(the resolved default configuration, config/*.yaml of the reference) and the launch lines of training/*.sh."""
import filecmp
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from learningsparsepreconditioner4gpu_b200.neural_cg.utils import config as hc  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def test_yaml_files_are_the_references_verbatim():
    ref = "/root/reference/config"
    if not os.path.isdir(ref):
        pytest.skip("reference tree not present on this box")
    names = sorted(f for f in os.listdir(ref) if f.endswith(".yaml"))
    assert names == sorted(os.listdir(hc.CONFIG_DIR))
    for f in names:
        assert filecmp.cmp(os.path.join(ref, f), os.path.join(hc.CONFIG_DIR, f), shallow=False), f


def test_default_composition_matches_survey_appendix_a():
    cfg = hc.compose("basic", ["exp_name=heat"])
    assert cfg.gnn_features == 16 and cfg.gnn_mlp_layers == 2
    for k in ("node_encoder", "edge_encoder", "node_decoder", "edge_decoder"):
        assert cfg.gnn[k].to_container() == {"pre_norm": "none", "hidden_channels": 16, "num_layers": 2}
    for k in ("node_mlp", "edge_mlp", "msg_mlp"):
        assert cfg.gnn[k].to_container() == {"pre_norm": "layer", "hidden_channels": 16, "num_layers": 2}
    g = cfg.gnn
    assert (g.num_mp_layers, g.node_residual, g.edge_residual, g.node_features, g.edge_features, g.msg_norm, g.aggr) == \
        (4, True, True, 16, 16, True, "add")
    assert cfg.epsilon == 3.0e-3 and cfg.batch_size == 4 and cfg.seed == 42 and cfg.check_converge is True
    assert cfg.check_methods == ["none", "ic", "ainv", "fsai", "diagonal"] and cfg.check_devices == ["cpu"]
    d = cfg.data
    assert d.prefix == "generated/heat" and d.is_fixed_topology is True and d.block_size == 1 and d.normalize_matrix == "mean"
    assert d.use_edge_features_as_node_feature == "disable" and d.use_random_rhs is True
    assert cfg.loss.name == "RelativeL2Loss_ANorm" and cfg.loss.params is None and cfg.batch_less is False
    assert cfg.optimizer.name == "adamw" and cfg.optimizer.params.lr == 1e-3 and isinstance(cfg.optimizer.params.lr, float)
    assert cfg.scheduler.params.gamma == 0.99 and cfg.trainer.max_epochs == 500 and cfg.test_max_iter == 1000
    assert cfg.split.train == 0.8 and cfg.checkpoint.every_n_epochs == 1
    # `_self_` last: basic.yaml's batch_size / epsilon replace trainer.yaml's `???`
    assert cfg.get("workspace", "simple") == "simple" and cfg.get("rtol") is None and "rtol" not in cfg


def test_overrides_and_errors():
    cfg = hc.compose("basic", ["exp_name=synthetic", "data.is_fixed_topology=false", "data.has_shared_features=false",
                               "data.use_node_features=false", "data.use_edge_features_as_node_feature=mean",
                               "+rtol=1e-8", "+workspace=scaled", "++epsilon=1e-2", "~pretrained", "gnn_features=32"])
    assert cfg.data.is_fixed_topology is False and cfg.data.use_edge_features_as_node_feature == "mean"   # training/synthetic.sh
    assert cfg.rtol == 1e-8 and cfg.workspace == "scaled" and cfg.epsilon == 1e-2 and "pretrained" not in cfg
    assert cfg.gnn.msg_mlp.hidden_channels == 32                      # interpolation follows the override
    with pytest.raises(hc.MissingMandatoryValue):
        hc.compose("basic").exp_name
    with pytest.raises(hc.MissingMandatoryValue):
        hc.compose("basic").data.prefix                               # generated/${exp_name}
    with pytest.raises(hc.ConfigKeyError):
        hc.compose("basic", ["rtol=1e-6"])                            # new keys need '+'
    with pytest.raises(hc.ConfigError):
        hc.compose("basic", ["+epsilon=1"])                           # '+' refuses existing keys
    with pytest.raises(hc.ConfigError):
        hc.compose("nope")
    assert hc.parse_cli(["--config-name=basic_multidata", "exp_name=x", "+rtol=1"]) == ("basic_multidata", None, ["exp_name=x", "+rtol=1"])
    assert "all_prefix" in hc.compose("basic_multidata", ["exp_name=x"]).data


def test_config_splats_into_the_workspace_ctor():
    """train.py:56-61 / infer.py:231-236: workspace_class(edge_features=..., node_features=..., block_size=..., **cfg)."""
    from learningsparsepreconditioner4gpu_b200.neural_cg.workspace import SimpleTrainingWorkspace
    from test_host_logic import make_workspace

    cfg = hc.compose("basic", ["exp_name=heat"])
    torch.manual_seed(42)
    ws = SimpleTrainingWorkspace(edge_features=1, node_features=5, block_size=1, **hc.workspace_kwargs(cfg))
    ref = make_workspace(5, 1, 1)                                       # the hand-written default config, seed 42
    a, b = ws.state_dict(), ref.state_dict()
    assert list(a) == list(b) and all(torch.equal(a[k], b[k]) for k in a)
    assert ws.epsilon == 3.0e-3 and ws.check_methods == ["none", "ic", "ainv", "fsai", "diagonal"]


def test_datamodule_split_and_loaders():
    from learningsparsepreconditioner4gpu_b200.neural_cg.utils.datamodule import FolderDataModule

    cfg = hc.compose("basic", ["exp_name=heat_fixed", "data.prefix=" + os.path.join(GOLD, "folders", "heat_fixed"),
                               "data.use_random_rhs=false"])
    dm = FolderDataModule(cfg.data, cfg.split, cfg.batch_size)
    n = len(dm.dataset)
    assert sorted(dm.train_idx + dm.val_idx) == list(range(n)) and len(dm.train_idx) == int(0.8 * n)
    from sklearn.model_selection import train_test_split

    tr, va = train_test_split(range(n), train_size=0.8, random_state=42, shuffle=True)      # utils/datamodule.py:25-27
    assert dm.train_idx == list(tr) and dm.val_idx == list(va)
    val = list(dm.val_dataloader())
    assert len(val) == len(va) and all(int(b.ptr.numel()) == 2 for b in val)
    tb = next(iter(dm.train_dataloader()))
    assert int(tb.ptr.numel()) == min(4, len(tr)) + 1 and int(tb.ptr[-1]) == tb.x.shape[0]
    assert len(list(dm.inference_dataloader())) == n


@pytest.mark.gpu
def test_hydra_style_infer_entry_point(tmp_path):
    from learningsparsepreconditioner4gpu_b200 import infer

    stats = infer.main(["exp_name=heat_fixed", "data.prefix=" + os.path.join(GOLD, "folders", "heat_fixed"),
                        "data.use_random_rhs=false", "+rtol=1e-6", "+warmup=1", "+repeat=1", "+dataloader=infer",
                        "+out_dir=" + str(tmp_path), "+infer_prefix=t_"])
    assert set(stats.stat_dict) == {"PCG-none-cuda", "PCG-diagonal-cuda", "Neural+CUDA"}
    assert (tmp_path / "infer_t_heat_fixed_6.csv").exists() and (tmp_path / "all_infer_t_heat_fixed_6.csv").exists()
