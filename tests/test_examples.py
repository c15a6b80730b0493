"""Every shipped example config parses, resolves against the C++ config model and — for the light ones — passes
`det experiment create --local --test` (BASELINE config #1 is the mnist_pytorch one) on CPU."""
import pathlib
import subprocess
import sys

import pytest
import yaml

ROOT = pathlib.Path(__file__).resolve().parent.parent
CONFIGS = sorted((ROOT / "examples").glob("*/*/*.yaml"))


def test_examples_present():
    names = {p.parent.name for p in CONFIGS}
    assert {"mnist_pytorch", "cifar10_pytorch", "resnet50_imagenet_pytorch", "bert_glue_pytorch", "gan_mnist_pytorch",
            "dcgan_pytorch", "fashion_mnist_pytorch", "iris_pytorch", "unets_pytorch", "data_layer_mnist_pytorch",
            "gbt_titanic_pytorch"} <= names


@pytest.mark.parametrize("path", CONFIGS, ids=lambda p: f"{p.parent.name}/{p.name}")
def test_example_config_resolves(path):
    from determined_b200.master import _core

    cfg = yaml.safe_load(path.read_text())
    assert (path.parent / "model_def.py").exists()
    mod, cls = cfg["entrypoint"].split(":")
    assert mod == "model_def" and cls.endswith("Trial")
    resolved = _core.resolve_config(cfg)
    assert resolved["searcher"]["name"] == cfg["searcher"]["name"]
    assert resolved["resources"]["slots_per_trial"] == cfg.get("resources", {}).get("slots_per_trial", 1)
    if cfg["searcher"]["name"] == "adaptive_asha":
        sim = _core.simulate(resolved["searcher"], resolved["hyperparameters"], seed=7)
        assert sim["trials"] == cfg["searcher"]["max_trials"]


@pytest.mark.parametrize("example", ["tutorials/mnist_pytorch", "gan/gan_mnist_pytorch", "computer_vision/cifar10_pytorch"])
def test_example_local_test_mode(example):
    d = ROOT / "examples" / example
    out = subprocess.run([sys.executable, "-m", "determined_b200.cli", "experiment", "create", str(d / "const.yaml"), str(d),
                          "--local", "--test"], cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "Model definition test succeeded" in out.stdout + out.stderr
