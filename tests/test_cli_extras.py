"""CLI pieces that need no cluster: ``experiment create --git`` metadata checks and CSV rendering
(reference cli/determined_cli/experiment.py:124-186, render.py)."""
import pathlib
import subprocess

import pytest

from determined_b200.cli import __main__ as cli


def _git(cwd, *args):
    subprocess.run(["git", "-C", str(cwd), *args], check=True, capture_output=True,
                   env={"GIT_AUTHOR_NAME": "t", "GIT_AUTHOR_EMAIL": "t@example.com", "GIT_COMMITTER_NAME": "t", "GIT_COMMITTER_EMAIL": "t@example.com",
                        "HOME": str(cwd), "PATH": "/usr/bin:/bin:/usr/local/bin"})


def test_git_metadata_requires_a_clean_pushed_tree(tmp_path):
    remote = tmp_path / "remote.git"
    work = tmp_path / "work"
    subprocess.run(["git", "init", "--bare", "-q", str(remote)], check=True)
    subprocess.run(["git", "init", "-q", str(work)], check=True)
    (work / "model_def.py").write_text("x = 1\n")
    with pytest.raises(SystemExit, match="uncommitted"):
        cli.git_metadata(str(work))                      # untracked file
    _git(work, "add", ".")
    _git(work, "commit", "-q", "-m", "first")
    with pytest.raises(SystemExit, match="remote|failed"):
        cli.git_metadata(str(work))                      # no remote configured
    _git(work, "remote", "add", "origin", str(remote))
    with pytest.raises(SystemExit, match="not been pushed"):
        cli.git_metadata(str(work))
    branch = subprocess.run(["git", "-C", str(work), "rev-parse", "--abbrev-ref", "HEAD"], capture_output=True, text=True).stdout.strip()
    _git(work, "push", "-q", "origin", branch)
    md = cli.git_metadata(str(work))
    assert md["git_remote"] == str(remote) and len(md["git_commit"]) == 40 and md["git_committer"] == "t <t@example.com>"
    assert md["git_commit_date"][:2] == "20"
    (work / "model_def.py").write_text("x = 2\n")
    with pytest.raises(SystemExit, match="uncommitted"):
        cli.git_metadata(str(work))


def test_csv_rendering_of_list_output():
    cli.OUTPUT["csv"] = True
    try:
        assert cli._table([[1, "a,b", None], [2, 'q"uote', 3.5]], ["id", "name", "v"]) == 'id,name,v\n1,"a,b",\n2,"q""uote",3.5'
    finally:
        cli.OUTPUT["csv"] = False
    assert "id" in cli._table([[1, "x"]], ["id", "name"])
    args = cli.build_parser().parse_args(["--csv", "experiment", "list"])
    assert args.csv and args.func is cli.experiment_list
    args = cli.build_parser().parse_args(["e", "create", "c.yaml", "dir", "--git"])
    assert args.git
    args = cli.build_parser().parse_args(["tunnel", "7", "--listener", "2222"])
    assert args.task_id == "7" and args.listener == 2222


def test_cli_doc_is_generated_from_the_parser():
    """docs/cli.md is the argparse tree, verb by verb; a new flag or verb without `python scripts/gen_cli_doc.py` fails here."""
    import subprocess
    import sys

    root = pathlib.Path(__file__).resolve().parent.parent
    out = subprocess.run([sys.executable, str(root / "scripts" / "gen_cli_doc.py"), "--check"], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr


def test_sdk_checkpoint_and_model_helpers(tmp_path):
    """Static helpers of the SDK objects (reference common/determined_common/experimental/checkpoint/_checkpoint.py,
    model.py): metadata.json parsing, framework detection, JSON round trips."""
    import json

    from determined_b200.experimental.client import Checkpoint, Model, ModelFramework, Session

    (tmp_path / "metadata.json").write_text(json.dumps({"framework": "torch-2.11.0", "format": "cloudpickle", "trial_id": 3}))
    meta = Checkpoint.parse_metadata(str(tmp_path))
    assert meta["trial_id"] == 3 and Checkpoint.get_type(meta) is ModelFramework.PYTORCH
    assert Checkpoint.get_type({"format": "saved_model"}) is ModelFramework.TENSORFLOW
    assert Checkpoint.get_type({"framework": "tensorflow-1.15"}) is ModelFramework.TENSORFLOW
    with pytest.raises(AssertionError):
        Checkpoint.get_type({"format": "onnx"})
    s = Session("127.0.0.1:1")
    c = Checkpoint.from_json({"uuid": "u-1", "trial_id": 3, "experiment_id": 2, "total_batches": 40, "state": "COMPLETED",
                              "resources": {"state_dict.pth": 10}, "framework": "torch-2.11.0", "format": "cloudpickle"}, s)
    assert (c.uuid, c.trial_id, c.batch_number, c.state.name) == ("u-1", 3, 40, "COMPLETED")
    m = Model.from_json({"name": "xor", "description": "d", "metadata": {"a": 1}}, s)
    assert Model.from_json(m.to_json(), s).to_json() == m.to_json() and m.metadata == {"a": 1}
