"""gRPC API (determined.api.v1.Determined) through the gateway in front of a real det-master, and the generated contract
files (reference: proto/src/determined/api/v1/api.proto, master/internal/grpc/api.go, webui/api-ts-sdk)."""
import os
import pathlib
import subprocess
import sys
import time

import pytest

grpc = pytest.importorskip("grpc")

from determined_b200.master import api_spec, grpc_gateway  # noqa: E402

ROOT = pathlib.Path(__file__).resolve().parent.parent
LIB = ROOT / "determined_b200" / "_lib"
MODEL_DIR = ROOT / "tests" / "fixtures" / "xor_model_def"


@pytest.fixture(scope="module")
def stack(tmp_path_factory):
    from determined_b200 import _build

    _build.build_core()
    tmp = tmp_path_factory.mktemp("grpc")
    port_file = tmp / "port"
    master = subprocess.Popen([str(LIB / "det-master"), "--port", "0", "--data-dir", str(tmp / "data"), "--artificial-slots", "2",
                               "--repo-root", str(ROOT), "--python", sys.executable, "--print-port-file", str(port_file)],
                              stdout=open(tmp / "master.log", "w"), stderr=subprocess.STDOUT, env=dict(os.environ, DET_CONFIG_DIR=str(tmp / "cfg")))
    for _ in range(100):
        if port_file.exists() and port_file.read_text().strip():
            break
        time.sleep(0.05)
    server, gport = grpc_gateway.serve(f"127.0.0.1:{int(port_file.read_text())}")
    yield {"client": grpc_gateway.Client(f"127.0.0.1:{gport}"), "tmp": tmp}
    server.stop(0)
    master.terminate()
    try:
        master.wait(timeout=5)
    except Exception:
        master.kill()


def test_contract_files_are_generated_from_the_table():
    out = subprocess.run([sys.executable, str(ROOT / "scripts" / "gen_api.py"), "--check"], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout
    proto = (ROOT / "proto" / "determined" / "api" / "v1" / "api.proto").read_text()
    assert proto.count("  rpc ") == len(api_spec.RPCS) >= 60 and "service Determined" in proto
    assert "returns (stream TrialLogsResponse)" in proto
    ts = (ROOT / "webui" / "api-ts-sdk" / "api.ts").read_text()
    assert "export class DeterminedApi" in ts and "async getExperiment(req: GetExperimentRequest)" in ts


def test_master_cluster_templates_and_registry_over_grpc(stack):
    c = stack["client"]
    m = c.GetMaster().master
    assert m.version.startswith("0.13") and m.cluster_id and m.raw["scheduler"] == m.scheduler
    agents = c.GetAgents().agents
    assert agents[0].id == "agent-local" and agents[0].num_slots == 2
    assert len(c.GetSlots().slots) == 2
    assert c.GetAgent(agent_id="agent-local").agent.num_slots == 2
    slot = c.GetSlot(agent_id="agent-local", slot_id="0").slot
    assert slot.agent_id == "agent-local" and slot.enabled
    c.DisableAgent(agent_id="agent-local")
    assert c.GetAgent(agent_id="agent-local").agent.raw["enabled"] is False
    c.EnableAgent(agent_id="agent-local")
    assert c.GetAgent(agent_id="agent-local").agent.raw["enabled"] is True
    with pytest.raises(grpc.RpcError) as e:
        c.GetAgent(agent_id="nope")
    assert e.value.code() == grpc.StatusCode.NOT_FOUND
    assert c.GetUser(name="determined").user.username == "determined"
    assert c.GetTelemetry().telemetry["enabled"] is False
    c.PutTemplate(name="small", config={"scheduling_unit": 7})
    assert c.GetTemplate(name="small").template.config["scheduling_unit"] == 7
    assert [t.name for t in c.GetTemplates().templates] == ["small"]
    c.PostModel(name="xor", description="demo", metadata={"k": "v"})
    assert c.GetModel(name="xor").model.description == "demo"
    sim = c.PreviewHPSearch(searcher={"name": "random", "metric": "m", "max_trials": 3, "max_length": {"batches": 5}},
                            hyperparameters={"global_batch_size": 1}).simulation
    assert dict(sim["results"]) == {"5B V": 3.0}
    token = c.Login(username="determined", password="").session["token"]
    assert token
    with pytest.raises(grpc.RpcError) as e:
        c.GetExperiment(id=4242)
    assert e.value.code() == grpc.StatusCode.NOT_FOUND
    with pytest.raises(grpc.RpcError) as e:
        c.CreateExperiment(config={"entrypoint": "m:T", "hyperparameters": {}, "searcher": {"name": "single", "metric": "x"}})
    assert e.value.code() == grpc.StatusCode.INVALID_ARGUMENT and "global_batch_size" in e.value.details()


def test_experiment_lifecycle_and_log_stream_over_grpc(stack):
    c = stack["client"]
    files = [{"path": p.name, "type": "file", "content": p.read_text(), "mode": 0o644}
             for p in sorted(MODEL_DIR.iterdir()) if p.is_file()]
    cfg = {"entrypoint": "model_def:XorTrial", "description": "grpc", "scheduling_unit": 10, "max_restarts": 0,
           "hyperparameters": {"global_batch_size": 8, "hidden_size": 4, "learning_rate": 0.5},
           "searcher": {"name": "single", "metric": "validation_loss", "max_length": {"batches": 20}},
           "checkpoint_storage": {"type": "shared_fs", "host_path": str(stack["tmp"] / "ckpts")}}
    exp = c.CreateExperiment(config=cfg, model_definition=files, activate=True).experiment
    assert exp.id >= 1
    deadline = time.time() + 180
    while time.time() < deadline:
        e = c.GetExperiment(id=exp.id).experiment
        if e.state in ("COMPLETED", "ERROR", "CANCELED"):
            break
        time.sleep(0.5)
    assert e.state == "COMPLETED" and e.progress == 1.0 and len(e.trials) == 1
    t = c.GetTrial(id=e.trials[0].id, metrics=True).trial
    assert t.total_batches_processed == 20 and len(t.steps) == 2 and t.hparams["hidden_size"] == 4
    cks = c.GetExperimentCheckpoints(id=exp.id).checkpoints
    assert cks and c.GetCheckpoint(uuid=cks[-1].uuid).checkpoint.trial_id == t.id
    c.PostCheckpointMetadata(uuid=cks[-1].uuid, metadata={"note": "via grpc"})
    assert c.GetCheckpoint(uuid=cks[-1].uuid).checkpoint.metadata["note"] == "via grpc"
    batches = list(c.TrialLogs(id=t.id, follow=True))
    lines = [l.message for b in batches for l in b.entries]
    assert lines and any("checkpoint" in l.lower() for l in lines)
    only = [l.message for b in c.TrialLogs(id=t.id, search="checkpoint") for l in b.entries]
    assert only and all("checkpoint" in l for l in only) and len(only) < len(lines)
    # per-experiment / per-trial views
    assert [x.id for x in c.GetExperimentTrials(id=exp.id).trials] == [t.id]
    assert [x.uuid for x in c.GetTrialCheckpoints(id=t.id).checkpoints] == [x.uuid for x in cks]
    hist = list(c.GetExperimentValidationHistory(id=exp.id).result["validation_history"])
    assert hist and hist[0]["trial_id"] == t.id and all(a["searcher_metric"] > b["searcher_metric"] for a, b in zip(hist, hist[1:]))
    fields = c.TrialLogsFields(id=t.id).fields
    assert list(fields["agent_ids"]) == ["agent-local"] and "stdout" in list(fields["stdtypes"]) and len(fields["container_ids"]) == 1
    c.PatchExperiment(id=exp.id, labels=["grpc", "demo"])
    labels = {x["label"]: x["count"] for x in c.GetExperimentLabels().result["labels"]}
    assert labels.get("grpc") == 1 and labels.get("demo") == 1
    c.PostModel(name="xor-grpc", description="", metadata={})
    c.PostModelVersion(name="xor-grpc", checkpoint_uuid=cks[-1].uuid)
    ver = c.GetModelVersion(name="xor-grpc", version=1).version_info
    assert ver["checkpoint_uuid"] == cks[-1].uuid and ver["model"]["name"] == "xor-grpc"
    # the reference's HP-visualisation RPCs
    names = c.MetricNames(id=exp.id).names
    assert names["searcher_metric"] == "validation_loss" and "loss" in list(names["training_metrics"])
    assert list(c.MetricBatches(id=exp.id, metric_name="loss", metric_type="training").result["batches"]) == [10, 20]
    snap = c.TrialsSnapshot(id=exp.id, batches_processed=20, metric_name="validation_loss", metric_type="validation").result["trials"]
    assert len(snap) == 1 and snap[0]["trial_id"] == t.id and snap[0]["hparams"]["hidden_size"] == 4
    sample = c.TrialsSample(id=exp.id, metric_name="loss", metric_type="training", max_trials=5, max_datapoints=10).result
    assert [p["batches"] for p in sample["trials"][0]["data"]] == [10, 20]
    c.ArchiveExperiment(id=exp.id)
    assert all(x.id != exp.id for x in c.GetExperiments().experiments)
    assert any(x.id == exp.id for x in c.GetExperiments(archived=True).experiments)
