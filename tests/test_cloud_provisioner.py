"""AWS / GCP instance providers of the provisioner (reference: master/internal/provisioner/aws.go, aws_spot.go, gcp.go):
the command lines they build, the listings they parse, and an end-to-end scale-up / scale-down through a stand-in `aws`
client whose instances are local det-agent processes."""
import json
import os
import pathlib
import signal
import subprocess
import sys
import time

from determined_b200.experimental.client import Determined
from determined_b200.master import _core

ROOT = pathlib.Path(__file__).resolve().parent.parent
LIB = ROOT / "determined_b200" / "_lib"
MODEL_DIR = str(ROOT / "tests" / "fixtures" / "xor_model_def")
FAKE = ROOT / "tests" / "fixtures" / "fake_cloud"

AWS = {"type": "aws", "region": "us-west-2", "image_id": "ami-123", "instance_type": "p5.48xlarge", "ssh_key_name": "k",
       "security_group_id": "sg-1", "subnet_id": "subnet-9", "iam_instance_profile_arn": "arn:aws:iam::1:instance-profile/det",
       "root_volume_size": 200, "tag_key": "managed-by", "tag_value": "det-42", "agent_command": "/opt/det/det-agent"}
GCP = {"type": "gcp", "project": "proj", "zone": "us-central1-a", "boot_disk_source_image": "projects/p/global/images/det",
       "boot_disk_size": 200, "instance_type": {"machine_type": "a3-highgpu-8g", "gpu_type": "nvidia-h100-80gb", "gpu_num": 8, "preemptible": True},
       "network_interface": {"network": "default", "subnetwork": "sub"}, "label_key": "managed-by", "label_value": "det-42"}


def _opt(argv, flag):
    return argv[argv.index(flag) + 1]


def test_aws_commands():
    a = _core.cloud_command(AWS, "launch", count=3, master_host="10.0.0.5", master_port=8080)["argv"]
    assert a[:6] == ["aws", "--region", "us-west-2", "--output", "json", "ec2"] and a[6] == "run-instances"
    assert _opt(a, "--image-id") == "ami-123" and _opt(a, "--instance-type") == "p5.48xlarge" and _opt(a, "--count") == "3"
    assert _opt(a, "--key-name") == "k" and _opt(a, "--security-group-ids") == "sg-1" and _opt(a, "--subnet-id") == "subnet-9"
    assert _opt(a, "--iam-instance-profile") == "Arn=arn:aws:iam::1:instance-profile/det"
    assert "VolumeSize=200" in _opt(a, "--block-device-mappings") and "--instance-market-options" not in a
    assert "{Key=managed-by,Value=det-42}" in _opt(a, "--tag-specifications")
    ud = _opt(a, "--user-data")
    assert ud.startswith("#!/bin/bash") and "/opt/det/det-agent --master-host 10.0.0.5 --master-port 8080 --agent-id $(curl -s http://169.254.169.254/latest/meta-data/instance-id)" in ud
    spot = _core.cloud_command(dict(AWS, spot=True, spot_max_price="12.5"), "launch", count=1)["argv"]
    assert _opt(spot, "--instance-market-options") == "MarketType=spot,SpotOptions={MaxPrice=12.5,SpotInstanceType=one-time}"
    ls = _core.cloud_command(AWS, "list")["argv"]
    assert "describe-instances" in ls and "Name=tag:managed-by,Values=det-42" in ls
    t = _core.cloud_command(AWS, "terminate", ids=["i-1", "i-2"])["argv"]
    assert t[-4:] == ["terminate-instances", "--instance-ids", "i-1", "i-2"]
    listing = {"Reservations": [{"Instances": [{"InstanceId": "i-1", "State": {"Name": "pending"}}, {"InstanceId": "i-2", "State": {"Name": "running"}}]},
                                {"Instances": [{"InstanceId": "i-3", "State": {"Name": "stopping"}}, {"InstanceId": "i-4", "State": {"Name": "terminated"}}]}]}
    got = _core.cloud_command(AWS, "parse", output=listing)["instances"]
    assert got == [{"id": "i-1", "state": "starting"}, {"id": "i-2", "state": "running"}, {"id": "i-3", "state": "stopped"}]


def test_gcp_commands():
    a = _core.cloud_command(GCP, "launch", ids=["det-agent-1", "det-agent-2"], master_host="10.0.0.5", master_port=8080)["argv"]
    assert a[:4] == ["gcloud", "compute", "instances", "create"] and a[4:6] == ["det-agent-1", "det-agent-2"]
    assert _opt(a, "--machine-type") == "a3-highgpu-8g" and _opt(a, "--accelerator") == "type=nvidia-h100-80gb,count=8"
    assert "--preemptible" in a and _opt(a, "--maintenance-policy") == "TERMINATE"
    assert _opt(a, "--image") == "projects/p/global/images/det" and _opt(a, "--boot-disk-size") == "200GB"
    assert _opt(a, "--network") == "default" and _opt(a, "--subnet") == "sub" and _opt(a, "--labels") == "managed-by=det-42"
    assert "--agent-id $(hostname)" in _opt(a, "--metadata") and _opt(a, "--project") == "proj" and _opt(a, "--zone") == "us-central1-a"
    ls = _core.cloud_command(GCP, "list")["argv"]
    assert ls[3] == "list" and _opt(ls, "--filter") == "labels.managed-by=det-42" and _opt(ls, "--format") == "json"
    t = _core.cloud_command(GCP, "terminate", ids=["det-agent-1"])["argv"]
    assert t[3:5] == ["delete", "det-agent-1"] and "--quiet" in t and _opt(t, "--zone") == "us-central1-a"
    got = _core.cloud_command(GCP, "parse", output=[{"name": "a", "status": "STAGING"}, {"name": "b", "status": "RUNNING"}, {"name": "c", "status": "TERMINATED"}])
    assert got["instances"] == [{"id": "a", "state": "starting"}, {"id": "b", "state": "running"}, {"id": "c", "state": "stopped"}]


def test_aws_provider_scales_up_and_down(tmp_path):
    port_file = tmp_path / "port"
    state = tmp_path / "aws_state.json"
    env = dict(os.environ, DET_CONFIG_DIR=str(tmp_path / "cfg"), FAKE_AWS_STATE=str(state), FAKE_AWS_AGENT_BIN=str(LIB / "det-agent"),
               FAKE_AWS_SLOTS="1")
    prov = {"max_instances": 2, "slots_per_instance": 1, "max_idle_agent_period": 2, "max_agent_starting_period": 30, "tick_ms": 300,
            "provider": dict(AWS, cli=str(FAKE / "aws"), spot=True)}
    proc = subprocess.Popen([str(LIB / "det-master"), "--port", "0", "--data-dir", str(tmp_path / "data"), "--slot-type", "none", "--repo-root", str(ROOT),
                             "--python", sys.executable, "--print-port-file", str(port_file), "--provisioner", json.dumps(prov)],
                            stdout=open(tmp_path / "master.log", "w"), stderr=subprocess.STDOUT, env=env, start_new_session=True)
    try:
        for _ in range(200):
            if port_file.exists() and port_file.read_text().strip():
                break
            time.sleep(0.05)
        port = int(port_file.read_text())
        os.environ["DET_CONFIG_DIR"] = str(tmp_path / "cfg")
        d = Determined(f"127.0.0.1:{port}")
        cfg = {"entrypoint": "model_def:XorTrial", "hyperparameters": {"global_batch_size": 8, "hidden_size": 4, "learning_rate": 0.5},
               "searcher": {"name": "single", "metric": "validation_loss", "max_length": {"batches": 20}}, "scheduling_unit": 10,
               "checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ckpts")}}
        exp = d.create_experiment(cfg, MODEL_DIR)
        assert exp.wait(timeout=180) == "COMPLETED"
        s = d._session.get("/api/v1/provisioner")
        assert s["launched_total"] == 1, s
        calls = json.load(open(state))["calls"]
        run = [c for c in calls if "run-instances" in c]
        assert len(run) == 1 and "--instance-market-options" in run[0] and _opt(run[0], "--count") == "1"
        for _ in range(100):            # idle for > 2 s → terminated through the CLI
            s = d._session.get("/api/v1/provisioner")
            if s["terminated_total"] >= 1 and not s["instances"]:
                break
            time.sleep(0.2)
        assert s["terminated_total"] == 1 and s["instances"] == []
        st = json.load(open(state))
        assert st["instances"] == {} and any("terminate-instances" in c for c in st["calls"])
    finally:
        os.killpg(proc.pid, signal.SIGTERM)
        try:
            proc.wait(timeout=5)
        except Exception:
            os.killpg(proc.pid, signal.SIGKILL)
        try:
            for rec in json.load(open(state))["instances"].values():
                os.killpg(rec["pid"], signal.SIGKILL)
        except Exception:
            pass
