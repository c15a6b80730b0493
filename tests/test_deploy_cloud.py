"""det-deploy aws / gcp (reference: deploy/determined_deploy/aws, gcp): generated CloudFormation / Terraform documents and
the commands that apply them."""
import json
import subprocess
import sys
import pathlib

import yaml

from determined_b200.deploy import cloud

ROOT = pathlib.Path(__file__).resolve().parent.parent


def _master_provisioner(userdata: str):
    start = userdata.index("--provisioner '") + len("--provisioner '")
    return json.loads(userdata[start:userdata.index("'", start)])


def test_aws_templates():
    t = cloud.aws_template("c1", "key", "simple", spot=True, spot_max_price="9.5", max_agents=3)
    r = t["Resources"]
    assert {"SecurityGroup", "MasterRole", "MasterProfile", "AgentRole", "AgentProfile", "Master"} <= set(r) and "Vpc" not in r
    acts = r["MasterRole"]["Properties"]["Policies"][0]["PolicyDocument"]["Statement"][0]["Action"]
    assert {"ec2:RunInstances", "ec2:TerminateInstances", "ec2:DescribeInstances", "ec2:CreateTags"} <= set(acts)
    prov = _master_provisioner(r["Master"]["Properties"]["UserData"]["Fn::Base64"]["Fn::Sub"])
    assert prov["max_instances"] == 3 and prov["provider"]["type"] == "aws" and prov["provider"]["spot"] and prov["provider"]["spot_max_price"] == "9.5"
    assert prov["provider"]["tag_value"] == "c1" and prov["provider"]["security_group_id"] == "${SecurityGroup.GroupId}"
    ports = {(i["FromPort"], i["ToPort"]) for i in r["SecurityGroup"]["Properties"]["SecurityGroupIngress"]}
    assert ports == {(22, 22), (8080, 8080)}
    v = cloud.aws_template("c2", "key", "vpc")["Resources"]
    assert {"Vpc", "Subnet", "Gateway", "Route"} <= set(v) and v["Master"]["Properties"]["SubnetId"] == {"Ref": "Subnet"}
    assert _master_provisioner(v["Master"]["Properties"]["UserData"]["Fn::Base64"]["Fn::Sub"])["provider"]["subnet_id"] == "${Subnet}"
    e = cloud.aws_template("c3", "key", "efs")["Resources"]
    assert {"FileSystem", "MountTarget"} <= set(e)
    assert "mount -t efs" in e["Master"]["Properties"]["UserData"]["Fn::Base64"]["Fn::Sub"]
    assert yaml.safe_load(yaml.safe_dump(t)) == t


def test_gcp_terraform():
    cfg = cloud.gcp_terraform("c9", "proj", region="europe-west4", gpu_num=8, preemptible=True, max_agents=2)
    res = cfg["resource"]
    assert res["google_compute_instance"]["master"]["zone"] == "europe-west4-a"
    prov = _master_provisioner(res["google_compute_instance"]["master"]["metadata_startup_script"])
    assert prov["provider"]["type"] == "gcp" and prov["provider"]["instance_type"]["preemptible"] and prov["slots_per_instance"] == 8
    assert prov["provider"]["label_value"] == "c9" and prov["max_instances"] == 2
    assert {"det_external", "det_internal"} == set(res["google_compute_firewall"])
    assert json.loads(json.dumps(cfg)) == cfg


def test_det_deploy_cli_dry_run(tmp_path):
    out = subprocess.run([sys.executable, str(ROOT / "det-deploy"), "aws", "up", "--cluster-id", "t1", "--keypair", "k", "--region", "us-east-1",
                          "--deployment-type", "vpc", "--template-out", str(tmp_path / "t.yaml"), "--dry-run"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    j = json.loads(out.stdout)
    assert j["steps"][0]["cmd"][:5] == ["aws", "--region", "us-east-1", "cloudformation", "deploy"] and j["steps"][0]["skipped"]
    assert "Vpc" in yaml.safe_load((tmp_path / "t.yaml").read_text())["Resources"]
    out = subprocess.run([sys.executable, str(ROOT / "det-deploy"), "gcp", "up", "--cluster-id", "t2", "--project-id", "p", "--local-state-path",
                          str(tmp_path / "tf"), "--dry-run"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    j = json.loads(out.stdout)
    assert [s["cmd"][2] for s in j["steps"]] == ["init", "apply"] and (tmp_path / "tf" / "main.tf.json").exists()
    out = subprocess.run([sys.executable, str(ROOT / "det-deploy"), "aws", "down", "--cluster-id", "t1", "--dry-run"], capture_output=True, text=True)
    assert out.returncode == 0 and "delete-stack" in out.stdout
