"""``det-deploy aws`` / ``det-deploy gcp``: cloud bring-up of a master whose provisioner launches agent instances.

Reference: ``deploy/determined_deploy/aws`` (CloudFormation templates ``simple`` / ``vpc`` / ``efs`` deployed with boto3) and
``deploy/determined_deploy/gcp`` (Terraform modules driven by ``terraform apply``). Here the templates are generated in
Python — a CloudFormation document and a Terraform JSON configuration — and applied with the vendors' command-line
clients (``aws cloudformation deploy``, ``terraform``), the same way the master's provisioner providers work
(``csrc/core/cloud_providers.hpp``). ``--dry-run`` writes the template and prints the commands without running them.
"""
from __future__ import annotations

import json
import pathlib
import shutil
import subprocess
from typing import Any, Dict, List, Optional

import yaml

AWS_DEPLOYMENT_TYPES = ("simple", "vpc", "efs")
DEFAULT_MASTER_PORT = 8080


def _provisioner_config(provider: Dict[str, Any], max_instances: int, slots_per_instance: int, idle_period_s: int) -> Dict[str, Any]:
    return {"max_instances": max_instances, "slots_per_instance": slots_per_instance, "max_idle_agent_period": idle_period_s,
            "max_agent_starting_period": 600, "tick_ms": 5000, "provider": provider}


def _master_boot_script(install_dir: str, port: int, provisioner: Dict[str, Any], scheduler: str, extra: str = "") -> str:
    return "\n".join([
        "#!/bin/bash", "set -e", extra,
        f"cd {install_dir}",
        f"exec ./determined_b200/_lib/det-master --host 0.0.0.0 --port {port} --slot-type none --scheduler {scheduler} "
        f"--repo-root {install_dir} --python python3 --data-dir /var/lib/determined-b200 "
        f"--provisioner '{json.dumps(provisioner)}'", ""])


# ---- AWS: CloudFormation ----------------------------------------------------------------------------------------------
def aws_template(cluster_id: str, keypair: str, deployment_type: str = "simple", master_instance_type: str = "m5.large",
                 agent_instance_type: str = "p5.48xlarge", image_id: str = "ami-determined-b200", inbound_cidr: str = "0.0.0.0/0",
                 max_agents: int = 5, slots_per_agent: int = 8, idle_period_s: int = 600, spot: bool = False,
                 spot_max_price: str = "", install_dir: str = "/opt/determined_b200", scheduler: str = "fair_share",
                 port: int = DEFAULT_MASTER_PORT) -> Dict[str, Any]:
    """The CloudFormation document of one cluster: security group, IAM role the master's provisioner runs under, the
    master instance; ``vpc`` adds a dedicated VPC + subnet, ``efs`` additionally a shared file system for checkpoints."""
    if deployment_type not in AWS_DEPLOYMENT_TYPES:
        raise ValueError(f"deployment type must be one of {AWS_DEPLOYMENT_TYPES}")
    res: Dict[str, Any] = {}
    own_vpc = deployment_type in ("vpc", "efs")
    if own_vpc:
        res["Vpc"] = {"Type": "AWS::EC2::VPC", "Properties": {"CidrBlock": "10.0.0.0/16", "EnableDnsSupport": True, "EnableDnsHostnames": True}}
        res["Subnet"] = {"Type": "AWS::EC2::Subnet", "Properties": {"VpcId": {"Ref": "Vpc"}, "CidrBlock": "10.0.0.0/24", "MapPublicIpOnLaunch": True}}
        res["Gateway"] = {"Type": "AWS::EC2::InternetGateway"}
        res["GatewayAttachment"] = {"Type": "AWS::EC2::VPCGatewayAttachment", "Properties": {"VpcId": {"Ref": "Vpc"}, "InternetGatewayId": {"Ref": "Gateway"}}}
        res["RouteTable"] = {"Type": "AWS::EC2::RouteTable", "Properties": {"VpcId": {"Ref": "Vpc"}}}
        res["Route"] = {"Type": "AWS::EC2::Route", "DependsOn": "GatewayAttachment",
                        "Properties": {"RouteTableId": {"Ref": "RouteTable"}, "DestinationCidrBlock": "0.0.0.0/0", "GatewayId": {"Ref": "Gateway"}}}
        res["SubnetRoute"] = {"Type": "AWS::EC2::SubnetRouteTableAssociation", "Properties": {"SubnetId": {"Ref": "Subnet"}, "RouteTableId": {"Ref": "RouteTable"}}}
    sg: Dict[str, Any] = {"GroupDescription": f"determined-b200 cluster {cluster_id}",
                          "SecurityGroupIngress": [{"IpProtocol": "tcp", "FromPort": 22, "ToPort": 22, "CidrIp": inbound_cidr},
                                                   {"IpProtocol": "tcp", "FromPort": port, "ToPort": port, "CidrIp": inbound_cidr}]}
    if own_vpc:
        sg["VpcId"] = {"Ref": "Vpc"}
    res["SecurityGroup"] = {"Type": "AWS::EC2::SecurityGroup", "Properties": sg}
    # members of the group talk freely to each other: master <-> agents, NCCL between agents
    res["SecurityGroupSelf"] = {"Type": "AWS::EC2::SecurityGroupIngress",
                                "Properties": {"GroupId": {"Fn::GetAtt": ["SecurityGroup", "GroupId"]}, "IpProtocol": "-1",
                                               "SourceSecurityGroupId": {"Fn::GetAtt": ["SecurityGroup", "GroupId"]}}}
    res["MasterRole"] = {"Type": "AWS::IAM::Role", "Properties": {
        "AssumeRolePolicyDocument": {"Version": "2012-10-17", "Statement": [{"Effect": "Allow", "Principal": {"Service": ["ec2.amazonaws.com"]},
                                                                               "Action": ["sts:AssumeRole"]}]},
        "Policies": [{"PolicyName": "provisioner", "PolicyDocument": {"Version": "2012-10-17", "Statement": [
            {"Effect": "Allow", "Resource": "*", "Action": ["ec2:RunInstances", "ec2:TerminateInstances", "ec2:DescribeInstances", "ec2:CreateTags",
                                                            "ec2:RequestSpotInstances", "ec2:CancelSpotInstanceRequests", "ec2:DescribeSpotInstanceRequests"]},
            {"Effect": "Allow", "Resource": {"Fn::GetAtt": ["AgentRole", "Arn"]}, "Action": ["iam:PassRole"]}]}}]}}
    res["MasterProfile"] = {"Type": "AWS::IAM::InstanceProfile", "Properties": {"Roles": [{"Ref": "MasterRole"}]}}
    res["AgentRole"] = {"Type": "AWS::IAM::Role", "Properties": {
        "AssumeRolePolicyDocument": {"Version": "2012-10-17", "Statement": [{"Effect": "Allow", "Principal": {"Service": ["ec2.amazonaws.com"]},
                                                                               "Action": ["sts:AssumeRole"]}]},
        "Policies": [{"PolicyName": "checkpoints", "PolicyDocument": {"Version": "2012-10-17", "Statement": [
            {"Effect": "Allow", "Resource": "*", "Action": ["s3:GetObject", "s3:PutObject", "s3:DeleteObject", "s3:ListBucket"]}]}}]}}
    res["AgentProfile"] = {"Type": "AWS::IAM::InstanceProfile", "Properties": {"Roles": [{"Ref": "AgentRole"}]}}
    mount = ""
    agent_startup = ""
    if deployment_type == "efs":
        res["FileSystem"] = {"Type": "AWS::EFS::FileSystem", "Properties": {"PerformanceMode": "generalPurpose"}}
        res["MountTarget"] = {"Type": "AWS::EFS::MountTarget", "Properties": {"FileSystemId": {"Ref": "FileSystem"}, "SubnetId": {"Ref": "Subnet"},
                                                                              "SecurityGroups": [{"Fn::GetAtt": ["SecurityGroup", "GroupId"]}]}}
        mount = "mkdir -p /mnt/efs && mount -t efs ${FileSystem}:/ /mnt/efs"
        agent_startup = mount
    provider: Dict[str, Any] = {"type": "aws", "region": "${AWS::Region}", "image_id": image_id, "instance_type": agent_instance_type,
                                "ssh_key_name": keypair, "security_group_id": "${SecurityGroup.GroupId}",
                                "iam_instance_profile_arn": "${AgentProfile.Arn}", "tag_key": "determined-cluster", "tag_value": cluster_id,
                                "agent_command": f"{install_dir}/determined_b200/_lib/det-agent", "root_volume_size": 200}
    if own_vpc:
        provider["subnet_id"] = "${Subnet}"
    if spot:
        provider["spot"] = True
        if spot_max_price:
            provider["spot_max_price"] = spot_max_price
    if agent_startup:
        provider["startup_script"] = agent_startup
    prov = _provisioner_config(provider, max_agents, slots_per_agent, idle_period_s)
    master: Dict[str, Any] = {"ImageId": image_id, "InstanceType": master_instance_type, "KeyName": keypair,
                              "IamInstanceProfile": {"Ref": "MasterProfile"},
                              "Tags": [{"Key": "Name", "Value": f"det-master-{cluster_id}"}, {"Key": "determined-master", "Value": cluster_id}],
                              "UserData": {"Fn::Base64": {"Fn::Sub": _master_boot_script(install_dir, port, prov, scheduler, mount)}}}
    if own_vpc:
        master["SubnetId"] = {"Ref": "Subnet"}
        master["SecurityGroupIds"] = [{"Fn::GetAtt": ["SecurityGroup", "GroupId"]}]
    else:
        master["SecurityGroups"] = [{"Ref": "SecurityGroup"}]
    res["Master"] = {"Type": "AWS::EC2::Instance", "Properties": master}
    return {"AWSTemplateFormatVersion": "2010-09-09", "Description": f"determined_b200 cluster {cluster_id} ({deployment_type})",
            "Resources": res,
            "Outputs": {"MasterAddress": {"Value": {"Fn::GetAtt": ["Master", "PublicDnsName"]}},
                        "MasterUrl": {"Value": {"Fn::Sub": "http://${Master.PublicDnsName}:" + str(port)}},
                        "AgentTag": {"Value": f"determined-cluster={cluster_id}"}}}


def aws_commands(cluster_id: str, template_path: str, region: Optional[str], verb: str, cli: str = "aws") -> List[List[str]]:
    base = [cli] + (["--region", region] if region else [])
    if verb == "up":
        return [base + ["cloudformation", "deploy", "--template-file", template_path, "--stack-name", cluster_id,
                        "--capabilities", "CAPABILITY_NAMED_IAM", "--no-fail-on-empty-changeset"],
                base + ["cloudformation", "describe-stacks", "--stack-name", cluster_id, "--query", "Stacks[0].Outputs", "--output", "json"]]
    if verb == "down":
        # agent instances are not stack resources: terminate the tagged ones first, then the stack
        return [base + ["ec2", "describe-instances", "--filters", f"Name=tag:determined-cluster,Values={cluster_id}",
                        "--query", "Reservations[].Instances[].InstanceId", "--output", "text"],
                base + ["cloudformation", "delete-stack", "--stack-name", cluster_id],
                base + ["cloudformation", "wait", "stack-delete-complete", "--stack-name", cluster_id]]
    raise ValueError(verb)


# ---- GCP: Terraform ------------------------------------------------------------------------------------------------------
def gcp_terraform(cluster_id: str, project_id: str, region: str = "us-central1", zone: Optional[str] = None,
                  master_machine_type: str = "n1-standard-4", agent_machine_type: str = "a3-highgpu-8g", gpu_type: str = "nvidia-b200",
                  gpu_num: int = 8, image: str = "determined-b200", max_agents: int = 5, idle_period_s: int = 600,
                  preemptible: bool = False, install_dir: str = "/opt/determined_b200", scheduler: str = "fair_share",
                  port: int = DEFAULT_MASTER_PORT) -> Dict[str, Any]:
    """Terraform JSON configuration: network + firewall, service account, master instance with its boot script."""
    zone = zone or region + "-a"
    net = f"det-{cluster_id}"
    provider = {"type": "gcp", "project": project_id, "zone": zone, "boot_disk_source_image": image, "boot_disk_size": 200,
                "instance_type": {"machine_type": agent_machine_type, "gpu_type": gpu_type, "gpu_num": gpu_num, "preemptible": preemptible},
                "network_interface": {"network": net, "subnetwork": net}, "label_key": "determined-cluster", "label_value": cluster_id,
                "name_prefix": f"det-agent-{cluster_id}-", "agent_command": f"{install_dir}/determined_b200/_lib/det-agent",
                "service_account": {"email": "${google_service_account.det.email}"}}
    prov = _provisioner_config(provider, max_agents, gpu_num, idle_period_s)
    return {
        "terraform": {"required_providers": {"google": {"source": "hashicorp/google"}}},
        "provider": {"google": {"project": project_id, "region": region, "zone": zone}},
        "resource": {
            "google_compute_network": {"det": {"name": net, "auto_create_subnetworks": False}},
            "google_compute_subnetwork": {"det": {"name": net, "network": "${google_compute_network.det.id}", "ip_cidr_range": "10.0.0.0/16", "region": region}},
            "google_compute_firewall": {
                "det_external": {"name": f"{net}-external", "network": "${google_compute_network.det.id}", "source_ranges": ["0.0.0.0/0"],
                                 "allow": [{"protocol": "tcp", "ports": ["22", str(port)]}]},
                "det_internal": {"name": f"{net}-internal", "network": "${google_compute_network.det.id}", "source_ranges": ["10.0.0.0/16"],
                                 "allow": [{"protocol": "tcp"}, {"protocol": "udp"}, {"protocol": "icmp"}]}},
            "google_service_account": {"det": {"account_id": f"det-{cluster_id}"[:30], "display_name": f"determined_b200 {cluster_id}"}},
            "google_project_iam_member": {"det_compute": {"project": project_id, "role": "roles/compute.admin",
                                                          "member": "serviceAccount:${google_service_account.det.email}"},
                                          "det_sa_user": {"project": project_id, "role": "roles/iam.serviceAccountUser",
                                                          "member": "serviceAccount:${google_service_account.det.email}"}},
            "google_compute_instance": {"master": {
                "name": f"det-master-{cluster_id}", "machine_type": master_machine_type, "zone": zone,
                "labels": {"determined-master": cluster_id},
                "boot_disk": {"initialize_params": {"image": image, "size": 100}},
                "network_interface": {"subnetwork": "${google_compute_subnetwork.det.id}", "access_config": {}},
                "service_account": {"email": "${google_service_account.det.email}", "scopes": ["cloud-platform"]},
                "metadata_startup_script": _master_boot_script(install_dir, port, prov, scheduler)}}},
        "output": {"master_address": {"value": "${google_compute_instance.master.network_interface[0].access_config[0].nat_ip}"},
                   "agent_label": {"value": f"determined-cluster={cluster_id}"}}}


def gcp_commands(state_dir: str, verb: str, cli: str = "terraform") -> List[List[str]]:
    if verb == "up":
        return [[cli, f"-chdir={state_dir}", "init", "-input=false"], [cli, f"-chdir={state_dir}", "apply", "-auto-approve", "-input=false"]]
    if verb == "down":
        return [[cli, f"-chdir={state_dir}", "destroy", "-auto-approve", "-input=false"]]
    raise ValueError(verb)


# ---- CLI glue --------------------------------------------------------------------------------------------------------------
def _run(cmds: List[List[str]], dry_run: bool) -> List[Dict[str, Any]]:
    out = []
    for c in cmds:
        if dry_run:
            out.append({"cmd": c, "skipped": True})
            continue
        if shutil.which(c[0]) is None:
            raise SystemExit(f"det-deploy: `{c[0]}` is not on PATH — install it or use --dry-run to inspect the generated template")
        p = subprocess.run(c, capture_output=True, text=True)
        out.append({"cmd": c, "rc": p.returncode, "stdout": p.stdout[-4000:], "stderr": p.stderr[-2000:]})
        if p.returncode != 0:
            break
    return out


def add_parsers(sub: Any) -> None:
    aws = sub.add_parser("aws", help="CloudFormation deployment (master + provisioned agents)")
    asub = aws.add_subparsers(dest="verb", required=True)
    up = asub.add_parser("up")
    up.add_argument("--cluster-id", required=True)
    up.add_argument("--keypair", required=True)
    up.add_argument("--region")
    up.add_argument("--deployment-type", default="simple", choices=AWS_DEPLOYMENT_TYPES)
    up.add_argument("--master-instance-type", default="m5.large")
    up.add_argument("--agent-instance-type", default="p5.48xlarge")
    up.add_argument("--image-id", default="ami-determined-b200")
    up.add_argument("--inbound-cidr", default="0.0.0.0/0")
    up.add_argument("--max-dynamic-agents", type=int, default=5)
    up.add_argument("--max-idle-agent-period", type=int, default=600)
    up.add_argument("--spot", action="store_true")
    up.add_argument("--spot-max-price", default="")
    up.add_argument("--template-out", default=None)
    up.add_argument("--dry-run", action="store_true")
    down = asub.add_parser("down")
    down.add_argument("--cluster-id", required=True)
    down.add_argument("--region")
    down.add_argument("--dry-run", action="store_true")
    gcp = sub.add_parser("gcp", help="Terraform deployment (master + provisioned agents)")
    gsub = gcp.add_subparsers(dest="verb", required=True)
    gup = gsub.add_parser("up")
    gup.add_argument("--cluster-id", required=True)
    gup.add_argument("--project-id", required=True)
    gup.add_argument("--region", default="us-central1")
    gup.add_argument("--zone")
    gup.add_argument("--master-instance-type", default="n1-standard-4")
    gup.add_argument("--agent-instance-type", default="a3-highgpu-8g")
    gup.add_argument("--gpu-type", default="nvidia-b200")
    gup.add_argument("--gpu-num", type=int, default=8)
    gup.add_argument("--max-dynamic-agents", type=int, default=5)
    gup.add_argument("--max-idle-agent-period", type=int, default=600)
    gup.add_argument("--preemptible", action="store_true")
    gup.add_argument("--local-state-path", default=".")
    gup.add_argument("--dry-run", action="store_true")
    gdown = gsub.add_parser("down")
    gdown.add_argument("--local-state-path", default=".")
    gdown.add_argument("--dry-run", action="store_true")


def run(args: Any) -> Dict[str, Any]:
    if args.flavour == "aws":
        if args.verb == "up":
            tpl = aws_template(args.cluster_id, args.keypair, args.deployment_type, args.master_instance_type, args.agent_instance_type,
                               args.image_id, args.inbound_cidr, args.max_dynamic_agents, 8, args.max_idle_agent_period, args.spot,
                               args.spot_max_price)
            path = pathlib.Path(args.template_out or f"det-{args.cluster_id}.cfn.yaml")
            path.write_text(yaml.safe_dump(tpl, sort_keys=False))
            return {"template": str(path), "steps": _run(aws_commands(args.cluster_id, str(path), args.region, "up"), args.dry_run)}
        cmds = aws_commands(args.cluster_id, "", args.region, "down")
        if args.dry_run:
            return {"steps": _run(cmds, True)}
        steps = _run(cmds[:1], False)
        ids = steps[0].get("stdout", "").split()
        if ids:
            steps += _run([[cmds[0][0]] + (["--region", args.region] if args.region else []) + ["ec2", "terminate-instances", "--instance-ids"] + ids], False)
        return {"steps": steps + _run(cmds[1:], False)}
    if args.verb == "up":
        cfg = gcp_terraform(args.cluster_id, args.project_id, args.region, args.zone, args.master_instance_type, args.agent_instance_type,
                            args.gpu_type, args.gpu_num, "determined-b200", args.max_dynamic_agents, args.max_idle_agent_period, args.preemptible)
        d = pathlib.Path(args.local_state_path)
        d.mkdir(parents=True, exist_ok=True)
        (d / "main.tf.json").write_text(json.dumps(cfg, indent=1))
        return {"terraform": str(d / "main.tf.json"), "steps": _run(gcp_commands(str(d), "up"), args.dry_run)}
    return {"steps": _run(gcp_commands(args.local_state_path, "down"), args.dry_run)}
