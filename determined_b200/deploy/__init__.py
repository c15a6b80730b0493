"""Cluster bring-up helpers (``det-deploy``). Only the single-box ``local`` flavour exists: the reference's AWS
CloudFormation / GCP Terraform flavours (``deploy/determined_deploy/aws``, ``gcp``) need cloud APIs that do not exist
offline."""
