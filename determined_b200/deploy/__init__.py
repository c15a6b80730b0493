"""Cluster bring-up helpers (``det-deploy``): ``local`` (processes on this box), ``aws`` (CloudFormation) and ``gcp``
(Terraform) — reference ``deploy/determined_deploy/{local,aws,gcp}``."""
