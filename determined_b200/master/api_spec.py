"""The master's API as data: messages and RPCs of ``determined.api.v1.Determined``.

The reference defines its API in protobuf (``proto/src/determined/api/v1/api.proto``, 73 RPCs, served by
``master/internal/grpc/api.go`` and re-exposed over REST by grpc-gateway) and generates the TypeScript SDK of the WebUI
from it. Here the REST routes of the C++ master are primary and this table binds every RPC to one of them; three things are
derived from it and nothing else:

* ``proto/determined/api/v1/api.proto`` — the published contract (``scripts/gen_api.py``, drift-tested);
* the descriptors the gRPC gateway (``grpc_gateway.py``) builds at start-up — no ``protoc`` needed at run time;
* ``webui/api-ts-sdk/api.ts`` — a typed ``fetch`` client for browser code.

A field is ``(name, type)`` or ``(name, type, "repeated")``. Scalar types are proto3 names; ``Struct`` / ``Value`` are the
``google.protobuf`` well-known types (experiment configs, hyperparameters and metrics are free-form JSON); anything else
names a message of this table. Every resource message ends with ``raw`` — the complete JSON object as the master returned
it — so that fields this table does not type are not lost.
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Tuple

PACKAGE = "determined.api.v1"
SERVICE = "Determined"

Field = Tuple[str, ...]

RESOURCES: Dict[str, List[Field]] = {
    "Master": [("version", "string"), ("master_id", "string"), ("cluster_id", "string"), ("cluster_name", "string"),
               ("uptime_s", "int64"), ("scheduler", "string")],
    "User": [("username", "string"), ("admin", "bool"), ("active", "bool")],
    "Experiment": [("id", "int32"), ("state", "string"), ("description", "string"), ("archived", "bool"), ("progress", "double"),
                   ("num_trials", "int32"), ("owner", "string"), ("start_time", "string"), ("end_time", "string"),
                   ("labels", "string", "repeated"), ("config", "Struct"), ("trials", "Trial", "repeated")],
    "Trial": [("id", "int32"), ("experiment_id", "int32"), ("state", "string"), ("hparams", "Struct"), ("seed", "int64"),
              ("restarts", "int32"), ("total_batches_processed", "int64"), ("num_steps", "int32"), ("start_time", "string"),
              ("end_time", "string"), ("latest_validation", "Struct"), ("steps", "Struct", "repeated"),
              ("validations", "Struct", "repeated"), ("checkpoints", "Checkpoint", "repeated")],
    "Checkpoint": [("uuid", "string"), ("experiment_id", "int32"), ("trial_id", "int32"), ("state", "string"),
                   ("total_batches", "int64"), ("framework", "string"), ("format", "string"), ("resources", "Struct"),
                   ("metadata", "Struct"), ("end_time", "string")],
    "LogEntry": [("id", "int64"), ("trial_id", "int32"), ("timestamp", "string"), ("level", "string"), ("message", "string"),
                 ("agent_id", "string"), ("container_id", "string"), ("rank", "Value"), ("source", "string"), ("stdtype", "string")],
    "Agent": [("id", "string"), ("label", "string"), ("num_slots", "int32"), ("num_containers", "int32"), ("registered_time", "string"),
              ("driver_version", "string"), ("addresses", "string", "repeated"), ("slots", "Slot", "repeated")],
    "Slot": [("id", "string"), ("agent_id", "string"), ("enabled", "bool"), ("device", "Struct"), ("container", "Value")],
    "Task": [("id", "string"), ("name", "string"), ("state", "string"), ("slots_needed", "int32"), ("group", "string"),
             ("containers", "Struct", "repeated")],
    "Template": [("name", "string"), ("config", "Struct")],
    "Model": [("name", "string"), ("description", "string"), ("metadata", "Struct"), ("creation_time", "string"),
              ("last_updated_time", "string"), ("num_versions", "int32")],
    "ModelVersion": [("version", "int32"), ("checkpoint_uuid", "string"), ("creation_time", "string"), ("checkpoint", "Checkpoint")],
    "Command": [("id", "string"), ("type", "string"), ("state", "string"), ("description", "string"), ("owner", "string"),
                ("is_ready", "bool"), ("exit_code", "Value"), ("proxy_path", "string"), ("argv", "string", "repeated"),
                ("logs", "Struct", "repeated")],
}


def _rpc(name: str, method: str, path: str, request: List[Field], response: Optional[Field] = None, *, body: str = "fields",
         stream: bool = False, doc: str = "") -> Dict[str, Any]:
    """``path`` uses ``{field}`` for request fields that become path segments; with ``body="fields"`` the remaining request
    fields form the JSON body of POST / PUT / PATCH calls and the query string of GET / DELETE calls; ``body="<field>"``
    sends that one field's value AS the body. ``response`` names the single field of ``<Name>Response`` the returned JSON
    (object or array) is placed in; ``None`` means an empty response."""
    return {"name": name, "method": method, "path": path, "request": request, "response": response, "body": body,
            "stream": stream, "doc": doc}


ID32 = ("id", "int32")

RPCS: List[Dict[str, Any]] = [
    # ---- master / auth / users -----------------------------------------------------------------------------------------
    _rpc("GetMaster", "GET", "/api/v1/master", [], ("master", "Master"), doc="Version, cluster id and scheduler of the master."),
    _rpc("GetMasterConfig", "GET", "/api/v1/master/config", [], ("config", "Struct"), doc="The master's effective configuration."),
    _rpc("MasterLogs", "GET", "/api/v1/master/logs", [("limit", "int32")], ("entries", "Struct", "repeated"), stream=True,
         doc="The master's own log ring."),
    _rpc("Login", "POST", "/api/v1/auth/login", [("username", "string"), ("password", "string")], ("session", "Struct"),
         doc="Exchange credentials for a bearer token."),
    _rpc("Logout", "POST", "/api/v1/auth/logout", [], None),
    _rpc("CurrentUser", "GET", "/api/v1/auth/user", [], ("user", "User")),
    _rpc("GetUsers", "GET", "/api/v1/users", [], ("users", "User", "repeated")),
    _rpc("GetUser", "GET", "/api/v1/users/{name}", [("name", "string")], ("user", "User")),
    _rpc("GetTelemetry", "GET", "/api/v1/master/telemetry", [], ("telemetry", "Struct"), doc="enabled, segment_key (always empty: the sink is a local file)."),
    _rpc("PostUser", "POST", "/api/v1/users", [("username", "string"), ("password", "string"), ("admin", "bool")], ("user", "User")),
    _rpc("SetUserPassword", "PATCH", "/api/v1/users/{name}", [("name", "string"), ("password", "string")], ("user", "User")),
    _rpc("SetUserActive", "PATCH", "/api/v1/users/{name}", [("name", "string"), ("active", "bool")], ("user", "User")),
    # ---- experiments -----------------------------------------------------------------------------------------------------
    _rpc("CreateExperiment", "POST", "/api/v1/experiments",
         [("config", "Struct"), ("model_definition", "Struct", "repeated"), ("template", "string"), ("activate", "bool"),
          ("validate_only", "bool"), ("parent_id", "int32")], ("experiment", "Experiment"),
         doc="Submit an experiment: config + model-definition files ({path, content (text), mode})."),
    _rpc("GetExperiments", "GET", "/api/v1/experiments",
         [("states", "string"), ("archived", "bool"), ("users", "string"), ("labels", "string"), ("description", "string"),
          ("sort_by", "string"), ("order_by", "string"), ("offset", "int32"), ("limit", "int32")],
         ("experiments", "Experiment", "repeated"),
         doc="Filter (comma-separated sets; description is a substring), sort (id, description, start_time, end_time, state, progress, user, num_trials) and page."),
    _rpc("GetExperiment", "GET", "/api/v1/experiments/{id}", [ID32, ("metrics", "bool")], ("experiment", "Experiment")),
    _rpc("ActivateExperiment", "POST", "/api/v1/experiments/{id}/activate", [ID32], None),
    _rpc("PauseExperiment", "POST", "/api/v1/experiments/{id}/pause", [ID32], None),
    _rpc("CancelExperiment", "POST", "/api/v1/experiments/{id}/cancel", [ID32], None),
    _rpc("KillExperiment", "POST", "/api/v1/experiments/{id}/kill", [ID32], None),
    _rpc("ArchiveExperiment", "POST", "/api/v1/experiments/{id}/archive", [ID32], None),
    _rpc("UnarchiveExperiment", "POST", "/api/v1/experiments/{id}/unarchive", [ID32], None),
    _rpc("PatchExperiment", "PATCH", "/api/v1/experiments/{id}",
         [ID32, ("description", "string"), ("labels", "string", "repeated"), ("resources", "Struct"), ("checkpoint_storage", "Struct")],
         ("experiment", "Experiment"), doc="Description, labels, max_slots / weight / priority, GC policy."),
    _rpc("DeleteExperiment", "DELETE", "/api/v1/experiments/{id}", [ID32], None),
    _rpc("GetExperimentCheckpoints", "GET", "/api/v1/experiments/{id}/checkpoints", [ID32], ("checkpoints", "Checkpoint", "repeated")),
    _rpc("GetExperimentTrials", "GET", "/api/v1/experiments/{id}/trials", [ID32, ("state", "string"), ("metrics", "bool")],
         ("trials", "Trial", "repeated")),
    _rpc("GetExperimentLabels", "GET", "/api/v1/experiment/labels", [], ("result", "Struct"), doc="labels[]: {label, count}, most used first."),
    _rpc("GetExperimentValidationHistory", "GET", "/api/v1/experiments/{id}/validation-history", [ID32], ("result", "Struct"),
         doc="validation_history[]: every validation that set a new best of the searcher metric, in time order."),
    _rpc("GetModelDef", "GET", "/api/v1/experiments/{id}/model_def", [ID32], ("model_def", "Struct")),
    _rpc("ExperimentMetrics", "GET", "/api/v1/experiments/{id}/metrics",
         [ID32, ("name", "string"), ("type", "string"), ("max_points", "int32")], ("metrics", "Struct"),
         doc="Per-trial metric series of an experiment, LTTB down-sampled to max_points."),
    _rpc("MetricNames", "GET", "/api/v1/experiments/{id}/metrics-stream/metric-names", [ID32], ("names", "Struct"),
         doc="searcher_metric, training_metrics[], validation_metrics[] recorded so far by the experiment's trials."),
    _rpc("MetricBatches", "GET", "/api/v1/experiments/{id}/metrics-stream/batches",
         [ID32, ("metric_name", "string"), ("metric_type", "string")], ("result", "Struct"),
         doc="batches[]: the milestones (batches processed) at which the metric has been recorded; metric_type training | validation."),
    _rpc("TrialsSnapshot", "GET", "/api/v1/experiments/{id}/metrics-stream/trials-snapshot",
         [ID32, ("batches_processed", "int32"), ("metric_name", "string"), ("metric_type", "string")], ("result", "Struct"),
         doc="trials[]: {trial_id, hparams, metric} of every trial that has a value of the metric at that point of progress."),
    _rpc("TrialsSample", "GET", "/api/v1/experiments/{id}/metrics-stream/trials-sample",
         [ID32, ("metric_name", "string"), ("metric_type", "string"), ("max_trials", "int32"), ("max_datapoints", "int32"),
          ("start_batches", "int32"), ("end_batches", "int32")], ("result", "Struct"),
         doc="trials[]: {trial_id, hparams, data[{batches, value}]} of the best max_trials trials, down-sampled; promoted_trials / demoted_trials."),
    _rpc("PreviewHPSearch", "POST", "/api/v1/searcher/preview", [("searcher", "Struct"), ("hyperparameters", "Struct"), ("seed", "int32")],
         ("simulation", "Struct"), doc="Simulate a searcher: multiset of per-trial operation sequences."),
    # ---- trials / checkpoints ------------------------------------------------------------------------------------------------
    _rpc("GetTrial", "GET", "/api/v1/trials/{id}", [ID32, ("metrics", "bool")], ("trial", "Trial")),
    _rpc("KillTrial", "POST", "/api/v1/trials/{id}/kill", [ID32], None),
    _rpc("TrialLogs", "GET", "/api/v1/trials/{id}/logs",
         [ID32, ("offset", "int64"), ("limit", "int32"), ("head", "int32"), ("tail", "int32"), ("follow", "bool"), ("level", "string"),
          ("agent_ids", "string"), ("container_ids", "string"), ("rank_ids", "string"), ("stdtypes", "string"), ("sources", "string"),
          ("timestamp_before", "string"), ("timestamp_after", "string"), ("search", "string")],
         ("entries", "LogEntry", "repeated"), stream=True,
         doc="Filterable trial logs (sets are comma-separated, level is a minimum); every entry carries its position `id`; with follow the stream stays open until the trial ends."),
    _rpc("TrialLogsFields", "GET", "/api/v1/trials/{id}/logs/fields", [ID32], ("fields", "Struct"),
         doc="Distinct agent_ids, container_ids, rank_ids, stdtypes, sources of a trial's log lines."),
    _rpc("GetTrialCheckpoints", "GET", "/api/v1/trials/{id}/checkpoints", [ID32, ("state", "string")], ("checkpoints", "Checkpoint", "repeated")),
    _rpc("GetCheckpoint", "GET", "/api/v1/checkpoints/{uuid}", [("uuid", "string")], ("checkpoint", "Checkpoint")),
    _rpc("PostCheckpointMetadata", "POST", "/api/v1/checkpoints/{uuid}/metadata",
         [("uuid", "string"), ("metadata", "Struct"), ("remove", "string", "repeated")], ("checkpoint", "Checkpoint"),
         doc="Merge keys into a checkpoint's user metadata; `remove` deletes keys."),
    # ---- cluster -------------------------------------------------------------------------------------------------------------
    _rpc("GetAgents", "GET", "/api/v1/agents", [], ("agents", "Agent", "repeated")),
    _rpc("GetAgent", "GET", "/api/v1/agents/{agent_id}", [("agent_id", "string")], ("agent", "Agent")),
    _rpc("EnableAgent", "POST", "/api/v1/agents/{agent_id}/enable", [("agent_id", "string")], None),
    _rpc("DisableAgent", "POST", "/api/v1/agents/{agent_id}/disable", [("agent_id", "string")], None),
    _rpc("GetSlots", "GET", "/api/v1/slots", [], ("slots", "Slot", "repeated")),
    _rpc("GetSlot", "GET", "/api/v1/agents/{agent_id}/slots/{slot_id}", [("agent_id", "string"), ("slot_id", "string")], ("slot", "Slot")),
    _rpc("EnableSlot", "POST", "/api/v1/agents/{agent_id}/slots/{slot_id}/enable", [("agent_id", "string"), ("slot_id", "string")], None),
    _rpc("DisableSlot", "POST", "/api/v1/agents/{agent_id}/slots/{slot_id}/disable", [("agent_id", "string"), ("slot_id", "string")], None),
    _rpc("GetTasks", "GET", "/api/v1/tasks", [], ("tasks", "Task", "repeated")),
    _rpc("GetProvisioner", "GET", "/api/v1/provisioner", [], ("provisioner", "Struct")),
    _rpc("GetKubernetesPods", "GET", "/api/v1/kubernetes/pods", [], ("pods", "Struct", "repeated")),
    # ---- templates / model registry ----------------------------------------------------------------------------------------
    _rpc("GetTemplates", "GET", "/api/v1/templates", [], ("templates", "Template", "repeated")),
    _rpc("GetTemplate", "GET", "/api/v1/templates/{name}", [("name", "string")], ("template", "Template")),
    _rpc("PutTemplate", "PUT", "/api/v1/templates/{name}", [("name", "string"), ("config", "Struct")], ("template", "Template")),
    _rpc("DeleteTemplate", "DELETE", "/api/v1/templates/{name}", [("name", "string")], None),
    _rpc("GetModels", "GET", "/api/v1/models", [], ("models", "Model", "repeated")),
    _rpc("PostModel", "POST", "/api/v1/models", [("name", "string"), ("description", "string"), ("metadata", "Struct")], ("model", "Model")),
    _rpc("GetModel", "GET", "/api/v1/models/{name}", [("name", "string")], ("model", "Model")),
    _rpc("PatchModel", "PATCH", "/api/v1/models/{name}", [("name", "string"), ("description", "string"), ("metadata", "Struct")], ("model", "Model")),
    _rpc("GetModelVersions", "GET", "/api/v1/models/{name}/versions", [("name", "string")], ("versions", "ModelVersion", "repeated")),
    _rpc("GetModelVersion", "GET", "/api/v1/models/{name}/versions/{version}", [("name", "string"), ("version", "int32")],
         ("version_info", "Struct")),
    _rpc("PostModelVersion", "POST", "/api/v1/models/{name}/versions", [("name", "string"), ("checkpoint_uuid", "string")],
         ("version", "ModelVersion")),
]

# commands / notebooks / shells / tensorboards: the same four verbs on four collections
for _kind, _coll in (("Command", "commands"), ("Notebook", "notebooks"), ("Shell", "shells"), ("Tensorboard", "tensorboards")):
    RPCS += [
        _rpc(f"Launch{_kind}", "POST", f"/api/v1/{_coll}",
             [("argv", "string", "repeated"), ("description", "string"), ("slots", "int32"), ("proxy", "bool"), ("config", "Struct"),
              ("experiment_ids", "int32", "repeated"), ("trial_ids", "int32", "repeated")], ("command", "Command")),
        _rpc(f"Get{_kind}s", "GET", f"/api/v1/{_coll}", [], ("commands", "Command", "repeated")),
        _rpc(f"Get{_kind}", "GET", f"/api/v1/{_coll}/{{id}}", [("id", "string")], ("command", "Command")),
        _rpc(f"Kill{_kind}", "POST", f"/api/v1/{_coll}/{{id}}/kill", [("id", "string")], None),
        _rpc(f"{_kind}Logs", "GET", f"/api/v1/{_coll}/{{id}}/logs", [("id", "string"), ("offset", "int64"), ("limit", "int32")],
             ("entries", "Struct", "repeated"), stream=True, doc="Output lines of the task."),
    ]


def path_fields(rpc: Dict[str, Any]) -> List[str]:
    import re

    return re.findall(r"\{(\w+)\}", rpc["path"])
