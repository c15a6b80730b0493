"""Checkpoint interchange with the UNMODIFIED reference harness (``baseline/_ref``), used by
``tests/test_reference_checkpoint_compat.py``. Runs as a subprocess with ``PYTHONPATH=baseline/shims:baseline/_ref`` and
a working directory outside the repository, so ``determined`` here is the reference package.

    save <model_dir> <ckpt_dir> <out.json>   train the model definition with the reference controller, checkpoint it
    load <model_dir> <ckpt_dir> <out.json>   start the reference controller FROM a checkpoint, report the loaded weights,
                                             train one more workload
"""
import det_ref_compat  # noqa: F401  (must precede the reference imports)

import json
import pathlib
import sys

import torch

from determined import _execution, load, workload


def main() -> None:
    mode, model_dir, ckpt_dir, out_path = sys.argv[1:5]
    sys.path.insert(0, model_dir)
    import model_def  # the same file our own harness runs: it only imports `determined.pytorch`

    hp = {"global_batch_size": 8, "hidden_size": 4, "learning_rate": 0.5}
    config = {"hyperparameters": hp,
              "searcher": {"name": "single", "metric": "validation_loss", "smaller_is_better": True, "max_length": {"batches": 100}},
              "optimizations": {"aggregation_frequency": 1, "average_aggregated_gradients": True, "average_training_metrics": False,
                                "gradient_compression": False, "mixed_precision": "O0", "tensor_fusion_threshold": 64,
                                "tensor_fusion_cycle_time": 5, "auto_tune_tensor_fusion": False},
              "reproducibility": {"experiment_seed": 99}}
    env, rendezvous_info, hvd_config = _execution._make_local_execution_env(True, config, hparams=hp, limit_gpus=0)
    report = {}
    box = {}

    def weights():
        return torch.cat([p.detach().reshape(-1) for m in box["c"].context.models for p in m.parameters()]).tolist()

    def workloads():
        icpt = workload.WorkloadResponseInterceptor()
        total = 0
        if mode == "load":
            report["loaded_weights"] = weights()
        for step in range(1, 3 if mode == "save" else 2):
            yield from icpt.send(workload.train_workload(step, num_batches=10, total_batches_processed=total), [])
            total += 10
            report.setdefault("losses", []).append(float(icpt.metrics_result()["metrics"]["avg_metrics"]["loss"]))
        if mode == "save":
            report["final_weights"] = weights()
            yield from icpt.send(workload.checkpoint_workload(), [pathlib.Path(ckpt_dir)])
            report["checkpoint_response"] = {k: str(v) for k, v in icpt.result.items()} if isinstance(getattr(icpt, "result", None), dict) else "ok"
        yield from icpt.send(workload.terminate_workload(), [])

    controller = load.load_controller_from_trial(model_def.XorTrial, env, workloads(), pathlib.Path(ckpt_dir) if mode == "load" else None,
                                                 rendezvous_info, hvd_config)
    box["c"] = controller
    controller.run()
    with open(out_path, "w") as f:
        json.dump(report, f)
    print("REF_CKPT_OK", flush=True)


if __name__ == "__main__":
    main()
