"""One rank of a multi-process trial used by tests/test_distributed.py (gloo on CPU) and tests/test_gpu_multi.py
(NCCL + peer-memory kernels on GPUs). Launched with RANK / WORLD_SIZE / MASTER_* in the environment; rank 0 writes a
JSON report."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests.fixtures import pytorch_models as M  # noqa: E402
from tests.harness_utils import TrainAndValidate, make_controller  # noqa: E402


def main() -> None:
    out_path, case = sys.argv[1], sys.argv[2]
    rank = int(os.environ["RANK"])
    if case == "onevar":
        trial, hp, steps, unit = M.OneVarTrial, {"learning_rate": 0.01, "global_batch_size": 8}, 4, 2
        cfg = {"optimizations": {"average_training_metrics": True}}
    elif case == "xor_agg":
        trial, hp, steps, unit = M.XorTrial, {"hidden_size": 4, "learning_rate": 0.05, "global_batch_size": 8, "shuffle": False}, 4, 2
        cfg = {"optimizations": {"aggregation_frequency": 2, "average_training_metrics": True}}
    elif case == "xor_compressed":
        trial, hp, steps, unit = M.XorTrial, {"hidden_size": 4, "learning_rate": 0.05, "global_batch_size": 8, "shuffle": False}, 4, 2
        cfg = {"optimizations": {"gradient_compression": True, "average_training_metrics": True}}
    elif case == "xor_plain":
        trial, hp, steps, unit = M.XorTrial, {"hidden_size": 4, "learning_rate": 0.05, "global_batch_size": 8, "shuffle": False}, 4, 2
        cfg = {"optimizations": {"average_training_metrics": True}}
    elif case == "xor_fp16_inf":
        # fp16 loss scaling on the library all-reduce path: rank 1 poisons one gradient on the second batch; every rank
        # must skip that update, halve the scale and stay bit-identical
        import torch as _t
        from determined_b200 import pytorch as dp

        class PoisonedXor(M.XorTrial):
            def __init__(self, context):
                super().__init__(context)
                context.configure_apex_amp(self.model, self.opt, cast_model_type=_t.float16, loss_scale=1024.0)
                self.snap = []

            def train_batch(self, batch, epoch_idx, batch_idx):
                out = self.model(batch["data"])
                loss = self.loss_fn(out, batch["label"])
                self.context.backward(loss)
                if batch_idx == 1 and rank == 1:
                    self.model.l1.weight.grad[0, 0] = float("inf")
                before = _t.cat([p.detach().reshape(-1).clone() for p in self.model.parameters()])
                self.context.step_optimizer(self.opt)
                after = _t.cat([p.detach().reshape(-1).clone() for p in self.model.parameters()])
                return {"loss": loss, "moved": (before != after).any().float(),
                        "scale": self.context._loss_scaler.state[0].clone()}

        trial, hp, steps, unit = PoisonedXor, {"hidden_size": 4, "learning_rate": 0.05, "global_batch_size": 8, "shuffle": False}, 2, 2
        cfg = {"optimizations": {"average_training_metrics": False}}
    elif case == "xor_val_short":
        # fewer validation batches than ranks: rank 1's shard is empty, every rank still runs the same collectives
        class ShortVal(M.XorTrial):
            def build_validation_data_loader(self):
                from determined_b200 import pytorch as dp
                return dp.DataLoader(M._XorData(4, 2), batch_size=self.context.get_per_slot_batch_size())

        trial, hp, steps, unit = ShortVal, {"hidden_size": 4, "learning_rate": 0.05, "global_batch_size": 8, "shuffle": False}, 2, 2
        cfg = {"optimizations": {"average_training_metrics": True}}
    elif case in ("mnist_gpu", "mnist_gpu_overlap"):
        from determined_b200.models import vision_trials as VT

        trial, hp, steps, unit = VT.MNistTrial, {"global_batch_size": 64, "learning_rate": 1.0, "dropout1": 0.0, "dropout2": 0.0,
                                                 "overlap_grad_allreduce": case.endswith("overlap")}, 4, 4
        cfg = {"data": {"train_size": 1024, "validation_size": 128}, "optimizations": {"average_training_metrics": True}}
    else:
        raise SystemExit(f"unknown case {case}")
    tv = TrainAndValidate()
    controller = make_controller(trial, hp, tv.send(steps=steps, validation_freq=steps, scheduling_unit=unit), exp_config=cfg)
    controller.run()
    params = torch.cat([p.detach().float().reshape(-1).cpu() for m in controller.context.models for p in m.parameters()])
    gathered = controller.context.comm.allgather_object(params.tolist())
    if rank == 0:
        ov = getattr(controller.context, "_overlap", None)
        report = {"overlap_launches": len(ov.launch_log) if ov is not None else 0, "training": tv.training_metrics, "validation": tv.validation_metrics, "params_by_rank": gathered,
                  "world": controller.context.comm.size,
                  "comm": controller.context.comm.describe() if hasattr(controller.context.comm, "describe") else {}}
        with open(out_path, "w") as f:
            json.dump(report, f, default=float)
    import torch.distributed as dist

    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
