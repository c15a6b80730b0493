"""TensorBoard support: event-file tracking + sync to checkpoint storage, per-batch metric writer."""
from .build import BatchMetricWriter, TensorboardManager, build_metric_writer, build_tensorboard_manager, get_sync_path

__all__ = ["BatchMetricWriter", "TensorboardManager", "build_metric_writer", "build_tensorboard_manager", "get_sync_path"]
