"""Exceptions raised by the harness (reference: ``harness/determined/errors.py``)."""


class InvalidExperimentException(Exception):
    """The experiment definition (model code or configuration) cannot work; restarting will not help."""


class InvalidModelException(InvalidExperimentException):
    pass


class InvalidConfigurationException(InvalidExperimentException):
    pass


class InvalidDataTypeException(InvalidExperimentException):
    """The trial handed the harness input data of a kind it cannot iterate."""


class InternalException(Exception):
    """A bug in the harness itself."""


class SkipWorkloadException(Exception):
    """Raised from inside a workload handler to answer the workload with ``Skipped`` (non-chief ranks)."""


class WorkerError(Exception):
    """A worker process of a multi-slot trial died or reported a failure."""
