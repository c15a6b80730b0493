"""Stand-in for the third-party ``horovod`` package, which cannot be built offline (it needs MPI/Gloo + NCCL sources and a
network). Only ``horovod.torch`` is provided — see that module. Lives under ``baseline/shims`` because it exists solely so
that the UNMODIFIED reference harness can take its own ``hvd_config.use == True`` code path in the benchmark's reference arm;
it is not part of the product."""
__version__ = "0.19.5+nccl-standin"
