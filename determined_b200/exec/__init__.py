"""Process entrypoints launched by the master: harness (trial runner), gc_checkpoints, tensorboard."""
