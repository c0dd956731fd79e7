"""B200-native data-parallel trainer (same capabilities/CLI/checkpoints as the reference
zahmedy/PyTorch-Distributed-Data-Parallel-DDP-Trainer; see SURVEY.md and DESIGN.md).

Sub-packages
  models/    SimpleCNN on a flat parameter arena (CPU oracle path + sm_100a kernel path)
  ops/       ctypes bindings of the native library (csrc/, include/ddp_b200.h)
  parallel/  process-group setup, launcher, sampler, symmetric-memory communicator, engine
  utils/     checkpoint I/O, logging/timing, synthetic + IDX datasets
"""
__version__ = "0.1.0"
