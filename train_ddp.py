#!/usr/bin/env python
"""Entry point with the reference's command line (reference train_ddp.py:215-228, README.md:59-79):

    torchrun --nproc_per_node=2 train_ddp.py --epochs 3 --batch_size 64
    python train_ddp.py --world_size 2            # spawns the workers itself

``--world_size`` is the flag the reference documents but never implemented.
Rerunning the same command resumes from the newest ``./checkpoints/epoch_N.pt``.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from ddp_b200.parallel.launcher import launch  # noqa: E402
from ddp_b200.trainer import ddp_train  # noqa: E402


def main(argv=None):
    parser = argparse.ArgumentParser(description="DDP Training Script (B200-native)")
    parser.add_argument("--epochs", type=int, default=10, help="Number of training epochs")
    parser.add_argument("--batch_size", type=int, default=32, help="Batch size for training (per rank)")
    parser.add_argument("--world_size", type=int, default=None,
                        help="Processes to spawn (default: torchrun's WORLD_SIZE, else GPU count, else 2)")
    parser.add_argument("--max_restarts", type=int, default=0,
                        help="Self-spawn mode: restart the worker group this many times if a rank dies "
                             "(each restart resumes from the newest checkpoint); under torchrun use its "
                             "own --max-restarts")
    args = parser.parse_args(argv)
    launch(ddp_train, args.world_size, args=(args.epochs, args.batch_size), max_restarts=args.max_restarts)


if __name__ == "__main__":
    main()
