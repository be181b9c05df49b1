#!/usr/bin/env python3
"""Drop-in entry point at the reference's path (``benchmarking/train_harness.py``): same flags,
same result JSON, executed by the b200bench engine."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from b200bench.cli import main  # noqa: E402

if __name__ == "__main__":
    main()
