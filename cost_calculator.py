#!/usr/bin/env python3
"""Cost of 1K tokens from a `latency_ms status` text file and a GPU hourly price.

Drop-in for kserve-vllm-mini's cost_calculator.py; the masked mean runs on the GPU.

Usage:
  python cost_calculator.py RESULTS_FILE GPU_HOURLY_COST [--requests-per-1k-tokens N]
"""

import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from kvmini_b200.cost_api import calculator_main as main  # noqa: E402

if __name__ == "__main__":
    main()
