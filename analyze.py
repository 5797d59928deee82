#!/usr/bin/env python3
"""Analyze a run directory (requests.csv) and write results.json + requests_classified.csv.

Drop-in for kserve-vllm-mini's analyze.py - same command line, same files, same
importable functions - with the per-request aggregation running as sm_100a CUDA kernels
(see kserve-vllm-mini_b200/analyze_api.py and DESIGN.md).

Usage:
  python analyze.py --run-dir runs/<id> --namespace <ns> --service <isvc> [--prom-url URL]
"""

import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from kvmini_b200.analyze_api import (  # noqa: E402,F401
    compute_histograms,
    main,
    percentile,
    read_requests_csv,
    window_bounds,
)

if __name__ == "__main__":
    main()
