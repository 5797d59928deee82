#!/usr/bin/env python3
"""Cost estimator: cost_per_request / cost_per_1k_tokens (+ cold/warm split) -> results.json.

Drop-in for kserve-vllm-mini's cost_estimator.py (same command line, same keys); the
reductions over the request log run on the GPU (kserve-vllm-mini_b200/cost_api.py).

Usage:
  python cost_estimator.py --run-dir runs/<id> --namespace <ns> --service <isvc> --cost-file cost.yaml
"""

import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

try:
    import yaml  # noqa: F401
except Exception:
    print("ERROR: Missing 'pyyaml'. Install with: pip install pyyaml", file=sys.stderr)
    sys.exit(2)

from kvmini_b200.cost_api import (  # noqa: E402,F401
    UnitPricing,
    calculate_cold_warm_costs,
    estimator_main as main,
    load_pricing,
    parse_k8s_quantity,
)

if __name__ == "__main__":
    main()
