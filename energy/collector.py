#!/usr/bin/env python3
"""Energy collector & integrator (drop-in for kserve-vllm-mini's energy/collector.py).

  python energy/collector.py collect --namespace NS --service S --prom-url URL \
      --interval 1 --duration 120 --out runs/<id>/power.json
  python energy/collector.py integrate --run-dir runs/<id> [--power P] [--include-warmup] \
      [--idle-tax baseline|series] [--merge-results]

Same files in (requests.csv, power.json) and out (energy.json, results.json merge), same exit
codes; integration runs as sm_100a CUDA kernels.
"""

import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from kvmini_b200.energy_api import (  # noqa: E402,F401
    PROM_QUERIES,
    PowerSample,
    collect_power,
    integrate_energy,
    load_power_samples,
    main,
    merge_results,
    read_requests_csv,
    run_window_bounds,
    trapezoidal_wh,
    write_json,
)

if __name__ == "__main__":
    sys.exit(main())
