"""Energy accounting for kserve-vllm-mini run directories (drop-in package name).

`energy.collector` keeps the reference's CLI (`collect`, `integrate`) and importable helpers;
the sample arithmetic - window clip, stable sort by timestamp, trapezoidal Wh, idle tax,
per-request interval join - runs on the GPU (kserve-vllm-mini_b200/energy_api.py).
"""

__all__ = ["__version__"]
__version__ = "0.1.0"
