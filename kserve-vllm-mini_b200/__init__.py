"""kvmini_b200 - B200-native run analysis for kserve-vllm-mini run directories.

Importable as `kvmini_b200` (see ../kvmini_b200.py; the directory name carries a hyphen).
Sub-modules:
  columns   requests.csv -> columnar log (host ingest)
  engine    device buffers + launches of libkvmini_b200.so (ctypes, no CPU fallback)
  report    raw device results -> results.json fields (reference key order)
  cluster   kubectl / Prometheus glue (host I/O, unchanged semantics)
  analyze_api / energy_api / cost_api   the reference's importable functions and CLIs
  synth     seeded synthetic inputs for the BASELINE.json configs
"""

__version__ = "0.1.0"
