"""Host-side cluster glue the analysis needs: kubectl pod facts and Prometheus averages.

None of this is on the device path (it is subprocess / HTTP I/O on a handful of pods);
it is kept so the drop-in scripts behave like the reference's when a cluster is present
and degrade the same way when it is not:

  analyze.py:358-395  cold-start times  -> [] on any failure
  analyze.py:192-309  Prometheus vector / matrix averages, utilisation queries
  analyze.py:312-355  cache hit ratio (PromQL, else `kubectl logs | grep`)
  cost_estimator.py:159-198,343-396  pod list (NOT guarded: a missing kubectl raises),
                      node GPU label, resource-seconds
"""

from __future__ import annotations

import datetime as dt
import json
import re
import subprocess
import urllib.parse
import urllib.request
from typing import Any, Dict, List, Optional, Tuple

ISVC_LABEL = "serving.kserve.io/inferenceservice"


def _sh(cmd: List[str]) -> str:
    return subprocess.check_output(cmd, text=True)


def _parse_ts(s: str) -> dt.datetime:
    return dt.datetime.fromisoformat(s.replace("Z", "+00:00"))


def kubectl_pods(namespace: str, service: str) -> Dict[str, Any]:
    """`kubectl get pods -l <isvc label> -o json`; raises like the reference when kubectl fails."""
    return json.loads(_sh(["kubectl", "get", "pods", "-n", namespace, "-l",
                           f"{ISVC_LABEL}={service}", "-o", "json"]))


def cold_start_times(namespace: str, service: str, start: float, end: float) -> List[float]:
    """First container `running.startedAt` per pod that falls inside [start, end]."""
    try:
        pods = kubectl_pods(namespace, service)
    except Exception:
        return []
    w0 = dt.datetime.fromtimestamp(start, tz=dt.timezone.utc)
    w1 = dt.datetime.fromtimestamp(end, tz=dt.timezone.utc)
    found: List[float] = []
    for pod in pods.get("items", []):
        for cs in pod.get("status", {}).get("containerStatuses", []):
            started = cs.get("state", {}).get("running", {}).get("startedAt")
            if started:
                ts = _parse_ts(started)
                if w0 <= ts <= w1:
                    found.append(ts.timestamp())
                    break  # one cold start per pod
    return sorted(found)


# --------------------------------------------------------------------------- Prometheus
def prom_query(prom_url: str, query: str, start: Optional[float] = None, end: Optional[float] = None,
               step: Optional[int] = None) -> Dict[str, Any]:
    if start is not None and end is not None and step is not None:
        path, params = "/api/v1/query_range", {"query": query, "start": str(start), "end": str(end),
                                               "step": str(step)}
    else:
        path, params = "/api/v1/query", {"query": query}
    url = urllib.parse.urljoin(prom_url, path) + "?" + urllib.parse.urlencode(params)
    with urllib.request.urlopen(url, timeout=15) as resp:
        return json.loads(resp.read())


def prom_vector_avg(result: Dict[str, Any]) -> Optional[float]:
    try:
        vals = [float(s["value"][1]) for s in result["data"]["result"]]
        return sum(vals) / len(vals) if vals else None
    except Exception:
        return None


def prom_matrix_timeavg(result: Dict[str, Any]) -> Optional[float]:
    try:
        pts = [float(v) for series in result["data"]["result"] for _t, v in series.get("values", [])]
        return sum(pts) / len(pts) if pts else None
    except Exception:
        return None


def prom_matrix_timeavg_body(body: bytes) -> Optional[float]:
    """prom_matrix_timeavg (analyze.py:233-247) straight from the response BODY: large range-query
    results (hours of 1 kHz power samples) are scanned into columns by the native parser instead
    of becoming one Python object per point; the mean is then taken exactly like the reference
    does (`sum(pts) / len(pts)` over the points in series order).  Falls back to the json module."""
    try:
        from . import energy_api
        cols = energy_api.prom_result_columns(body)
    except Exception:
        cols = None
    if cols is None:
        try:
            return prom_matrix_timeavg(json.loads(body))
        except Exception:
            return None
    pts: List[float] = []
    for _ts, vals in cols:
        pts.extend(vals.tolist())
    return sum(pts) / len(pts) if pts else None


def _first_avg(prom_url: str, queries: List[str]) -> Optional[float]:
    for q in queries:
        v = prom_vector_avg(prom_query(prom_url, q))
        if v is not None:
            return v
    return None


def utilization(prom_url: str, namespace: str, isvc: str, start: float, end: float
                ) -> Dict[str, Optional[float]]:
    """analyze.py:250-309: DCGM util / memory / power averaged over the window, container
    CPU cores and working-set bytes."""
    step = max(5, int((end - start) / 60))
    dur = f"{int(end - start)}s"
    sel = f'{{namespace="{namespace}",pod=~"{isvc}-predictor-.*"}}'
    csel = f'{{namespace="{namespace}",pod=~"{isvc}-predictor-.*",container!="",container!="POD"}}'

    def over(metrics):
        return [f"avg_over_time({m}{sel}[{dur}])" for m in metrics]

    return {
        "gpu_util_avg": _first_avg(prom_url, over(["DCGM_FI_DEV_GPU_UTIL", "nvidia_dcgm_gpu_utilization",
                                                   "nvidia_gpu_utilization"])),
        "gpu_mem_used_avg": _first_avg(prom_url, over(["DCGM_FI_DEV_FB_USED", "nvidia_dcgm_fb_used_bytes"])),
        "gpu_power_watts_avg": _first_avg(prom_url, over(["DCGM_FI_DEV_POWER_USAGE",
                                                          "nvidia_dcgm_power_usage_watts",
                                                          "nvidia_gpu_power_watts"])),
        "cpu_util_avg": prom_vector_avg(prom_query(
            prom_url, f"avg(rate(container_cpu_usage_seconds_total{csel}[2m]))", start, end, step)),
        "mem_used_avg": prom_vector_avg(prom_query(
            prom_url, f"avg(container_memory_working_set_bytes{csel})", start, end, step)),
    }


def cache_hit_ratio(prom_url: Optional[str], namespace: str, isvc: str, start: float, end: float
                    ) -> Optional[float]:
    """analyze.py:312-355."""
    if prom_url:
        dur = f"{int(end - start)}s"
        sel = f'{{namespace="{namespace}",pod=~"{isvc}-predictor-.*"}}'

        def rate(m):
            return f"sum(rate({m}{sel}[{dur}]))"

        for q in (f"{rate('vllm_prompt_cache_hits_total')} / {rate('vllm_prompt_cache_requests_total')}",
                  f"{rate('vllm_cache_hits_total')} / ({rate('vllm_cache_hits_total')} + "
                  f"{rate('vllm_cache_misses_total')})"):
            try:
                v = prom_vector_avg(prom_query(prom_url, q))
                if v is not None and 0 <= v <= 1:
                    return v
            except Exception:
                pass
    try:
        out = _sh(["bash", "-lc", f"kubectl -n {namespace} logs -l {ISVC_LABEL}={isvc} --tail=200 "
                                  "| grep -i 'cache' || true"])
        hits = sum("cache hit" in ln.lower() for ln in out.splitlines())
        misses = sum("cache miss" in ln.lower() for ln in out.splitlines())
        if hits + misses > 0:
            return hits / (hits + misses)
    except Exception:
        pass
    return None


# --------------------------------------------------------------------------- cost inputs
_SUFFIX = (("Ei", 2**60), ("Pi", 2**50), ("Ti", 2**40), ("Gi", 2**30), ("Mi", 2**20), ("Ki", 2**10),
           ("E", 10**18), ("P", 10**15), ("T", 10**12), ("G", 10**9), ("M", 10**6), ("K", 10**3))


def parse_k8s_quantity(q: Optional[str]) -> float:
    """cost_estimator.py:48-83: '500m' -> 0.5 cores, '2Gi' -> bytes, plain number as is."""
    if not q:
        return 0.0
    if q.endswith("m"):
        return float(q[:-1]) / 1000.0
    for suf, mult in _SUFFIX:
        if q.endswith(suf):
            return float(q[: -len(suf)]) * mult
    try:
        return float(q)
    except Exception:
        return 0.0


def container_resources(container: dict, use_requests: bool) -> Tuple[float, float, float]:
    """(cpu cores, memory GiB, gpus) of one container spec (cost_estimator.py:216-235)."""
    rs = container.get("resources", {})
    chosen = rs.get("requests" if use_requests else "limits", {}) or {}
    cpu = parse_k8s_quantity(chosen.get("cpu"))
    mem_b = parse_k8s_quantity(chosen.get("memory"))
    mem_gib = mem_b / (2**30) if mem_b else 0.0
    gpus = 0.0
    for scope in (chosen, rs.get("limits", {}) or {}):
        val = scope.get("nvidia.com/gpu")
        if val:
            try:
                gpus = float(val)
                break
            except Exception:
                pass
    return cpu, mem_gib, gpus


def node_gpu_label(namespace: str, pod: str) -> Optional[str]:
    """cost_estimator.py:177-198."""
    try:
        pod_json = json.loads(_sh(["kubectl", "get", "pod", pod, "-n", namespace, "-o", "json"]))
        node = pod_json["spec"].get("nodeName")
        if not node:
            return None
        labels = json.loads(_sh(["kubectl", "get", "node", node, "-o", "json"])
                            ).get("metadata", {}).get("labels", {})
        for key in ("nvidia.com/gpu.product", "nvidia.com/gpu.product.name", "nvidia.com/gpu.family"):
            if key in labels:
                return labels[key]
    except Exception:
        return None
    return None


def resource_seconds(pods: dict, window_start: float, window_end: float) -> Dict[str, float]:
    """cost_estimator.py:343-396, including its 0.2 GiB-second initial memory term."""
    w0 = dt.datetime.fromtimestamp(window_start, tz=dt.timezone.utc)
    w1 = dt.datetime.fromtimestamp(window_end, tz=dt.timezone.utc)
    cpu_s, mem_s, gpu_s = 0.0, 0.2, 0.0
    for pod in pods.get("items", []):
        status = pod.get("status", {})
        spec_by_name = {c.get("name"): c for c in pod.get("spec", {}).get("containers", [])}
        first_start: Optional[dt.datetime] = None
        last_end: Optional[dt.datetime] = None
        for cs in status.get("containerStatuses", []):
            state = cs.get("state", {})
            if state.get("running", {}).get("startedAt"):
                t = _parse_ts(state["running"]["startedAt"])
                first_start = min(first_start, t) if first_start else t
            if state.get("terminated", {}).get("finishedAt"):
                t = _parse_ts(state["terminated"]["finishedAt"])
                last_end = max(last_end, t) if last_end else t
        if not first_start and status.get("startTime"):
            first_start = _parse_ts(status["startTime"])
        if not last_end:
            last_end = w1
        if not first_start or not last_end:
            continue
        a, b = max(first_start, w0), min(last_end, w1)
        if b <= a:
            continue
        seconds = (b - a).total_seconds()
        for cs in status.get("containerStatuses", []):
            cpu, mem_gib, gpus = container_resources(spec_by_name.get(cs.get("name"), {}), True)
            cpu_s += cpu * seconds
            mem_s += mem_gib * seconds
            gpu_s += gpus * seconds
    return {"cpu_core_seconds": cpu_s, "mem_gib_seconds": mem_s, "gpu_seconds": gpu_s}


def pick_gpu_cost(gpu_default: float, gpu_map: Dict[str, float], label: Optional[str]) -> float:
    """cost_estimator.py:201-213."""
    if not label:
        return gpu_default
    if label in gpu_map:
        return gpu_map[label]
    key = "-".join(t for t in re.split(r"[^A-Za-z0-9]+", label) if t).lower()
    for name, price in gpu_map.items():
        if key in name.lower() or name.lower() in key:
            return price
    return gpu_default
