"""Health monitoring: system / training / inference monitors, a manager thread, alert callbacks and the JSON
report consumed by ``llmctl health drift``.

Public names and thresholds follow reference llmctl/metrics/health.py (system: cpu 80/95 %, memory 85/95 %,
disk 85/95 %, GPU memory 90/98 %; training: stall > 300 s, non-finite loss, grad norm outside (0.001, 100);
inference: error rate > 5 %, average latency > 10 s, queue > 100).  Differences: the disk and GPU-memory
thresholds actually fire (the reference stores those metrics under one name and looks them up under another,
SURVEY.md §2.10 D12), shared state is lock-protected, and the monitors are fed by the train loop and the
server (the reference never calls them).
"""
from __future__ import annotations

import json
import math
import threading
import time
from collections import deque
from dataclasses import dataclass, field
from enum import Enum
from typing import Callable, Dict, List, Optional

import psutil


class HealthStatus(Enum):
    HEALTHY = "healthy"
    WARNING = "warning"
    CRITICAL = "critical"
    UNKNOWN = "unknown"


_ORDER = {HealthStatus.HEALTHY: 0, HealthStatus.UNKNOWN: 1, HealthStatus.WARNING: 2, HealthStatus.CRITICAL: 3}


@dataclass
class HealthReport:
    component: str
    status: HealthStatus
    checks: Dict[str, bool] = field(default_factory=dict)
    metrics: Dict[str, float] = field(default_factory=dict)
    message: str = ""
    timestamp: float = field(default_factory=time.time)

    def to_dict(self) -> Dict:
        return {"component": self.component, "status": self.status.value, "checks": self.checks,
                "metrics": self.metrics, "message": self.message, "timestamp": self.timestamp}


def _grade(value: float, warn: float, crit: float) -> HealthStatus:
    if value >= crit:
        return HealthStatus.CRITICAL
    if value >= warn:
        return HealthStatus.WARNING
    return HealthStatus.HEALTHY


class SystemHealthMonitor:
    THRESHOLDS = {"cpu_usage": (80.0, 95.0), "memory_usage": (85.0, 95.0), "disk_space": (85.0, 95.0),
                  "gpu_memory": (90.0, 98.0)}

    def __init__(self, cpu_interval: float = 0.1, disk_path: str = "/"):
        self.cpu_interval, self.disk_path = cpu_interval, disk_path

    def collect(self) -> Dict[str, float]:
        m = {"cpu_usage": float(psutil.cpu_percent(interval=self.cpu_interval)),
             "memory_usage": float(psutil.virtual_memory().percent)}
        try:
            m["disk_space"] = float(psutil.disk_usage(self.disk_path).percent)
        except OSError:
            pass
        try:
            import torch
            if torch.cuda.is_available():
                worst = 0.0
                for i in range(torch.cuda.device_count()):
                    free, total = torch.cuda.mem_get_info(i)
                    worst = max(worst, 100.0 * (total - free) / total)
                m["gpu_memory"] = worst
        except Exception:  # noqa: BLE001
            pass
        return m

    def get_health_report(self) -> HealthReport:
        metrics = self.collect()
        status, checks, msgs = HealthStatus.HEALTHY, {}, []
        for name, (warn, crit) in self.THRESHOLDS.items():
            if name not in metrics:
                continue
            g = _grade(metrics[name], warn, crit)
            checks[name] = g == HealthStatus.HEALTHY
            if g != HealthStatus.HEALTHY:
                msgs.append(f"{name}={metrics[name]:.1f}% ({g.value})")
            if _ORDER[g] > _ORDER[status]:
                status = g
        return HealthReport("system", status, checks, metrics, "; ".join(msgs) or "all system checks passed")


class TrainingHealthMonitor:
    def __init__(self, stall_seconds: float = 300.0):
        self.stall_seconds = stall_seconds
        self._lock = threading.Lock()
        self.last_update: Optional[float] = None
        self.step, self.loss, self.grad_norm = 0, None, None
        self.history: deque = deque(maxlen=1000)

    def update_metrics(self, step: int, loss: float, grad_norm: Optional[float] = None) -> None:
        with self._lock:
            self.step, self.loss, self.grad_norm = step, loss, grad_norm
            self.last_update = time.time()
            self.history.append((self.last_update, step, loss, grad_norm))

    def get_health_report(self) -> HealthReport:
        with self._lock:
            last, loss, gn, step = self.last_update, self.loss, self.grad_norm, self.step
        if last is None:
            return HealthReport("training", HealthStatus.UNKNOWN, {}, {}, "no training metrics received")
        checks, msgs, status = {}, [], HealthStatus.HEALTHY
        stalled = time.time() - last > self.stall_seconds
        checks["progress"] = not stalled
        if stalled:
            status = HealthStatus.CRITICAL
            msgs.append(f"no step for {time.time() - last:.0f}s")
        finite = loss is not None and math.isfinite(loss)
        checks["loss_finite"] = finite
        if not finite:
            status = HealthStatus.CRITICAL
            msgs.append("loss is NaN/Inf")
        if gn is not None:
            ok = math.isfinite(gn) and 0.001 < gn < 100.0
            checks["grad_norm"] = ok
            if not ok:
                if _ORDER[HealthStatus.WARNING] > _ORDER[status]:
                    status = HealthStatus.WARNING
                msgs.append(f"grad_norm={gn:.4g} outside (0.001, 100)")
        metrics = {"step": float(step), "loss": float(loss) if loss is not None else float("nan"),
                   "seconds_since_update": time.time() - last}
        if gn is not None:
            metrics["grad_norm"] = float(gn)
        return HealthReport("training", status, checks, metrics, "; ".join(msgs) or "training is progressing")


class InferenceHealthMonitor:
    def __init__(self, window: int = 1000):
        self._lock = threading.Lock()
        self.latencies: deque = deque(maxlen=window)
        self.outcomes: deque = deque(maxlen=window)
        self.queue_size = 0
        self.active_requests = 0

    def record_request(self, latency: float, success: bool = True) -> None:
        with self._lock:
            self.latencies.append(latency)
            self.outcomes.append(bool(success))

    def update_queue(self, n: int) -> None:
        with self._lock:
            self.queue_size = n

    def update_active_requests(self, n: int) -> None:
        with self._lock:
            self.active_requests = n

    def get_health_report(self) -> HealthReport:
        with self._lock:
            lats, outs, q, act = list(self.latencies), list(self.outcomes), self.queue_size, self.active_requests
        if not outs:
            return HealthReport("inference", HealthStatus.UNKNOWN, {}, {"queue_size": float(q)}, "no requests served yet")
        err = 100.0 * (1.0 - sum(outs) / len(outs))
        avg = sum(lats) / len(lats)
        checks = {"error_rate": err < 5.0, "latency": avg < 10.0, "queue_size": q < 100}
        status = HealthStatus.HEALTHY
        msgs = []
        if not checks["error_rate"]:
            status = HealthStatus.CRITICAL
            msgs.append(f"error rate {err:.1f}%")
        for name, text in (("latency", f"avg latency {avg:.2f}s"), ("queue_size", f"queue {q}")):
            if not checks[name]:
                if _ORDER[HealthStatus.WARNING] > _ORDER[status]:
                    status = HealthStatus.WARNING
                msgs.append(text)
        return HealthReport("inference", status, checks,
                            {"error_rate": err, "avg_latency": avg, "queue_size": float(q), "active_requests": float(act),
                             "requests": float(len(outs))}, "; ".join(msgs) or "inference is healthy")


class HealthManager:
    def __init__(self, check_interval: float = 30.0):
        self.check_interval = check_interval
        self.monitors = {"system": SystemHealthMonitor(), "training": TrainingHealthMonitor(),
                         "inference": InferenceHealthMonitor()}
        self.alert_callbacks: List[Callable[[HealthReport], None]] = []
        self._thread: Optional[threading.Thread] = None
        self._stop = threading.Event()
        self.history: deque = deque(maxlen=1000)

    def add_alert_callback(self, cb: Callable[[HealthReport], None]) -> None:
        self.alert_callbacks.append(cb)

    def start_monitoring(self) -> None:
        if self._thread and self._thread.is_alive():
            return
        self._stop.clear()
        self._thread = threading.Thread(target=self._loop, daemon=True, name="llmctl-health")
        self._thread.start()

    def stop_monitoring(self) -> None:
        self._stop.set()
        if self._thread:
            self._thread.join(timeout=2)

    def _loop(self) -> None:
        while not self._stop.is_set():
            self._perform_health_checks()
            self._stop.wait(self.check_interval)

    def _perform_health_checks(self) -> Dict[str, HealthReport]:
        reports = self.get_current_health()
        self.history.append((time.time(), {k: r.status.value for k, r in reports.items()}))
        for r in reports.values():
            if r.status in (HealthStatus.WARNING, HealthStatus.CRITICAL):
                for cb in self.alert_callbacks:
                    try:
                        cb(r)
                    except Exception:  # noqa: BLE001
                        pass
        return reports

    def get_current_health(self, components: Optional[List[str]] = None) -> Dict[str, HealthReport]:
        names = components or list(self.monitors)
        return {n: self.monitors[n].get_health_report() for n in names if n in self.monitors}

    def get_overall_status(self) -> HealthStatus:
        reports = self.get_current_health()
        worst = HealthStatus.HEALTHY
        for r in reports.values():
            if r.status != HealthStatus.UNKNOWN and _ORDER[r.status] > _ORDER[worst]:
                worst = r.status
        return worst

    # feeds
    def update_training_metrics(self, step: int, loss: float, grad_norm: Optional[float] = None) -> None:
        self.monitors["training"].update_metrics(step, loss, grad_norm)

    def record_inference_request(self, latency: float, success: bool = True) -> None:
        self.monitors["inference"].record_request(latency, success)

    def update_inference_metrics(self, latency: float, error: bool = False, queue_size: int = 0) -> None:
        self.monitors["inference"].record_request(latency, not error)
        self.monitors["inference"].update_queue(queue_size)

    def update_active_requests(self, count: int) -> None:
        self.monitors["inference"].update_active_requests(count)

    def save_health_report(self, filepath: str) -> Dict:
        data = {n: r.to_dict() for n, r in self.get_current_health().items()}
        with open(filepath, "w") as f:
            json.dump(data, f, indent=2)
        return data


def create_default_alert_callback() -> Callable[[HealthReport], None]:
    def cb(report: HealthReport) -> None:
        tag = "CRITICAL" if report.status == HealthStatus.CRITICAL else "WARNING"
        print(f"[health] {tag}: {report.component} - {report.message}", flush=True)
    return cb


_global_health_manager: Optional[HealthManager] = None


def get_health_manager() -> HealthManager:
    """Global manager (created on first use so the train loop / server can always feed it)."""
    global _global_health_manager
    if _global_health_manager is None:
        _global_health_manager = HealthManager()
    return _global_health_manager


def setup_health_monitoring(**kwargs) -> HealthManager:
    global _global_health_manager
    _global_health_manager = HealthManager(**kwargs)
    _global_health_manager.add_alert_callback(create_default_alert_callback())
    return _global_health_manager


def compare_with_baseline(current: Dict, baseline: Dict, tolerance_pct: float = 10.0) -> List[Dict]:
    """Metric-by-metric drift of a health report against a saved one (``llmctl health drift``)."""
    drifts = []
    for comp, base in baseline.items():
        cur = current.get(comp, {})
        for name, bval in (base.get("metrics") or {}).items():
            cval = (cur.get("metrics") or {}).get(name)
            if cval is None or not isinstance(bval, (int, float)) or name in ("seconds_since_update", "step", "requests"):
                continue
            denom = abs(bval) if abs(bval) > 1e-9 else 1.0
            pct = 100.0 * abs(cval - bval) / denom
            drifts.append({"component": comp, "metric": name, "baseline": bval, "current": cval, "drift_pct": pct,
                           "drifted": pct > tolerance_pct})
    return drifts
