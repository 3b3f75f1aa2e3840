"""Observability: background system sampler, Prometheus exporter, OTLP span/metric exporter, and the manager
singleton the train loop and the server report into.

Metric names are the reference's (reference llmctl/metrics/observability.py:237-251):
``llmctl_training_loss``, ``llmctl_training_learning_rate``, ``llmctl_training_steps_total``,
``llmctl_training_tokens_per_second``, ``llmctl_inference_latency_seconds``, ``llmctl_inference_requests_total``,
``llmctl_inference_active_requests``, ``llmctl_inference_tokens_per_second``, ``llmctl_system_cpu_percent``,
``llmctl_system_memory_percent``, ``llmctl_system_gpu_memory_used_gb{gpu_id}``; added:
``llmctl_training_grad_norm``, ``llmctl_training_mfu``, ``llmctl_training_step_seconds``,
``llmctl_inference_ttft_seconds``, ``llmctl_system_gpu_utilization{gpu_id}`` (real NVML value — the
reference hard-codes 0.0).  Counters and the histogram are actually updated.
Exporters start only when asked (``setup_observability(prometheus_port=..., otlp_endpoint=...)``).
"""
from __future__ import annotations

import threading
import time
from collections import deque
from dataclasses import dataclass, field
from typing import Any, Deque, Dict, List, Optional

import psutil


@dataclass
class SystemSample:
    timestamp: float
    cpu_percent: float
    memory_percent: float
    memory_used_gb: float
    gpu_memory_used_gb: List[float] = field(default_factory=list)
    gpu_utilization: List[float] = field(default_factory=list)
    gpu_power_w: List[float] = field(default_factory=list)
    gpu_sm_clock_mhz: List[float] = field(default_factory=list)
    network_mbps: float = 0.0
    disk_mbps: float = 0.0


class MetricsCollector:
    """Daemon thread sampling CPU / memory / GPU (NVML) / network / disk into 1000-deep histories."""

    def __init__(self, interval: float = 1.0, history: int = 1000):
        self.interval = interval
        self.system: Deque[SystemSample] = deque(maxlen=history)
        self.training: Deque[Dict[str, Any]] = deque(maxlen=history)
        self.inference: Deque[Dict[str, Any]] = deque(maxlen=history)
        self._lock = threading.Lock()
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None
        self._nvml = None
        self._last_net = self._last_disk = None

    def _init_nvml(self):
        if self._nvml is None:
            try:
                import pynvml
                pynvml.nvmlInit()
                self._nvml = pynvml
            except Exception:  # noqa: BLE001
                self._nvml = False
        return self._nvml

    def sample_once(self) -> SystemSample:
        now = time.time()
        vm = psutil.virtual_memory()
        s = SystemSample(now, psutil.cpu_percent(interval=None), vm.percent, vm.used / 2 ** 30)
        nv = self._init_nvml()
        if nv:
            try:
                for i in range(nv.nvmlDeviceGetCount()):
                    h = nv.nvmlDeviceGetHandleByIndex(i)
                    s.gpu_memory_used_gb.append(nv.nvmlDeviceGetMemoryInfo(h).used / 2 ** 30)
                    s.gpu_utilization.append(float(nv.nvmlDeviceGetUtilizationRates(h).gpu))
                    s.gpu_power_w.append(nv.nvmlDeviceGetPowerUsage(h) / 1000.0)
                    s.gpu_sm_clock_mhz.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
            except Exception:  # noqa: BLE001
                pass
        try:
            net = psutil.net_io_counters()
            if self._last_net:
                t0, b0 = self._last_net
                s.network_mbps = 8e-6 * (net.bytes_sent + net.bytes_recv - b0) / max(now - t0, 1e-6)
            self._last_net = (now, net.bytes_sent + net.bytes_recv)
            dk = psutil.disk_io_counters()
            if dk:
                if self._last_disk:
                    t0, b0 = self._last_disk
                    s.disk_mbps = 1e-6 * (dk.read_bytes + dk.write_bytes - b0) / max(now - t0, 1e-6)
                self._last_disk = (now, dk.read_bytes + dk.write_bytes)
        except Exception:  # noqa: BLE001
            pass
        with self._lock:
            self.system.append(s)
        return s

    def start(self):
        if self._thread and self._thread.is_alive():
            return
        self._stop.clear()
        self._thread = threading.Thread(target=self._run, daemon=True, name="llmctl-metrics")
        self._thread.start()

    def _run(self):
        while not self._stop.is_set():
            self.sample_once()
            self._stop.wait(self.interval)

    def stop(self):
        self._stop.set()
        if self._thread:
            self._thread.join(timeout=2)

    def record_training(self, rec: Dict[str, Any]):
        with self._lock:
            self.training.append(rec)

    def record_inference(self, rec: Dict[str, Any]):
        with self._lock:
            self.inference.append(rec)

    def latest(self) -> Dict[str, Any]:
        with self._lock:
            return {"system": self.system[-1] if self.system else None,
                    "training": self.training[-1] if self.training else None,
                    "inference": self.inference[-1] if self.inference else None}


class PrometheusExporter:
    def __init__(self, port: int = 8000, registry=None):
        from prometheus_client import REGISTRY, Counter, Gauge, Histogram
        self.port = port
        reg = registry or REGISTRY
        self.registry = reg
        G = lambda n, d, l=(): Gauge(n, d, list(l), registry=reg)  # noqa: E731
        self.training_loss = G("llmctl_training_loss", "Training loss")
        self.training_lr = G("llmctl_training_learning_rate", "Learning rate")
        self.training_steps = Counter("llmctl_training_steps_total", "Optimizer steps", registry=reg)
        self.training_tps = G("llmctl_training_tokens_per_second", "Training throughput")
        self.training_grad_norm = G("llmctl_training_grad_norm", "Global gradient norm")
        self.training_mfu = G("llmctl_training_mfu", "Model FLOPs utilisation vs measured peak")
        self.training_step_s = G("llmctl_training_step_seconds", "Step time (max over ranks)")
        self.inf_latency = Histogram("llmctl_inference_latency_seconds", "Request latency", registry=reg)
        self.inf_ttft = Histogram("llmctl_inference_ttft_seconds", "Time to first token", registry=reg)
        self.inf_requests = Counter("llmctl_inference_requests_total", "Requests served", ["status"], registry=reg)
        self.inf_active = G("llmctl_inference_active_requests", "Active requests")
        self.inf_tps = G("llmctl_inference_tokens_per_second", "Decode throughput")
        self.sys_cpu = G("llmctl_system_cpu_percent", "CPU percent")
        self.sys_mem = G("llmctl_system_memory_percent", "Memory percent")
        self.sys_gpu_mem = G("llmctl_system_gpu_memory_used_gb", "GPU memory used", ("gpu_id",))
        self.sys_gpu_util = G("llmctl_system_gpu_utilization", "GPU utilisation percent", ("gpu_id",))
        self._started = False
        self._last_step = 0

    def start(self):
        if not self._started and self.port:
            from prometheus_client import start_http_server
            start_http_server(self.port, registry=self.registry)
            self._started = True

    def update_system(self, s: SystemSample):
        self.sys_cpu.set(s.cpu_percent)
        self.sys_mem.set(s.memory_percent)
        for i, v in enumerate(s.gpu_memory_used_gb):
            self.sys_gpu_mem.labels(gpu_id=str(i)).set(v)
        for i, v in enumerate(s.gpu_utilization):
            self.sys_gpu_util.labels(gpu_id=str(i)).set(v)


def _endpoint_reachable(endpoint: str, timeout: float = 0.5) -> bool:
    import socket
    from urllib.parse import urlparse
    u = urlparse(endpoint)
    host, port = u.hostname or "localhost", u.port or (443 if u.scheme == "https" else 80)
    try:
        with socket.create_connection((host, port), timeout=timeout):
            return True
    except OSError:
        return False


class OpenTelemetryExporter:
    """OTLP/HTTP spans (one per training step / inference request) and two histograms."""

    def __init__(self, endpoint: str = "http://localhost:4318", service_name: str = "llmctl"):
        from opentelemetry import metrics as otm
        from opentelemetry import trace
        from opentelemetry.sdk.resources import Resource
        from opentelemetry.sdk.trace import TracerProvider
        from opentelemetry.sdk.trace.export import BatchSpanProcessor
        self.endpoint = endpoint.rstrip("/")
        if not _endpoint_reachable(self.endpoint):
            # presets ship the reference's default collector address; with nobody listening the exporters would print
            # a connection traceback for every batch — say it once and keep the spans in-process instead
            import warnings
            warnings.warn(f"OTLP collector {self.endpoint} is not reachable: telemetry export disabled for this run")
            raise ConnectionError(f"OTLP collector {self.endpoint} is not reachable")
        res = Resource.create({"service.name": service_name})
        self.provider = TracerProvider(resource=res)
        try:
            from opentelemetry.exporter.otlp.proto.http.trace_exporter import OTLPSpanExporter
            self.provider.add_span_processor(BatchSpanProcessor(OTLPSpanExporter(endpoint=f"{self.endpoint}/v1/traces")))
        except Exception:  # noqa: BLE001 - exporter package missing: spans stay in-process
            pass
        self.tracer = self.provider.get_tracer("llmctl")
        self.meter = None
        try:
            from opentelemetry.exporter.otlp.proto.http.metric_exporter import OTLPMetricExporter
            from opentelemetry.sdk.metrics import MeterProvider
            from opentelemetry.sdk.metrics.export import PeriodicExportingMetricReader
            reader = PeriodicExportingMetricReader(OTLPMetricExporter(endpoint=f"{self.endpoint}/v1/metrics"),
                                                   export_interval_millis=5000)
            self.meter = MeterProvider(resource=res, metric_readers=[reader]).get_meter("llmctl")
            self.loss_hist = self.meter.create_histogram("llmctl.training.loss")
            self.lat_hist = self.meter.create_histogram("llmctl.inference.latency")
        except Exception:  # noqa: BLE001
            self.meter = None

    def record_training_step(self, **attrs):
        with self.tracer.start_as_current_span("training_step") as span:
            for k, v in attrs.items():
                if isinstance(v, (int, float, str, bool)):
                    span.set_attribute(f"llmctl.{k}", v)
        if self.meter and "loss" in attrs:
            self.loss_hist.record(float(attrs["loss"]))

    def record_inference_request(self, **attrs):
        with self.tracer.start_as_current_span("inference_request") as span:
            for k, v in attrs.items():
                if isinstance(v, (int, float, str, bool)):
                    span.set_attribute(f"llmctl.{k}", v)
        if self.meter and "latency" in attrs:
            self.lat_hist.record(float(attrs["latency"]))


class ObservabilityManager:
    def __init__(self, prometheus_port: Optional[int] = None, otlp_endpoint: Optional[str] = None,
                 sample_interval: float = 1.0, start_sampler: bool = False):
        self.collector = MetricsCollector(sample_interval)
        self.prometheus: Optional[PrometheusExporter] = None
        self.otel: Optional[OpenTelemetryExporter] = None
        self._refresh: Optional[threading.Thread] = None
        self._stop = threading.Event()
        if prometheus_port is not None:
            self.enable_prometheus(prometheus_port)
        if otlp_endpoint:
            try:
                self.otel = OpenTelemetryExporter(otlp_endpoint)
            except Exception:  # noqa: BLE001
                self.otel = None
        if start_sampler:
            self.start()

    def enable_prometheus(self, port: Optional[int] = 8000, registry=None) -> PrometheusExporter:
        if self.prometheus is None:
            self.prometheus = PrometheusExporter(port or 0, registry)
            if port:
                self.prometheus.start()
        return self.prometheus

    def start(self):
        self.collector.start()
        if self._refresh is None or not self._refresh.is_alive():
            self._stop.clear()
            self._refresh = threading.Thread(target=self._refresh_loop, daemon=True, name="llmctl-prom-refresh")
            self._refresh.start()

    def _refresh_loop(self):
        while not self._stop.is_set():
            s = self.collector.latest()["system"]
            if s is not None and self.prometheus is not None:
                self.prometheus.update_system(s)
            self._stop.wait(5.0)

    def stop(self):
        self._stop.set()
        self.collector.stop()

    # ---- feeds -----------------------------------------------------------------------------------------
    def record_training_step(self, step: int, loss: float, learning_rate: float = 0.0, tokens_per_second: float = 0.0,
                             grad_norm: Optional[float] = None, step_time: Optional[float] = None,
                             flops: Optional[float] = None, mfu: Optional[float] = None, **extra):
        rec = {"step": step, "loss": loss, "learning_rate": learning_rate, "tokens_per_second": tokens_per_second,
               "grad_norm": grad_norm, "step_time": step_time, "flops": flops, "mfu": mfu, "time": time.time(), **extra}
        self.collector.record_training(rec)
        p = self.prometheus
        if p is not None:
            p.training_loss.set(loss)
            p.training_lr.set(learning_rate)
            p.training_tps.set(tokens_per_second)
            p.training_steps.inc(max(step - p._last_step, 0))
            p._last_step = step
            if grad_norm is not None:
                p.training_grad_norm.set(grad_norm)
            if mfu is not None:
                p.training_mfu.set(mfu)
            if step_time is not None:
                p.training_step_s.set(step_time)
        if self.otel is not None:
            self.otel.record_training_step(**{k: v for k, v in rec.items() if v is not None})

    def record_inference_request(self, latency: float, tokens: int = 0, ttft: Optional[float] = None,
                                 queue_depth: int = 0, error: bool = False, **extra):
        rec = {"latency": latency, "tokens": tokens, "ttft": ttft, "queue_depth": queue_depth, "error": error,
               "time": time.time(), **extra}
        self.collector.record_inference(rec)
        p = self.prometheus
        if p is not None:
            p.inf_latency.observe(latency)
            if ttft is not None:
                p.inf_ttft.observe(ttft)
            p.inf_requests.labels(status="error" if error else "ok").inc()
            p.inf_active.set(queue_depth)
            if latency > 0:
                p.inf_tps.set(tokens / latency)
        if self.otel is not None:
            self.otel.record_inference_request(**{k: v for k, v in rec.items() if v is not None})

    def get_metrics_summary(self) -> Dict[str, Any]:
        return self.collector.latest()


_global_manager: Optional[ObservabilityManager] = None


def get_observability_manager() -> ObservabilityManager:
    global _global_manager
    if _global_manager is None:
        _global_manager = ObservabilityManager()
    return _global_manager


def setup_observability(prometheus_port: Optional[int] = 8000, otlp_endpoint: Optional[str] = None,
                        start_sampler: bool = True, **kwargs) -> ObservabilityManager:
    global _global_manager
    _global_manager = ObservabilityManager(prometheus_port, otlp_endpoint, start_sampler=start_sampler, **kwargs)
    return _global_manager
