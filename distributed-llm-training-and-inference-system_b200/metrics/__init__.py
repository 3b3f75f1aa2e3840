"""Metrics layer — same exports as reference llmctl/metrics/__init__.py:3-30, plus roofline reporting."""
from .health import (HealthManager, HealthReport, HealthStatus, InferenceHealthMonitor, SystemHealthMonitor,  # noqa: F401
                     TrainingHealthMonitor, get_health_manager, setup_health_monitoring)
from .observability import (MetricsCollector, ObservabilityManager, OpenTelemetryExporter, PrometheusExporter,  # noqa: F401
                            get_observability_manager, setup_observability)
from .roofline import RooflinePoint, measured_peaks, peak_flops, peak_hbm_bytes_per_s  # noqa: F401
