"""``llmctl hw probe`` / ``llmctl hw benchmark`` — discover the machine and *measure* its limits.

Surface of reference llmctl/cli/commands/hw.py:133-345 (``probe --emit --format --verbose``,
``benchmark --component memory|compute|network --duration``; profile sections ``system/cpu/memory/gpu/
interconnect/limits``).  The reference fills ``[limits]`` with A100 constants (312 TFLOP/s, 1935 GB/s,
600/32 GB/s: hw.py:179-184) and benchmarks numpy on the CPU; here, when GPUs are present, the limits are
measured: bf16 GEMM FLOP/s (tcgen05 kernel and cuBLAS), HBM copy bandwidth, per-pair P2P bandwidth and
NVLS/multicast availability (BASELINE.json item 4).  ``detected_at`` holds a timestamp (the reference
stores the cwd there: hw.py:173).
"""
from __future__ import annotations

import json
import os
import platform
import socket
import subprocess
import time
from pathlib import Path
from typing import Any, Dict, List, Optional

import psutil
import tomli_w
import typer
from rich.table import Table

from ..common import console, default_command

app = typer.Typer(help="Hardware discovery and profiling")

# nominal per-GPU constants used only when nothing can be measured
_NOMINAL = {
    "B200": dict(flops=2.25e15, hbm=7700.0), "H100": dict(flops=0.99e15, hbm=3350.0),
    "A100": dict(flops=312e12, hbm=1935.0),
}


def get_cpu_info() -> Dict[str, Any]:
    brand = platform.processor() or "unknown"
    try:
        import cpuinfo
        brand = cpuinfo.get_cpu_info().get("brand_raw", brand)
    except Exception:  # noqa: BLE001
        pass
    freq = psutil.cpu_freq()
    return {"brand": brand, "cores": psutil.cpu_count(logical=False) or 0, "threads": psutil.cpu_count() or 0,
            "frequency_mhz": float(freq.max or freq.current) if freq else 0.0, "architecture": platform.machine()}


def get_memory_info() -> Dict[str, Any]:
    vm = psutil.virtual_memory()
    return {"total_gb": round(vm.total / 2 ** 30, 2), "available_gb": round(vm.available / 2 ** 30, 2),
            "used_gb": round(vm.used / 2 ** 30, 2), "percentage": vm.percent}


def get_gpu_info() -> Dict[str, Any]:
    info: Dict[str, Any] = {"count": 0, "driver_version": "unknown", "cuda_version": "unknown",
                            "total_memory_gb": 0.0, "devices": []}
    try:
        import torch
        info["cuda_version"] = str(torch.version.cuda)
        if torch.cuda.is_available():
            info["count"] = torch.cuda.device_count()
            for i in range(info["count"]):
                p = torch.cuda.get_device_properties(i)
                info["devices"].append({"id": i, "name": p.name, "memory_gb": round(p.total_memory / 2 ** 30, 2),
                                        "compute_capability": f"{p.major}.{p.minor}",
                                        "multiprocessors": p.multi_processor_count})
                info["total_memory_gb"] += p.total_memory / 2 ** 30
            info["total_memory_gb"] = round(info["total_memory_gb"], 2)
    except Exception:  # noqa: BLE001
        pass
    try:
        import pynvml
        pynvml.nvmlInit()
        v = pynvml.nvmlSystemGetDriverVersion()
        info["driver_version"] = v.decode() if isinstance(v, bytes) else str(v)
    except Exception:  # noqa: BLE001
        pass
    return info


def detect_interconnect() -> Dict[str, Any]:
    out = {"intra_node": "PCIe", "inter_node": "Ethernet", "topology": "unknown"}
    try:
        r = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=10)
        if r.returncode == 0:
            if "NV" in r.stdout:
                out["intra_node"] = "NVLink"
                links = [tok for tok in r.stdout.split() if tok.startswith("NV") and tok[2:].isdigit()]
                if links:
                    out["topology"] = f"NVSwitch all-to-all ({max(links, key=lambda t: int(t[2:]))} per pair)"
            elif "PIX" in r.stdout or "PHB" in r.stdout:
                out["topology"] = "PCIe"
    except (OSError, subprocess.SubprocessError):
        pass
    try:
        if Path("/sys/class/infiniband").exists() and any(Path("/sys/class/infiniband").iterdir()):
            out["inter_node"] = "InfiniBand"
        else:
            r = subprocess.run(["ibstat"], capture_output=True, timeout=5)
            if r.returncode == 0:
                out["inter_node"] = "InfiniBand"
    except (OSError, subprocess.SubprocessError):
        pass
    return out


# ---- measurements -----------------------------------------------------------------------------------
def measure_gemm_tflops(n: int = 8192, iters: int = 10, native: bool = True) -> Optional[float]:
    import torch
    if not torch.cuda.is_available():
        return None
    a = torch.randn(n, n, device="cuda").to(torch.bfloat16)
    b = torch.randn(n, n, device="cuda").to(torch.bfloat16)
    if native:
        from ...ops import functional as ops
        fn = lambda: ops.gemm_raw(a, b)  # noqa: E731
    else:
        fn = lambda: torch.matmul(a, b.t())  # noqa: E731
    for _ in range(3):
        fn()
    best = float("inf")
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return 2.0 * n ** 3 / (best * 1e-3) / 1e12


def measure_hbm_gbps(nbytes: int = 2 << 30, iters: int = 10) -> Optional[float]:
    import torch
    if not torch.cuda.is_available():
        return None
    a = torch.empty(nbytes // 2, dtype=torch.bfloat16, device="cuda").normal_()
    b = torch.empty_like(a)
    for _ in range(2):
        b.copy_(a)
    best = float("inf")
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); b.copy_(a); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return 2.0 * nbytes / (best * 1e-3) / 1e9


def measure_p2p_matrix(nbytes: int = 256 << 20) -> Optional[List[List[float]]]:
    import torch
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    if n < 2:
        return None
    mat = [[0.0] * n for _ in range(n)]
    for i in range(n):
        src = torch.empty(nbytes, dtype=torch.uint8, device=f"cuda:{i}")
        for j in range(n):
            if i == j:
                continue
            dst = torch.empty(nbytes, dtype=torch.uint8, device=f"cuda:{j}")
            dst.copy_(src); torch.cuda.synchronize(j)
            with torch.cuda.device(j):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); dst.copy_(src); e1.record(); torch.cuda.synchronize(j)
                mat[i][j] = round(nbytes / (e0.elapsed_time(e1) * 1e-3) / 1e9, 1)
    return mat


def multicast_supported() -> Optional[bool]:
    try:
        from cuda.bindings import driver as cu
        cu.cuInit(0)
        err, dev = cu.cuDeviceGet(0)
        err, v = cu.cuDeviceGetAttribute(cu.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev)
        return bool(v) if int(err) == 0 else None
    except Exception:  # noqa: BLE001
        return None


def build_profile(measure: bool = True) -> Dict[str, Any]:
    gpu = get_gpu_info()
    inter = detect_interconnect()
    n = gpu["count"]
    name = gpu["devices"][0]["name"] if gpu["devices"] else ""
    nominal = next((v for k, v in _NOMINAL.items() if k in name), _NOMINAL["B200"])
    nvlink = inter["intra_node"] == "NVLink"
    limits: Dict[str, Any] = {"estimated_flops": max(n, 1) * nominal["flops"], "memory_bw_gbps": max(n, 1) * nominal["hbm"],
                              "intra_node_bw_gbps": 900.0 if nvlink else 32.0,
                              "inter_node_bw_gbps": 200.0 if inter["inter_node"] == "InfiniBand" else 10.0,
                              "measured": False}
    if measure and n > 0:
        try:
            tf_native = measure_gemm_tflops(native=True)
            tf_lib = measure_gemm_tflops(native=False)
            hbm = measure_hbm_gbps()
            limits.update(measured=True, gemm_tflops_per_gpu=round(max(tf_native or 0, tf_lib or 0), 1),
                          gemm_tflops_tcgen05_kernel=round(tf_native or 0, 1), gemm_tflops_cublas=round(tf_lib or 0, 1),
                          hbm_gbps_per_gpu=round(hbm or 0, 1), estimated_flops=n * max(tf_native or 0, tf_lib or 0) * 1e12,
                          memory_bw_gbps=n * (hbm or 0))
            p2p = measure_p2p_matrix()
            if p2p:
                inter["p2p_matrix_gbps"] = p2p
                vals = [v for row in p2p for v in row if v > 0]
                limits["intra_node_bw_gbps"] = round(sum(vals) / len(vals), 1)
            mc = multicast_supported()
            if mc is not None:
                inter["nvls_multicast"] = mc
        except Exception as e:  # noqa: BLE001
            limits["measure_error"] = f"{type(e).__name__}: {e}"[:200]
    return {
        "system": {"hostname": socket.gethostname(), "os": f"{platform.system()} {platform.release()}",
                   "python_version": platform.python_version(), "detected_at": time.strftime("%Y-%m-%dT%H:%M:%S%z"),
                   "cwd": os.getcwd()},
        "cpu": get_cpu_info(), "memory": get_memory_info(), "gpu": gpu, "interconnect": inter, "limits": limits,
    }


def _drop_none(o):
    if isinstance(o, dict):
        return {k: _drop_none(v) for k, v in o.items() if v is not None}
    if isinstance(o, list):
        return [_drop_none(v) for v in o]
    return o


def probe(
    emit: Optional[Path] = typer.Option(None, "--emit", help="Write the hardware profile to this file"),
    format: str = typer.Option("toml", "--format", help="Output format: toml | json"),
    verbose: bool = typer.Option(False, "--verbose", "-v", help="Show per-device details"),
    no_measure: bool = typer.Option(False, "--no-measure", help="Skip the GPU measurements (constants only)"),
):
    """Probe hardware and generate a hardware profile."""
    console.print("[bold blue]Hardware Profile[/bold blue]")
    prof = build_profile(measure=not no_measure)
    t = Table(title="System"); t.add_column("Property"); t.add_column("Value")
    for k, v in prof["system"].items():
        t.add_row(k, str(v))
    console.print(t)
    t = Table(title="CPU"); t.add_column("Property"); t.add_column("Value")
    for k, v in prof["cpu"].items():
        t.add_row(k, str(v))
    console.print(t)
    t = Table(title="Memory"); t.add_column("Property"); t.add_column("Value")
    for k, v in prof["memory"].items():
        t.add_row(k, str(v))
    console.print(t)
    g = prof["gpu"]
    t = Table(title="GPU"); t.add_column("Property"); t.add_column("Value")
    for k in ("count", "driver_version", "cuda_version", "total_memory_gb"):
        t.add_row(k, str(g[k]))
    console.print(t)
    if verbose and g["devices"]:
        t = Table(title="GPU devices")
        for c in ("id", "name", "memory_gb", "compute_capability", "multiprocessors"):
            t.add_column(c)
        for d in g["devices"]:
            t.add_row(*(str(d[c]) for c in ("id", "name", "memory_gb", "compute_capability", "multiprocessors")))
        console.print(t)
    t = Table(title="Interconnect"); t.add_column("Property"); t.add_column("Value")
    for k, v in prof["interconnect"].items():
        t.add_row(k, str(v) if k != "p2p_matrix_gbps" else f"{len(v)}x{len(v)} matrix")
    console.print(t)
    t = Table(title="Limits" + (" (measured)" if prof["limits"].get("measured") else " (nominal constants)"))
    t.add_column("Property"); t.add_column("Value")
    for k, v in prof["limits"].items():
        t.add_row(k, f"{v:.4g}" if isinstance(v, float) else str(v))
    console.print(t)
    if emit:
        emit = Path(emit)
        emit.parent.mkdir(parents=True, exist_ok=True)
        clean = _drop_none(prof)
        if format == "json":
            emit.write_text(json.dumps(clean, indent=2))
        else:
            emit.write_text(tomli_w.dumps(clean))
        console.print(f"[green]Hardware profile written to {emit}[/green]")
    return prof


def benchmark(
    component: str = typer.Option("memory", "--component", help="memory | compute | network"),
    duration: int = typer.Option(10, "--duration", help="Upper bound on seconds spent measuring"),
):
    """Run hardware micro-benchmarks (HBM copy, bf16 GEMM, GPU-to-GPU copy)."""
    import torch
    console.print(f"[bold blue]Hardware benchmark: {component}[/bold blue]")
    have_gpu = torch.cuda.is_available()
    if component == "memory":
        if have_gpu:
            console.print(f"HBM copy bandwidth: [green]{measure_hbm_gbps():.1f} GB/s[/green] (read+write bytes, CUDA events)")
        else:
            import numpy as np
            a = np.ones(50_000_000, dtype=np.float32); b = np.empty_like(a)
            t0 = time.perf_counter(); n = 0
            while time.perf_counter() - t0 < min(duration, 2):
                np.copyto(b, a); n += 1
            console.print(f"Host memory copy bandwidth: [green]{2 * a.nbytes * n / (time.perf_counter() - t0) / 1e9:.1f} GB/s[/green]")
    elif component == "compute":
        if have_gpu:
            console.print(f"bf16 GEMM 8192^3 (tcgen05 kernel): [green]{measure_gemm_tflops(native=True):.1f} TFLOP/s[/green]")
            console.print(f"bf16 GEMM 8192^3 (cuBLAS):         [green]{measure_gemm_tflops(native=False):.1f} TFLOP/s[/green]")
        else:
            a = torch.randn(1024, 1024); t0 = time.perf_counter(); n = 0
            while time.perf_counter() - t0 < min(duration, 2):
                a @ a; n += 1
            console.print(f"CPU fp32 GEMM 1024^3: [green]{2 * 1024 ** 3 * n / (time.perf_counter() - t0) / 1e9:.1f} GFLOP/s[/green]")
    elif component == "network":
        m = measure_p2p_matrix() if have_gpu else None
        if m:
            for i, row in enumerate(m):
                console.print(f"GPU{i} -> " + "  ".join(f"{v:7.1f}" for v in row) + "  GB/s")
        else:
            console.print("[yellow]network: fewer than 2 GPUs visible; nothing to measure (use `llmctl bench comms` under torchrun)[/yellow]")
    else:
        console.print(f"[red]Unknown component: {component}[/red]")
        raise typer.Exit(1)


app.command("probe")(probe)
app.command("benchmark")(benchmark)
default_command(app, probe)
