"""Minimal `gpuhunt` data types used on the offer-matching path.

The reference depends on the third-party package ``gpuhunt`` (pinned
``gpuhunt>=0.0.19,<0.1.0`` at reference ``setup.py:37``) for ``QueryFilter``,
``CatalogItem``, ``AcceleratorVendor`` and the table of known NVIDIA GPUs that maps a
GPU name to its compute capability.  The package is not vendored in the reference tree
and is not installable in this environment, so the *data types* are defined here (they
carry no matching logic).  When the real ``gpuhunt`` is importable it is used instead so
that objects stay interchangeable with the reference's.

What is deliberately NOT here: ``gpuhunt.matches`` / ``Catalog.query``.  Those are the
predicate loop this project moves to the GPU (``csrc/dsmatch.cu``); the CPU restatement
lives in ``oracle/`` and is test infrastructure only.

Reference call sites of these types: ``core/backends/base/offers.py:96-146``,
``core/models/resources.py:131-232``, ``core/models/instances.py:14-36``.
"""

from __future__ import annotations

import enum
from dataclasses import dataclass
from typing import List, Optional, Tuple, Union

try:  # pragma: no cover - exercised only where the real package exists
    import gpuhunt as _real  # type: ignore

    AcceleratorVendor = _real.AcceleratorVendor
    QueryFilter = _real.QueryFilter
    CatalogItem = _real.CatalogItem
    KNOWN_NVIDIA_GPUS = list(_real.KNOWN_NVIDIA_GPUS)
    HAVE_REAL_GPUHUNT = True
except Exception:  # ModuleNotFoundError in this image
    HAVE_REAL_GPUHUNT = False

    class AcceleratorVendor(str, enum.Enum):
        NVIDIA = "nvidia"
        AMD = "amd"
        GOOGLE = "google"
        INTEL = "intel"

        @classmethod
        def cast(cls, value: Union[str, "AcceleratorVendor"]) -> "AcceleratorVendor":
            if isinstance(value, cls):
                return value
            if isinstance(value, str):
                try:
                    return cls(value.lower())
                except ValueError:
                    pass
            raise ValueError(f"Unknown accelerator vendor: {value!r}")

    @dataclass
    class QueryFilter:
        """Field names must equal ``Catalog.query`` keyword names: the reference calls
        ``catalog.query(**asdict(q))`` (``core/backends/base/offers.py:34``)."""

        provider: Optional[List[str]] = None
        min_cpu: Optional[int] = None
        max_cpu: Optional[int] = None
        min_memory: Optional[float] = None
        max_memory: Optional[float] = None
        min_gpu_count: Optional[int] = None
        max_gpu_count: Optional[int] = None
        gpu_vendor: Optional[AcceleratorVendor] = None
        gpu_name: Optional[List[str]] = None
        min_gpu_memory: Optional[float] = None
        max_gpu_memory: Optional[float] = None
        min_total_gpu_memory: Optional[float] = None
        max_total_gpu_memory: Optional[float] = None
        min_disk_size: Optional[int] = None
        max_disk_size: Optional[int] = None
        min_price: Optional[float] = None
        max_price: Optional[float] = None
        min_compute_capability: Optional[Tuple[int, int]] = None
        max_compute_capability: Optional[Tuple[int, int]] = None
        spot: Optional[bool] = None

    @dataclass
    class CatalogItem:
        instance_name: str
        location: str
        price: float
        cpu: int
        memory: float
        gpu_count: int
        gpu_name: Optional[str]
        gpu_memory: Optional[float]
        spot: bool
        disk_size: Optional[float]
        gpu_vendor: Optional[AcceleratorVendor] = None
        provider: Optional[str] = None

    @dataclass(frozen=True)
    class NvidiaGPUInfo:
        name: str
        memory: int
        compute_capability: Tuple[int, int]

    # name -> compute capability, as published by NVIDIA (the only use on this path is the
    # `min_compute_capability` predicate).
    _CC = {
        "A10": (8, 6), "A10G": (8, 6), "A16": (8, 6), "A30": (8, 0), "A40": (8, 6),
        "A100": (8, 0), "A2": (8, 6), "A4000": (8, 6), "A4500": (8, 6), "A5000": (8, 6),
        "A6000": (8, 6), "B200": (10, 0), "H100": (9, 0), "H100NVL": (9, 0),
        "H200": (9, 0), "L4": (8, 9), "L40": (8, 9), "L40S": (8, 9), "P100": (6, 0),
        "P4": (6, 1), "P40": (6, 1), "RTX2000Ada": (8, 9), "RTX3060": (8, 6),
        "RTX3070": (8, 6), "RTX3080": (8, 6), "RTX3090": (8, 6), "RTX4000Ada": (8, 9),
        "RTX4070": (8, 9), "RTX4080": (8, 9), "RTX4090": (8, 9), "RTX5000Ada": (8, 9),
        "RTX6000": (7, 5), "RTX6000Ada": (8, 9), "RTX8000": (7, 5), "T4": (7, 5),
        "V100": (7, 0), "K80": (3, 7), "M60": (5, 2),
    }
    _MEM = {
        "A10": 24, "A10G": 24, "A16": 16, "A30": 24, "A40": 48, "A100": 80, "A2": 16,
        "A4000": 16, "A4500": 20, "A5000": 24, "A6000": 48, "B200": 180, "H100": 80,
        "H100NVL": 94, "H200": 141, "L4": 24, "L40": 48, "L40S": 48, "P100": 16, "P4": 8,
        "P40": 24, "RTX2000Ada": 16, "RTX3060": 12, "RTX3070": 8, "RTX3080": 10,
        "RTX3090": 24, "RTX4000Ada": 20, "RTX4070": 12, "RTX4080": 16, "RTX4090": 24,
        "RTX5000Ada": 32, "RTX6000": 24, "RTX6000Ada": 48, "RTX8000": 48, "T4": 16,
        "V100": 32, "K80": 12, "M60": 8,
    }
    KNOWN_NVIDIA_GPUS = [NvidiaGPUInfo(n, _MEM[n], cc) for n, cc in sorted(_CC.items())]


def compute_capability_of(gpu_name: Optional[str]) -> Optional[Tuple[int, int]]:
    """Case-insensitive lookup of a GPU's compute capability; ``None`` when unknown
    (an unknown capability fails a ``min_compute_capability`` bound)."""
    if gpu_name is None:
        return None
    return _CC_BY_LOWER.get(gpu_name.lower())


_CC_BY_LOWER = {g.name.lower(): tuple(g.compute_capability) for g in KNOWN_NVIDIA_GPUS}
