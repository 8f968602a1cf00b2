"""Host-side mirror of the reference's models that sit on the offer-matching path.

Only what the matcher reads or returns is modelled.  Field names, defaults and parsing
rules follow the reference so the functions in :mod:`dstack_b200.offers` and
:mod:`dstack_b200.pools` are drop-in; every consumer in this package is duck-typed, so
the reference's own pydantic objects work unchanged as inputs too.

Reference (``src/dstack/_internal/...``):
  * job-predicate source: ``core/models/resources.py:19-292`` (``Range``, ``Memory``,
    ``ComputeCapability``, ``GPUSpec``, ``DiskSpec``, ``ResourcesSpec``; defaults cpu ``2..``,
    memory ``8GB..``, gpu count ``1..1``, disk ``100GB..`` at ``:125-127,250``)
  * offer-row source: ``core/models/instances.py:14-145``
  * ``Requirements`` ``core/models/runs.py:148-162``; ``JobProvisioningData`` ``:201-232``
  * filter inputs: ``core/models/profiles.py:115-222``; ``BackendType``
    ``core/models/backends/base.py:24-39``
"""

from __future__ import annotations

import enum
import math
from typing import Any, Dict, Generic, List, Optional, Tuple, TypeVar, Union

try:  # pydantic 2.x ships the 1.x API the reference is written against
    from pydantic.v1 import BaseModel, root_validator, validator
    from pydantic.v1.generics import GenericModel
except ImportError:  # pragma: no cover - genuine pydantic 1.x
    from pydantic import BaseModel, root_validator, validator  # type: ignore
    from pydantic.generics import GenericModel  # type: ignore

from ._gpuhunt import AcceleratorVendor

# ----------------------------------------------------------------------------- enums


class BackendType(str, enum.Enum):
    AWS = "aws"
    AZURE = "azure"
    CUDO = "cudo"
    DATACRUNCH = "datacrunch"
    DSTACK = "dstack"
    GCP = "gcp"
    KUBERNETES = "kubernetes"
    LAMBDA = "lambda"
    LOCAL = "local"
    REMOTE = "remote"
    NEBIUS = "nebius"
    OCI = "oci"
    RUNPOD = "runpod"
    TENSORDOCK = "tensordock"
    VASTAI = "vastai"
    VULTR = "vultr"


# capability lists, reference core/backends/__init__.py:3-29
BACKENDS_WITH_MULTINODE_SUPPORT = [
    BackendType.AWS, BackendType.AZURE, BackendType.GCP,
    BackendType.REMOTE, BackendType.OCI, BackendType.VULTR,
]
BACKENDS_WITH_CREATE_INSTANCE_SUPPORT = [
    BackendType.AWS, BackendType.DSTACK, BackendType.AZURE, BackendType.CUDO,
    BackendType.DATACRUNCH, BackendType.GCP, BackendType.LAMBDA, BackendType.OCI,
    BackendType.TENSORDOCK, BackendType.VULTR,
]
BACKENDS_WITH_PLACEMENT_GROUPS_SUPPORT = [BackendType.AWS]
BACKENDS_WITH_RESERVATION_SUPPORT = [BackendType.AWS]


class InstanceAvailability(enum.Enum):
    UNKNOWN = "unknown"
    AVAILABLE = "available"
    NOT_AVAILABLE = "not_available"
    NO_QUOTA = "no_quota"
    IDLE = "idle"
    BUSY = "busy"

    def is_available(self) -> bool:
        # instances.py:126-131
        return self in _AVAILABLE_STATES


_AVAILABLE_STATES = frozenset(
    (InstanceAvailability.UNKNOWN, InstanceAvailability.AVAILABLE, InstanceAvailability.IDLE)
)


class InstanceRuntime(enum.Enum):
    SHIM = "shim"
    RUNNER = "runner"


class InstanceStatus(str, enum.Enum):
    PENDING = "pending"
    PROVISIONING = "provisioning"
    IDLE = "idle"
    BUSY = "busy"
    TERMINATING = "terminating"
    TERMINATED = "terminated"

    def is_available(self) -> bool:
        return self in (InstanceStatus.IDLE, InstanceStatus.BUSY)


class SpotPolicy(str, enum.Enum):
    SPOT = "spot"
    ONDEMAND = "on-demand"
    AUTO = "auto"


class CreationPolicy(str, enum.Enum):
    REUSE = "reuse"
    REUSE_OR_CREATE = "reuse-or-create"


class CoreModel(BaseModel):
    """The reference builds a strict request class and a lenient ``__response__`` twin with
    ``pydantic_duality`` (``core/models/common.py:14-15``).  Here one lenient class plays
    both roles; ``__response__`` / ``__request__`` point back at it so call sites written
    as ``Model.__response__.parse_raw(...)`` (``server/services/pools.py:499``) keep working."""

    class Config:
        extra = "ignore"

    def __init_subclass__(cls, **kw):
        super().__init_subclass__(**kw)
        cls.__response__ = cls
        cls.__request__ = cls


# ----------------------------------------------------------------------------- resource specs

N = TypeVar("N", bound=Union[int, float])


class Range(GenericModel, Generic[N]):
    """Closed interval with optional ends; ``"a..b"``, ``"a.."``, ``"..b"`` or a scalar."""

    min: Optional[N]
    max: Optional[N]

    class Config:
        extra = "forbid"

    @classmethod
    def __get_validators__(cls):
        yield cls._from_text
        yield cls.validate

    @classmethod
    def _from_text(cls, value: Any) -> Any:
        if isinstance(value, str) and ".." in value:
            lo, hi = value.replace(" ", "").split("..")
            return {"min": lo or None, "max": hi or None}
        if isinstance(value, (str, int, float)):
            return {"min": value, "max": value}
        return value

    @root_validator()
    def _check_bounds(cls, values):
        lo, hi = values.get("min"), values.get("max")
        if lo is None and hi is None:
            raise ValueError("Invalid empty range: ..")
        if lo is not None and hi is not None and lo > hi:
            raise ValueError(f"Invalid range order: {lo}..{hi}")
        return values

    def __str__(self) -> str:
        lo = "" if self.min is None else self.min
        hi = "" if self.max is None else self.max
        return str(lo) if lo == hi else f"{lo}..{hi}"

    def intersect(self, other: "Range") -> Optional["Range"]:
        """resources.py:59-73.  ``None`` ends are infinities; empty overlap gives ``None``."""
        lo = max(-math.inf if self.min is None else self.min,
                 -math.inf if other.min is None else other.min)
        hi = min(math.inf if self.max is None else self.max,
                 math.inf if other.max is None else other.max)
        if lo > hi:
            return None
        return Range(min=None if math.isinf(lo) else lo, max=None if math.isinf(hi) else hi)


class Memory(float):
    """Gigabytes as a float; accepts ``MB``/``GB``/``TB`` suffixes (resources.py:76-101)."""

    _UNITS = {"tb": 1024.0, "gb": 1.0, "mb": 1.0 / 1024}

    @classmethod
    def __get_validators__(cls):
        yield cls.parse

    @classmethod
    def parse(cls, value: Any) -> "Memory":
        if isinstance(value, (int, float)):
            return cls(value)
        if isinstance(value, str):
            text = value.replace(" ", "").lower()
            unit = text[-2:]
            if unit == "gb":
                return cls(text[:-2])
            if unit == "tb":
                return cls(float(text[:-2]) * 1024)
            if unit == "mb":
                return cls(float(text[:-2]) / 1024)
            return cls(text)
        raise ValueError(f"Invalid memory size: {value}")

    def __repr__(self) -> str:
        return f"{self:g}GB"


class ComputeCapability(Tuple[int, int]):
    @classmethod
    def __get_validators__(cls):
        yield cls.validate

    @classmethod
    def validate(cls, value: Any) -> Tuple[int, int]:
        if isinstance(value, float):
            value = str(value)
        if isinstance(value, str):
            value = value.strip().split(".")
        if isinstance(value, (tuple, list)) and len(value) == 2:
            return int(value[0]), int(value[1])
        raise ValueError(f"Invalid compute capability: {value}")

    def __str__(self) -> str:
        return f"{self[0]}.{self[1]}"


DEFAULT_CPU_COUNT = Range[int](min=2)
DEFAULT_MEMORY_SIZE = Range[Memory](min=Memory.parse("8GB"))
DEFAULT_GPU_COUNT = Range[int](min=1, max=1)


def _vendor_token(text: str) -> AcceleratorVendor:
    text = text.lower()
    return AcceleratorVendor.GOOGLE if text == "tpu" else AcceleratorVendor.cast(text)


class GPUSpec(CoreModel):
    """``"A10,A10G:2"``, ``"24GB..80GB:1..8"``, ``"MI300X:AMD"`` ... (resources.py:130-232)."""

    vendor: Optional[AcceleratorVendor] = None
    name: Optional[List[str]] = None
    count: Range[int] = DEFAULT_GPU_COUNT
    memory: Optional[Range[Memory]] = None
    total_memory: Optional[Range[Memory]] = None
    compute_capability: Optional[ComputeCapability] = None

    class Config:
        extra = "forbid"

    @classmethod
    def __get_validators__(cls):
        yield cls._from_text
        yield cls.validate

    @classmethod
    def _from_text(cls, value: Any) -> Any:
        if isinstance(value, int):
            value = str(value)
        if not isinstance(value, str):
            return value
        fields: Dict[str, Any] = {}

        def put(key: str, val: Any) -> None:
            if key in fields:
                raise ValueError(f"GPU spec {key} conflict: {value}")
            fields[key] = val

        for token in value.replace(" ", "").split(":"):
            if not token:
                raise ValueError(f"GPU spec contains empty token: {value}")
            try:
                vendor = _vendor_token(token)
            except ValueError:
                vendor = None
            if vendor is not None:
                put("vendor", vendor)
            elif token[0].isalpha():  # a GPU name starts with a letter
                names = token.split(",")
                if not all(names):
                    raise ValueError(f"GPU name can not be empty: {value}")
                put("name", names)
            elif any(ch.isalpha() for ch in token):  # a memory size carries a unit
                put("memory", token)
            else:
                put("count", token)
        return fields

    @validator("name", pre=True)
    def _strip_tpu_prefix(cls, value: Any) -> Any:
        if value is None:
            return None
        items = value if isinstance(value, list) else [value]
        return [n[4:] if isinstance(n, str) and n.startswith("tpu-") else n for n in items]

    @validator("vendor", pre=True)
    def _cast_vendor(cls, value: Any) -> Any:
        if value is None or isinstance(value, AcceleratorVendor):
            return value
        if isinstance(value, str):
            return _vendor_token(value)
        raise TypeError(f"Unsupported type: {value!r}")


class DiskSpec(CoreModel):
    size: Range[Memory]

    class Config:
        extra = "forbid"

    @classmethod
    def __get_validators__(cls):
        yield cls._from_scalar
        yield cls.validate

    @classmethod
    def _from_scalar(cls, value: Any) -> Any:
        return {"size": value} if isinstance(value, (str, int, float)) else value


DEFAULT_DISK = DiskSpec(size=Range[Memory](min=Memory.parse("100GB"), max=None))


def pretty_resources(cpus=None, memory=None, gpu_count=None, gpu_name=None, gpu_memory=None,
                     total_gpu_memory=None, compute_capability=None, disk_size=None) -> str:
    """One-line resource summary; feeds ``Resources.description`` (utils/common.py:79-125)."""
    out: List[str] = []
    if cpus is not None:
        out.append(f"{cpus}xCPU")
    if memory is not None:
        out.append(f"{memory}")
    if gpu_count:
        detail = [f"{gpu_memory}"] if gpu_memory is not None else []
        if total_gpu_memory is not None:
            detail.append(f"total {total_gpu_memory}")
        if compute_capability is not None:
            detail.append(f"{compute_capability}")
        text = f"{gpu_count}x{gpu_name or 'GPU'}"
        out.append(text + (f" ({', '.join(detail)})" if detail else ""))
    if disk_size:
        out.append(f"{disk_size} (disk)")
    return ", ".join(out)


class ResourcesSpec(CoreModel):
    cpu: Range[int] = DEFAULT_CPU_COUNT
    memory: Range[Memory] = DEFAULT_MEMORY_SIZE
    shm_size: Optional[Memory] = None
    gpu: Optional[GPUSpec] = None
    disk: Optional[DiskSpec] = DEFAULT_DISK

    class Config:
        extra = "forbid"

    def pretty_format(self) -> str:
        kw: Dict[str, Any] = dict(cpus=self.cpu, memory=self.memory)
        if self.gpu:
            g = self.gpu
            kw.update(gpu_name=",".join(g.name) if g.name else None, gpu_count=g.count,
                      gpu_memory=g.memory, total_gpu_memory=g.total_memory,
                      compute_capability=g.compute_capability)
        if self.disk:
            kw.update(disk_size=self.disk.size)
        return pretty_resources(**kw)


class Requirements(CoreModel):
    resources: ResourcesSpec
    max_price: Optional[float]
    spot: Optional[bool]
    reservation: Optional[str]

    def pretty_format(self, resources_only: bool = False) -> str:
        text = self.resources.pretty_format()
        if not resources_only:
            if self.spot is not None:
                text += f", {'spot' if self.spot else 'on-demand'}"
            if self.max_price is not None:
                text += f" under ${self.max_price:g} per hour"
        return text


# ----------------------------------------------------------------------------- offers


class Gpu(CoreModel):
    name: str
    memory_mib: int
    vendor: Optional[AcceleratorVendor] = None

    @root_validator(pre=True)
    def _fill_vendor(cls, values):
        # instances.py:21-36: a "tpu-" name prefix means Google; default vendor is NVIDIA
        name = values.get("name")
        tpu = bool(name) and name.startswith("tpu-")
        if tpu:
            values["name"] = name[4:]
        vendor = values.get("vendor")
        if vendor is None:
            values["vendor"] = AcceleratorVendor.GOOGLE if tpu else AcceleratorVendor.NVIDIA
        else:
            values["vendor"] = AcceleratorVendor.cast(vendor)
        return values


class Disk(CoreModel):
    size_mib: int


class Resources(CoreModel):
    cpus: int
    memory_mib: int
    gpus: List[Gpu]
    spot: bool
    disk: Disk = Disk(size_mib=102400)  # 100GB default kept for old stored offers
    description: str = ""

    def pretty_format(self, include_spot: bool = False) -> str:
        kw: Dict[str, Any] = {}
        if self.cpus > 0:
            kw["cpus"] = self.cpus
        if self.memory_mib > 0:
            kw["memory"] = f"{self.memory_mib / 1024:.0f}GB"
        if self.disk.size_mib > 0:
            kw["disk_size"] = f"{self.disk.size_mib / 1024:.1f}GB"
        if self.gpus:
            first = self.gpus[0]
            kw["gpu_name"] = first.name
            kw["gpu_count"] = len(self.gpus)
            if first.memory_mib > 0:
                kw["gpu_memory"] = f"{first.memory_mib / 1024:.0f}GB"
        text = pretty_resources(**kw)
        return text + ", SPOT" if include_spot and self.spot else text


class InstanceType(CoreModel):
    name: str
    resources: Resources


class InstanceOffer(CoreModel):
    backend: BackendType
    instance: InstanceType
    region: str
    price: float


class InstanceOfferWithAvailability(InstanceOffer):
    availability: InstanceAvailability
    instance_runtime: InstanceRuntime = InstanceRuntime.SHIM
    blocks: int = 1
    total_blocks: int = 1


class JobProvisioningData(CoreModel):
    backend: BackendType
    base_backend: Optional[BackendType] = None
    instance_type: InstanceType
    instance_id: str = ""
    hostname: Optional[str] = None
    region: str
    availability_zone: Optional[str] = None
    reservation: Optional[str] = None
    price: float = 0.0
    username: str = ""
    dockerized: bool = True

    def get_base_backend(self) -> BackendType:
        return self.backend if self.base_backend is None else self.base_backend


class Profile(CoreModel):
    """Matching-relevant subset of ``ProfileParams`` (profiles.py:115-222)."""

    name: str = "default"
    backends: Optional[List[BackendType]] = None
    regions: Optional[List[str]] = None
    instance_types: Optional[List[str]] = None
    reservation: Optional[str] = None
    spot_policy: Optional[SpotPolicy] = None
    max_price: Optional[float] = None
    creation_policy: Optional[CreationPolicy] = None
    pool_name: Optional[str] = None
    instance_name: Optional[str] = None
