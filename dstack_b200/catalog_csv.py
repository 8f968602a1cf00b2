"""Bulk ingest of gpuhunt's offline catalog format.

gpuhunt ships one CSV per provider inside ``catalog.zip`` (``contributing/GPUHUNT.md:18-27,93-95``)
with the columns

    instance_name,location,price,cpu,memory,gpu_count,gpu_name,gpu_memory,spot,disk_size[,gpu_vendor]

and turns every row into a ``CatalogItem`` before filtering it.  Here a provider's CSV goes
straight to :class:`~dstack_b200.encode.OfferColumns` with vectorised parsing -- no Python object
per row -- so a multi-million-row catalog can be encoded and uploaded in seconds.  Empty cells
mean ``None`` (no GPU name / memory / disk size / vendor), ``spot`` is ``True``/``False``.

``GpuCatalog.add_csv`` (in :mod:`dstack_b200.offers`) is the row-object counterpart used when the
caller wants ``CatalogItem`` objects back from ``query``.
"""

from __future__ import annotations

import io
from typing import List, Optional, Union

import numpy as np

from .encode import VENDOR_CODES, OfferColumns, ShapeTable

COLUMNS = ["instance_name", "location", "price", "cpu", "memory", "gpu_count", "gpu_name", "gpu_memory",
           "spot", "disk_size", "gpu_vendor"]


def _factorize(series):
    import pandas as pd

    codes, names = pd.factorize(series)  # first-appearance order
    return codes.astype(np.int32), [str(x) for x in names]


def columns_from_catalog_csv(source: Union[str, bytes, io.IOBase], *, backend_code: int = 0) -> OfferColumns:
    """One provider's CSV (text, bytes, path-like file object) -> ``OfferColumns``.

    Rows keep their file order (gpuhunt stores each provider's rows sorted by price)."""
    import pandas as pd

    if isinstance(source, bytes):
        source = io.BytesIO(source)
    elif isinstance(source, str) and "\n" in source:
        source = io.StringIO(source)
    text_cols = {c: str for c in ("instance_name", "location", "gpu_name", "spot", "gpu_vendor")}
    df = pd.read_csv(source, dtype=text_cols, keep_default_na=False, na_values={c: [""] for c in
                     ("price", "cpu", "memory", "gpu_count", "gpu_memory", "disk_size")})
    missing = [c for c in COLUMNS[:-1] if c not in df.columns]
    if missing:
        raise ValueError(f"catalog CSV lacks columns {missing}")
    n = len(df)
    price = df["price"].to_numpy(dtype=np.float64)
    cpu = df["cpu"].fillna(0).to_numpy(dtype=np.int64)
    memory = df["memory"].fillna(0.0).to_numpy(dtype=np.float64)
    gpu_count = df["gpu_count"].fillna(0).to_numpy(dtype=np.int64)
    gpu_memory = df["gpu_memory"].to_numpy(dtype=np.float64)
    disk = df["disk_size"].to_numpy(dtype=np.float64)
    spot = df["spot"].str.strip().str.lower().isin(["true", "1"]).to_numpy()
    vendor_txt = df["gpu_vendor"].str.lower() if "gpu_vendor" in df.columns else pd.Series([""] * n)
    vendor_codes, vendor_names = pd.factorize(vendor_txt)
    unknown = set(vendor_names) - set(k for k in VENDOR_CODES if k) - {""}
    if unknown:
        raise ValueError(f"unknown gpu_vendor values {sorted(unknown)}")
    vendor = np.array([VENDOR_CODES[v or None] for v in vendor_names], dtype=np.int8)[vendor_codes] \
        if n else np.zeros(0, np.int8)

    # distinct resource shapes (everything the resource predicates read)
    shape_key = pd.DataFrame({"cpu": cpu, "memory": memory, "disk": np.nan_to_num(disk, nan=-1.0),
                              "gpu_count": gpu_count, "vendor": vendor, "gpu_name": df["gpu_name"],
                              "gpu_memory": np.nan_to_num(gpu_memory, nan=-1.0)})
    inv = shape_key.groupby(list(shape_key.columns), sort=False).ngroup().to_numpy()
    first = np.full(int(inv.max()) + 1 if n else 0, n, dtype=np.int64)
    np.minimum.at(first, inv, np.arange(n))
    gname = df["gpu_name"].to_numpy(dtype=object)
    shapes = ShapeTable(
        cpus=cpu[first], memory_gib=memory[first], disk_gib=disk[first], gpu_count=gpu_count[first],
        gpu_vendor=vendor[first], gpu_name=[(g if g != "" else None) for g in gname[first]],
        gpu_memory_gib=gpu_memory[first])
    region_id, regions = _factorize(df["location"])
    itype_id, itypes = _factorize(df["instance_name"])
    code = np.full(n, backend_code, dtype=np.int8)
    return OfferColumns(shapes=shapes, shape_id=np.asarray(inv).reshape(-1).astype(np.int32), price=price, spot=spot,
                        available=np.ones(n, bool), backend=code, src_backend=code.copy(),
                        region_id=region_id, regions=regions, itype_id=itype_id, itypes=itypes)


def catalog_items_from_csv(text: str, provider: Optional[str] = None) -> List["object"]:
    """Row-object reader (``CatalogItem`` per row), same cell conventions as above."""
    import csv

    from ._gpuhunt import AcceleratorVendor, CatalogItem

    def cell(v, load=None):
        if v is None or v == "":
            return None
        return load(v) if load else v

    items = []
    for row in csv.DictReader(io.StringIO(text)):
        vendor = cell(row.get("gpu_vendor"))
        items.append(CatalogItem(
            instance_name=cell(row["instance_name"]), location=cell(row["location"]),
            price=cell(row["price"], float), cpu=cell(row["cpu"], int), memory=cell(row["memory"], float),
            gpu_count=cell(row["gpu_count"], int) or 0, gpu_name=cell(row["gpu_name"]),
            gpu_memory=cell(row["gpu_memory"], float), spot=cell(row["spot"], lambda s: s.strip().lower() in ("true", "1")),
            disk_size=cell(row["disk_size"], float),
            gpu_vendor=None if vendor is None else AcceleratorVendor.cast(vendor), provider=provider))
    return items
