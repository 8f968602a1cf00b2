"""Import the reference's own modules from ``/root/reference`` (build container only).

TEST INFRASTRUCTURE.  Used by ``tests/golden/make_golden.py`` to generate the committed
golden vectors and by the container-only cross-check tests; ``/root/reference`` does not
exist on the GPU box, so nothing marked ``gpu``, ``smoke()`` or ``bench.py`` calls this.

Recipe: alias ``pydantic`` to ``pydantic.v1`` (the reference targets
pydantic 1.x), stub ``pydantic_duality.DualBaseModel``, and provide ``gpuhunt`` from the
restated types + predicate.  With that the following import unmodified:
``core.models.{resources,instances,runs,profiles,backends.base}`` and
``core.backends.base.offers``.
"""

from __future__ import annotations

import importlib
import os
import sys
import types

REFERENCE_SRC = "/root/reference/src"


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_SRC, "dstack", "_internal"))


_loaded = None


def load():
    """Returns a namespace with the reference modules: ``offers``, ``resources``,
    ``instances``, ``runs``, ``profiles``, ``backends_base`` and the ``gpuhunt`` stand-in."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError("reference tree not present")

    import pydantic.v1 as pv1

    saved = {k: v for k, v in sys.modules.items() if k == "pydantic" or k.startswith("pydantic.")}
    sys.modules["pydantic"] = pv1
    for sub in ("generics", "fields", "main", "typing", "error_wrappers", "errors", "validators",
                "types", "class_validators", "json", "schema", "utils", "networks",
                "datetime_parse", "color", "tools", "config"):
        try:
            sys.modules[f"pydantic.{sub}"] = importlib.import_module(f"pydantic.v1.{sub}")
        except ImportError:
            pass

    duality = types.ModuleType("pydantic_duality")

    class DualBaseModel(pv1.BaseModel):
        def __init_subclass__(cls, **kw):
            super().__init_subclass__(**kw)
            cls.__response__ = cls
            cls.__request__ = cls

    duality.DualBaseModel = DualBaseModel
    sys.modules["pydantic_duality"] = duality

    from dstack_b200 import _gpuhunt as types_
    from oracle import gpuhunt_restated as pred

    gh = types.ModuleType("gpuhunt")
    gh.AcceleratorVendor = types_.AcceleratorVendor
    gh.QueryFilter = types_.QueryFilter
    gh.CatalogItem = types_.CatalogItem
    gh.KNOWN_NVIDIA_GPUS = types_.KNOWN_NVIDIA_GPUS
    gh.KNOWN_AMD_GPUS = []
    gh.KNOWN_TPUS = []
    gh.matches = pred.matches
    gh.Catalog = pred.Catalog
    gh.default_catalog = lambda: pred.Catalog()
    sys.modules["gpuhunt"] = gh

    sys.path.insert(0, REFERENCE_SRC)
    try:
        ns = types.SimpleNamespace(
            offers=importlib.import_module("dstack._internal.core.backends.base.offers"),
            resources=importlib.import_module("dstack._internal.core.models.resources"),
            instances=importlib.import_module("dstack._internal.core.models.instances"),
            runs=importlib.import_module("dstack._internal.core.models.runs"),
            profiles=importlib.import_module("dstack._internal.core.models.profiles"),
            backends_base=importlib.import_module("dstack._internal.core.models.backends.base"),
            gpuhunt=gh,
            pydantic=pv1,
        )
    finally:
        sys.path.remove(REFERENCE_SRC)
        # give the rest of the process its real pydantic back; the reference modules keep
        # their own bound references to pydantic.v1
        for k in [k for k in sys.modules if k == "pydantic" or k.startswith("pydantic.")]:
            if not k.startswith("pydantic.v1"):
                del sys.modules[k]
        sys.modules.update(saved)
    _loaded = ns
    return ns
