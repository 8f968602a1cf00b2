"""gpuhunt-style CSV ingest: the vectorised reader and the row-object
reader must describe the same catalog -- checked through the encoder and the emulated matcher."""

import numpy as np
import pytest

import device_emulator as emu
from dstack_b200 import models as M
from dstack_b200.catalog_csv import catalog_items_from_csv, columns_from_catalog_csv
from dstack_b200.encode import encode_table, pack_jobs
from dstack_b200.ingest import JobFilter, columns_from_catalog_items, job_columns
from dstack_b200.offers import requirements_to_query_filter
from oracle import gpuhunt_restated

HEADER = "instance_name,location,price,cpu,memory,gpu_count,gpu_name,gpu_memory,spot,disk_size,gpu_vendor"


def _csv(n, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    gpus = [("A100", 80.0, "nvidia"), ("T4", 16.0, "nvidia"), ("MI300X", 192.0, "amd"), ("v5litepod-8", "", "google"),
            ("H100", 80.0, ""), ("a100", 40.0, "nvidia")]
    rows = []
    price = 0.05
    for i in range(n):
        price = round(price + float(rng.uniform(0, 0.02)), 4)
        g = gpus[int(rng.integers(0, len(gpus)))]
        cnt = int(rng.choice([0, 0, 1, 2, 8]))
        rows.append(",".join(str(x) for x in [
            f"type-{int(rng.integers(0, 40))}", f"loc-{int(rng.integers(0, 9))}", price,
            int(rng.choice([2, 4, 8, 48, 96])), float(rng.choice([7.5, 16, 15.25, 61.0, 384])), cnt,
            g[0] if cnt else "", g[1] if cnt else "", "True" if rng.random() < 0.4 else "False",
            ["", "", "100.0", "500"][int(rng.integers(0, 4))], g[2] if cnt else ""]))
    return HEADER + "\n" + "\n".join(rows) + "\n"


REQS = [{"resources": {}}, {"resources": {"gpu": "A100"}}, {"resources": {"gpu": "16GB..:1..8"}},
        {"resources": {"gpu": "amd:8", "disk": "200GB.."}}, {"resources": {"gpu": "tpu:v5litepod-8"}},
        {"resources": {"cpu": "48..", "memory": "61GB.."}, "max_price": 3.0, "spot": True},
        {"resources": {"gpu": {"compute_capability": "8.0", "count": "1..8"}}, "spot": False}]


@pytest.mark.parametrize("n,seed", [(700, 1), (33, 2)])
def test_vectorised_csv_ingest_equals_row_objects(n, seed):
    text = _csv(n, seed)
    items = catalog_items_from_csv(text, provider="aws")
    assert len(items) == n and items[0].provider == "aws"
    t_vec = encode_table(columns_from_catalog_csv(text))
    t_obj = encode_table(columns_from_catalog_items(items))
    assert (t_vec.n_rows, t_vec.split_row) == (t_obj.n_rows, t_obj.split_row)
    assert np.array_equal(t_vec.perm, t_obj.perm)
    assert np.array_equal(np.nan_to_num(t_vec.price, nan=-1), np.nan_to_num(t_obj.price, nan=-1))
    reqs = [M.Requirements.parse_obj({"max_price": None, "spot": None, **r}) for r in REQS]
    filters = [JobFilter(query=requirements_to_query_filter(r)) for r in reqs]
    m_vec = emu.match_matrix(t_vec, pack_jobs(t_vec, job_columns(t_vec, filters)))
    m_obj = emu.match_matrix(t_obj, pack_jobs(t_obj, job_columns(t_obj, filters)))
    assert np.array_equal(m_vec, m_obj)
    # and both agree with the restated predicate applied row by row
    for j, r in enumerate(reqs):
        q = requirements_to_query_filter(r)
        want = [i for i, it in enumerate(items) if gpuhunt_restated.matches(it, q)]
        assert sorted(int(t_vec.perm[row]) for row in np.flatnonzero(m_vec[j])) == want, REQS[j]
    assert m_vec.any()


def test_csv_errors():
    with pytest.raises(ValueError, match="lacks columns"):
        columns_from_catalog_csv("instance_name,price\nx,1\n")
    bad = HEADER + "\nx,loc,1.0,2,8,1,A100,80,False,,voodoo\n"
    with pytest.raises(ValueError, match="gpu_vendor"):
        columns_from_catalog_csv(bad)
