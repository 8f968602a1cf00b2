"""Seeded synthetic workloads: the catalogs, pools and job batches of BASELINE.json's configs.

Used by ``bench.py``, ``__graft_entry__.smoke()``, the tests and the tools; not part of the product
package (``dstack_b200`` never imports it).
"""
