"""CPU oracle for the offer-matching path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``dstack_b200/`` may import this package.  It is used by ``tests/``,
``__graft_entry__.smoke()`` and the CPU-baseline / ``--impl reference`` legs of ``bench.py``
as the checker and as the timed CPU baseline, never as the product path.

Parity status (see DESIGN.md "Oracle"):
  * the reference's OWN half of the path (``requirements_to_query_filter``,
    ``offer_to_catalog_item``, ``match_requirements``, ``get_catalog_offers``,
    ``catalog_item_to_offer``, ``choose_disk_size_mib``, ``Range.intersect``, model parsing)
    is PINNED: ``tests/golden/*.json`` were produced by importing those functions from
    ``/root/reference`` (``tests/golden/make_golden.py``) and ``oracle.reference_path`` is
    checked against them.
  * ``gpuhunt.matches`` / ``Catalog.query`` (third party, ``gpuhunt>=0.0.19,<0.1.0``,
    reference ``setup.py:37``, source absent from ``/root/reference`` and not installable
    here) is restated from its published behaviour in ``oracle.gpuhunt_restated`` --
    PARITY UNPINNED for that part: no reference test holds a vector for it
    (SURVEY.md section 8c).  The known-answer table in ``tests/test_oracle_known_answers.py``
    is the de-facto pin and is labelled as such.
"""
