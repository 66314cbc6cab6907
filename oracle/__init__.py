"""CPU oracle for the learned-SPAI solve path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import anything from here; the product (``learningsparsepreconditioner4gpu_b200``) never does.
"""
