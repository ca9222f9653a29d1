"""oracle/ -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU restatement of the acoss Serra09 all-pairs scoring path
(benchmarking/CRPUtils.py, benchmarking/SequenceAlignment.c,
benchmarking/Serra09.py:158-196, benchmarking/CoverAlgorithm.py:166-183).

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import this package, and only as
the checker (or as the CPU arm being timed), never as part of the product path.

Parity pinning: the reference ships NO tests, golden vectors or fixtures for
this path.  The oracle is therefore pinned against
outputs of the reference code itself run in the build container
(``oracle/make_golden.py`` imports /root/reference/benchmarking/CRPUtils.py and
the compiled reference SequenceAlignment.c and writes ``tests/golden/*.npz``).
"""
