"""TEST STUB, not the upstream package: just enough of ``mamba_ssm``'s import surface for
``bench.py --impl reference`` to reach its timing loop on the CPU box (tests/test_bench_reference_arm.py).
Only ever on ``sys.path`` inside that test's subprocess."""
